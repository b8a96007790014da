#!/usr/bin/env python
"""Development probe: end-to-end ceres_render() time into a page-locked host frame vs number of row bands (in-kernel bands by
default; CERES_LAUNCH_BANDS=1 = one launch per band) and of the 8-bit payload path."""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import ceres_raytracer_b200 as cr

name, cam, sun, W, H, flags = cr.configs.WORKLOADS["dragon_4k_shadow"]
sun = np.asarray(sun, np.float32)
with np.load(os.path.join(ROOT, "tests", "golden", name + "_scene.npz")) as z:
    pod = {k: z[k] for k in z.files}
scene = cr.Scene.from_pod(pod)
hb = cr.PinnedBuffer((H, W, 3), np.float32)
hb8 = cr.PinnedBuffer((H, W, 3), np.uint8)
ref, ref_ids, ref_cnt = scene.render(cam, sun, W, H, flags=flags, want_ids=True)
ref = ref.copy()
subs = [int(x) for x in (sys.argv[1] if len(sys.argv) > 1 else "0,1,2,4,6,8,12,16").split(",")]
mode = "bands"
for sub in subs:
    scene.set_tuning(subframes=sub)
    for _ in range(3):
        scene.render(cam, sun, W, H, flags=flags, pixels=hb.array)
    ok = np.array_equal(hb.array.view(np.uint32), ref.view(np.uint32))
    ts, t8 = [], []
    for _ in range(30):
        t0 = time.perf_counter()
        r = scene.render(cam, sun, W, H, flags=flags, pixels=hb.array)
        ts.append(time.perf_counter() - t0)
        ok = ok and r[2] == ref_cnt
    tr = []
    scene.render(cam, sun, W, H, flags=flags | cr.FRAME_RETAINED, pixels=hb.array)
    for _ in range(30):
        t0 = time.perf_counter()
        r = scene.render(cam, sun, W, H, flags=flags | cr.FRAME_RETAINED, pixels=hb.array)
        tr.append(time.perf_counter() - t0)
    ok = ok and r[2] == ref_cnt and np.array_equal(hb.array.view(np.uint32), ref.view(np.uint32))
    for _ in range(30):
        t0 = time.perf_counter()
        scene.render_ppm8(cam, sun, W, H, flags=flags, out=hb8.array)
        t8.append(time.perf_counter() - t0)
    print(f"{mode}, {sub} bands: float frame median {np.median(ts)*1e3:.3f} ms  min {min(ts)*1e3:.3f} ms | retained {np.median(tr)*1e3:.3f} / {min(tr)*1e3:.3f} ms | "
          f"ppm8 median {np.median(t8)*1e3:.3f} ms  min {min(t8)*1e3:.3f} ms  frame_ok={ok}", flush=True)
