#!/usr/bin/env python
"""Development probe: end-to-end ceres_render() time into a page-locked host frame vs number of row bands
(sub-frames).  Run once with CERES_UNIFORM_BANDS=1 and once without to compare equal and progressive bands."""
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
ref = scene.render(cam, sun, W, H, flags=flags)[0].copy()
for sub in (1, 3, 4, 5, 6, 7, 8):
    scene.set_tuning(subframes=sub)
    for _ in range(3):
        scene.render(cam, sun, W, H, flags=flags, pixels=hb.array)
    ok = np.array_equal(hb.array.view(np.uint32), ref.view(np.uint32))
    ts = []
    for _ in range(30):
        t0 = time.perf_counter()
        scene.render(cam, sun, W, H, flags=flags, pixels=hb.array)
        ts.append(time.perf_counter() - t0)
    print(f"{'uniform' if os.environ.get('CERES_UNIFORM_BANDS') else 'progressive'} bands, {sub} sub-frames: "
          f"median {np.median(ts)*1e3:.3f} ms  min {min(ts)*1e3:.3f} ms  frame_ok={ok}", flush=True)
