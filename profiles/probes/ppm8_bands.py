#!/usr/bin/env python
"""Probe (development tool): ceres_render_ppm8 / ppm16 into page-locked memory vs the number of row bands."""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import ceres_raytracer_b200 as cr

name, cam, sun, W, H, flags = cr.configs.WORKLOADS["dragon_4k_shadow"]
sun = np.asarray(sun, np.float32)
with np.load(os.path.join(ROOT, "tests", "golden", name + "_scene.npz")) as z:
    pod = {k: z[k] for k in z.files}
scene = cr.Scene.from_pod(pod)
h8, h16 = cr.PinnedBuffer((H, W, 3), np.uint8), cr.PinnedBuffer((H, W, 3), np.uint16)
for sub in (1, 2, 3, 4, 6, 8):
    scene.set_tuning(subframes=sub)
    row = []
    for fn, hb in ((scene.render_ppm8, h8), (scene.render_ppm16, h16)):
        for _ in range(3):
            fn(cam, sun, W, H, flags=flags, out=hb.array)
        ts = []
        for _ in range(40):
            t0 = time.perf_counter()
            fn(cam, sun, W, H, flags=flags, out=hb.array)
            ts.append(time.perf_counter() - t0)
        row.append(np.median(ts) * 1e3)
    print(f"{sub} bands: ppm8 {row[0]:.3f} ms  ppm16 {row[1]:.3f} ms", flush=True)
