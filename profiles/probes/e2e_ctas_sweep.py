#!/usr/bin/env python
"""Probe (development tool): end-to-end ceres_render() into a page-locked frame vs persistent CTAs per SM of the band
kernels (fewer CTAs per kernel = two bands' kernels co-reside, so band 0's shadow wave does not have to wait for band 1's
whole primary wave) x number of row bands."""
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
hb = cr.PinnedBuffer((H, W, 3), np.float32)
ref = scene.render(cam, sun, W, H, flags=flags)[0].copy()
for sub in (8, 6, 4):
    for ctas in (0, 6, 5, 4, 3, 2):
        scene.set_tuning(subframes=sub, ctas_per_sm=ctas)
        for _ in range(3):
            scene.render(cam, sun, W, H, flags=flags, pixels=hb.array)
        ok = np.array_equal(hb.array.view(np.uint32), ref.view(np.uint32))
        ts = []
        for _ in range(40):
            t0 = time.perf_counter()
            scene.render(cam, sun, W, H, flags=flags, pixels=hb.array)
            ts.append(time.perf_counter() - t0)
        print(f"{sub} bands, {ctas or 'default'} CTAs/SM: median {np.median(ts)*1e3:.3f} ms  min {min(ts)*1e3:.3f} ms  frame_ok={ok}", flush=True)
