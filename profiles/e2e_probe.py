#!/usr/bin/env python
"""Development probe: end-to-end ceres_render() time into a page-locked host frame vs band size (run once per
CERES_HOST_FILL_THREADS value: the pool size is fixed at first use)."""
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
for variant in (0, 2):
    for band in (1 << 20, 2 << 20, 3 << 20, 6 << 20, 12 << 20, 64 << 20):
        scene.set_tuning(variant=variant, band_bytes=band)
        for _ in range(3):
            scene.render(cam, sun, W, H, flags=flags, pixels=hb.array)
        ts = []
        for _ in range(20):
            t0 = time.perf_counter()
            scene.render(cam, sun, W, H, flags=flags, pixels=hb.array)
            ts.append(time.perf_counter() - t0)
        print(f"fill threads {os.environ.get('CERES_HOST_FILL_THREADS', 'default')} variant {variant} band {band >> 20:3d} MiB: "
              f"median {np.median(ts)*1e3:.3f} ms  min {min(ts)*1e3:.3f} ms", flush=True)
