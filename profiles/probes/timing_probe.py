import os, sys, numpy as np
sys.path.insert(0, "/root/repo")
import ceres_raytracer_b200 as cr
name, cam, sun, W, H, flags = cr.configs.WORKLOADS["dragon_4k_shadow"]
pod = dict(np.load("/root/repo/tests/golden/dragon_scene.npz"))
scene = cr.Scene.from_pod(pod); hb = cr.PinnedBuffer((H, W, 3), np.float32)
for band in (6 << 20, 12 << 20, 64 << 20):
    scene.set_tuning(band_bytes=band)
    for _ in range(5): scene.render(cam, np.asarray(sun, np.float32), W, H, pixels=hb.array)
