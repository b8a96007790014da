import os, sys, numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import ceres_raytracer_b200 as cr
name, cam, sun, W, H, flags = cr.configs.WORKLOADS["dragon_4k_shadow"]
pod = dict(np.load(os.path.join(os.path.dirname(cr.__file__), "..", "tests", "golden", "dragon_scene.npz")))
scene = cr.Scene.from_pod(pod); hb = cr.PinnedBuffer((H, W, 3), np.float32)
for sub in (1, 2, 3, 4, 6, 8):
    scene.set_tuning(subframes=sub)
    print("subframes", sub, file=sys.stderr, flush=True)
    for _ in range(4): scene.render(cam, np.asarray(sun, np.float32), W, H, pixels=hb.array)
