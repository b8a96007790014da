#!/usr/bin/env python
"""Stress of the first frames of FRESH scenes (allocation + initialisation of the per-frame state, then immediately two frames in
flight from two threads): 60 x (create scene, render a 6-frame 4K sequence, compare every frame, destroy).  A queue wait that
expires (the legacy-stream memset of ready[] racing a frame kernel on a non-blocking stream, fixed in round 2) raises here."""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import ceres_raytracer_b200 as cr
import bench

pod = bench.load_scene_pod("dragon")
c = cr.configs.DRAGON_CAMERA
sun = np.array([-50, -20, 0], np.float32)
W, H, n = 3840, 2160, 6
cams = np.tile(np.array(list(c.eye) + list(c.dir) + list(c.up) + [c.fov] + list(sun), np.float32), (n, 1))
ref = None
t0 = time.time()
for it in range(int(sys.argv[1]) if len(sys.argv) > 1 else 60):
    with cr.Scene.from_pod(pod) as s:
        px, cnt = s.render_sequence(cams, W, H)
        if ref is None:
            ref = px[0].copy()
        for k in range(n):
            assert np.array_equal(px[k].view(np.uint32), ref.view(np.uint32)), (it, k)
        if it % 3 == 0:
            out, _ = s.render_sequence_payload(cams[:3], W, H, bytes_per_channel=1)
            assert np.array_equal(out[0], out[2])
print(f"fresh-scene stress: ok, {time.time() - t0:.1f} s")
