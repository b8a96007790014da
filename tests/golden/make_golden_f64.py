#!/usr/bin/env python
"""Writes tests/golden/f64_bunny.npz: the bunny loaded and built by the reference with Scalar = double (`anim -d`,
anim.cpp:154-155; obj_norms.hpp:120, rotate y -145, SweepSah + NodeLayoutOptimizer) and frames rendered from it by the
reference's own code (oracle/_ref/libceres_ref64_strict.so).  Needs /root/reference (this container only); the GPU box
and the CPU tests read the committed file.

    scene   nodes (64-byte Bvh<double>::Node), prim_indices, triangles float64[n, 12], norms float64[n, 9]
    frames  two cameras x {shadows, primary only}: ids uint32, state uint8, rgb (float32: render.hpp's colour
            accumulators are float also when Scalar = double, so the widened values are exact), counters
"""
import os
import sys
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from oracle import f64  # noqa: E402

W, H = 480, 270
CAMS = [((0, .1, -.3), (0, 0, 1), (0, 1, 0), 60.0), ((.12, .18, -.22), (-.35, -.25, 1), (0, 1, 0), 45.0)]
SUN = np.array([-50., -20., 0.])

rs = f64.RefScene64.load_obj("/root/reference/data/bunny.obj", 1, -145.0)
out = {k: v for k, v in rs.pod().items()}
out["size"] = np.array([W, H], np.int64)
out["sun"] = SUN
out["cams"] = np.stack([f64.cam_array(*c) for c in CAMS])
for k, c in enumerate(CAMS):
    cam = f64.cam_array(*c)
    plain, cnt_plain, _ = rs.render(cam, SUN, W, H)
    for flags in (0, 1):
        px, ids, st, cnt = rs.render_ex(cam, SUN, W, H, flags)
        if flags == 0:
            assert np.array_equal(px.view(np.uint64), plain.view(np.uint64)) and cnt == cnt_plain
        assert np.array_equal(px, px.astype(np.float32).astype(np.float64))
        out[f"ids_{k}_{flags}"], out[f"state_{k}_{flags}"] = ids, st
        out[f"rgb_{k}_{flags}"], out[f"counters_{k}_{flags}"] = px.astype(np.float32), np.array(cnt, np.uint64)
        print(f"camera {k} flags {flags}: counters {cnt}, hit pixels {(ids != f64.MISS).sum()}")
path = os.path.join(HERE, "f64_bunny.npz")
np.savez_compressed(path, **out)
print(path, os.path.getsize(path), "bytes")
