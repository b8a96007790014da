#!/usr/bin/env python
"""Writes the two scene dumps profiles/sanitize_driver.cpp reads (numpy only): scene.bin from tests/golden/bunny_scene.npz
(float) and scene64.bin from tests/golden/f64_bunny.npz (double).  Layout: u64 node_count, u64 tri_count, nodes, primitive
indices (u64), triangles, vertex normals."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
out = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "gpurun_out")
os.makedirs(out, exist_ok=True)
for src, dst, ft in (("bunny_scene.npz", "scene.bin", np.float32), ("f64_bunny.npz", "scene64.bin", np.float64)):
    with np.load(os.path.join(ROOT, "tests", "golden", src)) as z:
        with open(os.path.join(out, dst), "wb") as f:
            f.write(np.array([len(z["nodes"]), len(z["triangles"])], np.uint64).tobytes())
            f.write(np.ascontiguousarray(z["nodes"]).tobytes())
            f.write(np.ascontiguousarray(z["prim_indices"], np.uint64).tobytes())
            f.write(np.ascontiguousarray(z["triangles"], ft).tobytes())
            f.write(np.ascontiguousarray(z["norms"], ft).tobytes())
print("wrote", out)
