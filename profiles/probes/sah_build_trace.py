#!/usr/bin/env python
"""Build time of the top-down binned SAH tree (ceres_build_options algorithm 2) level by level (CERES_BUILD_TRACE=1 makes the
library print one line per level on stderr).  python profiles/probes/sah_build_trace.py [workload]"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
os.environ["CERES_BUILD_TRACE"] = "1"
import bench
import ceres_raytracer_b200 as cr

wl = sys.argv[1] if len(sys.argv) > 1 else "grid448_4k_shadow"
pod, cam, sun, W, H, flags, scene = bench.load_workload(wl, device=0)
if scene is not None:
    scene.close()
cr.build_options(int(sys.argv[2]) if len(sys.argv) > 2 else 0)
for k in range(2):
    print(f"--- build {k}", file=sys.stderr, flush=True)
    s = cr.Scene.from_lbvh(pod["triangles"], pod["norms"])
    print(f"{wl}: {pod['triangles'].shape[0]} triangles, variant {os.environ.get('CERES_B200_VARIANT', 'product')}: build kernels {s.build_ms:.3f} ms, {s.node_count} nodes, stack depth {s.stack_depth}", flush=True)
    s.close()
