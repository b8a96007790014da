#!/usr/bin/env python
"""Probe (development tool): device timeline of the host path's row bands (CERES_TRACE_TIMING=1 makes the library print
"kernels done / copy done" per band) next to the host-clock time of the whole call."""
import os, sys, time
os.environ["CERES_TRACE_TIMING"] = "1"
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import numpy as np
import bench
import ceres_raytracer_b200 as cr

pod, cam, sun, W, H, flags, scene = bench.load_workload("dragon_4k_shadow", device=0)
scene = scene or cr.Scene.from_pod(pod)
host = cr.PinnedBuffer((H, W, 3))
for _ in range(4):
    t0 = time.perf_counter()
    scene.render(cam, sun, W, H, flags=flags, pixels=host.array)
    print(f"call {1e6*(time.perf_counter()-t0):.0f} us", file=sys.stderr)
