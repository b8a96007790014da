#!/usr/bin/env python
"""Histogram of node-pair steps per ray (statistics build of the shipped kernels) and what it implies for SIMD
efficiency: with whole-warp refill a warp of 32 consecutive rays costs max(steps) per lane."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import bench
import ceres_raytracer_b200 as cr

wl = sys.argv[1] if len(sys.argv) > 1 else "dragon_4k_shadow"
pod, cam, sun, W, H, flags, scene = bench.load_workload(wl, device=0)
if scene is None:
    scene = cr.Scene.from_pod(pod)
dev = torch.device("cuda", 0)
px = torch.empty(H * W * 3, dtype=torch.float32, device=dev)
cnt = torch.zeros(2, dtype=torch.int64, device=dev)
scene.set_stats(True)
scene.render_device(cam, sun, W, H, flags=flags, pixels=px, counters=cnt)
torch.cuda.synchronize()
hp, hs = scene.step_histogram()
for name, h in (("primary (rays inside the culling rectangle)", hp), ("shadow (any-hit)", hs)):
    n = int(h.sum())
    mid = (np.arange(64) * 4 + 2).astype(np.float64)
    mean = float((h * mid).sum() / max(n, 1))
    cum = np.cumsum(h) / max(n, 1)
    print(f"{wl} {name}: rays {n}, mean steps ~{mean:.1f}, p50 <{4*(int(np.searchsorted(cum, .5))+1)}, p90 <{4*(int(np.searchsorted(cum, .9))+1)}, p99 <{4*(int(np.searchsorted(cum, .99))+1)}")
    print("  bins of 4 steps:", " ".join(str(int(x)) for x in h[: int(np.max(np.nonzero(h)[0])) + 1]))
    for K in (8, 16, 24, 32, 48):
        sl = np.ceil((mid + 2) / K) * K
        print(f"  time-sliced at K={K}: utilisation bound {float((h * mid).sum() / (h * sl).sum()):.2f}")
