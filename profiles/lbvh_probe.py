#!/usr/bin/env python
"""GPU BVH build (ceres_build_lbvh): build time, tree shape, and what the linear BVH costs at render time compared
with the reference's SweepSah tree of the same triangles (same frame; hit ids are compared)."""
import os, sys, time
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
n = pod["triangles"].shape[0]
t0 = time.perf_counter()
nodes, idx, ms = cr.build_lbvh(pod["triangles"])
t_call = time.perf_counter() - t0
nodes, idx, ms = cr.build_lbvh(pod["triangles"])          # second call: no context / allocator warm-up
print(f"{wl}: {n} triangles -> {nodes.shape[0]} nodes; build kernels {ms:.3f} ms (first call incl. transfers {t_call*1e3:.1f} ms)")
lb = dict(pod, nodes=nodes, prim_indices=idx)
t0 = time.perf_counter()
s2 = cr.Scene.from_pod(lb)
print(f"  ceres_scene_create (host flatten + upload) {1e3*(time.perf_counter()-t0):.1f} ms")
t0 = time.perf_counter()
s3 = cr.Scene.from_lbvh(pod["triangles"], pod["norms"])
print(f"  ceres_scene_create_lbvh (upload + build + device-side scene, tree never on the host) {1e3*(time.perf_counter()-t0):.1f} ms, build kernels {s3.build_ms:.3f} ms")
s3.close()
dev = torch.device("cuda", 0)
px = torch.empty(H * W * 3, dtype=torch.float32, device=dev)
ids = [torch.empty(H * W, dtype=torch.int32, device=dev) for _ in range(2)]
cnt = torch.zeros(2, dtype=torch.int64, device=dev)
for name, sc, d_ids in (("reference-built tree", scene, ids[0]), ("GPU linear BVH", s2, ids[1])):
    for _ in range(3):
        sc.render_device(cam, sun, W, H, flags=flags, pixels=px, ids=d_ids, counters=cnt)
    torch.cuda.synchronize()
    sc.set_profiling(True)
    for _ in range(10):
        sc.render_device(cam, sun, W, H, flags=flags, pixels=px, counters=cnt)
    torch.cuda.synchronize()
    k, f = sc.kernel_ms()
    sc.set_profiling(False)
    print(f"  {name}: {k['total']/f:.3f} ms/frame (primary {k['primary']/f:.3f}, shadow {k['shadow']/f:.3f}); counters {cnt.cpu().tolist()}")
a, b = ids[0].cpu().numpy(), ids[1].cpu().numpy()
print(f"  hit ids differing between the two trees: {(a != b).sum()} of {a.size} (exact-t ties and edge cases may legitimately differ)")
