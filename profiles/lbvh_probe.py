#!/usr/bin/env python
"""GPU BVH build (ceres_build_lbvh / ceres_scene_create_lbvh): build time and what each tree costs at render time compared
with the reference's SweepSah tree of the same triangles (same frame; hit ids are compared).
    python profiles/lbvh_probe.py [workload] [radius,radius,...]"""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import bench
import ceres_raytracer_b200 as cr

wl = sys.argv[1] if len(sys.argv) > 1 else "dragon_4k_shadow"
# configurations: algorithm:radius:max_leaf:triangle_cost, comma separated
cfgs = [tuple(float(v) if k == 3 else int(v) for k, v in enumerate(c.split(":"))) for c in
        (sys.argv[2] if len(sys.argv) > 2 else "1:0:1:1,1:0:8:1.5,3:16:1:1,3:8:8:1.5,2:0:1:1,2:0:8:1,2:0:8:1.5").split(",")]
pod, cam, sun, W, H, flags, scene = bench.load_workload(wl, device=0)
if scene is None:
    scene = cr.Scene.from_pod(pod)
n = pod["triangles"].shape[0]
dev = torch.device("cuda", 0)
px = torch.empty(H * W * 3, dtype=torch.float32, device=dev)
cnt = torch.zeros(2, dtype=torch.int64, device=dev)


def frame_ms(sc, d_ids):
    for _ in range(3):
        sc.render_device(cam, sun, W, H, flags=flags, pixels=px, ids=d_ids, counters=cnt)
    torch.cuda.synchronize()
    sc.set_profiling(True)
    for _ in range(10):
        sc.render_device(cam, sun, W, H, flags=flags, pixels=px, counters=cnt)
    torch.cuda.synchronize()
    k, f = sc.kernel_ms()
    sc.set_profiling(False)
    return k["total"] / f


ids_ref = torch.empty(H * W, dtype=torch.int32, device=dev)
t_ref = frame_ms(scene, ids_ref)
print(f"{wl}: {n} triangles; reference-built (SweepSah) tree: {t_ref:.3f} ms/frame, counters {cnt.cpu().tolist()}", flush=True)
for algo, radius, max_leaf, ci in cfgs:
    cr.build_options(algo, radius, max_leaf, ci)
    cr.Scene.from_lbvh(pod["triangles"], pod["norms"]).close()              # warm-up (context, allocator)
    t0 = time.perf_counter()
    s3 = cr.Scene.from_lbvh(pod["triangles"], pod["norms"])
    t_call = time.perf_counter() - t0
    ids = torch.empty(H * W, dtype=torch.int32, device=dev)
    t = frame_ms(s3, ids)
    diff = int((ids.cpu().numpy() != ids_ref.cpu().numpy()).sum())
    name = ("radix tree" if algo == 1 else "top-down binned SAH" if algo in (0, 2) else f"clustering r{radius or 8}") + f", leaves <= {max_leaf} (Ci {ci:g})"
    print(f"  {name:40s} {s3.node_count:9d} nodes: build kernels {s3.build_ms:8.3f} ms, upload + build + scene {t_call*1e3:8.1f} ms, stack depth {s3.stack_depth:2d}; "
          f"{t:.3f} ms/frame = {t / t_ref:.3f} x the reference tree; counters {cnt.cpu().tolist()}; ids differing {diff}", flush=True)
    s3.close()
cr.build_options()
