#!/usr/bin/env python
"""Scalar = double on N GPUs: one 4K double frame, row-block shards (ceres_render_f64_shard), every rank widening its rows
straight into ONE host frame shared by the ranks of the box (SharedHostFrame, float64).  Run under torchrun; N = 1 works too.
Rank 0 checks the assembled frame and the summed counters against its own unsharded ceres_render_f64."""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import torch.distributed as dist
import ceres_raytracer_b200 as cr
from ceres_raytracer_b200.shard import SharedHostFrame
from oracle import f64

rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
if world > 1:
    dist.init_process_group("gloo")
with np.load(os.path.join(ROOT, "tests", "golden", "dragon_scene.npz")) as z:
    pod = f64.widen({k: z[k] for k in z.files})
c = cr.configs.DRAGON_CAMERA
cam, sun = f64.cam_array(c.eye, c.dir, c.up, c.fov), np.array([-50., -20., 0.])
for W, H in ((1920, 1080), (3840, 2160)):
    with cr.SceneF64.from_pod(pod, device=local) as s:
        hf = SharedHostFrame(f"ceres_f64_{os.environ.get('MASTER_PORT', '0')}_{W}", (H, W, 3), rank, world, dtype=np.float64)
        try:
            for _ in range(3):
                cnt = hf.render(s, cam, sun, W, H)
            ts = []
            for _ in range(10):
                if world > 1:
                    dist.barrier()
                t0 = time.perf_counter()
                cnt = hf.render(s, cam, sun, W, H)
                ts.append((time.perf_counter() - t0) * 1e3)
            if rank == 0:
                ref = np.full((H, W, 3), np.nan)
                _, _, rcnt = s.render(cam, sun, W, H, pixels=ref)
                t1 = []
                for _ in range(5):
                    t0 = time.perf_counter(); s.render(cam, sun, W, H, pixels=ref); t1.append((time.perf_counter() - t0) * 1e3)
                same = np.array_equal(hf.array.view(np.uint64), ref.view(np.uint64))
                print(f"N={world} {W}x{H} double frame into the shared host frame: min {min(ts):.3f} ms, median {np.median(ts):.3f} ms per frame; "
                      f"one GPU, unsharded, private pageable frame: min {min(t1):.3f} ms; frame bytes equal {same}, counters {cnt} equal {cnt == rcnt}", flush=True)
            if world > 1:
                dist.barrier()
        finally:
            hf.close()
if world > 1:
    dist.destroy_process_group()
