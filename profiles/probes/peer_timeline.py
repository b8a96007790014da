#!/usr/bin/env python
"""Per-rank timeline of one peer frame (ceres_render_peer) on N GPUs: run under torchrun.  Every rank reports, relative to
its own kernel's first warp: when the gate opened, when the primary phase and the kernel ended, and the CUDA-event time
of its kernel; rank 0 prints one line per rank.

    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 --master-port 29533 profiles/probes/peer_timeline.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)


def main():
    import torch
    import torch.distributed as dist
    import ceres_raytracer_b200 as cr
    import bench
    from ceres_raytracer_b200.shard import PeerFrame
    world, rank, local = int(os.environ["WORLD_SIZE"]), int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    pod, cam, sun, W, H, flags, scene = bench.load_workload("dragon_4k_shadow", device=local)
    if scene is None:
        scene = cr.Scene.from_pod(pod, device=local)
    share = int(os.environ.get("SHARE", "0"))
    scene.set_tuning(share=share, shadow_block=int(os.environ.get("CHUNK", "0")))
    pf = PeerFrame(scene, W, H, rank, world)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    for _ in range(5):
        pf.render(cam, sun, flags=flags)
    torch.cuda.synchronize(); dist.barrier()
    rows = []
    for flushed in (1, 0):
        scene.set_timeline(True)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ms = []
        for _ in range(10):
            if flushed:
                flush.fill_(1)
            a.record(); pf.render(cam, sun, flags=flags); b.record()
            torch.cuda.synchronize()
            ms.append(a.elapsed_time(b))
        tl = scene.timeline().astype(np.float64)
        scene.set_timeline(False)
        t0 = tl[:, 0].min()
        row = [float(np.median(ms)) * 1e3, (tl[:, 11].max() - t0) / 1e3, (np.median(tl[:, 1]) - t0) / 1e3, (tl[:, 1].max() - t0) / 1e3,
               (np.median(tl[:, 2]) - t0) / 1e3, (tl[:, 2].max() - t0) / 1e3, float(tl[:, 5].max()) / 1e3, float(tl[:, 8].max()) / 1e3]
        out = [None] * world
        dist.all_gather_object(out, row)
        if rank == 0:
            print(f"N {world} share {share} {'L2 flushed' if flushed else 'back to back'}: per rank [us]  kernel(event)  gate open  primary end p50/max  end p50/max  longest tile  longest chunk")
            for r, q in enumerate(out):
                print(f"   rank {r}: " + "  ".join(f"{x:7.1f}" for x in q), flush=True)
        dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
