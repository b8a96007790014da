#!/usr/bin/env python
"""Compute-only time of ONE rank's share of the headline frame on one GPU (no peer stores, no cross-rank waits):
render_device with shard (16, k, N) into a local packed band.  Compared with kernels_ms_per_rank of the N-rank bench it
tells how much of a peer frame is NVLink stores / waiting rather than tracing.

    python profiles/probes/shard_alone.py [--workload dragon_4k_shadow] [--frames 20]
"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="dragon_4k_shadow")
    ap.add_argument("--frames", type=int, default=20)
    ap.add_argument("--shards", default="1,2,4,8,16")
    ap.add_argument("--all-ranks", action="store_true")
    ap.add_argument("--block-rows", default="16")
    ap.add_argument("--share", type=int, default=0)
    ap.add_argument("--ctas", default="0")
    ap.add_argument("--chunk", default="0", help="records per shadow work item (8/16/32; 0 = default)")
    args = ap.parse_args()
    import torch
    import ceres_raytracer_b200 as cr
    import bench
    pod, cam, sun, W, H, flags, scene = bench.load_workload(args.workload, device=0)
    if scene is None:
        scene = cr.Scene.from_pod(pod)
    dev = torch.device("cuda", 0)
    d_px = torch.empty(H * W * 3, dtype=torch.float32, device=dev)
    d_cnt = torch.zeros(2, dtype=torch.int64, device=dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    for br, n, ctas, chunk in [(int(b), int(x), int(c), int(h)) for b in args.block_rows.split(",") for x in args.shards.split(",") for c in args.ctas.split(",") for h in args.chunk.split(",")]:
      if True:
        scene.set_tuning(share=args.share, ctas_per_sm=ctas, shadow_block=chunk)
        for k in (range(n) if args.all_ranks else sorted({0, n - 1})):
            shard = (br, k, n) if n > 1 else None
            fl = flags | (cr.COMPACT_ROWS if n > 1 else 0)
            for _ in range(3):
                scene.render_device(cam, sun, W, H, flags=fl, shard=shard, pixels=d_px, counters=d_cnt)
            torch.cuda.synchronize()
            for flushed in (True,):
                scene.set_profiling(True)
                for _ in range(args.frames):
                    if flushed:
                        flush.fill_(1)
                    scene.render_device(cam, sun, W, H, flags=fl, shard=shard, pixels=d_px, counters=d_cnt)
                torch.cuda.synchronize()
                ms, cnt = scene.kernel_ms()
                scene.set_profiling(False)
                print(f"block rows {br:2d} ctas/SM {ctas} chunk {chunk:2d} N {n:2d} rank {k:2d} {'L2 flushed' if flushed else 'warm      '}: " +
                      "  ".join(f"{q} {ms[q] / cnt * 1e3:7.1f} us" for q in ms), flush=True)


if __name__ == "__main__":
    main()
