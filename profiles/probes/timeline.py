#!/usr/bin/env python
"""Per-warp timeline of the frame kernel for one rank's share of the headline frame on one GPU:
when the warps leave the primary phase, when they end, and how the work was spread.

    python profiles/probes/timeline.py [--workload dragon_4k_shadow] [--shards 1,8] [--share 1,2]
"""
import argparse
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="dragon_4k_shadow")
    ap.add_argument("--shards", default="1,8")
    ap.add_argument("--share", default="1,2")
    ap.add_argument("--rank", type=int, default=0)
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
    for n in [int(x) for x in args.shards.split(",")]:
        for share in [int(x) for x in args.share.split(",")]:
            shard = (cr.configs.TILE_BLOCK_ROWS, args.rank % n, n) if n > 1 else None
            fl = flags | (cr.COMPACT_ROWS if n > 1 else 0)
            scene.set_tuning(share=share)
            scene.set_stats(True)
            scene.render_device(cam, sun, W, H, flags=fl, shard=shard, pixels=d_px, counters=d_cnt)
            torch.cuda.synchronize()
            st = scene.get_stats()
            scene.set_stats(False)
            scene.set_timeline(True)
            for _ in range(3):
                scene.render_device(cam, sun, W, H, flags=fl, shard=shard, pixels=d_px, counters=d_cnt)
            torch.cuda.synchronize()
            tl = scene.timeline().astype(np.float64)
            scene.set_timeline(False)
            t0 = tl[:, 0].min()
            b, p, e = (tl[:, 0] - t0) / 1e3, (tl[:, 1] - t0) / 1e3, (tl[:, 2] - t0) / 1e3
            q = lambda a: " ".join(f"{np.percentile(a, x):7.1f}" for x in (0, 10, 50, 90, 99, 100))
            print(f"N {n} rank {args.rank % n} share {share}: warps {len(tl)}  S_shadow {st['S_shadow']} T_shadow {st['T_shadow']}")
            print(f"   percentiles 0/10/50/90/99/100 [us]   begin {q(b)}")
            print(f"                                  primary end {q(p)}")
            print(f"                                          end {q(e)}")
            print(f"   tiles per warp {q(tl[:, 3])}   chunks per warp {q(tl[:, 4])}")
            print(f"   longest tile per warp [us] {q(tl[:, 5] / 1e3)}   longest chunk [us] {q(tl[:, 8] / 1e3)}   wait for chunks [us] {q(tl[:, 10] / 1e3)}")
            k = int(np.argmax(tl[:, 1]))
            print(f"   last warp out of the primary phase: tiles {int(tl[k, 3])}, longest tile {tl[k, 5] / 1e3:.1f} us (item {int(tl[k, 6])}), last tile claimed at {(tl[k, 7] - t0) / 1e3:.1f} us, left at {p[k]:.1f} us")
            k = int(np.argmax(tl[:, 2]))
            print(f"   last warp out of the kernel: chunks {int(tl[k, 4])}, longest chunk {tl[k, 8] / 1e3:.1f} us (chunk {int(tl[k, 9])}), waited {tl[k, 10] / 1e3:.1f} us, primary end {p[k]:.1f} us, end {e[k]:.1f} us")
            # how many warps are still running at time t
            for t in np.linspace(0, e.max(), 11)[1:]:
                print(f"   t = {t:6.1f} us: {int((p > t).sum()):5d} warps in the primary phase, {int(((p <= t) & (e > t)).sum()):5d} in the shadow phase")
    scene.set_tuning()


if __name__ == "__main__":
    main()
