#!/usr/bin/env python
"""Kernel A/B bench (development tool): per-kernel CUDA-event times of the frame kernels for each tuning variant, with
the hit-id hash checked against the golden value so that a fast-but-wrong variant is caught in the same run.

    python profiles/kbench.py [--workload dragon_4k_shadow] [--variants 0,1,2] [--frames 20] [--ctas 0]
"""
import argparse
import hashlib
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="dragon_4k_shadow")
    ap.add_argument("--variants", default="0")
    ap.add_argument("--frames", type=int, default=20)
    ap.add_argument("--ctas", default="0")
    ap.add_argument("--refill", default="0")
    ap.add_argument("--block", type=int, default=0)
    ap.add_argument("--records", default="0", help="frame kernel: records per shadow work item (8/16/32)")
    ap.add_argument("--sub", default="1")
    ap.add_argument("--chunk", default="0")
    ap.add_argument("--order", default="0", help="frame kernel: shadow child order(s), 0 default / 1 near-first / 2 left-first")
    ap.add_argument("--share", default="0", help="frame kernel: shadow-subtree sharing, 0 default / 1 off / 2 on; 2:P:M = on with period P, min idle lanes M")
    args = ap.parse_args()
    import torch
    import ceres_raytracer_b200 as cr
    import bench
    pod, cam, sun, W, H, flags, scene = bench.load_workload(args.workload, device=0)
    name = cr.configs.WORKLOADS[args.workload][0]
    golden = json.load(open(os.path.join(ROOT, "tests", "golden", "frame_hashes.json")))
    key = f"{name}_{W}x{H}_{'primary' if flags & 1 else 'shadow'}"
    if scene is None:
        scene = cr.Scene.from_pod(pod)
    dev = torch.device("cuda", 0)
    d_px = torch.empty(H * W * 3, dtype=torch.float32, device=dev)
    d_ids = torch.empty(H * W, dtype=torch.int32, device=dev)
    d_cnt = torch.zeros(2, dtype=torch.int64, device=dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    ref_px = None
    ref_ids = None
    for ctas, refill, sub, chunk, order, share in [(int(c), int(r), int(q), int(h), int(o), sh) for c in args.ctas.split(",") for r in args.refill.split(",") for q in args.sub.split(",") for h in args.chunk.split(",") for o in args.order.split(",") for sh in args.share.split(",")]:
      for recs in [int(a) for a in args.records.split(",")]:
        shp = [int(x) for x in share.split(":")] + [0, 0]
        for v in [int(x) for x in args.variants.split(",")]:
            if v != 0 and os.environ.get("CERES_B200_EXPERIMENTS", "") in ("", "0"):
                print(f"variant {v}: skipped (wave pipelines need CERES_B200_EXPERIMENTS=1)", flush=True)
                continue
            scene.set_tuning(variant=v, ctas_per_sm=ctas, refill_min=refill, primary_block=args.block, shadow_block=args.block if v else recs, subframes=sub, chunk=chunk, shadow_order=order, share=shp[0], share_period=shp[1], share_min_idle=shp[2])
            d_px.fill_(float("nan"))
            scene.render_device(cam, sun, W, H, flags=flags, pixels=d_px, ids=d_ids, counters=d_cnt)
            torch.cuda.synchronize()
            ids = d_ids.cpu().numpy().view(np.uint32).reshape(H, W)
            ok = hashlib.sha256(ids.tobytes()).hexdigest() == golden[key]["ids_sha256"] if key in golden else None
            if key not in golden:       # no recorded hash (synthetic scenes): ids must equal the first variant's
                if ref_ids is None:
                    ref_ids = ids.copy()
                ok = "same-as-first" if np.array_equal(ids, ref_ids) else False
            cnt = tuple(d_cnt.cpu().tolist())
            cnt_ok = cnt == (golden[key]["stats"]["rays"], golden[key]["stats"]["hits"]) if key in golden else None
            px = d_px.cpu().numpy()
            if ref_px is None:
                ref_px = px
            same_px = bool(np.array_equal(px.view(np.uint32), ref_px.view(np.uint32)))
            for _ in range(3):
                scene.render_device(cam, sun, W, H, flags=flags, pixels=d_px, counters=d_cnt)
            torch.cuda.synchronize()
            scene.set_profiling(True)
            for _ in range(args.frames):
                flush.fill_(1)
                scene.render_device(cam, sun, W, H, flags=flags, pixels=d_px, counters=d_cnt)
            torch.cuda.synchronize()
            ms, n = scene.kernel_ms()
            scene.set_profiling(False)
            parts = "  ".join(f"{k} {ms[k]/n*1e3:7.1f} us" for k in ms)
            print(f"variant {v} ctas/SM {ctas} refill {refill} subframes {sub} chunk {chunk} order {order} share {share} records {recs}: {parts}   ids_ok={ok} counters_ok={cnt_ok} "
                  f"pixels_same_as_first={same_px}", flush=True)
    scene.set_tuning()


if __name__ == "__main__":
    main()
