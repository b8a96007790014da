#!/usr/bin/env python
"""Probe (development tool): how fast does the copy engine drain the 4K frame's screen rectangle (29.7 MB, 2-D D2H into
page-locked memory) while the trace kernels of following frames run on another stream -- one copy, or the same bytes
split over 2 / 4 / 8 streams (row bands)?"""
import ctypes as C
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import torch
import bench
import ceres_raytracer_b200 as cr

rt = C.CDLL("libcudart.so.12")
rt.cudaMemcpy2DAsync.argtypes = [C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t, C.c_size_t, C.c_size_t, C.c_int, C.c_void_p]
pod, cam, sun, W, H, flags, scene = bench.load_workload("dragon_4k_shadow", device=0)
scene = scene or cr.Scene.from_pod(pod)
dev = torch.device("cuda", 0)
px = torch.zeros(H * W * 3, dtype=torch.float32, device=dev)
cnt = torch.zeros(2, dtype=torch.int64, device=dev)
host = cr.PinnedBuffer((H, W, 3))
x0, y0, rw, rh = 876, 489, 2092, 1182
work = torch.cuda.Stream()
copies = [torch.cuda.Stream() for _ in range(8)]

def drain(nstreams):
    rows = (rh + nstreams - 1) // nstreams
    for k in range(nstreams):
        a, b = y0 + k * rows, min(y0 + rh, y0 + (k + 1) * rows)
        if b <= a:
            continue
        off = (a * W + x0) * 12
        rc = rt.cudaMemcpy2DAsync(host.array.ctypes.data + off, W * 12, px.data_ptr() + off, W * 12, rw * 12, b - a, 2, C.c_void_p(copies[k].cuda_stream))
        assert rc == 0, rc

for _ in range(3):
    scene.render_device(cam, sun, W, H, flags=flags, pixels=px, counters=cnt, stream=work.cuda_stream)
torch.cuda.synchronize()
for load in (0, 1):
    for ns in (1, 2, 4, 8):
        best = 1e9
        for rep in range(5):
            torch.cuda.synchronize()
            if load:
                for _ in range(6):      # ~3 ms of traversal kernels in flight
                    scene.render_device(cam, sun, W, H, flags=flags, pixels=px, counters=cnt, stream=work.cuda_stream)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            for s in copies[:ns]:
                pass
            e0.record(copies[0])
            for s in copies[1:ns]:
                s.wait_event(e0)
            drain(ns)
            for s in copies[1:ns]:
                ev = torch.cuda.Event(); ev.record(s); copies[0].wait_event(ev)
            e1.record(copies[0])
            torch.cuda.synchronize()
            best = min(best, e0.elapsed_time(e1))
        print(f"{'frames running' if load else 'idle GPU     '}: {ns} copy stream(s): {best:.3f} ms = {rw*rh*12/best/1e6:.1f} GB/s")
