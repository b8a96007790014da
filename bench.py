#!/usr/bin/env python
"""bench.py -- BASELINE.json's metric (Mrays/s, ms/frame; dragon.obj 3840x2160 with shadow rays) on N B200s.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

One "step" = one frame of the hot path (render.hpp:86-156) over the reference-built dragon scene.  Rays are the
reference's own counter: W*H + primary hits (render.hpp:115,138) = 9 450 006 per frame; `rays_traced` is the number of
rays the GPU actually walks through the BVH (pixels of the scene's conservative screen rectangle + shadow rays).
  value          device-resident throughput: scene in HBM, frame left in HBM (CUDA events on the launch stream; L2 is
                 flushed before every timed frame)
  sustained      >= 2 s of back-to-back frames (no flush; every frame rewrites the 99.5 MB frame buffer) with the SM
                 clock sampled in-process (NVML) every 10 ms
  e2e            the same frame through the reference-facing C-ABI call ceres_render() with a page-locked HOST frame:
                 the device->host copy of the float RGB frame and of the counters is inside the timed region;
                 e2e.floor_ms = host frame bytes / host store bandwidth measured in this run
  e2e_pageable   the same call into plain malloc'ed memory, i.e. what anim.cpp:76 / static.cpp:117 pass
  roofline       the frame kernel: warp instructions (committed ncu capture) against the SMs' issue slots at the
                 measured SM clock, next to its algorithmic bytes against the measured HBM and L2 peaks
  cpu_baseline   the reference's own render() (oracle/_ref, its flags, all host threads) on the same frame
N > 1 (torchrun): the frame is split in interleaved 16-row blocks, one shard per GPU, scene replicated; every rank's
frame kernel stores its pixels straight into rank 0's frame over NVLink (no collective on the data path); time = max
over ranks; scaling is strong (one frame, more GPUs).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOAD = "dragon_4k_shadow"
METRIC = "Mrays/s (dragon.obj 3840x2160, primary + shadow rays)"
UNIT = "Mrays/s"
DATA = "dragon.obj loaded + BVH built by the reference code (tests/golden/dragon_scene.npz); no synthetic substitute needed"


def metric_name(workload, W, H, flags):
    if workload == WORKLOAD:
        return METRIC
    return f"Mrays/s ({workload}: {W}x{H}, {'primary rays only' if flags & 1 else 'primary + shadow rays'})"


def load_scene_pod(name):
    """The reference-loaded, reference-built scene (obj_norms.hpp loader, rotate_triangles, SweepSah + layout
    optimiser), dumped by tests/golden/make_golden.py.  It is INPUT DATA for both arms."""
    p = os.path.join(ROOT, "tests", "golden", name + "_scene.npz")
    if not os.path.exists(p):
        raise SystemExit(f"{p} missing: run tests/golden/make_golden.py where /root/reference exists")
    with np.load(p) as z:
        return {k: z[k] for k in z.files}


def load_workload(workload, device=None):
    """-> (pod, camera, sun, W, H, flags, scene or None).  'grid:<base>:AxBxC' scenes are synthesised from the base
    scene (ceres_raytracer_b200.scenes); all their boxes come from the device refit, so the scene is created here and
    the pod gets the downloaded nodes (input data for the CPU arm)."""
    import ceres_raytracer_b200 as cr
    name, cam, sun, W, H, flags = cr.configs.WORKLOADS[workload]
    if not name.startswith("grid:"):
        return load_scene_pod(name), cam, np.asarray(sun, np.float32), W, H, flags, None
    _, base, dims = name.split(":")
    nx, ny, nz = (int(x) for x in dims.split("x"))
    pod, info = cr.scenes.instanced_grid(load_scene_pod(base), nx, ny, nz)
    cam, sun = cr.scenes.grid_camera(info)
    scene = None
    if device is not None:
        scene = cr.Scene.from_pod(pod, device=device)
        scene.refit()
        pod["nodes"], _ = scene.download()
    else:       # CPU arm only: the oracle's refit (bit-identical boxes, tests/test_gpu_prep.py)
        from oracle import port
        pod["nodes"] = port.refit(pod)
    return pod, cam, np.asarray(sun, np.float32), W, H, flags, scene


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """SM clock + clock-event reasons sampled every `period` seconds from a thread of this process through NVML (the
    nvidia-smi -lms recipe of B200_PROFILING.md misses regions of a few ms); falls back to nvidia-smi when NVML is not
    importable.  mark(tag) labels the samples that follow, so that the summary can tell the timed regions from the rest."""
    NAMES = {"hw_slowdown": 0x8, "hw_thermal_slowdown": 0x40, "sw_thermal_slowdown": 0x20, "sw_power_cap": 0x4}
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index, period=0.01):
        self.index, self.period, self.rows, self.tag, self.stop, self.proc, self.h = index, period, [], "idle", False, None, None
        self.max_mhz = None

    def __enter__(self):
        try:
            import pynvml
            import torch
            pynvml.nvmlInit()
            try:
                uuid = str(torch.cuda.get_device_properties(self.index).uuid)
                self.h = pynvml.nvmlDeviceGetHandleByUUID(("GPU-" + uuid) if not uuid.startswith("GPU-") else uuid)
            except Exception:
                self.h = pynvml.nvmlDeviceGetHandleByIndex(self.index)
            self.nv = pynvml
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
            self.t = threading.Thread(target=self._poll, daemon=True)
            self.t.start()
        except Exception:
            self.h = None
            try:
                self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                              "--format=csv,noheader,nounits", "-lms", "20"],
                                             stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
                self.t = threading.Thread(target=self._read, daemon=True)
                self.t.start()
            except Exception:
                self.proc = None
        return self

    def mark(self, tag):
        self.tag = tag

    def _poll(self):
        nv = self.nv
        while not self.stop:
            try:
                mhz = float(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                try:
                    bits = int(nv.nvmlDeviceGetCurrentClocksEventReasons(self.h))
                except Exception:
                    bits = int(nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h))
                self.rows.append((self.tag, mhz, bits))
            except Exception:
                pass
            time.sleep(self.period)

    def _read(self):
        for line in self.proc.stdout:
            r = [x.strip() for x in line.split(",")]
            try:
                bits = sum(v for (n, v), x in zip(self.NAMES.items(), r[2:6]) if x.lower().startswith("active"))
                self.max_mhz = max(self.max_mhz or 0.0, float(r[1]))
                self.rows.append((self.tag, float(r[0]), bits))
            except Exception:
                continue

    def __exit__(self, *a):
        self.stop = True
        if self.proc:
            time.sleep(0.05)
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except Exception:
                self.proc.kill()

    def summary(self, tags=("timed", "sustained")):
        rows = [r for r in self.rows if r[0] in tags] or self.rows
        sm = [r[1] for r in rows]
        bits = 0
        for r in rows:
            bits |= r[2]
        per_tag = {}
        for t in sorted({r[0] for r in self.rows}):
            v = [r[1] for r in self.rows if r[0] == t]
            per_tag[t] = {"samples": len(v), "sm_mhz": float(np.median(v))}
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": self.max_mhz,
                "reasons": sorted(n for n, v in self.NAMES.items() if bits & v), "samples": len(sm),
                "source": "NVML in-process, every %d ms" % int(self.period * 1e3) if self.h is not None else "nvidia-smi -lms 20",
                "regions": per_tag}


def use_all_host_threads():
    """torchrun exports OMP_NUM_THREADS=1 to its workers; the CPU arm is the reference with all the host threads it can
    use (its `#pragma omp parallel for`, render.hpp:104), so give libgomp the box's cores back."""
    import ctypes
    n = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    os.environ["OMP_NUM_THREADS"] = str(n)
    try:
        ctypes.CDLL("libgomp.so.1").omp_set_num_threads(n)
    except OSError:
        pass
    return n


def cpu_reference_frames(pod, cam, sun, W, H, frames, warmup=1, budget_s=None):
    """Times the reference's own render() on the host cores.  Returns (seconds per frame list, kind, cores, rays, hits)."""
    from oracle import ref, port
    use_all_host_threads()
    if ref.available("fast"):
        s = ref.RefScene.from_pod(pod, "fast")       # the reference's own flags (CMakeLists.txt:5,11-13)
        for _ in range(warmup):
            _, _, _, sec0 = s.render(cam, sun, W, H)
            if budget_s:
                frames = int(min(frames, max(2, budget_s / max(sec0, 1e-3))))
        secs, rays, hits = [], 0, 0
        for _ in range(frames):
            _, rays, hits, sec = s.render(cam, sun, W, H)
            secs.append(sec)
        cores = ref.threads("fast")
        s.close()
        return secs, "reference", cores, rays, hits
    secs, rays, hits = [], 0, 0
    for k in range(warmup + frames):
        t0 = time.perf_counter()
        r = port.render(pod, cam, sun, W, H, 0, want_pixels=True)
        if k >= warmup:
            secs.append(time.perf_counter() - t0)
        rays, hits = r["stats"]["rays"], r["stats"]["hits"]
    return secs, "port", port.threads(), rays, hits


def scene_device_bytes(pod):
    """Bytes of the flattened scene as ceres_scene_create lays it out (64 B per node pair, 48 B triangle + 48 B normals +
    4 B slot id per triangle): computed from the arrays, so the CPU arm can name the same config without a GPU."""
    n_nodes, n_tris = int(len(pod["nodes"])), int(len(pod["triangles"]))
    return max(1, (n_nodes - 1) // 2) * 64 + n_tris * (48 + 48 + 4)


def frame_config(workload, W, H, rays, scene_bytes, sharding, l2):
    """The `config` object of the JSON line; both arms build it here so that their keys AND values agree (the `hits`
    counter is not part of it: the reference's shipped -ffp-contract=fast build counts 1-2 more or fewer than the strict
    arithmetic the GPU reproduces)."""
    return {"workload": workload, "width": W, "height": H, "rays_per_frame": rays,
            "scene_device_bytes": scene_bytes, "sharding": sharding, "l2": l2}


def run_reference_arm(args):
    """--impl reference: the reference's CPU implementation of the path on the host cores (rank 0 only)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    pod, cam, sun, W, H, flags, _ = load_workload(args.workload)
    secs, kind, cores, rays, hits = cpu_reference_frames(pod, cam.as_array(), sun, W, H, frames=args.steps, warmup=args.warmup)
    ms = 1e3 * float(np.sum(secs)) / len(secs)
    val = rays / (ms * 1e-3) / 1e6
    line = {
        "impl": "reference", "metric": metric_name(args.workload, W, H, flags), "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "f32", "data": DATA,
        "config": frame_config(args.workload, W, H, rays, scene_device_bytes(pod), sharding_text(args.gpus), L2_TEXT),
        "counters": {"rays": rays, "hits": hits},
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": cores, "kind": kind,
                         "sample": f"{len(secs)} full {W}x{H} frames, unmodified render() (render.hpp:86-156), "
                                   f"best {min(secs)*1e3:.1f} ms, median {np.median(secs)*1e3:.1f} ms"},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)


L2_TEXT = "flushed (256 MiB fill) before every timed frame"


def sharding_text(world, peer=True):
    if world <= 1:
        return "none"
    import ceres_raytracer_b200.configs as cfg
    return (f"{world} x interleaved {cfg.TILE_BLOCK_ROWS}-row blocks, " +
            ("kernels store into rank 0's frame over NVLink peer memory (no collective)" if peer else "NCCL gather to rank 0"))


def algorithmic_bytes(st, W, rows, lit, traced_px):
    """Algorithmic bytes of one frame from the work the GPU executed (DESIGN.md 'roofline'): 64 B per node-pair step,
    48 B per triangle test, 16 B per queue record written and read, 48 B triangle re-read per shadow record, 36 B normals
    per lit pixel, 12 B per pixel written (frame kernel: the traced rectangle; clear_outside_kernel: everything else)."""
    P, Hh = W * rows, st["hits_primary"]
    return {
        "clear": 12 * (P - traced_px),
        "frame": 64 * (st["S_miss"] + st["S_hit"] + st["S_shadow"]) + 48 * (st["T_miss"] + st["T_hit"] + st["T_shadow"])
                 + 2 * 16 * Hh + 48 * Hh + 36 * lit + 12 * traced_px,
    }


def run_sequence(args):
    """BASELINE.json configs[3]: the 60-frame dragon turntable, frames sharded round-robin over the GPUs (no collective
    on the data path: anim.cpp:82-105 frames share only the read-only scene).  A step = one pass over the sequence."""
    import torch
    import torch.distributed as dist
    import ceres_raytracer_b200 as cr
    from ceres_raytracer_b200.shard import sequence_frames
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    warmup = max(args.warmup, 3)
    name, n_frames, W, H, flags = cr.configs.SEQUENCES[args.workload]
    pod = load_scene_pod(name)
    with np.load(os.path.join(ROOT, "tests", "golden", "kat.npz")) as z:
        path = np.ascontiguousarray(z["turntable"][:n_frames], np.float32)        # n x 13: eye, dir, up, fov, sun
    mine = sequence_frames(n_frames, rank, world)
    cams = [(cr.Camera(tuple(c[0:3]), tuple(c[3:6]), tuple(c[6:9]), float(c[9])), c[10:13].copy()) for c in path]
    scene = cr.Scene.from_pod(pod, device=local)
    d_fb = torch.empty(H * W * 3, dtype=torch.float32, device=dev)
    d_cnt = torch.zeros((n_frames, 2), dtype=torch.int64, device=dev)
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev)

    def one_pass():
        for f in mine:
            scene.render_device(cams[f][0], cams[f][1], W, H, flags=flags, pixels=d_fb, counters=d_cnt[f])

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def allsum(t):
        if world > 1:
            dist.all_reduce(t)
        return t

    for _ in range(warmup):
        one_pass()
    barrier()
    rays_total = int(allsum(d_cnt.sum(dim=0).clone())[0].item())
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    launches0 = cr.kernel_launch_count()
    scene.set_profiling(True)
    with ClockSampler(local) as clocks:
        barrier()
        clocks.mark("timed")
        for a, b in ev:
            flush.fill_(1)
            a.record()
            one_pass()
            b.record()
        barrier()
        clocks.mark("between")
    launches = cr.kernel_launch_count() - launches0
    kms, kframes = scene.kernel_ms()
    scene.set_profiling(False)
    t = torch.tensor([sum(a.elapsed_time(b) for a, b in ev) / args.steps], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_step = float(t.item())

    # ---- work executed by this rank (statistics build, untimed), summed over its frames --------------------------
    scene.set_stats(True)
    st = {k: 0 for k in ("S_miss", "T_miss", "S_hit", "T_hit", "S_shadow", "T_shadow", "rays", "hits")}
    traced_px = 0
    for f in mine:
        scene.render_device(cams[f][0], cams[f][1], W, H, flags=flags, pixels=d_fb, counters=d_cnt[f])
        torch.cuda.synchronize()
        for k, v in scene.get_stats().items():
            st[k] += v
        r_ = scene.last_rectangle
        traced_px += (r_[1] - r_[0]) * (r_[3] - r_[2])
    scene.set_stats(False)
    st["hits_primary"] = st["rays"] - W * H * len(mine)
    st["occluded"] = st["hits"] - st["hits_primary"]
    abytes = algorithmic_bytes(st, W, H * len(mine), st["hits_primary"] - st["occluded"], traced_px)
    per_pass = {k: kms[k] / max(args.steps, 1) for k in ("clear", "frame", "total")}
    dominant = "frame"
    peak, peak_src = measured_peak()
    achieved = abytes[dominant] / (per_pass[dominant] * 1e-3) / 1e9 if per_pass[dominant] > 0 else 0.0

    # ---- end to end: ceres_render_sequence() into page-locked HOST frames (D2H of frame k overlaps frame k+1) -------
    own = np.ascontiguousarray(path[mine])
    hb = cr.PinnedBuffer((len(mine), H, W, 3), np.float32)
    scene.render_sequence(own[:2], W, H, flags=flags, pixels_all=hb.array[:2])
    t_e2e = 0.0
    steps_e2e = max(1, min(args.steps, 3))
    for _ in range(steps_e2e):
        barrier()
        t0 = time.perf_counter()
        _, cnt_e2e = scene.render_sequence(own, W, H, flags=flags, pixels_all=hb.array)
        t_e2e += time.perf_counter() - t0
    t = torch.tensor([1e3 * t_e2e / steps_e2e], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_e2e = float(t.item())
    assert int(cnt_e2e[:, 0].sum()) == int(d_cnt[mine].sum(dim=0)[0].item())
    d2h_seq = scene.last_d2h_bytes * len(mine)       # turntable: the rectangle is (nearly) the same for every frame
    hb.close()
    # the same sequence as 8-bit payloads in anim's row order (ceres_render_sequence_payload): what apps/anim writes to disk
    hb8 = cr.PinnedBuffer((len(mine), H, W, 3), np.uint8)
    scene.render_sequence_payload(own[:2], W, H, flags=flags | cr.PAYLOAD_NO_FLIP, bytes_per_channel=1, out=hb8.array[:2])
    t_8 = 0.0
    for _ in range(steps_e2e):
        barrier()
        t0 = time.perf_counter()
        _, cnt_8 = scene.render_sequence_payload(own, W, H, flags=flags | cr.PAYLOAD_NO_FLIP, bytes_per_channel=1, out=hb8.array)
        t_8 += time.perf_counter() - t0
    t = torch.tensor([1e3 * t_8 / steps_e2e], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_8 = float(t.item())
    assert int(cnt_8[:, 0].sum()) == int(cnt_e2e[:, 0].sum())
    hb8.close()

    cpu = None
    if world == 1 and not args.no_cpu_baseline:
        secs, rays_cpu = [], 0
        sample = list(range(0, n_frames, max(1, n_frames // 4)))[:4]
        for f in sample:
            s_, kind, cores, r_, _h = cpu_reference_frames(pod, path[f, :10], path[f, 10:13], W, H, frames=1, warmup=1 if f == sample[0] else 0)
            secs += s_
            rays_cpu += r_
        cpu = {"value": rays_cpu / sum(secs) / 1e6, "unit": UNIT, "cores": cores, "kind": kind,
               "sample": f"frames {sample} of the {n_frames}-frame sequence, unmodified render() with the reference's flags"}
    if rank == 0:
        line = {
            "metric": f"Mrays/s ({args.workload}: {n_frames} frames {W}x{H}, primary + shadow rays)", "value": rays_total / (ms_step * 1e-3) / 1e6,
            "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": warmup, "ms_per_step": ms_step, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": DATA + "; camera path = the reference's turntable output (tests/golden/kat.npz)",
            "config": {"workload": args.workload, "frames": n_frames, "width": W, "height": H, "rays_per_sequence": rays_total,
                       "ms_per_frame": ms_step / max(1, len(sequence_frames(n_frames, 0, world))),
                       "sharding": f"frame f -> rank f mod {world}, no collective" if world > 1 else "none",
                       "l2": "flushed (256 MiB fill) before every timed pass; every frame rewrites the frame buffer"},
            "e2e": {"value": rays_total / (ms_e2e * 1e-3) / 1e6, "unit": UNIT, "ms_per_step": ms_e2e, "h2d_bytes_per_step": 13 * 4 * len(mine),
                    "d2h_bytes_per_step": d2h_seq, "host_frame_bytes": 3 * W * H * 4 * len(mine), "path": "ceres_render_sequence() -> page-locked host frames"},
            "e2e_ppm8": {"value": rays_total / (ms_8 * 1e-3) / 1e6, "unit": UNIT, "ms_per_step": ms_8,
                         "path": "ceres_render_sequence_payload(1 byte per channel, anim row order) -> page-locked 8-bit payloads"},
            "gpu_launches": launches, "kernels_ms": per_pass,
            "roofline": {"bound": "issue", "kernel": "frame_kernel", "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved / peak, "traffic": None, "peak_source": peak_src,
                         "algorithmic_bytes_per_launch": abytes[dominant] // max(1, len(mine)),
                         "note": "summed over this rank's frames; the 3 MB scene is L1/L2-resident, so the HBM figure is context only: "
                                 "the kernel is bound by instruction issue (see the single-frame bench line's roofline.issue)"},
            "clocks": clocks.summary(),
        }
        if cpu is not None:
            line["cpu_baseline"] = cpu
        emit(line)
    if world > 1:
        dist.destroy_process_group()


_RESULT = sys.stdout


def keep_stdout_for_the_result():
    """The contract is ONE JSON line on stdout.  Libraries write banners to file descriptor 1 from C (NCCL prints
    "NCCL version ..." when the box sets NCCL_DEBUG): point fd 1 at stderr for everything else and keep a private
    handle on the real stdout for the result line."""
    global _RESULT
    sys.stdout.flush()
    _RESULT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)


def emit(line):
    _RESULT.write(json.dumps(line) + "\n")
    _RESULT.flush()


def main():
    keep_stdout_for_the_result()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--workload", default=WORKLOAD)
    ap.add_argument("--sustain-s", type=float, default=2.0, help="length of the sustained block (seconds of back-to-back frames)")
    ap.add_argument("--verify", action="store_true",
                    help="after the measurement, render the same frame with the oracle (CPU) and compare hit ids, "
                         "counters and RGB at the full size (N=1)")
    ap.add_argument("--gather", default="peer", choices=["peer", "nccl"],
                    help="N>1: 'peer' = kernels store into rank 0's frame over NVLink (ceres_render_peer); "
                         "'nccl' = packed bands + NCCL gather + row scatter")
    args = ap.parse_args()
    import ceres_raytracer_b200.configs as _cfg
    if args.workload in _cfg.SEQUENCES:
        if args.impl == "reference":
            raise SystemExit("--impl reference runs single-frame workloads; the sequence bench times the reference on sample frames itself")
        return run_sequence(args)
    if args.impl == "reference":
        return run_reference_arm(args)

    import torch
    import torch.distributed as dist
    import ceres_raytracer_b200 as cr
    cfg = cr.configs

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world == 1 and args.gpus > 1:
        raise SystemExit("launch N>1 with: python -m torch.distributed.run --nproc-per-node N bench.py --gpus N ...")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    warmup = max(args.warmup, 3)

    pod, cam, sun, W, H, flags, scene = load_workload(args.workload, device=local)
    if scene is None:
        scene = cr.Scene.from_pod(pod, device=local)
    # row-block sharding + the gather that assembles the frame on rank 0 (ceres_raytracer_b200/shard.py)
    from ceres_raytracer_b200.shard import FrameGather
    fg = FrameGather(W, H, cfg.TILE_BLOCK_ROWS, rank, world, dev)
    shard, n_rows = fg.shard, fg.n_rows
    d_band, d_cnt = fg.band, fg.counters
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev)         # > 126 MB L2
    rflags = flags | (cr.COMPACT_ROWS if world > 1 else 0)

    pf = None
    if world > 1 and args.gather == "peer":
        from ceres_raytracer_b200.shard import PeerFrame
        pf = PeerFrame(scene, W, H, rank, world)          # the frame lives in rank 0's HBM, every rank's kernels write it

    def frame():
        if pf is not None:
            pf.render(cam, sun, flags=flags)        # ONE kernel per rank: trace + shade-into-rank-0 (+ rank 0 waits inside it)
            return
        scene.render_device(cam, sun, W, H, flags=rflags, shard=shard, pixels=d_band, counters=d_cnt)
        fg.assemble()               # N > 1: NCCL gather to rank 0 + row scatter, counters summed over shards

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def allmax(x):
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def allgather(x):
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        if world == 1:
            return [float(x)]
        out = [torch.zeros_like(t) for _ in range(world)]
        dist.all_gather(out, t)
        return [float(o.item()) for o in out]

    with ClockSampler(local) as clocks:          # samples from here to the end; the regions are tagged below
        clocks.mark("warmup")
        for _ in range(warmup):
            flush.fill_(1)
            frame()
        barrier()
        counters = (pf.counters if pf is not None and rank == 0 else d_cnt).cpu().tolist()
        rays_frame, hits_frame = int(counters[0]), int(counters[1])
        if pf is not None and rank == 0 and pf.error():
            raise SystemExit("peer frame: a rank did not deliver within the wait bound")

        # ---- timed region: K frames, each preceded by an L2 flush that is outside its event pair ------------------
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
        launches0 = cr.kernel_launch_count()
        scene.set_profiling(True)            # per-kernel CUDA events on the launch stream, inside the timed region
        barrier()
        clocks.mark("timed")
        for a, b in ev:
            flush.fill_(1)
            a.record()
            frame()
            b.record()
        barrier()
        clocks.mark("between")
        launches = cr.kernel_launch_count() - launches0
        kms, kframes = scene.kernel_ms()
        scene.set_profiling(False)
        ms_step = allmax(sum(a.elapsed_time(b) for a, b in ev) / args.steps)
        kavg = {k: kms[k] / max(kframes, 1) for k in ("clear", "frame", "total")}
        per_rank_kernel_ms = allgather(kavg["total"])

        # ---- sustained: >= args.sustain_s of back-to-back frames, clock sampled throughout (warm L2: no flush) -------
        barrier()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(args.steps):
            frame()
        b.record()
        barrier()
        ms_warm = allmax(a.elapsed_time(b) / args.steps)
        n_sus = int(max(args.steps, min(200000, args.sustain_s * 1e3 / max(ms_warm, 1e-3))))
        n_sus = int(allmax(n_sus))
        barrier()
        clocks.mark("sustained")
        a.record()
        done = 0
        while done < n_sus:                     # in slices, so that the host never runs far ahead of the device queue
            for _ in range(min(500, n_sus - done)):
                frame()
            done += min(500, n_sus - done)
            torch.cuda.current_stream().synchronize()
        b.record()
        barrier()
        clocks.mark("between")
        ms_sus = allmax(a.elapsed_time(b) / n_sus)
        if pf is not None and rank == 0 and pf.error():
            raise SystemExit("peer frame: a rank did not deliver within the wait bound")

        # ---- work actually executed by this rank's kernels (statistics build of the same kernel, untimed) ----------
        scene.set_stats(True)
        scene.render_device(cam, sun, W, H, flags=rflags, shard=shard, pixels=d_band, counters=d_cnt)
        torch.cuda.synchronize()
        st = scene.get_stats()
        scene.set_stats(False)
        rect = scene.last_rectangle
        traced_px = (rect[1] - rect[0]) * (rect[3] - rect[2])
        local_rays, local_hits = st["rays"], st["hits"]
        if flags & cr.NO_SHADOWS:      # primary-only: every hit is lit, hits counter = primary hits
            st["hits_primary"], st["occluded"] = local_hits, 0
        else:
            st["hits_primary"] = local_rays - W * n_rows
            st["occluded"] = local_hits - st["hits_primary"]
        lit = st["hits_primary"] - st["occluded"]
        rays_traced_local = traced_px + (0 if flags & cr.NO_SHADOWS else st["hits_primary"])
        t = torch.tensor([rays_traced_local], dtype=torch.int64, device=dev)
        if world > 1:
            dist.all_reduce(t)
        rays_traced = int(t.item())
        abytes = algorithmic_bytes(st, W, n_rows, lit, traced_px)
        peak, peak_src = measured_peak()
        achieved = abytes["frame"] / (kavg["frame"] * 1e-3) / 1e9 if kavg["frame"] > 0 else 0.0
        whole = sum(abytes.values()) / (kavg["total"] * 1e-3) / 1e9 if kavg["total"] > 0 else 0.0
        # second denominator: read bandwidth of an L2-resident working set, measured on this GPU
        try:
            l2_peak = cr.probe_read_bandwidth(48 << 20, repeats=20, device=local)
        except Exception:
            l2_peak = None

        # ---- end to end through the reference-facing call with a HOST framebuffer ------------------------------------
        clocks.mark("e2e")
        steps_e2e = args.steps
        extra = {}

        def time_host_call(fn, n=steps_e2e, pre=2):
            for _ in range(pre):
                fn()
            torch.cuda.synchronize()
            tot = 0.0
            for _ in range(n):
                flush.fill_(1)
                torch.cuda.synchronize()
                t0 = time.perf_counter()
                out = fn()
                tot += time.perf_counter() - t0
            return 1e3 * tot / n, out

        if world == 1:
            # N = 1: ceres_render() (the drop-in for render.hpp:86-156): launch + D2H of the frame + counters inside the timed region
            hb = cr.PinnedBuffer((H, W, 3), np.float32)
            ms_e2e, r = time_host_call(lambda: scene.render(cam, sun, W, H, flags=flags, pixels=hb.array))
            assert r[2] == (rays_frame, hits_frame)
            d2h = scene.last_d2h_bytes        # the rectangles that hold hit pixels + counters; the rest of the frame is host-filled zeros
            # opt-in contract CERES_FRAME_RETAINED (include/ceres_b200.h): the caller passes the buffer of the previous call,
            # untouched -- what anim.cpp does -- so the ~70 % of the frame outside the scene's rectangle is not zero-filled again
            ms_ret, r = time_host_call(lambda: scene.render(cam, sun, W, H, flags=flags | cr.FRAME_RETAINED, pixels=hb.array))
            assert r[2] == (rays_frame, hits_frame)
            extra["e2e_retained"] = {"value": rays_frame / (ms_ret * 1e-3) / 1e6, "unit": UNIT, "ms_per_step": ms_ret,
                                     "d2h_bytes_per_step": scene.last_d2h_bytes,
                                     "path": "ceres_render(CERES_FRAME_RETAINED) -> the same page-locked frame as in the previous call (opt-in: "
                                             "the caller promises not to write into it between calls); same frame, no host zero fill"}
            hb.close()
            # the same call into a plain malloc'ed frame -- what anim.cpp:76 / static.cpp:117 hand to render()
            plain = np.empty((H, W, 3), np.float32)
            ms_pg, r = time_host_call(lambda: scene.render(cam, sun, W, H, flags=flags, pixels=plain))
            assert r[2] == (rays_frame, hits_frame)
            extra["e2e_pageable"] = {"value": rays_frame / (ms_pg * 1e-3) / 1e6, "unit": UNIT, "ms_per_step": ms_pg,
                                     "d2h_bytes_per_step": scene.last_d2h_bytes,
                                     "path": "ceres_render() -> plain malloc'ed (pageable) frame, the buffer anim.cpp:76 / static.cpp:117 pass"}
            del plain
            # the same frame as 8-bit P6 payload (static.cpp:135-147 on the device; N1): a quarter of the bytes
            hb8 = cr.PinnedBuffer((H, W, 3), np.uint8)
            ms8, _ = time_host_call(lambda: scene.render_ppm8(cam, sun, W, H, flags=flags, out=hb8.array))
            ppm8 = {"value": rays_frame / (ms8 * 1e-3) / 1e6, "unit": UNIT, "ms_per_step": ms8,
                    "d2h_bytes_per_step": scene.last_d2h_bytes, "path": "ceres_render_ppm8() -> page-locked 8-bit P6 payload"}
            hb8.close()
            # and as 16-bit payload (the x 65535 scaling of anim.cpp:117-119 on the device): half of the bytes
            hb16 = cr.PinnedBuffer((H, W, 3), np.uint16)
            ms16, _ = time_host_call(lambda: scene.render_ppm16(cam, sun, W, H, flags=flags, out=hb16.array))
            ppm8["ppm16"] = {"value": rays_frame / (ms16 * 1e-3) / 1e6, "unit": UNIT, "ms_per_step": ms16,
                             "d2h_bytes_per_step": scene.last_d2h_bytes, "path": "ceres_render_ppm16() -> page-locked 16-bit P6 payload"}
            hb16.close()
            extra["e2e_ppm8"] = ppm8
            extra["e2e_shim"] = shim_bench(pod, cam, sun, W, H, steps_e2e, rays_frame, hits_frame)
            try:
                extra["device_built_tree"] = device_tree_block(scene, pod, cam, sun, W, H, flags, dev)
            except Exception as e:      # next-row measurement: never takes the headline down with it
                extra["device_built_tree"] = {"error": str(e)}
        else:
            # N > 1: one HOST frame shared by the ranks (page-locked in every process); each rank renders its row blocks and
            # DMA-writes them in place over its own PCIe link; the counter barrier in the shared control page completes it.
            from ceres_raytracer_b200.shard import SharedHostFrame
            port_tag = os.environ.get('MASTER_PORT', '0')

            def time_shared(hf, fl):
                t = 0.0
                cnt = None
                for k in range(steps_e2e + 2):
                    flush.fill_(1)
                    barrier()
                    t0 = time.perf_counter()
                    cnt = hf.render(scene, cam, sun, W, H, flags=fl)
                    if k >= 2:
                        t += time.perf_counter() - t0
                if rank == 0:
                    assert tuple(cnt) == (rays_frame, hits_frame)
                tt = torch.tensor([scene.last_d2h_bytes], dtype=torch.int64, device=dev)
                dist.all_reduce(tt)
                return allmax(1e3 * t / steps_e2e), int(tt.item())

            hf = SharedHostFrame(f"ceres_bench_{port_tag}", (H, W, 3), rank, world)
            ms_e2e, d2h = time_shared(hf, flags)
            # opt-in CERES_FRAME_RETAINED: every rank keeps the state of its own rows of the shared frame
            ms_ret, d2h_ret = time_shared(hf, flags | cr.FRAME_RETAINED)
            extra["e2e_retained"] = {"value": rays_frame / (ms_ret * 1e-3) / 1e6, "unit": UNIT, "ms_per_step": ms_ret, "d2h_bytes_per_step": d2h_ret,
                                     "path": "ceres_render_shard(CERES_FRAME_RETAINED) per rank -> the same shared page-locked frame as in the previous "
                                             "call (opt-in contract); no host zero fill"}
            hf.close()
            # the 8-bit P6 payload, sharded the same way (ceres_render_ppm8_shard): 24.9 MB of host memory instead of 99.5 MB
            hf8 = SharedHostFrame(f"ceres_bench8_{port_tag}", (H, W, 3), rank, world, dtype=np.uint8)
            ms8, d2h8 = time_shared(hf8, flags)
            extra["e2e_ppm8"] = {"value": rays_frame / (ms8 * 1e-3) / 1e6, "unit": UNIT, "ms_per_step": ms8, "d2h_bytes_per_step": d2h8,
                                 "path": "ceres_render_ppm8_shard() per rank -> one shared page-locked 8-bit P6 payload"}
            hf8.close()
        # host-side floor of a float host frame: every byte of it is written into host DRAM by somebody
        try:
            fill_gbs = cr.probe_host_fill_bandwidth(256 << 20, 3) if rank == 0 else 0.0
        except Exception:
            fill_gbs = 0.0
        clocks.mark("between")

    # ---- CPU baseline: the reference's own render() on this box's host cores (rank 0, N = 1 only) ------------------
    cpu = None
    if world == 1 and not args.no_cpu_baseline:
        try:
            secs, kind, cores, rays_cpu, _h = cpu_reference_frames(pod, cam.as_array(), sun, W, H, frames=8, warmup=1, budget_s=12.0)
            ms_cpu = 1e3 * float(np.mean(secs))
            cpu = {"value": rays_cpu / (ms_cpu * 1e-3) / 1e6, "unit": UNIT, "cores": cores, "kind": kind,
                   "ms_per_frame": ms_cpu,
                   "sample": f"{len(secs)} full {W}x{H} frames of the same workload, unmodified render() with the reference's flags "
                             f"(best {min(secs)*1e3:.1f} ms, median {np.median(secs)*1e3:.1f} ms)"}
        except Exception as e:      # the baseline is reporting only; never fail the GPU line for it
            cpu = {"value": None, "unit": UNIT, "cores": None, "kind": "unavailable", "sample": str(e)[:200]}

    # ---- --verify: the same frame through the oracle, full size ----------------------------------------------------
    verified = None
    if args.verify and world == 1:
        from oracle import port
        px, ids, cnt = scene.render(cam, sun, W, H, flags=flags, want_ids=True)
        o = port.render(pod, cam.as_array(), sun, W, H, flags)
        verified = {"oracle": "oracle/ceres_oracle.c (strict restatement)", "ids_equal": bool(np.array_equal(ids.reshape(-1), o["ids"].reshape(-1))),
                    "id_mismatches": int((ids.reshape(-1) != o["ids"].reshape(-1)).sum()),
                    "counters_equal": bool(cnt == (o["stats"]["rays"], o["stats"]["hits"])),
                    "max_abs_rgb_diff": float(np.abs(px.reshape(-1) - o["pixels"].reshape(-1)).max()),
                    "oracle_steps": {k: int(o["stats"][k]) for k in ("S_miss", "S_hit", "S_shadow", "T_miss", "T_hit", "T_shadow") if k in o["stats"]}}
        del px, ids, o

    if rank == 0:
        ck = clocks.summary()
        sm_count = torch.cuda.get_device_properties(local).multi_processor_count
        # issue roofline: warp instructions of the frame kernel (deterministic for a workload; counted by the committed
        # ncu capture) against the issue slots of the SMs -- 4 schedulers per SM, one warp instruction per cycle each
        issue = None
        ip = os.path.join(ROOT, "profiles", "ncu_frame_kernel.json")
        if os.path.exists(ip) and args.workload == WORKLOAD and world == 1:
            try:
                cap = json.load(open(ip))
                f_hz = (ck["sm_mhz"] or ck["sm_max_mhz"] or 1965.0) * 1e6
                slots = sm_count * 4 * f_hz * kavg["frame"] * 1e-3
                issue = {"warp_instructions": cap["warp_instructions"], "issue_slots": slots, "frac": cap["warp_instructions"] / slots,
                         "lanes_per_instruction": cap.get("lanes_per_instruction"), "sm_mhz": f_hz / 1e6,
                         "ncu_issue_active_pct": cap.get("issue_active_pct"), "source": cap.get("source")}
            except Exception:
                issue = None
        host_bytes = 3 * W * H * 4
        line = {
            "metric": metric_name(args.workload, W, H, flags), "value": rays_frame / (ms_step * 1e-3) / 1e6, "unit": UNIT, "n_gpus": world,
            "steps": args.steps, "warmup": warmup, "ms_per_step": ms_step, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": DATA,
            "config": frame_config(args.workload, W, H, rays_frame, scene_device_bytes(pod), sharding_text(world, pf is not None), L2_TEXT),
            "counters": {"rays": rays_frame, "hits": hits_frame},
            "rays_traced": {"per_frame": rays_traced, "value": rays_traced / (ms_step * 1e-3) / 1e6, "unit": "M traced rays/s",
                            "note": "rays actually walked through the BVH: pixels of the scene's conservative screen rectangle + shadow rays; "
                                    "`value` counts the reference's rays (W*H + primary hits), most of which miss the rectangle"},
            "sustained": {"ms_per_step": ms_sus, "value": rays_frame / (ms_sus * 1e-3) / 1e6, "unit": UNIT, "frames": n_sus,
                          "seconds": ms_sus * n_sus * 1e-3, "l2": "not flushed (back-to-back frames; every frame rewrites the frame buffer)",
                          "ms_per_step_same_length_as_timed": ms_warm,
                          "sm_mhz": ck["regions"].get("sustained", {}).get("sm_mhz"), "clock_samples": ck["regions"].get("sustained", {}).get("samples", 0)},
            "e2e": {"value": rays_frame / (ms_e2e * 1e-3) / 1e6, "unit": UNIT, "ms_per_step": ms_e2e,
                    "h2d_bytes_per_step": 13 * 4, "d2h_bytes_per_step": d2h, "host_frame_bytes": host_bytes,
                    "floor_ms": (host_bytes / (fill_gbs * 1e9) * 1e3) if fill_gbs else None, "host_fill_gbs": fill_gbs or None,
                    "floor_note": "host_frame_bytes / streaming-store bandwidth of the host-fill pool measured in this run: every byte of the "
                                  "frame is written into host DRAM by somebody (DMA for the hit rectangles, the pool for the zeros)",
                    "path": "ceres_render() -> page-locked host framebuffer" if world == 1 else
                            "ceres_render_shard() per rank -> one shared page-locked host frame, each GPU drains its rows over its own PCIe link"},
            "gpu_launches": launches,
            "kernels_ms": kavg, "kernels_ms_per_rank": per_rank_kernel_ms,
            "roofline": {"bound": "issue", "kernel": "frame_kernel", "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved / peak, "traffic": ncu_dram_traffic(args.workload), "peak_source": peak_src,
                         "algorithmic_bytes_per_launch": abytes["frame"],
                         "issue": issue,
                         "l2": {"peak": l2_peak, "frac": (achieved / l2_peak) if l2_peak else None, "unit": "GB/s",
                                "peak_source": "measured here: ceres_probe_read_bandwidth, 48 MiB working set, 128-bit __ldg reads"},
                         "whole_frame": {"achieved": whole, "frac": whole / peak, "bytes": sum(abytes.values())},
                         "work": {k: st[k] for k in ("S_miss", "T_miss", "S_hit", "T_hit", "S_shadow", "T_shadow", "hits_primary", "occluded")},
                         "note": ("the 3 MB scene is L1/L2-resident (ncu: DRAM traffic ~1 % of the algorithmic bytes), so achieved/peak/frac -- the "
                                  "HBM-denominated figures SURVEY.md 8(d) prescribes -- exceed 1 and are context only; the kernel is bound by "
                                  "instruction issue x SIMD efficiency: see `issue`"
                                  if scene_device_bytes(pod) < (64 << 20) else
                                  f"scene ({scene_device_bytes(pod) / 1e9:.2f} GB) exceeds the 126 MB L2: node/triangle fetches are HBM/L2-miss traffic")},
            "clocks": ck,
        }
        line.update(extra)
        if cpu is not None:
            line["cpu_baseline"] = cpu
        if verified is not None:
            line["verified"] = verified
        if args.workload != WORKLOAD:
            line["data"] = f"synthetic: {cfg.WORKLOADS[args.workload][0]} (ceres_raytracer_b200.scenes.instanced_grid of the reference-built dragon scene, boxes by device refit)"
        emit(line)
    if world > 1:
        dist.destroy_process_group()


def device_tree_block(scene, pod, cam, sun, W, H, flags, dev, frames=20):
    """N3: the same frame from the tree the LIBRARY builds on the device (ceres_scene_create_lbvh, default
    builder) next to the caller's reference-built tree, both device-resident, back-to-back frames, CUDA events."""
    import torch
    import ceres_raytracer_b200 as cr
    px = torch.empty(H * W * 3, dtype=torch.float32, device=dev)
    cnt = torch.zeros(2, dtype=torch.int64, device=dev)

    def ms_per_frame(sc, ids=None):
        for _ in range(3):
            sc.render_device(cam, sun, W, H, flags=flags, pixels=px, ids=ids, counters=cnt)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(frames):
            sc.render_device(cam, sun, W, H, flags=flags, pixels=px, counters=cnt)
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / frames

    ids_a = torch.empty(H * W, dtype=torch.int32, device=dev)
    ms_ref = ms_per_frame(scene, ids_a)
    cnt_ref = cnt.cpu().tolist()
    cr.Scene.from_lbvh(pod["triangles"], pod["norms"], device=dev.index or 0).close()      # allocator / module warm-up
    with cr.Scene.from_lbvh(pod["triangles"], pod["norms"], device=dev.index or 0) as s2:
        ids_b = torch.empty(H * W, dtype=torch.int32, device=dev)
        ms = ms_per_frame(s2, ids_b)
        return {"builder": "level-synchronous top-down binned SAH + SAH leaf collapsing (ceres_scene_create_lbvh, default options)",
                "triangles": int(pod["triangles"].shape[0]), "nodes": int(s2.node_count), "reference_tree_nodes": int(pod["nodes"].shape[0]),
                "build_kernels_ms": s2.build_ms, "stack_depth": s2.stack_depth,
                "ms_per_step": ms, "reference_tree_ms_per_step": ms_ref, "vs_reference_tree": ms / ms_ref,
                "counters_equal": cnt.cpu().tolist() == cnt_ref, "hit_ids_differing": int((ids_a != ids_b).sum().item()),
                "note": "frames back to back, L2 not flushed, both trees timed the same way here; ids may differ where two triangles are hit "
                        "at exactly the same t (SURVEY.md 7.3-H2): parity of this tree is against the oracle walking it (tests/test_gpu_lbvh.py)"}


def shim_bench(pod, cam, sun, W, H, frames, rays, hits):
    """The drop-in C++ call -- render<float>() of include/ceres/render.hpp, apps/shim_bench.cpp -- into the frame the reference's
    callers allocate (std::make_unique<float[]>, anim.cpp:76 / static.cpp:117): the shim page-locks it on first sight."""
    exe = os.path.join(ROOT, "apps", "_build", "shim_bench")
    if not os.path.exists(exe):
        return {"unavailable": "apps/_build/shim_bench not built (needs the reference's lib/bvh headers at build time)"}
    import subprocess
    import tempfile
    out = {"path": "apps/shim_bench: render<float>() of include/ceres/render.hpp, C++ caller, same scene arrays and camera", "unit": UNIT}
    try:
        with tempfile.TemporaryDirectory() as d:
            p = os.path.join(d, "scene.bin")
            with open(p, "wb") as f:
                f.write(np.array([len(pod["nodes"]), len(pod["triangles"])], np.uint64).tobytes())
                f.write(np.ascontiguousarray(pod["nodes"]).tobytes())
                f.write(np.ascontiguousarray(pod["prim_indices"], np.uint64).tobytes())
                f.write(np.ascontiguousarray(pod["triangles"], np.float32).tobytes())
                f.write(np.ascontiguousarray(pod["norms"], np.float32).tobytes())
            c = cam.as_array()
            argv = [exe, p, str(W), str(H), str(frames)] + [repr(float(x)) for x in list(c[:10]) + list(np.asarray(sun, np.float32))]
            for key, env in (("autopin", {}), ("no_autopin", {"CERES_NO_AUTOPIN": "1"})):
                r = subprocess.run(argv, env=dict(os.environ, **env), capture_output=True, text=True, timeout=120)
                if r.returncode != 0:
                    return {"unavailable": f"shim_bench exit {r.returncode}: {r.stderr[-200:]}"}
                for ln in r.stdout.splitlines():
                    q = json.loads(ln)
                    if (q["rays"], q["hits"]) != (rays, hits):
                        return {"unavailable": f"shim_bench counters {q['rays']}, {q['hits']} differ from the GPU arm's"}
                    tag = "make_unique" if q["frame"].startswith("make_unique") else "host_alloc"
                    if key == "no_autopin" and tag == "host_alloc":
                        continue
                    name = tag if key == "autopin" else "make_unique_no_autopin"
                    out[name] = {"ms_per_step": q["ms_mean"], "ms_best": q["ms_best"], "value": rays / (q["ms_mean"] * 1e-3) / 1e6, "frame": q["frame"]}
    except Exception as e:
        return {"unavailable": str(e)[:200]}
    return out


def ncu_dram_traffic(workload):
    """dram__bytes_read.sum + dram__bytes_write.sum of the frame kernel, per launch, from the committed ncu --set full capture."""
    tp = os.path.join(ROOT, "profiles", "ncu_frame_kernel.json")
    if workload != WORKLOAD or not os.path.exists(tp):
        return None
    try:
        return json.load(open(tp)).get("dram_bytes")
    except Exception:
        return None


if __name__ == "__main__":
    main()
