#!/usr/bin/env python
"""Scalar = double path (ceres_render_f64): kernel time and whole-call time for the dragon at 1080p / 4K, next to the
reference's own render<double>() on the host cores (oracle/_ref/libceres_ref64_strict.so) and the float path."""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import ceres_raytracer_b200 as cr
from oracle import f64

with np.load(os.path.join(ROOT, "tests", "golden", "dragon_scene.npz")) as z:
    pod32 = {k: z[k] for k in z.files}
pod = f64.widen(pod32)
c = cr.configs.DRAGON_CAMERA
cam, sun = f64.cam_array(c.eye, c.dir, c.up, c.fov), np.array([-50., -20., 0.])
ref = f64.RefScene64.from_pod(pod) if f64.ref_available() else None
with cr.SceneF64.from_pod(pod) as s, cr.Scene.from_pod(pod32) as s32:
    for W, H in ((1920, 1080), (3840, 2160)):
        s.render(cam, sun, W, H)
        ks, ts = [], []
        for _ in range(5):
            t0 = time.perf_counter()
            px, _, cnt = s.render(cam, sun, W, H)
            ts.append(time.perf_counter() - t0)
            ks.append(s.kernel_ms)
        k, t = min(ks), min(ts) * 1e3
        frame = np.full((H, W, 3), np.nan)
        tr = []
        for _ in range(8):
            t0 = time.perf_counter()
            px2, _, cnt2 = s.render(cam, sun, W, H, pixels=frame)
            tr.append(time.perf_counter() - t0)
        assert cnt2 == cnt and np.array_equal(px2, px)
        line = (f"{W}x{H}: rays {cnt[0]} hits {cnt[1]}; f64 device time (bands: kernels + copies) {k:.3f} ms; whole call into a fresh pageable numpy "
                f"frame {t:.1f} ms, into a REUSED pageable frame (anim.cpp:76) {min(tr)*1e3:.2f} ms (median {np.median(tr)*1e3:.2f})")
        s32.render(c, sun, W, H)
        t0 = time.perf_counter(); _, _, c32 = s32.render(c, sun, W, H); t32 = (time.perf_counter() - t0) * 1e3
        line += f"; float path whole call {t32:.2f} ms (hits {c32[1]})"
        if ref is not None:
            ref.render(cam, sun, W, H)
            best = min(ref.render(cam, sun, W, H)[2] for _ in range(3))
            rp, rc, _ = ref.render(cam, sun, W, H)
            line += f"; reference render<double>() on {os.cpu_count()} threads {best*1e3:.1f} ms = {rc[0]/best/1e6:.1f} Mrays/s, counters equal {rc == cnt}, max |rgb diff| {np.abs(rp - px).max():.2e}"
        print(line)
