"""The drop-in executables (apps/render, apps/anim -- C++ host code over the C ABI) against the reference's own code.

The reference tree does not travel to the GPU box, so the OBJ input is synthesised here (a bumpy torus with shared
vertices, quads, negative indices and CRLF line ends -- the loader features obj_norms.hpp:77-105 handles); the
expected output comes from oracle/_ref/libceres_ref_strict.so (the reference's loader, builders, rotate_triangles and
render() compiled unchanged), which does travel.  Bars as in test_gpu_parity.py: ids bit-exact, RGB 1e-6, PPM bytes
equal up to one LSB on <= 1e-5 of the bytes."""
import os
import subprocess

import numpy as np
import pytest

from oracle import ref

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
RENDER = os.path.join(ROOT, "apps", "_build", "render")
ANIM = os.path.join(ROOT, "apps", "_build", "anim")
SHIM_CHECK = os.path.join(ROOT, "apps", "_build", "shim_check")
RGB_ATOL = 1e-6


def write_torus_obj(path, nu=96, nv=48):
    u = np.linspace(0, 2 * np.pi, nu, endpoint=False)
    v = np.linspace(0, 2 * np.pi, nv, endpoint=False)
    uu, vv = np.meshgrid(u, v, indexing="ij")
    r = 0.35 + 0.05 * np.sin(5 * uu) * np.cos(3 * vv)
    x = (1.0 + r * np.cos(vv)) * np.cos(uu)
    y = r * np.sin(vv)
    z = (1.0 + r * np.cos(vv)) * np.sin(uu)
    nvert = nu * nv
    lines = ["# synthetic torus", "o torus", "s 1"]
    for a, b, c in zip(x.ravel(), y.ravel(), z.ravel()):
        lines.append(f"v {a:.7f} {b:.7f} {c:.7f}")
    lines.append("vn 0 1 0")            # ignored by the loader
    k = 0
    for i in range(nu):
        for j in range(nv):
            a, b = i * nv + j, i * nv + (j + 1) % nv
            c, d = ((i + 1) % nu) * nv + (j + 1) % nv, ((i + 1) % nu) * nv + j
            if k % 3 == 0:      # quad -> fan-triangulated by the loader
                lines.append(f"f {a+1} {b+1} {c+1} {d+1}")
            elif k % 3 == 1:    # two triangles with v/vt/vn style indices
                lines.append(f"f {a+1}//1 {b+1}//1 {c+1}//1")
                lines.append(f"f {a+1}//1 {c+1}//1 {d+1}//1")
            else:               # negative (relative) indices
                lines.append(f"f {a-nvert} {b-nvert} {c-nvert}")
                lines.append(f"f {a-nvert} {c-nvert} {d-nvert}")
            k += 1
    with open(path, "wb") as f:
        f.write(("\r\n".join(lines) + "\r\n").encode())


def read_ppm(path):
    with open(path, "rb") as f:
        head = f.readline().split()
        assert head[0] == b"P6"
        W, H, mx = int(head[1]), int(head[2]), int(head[3])
        data = np.frombuffer(f.read(), np.uint8 if mx < 256 else ">u2")
    return data.reshape(H, W, 3)


def quantise(px):
    x = px.astype(np.float32) * np.float32(255)
    x = np.where(np.float32(255) < x, np.float32(255), x)
    x = np.where(x < 0, np.float32(0), x)
    return x.astype(np.uint8)[::-1]


needs_binaries = pytest.mark.skipif(not (os.path.exists(RENDER) and os.path.exists(ANIM) and ref.available("strict")),
                                    reason="apps/_build or oracle/_ref not built (run __graft_entry__.build() where /root/reference exists)")


@pytest.mark.gpu
@needs_binaries
@pytest.mark.parametrize("builder,bcode,extra", [("binned", 1, []), ("sweep", 0, ["--dump-ids"]), ("sweep", 0, ["--no-shadow", "--dump-ids"])])
def test_render_binary_matches_reference(tmp_path, builder, bcode, extra):
    obj = str(tmp_path / "torus.obj")
    write_torus_obj(obj)
    W, H = 640, 360
    out = str(tmp_path / "render.ppm")
    cmd = [RENDER, obj, "--eye", "0.3", "1.2", "-3.0", "--dir", "-0.1", "-0.35", "1", "--rotate", "y", "-145", "--sun", "-50", "60", "-20",
           "--width", str(W), "--height", str(H), "-o", out, "--builder", builder, "--dump-rgb", str(tmp_path / "rgb.f32")]
    want_ids = "--dump-ids" in extra
    extra = [e for e in extra if e != "--dump-ids"]
    if want_ids:
        cmd += ["--dump-ids", str(tmp_path / "ids.u32")]
    r = subprocess.run(cmd + extra, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "BVH of" in r.stdout and "Rendering image (640x360)" in r.stdout and "Rays:" in r.stdout

    s = ref.RefScene.load_obj(obj, rot_axis=1, rot_deg=-145.0, builder=bcode)
    cam = ref.cam_array((0.3, 1.2, -3.0), (-0.1, -0.35, 1), (0, 1, 0), 60)
    sun = np.array([-50, 60, -20], np.float32)
    flags = 1 if "--no-shadow" in extra else 0
    o = s.render_ex(cam, sun, W, H, flags)
    rays, hits = o["stats"]["rays"], o["stats"]["hits"]
    assert f"Rays: {rays}\tHits: {hits}" in r.stdout
    assert 0.1 < (o["ids"] != ref.MISS).mean() < 0.9
    rgb = np.fromfile(str(tmp_path / "rgb.f32"), np.float32).reshape(H, W, 3)
    assert np.abs(rgb - o["pixels"]).max() <= RGB_ATOL
    if want_ids:
        ids = np.fromfile(str(tmp_path / "ids.u32"), np.uint32).reshape(H, W)
        assert np.array_equal(ids, o["ids"])
    img = read_ppm(out)
    d = img.astype(np.int16) - quantise(o["pixels"]).astype(np.int16)
    assert np.abs(d).max() <= 1 and (d != 0).mean() <= 1e-5
    if not flags:
        # the unmodified reference render() itself
        px, rrays, rhits, _ = s.render(cam, sun, W, H)
        assert (rrays, rhits) == (rays, hits) and np.abs(rgb - px).max() <= RGB_ATOL
    s.close()


@pytest.mark.gpu
@needs_binaries
def test_render_binary_with_the_device_built_tree(tmp_path):
    """--builder device: the tree comes from ceres_build_bvh (top-down binned SAH on the GPU) in the reference's own Bvh
    structure.  Another tree than SweepSah's, so the hit ids may differ where two triangles are hit at exactly the same t
    (SURVEY.md 7.3-H2); everything a tie cannot change -- the counters, which pixels are hit, the colour of every pixel
    whose id agrees -- must equal the reference's frame from its own tree."""
    obj = str(tmp_path / "torus.obj")
    write_torus_obj(obj)
    W, H = 640, 360
    cmd = [RENDER, obj, "--eye", "0.3", "1.2", "-3.0", "--dir", "-0.1", "-0.35", "1", "--rotate", "y", "-145", "--sun", "-50", "60", "-20",
           "--width", str(W), "--height", str(H), "-o", str(tmp_path / "render.ppm"), "--builder", "device",
           "--dump-rgb", str(tmp_path / "rgb.f32"), "--dump-ids", str(tmp_path / "ids.u32")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "B200 binned-SAH builder" in r.stdout and "device build kernels" in r.stdout and "BVH of" in r.stdout
    s = ref.RefScene.load_obj(obj, rot_axis=1, rot_deg=-145.0, builder=0)
    cam = ref.cam_array((0.3, 1.2, -3.0), (-0.1, -0.35, 1), (0, 1, 0), 60)
    o = s.render_ex(cam, np.array([-50, 60, -20], np.float32), W, H, 0)
    s.close()
    assert f"Rays: {o['stats']['rays']}\tHits: {o['stats']['hits']}" in r.stdout
    ids = np.fromfile(str(tmp_path / "ids.u32"), np.uint32).reshape(H, W)
    rgb = np.fromfile(str(tmp_path / "rgb.f32"), np.float32).reshape(H, W, 3)
    assert np.array_equal(ids == ref.MISS, o["ids"] == ref.MISS)
    same = ids == o["ids"]
    assert same.mean() > 0.999
    assert np.abs(rgb - o["pixels"])[same].max() <= RGB_ATOL


@pytest.mark.gpu
@needs_binaries
def test_render_binary_ppm8_writes_the_same_file(tmp_path):
    """--ppm8 (quantise + flip on the device, 8-bit rectangle DMA) and the default float path write identical PPMs."""
    obj = str(tmp_path / "torus.obj")
    write_torus_obj(obj)
    outs = []
    for k, extra in enumerate(([], ["--ppm8"])):
        out = str(tmp_path / f"render{k}.ppm")
        cmd = [RENDER, obj, "--eye", "0.3", "1.2", "-3.0", "--dir", "-0.1", "-0.35", "1", "--rotate", "y", "-145", "--sun", "-50", "60", "-20",
               "--width", "640", "--height", "360", "-o", out, "--builder", "sweep"] + extra
        r = subprocess.run(cmd, capture_output=True, text=True, timeout=300)
        assert r.returncode == 0, r.stdout + r.stderr
        outs.append((open(out, "rb").read(), [l for l in r.stdout.splitlines() if l.startswith("Rays:")][0].split("\t")[:2]))
    assert outs[0][0] == outs[1][0] and outs[0][1] == outs[1][1]


@pytest.mark.gpu
@needs_binaries
def test_anim_binary_matches_reference_turntable(tmp_path):
    obj = str(tmp_path / "torus.obj")
    write_torus_obj(obj, 64, 32)
    W, H, n = 200, 320, 5
    # upstream camera (-150,-20,0) looks along +x from far away; scale the torus up so that it fills the view
    with open(obj, "rb") as f:
        txt = f.read().decode().replace("\r\n", "\n").split("\n")
    with open(obj, "w") as f:
        for line in txt:
            if line.startswith("v "):
                a, b, c = (float(t) * 40 for t in line.split()[1:4])
                f.write(f"v {a:.5f} {b:.5f} {c:.5f}\n")
            else:
                f.write(line + "\n")
    out = str(tmp_path / "turn.mp4")
    r = subprocess.run([ANIM, "--input", obj, "--frames", str(n), "--width", str(W), "--height", str(H), out],
                       capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    s = ref.RefScene.load_obj(obj, builder=0)
    cams = ref.turntable(ref.cam_array((-150, -20, 0), (1, -0.1, 0), (0, -1, 0), 60), [-500, 300, 10], [0, 1, 0], n)
    total = 0
    for k in range(n):
        px, rays, hits, _ = s.render(cams[k, :10], cams[k, 10:13], W, H)
        assert f"Rays: {rays}\tHits: {hits}" in r.stdout
        total += rays
        img = read_ppm(str(tmp_path / f"turn_{k:04d}.ppm"))
        d = img.astype(np.int16) - quantise(px)[::-1].astype(np.int16)       # anim frames are NOT flipped (anim.cpp:117-119)
        assert np.abs(d).max() <= 1 and (d != 0).mean() <= 1e-5
        assert (img != 0).any()
    assert f"Total Rays: {total}" in r.stdout
    s.close()
    # --ppm16: the x 65535 scaling of anim.cpp:117-119, quantised on the GPU, big-endian P6 with maxval 65535
    out16 = str(tmp_path / "turn16.mp4")
    r16 = subprocess.run([ANIM, "--input", obj, "--frames", "2", "--width", str(W), "--height", str(H), "--ppm16", out16],
                         capture_output=True, text=True, timeout=300)
    assert r16.returncode == 0, r16.stdout + r16.stderr
    s = ref.RefScene.load_obj(obj, builder=0)
    cams2 = ref.turntable(ref.cam_array((-150, -20, 0), (1, -0.1, 0), (0, -1, 0), 60), [-500, 300, 10], [0, 1, 0], 2)
    for k in range(2):
        px, rays, hits, _ = s.render(cams2[k, :10], cams2[k, 10:13], W, H)
        x = px.astype(np.float32) * np.float32(65535)
        x = np.where(np.float32(65535) < x, np.float32(65535), x)
        want = np.where(x < 0, np.float32(0), x).astype(np.uint16)
        img = read_ppm(str(tmp_path / f"turn16_{k:04d}.ppm")).astype(np.int32)
        d = img - want.astype(np.int32)
        assert np.abs(d).max() <= 1 and (d != 0).mean() <= 1e-4 and img.max() > 255
    s.close()
    # -d (anim.cpp:154-155): loader, builder, camera path and render in double, against the reference's double build
    from oracle import f64
    if not f64.ref_available():
        return
    out = str(tmp_path / "turnd.mp4")
    r = subprocess.run([ANIM, "-d", "--input", obj, "--frames", "3", "--width", str(W), "--height", str(H), out],
                       capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    s = f64.RefScene64.load_obj(obj)
    cams = f64.ref_turntable(f64.cam_array((-150, -20, 0), (1, -0.1, 0), (0, -1, 0), 60), [-500, 300, 10], [0, 1, 0], 3)
    for k in range(3):
        px, (rays, hits), _ = s.render(cams[k, :10], cams[k, 10:13], W, H)
        assert f"Rays: {rays}\tHits: {hits}" in r.stdout
        img = read_ppm(str(tmp_path / f"turnd_{k:04d}.ppm"))
        d = img.astype(np.int16) - quantise(px.astype(np.float32))[::-1].astype(np.int16)      # unflipped, as anim.cpp:117-119
        assert np.abs(d).max() <= 1 and (d != 0).mean() <= 1e-5
        assert (img != 0).any()
    s.close()


@pytest.mark.gpu
@needs_binaries
@pytest.mark.skipif(not os.path.exists(SHIM_CHECK), reason="apps/_build/shim_check not built")
def test_drop_in_header_renders_float_and_double_like_the_reference(tmp_path):
    """render<float>() and render<double>() called through include/ceres/render.hpp by a caller that is otherwise the
    reference's own code (loader, SweepSah builder): counters equal, frame sums equal to 1e-6 relative."""
    import re
    from oracle import f64
    obj = str(tmp_path / "torus.obj")
    write_torus_obj(obj, 64, 32)
    with open(obj, "rb") as f:
        txt = f.read().decode().replace("\r\n", "\n").split("\n")
    with open(obj, "w") as f:
        for line in txt:
            if line.startswith("v "):
                a, b, c = (float(t) * 40 for t in line.split()[1:4])
                f.write(f"v {a:.5f} {b:.5f} {c:.5f}\n")
            else:
                f.write(line + "\n")
    W, H = 200, 320
    r = subprocess.run([SHIM_CHECK, obj, str(W), str(H)], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    got = {}
    for m in re.finditer(r"(float|double) pass (\d): rays (\d+) hits (\d+) lit (\d+) negative (\d+) sum (\S+)", r.stdout):
        got[(m.group(1), int(m.group(2)))] = (int(m.group(3)), int(m.group(4)), int(m.group(5)), int(m.group(6)), float(m.group(7)))
    assert len(got) == 4, r.stdout
    s32 = ref.RefScene.load_obj(obj, builder=0)
    s64 = f64.RefScene64.load_obj(obj) if f64.ref_available() else None
    sun = [-500, 300, 10]
    for k in range(2):
        eye = (-150 + 5 * k, -20, 0)
        px, rays, hits, _ = s32.render(ref.cam_array(eye, (1, -0.1, 0), (0, -1, 0), 60), sun, W, H)
        g = got[("float", k)]
        assert g[:2] == (rays, hits) and g[3] == 0 and g[2] == int((px > 0).sum())
        assert abs(g[4] - float(px.astype(np.float64).sum())) <= 1e-6 * max(1.0, abs(g[4]))
        if s64 is not None:
            px, (rays, hits), _ = s64.render(f64.cam_array(eye, (1, -0.1, 0), (0, -1, 0), 60), np.array(sun, np.float64), W, H)
            g = got[("double", k)]
            assert g[:2] == (rays, hits) and g[3] == 0 and g[2] == int((px > 0).sum())
            assert abs(g[4] - float(px.sum())) <= 1e-6 * max(1.0, abs(g[4]))
    assert got[("float", 0)][1] > 0 and got[("float", 0)] != got[("float", 1)]
    s32.close()


@needs_binaries
def test_binaries_fail_loudly_without_gpu_or_input(tmp_path):
    r = subprocess.run([RENDER, str(tmp_path / "missing.obj")], capture_output=True, text=True)
    assert r.returncode != 0 and "empty or cannot be loaded" in r.stderr
    r = subprocess.run([RENDER, "--bogus"], capture_output=True, text=True)
    assert r.returncode == 2
