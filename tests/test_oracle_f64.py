"""The Scalar = double checker (oracle/ceres_oracle_f64.c) against the reference: the committed fixture made by the
reference's own double-precision code (tests/golden/f64_bunny.npz) and, where oracle/_ref is present, that code itself."""
import os

import numpy as np
import pytest

from conftest import GOLDEN, load_npz
from oracle import f64


@pytest.fixture(scope="module")
def gold():
    return load_npz(os.path.join(GOLDEN, "f64_bunny.npz"))


def scene_of(g):
    return {k: g[k] for k in ("nodes", "prim_indices", "triangles", "norms")}


def test_double_layouts(gold):
    assert gold["nodes"].dtype.itemsize == 64 and gold["triangles"].dtype == np.float64 and gold["triangles"].shape[1] == 12
    assert gold["norms"].shape == (gold["triangles"].shape[0], 9)
    t = gold["triangles"]
    assert np.array_equal(np.cross(t[:, 3:6], t[:, 6:9]), t[:, 9:12])          # n = cross(e1, e2), triangle.hpp:30-34
    assert not np.array_equal(t, t.astype(np.float32).astype(np.float64))      # a genuine double scene


@pytest.mark.parametrize("k", [0, 1])
@pytest.mark.parametrize("flags", [0, 1])
def test_port_f64_equals_reference_fixture(gold, k, flags):
    W, H = (int(x) for x in gold["size"])
    px, ids, st, cnt = f64.port_render(scene_of(gold), gold["cams"][k], gold["sun"], W, H, flags)
    assert np.array_equal(ids, gold[f"ids_{k}_{flags}"]) and np.array_equal(st, gold[f"state_{k}_{flags}"])
    assert cnt == tuple(int(x) for x in gold[f"counters_{k}_{flags}"])
    assert np.array_equal(px, gold[f"rgb_{k}_{flags}"].astype(np.float64))     # bit-exact: the colours are floats


@pytest.mark.skipif(not f64.ref_available(), reason="oracle/_ref/libceres_ref64_strict.so not built")
def test_port_f64_equals_reference_on_a_widened_float_scene(bunny_pod):
    pod = f64.widen(bunny_pod)
    cam, sun = f64.cam_array((0.05, .12, -.28), (-.1, -.05, 1), (0, 1, 0), 50.0), np.array([-50., -20., 0.])
    rs = f64.RefScene64.from_pod(pod)
    for flags in (0, 1):
        a, b = rs.render_ex(cam, sun, 320, 200, flags), f64.port_render(pod, cam, sun, 320, 200, flags)
        assert np.array_equal(a[1], b[1]) and np.array_equal(a[2], b[2]) and a[3] == b[3]
        assert np.array_equal(a[0].view(np.uint64), b[0].view(np.uint64))
    plain, cnt, _ = rs.render(cam, sun, 320, 200)
    a = rs.render_ex(cam, sun, 320, 200, 0)
    assert np.array_equal(plain.view(np.uint64), a[0].view(np.uint64)) and cnt == a[3]


def test_frame_setup_f64_is_double_arithmetic():
    cam = f64.cam_array((0, 0, 0), (0.3, 0.1, 1), (0, 1, 0), 60.0)
    fs = f64.port_frame_setup(cam, 1920, 1080)
    d = np.array([0.3, 0.1, 1.0])
    inv = 1.0 / np.sqrt(d[0] * d[0] + d[1] * d[1] + d[2] * d[2])
    assert np.array_equal(fs[:3], d * inv)
    assert abs(np.linalg.norm(fs[3:6]) - np.tan(60.0 * (3.14159265 * (1.0 / 180.0) * 0.5))) < 1e-15
