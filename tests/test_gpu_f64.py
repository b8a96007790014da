"""GPU tests of the Scalar = double path (ceres_render_f64; `anim -d`, anim.cpp:154-155).
Bars: hit ids and the two counters bit-exact against the reference's own double-precision code (committed fixture
tests/golden/f64_bunny.npz) and against the C restatement (oracle/ceres_oracle_f64.c) on other scenes / cameras;
RGB within 1e-6 (the colours are float-valued; pow(x, 24) is evaluated by repeated squaring on the device)."""
import os

import numpy as np
import pytest

from conftest import GOLDEN, load_npz
from oracle import f64

pytestmark = pytest.mark.gpu
RGB_TOL = 1e-6


@pytest.fixture(scope="module")
def cr():
    import ceres_raytracer_b200 as cr
    return cr


@pytest.fixture(scope="module")
def gold():
    return load_npz(os.path.join(GOLDEN, "f64_bunny.npz"))


def compare(got, want):
    px, ids, cnt = got
    wpx, wids, wcnt = want
    assert np.array_equal(ids, wids)
    assert cnt == wcnt
    assert np.abs(px - wpx).max() <= RGB_TOL
    assert np.array_equal(px == 0, wpx == 0)                 # same lit / dark verdict per pixel
    return float((px == wpx).mean())


@pytest.mark.parametrize("k", [0, 1])
@pytest.mark.parametrize("flags", [0, 1])
def test_f64_frames_equal_the_reference_fixture(cr, gold, k, flags):
    W, H = (int(x) for x in gold["size"])
    with cr.SceneF64.from_pod(gold) as s:
        got = s.render(gold["cams"][k], gold["sun"], W, H, flags=flags, want_ids=True)
        assert s.kernel_ms > 0
    exact = compare(got, (gold[f"rgb_{k}_{flags}"].astype(np.float64), gold[f"ids_{k}_{flags}"], tuple(int(x) for x in gold[f"counters_{k}_{flags}"])))
    assert exact > 0.999


def test_f64_dragon_against_the_c_restatement(cr, dragon_pod):
    pod = f64.widen(dragon_pod)
    c = cr.configs.DRAGON_CAMERA
    cam = f64.cam_array(c.eye, c.dir, c.up, c.fov)
    sun = np.array([-50., -20., 0.])
    W, H = 640, 360
    with cr.SceneF64.from_pod(pod) as s:
        got = s.render(cam, sun, W, H, want_ids=True)
    px, ids, st, cnt = f64.port_render(pod, cam, sun, W, H)
    compare(got, (px, ids, cnt))
    assert (ids != f64.MISS).sum() > 10000 and (st == 2).sum() > 1000


def test_f64_is_not_the_float_path(cr, bunny_pod):
    """The same (float-representable) scene through both precisions: close, but double arithmetic decides some
    pixels differently -- the double path is a separate computation, not the float path widened."""
    cam, sun = cr.configs.BUNNY_CAMERA, (-50.0, -20.0, 0.0)
    with cr.Scene.from_pod(bunny_pod) as a, cr.SceneF64.from_pod(f64.widen(bunny_pod)) as b:
        pa, ia, ca = a.render(cam, sun, 480, 270, want_ids=True)
        pb, ib, cb = b.render(cam, sun, 480, 270, want_ids=True)
    assert (ia != ib).mean() < 0.01 and abs(ca[1] - cb[1]) < 0.01 * ca[1]
    assert np.abs(pa.astype(np.float64) - pb)[ia == ib].mean() < 1e-3


def test_f64_leaf_root_and_errors(cr):
    tri = np.zeros((1, 12))
    p0, p1, p2 = np.array([-1., -1, 5]), np.array([1., -1, 5]), np.array([0., 1, 5])
    tri[0, 0:3], tri[0, 3:6], tri[0, 6:9] = p0, p0 - p1, p2 - p0
    tri[0, 9:12] = np.cross(tri[0, 3:6], tri[0, 6:9])
    nodes = np.zeros(1, cr.NODE64_DT)
    nodes[0]["bounds"] = [-1, 1, -1, 1, 5, 5]
    nodes[0]["primitive_count"], nodes[0]["first_child_or_primitive"] = 1, 0
    pod = {"nodes": nodes, "prim_indices": np.zeros(1, np.uint64), "triangles": tri, "norms": np.tile([0., 0, -1], (1, 3))}
    cam, sun = f64.cam_array((0, 0, 0), (0, 0, 1), (0, 1, 0), 60.0), np.array([0., 0, -10])
    with cr.SceneF64.from_pod(pod) as s:
        got = s.render(cam, sun, 64, 48, want_ids=True)
    px, ids, st, cnt = f64.port_render(pod, cam, sun, 64, 48)
    compare(got, (px, ids, cnt))
    assert cnt[1] > 0
    with pytest.raises(ValueError):
        cr.SceneF64(np.zeros(1, cr.NODE_DT), np.zeros(1, np.uint64), tri)
    bad = nodes.copy()
    bad[0]["primitive_count"] = 0
    bad[0]["first_child_or_primitive"] = 7
    with pytest.raises(cr.CeresError):
        cr.SceneF64(bad, np.zeros(1, np.uint64), tri)


@pytest.mark.parametrize("count,block_rows,W,H", [(2, 16, 640, 360), (8, 16, 640, 360), (3, 5, 333, 97), (8, 64, 200, 100)])
def test_f64_row_block_shards_assemble_the_frame(cr, dragon_pod, count, block_rows, W, H):
    """The float path's row-block sharding on the double path (anim.cpp:146-155 at scale; SURVEY.md 8(e): render the N
    shards one after the other into ONE frame and require byte-equality with the unsharded frame).  Every shard writes all of
    its own rows (zeros included) and nothing else; the shards' counters add up to the frame's."""
    pod = f64.widen(dragon_pod)
    c = cr.configs.DRAGON_CAMERA
    cam = f64.cam_array(c.eye, c.dir, c.up, c.fov)
    sun = np.array([-50., -20., 0.])
    with cr.SceneF64.from_pod(pod) as s:
        ref_px, ref_ids, ref_cnt = s.render(cam, sun, W, H, want_ids=True)
        px = np.full((H, W, 3), np.nan)
        ids = np.full((H, W), 12345, np.uint32)
        rays = hits = 0
        for k in range(count):
            before_px, before_ids = px.copy(), ids.copy()
            _, _, (r, h) = s.render(cam, sun, W, H, pixels=px, ids=ids, shard=(block_rows, k, count))
            rows = cr.shard_rows(H, (block_rows, k, count))
            others = np.setdiff1d(np.arange(H), rows)
            assert np.array_equal(px[rows].view(np.uint64), ref_px[rows].view(np.uint64)) and np.array_equal(ids[rows], ref_ids[rows])
            assert np.array_equal(px[others].view(np.uint64), before_px[others].view(np.uint64)), "a shard wrote outside its rows"
            assert np.array_equal(ids[others], before_ids[others])
            assert r == len(rows) * W + int((ref_ids[rows] != cr.MISS).sum())
            rays += r
            hits += h
        assert np.array_equal(px.view(np.uint64), ref_px.view(np.uint64)) and np.array_equal(ids, ref_ids)
        assert (rays, hits) == ref_cnt
        with pytest.raises(cr.CeresError):
            s.render(cam, sun, W, H, pixels=px, shard=(0, 1, 2))
        with pytest.raises(cr.CeresError):
            s.render(cam, sun, W, H, pixels=px, shard=(16, 2, 2))
