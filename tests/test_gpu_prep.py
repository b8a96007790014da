"""GPU parity tests of the scene-preparation kernels (rotate_triangles, transform_triangles,
HierarchyRefitter::refit on the device) and of the instanced-grid stress scene (config 5) that is built with them.
Bars: triangles and node boxes BIT-EXACT against the oracle (which tests/test_oracle.py pins to the reference's own
functions); frames rendered from the prepared scene meet the bars of test_gpu_parity.py."""
import numpy as np
import pytest

from oracle import port
from test_gpu_parity import check_frame, SUN

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def cr():
    import ceres_raytracer_b200 as cr
    return cr


def bits(a):
    return np.ascontiguousarray(a).view(np.uint32)


def test_refit_of_the_untouched_scene_reproduces_the_builders_boxes(cr, dragon_pod):
    with cr.Scene.from_pod(dragon_pod) as s:
        s.refit()
        nodes, tris = s.download()
    assert np.array_equal(bits(tris), bits(dragon_pod["triangles"]))
    assert np.array_equal(bits(nodes), bits(dragon_pod["nodes"]))


@pytest.mark.parametrize("axis,deg", [(0, 90.0), (1, -145.0), (2, 33.0)])
def test_rotate_refit_render(cr, bunny_pod, axis, deg):
    cam = cr.configs.BUNNY_CAMERA
    with cr.Scene.from_pod(bunny_pod) as s:
        s.rotate(axis, deg)
        s.refit()
        nodes, tris = s.download()
        want_tris = port.rotate_triangles(axis, deg, bunny_pod["triangles"])
        assert np.array_equal(bits(tris), bits(want_tris))
        moved = dict(bunny_pod, triangles=np.asarray(want_tris, np.float32).reshape(-1, 12))
        want_nodes = port.refit(moved)
        assert np.array_equal(bits(nodes), bits(want_nodes))
        moved["nodes"] = want_nodes
        check_frame(cr, s, moved, cam, SUN, 640, 360, 0)


def test_transform_refit_render_twice(cr, dragon_pod):
    """Two successive Transform steps (a turntable step about z plus a translation), as the rotating-object mode of
    anim.cpp:83-84 would apply them frame after frame."""
    cam = cr.configs.DRAGON_CAMERA
    c, s_ = np.float32(np.cos(0.3)), np.float32(np.sin(0.3))
    a = np.array([[c, -s_, 0], [s_, c, 0], [0, 0, 1]], np.float32)
    v = np.array([0.25, -0.5, 0.125], np.float32)
    pod = dict(dragon_pod)
    with cr.Scene.from_pod(dragon_pod) as s:
        for step in range(2):
            s.transform(a, v)
            s.refit()
            nodes, tris = s.download()
            pod["triangles"] = port.transform_triangles(a, v, pod["triangles"])
            pod["nodes"] = port.refit(pod)
            assert np.array_equal(bits(tris), bits(pod["triangles"])), step
            assert np.array_equal(bits(nodes), bits(pod["nodes"])), step
            check_frame(cr, s, pod, cam, SUN, 960, 540, 0)


def test_instanced_grid_scene(cr, dragon_pod):
    grid, info = cr.scenes.instanced_grid(dragon_pod, 3, 2, 2)
    cam, sun = cr.scenes.grid_camera(info)
    sun = np.asarray(sun, np.float32)
    with cr.Scene.from_pod(grid) as s:
        s.refit()                                   # all boxes of the flat tree come from the device refit
        nodes, tris = s.download()
        assert np.array_equal(bits(tris), bits(grid["triangles"]))
        want = port.refit(grid)
        assert np.array_equal(bits(nodes), bits(want))
        grid["nodes"] = want
        check_frame(cr, s, grid, cam, sun, 1280, 720, 0)
        check_frame(cr, s, grid, cam, sun, 640, 360, cr.NO_SHADOWS)


def test_device_built_tree_of_the_10m_triangle_scene(cr, dragon_pod):
    """N3 at the size it exists for: the 10.5 M-triangle instanced-dragon scene (BASELINE config 5) built on the device by the
    default builder (top-down binned SAH).  Size-independent properties against the frame from the reference-built tree of the
    same triangles: everything a tie between two triangles at exactly the same t cannot change -- the two counters, which pixels
    hit, which pixels are lit -- is equal, and the ids differ on a few dozen pixels at most; the tree is small enough for the
    traversal stack and the build stays under the 50 ms bar (profiles/r2/sweeps.md: 38.5 ms)."""
    import bench
    pod, cam, sun, W, H, flags, scene = bench.load_workload("grid448_4k_shadow", device=0)
    try:
        n = len(pod["triangles"])
        assert n == 448 * len(dragon_pod["triangles"])
        px_a, ids_a, cnt_a = scene.render(cam, sun, W, H, flags=flags, want_ids=True)
        cr.build_options()
        with cr.Scene.from_lbvh(pod["triangles"], pod["norms"]) as s:
            assert s.node_count % 2 == 1 and s.node_count <= 2 * n - 1 and s.stack_depth <= 64
            assert 0 < s.build_ms < 50.0 * 2, s.build_ms          # twice the bar: a shared box may be slower than the measured 38.5 ms
            px_b, ids_b, cnt_b = s.render(cam, sun, W, H, flags=flags, want_ids=True)
        assert cnt_a == cnt_b and cnt_a[1] > 1_000_000
        assert np.array_equal(ids_a == cr.MISS, ids_b == cr.MISS)
        differ = ids_a != ids_b
        assert int(differ.sum()) <= 64, int(differ.sum())
        assert np.array_equal(px_a[~differ].view(np.uint32), px_b[~differ].view(np.uint32))
    finally:
        scene.close()
