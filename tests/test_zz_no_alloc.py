"""ceres_render_device_static (include/ceres_b200.h): the frame kernel over the library's static device scratch, for callers
that may not allocate.  No counterpart in the reference.

Checked through the C ABI against ceres_render_device on the same inputs, with the bar the other device-path tests use
(tests/test_gpu_parity.py): pixels, ids and the (rays, hits) counters BIT-EXACT.  The file sorts last on purpose
(a failure here hides nothing behind it) and does not depend on what ran before it."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

SUN = np.array([-50, -20, 0], np.float32)


@pytest.fixture(scope="module")
def cr():
    import ceres_raytracer_b200 as cr
    return cr


@pytest.fixture(scope="module")
def bunny(cr, bunny_pod):
    s = cr.Scene.from_pod(bunny_pod)
    yield s
    s.close()


@pytest.fixture(scope="module")
def dragon(cr, dragon_pod):
    s = cr.Scene.from_pod(dragon_pod)
    yield s
    s.close()


def bunny_cameras(cr, pod, n=4, seed=11):
    b = pod["nodes"][0]["bounds"]                    # root box: min/max per axis (bvh.hpp:25-30)
    centre = np.array([(b[0] + b[1]) / 2, (b[2] + b[3]) / 2, (b[4] + b[5]) / 2])
    size = float(max(b[1] - b[0], b[3] - b[2], b[5] - b[4]))
    rng = np.random.default_rng(seed)
    cams = []
    for k in range(n):
        d = rng.normal(size=3)
        d /= np.linalg.norm(d)
        eye = centre - d * size * float(rng.uniform(0.8, 2.5))
        cams.append(cr.Camera(eye, d + rng.normal(size=3) * (0.0 if k % 2 == 0 else 0.3), rng.normal(size=3), float(rng.choice([30, 60, 100]))))
    return cams


def both(scene, cam, W, H, flags, want_ids=True, want_counters=True):
    """(pixels, ids, counters) of ceres_render_device and of ceres_render_device_static on the same inputs."""
    import torch
    out = []
    for static in (False, True):
        d_px = torch.full((H * W * 3,), float("nan"), dtype=torch.float32, device="cuda")      # the call owns every pixel
        d_ids = torch.full((H * W,), 12345, dtype=torch.int32, device="cuda") if want_ids else None
        d_cnt = torch.full((2,), -7, dtype=torch.int64, device="cuda") if want_counters else None
        if static:
            scene.render_device_static(cam, SUN, W, H, flags=flags, pixels=d_px, ids=d_ids, counters=d_cnt)
        else:
            scene.render_device(cam, SUN, W, H, flags=flags, pixels=d_px, ids=d_ids, counters=d_cnt)
        torch.cuda.synchronize()
        out.append((d_px.cpu().numpy().view(np.uint32),
                    d_ids.cpu().numpy().view(np.uint32) if want_ids else None,
                    tuple(d_cnt.cpu().tolist()) if want_counters else None))
    return out


def assert_same(ref, got):
    assert np.array_equal(ref[0], got[0]), f"{int((ref[0] != got[0]).sum())} pixel words differ"
    if ref[1] is not None:
        assert np.array_equal(ref[1], got[1]), f"{int((ref[1] != got[1]).sum())} hit ids differ"
    assert ref[2] == got[2]


def test_two_calls_in_a_row_match_render_device(cr, bunny, bunny_pod):
    cams = bunny_cameras(cr, bunny_pod)
    # a large frame first, then a smaller one: the second call runs over whatever the first left in the scratch
    ref, got = both(bunny, cams[0], 640, 360, 0)
    assert_same(ref, got)
    assert ref[2][1] > 0, "the camera should see the model"
    ref, got = both(bunny, cams[1], 97, 61, 0)
    assert_same(ref, got)
    # and on: without shadows, without ids / counters, the first camera again
    ref, got = both(bunny, cams[2], 256, 144, cr.NO_SHADOWS)
    assert_same(ref, got)
    ref, got = both(bunny, cams[3], 333, 16, 0, want_ids=False, want_counters=False)
    assert_same(ref, got)
    ref, got = both(bunny, cams[0], 640, 360, 0)
    assert_same(ref, got)


def test_back_to_back_on_one_stream_without_a_sync(cr, bunny, bunny_pod):
    """Two static calls enqueued one behind the other: the second re-arms the scratch in stream order."""
    import torch
    cams = bunny_cameras(cr, bunny_pod, n=2, seed=5)
    W, H = 320, 200
    refs = []
    for cam in cams:
        d_px = torch.zeros(H * W * 3, dtype=torch.float32, device="cuda")
        d_cnt = torch.zeros(2, dtype=torch.int64, device="cuda")
        bunny.render_device(cam, SUN, W, H, pixels=d_px, counters=d_cnt)
        torch.cuda.synchronize()
        refs.append((d_px.cpu().numpy().view(np.uint32), tuple(d_cnt.cpu().tolist())))
    outs = [(torch.full((H * W * 3,), float("nan"), dtype=torch.float32, device="cuda"), torch.zeros(2, dtype=torch.int64, device="cuda")) for _ in cams]
    for _ in range(3):
        for cam, (d_px, d_cnt) in zip(cams, outs):
            bunny.render_device_static(cam, SUN, W, H, pixels=d_px, counters=d_cnt)
    torch.cuda.synchronize()
    for (rpx, rcnt), (d_px, d_cnt) in zip(refs, outs):
        assert np.array_equal(rpx, d_px.cpu().numpy().view(np.uint32)) and rcnt == tuple(d_cnt.cpu().tolist())
    assert bunny.frame_error() == 0


def test_largest_frame_and_the_limit(cr, dragon):
    """width * height == STATIC_MAX_PIXELS renders; one row more is refused with a message that names the limit."""
    import torch
    assert cr.STATIC_MAX_PIXELS == 1 << 20
    cam = cr.configs.DRAGON_CAMERA
    ref, got = both(dragon, cam, 1024, 1024, 0)
    assert_same(ref, got)
    ref, got = both(dragon, cam, 1280, 720, 0)
    assert_same(ref, got)
    d_px = torch.zeros(8, dtype=torch.float32, device="cuda")      # never touched: the size check comes first
    with pytest.raises(cr.CeresError, match="CERES_STATIC_MAX_PIXELS"):
        dragon.render_device_static(cam, SUN, 1024, 1025, pixels=d_px)
    with pytest.raises(cr.CeresError, match="CERES_STATIC_MAX_PIXELS"):
        dragon.render_device_static(cam, SUN, 3840, 2160, pixels=d_px)
    with pytest.raises(cr.CeresError, match="CERES_NO_SHADOWS"):
        dragon.render_device_static(cam, SUN, 64, 64, flags=cr.COMPACT_ROWS, pixels=d_px)
    # a refused call leaves the entry point usable
    ref, got = both(dragon, cam, 640, 480, cr.NO_SHADOWS)
    assert_same(ref, got)


def test_the_call_allocates_no_device_memory(cr, dragon):
    """Free device memory as the driver reports it does not move over a run of calls (after the first one, which loads
    what the CUDA runtime loads lazily)."""
    import torch
    cam, W, H = cr.configs.DRAGON_CAMERA, 1024, 768
    d_px = torch.zeros(H * W * 3, dtype=torch.float32, device="cuda")
    d_ids = torch.zeros(H * W, dtype=torch.int32, device="cuda")
    d_cnt = torch.zeros(2, dtype=torch.int64, device="cuda")
    dragon.render_device_static(cam, SUN, W, H, pixels=d_px, ids=d_ids, counters=d_cnt)
    dragon.render_device_static(cam, SUN, 200, 100, pixels=d_px, ids=d_ids, counters=d_cnt)
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    for k in range(20):
        w, h = (W, H) if k % 2 == 0 else (512 + 8 * k, 300 + k)
        dragon.render_device_static(cam, SUN, w, h, pixels=d_px, ids=d_ids, counters=d_cnt)
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    assert free1 == free0, f"free device memory moved by {free0 - free1} bytes over 20 calls"


def test_the_call_can_be_captured_in_a_cuda_graph(cr, dragon):
    """The reason the entry point exists: a caller under CUDA graph capture.  After one warm-up call (which takes the scratch's
    address) the call is captured on a stream -- nothing in it allocates or synchronises -- and the graph replayed three times
    gives the frame of ceres_render_device every time."""
    import torch
    cam, W, H = cr.configs.DRAGON_CAMERA, 1280, 720
    ref = both(dragon, cam, W, H, 0)[0]
    px = torch.full((H * W * 3,), float("nan"), dtype=torch.float32, device="cuda")
    ids = torch.zeros(H * W, dtype=torch.int32, device="cuda")
    cnt = torch.zeros(2, dtype=torch.int64, device="cuda")
    dragon.render_device_static(cam, SUN, W, H, pixels=px, ids=ids, counters=cnt)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        dragon.render_device_static(cam, SUN, W, H, pixels=px, ids=ids, counters=cnt)
    for _ in range(3):
        px.fill_(float("nan")); ids.zero_(); cnt.zero_()
        g.replay()
        torch.cuda.synchronize()
        assert_same(ref, (px.cpu().numpy().view(np.uint32), ids.cpu().numpy().view(np.uint32), tuple(cnt.cpu().tolist())))
    del g
