"""GPU parity tests: the CUDA path, called through the C ABI, against the oracle on the same inputs.

Bars (BASELINE.json north_star): per-pixel hit triangle IDs, hit/miss/occlusion states, t/u/v and the (rays, hits)
counters are BIT-EXACT against the strict oracle; shaded RGB within RGB_ATOL absolute per channel (the only
non-bit-exact step is pow(double, 24), evaluated by repeated squaring on the GPU and by glibc pow in the
reference; both are then narrowed to float); PPM bytes equal except for <= PPM_MAX_DIFF_FRACTION of bytes off by
one LSB."""
import hashlib
import os

import numpy as np
import pytest

from conftest import scene_pod
from oracle import port

pytestmark = pytest.mark.gpu

RGB_ATOL = 1e-6
PPM_MAX_DIFF_FRACTION = 1e-5
SUN = np.array([-50, -20, 0], np.float32)


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


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


def check_frame(cr, scene, pod, cam, sun, W, H, flags, golden=None):
    px, ids, (rays, hits) = scene.render(cam, sun, W, H, flags=flags, want_ids=True)
    o = port.render(pod, cam.as_array(), sun, W, H, flags)
    assert np.array_equal(ids, o["ids"]), f"{int((ids != o['ids']).sum())} hit ids differ"
    assert (rays, hits) == (o["stats"]["rays"], o["stats"]["hits"])
    lit_gpu = (px != 0).any(axis=2)
    # every lit pixel of the oracle has a non-zero colour (ambient term), so the state can be compared exactly
    assert np.array_equal(lit_gpu, o["state"] == 1)
    d = np.abs(px.astype(np.float64) - o["pixels"].astype(np.float64))
    assert d.max() <= RGB_ATOL, f"max |dRGB| = {d.max()}"
    exact = float((px.view(np.uint32) == o["pixels"].view(np.uint32)).mean())
    b_gpu = cr.write_ppm(os.devnull, px)
    b_cpu = cr.write_ppm(os.devnull, o["pixels"])
    diff = b_gpu.astype(np.int16) - b_cpu.astype(np.int16)
    assert np.abs(diff).max() <= 1 and (diff != 0).mean() <= PPM_MAX_DIFF_FRACTION
    if golden is not None:
        assert sha(ids) == golden["ids_sha256"]
        assert (rays, hits) == (golden["stats"]["rays"], golden["stats"]["hits"])
    return exact


def test_trace_rays_kat(bunny, kat):
    ids, tuv = bunny.trace_rays(kat["trace_rays"])
    assert np.array_equal(ids, kat["trace_ids"])
    assert np.array_equal(tuv.view(np.uint32), kat["trace_tuv"].view(np.uint32))


@pytest.mark.parametrize("tag,flags", [("shadow", 0), ("primary", 1)])
def test_bunny_golden_frames(cr, bunny, bunny_frames, tag, flags):
    cam = cr.Camera(bunny_frames["cam"][0:3], bunny_frames["cam"][3:6], bunny_frames["cam"][6:9], float(bunny_frames["cam"][9]))
    px, ids, (rays, hits) = bunny.render(cam, bunny_frames["sun"], 320, 180, flags=flags, want_ids=True)
    assert np.array_equal(ids, bunny_frames[f"{tag}_ids"])
    assert np.abs(px - bunny_frames[f"{tag}_rgb"]).max() <= RGB_ATOL
    st = bunny_frames[f"{tag}_stats"]
    assert (rays, hits) == (int(st[6]), int(st[7]))


@pytest.mark.parametrize("W,H,flags", [(1920, 1080, 1), (1920, 1080, 0), (1, 1, 0), (33, 7, 0), (1000, 9, 1)])
def test_bunny_vs_oracle(cr, bunny, bunny_pod, W, H, flags):
    check_frame(cr, bunny, bunny_pod, cr.configs.BUNNY_CAMERA, SUN, W, H, flags)


@pytest.mark.parametrize("W,H,flags,key", [
    (1920, 1080, 1, "dragon_1920x1080_primary"),   # BASELINE config 2
    (1920, 1080, 0, "dragon_1920x1080_shadow"),
    (3840, 2160, 0, "dragon_3840x2160_shadow"),    # BASELINE config 3 (metric config)
])
def test_dragon_vs_oracle(cr, dragon, dragon_pod, frame_hashes, W, H, flags, key):
    exact = check_frame(cr, dragon, dragon_pod, cr.configs.DRAGON_CAMERA, SUN, W, H, flags, frame_hashes[key])
    assert exact > 0.999


def test_inside_and_degenerate_cameras(cr, dragon, dragon_pod):
    # camera inside the model's bounding box, axis-aligned view (zero direction components -> safe_inverse path)
    for cam in (cr.Camera((0, 0, 2), (0, 0, -1), (0, 1, 0), 90), cr.Camera((0, -3, 2), (1, 0, 0), (0, 0, 1), 120),
                cr.Camera((0, -15, 2), (0, 1, 0), (0, 0, 1), 1.0)):
        check_frame(cr, dragon, dragon_pod, cam, np.array([10, -30, 40], np.float32), 640, 360, 0)


def test_random_cameras_screen_rectangle_culling(cr, bunny, bunny_pod):
    """The primary wave only traces the conservative screen rectangle of the scene and the host path only copies that
    rectangle (everything else is filled with zeros on the host): whole frames must still equal the oracle's for
    cameras that see the model fully, partly, from inside its box, grazing, or not at all."""
    rng = np.random.default_rng(2024)
    b = bunny_pod["nodes"][0]["bounds"]
    centre = np.array([(b[0] + b[1]) / 2, (b[2] + b[3]) / 2, (b[4] + b[5]) / 2])
    size = float(max(b[1] - b[0], b[3] - b[2], b[5] - b[4]))
    cams = []
    for k in range(28):
        d = rng.normal(size=3); d /= np.linalg.norm(d)
        dist = size * (0.2 + 3.0 * rng.random())
        eye = centre - d * dist
        look = d + rng.normal(size=3) * (0.0 if k % 4 == 0 else 0.5)       # some cameras look off to the side
        up = rng.normal(size=3)
        cams.append(cr.Camera(eye, look, up, float(rng.choice([20, 60, 100, 150]))))
    cams.append(cr.Camera(centre + np.array([0, 0, 5 * size]), (0, 0, 1), (0, 1, 0), 60))        # model behind the camera
    cams.append(cr.Camera(centre + np.array([0, 0, -5 * size]), (1, 0, 0), (0, 1, 0), 30))       # model off screen
    cams.append(cr.Camera(centre, (0, 0, 1), (0, 1, 0), 90))                                      # eye inside the model's box
    sun = np.array([3, 10, -4], np.float32)
    shapes = [(256, 144), (97, 61), (16, 333)]
    for k, cam in enumerate(cams):
        W, H = shapes[k % 3]
        check_frame(cr, bunny, bunny_pod, cam, sun, W, H, k % 2)
    # page-locked destination (band-pipelined DMA of the rectangle) and the device path give the same frame
    import torch
    with cr.PinnedBuffer((144, 256, 3)) as hb:
        for cam in cams[:6]:
            ref_px, ref_ids, ref_cnt = bunny.render(cam, sun, 256, 144, want_ids=True)
            hb.array[:] = np.nan
            px, _, cnt = bunny.render(cam, sun, 256, 144, pixels=hb.array)
            assert np.array_equal(px.view(np.uint32), ref_px.view(np.uint32)) and cnt == ref_cnt
            d_px = torch.full((144 * 256 * 3,), float("nan"), dtype=torch.float32, device="cuda")
            d_ids = torch.zeros(144 * 256, dtype=torch.int32, device="cuda")
            d_cnt = torch.zeros(2, dtype=torch.int64, device="cuda")
            bunny.render_device(cam, sun, 256, 144, pixels=d_px, ids=d_ids, counters=d_cnt)
            torch.cuda.synchronize()
            assert np.array_equal(d_px.cpu().numpy().view(np.uint32), ref_px.ravel().view(np.uint32))
            assert np.array_equal(d_ids.cpu().numpy().view(np.uint32), ref_ids.ravel())
            assert tuple(d_cnt.cpu().tolist()) == ref_cnt


@pytest.mark.parametrize("sub", [1, 3, 8])
def test_subframes_do_not_change_the_frame(cr, dragon, sub):
    cam, W, H = cr.configs.DRAGON_CAMERA, 1280, 720
    ref = dragon.render(cam, SUN, W, H, want_ids=True)
    dragon.set_tuning(subframes=sub)
    try:
        with cr.PinnedBuffer((H, W, 3)) as hb:
            px, ids, cnt = dragon.render(cam, SUN, W, H, pixels=hb.array, want_ids=True)
            assert np.array_equal(px.view(np.uint32), ref[0].view(np.uint32)) and np.array_equal(ids, ref[1]) and cnt == ref[2]
    finally:
        dragon.set_tuning()


@pytest.mark.parametrize("count", [2, 3, 8])
def test_row_block_shards_reassemble_exactly(cr, dragon, count):
    cam, W, H = cr.configs.DRAGON_CAMERA, 1280, 720
    full, ids_full, cnt_full = dragon.render(cam, SUN, W, H, want_ids=True)
    frame = np.full_like(full, np.nan)
    ids = np.zeros_like(ids_full)
    rays = hits = 0
    for k in range(count):
        sh = (16, k, count)
        # in-place rows
        px, idk, (r, h) = dragon.render(cam, SUN, W, H, shard=sh, want_ids=True)
        rows = cr.shard_rows(H, sh)
        frame[rows] = px[rows]
        ids[rows] = idk[rows]
        rays += r; hits += h
        # compact rows
        pc, ic, _ = dragon.render(cam, SUN, W, H, flags=cr.COMPACT_ROWS, shard=sh, want_ids=True)
        assert np.array_equal(pc.view(np.uint32), px[rows].view(np.uint32)) and np.array_equal(ic, idk[rows])
    assert np.array_equal(frame.view(np.uint32), full.view(np.uint32))
    assert np.array_equal(ids, ids_full)
    assert (rays, hits) == cnt_full


def test_device_api_sequence_and_quantize(cr, dragon, dragon_pod, kat):
    import torch
    W, H = 640, 360
    cams = np.concatenate([kat["turntable"][:4, :]], axis=0)
    px_all, cnt = dragon.render_sequence(cams, W, H)
    for k in range(cams.shape[0]):
        c = cams[k]
        cam = cr.Camera(c[0:3], c[3:6], c[6:9], float(c[9]))
        px, _, counters = dragon.render(cam, c[10:13], W, H)
        assert np.array_equal(px.view(np.uint32), px_all[k].view(np.uint32))
        assert counters == (int(cnt[k, 0]), int(cnt[k, 1]))
        o = port.render(dragon_pod, c[0:10], c[10:13], W, H, 0)
        assert counters == (o["stats"]["rays"], o["stats"]["hits"])
    # device-resident frame + on-device 8-bit flip/quantise == PPM payload of static.cpp:135-147
    d_px = torch.empty(H * W * 3, dtype=torch.float32, device="cuda")
    d_cnt = torch.zeros(2, dtype=torch.int64, device="cuda")
    d_q = torch.empty(H * W * 3, dtype=torch.uint8, device="cuda")
    cam = cr.configs.DRAGON_CAMERA
    dragon.render_device(cam, SUN, W, H, pixels=d_px, counters=d_cnt)
    dragon.quantize_device(d_px, W, H, d_q)
    torch.cuda.synchronize()
    px, _, counters = dragon.render(cam, SUN, W, H)
    assert np.array_equal(d_px.cpu().numpy().view(np.uint32), px.ravel().view(np.uint32))
    assert tuple(d_cnt.cpu().tolist()) == counters
    assert np.array_equal(d_q.cpu().numpy().reshape(H, W, 3), cr.write_ppm(os.devnull, px))


def test_gpu_statistics_match_oracle_for_primary_rays(cr, dragon, dragon_pod):
    W, H = 960, 540
    o = port.render(dragon_pod, cr.configs.DRAGON_CAMERA.as_array(), SUN, W, H, 0, want_pixels=False)["stats"]
    got = {}
    for order in (1, 2):       # any-hit shadow rays: nearer child first (the reference's order) / left child first
        dragon.set_tuning(shadow_order=order, share=1)      # sharing off: a shared walk may look into subtrees a closest-hit walk prunes
        dragon.set_stats(True)
        try:
            dragon.render(cr.configs.DRAGON_CAMERA, SUN, W, H)
            got[order] = st = dragon.get_stats()
        finally:
            dragon.set_stats(False)
            dragon.set_tuning()
        # same tree walked in the same order: every primary ray that hits does exactly the oracle's work
        for k in ("T_miss", "S_hit", "T_hit", "rays", "hits"):
            assert st[k] == o[k], (order, k)
        assert 0 < st["S_shadow"] <= o["S_shadow"] and 0 < st["T_shadow"] <= o["T_shadow"]   # any-hit exits early
        # culling only removes rays that fail at the root: at most one root step per missing pixel is skipped
        assert st["S_miss"] < o["S_miss"]
        assert o["S_miss"] - st["S_miss"] <= W * H - (o["rays"] - W * H)
    assert got[1]["S_shadow"] != got[2]["S_shadow"]       # the order knob does change the walk ...


@pytest.mark.parametrize("order", [1, 2])
@pytest.mark.parametrize("ctas", [0, 3])
def test_shadow_order_and_grid_size_do_not_change_the_frame(cr, dragon, dragon_pod, order, ctas):
    """... but not the frame: the boolean of an any-hit ray is independent of the child order (a leaf is reached iff its
    own box and every box above it pass the slab test), and the persistent grid size only changes who traces what."""
    cam, W, H = cr.configs.DRAGON_CAMERA, 1280, 720
    ref = dragon.render(cam, SUN, W, H, want_ids=True)
    dragon.set_tuning(shadow_order=order, ctas_per_sm=ctas)
    try:
        px, ids, cnt = dragon.render(cam, SUN, W, H, want_ids=True)
        assert np.array_equal(px.view(np.uint32), ref[0].view(np.uint32)) and np.array_equal(ids, ref[1]) and cnt == ref[2]
        check_frame(cr, dragon, dragon_pod, cam, SUN, 640, 360, 0)
    finally:
        dragon.set_tuning()


@pytest.mark.parametrize("share,period,min_idle", [(1, 0, 0), (2, 0, 0), (2, 1, 1), (2, 2, 31), (2, 7, 12)])
def test_shadow_subtree_sharing_does_not_change_the_frame(cr, dragon, dragon_pod, bunny, bunny_pod, share, period, min_idle):
    """Idle lanes of a shadow chunk walk subtrees taken from the busy lanes' stacks (frame_fused.cuh, SHARE): the boolean
    of render.hpp:137-140 cannot depend on who walks what, so frames, ids and counters are those of the unshared walk
    and of the oracle, whatever the hand-over period and threshold."""
    cam, W, H = cr.configs.DRAGON_CAMERA, 1280, 720
    dragon.set_tuning(share=1)
    ref = dragon.render(cam, SUN, W, H, want_ids=True)
    dragon.set_tuning(share=share, share_period=period, share_min_idle=min_idle)
    bunny.set_tuning(share=share, share_period=period, share_min_idle=min_idle)
    try:
        px, ids, cnt = dragon.render(cam, SUN, W, H, want_ids=True)
        assert np.array_equal(px.view(np.uint32), ref[0].view(np.uint32)) and np.array_equal(ids, ref[1]) and cnt == ref[2]
        check_frame(cr, dragon, dragon_pod, cam, SUN, 640, 360, 0)
        check_frame(cr, bunny, bunny_pod, cr.configs.BUNNY_CAMERA, np.array(cr.configs.BUNNY_SUN, np.float32), 480, 270, 0)
    finally:
        dragon.set_tuning()
        bunny.set_tuning()


def test_shared_shadow_walk_statistics(cr, dragon):
    """Statistics of the shared walk: it does at least the unshared walk's node steps for the rays that reach the sun (every
    reachable node is visited exactly once by somebody) and its counters are the same."""
    cam, W, H = cr.configs.DRAGON_CAMERA, 960, 540
    got = {}
    for share in (1, 2):
        dragon.set_tuning(share=share, subframes=1)
        dragon.set_stats(True)
        try:
            dragon.render(cam, SUN, W, H)
            got[share] = dragon.get_stats()
            dragon.set_stats(False)
            dragon.set_timeline(True)
            dragon.render(cam, SUN, W, H)
            tl = dragon.timeline()
        finally:
            dragon.set_stats(False)
            dragon.set_timeline(False)
            dragon.set_tuning()
        assert tl.shape[1] == 12 and len(tl) > 0 and (tl[:, 2] >= tl[:, 1]).all() and (tl[:, 1] >= tl[:, 0]).all()
        assert int(tl[:, 4].sum()) == (got[share]["rays"] - W * H + 31) // 32       # every queue chunk was taken by one warp
    for k in ("S_miss", "T_miss", "S_hit", "T_hit", "rays", "hits"):
        assert got[1][k] == got[2][k], k
    assert got[2]["S_shadow"] > 0 and got[2]["T_shadow"] > 0


def test_wave_pipeline_variants_are_not_in_the_product_library(cr, dragon):
    """ceres_tuning.variant 1-6 (the wave pipelines of round 1) exist only in the -DCERES_EXPERIMENTS build."""
    if os.environ.get("CERES_B200_EXPERIMENTS", "") not in ("", "0"):
        pytest.skip("running against the experiments build")
    dragon.set_tuning(variant=6)
    try:
        with pytest.raises(cr.CeresError):
            dragon.render(cr.configs.DRAGON_CAMERA, SUN, 320, 180)
    finally:
        dragon.set_tuning()
    dragon.render(cr.configs.DRAGON_CAMERA, SUN, 320, 180)       # and the scene is still usable afterwards


def test_tiny_scenes_root_leaf_and_single_pair(cr):
    # 1 triangle -> the root is a leaf (single_ray_traverser.hpp:72-73); hand-built POD, no builder needed
    tri = port.make_triangles(np.array([[-1, -1, 5, 1, -1, 5, 0, 1, 5]], np.float32))
    nodes = np.zeros(1, cr.NODE_DT)
    nodes[0]["bounds"] = [-1, 1, -1, 1, 5, 5]; nodes[0]["primitive_count"] = 1
    pod = {"nodes": nodes, "prim_indices": np.array([0], np.uint64), "triangles": tri,
           "norms": np.tile(np.array([0, 0, -1], np.float32), (1, 3))}
    cam = cr.Camera((0, 0, 0), (0, 0, 1), (0, 1, 0), 60)
    with cr.Scene.from_pod(pod) as s:
        check_frame(cr, s, pod, cam, np.array([0, 0, -10], np.float32), 64, 48, 0)
    # 2 triangles under one inner root; the second exactly behind the first (shadowing) and coplanar duplicates
    tri2 = port.make_triangles(np.array([[-1, -1, 5, 1, -1, 5, 0, 1, 5], [-1, -1, 5, 1, -1, 5, 0, 1, 5],
                                         [-2, -2, 3, 2, -2, 3, 0, 2, 3]], np.float32))
    nodes = np.zeros(3, cr.NODE_DT)
    nodes[0]["bounds"] = [-2, 2, -2, 2, 3, 5]; nodes[0]["first_child_or_primitive"] = 1
    nodes[1]["bounds"] = [-1, 1, -1, 1, 5, 5]; nodes[1]["primitive_count"] = 2; nodes[1]["first_child_or_primitive"] = 0
    nodes[2]["bounds"] = [-2, 2, -2, 2, 3, 3]; nodes[2]["primitive_count"] = 1; nodes[2]["first_child_or_primitive"] = 2
    pod = {"nodes": nodes, "prim_indices": np.array([1, 0, 2], np.uint64), "triangles": tri2,
           "norms": np.tile(np.array([0, 0, -1], np.float32), (3, 3))}
    with cr.Scene.from_pod(pod) as s:
        check_frame(cr, s, pod, cam, np.array([0, 0, 100], np.float32), 64, 48, 0)
        check_frame(cr, s, pod, cr.Camera((0, 0, 10), (0, 0, -1), (0, 1, 0), 60), np.array([0, 0, -100], np.float32), 64, 48, 0)


def test_invalid_arguments_fail_loudly(cr, bunny_pod):
    bad = dict(bunny_pod)
    bad["prim_indices"] = bunny_pod["prim_indices"].copy()
    bad["prim_indices"][0] = 10 ** 9
    with pytest.raises(cr.CeresError):
        cr.Scene.from_pod(bad)
    with cr.Scene.from_pod(bunny_pod) as s:
        with pytest.raises(cr.CeresError):
            s.render(cr.configs.BUNNY_CAMERA, SUN, 0, 10)
        rays = np.zeros((1, 8), np.float32); rays[0, 7] = np.nan
        with pytest.raises(cr.CeresError):
            s.trace_rays(rays)


@pytest.mark.parametrize("bands", [1, 2, 5])
def test_band_pipelined_host_render_is_identical(cr, dragon, bands):
    """ceres_render() renders the frame in row bands and copies each finished band while the next one renders; the
    result must not depend on the band count, on the destination being page-locked or pageable, or on sharding."""
    cam, W, H = cr.configs.DRAGON_CAMERA, 1000, 531
    ref_px, ref_ids, ref_cnt = dragon.render(cam, SUN, W, H, want_ids=True)       # pageable numpy buffer: one band
    with cr.PinnedBuffer((H, W, 3)) as hb:
        dragon.set_tuning(subframes=bands)
        try:
            hb.array[:] = np.nan
            px, ids, cnt = dragon.render(cam, SUN, W, H, pixels=hb.array, want_ids=True)
            assert np.array_equal(px.view(np.uint32), ref_px.view(np.uint32))
            assert np.array_equal(ids, ref_ids) and cnt == ref_cnt
            # sharded, rows scattered in place into the page-locked frame
            hb.array[:] = np.nan
            tot = [0, 0]
            for k in range(3):
                _, _, c = dragon.render(cam, SUN, W, H, pixels=hb.array, shard=(16, k, 3))
                tot[0] += c[0]; tot[1] += c[1]
            assert np.array_equal(hb.array.view(np.uint32), ref_px.view(np.uint32)) and tuple(tot) == ref_cnt
        finally:
            dragon.set_tuning()


@pytest.mark.parametrize("W,H", [(1280, 720), (333, 217)])
def test_ppm8_host_path_equals_quantised_float_frame(cr, dragon, bunny, W, H):
    """ceres_render_ppm8 (device quantise + flip, rectangle-only DMA, host zero fill) against the float frame pushed
    through the reference's PPM arithmetic (static.cpp:135-147) on the host: byte-identical."""
    for scene, cam in ((dragon, cr.configs.DRAGON_CAMERA), (bunny, cr.configs.BUNNY_CAMERA)):
        px, _, cnt = scene.render(cam, SUN, W, H)
        want = cr.write_ppm(os.devnull, px)
        with cr.PinnedBuffer((H, W, 3), np.uint8) as hb:
            hb.array[:] = 0xAB
            got, cnt8 = scene.render_ppm8(cam, SUN, W, H, out=hb.array)
            assert np.array_equal(got, want) and cnt8 == cnt
        got2, _ = scene.render_ppm8(cam, SUN, W, H)               # pageable destination
        assert np.array_equal(got2, want)


@pytest.mark.parametrize("W,H", [(1280, 720), (333, 217)])
def test_ppm16_host_path_equals_quantised_float_frame(cr, dragon, W, H):
    """ceres_render_ppm16 against the float frame pushed through the 16-bit arithmetic on the host
    (x 65535 as anim.cpp:117-119, clamped, truncated, rows flipped, big-endian): byte-identical."""
    cam = cr.configs.DRAGON_CAMERA
    px, _, cnt = dragon.render(cam, SUN, W, H)
    x = px.astype(np.float32) * np.float32(65535)
    x = np.where(np.float32(65535) < x, np.float32(65535), x)
    x = np.where(x < 0, np.float32(0), x)
    want = x.astype(np.uint16)[::-1].astype(">u2")
    got, cnt16 = dragon.render_ppm16(cam, SUN, W, H)
    assert cnt16 == cnt and np.array_equal(got.view(np.uint8), want.view(np.uint8))
    assert (got != 0).any() and dragon.last_d2h_bytes < got.nbytes
    # CERES_PAYLOAD_NO_FLIP: image row = frame row, as anim.cpp:117-119 fills its frames
    got_nf, _ = dragon.render_ppm16(cam, SUN, W, H, flags=cr.PAYLOAD_NO_FLIP)
    assert np.array_equal(got_nf.view(np.uint8), want[::-1].view(np.uint8))
    got8_nf, _ = dragon.render_ppm8(cam, SUN, W, H, flags=cr.PAYLOAD_NO_FLIP, shard=(16, 1, 3))
    full8, _ = dragon.render_ppm8(cam, SUN, W, H)
    rows = cr.shard_rows(H, (16, 1, 3))
    assert np.array_equal(got8_nf[rows], full8[::-1][rows]) and not got8_nf[np.setdiff1d(np.arange(H), rows)].any()
    # and the 8-bit path still agrees with its own definition after sharing the kernel
    got8, _ = dragon.render_ppm8(cam, SUN, W, H)
    assert np.array_equal(got8, cr.write_ppm(os.devnull, px))


def _random_rays(rng, n, lo, hi):
    """Same recipe as tests/golden/make_golden.py: origins around the scene box, directions on the sphere plus
    axis-aligned, signed-zero and tiny components (safe_inverse / signbit, vector.hpp:69-74), half of them aimed at
    the scene; some finite tmax and some tmin > 0."""
    ext = hi - lo
    o = (lo - ext) + rng.random((n, 3)) * (3 * ext)
    d = rng.normal(size=(n, 3))
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    k = n // 8
    d[:k, 0] = 0.0
    d[k:2 * k, 1] = -0.0
    d[2 * k:3 * k, 2] = 1e-9 * rng.choice([-1, 1], size=k)
    c = (lo + hi) / 2 + (rng.random((n, 3)) - 0.5) * ext * 0.8
    aim = rng.random(n) < 0.5
    v = c - o
    v /= np.linalg.norm(v, axis=1, keepdims=True)
    d[aim] = v[aim]
    rays = np.zeros((n, 8), np.float32)
    rays[:, 0:3], rays[:, 3:6] = o, d
    rays[:, 7] = np.finfo(np.float32).max
    rays[::7, 7] = (rng.random(len(rays[::7])) * np.linalg.norm(ext) * 2).astype(np.float32)
    rays[::11, 6] = (rng.random(len(rays[::11])) * np.linalg.norm(ext) * 0.5).astype(np.float32)
    return rays


@pytest.mark.parametrize("which", ["bunny", "dragon"])
def test_trace_rays_2p20_random_rays_bit_exact(cr, bunny, bunny_pod, dragon, dragon_pod, which):
    """SURVEY.md 8(d) unit KATs at full size: 2^20 seeded random rays through the slab test, the ordered traversal and
    the triangle test; hit primitive and the raw bit patterns of (t, u, v) against the strict C restatement."""
    scene, pod = (bunny, bunny_pod) if which == "bunny" else (dragon, dragon_pod)
    b = pod["nodes"][0]["bounds"].astype(np.float64)
    rays = _random_rays(np.random.default_rng(12345), 1 << 20, b[0::2], b[1::2])
    ids, tuv = scene.trace_rays(rays)
    want_ids, want_tuv = port.trace_rays(pod, rays)
    assert np.array_equal(ids, want_ids)
    assert np.array_equal(tuv.view(np.uint32), want_tuv.view(np.uint32))
    hit = ids != cr.MISS
    assert 0.2 < hit.mean() < 0.9


def test_sequence_two_frames_in_flight_equals_single_frames(cr, dragon_pod, kat):
    """ceres_render_sequence keeps two frames in flight (a second set of per-frame resources over the same geometry, odd
    frames on a second host thread): every frame must equal the same frame rendered alone, also after the geometry was
    moved and refitted (the second set is refreshed from the scene on every call), for an odd number of frames."""
    W, H = 512, 288
    cams = np.ascontiguousarray(kat["turntable"][:7, :])
    with cr.Scene.from_pod(dragon_pod) as s:
        for moved in (False, True):
            if moved:
                s.rotate(2, 11.0)
                s.refit()
            px_all, cnt = s.render_sequence(cams, W, H)
            for k in range(cams.shape[0]):
                c = cams[k]
                px, _, counters = s.render(cr.Camera(c[0:3], c[3:6], c[6:9], float(c[9])), c[10:13], W, H)
                assert np.array_equal(px.view(np.uint32), px_all[k].view(np.uint32)), (moved, k)
                assert counters == (int(cnt[k, 0]), int(cnt[k, 1]))
            assert (px_all != 0).any()


def test_retained_frame_contract(cr, dragon):
    """CERES_FRAME_RETAINED (opt-in): the caller hands in the buffers of the previous call, untouched; the library zeroes only
    what the previous frame's rectangles leave uncovered.  Every frame of a sequence with moving, shrinking, growing and
    disappearing rectangles must equal the frame rendered into a NaN-poisoned buffer WITHOUT the flag; the default path
    never trusts the buffer's contents."""
    W, H = 1280, 720
    cams = [cr.configs.DRAGON_CAMERA, cr.configs.DRAGON_CAMERA,
            cr.Camera((3.0, -15.0, 2.5), (-0.1, 1.0, 0.0), (0.0, 0.0, 1.0), 45.0),          # other rectangle, other band layout
            cr.Camera((0.5, -15.0, 2.0), (0.0, 1.0, 0.0), (0.0, 0.0, 1.0), 60.0),           # shifted by a few pixels
            cr.Camera((0.0, -40.0, 2.0), (0.0, 1.0, 0.0), (0.0, 0.0, 1.0), 60.0),           # far away: small rectangle
            cr.Camera((0.0, -15.0, 2.0), (0.0, -1.0, 0.0), (0.0, 0.0, 1.0), 60.0),          # looking away: no rectangle at all
            cr.configs.DRAGON_CAMERA,
            cr.Camera((0.0, -6.0, 2.0), (0.0, 1.0, 0.0), (0.0, 0.0, 1.0), 60.0)]            # close up: the rectangle is the frame
    hb = cr.PinnedBuffer((H, W, 3), np.float32)
    hi = cr.PinnedBuffer((H, W), np.uint32)
    try:
        hb.array.fill(np.nan)
        hi.array.fill(7)
        for k, cam in enumerate(cams):
            for flags in (0, cr.NO_SHADOWS):
                fresh = np.full((H, W, 3), np.nan, np.float32)
                want = dragon.render(cam, SUN, W, H, flags=flags, pixels=fresh, want_ids=True)
                got = dragon.render(cam, SUN, W, H, flags=flags | cr.FRAME_RETAINED, pixels=hb.array, ids=hi.array)
                assert np.array_equal(got[0].view(np.uint32), want[0].view(np.uint32)), (k, flags)
                assert np.array_equal(got[1], want[1]) and got[2] == want[2], (k, flags)
        # the flag is a promise: a caller who breaks it gets his own scribbling back (outside the rectangles) ...
        dragon.render(cams[0], SUN, W, H, flags=cr.FRAME_RETAINED, pixels=hb.array, ids=hi.array)
        hb.array[0, 0, :] = 5.0
        dragon.render(cams[0], SUN, W, H, flags=cr.FRAME_RETAINED, pixels=hb.array, ids=hi.array)
        assert hb.array[0, 0, 0] == 5.0
        # ... and the default call, which promises nothing, overwrites every pixel (render.hpp:107,117)
        dragon.render(cams[0], SUN, W, H, pixels=hb.array, ids=hi.array)
        assert hb.array[0, 0, 0] == 0.0
        # a call without the flag forgets the retained state: the next retained call re-establishes the whole frame
        hb.array[1, 1, :] = 9.0
        got = dragon.render(cams[0], SUN, W, H, flags=cr.FRAME_RETAINED, pixels=hb.array, ids=hi.array)
        assert hb.array[1, 1, 0] == 0.0
        # another size in the same memory: re-established as well
        small = hb.array.reshape(-1)[:3 * 640 * 360].reshape(360, 640, 3)
        small_ids = hi.array.reshape(-1)[:640 * 360].reshape(360, 640)
        small.fill(np.nan)
        got = dragon.render(cams[0], SUN, 640, 360, flags=cr.FRAME_RETAINED, pixels=small, ids=small_ids)
        want = dragon.render(cams[0], SUN, 640, 360, want_ids=True)
        assert np.array_equal(got[0].view(np.uint32), want[0].view(np.uint32)) and np.array_equal(got[1], want[1])
    finally:
        hb.close()
        hi.close()


def chain_scene(cr, depth):
    """A degenerate BVH of `depth` + 1 triangles: every inner node has a leaf (left) and another inner node (right), so the
    traversal stack of a ray that enters every box grows to `depth` entries (single_ray_traverser.hpp:14: 64 at most).  The
    triangles overlap along z so that many rays pass through every box."""
    n = depth + 1
    v = np.zeros((n, 9), np.float32)
    for k in range(n):
        z, s = 5.0 + 0.25 * k, 1.0 + 0.05 * k
        v[k] = [-s, -s, z, s, -s, z, 0.0, s, z + 0.1]
    tri = port.make_triangles(v)
    lo = v.reshape(n, 3, 3).min(axis=1)
    hi = v.reshape(n, 3, 3).max(axis=1)
    nodes = np.zeros(2 * n - 1, cr.NODE_DT)

    def box(a, b):      # union of triangles a..b-1
        l, h = lo[a:b].min(axis=0), hi[a:b].max(axis=0)
        return [l[0], h[0], l[1], h[1], l[2], h[2]]

    # node 0 = root over all; inner node k (index idx) has children idx' (leaf: triangle k) and idx'+1 (rest)
    idx = 0
    for k in range(n - 1):
        nodes[idx]["bounds"] = box(k, n)
        nodes[idx]["primitive_count"] = 0
        child = 2 * k + 1
        nodes[idx]["first_child_or_primitive"] = child
        nodes[child]["bounds"] = box(k, k + 1); nodes[child]["primitive_count"] = 1; nodes[child]["first_child_or_primitive"] = k
        idx = child + 1
    nodes[idx]["bounds"] = box(n - 1, n); nodes[idx]["primitive_count"] = 1; nodes[idx]["first_child_or_primitive"] = n - 1
    return {"nodes": nodes, "prim_indices": np.arange(n, dtype=np.uint64), "triangles": tri,
            "norms": np.tile(np.array([0, 0, -1], np.float32), (n, 3))}


@pytest.mark.parametrize("depth", [20, 49, 63])
def test_deep_trees_up_to_the_reference_stack(cr, depth):
    """Trees deeper than 48 need more than the default 48 KB of dynamic shared memory in the 256-thread kernels
    (ceres_trace_rays: 64 entries x 4 B x 256 threads = 64 KB); every kernel opts in or sizes its CTA accordingly."""
    pod = chain_scene(cr, depth)
    cam = cr.Camera((0, 0, 0), (0, 0, 1), (0, 1, 0), 60)
    rng = np.random.default_rng(depth)
    rays = np.zeros((4096, 8), np.float32)
    rays[:, :3] = rng.uniform(-0.5, 0.5, (4096, 3)) * [1, 1, 0]
    d = rng.normal(size=(4096, 3)) * [0.08, 0.08, 0] + [0, 0, 1]
    rays[:, 3:6] = d / np.linalg.norm(d, axis=1, keepdims=True)
    rays[:, 7] = np.finfo(np.float32).max
    with cr.Scene.from_pod(pod) as s:
        ids, tuv = s.trace_rays(rays)
        o_ids, o_tuv = port.trace_rays(pod, rays)
        assert np.array_equal(ids, o_ids) and np.array_equal(tuv.view(np.uint32), o_tuv.view(np.uint32))
        assert (ids != cr.MISS).sum() > 1000
        check_frame(cr, s, pod, cam, np.array([0, 0, -10], np.float32), 320, 240, 0)
        check_frame(cr, s, pod, cr.Camera((0.3, 0.2, 40), (0, 0, -1), (0, 1, 0), 30), np.array([5, 5, 100], np.float32), 320, 240, 0)
        for share in (1, 2):
            s.set_tuning(share=share)
            check_frame(cr, s, pod, cam, np.array([3, 0, -10], np.float32), 160, 120, 0)
        assert s.frame_error() == 0
    too_deep = chain_scene(cr, 65)
    with pytest.raises(cr.CeresError):
        cr.Scene.from_pod(too_deep)


def test_culling_rectangle_adversarial(cr, bunny, bunny_pod):
    """The conservative screen rectangle (cull_rectangle): pixels outside it are never traced, so it must contain every pixel
    whose ray can hit.  (a) cameras placed so that the scene's root box ends within a fraction of a pixel of a pixel centre
    row / column, from both sides; (b) the whole scene scaled by 1e4 (coordinates ~1e4, the margins are relative); (c) the eye
    inside the root box (culling must switch itself off).  Every frame is compared with the oracle, which traces every pixel."""
    W, H = 640, 360
    b = bunny_pod["nodes"][0]["bounds"].astype(np.float64)
    cx, cy = 0.5 * (b[0] + b[1]), 0.5 * (b[2] + b[3])
    tan_h = np.tan(np.radians(30.0))
    for k, frac in enumerate([0.0, 0.02, -0.02, 0.1, -0.1, 0.49, -0.49, 0.5]):
        # distance d such that the box's right edge (front face) projects exactly frac pixels off the centre of column 500
        u = 2.0 * (500 + 0.5 + frac) / W - 1.0
        d = (b[1] - cx) / (u * tan_h)
        cam = cr.Camera((cx, cy, b[4] - d), (0, 0, 1), (0, 1, 0), 60.0)
        check_frame(cr, bunny, bunny_pod, cam, np.array([-50, -20, 0], np.float32), W, H, 0)
        # and the top edge against a pixel row
        v = 2.0 * (300 + 0.5 + frac) / H - 1.0
        d = (b[3] - cy) / (v * tan_h * H / W)
        cam = cr.Camera((cx, cy, b[4] - d), (0, 0, 1), (0, 1, 0), 60.0)
        check_frame(cr, bunny, bunny_pod, cam, np.array([-50, -20, 0], np.float32), W, H, cr.NO_SHADOWS)
    # (b) the scene 1e4 times larger (rebuilt boxes: scale is exact for powers of two only, so refit the boxes on the oracle side)
    big = {k: v.copy() for k, v in bunny_pod.items()}
    big["triangles"] = (bunny_pod["triangles"].astype(np.float32) * np.float32(8192.0)).astype(np.float32)       # p0, e1, e2, n all scale
    big["triangles"][:, 9:12] = bunny_pod["triangles"][:, 9:12] * np.float32(8192.0 * 8192.0)                  # n = e1 x e2
    big["nodes"] = bunny_pod["nodes"].copy()
    big["nodes"]["bounds"] = bunny_pod["nodes"]["bounds"] * np.float32(8192.0)
    with cr.Scene.from_pod(big) as s:
        for eye in ((0.0, 0.1 * 8192, -0.3 * 8192), (0.02 * 8192, 0.1 * 8192, -0.12 * 8192), (0.0, 0.1 * 8192, -3.0 * 8192)):
            check_frame(cr, s, big, cr.Camera(eye, (0, 0, 1), (0, 1, 0), 60.0), np.array([-50 * 8192, -20 * 8192, 0], np.float32), W, H, 0)
    # (c) eye inside the root box, looking along each axis
    for dirv, up in (((0, 0, 1), (0, 1, 0)), ((1, 0, 0), (0, 1, 0)), ((0, -1, 0), (0, 0, 1))):
        cam = cr.Camera((cx, cy, 0.5 * (b[4] + b[5])), dirv, up, 90.0)
        check_frame(cr, bunny, bunny_pod, cam, np.array([-50, -20, 0], np.float32), W, H, 0)


def test_config5_full_size_against_the_oracle(cr, dragon_pod):
    """BASELINE config 5 at its FULL size: the 10.5 M-triangle instanced-dragon scene (8 x 8 x 7 copies, one flat tree, boxes by
    the device refit) at 7680 x 4320 with shadow rays; hit ids, counters and RGB against the oracle's render of the same
    33 M pixels (~1 minute of host work)."""
    import bench
    pod, cam, sun, W, H, flags, scene = bench.load_workload("grid448_8k_shadow", device=0)
    try:
        assert len(pod["triangles"]) == 448 * len(dragon_pod["triangles"]) >= 10_000_000 and (W, H) == (7680, 4320)
        px, ids, cnt = scene.render(cam, sun, W, H, flags=flags, want_ids=True)
        o = port.render(pod, cam.as_array(), sun, W, H, flags)
        assert int((ids != o["ids"]).sum()) == 0
        assert cnt == (o["stats"]["rays"], o["stats"]["hits"])
        assert float(np.abs(px - o["pixels"]).max()) <= RGB_ATOL
        assert cnt[0] > 50_000_000
    finally:
        scene.close()


@pytest.mark.parametrize("count", [2, 3, 8])
@pytest.mark.parametrize("W,H", [(1280, 720), (333, 217)])
def test_sharded_payloads_fill_one_shared_frame(cr, dragon, count, W, H):
    """ceres_render_ppm8_shard / ppm16_shard: every shard writes only its rows of the full-size flipped payload; the shards of
    one frame together give exactly the unsharded payload, bytes outside a shard's rows are left alone."""
    cam = cr.configs.DRAGON_CAMERA
    want8, cnt = dragon.render_ppm8(cam, SUN, W, H)
    want16, _ = dragon.render_ppm16(cam, SUN, W, H)
    with cr.PinnedBuffer((H, W, 3), np.uint8) as h8:
        h8.array[:] = 0xAB
        got16 = np.full((H, W, 3), 0xABAB, ">u2")
        rays = hits = 0
        for k in range(count):
            before = h8.array.copy()
            _, (r, h) = dragon.render_ppm8(cam, SUN, W, H, out=h8.array, shard=(16, k, count))
            rays += r; hits += h
            rows = H - 1 - cr.shard_rows(H, (16, k, count))            # output rows of this shard
            other = np.ones(H, bool); other[rows] = False
            assert np.array_equal(h8.array[other], before[other])
            dragon.render_ppm16(cam, SUN, W, H, out=got16, shard=(16, k, count))
        assert np.array_equal(h8.array, want8) and (rays, hits) == cnt
        assert np.array_equal(got16.view(np.uint8), want16.view(np.uint8))


@pytest.mark.parametrize("count", [2, 8])
def test_scattered_shards_tight_copies_and_retained_frames(cr, dragon, count):
    """Scattered shards (every rank writes its row blocks in place into one host frame) use the tight band copies too, and the
    CERES_FRAME_RETAINED contract holds per shard: a sequence of views into the same frame equals the unsharded frames."""
    W, H = 1280, 720
    cams = [cr.configs.DRAGON_CAMERA, cr.Camera((3.0, -15.0, 2.5), (-0.1, 1.0, 0.0), (0.0, 0.0, 1.0), 45.0),
            cr.Camera((0.0, -15.0, 2.0), (0.0, -1.0, 0.0), (0.0, 0.0, 1.0), 60.0), cr.configs.DRAGON_CAMERA, cr.configs.DRAGON_CAMERA]
    scenes = [dragon] + [cr.Scene.from_pod(scene_pod("dragon")) for _ in range(count - 1)]     # one scene per shard keeps its own state
    hb = cr.PinnedBuffer((H, W, 3), np.float32)
    hi = cr.PinnedBuffer((H, W), np.uint32)
    try:
        hb.array.fill(np.nan); hi.array.fill(7)
        for k, cam in enumerate(cams):
            want = dragon.render(cam, SUN, W, H, want_ids=True)
            rays = hits = 0
            for r in range(count):
                _, _, (a, b) = scenes[r].render(cam, SUN, W, H, flags=cr.FRAME_RETAINED, pixels=hb.array, ids=hi.array, shard=(16, r, count))
                rays += a; hits += b
            assert np.array_equal(hb.array.view(np.uint32), want[0].view(np.uint32)), k
            assert np.array_equal(hi.array, want[1]) and (rays, hits) == want[2], k
            if k == 0:      # tight copies: fewer bytes than the conservative rectangles of the shards (pixels + ids)
                conservative = sum((r[1] - r[0]) * (r[3] - r[2]) * 16 for r in (s.last_rectangle for s in scenes))
                assert sum(s.last_d2h_bytes for s in scenes) < 0.85 * conservative
    finally:
        hb.close(); hi.close()
        for s in scenes[1:]:
            s.close()


def test_calls_during_a_sequence_fail_loudly(cr, dragon_pod):
    """ceres_render_sequence is the one call in which the LIBRARY runs a second thread over a scene; a refit or a render issued
    on that scene from another thread meanwhile is refused (it would race with the sequence's frames), and the sequence's
    frames are untouched by the attempt."""
    import threading
    W, H, n = 1920, 1080, 40
    c = cr.configs.DRAGON_CAMERA
    cams = np.tile(np.array(list(c.eye) + list(c.dir) + list(c.up) + [c.fov] + list(SUN), np.float32), (n, 1))
    with cr.Scene.from_pod(dragon_pod) as s:
        want = s.render(c, SUN, W, H)
        out = {}
        t = threading.Thread(target=lambda: out.update(seq=s.render_sequence(cams, W, H)))
        refused = other = 0
        t.start()
        while t.is_alive():
            try:
                s.refit()
                other += 1
            except cr.CeresError as e:
                assert "ceres_render_sequence is running" in str(e)
                refused += 1
        t.join()
        px_all, cnt = out["seq"]
        for k in range(n):
            assert np.array_equal(px_all[k].view(np.uint32), want[0].view(np.uint32)) and tuple(int(x) for x in cnt[k]) == want[2]
        assert refused > 0
        s.refit()                                   # and the scene is free again afterwards


def test_sequence_as_payloads(cr, dragon):
    """ceres_render_sequence_payload: every frame of a camera path as 8- / 16-bit P6 payload equals the single-frame calls."""
    W, H, n = 640, 360, 5
    c = cr.configs.DRAGON_CAMERA
    cams = np.tile(np.array(list(c.eye) + list(c.dir) + list(c.up) + [c.fov] + list(SUN), np.float32), (n, 1))
    cams[:, 0] += np.linspace(0.0, 2.0, n, dtype=np.float32)             # the eye moves: the frames differ
    for bpc, single in ((1, dragon.render_ppm8), (2, dragon.render_ppm16)):
        for flags in (0, cr.PAYLOAD_NO_FLIP):
            out, cnt = dragon.render_sequence_payload(cams, W, H, flags=flags, bytes_per_channel=bpc)
            for k in range(n):
                cam = cr.Camera(tuple(cams[k, 0:3]), tuple(cams[k, 3:6]), tuple(cams[k, 6:9]), float(cams[k, 9]))
                want, wc = single(cam, cams[k, 10:13], W, H, flags=flags)
                assert np.array_equal(out[k].view(np.uint8), want.view(np.uint8)) and tuple(int(x) for x in cnt[k]) == wc
            assert not np.array_equal(out[0], out[-1])
