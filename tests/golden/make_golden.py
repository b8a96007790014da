"""Generates the committed golden fixtures from the REFERENCE's own code (oracle/_ref/libceres_ref_strict.so =
/root/reference headers compiled unchanged with -ffp-contract=off; see oracle/Makefile, oracle/ref_driver.cpp).

Run in the build container (needs /root/reference):   python tests/golden/make_golden.py
Outputs (all small, committed):
  dragon_scene.npz       same for dragon.obj after rotate x 90 (static.cpp:72-73,83-88)
  bunny_scene.npz        the reference-loaded, reference-built bunny scene (POD arrays), so golden tests need no _ref
  bunny_frames.npz       320x180 frames (ids, state, tuv, rgb, stats) in shadow and primary-only mode
  kat.npz                known-answer vectors: slab test, triangle test, smooth_shading, 20k traced rays, turntable
  frame_hashes.json      sha256 of the id/state buffers + counters + S/T statistics for the BASELINE configs
"""
import hashlib
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
from oracle import ref  # noqa: E402

BUNNY_CAM = ref.cam_array((0.0, 0.1, -0.3), (0, 0, 1), (0, 1, 0), 60)
DRAGON_CAM = ref.cam_array((0.0, -15.0, 2.0), (0, 1, 0), (0, 0, 1), 60)
SUN = np.array([-50.0, -20.0, 0.0], np.float32)


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def random_rays(rng, n, lo, hi):
    """Origins around the scene box, directions on the sphere plus axis-aligned / signed-zero / tiny components."""
    ext = hi - lo
    o = (lo - ext) + rng.random((n, 3)) * (3 * ext)
    d = rng.normal(size=(n, 3))
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    k = n // 8
    d[:k, 0] = 0.0
    d[k:2 * k, 1] = -0.0
    d[2 * k:3 * k, 2] = 1e-9 * rng.choice([-1, 1], size=k)
    # aim half of the rays at the scene centre so that many hit
    c = (lo + hi) / 2 + (rng.random((n, 3)) - 0.5) * ext * 0.8
    aim = rng.random(n) < 0.5
    v = c - o
    v /= np.linalg.norm(v, axis=1, keepdims=True)
    d[aim] = v[aim]
    rays = np.zeros((n, 8), np.float32)
    rays[:, 0:3], rays[:, 3:6] = o, d
    rays[:, 6] = 0.0
    rays[:, 7] = np.finfo(np.float32).max
    rays[::7, 7] = (rng.random(len(rays[::7])) * np.linalg.norm(ext) * 2).astype(np.float32)  # finite tmax
    rays[::11, 6] = (rng.random(len(rays[::11])) * np.linalg.norm(ext) * 0.5).astype(np.float32)  # tmin > 0
    return rays


def main():
    rng = np.random.default_rng(12345)
    ref_root = os.environ.get("CERES_REFERENCE", "/root/reference")
    bunny = ref.RefScene.load_obj(os.path.join(ref_root, "data/bunny.obj"), 1, -145.0, 0, "strict")
    pod = bunny.pod()
    np.savez_compressed(os.path.join(HERE, "bunny_scene.npz"), **pod)
    # dragon: static.cpp scene (rotate x 90), same loader/builder; input data of bench.py and the dragon parity tests
    dragon = ref.RefScene.load_obj(os.path.join(ref_root, "data/dragon.obj"), 0, 90.0, 0, "strict")
    np.savez_compressed(os.path.join(HERE, "dragon_scene.npz"), **dragon.pod())
    dragon.close()

    frames = {}
    for tag, flags in (("shadow", 0), ("primary", 1)):
        r = bunny.render_ex(BUNNY_CAM, SUN, 320, 180, flags)
        frames[f"{tag}_ids"] = r["ids"]
        frames[f"{tag}_state"] = r["state"]
        frames[f"{tag}_tuv"] = r["tuv"]
        frames[f"{tag}_rgb"] = r["pixels"]
        frames[f"{tag}_stats"] = np.array(list(r["stats"].values()), np.uint64)
    px, rays, hits, _ = bunny.render(BUNNY_CAM, SUN, 320, 180)   # the unmodified render()
    assert np.array_equal(px, frames["shadow_rgb"])
    frames["render_counters"] = np.array([rays, hits], np.int64)
    frames["cam"], frames["sun"] = BUNNY_CAM, SUN
    np.savez_compressed(os.path.join(HERE, "bunny_frames.npz"), **frames)

    # ---- KATs
    tris = pod["triangles"]
    lo = tris[:, 0:3].min(0).astype(np.float64)
    hi = tris[:, 0:3].max(0).astype(np.float64)
    n = 8192
    rays = random_rays(rng, n, lo, hi)
    boxes = pod["nodes"]["bounds"][rng.integers(0, len(pod["nodes"]), n)]
    slab = ref.slab(boxes, rays)
    tsel = rng.integers(0, len(tris), n)
    p0 = tris[tsel, 0:3]
    verts = np.concatenate([p0, p0 - tris[tsel, 3:6], p0 + tris[tsel, 6:9]], axis=1).astype(np.float32)
    # aim the triangle-test rays at points inside their triangle half of the time
    bar = rng.random((n, 2)); bar[bar.sum(1) > 1] = 1 - bar[bar.sum(1) > 1]
    tgt = verts[:, 0:3] + bar[:, :1] * (verts[:, 3:6] - verts[:, 0:3]) + bar[:, 1:] * (verts[:, 6:9] - verts[:, 0:3])
    trays = rays.copy()
    aim = rng.random(n) < 0.6
    v = tgt - trays[:, 0:3]
    v /= np.linalg.norm(v, axis=1, keepdims=True)
    trays[aim, 3:6] = v[aim].astype(np.float32)
    trays[:, 6], trays[:, 7] = 0.0, np.finfo(np.float32).max
    thit, ttuv = ref.tri_intersect(verts, trays)
    sh_in = np.zeros((n, 17), np.float32)
    sl = rng.normal(size=(n, 3)); sl /= np.linalg.norm(sl, axis=1, keepdims=True)
    vw = rng.normal(size=(n, 3)); vw /= np.linalg.norm(vw, axis=1, keepdims=True)
    sh_in[:, 0:3], sh_in[:, 12:15] = sl, vw
    sh_in[:, 3:12] = pod["norms"][rng.integers(0, len(tris), n)]
    uv = rng.random((n, 2)); uv[uv.sum(1) > 1] = 1 - uv[uv.sum(1) > 1]
    sh_in[:, 15:17] = uv
    sh_out = ref.smooth_shading(sh_in)
    trace = random_rays(rng, 20000, lo, hi)
    tr_ids, tr_tuv = bunny.trace_rays(trace)
    turn = ref.turntable(DRAGON_CAM, SUN, np.array([0, 0, 1], np.float32), 60)
    np.savez_compressed(os.path.join(HERE, "kat.npz"), slab_boxes=boxes, slab_rays=rays, slab_out=slab,
                        tri_verts=verts, tri_rays=trays, tri_hit=thit, tri_tuv=ttuv, shade_in=sh_in, shade_out=sh_out,
                        trace_rays=trace, trace_ids=tr_ids, trace_tuv=tr_tuv, turntable=turn)
    print("KAT: tri hits", int(thit.sum()), "/", n, "; traced hits", int((tr_ids != ref.MISS).sum()), "/ 20000")

    # ---- frame hashes for the BASELINE configs (scenes = the *_scene.npz fixtures above)
    hashes = {}
    for name, cam, confs in (("bunny", BUNNY_CAM, [(1920, 1080, 1), (1920, 1080, 0)]),
                             ("dragon", DRAGON_CAM, [(1920, 1080, 1), (1920, 1080, 0), (3840, 2160, 0)])):
        spod = dict(np.load(os.path.join(HERE, name + "_scene.npz")))
        s = ref.RefScene.from_pod(spod, "strict")
        sf = ref.RefScene.from_pod(spod, "fast")
        for W, H, flags in confs:
            r = s.render_ex(cam, SUN, W, H, flags, want_pixels=False)
            rf = sf.render_ex(cam, SUN, W, H, flags, want_pixels=False)
            key = f"{name}_{W}x{H}_{'primary' if flags else 'shadow'}"
            hashes[key] = {
                "scene_sha256": sha(spod["nodes"]) [:16] + sha(spod["prim_indices"])[:16],
                "ids_sha256": sha(r["ids"]), "state_sha256": sha(r["state"]), "stats": r["stats"],
                "primary_hits": int((r["ids"] != ref.MISS).sum()), "lit": int((r["state"] == 1).sum()),
                "fast_build_id_mismatches": int((r["ids"] != rf["ids"]).sum()),
                "fast_build_state_mismatches": int((r["state"] != rf["state"]).sum()),
                "fast_build_stats": rf["stats"],
            }
            print(key, hashes[key]["stats"], "fast-vs-strict id diffs", hashes[key]["fast_build_id_mismatches"])
        s.close(); sf.close()
    with open(os.path.join(HERE, "frame_hashes.json"), "w") as f:
        json.dump(hashes, f, indent=1, sort_keys=True)


if __name__ == "__main__":
    main()
