"""Run with CERES_B200_EXPERIMENTS=1 (tests/test_gpu_experiments.py does): every wave-pipeline variant of the experiments
build must render the frame kernel's frame -- ids, pixels, counters -- and do the oracle's primary work."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    assert os.environ.get("CERES_B200_EXPERIMENTS") == "1"
    import ceres_raytracer_b200 as cr
    from conftest import scene_pod
    from oracle import port
    pod = scene_pod("dragon")
    cam, sun, W, H = cr.configs.DRAGON_CAMERA, np.array([-50, -20, 0], np.float32), 1280, 720
    o = port.render(pod, cam.as_array(), sun, W, H, 0)
    with cr.Scene.from_pod(pod) as s:
        ref = s.render(cam, sun, W, H, want_ids=True)
        assert np.array_equal(ref[1], o["ids"]) and ref[2] == (o["stats"]["rays"], o["stats"]["hits"])
        s.set_stats(True)
        s.render(cam, sun, W, H)
        ref_stats = s.get_stats()
        for variant, refill in ((1, 0), (2, 0), (3, 32), (3, 8), (4, 0), (5, 0), (6, 0), (6, 8)):
            s.set_tuning(variant=variant, refill_min=refill)
            px, ids, cnt = s.render(cam, sun, W, H, want_ids=True)
            assert np.array_equal(px.view(np.uint32), ref[0].view(np.uint32)), variant
            assert np.array_equal(ids, ref[1]) and cnt == ref[2], variant
            st = s.get_stats()
            for k in ("T_miss", "S_hit", "T_hit", "rays", "hits"):
                assert st[k] == ref_stats[k], (variant, refill, k)
            if variant not in (1, 2):      # 1 and 2 do not cull the columns outside the screen rectangle
                assert st["S_miss"] == ref_stats["S_miss"], (variant, refill)
            print(f"variant {variant} refill {refill}: ok", flush=True)
        s.set_stats(False)
        s.set_tuning()
    print("experiments ok")


if __name__ == "__main__":
    main()
