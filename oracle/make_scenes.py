"""Builds oracle/_ref/scenes/{bunny,dragon}.npz with the REFERENCE's own loader and BVH builder (strict build):
obj::load_from_file -> rotate_triangles -> SweepSahBuilder + NodeLayoutOptimizer (anim.cpp:62-65).
Run here (needs /root/reference); the .npz files travel to the GPU box with oracle/_ref/ (git-ignored).
TEST INFRASTRUCTURE."""
import os
import sys
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
from oracle import ref  # noqa: E402

REF = os.environ.get("CERES_REFERENCE", "/root/reference")
OUT = os.path.join(HERE, "_ref", "scenes")
SCENES = {"bunny": ("data/bunny.obj", 1, -145.0), "dragon": ("data/dragon.obj", 0, 90.0)}


def main(force=False):
    os.makedirs(OUT, exist_ok=True)
    for name, (rel, axis, deg) in SCENES.items():
        dst = os.path.join(OUT, name + ".npz")
        if os.path.exists(dst) and not force:
            continue
        s = ref.RefScene.load_obj(os.path.join(REF, rel), axis, deg, 0, "strict")
        pod = s.pod()
        np.savez(dst, **pod)
        print(f"{name}: {s.node_count} nodes, {s.tri_count} triangles -> {dst}")
        s.close()


if __name__ == "__main__":
    main(force="--force" in sys.argv)
