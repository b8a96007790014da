"""The workloads BASELINE.json names, made concrete as in SURVEY.md 8(d).  Cameras/suns are the constants of the
reference drivers (static.cpp:38-47,72-73) and of the README example (README.md:11)."""
from .render import Camera

# static.cpp:38-47: dragon.obj rotated 90 deg about x
DRAGON_CAMERA = Camera(eye=(0.0, -15.0, 2.0), dir=(0.0, 1.0, 0.0), up=(0.0, 0.0, 1.0), fov=60.0)
DRAGON_SUN = (-50.0, -20.0, 0.0)
DRAGON_ROTATE = (0, 90.0)          # axis x, degrees

# README.md:11: ./render ../data/bunny.obj --eye 0 .1 -.3 --rotate y -145 ; dir/up are decision D1 of SURVEY.md 8(b)
BUNNY_CAMERA = Camera(eye=(0.0, 0.1, -0.3), dir=(0.0, 0.0, 1.0), up=(0.0, 1.0, 0.0), fov=60.0)
BUNNY_SUN = (-50.0, -20.0, 0.0)
BUNNY_ROTATE = (1, -145.0)         # axis y

WORKLOADS = {
    # name: (scene, camera, sun, width, height, flags)
    "bunny_1080p_primary": ("bunny", BUNNY_CAMERA, BUNNY_SUN, 1920, 1080, 1),
    "bunny_1080p_shadow": ("bunny", BUNNY_CAMERA, BUNNY_SUN, 1920, 1080, 0),
    "dragon_1080p_primary": ("dragon", DRAGON_CAMERA, DRAGON_SUN, 1920, 1080, 1),
    "dragon_1080p_shadow": ("dragon", DRAGON_CAMERA, DRAGON_SUN, 1920, 1080, 0),
    "dragon_4k_shadow": ("dragon", DRAGON_CAMERA, DRAGON_SUN, 3840, 2160, 0),
    # BASELINE.json configs[4] / SURVEY.md 8(d) config 5: the dragon on an 8x8x7 grid (448 copies = 10 523 520
    # triangles), one flat BVH; camera and sun come from scenes.grid_camera (None here)
    "grid448_8k_shadow": ("grid:dragon:8x8x7", None, None, 7680, 4320, 0),
    "grid448_4k_shadow": ("grid:dragon:8x8x7", None, None, 3840, 2160, 0),
    "grid12_1080p_shadow": ("grid:dragon:3x2x2", None, None, 1920, 1080, 0),
}

# BASELINE.json configs[3] / SURVEY.md 8(d) config 4: 60-frame turntable of the dragon (anim.cpp:77-89,142), frames
# round-robin over the GPUs.  name: (scene, frames, width, height, flags); the camera path is the reference's own
# output for the static.cpp camera (tests/golden/kat.npz "turntable", made by tests/golden/make_golden.py)
SEQUENCES = {
    "dragon_turntable_1080p": ("dragon", 60, 1920, 1080, 0),
    "dragon_turntable_4k": ("dragon", 60, 3840, 2160, 0),
}

TILE_BLOCK_ROWS = 16   # multi-GPU row-block interleave
