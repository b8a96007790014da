import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")
SCENES = os.path.join(ROOT, "oracle", "_ref", "scenes")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu on the GPU box)")


def load_npz(path):
    with np.load(path) as z:
        return {k: z[k] for k in z.files}


@pytest.fixture(scope="session")
def bunny_pod():
    return load_npz(os.path.join(GOLDEN, "bunny_scene.npz"))


@pytest.fixture(scope="session")
def bunny_frames():
    return load_npz(os.path.join(GOLDEN, "bunny_frames.npz"))


@pytest.fixture(scope="session")
def kat():
    return load_npz(os.path.join(GOLDEN, "kat.npz"))


@pytest.fixture(scope="session")
def frame_hashes():
    import json
    with open(os.path.join(GOLDEN, "frame_hashes.json")) as f:
        return json.load(f)


def scene_pod(name):
    """Reference-built scene dumps (oracle/_ref/scenes, made by oracle/make_scenes.py); None when absent."""
    for p in (os.path.join(GOLDEN, name + "_scene.npz"), os.path.join(SCENES, name + ".npz")):
        if os.path.exists(p):
            return load_npz(p)
    return None


@pytest.fixture(scope="session")
def dragon_pod():
    pod = scene_pod("dragon")
    if pod is None:
        pytest.skip("tests/golden/dragon_scene.npz missing (tests/golden/make_golden.py)")
    return pod
