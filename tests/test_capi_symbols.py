"""The C-ABI library loads on a CPU-only box and exports every symbol include/ceres_b200.h declares.  No compute
calls here; on a box without a GPU the compute entry points must FAIL (no CPU fallback)."""
import ctypes
import os
import re

import numpy as np
import pytest

from conftest import ROOT
from ceres_raytracer_b200 import _capi


def declared_symbols():
    src = open(os.path.join(ROOT, "include", "ceres_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(ceres_[a-z_0-9]+)\s*\(", src)))


def test_header_symbols_exported():
    L = _capi.lib()
    names = declared_symbols()
    assert len(names) >= 15
    for n in names:
        assert hasattr(L, n), f"{n} declared in include/ceres_b200.h but not exported"
    assert sorted(_capi.SYMBOLS) == names
    assert L.ceres_abi_version() == 2


def test_frame_setup_is_host_only_and_matches_oracle():
    from oracle import port
    from ceres_raytracer_b200 import Camera, frame_setup
    cam = Camera((0, -15, 2), (0, 1, 0.1), (0, 0, 1), 60)
    d, u, v = frame_setup(cam, 1920, 1080)
    o = port.frame_setup(cam.as_array(), 1920, 1080)
    assert np.array_equal(np.concatenate([d, u, v]).view(np.uint32), o.view(np.uint32))


def test_no_cpu_fallback(bunny_pod):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from ceres_raytracer_b200 import Scene, CeresError
    with pytest.raises(CeresError):
        Scene.from_pod(bunny_pod)


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "ceres_raytracer_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cpp", ".h", ".cuh")):
                txt = open(os.path.join(dirpath, f)).read()
                assert "oracle" not in txt.replace("no CPU fallback", ""), f"{f} mentions the oracle"
