"""The wave pipelines of round 1 (ceres_tuning.variant 1-6) live in the -DCERES_EXPERIMENTS build only
(libceres_b200_exp.so); they are negative results kept for A/B measurements (profiles/kbench.py), and this test keeps
them honest: in a child process that loads the experiments library they must reproduce the frame kernel's frame."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_wave_pipeline_variants_render_the_same_frame():
    env = dict(os.environ, CERES_B200_EXPERIMENTS="1")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tests", "experiments_check.py")], env=env, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "experiments ok" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]
