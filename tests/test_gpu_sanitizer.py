"""The repository's memcheck / racecheck (profiles/run_sanitizers.sh; compute-sanitizer is closed on the GPU pool): a
reduced workload that launches every shipped kernel, run against the product library (repeat-and-compare), against the
CHECKED build (-DCERES_CHECKED: device-side bounds checks + protocol invariants between frames) and against the checked
build with a deliberately wrong bound (the checks must fire).  Fails on any report."""
import os
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.gpu
def test_checked_build_and_repeat_and_compare_report_nothing(tmp_path):
    if not (shutil.which("nvcc") or os.path.exists("/usr/local/cuda/bin/nvcc")):
        pytest.skip("nvcc not installed (the driver program is compiled on the box)")
    r = subprocess.run(["bash", os.path.join(ROOT, "profiles", "run_sanitizers.sh"), str(tmp_path)],
                       capture_output=True, text=True, timeout=1700, cwd=ROOT,
                       env=dict(os.environ, PATH=os.environ.get("PATH", "") + ":/usr/local/cuda/bin"))
    logs = ""
    for name in ("sanitizer_product.log", "sanitizer_checked.log", "sanitizer_selftest.log"):
        p = tmp_path / name
        if p.exists():
            logs += f"--- {name}\n" + p.read_text()[-1500:]
    assert r.returncode == 0, r.stdout + r.stderr + logs
    assert "0 failed check(s)" in r.stdout and "the checks are live" in r.stdout, r.stdout
