"""In-tree build of the CUDA library (no JIT cache: the built .so travels with the source tree)."""
import os
import shutil
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
SRC = os.path.join(HERE, "csrc", "ceres_b200.cu")
LIB = os.path.join(HERE, "libceres_b200.so")
# The same sources with -DCERES_EXPERIMENTS: the product plus the wave pipelines of round 1 (ceres_tuning.variant 1-6),
# kept for A/B measurements (profiles/kbench.py).  Loaded only when CERES_B200_EXPERIMENTS=1 is set; never shipped as
# the product library.
LIB_EXP = os.path.join(HERE, "libceres_b200_exp.so")
HEADER = os.path.join(ROOT, "include", "ceres_b200.h")

# sm_100a only.  -fmad=false: belt and braces -- the kernels already issue every operation of the path through
# the non-contracting *_rn intrinsics.  Host code of the same file does the frame setup of render.hpp:91-97 and
# must not contract either.
NVCC_FLAGS = [
    "-O3", "-std=c++17", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-fmad=false",
    "-Xcompiler", "-fPIC,-ffp-contract=off,-mavx2", "-shared",
]


def nvcc():
    exe = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(exe):
        raise RuntimeError("nvcc not found: cannot build ceres_raytracer_b200/libceres_b200.so")
    return exe


def _deps():
    out = [SRC, HEADER]
    for d, _, names in os.walk(os.path.join(HERE, "csrc")):
        out += [os.path.join(d, n) for n in names]
    return out


def needs_build(lib=LIB):
    if not os.path.exists(lib):
        return True
    t = os.path.getmtime(lib)
    return any(os.path.getmtime(p) > t for p in _deps())


def variant_path(name):
    return os.path.join(HERE, f"libceres_b200_{name}.so")


def build(force=False, verbose=False, experiments=False, variant=None, extra=()):
    lib = variant_path(variant) if variant else LIB_EXP if experiments else LIB
    if not force and not variant and not needs_build(lib):
        return lib
    cmd = [nvcc()] + NVCC_FLAGS + (["-DCERES_EXPERIMENTS", "-diag-suppress", "177"] if experiments else []) \
        + (["-Xptxas", "-v"] if verbose else []) + list(extra) + ["-o", lib, SRC]
    subprocess.check_call(cmd, env=dict(os.environ))
    return lib


if __name__ == "__main__":
    import sys
    # --variant NAME [nvcc flags...]: libceres_b200_NAME.so with extra compile-time options, loaded with CERES_B200_VARIANT=NAME
    if "--variant" in sys.argv:
        k = sys.argv.index("--variant")
        print(build(variant=sys.argv[k + 1], extra=sys.argv[k + 2:]))
    else:
        print(build(force="--force" in sys.argv, verbose="-v" in sys.argv, experiments="--experiments" in sys.argv))
