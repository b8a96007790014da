"""In-tree build of the CUDA library (no JIT cache: the built .so travels with the source tree)."""
import os
import shutil
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
SRC = os.path.join(HERE, "csrc", "ceres_b200.cu")
LIB = os.path.join(HERE, "libceres_b200.so")
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


def needs_build():
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    deps = [SRC, HEADER] + [os.path.join(HERE, 'csrc', n) for n in os.listdir(os.path.join(HERE, 'csrc'))]
    return any(os.path.getmtime(p) > t for p in deps)


def build(force=False, verbose=False):
    if not force and not needs_build():
        return LIB
    cmd = [nvcc()] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-o", LIB, SRC]
    env = dict(os.environ)
    subprocess.check_call(cmd, env=env)
    return LIB


if __name__ == "__main__":
    import sys
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
