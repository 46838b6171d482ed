"""Build the CUDA library (sm_100a only) in-tree: hyteg_b200/libhyteg_b200.so."""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "csrc")
OUT = os.path.join(HERE, "libhyteg_b200.so")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "--use_fast_math=false",
         "-Xcompiler", "-fPIC", "-shared", "-Xptxas", "-v"]
SOURCES = ["kernels.cu", "shuffle_kernel.cu", "p2_kernels.cu", "runtime.cu", "topology.cpp", "stencil_assembly.cpp", "solvers_capi.cpp"]


def needs_build():
    if not os.path.exists(OUT):
        return True
    t = os.path.getmtime(OUT)
    deps = [os.path.join(SRC, f) for f in os.listdir(SRC)]
    inc = os.path.join(HERE, "..", "include")
    deps += [os.path.join(inc, "hyteg_b200.h"), os.path.join(inc, "hyteg_b200", "hyteg.hpp"), os.path.join(inc, "hyteg_b200", "hyteg_solvers.hpp")]
    return any(os.path.getmtime(d) > t for d in deps)


def build(force=False, verbose=False):
    if not force and not needs_build():
        return OUT
    flags = [f for f in FLAGS if f != "--use_fast_math=false"]
    cmd = [NVCC] + flags + [os.path.join(SRC, s) for s in SOURCES] + ["-o", OUT]
    res = subprocess.run(cmd, capture_output=True, text=True)
    log = res.stdout + res.stderr
    with open(os.path.join(HERE, "build.log"), "w") as f:
        f.write(" ".join(cmd) + "\n" + log)
    if verbose or res.returncode != 0:
        print(log)
    if res.returncode != 0:
        raise RuntimeError("nvcc failed, see hyteg_b200/build.log")
    return OUT


if __name__ == "__main__":
    build(force="--force" in sys.argv, verbose=True)
