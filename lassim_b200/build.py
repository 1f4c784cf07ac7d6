"""In-tree build of liblassim_b200.so (hand-written CUDA for sm_100a + the C
ABI of include/lassim_b200.h).  nvcc cross-compiles without a GPU."""
import os
import subprocess

PKG = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(PKG, "liblassim_b200.so")
SOURCES = [os.path.join(PKG, "csrc", "lassim_capi.cu")]
DEPS = SOURCES + [os.path.join(PKG, "csrc", f) for f in sorted(os.listdir(os.path.join(PKG, "csrc")))
                  if f.endswith(".cuh")] + [os.path.join(os.path.dirname(PKG), "include", "lassim_b200.h")]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo",
              "-std=c++17", "-shared", "-Xcompiler", "-fPIC"]


def is_stale():
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    return any(os.path.getmtime(d) > t for d in DEPS if os.path.exists(d))


def build_library(force=False, verbose=False):
    """Compile the library if it is missing or older than its sources.
    Returns the path.  Raises if nvcc is missing or compilation fails."""
    if not force and not is_stale():
        return LIB
    nvcc = os.environ.get("NVCC", "nvcc")
    cmd = [nvcc] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + \
        ["-o", LIB] + SOURCES
    subprocess.check_call(cmd)
    return LIB


if __name__ == "__main__":
    print(build_library(force=True, verbose=True))
