"""Builds libcurl_b200.so (the sm_100a kernels + C ABI) in-tree with nvcc.

    python -m curl_b200.build [--force]

The library is built next to this file so that it travels with the repo snapshot to the GPU box.
nvcc cross-compiles for sm_100a without a GPU.  The CUDA runtime is linked statically, so the
library has no dependency on a particular libcudart.so being on the loader path.
"""
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libcurl_b200.so")
SOURCES = [os.path.join(CSRC, "curl_kernels.cu"), os.path.join(CSRC, "curl_trispace.cu"),
           os.path.join(CSRC, "curl_loss.cu"), os.path.join(CSRC, "curl_segclamp.cu"), os.path.join(CSRC, "curl_convert.cu")]
DEPS = SOURCES + [os.path.join(CSRC, "curl_pixel.cuh"), os.path.join(CSRC, "curl_vec.cuh"), os.path.join(CSRC, "curl_poly.cuh"),
                  os.path.join(CSRC, "curl_loss.cuh"), os.path.join(CSRC, "curl_seg.cuh"), os.path.join(HERE, "..", "include", "curl_b200.h")]

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-std=c++17", "-lineinfo",
    "-Xcompiler", "-fPIC", "-shared", "--cudart", "static",
]


def nvcc_path():
    cand = os.environ.get("NVCC") or shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(cand):
        raise RuntimeError("nvcc not found (set NVCC=/path/to/nvcc)")
    return cand


def is_stale():
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    return any(os.path.getmtime(d) > t for d in DEPS)


def build(force=False, verbose=False):
    """Compile the library if missing or older than its sources.  Returns the path."""
    if not force and not is_stale():
        return LIB
    extra = os.environ.get("CURL_NVCC_EXTRA", "").split()   # experiments only, e.g. -DBWD_MIN_BLOCKS=2
    cmd = [nvcc_path()] + NVCC_FLAGS + extra + (["-Xptxas", "-v"] if verbose else []) + ["-o", LIB] + SOURCES
    proc = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if proc.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + " ".join(cmd) + "\n" + proc.stdout)
    if verbose:
        print(proc.stdout)
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
