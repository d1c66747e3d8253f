"""Builds libfec_b200.so (CUDA kernels + C ABI) in-tree with nvcc for sm_100a. No GPU is needed to build."""
import os
import shutil
import subprocess

PKG_DIR = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(PKG_DIR, "csrc")
LIB_PATH = os.path.join(PKG_DIR, "libfec_b200.so")
SOURCES = ["fec_engine.cu", "fec_harq_pool.cpp"]
HEADERS = ["fec_device.cuh", "fec_kernel_generic.cuh", "fec_kernel_packed.cuh", "fec_tb.cuh", "fec_encode.cuh", "fec_demod.cuh", "ldpc_tables.h",
           os.path.join("..", "..", "include", "fec_b200.h")]
NVCC_FLAGS = (["-DFEC_PROFILE=" + os.environ["FEC_PROFILE"]] if os.environ.get("FEC_PROFILE") else []) + \
    ["-D" + d for d in os.environ.get("FEC_DEFS", "").split()] + ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17", "-shared",
              "-Xcompiler", "-fPIC", "-Xptxas", "-v"]


def _nvcc():
    for cand in (shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    return None


def needs_build():
    if not os.path.exists(LIB_PATH):
        return True
    t = os.path.getmtime(LIB_PATH)
    deps = [os.path.join(CSRC, f) for f in SOURCES + HEADERS]
    return any(os.path.exists(d) and os.path.getmtime(d) > t for d in deps)


def build(force=False, verbose=False):
    """Compiles the extension if it is missing or older than its sources. Returns the path of the .so."""
    if not force and not needs_build():
        return LIB_PATH
    nvcc = _nvcc()
    if nvcc is None:
        if os.path.exists(LIB_PATH):
            return LIB_PATH  # prebuilt file shipped with the snapshot
        raise RuntimeError("nvcc not found and libfec_b200.so is not built")
    cmd = [nvcc] + NVCC_FLAGS + ["-o", LIB_PATH] + [os.path.join(CSRC, s) for s in SOURCES]
    res = subprocess.run(cmd, cwd=CSRC, capture_output=True, text=True)
    if verbose or res.returncode != 0:
        print(res.stdout)
        print(res.stderr)
    if res.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + res.stderr[-4000:])
    with open(os.path.join(PKG_DIR, "build.log"), "w") as f:
        f.write(" ".join(cmd) + "\n" + res.stdout + res.stderr)
    return LIB_PATH
