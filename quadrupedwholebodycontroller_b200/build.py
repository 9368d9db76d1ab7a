"""Build recipe of the CUDA library (in-tree, sm_100a only)."""
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB_PATH = os.environ.get("WBC_B200_LIB") or os.path.join(HERE, "libwbc_b200.so")     # override = DEV experiments with another build of the same CUDA library
EXACT_LIB_PATH = os.path.join(HERE, "libwbc_b200_exact.so")   # same sources with -DWBC_EXACT -fmad=false: qpOASES' arithmetic operation by operation (verification only, slower)
F32_LIB_PATH = os.path.join(HERE, "libwbc_b200_f32.so")       # same sources with -DWBC_FP32: the QP solver in single precision (qpOASES' __USE_SINGLE_PRECISION__ tolerances); RBD / assembly stay fp64
PROF_LIB_PATH = os.path.join(HERE, "libwbc_b200_prof.so")     # same sources with -DWBC_PROFILE: clock64 phase timers + algorithmic-flop counter (measurement only)
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17", "-Xcompiler", "-fPIC", "-shared"]


def _stale(path=None):
    path = path or LIB_PATH
    if not os.path.exists(path):
        return True
    t = os.path.getmtime(path)
    srcs = [os.path.join(CSRC, f) for f in os.listdir(CSRC)] + [os.path.join(HERE, "..", "include", "wbc_b200.h")]
    return any(os.path.getmtime(s) > t for s in srcs)


def build_variant(out_path, defines=(), verbose=False):
    """DEV: build the same library with extra -D switches to another path (tools/gpu_* experiments)."""
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    cmd = [nvcc] + NVCC_FLAGS + [f"-D{d}" for d in defines] + (["-Xptxas", "-v"] if verbose else []) + [os.path.join(CSRC, "wbc_kernels.cu"), "-o", out_path]
    subprocess.check_call(cmd)
    return out_path


def build_library(force=False, verbose=False):
    """nvcc -gencode arch=compute_100a,code=sm_100a ... csrc/wbc_kernels.cu -> libwbc_b200.so (cross-compiles without a GPU)."""
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    default = os.path.join(HERE, "libwbc_b200.so")
    jobs = []
    if force or _stale(default):
        jobs.append(subprocess.Popen([nvcc] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + [os.path.join(CSRC, "wbc_kernels.cu"), "-o", default]))
    if force or _stale(PROF_LIB_PATH):
        jobs.append(subprocess.Popen([nvcc] + NVCC_FLAGS + ["-DWBC_PROFILE", os.path.join(CSRC, "wbc_kernels.cu"), "-o", PROF_LIB_PATH]))
    if force or _stale(EXACT_LIB_PATH):
        jobs.append(subprocess.Popen([nvcc] + NVCC_FLAGS + ["-DWBC_EXACT", "-fmad=false", os.path.join(CSRC, "wbc_kernels.cu"), "-o", EXACT_LIB_PATH]))
    if force or _stale(F32_LIB_PATH):
        jobs.append(subprocess.Popen([nvcc] + NVCC_FLAGS + ["-DWBC_FP32", os.path.join(CSRC, "wbc_kernels.cu"), "-o", F32_LIB_PATH]))
    for j in jobs:
        if j.wait() != 0:
            raise RuntimeError("nvcc failed building the CUDA library")
    return LIB_PATH
