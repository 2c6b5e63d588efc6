"""Builds librpt_gpu.so (sm_100a only) in-tree with nvcc.  No GPU is needed to build."""
import os
import shutil
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.environ.get("RPT_GPU_LIB") or os.path.join(HERE, "librpt_gpu.so")
SOURCES = ["rpt_kernels.cu", "rpt_stats.cu", "rpt_acero.cu", "rpt_capi.cu", "rpt_host.cpp"]
HEADERS = ["rpt_hash.cuh", "rpt_kernels.cuh", "rpt_host.hpp", os.path.join("..", "..", "include", "rpt_gpu.h")]

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC,-fvisibility=hidden,-Wall",
    "-shared",
]


def _nvcc():
    path = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(path):
        raise RuntimeError("nvcc not found; librpt_gpu.so cannot be built")
    return path


def needs_build():
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    files = [os.path.join(CSRC, f) for f in SOURCES + HEADERS if os.path.exists(os.path.join(CSRC, f))]
    return any(os.path.getmtime(f) > t for f in files)


def build(force=False, verbose=False, extra_flags=(), out=None):
    """Compile every CUDA source for sm_100a into librpt_gpu.so next to this file.
    extra_flags / out build an experimental variant (e.g. -DRPT_PROBE_ROUNDS=4) under another name; RPT_GPU_LIB
    selects which file the Python driver loads."""
    if out is not None:
        srcs = [os.path.join(CSRC, f) for f in SOURCES if os.path.exists(os.path.join(CSRC, f))]
        cmd = [_nvcc()] + NVCC_FLAGS + list(extra_flags) + ["-o", out] + srcs
        proc = subprocess.run(cmd, cwd=CSRC, capture_output=True, text=True)
        if proc.returncode != 0:
            raise RuntimeError("nvcc failed:\n" + proc.stdout + proc.stderr)
        return out
    if not force and not needs_build():
        return LIB
    srcs = [os.path.join(CSRC, f) for f in SOURCES if os.path.exists(os.path.join(CSRC, f))]
    cmd = [_nvcc()] + NVCC_FLAGS + ["-o", LIB + ".tmp"] + srcs
    if verbose:
        print(" ".join(cmd))
    proc = subprocess.run(cmd, cwd=CSRC, capture_output=True, text=True)
    if proc.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + proc.stdout + proc.stderr)
    os.replace(LIB + ".tmp", LIB)
    return LIB


if __name__ == "__main__":
    print(build(force=True, verbose=True))
