"""Builds librpt_gpu.so (sm_100a only) in-tree with nvcc.  No GPU is needed to build."""
import os
import shutil
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.environ.get("RPT_GPU_LIB") or os.path.join(HERE, "librpt_gpu.so")
SOURCES = ["rpt_kernels.cu", "rpt_stats.cu", "rpt_acero.cu", "rpt_capi.cu", "rpt_noalloc.cu"]
HEADERS = ["rpt_hash.cuh", "rpt_kernels.cuh", os.path.join("..", "..", "include", "rpt_gpu.h")]

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


def _compile_and_link(out, extra_flags=(), verbose=False):
    """One nvcc -c per source, in parallel (the sources share no device code), then one link."""
    from concurrent.futures import ThreadPoolExecutor
    objdir = os.path.join(HERE, "build", os.path.basename(out) + ".obj")
    os.makedirs(objdir, exist_ok=True)
    compile_flags = [f for f in NVCC_FLAGS if f != "-shared"]

    def one(src):
        obj = os.path.join(objdir, src.replace(".cu", ".o"))
        cmd = [_nvcc()] + compile_flags + list(extra_flags) + ["-c", os.path.join(CSRC, src), "-o", obj]
        if verbose:
            print(" ".join(cmd))
        proc = subprocess.run(cmd, cwd=CSRC, capture_output=True, text=True)
        if proc.returncode != 0:
            raise RuntimeError("nvcc failed:\n" + proc.stdout + proc.stderr)
        return obj

    with ThreadPoolExecutor(max_workers=len(SOURCES)) as pool:
        objs = list(pool.map(one, SOURCES))
    cmd = [_nvcc(), "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", out + ".tmp"] + objs
    proc = subprocess.run(cmd, cwd=CSRC, capture_output=True, text=True)
    if proc.returncode != 0:
        raise RuntimeError("nvcc link failed:\n" + proc.stdout + proc.stderr)
    os.replace(out + ".tmp", out)
    return out


def build(force=False, verbose=False, extra_flags=(), out=None):
    """Compile every CUDA source for sm_100a into librpt_gpu.so next to this file.
    extra_flags / out build an experimental variant (e.g. -DRPT_PROBE_ROUNDS=4) under another name; RPT_GPU_LIB
    selects which file the Python driver loads."""
    if out is not None:
        return _compile_and_link(out, extra_flags, verbose)
    if not force and not needs_build():
        return LIB
    return _compile_and_link(LIB, (), verbose)


if __name__ == "__main__":
    print(build(force=True, verbose=True))
