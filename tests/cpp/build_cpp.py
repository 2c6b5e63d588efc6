"""TEST INFRASTRUCTURE: builds tests/cpp/host_mirror_test, the driver of the C++ host mirror (include/rpt_gpu.hpp).
It links the product library (librpt_gpu.so) and, as the checker, the oracle (oracle/librpt_oracle.so)."""
import os
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
PKG = os.path.join(ROOT, "duckdb-robust-predicate-transfer_b200")


def build_host_mirror_test(force=False):
    """Compiles tests/cpp/host_mirror_test (g++; the C++ host mirror include/rpt_gpu.hpp against librpt_gpu.so,
    with the oracle library linked in as the checker).  Returns the path of the binary."""
    root = ROOT
    src = os.path.join(root, "tests", "cpp", "host_mirror_test.cpp")
    out = os.path.join(root, "tests", "cpp", "host_mirror_test")
    deps = [src, os.path.join(root, "include", "rpt_gpu.hpp"), os.path.join(root, "include", "rpt_gpu.h"),
            os.path.join(root, "oracle", "rpt_oracle.h")]
    if not force and os.path.exists(out) and all(os.path.getmtime(d) <= os.path.getmtime(out) for d in deps):
        return out
    cmd = ["g++", "-std=c++17", "-O2", "-Wall", "-I" + os.path.join(root, "include"), "-I" + os.path.join(root, "oracle"),
           src, "-o", out, "-L" + PKG, "-L" + os.path.join(root, "oracle"), "-lrpt_gpu", "-lrpt_oracle",
           "-Wl,-rpath,$ORIGIN/../../" + os.path.basename(PKG) + ":$ORIGIN/../../oracle"]
    proc = subprocess.run(cmd, capture_output=True, text=True)
    if proc.returncode != 0:
        raise RuntimeError("g++ failed:\n" + proc.stdout + proc.stderr)
    return out
