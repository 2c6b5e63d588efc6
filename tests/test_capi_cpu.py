"""CPU suite: the C-ABI library loads, exports exactly what include/rpt_gpu.h declares, restates the
reference's sizing arithmetic, and fails loudly (no CPU fallback) when there is no CUDA device."""
import ctypes as C
import os
import re
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "rpt_gpu.h")


def declared_symbols():
    text = open(HEADER).read()
    return sorted(set(re.findall(r"RPT_API\s+[\w\s\*]+?\b(rpt_\w+)\s*\(", text)))


def test_header_compiles_as_c(tmp_path):
    src = tmp_path / "t.c"
    src.write_text('#include "rpt_gpu.h"\nint main(void){rpt_col_t c; (void)c; return sizeof(rpt_col_t)==56?0:1;}\n')
    exe = tmp_path / "t"
    subprocess.run(["gcc", "-std=c99", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"), str(src), "-o",
                    str(exe)], check=True)
    assert subprocess.run([str(exe)]).returncode == 0


def test_exports_every_declared_symbol(rpt):
    decl = declared_symbols()
    assert len(decl) >= 30
    out = subprocess.run(["nm", "-D", "--defined-only", rpt.LIB], check=True, capture_output=True, text=True).stdout
    exported = sorted(set(l.split()[-1] for l in out.splitlines() if " T rpt_" in l))
    assert exported == decl
    for name in decl:  # and the Python binding knows every one of them
        assert name in rpt._SIGNATURES, name
        assert getattr(rpt.lib(), name) is not None


def test_col_struct_layout_matches(rpt):
    import oracle_lib
    assert C.sizeof(rpt.RptCol) == C.sizeof(oracle_lib.Col) == 56
    for (n1, _), (n2, _) in zip(rpt.RptCol._fields_, oracle_lib.Col._fields_):
        assert n1 == n2
        assert getattr(rpt.RptCol, n1).offset == getattr(oracle_lib.Col, n2).offset


def test_sizing_matches_oracle(rpt, oracle):
    for rows in [0, 1, 42, 43, 50000, 100000, 5 * 10**7, 10**8, 357913941, 357913942, 2**32 - 1, 2**40]:
        assert rpt.sectors_for_rows(rows) == oracle.num_sectors(rows), rows


def test_sizing_overflow_is_an_error(rpt):
    # NextPowerOfTwo throws OutOfRangeException when the result wraps (constants.cpp:43-45)
    with pytest.raises(rpt.RptError):
        rpt.sectors_for_rows((2**63 + 12) // 12 + 1)


def test_no_cpu_fallback(rpt):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(rpt.RptError, match="no CPU fallback"):
        rpt.PTBloomFilter().Initialize(1000)


def test_product_does_not_reference_oracle():
    """The oracle is test infrastructure: nothing under the package or include/ may mention it."""
    pkg = os.path.join(ROOT, "duckdb-robust-predicate-transfer_b200")
    for base, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".hpp", ".h")):
                text = open(os.path.join(base, f), errors="replace").read()
                assert "rpt_oracle" not in text and "librpt_ref" not in text and "oracle_lib" not in text, f


def test_cpp_host_mirror_compiles(rpt):
    """include/rpt_gpu.hpp (the C++ mirror of PTBloomFilter + the CREATE_BF / USE_BF glue) and its test driver
    build with g++ alone against the C ABI; the driver itself runs in the gpu suite."""
    import os
    import oracle_lib
    oracle_lib.oracle()  # makes sure oracle/librpt_oracle.so exists (the driver links the checker)
    from cpp import build_cpp
    path = build_cpp.build_host_mirror_test(force=True)
    assert os.path.exists(path) and os.access(path, os.X_OK)
