"""The drop-in boundary against the REAL reference.

CPU (where /root/reference is mounted): integration/bloom_filter_gpu.cpp — the replacement of the reference's
src/bloom_filter.cpp — and the reference's own, unmodified operator sources compile against
integration/bloom_filter.hpp + DuckDB's headers (g++ -fsyntax-only: seconds, not the 9-minute engine build).

GPU (-m gpu): BASELINE configs 1 (3-table variant), 4 (TPC-H-shaped star joins) and 5 (JOB-shaped synthetic IMDB) run
through DuckDB's shell with the reference extension whose PTBloomFilter body is librpt_gpu.so, and are diffed against
plain DuckDB — the method of the reference's test_job_queries.sh:229-300.  Needs baseline/_ref/bin (built by
integration/build_extension.sh in the container where the reference is mounted; git-ignored, travels with gpurun)."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference"
BIN = os.path.join(ROOT, "baseline", "_ref", "bin")
INC = ["-I" + os.path.join(ROOT, "integration"), "-I" + os.path.join(ROOT, "include"),
       "-I" + os.path.join(REF, "duckdb", "src", "include"), "-I" + os.path.join(REF, "src", "include"),
       "-I" + os.path.join(REF, "src")]

needs_reference = pytest.mark.skipif(not os.path.isdir(os.path.join(REF, "duckdb", "src", "include")),
                                     reason="/root/reference is not mounted here")


@needs_reference
def test_integration_tu_compiles_against_the_reference_headers():
    subprocess.run(["g++", "-std=c++17", "-fsyntax-only", "-Wall"] + INC +
                   [os.path.join(ROOT, "integration", "bloom_filter_gpu.cpp")], check=True)


@needs_reference
@pytest.mark.parametrize("src", ["operators/physical_create_bf.cpp", "operators/physical_use_bf.cpp"])
def test_reference_operators_compile_unmodified_against_the_replacement_header(src):
    """`#include "bloom_filter.hpp"` in the reference's operator sources resolves to integration/bloom_filter.hpp."""
    subprocess.run(["g++", "-std=c++17", "-fsyntax-only"] + INC + [os.path.join(REF, "src", src)], check=True)


def test_golden_outputs_of_plain_duckdb_are_committed():
    for name in ("config1_three_tables", "config4_tpch_star", "config5_job"):
        path = os.path.join(ROOT, "tests", "golden", "e2e", name + ".out")
        assert os.path.getsize(path) > 0, path


@pytest.mark.gpu
def test_e2e_configs_through_the_duckdb_operators(tmp_path):
    have = all(os.path.exists(os.path.join(BIN, p)) for p in ("duckdb", "gpu/rpt.duckdb_extension"))
    if not have:
        pytest.skip("baseline/_ref/bin is missing: run integration/build_extension.sh where /root/reference is mounted")
    proc = subprocess.run([sys.executable, os.path.join(ROOT, "integration", "e2e", "run_e2e.py"), "--ext", "gpu",
                           "--out", str(tmp_path)], capture_output=True, text=True, timeout=1500)
    report = [json.loads(l) for l in proc.stdout.splitlines() if l.startswith("{")]
    assert proc.returncode == 0, proc.stdout[-3000:] + proc.stderr[-2000:]
    assert len(report) == 3
    for e in report:
        assert e["rc"] == 0 and e["identical_to_plain_duckdb"] is True, e
        # the filters were really built and probed on the device (RPT_TRACE lines of librpt_gpu.so)
        assert e["gpu_insert_calls"] > 0 and e["gpu_rows_inserted"] > 0, e
        assert e["gpu_probe_calls"] > 0 or e["workload"] != "config1_three_tables", e
