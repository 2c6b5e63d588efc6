#!/usr/bin/env python
"""End-to-end runs of BASELINE configs 1 (3-table variant), 4 (TPC-H-shaped star joins) and 5 (JOB-shaped synthetic
IMDB) through the REAL operators: DuckDB's shell + the reference's rpt extension, whose PTBloomFilter body is either
the reference's own (CPU: `--ext stock`) or integration/bloom_filter_gpu.cpp over librpt_gpu.so (`--ext gpu`).

The method is the reference's own (test_job_queries.sh:229-300): run every query without the extension and with it,
`-noheader -list`, and diff the outputs.  `--ext none` writes / checks the golden outputs of plain DuckDB
(tests/golden/e2e/*.out, generated on the CPU box); `--ext gpu` needs a B200 (there is no CPU fallback).

    python integration/e2e/run_e2e.py --ext none --write-golden          # CPU box, once
    python integration/e2e/run_e2e.py --ext stock                        # CPU: the unmodified reference extension
    python integration/e2e/run_e2e.py --ext gpu [--tpch-sf 10] [--job-scale 4] [--live-baseline]   # GPU box

Binaries come from integration/build_extension.sh (baseline/_ref/bin/, git-ignored, travels with gpurun).
"""
import argparse
import json
import os
import subprocess
import sys
import time

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
BIN = os.path.join(ROOT, "baseline", "_ref", "bin")
GOLDEN = os.path.join(ROOT, "tests", "golden", "e2e")


def workloads(args):
    return [
        ("config1_three_tables", [], ["config1_three_tables.sql"], {}),
        ("config4_tpch_star", [f"CALL dbgen(sf={args.tpch_sf});"], ["config4_tpch_star.sql"], {"sf": args.tpch_sf}),
        ("config5_job", [f"SET VARIABLE scale={args.job_scale};"], ["config5_job_setup.sql", "config5_job_queries.sql"],
         {"scale": args.job_scale}),
    ]


def run_duckdb(ext, pre, files, settings, threads, trace_path=None):
    cmd = [os.path.join(BIN, "duckdb"), "-unsigned", "-noheader", "-list"]
    if threads:
        cmd += ["-c", f"SET threads={threads};"]
    for p in pre:
        cmd += ["-c", p]
    if ext != "none":
        cmd += ["-c", f"LOAD '{os.path.join(BIN, ext, 'rpt.duckdb_extension')}';"]
        for s in settings:
            cmd += ["-c", s]
    for f in files:
        cmd += ["-c", f".read {os.path.join(HERE, f)}"]
    env = dict(os.environ)
    if ext == "gpu":
        env["RPT_TRACE"] = "1"  # the library reports every host-column insert / probe call on stderr
    t0 = time.time()
    proc = subprocess.run(cmd, capture_output=True, text=True, env=env, cwd="/tmp")
    dt = time.time() - t0
    if trace_path:
        with open(trace_path, "w") as f:
            f.write(proc.stderr)
    return proc.returncode, proc.stdout, proc.stderr, dt


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--ext", default="gpu", choices=["none", "stock", "gpu"])
    ap.add_argument("--tpch-sf", type=float, default=1)
    ap.add_argument("--job-scale", type=int, default=1)
    ap.add_argument("--threads", type=int, default=0)
    ap.add_argument("--write-golden", action="store_true")
    ap.add_argument("--live-baseline", action="store_true", help="also run plain DuckDB here and diff against it "
                    "(needed when the scale differs from the golden files')")
    ap.add_argument("--setting", action="append", default=[], help="extra SET after LOAD, e.g. \"SET rpt_pass_mode='forward_only';\"")
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out"))
    ap.add_argument("--only", default="")
    args = ap.parse_args()
    os.makedirs(args.out, exist_ok=True)
    os.makedirs(GOLDEN, exist_ok=True)
    report, failed = [], False
    for name, pre, files, params in workloads(args):
        if args.only and args.only not in name:
            continue
        golden_path = os.path.join(GOLDEN, name + ".out")
        default_scale = (name != "config4_tpch_star" or args.tpch_sf == 1) and (name != "config5_job" or args.job_scale == 1)
        trace = os.path.join(args.out, f"e2e_{args.ext}_{name}.stderr")
        rc, out, err, dt = run_duckdb(args.ext, pre, files, args.setting, args.threads, trace)
        with open(os.path.join(args.out, f"e2e_{args.ext}_{name}.out"), "w") as f:
            f.write(out)
        entry = {"workload": name, "ext": args.ext, "params": params, "rc": rc, "seconds": round(dt, 2),
                 "result_rows": len(out.splitlines())}
        if rc != 0:
            entry["error"] = err[-2000:]
            failed = True
        if args.ext == "none" and args.write_golden and default_scale and rc == 0:
            with open(golden_path, "w") as f:
                f.write(out)
            entry["golden"] = "written"
        elif rc == 0:
            want = None
            if args.live_baseline or not default_scale:
                rc0, want, _, dt0 = run_duckdb("none", pre, files, [], args.threads)
                entry["baseline"] = f"plain DuckDB run here ({dt0:.1f} s)"
            elif os.path.exists(golden_path):
                want = open(golden_path).read()
                entry["baseline"] = "tests/golden/e2e (plain DuckDB, CPU box)"
            if want is not None:
                got_l, want_l = out.splitlines(), want.splitlines()
                diff = [(g, w) for g, w in zip(got_l, want_l) if g != w]
                entry["identical_to_plain_duckdb"] = (out == want)
                if out != want:
                    entry["first_differences"] = diff[:5] + ([("len", len(got_l), len(want_l))] if len(got_l) != len(want_l) else [])
                    failed = True
        if args.ext == "gpu":
            lines = err.splitlines()
            entry["gpu_insert_calls"] = sum(1 for l in lines if "[rpt trace] insert host" in l)
            entry["gpu_probe_calls"] = sum(1 for l in lines if "[rpt trace] probe host" in l)
            rows_in = sum(int(l.split("insert host:")[1].split("rows")[0]) for l in lines if "[rpt trace] insert host" in l)
            rows_pr = sum(int(l.split("probe host:")[1].split("rows")[0]) for l in lines if "[rpt trace] probe host" in l)
            entry["gpu_rows_inserted"], entry["gpu_rows_probed"] = rows_in, rows_pr
        report.append(entry)
        print(json.dumps(entry), flush=True)
    with open(os.path.join(args.out, f"e2e_{args.ext}_report.json"), "w") as f:
        json.dump(report, f, indent=1)
    return 1 if failed else 0


if __name__ == "__main__":
    sys.exit(main())
