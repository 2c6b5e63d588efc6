#!/usr/bin/env python
"""bench.py — BASELINE.json's metric on BASELINE.json's config 2 (single-key int64 bloom build 100M keys /
probe 1B keys, synthetic uniform, SURVEY §8d generators).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--sel-ppm 100000] [--impl reference]

A step = one pass of the hot path: zero the filter, build it from the build column, (N>1: OR-merge the partial
filters over NVLink), probe the probe column into a compacted, ordered selection vector.
  value   whole-job Gkeys/s = (build keys + probe keys of all ranks) / step time, inputs resident in HBM,
          CUDA-event timed, max over ranks.
  e2e     the same step through the C ABI with HOST (pinned) key buffers and host selection-vector output:
          H2D and D2H copies inside the timed region.
  roofline  the probe (dominant; all launches of one rpt_bf_probe call): algorithmic bytes 8·m + 4·found + F ÷ the
          CUDA-event duration of the call, against MEASURED_PEAKS.json's HBM copy bandwidth.
  cpu_baseline  the reference's own loops (oracle/_ref, compiled from the reference sources) or the C oracle port,
          on the box's host cores, bounded sample (N=1, rank 0 only).
N>1 (torchrun, one rank per GPU): the 100M-row build column is sharded by row range across ranks (each builds a
partial filter, OR-merged over NVLink peer memory); every rank probes its own 1B-row shard of an N·1B-row probe
column against the replicated filter.  Per-GPU probe work is fixed → "weak".

--impl reference times the reference CPU implementation (same metric, bounded sample per step).
The oracle/ directory is used only as the checker / CPU baseline, never on the measured GPU path.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

METRIC = "bloom build+probe throughput (build 100M int64 keys, probe 1B int64 keys per GPU)"
UNIT = "Gkeys/s"
_REAL_STDOUT = sys.stdout


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--build-rows", type=int, default=100_000_000)
    ap.add_argument("--probe-rows", type=int, default=1_000_000_000, help="probe rows per GPU")
    ap.add_argument("--sel-ppm", type=int, default=100_000, help="member fraction of the probe column in ppm (10 %%)")
    ap.add_argument("--e2e-steps", type=int, default=3)
    ap.add_argument("--e2e-warmup", type=int, default=3, help="untimed e2e steps before the timed ones (the first "
                    "host-column calls after the device-column legs grow the stream-ordered pool and the pipeline's "
                    "buffers: 1.3 s, then 0.2 s, then the steady 0.17 s per step — profiles/bench_r2y.json)")
    ap.add_argument("--clock-warmup-s", type=float, default=0.0, help="extra untimed steps for this many seconds "
                    "before the timed region (sustained-load measurement: the step then runs under sw_power_cap "
                    "at ~1.71 GHz, 114.5 instead of 121 Gkeys/s; profiles/bench_r1h_sustained.json)")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--cpu-probe-sample", type=int, default=200_000_000, help="probe rows of the cpu_baseline leg "
                    "inside the GPU arm (a bounded sample); --impl reference always probes ALL probe rows every step")
    ap.add_argument("--no-sweep", action="store_true", help="skip detail.sweep (1 %% / 50 %% members, the 16 MiB "
                    "L2-resident companion, config 3); N=1 only, outside the timed steps")
    ap.add_argument("--no-config3", action="store_true")
    ap.add_argument("--sustain-s", type=float, default=2.0, help="detail.sustained: after the timed steps, this many "
                    "seconds of untimed back-to-back steps, then the same K steps timed again (power-capped clocks)")
    ap.add_argument("--merge", default="auto", choices=["auto", "peer", "allgather"])
    ap.add_argument("--sweep", action="store_true", help="(default since round 2; kept for old command lines)")
    return ap.parse_args()


def load_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            p = json.load(f)
        return float(p["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler:
    """SM clock / throttle reasons sampled DURING the timed region through NVML (nvidia_ml_py) every few
    milliseconds from a background thread — the timed region of this bench is tens of milliseconds, too short for
    `nvidia-smi -lms`.  Falls back to one nvidia-smi query when NVML is not importable."""
    REASONS = {"hw_slowdown": 0x8, "sw_power_cap": 0x4, "sw_thermal_slowdown": 0x20, "hw_thermal_slowdown": 0x40}

    def __init__(self, index):
        self.index, self.rows, self.thread, self.stop_flag = index, [], None, False
        self.nvml, self.handle = None, None
        try:
            import pynvml
            pynvml.nvmlInit()
            # NVML enumerates physical GPUs; honour CUDA_VISIBLE_DEVICES when it is a plain index list
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            phys = index
            if vis:
                ids = [v.strip() for v in vis.split(",") if v.strip()]
                if index < len(ids) and ids[index].isdigit():
                    phys = int(ids[index])
            self.handle = pynvml.nvmlDeviceGetHandleByIndex(phys)
            self.nvml = pynvml
        except Exception:
            self.nvml = None

    def _loop(self):
        n = self.nvml
        while not self.stop_flag:
            try:
                sm = n.nvmlDeviceGetClockInfo(self.handle, n.NVML_CLOCK_SM)
                reasons = n.nvmlDeviceGetCurrentClocksEventReasons(self.handle) \
                    if hasattr(n, "nvmlDeviceGetCurrentClocksEventReasons") \
                    else n.nvmlDeviceGetCurrentClocksThrottleReasons(self.handle)
                power = n.nvmlDeviceGetPowerUsage(self.handle) / 1000.0
                self.rows.append((sm, reasons, power))
            except Exception:
                pass
            time.sleep(0.002)

    def start(self):
        if self.nvml:
            self.stop_flag = False
            self.thread = threading.Thread(target=self._loop, daemon=True)
            self.thread.start()

    def stop(self):
        if not self.nvml:
            return self._smi_once()
        self.stop_flag = True
        self.thread.join(timeout=2)
        n = self.nvml
        try:
            mx = n.nvmlDeviceGetMaxClockInfo(self.handle, n.NVML_CLOCK_SM)
        except Exception:
            mx = None
        sm = sorted(r[0] for r in self.rows)
        mask = 0
        for r in self.rows:
            mask |= int(r[1])
        reasons = sorted(name for name, bit in self.REASONS.items() if mask & bit)
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx,
                "power_w_max": max((r[2] for r in self.rows), default=None), "samples": len(sm), "reasons": reasons,
                "how": "NVML, 2 ms period, during the timed region"}

    def _smi_once(self):
        try:
            out = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=clocks.sm,clocks.max.sm,power.draw",
                                  "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=10).stdout
            sm, mx, pw = [float(x) for x in out.strip().split(",")]
            return {"sm_mhz": sm, "sm_max_mhz": mx, "power_w_max": pw, "samples": 1, "reasons": [],
                    "how": "one nvidia-smi query after the timed region (NVML not importable)"}
        except Exception:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"], "samples": 0}


# ---------------------------------------------------------------------------------------------------------
# CPU arm: the reference's own loops (oracle/_ref) or the oracle port, on the host cores
# ---------------------------------------------------------------------------------------------------------
def cpu_run(args, steps, warmup, full_probe):
    """Returns (value Gkeys/s, seconds per step, info dict).  One step = build ALL build rows into a fresh filter +
    probe the probe column: all of it (full_probe, the --impl reference arm: nothing is extrapolated) or its first
    --cpu-probe-sample rows scaled to the full row count (the bounded cpu_baseline leg of the GPU arm)."""
    import numpy as np
    import oracle_lib
    oracle = oracle_lib.oracle()
    ref = oracle_lib.ref()
    kind = "reference" if ref is not None else "port"
    eng = ref if ref is not None else oracle
    cores = os.cpu_count() or 1
    n, m = args.build_rows, args.probe_rows
    ms = m if full_probe else min(m, args.cpu_probe_sample)
    keys = oracle.synth_build_keys(0, n).view(np.int64)
    probe = oracle.synth_probe_keys(0, ms, n, args.sel_ppm).view(np.int64)
    S = oracle.num_sectors(n)
    tb, tp, tw, found = [], [], [], 0
    for it in range(warmup + steps):
        words = np.zeros(S, dtype=np.uint64)
        t0 = time.perf_counter()
        b = eng.bench_build(words, keys, cores)
        p, found = eng.bench_probe(words, probe, cores)
        wall = time.perf_counter() - t0
        if it >= warmup:
            tb.append(b)
            tp.append(p)
            tw.append(wall)
    tb_med = sorted(tb)[len(tb) // 2]
    tp_med = sorted(tp)[len(tp) // 2]
    if full_probe:
        t_step = sum(tw) / len(tw)  # wall clock of a whole step, averaged over the K steps actually run
        sample = (f"every step builds all {n} keys on {cores} threads into one shared {S * 8 >> 20} MiB filter (median "
                  f"{tb_med:.3f} s) and probes ALL {m} probe rows (median {tp_med:.3f} s); nothing extrapolated; "
                  "2048-row chunks, relaxed fetch_or, private selection buffers")
    else:
        t_step = tb_med + tp_med * (m / ms)
        sample = (f"build all {n} keys on {cores} threads into one shared {S * 8 >> 20} MiB filter (median "
                  f"{tb_med:.3f} s) + probe the first {ms} of {m} probe rows (median {tp_med:.3f} s, scaled x"
                  f"{m / ms:.1f}); 2048-row chunks, relaxed fetch_or, private selection buffers")
    value = (n + m) / t_step / 1e9
    info = {"value": value, "unit": UNIT, "cores": cores, "kind": kind, "sample": sample,
            "build_gkeys_s": n / tb_med / 1e9, "probe_gkeys_s": ms / tp_med / 1e9,
            "probe_rows_timed": int(ms), "probe_found": int(found)}
    return value, t_step, info


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    steps, warmup = max(1, args.steps), max(0, args.warmup)
    value, t_full, info = cpu_run(args, steps, warmup, full_probe=True)
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": steps, "warmup": warmup,
        "ms_per_step": t_full * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u64", "data": "synthetic",
        "config": workload_config(args, max(1, args.gpus)), "cpu_baseline": info,
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    _REAL_STDOUT.write(json.dumps(line) + "\n")
    _REAL_STDOUT.flush()
    return 0


def filter_sectors(rows):
    """S of BloomFilter::Initialize (duckdb/src/planner/filter/bloom_filter.cpp:7-9,17-19) — plain arithmetic, so that
    both arms can describe the workload without loading anything."""
    bits = max(512, 12 * rows)
    p2 = 1 << (bits - 1).bit_length()
    return min(p2 >> 6, 1 << 26)


def workload_config(args, world):
    """The SAME dict from both arms (the driver compares them)."""
    S = filter_sectors(args.build_rows)
    merge = "none" if world == 1 else ("peer" if args.merge == "auto" else args.merge)
    return {"workload": "config 2: single-key int64 bloom build 100M keys / probe 1B keys, synthetic uniform "
                        "(SURVEY §8d splitmix64 generators)",
            "build_rows_total": args.build_rows, "probe_rows_per_gpu": args.probe_rows,
            "member_fraction_ppm": args.sel_ppm, "filter_sectors": S, "filter_mib": (S * 8) >> 20,
            "sharding": f"build column row-range sharded over {world} rank(s), OR-merged; probe shard-local",
            "merge": merge,
            "l2_policy": "inputs larger than L2 (8 GB key column + 256 MiB filter per step); no flush needed"}


# ---------------------------------------------------------------------------------------------------------
# GPU arm
# ---------------------------------------------------------------------------------------------------------
def run_b200(args):
    import numpy as np
    import torch
    import torch.distributed as dist

    import rpt_b200 as rpt

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus and world > 1:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}")
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: librpt_gpu has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    rpt.lib()  # fails loudly when the extension is missing

    n, m, ppm = args.build_rows, args.probe_rows, args.sel_ppm
    lo, hi = rpt.shard_range(n, world, rank)
    n_local = hi - lo
    probe_first = rank * m

    # a non-default stream: a NULL rpt_stream_t would mean "the handle's own stream", which torch events cannot see
    stream = torch.cuda.Stream(device=dev)
    torch.cuda.set_stream(stream)
    sh = stream.cuda_stream
    assert sh != 0
    dk = torch.empty(max(n_local, 2), dtype=torch.int64, device=dev)
    dp = torch.empty(m, dtype=torch.int64, device=dev)
    rpt.synth_build_keys(dk, lo, n_local, device=local_rank, stream=sh)
    rpt.synth_probe_keys(dp, probe_first, m, n, ppm, device=local_rank, stream=sh)
    dsel = torch.empty(m, dtype=torch.int32, device=dev)
    dfound = torch.zeros(1, dtype=torch.int64, device=dev)
    torch.cuda.synchronize()

    sb = rpt.ShardedBloomFilter(device=local_rank, merge=args.merge).Initialize(n)
    S = sb.num_sectors
    F = S * 8
    build_col = rpt.device_col(dk, rpt.INT64)
    probe_col = rpt.device_col(dp, rpt.INT64)

    def step(ev=None):
        sb.Reset(n, sh)                       # zero-fill (BloomFilter::Initialize)
        sb.Insert(build_col, n_local, sh)     # build kernel
        if ev:
            ev[1].record(stream)
        sb.Merge(sh)                          # N>1: OR-allreduce over NVLink
        if ev:
            ev[2].record(stream)
        sb.LookupSelDevice(probe_col, m, dsel, dfound, sh)  # probe kernel
        if ev:
            ev[3].record(stream)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    sampler = ClockSampler(local_rank)  # NVML initialised before the warm-up: no idle gap in front of the timed region
    for _ in range(args.warmup):
        step()
    barrier()
    # Optional sustained-load mode (--clock-warmup-s): further UNTIMED steps so that the K timed steps run in the
    # power/clock state of a long job (on this pool: ~830 W, sw_power_cap, 1.71 GHz) instead of the burst state.
    spin_steps = 0
    if args.clock_warmup_s > 0:
        t_spin = time.perf_counter()
        for _ in range(3):
            step()
        torch.cuda.synchronize()
        t_step = (time.perf_counter() - t_spin) / 3
        if world > 1:  # every rank must run the same number of steps (the step holds collectives)
            t = torch.tensor([t_step], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            t_step = float(t.item())
        spin_steps = int(min(1000, max(0, args.clock_warmup_s / max(t_step, 1e-4))))
        for _ in range(spin_steps):
            step()
        spin_steps += 3
    barrier()
    found = int(dfound.item())
    launches0 = sb.launch_count

    sampler.start()
    evs = [[torch.cuda.Event(enable_timing=True) for _ in range(4)] for _ in range(args.steps)]
    t_start = torch.cuda.Event(enable_timing=True)
    t_end = torch.cuda.Event(enable_timing=True)
    barrier()
    t_start.record(stream)
    for k in range(args.steps):
        evs[k][0].record(stream)
        step(evs[k])
    t_end.record(stream)
    barrier()
    clocks = sampler.stop()
    total_ms = t_start.elapsed_time(t_end)
    build_ms = sum(e[0].elapsed_time(e[1]) for e in evs) / args.steps
    merge_ms = sum(e[1].elapsed_time(e[2]) for e in evs) / args.steps
    probe_ms = sum(e[2].elapsed_time(e[3]) for e in evs) / args.steps
    launches = sb.launch_count - launches0

    # max over ranks
    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    total_ms = max_over_ranks(total_ms)
    probe_ms_max = max_over_ranks(probe_ms)
    build_ms_max = max_over_ranks(build_ms)
    merge_ms_max = max_over_ranks(merge_ms)
    ms_per_step = total_ms / args.steps
    keys_per_step = n + world * m
    value = keys_per_step / (ms_per_step * 1e-3) / 1e9

    # ---- parity spot-check inside the bench run (not timed): property checks on this rank's results ----
    sel = dsel[:found]
    ordered = bool((sel[1:] > sel[:-1]).all()) if found > 1 else True
    member = dp >= 0
    members = int(member.sum().item())
    members_kept = int(member[sel.long()].sum().item())
    del member, sel
    parity = {"sel_strictly_ascending": ordered, "members": members, "members_kept": members_kept,
              "found": found, "false_positive_rate": (found - members_kept) / max(m - members, 1)}
    if not ordered or members != members_kept:
        raise SystemExit(f"bench parity check failed on rank {rank}: {parity}")

    # ---- bit-exactness of the (merged) filter, not timed: FNV-1a-64 of the words (SURVEY §8c recipe) must be the
    # same on every rank and equal to a filter built on THIS GPU alone from all n build keys (SURVEY §8e:
    # "N-shard filter bytes == 1-shard filter bytes") ----
    fnv = sb.bf.Checksum()
    if world > 1:
        dk_all = torch.empty(n, dtype=torch.int64, device=dev)
        rpt.synth_build_keys(dk_all, 0, n, device=local_rank, stream=sh)
        single = rpt.PTBloomFilter(local_rank).Initialize(n)
        single.Insert(rpt.device_col(dk_all, rpt.INT64), n, sh)
        torch.cuda.synchronize()
        fnv_single = single.Checksum()
        single.close()
        del dk_all
        halves = torch.tensor([fnv >> 32, fnv & 0xffffffff, fnv_single >> 32, fnv_single & 0xffffffff],
                              dtype=torch.int64, device=dev)
        lo_t, hi_t = halves.clone(), halves.clone()
        dist.all_reduce(lo_t, op=dist.ReduceOp.MIN)
        dist.all_reduce(hi_t, op=dist.ReduceOp.MAX)
        ranks_agree = bool((lo_t == hi_t).all().item())
    else:
        fnv_single, ranks_agree = fnv, True
    parity["filter_fnv1a"] = f"0x{fnv:016x}"
    parity["filter_fnv1a_single_gpu_build"] = f"0x{fnv_single:016x}"
    parity["ranks_agree"] = ranks_agree
    if fnv != fnv_single or not ranks_agree:
        raise SystemExit(f"bench parity check failed on rank {rank}: merged filter checksum 0x{fnv:016x}, single-GPU "
                         f"build 0x{fnv_single:016x}, ranks agree: {ranks_agree}")

    # ---- roofline of the dominant kernel (probe) ----
    peak, peak_src = load_peaks()
    algo_bytes = 8 * m + 4 * found + F
    achieved = algo_bytes / (probe_ms * 1e-3) / 1e9
    pre_state, pre_pass = sb.bf.PrefilterInfo()
    kernels = ("rpt_bf_probe = fold_kernel (once per build) + probe_fold_kernel<FlatU64SrcT<0>,24,4> + 4 x probe_bucket_half_kernel<4> + "
               "count_segments/scan_blocks/scan_totals + compact_bitmap_kernel" if pre_state == 1 else
               "rpt_bf_probe = probe_partition_kernel<FlatU64SrcT<0>,256,4> + 3 x probe_bucket_kernel<168> + "
               "count_segments/scan_blocks/scan_totals + compact_bitmap_kernel")
    roofline = {"kernel": kernels, "bound": "hbm", "achieved": achieved, "peak": peak,
                "unit": "GB/s", "frac": achieved / peak, "traffic": None, "peak_source": peak_src,
                "algorithmic_bytes_per_launch": algo_bytes, "launch_ms": probe_ms,
                "bytes_per_key": algo_bytes / m,
                "build": {"algorithmic_bytes_per_launch": 8 * n_local + 2 * F, "ms": build_ms,
                          "achieved": (8 * n_local + 2 * F) / (build_ms * 1e-3) / 1e9,
                          "frac": (8 * n_local + 2 * F) / (build_ms * 1e-3) / 1e9 / peak,
                          "note": "zero-fill + build kernel; random 8-byte atomics, not a streaming kernel"}}
    traffic_path = os.path.join(ROOT, "profiles", "probe_traffic.json")
    if os.path.exists(traffic_path):
        try:
            with open(traffic_path) as f:
                tr = json.load(f)
            if pre_state != 1 and "without_prefilter" in tr:
                tr = tr["without_prefilter"]
            roofline["traffic"] = tr.get("dram_bytes_per_launch")
            roofline["traffic_source"] = tr.get("source")
        except Exception:
            pass

    detail = {"build_ms": build_ms_max, "merge_ms": merge_ms_max, "probe_ms": probe_ms_max,
              "build_gkeys_s": n / (build_ms_max * 1e-3) / 1e9,
              "probe_gkeys_s": world * m / (probe_ms_max * 1e-3) / 1e9,
              "found_rank0": found, "filter_sectors": S, "filter_mib": F >> 20, "parity": parity,
              "prefilter": {"state": pre_state, "estimated_pass_rate_of_non_members": pre_pass,
                            "note": "folded pre-filter of the probe (rpt_gpu.h rpt_bf_set_prefilter): 1 = in use, "
                                    "0 = none, 2 = made but too dense"}}

    # ---- sustained figure: the timed region above is a fraction of a second (burst clocks); the same K steps after
    # --sustain-s seconds of back-to-back untimed steps run in the power/clock state of a long job ----
    if args.sustain_s > 0:
        spin = int(min(2000, max(1, args.sustain_s / max(ms_per_step * 1e-3, 1e-4))))
        sampler2 = ClockSampler(local_rank)
        for _ in range(spin):
            step()
        barrier()
        sampler2.start()
        s0, s1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s0.record(stream)
        for _ in range(args.steps):
            step()
        s1.record(stream)
        barrier()
        clocks2 = sampler2.stop()
        sus_ms = max_over_ranks(s0.elapsed_time(s1)) / args.steps
        detail["sustained"] = {"value": keys_per_step / (sus_ms * 1e-3) / 1e9, "unit": UNIT, "ms_per_step": sus_ms,
                               "untimed_spin_steps": spin, "clocks": clocks2}

    # ---- sweep (N=1, outside the timed steps): BASELINE config 2's other selectivities, the L2-resident companion
    # (SURVEY §8d "2-L2"), and config 3 ----
    if world == 1 and not args.no_sweep:
        detail["sweep"] = sweep(rpt, torch, dev, local_rank, sb, dp, dsel, dfound, n, m, sh, stream, peak)

    # ---- e2e: host buffers through the C ABI, copies inside the timed region ----
    e2e = None
    if not args.no_e2e:
        e2e = run_e2e(args, rpt, torch, dist, dev, local_rank, world, rank, sb, dk, dp, n, m, n_local, sh, barrier,
                      max_over_ranks, found)

    if world == 1 and not args.no_sweep and not args.no_config3:
        del dk, dp, dsel
        torch.cuda.empty_cache()
        detail["config3"] = config3(rpt, torch, dev, local_rank, sh, stream, peak)
        dk = dp = dsel = None
        torch.cuda.empty_cache()

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        del dk, dp, dsel
        torch.cuda.empty_cache()
        try:
            _, _, cpu = cpu_run(args, steps=1, warmup=0, full_probe=False)
        except Exception as exc:  # the baseline must not take the GPU numbers down with it
            cpu = {"value": None, "unit": UNIT, "cores": os.cpu_count(), "kind": "port", "sample": f"failed: {exc}"}

    if rank == 0:
        cfg = workload_config(args, world)
        assert cfg["filter_sectors"] == S and cfg["merge"] == (sb.merge if world > 1 else "none")
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u64", "data": "synthetic", "config": cfg,
            "clocks": clocks, "clock_warmup_steps": spin_steps, "e2e": e2e, "gpu_launches": int(launches), "roofline": roofline,
            "cpu_baseline": cpu, "detail": detail,
        }
        _REAL_STDOUT.write(json.dumps(line) + "\n")
        _REAL_STDOUT.flush()
    sb.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


def sweep(rpt, torch, dev, local_rank, sb, dp, dsel, dfound, n, m, sh, stream, peak):
    out = {}
    probe_col = rpt.device_col(dp, rpt.INT64)
    F = sb.num_sectors * 8

    def time_probe(bf, reps=5):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        for _ in range(2):
            bf.LookupSelDevice(probe_col, m, dsel, dfound, sh)
        torch.cuda.synchronize()
        e0.record(stream)
        for _ in range(reps):
            bf.LookupSelDevice(probe_col, m, dsel, dfound, sh)
        e1.record(stream)
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / reps, int(dfound.item())

    for ppm in (10_000, 500_000):
        rpt.synth_probe_keys(dp, 0, m, n, ppm, device=local_rank, stream=sh)
        ms, found = time_probe(sb)
        b = 8 * m + 4 * found + F
        out[f"probe_sel_{ppm // 10000}pct"] = {"ms": ms, "gkeys_s": m / ms / 1e6, "found": found,
                                               "achieved_gbs": b / ms / 1e6, "frac": b / ms / 1e6 / peak}
    # L2-resident companion (SURVEY §8d "2-L2"): n = 10^7 → S = 2^21, F = 16 MiB
    n2 = 10_000_000
    dk2 = torch.empty(n2, dtype=torch.int64, device=dev)
    rpt.synth_build_keys(dk2, 0, n2, device=local_rank, stream=sh)
    rpt.synth_probe_keys(dp, 0, m, n2, 100_000, device=local_rank, stream=sh)
    small = rpt.PTBloomFilter(local_rank).Initialize(n2)
    small.Insert(rpt.device_col(dk2, rpt.INT64), n2, sh)
    ms, found = time_probe(small)
    b = 8 * m + 4 * found + small.num_sectors * 8
    out["probe_l2_resident_16MiB_10pct"] = {"ms": ms, "gkeys_s": m / ms / 1e6, "found": found,
                                            "achieved_gbs": b / ms / 1e6, "frac": b / ms / 1e6 / peak}
    small.close()
    rpt.synth_probe_keys(dp, 0, m, n, 100_000, device=local_rank, stream=sh)
    torch.cuda.synchronize()
    return out


def config3(rpt, torch, dev, local_rank, sh, stream, peak, n=50_000_000, m=500_000_000, ppm=100_000):
    """BASELINE config 3 as the reference executes it: a two-column join key (int64 k1, varchar k2 =
    decimal text of k1) is TWO independent single-column filters (physical_create_bf.cpp:221-227); a USE_BF probes
    filter 1 with the whole chunk and filter 2 with filter 1's survivors (physical_use_bf.cpp:137-180).  Run twice:
    est_rows = n (right-sized) and est_rows = n/5, where the resize rule of Finalize (physical_create_bf.cpp:386-406)
    fires and both filters are re-allocated and re-hashed (src/bloom_filter.cpp:34-58) — the reference has no fold."""
    import ctypes as C
    out = {"build_rows": n, "probe_rows": m, "member_fraction_ppm": ppm,
           "shape": "2 x single-column filter (int64, varchar decimal text); filter 2 probes filter 1's survivors"}
    BATCH = 200_000_000  # the VARCHAR layout has 32-bit offsets: a call carries < 4 GiB of string bytes
    k1 = torch.empty(n, dtype=torch.int64, device=dev)
    p1 = torch.empty(m, dtype=torch.int64, device=dev)
    rpt.synth_build_keys(k1, 0, n, device=local_rank, stream=sh)
    rpt.synth_probe_keys(p1, 0, m, n, ppm, device=local_rank, stream=sh)
    torch.cuda.synchronize()
    ko, kb, kbytes = rpt.synth_decimal_text(k1)
    c_k1 = rpt.device_col(k1, rpt.INT64)
    c_k2 = rpt.device_col(ko, rpt.VARCHAR, aux=kb, aux_bytes=kbytes, data_len=n)
    batches, pbytes = [], 0
    for r0 in range(0, m, BATCH):
        r1 = min(r0 + BATCH, m)
        po, pb, nb = rpt.synth_decimal_text(p1[r0:r1])
        batches.append((r0, r1 - r0, rpt.device_col(p1[r0:r1], rpt.INT64),
                        rpt.device_col(po, rpt.VARCHAR, aux=pb, aux_bytes=nb, data_len=r1 - r0), nb))
        pbytes += nb
    torch.cuda.synchronize()
    out["probe_batches"] = [b[1] for b in batches]
    sel1 = torch.empty(BATCH, dtype=torch.int32, device=dev)
    sel2 = torch.empty(m, dtype=torch.int32, device=dev)
    found = torch.zeros(2, dtype=torch.int64, device=dev)
    f1, f2 = rpt.PTBloomFilter(local_rank), rpt.PTBloomFilter(local_rank)
    for label, est in (("est_rows_exact", n), ("est_rows_under_5x", n // 5)):
        f1.Initialize(est)
        f2.Initialize(est)

        def build():
            f1.Reset(est, sh)
            f2.Reset(est, sh)
            f1.Insert(c_k1, n, sh)
            f2.Insert(c_k2, n, sh)
            r1 = f1.Finalize(n, c_k1, n, sh)   # resize rule; rebuilds from the whole column when it fires
            r2 = f2.Finalize(n, c_k2, n, sh)
            return r1, r2

        def probe(times=None):
            n1_total = n2_total = 0
            for r0, mb, c_p1, c_p2, _ in batches:
                ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
                ev[0].record(stream)
                f1.LookupSelDevice(c_p1, mb, sel1, found[0:1], sh)
                ev[1].record(stream)
                n1 = int(found[0].item())   # the operator learns the survivor count between the two filters
                rpt._check(rpt.lib().rpt_bf_filter_sel(f2._h, C.byref(c_p2.raw), sel1.data_ptr(), n1,
                                                       sel2[n2_total:].data_ptr(), found[1:2].data_ptr(), rpt.MEM_DEVICE,
                                                       sh))
                ev[2].record(stream)
                n2 = int(found[1].item())
                sel2[n2_total:n2_total + n2] += r0     # batch-relative → column row ids (not timed as probe work)
                if times is not None:
                    torch.cuda.synchronize()
                    times[0] += ev[0].elapsed_time(ev[1])
                    times[1] += ev[1].elapsed_time(ev[2])
                n1_total += n1
                n2_total += n2
            return n1_total, n2_total

        build()
        probe()
        torch.cuda.synchronize()
        reps = 3
        tb = 0.0
        tp = [0.0, 0.0]
        for _ in range(reps):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            r1, r2 = build()
            e1.record(stream)
            torch.cuda.synchronize()
            tb += e0.elapsed_time(e1) / reps
            n1, n2 = probe(tp)
        tp1, tp2 = tp[0] / reps, tp[1] / reps
        s2 = sel2[:n2].long() & 0xFFFFFFFF
        members = p1 >= 0
        ok = bool((s2[1:] > s2[:-1]).all().item()) and int(members[s2].sum().item()) == int(members.sum().item())
        del members, s2
        if not ok:
            raise SystemExit(f"config 3 ({label}): selection not ascending or a member row was dropped")
        F1, F2 = f1.num_sectors * 8, f2.num_sectors * 8
        nbat = len(batches)
        b1 = 8 * m + 4 * n1 + nbat * F1                       # keys + selection + filter once per call
        b2 = 4 * n1 + 4 * n1 + int(pbytes * (n1 / m)) + 4 * n2 + nbat * F2   # in-selection + offsets + bytes of the survivors
        bb = 8 * n + 4 * (n + 1) + kbytes + 2 * (F1 + F2)
        out[label] = {
            "est_rows": est, "rebuilt": [bool(r1), bool(r2)], "filter_mib": [F1 >> 20, F2 >> 20],
            "filter_fnv1a": [f"0x{f1.Checksum():016x}", f"0x{f2.Checksum():016x}"],
            "build_ms": tb, "build_gkeys_s": 2 * n / tb / 1e6,
            "probe_int64_ms": tp1, "probe_int64_gkeys_s": m / tp1 / 1e6, "found_filter1": n1,
            "probe_varchar_survivors_ms": tp2, "found_both": n2,
            "probe_gkeys_s": m / (tp1 + tp2) / 1e6,
            "roofline": {"bound": "hbm", "unit": "GB/s", "peak": peak,
                         "probe_int64": {"algorithmic_bytes": b1, "achieved": b1 / tp1 / 1e6, "frac": b1 / tp1 / 1e6 / peak},
                         "probe_varchar_survivors": {"algorithmic_bytes": b2, "achieved": b2 / tp2 / 1e6,
                                                     "frac": b2 / tp2 / 1e6 / peak},
                         "build": {"algorithmic_bytes": bb, "achieved": bb / tb / 1e6, "frac": bb / tb / 1e6 / peak}},
        }
    # the varchar column probed on its own, all m rows (the staged string-hash kernel + the flat pre-hashed path)
    def probe_varchar():
        total = 0
        for r0, mb, _, c_p2, _ in batches:
            f2.LookupSelDevice(c_p2, mb, sel2, found[1:2], sh)
            total += int(found[1].item())
        return total
    probe_varchar()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record(stream)
    for _ in range(3):
        nv = probe_varchar()
    e1.record(stream)
    torch.cuda.synchronize()
    tv = e0.elapsed_time(e1) / 3
    bv = 4 * (m + len(batches)) + pbytes + 4 * nv + len(batches) * f2.num_sectors * 8
    out["probe_varchar_all_rows"] = {"ms": tv, "gkeys_s": m / tv / 1e6, "found": nv, "algorithmic_bytes": bv,
                                     "achieved_gbs": bv / tv / 1e6, "frac": bv / tv / 1e6 / peak}
    f1.close()
    f2.close()
    return out


def run_e2e(args, rpt, torch, dist, dev, local_rank, world, rank, sb, dk, dp, n, m, n_local, sh, barrier,
            max_over_ranks, found_device):
    """Host (pinned) key buffers in, host selection vector out, every step: the call a DuckDB operator makes."""
    import numpy as np
    hk = rpt.PinnedArray(max(n_local, 1), np.int64)
    hp = rpt.PinnedArray(m, np.int64)
    hsel = rpt.PinnedArray(m, np.uint32)
    # fill the host buffers from the device-generated columns (not timed)
    cudart = torch.cuda.cudart()
    torch.cuda.synchronize()
    hk_t = torch.from_numpy(hk.array)
    hp_t = torch.from_numpy(hp.array)
    if n_local:
        hk_t.copy_(dk[:n_local])
    hp_t.copy_(dp)
    torch.cuda.synchronize()
    del cudart
    build_col = rpt.host_col(hk.array[:n_local] if n_local else hk.array[:0])
    probe_col = rpt.KeyCol(rpt.RptCol(hp.ptr, None, None, None, m, 0, rpt.INT64, rpt.MEM_HOST), (hp,))
    import ctypes as C
    found = C.c_uint64(0)

    def step():
        sb.Reset(n, sh)
        torch.cuda.current_stream().synchronize()
        sb.Insert(build_col, n_local)                 # H2D of the build shard + build kernels
        sb.Merge(None)
        rpt._check(rpt.lib().rpt_bf_probe(sb.bf._h, C.byref(probe_col.raw), 1, m, hsel.ptr, C.addressof(found),
                                          rpt.MEM_HOST, None))  # H2D keys, probe kernels, D2H selection vector

    for _ in range(max(1, args.e2e_warmup)):  # warm-up (pinned staging, allocator pools, the pipeline's slots)
        step()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.e2e_steps):
        step()
    barrier()
    dt = (time.perf_counter() - t0) / args.e2e_steps
    dt = max_over_ranks(dt)
    if found.value != found_device:
        raise SystemExit(f"e2e parity: host path found {found.value}, device path {found_device}")
    keys_per_step = n + world * m
    out = {"value": keys_per_step / dt / 1e9, "unit": UNIT, "ms_per_step": dt * 1e3,
           "h2d_bytes_per_step": 8 * n_local + 8 * m, "d2h_bytes_per_step": 4 * int(found.value) + 8,
           "steps": args.e2e_steps, "warmup": max(1, args.e2e_warmup),
           "note": "per rank: pinned host key columns → rpt_bf_insert / rpt_bf_probe (C ABI, RPT_MEM_HOST) → host "
                   "selection vector; PCIe-bound"}
    hk.free()
    hp.free()
    hsel.free()
    return out


def main():
    # stdout carries exactly ONE JSON line: libraries that print there (NCCL's version banner) go to stderr
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    args = parse_args()
    if args.impl == "reference":
        return run_reference(args)
    return run_b200(args)


if __name__ == "__main__":
    sys.exit(main())
