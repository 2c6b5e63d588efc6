"""Development helper: README-variant filter (rpt_abf_*) build / fold / probe timings on cuda:0 next to the
reference-exact filter (rpt_bf_*) on the same columns.  CUDA events on the launching stream, after warm-up.
    python tools/acero_sweep.py [probe_rows=1e9] [reps=5] [cases=8e4,1e7,1e8] [oversize=1,16]
oversize = CreateEmpty is called with oversize x the true row count (a planner over-estimate), which is what gives
Fold something to do."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import rpt_b200 as rpt

m = int(float(sys.argv[1])) if len(sys.argv) > 1 else 1_000_000_000
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 5
cases = [int(float(x)) for x in (sys.argv[3] if len(sys.argv) > 3 else "8e4,1e7,1e8").split(",")]
oversizes = [int(x) for x in (sys.argv[4] if len(sys.argv) > 4 else "1,16").split(",")]
ppm = 100000

dev = torch.device("cuda", 0)
stream = torch.cuda.Stream(device=dev)
torch.cuda.set_stream(stream)
sh = stream.cuda_stream


def timed(fn, reps):
    fn()
    stream.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(stream)
    for _ in range(reps):
        fn()
    b.record(stream)
    stream.synchronize()
    return a.elapsed_time(b) / reps


print(f"# m={m} reps={reps} ppm={ppm}", flush=True)
dp = torch.empty(m, dtype=torch.int64, device=dev)
dsel = torch.empty(m, dtype=torch.int32, device=dev)
dfound = torch.zeros(1, dtype=torch.int64, device=dev)
for n in cases:
    dk = torch.empty(n, dtype=torch.int64, device=dev)
    rpt.synth_build_keys(dk, 0, n, stream=sh)
    rpt.synth_probe_keys(dp, 0, m, n, ppm, stream=sh)
    stream.synchronize()
    kc, pc = rpt.device_col(dk, rpt.INT64), rpt.device_col(dp, rpt.INT64)
    # reference-exact filter on the same columns
    bf = rpt.PTBloomFilter(0).Initialize(n)
    t_b = timed(lambda: bf.Insert(kc, n, stream=sh), reps)
    t_p = timed(lambda: bf.LookupSelDevice(pc, m, dsel, dfound, stream=sh), reps)
    f_ref = int(dfound.item())
    print(f"n={n:>11} rpt_bf  F={bf.num_sectors * 8 / 2**20:8.2f} MiB build {t_b:7.3f} ms {n / t_b / 1e6:7.1f} Gk/s | "
          f"probe {t_p:7.3f} ms {m / t_p / 1e6:7.1f} Gk/s found={f_ref}", flush=True)
    bf.close()
    for ov in oversizes:
        f = rpt.BlockedBloomFilter(0).CreateEmpty(n * ov)
        log0 = f.log_num_blocks
        t_b = timed(lambda: f.Insert(kc, n, stream=sh), reps)
        t_p0 = timed(lambda: f.FindDevice(pc, m, dsel, dfound, stream=sh), reps)
        f0 = int(dfound.item())
        t0 = time.perf_counter()
        k = f.Fold(stream=sh)
        t_f = (time.perf_counter() - t0) * 1e3
        t_p1 = timed(lambda: f.FindDevice(pc, m, dsel, dfound, stream=sh), reps)
        f1 = int(dfound.item())
        print(f"n={n:>11} rpt_abf x{ov:<3} F={2**log0 * 8 / 2**20:8.2f} MiB build {t_b:7.3f} ms {n / t_b / 1e6:7.1f} Gk/s | "
              f"probe {t_p0:7.3f} ms {m / t_p0 / 1e6:7.1f} Gk/s found={f0} | fold x{k} {t_f:6.3f} ms (host wall) -> "
              f"F={2**f.log_num_blocks * 8 / 2**20:8.2f} MiB probe {t_p1:7.3f} ms {m / t_p1 / 1e6:7.1f} Gk/s found={f1}",
              flush=True)
        f.close()
    del dk
