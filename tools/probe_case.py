"""Development helper: one build + a few probes of a given shape on cuda:0 (for ncu captures and quick timings).
    python tools/probe_case.py <build_rows> <probe_rows> [sel_ppm=100000] [reps=3]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import rpt_b200 as rpt

n, m = int(float(sys.argv[1])), int(float(sys.argv[2]))
ppm = int(sys.argv[3]) if len(sys.argv) > 3 else 100000
reps = int(sys.argv[4]) if len(sys.argv) > 4 else 3
dev = torch.device("cuda", 0)
stream = torch.cuda.Stream(device=dev)
torch.cuda.set_stream(stream)
sh = stream.cuda_stream
dk = torch.empty(n, dtype=torch.int64, device=dev)
dp = torch.empty(m, dtype=torch.int64, device=dev)
rpt.synth_build_keys(dk, 0, n, stream=sh)
rpt.synth_probe_keys(dp, 0, m, n, ppm, stream=sh)
dsel = torch.empty(m, dtype=torch.int32, device=dev)
dfound = torch.zeros(1, dtype=torch.int64, device=dev)
bf = rpt.PTBloomFilter(0).Initialize(n)
e = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
e[0].record(stream)
bf.Insert(rpt.device_col(dk, rpt.INT64), n, sh)
e[1].record(stream)
bf.LookupSelDevice(rpt.device_col(dp, rpt.INT64), m, dsel, dfound, sh)
torch.cuda.synchronize()
e[2].record(stream)
for _ in range(reps):
    bf.LookupSelDevice(rpt.device_col(dp, rpt.INT64), m, dsel, dfound, sh)
e[3].record(stream)
torch.cuda.synchronize()
F = bf.num_sectors * 8
pm = e[2].elapsed_time(e[3]) / reps
print(f"n={n} m={m} F={F >> 20} MiB build {e[0].elapsed_time(e[1]):.3f} ms ({n / e[0].elapsed_time(e[1]) / 1e6:.1f} Gkeys/s) "
      f"probe {pm:.3f} ms ({m / pm / 1e6:.1f} Gkeys/s) found {int(dfound.item())}")
