"""Development helper for ncu captures: one build + a few probes of one configuration on cuda:0.
    python tools/probe_one.py <build_rows> <probe_rows> [reps=3] [ppm=100000]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import rpt_b200 as rpt
n, m = int(float(sys.argv[1])), int(float(sys.argv[2]))
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 3
ppm = int(sys.argv[4]) if len(sys.argv) > 4 else 100_000
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
bf.Insert(rpt.device_col(dk, rpt.INT64), n, sh)
for _ in range(reps):
    bf.LookupSelDevice(rpt.device_col(dp, rpt.INT64), m, dsel, dfound, sh)
torch.cuda.synchronize()
print("found", int(dfound.item()))
