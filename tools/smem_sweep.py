"""Development helper: regime-1 (shared-memory filter) probe rate on cuda:0 for int64 / int32 keys, small filters.
    python tools/smem_sweep.py [probe_rows=1e9]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import rpt_b200 as rpt

m = int(float(sys.argv[1])) if len(sys.argv) > 1 else 1_000_000_000
dev = torch.device("cuda", 0)
stream = torch.cuda.Stream(device=dev)
torch.cuda.set_stream(stream)
sh = stream.cuda_stream
dp = torch.empty(m, dtype=torch.int64, device=dev)
dsel = torch.empty(m, dtype=torch.int32, device=dev)
dfound = torch.zeros(1, dtype=torch.int64, device=dev)
for n in (1000, 10_000, 40_000, 80_000):
    dk = torch.empty(n, dtype=torch.int64, device=dev)
    rpt.synth_build_keys(dk, 0, n, stream=sh)
    rpt.synth_probe_keys(dp, 0, m, n, 100_000, stream=sh)
    for kt in ("int64", "int32"):
        if kt == "int32":
            k, p = (dk & 0x7fffffff).to(torch.int32), (dp & 0x7fffffff).to(torch.int32)
            kc, pc, kb = rpt.device_col(k, rpt.INT32), rpt.device_col(p, rpt.INT32), 4
        else:
            kc, pc, kb = rpt.device_col(dk, rpt.INT64), rpt.device_col(dp, rpt.INT64), 8
        torch.cuda.synchronize()
        bf = rpt.PTBloomFilter(0).Initialize(n)
        bf.Insert(kc, n, sh)
        for _ in range(2):
            bf.LookupSelDevice(pc, m, dsel, dfound, sh)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record(stream)
        for _ in range(5):
            bf.LookupSelDevice(pc, m, dsel, dfound, sh)
        e1.record(stream)
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 5
        found = int(dfound.item())
        F = bf.num_sectors * 8
        algo = kb * m + 4 * found + F
        print(f"n={n:>6} F={F / 1024:7.1f} KiB {kt}: probe {ms:6.3f} ms {m / ms / 1e6:7.1f} Gkeys/s  {algo / ms / 1e6:7.1f} GB/s "
              f"algorithmic ({algo / ms / 1e6 / 6538.6:.3f} of copy peak, {algo / ms / 1e6 / 7170:.3f} of the 7.17 TB/s read stream)  found {found}", flush=True)
        bf.close()
