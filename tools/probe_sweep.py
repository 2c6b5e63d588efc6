"""Development helper: build + probe timings over filter sizes on cuda:0 (CUDA events, after warm-up).
    python tools/probe_sweep.py [probe_rows=1e9] [reps=5] [cases=80000,1e7,4e7,1e8]
Prints one line per case; env knobs (RPT_PROBE_UNITS, RPT_SLICE_MIB, RPT_GPU_LIB) are echoed."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import rpt_b200 as rpt

m = int(float(sys.argv[1])) if len(sys.argv) > 1 else 1_000_000_000
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 5
cases = [int(float(x)) for x in (sys.argv[3] if len(sys.argv) > 3 else "80000,1e7,4e7,1e8").split(",")]
ppm = int(os.environ.get("SWEEP_PPM", "100000"))
ktype = os.environ.get("SWEEP_TYPE", "int64")  # int64 | int32 (low 31 bits of the synthetic keys) | varchar (decimal text of the keys)


def decimal_text(x):
    """int64 tensor (non-negative after masking) → Arrow utf8 layout on the device: (int32 offsets[n+1], uint8 bytes):
    k2 = decimal text of k1.  Column by column to bound the temporaries."""
    x = x & 0x7fffffffffffffff
    n = x.numel()
    ndig = torch.ones(n, dtype=torch.int64, device=x.device)
    p = 10
    for _ in range(18):
        ndig += (x >= p).to(torch.int64)
        p *= 10
    offsets = torch.zeros(n + 1, dtype=torch.int64, device=x.device)
    torch.cumsum(ndig, 0, out=offsets[1:])
    total = int(offsets[-1].item())
    blob = torch.empty(total + 8, dtype=torch.uint8, device=x.device)
    p = 1
    for d in range(19):  # digit d counted from the right goes to offsets[i+1] - 1 - d
        has = ndig > d
        pos = (offsets[1:] - 1 - d)[has]
        blob[pos] = ((x[has] // p) % 10 + 48).to(torch.uint8)
        p *= 10
        del has, pos
    return offsets.to(torch.int32), blob, total


dev = torch.device("cuda", 0)
stream = torch.cuda.Stream(device=dev)
torch.cuda.set_stream(stream)
sh = stream.cuda_stream
knobs = {k: os.environ[k] for k in ("RPT_PROBE_UNITS", "RPT_SLICE_MIB", "RPT_GPU_LIB", "RPT_PROBE_PARTITION", "RPT_PART_MINB", "RPT_PART_STORE", "RPT_L2_PERSIST_MIB", "RPT_PREFILTER", "RPT_PRE_INLINE0", "RPT_PREFILTER_MAX_PASS") if k in os.environ}
print(f"# m={m} reps={reps} ppm={ppm} type={ktype} knobs={knobs}", flush=True)
dp = torch.empty(m, dtype=torch.int64, device=dev)
dsel = torch.empty(m, dtype=torch.int32, device=dev)
dfound = torch.zeros(1, dtype=torch.int64, device=dev)
for n in cases:
    dk = torch.empty(n, dtype=torch.int64, device=dev)
    rpt.synth_build_keys(dk, 0, n, stream=sh)
    rpt.synth_probe_keys(dp, 0, m, n, ppm, stream=sh)
    bf = rpt.PTBloomFilter(0).Initialize(n)
    if ktype == "int32":
        dk32, dp32 = (dk & 0x7fffffff).to(torch.int32), (dp & 0x7fffffff).to(torch.int32)
        torch.cuda.synchronize()
        kc, pc = rpt.device_col(dk32, rpt.INT32), rpt.device_col(dp32, rpt.INT32)
    elif ktype == "two":  # two key columns (generic source): (k, low 31 bits of k)
        dk32, dp32 = (dk & 0x7fffffff).to(torch.int32), (dp & 0x7fffffff).to(torch.int32)
        torch.cuda.synchronize()
        kc = [rpt.device_col(dk, rpt.INT64), rpt.device_col(dk32, rpt.INT32)]
        pc = [rpt.device_col(dp, rpt.INT64), rpt.device_col(dp32, rpt.INT32)]
    elif ktype == "varchar":
        ko, kb, kt = decimal_text(dk)
        po, pb, pt = decimal_text(dp)
        torch.cuda.synchronize()
        kc = rpt.device_col(ko, rpt.VARCHAR, aux=kb, aux_bytes=kt, data_len=n)
        pc = rpt.device_col(po, rpt.VARCHAR, aux=pb, aux_bytes=pt, data_len=m)
    else:
        kc, pc = rpt.device_col(dk, rpt.INT64), rpt.device_col(dp, rpt.INT64)
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
    for _ in range(2):
        bf.Reset(n, sh)
        bf.Insert(kc, n, sh)
        bf.LookupSelDevice(pc, m, dsel, dfound, sh)
    torch.cuda.synchronize()
    ev[0].record(stream)
    for _ in range(reps):
        bf.Reset(n, sh)
        bf.Insert(kc, n, sh)
    ev[1].record(stream)
    torch.cuda.synchronize()
    ev[2].record(stream)
    for _ in range(reps):
        bf.LookupSelDevice(pc, m, dsel, dfound, sh)
    ev[3].record(stream)
    torch.cuda.synchronize()
    F = bf.num_sectors * 8
    bm, pm = ev[0].elapsed_time(ev[1]) / reps, ev[2].elapsed_time(ev[3]) / reps
    found = int(dfound.item())
    algo = (4 if ktype == "int32" else (12 if ktype == "two" else 8)) * m + 4 * found + F
    if ktype == "varchar":
        algo = 4 * (m + 1) + pt + 4 * found + F  # offsets + string bytes replace 8·m
    # whole steps (reset + build + probe): what bench.py times; includes the fold of a pre-filtered probe
    ev2 = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    ev2[0].record(stream)
    for _ in range(reps):
        bf.Reset(n, sh)
        bf.Insert(kc, n, sh)
        bf.LookupSelDevice(pc, m, dsel, dfound, sh)
    ev2[1].record(stream)
    torch.cuda.synchronize()
    sm = ev2[0].elapsed_time(ev2[1]) / reps
    pre = bf.PrefilterInfo()
    print(f"n={n:>10} F={F / 2**20:8.3f} MiB  build {bm:7.3f} ms {n / bm / 1e6:7.1f} Gkeys/s | probe {pm:7.3f} ms "
          f"{m / pm / 1e6:7.1f} Gkeys/s  {algo / pm / 1e6:7.1f} GB/s algorithmic  found {found} | step {sm:7.3f} ms "
          f"prefilter state {pre[0]} pass {pre[1]:.4f}", flush=True)
    bf.close()
    del dk
