#!/bin/bash
# bitmap-pass probe: parity, then sweeps over staging variants / slice sizes, PCIe bandwidth
set -u
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_b.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu_b.log
{
RPT_PROBE_UNITS=8 python tools/probe_sweep.py 1e9 5
RPT_PROBE_UNITS=4 python tools/probe_sweep.py 1e9 5
RPT_PROBE_UNITS=8 RPT_SLICE_MIB=88 python tools/probe_sweep.py 1e9 5 1e8
RPT_PROBE_UNITS=4 RPT_SLICE_MIB=88 python tools/probe_sweep.py 1e9 5 1e8
RPT_PROBE_UNITS=4 RPT_SLICE_MIB=48 python tools/probe_sweep.py 1e9 5 1e8
RPT_PROBE_UNITS=4 SWEEP_PPM=10000 python tools/probe_sweep.py 1e9 5 1e7,1e8
RPT_PROBE_UNITS=4 SWEEP_PPM=500000 python tools/probe_sweep.py 1e9 5 1e7,1e8
} > gpurun_out/sweep_b.txt 2>&1
cat gpurun_out/sweep_b.txt
python - > gpurun_out/pcie.txt 2>&1 <<'PY'
import torch, time
for mb in (16, 128, 1024):
    n = mb << 20
    h = torch.empty(n, dtype=torch.uint8).pin_memory()
    d = torch.empty(n, dtype=torch.uint8, device="cuda")
    for name, fn in (("H2D", lambda: d.copy_(h, non_blocking=True)), ("D2H", lambda: h.copy_(d, non_blocking=True))):
        fn(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(5): fn()
        e1.record(); torch.cuda.synchronize()
        print(f"{name} {mb} MiB pinned: {5 * n / e0.elapsed_time(e1) / 1e6:.1f} GB/s")
    # both directions at once on two streams
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
    h2 = torch.empty(n, dtype=torch.uint8).pin_memory(); d2 = torch.empty(n, dtype=torch.uint8, device="cuda")
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(5):
        with torch.cuda.stream(s1): d.copy_(h, non_blocking=True)
        with torch.cuda.stream(s2): h2.copy_(d2, non_blocking=True)
    torch.cuda.synchronize(); dt = time.perf_counter() - t0
    print(f"bidirectional {mb} MiB: {5 * n / dt / 1e9:.1f} GB/s each way")
PY
cat gpurun_out/pcie.txt
