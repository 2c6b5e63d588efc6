"""Debug helper (torchrun, N ranks): overlapped slice-wise merge vs a single-GPU build, then the probe behind the gates."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.distributed as dist
import rpt_b200 as rpt

world, rank, lr = int(os.environ["WORLD_SIZE"]), int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
dev = torch.device("cuda", lr)
dist.init_process_group("nccl", device_id=dev)
n, m = 100_000_000, 200_000_000
lo, hi = rpt.shard_range(n, world, rank)
stream = torch.cuda.Stream(device=dev)
torch.cuda.set_stream(stream)
sh = stream.cuda_stream
dk = torch.empty(hi - lo, dtype=torch.int64, device=dev)
dall = torch.empty(n, dtype=torch.int64, device=dev)
dp = torch.empty(m, dtype=torch.int64, device=dev)
rpt.synth_build_keys(dk, lo, hi - lo, device=lr, stream=sh)
rpt.synth_build_keys(dall, 0, n, device=lr, stream=sh)
rpt.synth_probe_keys(dp, rank * m, m, n, 100_000, device=lr, stream=sh)
dsel = torch.empty(m, dtype=torch.int32, device=dev)
dfound = torch.zeros(1, dtype=torch.int64, device=dev)
torch.cuda.synchronize()
single = rpt.PTBloomFilter(lr).Initialize(n)
single.Insert(rpt.device_col(dall, rpt.INT64), n, sh)
torch.cuda.synchronize()
want = single.Checksum()
single.LookupSelDevice(rpt.device_col(dp, rpt.INT64), m, dsel, dfound, sh)
torch.cuda.synchronize()
want_found = int(dfound.item())
want_sel = dsel[:want_found].clone()
sb = rpt.ShardedBloomFilter(device=lr).Initialize(n)
for it in range(6):
    mode = os.environ.get("DBG_SYNC", "0")
    sb.Reset(n, sh)
    sb.Insert(rpt.device_col(dk, rpt.INT64), hi - lo, sh)
    sb.Merge(sh)
    if mode == "1":
        torch.cuda.synchronize()
        got = sb.bf.Checksum()
        print(f"rank {rank} it {it}: merged checksum {'OK' if got == want else 'DIFFERS'}", flush=True)
    sb.LookupSelDevice(rpt.device_col(dp, rpt.INT64), m, dsel, dfound, sh)
    torch.cuda.synchronize()
    f = int(dfound.item())
    if f != want_found and it == 5:
        have = torch.zeros(m, dtype=torch.bool, device=dev)
        have[dsel[:f].long()] = True
        missing = want_sel[~have[want_sel.long()]].long()
        segs = torch.unique(missing // 512)
        per = torch.bincount((missing // 512) - segs.min())
        print(f"rank {rank}: {missing.numel()} missing rows in {segs.numel()} segments of 512 rows; first segments "
              f"{segs[:8].tolist()}; max per segment {int(per.max())}; rows {missing[:12].tolist()}", flush=True)
    got = sb.bf.Checksum()
    print(f"rank {rank} it {it}: found {f} want {want_found} {'OK' if f == want_found else 'MISMATCH'}; checksum after probe "
          f"{'OK' if got == want else 'DIFFERS'}", flush=True)
sb.close()
dist.barrier()
dist.destroy_process_group()
