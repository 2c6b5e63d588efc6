"""Two (or more) processes, each with its own CUDA context — on ONE GPU if that is all there is — build row-range
shards of a key column and merge them with rpt_bf_peer_allreduce (the library-owned OR-allreduce: barrier words in
device memory written over CUDA IPC, waits through cuStreamWaitValue32).  torch.distributed (gloo) only carries the
64-byte IPC handles.  Every rank compares the merged words with a single build of the whole column, over several
epochs (reset → insert → merge) so that the deferred "everybody has pulled" barrier is exercised too.
    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 tests/multi/peer_allreduce_check.py"""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import rpt_b200 as rpt  # noqa: E402
from helpers import synth_build_keys  # noqa: E402


def main():
    dist.init_process_group("gloo")
    rank, world = dist.get_rank(), dist.get_world_size()
    device = rank % torch.cuda.device_count()
    torch.cuda.set_device(device)
    est = int(os.environ.get("PEER_CHECK_ROWS", "3000000"))
    bf = rpt.PTBloomFilter(device).Initialize(est)
    stream = torch.cuda.Stream()
    sh = stream.cuda_stream

    def exchange(blob):
        out = [None] * world
        dist.all_gather_object(out, blob)
        return out

    bf.PeerOpen(exchange(bf.IpcExport()), rank)
    bf.PeerFlagsOpen(exchange(bf.PeerFlagsExport()), rank)
    for epoch in range(4):
        n = est - 1000 * epoch
        keys = synth_build_keys(17 * epoch, n)
        lo, hi = n * rank // world, n * (rank + 1) // world
        if epoch == 2 and rank == 0:
            lo = hi  # a rank without rows: it learns from the barrier words that the others had some
        bf.PeerRelease(sh)  # everybody has pulled my slice of the previous merge: I may change my words
        bf.Reset(est, sh)
        if hi > lo:
            bf.Insert(rpt.host_col(keys[lo:hi]), hi - lo, sh)
        torch.cuda.synchronize()
        bf.PeerAllreduce(hi > lo, sh)
        assert bf.PeerAnyData(sh)
        stream.synchronize()
        got = bf.ExportWords()
        single = rpt.PTBloomFilter(device).Initialize(est)
        if epoch == 2:
            cut = n // world  # rank 0's shard was left out
            single.Insert(rpt.host_col(keys[cut:]), n - cut)
        else:
            single.Insert(rpt.host_col(keys), n)
        want = single.ExportWords()
        single.close()
        assert np.array_equal(got, want), f"rank {rank} epoch {epoch}: merged words differ from the single build"
        assert not bf.IsEmpty()
    dist.barrier()
    bf.PeerRelease(sh)
    stream.synchronize()
    dist.barrier()
    bf.close()
    if rank == 0:
        print(f"peer allreduce ok: {world} ranks on {torch.cuda.device_count()} device(s), 4 epochs, bit-exact")
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
