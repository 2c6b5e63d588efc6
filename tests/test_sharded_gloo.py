"""CPU suite: the multi-rank host logic (row-range shards → OR-merge → shard-local probe → ordered gather,
global-row-count resize rule) under gloo with world_size 2 and 3.  The filter arithmetic is the
oracle's (tests/oracle_engine.py); what is under test is rpt_b200.ShardedBloomFilter's plumbing."""
import os
import socket
import sys
import traceback

import numpy as np
import pytest
import torch.multiprocessing as mp

HERE = os.path.dirname(os.path.abspath(__file__))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, case, q):
    try:
        sys.path.insert(0, HERE)
        sys.path.insert(0, os.path.dirname(HERE))
        import torch.distributed as dist
        os.environ["MASTER_ADDR"] = "127.0.0.1"
        os.environ["MASTER_PORT"] = str(port)
        dist.init_process_group("gloo", rank=rank, world_size=world)
        import oracle_engine
        import oracle_lib
        import rpt_b200
        from helpers import synth_build_keys, synth_probe_keys
        oracle = oracle_lib.oracle()

        n, m, est = case["n"], case["m"], case["est"]
        keys = synth_build_keys(0, n)
        probe = synth_probe_keys(0, m, max(n, 1), 200000)
        # single-process answer
        one = oracle_engine.PTBloomFilter().Initialize(est)
        one.Insert(rpt_b200.host_col(keys), n)
        rebuilt_one = one.Finalize(n, rpt_b200.host_col(keys), n)
        want_sel = one.LookupSel(rpt_b200.host_col(probe), m)

        lo, hi = rpt_b200.shard_range(n, world, rank)
        plo, phi = rpt_b200.shard_range(m, world, rank)
        sb = rpt_b200.ShardedBloomFilter(engine=oracle_engine).Initialize(est)
        assert sb.merge == "allgather"
        shard = rpt_b200.host_col(keys[lo:hi])
        sb.Insert(shard, hi - lo)
        rebuilt = sb.Finalize(hi - lo, shard, hi - lo)
        assert rebuilt == rebuilt_one
        assert sb.finalized_ and sb.num_sectors == one.num_sectors
        assert sb.IsEmpty() == (n == 0)
        assert np.array_equal(sb.ExportWords(), one.ExportWords()), "merged words differ from the 1-shard build"
        local = sb.LookupSel(rpt_b200.host_col(probe[plo:phi]), phi - plo)
        glob = sb.GatherSelection(local, plo)
        assert np.array_equal(glob, want_sel.astype(np.uint64)), "gathered selection differs"
        dist.barrier()
        dist.destroy_process_group()
        q.put((rank, None))
    except Exception:
        q.put((rank, traceback.format_exc()))


CASES = [
    {"n": 100003, "m": 300007, "est": 100003},   # good estimate
    {"n": 50000, "m": 100000, "est": 1000},      # under-estimate: the GLOBAL count fires the resize rule
    {"n": 70, "m": 5000, "est": 64},             # tiny: shard boundaries at multiples of 64, rank 0 gets 64/0 rows
    {"n": 0, "m": 1000, "est": 10},              # empty build side: filter stays empty, nothing survives
]


@pytest.mark.parametrize("world", [2, 3])
@pytest.mark.parametrize("case", CASES, ids=lambda c: f"n{c['n']}_est{c['est']}")
def test_sharded_matches_single(world, case):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, case, q)) for r in range(world)]
    for p in procs:
        p.start()
    results = [q.get(timeout=180) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    for rank, err in results:
        assert err is None, f"rank {rank}:\n{err}"


def test_shard_range_properties():
    import rpt_b200
    for total in (0, 1, 63, 64, 65, 1000, 10**8, 10**9 + 7):
        for world in (1, 2, 3, 4, 8):
            prev = 0
            for r in range(world):
                lo, hi = rpt_b200.shard_range(total, world, r)
                assert lo == prev and hi >= lo
                assert lo % 64 == 0 or lo == total
                prev = hi
            assert prev == total
