"""Row-range sharded build / replicated probe across ranks: one process per GPU.

The reference has no distributed path (single process, N sink threads sharing one filter through relaxed
fetch_or, duckdb/src/planner/filter/bloom_filter.cpp:63-70).  fetch_or is commutative, associative and
idempotent, so the same words result when every rank inserts its contiguous row range into a private
full-size filter and the partial filters are OR-merged:

    rank r: Insert(rows [n·r/N, n·(r+1)/N))  →  OR-allreduce of the S words  →  every rank holds the
    filter a single process would have built  →  each rank probes its own row range; selection vectors
    are row-ordered inside a shard, and shards are contiguous, so concatenating them with the shard's base
    row offset reproduces LookupHashes' ascending order (bloom_filter.cpp:56-59).

The resize rule (physical_create_bf.cpp:386-406) needs the GLOBAL row count: a SUM all-reduce of one scalar.

torch.distributed is plumbing only (rendezvous, barriers, the scalar all-reduce, the handle exchange).
The merge itself is the library's: `peer` = two-phase reduce-scatter-OR + all-gather over NVLink peer
memory (rpt_bf_peer_*), `allgather` = NCCL all-gather of the partial filters + one OR kernel
(rpt_bf_or_parts).  NCCL has no bitwise-OR reduction, hence no plain all_reduce.

`engine` is the object that provides PTBloomFilter; the default is this package (CUDA, no CPU fallback).
The CPU test-suite injects an oracle-backed stand-in so that the host logic here runs under gloo.
"""
import os

import numpy as np


def shard_range(total_rows, world_size, rank):
    """Contiguous row range [lo, hi) of `rank`; lo is a multiple of 64 (validity words, 128-bit key loads and
    probe tiles stay aligned) except that the last shard ends at total_rows."""
    if not 0 <= rank < world_size:
        raise ValueError("rank out of range")
    def cut(r):
        if r >= world_size:
            return total_rows
        return min(total_rows, (total_rows * r // world_size) & ~63)
    return cut(rank), cut(rank + 1)


class _DeviceWords:
    """Zero-copy view of the filter words for torch (CUDA array interface)."""

    def __init__(self, ptr, n):
        self.__cuda_array_interface__ = {"shape": (int(n),), "typestr": "<i8", "data": (int(ptr), False),
                                         "version": 2}


class ShardedBloomFilter:
    """One rank's replica of a filter that is built from row-range shards and probed shard-locally.
    Method names follow the reference's PTBloomFilter (src/include/bloom_filter.hpp:22-57)."""

    def __init__(self, group=None, device=0, engine=None, merge="auto"):
        import torch.distributed as dist
        self.dist = dist
        self.group = group
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        if engine is None:
            import importlib
            engine = importlib.import_module(__package__)
        self.engine = engine
        self.device = device
        self.bf = engine.PTBloomFilter(device)
        if merge == "auto":
            merge = "peer" if (self.world > 1 and dist.get_backend(group) == "nccl") else "allgather"
        if merge not in ("peer", "allgather"):
            raise ValueError("merge must be 'auto', 'peer' or 'allgather'")
        self.merge = merge
        self._peer_open_for = None
        self.merged = False
        self._release_pending = False  # peers may still be pulling my reduced slice (see _release)
        # The library owns the whole collective (rpt_bf_peer_allreduce: barrier words in device memory, stream memory
        # operations) when the engine has it; RPT_MERGE_LIB=0 keeps the phase-by-phase form with NCCL barriers below.
        self._lib_merge = hasattr(self.bf, "PeerAllreduce") and os.environ.get("RPT_MERGE_LIB", "1") != "0"
        self._flags_open = False

    # -- plumbing ------------------------------------------------------------------------------------
    def _scalar_device(self):
        import torch
        return torch.device("cuda", self.device) if self.dist.get_backend(self.group) == "nccl" else torch.device("cpu")

    def _allreduce_sum(self, value):
        if self.world == 1:
            return int(value)
        import torch
        t = torch.tensor([int(value)], dtype=torch.int64, device=self._scalar_device())
        self.dist.all_reduce(t, op=self.dist.ReduceOp.SUM, group=self.group)
        return int(t.item())

    def device_barrier(self):
        """Cross-rank barrier ordered on the current torch stream, without blocking the host: every rank's
        earlier stream work is complete before any rank's later stream work starts."""
        if self.world == 1:
            return
        import torch
        if not hasattr(self, "_token"):
            self._token = torch.zeros(1, dtype=torch.int32, device=self._scalar_device())
        self.dist.all_reduce(self._token, group=self.group)

    def _flag_barrier(self, flag):
        """device_barrier() that also sums a 0/1 flag over the ranks.  Returns the device tensor holding the sum:
        reading it (`.item()`) is the only thing that blocks the host, and only a rank whose own flag is 0 has to."""
        import torch
        if not hasattr(self, "_flag"):
            self._flag = torch.zeros(1, dtype=torch.int32, device=self._scalar_device())
        self._flag.fill_(int(flag))
        self.dist.all_reduce(self._flag, group=self.group)
        return self._flag

    def _release(self, stream=None):
        """The last barrier of a stream-ordered peer merge ("nobody still reads my words") is not needed by the
        probes that follow — they only read — so it is taken here, right before the next mutation of the words."""
        if not self._release_pending:
            return
        import torch
        self._release_pending = False
        if getattr(self, "_side_pending", False):  # an overlapped merge may still be pulling slices on its stream
            torch.cuda.current_stream().wait_stream(self._merge_stream)
            self._side_pending = False
        if getattr(self, "_lib_release", False):   # the merge was the library's: so is its last barrier
            self._lib_release = False
            self.bf.PeerRelease(torch.cuda.current_stream().cuda_stream)
        else:
            self.device_barrier()
        if stream is None:
            torch.cuda.current_stream().synchronize()

    # -- PTBloomFilter surface ------------------------------------------------------------------------
    def Initialize(self, est_num_rows):
        """est_num_rows is the GLOBAL estimate (the optimizer's cardinality), the same on every rank, so S is too."""
        self._release(None)
        self.bf.Initialize(est_num_rows)
        self._peer_open_for = None
        self.merged = False
        return self

    def Reset(self, est_num_rows, stream=None):
        self._release(stream)
        # a reallocation (even one that lands on the same address) makes every peer's IPC mapping of the old words
        # stale: compare the library's generation counter, not the pointer
        before = getattr(self.bf, "words_generation", None)
        self.bf.Reset(est_num_rows, stream)
        after = getattr(self.bf, "words_generation", None)
        if before != after or before is None:
            self._peer_open_for = None
        self.merged = False
        return self

    def Insert(self, cols, count, stream=None):
        """Inserts this rank's shard of the build column."""
        self._release(stream)
        self.bf.Insert(cols, count, stream)
        self.merged = False

    def Merge(self, stream=None):
        """OR-allreduce of the partial filters.  stream=None: blocking (returns when the merged words are in
        place).  Otherwise `stream` must be the raw handle of torch's current, non-default CUDA stream: barriers,
        kernels and copies are all enqueued on it and nothing blocks the host."""
        if self.world == 1:
            self.merged = True
            return
        import torch
        has_local = 0 if self.bf.IsEmpty() else 1
        on_gpu = self.dist.get_backend(self.group) == "nccl"
        blocking = stream is None
        self._release(stream)
        # "does any rank hold data": the peer merge sums the flag inside its first barrier (no extra collective, no
        # host sync for a rank that has rows itself); the all-gather merge asks with a scalar all-reduce
        self._any_data = None if (on_gpu and self.merge == "peer") else self._allreduce_sum(has_local)
        if on_gpu:
            if blocking:
                # own side stream: a NULL rpt_stream_t means "the handle's stream", which NCCL cannot order against
                self.bf.Sync()
                torch.cuda.current_stream().synchronize()
                if not hasattr(self, "_side"):
                    self._side = torch.cuda.Stream(device=torch.device("cuda", self.device))
                run_on = self._side
            else:
                run_on = torch.cuda.current_stream()
                if run_on.cuda_stream != stream:
                    raise ValueError("Merge(stream): stream must be torch's current (non-default) CUDA stream")
            with torch.cuda.stream(run_on):
                self._merge_on(run_on.cuda_stream, has_local, defer_release=not blocking)
            if blocking:
                run_on.synchronize()
        else:
            self._merge_on(None, has_local)
        if self._any_data:
            self.bf.SetHasData(True)
        self.merged = True

    def _merge_on(self, stream, has_local=0, defer_release=False):
        import torch
        if self.merge == "peer":
            words = (self.bf.device_words, getattr(self.bf, "words_generation", 0))
            if self._peer_open_for != words:
                blob = torch.frombuffer(bytearray(self.bf.IpcExport()), dtype=torch.uint8).cuda(self.device)
                allb = torch.empty(self.world * 64, dtype=torch.uint8, device=blob.device)
                self.dist.all_gather_into_tensor(allb, blob, group=self.group)
                raw = allb.cpu().numpy().tobytes()
                self.bf.PeerOpen([raw[64 * r:64 * (r + 1)] for r in range(self.world)], self.rank)
                self._peer_open_for = words
            S = self.bf.num_sectors
            slice_sectors = self.engine.slice_sectors() if hasattr(self.engine, "slice_sectors") else S
            nslices = -(-S // slice_sectors)
            # Overlapped variant (RPT_MERGE_OVERLAP=1; OFF by default — measured, it does not pay: the probe's partition
            # kernel is DRAM-bound and fills every SM, so the merge of the later slices only moves into the probe:
            # N=2 8.34-8.37 ms per step overlapped vs 8.36 ms plain, N=8 9.11 vs 8.50 ms;
            # profiles/merge_overlap_r2.md).
            overlap = defer_release and nslices > 1 and hasattr(self.bf, "PeerMergeRange") and \
                os.environ.get("RPT_MERGE_OVERLAP", "0") == "1"
            if self._lib_merge and not overlap and self._any_data is None:
                if not self._flags_open:  # once per handle: the barrier words do not move when the filter is re-sized
                    blob = torch.frombuffer(bytearray(self.bf.PeerFlagsExport()), dtype=torch.uint8).cuda(self.device)
                    allb = torch.empty(self.world * 64, dtype=torch.uint8, device=blob.device)
                    self.dist.all_gather_into_tensor(allb, blob, group=self.group)
                    raw = allb.cpu().numpy().tobytes()
                    self.bf.PeerFlagsOpen([raw[64 * r:64 * (r + 1)] for r in range(self.world)], self.rank)
                    self._flags_open = True
                run_stream = stream if stream is not None else torch.cuda.current_stream().cuda_stream
                self.bf.PeerAllreduce(has_local, run_stream)
                # only a rank without rows of its own has to wait to learn whether anybody had some
                self._any_data = True if has_local else self.bf.PeerAnyData(run_stream)
                if defer_release:
                    self._release_pending = True
                    self._lib_release = True
                else:
                    self.bf.PeerRelease(run_stream)
                return
            if overlap:
                # Slice by slice on a side stream; the probe that follows on the main stream waits per slice
                # (rpt_bf_set_slice_events): its partition kernel needs slice 0 only, so the merge of slices 1.. is
                # enqueued underneath it instead of in front of it.
                main = torch.cuda.current_stream()
                if not hasattr(self, "_merge_stream"):
                    self._merge_stream = torch.cuda.Stream(device=torch.device("cuda", self.device))
                    self._slice_events = [torch.cuda.Event() for _ in range(8)]
                side = self._merge_stream
                side.wait_stream(main)            # my partial filter is complete
                with torch.cuda.stream(side):
                    if self._any_data is None:    # every rank's partial filter is complete (+ the has-data flag)
                        total = self._flag_barrier(has_local)
                    else:
                        total = None
                        self.device_barrier()
                    for sl in range(nslices):
                        lo = sl * slice_sectors
                        ln = min(slice_sectors, S - lo)
                        self.bf.PeerMergeRange(0, lo, ln, side.cuda_stream)
                        self.device_barrier()     # slice sl is reduced on every rank
                        self.bf.PeerMergeRange(1, lo, ln, side.cuda_stream)
                        self._slice_events[sl].record(side)
                self.bf.SetSliceEvents([self._slice_events[sl].cuda_event for sl in range(nslices)])
                self._side_pending = True         # the next mutation of the words waits for the side stream
                self._release_pending = True
                if total is not None:
                    self._any_data = True if has_local else int(total.item()) > 0
                return
            if self._any_data is None:            # every partial filter is complete (+ the has-data flag)
                total = self._flag_barrier(has_local)
                self._any_data = True if has_local else int(total.item()) > 0
            else:
                self.device_barrier()
            self.bf.PeerMergePhase(0, stream)     # my slice |= peers' copies of my slice (NVLink reads)
            self.device_barrier()                 # every slice is reduced
            self.bf.PeerMergePhase(1, stream)     # pull the other ranks' reduced slices
            if defer_release:                     # nobody still reads my words: due before my next mutation
                self._release_pending = True
            else:
                self.device_barrier()
        else:
            mine = self._words_tensor()
            parts = torch.empty(self.world * mine.numel(), dtype=mine.dtype, device=mine.device)
            self.dist.all_gather_into_tensor(parts, mine, group=self.group)
            self.bf.OrParts(parts, self.world, stream)
            self._parts = parts  # keep alive until the OR kernel has run

    def _words_tensor(self):
        """The S words as a tensor the collective backend can send (device view for NCCL; engines that keep
        their words on the host expose `words_tensor`)."""
        if hasattr(self.bf, "words_tensor"):
            return self.bf.words_tensor()
        import torch
        return torch.as_tensor(_DeviceWords(self.bf.device_words, self.bf.num_sectors),
                               device=torch.device("cuda", self.device))

    def Finalize(self, local_rows, cols=None, count=0, stream=None):
        """Resize rule on the GLOBAL row count, rebuild of every shard when it fires, merge, finalize.
        Returns True when the filter was rebuilt (rpt_bf_finalize's rebuilt_out)."""
        actual_rows = self._allreduce_sum(local_rows)
        rebuilt = self.bf.NeedsRebuild(actual_rows)
        if rebuilt:
            if cols is None:
                raise self.engine.RptError("the resize rule fires but no key columns were given to rehash from")
            # ReinitializeAndRehash(actual_rows) on every rank: same new S everywhere
            self._release(stream)
            self.bf.ReinitializeAndRehash(actual_rows, cols, count, stream)
            self._peer_open_for = None
            self.merged = False
        if not self.merged:
            self.Merge(stream)
        self.bf.Finalize(actual_rows)  # rule cannot fire any more; sets finalized_
        return rebuilt

    def LookupSel(self, cols, count):
        """Shard-local probe: row ids relative to the shard."""
        return self.bf.LookupSel(cols, count)

    def GatherSelection(self, local_sel, row_base):
        """Concatenates the shards' selection vectors in rank order with their base offsets → the global
        ascending selection (as uint64 row ids; the global row count may exceed sel_t's 2^32)."""
        import torch
        mine = torch.from_numpy(np.asarray(local_sel, dtype=np.int64) + int(row_base))
        if self.world == 1:
            return mine.numpy().astype(np.uint64)
        dev = self._scalar_device()
        counts = torch.zeros(self.world, dtype=torch.int64, device=dev)
        counts[self.rank] = mine.numel()
        self.dist.all_reduce(counts, group=self.group)
        counts = counts.cpu().tolist()
        cap = max(max(counts), 1)
        padded = torch.zeros(cap, dtype=torch.int64, device=dev)
        padded[:mine.numel()] = mine.to(dev)
        allp = torch.empty(self.world * cap, dtype=torch.int64, device=dev)
        self.dist.all_gather_into_tensor(allp, padded, group=self.group)
        allp = allp.cpu().numpy()
        return np.concatenate([allp[r * cap:r * cap + counts[r]] for r in range(self.world)]).astype(np.uint64)

    # passthroughs
    def __getattr__(self, name):
        if name in ("num_sectors", "finalized_", "launch_count", "device_words", "IsEmpty", "SizedForRows",
                    "ExportWords", "LookupSelDevice", "FilterSel", "LookupHashes", "Sync", "NeedsRebuild"):
            return getattr(self.bf, name)
        raise AttributeError(name)

    def close(self):
        if self._release_pending:  # peers may still be pulling my reduced slice: wait for them before freeing it
            try:
                self._release(None)
            except Exception:
                self._release_pending = False
        if self._peer_open_for is not None and hasattr(self.bf, "PeerClose"):
            try:
                self.bf.PeerClose()
            except Exception:
                pass
        self.bf.close()
