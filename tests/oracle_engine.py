"""TEST INFRASTRUCTURE: an oracle-backed stand-in for the CUDA engine's PTBloomFilter, used ONLY to run the
host-side sharding logic (rpt_b200.ShardedBloomFilter) under gloo on CPU.  The product never imports this."""
import numpy as np
import torch

import oracle_lib


class RptError(RuntimeError):
    pass


class PTBloomFilter:
    def __init__(self, device=0):
        self.o = oracle_lib.oracle()
        self.words = None

    def Initialize(self, est):
        est = int(est) & 0xFFFFFFFF
        self.words = self.o.new_filter(est)
        self.sized, self.has, self.fin = est, False, False
        return self

    def Reset(self, est, stream=None):
        return self.Initialize(est)

    def close(self):
        self.words = None

    @property
    def num_sectors(self):
        return self.words.shape[0]

    @property
    def finalized_(self):
        return self.fin

    def SizedForRows(self):
        return self.sized

    def IsEmpty(self):
        return not self.has

    def SetHasData(self, flag):
        self.has = bool(flag)

    def Sync(self):
        pass

    def Insert(self, cols, count, stream=None):
        if count == 0:
            return
        self.has = True
        self.o.insert(self.words, cols, count)

    def LookupSel(self, cols, count):
        return self.o.probe(self.words, cols, count)

    def NeedsRebuild(self, actual):
        return self.o.needs_rebuild(self.sized, actual)

    def ReinitializeAndRehash(self, actual_rows, cols, count, stream=None):
        self.Initialize(actual_rows)
        self.has = count > 0
        if count:
            self.o.insert(self.words, cols, count)

    def Finalize(self, actual, cols=None, count=0, stream=None):
        rebuilt = self.NeedsRebuild(actual)
        if rebuilt:
            if cols is None:
                raise RptError("resize rule fires but no key columns were given")
            self.ReinitializeAndRehash(actual, cols, count)
        self.fin = True
        return rebuilt

    def ExportWords(self):
        return self.words.copy()

    # collective plumbing: words as a CPU tensor; OR of gathered parts
    def words_tensor(self):
        return torch.from_numpy(self.words.view(np.int64))

    def OrParts(self, parts, nparts, stream=None):
        p = parts.numpy().view(np.uint64).reshape(nparts, -1)
        self.words |= np.bitwise_or.reduce(p, axis=0)
