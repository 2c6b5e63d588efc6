"""Shared test helpers: column builders that work for the oracle, the reference harness and the GPU library."""
import json
import os

import numpy as np

import oracle_lib

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")

NP = {1: np.int8, 2: np.uint8, 3: np.int8, 4: np.uint16, 5: np.int16, 6: np.uint32, 7: np.int32, 8: np.uint64,
      9: np.int64, 11: np.float32, 12: np.float64}


def load_json(name):
    with open(os.path.join(GOLDEN, name)) as f:
        return json.load(f)


class RawCol:
    """A host key column: same memory layout as rpt_col_t / rpto_col_t (usable by all three libraries)."""

    def __init__(self, type_id, data, aux=None, sel=None, validity=None, data_len=0):
        self.keep = (data, aux, sel, validity)
        self.raw = oracle_lib.Col(data.ctypes.data if data is not None and data.size else None,
                                  None if aux is None or aux.size == 0 else aux.ctypes.data,
                                  None if sel is None else sel.ctypes.data,
                                  None if validity is None else validity.ctypes.data, data_len,
                                  0 if aux is None else aux.size, type_id, 0)


def from_hex(h, dtype):
    return np.frombuffer(bytes.fromhex(h), dtype=dtype).copy()


def hx(v):
    return "0x%016x" % (int(v) & 0xFFFFFFFFFFFFFFFF)


def fnv_words(words):
    """FNV-1a-64 per word, vectorised is impossible (sequential); used on small arrays only."""
    h = 0xcbf29ce484222325
    for w in np.asarray(words, dtype=np.uint64).tolist():
        h = ((h ^ w) * 0x100000001b3) & 0xFFFFFFFFFFFFFFFF
    return h


def splitmix(x):
    with np.errstate(over="ignore"):
        x = x + np.uint64(0x9e3779b97f4a7c15)
        z = (x ^ (x >> np.uint64(30))) * np.uint64(0xbf58476d1ce4e5b9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94d049bb133111eb)
        return z ^ (z >> np.uint64(31))


def synth_build_keys(first, count):
    with np.errstate(over="ignore"):
        idx = np.arange(first, first + count, dtype=np.uint64)
        return (splitmix(np.uint64(0xB10F0001) + idx) & np.uint64(0x7fffffffffffffff)).astype(np.uint64)


def synth_probe_keys(first, count, n_build, sel_ppm):
    with np.errstate(over="ignore"):
        j = np.arange(first, first + count, dtype=np.uint64)
        a = splitmix(np.uint64(0x9B0BE001) + np.uint64(2) * j)
        b = splitmix(np.uint64(0x9B0BE001) + np.uint64(2) * j + np.uint64(1))
        member = (a % np.uint64(1000000)) < np.uint64(sel_ppm)
        return np.where(member,
                        splitmix(np.uint64(0xB10F0001) + (b % np.uint64(n_build))) & np.uint64(0x7fffffffffffffff),
                        b | np.uint64(0x8000000000000000)).astype(np.uint64)
