"""ctypes bindings to the CHECKER libraries (oracle/): test infrastructure only.

oracle()  → oracle/librpt_oracle.so   plain-C restatement (built on demand with gcc)
ref()     → oracle/_ref/librpt_ref.so the reference's own sources (prebuilt where /root/reference exists)
acero()   → oracle/librpt_oracle_acero.so  restatement of the README-described Arrow-style filter (parity unpinned)
"""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
ORACLE_LIB = os.path.join(ORACLE_DIR, "librpt_oracle.so")
REF_LIB = os.path.join(ORACLE_DIR, "_ref", "librpt_ref.so")

VARCHAR = 200
TYPE_WIDTH = {1: 1, 2: 1, 3: 1, 4: 2, 5: 2, 6: 4, 7: 4, 8: 8, 9: 8, 11: 4, 12: 8, 21: 16, 203: 16, 204: 16}


class Col(C.Structure):  # rpto_col_t == rpt_col_t layout
    _fields_ = [("data", C.c_void_p), ("aux", C.c_void_p), ("sel", C.c_void_p), ("validity", C.c_void_p),
                ("data_len", C.c_uint64), ("aux_bytes", C.c_uint64), ("type", C.c_int32), ("mem", C.c_int32)]


def build_oracle():
    src = os.path.join(ORACLE_DIR, "rpt_oracle.c")
    if (not os.path.exists(ORACLE_LIB)) or os.path.getmtime(src) > os.path.getmtime(ORACLE_LIB):
        subprocess.run(["make", "-C", ORACLE_DIR, "librpt_oracle.so"], check=True, capture_output=True)
    return ORACLE_LIB


def build_ref():
    """Only possible where /root/reference is mounted; elsewhere the prebuilt file is used as is."""
    if os.path.isdir("/root/reference/duckdb/src"):
        src = os.path.join(ORACLE_DIR, "ref_harness.cpp")
        if (not os.path.exists(REF_LIB)) or os.path.getmtime(src) > os.path.getmtime(REF_LIB):
            subprocess.run(["make", "-C", ORACLE_DIR, "_ref/librpt_ref.so"], check=True, capture_output=True)
    return REF_LIB if os.path.exists(REF_LIB) else None


_u64, _vp, _int = C.c_uint64, C.c_void_p, C.c_int
_oracle = None
_ref = None


def _p(a):
    return None if a is None else a.ctypes.data


class Oracle:
    def __init__(self, path):
        L = self.L = C.CDLL(path)
        sig = {
            "rpto_murmur64": (_u64, [_u64]), "rpto_hash_bytes": (_u64, [C.c_char_p, _u64]),
            "rpto_hash_value": (_u64, [_int, _vp]), "rpto_combine": (_u64, [_u64, _u64]),
            "rpto_next_pow2": (_u64, [_u64]), "rpto_num_sectors": (_u64, [_u64]), "rpto_mask": (_u64, [_u64]),
            "rpto_needs_rebuild": (_int, [_u64, _u64]),
            "rpto_hash_column": (None, [C.POINTER(Col), _u64, _vp, _int]),
            "rpto_hash_columns": (None, [C.POINTER(Col), _int, _u64, _vp]),
            "rpto_insert_hashes": (None, [_vp, _u64, _vp, _u64]),
            "rpto_lookup_hashes": (_u64, [_vp, _u64, _vp, _u64, _vp]),
            "rpto_insert": (None, [_vp, _u64, C.POINTER(Col), _int, _u64]),
            "rpto_probe": (_u64, [_vp, _u64, C.POINTER(Col), _int, _u64, _vp]),
            "rpto_check_statistics": (_int, [_vp, _u64, _int, _vp, _vp, _int]),
            "rpto_fnv1a_words": (_u64, [_vp, _u64]), "rpto_splitmix64": (_u64, [_u64]),
            "rpto_synth_build_keys": (None, [_vp, _u64, _u64]),
            "rpto_synth_probe_keys": (None, [_vp, _u64, _u64, _u64, C.c_uint32]),
            "rpto_bench_build_i64": (C.c_double, [_vp, _u64, _vp, _u64, _int]),
            "rpto_bench_probe_i64": (C.c_double, [_vp, _u64, _vp, _u64, _int, C.POINTER(_u64)]),
        }
        for name, (res, args) in sig.items():
            fn = getattr(L, name)
            fn.restype, fn.argtypes = res, args

    # -- scalars ------------------------------------------------------------------------------
    def hash_value(self, type_id, arr):
        return self.L.rpto_hash_value(type_id, arr.ctypes.data)

    def hash_bytes(self, b):
        return self.L.rpto_hash_bytes(b, len(b))

    def mask(self, h):
        return self.L.rpto_mask(h)

    def combine(self, a, b):
        return self.L.rpto_combine(a, b)

    def num_sectors(self, rows):
        return self.L.rpto_num_sectors(rows & 0xFFFFFFFFFFFFFFFF)

    def needs_rebuild(self, sized_for, actual):
        return bool(self.L.rpto_needs_rebuild(sized_for, actual))

    # -- columns: cols are rpt_b200.KeyCol (same struct layout) or (Col, keep) ----------------
    @staticmethod
    def _cols(cols):
        if not isinstance(cols, (list, tuple)):
            cols = [cols]
        arr = (Col * len(cols))()
        for i, c in enumerate(cols):
            C.memmove(C.byref(arr[i]), C.byref(c.raw), C.sizeof(Col))
        return arr, len(cols)

    def hash_columns(self, cols, count):
        arr, n = self._cols(cols)
        out = np.empty(count, dtype=np.uint64)
        self.L.rpto_hash_columns(arr, n, count, out.ctypes.data)
        return out

    def new_filter(self, rows):
        return np.zeros(self.num_sectors(rows), dtype=np.uint64)

    def insert(self, words, cols, count):
        arr, n = self._cols(cols)
        self.L.rpto_insert(words.ctypes.data, words.shape[0], arr, n, count)

    def probe(self, words, cols, count):
        arr, n = self._cols(cols)
        sel = np.empty(count + 1, dtype=np.uint32)
        f = self.L.rpto_probe(words.ctypes.data, words.shape[0], arr, n, count, sel.ctypes.data)
        return sel[:f].copy()

    def insert_hashes(self, words, hashes):
        self.L.rpto_insert_hashes(words.ctypes.data, words.shape[0], hashes.ctypes.data, hashes.shape[0])

    def lookup_hashes(self, words, hashes):
        sel = np.empty(hashes.shape[0] + 1, dtype=np.uint32)
        f = self.L.rpto_lookup_hashes(words.ctypes.data, words.shape[0], hashes.ctypes.data, hashes.shape[0],
                                      sel.ctypes.data)
        return sel[:f].copy()

    def check_statistics(self, words, type_id, mins, maxs, has_minmax=None):
        """BFTableFilter::CheckStatistics per row group (duckdb bloom_filter.cpp:138-196): uint8 codes."""
        mins, maxs = np.ascontiguousarray(mins), np.ascontiguousarray(maxs)
        count = mins.shape[0]
        step = mins.nbytes // count if count else 0
        out = np.zeros(count, dtype=np.uint8)
        for i in range(count):
            out[i] = self.L.rpto_check_statistics(words.ctypes.data, words.shape[0], type_id,
                                                  mins.ctypes.data + i * step, maxs.ctypes.data + i * step,
                                                  1 if has_minmax is None else int(has_minmax[i]))
        return out

    def fnv(self, words):
        return self.L.rpto_fnv1a_words(words.ctypes.data, words.shape[0])

    def synth_build_keys(self, first, count):
        out = np.empty(count, dtype=np.uint64)
        self.L.rpto_synth_build_keys(out.ctypes.data, first, count)
        return out

    def synth_probe_keys(self, first, count, n_build, sel_ppm):
        out = np.empty(count, dtype=np.uint64)
        self.L.rpto_synth_probe_keys(out.ctypes.data, first, count, n_build, sel_ppm)
        return out

    def bench_build(self, words, keys, threads):
        return self.L.rpto_bench_build_i64(words.ctypes.data, words.shape[0], keys.ctypes.data, keys.shape[0], threads)

    def bench_probe(self, words, keys, threads):
        found = _u64(0)
        t = self.L.rpto_bench_probe_i64(words.ctypes.data, words.shape[0], keys.ctypes.data, keys.shape[0], threads,
                                        C.byref(found))
        return t, found.value


class Ref:
    """The reference's own code (oracle/ref_harness.cpp)."""

    def __init__(self, path):
        L = self.L = C.CDLL(path)
        sig = {
            "ref_hash_value": (_u64, [_int, _vp]), "ref_hash_string": (_u64, [C.c_char_p, C.c_uint32]),
            "ref_hash_bytes": (_u64, [C.c_char_p, _u64]), "ref_null_hash": (_u64, []),
            "ref_combine": (_u64, [_u64, _u64]), "ref_next_pow2": (_u64, [_u64]), "ref_mask": (_u64, [_u64]),
            "ref_vector_hash": (_int, [_int, _vp, _vp, _vp, _vp, _u64, _u64, _vp, _int]),
            "ref_insert_hashes": (None, [_vp, _u64, _vp, _u64]),
            "ref_lookup_hashes": (_u64, [_vp, _u64, _vp, _u64, _vp]),
            "ref_lookup_one": (_int, [_vp, _u64, _u64]),
            "ref_check_statistics": (_int, [_vp, _u64, _int, _vp, _vp, _int]),
            "ref_bench_build_i64": (C.c_double, [_vp, _u64, _vp, _u64, _int]),
            "ref_bench_probe_i64": (C.c_double, [_vp, _u64, _vp, _u64, _int, C.POINTER(_u64)]),
        }
        for name, (res, args) in sig.items():
            fn = getattr(L, name)
            fn.restype, fn.argtypes = res, args

    def hash_value(self, type_id, arr):
        return self.L.ref_hash_value(type_id, arr.ctypes.data)

    def check_statistics(self, words, type_id, mins, maxs, has_minmax=None):
        """The reference's own TemplatedCheckStatistics<T> (duckdb bloom_filter.cpp:138-196), one row group at a time."""
        mins, maxs = np.ascontiguousarray(mins), np.ascontiguousarray(maxs)
        count = mins.shape[0]
        step = mins.nbytes // count if count else 0
        out = np.zeros(count, dtype=np.uint8)
        for i in range(count):
            out[i] = self.L.ref_check_statistics(words.ctypes.data, words.shape[0], type_id,
                                                 mins.ctypes.data + i * step, maxs.ctypes.data + i * step,
                                                 1 if has_minmax is None else int(has_minmax[i]))
        return out

    def hash_string(self, b):
        return self.L.ref_hash_string(b, len(b))

    def hash_bytes(self, b):
        return self.L.ref_hash_bytes(b, len(b))

    def mask(self, h):
        return self.L.ref_mask(h)

    def combine(self, a, b):
        return self.L.ref_combine(a, b)

    def null_hash(self):
        return self.L.ref_null_hash()

    def next_pow2(self, v):
        return self.L.ref_next_pow2(v)

    def hash_columns(self, cols, count):
        """HashColumns = Hash(col0) + CombineHash(col j) through the reference's TightLoop* templates."""
        if not isinstance(cols, (list, tuple)):
            cols = [cols]
        out = np.zeros(count, dtype=np.uint64)
        for j, c in enumerate(cols):
            r = c.raw
            rc = self.L.ref_vector_hash(r.type, r.data, r.aux, r.sel, r.validity, r.data_len or count, count,
                                        out.ctypes.data, 1 if j else 0)
            assert rc == 0
        return out

    def insert_hashes(self, words, hashes):
        self.L.ref_insert_hashes(words.ctypes.data, words.shape[0], hashes.ctypes.data, hashes.shape[0])

    def lookup_hashes(self, words, hashes):
        sel = np.empty(hashes.shape[0] + 1, dtype=np.uint32)
        f = self.L.ref_lookup_hashes(words.ctypes.data, words.shape[0], hashes.ctypes.data, hashes.shape[0],
                                     sel.ctypes.data)
        return sel[:f].copy()

    def bench_build(self, words, keys, threads):
        return self.L.ref_bench_build_i64(words.ctypes.data, words.shape[0], keys.ctypes.data, keys.shape[0], threads)

    def bench_probe(self, words, keys, threads):
        found = _u64(0)
        t = self.L.ref_bench_probe_i64(words.ctypes.data, words.shape[0], keys.ctypes.data, keys.shape[0], threads,
                                       C.byref(found))
        return t, found.value


def oracle():
    global _oracle
    if _oracle is None:
        _oracle = Oracle(build_oracle())
    return _oracle


def ref():
    global _ref
    if _ref is None:
        path = build_ref()
        if path is None:
            return None
        _ref = Ref(path)
    return _ref


ACERO_LIB = os.path.join(ORACLE_DIR, "librpt_oracle_acero.so")
_acero = None


class AceroOracle:
    """oracle/acero_oracle.cpp — checker of the rpt_abf_* variant; works on 64-bit hashes (from Oracle.hash_columns)."""

    def __init__(self, path):
        L = self.L = C.CDLL(path)
        sig = {
            "acero_mask_bits": (None, [_vp]), "acero_mask": (_u64, [_u64]), "acero_log_blocks": (_int, [_u64]),
            "acero_insert": (None, [_vp, _int, _vp, _u64]), "acero_find": (_u64, [_vp, _int, _vp, _u64, _vp]),
            "acero_fold": (_int, [_vp, C.POINTER(_int)]),
        }
        for name, (res, args) in sig.items():
            fn = getattr(L, name)
            fn.restype, fn.argtypes = res, args

    def mask_bits(self):
        out = np.zeros(136, dtype=np.uint8)
        self.L.acero_mask_bits(out.ctypes.data)
        return out

    def mask(self, h):
        return self.L.acero_mask(h)

    def log_blocks(self, rows):
        return self.L.acero_log_blocks(rows)

    def new_filter(self, rows):
        return np.zeros(1 << self.log_blocks(rows), dtype=np.uint64)

    @staticmethod
    def _log(blocks):
        return int(blocks.shape[0]).bit_length() - 1

    def insert(self, blocks, hashes):
        hashes = np.ascontiguousarray(hashes, dtype=np.uint64)
        self.L.acero_insert(blocks.ctypes.data, self._log(blocks), hashes.ctypes.data, hashes.shape[0])

    def find(self, blocks, hashes):
        hashes = np.ascontiguousarray(hashes, dtype=np.uint64)
        sel = np.empty(hashes.shape[0] + 1, dtype=np.uint32)
        f = self.L.acero_find(blocks.ctypes.data, self._log(blocks), hashes.ctypes.data, hashes.shape[0],
                              sel.ctypes.data)
        return sel[:f].copy()

    def fold(self, blocks):
        """Returns (folded blocks, number of halvings); the input array is modified like the in-place original."""
        log = _int(self._log(blocks))
        k = self.L.acero_fold(blocks.ctypes.data, C.byref(log))
        return blocks[: 1 << log.value].copy(), k


def acero():
    global _acero
    if _acero is None:
        src = os.path.join(ORACLE_DIR, "acero_oracle.cpp")
        if (not os.path.exists(ACERO_LIB)) or os.path.getmtime(src) > os.path.getmtime(ACERO_LIB):
            subprocess.run(["make", "-C", ORACLE_DIR, "librpt_oracle_acero.so"], check=True, capture_output=True)
        _acero = AceroOracle(ACERO_LIB)
    return _acero
