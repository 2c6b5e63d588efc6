"""rpt_b200 — Python driver over librpt_gpu.so (the C ABI in include/rpt_gpu.h).

Python is plumbing here (tests, bench, torch.distributed); the product is the shared library.
The class below mirrors the reference's ``PTBloomFilter`` (reference src/include/bloom_filter.hpp:22-57)
method for method so that the parity tests read like tests of the reference wrapper.

There is no CPU fallback: importing works without a GPU (so symbols can be checked), but every
compute call goes to the CUDA library and raises ``RptError`` when it fails.
"""
import ctypes as C
import os

import numpy as np

from .build import LIB, build  # noqa: F401

# duckdb::PhysicalType values (include/rpt_gpu.h enum rpt_type)
BOOL, UINT8, INT8, UINT16, INT16, UINT32, INT32, UINT64, INT64 = 1, 2, 3, 4, 5, 6, 7, 8, 9
FLOAT, DOUBLE, INTERVAL, VARCHAR, UINT128, INT128, STRING_T = 11, 12, 21, 200, 203, 204, 1200
MEM_HOST, MEM_DEVICE = 0, 1
MAX_KEY_COLS = 8

NUMPY_TYPES = {
    np.dtype(np.bool_): BOOL, np.dtype(np.uint8): UINT8, np.dtype(np.int8): INT8,
    np.dtype(np.uint16): UINT16, np.dtype(np.int16): INT16, np.dtype(np.uint32): UINT32,
    np.dtype(np.int32): INT32, np.dtype(np.uint64): UINT64, np.dtype(np.int64): INT64,
    np.dtype(np.float32): FLOAT, np.dtype(np.float64): DOUBLE,
}
_NUMPY_OF_TYPE = {v: k for k, v in NUMPY_TYPES.items()}
TYPE_WIDTH = {BOOL: 1, UINT8: 1, INT8: 1, UINT16: 2, INT16: 2, UINT32: 4, INT32: 4, UINT64: 8, INT64: 8,
              FLOAT: 4, DOUBLE: 8, INTERVAL: 16, UINT128: 16, INT128: 16}


class RptError(RuntimeError):
    pass


class RptCol(C.Structure):
    """struct rpt_col (include/rpt_gpu.h)."""
    _fields_ = [
        ("data", C.c_void_p), ("aux", C.c_void_p), ("sel", C.c_void_p), ("validity", C.c_void_p),
        ("data_len", C.c_uint64), ("aux_bytes", C.c_uint64), ("type", C.c_int32), ("mem", C.c_int32),
    ]


_lib = None

_SIGNATURES = {
    "rpt_gpu_last_error": (C.c_char_p, []),
    "rpt_gpu_version": (C.c_char_p, []),
    "rpt_gpu_device_count": (C.c_int, [C.POINTER(C.c_int)]),
    "rpt_host_alloc": (C.c_int, [C.POINTER(C.c_void_p), C.c_uint64]),
    "rpt_host_free": (C.c_int, [C.c_void_p]),
    "rpt_bf_create": (C.c_int, [C.POINTER(C.c_void_p), C.c_uint64, C.c_int]),
    "rpt_bf_destroy": (C.c_int, [C.c_void_p]),
    "rpt_bf_reset": (C.c_int, [C.c_void_p, C.c_uint64, C.c_void_p]),
    "rpt_bf_num_sectors": (C.c_int, [C.c_void_p, C.POINTER(C.c_uint64)]),
    "rpt_bf_sized_for_rows": (C.c_int, [C.c_void_p, C.POINTER(C.c_uint64)]),
    "rpt_bf_is_empty": (C.c_int, [C.c_void_p, C.POINTER(C.c_int)]),
    "rpt_bf_is_finalized": (C.c_int, [C.c_void_p, C.POINTER(C.c_int)]),
    "rpt_bf_device": (C.c_int, [C.c_void_p, C.POINTER(C.c_int)]),
    "rpt_bf_sectors_for_rows": (C.c_int, [C.c_uint64, C.POINTER(C.c_uint64)]),
    "rpt_bf_insert": (C.c_int, [C.c_void_p, C.POINTER(RptCol), C.c_int, C.c_uint64, C.c_void_p]),
    "rpt_bf_needs_rebuild": (C.c_int, [C.c_void_p, C.c_uint64, C.POINTER(C.c_int)]),
    "rpt_bf_reinitialize_and_rehash": (C.c_int, [C.c_void_p, C.c_uint64, C.POINTER(RptCol), C.c_int, C.c_uint64,
                                                 C.c_void_p]),
    "rpt_bf_finalize": (C.c_int, [C.c_void_p, C.c_uint64, C.POINTER(RptCol), C.c_int, C.c_uint64,
                                  C.POINTER(C.c_int), C.c_void_p]),
    "rpt_bf_probe": (C.c_int, [C.c_void_p, C.POINTER(RptCol), C.c_int, C.c_uint64, C.c_void_p, C.c_void_p, C.c_int,
                               C.c_void_p]),
    "rpt_bf_probe_u64_noalloc": (C.c_int, [C.c_void_p, C.c_uint64, C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p,
                                           C.c_int, C.c_void_p]),
    "rpt_bf_filter_sel": (C.c_int, [C.c_void_p, C.POINTER(RptCol), C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p,
                                    C.c_int, C.c_void_p]),
    "rpt_bf_lookup_hashes_host": (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p]),
    "rpt_bf_check_statistics": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64,
                                          C.c_void_p]),
    "rpt_bf_export_words": (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint64]),
    "rpt_bf_import_words": (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint64]),
    "rpt_bf_device_words": (C.c_int, [C.c_void_p, C.POINTER(C.c_void_p)]),
    "rpt_bf_sync": (C.c_int, [C.c_void_p]),
    "rpt_bf_set_has_data": (C.c_int, [C.c_void_p, C.c_int]),
    "rpt_bf_launch_count": (C.c_int, [C.c_void_p, C.POINTER(C.c_longlong)]),
    "rpt_bf_minmax_begin": (C.c_int, [C.c_void_p, C.c_int]),
    "rpt_bf_minmax_end": (C.c_int, [C.c_void_p]),
    "rpt_bf_minmax_get": (C.c_int, [C.c_void_p, C.POINTER(C.c_int), C.POINTER(C.c_int), C.c_void_p, C.c_void_p]),
    "rpt_bf_or_parts": (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p]),
    "rpt_bf_ipc_export": (C.c_int, [C.c_void_p, C.c_void_p]),
    "rpt_bf_peer_open": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int]),
    "rpt_bf_peer_merge_phase": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p]),
    "rpt_bf_peer_close": (C.c_int, [C.c_void_p]),
    "rpt_bf_peer_flags_export": (C.c_int, [C.c_void_p, C.c_void_p]),
    "rpt_bf_peer_flags_open": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int]),
    "rpt_bf_peer_allreduce": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p]),
    "rpt_bf_peer_release": (C.c_int, [C.c_void_p, C.c_void_p]),
    "rpt_bf_peer_any_data": (C.c_int, [C.c_void_p, C.c_void_p, C.POINTER(C.c_int)]),
    "rpt_bf_allreduce_or": (C.c_int, [C.POINTER(C.c_void_p), C.c_int]),
    "rpt_abf_log_blocks_for_rows": (C.c_int, [C.c_uint64, C.POINTER(C.c_int)]),
    "rpt_abf_create": (C.c_int, [C.POINTER(C.c_void_p), C.c_uint64, C.c_int]),
    "rpt_abf_destroy": (C.c_int, [C.c_void_p]),
    "rpt_abf_log_num_blocks": (C.c_int, [C.c_void_p, C.POINTER(C.c_int)]),
    "rpt_abf_mask_bits": (C.c_int, [C.c_void_p]),
    "rpt_abf_insert": (C.c_int, [C.c_void_p, C.POINTER(RptCol), C.c_int, C.c_uint64, C.c_void_p]),
    "rpt_abf_fold": (C.c_int, [C.c_void_p, C.c_void_p, C.POINTER(C.c_int)]),
    "rpt_abf_probe": (C.c_int, [C.c_void_p, C.POINTER(RptCol), C.c_int, C.c_uint64, C.c_void_p, C.c_void_p, C.c_int,
                                C.c_void_p]),
    "rpt_abf_export_blocks": (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint64]),
    "rpt_abf_or_parts": (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p]),
    "rpt_abf_device_blocks": (C.c_int, [C.c_void_p, C.POINTER(C.c_void_p)]),
    "rpt_abf_sync": (C.c_int, [C.c_void_p]),
    "rpt_abf_launch_count": (C.c_int, [C.c_void_p, C.POINTER(C.c_longlong)]),
    "rpt_bf_peer_merge_range": (C.c_int, [C.c_void_p, C.c_int, C.c_uint64, C.c_uint64, C.c_void_p]),
    "rpt_bf_slice_sectors": (C.c_int, [C.POINTER(C.c_uint64)]),
    "rpt_bf_set_slice_events": (C.c_int, [C.c_void_p, C.POINTER(C.c_void_p), C.c_int]),
    "rpt_bf_checksum": (C.c_int, [C.c_void_p, C.POINTER(C.c_uint64)]),
    "rpt_bf_words_generation": (C.c_int, [C.c_void_p, C.POINTER(C.c_uint64)]),
    "rpt_bf_set_prefilter": (C.c_int, [C.c_void_p, C.c_int]),
    "rpt_bf_prefilter_info": (C.c_int, [C.c_void_p, C.POINTER(C.c_int), C.POINTER(C.c_double)]),
    "rpt_synth_build_keys": (C.c_int, [C.c_void_p, C.c_uint64, C.c_uint64, C.c_int, C.c_void_p]),
    "rpt_synth_probe_keys": (C.c_int, [C.c_void_p, C.c_uint64, C.c_uint64, C.c_uint64, C.c_uint32, C.c_int,
                                       C.c_void_p]),
}


def lib():
    """Loads librpt_gpu.so (never builds implicitly; fails loudly when it is missing)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB):
            raise RptError(f"{LIB} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                           "(there is no CPU fallback)")
        handle = C.CDLL(LIB)
        for name, (res, args) in _SIGNATURES.items():
            fn = getattr(handle, name)
            fn.restype = res
            fn.argtypes = args
        _lib = handle
    return _lib


def _check(rc):
    if rc != 0:
        raise RptError(lib().rpt_gpu_last_error().decode("utf-8", "replace"))


def slice_sectors():
    out = C.c_uint64(0)
    _check(lib().rpt_bf_slice_sectors(C.byref(out)))
    return out.value


def device_count():
    n = C.c_int(0)
    _check(lib().rpt_gpu_device_count(C.byref(n)))
    return n.value


def sectors_for_rows(rows):
    out = C.c_uint64(0)
    _check(lib().rpt_bf_sectors_for_rows(C.c_uint64(rows & 0xFFFFFFFFFFFFFFFF), C.byref(out)))
    return out.value


class KeyCol:
    """One key column plus the Python objects that keep its buffers alive."""

    def __init__(self, raw, keep):
        self.raw = raw
        self.keep = keep


def _ptr(x):
    if x is None:
        return None
    if isinstance(x, np.ndarray):
        return x.ctypes.data
    if hasattr(x, "data_ptr"):  # torch tensor
        return x.data_ptr()
    return int(x)


def host_col(values, type=None, sel=None, validity=None, data_len=None):
    """Key column over numpy arrays.  ``values``: 1-D array of a fixed-width type (or an (n,2) uint64
    array with type=INT128/UINT128/INTERVAL raw 16-byte entries)."""
    values = np.ascontiguousarray(values)
    t = type if type is not None else NUMPY_TYPES[values.dtype]
    n = values.shape[0]
    raw = RptCol(_ptr(values), None, _ptr(sel), _ptr(validity), data_len if data_len is not None else n, 0, t,
                 MEM_HOST)
    return KeyCol(raw, (values, sel, validity))


def host_varchar_col(offsets, data, sel=None, validity=None):
    """Arrow-utf8 layout: uint32 offsets[n+1] + bytes."""
    offsets = np.ascontiguousarray(offsets, dtype=np.uint32)
    data = np.ascontiguousarray(data, dtype=np.uint8)
    raw = RptCol(_ptr(offsets), _ptr(data) if data.size else None, _ptr(sel), _ptr(validity), offsets.shape[0] - 1,
                 data.size, VARCHAR, MEM_HOST)
    return KeyCol(raw, (offsets, data, sel, validity))


def strings_to_arrow(strings):
    """list[bytes] → (uint32 offsets, uint8 bytes)."""
    lens = np.fromiter((len(s) for s in strings), dtype=np.uint64, count=len(strings))
    offsets = np.zeros(len(strings) + 1, dtype=np.uint32)
    np.cumsum(lens, out=offsets[1:], dtype=np.uint64) if len(strings) else None
    data = np.frombuffer(b"".join(strings), dtype=np.uint8).copy()
    return offsets, data


def device_col(tensor_or_ptr, type, sel=None, validity=None, data_len=0, aux=None, aux_bytes=0):
    """Key column over device memory (torch CUDA tensors or raw device pointers)."""
    raw = RptCol(_ptr(tensor_or_ptr), _ptr(aux), _ptr(sel), _ptr(validity), data_len, aux_bytes, type, MEM_DEVICE)
    return KeyCol(raw, (tensor_or_ptr, sel, validity, aux))


def _col_array(cols):
    if isinstance(cols, KeyCol):
        cols = [cols]
    arr = (RptCol * len(cols))(*[c.raw for c in cols])
    return arr, len(cols)


class PTBloomFilter:
    """Mirror of the reference's PTBloomFilter (src/include/bloom_filter.hpp:22-57) over the C ABI."""

    def __init__(self, device=0):
        self._h = C.c_void_p(None)
        self.device = device

    # -- Initialize(ClientContext&, uint32_t est_num_rows): src/bloom_filter.cpp:27-32 ---------
    def Initialize(self, est_num_rows):
        self.close()
        est = int(est_num_rows) & 0xFFFFFFFF  # the reference's parameter is uint32_t: estimates wrap
        _check(lib().rpt_bf_create(C.byref(self._h), est, self.device))
        return self

    def Reset(self, est_num_rows, stream=None):
        """Initialize again on the live handle (zero-fill is asynchronous when the size is unchanged)."""
        _check(lib().rpt_bf_reset(self._h, int(est_num_rows) & 0xFFFFFFFF, stream))
        return self

    def close(self):
        if self._h:
            lib().rpt_bf_destroy(self._h)
            self._h = C.c_void_p(None)

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _get(self, fn, ctype):
        out = ctype(0)
        _check(fn(self._h, C.byref(out)))
        return out.value

    @property
    def num_sectors(self):
        return self._get(lib().rpt_bf_num_sectors, C.c_uint64)

    def SizedForRows(self):
        return self._get(lib().rpt_bf_sized_for_rows, C.c_uint64)

    def IsEmpty(self):
        return bool(self._get(lib().rpt_bf_is_empty, C.c_int))

    @property
    def finalized_(self):
        return bool(self._get(lib().rpt_bf_is_finalized, C.c_int))

    @property
    def launch_count(self):
        return self._get(lib().rpt_bf_launch_count, C.c_longlong)

    @property
    def device_words(self):
        return self._get(lib().rpt_bf_device_words, C.c_void_p)

    # -- Insert(DataChunk&, cols): src/bloom_filter.cpp:70-78 ----------------------------------
    def Insert(self, cols, count, stream=None):
        arr, n = _col_array(cols)
        _check(lib().rpt_bf_insert(self._h, arr, n, count, stream))

    # -- LookupSel(DataChunk&, SelectionVector&, cols, buf): src/bloom_filter.cpp:60-68 --------
    def LookupSel(self, cols, count):
        """Returns the selection vector (uint32 numpy array of length `found`)."""
        arr, n = _col_array(cols)
        sel = np.empty(max(count, 1), dtype=np.uint32)
        found = C.c_uint64(0)
        _check(lib().rpt_bf_probe(self._h, arr, n, count, sel.ctypes.data, C.addressof(found), MEM_HOST, None))
        return sel[: found.value]

    def LookupSelDevice(self, cols, count, sel_out_ptr, found_out_ptr, stream=None):
        """Asynchronous device→device probe (sel_out: uint32[count], found_out: uint64[1] on the device)."""
        arr, n = _col_array(cols)
        _check(lib().rpt_bf_probe(self._h, arr, n, count, _ptr(sel_out_ptr), _ptr(found_out_ptr), MEM_DEVICE,
                                  stream))

    def LookupSelNoAlloc(self, keys_ptr, count, sel_out_ptr, found_out_ptr, stream=None):
        """rpt_bf_probe_u64_noalloc: the allocation-free probe of one non-NULL 64-bit device key column (no reference
        counterpart; count <= NOALLOC_MAX_ROWS; calls on one device share a static scratch and must not overlap)."""
        _check(lib().rpt_bf_probe_u64_noalloc(self.device_words, self.num_sectors, _ptr(keys_ptr), count,
                                              _ptr(sel_out_ptr), _ptr(found_out_ptr), self.device, stream))

    def FilterSel(self, key, in_sel, approved_count):
        """BFTableFilter::Filter (duckdb bloom_filter.cpp:94-131) on host buffers."""
        sel = np.empty(max(approved_count, 1), dtype=np.uint32)
        found = C.c_uint64(0)
        _check(lib().rpt_bf_filter_sel(self._h, C.byref(key.raw), _ptr(in_sel), approved_count, sel.ctypes.data,
                                       C.addressof(found), MEM_HOST, None))
        return sel[: found.value]

    def LookupHashes(self, hashes):
        hashes = np.ascontiguousarray(hashes, dtype=np.uint64)
        hits = np.zeros(hashes.shape[0], dtype=np.uint8)
        _check(lib().rpt_bf_lookup_hashes_host(self._h, hashes.ctypes.data, hashes.shape[0], hits.ctypes.data))
        return hits.astype(bool)

    def CheckStatistics(self, type_code, mins, maxs, has_minmax=None):
        """BFTableFilter::CheckStatistics (duckdb bloom_filter.cpp:138-196) over a batch of row groups.
        mins/maxs: arrays of the column's physical type (128-bit types: uint64 pairs {lower, upper}, shape (k, 2)).
        Returns uint8 codes: 0 NO_PRUNING_POSSIBLE, 1 FILTER_ALWAYS_TRUE, 2 FILTER_ALWAYS_FALSE."""
        mins = np.ascontiguousarray(mins)
        maxs = np.ascontiguousarray(maxs)
        count = mins.shape[0]
        if maxs.shape != mins.shape or mins.dtype != maxs.dtype:
            raise ValueError("mins and maxs must have the same shape and dtype")
        width = TYPE_WIDTH.get(type_code)
        if width is not None and mins.nbytes != count * width:
            raise ValueError("statistics arrays do not match the physical type's width")
        has = None
        if has_minmax is not None:
            has = np.ascontiguousarray(has_minmax, dtype=np.uint8)
            if has.shape[0] != count:
                raise ValueError("has_minmax must have one entry per row group")
        out = np.zeros(count, dtype=np.uint8)
        _check(lib().rpt_bf_check_statistics(self._h, int(type_code), mins.ctypes.data, maxs.ctypes.data,
                                             has.ctypes.data if has is not None else None, count, out.ctypes.data))
        return out

    # -- resize rule + ReinitializeAndRehash: physical_create_bf.cpp:386-406, bloom_filter.cpp:34-58
    def NeedsRebuild(self, actual_rows):
        out = C.c_int(0)
        _check(lib().rpt_bf_needs_rebuild(self._h, actual_rows, C.byref(out)))
        return bool(out.value)

    def ReinitializeAndRehash(self, actual_rows, cols, count, stream=None):
        arr, n = _col_array(cols)
        _check(lib().rpt_bf_reinitialize_and_rehash(self._h, actual_rows, arr, n, count, stream))

    def Finalize(self, actual_rows, cols=None, count=0, stream=None):
        rebuilt = C.c_int(0)
        if cols is None:
            _check(lib().rpt_bf_finalize(self._h, actual_rows, None, 0, 0, C.byref(rebuilt), stream))
        else:
            arr, n = _col_array(cols)
            _check(lib().rpt_bf_finalize(self._h, actual_rows, arr, n, count, C.byref(rebuilt), stream))
        return bool(rebuilt.value)

    # -- UpdateMinMax next to the insert: physical_create_bf.cpp:82-176, 229-239 -----------------
    def MinMaxBegin(self, type):
        _check(lib().rpt_bf_minmax_begin(self._h, type))
        self._mm_type = type

    def MinMaxEnd(self):
        _check(lib().rpt_bf_minmax_end(self._h))

    def MinMaxGet(self):
        """Returns (has_value, has_nan, min, max) with min/max as numpy scalars of the column's physical type."""
        dt = _NUMPY_OF_TYPE[self._mm_type]
        lo, hi = np.zeros(1, dtype=dt), np.zeros(1, dtype=dt)
        has, nan = C.c_int(0), C.c_int(0)
        _check(lib().rpt_bf_minmax_get(self._h, C.byref(has), C.byref(nan), lo.ctypes.data, hi.ctypes.data))
        return bool(has.value), bool(nan.value), lo[0], hi[0]

    # -- filter words ---------------------------------------------------------------------------
    def ExportWords(self):
        words = np.empty(self.num_sectors, dtype=np.uint64)
        _check(lib().rpt_bf_export_words(self._h, words.ctypes.data, words.shape[0]))
        return words

    def Checksum(self):
        """FNV-1a-64 over the filter words (SURVEY §8c recipe)."""
        return self._get(lib().rpt_bf_checksum, C.c_uint64)

    @property
    def words_generation(self):
        return self._get(lib().rpt_bf_words_generation, C.c_uint64)

    def SetPrefilter(self, mode):
        """Folded pre-filter of large probes (rpt_gpu.h): -1 library default, 0 off, 1 by density, 2 always."""
        _check(lib().rpt_bf_set_prefilter(self._h, int(mode)))

    def PrefilterInfo(self):
        """(state, pass_rate): state 0 = no fold for the current words, 1 = in use, 2 = made but too dense."""
        state, rate = C.c_int(0), C.c_double(1.0)
        _check(lib().rpt_bf_prefilter_info(self._h, C.byref(state), C.byref(rate)))
        return state.value, rate.value

    def ImportWords(self, words):
        words = np.ascontiguousarray(words, dtype=np.uint64)
        _check(lib().rpt_bf_import_words(self._h, words.ctypes.data, words.shape[0]))

    def Sync(self):
        _check(lib().rpt_bf_sync(self._h))

    def SetHasData(self, flag):
        _check(lib().rpt_bf_set_has_data(self._h, int(bool(flag))))

    # -- multi-GPU merge --------------------------------------------------------------------------
    def OrParts(self, dev_parts_ptr, nparts, stream=None):
        _check(lib().rpt_bf_or_parts(self._h, _ptr(dev_parts_ptr), nparts, stream))

    def IpcExport(self):
        buf = C.create_string_buffer(64)
        _check(lib().rpt_bf_ipc_export(self._h, buf))
        return buf.raw

    def PeerOpen(self, handles, rank):
        blob = b"".join(handles)
        _check(lib().rpt_bf_peer_open(self._h, blob, len(handles), rank))

    def PeerMergePhase(self, phase, stream=None):
        _check(lib().rpt_bf_peer_merge_phase(self._h, phase, stream))

    def PeerMergeRange(self, phase, first_sector, nsectors, stream=None):
        _check(lib().rpt_bf_peer_merge_range(self._h, phase, first_sector, nsectors, stream))

    def SetSliceEvents(self, events):
        """events: raw cudaEvent_t handles (ints), one per 64 MiB slice; [] clears the gates."""
        arr = (C.c_void_p * max(len(events), 1))(*[C.c_void_p(int(e)) for e in events])
        _check(lib().rpt_bf_set_slice_events(self._h, arr, len(events)))

    def PeerClose(self):
        _check(lib().rpt_bf_peer_close(self._h))

    # the whole OR-allreduce owned by the library (barrier words + stream memory operations, rpt_gpu.h)
    def PeerFlagsExport(self):
        buf = C.create_string_buffer(64)
        _check(lib().rpt_bf_peer_flags_export(self._h, buf))
        return buf.raw

    def PeerFlagsOpen(self, handles, rank):
        blob = b"".join(handles)
        _check(lib().rpt_bf_peer_flags_open(self._h, blob, len(handles), rank))

    def PeerAllreduce(self, has_local_data, stream=None):
        _check(lib().rpt_bf_peer_allreduce(self._h, int(bool(has_local_data)), stream))

    def PeerRelease(self, stream=None):
        _check(lib().rpt_bf_peer_release(self._h, stream))

    def PeerAnyData(self, stream=None):
        out = C.c_int(0)
        _check(lib().rpt_bf_peer_any_data(self._h, stream, C.byref(out)))
        return bool(out.value)


class BlockedBloomFilter:
    """The filter the reference's README.md:23-32 describes (Arrow Acero BlockedBloomFilter: 1024-entry mask table,
    8 bits per key, Fold) over rpt_abf_*.  NOT the reference's code path — parity unpinned; the
    reference-exact filter is PTBloomFilter above."""

    def __init__(self, device=0):
        self._h = C.c_void_p(None)
        self.device = device

    def CreateEmpty(self, num_rows_to_insert):
        self.close()
        _check(lib().rpt_abf_create(C.byref(self._h), int(num_rows_to_insert), self.device))
        return self

    def close(self):
        if self._h:
            lib().rpt_abf_destroy(self._h)
            self._h = C.c_void_p(None)

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def log_num_blocks(self):
        out = C.c_int(0)
        _check(lib().rpt_abf_log_num_blocks(self._h, C.byref(out)))
        return out.value

    @property
    def launch_count(self):
        out = C.c_longlong(0)
        _check(lib().rpt_abf_launch_count(self._h, C.byref(out)))
        return out.value

    @property
    def device_blocks(self):
        out = C.c_void_p(None)
        _check(lib().rpt_abf_device_blocks(self._h, C.byref(out)))
        return out.value

    def Insert(self, cols, count, stream=None):
        arr, n = _col_array(cols)
        _check(lib().rpt_abf_insert(self._h, arr, n, count, stream))

    def Fold(self, stream=None):
        """Returns the number of halvings."""
        out = C.c_int(0)
        _check(lib().rpt_abf_fold(self._h, stream, C.byref(out)))
        return out.value

    def Find(self, cols, count):
        """Ascending row ids of the rows that pass (uint32 numpy array)."""
        arr, n = _col_array(cols)
        sel = np.empty(max(count, 1), dtype=np.uint32)
        found = C.c_uint64(0)
        _check(lib().rpt_abf_probe(self._h, arr, n, count, sel.ctypes.data, C.addressof(found), MEM_HOST, None))
        return sel[: found.value]

    def FindDevice(self, cols, count, sel_out_ptr, found_out_ptr, stream=None):
        arr, n = _col_array(cols)
        _check(lib().rpt_abf_probe(self._h, arr, n, count, _ptr(sel_out_ptr), _ptr(found_out_ptr), MEM_DEVICE,
                                   stream))

    def OrParts(self, dev_parts_ptr, nparts, stream=None):
        _check(lib().rpt_abf_or_parts(self._h, _ptr(dev_parts_ptr), nparts, stream))

    def ExportBlocks(self):
        out = np.empty(1 << self.log_num_blocks, dtype=np.uint64)
        _check(lib().rpt_abf_export_blocks(self._h, out.ctypes.data, out.shape[0]))
        return out

    def Sync(self):
        _check(lib().rpt_abf_sync(self._h))


def acero_mask_bits():
    out = np.zeros(136, dtype=np.uint8)
    _check(lib().rpt_abf_mask_bits(out.ctypes.data))
    return out


def acero_log_blocks_for_rows(rows):
    out = C.c_int(0)
    _check(lib().rpt_abf_log_blocks_for_rows(int(rows), C.byref(out)))
    return out.value


def allreduce_or(filters):
    """Single-process multi-GPU OR-allreduce over peer access (rpt_bf_allreduce_or)."""
    arr = (C.c_void_p * len(filters))(*[f._h for f in filters])
    _check(lib().rpt_bf_allreduce_or(arr, len(filters)))


def synth_build_keys(dev_ptr, first, count, device=0, stream=None):
    _check(lib().rpt_synth_build_keys(_ptr(dev_ptr), first, count, device, stream))


def synth_probe_keys(dev_ptr, first, count, n_build, sel_ppm, device=0, stream=None):
    _check(lib().rpt_synth_probe_keys(_ptr(dev_ptr), first, count, n_build, sel_ppm, device, stream))


class PinnedArray:
    """numpy view over rpt_host_alloc'd (page-locked) memory."""

    def __init__(self, count, dtype):
        self.dtype = np.dtype(dtype)
        self.nbytes = int(count) * self.dtype.itemsize
        p = C.c_void_p(None)
        _check(lib().rpt_host_alloc(C.byref(p), self.nbytes))
        self.ptr = p.value
        buf = (C.c_uint8 * max(self.nbytes, 1)).from_address(self.ptr)
        self.array = np.frombuffer(buf, dtype=self.dtype, count=int(count))

    def free(self):
        if self.ptr:
            self.array = None
            lib().rpt_host_free(self.ptr)
            self.ptr = None


from .sharded import ShardedBloomFilter, shard_range  # noqa: E402,F401


def synth_decimal_text(x):
    """SURVEY §8(d) config 3's second key column: k2 = decimal text of k1 (1–19 bytes).  int64 torch tensor on a CUDA
    device (bit 63 is masked off) → Arrow utf8 layout on the same device: (int32 offsets[n+1], uint8 bytes, total
    bytes).  Digit column by digit column, to bound the temporaries; benchmark / test input only."""
    import torch
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
    if total >= 2**32:
        raise RptError("string bytes exceed the 32-bit offsets of the VARCHAR layout")
    blob = torch.zeros(total + 8, dtype=torch.uint8, device=x.device)
    p = 1
    for d in range(19):  # digit d counted from the right goes to offsets[i+1] - 1 - d
        has = ndig > d
        pos = (offsets[1:] - 1 - d)[has]
        blob[pos] = ((x[has] // p) % 10 + 48).to(torch.uint8)
        p *= 10
        del has, pos
    return offsets.to(torch.int32), blob, total
