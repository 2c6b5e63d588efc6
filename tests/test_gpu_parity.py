"""GPU suite (-m gpu): the CUDA path, called through the C ABI (librpt_gpu.so), against the CPU oracle.

Bit-exact everywhere: filter words are compared with array_equal (memcmp), selection vectors element-wise.
The oracle itself is pinned to the reference by tests/test_oracle_cpu.py; the committed golden vectors
(tests/golden/ref_vectors.json, produced by the reference's own sources) are checked here directly as well.
Nothing here reads /root/reference.
"""
import numpy as np
import pytest

from helpers import NP, fnv_words, from_hex, hx, load_json, synth_build_keys, synth_probe_keys

pytestmark = pytest.mark.gpu

KA = load_json("known_answers.json")
RV = load_json("ref_vectors.json")


@pytest.fixture(scope="module")
def torch_cuda():
    import torch
    assert torch.cuda.is_available(), "the gpu suite needs a CUDA device"
    torch.cuda.set_device(0)
    return torch


def gpu_filter(rpt, est):
    return rpt.PTBloomFilter(0).Initialize(est)


def check_build_probe(rpt, oracle, est, build_cols, n, probe_cols, m):
    """Insert n rows, compare the words; probe m rows, compare the selection vector."""
    bf = gpu_filter(rpt, est)
    words = oracle.new_filter(est & 0xFFFFFFFF)
    assert bf.num_sectors == words.shape[0]
    assert bf.IsEmpty()
    bf.Insert(build_cols, n)
    oracle.insert(words, build_cols, n)
    assert bf.IsEmpty() == (n == 0)
    got = bf.ExportWords()
    assert np.array_equal(got, words), "filter words differ from the oracle"
    sel = bf.LookupSel(probe_cols, m)
    want = oracle.probe(words, probe_cols, m)
    assert sel.shape == want.shape and np.array_equal(sel, want), "selection vector differs from the oracle"
    bf.close()
    return words, want


# ---------------------------------------------------------------------------------------------------------
# known answers and golden vectors of the reference
# ---------------------------------------------------------------------------------------------------------
def test_known_masks_single_key(rpt):
    """One key → exactly GetMask(Hash(key)) in word Hash(key) & (S-1)."""
    for k, m in KA["mask_int64"].items():
        h = int(KA["hash_int64"][k], 16)
        bf = gpu_filter(rpt, 1)  # S = 8
        bf.Insert(rpt.host_col(np.array([int(k)], dtype=np.int64)), 1)
        words = bf.ExportWords()
        want = np.zeros(8, dtype=np.uint64)
        want[h & 7] = int(m, 16)
        assert np.array_equal(words, want), k
        bf.close()


def test_known_null_hash(rpt):
    bf = gpu_filter(rpt, 1)
    col = rpt.host_col(np.array([7], dtype=np.int64), validity=np.array([0], dtype=np.uint64))
    bf.Insert(col, 1)
    words = bf.ExportWords()
    h = int(KA["null_hash"], 16)
    hi = h >> 32
    mask = (1 << (hi & 63)) | (1 << ((hi >> 8) & 63)) | (1 << ((hi >> 16) & 63)) | (1 << ((hi >> 24) & 63))
    want = np.zeros(8, dtype=np.uint64)
    want[h & 7] = mask
    assert np.array_equal(words, want)
    bf.close()


def test_known_filter_100k(rpt, oracle):
    """est 100000, keys 2i → S=32768; probing 0..10^6-1 leaves 102353 rows."""
    f = KA["filters"][0]
    keys = 2 * np.arange(f["n_keys"], dtype=np.int64)
    probe = np.arange(f["probe_range"], dtype=np.int64)
    words, sel = check_build_probe(rpt, oracle, f["est_rows"], rpt.host_col(keys), keys.size, rpt.host_col(probe),
                                   probe.size)
    assert words.shape[0] == f["num_sectors"]
    assert sel.size == f["survivors"]


def test_golden_filters(rpt):
    """Filters and selection vectors made by the reference's own InsertHashes/LookupHashes (ref_vectors.json)."""
    for f in RV["filters"]:
        keys = synth_build_keys(0, f["n_keys"])
        probe = synth_probe_keys(0, f["n_probe"], max(f["n_keys"], 1), f["sel_ppm"])
        bf = gpu_filter(rpt, f["est_rows"])
        assert bf.num_sectors == f["num_sectors"]
        bf.Insert(rpt.host_col(keys), keys.size)
        assert hx(fnv_words(bf.ExportWords())) == f["fnv1a_words"]
        sel = bf.LookupSel(rpt.host_col(probe), probe.size)
        assert sel.size == f["survivors"]
        assert hx(fnv_words(sel.astype(np.uint64))) == f["sel_fnv"]
        assert sel[:8].tolist() == f["sel_head"] and sel[-8:].tolist() == f["sel_tail"]
        bf.close()


def golden_vector_cols(rpt):
    v = RV["vector"]
    n, dlen = v["n"], v["dlen"]
    c0 = rpt.host_col(from_hex(v["data_i64"], np.int64), sel=from_hex(v["sel"], np.uint32),
                      validity=from_hex(v["validity"], np.uint64), data_len=dlen)
    c1 = rpt.host_col(from_hex(v["data_i32"], np.int32), validity=from_hex(v["validity_i32"], np.uint64))
    c2 = rpt.host_varchar_col(from_hex(v["str_offsets"], np.uint32), from_hex(v["str_bytes"], np.uint8))
    return n, c0, c1, c2


def test_golden_vector_hashes(rpt, oracle):
    """The reference's TightLoopHash/CombineHash outputs (sel + validity + 3-column combine) → the filter
    built from those hashes must equal the filter the GPU builds from the raw columns."""
    v = RV["vector"]
    n, c0, c1, c2 = golden_vector_cols(rpt)
    for cols, key in (([c0], "hash_c0"), ([c0, c1], "hash_c0_c1"), ([c2], "hash_c2"), ([c1, c2, c0], "hash_c1_c2_c0")):
        hashes = from_hex(v[key], np.uint64)
        want = oracle.new_filter(n)
        oracle.insert_hashes(want, hashes)
        bf = gpu_filter(rpt, n)
        bf.Insert(cols, n)
        assert np.array_equal(bf.ExportWords(), want), key
        # and LookupHashes on the pre-hashed values agrees: every inserted hash is found
        assert bf.LookupHashes(hashes).all()
        bf.close()


def test_golden_scalar_hashes(rpt, oracle):
    """Per-type Hash<T> golden values: a filter of S=8 sectors built from value i alone has word/mask of hash i."""
    for t, entry in RV["scalars"].items():
        t = int(t)
        raw = bytes.fromhex(entry["bytes"])
        hashes = np.array([int(h, 16) for h in entry["hashes"]], dtype=np.uint64)
        if t in NP:
            vals = np.frombuffer(raw, dtype=NP[t]).copy()
            col = rpt.host_col(vals, type=t)
        else:
            vals = np.frombuffer(raw, dtype=np.uint64).reshape(-1, 2).copy()
            col = rpt.host_col(vals, type=t)
        n = vals.shape[0]
        want = oracle.new_filter(n)
        oracle.insert_hashes(want, hashes)
        bf = gpu_filter(rpt, n)
        bf.Insert(col, n)
        assert np.array_equal(bf.ExportWords(), want), f"physical type {t}"
        bf.close()


def test_golden_strings(rpt, oracle):
    strs = [bytes.fromhex(s["hex"]) for s in RV["strings"]]
    hashes = np.array([int(s["hash"], 16) for s in RV["strings"]], dtype=np.uint64)
    offs, data = rpt.strings_to_arrow(strs)
    n = len(strs)
    want = oracle.new_filter(n)
    oracle.insert_hashes(want, hashes)
    bf = gpu_filter(rpt, n)
    bf.Insert(rpt.host_varchar_col(offs, data), n)
    assert np.array_equal(bf.ExportWords(), want)
    bf.close()


# ---------------------------------------------------------------------------------------------------------
# differential against the oracle: every physical type, selections, validity, multi-column keys
# ---------------------------------------------------------------------------------------------------------
def random_values(rng, t, n):
    dt = NP[t]
    if t == 1:
        return rng.integers(0, 2, size=n, dtype=np.int8)
    if np.issubdtype(dt, np.integer):
        ii = np.iinfo(dt)
        return rng.integers(ii.min, ii.max, size=n, dtype=dt, endpoint=True)
    bits = rng.integers(0, 2 ** (8 * np.dtype(dt).itemsize), size=n, dtype=np.uint32 if dt == np.float32 else np.uint64)
    # random bit patterns include NaNs, denormals, infinities; sprinkle ±0
    bits[::17] = 0
    bits[5::17] = 1 << (8 * np.dtype(dt).itemsize - 1)
    return bits.view(dt)


@pytest.mark.parametrize("t", sorted(NP))
def test_types_flat(rpt, oracle, t):
    rng = np.random.default_rng(100 + t)
    n, m = 20011, 50021
    keys = random_values(rng, t, n)
    probe = np.concatenate([keys[rng.integers(0, n, size=m // 3)], random_values(rng, t, m - m // 3)])
    rng.shuffle(probe)
    check_build_probe(rpt, oracle, n, rpt.host_col(keys, type=t), n, rpt.host_col(probe, type=t), m)


@pytest.mark.parametrize("t", [203, 204, 21])
def test_types_16_byte(rpt, oracle, t):
    rng = np.random.default_rng(200 + t)
    n, m = 9001, 30011
    keys = rng.integers(0, 2**64, size=(n, 2), dtype=np.uint64)
    if t == 21:
        iv = np.zeros(n, dtype=[("m", "<i4"), ("d", "<i4"), ("us", "<i8")])
        iv["m"] = rng.integers(-2**31, 2**31 - 1, n)
        iv["d"] = rng.integers(-2**31, 2**31 - 1, n)
        iv["us"] = rng.integers(-2**62, 2**62, n)
        keys = iv.view(np.uint64).reshape(n, 2).copy()
    probe = np.concatenate([keys[rng.integers(0, n, size=m // 2)],
                            rng.integers(0, 2**64, size=(m - m // 2, 2), dtype=np.uint64)])
    check_build_probe(rpt, oracle, n, rpt.host_col(keys, type=t), n, rpt.host_col(probe, type=t), m)


@pytest.mark.parametrize("t", [9, 7, 5, 12])
def test_sel_and_validity(rpt, oracle, t):
    """Dictionary-style vectors: rows read data[sel[i]], NULL where the validity bit of sel[i] is clear;
    NULL rows are inserted and probed with NULL_HASH like any value (vector_hash.cpp:16-23)."""
    rng = np.random.default_rng(300 + t)
    dlen, n, m = 777, 10007, 40009
    data = random_values(rng, t, dlen)
    sel = rng.integers(0, dlen, size=n, dtype=np.uint32)
    validity = rng.integers(0, 2**64, size=(dlen + 63) // 64, dtype=np.uint64)
    build = rpt.host_col(data, type=t, sel=sel, validity=validity, data_len=dlen)
    pdata = np.concatenate([data, random_values(rng, t, dlen)])
    psel = rng.integers(0, 2 * dlen, size=m, dtype=np.uint32)
    pvalid = rng.integers(0, 2**64, size=(2 * dlen + 63) // 64, dtype=np.uint64)
    probe = rpt.host_col(pdata, type=t, sel=psel, validity=pvalid, data_len=2 * dlen)
    check_build_probe(rpt, oracle, n, build, n, probe, m)
    # validity without a selection (flat vector with NULLs), chunk-unaligned sizes
    v2 = rng.integers(0, 2**64, size=(n + 63) // 64, dtype=np.uint64)
    k2 = random_values(rng, t, n)
    check_build_probe(rpt, oracle, n, rpt.host_col(k2, type=t, validity=v2), n, rpt.host_col(k2, type=t), n)


def random_strings(rng, n, max_len=40):
    lens = rng.integers(0, max_len + 1, size=n)
    blob = rng.integers(0, 256, size=int(lens.sum()), dtype=np.uint8)
    offs = np.zeros(n + 1, dtype=np.uint32)
    offs[1:] = np.cumsum(lens)
    return offs, blob


def test_varchar(rpt, oracle):
    rng = np.random.default_rng(41)
    n, m = 30011, 60013
    offs, blob = random_strings(rng, n)
    poffs, pblob = random_strings(rng, m, max_len=3)  # short strings collide with short build strings
    check_build_probe(rpt, oracle, n, rpt.host_varchar_col(offs, blob), n, rpt.host_varchar_col(poffs, pblob), m)
    # probing the build column itself: no false negatives, identity selection
    bf = gpu_filter(rpt, n)
    bf.Insert(rpt.host_varchar_col(offs, blob), n)
    assert np.array_equal(bf.LookupSel(rpt.host_varchar_col(offs, blob), n), np.arange(n, dtype=np.uint32))
    bf.close()


def make_string_t(strs):
    """duckdb::string_t array (string_type.hpp:24): {uint32 len, char prefix[4], char* ptr} or 12 inlined bytes."""
    n = len(strs)
    arr = np.zeros(n * 16, dtype=np.uint8)
    keep = []
    for i, s in enumerate(strs):
        arr[i * 16:i * 16 + 4] = np.frombuffer(np.uint32(len(s)).tobytes(), dtype=np.uint8)
        if len(s) <= 12:
            arr[i * 16 + 4:i * 16 + 4 + len(s)] = np.frombuffer(s, dtype=np.uint8)
        else:
            buf = np.frombuffer(s, dtype=np.uint8).copy()
            keep.append(buf)
            arr[i * 16 + 4:i * 16 + 8] = buf[:4]
            arr[i * 16 + 8:i * 16 + 16] = np.frombuffer(np.uint64(buf.ctypes.data).tobytes(), dtype=np.uint8)
    return arr, keep


def test_string_t_host_layout(rpt, oracle):
    """RPT_STRING_T: the library gathers DuckDB's 16-byte string_t entries itself (hash.cpp:115-151 hashes the
    inlined and the pointer form identically)."""
    rng = np.random.default_rng(43)
    n = 5003
    offs, blob = random_strings(rng, n, max_len=30)
    strs = [blob[offs[i]:offs[i + 1]].tobytes() for i in range(n)]
    arr, keep = make_string_t(strs)
    col = rpt.KeyCol(rpt.RptCol(arr.ctypes.data, None, None, None, n, 0, rpt.STRING_T, rpt.MEM_HOST), (arr, keep))
    want = oracle.new_filter(n)
    oracle.insert(want, [rpt.host_varchar_col(offs, blob)], n)
    bf = gpu_filter(rpt, n)
    bf.Insert(col, n)
    assert np.array_equal(bf.ExportWords(), want)
    assert np.array_equal(bf.LookupSel(col, n), np.arange(n, dtype=np.uint32))
    bf.close()


def test_multi_column_keys(rpt, oracle):
    """HashColumns with >1 column (src/bloom_filter.cpp:11-24): Hash(col0) then CombineHash per column."""
    rng = np.random.default_rng(47)
    n, m = 15013, 45007
    a = rng.integers(-50, 50, size=n, dtype=np.int64)
    b = rng.integers(0, 100, size=n, dtype=np.int32)
    offs, blob = random_strings(rng, n, max_len=2)
    va = rng.integers(0, 2**64, size=(n + 63) // 64, dtype=np.uint64)
    build = [rpt.host_col(a, validity=va), rpt.host_col(b), rpt.host_varchar_col(offs, blob)]
    pa = rng.integers(-50, 50, size=m, dtype=np.int64)
    pb = rng.integers(0, 100, size=m, dtype=np.int32)
    poffs, pblob = random_strings(rng, m, max_len=2)
    probe = [rpt.host_col(pa), rpt.host_col(pb), rpt.host_varchar_col(poffs, pblob)]
    check_build_probe(rpt, oracle, n, build, n, probe, m)
    check_build_probe(rpt, oracle, n, build[:2], n, probe[:2], m)


# ---------------------------------------------------------------------------------------------------------
# sizes and edges
# ---------------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("n", [0, 1, 2, 3, 31, 32, 33, 63, 64, 65, 511, 512, 513, 4095, 4096, 4097, 8191, 8192, 8193])
def test_ragged_counts(rpt, oracle, n):
    rng = np.random.default_rng(500 + n)
    keys = rng.integers(-2**63, 2**63 - 1, size=max(n, 1), dtype=np.int64)[:n]
    m = n + 5
    probe = np.concatenate([keys, rng.integers(-2**63, 2**63 - 1, size=5, dtype=np.int64)])
    check_build_probe(rpt, oracle, 64, rpt.host_col(keys) if n else rpt.host_col(np.zeros(1, np.int64)), n,
                      rpt.host_col(probe), m)


def test_empty_probe_and_empty_filter(rpt):
    bf = gpu_filter(rpt, 1000)
    assert bf.IsEmpty() and not bf.finalized_
    keys = np.arange(100, dtype=np.int64)
    assert bf.LookupSel(rpt.host_col(keys), 0).size == 0           # count == 0
    assert bf.LookupSel(rpt.host_col(keys), 100).size == 0         # all-zero words pass nothing
    bf.Insert(rpt.host_col(keys), 0)                               # no-op keeps has_data_ false
    assert bf.IsEmpty()
    assert not bf.Finalize(0)
    assert bf.finalized_
    bf.close()


@pytest.mark.parametrize("est,n", [(10, 10), (1300, 1300), (21845, 21845), (21846, 21846), (200000, 200000),
                                   (3000000, 3000000)])
def test_filter_sizes(rpt, oracle, est, n):
    """Shared-memory-staged filters (S ≤ 2^14), the boundary to global filters, and L2-sized filters."""
    keys = synth_build_keys(0, n)
    m = min(4 * n + 1000, 4000003)
    probe = synth_probe_keys(0, m, n, 250000)
    check_build_probe(rpt, oracle, est, rpt.host_col(keys), n, rpt.host_col(probe), m)


@pytest.mark.parametrize("est,slices", [(50_000_000, 2), (100_000_000, 4), (300_000_000, 8)])
def test_filters_beyond_one_l2_slice(rpt, oracle, est, slices):
    """F = 128 / 256 / 512 MiB: the build runs in sector-range passes; the probe runs in sector-range passes (2
    slices; 4-byte keys at 4 slices) or partitions the keys by slice (probe_partition_kernel / probe_bucket_kernel:
    4 and 8 slices; test_probe_dispatch_forced_either_way repeats this test with either path forced at every
    size).  Ragged probe length;
    a probe column whose keys all fall into one slice overflows the fixed (segment, slice) regions and takes the
    on-the-spot lookups instead."""
    assert oracle.num_sectors(est) == slices << 23
    n = 300_001
    keys = synth_build_keys(0, n)
    m = 1_300_007
    probe = synth_probe_keys(0, m, n, 250000)
    check_build_probe(rpt, oracle, est, rpt.host_col(keys), n, rpt.host_col(probe), m)
    # skew: many copies of few keys → every key of a segment in the same slice
    rng = np.random.default_rng(est)
    few = np.concatenate([keys[:5], rng.integers(-2**62, 2**62, 5)]).astype(np.int64)
    skew = few[rng.integers(0, few.size, 700_001)]
    check_build_probe(rpt, oracle, est, rpt.host_col(keys), n, rpt.host_col(skew), skew.size)
    # other key sources go through the same kernels (generic path): int32 keys with NULLs
    k32 = rng.integers(-2**31, 2**31 - 1, 200_000).astype(np.int32)
    p32 = np.concatenate([k32[:50_000], rng.integers(-2**31, 2**31 - 1, 450_003).astype(np.int32)])
    val = np.packbits(rng.random(p32.size) < 0.9, bitorder="little")
    val = np.concatenate([val, np.zeros((-len(val)) % 8, dtype=np.uint8)]).view(np.uint64)
    check_build_probe(rpt, oracle, est, rpt.host_col(k32), k32.size, rpt.host_col(p32, validity=val), p32.size)


def random_validity(rng, n, p_valid):
    bits = np.packbits(rng.random(n) < p_valid, bitorder="little")
    return np.concatenate([bits, np.zeros((-len(bits)) % 8, dtype=np.uint8)]).view(np.uint64)


@pytest.mark.parametrize("nulls", [False, True])
@pytest.mark.parametrize("dt", [np.int32, np.uint32, np.int64])
@pytest.mark.parametrize("est,m", [(2000, 300_001), (3_000_000, 2_000_003), (100_000_000, 1_000_001),
                                   (300_000_000, 700_001)])
def test_flat_columns_every_regime(rpt, oracle, dt, nulls, est, m):
    """Flat INT32 / UINT32 / INT64 columns, with and without a validity bitmap, take the staged paths (FlatU32Src /
    FlatU64Src): shared-memory filter, L2-resident filter, sector-range passes, partitioned probe; ragged lengths
    leave a partial last segment.  NULL rows hash to NULL_HASH on both sides (they are inserted and probed)."""
    rng = np.random.default_rng(est + m)
    info = np.iinfo(dt)
    n = min(est, 250_000)
    keys = rng.integers(info.min, info.max, n, dtype=dt, endpoint=True)
    members = keys[rng.integers(0, n, m // 5)]
    probe = np.concatenate([members, rng.integers(info.min, info.max, m - members.size, dtype=dt, endpoint=True)])
    rng.shuffle(probe)
    assert probe.size == m
    kv = random_validity(rng, n, 0.95) if nulls else None
    pv = random_validity(rng, m, 0.8) if nulls else None
    check_build_probe(rpt, oracle, est, rpt.host_col(keys, validity=kv), n, rpt.host_col(probe, validity=pv), m)


def test_duplicate_heavy_keys_small_filter(rpt, oracle):
    """Warp-aggregated atomics path: many lanes of a warp hit the same sector."""
    rng = np.random.default_rng(59)
    keys = rng.integers(0, 50, size=100003, dtype=np.int64)
    probe = rng.integers(0, 200, size=100003, dtype=np.int64)
    check_build_probe(rpt, oracle, 40, rpt.host_col(keys), keys.size, rpt.host_col(probe), probe.size)


def test_host_chunk_boundaries(rpt, oracle):
    """Host columns are uploaded in 2^24-row chunks; survivors of chunk c are offset by c·2^24."""
    n = (1 << 24) + 12345
    keys = synth_build_keys(0, n)
    m = (1 << 25) + 777
    probe = synth_probe_keys(0, m, n, 100000)
    v = np.full((n + 63) // 64, 0xFFFFFFFFFFFFFFFF, dtype=np.uint64)
    v[::5] = 0xFFFF0000FFFF0000
    check_build_probe(rpt, oracle, n, rpt.host_col(keys, validity=v), n, rpt.host_col(probe), m)


def test_unaligned_key_pointer(rpt, oracle):
    """An int64 column whose base is only 8-byte aligned cannot take the 128-bit path."""
    base = synth_build_keys(0, 100001)
    keys = base[1:]
    assert keys.ctypes.data % 16 == 8 or base.ctypes.data % 16 == 8
    check_build_probe(rpt, oracle, keys.size, rpt.host_col(keys), keys.size, rpt.host_col(base[3:]), base.size - 3)


# ---------------------------------------------------------------------------------------------------------
# resize rule, scan-side probe, pre-hashed lookups
# ---------------------------------------------------------------------------------------------------------
def test_finalize_resize_rule(rpt, oracle):
    """physical_create_bf.cpp:386-406 + ReinitializeAndRehash (src/bloom_filter.cpp:34-58)."""
    n = 50000
    keys = synth_build_keys(0, n)
    col = rpt.host_col(keys)
    bf = gpu_filter(rpt, 1000)                       # under-estimate: 16384 bits < 8 bits/key · 50000
    bf.Insert(col, n)
    assert bf.NeedsRebuild(n) == oracle.needs_rebuild(1000, n) == True  # noqa: E712
    with pytest.raises(rpt.RptError):
        bf.Finalize(n)                               # the rule fires but there is nothing to rehash from
    assert bf.Finalize(n, col, n)
    assert bf.finalized_ and bf.SizedForRows() == n
    want = oracle.new_filter(n)
    oracle.insert(want, [col], n)
    assert bf.num_sectors == want.shape[0]
    assert np.array_equal(bf.ExportWords(), want)
    bf.close()
    bf = gpu_filter(rpt, n)                          # good estimate: untouched
    bf.Insert(col, n)
    assert not bf.Finalize(n, col, n)
    assert np.array_equal(bf.ExportWords(), want)
    # never shrinks: far fewer rows than estimated keeps the allocation
    assert not bf.NeedsRebuild(10)
    bf.close()


@pytest.mark.parametrize("t", sorted(NP))
def test_minmax_next_to_insert(rpt, t):
    """UpdateMinMax (physical_create_bf.cpp:82-176): range of the valid rows, accumulated over chunked inserts
    (host chunks with a validity mask, then a device-style flat chunk), against numpy on the same rows."""
    rng = np.random.default_rng(1000 + t)
    dt = NP[t]
    n = 70_001

    def column(k):
        if t == 1:  # BOOL, held as one byte 0/1
            return rng.integers(0, 2, k).astype(dt)
        if np.issubdtype(dt, np.floating):
            return (rng.standard_normal(k) * 1e6).astype(dt)
        info = np.iinfo(dt)
        return rng.integers(info.min, info.max, k, dtype=dt, endpoint=True)

    bf = gpu_filter(rpt, 2 * n)
    bf.MinMaxBegin(t)
    has, nan, _, _ = bf.MinMaxGet()
    assert not has and not nan                      # ColumnMinMax::has_value starts false
    a = column(n)
    valid = rng.random(n) < 0.7
    validity = np.packbits(valid, bitorder="little")
    validity = np.concatenate([validity, np.zeros((-len(validity)) % 8, dtype=np.uint8)]).view(np.uint64)
    bf.Insert(rpt.host_col(a, type=t, validity=validity), n)
    b = column(n)
    bf.Insert(rpt.host_col(b, type=t), n)
    has, nan, lo, hi = bf.MinMaxGet()
    ref = np.concatenate([a[valid], b])
    assert has and not nan
    assert lo == ref.min() and hi == ref.max()
    assert t == 1 or (lo.dtype == dt and hi.dtype == dt)
    # an all-NULL chunk changes nothing; after MinMaxEnd inserts are no longer folded in
    bf.Insert(rpt.host_col(column(64), type=t, validity=np.zeros(1, dtype=np.uint64)), 64)
    bf.MinMaxEnd()
    bf.Insert(rpt.host_col(column(n), type=t), n)
    assert bf.MinMaxGet()[2:] == (lo, hi)
    bf.close()


def test_minmax_nan_and_unsupported(rpt):
    bf = gpu_filter(rpt, 100)
    bf.MinMaxBegin(rpt.DOUBLE)
    bf.Insert(rpt.host_col(np.array([3.0, np.nan, -0.0, 7.5], dtype=np.float64)), 4)
    has, nan, lo, hi = bf.MinMaxGet()
    assert has and nan and lo == 0.0 and np.signbit(lo) and hi == 7.5
    with pytest.raises(rpt.RptError):
        bf.MinMaxBegin(rpt.VARCHAR)
    bf.close()


def test_uint32_estimate_wrap(rpt, oracle):
    """PTBloomFilter::Initialize takes uint32_t est_num_rows (bloom_filter.hpp:25): 2^32+100 sizes like 100."""
    bf = gpu_filter(rpt, 2**32 + 100)
    assert bf.num_sectors == oracle.num_sectors(100) == 32
    assert bf.SizedForRows() == 100
    bf.close()


def test_filter_sel_scan_side(rpt, oracle):
    """BFTableFilter::Filter (duckdb bloom_filter.cpp:94-131): keys read through an incoming selection, output
    translated back through it."""
    rng = np.random.default_rng(61)
    n, rows = 20000, 70001
    keys = synth_build_keys(0, n)
    bf = gpu_filter(rpt, n)
    bf.Insert(rpt.host_col(keys), n)
    words = bf.ExportWords()
    column = synth_probe_keys(0, rows, n, 300000)
    in_sel = np.sort(rng.choice(rows, size=30011, replace=False)).astype(np.uint32)
    got = bf.FilterSel(rpt.host_col(column), in_sel, in_sel.size)
    flat = oracle.probe(words, [rpt.host_col(column[in_sel])], in_sel.size)
    assert np.array_equal(got, in_sel[flat])
    got = bf.FilterSel(rpt.host_col(column), None, rows)
    assert np.array_equal(got, oracle.probe(words, [rpt.host_col(column)], rows))
    bf.close()


def test_lookup_prehashed(rpt, oracle):
    """BloomFilter::LookupOne on hashes (BFTableFilter::FilterValue / CheckStatistics :138-169)."""
    n = 3000
    keys = np.arange(1000, 1000 + n, dtype=np.int64)
    bf = gpu_filter(rpt, n)
    bf.Insert(rpt.host_col(keys), n)
    words = bf.ExportWords()
    span = np.arange(0, 2049, dtype=np.int64)
    hashes = oracle.hash_columns([rpt.host_col(span)], span.size)
    hits = bf.LookupHashes(hashes)
    sel = oracle.lookup_hashes(words, hashes)
    want = np.zeros(span.size, dtype=bool)
    want[sel] = True
    assert np.array_equal(hits, want)
    bf.close()


def test_import_export_roundtrip(rpt, oracle):
    """A filter built by CPU sink threads can be uploaded and probed on the GPU, and vice versa."""
    n = 40000
    keys = synth_build_keys(0, n)
    words = oracle.new_filter(n)
    oracle.insert(words, [rpt.host_col(keys)], n)
    bf = gpu_filter(rpt, n)
    bf.ImportWords(words)
    assert not bf.IsEmpty()
    probe = synth_probe_keys(0, 100000, n, 100000)
    assert np.array_equal(bf.LookupSel(rpt.host_col(probe), probe.size),
                          oracle.probe(words, [rpt.host_col(probe)], probe.size))
    assert np.array_equal(bf.ExportWords(), words)
    bf.close()


# ---------------------------------------------------------------------------------------------------------
# device-resident columns (the benchmark path) and the synthetic generators
# ---------------------------------------------------------------------------------------------------------
def test_device_columns_and_generators(rpt, oracle, torch_cuda):
    torch = torch_cuda
    n, m = 1000003, 5000011
    dk = torch.empty(n, dtype=torch.int64, device="cuda")
    dp = torch.empty(m, dtype=torch.int64, device="cuda")
    rpt.synth_build_keys(dk, 0, n)
    rpt.synth_probe_keys(dp, 0, m, n, 100000)
    torch.cuda.synchronize()
    keys = dk.cpu().numpy().view(np.uint64)
    probe = dp.cpu().numpy().view(np.uint64)
    assert np.array_equal(keys, oracle.synth_build_keys(0, n))
    assert np.array_equal(probe, oracle.synth_probe_keys(0, m, n, 100000))
    bf = gpu_filter(rpt, n)
    torch.cuda.synchronize()  # inputs were produced on torch's stream; the handle's stream is non-blocking
    bf.Insert(rpt.device_col(dk, rpt.INT64), n)
    want = oracle.new_filter(n)
    oracle.insert(want, [rpt.host_col(keys)], n)
    assert np.array_equal(bf.ExportWords(), want)
    dsel = torch.empty(m, dtype=torch.int32, device="cuda")
    dfound = torch.zeros(1, dtype=torch.int64, device="cuda")
    torch.cuda.synchronize()  # inputs were produced on torch's stream; the handle's stream is non-blocking
    bf.LookupSelDevice(rpt.device_col(dp, rpt.INT64), m, dsel, dfound)
    bf.Sync()
    found = int(dfound.item())
    sel = dsel[:found].cpu().numpy().view(np.uint32)
    assert np.array_equal(sel, oracle.probe(want, [rpt.host_col(probe)], m))
    # device keys → host outputs
    assert np.array_equal(bf.LookupSel(rpt.device_col(dp, rpt.INT64), m), sel)
    # a device column whose base is only 8-byte aligned cannot take the 128-bit loads
    got = bf.LookupSel(rpt.device_col(dp[1:], rpt.INT64), m - 1)
    assert np.array_equal(got, sel[sel > 0] - 1)
    bf2 = gpu_filter(rpt, n)
    torch.cuda.synchronize()  # inputs were produced on torch's stream; the handle's stream is non-blocking
    bf2.Insert(rpt.device_col(dk[1:], rpt.INT64), n - 1)
    torch.cuda.synchronize()  # inputs were produced on torch's stream; the handle's stream is non-blocking
    bf2.Insert(rpt.device_col(dk[:1], rpt.INT64), 1)
    assert np.array_equal(bf2.ExportWords(), want)
    bf2.close()
    bf.close()


def test_device_varchar_and_validity(rpt, oracle, torch_cuda):
    torch = torch_cuda
    rng = np.random.default_rng(71)
    n = 200003
    offs, blob = random_strings(rng, n, max_len=25)
    valid = rng.integers(0, 2**64, size=(n + 63) // 64, dtype=np.uint64)
    pad = np.concatenate([blob, np.zeros(8, dtype=np.uint8)])
    d_off = torch.from_numpy(offs.view(np.int32)).cuda()
    d_blob = torch.from_numpy(pad).cuda()
    d_val = torch.from_numpy(valid.view(np.int64)).cuda()
    col = rpt.device_col(d_off, rpt.VARCHAR, validity=d_val, aux=d_blob, aux_bytes=blob.size, data_len=n)
    hcol = rpt.host_varchar_col(offs, blob, validity=valid)
    bf = gpu_filter(rpt, n)
    bf.Insert(col, n)
    want = oracle.new_filter(n)
    oracle.insert(want, [hcol], n)
    assert np.array_equal(bf.ExportWords(), want)
    assert np.array_equal(bf.LookupSel(col, n), np.arange(n, dtype=np.uint32))
    bf.close()


# ---------------------------------------------------------------------------------------------------------
# OR-merge of partial filters (multi-GPU path) on one device
# ---------------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("nparts", [2, 3, 8])
def test_sharded_build_equals_single_build(rpt, oracle, torch_cuda, nparts):
    """N row-range shards → N partial filters → OR → byte-identical to the 1-shard build."""
    torch = torch_cuda
    n = 300007
    keys = synth_build_keys(0, n)
    want = oracle.new_filter(n)
    oracle.insert(want, [rpt.host_col(keys)], n)
    bounds = [n * r // nparts for r in range(nparts + 1)]
    parts = []
    for r in range(nparts):
        p = gpu_filter(rpt, n)
        shard = keys[bounds[r]:bounds[r + 1]]
        p.Insert(rpt.host_col(shard), shard.size)
        parts.append(p)
    S = want.shape[0]
    # all-gather layout: nparts·S contiguous words
    gathered = torch.empty(nparts * S, dtype=torch.int64, device="cuda")
    for r, p in enumerate(parts):
        gathered[r * S:(r + 1) * S] = torch.from_numpy(p.ExportWords().view(np.int64)).cuda()
    merged = gpu_filter(rpt, n)
    merged.OrParts(gathered, nparts)
    assert np.array_equal(merged.ExportWords(), want)
    # in-place peer merge between handles that live on the same device
    rpt.allreduce_or(parts)
    for p in parts:
        assert np.array_equal(p.ExportWords(), want)
        assert not p.IsEmpty()
        p.close()
    merged.close()


# ---------------------------------------------------------------------------------------------------------
# full-size configuration (BASELINE.json config 2): known answer + size-independent properties
# ---------------------------------------------------------------------------------------------------------
def test_config2_known_answer_full_size(rpt, torch_cuda):
    """est n=10^8, keys 2i (i<10^8) → S=2^25; probing 0..2·10^8-1 leaves 100 242 750 rows —
    a number produced by the compiled reference itself (PTBloomFilter through libduckdb)."""
    torch = torch_cuda
    f = KA["filters"][1]
    n, m = f["n_keys"], f["probe_range"]
    dk = torch.arange(0, 2 * n, 2, dtype=torch.int64, device="cuda")
    dp = torch.arange(0, m, dtype=torch.int64, device="cuda")
    bf = gpu_filter(rpt, f["est_rows"])
    assert bf.num_sectors == f["num_sectors"]
    torch.cuda.synchronize()  # inputs were produced on torch's stream; the handle's stream is non-blocking
    bf.Insert(rpt.device_col(dk, rpt.INT64), n)
    dsel = torch.empty(m, dtype=torch.int32, device="cuda")
    dfound = torch.zeros(1, dtype=torch.int64, device="cuda")
    torch.cuda.synchronize()  # inputs were produced on torch's stream; the handle's stream is non-blocking
    bf.LookupSelDevice(rpt.device_col(dp, rpt.INT64), m, dsel, dfound)
    bf.Sync()
    found = int(dfound.item())
    assert found == f["survivors"]
    sel = dsel[:found]
    # properties: strictly ascending; every member (even row id) survives; probing the survivors again keeps all
    assert bool((sel[1:] > sel[:-1]).all())
    assert int((sel % 2 == 0).sum().item()) == n
    survivors = dp[sel.long()]
    dsel2 = torch.empty(found, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()  # inputs were produced on torch's stream; the handle's stream is non-blocking
    bf.LookupSelDevice(rpt.device_col(survivors, rpt.INT64), found, dsel2, dfound)
    bf.Sync()
    assert int(dfound.item()) == found
    assert bool((dsel2 == torch.arange(found, dtype=torch.int32, device="cuda")).all())
    bf.close()


def test_config2_full_size_against_oracle(rpt, oracle, torch_cuda):
    """Config 2 at BASELINE.json's sizes: 10^8-key build compared word for word with the (threaded) oracle;
    the 10^9-row probe compared exactly on two 2·10^7-row windows, and through properties on the whole."""
    torch = torch_cuda
    n, m, ppm = 10**8, 10**9, 100000
    dk = torch.empty(n, dtype=torch.int64, device="cuda")
    rpt.synth_build_keys(dk, 0, n)
    torch.cuda.synchronize()  # the generators ran on the NULL stream; the handle's stream is non-blocking
    bf = gpu_filter(rpt, n)
    torch.cuda.synchronize()  # inputs were produced on torch's stream; the handle's stream is non-blocking
    bf.Insert(rpt.device_col(dk, rpt.INT64), n)
    got = bf.ExportWords()
    keys = oracle.synth_build_keys(0, n)
    want = oracle.new_filter(n)
    oracle.bench_build(want, keys.view(np.int64), 8)  # same fetch_or loop on 8 threads
    assert np.array_equal(got, want)
    del keys
    # sharded build of the same column (4 row ranges, separate handles, OR-merged) is byte-identical
    parts = []
    for r in range(4):
        p = gpu_filter(rpt, n)
        lo, hi = n * r // 4, n * (r + 1) // 4
        torch.cuda.synchronize()  # inputs were produced on torch's stream; the handle's stream is non-blocking
        p.Insert(rpt.device_col(dk[lo:hi], rpt.INT64), hi - lo)
        parts.append(p)
    rpt.allreduce_or(parts)
    for p in parts:
        assert np.array_equal(p.ExportWords(), got)
        p.close()
    del dk, parts
    # probe
    dp = torch.empty(m, dtype=torch.int64, device="cuda")
    rpt.synth_probe_keys(dp, 0, m, n, ppm)
    torch.cuda.synchronize()
    dsel = torch.empty(m, dtype=torch.int32, device="cuda")
    dfound = torch.zeros(1, dtype=torch.int64, device="cuda")
    torch.cuda.synchronize()  # inputs were produced on torch's stream; the handle's stream is non-blocking
    bf.LookupSelDevice(rpt.device_col(dp, rpt.INT64), m, dsel, dfound)
    bf.Sync()
    found = int(dfound.item())
    sel = dsel[:found]
    assert bool((sel[1:] > sel[:-1]).all())
    # members (bit 63 clear) all survive; the rest are false positives at the filter's rate (≈0.24 %)
    member = dp >= 0
    n_member = int(member.sum().item())
    assert int(member[sel.long()].sum().item()) == n_member
    fp = found - n_member
    assert 0.0015 < fp / (m - n_member) < 0.0035
    # exact windows against the oracle
    W = 2 * 10**7
    for start in (0, m - W):
        probe = oracle.synth_probe_keys(start, W, n, ppm)
        want_sel = oracle.probe(want, [rpt.host_col(probe)], W).astype(np.int64) + start
        sel64 = sel.long() & 0xFFFFFFFF
        mask = (sel64 >= start) & (sel64 < start + W)
        got_sel = sel64[mask].cpu().numpy()
        assert np.array_equal(got_sel, want_sel), start
    # the whole probe split in two calls concatenates to the same selection vector
    half = m // 2 + 12345
    dsel_a = torch.empty(half, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()  # inputs were produced on torch's stream; the handle's stream is non-blocking
    bf.LookupSelDevice(rpt.device_col(dp[:half], rpt.INT64), half, dsel_a, dfound)
    bf.Sync()
    fa = int(dfound.item())
    dsel_b = torch.empty(m - half, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()  # inputs were produced on torch's stream; the handle's stream is non-blocking
    bf.LookupSelDevice(rpt.device_col(dp[half:], rpt.INT64), m - half, dsel_b, dfound)
    bf.Sync()
    fb = int(dfound.item())
    assert fa + fb == found
    assert bool((dsel_a[:fa] == sel[:fa]).all())
    assert bool(((dsel_b[:fb].long() & 0xFFFFFFFF) + half == (sel[fa:].long() & 0xFFFFFFFF)).all())
    bf.close()


def test_cpp_host_mirror(rpt, oracle):
    """tests/cpp/host_mirror_test.cpp: CreateBFBatch / UseBFBatch / PTBloomFilter (include/rpt_gpu.hpp) driven like
    the reference's Sink / Finalize / ExecuteInternal, compared with the oracle inside the C++ driver."""
    import subprocess
    from cpp import build_cpp
    path = build_cpp.build_host_mirror_test()
    proc = subprocess.run([path], capture_output=True, text=True, timeout=600)
    assert proc.returncode == 0, proc.stdout + proc.stderr
    assert "host mirror ok" in proc.stdout


@pytest.mark.parametrize("type_id", [2, 3, 4, 5, 6, 7, 8, 9, 203, 204])
def test_check_statistics_zone_maps(rpt, oracle, type_id):
    """rpt_bf_check_statistics (BFTableFilter::CheckStatistics, duckdb bloom_filter.cpp:138-196) over a batch of
    row groups vs the oracle, every integral physical type: always-true, always-false and mixed ranges, the
    type's limits, invalid (min > max) and too-wide ranges, missing statistics, 128-bit carry."""
    from test_oracle_cpu import INT_TYPES, stats_arrays, stats_cases
    rng = np.random.default_rng(100 + type_id)
    _, bits, signed = INT_TYPES[type_id]
    tmax = (1 << (bits - 1)) - 1 if signed else (1 << bits) - 1
    # build column: every value of 0..min(3000, tmax) except multiples of 7, plus the type's last values;
    # values 3001..5000 stay out (always-false ranges, up to false positives — the oracle sees the same words)
    vals = [v for v in range(0, min(3000, tmax) + 1) if v % 7] + [tmax - k for k in range(min(4, tmax))]
    if bits == 128:
        u = [v & ((1 << 128) - 1) for v in vals]
        data = np.array([[x & 0xFFFFFFFFFFFFFFFF, x >> 64] for x in u], dtype=np.uint64)
        col = rpt.host_col(data, type=type_id)
    else:
        col = rpt.host_col(np.array(vals, dtype=INT_TYPES[type_id][0]))
    bf = rpt.PTBloomFilter(0).Initialize(len(vals))
    bf.Insert(col, len(vals))
    words = bf.ExportWords()
    pairs = stats_cases(type_id, rng, n_random=400)
    mins, maxs = stats_arrays(type_id, pairs)
    has = (rng.random(len(pairs)) < 0.95).astype(np.uint8)
    got = bf.CheckStatistics(type_id, mins, maxs, has)
    want = oracle.check_statistics(words, type_id, mins, maxs, has)
    assert np.array_equal(got, want)
    assert np.array_equal(bf.CheckStatistics(type_id, mins, maxs), oracle.check_statistics(words, type_id, mins, maxs))
    if tmax > 5000:
        assert set(want.tolist()) == {0, 1, 2}
    assert bf.CheckStatistics(type_id, mins[:0], maxs[:0]).size == 0
    # a physical type the reference does not prune on
    f = np.array([1.0, 2.0], dtype=np.float64)
    assert bf.CheckStatistics(rpt.DOUBLE, f, f).tolist() == [0, 0]
    bf.close()


@pytest.mark.parametrize("mode", ["0", "1"])
def test_probe_dispatch_forced_either_way(mode):
    """The dispatch between sector-range passes and the partitioned probe follows measurements; both
    paths must be right at every sliced filter size, so the large-filter parity tests run again in a child process
    with RPT_PROBE_PARTITION=0 (passes everywhere) and =1 (partition everywhere).  The knob is read once per
    process, hence the child."""
    import os
    import subprocess
    import sys
    if os.environ.get("RPT_TEST_NESTED"):
        pytest.skip("already inside the child run")
    env = dict(os.environ, RPT_PROBE_PARTITION=mode, RPT_TEST_NESTED="1")
    here = os.path.dirname(os.path.abspath(__file__))
    proc = subprocess.run([sys.executable, "-m", "pytest", os.path.join(here, "test_gpu_parity.py"), "-x", "-q", "-m", "gpu",
                           "-p", "no:cacheprovider", "-k", "beyond_one or flat_columns_every_regime"],
                          env=env, capture_output=True, text=True, timeout=900)
    assert proc.returncode == 0, proc.stdout[-3000:] + proc.stderr[-2000:]
    assert " passed" in proc.stdout


def test_dictionary_varchar_hashed_once(rpt, oracle):
    """Dictionary vectors (a VARCHAR dictionary seen through a selection, vector_hash.cpp:484-491): the library
    hashes the dictionary once and the rows gather the hashes.  Same words and selections as hashing row by row
    (the oracle), NULL dictionary entries included; host columns, device columns, and duckdb::string_t dictionaries;
    the extra launch shows that the short path was taken, and a dictionary longer than half the rows does not take it."""
    import torch
    rng = np.random.default_rng(77)
    words_list = [b"k%d-%s" % (i, b"x" * int(rng.integers(0, 40))) for i in range(3000)]
    off, blob = rpt.strings_to_arrow(words_list)
    dval = random_validity(rng, len(words_list), 0.97)
    n, m = 120_000, 400_003
    bsel = rng.integers(0, 1500, n).astype(np.uint32)            # build rows use the first half of the dictionary
    psel = rng.integers(0, len(words_list), m).astype(np.uint32)
    build = rpt.host_varchar_col(off, blob, sel=bsel, validity=dval)
    probe = rpt.host_varchar_col(off, blob, sel=psel, validity=dval)
    bf = gpu_filter(rpt, n)
    l0 = bf.launch_count
    bf.Insert(build, n)
    assert bf.launch_count - l0 == 2  # hash_entries_kernel + build_kernel
    words = oracle.new_filter(n)
    oracle.insert(words, build, n)
    assert np.array_equal(bf.ExportWords(), words)
    want = oracle.probe(words, probe, m)
    assert np.array_equal(bf.LookupSel(probe, m), want)
    assert 0 < want.size < m
    # device-resident dictionary column
    d_off, d_blob = torch.from_numpy(off.astype(np.int32)).cuda(), torch.from_numpy(np.concatenate([blob, np.zeros(8, np.uint8)])).cuda()
    d_sel, d_val = torch.from_numpy(psel.astype(np.int32)).cuda(), torch.from_numpy(dval.view(np.int64)).cuda()
    dcol = rpt.device_col(d_off, rpt.VARCHAR, sel=d_sel, validity=d_val, data_len=len(words_list), aux=d_blob,
                          aux_bytes=blob.size)
    dsel = torch.zeros(m, dtype=torch.int32, device="cuda")
    dfound = torch.zeros(1, dtype=torch.int64, device="cuda")
    bf.LookupSelDevice(dcol, m, dsel, dfound)
    bf.Sync()
    torch.cuda.synchronize()
    f = int(dfound.item())
    assert f == want.size and np.array_equal(dsel[:f].cpu().numpy().view(np.uint32), want)
    # scan-side form through an incoming selection: rows are dictionary ids here, output = surviving in_sel entries
    got = bf.FilterSel(rpt.host_varchar_col(off, blob, validity=dval), psel, m)
    assert np.array_equal(got, psel[want])
    # a dictionary longer than half the rows is hashed row by row (one launch fewer); same answer
    short = 2 * len(words_list) - 1
    l1 = bf.launch_count
    got_short = bf.LookupSel(rpt.host_varchar_col(off, blob, sel=psel[:short], validity=dval), short)
    took = bf.launch_count - l1
    l2 = bf.launch_count
    got_long = bf.LookupSel(rpt.host_varchar_col(off, blob, sel=psel[:2 * len(words_list)], validity=dval),
                            2 * len(words_list))
    assert bf.launch_count - l2 == took + 1
    assert np.array_equal(got_short, want[want < short])
    assert np.array_equal(got_long, want[want < 2 * len(words_list)])
    bf.close()


def test_concurrent_sink_and_probe_threads(rpt, oracle):
    """The reference's threading contract: N sink threads Insert into ONE shared filter at the same
    time (physical_create_bf.cpp:220-227, relaxed fetch_or), later N threads probe the finalized, read-only filter
    with their own selection buffers (physical_use_bf.hpp:13-25).  Eight host threads do both through one handle:
    the words equal a sequential build and every thread's selection equals the oracle's."""
    import threading
    rng = np.random.default_rng(808)
    T, per = 8, 150_000
    keys = rng.integers(-2**62, 2**62, T * per).astype(np.int64)
    est = T * per
    bf = gpu_filter(rpt, est)
    errors = []

    def sink(t):
        try:
            part = keys[t * per:(t + 1) * per]
            for r0 in range(0, per, 30_000):  # several Insert calls per thread, interleaving with the others
                bf.Insert(rpt.host_col(part[r0:r0 + 30_000]), min(30_000, per - r0))
        except Exception as e:  # noqa: BLE001
            errors.append(e)

    threads = [threading.Thread(target=sink, args=(t,)) for t in range(T)]
    [t.start() for t in threads]
    [t.join() for t in threads]
    assert not errors, errors
    words = oracle.new_filter(est)
    oracle.insert(words, rpt.host_col(keys), keys.size)
    assert np.array_equal(bf.ExportWords(), words)
    assert not bf.Finalize(keys.size)

    probes = [np.where(rng.random(200_003) < 0.2, keys[rng.integers(0, keys.size, 200_003)],
                       rng.integers(-2**62, 2**62, 200_003)).astype(np.int64) for _ in range(T)]
    got = [None] * T

    def probe(t):
        try:
            for _ in range(3):
                got[t] = bf.LookupSel(rpt.host_col(probes[t]), probes[t].size)
        except Exception as e:  # noqa: BLE001
            errors.append(e)

    threads = [threading.Thread(target=probe, args=(t,)) for t in range(T)]
    [t.start() for t in threads]
    [t.join() for t in threads]
    assert not errors, errors
    for t in range(T):
        assert np.array_equal(got[t], oracle.probe(words, rpt.host_col(probes[t]), probes[t].size)), t
    bf.close()


@pytest.mark.parametrize("shape", ["short", "mixed", "long"])
def test_flat_varchar_staged_hashing(rpt, oracle, shape, torch_cuda):
    """Flat VARCHAR columns of ≥ 2^15 rows are hashed by hash_varchar_kernel (128-row steps staged through shared
    memory) ahead of the build / probe kernels.  short: decimal-text keys (every step staged); mixed: lengths 0..70
    with empty strings and NULLs (steps whose bytes exceed the 4 KiB buffer fall back to global loads); long: 100+
    byte strings (always the fallback).  Ragged row counts leave a partial last step.  Host and device columns."""
    import torch
    rng = np.random.default_rng({"short": 1, "mixed": 2, "long": 3}[shape])
    n, m = 70_001, 333_337

    def make(count, pool):
        if shape == "short":
            strs = [b"%d" % int(v) for v in rng.integers(0, 2**62, count)]
        elif shape == "mixed":
            strs = [bytes(rng.integers(97, 123, int(l), dtype=np.uint8)) for l in rng.integers(0, 71, count)]
        else:
            strs = [bytes(rng.integers(97, 123, int(l), dtype=np.uint8)) for l in rng.integers(100, 180, count)]
        if pool is not None:  # every fifth row is a build-side string
            for i in range(0, count, 5):
                strs[i] = pool[int(rng.integers(0, len(pool)))]
        return strs

    bstr = make(n, None)
    pstr = make(m, bstr)
    boff, bblob = rpt.strings_to_arrow(bstr)
    poff, pblob = rpt.strings_to_arrow(pstr)
    bval = random_validity(rng, n, 0.98)
    pval = random_validity(rng, m, 0.9)
    build = rpt.host_varchar_col(boff, bblob, validity=bval)
    probe = rpt.host_varchar_col(poff, pblob, validity=pval)
    bf = gpu_filter(rpt, n)
    l0 = bf.launch_count
    bf.Insert(build, n)
    assert bf.launch_count - l0 == 2  # hash_varchar_kernel + build_kernel
    words = oracle.new_filter(n)
    oracle.insert(words, build, n)
    assert np.array_equal(bf.ExportWords(), words)
    want = oracle.probe(words, probe, m)
    assert np.array_equal(bf.LookupSel(probe, m), want)
    assert m // 6 < want.size < m
    # device-resident column, unaligned byte buffer start (the kernel aligns its copies itself)
    d_off = torch.from_numpy((poff.astype(np.int64) + 3).astype(np.int32)).cuda()
    d_blob = torch.from_numpy(np.concatenate([np.zeros(3, np.uint8), pblob, np.zeros(16, np.uint8)])).cuda()
    d_val = torch.from_numpy(pval.view(np.int64)).cuda()
    dcol = rpt.device_col(d_off, rpt.VARCHAR, validity=d_val, data_len=m, aux=d_blob, aux_bytes=pblob.size + 3)
    dsel = torch.zeros(m, dtype=torch.int32, device="cuda")
    dfound = torch.zeros(1, dtype=torch.int64, device="cuda")
    bf.LookupSelDevice(dcol, m, dsel, dfound)
    bf.Sync()
    torch.cuda.synchronize()
    f = int(dfound.item())
    assert f == want.size and np.array_equal(dsel[:f].cpu().numpy().view(np.uint32), want)
    bf.close()


# ---------------------------------------------------------------------------------------------------------
# round 2: NULL / unreferenced string_t payloads, concurrent handles of different sizes, fused min/max, checksum
# ---------------------------------------------------------------------------------------------------------
def test_string_t_null_rows_with_poisoned_payload(rpt, oracle):
    """A NULL row's string_t is undefined in DuckDB and the reference never reads it (TightLoopHash:
    `is_null ? NULL_HASH : Hash(input)`, vector_hash.cpp:43-61).  The NULL entries here carry 0xFF in every byte
    (a 4 GiB length and a wild pointer): the library must not dereference them, flat or through a selection."""
    rng = np.random.default_rng(4301)
    n = 6007
    offs, blob = random_strings(rng, n, max_len=30)
    strs = [blob[offs[i]:offs[i + 1]].tobytes() for i in range(n)]
    arr, keep = make_string_t(strs)
    valid_bits = rng.random(n) < 0.8
    validity = np.zeros((n + 63) // 64, dtype=np.uint64)
    for i in np.nonzero(valid_bits)[0]:
        validity[i >> 6] |= np.uint64(1) << np.uint64(i & 63)
    for i in np.nonzero(~valid_bits)[0]:
        arr[i * 16:(i + 1) * 16] = 0xFF
    col = rpt.KeyCol(rpt.RptCol(arr.ctypes.data, None, None, validity.ctypes.data, n, 0, rpt.STRING_T, rpt.MEM_HOST),
                     (arr, keep, validity))
    ref_col = rpt.host_varchar_col(offs, blob, validity=validity)
    want = oracle.new_filter(n)
    oracle.insert(want, [ref_col], n)
    bf = gpu_filter(rpt, n)
    bf.Insert(col, n)
    assert np.array_equal(bf.ExportWords(), want)
    assert np.array_equal(bf.LookupSel(col, n), oracle.probe(want, [ref_col], n))
    # through a selection that references only some entries: the unreferenced ones are poisoned as well
    m = 3001
    sel = rng.choice(np.nonzero(valid_bits)[0], size=m).astype(np.uint32)
    arr2 = arr.copy()
    unref = np.ones(n, dtype=bool)
    unref[sel] = False
    for i in np.nonzero(unref)[0]:
        arr2[i * 16:(i + 1) * 16] = 0xFF
    col2 = rpt.KeyCol(rpt.RptCol(arr2.ctypes.data, None, sel.ctypes.data, None, n, 0, rpt.STRING_T, rpt.MEM_HOST),
                      (arr2, keep, sel))
    ref2 = rpt.host_varchar_col(offs, blob, sel=sel)
    assert np.array_equal(bf.LookupSel(col2, m), oracle.probe(want, [ref2], m))
    # a selection entry beyond data_len is an error, not a wild read
    bad = sel.copy()
    bad[7] = n + 5
    col3 = rpt.KeyCol(rpt.RptCol(arr2.ctypes.data, None, bad.ctypes.data, None, n, 0, rpt.STRING_T, rpt.MEM_HOST),
                      (arr2, keep, bad))
    with pytest.raises(rpt.RptError):
        bf.LookupSel(col3, m)
    bf.close()


def test_concurrent_probes_on_handles_of_different_sizes(rpt, oracle):
    """Two README-sized filters of different S (both shared-memory staged) probed from two host threads at the same
    time: the dynamic shared-memory limit of the kernels is raised once to the device maximum, never per launch to
    this filter's size, so one thread cannot lower it under the other's launch."""
    import threading
    rng = np.random.default_rng(909)
    sizes = [3000, 80000]  # S = 2^10 and 2^14 sectors
    filters, words, probes = [], [], []
    for est in sizes:
        keys = rng.integers(-2**62, 2**62, est).astype(np.int64)
        bf = gpu_filter(rpt, est)
        bf.Insert(rpt.host_col(keys), est)
        w = oracle.new_filter(est)
        oracle.insert(w, rpt.host_col(keys), est)
        assert np.array_equal(bf.ExportWords(), w)
        filters.append(bf)
        words.append(w)
        probes.append(np.where(rng.random(600_011) < 0.3, keys[rng.integers(0, est, 600_011)],
                               rng.integers(-2**62, 2**62, 600_011)).astype(np.int64))
    errors, got = [], [None, None]

    def run(t):
        try:
            for _ in range(20):
                got[t] = filters[t].LookupSel(rpt.host_col(probes[t]), probes[t].size)
        except Exception as e:  # noqa: BLE001
            errors.append(e)

    threads = [threading.Thread(target=run, args=(t,)) for t in range(2)]
    [t.start() for t in threads]
    [t.join() for t in threads]
    assert not errors, errors
    for t in range(2):
        assert np.array_equal(got[t], oracle.probe(words[t], rpt.host_col(probes[t]), probes[t].size))
        filters[t].close()


def test_minmax_is_fused_into_the_build_kernel(rpt, torch_cuda):
    """UpdateMinMax runs next to the insert on the same chunk (physical_create_bf.cpp:220-239): with range tracking on,
    a single-column insert launches exactly as many kernels as without (the range is folded in by build_kernel)."""
    torch = torch_cuda
    n = 3_000_017
    keys = torch.randint(-2**62, 2**62, (n,), dtype=torch.int64, device="cuda:0")
    plain = gpu_filter(rpt, n)
    plain.Insert(rpt.device_col(keys, rpt.INT64), n)
    plain.Sync()
    tracked = gpu_filter(rpt, n)
    tracked.MinMaxBegin(rpt.INT64)
    tracked.Insert(rpt.device_col(keys, rpt.INT64), n)
    has, nan, lo, hi = tracked.MinMaxGet()
    assert has and not nan and int(lo) == int(keys.min().item()) and int(hi) == int(keys.max().item())
    assert tracked.launch_count == plain.launch_count
    assert np.array_equal(tracked.ExportWords(), plain.ExportWords())
    # a filter large enough for sector-range passes: the range is folded in once, by the first pass
    big = gpu_filter(rpt, 100_000_000)
    big.MinMaxBegin(rpt.INT64)
    big.Insert(rpt.device_col(keys, rpt.INT64), n)
    has, nan, lo, hi = big.MinMaxGet()
    assert has and int(lo) == int(keys.min().item()) and int(hi) == int(keys.max().item())
    for f in (plain, tracked, big):
        f.close()


def test_checksum_known_answers(rpt):
    """rpt_bf_checksum = FNV-1a-64 per word (xor, then multiply), the recipe of SURVEY §8(c) and of the golden
    vectors' "fnv1a_words" (test_golden_filters ties that fold to reference-made filters) — equal to the test
    helper's fold of the exported words.  (The survey's own checksum value
    for this filter could not be reproduced by anyone, see tests/test_oracle_cpu.py; the words are pinned directly.)"""
    n = 100_000
    bf = gpu_filter(rpt, n)
    bf.Insert(rpt.host_col(np.arange(n, dtype=np.int64) * 2), n)
    assert bf.Checksum() == fnv_words(bf.ExportWords())
    g0 = bf.words_generation
    bf.Reset(n)
    assert bf.words_generation == g0          # same S: nothing reallocated
    bf.Reset(10 * n)
    assert bf.words_generation == g0 + 1      # new S: peers must re-open their mappings
    bf.close()
