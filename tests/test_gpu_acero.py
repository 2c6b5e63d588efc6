"""GPU suite (-m gpu) for the README-described filter variant (rpt_abf_*: Arrow-style mask table, 8 bits per key, Fold).

PARITY UNPINNED — the reference holds no code for this filter.  The CUDA kernels are compared bit for
bit with oracle/acero_oracle.cpp (an independent CPU restatement of the published Arrow algorithm) on the DuckDB key
hashes of the pinned oracle, and through size-independent properties at larger sizes.  Nothing here reads
/root/reference.
"""
import numpy as np
import pytest

from helpers import synth_build_keys, synth_probe_keys

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def torch_cuda():
    import torch
    assert torch.cuda.is_available(), "the gpu suite needs a CUDA device"
    torch.cuda.set_device(0)
    return torch


def check(rpt, oracle, acero, rows_for_sizing, build_cols, n, probe_cols, m, fold=True):
    f = rpt.BlockedBloomFilter(0).CreateEmpty(rows_for_sizing)
    blocks = acero.new_filter(rows_for_sizing)
    assert (1 << f.log_num_blocks) == blocks.shape[0]
    hb = oracle.hash_columns(build_cols, n)
    hp = oracle.hash_columns(probe_cols, m)
    f.Insert(build_cols, n)
    acero.insert(blocks, hb)
    assert np.array_equal(f.ExportBlocks(), blocks), "blocks differ from the checker after Insert"
    assert np.array_equal(f.Find(probe_cols, m), acero.find(blocks, hp)), "Find differs before Fold"
    if fold:
        k = f.Fold()
        folded, want_k = acero.fold(blocks)
        assert k == want_k and (1 << f.log_num_blocks) == folded.shape[0]
        assert np.array_equal(f.ExportBlocks(), folded), "blocks differ from the checker after Fold"
        assert np.array_equal(f.Find(probe_cols, m), acero.find(folded, hp)), "Find differs after Fold"
        assert f.Find(build_cols, n).shape[0] == n, "Fold lost a member"
    assert f.launch_count > 0
    f.close()


@pytest.mark.parametrize("n,m,oversize", [(50_000, 100_000, 1), (50_001, 99_999, 16), (1, 7, 1), (999, 1025, 64),
                                           (300_000, 1_000_003, 4),
                                           (20_000, 1_500_001, 1),      # 32 KiB: the largest filter staged in shared memory
                                           (3_000, 1_200_000, 1),       # 4 KiB, staged
                                           (100_000, 1_200_000, 1)])    # 128 KiB: L1/L2 hits
def test_int64_host_columns(rpt, oracle, acero, n, m, oversize):
    keys = synth_build_keys(0, n)
    probe = synth_probe_keys(0, m, n, 100_000)
    check(rpt, oracle, acero, n * oversize, rpt.host_col(keys), n, rpt.host_col(probe), m)


def test_empty_calls(rpt, oracle, acero):
    f = rpt.BlockedBloomFilter(0).CreateEmpty(0)
    assert f.log_num_blocks == 3
    empty = np.zeros(0, dtype=np.uint64)
    f.Insert(rpt.host_col(empty), 0)
    assert not f.ExportBlocks().any()
    assert f.Find(rpt.host_col(empty), 0).shape[0] == 0
    probe = synth_probe_keys(0, 1000, 10, 500_000)
    assert f.Find(rpt.host_col(probe), 1000).shape[0] == 0       # empty filter: nothing passes
    assert f.Fold() == 0
    f.close()


@pytest.mark.parametrize("dtype", [np.int8, np.uint16, np.int32, np.uint32, np.float32, np.float64, np.uint64])
def test_other_key_types(rpt, oracle, acero, dtype):
    rng = np.random.default_rng(11)
    n, m = 20_000, 60_000
    if np.issubdtype(dtype, np.floating):
        keys = rng.standard_normal(n).astype(dtype)
        probe = np.concatenate([keys[: m // 3], rng.standard_normal(m - m // 3).astype(dtype)])
        probe[::97] = np.nan
        probe[1::97] = -0.0
    else:
        info = np.iinfo(dtype)
        keys = rng.integers(info.min, info.max, n, dtype=dtype, endpoint=True)
        probe = rng.integers(info.min, info.max, m, dtype=dtype, endpoint=True)
    check(rpt, oracle, acero, n * 8, rpt.host_col(keys), n, rpt.host_col(probe), m)


def test_nullable_selected_and_composite(rpt, oracle, acero):
    rng = np.random.default_rng(5)
    n, m = 30_000, 80_000
    a = rng.integers(0, 1 << 40, n, dtype=np.int64)
    validity = rng.integers(0, 1 << 63, (n + 63) // 64, dtype=np.uint64) | rng.integers(0, 1 << 63, (n + 63) // 64,
                                                                                       dtype=np.uint64)
    strings = [b"key-%d" % (i * 7919 % 100_003) for i in range(n)]
    off, data = rpt.strings_to_arrow(strings)
    build = [rpt.host_col(a, validity=validity), rpt.host_varchar_col(off, data)]
    sel = rng.integers(0, n, m, dtype=np.uint32)
    probe = [rpt.host_col(a, sel=sel, validity=validity, data_len=n), rpt.host_varchar_col(off, data, sel=sel)]
    check(rpt, oracle, acero, n * 32, build, n, probe, m)


def test_block_range_passes(rpt, oracle, acero):
    """256 MiB filter (2^25 blocks): build and probe run in four L2-resident block-range passes."""
    n, m = 20_000_000, 4_000_000
    keys = synth_build_keys(0, n)
    probe = synth_probe_keys(0, m, n, 300_000)
    f = rpt.BlockedBloomFilter(0).CreateEmpty(n * 8)
    assert f.log_num_blocks == 25
    f.Insert(rpt.host_col(keys), n)
    assert f.launch_count == 4 + 1     # host chunk of 2^24 rows: 4 passes; the 3.2M-row rest is too small to slice
    blocks = acero.new_filter(n * 8)
    acero.insert(blocks, oracle.hash_columns([rpt.host_col(keys)], n))
    assert np.array_equal(f.ExportBlocks(), blocks)
    hp = oracle.hash_columns([rpt.host_col(probe)], m)
    assert np.array_equal(f.Find(rpt.host_col(probe), m), acero.find(blocks, hp))
    k = f.Fold()
    folded, want_k = acero.fold(blocks)
    assert k == want_k and k >= 2
    assert np.array_equal(f.ExportBlocks(), folded)
    assert np.array_equal(f.Find(rpt.host_col(probe), m), acero.find(folded, hp))
    f.close()


def test_row_range_shards_merge_to_the_single_build(rpt, torch_cuda):
    """Two partial filters built from the two halves of the build column, OR-merged, equal the single build."""
    torch = torch_cuda
    n = 3_000_001
    keys = synth_build_keys(0, n)
    whole = rpt.BlockedBloomFilter(0).CreateEmpty(n * 4)
    whole.Insert(rpt.host_col(keys), n)
    half = n // 2
    a = rpt.BlockedBloomFilter(0).CreateEmpty(n * 4)
    b = rpt.BlockedBloomFilter(0).CreateEmpty(n * 4)
    a.Insert(rpt.host_col(keys[:half]), half)
    b.Insert(rpt.host_col(keys[half:]), n - half)
    part = torch.from_numpy(b.ExportBlocks().view(np.int64)).to("cuda:0")
    a.OrParts(part, 1)
    a.Sync()
    assert np.array_equal(a.ExportBlocks(), whole.ExportBlocks())
    assert a.Fold() == whole.Fold()
    assert np.array_equal(a.ExportBlocks(), whole.ExportBlocks())
    for f in (whole, a, b):
        f.close()


def test_device_columns_and_property_at_scale(rpt, oracle, acero, torch_cuda):
    """10^7 build / 5·10^7 probe rows on the device; checker on a prefix, Fold == direct small build at full size."""
    torch = torch_cuda
    n, m = 10_000_000, 50_000_000
    dk = torch.empty(n, dtype=torch.int64, device="cuda:0")
    dp = torch.empty(m, dtype=torch.int64, device="cuda:0")
    rpt.synth_build_keys(dk, 0, n)
    rpt.synth_probe_keys(dp, 0, m, n, 100_000)
    torch.cuda.synchronize()
    big = rpt.BlockedBloomFilter(0).CreateEmpty(n * 64)        # sparse: Fold has 5-6 halvings to do
    big.Insert(rpt.device_col(dk, rpt.INT64), n)
    big.Sync()
    k = big.Fold()
    assert k >= 4
    small = rpt.BlockedBloomFilter(0).CreateEmpty((n * 64) >> k)
    assert small.log_num_blocks == big.log_num_blocks
    small.Insert(rpt.device_col(dk, rpt.INT64), n)
    small.Sync()
    folded = big.ExportBlocks()
    assert np.array_equal(folded, small.ExportBlocks()), "Fold differs from building the small filter directly"
    bits = int(np.unpackbits(folded.view(np.uint8)).sum())
    assert 4 * bits >= folded.shape[0] * 64

    dsel = torch.empty(m, dtype=torch.int32, device="cuda:0")
    dfound = torch.zeros(1, dtype=torch.int64, device="cuda:0")
    big.FindDevice(rpt.device_col(dp, rpt.INT64), m, dsel, dfound)
    big.Sync()
    found = int(dfound.item())
    sel = dsel[:found].cpu().numpy().view(np.uint32)
    assert np.all(np.diff(sel.astype(np.int64)) > 0), "selection vector is not strictly ascending"
    # every member row passes (no false negatives): members have bit 63 clear
    probe_host = dp.cpu().numpy().view(np.uint64)
    members = np.flatnonzero((probe_host >> np.uint64(63)) == 0).astype(np.uint32)
    assert np.isin(members, sel, assume_unique=True).all()
    # the checker on the first 2·10^6 probe rows
    pm = 2_000_000
    blocks = acero.new_filter(n * 64)
    acero.insert(blocks, oracle.hash_columns([rpt.host_col(synth_build_keys(0, n))], n))
    want_folded, want_k = acero.fold(blocks)
    assert want_k == k and np.array_equal(folded, want_folded)
    want = acero.find(want_folded, oracle.hash_columns([rpt.host_col(probe_host[:pm].copy())], pm))
    assert np.array_equal(sel[sel < pm], want)
    big.close()
    small.close()
