"""GPU suite (-m gpu): the folded pre-filter of large probes (rpt_bf_set_prefilter, DESIGN.md §4 regime 3c).

Filters of 256 / 512 MiB are probed through an L2-resident OR-fold of their slices first; only the fold's survivors
meet the real words.  The answers must not depend on it: every case here runs the same probe with the pre-filter
off (mode 0: the partitioned probe the other suites pin to the oracle), forced on (mode 2) and by density (mode 1),
compares the selection vectors element-wise, and checks the oracle's answer on a window of the rows.
"""
import numpy as np
import pytest

from helpers import synth_build_keys, synth_probe_keys

pytestmark = pytest.mark.gpu

EST_256 = 10**8          # S = 2^25 sectors = 256 MiB (4 slices)
EST_512 = 2 * 10**8      # S = 2^26 sectors = 512 MiB (8 slices)
M = (1 << 23) + 12345    # probe rows: above the fold threshold (2^22), not a multiple of a 512-row segment


@pytest.fixture(scope="module")
def torch_cuda():
    import torch
    assert torch.cuda.is_available(), "the gpu suite needs a CUDA device"
    torch.cuda.set_device(0)
    return torch


def probe_device(torch, bf, col, m):
    dsel = torch.empty(max(m, 1), dtype=torch.int32, device="cuda")
    dfound = torch.zeros(1, dtype=torch.int64, device="cuda")
    torch.cuda.synchronize()  # inputs were produced on torch's stream; the handle's stream is non-blocking
    bf.LookupSelDevice(col, m, dsel, dfound)
    bf.Sync()
    found = int(dfound.item())
    return dsel[:found].clone()


def three_modes(torch, bf, col, m, expect_auto_state=None):
    """Selection vectors with the pre-filter off / forced / by density; asserts they are identical."""
    bf.SetPrefilter(0)
    plain = probe_device(torch, bf, col, m)
    assert bf.PrefilterInfo()[0] in (0, 2)  # off: never reported as in use
    bf.SetPrefilter(2)
    forced = probe_device(torch, bf, col, m)
    state, rate = bf.PrefilterInfo()
    assert state == 1 and 0.0 <= rate <= 1.0
    bf.SetPrefilter(1)
    auto = probe_device(torch, bf, col, m)
    if expect_auto_state is not None:
        assert bf.PrefilterInfo()[0] == expect_auto_state
    assert plain.shape == forced.shape and bool((plain == forced).all()), "forced pre-filter changed the answer"
    assert plain.shape == auto.shape and bool((plain == auto).all()), "density-driven pre-filter changed the answer"
    bf.SetPrefilter(-1)
    return plain


def oracle_window(rpt, oracle, words, probe_host, sel, start, width):
    want = oracle.probe(words, [rpt.host_col(probe_host[start:start + width])], width).astype(np.int64) + start
    s64 = sel.long() & 0xFFFFFFFF
    got = s64[(s64 >= start) & (s64 < start + width)].cpu().numpy()
    assert np.array_equal(got, want)


@pytest.mark.parametrize("est,n", [(EST_256, 3_000_000), (EST_512, 5_000_000)])
def test_sparse_filter_int64(rpt, oracle, torch_cuda, est, n):
    """A sparse large filter (the fold rejects almost every non-member): 4-slice and 8-slice partitions."""
    torch = torch_cuda
    keys = synth_build_keys(0, n)
    probe = synth_probe_keys(0, M, n, 100000)
    bf = rpt.PTBloomFilter(0).Initialize(est)
    bf.Insert(rpt.host_col(keys), n)
    words = oracle.new_filter(est)
    oracle.insert(words, [rpt.host_col(keys)], n)
    assert np.array_equal(bf.ExportWords(), words)
    dp = torch.from_numpy(probe.view(np.int64)).cuda()
    sel = three_modes(torch, bf, rpt.device_col(dp, rpt.INT64), M, expect_auto_state=1)
    oracle_window(rpt, oracle, words, probe, sel, 0, 1 << 20)
    oracle_window(rpt, oracle, words, probe, sel, M - (1 << 20), 1 << 20)
    # host columns (the chunked three-stream pipeline) take the same path
    bf.SetPrefilter(2)
    host_sel = bf.LookupSel(rpt.host_col(probe), M)
    assert np.array_equal(host_sel, sel.cpu().numpy().view(np.uint32))
    bf.close()


def test_fold_follows_the_words(rpt, oracle, torch_cuda):
    """The fold is remade after every change of the words: keys inserted after a probe are found by the next one;
    a reset empties it; an imported filter is folded anew."""
    torch = torch_cuda
    n = 1_000_000
    first = synth_build_keys(0, n)
    second = synth_build_keys(n, n)
    both = np.concatenate([first, second])
    pad = np.resize(both, M)  # every probe row is a key of the first or the second batch
    dp = torch.from_numpy(pad.view(np.int64)).cuda()
    col = rpt.device_col(dp, rpt.INT64)
    bf = rpt.PTBloomFilter(0).Initialize(EST_256)
    bf.SetPrefilter(2)
    bf.Insert(rpt.host_col(first), n)
    sel1 = probe_device(torch, bf, col, M)
    assert bf.PrefilterInfo()[0] == 1
    words1 = oracle.new_filter(EST_256)
    oracle.insert(words1, [rpt.host_col(first)], n)
    assert np.array_equal(sel1.cpu().numpy().view(np.uint32), oracle.probe(words1, [rpt.host_col(pad)], M))
    bf.Insert(rpt.host_col(second), n)
    assert bf.PrefilterInfo()[0] == 0  # stale until the next large probe
    sel2 = probe_device(torch, bf, col, M)
    assert sel2.shape[0] == M  # all rows are members now
    exported = bf.ExportWords()
    bf.Reset(EST_256)
    assert probe_device(torch, bf, col, M).shape[0] == 0
    bf.ImportWords(exported)
    assert probe_device(torch, bf, col, M).shape[0] == M
    bf.close()


def test_dense_filter_all_members_and_skew(rpt, oracle, torch_cuda):
    """Config-2 density (10^8 keys in 256 MiB): forced and automatic pre-filter against the plain path on (a) the
    config-2 probe mix, (b) rows that are all members (every region as full as without the pre-filter) and (c) one
    key repeated (its region overflows: the segment is looked up on the spot)."""
    torch = torch_cuda
    n = 10**8
    dk = torch.empty(n, dtype=torch.int64, device="cuda")
    rpt.synth_build_keys(dk, 0, n)
    torch.cuda.synchronize()
    bf = rpt.PTBloomFilter(0).Initialize(n)
    bf.Insert(rpt.device_col(dk, rpt.INT64), n)
    bf.Sync()
    m = (1 << 24) + 777
    dp = torch.empty(m, dtype=torch.int64, device="cuda")
    rpt.synth_probe_keys(dp, 0, m, n, 100000)
    torch.cuda.synchronize()
    sel = three_modes(torch, bf, rpt.device_col(dp, rpt.INT64), m, expect_auto_state=1)
    state, rate = bf.PrefilterInfo()
    assert 0.04 < rate < 0.15  # ≈ 8 % of the non-members survive the fold at this density
    members = dk[:m].contiguous()
    all_sel = three_modes(torch, bf, rpt.device_col(members, rpt.INT64), m)
    assert all_sel.shape[0] == m
    one = torch.full((m,), int(dk[12345].item()), dtype=torch.int64, device="cuda")
    assert three_modes(torch, bf, rpt.device_col(one, rpt.INT64), m).shape[0] == m
    absent = torch.full((m,), -7, dtype=torch.int64, device="cuda")
    three_modes(torch, bf, rpt.device_col(absent, rpt.INT64), m)
    # exact window against the oracle (threaded build of the same 10^8 keys)
    keys = oracle.synth_build_keys(0, n)
    want = oracle.new_filter(n)
    oracle.bench_build(want, keys.view(np.int64), 8)
    W = 1 << 21
    probe = oracle.synth_probe_keys(m - W, W, n, 100000)
    want_sel = oracle.probe(want, [rpt.host_col(probe)], W).astype(np.int64) + (m - W)
    s64 = sel.long() & 0xFFFFFFFF
    assert np.array_equal(s64[s64 >= m - W].cpu().numpy(), want_sel)
    bf.close()


def test_other_key_sources(rpt, oracle, torch_cuda):
    """int32, nullable int64, two-column and VARCHAR keys through the pre-filtered partition (each key source is
    its own kernel instantiation)."""
    torch = torch_cuda
    rng = np.random.default_rng(2025)
    n, m = 2_000_000, (1 << 22) + 4321
    # int32
    k32 = rng.integers(-2**31, 2**31 - 1, size=n, dtype=np.int64).astype(np.int32)
    p32 = np.concatenate([k32[: m // 4], rng.integers(-2**31, 2**31 - 1, size=m - m // 4, dtype=np.int64).astype(np.int32)])
    bf = rpt.PTBloomFilter(0).Initialize(EST_256)
    bf.Insert(rpt.host_col(k32), n)
    d32 = torch.from_numpy(p32).cuda()
    sel = three_modes(torch, bf, rpt.device_col(d32, rpt.INT32), m)
    words = oracle.new_filter(EST_256)
    oracle.insert(words, [rpt.host_col(k32)], n)
    assert np.array_equal(sel.cpu().numpy().view(np.uint32), oracle.probe(words, [rpt.host_col(p32)], m))
    bf.close()
    # nullable int64 (NULL rows hash to NULL_HASH on both sides)
    k64 = synth_build_keys(0, n)
    p64 = synth_probe_keys(0, m, n, 250000)
    kval = rng.integers(0, 2**64, size=(n + 63) // 64, dtype=np.uint64)
    pval = rng.integers(0, 2**64, size=(m + 63) // 64, dtype=np.uint64)
    bf = rpt.PTBloomFilter(0).Initialize(EST_256)
    bf.Insert(rpt.host_col(k64, validity=kval), n)
    dp = torch.from_numpy(p64.view(np.int64)).cuda()
    dv = torch.from_numpy(pval.view(np.int64)).cuda()
    sel = three_modes(torch, bf, rpt.device_col(dp, rpt.INT64, validity=dv), m)
    words = oracle.new_filter(EST_256)
    oracle.insert(words, [rpt.host_col(k64, validity=kval)], n)
    assert np.array_equal(sel.cpu().numpy().view(np.uint32),
                          oracle.probe(words, [rpt.host_col(p64, validity=pval)], m))
    bf.close()
    # two key columns (generic source)
    a, b = synth_build_keys(0, n), rng.integers(0, 1000, size=n, dtype=np.int64).astype(np.int32)
    pa = synth_probe_keys(0, m, n, 500000)
    pb = np.resize(b, m)
    bf = rpt.PTBloomFilter(0).Initialize(EST_256)
    bf.Insert([rpt.host_col(a), rpt.host_col(b)], n)
    da, db = torch.from_numpy(pa.view(np.int64)).cuda(), torch.from_numpy(pb).cuda()
    sel = three_modes(torch, bf, [rpt.device_col(da, rpt.INT64), rpt.device_col(db, rpt.INT32)], m)
    words = oracle.new_filter(EST_256)
    oracle.insert(words, [rpt.host_col(a), rpt.host_col(b)], n)
    assert np.array_equal(sel.cpu().numpy().view(np.uint32),
                          oracle.probe(words, [rpt.host_col(pa), rpt.host_col(pb)], m))
    bf.close()
    # VARCHAR (hashed ahead, then the pre-hashed flat source)
    txt = [str(int(v)).encode() for v in synth_build_keys(0, 300000).view(np.int64)]
    ptxt = [str(int(v)).encode() for v in synth_probe_keys(0, 300000, 300000, 300000).view(np.int64)]
    reps = (m + len(ptxt) - 1) // len(ptxt)
    ptxt = (ptxt * reps)[:m]

    def arrow(strings):
        lens = np.fromiter((len(s) for s in strings), dtype=np.uint32, count=len(strings))
        offs = np.zeros(len(strings) + 1, dtype=np.uint32)
        np.cumsum(lens, out=offs[1:])
        return offs, np.frombuffer(b"".join(strings), dtype=np.uint8).copy()

    koffs, kblob = arrow(txt)
    poffs, pblob = arrow(ptxt)
    bf = rpt.PTBloomFilter(0).Initialize(EST_256)
    bf.Insert(rpt.host_varchar_col(koffs, kblob), len(txt))
    d_off = torch.from_numpy(poffs.view(np.int32)).cuda()
    d_blob = torch.from_numpy(np.concatenate([pblob, np.zeros(8, dtype=np.uint8)])).cuda()
    col = rpt.device_col(d_off, rpt.VARCHAR, aux=d_blob, aux_bytes=pblob.size, data_len=m)
    sel = three_modes(torch, bf, col, m)
    words = oracle.new_filter(EST_256)
    oracle.insert(words, [rpt.host_varchar_col(koffs, kblob)], len(txt))
    assert np.array_equal(sel.cpu().numpy().view(np.uint32),
                          oracle.probe(words, [rpt.host_varchar_col(poffs, pblob)], m))
    bf.close()


def test_concurrent_probes_share_one_fold(rpt, oracle, torch_cuda):
    """Eight host threads probe one large filter at the same time, each on its own stream (the reference's
    ExecuteInternal threading contract): the first one to arrive folds the filter, the others wait for the fold's
    event; every thread gets the selection vector of the plain path."""
    import threading
    torch = torch_cuda
    n = 4_000_000
    keys = synth_build_keys(0, n)
    bf = rpt.PTBloomFilter(0).Initialize(EST_256)
    bf.Insert(rpt.host_col(keys), n)
    m = (1 << 22) + 99
    probes, cols, want = [], [], []
    for t in range(8):
        p = synth_probe_keys(t * m, m, n, 200000)
        d = torch.from_numpy(p.view(np.int64)).cuda()
        probes.append(d)
        cols.append(rpt.device_col(d, rpt.INT64))
    bf.SetPrefilter(0)
    for t in range(8):
        want.append(probe_device(torch, bf, cols[t], m))
    bf.SetPrefilter(1)
    bf.Insert(rpt.host_col(keys[:1000]), 1000)  # same words, new version: the fold has to be made again
    torch.cuda.synchronize()
    got, errors = [None] * 8, []

    def work(t):
        try:
            s = torch.cuda.Stream()
            dsel = torch.empty(m, dtype=torch.int32, device="cuda")
            dfound = torch.zeros(1, dtype=torch.int64, device="cuda")
            torch.cuda.synchronize()
            for _ in range(3):
                bf.LookupSelDevice(cols[t], m, dsel, dfound, s.cuda_stream)
            s.synchronize()
            got[t] = dsel[:int(dfound.item())].clone()
        except Exception as e:  # noqa: BLE001 - reported below
            errors.append(repr(e))

    threads = [threading.Thread(target=work, args=(t,)) for t in range(8)]
    for th in threads:
        th.start()
    for th in threads:
        th.join()
    assert not errors, errors
    assert bf.PrefilterInfo()[0] == 1
    for t in range(8):
        assert got[t].shape == want[t].shape and bool((got[t] == want[t]).all()), t
    bf.close()
