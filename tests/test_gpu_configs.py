"""GPU suite (-m gpu): the hot path driven in the SHAPES of BASELINE.json's configs 1, 3 and 4 (reduced sizes), the
way the reference's operators drive it — CREATE_BF: Initialize(estimate) → Insert per chunk → Finalize (resize
rule, physical_create_bf.cpp:386-406); USE_BF: one LookupSel per filter, each on the survivors of the previous one
(`input.Slice(sel, result_count)`, physical_use_bf.cpp:137-190) — and compared with the CPU oracle doing the same.
Bit-exact: filter words and every intermediate selection vector."""
import numpy as np
import pytest

from helpers import synth_build_keys

pytestmark = pytest.mark.gpu

CHUNK = 2048  # STANDARD_VECTOR_SIZE


def decimal_text(keys):
    """k2 = decimal text of k1, Arrow layout."""
    strs = [b"%d" % int(k) for k in keys]
    lens = np.fromiter((len(s) for s in strs), dtype=np.uint64, count=len(strs))
    offsets = np.zeros(len(strs) + 1, dtype=np.uint32)
    np.cumsum(lens, out=offsets[1:], dtype=np.uint64)
    return offsets, np.frombuffer(b"".join(strs), dtype=np.uint8).copy()


def create_bf(rpt, oracle, est, make_col, n):
    """CREATE_BF on both engines: per-chunk Insert into a filter sized for `est`, then Finalize(actual rows) with the
    whole column at hand for the rebuild.  Returns (gpu filter, oracle words)."""
    bf = rpt.PTBloomFilter(0).Initialize(est)
    words = oracle.new_filter(est)
    for r0 in range(0, n, CHUNK):
        r1 = min(r0 + CHUNK, n)
        bf.Insert(make_col(r0, r1), r1 - r0)
        oracle.insert(words, [make_col(r0, r1)], r1 - r0)
    rebuilt = bf.Finalize(n, make_col(0, n), n)
    assert rebuilt == oracle.needs_rebuild(est, n)
    if rebuilt:
        words = oracle.new_filter(n)
        oracle.insert(words, [make_col(0, n)], n)
    assert bf.finalized_ and not bf.IsEmpty()
    assert bf.num_sectors == words.shape[0]
    assert np.array_equal(bf.ExportWords(), words)
    return bf, words


def test_config1_readme_tables(rpt, oracle):
    """README example: t1 = range(100000), t2 = range(50000) on id (count = 50000): the filter built on t2.id
    (S = 2^14, shared-memory regime) keeps every matching t1 row and the known number of false positives."""
    t1 = np.arange(100000, dtype=np.int64)
    t2 = np.arange(50000, dtype=np.int64)
    bf, words = create_bf(rpt, oracle, t2.size, lambda a, b: rpt.host_col(t2[a:b]), t2.size)
    assert bf.num_sectors == 1 << 14
    got = bf.LookupSel(rpt.host_col(t1), t1.size)
    want = oracle.probe(words, [rpt.host_col(t1)], t1.size)
    assert np.array_equal(got, want)
    assert np.array_equal(got[:50000], np.arange(50000, dtype=np.uint32))  # no false negatives: the join's 50000 rows
    bf.close()


@pytest.mark.parametrize("under_estimate", [False, True])
def test_config3_two_filters_int64_and_varchar(rpt, oracle, under_estimate):
    """Config 3 shape: build rows (k1 int64, k2 varchar = decimal text of k1) → TWO single-column filters (what the
    reference builds); run once with a good estimate and once with est = rows/5, which makes the
    resize rule rebuild both filters.  USE_BF: filter 1 on k1, then filter 2 on k2 of the survivors only."""
    n, m = 60_000, 300_000
    k1 = synth_build_keys(0, n).view(np.int64)
    off, blob = decimal_text(k1)
    est = n // 5 if under_estimate else n
    bf1, w1 = create_bf(rpt, oracle, est, lambda a, b: rpt.host_col(k1[a:b]), n)
    bf2, w2 = create_bf(rpt, oracle, est,
                        lambda a, b: rpt.host_varchar_col(off[a:b + 1] - off[a], blob[off[a]:off[b]]), n)
    rng = np.random.default_rng(3)
    member = rng.random(m) < 0.10
    p1 = np.where(member, k1[rng.integers(0, n, m)], rng.integers(0, 2**62, m) | (1 << 62)).astype(np.int64)
    # the varchar column mostly agrees with k1, sometimes not (a row can pass filter 1 and fail filter 2)
    p2_keys = np.where(rng.random(m) < 0.9, p1, p1 + 1)
    poff, pblob = decimal_text(p2_keys)
    total = 0
    for r0 in range(0, m, 50 * CHUNK):  # USE_BF over batches of chunks
        r1 = min(r0 + 50 * CHUNK, m)
        cnt = r1 - r0
        c1 = rpt.host_col(p1[r0:r1])
        sel1 = bf1.LookupSel(c1, cnt)
        assert np.array_equal(sel1, oracle.probe(w1, [c1], cnt))
        # input.Slice(sel1): the key vector of filter 2 is the varchar column seen through sel1
        lo, lb = poff[r0:r1 + 1] - poff[r0], pblob[poff[r0]:poff[r1]]
        c2 = rpt.host_varchar_col(lo, lb, sel=sel1)
        sel2 = bf2.LookupSel(c2, sel1.size)
        assert np.array_equal(sel2, oracle.probe(w2, [c2], sel1.size))
        final = sel1[sel2]  # chunk-relative row ids that reach the join
        true_rows = np.nonzero(member[r0:r1] & (p2_keys[r0:r1] == p1[r0:r1]))[0]
        assert np.isin(true_rows, final).all()  # no false negatives through the cascade
        total += final.size
    assert 0 < total < m
    bf1.close()
    bf2.close()


def test_config4_star_join_cascade(rpt, oracle):
    """Config 4 shape (5-way star, TPC-H-like): a fact table with four foreign keys, one filter per dimension table
    (different sizes: shared-memory, L2-resident; different key types), cascaded over the fact rows in plan order.
    Every intermediate selection vector equals the oracle's; the surviving set contains every true join row."""
    rng = np.random.default_rng(44)
    dims = [  # (rows kept after the dimension's own predicate, key domain, dtype)
        (20_000, 200_000, np.int32),      # part-like: 10 % of the domain
        (150_000, 300_000, np.int64),     # customer-like: 50 %
        (5, 25, np.int8),                 # nation-like: tiny
        (400_000, 1_500_000, np.int64),   # orders-like
    ]
    m = 1_000_000
    filters, fact_cols, kept_sets = [], [], []
    for rows, domain, dt in dims:
        kept = rng.choice(domain, size=rows, replace=False).astype(dt)
        bf, words = create_bf(rpt, oracle, rows, lambda a, b, k=kept: rpt.host_col(k[a:b]), rows)
        filters.append((bf, words))
        fact_cols.append(rng.integers(0, domain, m).astype(dt))
        kept_sets.append(kept)
    sel = None  # rows of the fact batch still alive, ascending
    for (bf, words), col in zip(filters, fact_cols):
        keycol = rpt.host_col(col) if sel is None else rpt.host_col(col, sel=sel, data_len=m)
        cnt = m if sel is None else sel.size
        s = bf.LookupSel(keycol, cnt)
        assert np.array_equal(s, oracle.probe(words, [keycol], cnt))
        sel = s if sel is None else sel[s]
        assert (np.diff(sel.astype(np.int64)) > 0).all()
    joined = np.ones(m, dtype=bool)
    for col, kept in zip(fact_cols, kept_sets):
        joined &= np.isin(col, kept)
    assert np.isin(np.nonzero(joined)[0], sel).all()
    assert sel.size < m // 20  # the cascade prunes most of the fact table before the joins
    for bf, _ in filters:
        bf.close()


def test_config3_full_size_against_oracle(rpt, oracle):
    """BASELINE config 3 at full size: 5·10^7 build rows (k1 int64, k2 = decimal text of k1) → two
    single-column 128 MiB filters, built once with est_rows = 5·10^7 and once with est_rows = 10^7 (the resize rule
    of PhysicalCreateBF::Finalize, physical_create_bf.cpp:386-406, re-allocates and re-hashes both,
    src/bloom_filter.cpp:34-58); 5·10^8 probe rows at 10 % members, filter 2 on filter 1's survivors.
    Words of both filters equal the oracle's; the selection vectors equal the oracle's on two 2·10^7-row windows;
    size-independent properties (ascending, every member kept) on the whole."""
    import ctypes as C
    import torch
    assert torch.cuda.is_available()
    n, m, ppm = 50_000_000, 500_000_000, 100_000
    dev = "cuda:0"
    k1 = torch.empty(n, dtype=torch.int64, device=dev)
    rpt.synth_build_keys(k1, 0, n)
    torch.cuda.synchronize()
    ko, kb, kbytes = rpt.synth_decimal_text(k1)
    torch.cuda.synchronize()
    c_k1 = rpt.device_col(k1, rpt.INT64)
    c_k2 = rpt.device_col(ko, rpt.VARCHAR, aux=kb, aux_bytes=kbytes, data_len=n)
    # the oracle's copy of the build side (the generated text is spot-checked against Python's formatting)
    h_k1 = k1.cpu().numpy()
    h_ko, h_kb = ko.cpu().numpy().view(np.uint32), kb.cpu().numpy()[:kbytes]
    for i in (0, 1, 12345, n - 1):
        assert h_kb[h_ko[i]:h_ko[i + 1]].tobytes() == b"%d" % int(h_k1[i])
    w1 = oracle.new_filter(n)
    oracle.bench_build(w1, h_k1, 8)
    w2 = oracle.new_filter(n)
    oracle.insert(w2, [rpt.host_varchar_col(h_ko, h_kb)], n)

    f1, f2 = rpt.PTBloomFilter(0), rpt.PTBloomFilter(0)
    for est in (n, n // 5):
        f1.Initialize(est)
        f2.Initialize(est)
        f1.Insert(c_k1, n)
        f2.Insert(c_k2, n)
        rebuilt = (f1.Finalize(n, c_k1, n), f2.Finalize(n, c_k2, n))
        assert rebuilt == (est != n, est != n) and rebuilt[0] == oracle.needs_rebuild(est, n)
        assert f1.num_sectors == f2.num_sectors == 1 << 24
        assert np.array_equal(f1.ExportWords(), w1), est
        assert np.array_equal(f2.ExportWords(), w2), est
    del h_k1, h_ko, h_kb

    p1 = torch.empty(m, dtype=torch.int64, device=dev)
    rpt.synth_probe_keys(p1, 0, m, n, ppm)
    torch.cuda.synchronize()
    BATCH = 200_000_000  # 32-bit VARCHAR offsets: a call carries < 4 GiB of string bytes
    sel1 = torch.empty(BATCH, dtype=torch.int32, device=dev)
    sel2 = torch.empty(BATCH, dtype=torch.int32, device=dev)
    found = torch.zeros(2, dtype=torch.int64, device=dev)
    all1, all2, texts = [], [], {}
    for r0 in range(0, m, BATCH):
        r1 = min(r0 + BATCH, m)
        mb = r1 - r0
        po, pb, pbytes = rpt.synth_decimal_text(p1[r0:r1])
        texts[r0] = (po, pb)
        c_p1 = rpt.device_col(p1[r0:r1], rpt.INT64)
        c_p2 = rpt.device_col(po, rpt.VARCHAR, aux=pb, aux_bytes=pbytes, data_len=mb)
        torch.cuda.synchronize()
        f1.LookupSelDevice(c_p1, mb, sel1, found[0:1])
        f1.Sync()
        n1 = int(found[0].item())
        rpt._check(rpt.lib().rpt_bf_filter_sel(f2._h, C.byref(c_p2.raw), sel1.data_ptr(), n1, sel2.data_ptr(),
                                               found[1:2].data_ptr(), rpt.MEM_DEVICE, None))
        f2.Sync()
        n2 = int(found[1].item())
        all1.append((sel1[:n1].long() & 0xFFFFFFFF) + r0)
        all2.append((sel2[:n2].long() & 0xFFFFFFFF) + r0)
    s1, s2 = torch.cat(all1), torch.cat(all2)
    n1, n2 = s1.numel(), s2.numel()
    assert bool((s1[1:] > s1[:-1]).all()) and bool((s2[1:] > s2[:-1]).all())
    member = p1 >= 0
    n_member = int(member.sum().item())
    assert int(member[s2].sum().item()) == n_member          # no false negatives through the cascade
    assert n_member <= n2 <= n1 < n_member + (m - n_member) // 200
    del member
    W = 2 * 10**7
    for start in (0, m - W):
        r0 = start // BATCH * BATCH
        po, pb = texts[r0]
        a0 = start - r0
        hp1 = p1[start:start + W].cpu().numpy()
        b0, b1 = int(po[a0].item()), int(po[a0 + W].item())
        hoff = (po[a0:a0 + W + 1].cpu().numpy().view(np.uint32).astype(np.int64) - b0).astype(np.uint32)
        hblob = pb[b0:b1].cpu().numpy()
        c1 = rpt.host_col(hp1)
        want1 = oracle.probe(w1, [c1], W)
        got1 = (s1[(s1 >= start) & (s1 < start + W)] - start).cpu().numpy().astype(np.uint32)
        assert np.array_equal(got1, want1), start
        c2 = rpt.host_varchar_col(hoff, hblob, sel=want1)
        want2 = want1[oracle.probe(w2, [c2], want1.size)]   # BFTableFilter::Filter translates through the incoming sel
        got2 = (s2[(s2 >= start) & (s2 < start + W)] - start).cpu().numpy().astype(np.uint32)
        assert np.array_equal(got2, want2), start
    f1.close()
    f2.close()
