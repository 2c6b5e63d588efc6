"""rpt_bf_probe_u64_noalloc (include/rpt_gpu.h) against rpt_bf_probe through the C ABI: bit-exact selection vectors
on two calls in a row (the second smaller and with another selectivity, so stale scratch of the first would show),
the stated limit, and the empty input.  The entry point has no counterpart in the reference (added
for callers that cannot call cudaMalloc).  Collected last on purpose (file name)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def torch_cuda():
    import torch
    assert torch.cuda.is_available(), "the gpu suite needs a CUDA device"
    torch.cuda.set_device(0)
    return torch


def probe_both(rpt, torch, bf, dp, m):
    """(sel, found) of the existing entry point and of the allocation-free one on the same device column."""
    dsel = torch.empty(max(m, 1), dtype=torch.int32, device="cuda")
    dfound = torch.zeros(1, dtype=torch.int64, device="cuda")
    torch.cuda.synchronize()  # inputs were produced on torch's stream; the handle's stream is non-blocking
    bf.LookupSelDevice(rpt.device_col(dp, rpt.INT64), m, dsel, dfound)
    bf.Sync()
    want_found = int(dfound.item())
    want = dsel[:want_found].cpu().numpy().view(np.uint32).copy()
    nsel = torch.full((max(m, 1),), -1, dtype=torch.int32, device="cuda")
    nfound = torch.full((1,), -7, dtype=torch.int64, device="cuda")
    torch.cuda.synchronize()
    free_before, _ = torch.cuda.mem_get_info()
    bf.LookupSelNoAlloc(dp, m, nsel, nfound)  # stream NULL: ordered with torch's default stream
    torch.cuda.synchronize()
    free_after, _ = torch.cuda.mem_get_info()
    got_found = int(nfound.item())
    got = nsel[:got_found].cpu().numpy().view(np.uint32).copy()
    return want, got, free_before, free_after


def test_noalloc_probe_matches_probe_on_two_calls_in_a_row(rpt, oracle, torch_cuda):
    torch = torch_cuda
    n = 300_000
    dk = torch.empty(n, dtype=torch.int64, device="cuda")
    rpt.synth_build_keys(dk, 0, n)
    bf = rpt.PTBloomFilter(0).Initialize(n)
    torch.cuda.synchronize()
    bf.Insert(rpt.device_col(dk, rpt.INT64), n)
    bf.Sync()
    calls = [(5_000_003, 500_000), (1_234_567, 20_000)]  # (rows, selectivity in ppm): ragged tails, 50 % then 2 %
    for i, (m, ppm) in enumerate(calls):
        dp = torch.empty(m, dtype=torch.int64, device="cuda")
        rpt.synth_probe_keys(dp, 17 * i, m, n, ppm)
        want, got, free_before, free_after = probe_both(rpt, torch, bf, dp, m)
        assert want.shape[0] > 0
        assert np.array_equal(got, want), f"call {i}"
        if i > 0:  # the first call may load the kernels' module; after that the device's free memory must not move
            assert free_after == free_before
    # the oracle on the last batch as well (host copy of the same keys)
    keys = dp.cpu().numpy()
    words = oracle.new_filter(n)
    oracle.insert(words, [rpt.host_col(dk.cpu().numpy())], n)
    assert np.array_equal(got, oracle.probe(words, [rpt.host_col(keys)], m))
    bf.close()


def test_noalloc_probe_limit_and_empty_input(rpt, torch_cuda):
    torch = torch_cuda
    n = 1000
    dk = torch.arange(n, dtype=torch.int64, device="cuda")
    bf = rpt.PTBloomFilter(0).Initialize(n)
    torch.cuda.synchronize()
    bf.Insert(rpt.device_col(dk, rpt.INT64), n)
    bf.Sync()
    dsel = torch.empty(16, dtype=torch.int32, device="cuda")
    dfound = torch.full((1,), 99, dtype=torch.int64, device="cuda")
    with pytest.raises(RuntimeError, match="RPT_NOALLOC_MAX_ROWS"):
        bf.LookupSelNoAlloc(dk, (1 << 24) + 1, dsel, dfound)  # rejected before anything is launched
    bf.LookupSelNoAlloc(dk, 0, dsel, dfound)
    torch.cuda.synchronize()
    assert int(dfound.item()) == 0
    # exactly the maximum is accepted: members only, every row survives
    m = 1 << 24
    dp = torch.arange(m, dtype=torch.int64, device="cuda") % n
    big = torch.empty(m, dtype=torch.int32, device="cuda")
    bf.LookupSelNoAlloc(dp, m, big, dfound)
    torch.cuda.synchronize()
    assert int(dfound.item()) == m
    assert bool((big == torch.arange(m, dtype=torch.int32, device="cuda")).all())
    bf.close()
