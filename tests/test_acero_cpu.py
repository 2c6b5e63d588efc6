"""CPU suite for the README-described filter variant (rpt_abf_*, reference README.md:23-32).

PARITY UNPINNED: the reference holds no code, test or vector for this filter, so these tests check the
checker (oracle/acero_oracle.cpp) through properties of the published Arrow algorithm, and the host-side pieces of
the product (mask table, sizing) against the checker.  No GPU compute here.
"""
import hashlib

import numpy as np

from helpers import splitmix


def hashes(first, count):
    return splitmix(np.arange(first, first + count, dtype=np.uint64))


def test_mask_table_windows_hold_4_or_5_bits(acero):
    bits = np.unpackbits(acero.mask_bits(), bitorder="little")
    counts = [int(bits[i:i + 57].sum()) for i in range(1024)]
    assert min(counts) == 4 and max(counts) == 5
    assert not bits[1024 + 56:].any()          # nothing beyond the last window
    # the table is deterministic (mt19937 seeded with eight zeros through libstdc++'s uniform_int_distribution);
    # self-generated fingerprint, NOT a reference vector
    assert hashlib.sha256(acero.mask_bits().tobytes()).hexdigest()[:16] == TABLE_SHA16


def test_product_mask_table_equals_checker(rpt, acero):
    assert np.array_equal(rpt.acero_mask_bits(), acero.mask_bits())


def test_mask_is_rotated_window(acero):
    bits = np.unpackbits(acero.mask_bits(), bitorder="little").astype(np.uint64)
    for h in [0, 1, 1023, 1024, 0xFFFF, 0x123456789ABCDEF0, (1 << 64) - 1]:
        mid, rot = h & 1023, (h >> 10) & 63
        w = sum(int(bits[mid + j]) << j for j in range(57))
        want = ((w << rot) | (w >> (64 - rot))) & ((1 << 64) - 1) if rot else w
        assert acero.mask(h) == want
        assert bin(want).count("1") in (4, 5)


def test_sizing(rpt, acero):
    for rows in [0, 1, 63, 64, 65, 1000, 50_000, 100_000, 10 ** 7, 10 ** 8, 3 * 10 ** 8, (1 << 28) + 1]:
        log = acero.log_blocks(rows)
        assert (1 << (log + 6)) >= max(512, 8 * rows) and (log == 3 or (1 << (log + 5)) < 8 * rows)
        assert rpt.acero_log_blocks_for_rows(rows) == log


def test_no_false_negatives_and_fold_keeps_members(acero):
    n = 20_000
    h = hashes(0, n)
    blocks = acero.new_filter(n * 64)              # 64x oversized: sparse, so Fold has work to do
    acero.insert(blocks, h)
    assert acero.find(blocks, h).shape[0] == n
    before = acero.find(blocks, hashes(10 ** 6, 50_000)).shape[0]
    folded, k = acero.fold(blocks.copy())
    assert k >= 4
    log = folded.shape[0].bit_length() - 1
    set_bits = int(np.unpackbits(folded.view(np.uint8)).sum())
    assert log <= 4 or 4 * set_bits >= folded.shape[0] * 64      # Fold's stop condition
    assert 4 * set_bits < 2 * folded.shape[0] * 64 * 2            # and it did not overshoot wildly
    assert acero.find(folded, h).shape[0] == n
    assert acero.find(folded, hashes(10 ** 6, 50_000)).shape[0] >= before


def test_fold_equals_building_the_small_filter_directly(acero):
    # block_id keeps the low bits of (hash >> 16), so OR-ing the halves is the filter a smaller build would give
    n = 5_000
    h = hashes(77, n)
    big = acero.new_filter(n * 128)
    acero.insert(big, h)
    folded, _ = acero.fold(big.copy())
    small = np.zeros(folded.shape[0], dtype=np.uint64)
    acero.insert(small, h)
    assert np.array_equal(folded, small)


def test_fold_leaves_dense_and_tiny_filters_alone(acero):
    h = hashes(5, 4096)
    blocks = acero.new_filter(4096)                # 8 bits per key with 4-5 bits per key set: > 25 % dense
    acero.insert(blocks, h)
    folded, k = acero.fold(blocks.copy())
    assert k == 0 and np.array_equal(folded, blocks)
    tiny = acero.new_filter(0)                     # 512 bits = 2^3 blocks, below Fold's floor of 2^4
    folded, k = acero.fold(tiny.copy())
    assert k == 0 and folded.shape[0] == 8


TABLE_SHA16 = "16a857dcff60ee39"
