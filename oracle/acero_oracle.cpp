// TEST INFRASTRUCTURE — CPU checker for the README-described filter variant (rpt_abf_*).  Never linked into
// the product; only tests/ may load it.
//
// PARITY UNPINNED.  The reference's README.md:23-32 describes an Apache Arrow Acero BlockedBloomFilter port
// (1024-entry table of 57-bit masks, 8 bits per key, Fold), but the reference tree holds no source, test or
// golden vector for it.  The algorithm therefore comes from a third-party project that is absent from
// /root/reference: Apache Arrow, cpp/src/arrow/acero/bloom_filter.{h,cc} ("Arrow 21+" per README.md:25).  This
// file restates that published algorithm in scalar C++:
//   BloomFilterMasks::BloomFilterMasks()   table generation (mt19937 seeded with eight zeros)
//   BloomFilterMasks::mask(bit_offset)     57-bit window of the bit vector
//   BlockedBloomFilter::mask / block_id    hash bits 0-9 pick the mask, 10-15 rotate it, 16+ pick the block
//   BlockedBloomFilter::CreateEmpty        8 bits per key, min 512 bits, power of two
//   BlockedBloomFilter::Insert / Find      block |= mask ; (block & mask) == mask
//   BlockedBloomFilter::Fold / SingleFold  OR 2^k slices into the first while < 1/4 of the bits are set
// Hashes are inputs here (the DuckDB key hashes come from rpt_oracle.c, which IS pinned against the reference).
#include <cstdint>
#include <cstring>
#include <random>

namespace {
const int BITS_PER_MASK = 57;
const int MIN_SET = 4;
const int MAX_SET = 5;
const int LOG_MASKS = 10;
const int NUM_MASKS = 1 << LOG_MASKS;
const int TABLE_BYTES = (NUM_MASKS + 64) / 8;
const uint64_t WINDOW = (uint64_t(1) << BITS_PER_MASK) - 1;

struct Table {
	uint8_t bytes[TABLE_BYTES + 8];
	Table() {
		std::memset(bytes, 0, sizeof(bytes));
		std::seed_seq seq {0, 0, 0, 0, 0, 0, 0, 0};
		std::mt19937 gen(seq);
		std::uniform_int_distribution<uint64_t> draw;
		auto pick = [&](int lo, int hi) -> uint64_t { return lo + draw(gen) % uint64_t(hi - lo + 1); };
		auto test = [&](int64_t b) { return (bytes[b / 8] >> (b % 8)) & 1; };
		auto put = [&](int64_t b) { bytes[b / 8] |= uint8_t(1 << (b % 8)); };

		int live = int(pick(MIN_SET, MAX_SET));
		for (int k = 0; k < live; k++) {
			for (;;) {
				int pos = int(pick(0, BITS_PER_MASK - 1));
				if (!test(pos)) {
					put(pos);
					break;
				}
			}
		}
		for (int64_t b = BITS_PER_MASK; b < NUM_MASKS + BITS_PER_MASK - 1; b++) {
			int out_bit = test(b - BITS_PER_MASK);
			if (out_bit && live == MIN_SET) {
				put(b);
			} else if (!out_bit && live == MAX_SET) {
				// stays clear
			} else if (pick(0, 2 * BITS_PER_MASK - 1) < uint64_t(MIN_SET + MAX_SET)) {
				put(b);
				live += out_bit ? 0 : 1;
			} else {
				live -= out_bit ? 1 : 0;
			}
		}
	}
	uint64_t window(int offset) const {
		uint64_t raw;
		std::memcpy(&raw, bytes + offset / 8, sizeof(raw));
		return (raw >> (offset % 8)) & WINDOW;
	}
};

const Table &table() {
	static const Table t;
	return t;
}

inline uint64_t mask_of(uint64_t h) {
	uint64_t m = table().window(int(h & (NUM_MASKS - 1)));
	int rot = int((h >> LOG_MASKS) & 63);
	return rot ? (m << rot) | (m >> (64 - rot)) : m;
}
inline uint64_t block_of(uint64_t h, int log_blocks) {
	return (h >> (LOG_MASKS + 6)) & ((uint64_t(1) << log_blocks) - 1);
}
} // namespace

extern "C" {

void acero_mask_bits(uint8_t *out136) {
	std::memcpy(out136, table().bytes, TABLE_BYTES);
}

uint64_t acero_mask(uint64_t hash) {
	return mask_of(hash);
}

int acero_log_blocks(uint64_t rows) {
	uint64_t want = rows * 8 < 512 ? 512 : rows * 8;
	int log_bits = 0;
	while ((uint64_t(1) << log_bits) < want) {
		log_bits++;
	}
	return log_bits - 6;
}

void acero_insert(uint64_t *blocks, int log_blocks, const uint64_t *hashes, uint64_t n) {
	for (uint64_t i = 0; i < n; i++) {
		blocks[block_of(hashes[i], log_blocks)] |= mask_of(hashes[i]);
	}
}

uint64_t acero_find(const uint64_t *blocks, int log_blocks, const uint64_t *hashes, uint64_t n, uint32_t *sel) {
	uint64_t found = 0;
	for (uint64_t i = 0; i < n; i++) {
		uint64_t m = mask_of(hashes[i]);
		if ((blocks[block_of(hashes[i], log_blocks)] & m) == m) {
			sel[found++] = uint32_t(i);
		}
	}
	return found;
}

// Returns the number of halvings; *log_blocks is updated; the surviving blocks are blocks[0, 2^*log_blocks).
int acero_fold(uint64_t *blocks, int *log_blocks) {
	int halvings = 0;
	for (;;) {
		if (*log_blocks <= 4) {
			break;
		}
		const int64_t nblocks = int64_t(1) << *log_blocks;
		const int64_t nbits = nblocks * 64;
		int64_t set = 0;
		for (int64_t i = 0; i < nblocks; i++) {
			set += __builtin_popcountll(blocks[i]);
		}
		if (4 * set >= nbits) {
			break;
		}
		int k = 1;
		while ((*log_blocks - k) > 4 && 4 * set < (nbits >> k)) {
			k++;
		}
		const int64_t slice = nblocks >> k;
		for (int64_t s = 1; s < (int64_t(1) << k); s++) {
			for (int64_t i = 0; i < slice; i++) {
				blocks[i] |= blocks[s * slice + i];
			}
		}
		*log_blocks -= k;
		halvings += k;
	}
	return halvings;
}

} // extern "C"
