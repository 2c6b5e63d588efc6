// rpt_hash.cuh — device-side DuckDB-compatible key hashing, fused into the build / probe kernels.
//
// Bit-exact with the reference's hash path (all unsigned 64-bit integer arithmetic):
//   MurmurHash64            duckdb/src/include/duckdb/common/types/hash.hpp:36-43
//   Hash<T> (narrow ints)   hash.hpp:49-52   → MurmurHash64((uint64)(uint32_t)value)
//   Hash<int64/uint64>      hash.hpp:54-61
//   hugeint / uhugeint      duckdb/src/common/types/hash.cpp:22-30
//   float / double          hash.cpp:32-58   (−0 → +0, NaN → quiet NaN, hash the bit pattern)
//   interval                hash.cpp:60-65 + types/interval.hpp:178-193 (Normalize)
//   string                  hash.cpp:77-151  (length-seeded, 8-byte LE blocks, zero-padded tail)
//   NULL_HASH, CombineHash  duckdb/src/common/vector_operations/vector_hash.cpp:16-23,37-41
#pragma once
#include <cstdint>

#include "../../include/rpt_gpu.h"

namespace rpt {

constexpr uint64_t kMurmurK = 0xd6e8feb86659fd93ULL;
constexpr uint64_t kNullHash = 0xbf58476d1ce4e5b9ULL;
constexpr uint64_t kShiftMask = 0x3F3F3F3F3F3F3F3FULL;

__host__ __device__ __forceinline__ uint64_t murmur64(uint64_t x) {
	x ^= x >> 32;
	x *= kMurmurK;
	x ^= x >> 32;
	x *= kMurmurK;
	x ^= x >> 32;
	return x;
}

__host__ __device__ __forceinline__ uint64_t combine_hash(uint64_t a, uint64_t b) {
	a ^= a >> 32;
	a *= kMurmurK;
	return a ^ b;
}

// GetMask (duckdb/src/planner/filter/bloom_filter.cpp:29-41): bytes 4..7 of the hash, each & 63,
// select up to four bit positions of the 64-bit sector.
__host__ __device__ __forceinline__ uint64_t sector_mask_of(uint64_t h) {
	const uint32_t hi = static_cast<uint32_t>(h >> 32);
	return (1ULL << (hi & 63u)) | (1ULL << ((hi >> 8) & 63u)) | (1ULL << ((hi >> 16) & 63u)) |
	       (1ULL << ((hi >> 24) & 63u));
}

#ifdef __CUDACC__
// LookupOne (bloom_filter.cpp:72-78) without materialising the mask: (w & mask) == mask  ⇔  every one of the
// four selected bits of w is set.  Each (uint32_t)(w >> p) is one funnel shift (SHF.R.U64), so the test costs
// ≈10 integer instructions instead of ≈30 for building a 64-bit mask out of four 64-bit shifts.
__device__ __forceinline__ uint32_t sector_test(uint64_t w, uint32_t hi) {
	// one AND clears the two unused bits of every position byte; PRMT then isolates a byte in one instruction
	const uint32_t hm = hi & 0x3f3f3f3fu;
	const uint32_t p0 = __byte_perm(hm, 0u, 0x4440);
	const uint32_t p1 = __byte_perm(hm, 0u, 0x4441);
	const uint32_t p2 = __byte_perm(hm, 0u, 0x4442);
	const uint32_t p3 = hm >> 24;
	const uint32_t b = static_cast<uint32_t>(w >> p0) & static_cast<uint32_t>(w >> p1) & static_cast<uint32_t>(w >> p2) &
	                   static_cast<uint32_t>(w >> p3);
	return b & 1u;
}

// GetMask for the build: the three upper bit positions are isolated on the FMA pipe ((hi << k) >> 26 as IMAD.SHL +
// IMAD.HI), each 1 << p is one 64-bit funnel-shift pair and the four one-hot words meet in two LOP3 per half —
// ≈12 ALU-pipe instructions instead of ≈26 for the byte-wise form (the build kernel is ALU-bound: ncu r1a 76 %).
__device__ __forceinline__ uint64_t sector_mask_fast(uint32_t hi) {
	uint32_t t1, t2, t3, p1, p2, p3;
	asm("mul.lo.u32 %0, %1, 262144;" : "=r"(t1) : "r"(hi)); // << 18: byte 1 to the top, its two spare bits gone
	asm("mul.lo.u32 %0, %1, 1024;" : "=r"(t2) : "r"(hi));   // << 10
	asm("mul.lo.u32 %0, %1, 4;" : "=r"(t3) : "r"(hi));      // << 2
	asm("mul.hi.u32 %0, %1, 64;" : "=r"(p1) : "r"(t1));     // >> 26
	asm("mul.hi.u32 %0, %1, 64;" : "=r"(p2) : "r"(t2));
	asm("mul.hi.u32 %0, %1, 64;" : "=r"(p3) : "r"(t3));
	const uint32_t p0 = hi & 63u;
	return (1ULL << p0) | (1ULL << p1) | (1ULL << p2) | (1ULL << p3);
}
#endif

// One key column as the kernels see it (device pointers).
struct DevCol {
	const void *data;
	const void *aux;
	const uint32_t *sel;
	const uint64_t *validity;
	int32_t type;
	int32_t pad;
};

struct DevCols {
	DevCol c[RPT_MAX_KEY_COLS];
	int32_t n;
	int32_t pad;
};

#ifdef __CUDACC__

__device__ __forceinline__ uint64_t hash_f32(uint32_t bits) {
	if ((bits & 0x7fffffffu) == 0u) {
		bits = 0u; // ±0 → +0
	} else if ((bits & 0x7fffffffu) > 0x7f800000u) {
		bits = 0x7fc00000u; // any NaN → std::numeric_limits<float>::quiet_NaN()
	}
	return murmur64(static_cast<uint64_t>(bits));
}

__device__ __forceinline__ uint64_t hash_f64(uint64_t bits) {
	if ((bits & 0x7fffffffffffffffULL) == 0ULL) {
		bits = 0ULL;
	} else if ((bits & 0x7fffffffffffffffULL) > 0x7ff0000000000000ULL) {
		bits = 0x7ff8000000000000ULL;
	}
	return murmur64(bits);
}

__device__ __forceinline__ uint64_t hash_interval(int32_t in_months, int32_t in_days, int64_t in_micros) {
	const int64_t kMicrosPerDay = 86400000000LL;
	int64_t micros = in_micros;
	const int64_t carry_days = micros / kMicrosPerDay;
	micros -= carry_days * kMicrosPerDay;
	int64_t days = static_cast<int64_t>(in_days) + carry_days;
	const int64_t carry_months = days / 30;
	days -= carry_months * 30;
	const int64_t months = static_cast<int64_t>(in_months) + carry_months;
	return murmur64(static_cast<uint64_t>(days)) ^ murmur64(static_cast<uint64_t>(months)) ^
	       murmur64(static_cast<uint64_t>(micros));
}

// String hash over bytes [p, p+len).  The bytes are fetched as aligned 8-byte words and funnel-
// shifted, so a key costs ceil(len/8)+1 loads instead of len byte loads.  The word that holds the
// last byte may extend ≤7 bytes past the string; it is only read when it holds at least one
// needed byte, and buffers are readable up to the next 8-byte multiple (rpt_gpu.h, aux_bytes).
__device__ __forceinline__ uint64_t hash_bytes(const uint8_t *p, uint32_t len) {
	uint64_t h = 0xe17a1465ULL ^ (static_cast<uint64_t>(len) * 0xc6a4a7935bd1e995ULL);
	if (len != 0) {
		const uintptr_t addr = reinterpret_cast<uintptr_t>(p);
		const uint64_t *w = reinterpret_cast<const uint64_t *>(addr & ~static_cast<uintptr_t>(7));
		const uint32_t a = static_cast<uint32_t>(addr & 7u) * 8u; // bit offset of the first byte
		const uint32_t nwords = (static_cast<uint32_t>(addr & 7u) + len + 7u) >> 3; // aligned words spanned
		uint64_t cur = __ldg(w);
		uint32_t wi = 1;
		uint32_t remaining = len;
		while (remaining >= 8u) {
			uint64_t blk = cur >> a;
			if (a != 0) {
				const uint64_t nxt = __ldg(w + wi);
				blk |= nxt << (64u - a);
				cur = nxt;
			} else if (wi < nwords) {
				cur = __ldg(w + wi);
			}
			wi++;
			h ^= blk;
			h *= kMurmurK;
			remaining -= 8u;
		}
		if (remaining != 0u) {
			uint64_t blk = cur >> a;
			if (a != 0 && wi < nwords) {
				blk |= __ldg(w + wi) << (64u - a);
			}
			blk &= (~0ULL) >> (64u - remaining * 8u);
			h ^= blk;
			h *= kMurmurK;
		}
	}
	return murmur64(h);
}

// Internal column type (never crosses the C ABI): `data` holds the entries' 64-bit hashes.  A dictionary vector's
// entries are hashed once and the rows gather the hashes through the selection — what the reference does with a
// dictionary's cached hashes (vector_hash.cpp:484-491); same values, fewer string walks.
constexpr int kTypePrehashed = 0x7001;

// TightLoopHash body for one row (vector_hash.cpp:43-61): NULL rows hash to NULL_HASH.
__device__ __forceinline__ uint64_t hash_col_row(const DevCol &c, uint64_t row) {
	const uint64_t idx = c.sel ? static_cast<uint64_t>(__ldg(c.sel + row)) : row;
	if (c.validity && !((__ldg(c.validity + (idx >> 6)) >> (idx & 63)) & 1ULL)) {
		return kNullHash;
	}
	switch (c.type) {
	case RPT_BOOL:
	case RPT_INT8:
		return murmur64(static_cast<uint64_t>(static_cast<uint32_t>(
		    static_cast<int32_t>(__ldg(reinterpret_cast<const signed char *>(c.data) + idx)))));
	case RPT_UINT8:
		return murmur64(static_cast<uint64_t>(__ldg(reinterpret_cast<const unsigned char *>(c.data) + idx)));
	case RPT_INT16:
		return murmur64(static_cast<uint64_t>(
		    static_cast<uint32_t>(static_cast<int32_t>(__ldg(reinterpret_cast<const short *>(c.data) + idx)))));
	case RPT_UINT16:
		return murmur64(static_cast<uint64_t>(__ldg(reinterpret_cast<const unsigned short *>(c.data) + idx)));
	case RPT_INT32:
	case RPT_UINT32:
		return murmur64(static_cast<uint64_t>(__ldg(reinterpret_cast<const uint32_t *>(c.data) + idx)));
	case RPT_INT64:
	case RPT_UINT64:
		return murmur64(__ldg(reinterpret_cast<const unsigned long long *>(c.data) + idx));
	case RPT_FLOAT:
		return hash_f32(__ldg(reinterpret_cast<const uint32_t *>(c.data) + idx));
	case RPT_DOUBLE:
		return hash_f64(__ldg(reinterpret_cast<const unsigned long long *>(c.data) + idx));
	case RPT_INT128:
	case RPT_UINT128: {
		const ulonglong2 v = __ldg(reinterpret_cast<const ulonglong2 *>(c.data) + idx);
		return murmur64(v.x) ^ murmur64(v.y);
	}
	case RPT_INTERVAL: {
		const ulonglong2 v = __ldg(reinterpret_cast<const ulonglong2 *>(c.data) + idx);
		return hash_interval(static_cast<int32_t>(static_cast<uint32_t>(v.x)),
		                     static_cast<int32_t>(static_cast<uint32_t>(v.x >> 32)), static_cast<int64_t>(v.y));
	}
	case RPT_VARCHAR: {
		const uint32_t *off = reinterpret_cast<const uint32_t *>(c.data);
		const uint32_t b = __ldg(off + idx);
		const uint32_t e = __ldg(off + idx + 1);
		return hash_bytes(reinterpret_cast<const uint8_t *>(c.aux) + b, e - b);
	}
	case kTypePrehashed: // dictionary entries hashed once (launch_hash_entries); validity is already folded in
		return __ldg(reinterpret_cast<const unsigned long long *>(c.data) + idx);
	default:
		return 0;
	}
}

// HashColumns (src/bloom_filter.cpp:11-24): Hash(col0) then CombineHash for further columns.
__device__ __forceinline__ uint64_t hash_row(const DevCols &cols, uint64_t row) {
	uint64_t h = hash_col_row(cols.c[0], row);
	for (int j = 1; j < cols.n; j++) {
		h = combine_hash(h, hash_col_row(cols.c[j], row));
	}
	return h;
}

#endif // __CUDACC__

} // namespace rpt
