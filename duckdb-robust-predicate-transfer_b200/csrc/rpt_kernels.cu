// rpt_kernels.cu — hand-written sm_100a kernels for the rpt bloom-filter hot path.
//
//   build_kernel   PTBloomFilter::Insert  = HashColumns + BloomFilter::InsertHashes
//                  (reference src/bloom_filter.cpp:70-78, duckdb/src/planner/filter/bloom_filter.cpp:43-48,63-70)
//   probe_kernel   PTBloomFilter::LookupSel = HashColumns + BloomFilter::LookupHashes
//                  (src/bloom_filter.cpp:60-68, bloom_filter.cpp:50-61,72-78)
//   or_*_kernel    the OR-merge of partial filters (multi-GPU)
//
// The path is HBM / random-access bound integer work (≈15 integer ops per key, one random 8-byte
// filter word per key): no tensor cores, no GEMM shape.  What matters is (1) coalesced 128-bit
// streaming key loads that do not pollute L2, (2) as many independent random filter loads in flight
// per thread as registers allow, (3) keeping the filter — not the key stream — resident in the
// 126 MB L2 (createpolicy evict_last / evict_first), and (4) a single-pass, order-stable compaction
// so the selection vector is written once.
#include "rpt_kernels.cuh"

#include <algorithm>
#include <cstdlib>
#include <type_traits>

namespace rpt {

// ---------------------------------------------------------------------------------------------
// cache-policy helpers (PTX): keys are read once → L1 no_allocate + L2 evict_first; filter words are
// re-read at random → L2 evict_last.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ uint64_t policy_evict_first() {
	uint64_t p;
	asm("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
	return p;
}
__device__ __forceinline__ uint64_t policy_evict_last() {
	uint64_t p;
	asm("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(p));
	return p;
}
__device__ __forceinline__ ulonglong2 ld_stream_v2(const void *ptr, uint64_t pol) {
	ulonglong2 v;
	asm("ld.global.nc.L1::no_allocate.L2::cache_hint.v2.u64 {%0, %1}, [%2], %3;"
	    : "=l"(v.x), "=l"(v.y)
	    : "l"(ptr), "l"(pol));
	return v;
}
__device__ __forceinline__ uint64_t ld_stream_u64(const void *ptr, uint64_t pol) {
	uint64_t v;
	asm("ld.global.nc.L1::no_allocate.L2::cache_hint.u64 %0, [%1], %2;" : "=l"(v) : "l"(ptr), "l"(pol));
	return v;
}
__device__ __forceinline__ uint64_t ld_filter_word(const uint64_t *ptr, uint64_t pol) {
	uint64_t v;
	asm("ld.global.nc.L1::no_allocate.L2::cache_hint.u64 %0, [%1], %2;" : "=l"(v) : "l"(ptr), "l"(pol));
	return v;
}
__device__ __forceinline__ uint64_t ld_relaxed_u64(const uint64_t *ptr) {
	uint64_t v;
	asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(ptr) : "memory");
	return v;
}
__device__ __forceinline__ void st_relaxed_u64(uint64_t *ptr, uint64_t v) {
	asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(ptr), "l"(v) : "memory");
}
__device__ __forceinline__ void st_stream_v4(void *ptr, const uint4 v, uint64_t pol) {
	asm volatile("st.global.L2::cache_hint.v4.b32 [%0], {%1, %2, %3, %4}, %5;" ::"l"(ptr), "r"(v.x), "r"(v.y), "r"(v.z),
	             "r"(v.w), "l"(pol)
	             : "memory");
}
__device__ __forceinline__ void st_stream_v2(void *ptr, const uint2 v, uint64_t pol) {
	asm volatile("st.global.L2::cache_hint.v2.b32 [%0], {%1, %2}, %3;" ::"l"(ptr), "r"(v.x), "r"(v.y), "l"(pol)
	             : "memory");
}
__device__ __forceinline__ uint32_t lanemask_lt() {
	uint32_t m;
	asm("mov.u32 %0, %%lanemask_lt;" : "=r"(m));
	return m;
}

// ---------------------------------------------------------------------------------------------
// min / max of a key column (UpdateMinMax, reference src/operators/physical_create_bf.cpp:82-176): the build
// side's per-column range, which the reference computes chunk by chunk next to the insert loop (:229-239).
// Here it is FUSED into build_kernel (template flag MM): the keys are read once, every thread folds its valid
// (non-NULL, :93-95) rows into a running range, and a CTA issues one atomicMin/atomicMax pair at exit.
// Values are mapped to order-preserving unsigned 64-bit keys so that one pair serves every numeric type;
// NaNs are counted, not ordered (see rpt_gpu.h).
// ---------------------------------------------------------------------------------------------
struct MinMaxAcc {
	uint64_t lo = ~0ULL, hi = 0;
	uint32_t any = 0, nan = 0;
	__device__ __forceinline__ void fold(uint64_t key) {
		lo = key < lo ? key : lo;
		hi = key > hi ? key : hi;
		any = 1;
	}
};
constexpr uint64_t kSignBit = 1ULL << 63;

__device__ __forceinline__ bool minmax_key(const DevCol &c, uint64_t idx, uint64_t *key, bool *nan) {
	constexpr uint64_t kSign = 1ULL << 63;
	*nan = false;
	switch (c.type) {
	case RPT_BOOL:
	case RPT_UINT8:
		*key = __ldg(reinterpret_cast<const unsigned char *>(c.data) + idx);
		return true;
	case RPT_INT8:
		*key = static_cast<uint64_t>(static_cast<int64_t>(__ldg(reinterpret_cast<const signed char *>(c.data) + idx))) ^ kSign;
		return true;
	case RPT_UINT16:
		*key = __ldg(reinterpret_cast<const unsigned short *>(c.data) + idx);
		return true;
	case RPT_INT16:
		*key = static_cast<uint64_t>(static_cast<int64_t>(__ldg(reinterpret_cast<const short *>(c.data) + idx))) ^ kSign;
		return true;
	case RPT_UINT32:
		*key = __ldg(reinterpret_cast<const uint32_t *>(c.data) + idx);
		return true;
	case RPT_INT32:
		*key = static_cast<uint64_t>(static_cast<int64_t>(__ldg(reinterpret_cast<const int *>(c.data) + idx))) ^ kSign;
		return true;
	case RPT_UINT64:
		*key = __ldg(reinterpret_cast<const unsigned long long *>(c.data) + idx);
		return true;
	case RPT_INT64:
		*key = __ldg(reinterpret_cast<const unsigned long long *>(c.data) + idx) ^ kSign;
		return true;
	case RPT_FLOAT: {
		const uint32_t b = __ldg(reinterpret_cast<const uint32_t *>(c.data) + idx);
		if ((b & 0x7fffffffu) > 0x7f800000u) {
			*nan = true;
			return false;
		}
		const uint32_t o = (b & 0x80000000u) ? ~b : (b | 0x80000000u);
		*key = o;
		return true;
	}
	case RPT_DOUBLE: {
		const uint64_t b = __ldg(reinterpret_cast<const unsigned long long *>(c.data) + idx);
		if ((b & ~kSign) > 0x7ff0000000000000ULL) {
			*nan = true;
			return false;
		}
		*key = (b & kSign) ? ~b : (b | kSign);
		return true;
	}
	default:
		return false;
	}
}


// CTA-wide reduction of the threads' ranges, one atomic pair per CTA
template <int BLOCK>
__device__ __forceinline__ void minmax_commit(MinMaxAcc acc, MinMaxState *state) {
	__shared__ uint64_t s_lo[BLOCK / 32], s_hi[BLOCK / 32];
	__shared__ uint32_t s_flags[BLOCK / 32];
#pragma unroll
	for (int d = 16; d > 0; d >>= 1) {
		const uint64_t olo = __shfl_xor_sync(0xffffffffu, acc.lo, d);
		const uint64_t ohi = __shfl_xor_sync(0xffffffffu, acc.hi, d);
		acc.lo = olo < acc.lo ? olo : acc.lo;
		acc.hi = ohi > acc.hi ? ohi : acc.hi;
	}
	const uint32_t any = __any_sync(0xffffffffu, acc.any) ? 1u : 0u;
	const uint32_t nan = __any_sync(0xffffffffu, acc.nan) ? 2u : 0u;
	const uint32_t warp = threadIdx.x >> 5;
	if ((threadIdx.x & 31u) == 0) {
		s_lo[warp] = acc.lo;
		s_hi[warp] = acc.hi;
		s_flags[warp] = any | nan;
	}
	__syncthreads();
	if (threadIdx.x == 0) {
		uint64_t lo = ~0ULL, hi = 0;
		uint32_t flags = 0;
#pragma unroll
		for (int w = 0; w < BLOCK / 32; w++) {
			lo = s_lo[w] < lo ? s_lo[w] : lo;
			hi = s_hi[w] > hi ? s_hi[w] : hi;
			flags |= s_flags[w];
		}
		if (flags & 1u) {
			atomicMin(reinterpret_cast<unsigned long long *>(&state->min_key), static_cast<unsigned long long>(lo));
			atomicMax(reinterpret_cast<unsigned long long *>(&state->max_key), static_cast<unsigned long long>(hi));
			atomicOr(&state->has_value, 1u);
		}
		if (flags & 2u) {
			atomicOr(&state->has_nan, 1u);
		}
	}
}

// ---------------------------------------------------------------------------------------------
// key sources.  load2(row0, count, h0, h1) hashes rows row0 and row0+1 (row0 is even) and
// returns a 2-bit validity (rows ≥ count are invalid and must not be read).
// ---------------------------------------------------------------------------------------------
// Validity of flat columns (bit set = valid, row-indexed): NULL rows hash to NULL_HASH (vector_hash.cpp:16-23).
__device__ __forceinline__ uint32_t valid_pair(const uint64_t *validity, uint64_t row0) { // rows row0 (even), row0+1
	return static_cast<uint32_t>(__ldg(validity + (row0 >> 6)) >> (row0 & 63)) & 3u;
}
__device__ __forceinline__ bool valid_row(const uint64_t *validity, uint64_t row) {
	return (__ldg(validity + (row >> 6)) >> (row & 63)) & 1ULL;
}

// NULLABLE is a compile-time property: the all-valid variant carries no validity code at all (folding the check in
// at run time cost the sector-range pass 8 registers and 7 % of its time, ncu r1 final vs r1c).
// PREHASHED: the column holds 64-bit hashes (strings hashed ahead by hash_varchar_kernel), not keys
template <bool NULLABLE, bool PREHASHED = false>
struct FlatU64SrcT { // one flat int64 / uint64 column (validity bitmap iff NULLABLE, no selection), base 16-byte aligned
	static constexpr bool kNullable = NULLABLE;
	static constexpr bool kStageable = true; // a segment is 512 contiguous 8-byte keys: cp.async-able
	static constexpr int kKeyBytes = 8;
	const uint64_t *keys;
	const uint64_t *validity; // read iff NULLABLE
	__device__ __forceinline__ static uint64_t key_hash(uint64_t k) {
		return PREHASHED ? k : murmur64(k);
	}
	// KPU hashes out of one staged 16-byte unit
	__device__ __forceinline__ static void hash_unit(const uint4 &v, uint64_t *h) {
		h[0] = key_hash((static_cast<uint64_t>(v.y) << 32) | v.x);
		h[1] = key_hash((static_cast<uint64_t>(v.w) << 32) | v.z);
	}
	__device__ __forceinline__ const uint64_t *key_ptr() const {
		return keys;
	}
	__device__ __forceinline__ uint64_t hash_at(uint64_t row, uint64_t pol) const {
		if (NULLABLE && !valid_row(validity, row)) {
			return kNullHash;
		}
		return key_hash(ld_stream_u64(keys + row, pol));
	}
	// unchecked: rows row0 + r*stride (+0,+1), r < R, all in range.  All R 128-bit loads are issued
	// before the first hash so that they are in flight together.
	// MM: the raw keys are also folded into the running column range (mm_signed: INT64, else UINT64)
	template <int R, bool MM = false>
	__device__ __forceinline__ void load_rounds(uint64_t row0, uint64_t stride, uint64_t pol, uint64_t (&h)[2 * R],
	                                            MinMaxAcc *acc = nullptr, bool mm_signed = false) const {
		ulonglong2 k[R];
#pragma unroll
		for (int r = 0; r < R; r++) {
			k[r] = ld_stream_v2(keys + row0 + r * stride, pol);
		}
#pragma unroll
		for (int r = 0; r < R; r++) {
			h[2 * r] = key_hash(k[r].x);
			h[2 * r + 1] = key_hash(k[r].y);
		}
		const uint64_t flip = mm_signed ? kSignBit : 0ULL;
		if (NULLABLE) {
#pragma unroll
			for (int r = 0; r < R; r++) {
				const uint32_t v = valid_pair(validity, row0 + r * stride);
				h[2 * r] = (v & 1u) ? h[2 * r] : kNullHash;
				h[2 * r + 1] = (v & 2u) ? h[2 * r + 1] : kNullHash;
				if (MM) {
					if (v & 1u) {
						acc->fold(k[r].x ^ flip);
					}
					if (v & 2u) {
						acc->fold(k[r].y ^ flip);
					}
				}
			}
		} else if (MM) {
#pragma unroll
			for (int r = 0; r < R; r++) {
				acc->fold(k[r].x ^ flip);
				acc->fold(k[r].y ^ flip);
			}
		}
	}
	__device__ __forceinline__ void fold_minmax(uint64_t row, uint64_t pol, MinMaxAcc *acc, bool mm_signed) const {
		if (NULLABLE && !valid_row(validity, row)) {
			return;
		}
		acc->fold(ld_stream_u64(keys + row, pol) ^ (mm_signed ? kSignBit : 0ULL));
	}
	__device__ __forceinline__ uint32_t load2(uint64_t row0, uint64_t count, uint64_t pol, uint64_t &h0,
	                                          uint64_t &h1) const {
		h0 = h1 = 0;
		uint32_t v = 0;
		if (row0 < count) {
			h0 = hash_at(row0, pol);
			v = 1u;
		}
		if (row0 + 1 < count) {
			h1 = hash_at(row0 + 1, pol);
			v |= 2u;
		}
		return v;
	}
};
using FlatU64Src = FlatU64SrcT<false>;

// One flat 4-byte integer column (INT32 / UINT32 — also DATE), base 16-byte aligned.
// Hash<int32/uint32> = MurmurHash64 of the zero-extended 32-bit pattern (hash.hpp:49-52).
__device__ __forceinline__ uint2 ld_stream_v2_u32(const void *ptr, uint64_t pol) {
	uint2 v;
	asm("ld.global.nc.L1::no_allocate.L2::cache_hint.v2.u32 {%0, %1}, [%2], %3;" : "=r"(v.x), "=r"(v.y) : "l"(ptr), "l"(pol));
	return v;
}
__device__ __forceinline__ uint32_t ld_stream_u32(const void *ptr, uint64_t pol) {
	uint32_t v;
	asm("ld.global.nc.L1::no_allocate.L2::cache_hint.u32 %0, [%1], %2;" : "=r"(v) : "l"(ptr), "l"(pol));
	return v;
}
template <bool NULLABLE>
struct FlatU32SrcT {
	static constexpr bool kNullable = NULLABLE;
	static constexpr bool kStageable = true; // a segment is 512 contiguous 4-byte keys = 2 KiB
	static constexpr int kKeyBytes = 4;
	const uint32_t *keys;
	const uint64_t *validity; // read iff NULLABLE
	__device__ __forceinline__ static void hash_unit(const uint4 &v, uint64_t *h) {
		h[0] = murmur64(v.x);
		h[1] = murmur64(v.y);
		h[2] = murmur64(v.z);
		h[3] = murmur64(v.w);
	}
	__device__ __forceinline__ const uint32_t *key_ptr() const {
		return keys;
	}
	__device__ __forceinline__ uint64_t hash_at(uint64_t row, uint64_t pol) const {
		if (NULLABLE && !valid_row(validity, row)) {
			return kNullHash;
		}
		return murmur64(ld_stream_u32(keys + row, pol));
	}
	// order-preserving 64-bit image of a 4-byte key: sign-extended ^ 2^63 (INT32) or zero-extended (UINT32)
	__device__ __forceinline__ static uint64_t mm_key(uint32_t k, bool mm_signed) {
		return mm_signed ? (static_cast<uint64_t>(static_cast<int64_t>(static_cast<int32_t>(k))) ^ kSignBit)
		                 : static_cast<uint64_t>(k);
	}
	template <int R, bool MM = false>
	__device__ __forceinline__ void load_rounds(uint64_t row0, uint64_t stride, uint64_t pol, uint64_t (&h)[2 * R],
	                                            MinMaxAcc *acc = nullptr, bool mm_signed = false) const {
		uint2 k[R];
#pragma unroll
		for (int r = 0; r < R; r++) {
			k[r] = ld_stream_v2_u32(keys + row0 + r * stride, pol);
		}
#pragma unroll
		for (int r = 0; r < R; r++) {
			h[2 * r] = murmur64(k[r].x);
			h[2 * r + 1] = murmur64(k[r].y);
		}
		if (NULLABLE) {
#pragma unroll
			for (int r = 0; r < R; r++) {
				const uint32_t v = valid_pair(validity, row0 + r * stride);
				h[2 * r] = (v & 1u) ? h[2 * r] : kNullHash;
				h[2 * r + 1] = (v & 2u) ? h[2 * r + 1] : kNullHash;
				if (MM) {
					if (v & 1u) {
						acc->fold(mm_key(k[r].x, mm_signed));
					}
					if (v & 2u) {
						acc->fold(mm_key(k[r].y, mm_signed));
					}
				}
			}
		} else if (MM) {
#pragma unroll
			for (int r = 0; r < R; r++) {
				acc->fold(mm_key(k[r].x, mm_signed));
				acc->fold(mm_key(k[r].y, mm_signed));
			}
		}
	}
	__device__ __forceinline__ void fold_minmax(uint64_t row, uint64_t pol, MinMaxAcc *acc, bool mm_signed) const {
		if (NULLABLE && !valid_row(validity, row)) {
			return;
		}
		acc->fold(mm_key(ld_stream_u32(keys + row, pol), mm_signed));
	}
	__device__ __forceinline__ uint32_t load2(uint64_t row0, uint64_t count, uint64_t pol, uint64_t &h0,
	                                          uint64_t &h1) const {
		h0 = h1 = 0;
		uint32_t v = 0;
		if (row0 < count) {
			h0 = hash_at(row0, pol);
			v = 1u;
		}
		if (row0 + 1 < count) {
			h1 = hash_at(row0 + 1, pol);
			v |= 2u;
		}
		return v;
	}
};

using FlatU32Src = FlatU32SrcT<false>;

struct GenericSrc { // any type / sel / validity / multi-column combination
	static constexpr bool kStageable = false;
	static constexpr int kKeyBytes = 8; // unused (never staged)
	static constexpr bool kNullable = false;             // NULLs are handled per row by hash_row
	static constexpr const uint64_t *validity = nullptr;
	__device__ __forceinline__ static void hash_unit(const uint4 &, uint64_t *) {
	}
	DevCols cols;
	__device__ __forceinline__ const uint64_t *key_ptr() const {
		return nullptr;
	}
	__device__ __forceinline__ uint64_t hash_at(uint64_t row, uint64_t) const {
		return hash_row(cols, row);
	}
	template <int R, bool MM = false>
	__device__ __forceinline__ void load_rounds(uint64_t row0, uint64_t stride, uint64_t pol, uint64_t (&h)[2 * R],
	                                            MinMaxAcc *acc = nullptr, bool mm_signed = false) const {
#pragma unroll
		for (int r = 0; r < R; r++) {
			h[2 * r] = hash_row(cols, row0 + r * stride);
			h[2 * r + 1] = hash_row(cols, row0 + r * stride + 1);
			if (MM) { // the value was just read by hash_row: this re-read is an L1 hit
				fold_minmax(row0 + r * stride, pol, acc, mm_signed);
				fold_minmax(row0 + r * stride + 1, pol, acc, mm_signed);
			}
		}
	}
	__device__ __forceinline__ void fold_minmax(uint64_t row, uint64_t, MinMaxAcc *acc, bool) const {
		const DevCol &c = cols.c[0];
		const uint64_t idx = c.sel ? static_cast<uint64_t>(__ldg(c.sel + row)) : row;
		if (c.validity && !((__ldg(c.validity + (idx >> 6)) >> (idx & 63)) & 1ULL)) {
			return; // NULL rows are skipped (physical_create_bf.cpp:93-95)
		}
		uint64_t key;
		bool nan;
		if (minmax_key(c, idx, &key, &nan)) {
			acc->fold(key);
		}
		acc->nan |= nan ? 1u : 0u;
	}
	__device__ __forceinline__ uint32_t load2(uint64_t row0, uint64_t count, uint64_t, uint64_t &h0,
	                                          uint64_t &h1) const {
		uint32_t v = 0;
		h0 = h1 = 0;
		if (row0 < count) {
			h0 = hash_row(cols, row0);
			v = 1u;
		}
		if (row0 + 1 < count) {
			h1 = hash_row(cols, row0 + 1);
			v |= 2u;
		}
		return v;
	}
};

// ---------------------------------------------------------------------------------------------
// build
//
// Sector-range passes.  A miss in L2 costs a 128-byte DRAM line for 8 useful bytes (measured:
// profiles/mem_lab_r1.txt), so a filter that does not fit L2 is built in ceil(F / slice) passes over the
// key column: pass p only inserts the keys whose sector lies in [lo, lo+len).  The pass's slice of the
// filter stays L2-resident (evict_last on the reductions) while the keys stream through (evict_first),
// and re-reading 8 B/key per pass is far cheaper than ≥128 B/key of line fills and write-backs.
// ---------------------------------------------------------------------------------------------
constexpr int kBuildBlock = 256;
constexpr int kBuildUnroll = 4;

__device__ __forceinline__ void red_or_u64(unsigned long long *ptr, uint64_t mask, uint64_t pol) {
	asm volatile("red.global.or.L2::cache_hint.b64 [%0], %1, %2;" ::"l"(ptr), "l"(mask), "l"(pol) : "memory");
}

// AGG: lanes of a warp that hit the same sector OR their masks first and issue one reduction
// (worth it only for small filters or duplicate-heavy keys).
// WH: the first sector-range pass of a sliced build also writes every row's hash to hash_out; the later passes then
// read that column (FlatU64SrcT<false, true>: 8 bytes per row, no key decoding, no murmur, no NULL handling) instead of
// hashing the key columns again — a pass of a sliced build is issue/ALU-bound on exactly that work.
template <class Src, bool AGG, bool MM, bool WH = false>
__global__ void __launch_bounds__(kBuildBlock)
    build_kernel(const Src src, unsigned long long *__restrict__ filter, const uint32_t sector_mask,
                 const uint64_t count, const uint32_t slice_lo, const uint32_t slice_len, MinMaxState *mm_state,
                 const bool mm_signed, uint64_t *__restrict__ hash_out = nullptr) {
	const uint64_t pol = policy_evict_first();
	const uint64_t pol_filter = policy_evict_last();
	const uint64_t step = static_cast<uint64_t>(gridDim.x) * kBuildBlock * 2 * kBuildUnroll;
	// (ncu of a sector-range pass, profiles/ncu_build_r2.md: 43 instructions per key, issue 57 %, the largest stall the
	// wait for the tile's key loads — but loading the next tile's keys ahead changed nothing: 0.968 against 0.972 ms)
	MinMaxAcc acc;
	for (uint64_t base = static_cast<uint64_t>(blockIdx.x) * kBuildBlock * 2 * kBuildUnroll; base < count;
	     base += step) {
		uint64_t h[2 * kBuildUnroll];
		uint32_t valid = 0;
		if (base + static_cast<uint64_t>(kBuildBlock) * 2 * kBuildUnroll <= count) { // block-uniform
			src.template load_rounds<kBuildUnroll, MM>(base + threadIdx.x * 2, kBuildBlock * 2, pol, h, &acc, mm_signed);
			valid = (1u << (2 * kBuildUnroll)) - 1u;
			if (WH) {
#pragma unroll
				for (int u = 0; u < kBuildUnroll; u++) {
					const uint64_t row0 = base + static_cast<uint64_t>(u) * kBuildBlock * 2 + threadIdx.x * 2;
					*reinterpret_cast<ulonglong2 *>(hash_out + row0) = make_ulonglong2(h[2 * u], h[2 * u + 1]);
				}
			}
		} else {
#pragma unroll
			for (int u = 0; u < kBuildUnroll; u++) {
				const uint64_t row0 = base + static_cast<uint64_t>(u) * kBuildBlock * 2 + threadIdx.x * 2;
				const uint32_t v = src.load2(row0, count, pol, h[2 * u], h[2 * u + 1]);
				valid |= v << (2 * u);
				if (WH) {
					if (v & 1u) {
						hash_out[row0] = h[2 * u];
					}
					if (v & 2u) {
						hash_out[row0 + 1] = h[2 * u + 1];
					}
				}
				if (MM) {
					if (v & 1u) {
						src.fold_minmax(row0, pol, &acc, mm_signed);
					}
					if (v & 2u) {
						src.fold_minmax(row0 + 1, pol, &acc, mm_signed);
					}
				}
			}
		}
#pragma unroll
		for (int i = 0; i < 2 * kBuildUnroll; i++) {
			const uint32_t sector = static_cast<uint32_t>(h[i]) & sector_mask;
			const bool v = ((valid >> i) & 1u) && (sector - slice_lo < slice_len);
			uint64_t mask = sector_mask_fast(static_cast<uint32_t>(h[i] >> 32));
			if (AGG) {
				// every lane takes part; invalid lanes use a sector id no valid lane can have
				const uint32_t key = v ? sector : 0xffffffffu;
				const uint32_t peers = __match_any_sync(0xffffffffu, key);
				const uint32_t lo = __reduce_or_sync(peers, static_cast<uint32_t>(mask));
				const uint32_t hi = __reduce_or_sync(peers, static_cast<uint32_t>(mask >> 32));
				mask = (static_cast<uint64_t>(hi) << 32) | lo;
				if (v && (__ffs(peers) - 1) == static_cast<int>(threadIdx.x & 31)) {
					red_or_u64(filter + sector, mask, pol_filter);
				}
			} else if (v) {
				red_or_u64(filter + sector, mask, pol_filter);
			}
		}
	}
	if (MM) {
		minmax_commit<kBuildBlock>(acc, mm_state);
	}
}

// ---------------------------------------------------------------------------------------------
// probe: persistent CTAs take tiles in ticket order; per tile every thread owns ROUNDS×2 keys laid
// out so that each warp reads ROUNDS contiguous 512-byte segments (fully coalesced 128-bit loads);
// all 2·ROUNDS random filter loads of a thread are independent and in flight together; the tile's
// survivor count is chained to its predecessors with a decoupled look-back (one 64-bit descriptor
// per tile: status in bits 63:62, count in the low bits), which makes the compaction single-pass
// and globally order-stable, as LookupHashes' scalar loop is.
// ---------------------------------------------------------------------------------------------
constexpr uint64_t kDescAggregate = 1ULL << 62;
constexpr uint64_t kDescPrefix = 2ULL << 62;
constexpr uint64_t kDescValue = (1ULL << 62) - 1;

__device__ __forceinline__ uint32_t tile_lookback(uint64_t *desc, const uint64_t tile, const uint32_t lane,
                                                  const uint32_t tile_count) {
	if (lane == 0) {
		st_relaxed_u64(desc + tile, (tile == 0 ? kDescPrefix : kDescAggregate) | tile_count);
	}
	if (tile == 0) {
		return 0;
	}
	uint32_t exclusive = 0;
	int64_t pos = static_cast<int64_t>(tile) - 1;
	while (true) {
		const int64_t idx = pos - static_cast<int64_t>(lane);
		uint64_t d = kDescPrefix; // virtual predecessor of tile 0: inclusive prefix 0
		if (idx >= 0) {
			do {
				d = ld_relaxed_u64(desc + idx);
			} while ((d >> 62) == 0);
		}
		const uint32_t prefix_lanes = __ballot_sync(0xffffffffu, (d >> 62) == 2);
		uint32_t v = static_cast<uint32_t>(d & kDescValue);
		if (prefix_lanes) {
			const uint32_t first = __ffs(prefix_lanes) - 1;
			if (lane > first) {
				v = 0;
			}
		}
		exclusive += __reduce_add_sync(0xffffffffu, v);
		if (prefix_lanes) {
			break;
		}
		pos -= 32;
	}
	if (lane == 0) {
		st_relaxed_u64(desc + tile, kDescPrefix | (static_cast<uint64_t>(exclusive) + tile_count));
	}
	return exclusive;
}

// Lookup + test of 2·ROUNDS hashed keys.  Thread-local result: bit i is the hit flag of this thread's i-th key
// (i = 2·round + element).  SLICED: only sectors in [slice_lo, slice_lo + slice_len) are looked up (the other
// keys' bits stay 0 in this pass).
template <int ROUNDS, bool SMEM_FILTER, bool SLICED>
__device__ __forceinline__ uint32_t test_hashes(const uint64_t (&h)[2 * ROUNDS], const uint64_t *__restrict__ filter,
                                                const uint64_t *s_filter, const uint32_t sector_mask,
                                                const uint64_t pol_filter, const uint32_t slice_lo,
                                                const uint32_t slice_len) {
	// all random filter loads of this thread, independent of each other
	uint64_t w[2 * ROUNDS];
#pragma unroll
	for (int i = 0; i < 2 * ROUNDS; i++) {
		const uint32_t sector = static_cast<uint32_t>(h[i]) & sector_mask;
		if (SMEM_FILTER) {
			w[i] = s_filter[sector];
		} else if (SLICED) {
			w[i] = 0; // an all-zero word passes nothing
			if (sector - slice_lo < slice_len) {
				w[i] = ld_filter_word(filter + sector, pol_filter);
			}
		} else {
			w[i] = ld_filter_word(filter + sector, pol_filter);
		}
	}
	uint32_t hits = 0;
#pragma unroll
	for (int i = 2 * ROUNDS - 1; i >= 0; i--) { // hits·2 + bit is one IMAD on the FMA pipe; the ALU pipe is the busy one
		hits = hits * 2u + sector_test(w[i], static_cast<uint32_t>(h[i] >> 32));
	}
	return hits;
}

// Hash a tile's keys (read straight from global memory) and test them against the filter.
template <class Src, int ROUNDS, bool SMEM_FILTER, bool SLICED>
__device__ __forceinline__ uint32_t probe_tile(const Src &src, const uint64_t *__restrict__ filter,
                                               const uint64_t *s_filter, const uint32_t sector_mask,
                                               const uint64_t count, const uint64_t warp_base, const bool full_tile,
                                               const uint32_t lane, const uint64_t pol_keys, const uint64_t pol_filter,
                                               const uint32_t slice_lo, const uint32_t slice_len) {
	uint64_t h[2 * ROUNDS];
	uint32_t valid = 0;
	if (full_tile) { // block-uniform: every tile but the last
		src.template load_rounds<ROUNDS>(warp_base + lane * 2, 64, pol_keys, h);
		valid = static_cast<uint32_t>((1ull << (2 * ROUNDS)) - 1ull);
	} else {
#pragma unroll
		for (int r = 0; r < ROUNDS; r++) {
			const uint64_t row0 = warp_base + static_cast<uint64_t>(r) * 64 + lane * 2;
			valid |= src.load2(row0, count, pol_keys, h[2 * r], h[2 * r + 1]) << (2 * r);
		}
	}
	return test_hashes<ROUNDS, SMEM_FILTER, SLICED>(h, filter, s_filter, sector_mask, pol_filter, slice_lo, slice_len) &
	       valid;
}

// Pulls a slice of the filter into L2 with evict_last priority, one prefetch per 128-byte line, so that the
// random lookups of the pass start on a resident slice instead of faulting it in line by line.
__device__ __forceinline__ void warm_slice(const uint64_t *__restrict__ filter, const uint32_t slice_lo,
                                           const uint32_t slice_len, const uint32_t block) {
	const char *g = reinterpret_cast<const char *>(filter + slice_lo);
	const uint64_t nlines = (static_cast<uint64_t>(slice_len) * 8 + 127) >> 7;
	for (uint64_t i = static_cast<uint64_t>(blockIdx.x) * block + threadIdx.x; i < nlines;
	     i += static_cast<uint64_t>(gridDim.x) * block) {
		asm volatile("prefetch.global.L2::evict_last [%0];" ::"l"(g + (i << 7)));
	}
}

// ---- large probes: bitmap pass kernel(s) → segment-count scan → compaction ------------------------------
// The fused single-kernel probe further down chains tiles through a decoupled look-back; with persistent CTAs
// that puts a ticket fetch, three block barriers and the look-back latency on every tile's critical path (ncu:
// 46 % barrier + 34 % long-scoreboard stalls, issue slots 22 % busy).  Large probes therefore run as
// independent phases with no inter-CTA and no intra-CTA dependency inside the hot kernel:
//   probe_bitmap_kernel   hashes and tests the keys (once per sector-range slice) and ORs the hit flags into a
//                         ROW-ORDERED bitmap (bit i of the bitmap = row i passes); the last pass also leaves
//                         the hit count of every segment (512 consecutive rows)
//   scan_blocks_kernel    exclusive prefix sum inside blocks of 4096 segment counts + block totals
//   scan_totals_kernel    exclusive prefix sum of the block totals (one CTA), writes `found`
//   compact_bitmap_kernel bitmap words + offsets → ordered selection vector
//
// Work unit = one warp × one segment of 512 consecutive rows; lane l owns rows 16·l … 16·l+15 of the segment,
// so its 16 hit flags are one uint16 of the bitmap and a warp stores 64 contiguous bytes.
//
// Key staging (flat 8-byte column): the segment's 4 KiB arrive in shared memory through 8 coalesced cp.async
// (LDGSTS, 16 bytes per lane, L2 evict_first) into a warp-private double buffer while the previous segment is
// hashed and tested, so the DRAM latency of the key stream is off the critical path and no block barrier
// exists.  The layout in shared memory is linear; lane l reads its eight 16-byte units in the rotated order
// (p + l) mod 8, p = 0..7, which makes every quarter-warp's LDS.128 hit eight distinct bank groups.  The hit
// mask collected in that rotated order is rotated back by 2·(l mod 8) bits at the end.
constexpr uint64_t kScanBlock = 4096;
constexpr uint64_t kSegRows = 512;

// Staging geometry of a flat column with KB-byte keys: a 16-byte unit holds KPU keys, a lane owns U units
// (its 16 rows), a segment has SEG_UNITS units.  Lane l reads its units in the order (p + rot(l)) mod U:
//   8-byte keys: U = 8, rot = l mod 8        → unit index mod 8 = (p + l) mod 8, distinct over a quarter-warp
//   4-byte keys: U = 4, rot = (l / 2) mod 4  → unit index mod 8 = 4·(l mod 2) + (p + l/2) mod 4, distinct too
template <int KB>
struct StageGeo {
	static constexpr int KPU = 16 / KB;
	static constexpr int U = 16 / KPU;
	static constexpr int SEG_UNITS = static_cast<int>(kSegRows) / KPU;
	static constexpr int COPIES = SEG_UNITS / 32; // cp.async per lane and segment
	__device__ __forceinline__ static uint32_t rot(uint32_t lane) {
		return KB == 8 ? (lane & 7u) : ((lane >> 1) & 3u);
	}
};

__device__ __forceinline__ void cp_async_16(void *smem_dst, const void *gmem_src, uint64_t pol) {
	const uint32_t d = static_cast<uint32_t>(__cvta_generic_to_shared(smem_dst));
	asm volatile("cp.async.cg.shared.global.L2::cache_hint [%0], [%1], 16, %2;" ::"r"(d), "l"(gmem_src), "l"(pol)
	             : "memory");
}
__device__ __forceinline__ void cp_async_commit() {
	asm volatile("cp.async.commit_group;" ::: "memory");
}
template <int N>
__device__ __forceinline__ void cp_async_wait() {
	asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}

// Bit positions on the FMA pipe: (hi << k) >> 26 as IMAD.SHL + IMAD.HI, leaving the ALU pipe (the busiest one:
// ncu r1a 55 % vs 18 % FMA) to the four funnel shifts and two LOP3 of the test itself.
__device__ __forceinline__ uint32_t mul_lo_u32(uint32_t a, uint32_t b) {
	uint32_t d;
	asm("mul.lo.u32 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(b));
	return d;
}
__device__ __forceinline__ uint32_t mul_hi_u32(uint32_t a, uint32_t b) {
	uint32_t d;
	asm("mul.hi.u32 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(b));
	return d;
}
__device__ __forceinline__ uint32_t sector_test_fma(uint64_t w, uint32_t hi) {
	const uint32_t p0 = hi & 63u;
	const uint32_t p1 = mul_hi_u32(mul_lo_u32(hi, 1u << 18), 64u);
	const uint32_t p2 = mul_hi_u32(mul_lo_u32(hi, 1u << 10), 64u);
	const uint32_t p3 = mul_hi_u32(mul_lo_u32(hi, 4u), 64u);
	const uint32_t b = static_cast<uint32_t>(w >> p0) & static_cast<uint32_t>(w >> p1) & static_cast<uint32_t>(w >> p2) &
	                   static_cast<uint32_t>(w >> p3);
	return b & 1u;
}

// Staged segments of a nullable flat column: the lane's 16 validity bits (logical row order) are rotated into
// slot order and NULL rows take NULL_HASH.  N keys starting at slot `first`.
template <int N>
__device__ __forceinline__ void null_fix(uint64_t (&h)[N], uint32_t valid_slots, int first) {
#pragma unroll
	for (int i = 0; i < N; i++) {
		h[i] = ((valid_slots >> (first + i)) & 1u) ? h[i] : kNullHash;
	}
}
__device__ __forceinline__ uint32_t valid_slots_of(const uint64_t *validity, uint64_t seg, uint32_t lane,
                                                   uint32_t rot_bits) {
	const uint32_t v = __ldg(reinterpret_cast<const unsigned short *>(validity) + seg * 32 + lane);
	const uint32_t twice = v | (v << 16); // slot i holds logical bit (i + rot_bits) mod 16: rotate right
	return (twice >> rot_bits) & 0xffffu;
}

// Lookup + test of N hashed keys; bit i of the result is the hit flag of h[i].
template <int N, bool SMEM_FILTER, bool SLICED>
__device__ __forceinline__ uint32_t test_batch(const uint64_t (&h)[N], const uint64_t *__restrict__ filter,
                                               const uint64_t *s_filter, const uint32_t sector_mask,
                                               const uint64_t pol_filter, const uint32_t slice_lo,
                                               const uint32_t slice_len) {
	uint64_t w[N];
#pragma unroll
	for (int i = 0; i < N; i++) { // all random filter loads of this thread, independent of each other
		const uint32_t sector = static_cast<uint32_t>(h[i]) & sector_mask;
		if (SMEM_FILTER) {
			w[i] = s_filter[sector];
		} else if (SLICED) {
			w[i] = 0; // an all-zero word passes nothing
			if (sector - slice_lo < slice_len) {
				w[i] = ld_filter_word(filter + sector, pol_filter);
			}
		} else {
			w[i] = ld_filter_word(filter + sector, pol_filter);
		}
	}
	uint32_t hits = 0;
#pragma unroll
	for (int i = N - 1; i >= 0; i--) { // hits·2 + bit is one IMAD
		hits = hits * 2u + sector_test_fma(w[i], static_cast<uint32_t>(h[i] >> 32));
	}
	return hits;
}

// UNITS = 16-byte units (key pairs) per batch: 8 → all 16 keys of a lane in flight at once, 4 → two batches of 8.
template <class Src, int BLOCK, int MINB, int UNITS, bool SMEM_FILTER, bool SLICED>
__global__ void __launch_bounds__(BLOCK, MINB)
    probe_bitmap_kernel(const Src src, const uint64_t *__restrict__ filter, const uint32_t sector_mask,
                        const uint64_t count, const uint64_t num_segs, const uint32_t slice_lo,
                        const uint32_t slice_len, uint16_t *__restrict__ bitmap16, const int first_pass,
                        uint32_t *__restrict__ seg_counts /* non-null on the last pass */) {
	static_assert(UNITS == 8 || UNITS == 4, "a lane owns 8 units, processed in one or two batches");
	constexpr int NWARPS = BLOCK / 32;
	constexpr bool STAGED = Src::kStageable;
	using Geo = StageGeo<Src::kKeyBytes>;
	constexpr int SEG_UNITS = Geo::SEG_UNITS; // 16-byte units per segment
	constexpr int BATCH = 2 * UNITS;          // keys per batch
	// dynamic shared memory: [warp-private key double buffers (STAGED)][filter copy (SMEM_FILTER)]
	extern __shared__ __align__(128) unsigned char s_dyn[];
	uint4 *const s_keys = reinterpret_cast<uint4 *>(s_dyn);
	uint64_t *const s_filter = reinterpret_cast<uint64_t *>(s_dyn + (STAGED ? NWARPS * 2 * SEG_UNITS * 16 : 0));
	const uint32_t lane = threadIdx.x & 31u;
	const uint32_t warp = threadIdx.x >> 5;
	const uint64_t pol_keys = policy_evict_first();
	const uint64_t pol_filter = policy_evict_last();
	if (SMEM_FILTER) {
		const uint64_t nvec = (static_cast<uint64_t>(sector_mask) + 1) >> 1;
		const ulonglong2 *g = reinterpret_cast<const ulonglong2 *>(filter);
		ulonglong2 *sm = reinterpret_cast<ulonglong2 *>(s_filter);
		for (uint64_t i = threadIdx.x; i < nvec; i += BLOCK) {
			sm[i] = __ldg(g + i);
		}
		__syncthreads();
	}
	if (SLICED) {
		warm_slice(filter, slice_lo, slice_len, BLOCK);
	}
	const uint64_t stride = static_cast<uint64_t>(gridDim.x) * NWARPS;
	const uint64_t full_segs = count / kSegRows; // segments below this one are complete
	uint4 *const my_buf = s_keys + warp * (2 * SEG_UNITS);
	const uint32_t rot = Geo::rot(lane);

	// Software pipeline with ONE staging call site: iteration i finishes segment `cur` (staged in iteration
	// i-1 into buffer `par`) and stages segment `nxt` into the other buffer.  (A separate prologue call made
	// ptxas 12.9 emit LDGSTS with unset uniform registers.)
	constexpr uint64_t kNone = ~0ULL;
	uint64_t cur = kNone;
	uint32_t par = 1;
	for (uint64_t nxt = static_cast<uint64_t>(blockIdx.x) * NWARPS + warp;; nxt += stride) {
		if (STAGED) {
			__syncwarp(); // every lane is done reading the buffer that is overwritten next
			if (nxt < full_segs) {
				const uint4 *g = reinterpret_cast<const uint4 *>(src.key_ptr()) + nxt * SEG_UNITS + lane;
				uint4 *s = my_buf + (par ^ 1u) * SEG_UNITS + lane;
#pragma unroll
				for (int k = 0; k < Geo::COPIES; k++) {
					cp_async_16(s + k * 32, g + k * 32, pol_keys);
				}
			}
			cp_async_commit();
		}
		if (cur != kNone) {
			uint16_t *slot = bitmap16 + cur * 32 + lane;
			uint32_t prior = 0;
			if (!first_pass) {
				prior = *slot; // issued before the key / filter latency chain, consumed after it
			}
			const uint64_t row0 = cur * kSegRows + lane * 16;
			uint32_t hits = 0;
			if (STAGED && cur < full_segs) {
				cp_async_wait<1>();
				__syncwarp();
				const uint4 *mine = my_buf + par * SEG_UNITS + lane * Geo::U;
				const uint32_t vslots = Src::kNullable ? valid_slots_of(src.validity, cur, lane, Geo::KPU * rot) : 0xffffu;
#pragma unroll
				for (int b = 16 / BATCH - 1; b >= 0; b--) {
					uint64_t h[BATCH];
#pragma unroll
					for (int u = 0; u < BATCH / Geo::KPU; u++) {
						Src::hash_unit(mine[(b * (BATCH / Geo::KPU) + u + rot) & (Geo::U - 1)], h + u * Geo::KPU);
					}
					if (Src::kNullable) {
						null_fix<BATCH>(h, vslots, b * BATCH);
					}
					hits = (hits << BATCH) | test_batch<BATCH, SMEM_FILTER, SLICED>(h, filter, s_filter, sector_mask,
					                                                              pol_filter, slice_lo, slice_len);
				}
				// slot p held logical unit (p + rot) mod U: rotate the 16-bit mask left by KPU·rot
				const uint32_t twice = hits | (hits << 16);
				hits = (twice >> (16u - Geo::KPU * rot)) & 0xffffu;
			} else {
				const bool full = cur < full_segs;
#pragma unroll
				for (int b = 8 / UNITS - 1; b >= 0; b--) {
					uint64_t h[2 * UNITS];
					uint32_t valid = 0;
#pragma unroll
					for (int i = 0; i < 2 * UNITS; i++) {
						const uint64_t row = row0 + b * 2 * UNITS + i;
						h[i] = 0;
						if (full || row < count) {
							h[i] = src.hash_at(row, pol_keys);
							valid |= 1u << i;
						}
					}
					hits = (hits << (2 * UNITS)) |
					       (test_batch<2 * UNITS, SMEM_FILTER, SLICED>(h, filter, s_filter, sector_mask, pol_filter,
					                                                   slice_lo, slice_len) &
					        valid);
				}
			}
			hits |= prior;
			*slot = static_cast<uint16_t>(hits);
			if (seg_counts) { // kernel-uniform
				const uint32_t wc = __reduce_add_sync(0xffffffffu, __popc(hits));
				if (lane == 0) {
					seg_counts[cur] = wc;
				}
			}
		}
		if (nxt >= num_segs) {
			break;
		}
		cur = nxt;
		par ^= 1u;
	}
	if (STAGED) {
		cp_async_wait<0>();
	}
}

// ---- shared-memory filters, flat key columns: bulk-copy (TMA) staging, 24 warps per SM ---------------------------
// Regime 1 (F ≤ 128 KiB) of a staged source.  probe_bitmap_kernel's 8 KiB of cp.async double buffer per warp leave
// room for 12 warps next to a 128 KiB filter, and with 3 warps per scheduler the kernel was issue/latency-bound
// (ncu r1w: 59 % issue, 4.8 TB/s).  Here the filter arrives through ONE bulk copy per CTA (cp.async.bulk, SASS
// UBLKCP) and a warp's keys through bulk copies of HALF segments (256 rows) into two 2 KiB buffers, each guarded by
// its own mbarrier and refilled by one elected lane as soon as the warp has consumed it — 4 KiB per warp, 24 warps.
// Lane l owns rows 8l … 8l+7 of a half, i.e. one BYTE of the row-ordered bitmap; its 16-byte units are read in the
// rotated order of StageGeo (bank-conflict-free LDS.128) and the 8-bit hit mask is rotated back.
__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
	asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(static_cast<uint32_t>(__cvta_generic_to_shared(bar))),
	             "r"(count)
	             : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
	asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(
	                 static_cast<uint32_t>(__cvta_generic_to_shared(bar))),
	             "r"(bytes)
	             : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
	asm volatile("{\n"
	             ".reg .pred p;\n"
	             "WAIT_%=:\n"
	             "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
	             "@p bra DONE_%=;\n"
	             "bra WAIT_%=;\n"
	             "DONE_%=:\n"
	             "}" ::"r"(static_cast<uint32_t>(__cvta_generic_to_shared(bar))),
	             "r"(parity)
	             : "memory");
}
__device__ __forceinline__ void bulk_g2s(void *smem_dst, const void *gmem_src, uint32_t bytes, uint64_t *bar,
                                         uint64_t pol) {
	asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;" ::
	                 "r"(static_cast<uint32_t>(__cvta_generic_to_shared(smem_dst))),
	             "l"(gmem_src), "r"(bytes), "r"(static_cast<uint32_t>(__cvta_generic_to_shared(bar))), "l"(pol)
	             : "memory");
}

constexpr int kSmemProbeWarps = 24;
constexpr uint32_t kHalfRows = 256;

// GLOBAL: the same staging with the filter left in global memory (an L2-resident filter of up to one slice, regime 2):
// only the key buffers take shared memory, so the L1 keeps 156 KiB for the gathers in flight.
// SLICED (with GLOBAL): one sector-range pass of regime 3 — only the keys whose sector lies in [slice_lo, slice_lo +
// slice_len) are looked up, their hits are ORed into the bitmap bytes of the earlier passes (a byte belongs to the
// same lane in every pass), and the last pass (seg_counts != nullptr) leaves the segment counts.
template <class Src, int NWARPS, bool GLOBAL = false, bool SLICED = false>
__global__ void __launch_bounds__(NWARPS * 32, 1)
    probe_smem_kernel(const Src src, const uint64_t *__restrict__ filter, const uint32_t sector_mask, const uint64_t count,
                      const uint64_t num_segs, uint8_t *__restrict__ bitmap8, uint32_t *__restrict__ seg_counts,
                      const uint32_t slice_lo = 0, const uint32_t slice_len = 0, const int first_pass = 1) {
	static_assert(!SLICED || GLOBAL, "sector-range passes are for filters in global memory");
	static_assert(Src::kStageable, "flat key columns only");
	constexpr int KB = Src::kKeyBytes;
	constexpr int KPU = 16 / KB;                     // keys per 16-byte unit
	constexpr uint32_t HALF_BYTES = kHalfRows * KB;  // 2048 / 1024
	constexpr int LU = 8 / KPU;                      // 16-byte units a lane owns per half: 4 / 2
	extern __shared__ __align__(128) unsigned char s_dyn[];
	__shared__ __align__(8) uint64_t s_bars[NWARPS][2];
	__shared__ __align__(8) uint64_t s_fbar;
	const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
	const uint32_t filter_bytes = GLOBAL ? 0u : (sector_mask + 1u) * 8u;
	const uint64_t *const s_filter = reinterpret_cast<const uint64_t *>(s_dyn);
	unsigned char *const kbuf = s_dyn + filter_bytes + warp * 2u * HALF_BYTES;
	const uint64_t pol_keys = policy_evict_first();
	const uint64_t pol_filter = policy_evict_last();
	if (threadIdx.x == 0) {
		mbar_init(&s_fbar, 1);
	}
	if (lane == 0) {
		mbar_init(&s_bars[warp][0], 1);
		mbar_init(&s_bars[warp][1], 1);
	}
	asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
	__syncthreads();
	if (!GLOBAL && threadIdx.x == 0) { // the whole filter: one bulk copy
		mbar_expect_tx(&s_fbar, filter_bytes);
		bulk_g2s(s_dyn, filter, filter_bytes, &s_fbar, pol_filter);
	}
	if (GLOBAL) {
		warm_slice(filter, SLICED ? slice_lo : 0u, SLICED ? slice_len : sector_mask + 1u, NWARPS * 32);
	}
	const uint64_t full_segs = count / kSegRows;
	const uint64_t stride = static_cast<uint64_t>(gridDim.x) * NWARPS;
	const unsigned char *const keys = reinterpret_cast<const unsigned char *>(src.key_ptr());
	auto issue = [&](uint64_t seg, uint32_t h) { // elected lane: stage half h of segment seg
		if (seg < full_segs) {
			mbar_expect_tx(&s_bars[warp][h], HALF_BYTES);
			bulk_g2s(kbuf + h * HALF_BYTES, keys + (seg * kSegRows + h * kHalfRows) * KB, HALF_BYTES, &s_bars[warp][h],
			         pol_keys);
		}
	};
	const uint64_t seg0 = static_cast<uint64_t>(blockIdx.x) * NWARPS + warp;
	if (lane == 0) {
		issue(seg0, 0);
		issue(seg0, 1);
	}
	if (!GLOBAL) {
		mbar_wait(&s_fbar, 0);
	}
	const uint32_t rot = LU == 4 ? ((lane >> 1) & 3u) : ((lane >> 2) & 1u);
	uint32_t phase = 0;
	for (uint64_t seg = seg0; seg < num_segs; seg += stride) {
		uint32_t total = 0;
#pragma unroll
		for (uint32_t h = 0; h < 2; h++) {
			uint32_t hits = 0;
			uint32_t earlier = 0; // hits of the earlier sector-range passes
			if (SLICED && !first_pass) {
				earlier = bitmap8[seg * 64 + h * 32 + lane];
			}
			if (seg < full_segs) {
				mbar_wait(&s_bars[warp][h], (phase >> h) & 1u);
				phase ^= 1u << h;
				const uint4 *mine = reinterpret_cast<const uint4 *>(kbuf + h * HALF_BYTES) + lane * LU;
				uint64_t hh[8];
#pragma unroll
				for (int u = 0; u < LU; u++) {
					Src::hash_unit(mine[(u + rot) & (LU - 1)], hh + u * KPU);
				}
				if (Src::kNullable) {
					const uint32_t v = __ldg(reinterpret_cast<const unsigned char *>(src.validity) + seg * 64 + h * 32 + lane);
					const uint32_t twice = v | (v << 8); // slot i holds logical bit (i + KPU·rot) mod 8: rotate right
					null_fix<8>(hh, (twice >> (KPU * rot)) & 0xffu, 0);
				}
				hits = test_batch<8, !GLOBAL, SLICED>(hh, filter, s_filter, sector_mask, pol_filter, slice_lo, slice_len);
				const uint32_t twice = hits | (hits << 8); // slot order → row order: rotate left by KPU·rot
				hits = (twice >> (8u - KPU * rot)) & 0xffu;
				__syncwarp(); // every lane's keys are consumed: the buffer may be refilled
				if (lane == 0) {
					issue(seg + stride, h);
				}
			} else { // the partial last segment: straight from global memory
				const uint64_t row0 = seg * kSegRows + h * kHalfRows + lane * 8u;
#pragma unroll
				for (int i = 0; i < 8; i++) {
					if (row0 + i < count) {
						const uint64_t hv = src.hash_at(row0 + i, pol_keys);
						const uint32_t sector = static_cast<uint32_t>(hv) & sector_mask;
						if (!SLICED || sector - slice_lo < slice_len) {
							const uint64_t word = GLOBAL ? ld_filter_word(filter + sector, pol_filter) : s_filter[sector];
							hits |= sector_test(word, static_cast<uint32_t>(hv >> 32)) << i;
						}
					}
				}
			}
			hits |= earlier;
			bitmap8[seg * 64 + h * 32 + lane] = static_cast<uint8_t>(hits);
			total += __popc(hits);
		}
		if (!SLICED || seg_counts != nullptr) {
			total = __reduce_add_sync(0xffffffffu, total);
			if (lane == 0) {
				seg_counts[seg] = total;
			}
		}
	}
}

// ---- filters beyond one L2-resident slice: partition once, then probe slice by slice -----------------------
// The sector-range passes above re-read the 8-byte key column once per slice (4× at F = 256 MiB: ncu r1c 42 GB
// of DRAM traffic for 8.7 GB of algorithmic bytes, each pass DRAM-bound).  This path reads the keys ONCE:
//   probe_partition_kernel  hashes every key; keys of slice 0 are looked up on the spot; keys of slice s ≥ 1
//                           become 8-byte records {sector-in-slice : 23 | row-in-segment : 9 | hash bits 63..32}
//                           in bucket s.  Every (segment, slice) pair owns a fixed region of CAP records
//                           (≥ 4σ above the expected 512/P), so no atomics, no directory and no inter-warp
//                           dependency exist.  Placement is a counting sort inside the warp: per-lane counts
//                           of all slices live as 8-bit fields of one register, one 5-step warp scan on
//                           16-bit fields gives every lane its base in every region, a key's position is base
//                           + its rank inside the lane.  The records go through warp-private shared memory
//                           (aliased onto the key buffer just consumed) and leave the SM as contiguous
//                           16-byte stores.  A segment whose keys overflow a region (skew) is looked up on the
//                           spot instead — slower, never wrong.
//                           ncu (profiles/ncu_part_r1g.md): 73 instructions per key, 20.8 GB of DRAM traffic
//                           at 88 % of peak — 6 GB of it because slice 0 does NOT stay in L2 next to the
//                           record write stream (80 % of the inline lookups miss; store flavours, evict hints,
//                           the persisting set-aside and a smaller inline region were all tried: no gain).
//   probe_bucket_kernel     once per slice s ≥ 1 (slice prefetched into L2): record → filter word → test →
//                           RED.OR of the row's bit into the row-ordered bitmap (hits only).
//   count_segments_kernel   popcount of every segment's 512 bitmap bits → the scan's input.
// DRAM traffic per key: 8 (keys) + 6 + 6 (records out and in) instead of 32.
constexpr uint32_t kSliceShift = 23; // log2(kSliceSectors)
static_assert((1ULL << kSliceShift) == kSliceSectors, "record layout assumes 2^23-sector slices");

// INLINE0: the keys of slice 0 are looked up by the partition kernel itself (P-1 bucket regions per segment);
// otherwise slice 0 is a bucket like the others (P regions, no filter access in the partition kernel).
template <int P, bool INLINE0 = true>
struct PartCfg {
	static_assert(P == 2 || P == 4 || P == 8, "S is a power of two: 2, 4 or 8 slices");
	// records per (segment, slice) region; expectation 512/P, σ = sqrt(512·(1/P)(1-1/P)): +5.6σ / +4.1σ / +4.3σ
	static constexpr int CAP = P == 2 ? 320 : (P == 4 ? 168 : 96);
	static constexpr int FIRST = INLINE0 ? 1 : 0; // first slice that has a bucket
	static constexpr int REGIONS = P - FIRST;
	static constexpr int STAGE_BYTES = REGIONS * CAP * 8;
	static constexpr int NCNT = P <= 4 ? 1 : 2; // per-lane counters: 8 bits per slice, four slices per register
};

// Shared memory of one warp of the partition kernel: the key double buffer (staged sources) and the record
// stage.  When the stage fits into one key buffer it ALIASES the buffer of the segment being processed (its
// keys are in registers by then), which keeps a warp at 8 KiB.  Two CTAs per SM: at three (80 registers) the
// sixteen hashes spill, and the spill traffic on the LSU costs more than the extra warps hide (9.0 vs 7.3 ms).
template <class Src, int P, bool INLINE0 = true>
struct PartSmem {
	static constexpr int KEY_SEG_BYTES = Src::kStageable ? static_cast<int>(kSegRows) * Src::kKeyBytes : 0;
	static constexpr bool ALIAS = Src::kStageable && PartCfg<P, INLINE0>::STAGE_BYTES <= KEY_SEG_BYTES;
	static constexpr int WARP_BYTES =
	    2 * KEY_SEG_BYTES + (ALIAS ? 0 : (PartCfg<P, INLINE0>::STAGE_BYTES + 15) / 16 * 16);
};

__device__ __forceinline__ uint32_t shl_clamp(uint32_t v, uint32_t s) { // PTX shl.b32: a shift ≥ 32 yields 0
	uint32_t d;
	asm("shl.b32 %0, %1, %2;" : "=r"(d) : "r"(v), "r"(s));
	return d;
}

template <class Src, int BLOCK, int P, int SHIFT = kSliceShift, bool INLINE0 = true>
__global__ void __launch_bounds__(BLOCK, 2)
    probe_partition_kernel(const Src src, const uint64_t *__restrict__ filter, const uint32_t sector_mask,
                           const uint64_t count, const uint64_t num_segs, uint16_t *__restrict__ bitmap16,
                           uint2 *__restrict__ buckets, const uint64_t bucket_stride /* records per slice array */,
                           uint16_t *__restrict__ bucket_counts /* (P-1) × num_segs */,
                           uint32_t *__restrict__ ovf_list /* [0] = count, [1 + i] = i-th overflowed segment */) {
	constexpr int NWARPS = BLOCK / 32;
	constexpr bool STAGED = Src::kStageable;
	using Geo = StageGeo<Src::kKeyBytes>;
	using Smem = PartSmem<Src, P, INLINE0>;
	using Cfg = PartCfg<P, INLINE0>;
	constexpr int SEG_UNITS = Geo::SEG_UNITS;
	constexpr int CAP = Cfg::CAP;
	constexpr int NCNT = Cfg::NCNT;
	constexpr int FIRST = Cfg::FIRST;
	extern __shared__ __align__(128) unsigned char s_dyn[];
	const uint32_t lane = threadIdx.x & 31u;
	const uint32_t warp = threadIdx.x >> 5;
	unsigned char *const my_smem = s_dyn + warp * Smem::WARP_BYTES;
	uint4 *const my_buf = reinterpret_cast<uint4 *>(my_smem);
	const uint64_t pol_keys = policy_evict_first();
	const uint64_t pol_filter = policy_evict_last();
	if (INLINE0) {
		warm_slice(filter, 0u, 1u << SHIFT, BLOCK);
	}
	constexpr bool MASKED = !INLINE0; // slice 0 is a bucket: rows past the end must not become slice-0 records
	const uint64_t stride = static_cast<uint64_t>(gridDim.x) * NWARPS;
	const uint64_t full_segs = count / kSegRows;
	const uint32_t rot = Geo::rot(lane);

	constexpr uint64_t kNone = ~0ULL;
	uint64_t cur = kNone;
	uint32_t par = 1;
	for (uint64_t nxt = static_cast<uint64_t>(blockIdx.x) * NWARPS + warp;; nxt += stride) {
		if (STAGED) {
			__syncwarp(); // the buffer overwritten next (keys and, when aliased, records) is no longer read
			if (nxt < full_segs) {
				const uint4 *g = reinterpret_cast<const uint4 *>(src.key_ptr()) + nxt * SEG_UNITS + lane;
				uint4 *s = my_buf + (par ^ 1u) * SEG_UNITS + lane;
#pragma unroll
				for (int k = 0; k < Geo::COPIES; k++) {
					cp_async_16(s + k * 32, g + k * 32, pol_keys);
				}
			}
			cp_async_commit();
		}
		if (cur != kNone) {
			const bool staged_seg = STAGED && cur < full_segs;
			uint64_t h[16];
			uint32_t valid = 0xffffu; // slot order
			if (staged_seg) {
				cp_async_wait<1>();
				__syncwarp();
				const uint4 *mine = my_buf + par * SEG_UNITS + lane * Geo::U;
#pragma unroll
				for (int u = 0; u < Geo::U; u++) {
					Src::hash_unit(mine[(u + rot) & (Geo::U - 1)], h + u * Geo::KPU);
				}
				if (Src::kNullable) {
					null_fix<16>(h, valid_slots_of(src.validity, cur, lane, Geo::KPU * rot), 0);
				}
				if (Smem::ALIAS) {
					__syncwarp(); // every lane has its keys in registers: the buffer becomes the record stage
				}
			} else {
				const bool full = cur < full_segs;
				const uint64_t row0 = cur * kSegRows + lane * 16;
				valid = 0;
#pragma unroll
				for (int i = 0; i < 16; i++) {
					h[i] = 0; // a row past the end behaves like a slice-0 key whose hit is masked out below
					if (full || row0 + i < count) {
						h[i] = src.hash_at(row0 + i, pol_keys);
						valid |= 1u << i;
					}
				}
			}
			uint2 *const stage = reinterpret_cast<uint2 *>(
			    my_smem + (Smem::ALIAS ? par * Smem::KEY_SEG_BYTES : 2 * Smem::KEY_SEG_BYTES));
			const uint32_t unit_rot = staged_seg ? rot : 0u;
			const uint32_t alive = valid; // slot-order mask of the rows that exist

			// pass 1: per-lane counts of the keys of every slice (8-bit fields); the slice-0 lookups of the first
			// eight keys are issued (the second eight follow in pass 2, behind the records of the first eight)
			uint64_t w0[8];
			auto lookups = [&](const int lo) {
#pragma unroll
				for (int i = 0; i < 8; i++) {
					const uint32_t sector = static_cast<uint32_t>(h[lo + i]) & sector_mask;
					w0[i] = 0;
					if (INLINE0 && (sector >> SHIFT) == 0u) {
						w0[i] = ld_filter_word(filter + sector, pol_filter);
					}
				}
			};
			auto tests = [&](const int lo) -> uint32_t {
				uint32_t r = 0;
				if (!INLINE0) {
					return 0u;
				}
#pragma unroll
				for (int i = 7; i >= 0; i--) {
					r = r * 2u + sector_test_fma(w0[i], static_cast<uint32_t>(h[lo + i] >> 32));
				}
				return r << lo;
			};
			lookups(0);
			uint32_t cnt[NCNT];
#pragma unroll
			for (int j = 0; j < NCNT; j++) {
				cnt[j] = 0;
			}
#pragma unroll
			for (int i = 0; i < 16; i++) {
				const uint32_t sector = static_cast<uint32_t>(h[i]) & sector_mask;
				const uint32_t sh = (sector >> (SHIFT - 3)) & ((P - 1u) << 3); // 8 · slice
				const uint32_t one = MASKED ? (alive >> i) & 1u : 1u; // dropped keys and rows past the end are not counted
				if (NCNT == 1) {
					cnt[0] += one << sh;
				} else {
					cnt[0] += shl_clamp(one, sh);
					cnt[1] += shl_clamp(one, sh ^ 32u);
				}
			}
			// warp scan on 16-bit fields; base[j] = exclusive prefixes of slices 4j..4j+3 as one 64-bit word
			uint64_t base[NCNT];
			uint32_t tot[NCNT][2]; // warp totals, 16-bit fields: [j][0] = slices 4j, 4j+2; [j][1] = slices 4j+1, 4j+3
			bool overflow = false;
#pragma unroll
			for (int j = 0; j < NCNT; j++) {
				const uint32_t own0 = cnt[j] & 0x00ff00ffu, own1 = (cnt[j] >> 8) & 0x00ff00ffu;
				uint32_t e0 = own0, e1 = own1;
#pragma unroll
				for (int d = 1; d < 32; d <<= 1) {
					const uint32_t t0 = __shfl_up_sync(0xffffffffu, e0, d);
					const uint32_t t1 = __shfl_up_sync(0xffffffffu, e1, d);
					e0 += lane >= static_cast<uint32_t>(d) ? t0 : 0u;
					e1 += lane >= static_cast<uint32_t>(d) ? t1 : 0u;
				}
				tot[j][0] = __shfl_sync(0xffffffffu, e0, 31);
				tot[j][1] = __shfl_sync(0xffffffffu, e1, 31);
				e0 -= own0;
				e1 -= own1;
				base[j] = (static_cast<uint64_t>((e0 >> 16) | (e1 & 0xffff0000u)) << 32) | ((e0 & 0xffffu) | (e1 << 16));
				cnt[j] = 0; // reused below as the lane's running per-slice rank
			}
			// total of slice s (warp-uniform)
			auto slice_total = [&](int s) -> uint32_t {
				const uint32_t t = tot[s >> 2][s & 1];
				return (s & 2) ? t >> 16 : t & 0xffffu;
			};
#pragma unroll
			for (int s = FIRST; s < P; s++) { // a bucket region holds CAP records (INLINE0: slice 0 has none)
				overflow |= slice_total(s) > static_cast<uint32_t>(CAP);
			}
			if (lane == 0) {
#pragma unroll
				for (int s = FIRST; s < P; s++) {
					bucket_counts[(s - FIRST) * num_segs + cur] =
					    overflow ? uint16_t(0) : static_cast<uint16_t>(slice_total(s));
				}
			}
			uint32_t hits = 0;
			if (!overflow) { // warp-uniform; the usual case
				// pass 2: records into the stage.  position = exclusive base of the slice + rank inside the lane
				auto records = [&](const int lo) {
#pragma unroll
					for (int k = 0; k < 8; k++) {
						const int i = lo + k;
						const uint32_t sector = static_cast<uint32_t>(h[i]) & sector_mask;
						const uint32_t sh = (sector >> (SHIFT - 3)) & ((P - 1u) << 3);
						const uint32_t shr = sh & 31u;
						const uint32_t one = MASKED ? (alive >> i) & 1u : 1u;
						uint32_t rank, pos;
						if (NCNT == 1) {
							rank = (cnt[0] >> shr) & 0xffu;
							cnt[0] += one << shr;
							pos = static_cast<uint32_t>(base[0] >> (2u * shr)) & 0xffffu;
						} else {
							const bool up = sh >= 32u;
							rank = ((up ? cnt[1] : cnt[0]) >> shr) & 0xffu;
							cnt[0] += up ? 0u : one << shr;
							cnt[1] += up ? one << shr : 0u;
							pos = static_cast<uint32_t>((up ? base[1] : base[0]) >> (2u * shr)) & 0xffffu;
						}
						pos += rank;
						if ((!INLINE0 || sh != 0u) && one) {
							const uint32_t row9 =
							    lane * 16u + (((i / Geo::KPU + unit_rot) & (Geo::U - 1)) * Geo::KPU) + (i % Geo::KPU);
							// byte offset of record `pos` of slice s: ((s-FIRST)·CAP + pos)·8 with 8·s = sh
							const uint32_t off = mul_lo_u32(sh, CAP) + pos * 8u - FIRST * CAP * 8u;
							*reinterpret_cast<uint2 *>(reinterpret_cast<unsigned char *>(stage) + off) =
							    make_uint2((sector & ((1u << SHIFT) - 1u)) | (row9 << SHIFT),
							               static_cast<uint32_t>(h[i] >> 32));
						}
					}
				};
				records(0);
				hits = tests(0);
				lookups(8);
				records(8);
				hits |= tests(8);
			} else { // skewed keys: some region would overflow — this segment is looked up on the spot instead
#pragma unroll
				for (int i = 0; i < 16; i++) {
					const uint32_t sector = static_cast<uint32_t>(h[i]) & sector_mask;
					hits |= sector_test(__ldg(filter + sector), static_cast<uint32_t>(h[i] >> 32)) << i;
				}
				// Under an overlapped multi-GPU merge (slice gates) this kernel may only trust slice 0: the lookups
				// above may have read slices whose merge is still in flight, so the segment is also put on the list
				// that probe_overflow_kernel re-evaluates behind every gate.
				if (ovf_list != nullptr && lane == 0) {
					ovf_list[1u + atomicAdd(ovf_list, 1u)] = static_cast<uint32_t>(cur);
				}
			}
			hits &= valid;
			if (staged_seg) {
				const uint32_t twice = hits | (hits << 16);
				hits = (twice >> (16u - Geo::KPU * rot)) & 0xffffu;
			}
			bitmap16[cur * 32 + lane] = static_cast<uint16_t>(hits);
			// copy-out: the warp's records of every slice leave as contiguous 16-byte stores
			__syncwarp();
			if (!overflow) {
#pragma unroll
				for (int s = FIRST; s < P; s++) {
					const uint32_t total = slice_total(s);
					const uint4 *sv = reinterpret_cast<const uint4 *>(stage + (s - FIRST) * CAP);
					uint4 *dv = reinterpret_cast<uint4 *>(buckets + (s - FIRST) * bucket_stride + cur * CAP);
					for (uint32_t j = lane; 2u * j < total; j += 32u) {
						// an odd count copies one stale record: it stays inside the region, unread.  evict_first: the
						// record traffic must not push slice 0 of the filter out of L2
						st_stream_v4(dv + j, sv[j], pol_keys);
					}
				}
			}
			__syncwarp(); // the stage is rewritten by the next segment (or, aliased, by the key prefetch after it)
		}
		if (nxt >= num_segs) {
			break;
		}
		cur = nxt;
		par ^= 1u;
	}
	if (STAGED) {
		cp_async_wait<0>();
	}
}

// Segments the partition kernel could not place (a region would have overflowed: skewed keys): warp per listed
// segment, every key hashed again from the source and looked up wherever its sector lies.  Runs after the bucket
// kernels, i.e. behind every slice gate; the hot sectors of skewed keys are L2 hits.
template <class Src>
__global__ void __launch_bounds__(256)
    probe_overflow_kernel(const Src src, const uint64_t *__restrict__ filter, const uint32_t sector_mask,
                          const uint64_t count, const uint32_t *__restrict__ overflow, uint16_t *__restrict__ bitmap16) {
	const uint32_t lane = threadIdx.x & 31u;
	const uint64_t pol_keys = policy_evict_first();
	const uint32_t n = __ldg(overflow);
	const uint32_t warp_global = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
	const uint32_t num_warps = (gridDim.x * blockDim.x) >> 5;
	for (uint32_t i = warp_global; i < n; i += num_warps) {
		const uint64_t seg = __ldg(overflow + 1 + i);
		const uint64_t row0 = seg * kSegRows + lane * 16;
		uint32_t hits = 0;
#pragma unroll 4
		for (int k = 0; k < 16; k++) {
			if (row0 + k < count) {
				const uint64_t h = src.hash_at(row0 + k, pol_keys);
				const uint32_t sector = static_cast<uint32_t>(h) & sector_mask;
				hits |= sector_test(__ldg(filter + sector), static_cast<uint32_t>(h >> 32)) << k;
			}
		}
		bitmap16[seg * 32 + lane] = static_cast<uint16_t>(hits);
	}
}

// One slice's bucket: warp per segment region; all record loads, then all filter loads, are issued before the
// first test.  ITER = ceil(CAP / 32) records per lane.
template <int CAP, int SHIFT = kSliceShift>
__global__ void __launch_bounds__(256)
    probe_bucket_kernel(const uint2 *__restrict__ bucket, const uint16_t *__restrict__ counts, const uint64_t num_segs,
                        const uint64_t *__restrict__ filter, const uint32_t slice_lo, uint32_t *__restrict__ bitmap32) {
	constexpr int ITER = (CAP + 31) / 32;
	const uint32_t lane = threadIdx.x & 31u;
	const uint64_t pol_rec = policy_evict_first();
	const uint64_t pol_filter = policy_evict_last();
	warm_slice(filter, slice_lo, 1u << SHIFT, 256);
	const uint64_t *const slice = filter + slice_lo;
	const uint64_t warp_global = (static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
	const uint64_t num_warps = (static_cast<uint64_t>(gridDim.x) * blockDim.x) >> 5;
	for (uint64_t seg = warp_global; seg < num_segs; seg += num_warps) {
		const uint32_t c = __ldg(counts + seg);
		const uint2 *rec_base = bucket + seg * CAP;
		uint2 rec[ITER];
#pragma unroll
		for (int k = 0; k < ITER; k++) {
			const uint32_t j = lane + 32u * k;
			rec[k] = make_uint2(0u, 0u);
			if (j < c) {
				const uint64_t v = ld_stream_u64(rec_base + j, pol_rec);
				rec[k] = make_uint2(static_cast<uint32_t>(v), static_cast<uint32_t>(v >> 32));
			}
		}
		uint64_t w[ITER];
#pragma unroll
		for (int k = 0; k < ITER; k++) {
			w[k] = 0;
			if (lane + 32u * k < c) {
				w[k] = ld_filter_word(slice + (rec[k].x & ((1u << SHIFT) - 1u)), pol_filter);
			}
		}
#pragma unroll
		for (int k = 0; k < ITER; k++) {
			if (sector_test_fma(w[k], rec[k].y)) { // an all-zero word (no record) passes nothing
				const uint32_t row9 = rec[k].x >> SHIFT;
				atomicOr(bitmap32 + seg * 16 + (row9 >> 5), 1u << (row9 & 31u));
			}
		}
	}
}

// folded[i] = OR over j of words[i + j·fold_sectors] (i < fold_sectors), and the number of set bits of the folded
// filter (→ the share of non-members that will survive it).  128-bit loads, one slice after the other per thread.
__global__ void __launch_bounds__(256)
    fold_kernel(const ulonglong2 *__restrict__ words, const uint64_t fold_vecs, const uint32_t nslices,
                ulonglong2 *__restrict__ folded, unsigned long long *__restrict__ popcount_out) {
	__shared__ unsigned long long s_pop[8];
	unsigned long long pop = 0;
	const uint64_t step = static_cast<uint64_t>(gridDim.x) * blockDim.x;
	for (uint64_t i = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < fold_vecs; i += step) {
		ulonglong2 acc = make_ulonglong2(0ULL, 0ULL);
#pragma unroll 4
		for (uint32_t j = 0; j < nslices; j++) {
			const ulonglong2 v = __ldcs(words + i + j * fold_vecs);
			acc.x |= v.x;
			acc.y |= v.y;
		}
		folded[i] = acc;
		pop += __popcll(acc.x) + __popcll(acc.y);
	}
#pragma unroll
	for (int d = 16; d > 0; d >>= 1) {
		pop += __shfl_xor_sync(0xffffffffu, pop, d);
	}
	if ((threadIdx.x & 31u) == 0) {
		s_pop[threadIdx.x >> 5] = pop;
	}
	__syncthreads();
	if (threadIdx.x == 0) {
		unsigned long long t = 0;
#pragma unroll
		for (int w = 0; w < 8; w++) {
			t += s_pop[w];
		}
		atomicAdd(popcount_out, t);
	}
}

// ---- pre-filtered probe of flat key columns (regime 3c): fold → survivors → per-slice lookups --------------------
// probe_fold_kernel   the staging of probe_smem_kernel (bulk copies of HALF segments, one mbarrier per buffer, 4 KiB
//                     per warp) with the lookup going to the L2-resident FOLDED filter instead of shared memory.  A
//                     lane owns 8 keys of a half; the keys that survive the folded filter (members + its false
//                     positives, ≈ 17 % at config 2) become 8-byte records {sector-in-slice : 23 | row-in-half : 8,
//                     hash bits 63..32}, counting-sorted by slice inside the warp (8-bit per-lane counts, one warp scan
//                     on 16-bit fields) into the half buffer just consumed, PACKED slice after slice, and copied out to
//                     the half's 256-record region; ends[half·P + s] = end of slice s's records inside the region.
//                     No region can overflow (a half has 256 keys), so there is no fallback path.
// probe_bucket_half_kernel  once per slice: the slice's records of four halves at a time → real filter word → test →
//                     RED.OR into the row-ordered bitmap (which probe_fold_kernel zeroed).
// ncu of the first pre-filtered version (probe_partition_kernel with a folded test in front, commit 4d0fd14): 72
// instructions per key, 15 warps per SM, issue 36 % — latency-bound, 5.4 ms per 10^9 keys.  This one: 8 keys per
// lane and no 16-hash register block, so 24 warps per SM fit: 4.7 ms, L1TEX 81 % busy (profiles/ncu_fold_r2u.md).
template <int P>
struct FoldCfg {
	static constexpr int NCNT = P <= 4 ? 1 : 2; // per-lane counters: 8 bits per slice, four slices per register
};

template <class Src, int NWARPS, int P>
__global__ void __launch_bounds__(NWARPS * 32, 1)
    probe_fold_kernel(const Src src, const uint64_t *__restrict__ folded, const uint32_t fold_mask,
                      const uint32_t sector_mask, const uint64_t count, const uint64_t num_segs,
                      uint8_t *__restrict__ bitmap8, uint2 *__restrict__ records /* 256 per half */,
                      uint16_t *__restrict__ ends /* P per half */, unsigned long long *__restrict__ pre_stats) {
	static_assert(Src::kStageable, "flat key columns only");
	constexpr int KB = Src::kKeyBytes;
	constexpr int KPU = 16 / KB;                    // keys per 16-byte unit
	constexpr uint32_t HALF_BYTES = kHalfRows * KB; // 2048 / 1024
	constexpr int LU = 8 / KPU;                     // 16-byte units a lane owns per half: 4 / 2
	constexpr uint32_t STAGE_BYTES = kHalfRows * 8; // up to 256 records
	constexpr bool ALIAS = HALF_BYTES >= STAGE_BYTES; // the records overwrite the keys they came from
	constexpr uint32_t WARP_BYTES = 2 * HALF_BYTES + (ALIAS ? 0 : STAGE_BYTES);
	constexpr int NCNT = FoldCfg<P>::NCNT;
	constexpr uint32_t SHIFT = kSliceShift;
	extern __shared__ __align__(128) unsigned char s_dyn[];
	__shared__ __align__(8) uint64_t s_bars[NWARPS][2];
	const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
	unsigned char *const kbuf = s_dyn + warp * WARP_BYTES;
	const uint64_t pol_keys = policy_evict_first();
	const uint64_t pol_filter = policy_evict_last();
	if (lane == 0) {
		mbar_init(&s_bars[warp][0], 1);
		mbar_init(&s_bars[warp][1], 1);
	}
	asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
	__syncthreads();
	warm_slice(folded, 0u, fold_mask + 1u, NWARPS * 32);
	const uint64_t full_segs = count / kSegRows;
	const uint64_t stride = static_cast<uint64_t>(gridDim.x) * NWARPS;
	const unsigned char *const keys = reinterpret_cast<const unsigned char *>(src.key_ptr());
	auto issue = [&](uint64_t seg, uint32_t h) { // elected lane: stage half h of segment seg
		if (seg < full_segs) {
			mbar_expect_tx(&s_bars[warp][h], HALF_BYTES);
			bulk_g2s(kbuf + h * HALF_BYTES, keys + (seg * kSegRows + h * kHalfRows) * KB, HALF_BYTES, &s_bars[warp][h],
			         pol_keys);
		}
	};
	const uint64_t seg0 = static_cast<uint64_t>(blockIdx.x) * NWARPS + warp;
	if (lane == 0) {
		issue(seg0, 0);
		issue(seg0, 1);
	}
	const uint32_t rot = LU == 4 ? ((lane >> 1) & 3u) : ((lane >> 2) & 1u);
	uint32_t phase = 0;
	uint32_t seen_rows = 0, seen_alive = 0;
	for (uint64_t seg = seg0; seg < num_segs; seg += stride) {
#pragma unroll
		for (uint32_t h = 0; h < 2; h++) {
			const bool staged = seg < full_segs;
			uint64_t hh[8];
			uint32_t valid = 0xffu; // slot order
			if (staged) {
				mbar_wait(&s_bars[warp][h], (phase >> h) & 1u);
				phase ^= 1u << h;
				const uint4 *mine = reinterpret_cast<const uint4 *>(kbuf + h * HALF_BYTES) + lane * LU;
#pragma unroll
				for (int u = 0; u < LU; u++) {
					Src::hash_unit(mine[(u + rot) & (LU - 1)], hh + u * KPU);
				}
				if (Src::kNullable) {
					const uint32_t v = __ldg(reinterpret_cast<const unsigned char *>(src.validity) + seg * 64 + h * 32 + lane);
					const uint32_t twice = v | (v << 8); // slot i holds logical bit (i + KPU·rot) mod 8: rotate right
					null_fix<8>(hh, (twice >> (KPU * rot)) & 0xffu, 0);
				}
			} else { // the partial last segment: straight from global memory, slots in row order
				const uint64_t row0 = seg * kSegRows + h * kHalfRows + lane * 8u;
				valid = 0;
#pragma unroll
				for (int i = 0; i < 8; i++) {
					hh[i] = 0;
					if (row0 + i < count) {
						hh[i] = src.hash_at(row0 + i, pol_keys);
						valid |= 1u << i;
					}
				}
			}
			const uint32_t alive = test_batch<8, false, false>(hh, folded, nullptr, fold_mask, pol_filter, 0u, 0u) & valid;
			seen_rows += __popc(valid);
			seen_alive += __popc(alive);
			const uint64_t half = seg * 2 + h;
			bitmap8[half * 32 + lane] = 0; // the bucket kernels OR the hits in
			if (ALIAS && staged) {
				__syncwarp(); // every lane has its keys in registers: the buffer becomes the record stage
			}
			uint2 *const stage = reinterpret_cast<uint2 *>(kbuf + (ALIAS ? h * HALF_BYTES : 2 * HALF_BYTES));
			// per-lane counts of the survivors of every slice (8-bit fields)
			uint32_t cnt[NCNT];
#pragma unroll
			for (int j = 0; j < NCNT; j++) {
				cnt[j] = 0;
			}
#pragma unroll
			for (int i = 0; i < 8; i++) {
				const uint32_t sector = static_cast<uint32_t>(hh[i]) & sector_mask;
				const uint32_t sh = (sector >> (SHIFT - 3)) & ((P - 1u) << 3); // 8 · slice
				const uint32_t one = (alive >> i) & 1u;
				if (NCNT == 1) {
					cnt[0] += one << sh;
				} else {
					cnt[0] += shl_clamp(one, sh);
					cnt[1] += shl_clamp(one, sh ^ 32u);
				}
			}
			// warp scan on 16-bit fields; base[j] = exclusive prefixes of slices 4j..4j+3 as one 64-bit word
			uint64_t base[NCNT];
			uint32_t tot[NCNT][2];
#pragma unroll
			for (int j = 0; j < NCNT; j++) {
				const uint32_t own0 = cnt[j] & 0x00ff00ffu, own1 = (cnt[j] >> 8) & 0x00ff00ffu;
				uint32_t e0 = own0, e1 = own1;
#pragma unroll
				for (int d = 1; d < 32; d <<= 1) {
					const uint32_t t0 = __shfl_up_sync(0xffffffffu, e0, d);
					const uint32_t t1 = __shfl_up_sync(0xffffffffu, e1, d);
					e0 += lane >= static_cast<uint32_t>(d) ? t0 : 0u;
					e1 += lane >= static_cast<uint32_t>(d) ? t1 : 0u;
				}
				tot[j][0] = __shfl_sync(0xffffffffu, e0, 31);
				tot[j][1] = __shfl_sync(0xffffffffu, e1, 31);
				e0 -= own0;
				e1 -= own1;
				base[j] = (static_cast<uint64_t>((e0 >> 16) | (e1 & 0xffff0000u)) << 32) | ((e0 & 0xffffu) | (e1 << 16));
				cnt[j] = 0; // reused below as the lane's running per-slice rank
			}
			auto slice_total = [&](int s) -> uint32_t {
				const uint32_t t = tot[s >> 2][s & 1];
				return (s & 2) ? t >> 16 : t & 0xffffu;
			};
			// the slices' records are packed one behind the other: slice s starts where slice s-1 ends (warp-uniform)
			uint32_t run = 0;
			uint64_t endw[NCNT];
#pragma unroll
			for (int j = 0; j < NCNT; j++) {
				uint64_t add = 0;
				endw[j] = 0;
#pragma unroll
				for (int k = 0; k < 4 && 4 * j + k < P; k++) {
					add |= static_cast<uint64_t>(run) << (16 * k);
					run += slice_total(4 * j + k);
					endw[j] |= static_cast<uint64_t>(run) << (16 * k);
				}
				base[j] += add; // no carries: prefix + start ≤ 256
			}
			if (lane == 0) {
				uint64_t *e = reinterpret_cast<uint64_t *>(ends + half * P); // P = 4 or 8: 8-byte aligned
#pragma unroll
				for (int j = 0; j < NCNT; j++) {
					e[j] = endw[j];
				}
			}
			const uint32_t unit_rot = staged ? rot : 0u;
#pragma unroll
			for (int i = 0; i < 8; i++) {
				const uint32_t sector = static_cast<uint32_t>(hh[i]) & sector_mask;
				const uint32_t sh = (sector >> (SHIFT - 3)) & ((P - 1u) << 3);
				const uint32_t shr = sh & 31u;
				const uint32_t one = (alive >> i) & 1u;
				uint32_t rank, pos;
				if (NCNT == 1) {
					rank = (cnt[0] >> shr) & 0xffu;
					cnt[0] += one << shr;
					pos = static_cast<uint32_t>(base[0] >> (2u * shr)) & 0xffffu;
				} else {
					const bool up = sh >= 32u;
					rank = ((up ? cnt[1] : cnt[0]) >> shr) & 0xffu;
					cnt[0] += up ? 0u : one << shr;
					cnt[1] += up ? one << shr : 0u;
					pos = static_cast<uint32_t>((up ? base[1] : base[0]) >> (2u * shr)) & 0xffffu;
				}
				if (one) {
					// (holding a lane's first two survivors back and storing them with two warp-wide instructions instead
					// of eight nearly empty ones was measured: 6.46 against 6.26 ms per 10^9 keys — not adopted)
					const uint32_t row8 = lane * 8u + (((i / KPU + unit_rot) & (LU - 1)) * KPU) + (i % KPU);
					stage[pos + rank] =
					    make_uint2((sector & ((1u << SHIFT) - 1u)) | (row8 << SHIFT), static_cast<uint32_t>(hh[i] >> 32));
				}
			}
			__syncwarp();
			{
				// (copying every slice's run to its own 64-byte-aligned start — to save the bucket kernels the DRAM atom
				// a misaligned run costs — makes this kernel pay more than they save: 6.64 against 6.26 ms per 10^9 keys)
				uint2 *dv = records + half * kHalfRows;
				for (uint32_t j = lane; j < run; j += 32u) {
					st_stream_v2(dv + j, stage[j], pol_keys);
				}
			}
			__syncwarp(); // the stage has been read: the buffer may be refilled (aliased) / rewritten by the next half
			if (staged && lane == 0) {
				issue(seg + stride, h);
			}
		}
	}
	if (pre_stats != nullptr) {
		seen_rows = __reduce_add_sync(0xffffffffu, seen_rows);
		seen_alive = __reduce_add_sync(0xffffffffu, seen_alive);
		if (lane == 0) {
			atomicAdd(pre_stats, static_cast<unsigned long long>(seen_rows));
			atomicAdd(pre_stats + 1, static_cast<unsigned long long>(seen_alive));
		}
	}
}

// probe_fold_direct_kernel: the same step with NO shared memory at all.  What bounds the folded-filter gathers is the
// number of them an SM keeps in flight, and that number follows the L1's size: every outstanding line holds an L1
// line (measured: 217 KiB of shared memory per SM → 16.5 ms per 10^9 keys, 128 KiB → 5.4 ms, 96 KiB → 4.7 ms —
// each time ≈ (L1 lines) / latency).  So the keys come straight into registers (lane l owns rows l, l+32, …, l+224
// of a half: eight coalesced loads, all in flight before the first hash) and the survivors' records leave as
// 8-byte stores at their sorted positions (the L2 merges a region's stores).  Any key source; used for the generic
// ones (several columns, selections, dictionaries), which cannot be staged by bulk copies: two key columns, 4·10^8
// rows, F = 256 MiB: 6.3 ms against 9.2 ms without the pre-filter (profiles/prefilter_r2v.txt).  For flat columns
// probe_fold_kernel is faster (6.2 against 6.5 ms per 10^9 rows).
template <class Src, int P, int MINB>
__global__ void __launch_bounds__(256, MINB)
    probe_fold_direct_kernel(const Src src, const uint64_t *__restrict__ folded, const uint32_t fold_mask,
                             const uint32_t sector_mask, const uint64_t count, const uint64_t num_halves,
                             uint32_t *__restrict__ bitmap32, uint2 *__restrict__ records /* 256 per half */,
                             uint16_t *__restrict__ ends /* P per half */, unsigned long long *__restrict__ pre_stats) {
	constexpr int NCNT = FoldCfg<P>::NCNT;
	constexpr uint32_t SHIFT = kSliceShift;
	const uint32_t lane = threadIdx.x & 31u;
	const uint64_t pol_keys = policy_evict_first();
	const uint64_t pol_filter = policy_evict_last();
	warm_slice(folded, 0u, fold_mask + 1u, 256);
	const uint64_t warp_global = (static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
	const uint64_t num_warps = (static_cast<uint64_t>(gridDim.x) * blockDim.x) >> 5;
	uint32_t seen_rows = 0, seen_alive = 0;
	for (uint64_t half = warp_global; half < num_halves; half += num_warps) {
		const uint64_t row0 = half * kHalfRows + lane;
		uint64_t hh[8];
		uint32_t valid = 0xffu;
		if ((half + 1) * kHalfRows <= count) { // warp-uniform
#pragma unroll
			for (int i = 0; i < 8; i++) {
				hh[i] = src.hash_at(row0 + 32u * i, pol_keys);
			}
		} else {
			valid = 0;
#pragma unroll
			for (int i = 0; i < 8; i++) {
				hh[i] = 0;
				if (row0 + 32u * i < count) {
					hh[i] = src.hash_at(row0 + 32u * i, pol_keys);
					valid |= 1u << i;
				}
			}
		}
		const uint32_t alive = test_batch<8, false, false>(hh, folded, nullptr, fold_mask, pol_filter, 0u, 0u) & valid;
		seen_rows += __popc(valid);
		seen_alive += __popc(alive);
		if (lane < 8) {
			bitmap32[half * 8 + lane] = 0; // the bucket kernels OR the hits in
		}
		uint32_t cnt[NCNT];
#pragma unroll
		for (int j = 0; j < NCNT; j++) {
			cnt[j] = 0;
		}
#pragma unroll
		for (int i = 0; i < 8; i++) {
			const uint32_t sector = static_cast<uint32_t>(hh[i]) & sector_mask;
			const uint32_t sh = (sector >> (SHIFT - 3)) & ((P - 1u) << 3); // 8 · slice
			const uint32_t one = (alive >> i) & 1u;
			if (NCNT == 1) {
				cnt[0] += one << sh;
			} else {
				cnt[0] += shl_clamp(one, sh);
				cnt[1] += shl_clamp(one, sh ^ 32u);
			}
		}
		uint64_t base[NCNT];
		uint32_t tot[NCNT][2];
#pragma unroll
		for (int j = 0; j < NCNT; j++) {
			const uint32_t own0 = cnt[j] & 0x00ff00ffu, own1 = (cnt[j] >> 8) & 0x00ff00ffu;
			uint32_t e0 = own0, e1 = own1;
#pragma unroll
			for (int d = 1; d < 32; d <<= 1) {
				const uint32_t t0 = __shfl_up_sync(0xffffffffu, e0, d);
				const uint32_t t1 = __shfl_up_sync(0xffffffffu, e1, d);
				e0 += lane >= static_cast<uint32_t>(d) ? t0 : 0u;
				e1 += lane >= static_cast<uint32_t>(d) ? t1 : 0u;
			}
			tot[j][0] = __shfl_sync(0xffffffffu, e0, 31);
			tot[j][1] = __shfl_sync(0xffffffffu, e1, 31);
			e0 -= own0;
			e1 -= own1;
			base[j] = (static_cast<uint64_t>((e0 >> 16) | (e1 & 0xffff0000u)) << 32) | ((e0 & 0xffffu) | (e1 << 16));
			cnt[j] = 0;
		}
		auto slice_total = [&](int s) -> uint32_t {
			const uint32_t t = tot[s >> 2][s & 1];
			return (s & 2) ? t >> 16 : t & 0xffffu;
		};
		uint32_t run = 0;
		uint64_t endw[NCNT];
#pragma unroll
		for (int j = 0; j < NCNT; j++) {
			uint64_t add = 0;
			endw[j] = 0;
#pragma unroll
			for (int k = 0; k < 4 && 4 * j + k < P; k++) {
				add |= static_cast<uint64_t>(run) << (16 * k);
				run += slice_total(4 * j + k);
				endw[j] |= static_cast<uint64_t>(run) << (16 * k);
			}
			base[j] += add;
		}
		if (lane == 0) {
			uint64_t *e = reinterpret_cast<uint64_t *>(ends + half * P);
#pragma unroll
			for (int j = 0; j < NCNT; j++) {
				e[j] = endw[j];
			}
		}
		uint2 *const region = records + half * kHalfRows;
#pragma unroll
		for (int i = 0; i < 8; i++) {
			const uint32_t sector = static_cast<uint32_t>(hh[i]) & sector_mask;
			const uint32_t sh = (sector >> (SHIFT - 3)) & ((P - 1u) << 3);
			const uint32_t shr = sh & 31u;
			const uint32_t one = (alive >> i) & 1u;
			uint32_t rank, pos;
			if (NCNT == 1) {
				rank = (cnt[0] >> shr) & 0xffu;
				cnt[0] += one << shr;
				pos = static_cast<uint32_t>(base[0] >> (2u * shr)) & 0xffffu;
			} else {
				const bool up = sh >= 32u;
				rank = ((up ? cnt[1] : cnt[0]) >> shr) & 0xffu;
				cnt[0] += up ? 0u : one << shr;
				cnt[1] += up ? one << shr : 0u;
				pos = static_cast<uint32_t>((up ? base[1] : base[0]) >> (2u * shr)) & 0xffffu;
			}
			if (one) {
				const uint32_t row8 = lane + 32u * i;
				st_stream_v2(region + pos + rank,
				             make_uint2((sector & ((1u << SHIFT) - 1u)) | (row8 << SHIFT), static_cast<uint32_t>(hh[i] >> 32)),
				             pol_keys);
			}
		}
	}
	if (pre_stats != nullptr) {
		seen_rows = __reduce_add_sync(0xffffffffu, seen_rows);
		seen_alive = __reduce_add_sync(0xffffffffu, seen_alive);
		if (lane == 0) {
			atomicAdd(pre_stats, static_cast<unsigned long long>(seen_rows));
			atomicAdd(pre_stats + 1, static_cast<unsigned long long>(seen_alive));
		}
	}
}

// One slice's survivors.  A slice's share of a half is a dozen records, and the kernel is nothing but a chain of
// three dependent loads (bounds → records → filter words), so a warp step covers SIXTEEN halves: eight lanes per half,
// four halves per round, four rounds, up to 24 records of each half in flight (K = 3); a half with more records of
// the slice (rich in members) loops.  Earlier versions, one half per 32 lanes and four halves per step: 0.56 ms per
// slice at 32 warps per SM, 0.46 ms at 64 (profiles/prefilter_r2r.txt, prefilter_r2s.txt) — latency-bound.
template <int P>
__global__ void __launch_bounds__(256, 3)
    probe_bucket_half_kernel(const uint2 *__restrict__ records, const uint16_t *__restrict__ ends, const uint64_t num_halves,
                             const uint64_t *__restrict__ filter, const uint32_t s /* slice */,
                             uint32_t *__restrict__ bitmap32) {
	constexpr int T = 4, K = 3;
	constexpr uint32_t SHIFT = kSliceShift;
	const uint32_t lane = threadIdx.x & 31u;
	const uint32_t group = lane >> 3, q = lane & 7u;
	const uint64_t pol_rec = policy_evict_first();
	const uint64_t pol_filter = policy_evict_last();
	warm_slice(filter, s << SHIFT, 1u << SHIFT, 256);
	const uint64_t *const slice = filter + (static_cast<uint64_t>(s) << SHIFT);
	const uint64_t warp_global = (static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
	const uint64_t num_warps = (static_cast<uint64_t>(gridDim.x) * blockDim.x) >> 5;
	auto apply = [&](uint64_t half, uint64_t v, uint64_t word) {
		const uint32_t x = static_cast<uint32_t>(v);
		if (sector_test_fma(word, static_cast<uint32_t>(v >> 32))) { // an all-zero word (no record) passes nothing
			const uint32_t row8 = x >> SHIFT;
			atomicOr(bitmap32 + half * 8 + (row8 >> 5), 1u << (row8 & 31u));
		}
	};
	// the bounds of the NEXT step are loaded one step ahead (lo | hi << 16): two dependent loads per step, not three
	uint32_t lh[T];
	auto bounds = [&](uint64_t h0, uint32_t (&b)[T]) {
#pragma unroll
		for (int t = 0; t < T; t++) {
			const uint64_t half = h0 + t * 4 + group;
			b[t] = 0;
			if (half < num_halves) {
				const uint32_t h = __ldg(ends + half * P + s);
				const uint32_t l = s ? __ldg(ends + half * P + s - 1) : 0u;
				b[t] = l | (h << 16);
			}
		}
	};
	bounds(warp_global * (4 * T), lh);
	for (uint64_t h0 = warp_global * (4 * T); h0 < num_halves; h0 += num_warps * (4 * T)) {
		uint32_t next[T];
		bounds(h0 + num_warps * (4 * T), next);
		uint64_t rec[T][K];
#pragma unroll
		for (int t = 0; t < T; t++) {
#pragma unroll
			for (int k = 0; k < K; k++) {
				const uint32_t j = (lh[t] & 0xffffu) + q + 8u * k;
				rec[t][k] = 0;
				if (j < (lh[t] >> 16)) {
					rec[t][k] = ld_stream_u64(records + (h0 + t * 4 + group) * kHalfRows + j, pol_rec);
				}
			}
		}
		uint64_t w[T][K];
#pragma unroll
		for (int t = 0; t < T; t++) {
#pragma unroll
			for (int k = 0; k < K; k++) {
				w[t][k] = 0;
				if ((lh[t] & 0xffffu) + q + 8u * k < (lh[t] >> 16)) {
					w[t][k] = ld_filter_word(slice + (static_cast<uint32_t>(rec[t][k]) & ((1u << SHIFT) - 1u)), pol_filter);
				}
			}
		}
#pragma unroll
		for (int t = 0; t < T; t++) {
#pragma unroll
			for (int k = 0; k < K; k++) {
				apply(h0 + t * 4 + group, rec[t][k], w[t][k]);
			}
		}
#pragma unroll
		for (int t = 0; t < T; t++) {
			for (uint32_t j = (lh[t] & 0xffffu) + q + 8u * K; j < (lh[t] >> 16); j += 8u) { // halves rich in members
				const uint64_t half = h0 + t * 4 + group;
				const uint64_t v = ld_stream_u64(records + half * kHalfRows + j, pol_rec);
				apply(half, v, ld_filter_word(slice + (static_cast<uint32_t>(v) & ((1u << SHIFT) - 1u)), pol_filter));
			}
		}
#pragma unroll
		for (int t = 0; t < T; t++) {
			lh[t] = next[t];
		}
	}
}

// seg_counts[seg] = number of set bits among the segment's 512 bitmap bits
__global__ void __launch_bounds__(256)
    count_segments_kernel(const uint4 *__restrict__ bitmap128, const uint64_t num_segs, uint32_t *__restrict__ seg_counts) {
	const uint64_t step = static_cast<uint64_t>(gridDim.x) * blockDim.x;
	for (uint64_t seg = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x; seg < num_segs; seg += step) {
		uint32_t n = 0;
#pragma unroll
		for (int q = 0; q < 4; q++) {
			const uint4 v = __ldg(bitmap128 + seg * 4 + q);
			n += __popc(v.x) + __popc(v.y) + __popc(v.z) + __popc(v.w);
		}
		seg_counts[seg] = n;
	}
}

// Exclusive prefix sums inside blocks of kScanBlock counts (in place) and each block's total.
__global__ void __launch_bounds__(1024) scan_blocks_kernel(uint32_t *__restrict__ counts, const uint64_t n,
                                                           uint32_t *__restrict__ block_totals) {
	__shared__ uint32_t s_part[32];
	const uint64_t base = static_cast<uint64_t>(blockIdx.x) * kScanBlock + threadIdx.x * 4ull;
	uint32_t v[4];
#pragma unroll
	for (int j = 0; j < 4; j++) {
		v[j] = base + j < n ? counts[base + j] : 0u;
	}
	const uint32_t sum = v[0] + v[1] + v[2] + v[3];
	const uint32_t lane = threadIdx.x & 31u;
	const uint32_t warp = threadIdx.x >> 5;
	uint32_t incl = sum;
#pragma unroll
	for (int d = 1; d < 32; d <<= 1) {
		const uint32_t t = __shfl_up_sync(0xffffffffu, incl, d);
		if (lane >= static_cast<uint32_t>(d)) {
			incl += t;
		}
	}
	if (lane == 31) {
		s_part[warp] = incl;
	}
	__syncthreads();
	if (warp == 0) {
		const uint32_t p = s_part[lane];
		uint32_t w = p;
#pragma unroll
		for (int d = 1; d < 32; d <<= 1) {
			const uint32_t t = __shfl_up_sync(0xffffffffu, w, d);
			if (lane >= static_cast<uint32_t>(d)) {
				w += t;
			}
		}
		s_part[lane] = w - p;
		if (lane == 31) {
			block_totals[blockIdx.x] = w;
		}
	}
	__syncthreads();
	uint32_t run = s_part[warp] + incl - sum;
#pragma unroll
	for (int j = 0; j < 4; j++) {
		if (base + j < n) {
			counts[base + j] = run;
		}
		run += v[j];
	}
}

// Exclusive prefix sum of the block totals in place (nblocks ≤ 2048: count < 2^32 keys), total → *found_out.
__global__ void __launch_bounds__(1024) scan_totals_kernel(uint32_t *__restrict__ block_totals, const uint32_t nblocks,
                                                           uint64_t *__restrict__ found_out) {
	__shared__ uint32_t s_part[32];
	const uint32_t i0 = threadIdx.x * 2u;
	const uint32_t a = i0 < nblocks ? block_totals[i0] : 0u;
	const uint32_t b = i0 + 1 < nblocks ? block_totals[i0 + 1] : 0u;
	const uint32_t sum = a + b;
	const uint32_t lane = threadIdx.x & 31u;
	const uint32_t warp = threadIdx.x >> 5;
	uint32_t incl = sum;
#pragma unroll
	for (int d = 1; d < 32; d <<= 1) {
		const uint32_t t = __shfl_up_sync(0xffffffffu, incl, d);
		if (lane >= static_cast<uint32_t>(d)) {
			incl += t;
		}
	}
	if (lane == 31) {
		s_part[warp] = incl;
	}
	__syncthreads();
	if (warp == 0) {
		const uint32_t p = s_part[lane];
		uint32_t w = p;
#pragma unroll
		for (int d = 1; d < 32; d <<= 1) {
			const uint32_t t = __shfl_up_sync(0xffffffffu, w, d);
			if (lane >= static_cast<uint32_t>(d)) {
				w += t;
			}
		}
		s_part[lane] = w - p;
		if (lane == 31) {
			*found_out = w;
		}
	}
	__syncthreads();
	const uint32_t excl = s_part[warp] + incl - sum;
	if (i0 < nblocks) {
		block_totals[i0] = excl;
	}
	if (i0 + 1 < nblocks) {
		block_totals[i0 + 1] = excl + a;
	}
}

// Bitmap → ordered selection vector.  A warp takes chunks of 1024 rows (two segments, 32 bitmap words, one per
// lane).  Each lane expands the set bits of its word into a warp-private shared-memory run (position = warp
// scan of the popcounts), then the warp copies the run to sel_out with full-width contiguous stores.  Writing
// sel_out directly from the per-lane loops costs one LSU tag cycle per lane and store (measured: 2.6 ms extra
// at 50 % selectivity); a per-word ballot loop keeps the stores contiguous but needs ≈30 instructions per word
// (measured 1.05 ms at any selectivity).  The run is XOR-swizzled by its 32-entry row so that regularly spaced
// lane offsets (≈16 apart at 50 % selectivity) do not collapse onto two banks.
__device__ __forceinline__ uint32_t run_swizzle(uint32_t i) {
	return i ^ ((i >> 5) & 31u);
}

__global__ void __launch_bounds__(256)
    compact_bitmap_kernel(const uint32_t *__restrict__ bitmap32, const uint32_t *__restrict__ seg_offsets,
                          const uint32_t *__restrict__ block_offsets, const uint64_t num_segs,
                          uint32_t *__restrict__ sel_out, const uint32_t row_offset,
                          const uint32_t *__restrict__ translate) {
	__shared__ uint32_t s_run[8][1024];
	const uint32_t lane = threadIdx.x & 31u;
	uint32_t *const run = s_run[threadIdx.x >> 5];
	const uint64_t num_chunks = (num_segs + 1) >> 1;
	const uint64_t warp_global = (static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
	const uint64_t num_warps = (static_cast<uint64_t>(gridDim.x) * blockDim.x) >> 5;
	// software prefetch: the next chunk's word and offsets are loaded while the current one is expanded
	auto load_word = [&](uint64_t c) -> uint32_t {
		// a chunk's second segment may not exist (odd num_segs): its words are not initialised
		return (c < num_chunks && (2 * c + (lane >> 4)) < num_segs) ? __ldg(bitmap32 + c * 32 + lane) : 0u;
	};
	auto load_off = [&](uint64_t c) -> uint32_t {
		return c < num_chunks ? __ldg(seg_offsets + 2 * c) + __ldg(block_offsets + (2 * c) / kScanBlock) : 0u;
	};
	uint64_t c = warp_global;
	uint32_t word = load_word(c), off = load_off(c);
	while (c < num_chunks) {
		const uint64_t cn = c + num_warps;
		const uint32_t word_n = load_word(cn), off_n = load_off(cn);
		const uint32_t cnt = __popc(word);
		uint32_t incl = cnt;
#pragma unroll
		for (int d = 1; d < 32; d <<= 1) {
			const uint32_t t = __shfl_up_sync(0xffffffffu, incl, d);
			if (lane >= static_cast<uint32_t>(d)) {
				incl += t;
			}
		}
		const uint32_t total = __shfl_sync(0xffffffffu, incl, 31);
		if (total) { // warp-uniform
			uint32_t p = incl - cnt;
			const uint32_t row_base = static_cast<uint32_t>(c * 1024) + lane * 32 + row_offset;
			uint32_t w = word;
			while (w) {
				run[run_swizzle(p++)] = row_base + (__ffs(w) - 1);
				w &= w - 1;
			}
			__syncwarp();
			for (uint32_t i = lane; i < total; i += 32) {
				const uint32_t row = run[run_swizzle(i)];
				sel_out[off + i] = translate ? __ldg(translate + row) : row;
			}
			__syncwarp();
		}
		c = cn;
		word = word_n;
		off = off_n;
	}
}

static inline uint64_t ceil_div(uint64_t a, uint64_t b);
// bitmap → selection vector: segment counts are in seg_offsets (exclusive, per scan block) and block_totals
static void launch_compact(const DeviceInfo &dev, const uint32_t *bitmap, const uint32_t *seg_offsets,
                           const uint32_t *block_totals, uint64_t num_segs, uint32_t *sel_out, uint32_t row_offset,
                           const uint32_t *translate, cudaStream_t stream) {
	// 8 warps per CTA, one chunk per warp iteration; 32 KiB of runs per CTA: 6 CTAs per SM.  (A variant that stores
	// the row ids straight from the lanes, without the staged run: 0.31 against 0.26 ms at 10 % survivors and 2.4 ms
	// slower at 50 % — profiles/prefilter_r2v.txt.)
	const uint64_t cwant = ceil_div((num_segs + 1) / 2, 8ull);
	const uint64_t ccap = static_cast<uint64_t>(dev.sm_count) * 6;
	compact_bitmap_kernel<<<static_cast<unsigned>(cwant < ccap ? cwant : ccap), 256, 0, stream>>>(
	    bitmap, seg_offsets, block_totals, num_segs, sel_out, row_offset, translate);
}

// Small probes (≤ kFusedMaxKeys): one launch; tiles are chained with a decoupled look-back.
template <class Src, int ROUNDS, int BLOCK, bool SMEM_FILTER>
__global__ void __launch_bounds__(BLOCK)
    probe_kernel(const Src src, const uint64_t *__restrict__ filter, const uint32_t sector_mask, const uint64_t count,
                 const uint64_t num_tiles, uint32_t *__restrict__ sel_out, uint64_t *__restrict__ found_out,
                 const uint32_t row_offset, const uint32_t *__restrict__ translate, uint64_t *desc,
                 unsigned int *ticket) {
	constexpr int NWARPS = BLOCK / 32;
	constexpr uint64_t WARP_KEYS = 64ull * ROUNDS;
	constexpr uint64_t TILE_KEYS = WARP_KEYS * NWARPS;

	extern __shared__ __align__(16) uint64_t s_filter[];
	__shared__ uint32_t s_warp_off[NWARPS];
	__shared__ uint32_t s_tile_excl;
	__shared__ uint32_t s_tile;

	const uint32_t lane = threadIdx.x & 31u;
	const uint32_t warp = threadIdx.x >> 5;
	const uint64_t pol_keys = policy_evict_first();
	const uint64_t pol_filter = policy_evict_last();

	if (SMEM_FILTER) {
		const uint64_t nvec = (static_cast<uint64_t>(sector_mask) + 1) >> 1;
		const ulonglong2 *g = reinterpret_cast<const ulonglong2 *>(filter);
		ulonglong2 *s = reinterpret_cast<ulonglong2 *>(s_filter);
		for (uint64_t i = threadIdx.x; i < nvec; i += BLOCK) {
			s[i] = __ldg(g + i);
		}
	}
	while (true) {
		if (threadIdx.x == 0) {
			s_tile = atomicAdd(ticket, 1u);
		}
		__syncthreads(); // also publishes s_filter on the first pass
		const uint64_t tile = s_tile;
		if (tile >= num_tiles) {
			break;
		}
		const uint64_t warp_base = tile * TILE_KEYS + warp * WARP_KEYS;
		const uint32_t hits = probe_tile<Src, ROUNDS, SMEM_FILTER, false>(src, filter, s_filter, sector_mask, count,
		                                                                  warp_base, (tile + 1) * TILE_KEYS <= count, lane,
		                                                                  pol_keys, pol_filter, 0u, 0u);
		// 4. survivors per warp → per tile → global exclusive offset
		uint32_t warp_count = __reduce_add_sync(0xffffffffu, __popc(hits));
		if (lane == 0) {
			s_warp_off[warp] = warp_count;
		}
		__syncthreads();
		if (warp == 0) {
			const uint32_t c = lane < NWARPS ? s_warp_off[lane] : 0u;
			uint32_t incl = c;
#pragma unroll
			for (int d = 1; d < 32; d <<= 1) {
				const uint32_t t = __shfl_up_sync(0xffffffffu, incl, d);
				if (lane >= static_cast<uint32_t>(d)) {
					incl += t;
				}
			}
			const uint32_t tile_count = __shfl_sync(0xffffffffu, incl, 31);
			if (lane < NWARPS) {
				s_warp_off[lane] = incl - c;
			}
			const uint32_t excl = tile_lookback(desc, tile, lane, tile_count);
			if (lane == 0) {
				s_tile_excl = excl;
				if (tile == num_tiles - 1) {
					*found_out = static_cast<uint64_t>(excl) + tile_count;
				}
			}
		}
		__syncthreads();
		// 5. ordered write: row order inside a warp segment is (round, lane, element)
		uint32_t run = s_tile_excl + s_warp_off[warp];
		const uint32_t lt = lanemask_lt();
		const uint32_t rel_base = static_cast<uint32_t>(warp_base) + row_offset + lane * 2;
		if (__any_sync(0xffffffffu, hits != 0)) {
#pragma unroll
			for (int r = 0; r < ROUNDS; r++) {
				const uint32_t h0 = (hits >> (2 * r)) & 1u;
				const uint32_t h1 = (hits >> (2 * r + 1)) & 1u;
				const uint32_t b0 = __ballot_sync(0xffffffffu, h0);
				const uint32_t b1 = __ballot_sync(0xffffffffu, h1);
				uint32_t pos = run + __popc(b0 & lt) + __popc(b1 & lt);
				const uint32_t rel = rel_base + r * 64;
				if (h0) {
					sel_out[pos] = translate ? __ldg(translate + rel) : rel;
				}
				pos += h0;
				if (h1) {
					sel_out[pos] = translate ? __ldg(translate + rel + 1) : rel + 1;
				}
				run += __popc(b0) + __popc(b1);
			}
		}
	}
}

// ---------------------------------------------------------------------------------------------
// merge / misc kernels
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
    or_parts_kernel(ulonglong2 *__restrict__ words, const ulonglong2 *__restrict__ parts, const uint64_t nvec,
                    const uint64_t nparts) {
	const uint64_t step = static_cast<uint64_t>(gridDim.x) * blockDim.x;
	for (uint64_t i = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < nvec; i += step) {
		ulonglong2 acc = words[i];
		for (uint64_t p = 0; p < nparts; p++) {
			const ulonglong2 v = __ldg(parts + p * nvec + i);
			acc.x |= v.x;
			acc.y |= v.y;
		}
		words[i] = acc;
	}
}

// Peer-memory leg: srcs.p[k] are other GPUs' filters mapped through CUDA IPC / peer access; the
// loads cross NVLink (they bypass the local L2), the OR happens in registers, one local store.
__global__ void __launch_bounds__(256)
    or_ptrs_kernel(uint64_t *__restrict__ words, const PtrList srcs, const uint64_t off, const uint64_t len) {
	const uint64_t nvec = len >> 1;
	ulonglong2 *dst = reinterpret_cast<ulonglong2 *>(words + off);
	const uint64_t step = static_cast<uint64_t>(gridDim.x) * blockDim.x;
	for (uint64_t i = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < nvec; i += step) {
		ulonglong2 acc = dst[i];
		ulonglong2 v[16];
#pragma unroll
		for (int k = 0; k < 16; k++) {
			if (k < srcs.n) {
				v[k] = *(reinterpret_cast<const ulonglong2 *>(srcs.p[k] + off) + i);
			}
		}
#pragma unroll
		for (int k = 0; k < 16; k++) {
			if (k < srcs.n) {
				acc.x |= v[k].x;
				acc.y |= v[k].y;
			}
		}
		dst[i] = acc;
	}
}

__global__ void __launch_bounds__(256)
    lookup_hashes_kernel(const uint64_t *__restrict__ words, const uint64_t sector_mask,
                         const uint64_t *__restrict__ hashes, const uint64_t count, uint8_t *__restrict__ hits) {
	const uint64_t i = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
	if (i < count) {
		const uint64_t h = hashes[i];
		hits[i] = static_cast<uint8_t>(sector_test(words[h & sector_mask], static_cast<uint32_t>(h >> 32)));
	}
}

__host__ __device__ __forceinline__ uint64_t splitmix64(uint64_t x) {
	uint64_t z = x + 0x9e3779b97f4a7c15ULL;
	z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ULL;
	z = (z ^ (z >> 27)) * 0x94d049bb133111ebULL;
	return z ^ (z >> 31);
}

__global__ void __launch_bounds__(256) synth_build_kernel(uint64_t *out, const uint64_t first, const uint64_t count) {
	const uint64_t step = static_cast<uint64_t>(gridDim.x) * blockDim.x;
	for (uint64_t i = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < count; i += step) {
		out[i] = splitmix64(0xB10F0001ULL + first + i) & 0x7fffffffffffffffULL;
	}
}

__global__ void __launch_bounds__(256) synth_probe_kernel(uint64_t *out, const uint64_t first, const uint64_t count,
                                                          const uint64_t n_build, const uint32_t sel_ppm) {
	const uint64_t step = static_cast<uint64_t>(gridDim.x) * blockDim.x;
	for (uint64_t i = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < count; i += step) {
		const uint64_t j = first + i;
		const uint64_t a = splitmix64(0x9B0BE001ULL + 2 * j);
		const uint64_t b = splitmix64(0x9B0BE001ULL + 2 * j + 1);
		out[i] = (a % 1000000ULL < sel_ppm) ? (splitmix64(0xB10F0001ULL + (b % n_build)) & 0x7fffffffffffffffULL)
		                                    : (b | 0x8000000000000000ULL);
	}
}

// ---------------------------------------------------------------------------------------------
// launchers
// ---------------------------------------------------------------------------------------------
#ifndef RPT_PROBE_ROUNDS
#define RPT_PROBE_ROUNDS 8
#endif
constexpr int kProbeRounds = RPT_PROBE_ROUNDS; // 2·ROUNDS keys per thread
constexpr int kProbeBlock = 256;   // global-filter variant: several CTAs per SM
constexpr int kProbeBlockSmemPhased = 384; // phased shared-memory variant: 12 warps' key buffers + ≤128 KiB filter
constexpr int kProbeBlockSmem = 512; // shared-memory variant: ≤128 KiB filter + 1 CTA per SM at the top size
constexpr uint64_t kProbeTileKeys = 64ull * kProbeRounds * (kProbeBlock / 32);
constexpr uint64_t kFusedMaxKeys = 1ull << 18; // up to 64 tiles: launch-latency territory, one fused kernel

static inline uint64_t ceil_div(uint64_t a, uint64_t b) {
	return (a + b - 1) / b;
}

// A single flat 8-byte integer column, 16-byte aligned, takes the vectorised source.
static inline bool is_flat_u64(const DevCols &cols) {
	if (cols.n != 1) {
		return false;
	}
	const DevCol &c = cols.c[0];
	return (c.type == RPT_INT64 || c.type == RPT_UINT64) && !c.sel &&
	       (reinterpret_cast<uintptr_t>(c.data) & 15u) == 0 && (reinterpret_cast<uintptr_t>(c.validity) & 7u) == 0;
}
// A flat column of 64-bit hashes (a string column hashed ahead, rpt_capi.cu plan_prehash) takes the same path
static inline bool is_flat_prehashed(const DevCols &cols) {
	const DevCol &c = cols.c[0];
	return cols.n == 1 && c.type == kTypePrehashed && !c.sel && !c.validity &&
	       (reinterpret_cast<uintptr_t>(c.data) & 15u) == 0;
}
// ... and a single flat 4-byte integer column (INT32 / UINT32; DATE is INT32 physically) its own.
static inline bool is_flat_u32(const DevCols &cols) {
	if (cols.n != 1) {
		return false;
	}
	const DevCol &c = cols.c[0];
	return (c.type == RPT_INT32 || c.type == RPT_UINT32) && !c.sel &&
	       (reinterpret_cast<uintptr_t>(c.data) & 15u) == 0 && (reinterpret_cast<uintptr_t>(c.validity) & 7u) == 0;
}

static inline bool use_smem_filter(const DeviceInfo &dev, uint64_t num_sectors) {
	// the phased kernel also keeps 8 KiB of key double buffers per warp in shared memory
	return num_sectors <= kSmemFilterMaxSectors &&
	       num_sectors * 8 + (kProbeBlockSmemPhased / 32) * 8192 + 1024 <= dev.smem_optin;
}

// Sector-range slicing (see the build section): filters above the threshold are processed in passes of
// `slice` sectors.  RPT_SLICE_MIB overrides the slice size (MiB; 0 disables slicing) for experiments.
static uint64_t slice_sectors_for(uint64_t num_sectors) {
	static const long env_mib = [] {
		const char *e = getenv("RPT_SLICE_MIB");
		return e ? atol(e) : -1L;
	}();
	uint64_t slice = kSliceSectors;
	if (env_mib == 0) {
		return num_sectors;
	}
	if (env_mib > 0) {
		slice = (static_cast<uint64_t>(env_mib) << 20) / 8;
	}
	return num_sectors > slice ? slice : num_sectors;
}

static inline uint64_t align_up(uint64_t v, uint64_t a) {
	return (v + a - 1) / a * a;
}

static inline unsigned stream_grid(const DeviceInfo &dev, uint64_t items, unsigned block) {
	const uint64_t want = ceil_div(items, block);
	const uint64_t cap = static_cast<uint64_t>(dev.sm_count > 0 ? dev.sm_count : 148) * 8;
	return static_cast<unsigned>(want < cap ? (want ? want : 1) : cap);
}

// Scratch layout of a phased probe: [segment offsets u32 × cap][block totals u32 × blocks][row bitmap, 64 B × cap]
struct PhasedLayout {
	uint64_t seg_cap;    // segments (512 rows each), with slack so that the last compaction chunk stays in bounds
	uint64_t block_cap;  // scan blocks of kScanBlock segments
	size_t off_totals, off_hitbits, bytes;
};
static PhasedLayout phased_layout(uint64_t count) {
	PhasedLayout l;
	l.seg_cap = ceil_div(count, kSegRows) + 32;
	l.block_cap = ceil_div(l.seg_cap, kScanBlock);
	l.off_totals = align_up(l.seg_cap * sizeof(uint32_t), 256);
	l.off_hitbits = l.off_totals + align_up(l.block_cap * sizeof(uint32_t), 256);
	l.bytes = l.off_hitbits + l.seg_cap * 32 * sizeof(uint16_t);
	return l;
}

// Which filters take the partitioned probe.  Measured on B200 (profiles/sweep_r1f.txt, 10^9 probe rows):
//   F = 512 MiB (8 slices): partitioned 10.4 ms vs 15.9 ms for 8 sector-range passes  → partitioned
//   F = 256 MiB (4 slices): 10.3 ms vs 7.9–8.7 ms;  F = 128 MiB (2 slices): 8.0 ms vs 4.9 ms → passes
// (the partition kernel is issue/latency-bound at 88 instructions per key, ncu r1f; re-reading 8-byte keys two or
// four times is cheaper than that).  RPT_PROBE_PARTITION=1 forces it for every sliced filter, =0 disables it.
static int partition_mode() {
	static const int mode = [] {
		const char *e = getenv("RPT_PROBE_PARTITION");
		if (getenv("RPT_SLICE_MIB") != nullptr) {
			return 0;
		}
		return e ? (atoi(e) != 0 ? 2 : 0) : 1;
	}();
	return mode;
}
// Slices the scratch buffer is sized for (the choice between passes and partition depends on the key source, which
// the sizing call does not know): every sliced filter unless the partitioned probe is switched off.
static inline int partition_slices(uint64_t num_sectors) {
	const int slices = num_sectors > kSliceSectors ? static_cast<int>(num_sectors / kSliceSectors) : 0;
	return partition_mode() == 0 ? 0 : slices;
}
// Passes or partition, by measurement (profiles/sweep_r1h.txt, profiles/sweep_r1o_varchar.txt).
// Returns the number of slices to partition into, 0 for sector-range passes.
template <class Src>
static inline int partition_choice(uint64_t num_sectors) {
	const int slices = partition_slices(num_sectors);
	if (slices < 2 || partition_mode() == 2) {
		return slices < 2 ? 0 : slices;
	}
	if (!Src::kStageable) {
		return slices; // generic sources (strings, selections, several columns) hash each row once, not once per pass
	}
	if (slices == 2) {
		return 0; // two passes over a flat column: 4.89 vs 6.08 ms per 1e9 int64 rows
	}
	if (slices == 4 && Src::kKeyBytes == 4) {
		return 0; // four passes over a 4-byte column re-read 16 B per key: 6.92 vs 7.25 ms
	}
	return slices;
}
// Filters the folded pre-filter exists for: 4 or 8 slices of 64 MiB (256 / 512 MiB)
bool prefilter_applicable(uint64_t num_sectors) {
	return num_sectors == 4 * kSliceSectors || num_sectors == 8 * kSliceSectors;
}
static inline int partition_cap(int slices) {
	return slices == 2 ? PartCfg<2>::CAP : (slices == 4 ? PartCfg<4>::CAP : PartCfg<8>::CAP);
}
// Variants of the partitioned probe (experiments): RPT_PART_SLICE_MIB=32 partitions a filter of up
// to 256 MiB into 32-MiB slices (twice as many, each certain to stay L2-resident); RPT_PART_INLINE0=0 makes slice 0 a
// bucket like the others (no filter access, no L2 competition in the partition kernel).
struct PartPlan {
	int slices; // P
	int shift;  // log2(sectors per slice)
	bool inline0;
};
static PartPlan partition_plan(uint64_t num_sectors, int slices64) {
	static const int env_mib = [] {
		const char *e = getenv("RPT_PART_SLICE_MIB");
		return e ? atoi(e) : 64;
	}();
	static const bool env_inline0 = [] {
		const char *e = getenv("RPT_PART_INLINE0");
		return e ? atoi(e) != 0 : true;
	}();
	PartPlan plan {slices64, static_cast<int>(kSliceShift), env_inline0};
	if (env_mib == 32 && slices64 >= 2 && slices64 <= 4 && (num_sectors >> (kSliceShift - 1)) <= 8) {
		plan.slices = static_cast<int>(num_sectors >> (kSliceShift - 1));
		plan.shift = static_cast<int>(kSliceShift) - 1;
	}
	return plan;
}
// extra scratch behind the phased layout (R = bucket regions per segment, P or P-1):
// [bucket counts u16 × R·segs][overflow list u32 × (1 + segs)][records 8 B × R·segs·CAP]
static size_t partition_scratch_bytes(uint64_t count, uint64_t num_sectors, int slices64) {
	const uint64_t num_segs = ceil_div(count, kSegRows);
	const PartPlan plan = partition_plan(num_sectors, slices64);
	const uint64_t regions = plan.slices - (plan.inline0 ? 1 : 0);
	return align_up(regions * num_segs * sizeof(uint16_t), 256) + align_up((num_segs + 1) * sizeof(uint32_t), 256) +
	       regions * num_segs * partition_cap(plan.slices) * sizeof(uint2);
}
// extra scratch of a pre-filtered probe behind the phased layout: [ends u16 × P × halves][records 8 B × 256 × halves]
static size_t folded_scratch_bytes(uint64_t count, uint64_t num_sectors) {
	const uint64_t num_halves = 2 * ceil_div(count, kSegRows);
	const uint64_t slices = num_sectors / kSliceSectors;
	return align_up(num_halves * slices * sizeof(uint16_t), 256) + num_halves * kHalfRows * sizeof(uint2);
}

size_t probe_scratch_bytes(uint64_t count, uint64_t num_sectors, bool prefilter) {
	if (count <= kFusedMaxKeys) {
		return (ceil_div(count, kProbeTileKeys) + 2) * sizeof(uint64_t); // look-back descriptors + ticket
	}
	const int slices = partition_slices(num_sectors);
	size_t extra = slices ? partition_scratch_bytes(count, num_sectors, slices) : 0;
	if (prefilter_applicable(num_sectors)) {
		// One size for the pre-filtered and the plain probe of a filter, whichever a call takes: a handle whose probes
		// alternate between them (feedback, rpt_bf_set_prefilter) then asks the stream-ordered pool for the same block
		// every time.  With two sizes (8.03 / 8.33 GB per 10^9 rows) a long run of one followed by the other left calls
		// 20 ms slower for dozens of calls (bench.py after its sustained spin: profiles/bench_r2y.json).
		const size_t folded = folded_scratch_bytes(count, num_sectors);
		extra = extra > folded ? extra : folded;
	}
	(void)prefilter;
	return align_up(phased_layout(count).bytes, 256) + extra;
}

// scratch of a sliced build: one 64-bit hash per row, written by the first sector-range pass, read by the others
// Generic sources only (several key columns, selections, dictionaries: hashing a row is the expensive part — two key
// columns, 10^8 rows, F = 256 MiB: 1.64 against 3.15 ms).  For a flat column the extra 8 bytes per row written and
// read back cost more than the murmur they save (int64: 1.32 against 0.97 ms; profiles/build_hash_once_r2.txt).
static thread_local uint64_t *t_hash_scratch = nullptr;
size_t build_scratch_bytes(const DevCols &cols, uint64_t count, uint64_t num_sectors) {
	static const bool on = [] {
		const char *e = getenv("RPT_BUILD_HASH_ONCE");
		return !e || atoi(e) != 0;
	}();
	if (!on || slice_sectors_for(num_sectors) >= num_sectors || count < (1ULL << 20)) {
		return 0;
	}
	if (is_flat_prehashed(cols) || is_flat_u64(cols) || is_flat_u32(cols)) {
		return 0;
	}
	if (cols.n == 1 && cols.c[0].type == RPT_VARCHAR && !cols.c[0].sel) {
		return 0; // a flat string column is hashed ahead (hash_varchar_kernel) and then is a flat column of hashes
	}
	return count * sizeof(uint64_t);
}

template <class Src, bool AGG, bool MM>
static cudaError_t launch_build_mm(const DeviceInfo &dev, const Src &src, uint64_t count, uint64_t *words,
                                   uint64_t num_sectors, MinMaxState *mm, bool mm_signed, cudaStream_t stream,
                                   int *launches) {
	const uint64_t per_block = static_cast<uint64_t>(kBuildBlock) * 2 * kBuildUnroll;
	if constexpr (!AGG && !Src::kStageable) {
		if (t_hash_scratch != nullptr && slice_sectors_for(num_sectors) < num_sectors) {
			// hash once: pass 0 reduces its slice and writes the hashes, the other passes run on the hash column
			using HashSrc = FlatU64SrcT<false, true>;
			int p0 = 0, p1 = 0;
			cudaError_t e0 = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&p0, build_kernel<Src, false, MM, true>, kBuildBlock, 0);
			if (e0 == cudaSuccess) {
				e0 = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&p1, build_kernel<HashSrc, false, false>, kBuildBlock, 0);
			}
			if (e0 != cudaSuccess) {
				return e0;
			}
			const uint64_t want = ceil_div(count, per_block);
			const uint64_t g0 = std::min<uint64_t>(want, static_cast<uint64_t>(dev.sm_count) * (p0 > 0 ? p0 : 1) * 4);
			const uint64_t g1 = std::min<uint64_t>(want, static_cast<uint64_t>(dev.sm_count) * (p1 > 0 ? p1 : 1) * 4);
			const uint64_t slice = slice_sectors_for(num_sectors);
			const HashSrc hsrc {t_hash_scratch, nullptr};
			for (uint64_t lo = 0; lo < num_sectors; lo += slice) {
				const uint64_t len = lo + slice < num_sectors ? slice : num_sectors - lo;
				if (lo == 0) {
					build_kernel<Src, false, MM, true><<<static_cast<unsigned>(g0), kBuildBlock, 0, stream>>>(
					    src, reinterpret_cast<unsigned long long *>(words), static_cast<uint32_t>(num_sectors - 1), count, 0u,
					    static_cast<uint32_t>(len), mm, mm_signed, t_hash_scratch);
				} else {
					build_kernel<HashSrc, false, false><<<static_cast<unsigned>(g1), kBuildBlock, 0, stream>>>(
					    hsrc, reinterpret_cast<unsigned long long *>(words), static_cast<uint32_t>(num_sectors - 1), count,
					    static_cast<uint32_t>(lo), static_cast<uint32_t>(len), nullptr, false);
				}
				if (launches) {
					(*launches)++;
				}
			}
			return cudaGetLastError();
		}
	}
	int per_sm = 0;
	cudaError_t e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, build_kernel<Src, AGG, MM>, kBuildBlock, 0);
	if (e != cudaSuccess) {
		return e;
	}
	const uint64_t wave = static_cast<uint64_t>(dev.sm_count) * (per_sm > 0 ? per_sm : 1);
	uint64_t grid = ceil_div(count, per_block);
	if (grid > wave * 4) {
		grid = wave * 4; // grid-stride over the rest; a multiple of SMs × resident CTAs
	}
	const uint64_t slice = slice_sectors_for(num_sectors);
	for (uint64_t lo = 0; lo < num_sectors; lo += slice) {
		const uint64_t len = lo + slice < num_sectors ? slice : num_sectors - lo;
		if (MM && lo != 0) { // the column range is folded in by the first sector-range pass only
			build_kernel<Src, AGG, false><<<static_cast<unsigned>(grid), kBuildBlock, 0, stream>>>(
			    src, reinterpret_cast<unsigned long long *>(words), static_cast<uint32_t>(num_sectors - 1), count,
			    static_cast<uint32_t>(lo), static_cast<uint32_t>(len), nullptr, false);
		} else {
			build_kernel<Src, AGG, MM><<<static_cast<unsigned>(grid), kBuildBlock, 0, stream>>>(
			    src, reinterpret_cast<unsigned long long *>(words), static_cast<uint32_t>(num_sectors - 1), count,
			    static_cast<uint32_t>(lo), static_cast<uint32_t>(len), mm, mm_signed);
		}
		if (launches) {
			(*launches)++;
		}
	}
	return cudaGetLastError();
}
template <class Src, bool AGG>
static cudaError_t launch_build_t(const DeviceInfo &dev, const Src &src, uint64_t count, uint64_t *words,
                                  uint64_t num_sectors, MinMaxState *mm, bool mm_signed, cudaStream_t stream,
                                  int *launches) {
	return mm ? launch_build_mm<Src, AGG, true>(dev, src, count, words, num_sectors, mm, mm_signed, stream, launches)
	          : launch_build_mm<Src, AGG, false>(dev, src, count, words, num_sectors, nullptr, false, stream, launches);
}

// mm (optional): the build also folds the (single) key column's valid rows into *mm — UpdateMinMax fused into the
// insert pass (physical_create_bf.cpp:220-239).  Ignored for several key columns and for pre-hashed columns.
static cudaError_t launch_build_inner(const DeviceInfo &dev, const DevCols &cols, uint64_t count, uint64_t *words,
                                      uint64_t num_sectors, cudaStream_t stream, int *launches, MinMaxState *mm);
cudaError_t launch_build(const DeviceInfo &dev, const DevCols &cols, uint64_t count, uint64_t *words,
                         uint64_t num_sectors, cudaStream_t stream, int *launches, MinMaxState *mm,
                         uint64_t *hash_scratch) {
	t_hash_scratch = hash_scratch;
	const cudaError_t e = launch_build_inner(dev, cols, count, words, num_sectors, stream, launches, mm);
	t_hash_scratch = nullptr;
	return e;
}
static cudaError_t launch_build_inner(const DeviceInfo &dev, const DevCols &cols, uint64_t count, uint64_t *words,
                                      uint64_t num_sectors, cudaStream_t stream, int *launches, MinMaxState *mm) {
	if (count == 0) {
		return cudaSuccess;
	}
	if (cols.n != 1) {
		mm = nullptr;
	}
	// warp-aggregate only where same-sector collisions inside a warp are likely
	const bool agg = num_sectors <= 4096;
	const bool mm_signed = cols.c[0].type == RPT_INT64 || cols.c[0].type == RPT_INT32;
#define RPT_BUILD_FLAT(SRC, PTR_T, MMPTR)                                                                               \
	{                                                                                                                  \
		SRC src {reinterpret_cast<const PTR_T *>(cols.c[0].data), cols.c[0].validity};                                 \
		return agg ? launch_build_t<SRC, true>(dev, src, count, words, num_sectors, MMPTR, mm_signed, stream, launches) \
		           : launch_build_t<SRC, false>(dev, src, count, words, num_sectors, MMPTR, mm_signed, stream,         \
		                                        launches);                                                             \
	}
	if (is_flat_prehashed(cols)) {
		using PrehashedSrc = FlatU64SrcT<false, true>;
		RPT_BUILD_FLAT(PrehashedSrc, uint64_t, nullptr)
	}
	if (is_flat_u64(cols)) {
		if (cols.c[0].validity) {
			RPT_BUILD_FLAT(FlatU64SrcT<true>, uint64_t, mm)
		}
		RPT_BUILD_FLAT(FlatU64SrcT<false>, uint64_t, mm)
	}
	if (is_flat_u32(cols)) {
		if (cols.c[0].validity) {
			RPT_BUILD_FLAT(FlatU32SrcT<true>, uint32_t, mm)
		}
		RPT_BUILD_FLAT(FlatU32SrcT<false>, uint32_t, mm)
	}
#undef RPT_BUILD_FLAT
	GenericSrc src {cols};
	return agg ? launch_build_t<GenericSrc, true>(dev, src, count, words, num_sectors, mm, mm_signed, stream, launches)
	           : launch_build_t<GenericSrc, false>(dev, src, count, words, num_sectors, mm, mm_signed, stream, launches);
}

// Slice gates (multi-GPU merge overlapped with the probe, rpt_bf_set_slice_events): events[s] fires when sectors
// [s·kSliceSectors, (s+1)·kSliceSectors) of the filter hold their final (merged) words.  A probe launcher makes the
// stream wait for slice s right before the first kernel that reads it: the partition kernel only looks slice 0 up,
// bucket kernel s only slice s, sector-range pass p only slice p — so the merge of the later slices runs on another
// stream underneath the partition kernel.
static thread_local const SliceGate *t_gate = nullptr;
static cudaError_t wait_slices(cudaStream_t stream, int first, int last /* inclusive; <0: all */) {
	if (!t_gate || t_gate->n == 0) {
		return cudaSuccess;
	}
	if (last < 0 || last >= t_gate->n) {
		last = t_gate->n - 1;
	}
	static const bool gate_all = getenv("RPT_GATE_ALL") != nullptr; // debugging: every kernel waits for every slice
	if (gate_all) {
		first = 0;
		last = t_gate->n - 1;
	}
	for (int s = first; s <= last; s++) {
		if (t_gate->ev[s]) {
			const cudaError_t e = cudaStreamWaitEvent(stream, t_gate->ev[s], 0);
			if (e != cudaSuccess) {
				return e;
			}
		}
	}
	return cudaSuccess;
}

// Small probes: the fused look-back kernel (one launch + one memset).
template <class Src, int BLOCK, bool SMEM>
static cudaError_t launch_probe_fused(const DeviceInfo &dev, const Src &src, uint64_t count, const uint64_t *words,
                                      uint64_t num_sectors, uint32_t *sel_out, uint64_t *found_out,
                                      uint32_t row_offset, const uint32_t *translate, void *scratch,
                                      cudaStream_t stream, int *launches) {
	auto kern = probe_kernel<Src, kProbeRounds, BLOCK, SMEM>;
	const uint64_t tile_keys = 64ull * kProbeRounds * (BLOCK / 32);
	const uint64_t num_tiles = ceil_div(count, tile_keys);
	const size_t smem = SMEM ? num_sectors * sizeof(uint64_t) : 0;
	cudaError_t e;
	if (SMEM) {
		e = ensure_max_dyn_smem<probe_kernel<Src, kProbeRounds, BLOCK, SMEM>>(dev);
		if (e != cudaSuccess) {
			return e;
		}
	}
	int per_sm = 0;
	e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, BLOCK, smem);
	if (e != cudaSuccess) {
		return e;
	}
	if (per_sm < 1) {
		return cudaErrorLaunchOutOfResources;
	}
	uint64_t grid = static_cast<uint64_t>(dev.sm_count) * per_sm; // persistent: every CTA is resident
	if (grid > num_tiles) {
		grid = num_tiles;
	}
	uint64_t *desc = reinterpret_cast<uint64_t *>(scratch);
	unsigned int *ticket = reinterpret_cast<unsigned int *>(desc + num_tiles);
	e = wait_slices(stream, 0, -1);
	if (e != cudaSuccess) {
		return e;
	}
	e = cudaMemsetAsync(scratch, 0, (num_tiles + 1) * sizeof(uint64_t), stream);
	if (e != cudaSuccess) {
		return e;
	}
	kern<<<static_cast<unsigned>(grid), BLOCK, smem, stream>>>(src, words, static_cast<uint32_t>(num_sectors - 1), count,
	                                                          num_tiles, sel_out, found_out, row_offset, translate,
	                                                          desc, ticket);
	if (launches) {
		(*launches)++;
	}
	return cudaGetLastError();
}

// Large probes: bitmap pass kernel per sector-range slice → two-level scan of segment counts → compaction.
// UNITS/MINB: 16 keys in flight per lane at 2 CTAs/SM, or 8 keys in flight at 3 CTAs/SM (RPT_PROBE_UNITS=8|4).
template <class Src, int UNITS, int MINB, bool SMEM, bool SLICED>
static cudaError_t launch_probe_phased(const DeviceInfo &dev, const Src &src, uint64_t count, const uint64_t *words,
                                       uint64_t num_sectors, uint32_t *sel_out, uint64_t *found_out,
                                       uint32_t row_offset, const uint32_t *translate, void *scratch,
                                       cudaStream_t stream, int *launches) {
	const uint64_t num_segs = ceil_div(count, kSegRows);
	// Regimes 2 and 3 (one L2-resident slice; sector-range passes) of a flat column also take the bulk-copy staged
	// kernel, with the filter left in global memory: 4.02 against 4.16 ms per 10^9 rows at 64 MiB, equal at 16 MiB
	// (both a few per cent above the one-gather-per-clock floor plus the compaction); two passes at 128 MiB 4.69
	// against 4.88 ms.  RPT_L2_KERNEL=0: probe_bitmap_kernel (cp.async staging).
	static const bool l2_bulk = [] {
		const char *e = getenv("RPT_L2_KERNEL");
		return !e || atoi(e) != 0;
	}();
	if constexpr (Src::kStageable) {
	if (SMEM || l2_bulk) {
		// regime 1 of a flat column: bulk-copy staged kernel, 24 warps next to the filter (regimes 2 and 3: the filter
		// stays in global memory, one launch per sector-range slice)
		constexpr int NW = kSmemProbeWarps;
		constexpr bool GLOBAL = !SMEM;
		constexpr bool PASSES = SLICED && !SMEM;
		cudaError_t e = ensure_max_dyn_smem<probe_smem_kernel<Src, NW, GLOBAL, PASSES>>(dev);
		if (e != cudaSuccess) {
			return e;
		}
		const size_t smem = (GLOBAL ? 0 : num_sectors * sizeof(uint64_t)) +
		                    static_cast<size_t>(NW) * 2 * kHalfRows * Src::kKeyBytes;
		int per_sm = 0;
		e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, probe_smem_kernel<Src, NW, GLOBAL, PASSES>, NW * 32,
		                                                  smem);
		if (e != cudaSuccess) {
			return e;
		}
		if (per_sm < 1) {
			return cudaErrorLaunchOutOfResources;
		}
		uint64_t grid = static_cast<uint64_t>(dev.sm_count) * per_sm;
		const uint64_t want = ceil_div(num_segs, NW);
		grid = grid < want ? grid : want;
		const PhasedLayout lay = phased_layout(count);
		char *base = reinterpret_cast<char *>(scratch);
		uint32_t *seg_offsets = reinterpret_cast<uint32_t *>(base);
		uint32_t *block_totals = reinterpret_cast<uint32_t *>(base + lay.off_totals);
		uint8_t *bitmap = reinterpret_cast<uint8_t *>(base + lay.off_hitbits);
		const uint64_t slice = PASSES ? slice_sectors_for(num_sectors) : num_sectors;
		int passes = 0;
		for (uint64_t lo = 0; lo < num_sectors; lo += slice, passes++) {
			const uint64_t len = lo + slice < num_sectors ? slice : num_sectors - lo;
			const bool last = lo + slice >= num_sectors;
			e = wait_slices(stream, static_cast<int>(lo / kSliceSectors),
			                PASSES ? static_cast<int>((lo + len - 1) / kSliceSectors) : -1);
			if (e != cudaSuccess) {
				return e;
			}
			probe_smem_kernel<Src, NW, GLOBAL, PASSES><<<static_cast<unsigned>(grid), NW * 32, smem, stream>>>(
			    src, words, static_cast<uint32_t>(num_sectors - 1), count, num_segs, bitmap, last ? seg_offsets : nullptr,
			    static_cast<uint32_t>(lo), static_cast<uint32_t>(len), lo == 0 ? 1 : 0);
		}
		const uint64_t nblocks = ceil_div(num_segs, kScanBlock);
		scan_blocks_kernel<<<static_cast<unsigned>(nblocks), 1024, 0, stream>>>(seg_offsets, num_segs, block_totals);
		scan_totals_kernel<<<1, 1024, 0, stream>>>(block_totals, static_cast<uint32_t>(nblocks), found_out);
		launch_compact(dev, reinterpret_cast<const uint32_t *>(bitmap), seg_offsets, block_totals, num_segs, sel_out, row_offset, translate, stream);
		if (launches) {
			(*launches) += 3 + passes;
		}
		return cudaGetLastError();
	}
	}
	{
	// shared-memory filter of a generic source: one CTA per SM at the top size
	constexpr int BLOCK = SMEM ? kProbeBlockSmemPhased : kProbeBlock;
	auto pass = probe_bitmap_kernel<Src, BLOCK, MINB, UNITS, SMEM, SLICED>;
	const size_t smem = (Src::kStageable ? static_cast<size_t>(BLOCK / 32) * 2 * kSegRows * Src::kKeyBytes : 0) +
	                    (SMEM ? num_sectors * sizeof(uint64_t) : 0);
	cudaError_t e = ensure_max_dyn_smem<probe_bitmap_kernel<Src, BLOCK, MINB, UNITS, SMEM, SLICED>>(dev);
	if (e != cudaSuccess) {
		return e;
	}
	int per_sm = 0;
	e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, pass, BLOCK, smem);
	if (e != cudaSuccess) {
		return e;
	}
	if (per_sm < 1) {
		return cudaErrorLaunchOutOfResources;
	}
	uint64_t grid = static_cast<uint64_t>(dev.sm_count) * per_sm; // persistent warps, a multiple of the SM count
	const uint64_t want = ceil_div(num_segs, BLOCK / 32);
	if (grid > want) {
		grid = want;
	}
	const PhasedLayout lay = phased_layout(count);
	char *base = reinterpret_cast<char *>(scratch);
	uint32_t *seg_offsets = reinterpret_cast<uint32_t *>(base);
	uint32_t *block_totals = reinterpret_cast<uint32_t *>(base + lay.off_totals);
	uint16_t *bitmap = reinterpret_cast<uint16_t *>(base + lay.off_hitbits);
	const uint32_t sector_mask = static_cast<uint32_t>(num_sectors - 1);
	const uint64_t slice = SLICED ? slice_sectors_for(num_sectors) : num_sectors;
	for (uint64_t lo = 0; lo < num_sectors; lo += slice) {
		const uint64_t len = lo + slice < num_sectors ? slice : num_sectors - lo;
		const bool last = lo + slice >= num_sectors;
		// a pass reads the sectors [lo, lo + len): the gates of the 64 MiB slices it touches
		e = wait_slices(stream, static_cast<int>(lo / kSliceSectors),
		                SLICED ? static_cast<int>((lo + len - 1) / kSliceSectors) : -1);
		if (e != cudaSuccess) {
			return e;
		}
		pass<<<static_cast<unsigned>(grid), BLOCK, smem, stream>>>(src, words, sector_mask, count, num_segs,
		                                                          static_cast<uint32_t>(lo), static_cast<uint32_t>(len),
		                                                          bitmap, lo == 0 ? 1 : 0, last ? seg_offsets : nullptr);
		if (launches) {
			(*launches)++;
		}
	}
	const uint64_t nblocks = ceil_div(num_segs, kScanBlock);
	scan_blocks_kernel<<<static_cast<unsigned>(nblocks), 1024, 0, stream>>>(seg_offsets, num_segs, block_totals);
	scan_totals_kernel<<<1, 1024, 0, stream>>>(block_totals, static_cast<uint32_t>(nblocks), found_out);
	launch_compact(dev, reinterpret_cast<const uint32_t *>(bitmap), seg_offsets, block_totals, num_segs, sel_out, row_offset, translate, stream);
	if (launches) {
		(*launches) += 3;
	}
	return cudaGetLastError();
	}
}

static thread_local const Prefilter *t_pre = nullptr;

template <class Src, int P, int SHIFT = kSliceShift, bool INLINE0 = true>
static cudaError_t launch_probe_partitioned(const DeviceInfo &dev, const Src &src, uint64_t count,
                                            const uint64_t *words, uint64_t num_sectors, uint32_t *sel_out,
                                            uint64_t *found_out, uint32_t row_offset, const uint32_t *translate,
                                            void *scratch, cudaStream_t stream, int *launches) {
	constexpr int BLOCK = kProbeBlock;
	using Cfg = PartCfg<P, INLINE0>;
	constexpr int CAP = Cfg::CAP;
	constexpr int FIRST = Cfg::FIRST;
	constexpr int REGIONS = Cfg::REGIONS;
	auto part = probe_partition_kernel<Src, BLOCK, P, SHIFT, INLINE0>;
	auto bucket = probe_bucket_kernel<CAP, SHIFT>;
	const uint64_t num_segs = ceil_div(count, kSegRows);
	const size_t smem = static_cast<size_t>(BLOCK / 32) * PartSmem<Src, P, INLINE0>::WARP_BYTES;
	cudaError_t e = ensure_max_dyn_smem<probe_partition_kernel<Src, BLOCK, P, SHIFT, INLINE0>>(dev);
	if (e != cudaSuccess) {
		return e;
	}
	int per_sm = 0;
	e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, part, BLOCK, smem);
	if (e != cudaSuccess) {
		return e;
	}
	if (per_sm < 1) {
		return cudaErrorLaunchOutOfResources;
	}
	uint64_t grid = static_cast<uint64_t>(dev.sm_count) * per_sm;
	const bool gated = t_gate != nullptr && t_gate->n > 0;
	if (gated) {
		// The persistent CTAs of this kernel fill every SM's register file (2 × 256 threads × 128 registers), so the
		// merge kernels of the later slices, which become runnable while it runs, would find no SM to start on and
		// the "overlap" would serialise behind it.  A slightly smaller grid leaves a few SMs half empty for them.
		static const int pct = [] {
			const char *e = getenv("RPT_GATED_GRID_PCT");
			const int v = e ? atoi(e) : 94;
			return v < 10 ? 10 : (v > 100 ? 100 : v);
		}();
		grid = grid * pct / 100;
	}
	const uint64_t want = ceil_div(num_segs, BLOCK / 32);
	grid = grid < want ? grid : want;
	const PhasedLayout lay = phased_layout(count);
	char *base = reinterpret_cast<char *>(scratch);
	uint32_t *seg_offsets = reinterpret_cast<uint32_t *>(base);
	uint32_t *block_totals = reinterpret_cast<uint32_t *>(base + lay.off_totals);
	uint16_t *bitmap = reinterpret_cast<uint16_t *>(base + lay.off_hitbits);
	char *extra = base + align_up(lay.bytes, 256);
	uint16_t *bucket_counts = reinterpret_cast<uint16_t *>(extra);
	extra += align_up(REGIONS * num_segs * sizeof(uint16_t), 256);
	uint32_t *overflow = reinterpret_cast<uint32_t *>(extra);
	uint2 *buckets = reinterpret_cast<uint2 *>(extra + align_up((num_segs + 1) * sizeof(uint32_t), 256));
	const uint64_t bucket_stride = num_segs * CAP;
	if (gated) {
		e = cudaMemsetAsync(overflow, 0, sizeof(uint32_t), stream);
		if (e != cudaSuccess) {
			return e;
		}
	}
	// gates are per 64-MiB slice (kSliceShift); slice s of this plan lies in gate slice (s << SHIFT) >> kSliceShift
	auto gate_of = [](int s) { return static_cast<int>((static_cast<uint64_t>(s) << SHIFT) >> kSliceShift); };
	if (INLINE0) {
		e = wait_slices(stream, 0, 0); // the partition kernel looks up slice-0 keys only
		if (e != cudaSuccess) {
			return e;
		}
	}
	part<<<static_cast<unsigned>(grid), BLOCK, smem, stream>>>(src, words, static_cast<uint32_t>(num_sectors - 1), count,
	                                                          num_segs, bitmap, buckets, bucket_stride, bucket_counts,
	                                                          gated ? overflow : nullptr);
	int bper_sm = 0;
	e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bper_sm, bucket, 256, 0);
	if (e != cudaSuccess) {
		return e;
	}
	uint64_t bgrid = static_cast<uint64_t>(dev.sm_count) * (bper_sm > 0 ? bper_sm : 1);
	const uint64_t bwant = ceil_div(num_segs, 8);
	bgrid = bgrid < bwant ? bgrid : bwant;
	for (int s = FIRST; s < P; s++) {
		e = wait_slices(stream, gate_of(s), gate_of(s));
		if (e != cudaSuccess) {
			return e;
		}
		bucket<<<static_cast<unsigned>(bgrid), 256, 0, stream>>>(buckets + (s - FIRST) * bucket_stride,
		                                                        bucket_counts + (s - FIRST) * num_segs, num_segs, words,
		                                                        static_cast<uint32_t>(s) << SHIFT,
		                                                        reinterpret_cast<uint32_t *>(bitmap));
	}
	if (gated) { // overflowed segments (normally a handful) once more, behind every slice gate by now
		probe_overflow_kernel<Src><<<static_cast<unsigned>(dev.sm_count), 256, 0, stream>>>(
		    src, words, static_cast<uint32_t>(num_sectors - 1), count, overflow, bitmap);
	}
	count_segments_kernel<<<stream_grid(dev, num_segs, 256), 256, 0, stream>>>(
	    reinterpret_cast<const uint4 *>(bitmap), num_segs, seg_offsets);
	const uint64_t nblocks = ceil_div(num_segs, kScanBlock);
	scan_blocks_kernel<<<static_cast<unsigned>(nblocks), 1024, 0, stream>>>(seg_offsets, num_segs, block_totals);
	scan_totals_kernel<<<1, 1024, 0, stream>>>(block_totals, static_cast<uint32_t>(nblocks), found_out);
	launch_compact(dev, reinterpret_cast<const uint32_t *>(bitmap), seg_offsets, block_totals, num_segs, sel_out, row_offset, translate, stream);
	if (launches) {
		(*launches) += 1 + REGIONS + 4 + (gated ? 1 : 0);
	}
	return cudaGetLastError();
}

// Pre-filtered probe (regime 3c): probe_fold_kernel (flat key columns) or probe_fold_direct_kernel (generic sources),
// then one probe_bucket_half_kernel per slice, then counts / scans / compaction.
// Scratch behind the phased layout: [ends u16 × P × halves][records 8 B × 256 × halves] (folded_scratch_bytes).
constexpr int kFoldWarps = 24; // per CTA, one CTA per SM: measured 16 / 24 / 32 → 7.56 / 7.23 / 7.58 ms per probe (r2r)
template <class Src, int P>
static cudaError_t launch_probe_folded(const DeviceInfo &dev, const Src &src, uint64_t count, const uint64_t *words,
                                       uint64_t num_sectors, uint32_t *sel_out, uint64_t *found_out, uint32_t row_offset,
                                       const uint32_t *translate, void *scratch, cudaStream_t stream, int *launches) {
	const uint64_t num_segs = ceil_div(count, kSegRows);
	const uint64_t num_halves = num_segs * 2;
	const PhasedLayout lay = phased_layout(count);
	char *base = reinterpret_cast<char *>(scratch);
	uint32_t *seg_offsets = reinterpret_cast<uint32_t *>(base);
	uint32_t *block_totals = reinterpret_cast<uint32_t *>(base + lay.off_totals);
	uint8_t *bitmap = reinterpret_cast<uint8_t *>(base + lay.off_hitbits);
	char *extra = base + align_up(lay.bytes, 256);
	uint16_t *ends = reinterpret_cast<uint16_t *>(extra);
	uint2 *records = reinterpret_cast<uint2 *>(extra + align_up(num_halves * P * sizeof(uint16_t), 256));
	const uint32_t fold_mask = static_cast<uint32_t>(t_pre->sectors - 1);
	const uint32_t sector_mask = static_cast<uint32_t>(num_sectors - 1);
	unsigned long long *stats = reinterpret_cast<unsigned long long *>(t_pre->stats);
	cudaError_t e;
	int per_sm = 0;
	if constexpr (Src::kStageable) {
		constexpr int NW = kFoldWarps;
		constexpr uint32_t HALF_BYTES = kHalfRows * Src::kKeyBytes;
		constexpr uint32_t WARP_BYTES = 2 * HALF_BYTES + (HALF_BYTES >= kHalfRows * 8 ? 0 : kHalfRows * 8);
		auto kern = probe_fold_kernel<Src, NW, P>;
		const size_t smem = static_cast<size_t>(NW) * WARP_BYTES;
		e = ensure_max_dyn_smem<probe_fold_kernel<Src, NW, P>>(dev);
		if (e != cudaSuccess) {
			return e;
		}
		e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, NW * 32, smem);
		if (e != cudaSuccess) {
			return e;
		}
		if (per_sm < 1) {
			return cudaErrorLaunchOutOfResources;
		}
		uint64_t grid = static_cast<uint64_t>(dev.sm_count) * per_sm;
		const uint64_t want = ceil_div(num_segs, NW);
		grid = grid < want ? grid : want;
		kern<<<static_cast<unsigned>(grid), NW * 32, smem, stream>>>(src, t_pre->folded, fold_mask, sector_mask, count,
		                                                           num_segs, bitmap, records, ends, stats);
	} else {
		auto kern = probe_fold_direct_kernel<Src, P, 4>;
		e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, 256, 0);
		if (e != cudaSuccess) {
			return e;
		}
		if (per_sm < 1) {
			return cudaErrorLaunchOutOfResources;
		}
		uint64_t grid = static_cast<uint64_t>(dev.sm_count) * per_sm;
		const uint64_t want = ceil_div(num_halves, 8);
		grid = grid < want ? grid : want;
		kern<<<static_cast<unsigned>(grid), 256, 0, stream>>>(src, t_pre->folded, fold_mask, sector_mask, count, num_halves,
		                                                     reinterpret_cast<uint32_t *>(bitmap), records, ends, stats);
	}
	auto bucket = probe_bucket_half_kernel<P>;
	int bper_sm = 0;
	e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bper_sm, bucket, 256, 0);
	if (e != cudaSuccess) {
		return e;
	}
	uint64_t bgrid = static_cast<uint64_t>(dev.sm_count) * (bper_sm > 0 ? bper_sm : 1);
	const uint64_t bwant = ceil_div(num_halves, 128); // sixteen halves per warp step, eight warps per CTA
	bgrid = bgrid < bwant ? bgrid : bwant;
	for (int s = 0; s < P; s++) {
		bucket<<<static_cast<unsigned>(bgrid), 256, 0, stream>>>(records, ends, num_halves, words, static_cast<uint32_t>(s),
		                                                        reinterpret_cast<uint32_t *>(bitmap));
	}
	count_segments_kernel<<<stream_grid(dev, num_segs, 256), 256, 0, stream>>>(
	    reinterpret_cast<const uint4 *>(bitmap), num_segs, seg_offsets);
	const uint64_t nblocks = ceil_div(num_segs, kScanBlock);
	scan_blocks_kernel<<<static_cast<unsigned>(nblocks), 1024, 0, stream>>>(seg_offsets, num_segs, block_totals);
	scan_totals_kernel<<<1, 1024, 0, stream>>>(block_totals, static_cast<uint32_t>(nblocks), found_out);
	launch_compact(dev, reinterpret_cast<const uint32_t *>(bitmap), seg_offsets, block_totals, num_segs, sel_out, row_offset,
	               translate, stream);
	if (launches) {
		(*launches) += 1 + P + 4;
	}
	return cudaGetLastError();
}

static int probe_units() {
	static const int units = [] {
		const char *e = getenv("RPT_PROBE_UNITS");
		return e && atoi(e) == 4 ? 4 : 8;
	}();
	return units;
}

template <class Src>
static cudaError_t launch_probe_src(const DeviceInfo &dev, const Src &src, uint64_t count, const uint64_t *words,
                                    uint64_t num_sectors, uint32_t *sel_out, uint64_t *found_out, uint32_t row_offset,
                                    const uint32_t *translate, void *scratch, cudaStream_t stream, int *launches) {
#define RPT_PROBE_ARGS dev, src, count, words, num_sectors, sel_out, found_out, row_offset, translate, scratch, stream, launches
	const bool smem = use_smem_filter(dev, num_sectors);
	if (count <= kFusedMaxKeys) {
		return smem ? launch_probe_fused<Src, kProbeBlockSmem, true>(RPT_PROBE_ARGS)
		            : launch_probe_fused<Src, kProbeBlock, false>(RPT_PROBE_ARGS);
	}
	if (t_pre != nullptr && prefilter_applicable(num_sectors) && t_pre->sectors <= kSliceSectors &&
	    (t_gate == nullptr || t_gate->n == 0)) {
		if (num_sectors == 4 * kSliceSectors) {
			return launch_probe_folded<Src, 4>(RPT_PROBE_ARGS);
		}
		return launch_probe_folded<Src, 8>(RPT_PROBE_ARGS);
	}
	const int slices64 = smem ? 0 : partition_choice<Src>(num_sectors);
	if constexpr (Src::kStageable && Src::kKeyBytes == 8 && !Src::kNullable) { // experiment variants: flat 8-byte keys only
		if (slices64 >= 2) {
			const PartPlan plan = partition_plan(num_sectors, slices64);
			constexpr int kHalf = static_cast<int>(kSliceShift) - 1;
			if (plan.shift == kHalf && plan.slices == 8) {
				return plan.inline0 ? launch_probe_partitioned<Src, 8, kHalf, true>(RPT_PROBE_ARGS)
				                    : launch_probe_partitioned<Src, 8, kHalf, false>(RPT_PROBE_ARGS);
			}
			if (plan.shift == kHalf && plan.slices == 4 && plan.inline0) {
				return launch_probe_partitioned<Src, 4, kHalf, true>(RPT_PROBE_ARGS);
			}
			if (plan.shift == static_cast<int>(kSliceShift) && plan.slices == 4 && !plan.inline0) {
				return launch_probe_partitioned<Src, 4, static_cast<int>(kSliceShift), false>(RPT_PROBE_ARGS);
			}
		}
	}
	switch (slices64) {
	case 2:
		return launch_probe_partitioned<Src, 2>(RPT_PROBE_ARGS);
	case 4:
		return launch_probe_partitioned<Src, 4>(RPT_PROBE_ARGS);
	case 8:
		return launch_probe_partitioned<Src, 8>(RPT_PROBE_ARGS);
	default:
		break;
	}
	const bool wide = probe_units() == 8;
	if (smem) {
		return wide ? launch_probe_phased<Src, 8, 1, true, false>(RPT_PROBE_ARGS)
		            : launch_probe_phased<Src, 4, 1, true, false>(RPT_PROBE_ARGS);
	}
	if (slice_sectors_for(num_sectors) < num_sectors) {
		return wide ? launch_probe_phased<Src, 8, 2, false, true>(RPT_PROBE_ARGS)
		            : launch_probe_phased<Src, 4, 3, false, true>(RPT_PROBE_ARGS);
	}
	return wide ? launch_probe_phased<Src, 8, 2, false, false>(RPT_PROBE_ARGS)
	            : launch_probe_phased<Src, 4, 3, false, false>(RPT_PROBE_ARGS);
#undef RPT_PROBE_ARGS
}

static cudaError_t launch_probe_gated(const DeviceInfo &dev, const DevCols &cols, uint64_t count, const uint64_t *words,
                                      uint64_t num_sectors, uint32_t *sel_out, uint64_t *found_out, uint32_t row_offset,
                                      const uint32_t *translate, void *scratch, cudaStream_t stream, int *launches);

cudaError_t launch_probe(const DeviceInfo &dev, const DevCols &cols, uint64_t count, const uint64_t *words,
                         uint64_t num_sectors, uint32_t *sel_out, uint64_t *found_out, uint32_t row_offset,
                         const uint32_t *translate, void *scratch, cudaStream_t stream, int *launches,
                         const SliceGate *gate, const Prefilter *pre) {
	t_gate = gate;
	t_pre = pre;
	const cudaError_t e = launch_probe_gated(dev, cols, count, words, num_sectors, sel_out, found_out, row_offset,
	                                         translate, scratch, stream, launches);
	t_gate = nullptr;
	t_pre = nullptr;
	return e;
}

static cudaError_t launch_probe_gated(const DeviceInfo &dev, const DevCols &cols, uint64_t count, const uint64_t *words,
                                      uint64_t num_sectors, uint32_t *sel_out, uint64_t *found_out, uint32_t row_offset,
                                      const uint32_t *translate, void *scratch, cudaStream_t stream, int *launches) {
	if (count == 0) {
		return cudaMemsetAsync(found_out, 0, sizeof(uint64_t), stream);
	}
#define RPT_PROBE_FLAT(SRC, PTR_T)                                                                                      \
	{                                                                                                                  \
		SRC src {reinterpret_cast<const PTR_T *>(cols.c[0].data), cols.c[0].validity};                                 \
		return launch_probe_src(dev, src, count, words, num_sectors, sel_out, found_out, row_offset, translate,       \
		                        scratch, stream, launches);                                                            \
	}
	if (is_flat_prehashed(cols)) {
		using PrehashedSrc = FlatU64SrcT<false, true>;
		RPT_PROBE_FLAT(PrehashedSrc, uint64_t)
	}
	if (is_flat_u64(cols)) {
		if (cols.c[0].validity) {
			RPT_PROBE_FLAT(FlatU64SrcT<true>, uint64_t)
		}
		RPT_PROBE_FLAT(FlatU64SrcT<false>, uint64_t)
	}
	if (is_flat_u32(cols)) {
		if (cols.c[0].validity) {
			RPT_PROBE_FLAT(FlatU32SrcT<true>, uint32_t)
		}
		RPT_PROBE_FLAT(FlatU32SrcT<false>, uint32_t)
	}
#undef RPT_PROBE_FLAT
	GenericSrc src {cols};
	return launch_probe_src(dev, src, count, words, num_sectors, sel_out, found_out, row_offset, translate, scratch,
	                        stream, launches);
}

// Row-ordered bitmap → ordered selection vector, for producers outside this file (rpt_acero.cu): the scratch is
// laid out like a phased probe's (segment offsets, block totals, bitmap); the producer fills bitmap_words() for
// whole 512-row segments, this counts the segments, scans and compacts.
size_t bitmap_scratch_bytes(uint64_t count) {
	return align_up(phased_layout(count).bytes, 256);
}
uint32_t *bitmap_words(void *scratch, uint64_t count) {
	return reinterpret_cast<uint32_t *>(reinterpret_cast<char *>(scratch) + phased_layout(count).off_hitbits);
}
cudaError_t launch_bitmap_to_sel(const DeviceInfo &dev, uint64_t count, uint32_t *sel_out, uint64_t *found_out,
                                 uint32_t row_offset, void *scratch, cudaStream_t stream, int *launches) {
	const PhasedLayout lay = phased_layout(count);
	char *base = reinterpret_cast<char *>(scratch);
	uint32_t *seg_offsets = reinterpret_cast<uint32_t *>(base);
	uint32_t *block_totals = reinterpret_cast<uint32_t *>(base + lay.off_totals);
	const uint32_t *bitmap = reinterpret_cast<const uint32_t *>(base + lay.off_hitbits);
	const uint64_t num_segs = ceil_div(count, kSegRows);
	count_segments_kernel<<<stream_grid(dev, num_segs, 256), 256, 0, stream>>>(
	    reinterpret_cast<const uint4 *>(bitmap), num_segs, seg_offsets);
	const uint64_t nblocks = ceil_div(num_segs, kScanBlock);
	scan_blocks_kernel<<<static_cast<unsigned>(nblocks), 1024, 0, stream>>>(seg_offsets, num_segs, block_totals);
	scan_totals_kernel<<<1, 1024, 0, stream>>>(block_totals, static_cast<uint32_t>(nblocks), found_out);
	launch_compact(dev, bitmap, seg_offsets, block_totals, num_segs, sel_out, row_offset, nullptr, stream);
	if (launches) {
		(*launches) += 4;
	}
	return cudaGetLastError();
}

cudaError_t launch_fold(const DeviceInfo &dev, const uint64_t *words, uint64_t num_sectors, uint64_t *folded,
                        uint64_t fold_sectors, uint64_t *popcount_out, cudaStream_t stream) {
	cudaError_t e = cudaMemsetAsync(popcount_out, 0, sizeof(uint64_t), stream);
	if (e != cudaSuccess) {
		return e;
	}
	const uint64_t vecs = fold_sectors >> 1;
	fold_kernel<<<stream_grid(dev, vecs, 256), 256, 0, stream>>>(
	    reinterpret_cast<const ulonglong2 *>(words), vecs, static_cast<uint32_t>(num_sectors / fold_sectors),
	    reinterpret_cast<ulonglong2 *>(folded), reinterpret_cast<unsigned long long *>(popcount_out));
	return cudaGetLastError();
}

cudaError_t launch_or_parts(const DeviceInfo &dev, uint64_t *words, const uint64_t *parts, uint64_t num_sectors,
                            uint64_t nparts, cudaStream_t stream) {
	const uint64_t nvec = num_sectors >> 1;
	or_parts_kernel<<<stream_grid(dev, nvec, 256), 256, 0, stream>>>(
	    reinterpret_cast<ulonglong2 *>(words), reinterpret_cast<const ulonglong2 *>(parts), nvec, nparts);
	return cudaGetLastError();
}

// Cross-process barrier words of the library-owned merge (rpt_bf_peer_allreduce): one thread per peer stores `value`
// into that peer's flag word with system-scope release semantics (everything this GPU wrote before — earlier kernels
// on the stream — is visible to whoever sees the flag), over the peer's IPC mapping.
__global__ void signal_flags_kernel(const FlagList dst, const uint32_t value) {
	if (threadIdx.x < static_cast<unsigned>(dst.n)) {
		__threadfence_system();
		asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(dst.p[threadIdx.x]), "r"(value) : "memory");
	}
}
// Fallback of cuStreamWaitValue32 (driver entry point not available): one thread per flag polls its LOCAL word until
// it is >= value.  Peers signal before they wait, so the poll cannot deadlock; it occupies one small CTA.
__global__ void wait_flags_kernel(const uint32_t *flags, const int n, const int skip, const uint32_t value) {
	const int t = threadIdx.x;
	if (t < n && t != skip) {
		uint32_t v;
		do {
			asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(flags + t) : "memory");
		} while (static_cast<int32_t>(v - value) < 0);
	}
}
cudaError_t launch_signal_flags(const FlagList &dst, uint32_t value, cudaStream_t stream) {
	if (dst.n == 0) {
		return cudaSuccess;
	}
	signal_flags_kernel<<<1, 32, 0, stream>>>(dst, value);
	return cudaGetLastError();
}
cudaError_t launch_wait_flags(const uint32_t *flags, int n, int skip, uint32_t value, cudaStream_t stream) {
	wait_flags_kernel<<<1, 32, 0, stream>>>(flags, n, skip, value);
	return cudaGetLastError();
}

cudaError_t launch_or_ptrs(const DeviceInfo &dev, uint64_t *words, const PtrList &srcs, uint64_t off, uint64_t len,
                           cudaStream_t stream) {
	if (len == 0 || srcs.n == 0) {
		return cudaSuccess;
	}
	or_ptrs_kernel<<<stream_grid(dev, len >> 1, 256), 256, 0, stream>>>(words, srcs, off, len);
	return cudaGetLastError();
}

cudaError_t launch_lookup_hashes(const uint64_t *words, uint64_t num_sectors, const uint64_t *hashes, uint64_t count,
                                 uint8_t *hits, cudaStream_t stream) {
	if (count == 0) {
		return cudaSuccess;
	}
	lookup_hashes_kernel<<<static_cast<unsigned>(ceil_div(count, 256)), 256, 0, stream>>>(words, num_sectors - 1,
	                                                                                      hashes, count, hits);
	return cudaGetLastError();
}

cudaError_t launch_synth_build(uint64_t *out, uint64_t first, uint64_t count, cudaStream_t stream) {
	if (count == 0) {
		return cudaSuccess;
	}
	const uint64_t want = ceil_div(count, 256);
	synth_build_kernel<<<static_cast<unsigned>(want < 148 * 16 ? want : 148 * 16), 256, 0, stream>>>(out, first, count);
	return cudaGetLastError();
}

cudaError_t launch_synth_probe(uint64_t *out, uint64_t first, uint64_t count, uint64_t n_build, uint32_t sel_ppm,
                               cudaStream_t stream) {
	if (count == 0) {
		return cudaSuccess;
	}
	const uint64_t want = ceil_div(count, 256);
	synth_probe_kernel<<<static_cast<unsigned>(want < 148 * 16 ? want : 148 * 16), 256, 0, stream>>>(
	    out, first, count, n_build, sel_ppm);
	return cudaGetLastError();
}

} // namespace rpt
