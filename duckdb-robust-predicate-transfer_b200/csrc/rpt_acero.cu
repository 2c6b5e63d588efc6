// rpt_acero.cu — the README-described filter variant (reference README.md:23-32): an Acero-style blocked bloom
// filter with a 1024-entry table of 57-bit masks (4–5 bits set), 8 bits per key and Fold (OR the halves together
// until more than a quarter of the bits are set).
//
// NOT ON THE PARITY PATH.  No source, test or vector of this filter exists in the reference at this commit
// (SURVEY §0): src/bloom_filter.cpp wraps DuckDB's native BloomFilter, which rpt_kernels.cu reproduces bit for
// bit.  This file follows the published Apache Arrow algorithm the README names (arrow/acero/bloom_filter.{h,cc},
// Arrow 21: BloomFilterMasks, BlockedBloomFilter::CreateEmpty / mask / block_id / Insert / Find / Fold /
// SingleFold), fed with the same DuckDB key hashes as the reference path.  Parity is "unpinned": it is checked
// against oracle/acero_oracle.cpp, an independent CPU restatement of the same published algorithm.
//
// Kernels (sm_100a): build (128-bit key loads, red.or.b64 into 64-bit blocks; block-range passes for filters beyond
// L2), popcount, fold, and a row-ordered bitmap probe (filter staged in shared memory when it fits, L2-resident
// block-range passes when it exceeds L2) whose output goes through the same scan + compaction kernels as the
// reference path.
#include <cuda_runtime.h>

#include <cstdint>
#include <cstring>
#include <random>

#include "rpt_kernels.cuh"

namespace rpt {
namespace acero {

// ---------------------------------------------------------------------------------------------
// mask table (host): a 1024+56-bit vector in which every 57-bit window holds 4 or 5 set bits; mask i is the
// window starting at bit i.  Generated like BloomFilterMasks::BloomFilterMasks() — mt19937 seeded with eight
// zeros, uniform_int_distribution<uint64_t> — so the table equals the one a libstdc++ Arrow build holds.
// ---------------------------------------------------------------------------------------------
static inline bool get_bit(const uint8_t *bits, int64_t i) {
	return (bits[i >> 3] >> (i & 7)) & 1;
}
static inline void set_bit(uint8_t *bits, int64_t i) {
	bits[i >> 3] |= static_cast<uint8_t>(1u << (i & 7));
}

void generate_mask_bits(uint8_t *out) {
	std::seed_seq seed {0, 0, 0, 0, 0, 0, 0, 0};
	std::mt19937 engine(seed);
	std::uniform_int_distribution<uint64_t> dist;
	auto random = [&](int lo, int hi) { return lo + dist(engine) % static_cast<uint64_t>(hi - lo + 1); };

	memset(out, 0, kMaskBytes);
	// the first window: 4 or 5 distinct positions
	int in_window = static_cast<int>(random(kMinBitsSet, kMaxBitsSet));
	for (int placed = 0; placed < in_window;) {
		const int pos = static_cast<int>(random(0, kBitsPerMask - 1));
		if (!get_bit(out, pos)) {
			set_bit(out, pos);
			placed++;
		}
	}
	// slide the window one bit at a time; in_window = set bits of the window that ends at the current bit
	const int64_t total_bits = kNumMasks + kBitsPerMask - 1;
	for (int64_t i = kBitsPerMask; i < total_bits; i++) {
		const int leaving = get_bit(out, i - kBitsPerMask) ? 1 : 0;
		if (leaving == 1 && in_window == kMinBitsSet) {
			set_bit(out, i); // forced 1: the window would drop below the minimum
			continue;
		}
		if (leaving == 0 && in_window == kMaxBitsSet) {
			continue; // forced 0: the window is full
		}
		if (random(0, kBitsPerMask * 2 - 1) < static_cast<uint64_t>(kMinBitsSet + kMaxBitsSet)) {
			set_bit(out, i);
			if (leaving == 0) {
				in_window++;
			}
		} else if (leaving == 1) {
			in_window--;
		}
	}
}

void expand_masks(const uint8_t *bits, uint64_t *out) {
	uint8_t padded[kMaskBytes + 8];
	memset(padded, 0, sizeof(padded));
	memcpy(padded, bits, kMaskBytes);
	for (int i = 0; i < kNumMasks; i++) {
		uint64_t w;
		memcpy(&w, padded + i / 8, 8);
		out[i] = (w >> (i % 8)) & kFullMask;
	}
}

// ---------------------------------------------------------------------------------------------
// device side
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void stage_masks(uint64_t *s_masks, const uint64_t *__restrict__ masks) {
	for (uint32_t i = threadIdx.x; i < kNumMasks; i += blockDim.x) {
		s_masks[i] = __ldg(masks + i);
	}
	__syncthreads();
}

// BlockedBloomFilter::mask(uint64_t hash): low 10 bits pick the mask, the next 6 rotate it
__device__ __forceinline__ uint64_t key_mask(const uint64_t *s_masks, uint64_t h) {
	const uint64_t m = s_masks[static_cast<uint32_t>(h) & (kNumMasks - 1)];
	const uint32_t r = (static_cast<uint32_t>(h) >> kLogNumMasks) & 63u;
	return (m << r) | (m >> ((64u - r) & 63u));
}
// BlockedBloomFilter::block_id: the bits above those
__device__ __forceinline__ uint64_t key_block(uint64_t h, uint64_t block_mask) {
	return (h >> (kLogNumMasks + 6)) & block_mask;
}

// Cache policies as on the reference path (rpt_kernels.cu): keys are read once → L1 no_allocate + L2 evict_first;
// blocks are re-read at random → L2 evict_last, so that a slice survives the key stream passing through L2
// (ncu r1z2 without hints: 17.3 GB of DRAM reads per 10^9-row pass for 8 GB of keys, L2 hit rate 38 %).
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
__device__ __forceinline__ ulonglong2 ld_keys2(const void *ptr, uint64_t pol) {
	ulonglong2 v;
	asm("ld.global.nc.L1::no_allocate.L2::cache_hint.v2.u64 {%0, %1}, [%2], %3;"
	    : "=l"(v.x), "=l"(v.y)
	    : "l"(ptr), "l"(pol));
	return v;
}
__device__ __forceinline__ uint64_t ld_key(const void *ptr, uint64_t pol) {
	uint64_t v;
	asm("ld.global.nc.L1::no_allocate.L2::cache_hint.u64 %0, [%1], %2;" : "=l"(v) : "l"(ptr), "l"(pol));
	return v;
}
__device__ __forceinline__ uint64_t ld_block(const uint64_t *ptr, uint64_t pol) {
	uint64_t v;
	asm("ld.global.nc.L1::no_allocate.L2::cache_hint.u64 %0, [%1], %2;" : "=l"(v) : "l"(ptr), "l"(pol));
	return v;
}
__device__ __forceinline__ void red_or(uint64_t *addr, uint64_t v, uint64_t pol) {
	asm volatile("red.global.or.L2::cache_hint.b64 [%0], %1, %2;" ::"l"(addr), "l"(v), "l"(pol) : "memory");
}

constexpr int kBlock = 256;

// Build.  FLAT64: one flat, all-valid 8-byte integer column — two keys per 128-bit load, four keys in flight.
template <bool FLAT64>
__global__ void __launch_bounds__(kBlock)
    abf_build_kernel(const DevCols cols, const uint64_t count, uint64_t *__restrict__ blocks, const uint64_t block_mask,
                     const uint64_t slice_lo, const uint64_t slice_len, const uint64_t *__restrict__ masks) {
	__shared__ uint64_t s_masks[kNumMasks];
	stage_masks(s_masks, masks);
	// a pass only inserts the keys whose block lies in [slice_lo, slice_lo + slice_len): the slice stays L2-resident
	const uint64_t pol_keys = policy_evict_first();
	const uint64_t pol_blocks = policy_evict_last();
	auto insert = [&](uint64_t h) {
		const uint64_t b = key_block(h, block_mask);
		if (b - slice_lo < slice_len) {
			red_or(blocks + b, key_mask(s_masks, h), pol_blocks);
		}
	};
	const uint64_t tid = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
	const uint64_t nthreads = static_cast<uint64_t>(gridDim.x) * blockDim.x;
	if (FLAT64) {
		const ulonglong2 *keys2 = reinterpret_cast<const ulonglong2 *>(cols.c[0].data);
		const uint64_t pairs = count >> 1;
		uint64_t i = tid;
		for (; i + nthreads < pairs; i += 2 * nthreads) {
			const ulonglong2 a = ld_keys2(keys2 + i, pol_keys);
			const ulonglong2 b = ld_keys2(keys2 + i + nthreads, pol_keys);
			const uint64_t h0 = murmur64(a.x), h1 = murmur64(a.y), h2 = murmur64(b.x), h3 = murmur64(b.y);
			insert(h0);
			insert(h1);
			insert(h2);
			insert(h3);
		}
		for (; i < pairs; i += nthreads) {
			const ulonglong2 a = ld_keys2(keys2 + i, pol_keys);
			const uint64_t h0 = murmur64(a.x), h1 = murmur64(a.y);
			insert(h0);
			insert(h1);
		}
		if ((count & 1) && tid == 0) {
			const uint64_t h = murmur64(__ldg(reinterpret_cast<const unsigned long long *>(cols.c[0].data) + count - 1));
			insert(h);
		}
	} else {
		for (uint64_t row = tid; row < count; row += nthreads) {
			const uint64_t h = hash_row(cols, row);
			insert(h);
		}
	}
}

// Number of set bits of blocks[0, n) added to *total (Fold's density test)
__global__ void __launch_bounds__(kBlock)
    abf_popcount_kernel(const ulonglong2 *__restrict__ blocks2, const uint64_t n2, unsigned long long *total) {
	uint64_t acc = 0;
	const uint64_t step = static_cast<uint64_t>(gridDim.x) * blockDim.x;
	for (uint64_t i = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n2; i += step) {
		const ulonglong2 v = __ldg(blocks2 + i);
		acc += __popcll(v.x) + __popcll(v.y);
	}
#pragma unroll
	for (int d = 16; d > 0; d >>= 1) {
		acc += __shfl_down_sync(0xffffffffu, acc, d);
	}
	if ((threadIdx.x & 31) == 0 && acc) {
		atomicAdd(total, static_cast<unsigned long long>(acc));
	}
}

// SingleFold(num_folds): slice 0 |= slices 1 .. 2^num_folds - 1, in place.  Entry i of slice 0 is read and written
// by one thread only, so there is no ordering to respect between threads.
__global__ void __launch_bounds__(kBlock)
    abf_fold_kernel(ulonglong2 *__restrict__ blocks2, const uint64_t slice2, const uint32_t num_slices) {
	const uint64_t step = static_cast<uint64_t>(gridDim.x) * blockDim.x;
	for (uint64_t i = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < slice2; i += step) {
		ulonglong2 acc = blocks2[i];
		uint32_t s = 1;
		for (; s + 1 < num_slices; s += 2) { // two independent loads in flight
			const ulonglong2 a = blocks2[s * slice2 + i];
			const ulonglong2 b = blocks2[(s + 1) * slice2 + i];
			acc.x |= a.x | b.x;
			acc.y |= a.y | b.y;
		}
		if (s < num_slices) {
			const ulonglong2 a = blocks2[s * slice2 + i];
			acc.x |= a.x;
			acc.y |= a.y;
		}
		blocks2[i] = acc;
	}
}

// Probe: Find for every row, answers as a row-ordered bitmap (bit r of the bitmap = row r).  A warp takes kUnroll
// consecutive 32-row words per iteration: all key loads first, then all block loads, then the ballots.
// words_padded covers whole 512-row segments so that the compaction kernels never see an uninitialised word.
//   SMEM   the whole filter (≤ kSmemBlocksMax blocks — what Fold leaves of a small selective build) is staged in
//          shared memory next to the mask table;
//   sliced passes (slice_len < num_blocks): a pass answers the rows whose block lies in its slice, which stays
//          L2-resident, and ORs into the words of the earlier passes (a word belongs to one warp in every pass).
constexpr int kUnroll = 4;
constexpr int kBlockSmem = 512;
template <bool FLAT64, bool SMEM>
__global__ void __launch_bounds__(SMEM ? kBlockSmem : kBlock)
    abf_probe_bitmap_kernel(const DevCols cols, const uint64_t count, const uint64_t words_padded,
                            const uint64_t *__restrict__ blocks, const uint64_t block_mask, const uint64_t slice_lo,
                            const uint64_t slice_len, const int first_pass, const uint64_t *__restrict__ masks,
                            uint32_t *__restrict__ bitmap32) {
	__shared__ uint64_t s_masks[kNumMasks];
	extern __shared__ __align__(16) uint64_t s_blocks[];
	if (SMEM) {
		const ulonglong2 *src = reinterpret_cast<const ulonglong2 *>(blocks);
		ulonglong2 *dst = reinterpret_cast<ulonglong2 *>(s_blocks);
		for (uint64_t i = threadIdx.x; i < (block_mask + 1) / 2; i += blockDim.x) {
			dst[i] = __ldg(src + i);
		}
	}
	stage_masks(s_masks, masks);
	const uint64_t pol_keys = policy_evict_first();
	const uint64_t pol_blocks = policy_evict_last();
	// a filter of at most 256 KiB lives in L1: bypassing it (no_allocate) measured 5.70 ms against 3.64 ms per 10^9
	// rows at 128 KiB, while from 2 MiB upwards the bypass is the faster one (profiles/sweep_r1z3_acero_hints.txt)
	const bool via_l1 = block_mask < (1ULL << 15);
	const uint32_t lane = threadIdx.x & 31u;
	const uint64_t warp = (static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
	const uint64_t nwarps = (static_cast<uint64_t>(gridDim.x) * blockDim.x) >> 5;
	const uint64_t groups = (words_padded + kUnroll - 1) / kUnroll;
	for (uint64_t g = warp; g < groups; g += nwarps) {
		uint64_t h[kUnroll];
		bool live[kUnroll];
#pragma unroll
		for (int u = 0; u < kUnroll; u++) {
			const uint64_t row = (g * kUnroll + u) * 32 + lane;
			live[u] = row < count;
			h[u] = 0;
			if (live[u]) {
				h[u] = FLAT64 ? murmur64(ld_key(reinterpret_cast<const unsigned long long *>(cols.c[0].data) + row, pol_keys))
				              : hash_row(cols, row);
			}
		}
		uint32_t prev = 0; // lane u < kUnroll holds the word of the earlier passes
		if (!first_pass && lane < kUnroll && g * kUnroll + lane < words_padded) {
			prev = bitmap32[g * kUnroll + lane];
		}
		uint64_t b[kUnroll];
#pragma unroll
		for (int u = 0; u < kUnroll; u++) {
			const uint64_t blk = key_block(h[u], block_mask);
			live[u] = live[u] && (blk - slice_lo < slice_len);
			b[u] = live[u] ? (SMEM ? s_blocks[blk] : (via_l1 ? __ldg(blocks + blk) : ld_block(blocks + blk, pol_blocks))) : 0ull;
		}
		uint32_t mine = prev;
#pragma unroll
		for (int u = 0; u < kUnroll; u++) {
			const uint64_t m = key_mask(s_masks, h[u]);
			const uint32_t word = __ballot_sync(0xffffffffu, live[u] && (b[u] & m) == m);
			if (lane == u) {
				mine |= word;
			}
		}
		if (lane < kUnroll && g * kUnroll + lane < words_padded) {
			bitmap32[g * kUnroll + lane] = mine;
		}
	}
}

// Filters beyond this many blocks are built / probed in block-range passes of this size (64 MiB, the L2-resident
// slice size measured for the reference path: profiles/mem_lab_r1.txt)
constexpr uint64_t kSliceBlocks = kSliceSectors;
// ... up to this many passes: every pass re-reads the keys, and measured on B200 (profiles/sweep_r1z_acero.txt) 32
// passes over a 2 GiB filter cost 103 ms per 10^9 probe rows against 26.9 ms for one pass of HBM gathers, while 2
// passes over 128 MiB cost 8.1 ms against 10.7 ms.  Filters that over-sized are what Fold exists for.
constexpr uint64_t kMaxPasses = 4;
// Largest filter the probe stages in shared memory: 32 KiB next to the 8 KiB mask table keeps four 512-thread CTAs
// (all 64 warps) on an SM.  This kernel hides latency with resident warps, not with staged key buffers: at 128 KiB
// (one CTA, 16 warps per SM) staging measured 185 Gkeys/s against 301 Gkeys/s for plain L1/L2 hits.
constexpr uint64_t kSmemBlocksMax = 1ULL << 12;

static inline unsigned grid_for(const DeviceInfo &dev, uint64_t items, int per_sm) {
	const uint64_t want = (items + kBlock - 1) / kBlock;
	const uint64_t cap = static_cast<uint64_t>(dev.sm_count > 0 ? dev.sm_count : 148) * per_sm;
	return static_cast<unsigned>(want < cap ? (want ? want : 1) : cap);
}

static inline bool flat64(const DevCols &cols) {
	const DevCol &c = cols.c[0];
	return cols.n == 1 && !c.sel && !c.validity && (c.type == RPT_INT64 || c.type == RPT_UINT64) &&
	       (reinterpret_cast<uintptr_t>(c.data) & 15u) == 0;
}

cudaError_t launch_build(const DeviceInfo &dev, const DevCols &cols, uint64_t count, uint64_t *blocks,
                         uint64_t num_blocks, const uint64_t *masks, cudaStream_t stream, int *launches) {
	if (count == 0) {
		return cudaSuccess;
	}
	// small builds into large filters gain nothing from passes: every touched line is a compulsory miss anyway
	const bool sliced = num_blocks > kSliceBlocks && num_blocks <= kMaxPasses * kSliceBlocks && count >= num_blocks / 8;
	const uint64_t slice = sliced ? kSliceBlocks : num_blocks;
	const bool flat = flat64(cols);
	for (uint64_t lo = 0; lo < num_blocks; lo += slice) {
		if (flat) {
			abf_build_kernel<true><<<grid_for(dev, count / 4 + 1, 8), kBlock, 0, stream>>>(cols, count, blocks,
			                                                                              num_blocks - 1, lo, slice, masks);
		} else {
			abf_build_kernel<false><<<grid_for(dev, count, 8), kBlock, 0, stream>>>(cols, count, blocks,
			                                                                        num_blocks - 1, lo, slice, masks);
		}
		if (launches) {
			(*launches)++;
		}
	}
	return cudaGetLastError();
}

cudaError_t launch_popcount(const DeviceInfo &dev, const uint64_t *blocks, uint64_t num_blocks, uint64_t *total,
                            cudaStream_t stream) {
	cudaError_t e = cudaMemsetAsync(total, 0, sizeof(uint64_t), stream);
	if (e != cudaSuccess) {
		return e;
	}
	abf_popcount_kernel<<<grid_for(dev, num_blocks / 2, 8), kBlock, 0, stream>>>(
	    reinterpret_cast<const ulonglong2 *>(blocks), num_blocks / 2, reinterpret_cast<unsigned long long *>(total));
	return cudaGetLastError();
}

cudaError_t launch_fold(const DeviceInfo &dev, uint64_t *blocks, uint64_t num_blocks, int num_folds,
                        cudaStream_t stream) {
	const uint64_t slice = num_blocks >> num_folds;
	abf_fold_kernel<<<grid_for(dev, slice / 2, 8), kBlock, 0, stream>>>(reinterpret_cast<ulonglong2 *>(blocks),
	                                                                   slice / 2, 1u << num_folds);
	return cudaGetLastError();
}

template <bool FLAT64>
static cudaError_t launch_probe_t(const DeviceInfo &dev, const DevCols &cols, uint64_t count, const uint64_t *blocks,
                                  uint64_t num_blocks, const uint64_t *masks, uint32_t *bitmap32, cudaStream_t stream,
                                  int *launches) {
	const uint64_t words_padded = (count + 511) / 512 * 16;
	const uint64_t warps_wanted = (words_padded + kUnroll - 1) / kUnroll;
	if (num_blocks <= kSmemBlocksMax && count >= (1u << 20)) {
		auto kern = abf_probe_bitmap_kernel<FLAT64, true>;
		const size_t smem = num_blocks * sizeof(uint64_t);
		cudaError_t e = ensure_max_dyn_smem<abf_probe_bitmap_kernel<FLAT64, true>>(dev);
		if (e != cudaSuccess) {
			return e;
		}
		int per_sm = 0;
		e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, kBlockSmem, smem);
		if (e != cudaSuccess) {
			return e;
		}
		if (per_sm < 1) {
			return cudaErrorLaunchOutOfResources;
		}
		uint64_t grid = static_cast<uint64_t>(dev.sm_count) * per_sm; // persistent CTAs, a multiple of the SM count
		const uint64_t want = (warps_wanted + kBlockSmem / 32 - 1) / (kBlockSmem / 32);
		grid = grid < want ? grid : want;
		kern<<<static_cast<unsigned>(grid), kBlockSmem, smem, stream>>>(cols, count, words_padded, blocks,
		                                                              num_blocks - 1, 0, num_blocks, 1, masks, bitmap32);
		if (launches) {
			(*launches)++;
		}
		return cudaGetLastError();
	}
	const bool sliced = num_blocks > kSliceBlocks && num_blocks <= kMaxPasses * kSliceBlocks;
	const uint64_t slice = sliced ? kSliceBlocks : num_blocks;
	const unsigned grid = grid_for(dev, warps_wanted * 32, 8);
	for (uint64_t lo = 0; lo < num_blocks; lo += slice) {
		abf_probe_bitmap_kernel<FLAT64, false><<<grid, kBlock, 0, stream>>>(
		    cols, count, words_padded, blocks, num_blocks - 1, lo, slice, lo == 0 ? 1 : 0, masks, bitmap32);
		if (launches) {
			(*launches)++;
		}
	}
	return cudaGetLastError();
}

cudaError_t launch_probe(const DeviceInfo &dev, const DevCols &cols, uint64_t count, const uint64_t *blocks,
                         uint64_t num_blocks, const uint64_t *masks, uint32_t *sel_out, uint64_t *found_out,
                         uint32_t row_offset, void *scratch, cudaStream_t stream, int *launches) {
	if (count == 0) {
		return cudaMemsetAsync(found_out, 0, sizeof(uint64_t), stream);
	}
	uint32_t *bitmap32 = bitmap_words(scratch, count);
	const cudaError_t e = flat64(cols)
	                          ? launch_probe_t<true>(dev, cols, count, blocks, num_blocks, masks, bitmap32, stream, launches)
	                          : launch_probe_t<false>(dev, cols, count, blocks, num_blocks, masks, bitmap32, stream, launches);
	if (e != cudaSuccess) {
		return e;
	}
	return launch_bitmap_to_sel(dev, count, sel_out, found_out, row_offset, scratch, stream, launches);
}

} // namespace acero
} // namespace rpt
