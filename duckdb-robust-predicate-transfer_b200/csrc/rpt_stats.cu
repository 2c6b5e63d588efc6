// Zone-map pruning against the filter: the lookups of BFTableFilter::CheckStatistics /
// TemplatedCheckStatistics<T> (reference duckdb/src/planner/filter/bloom_filter.cpp:138-196), batched over row
// groups.  The typed range arithmetic (min > max, T-typed subtraction overflow, the 2048 limit) is scalar work the
// C ABI does on the host (rpt_capi.cu); what arrives here is, per row group, the first value as a 128-bit integer,
// the number of further values and how the type hashes.  One warp per row group: lane l looks up values
// l, l+32, …; hits are summed with a warp reduction.
#include "rpt_hash.cuh"
#include "rpt_kernels.cuh"

namespace rpt {

__global__ void __launch_bounds__(256)
    check_statistics_kernel(const uint64_t *__restrict__ words, const uint64_t sector_mask,
                            const StatsRange *__restrict__ ranges, const uint64_t count, uint8_t *__restrict__ results) {
	const uint32_t lane = threadIdx.x & 31u;
	const uint64_t warp = (static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
	if (warp >= count) {
		return;
	}
	const StatsRange r = ranges[warp];
	if (r.hash_class == kStatsSkip) { // lane-uniform: NO_PRUNING_POSSIBLE was decided on the host
		if (lane == 0) {
			results[warp] = kStatsNoPruning;
		}
		return;
	}
	uint32_t hits = 0;
	for (uint32_t i = lane; i <= r.range; i += 32u) {
		const uint64_t lower = r.first_lower + i;
		uint64_t h;
		if (r.hash_class == kStatsHash32) { // hash.hpp:49-61: every integral type up to 32 bits hashes (uint32_t)value
			h = murmur64(static_cast<uint64_t>(static_cast<uint32_t>(lower)));
		} else if (r.hash_class == kStatsHash64) {
			h = murmur64(lower);
		} else { // hash.cpp:22-30: hugeint / uhugeint = mm(lower) ^ mm(upper)
			h = murmur64(lower) ^ murmur64(r.first_upper + (lower < r.first_lower ? 1ULL : 0ULL));
		}
		hits += sector_test(__ldg(words + (h & sector_mask)), static_cast<uint32_t>(h >> 32));
	}
	hits = __reduce_add_sync(0xffffffffu, hits);
	if (lane == 0) {
		results[warp] = hits == 0 ? kStatsAlwaysFalse : (hits == r.range + 1u ? kStatsAlwaysTrue : kStatsNoPruning);
	}
}

cudaError_t launch_check_statistics(const uint64_t *words, uint64_t num_sectors, const StatsRange *ranges, uint64_t count,
                                    uint8_t *results, cudaStream_t stream) {
	if (count == 0) {
		return cudaSuccess;
	}
	const uint64_t blocks = (count + 7) / 8; // 8 warps per CTA
	check_statistics_kernel<<<static_cast<unsigned>(blocks), 256, 0, stream>>>(words, num_sectors - 1, ranges, count,
	                                                                         results);
	return cudaGetLastError();
}

// out[i] = hash of entry i of one column (NULL entries → NULL_HASH): the dictionary of a dictionary vector
__global__ void __launch_bounds__(256)
    hash_entries_kernel(const DevCol col, const uint64_t entries, uint64_t *__restrict__ out) {
	const uint64_t step = static_cast<uint64_t>(gridDim.x) * blockDim.x;
	for (uint64_t i = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < entries; i += step) {
		out[i] = hash_col_row(col, i);
	}
}

// ---- flat VARCHAR columns: hash once, staged through shared memory ---------------------------------------------
// A thread that hashes its own string straight from global memory walks a chain of dependent, badly coalesced
// loads (offset → first word → next word …): measured 32 Gkeys/s on 19-byte strings.  Here a warp takes 128
// consecutive rows per step: their 129 offsets and the (contiguous) bytes of all 128 strings arrive through
// asynchronous coalesced copies (LDGSTS) in warp-private shared memory, then lane l hashes rows 4l…4l+3 out of
// shared memory and writes four consecutive 64-bit hashes.  Software pipeline per warp: while step i is hashed,
// the bytes of step i+1 and the offsets of step i+2 are in flight (two byte buffers, three offset buffers).
// The hashes feed the ordinary build / probe kernels as a pre-hashed flat column (FlatU64SrcT<false, true>).
// Steps whose bytes do not fit the buffer (strings averaging more than 31 bytes), and the ragged last step, are
// hashed straight from global memory — for long strings that path streams through cache lines anyway.
constexpr int kVarRows = 128;                 // rows per warp step
constexpr int kVarBufWords = 512;             // 4 KiB of string bytes per buffer
constexpr int kVarOffWords = kVarRows + 1 + (kVarRows / 4) + 3; // offsets, padded by one word per 4 (conflict-free)
constexpr int kVarWarps = 4;                  // ~40 KiB of static shared memory per CTA, 5 CTAs per SM

__device__ __forceinline__ void cp_async_4(void *smem, const void *gmem) {
	asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(static_cast<uint32_t>(__cvta_generic_to_shared(smem))),
	             "l"(gmem)
	             : "memory");
}
__device__ __forceinline__ void cp_async_8(void *smem, const void *gmem) {
	asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(static_cast<uint32_t>(__cvta_generic_to_shared(smem))),
	             "l"(gmem)
	             : "memory");
}

__device__ __forceinline__ void cp_async_16_cg(void *smem, const void *gmem) {
	asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(static_cast<uint32_t>(__cvta_generic_to_shared(smem))),
	             "l"(gmem)
	             : "memory");
}

// Hash(string_t) (hash.cpp:77-151) of a string that starts byte_pos bytes into buf (shared memory).  Branch-free
// on the alignment: block j is cut out of three consecutive 32-bit words with two byte permutes (the selector
// depends on byte_pos & 3 only); the last, partial block is masked — zero padding, as the reference's tail step.
// Reads up to 8 bytes past the string (masked out; the buffer is padded).
__device__ __forceinline__ uint64_t hash_bytes_smem(const uint64_t *buf, uint32_t byte_pos, uint32_t len) {
	uint64_t h = 0xe17a1465ULL ^ (static_cast<uint64_t>(len) * 0xc6a4a7935bd1e995ULL);
	const uint32_t *w = reinterpret_cast<const uint32_t *>(buf) + (byte_pos >> 2);
	const uint32_t sel = 0x3210u + 0x1111u * (byte_pos & 3u);
	uint32_t x0 = w[0];
	for (uint32_t done = 0; done < len; done += 8u) {
		const uint32_t x1 = w[1], x2 = w[2];
		uint64_t blk = (static_cast<uint64_t>(__byte_perm(x1, x2, sel)) << 32) | __byte_perm(x0, x1, sel);
		const uint32_t rem = len - done; // ≥ 1
		blk &= (~0ULL) >> (64u - (rem < 8u ? rem : 8u) * 8u);
		h ^= blk;
		h *= kMurmurK;
		x0 = x2;
		w += 2;
	}
	return murmur64(h);
}

__device__ __forceinline__ uint32_t var_off_index(uint32_t j) {
	return j + (j >> 2);
}
__device__ __forceinline__ void var_issue_offsets(uint32_t *offs, const uint32_t *goff, uint64_t row0, uint64_t count,
                                                  uint32_t lane) {
	const uint32_t rows = count - row0 < kVarRows ? static_cast<uint32_t>(count - row0) : kVarRows;
	for (uint32_t j = lane; j <= rows; j += 32u) {
		cp_async_4(offs + var_off_index(j), goff + row0 + j);
	}
}
// Issues the copy of a step's string bytes (its offsets are in shared memory already).  Returns whether the step is
// staged (warp-uniform); skew: byte position in buf = string offset + skew.
__device__ __forceinline__ bool var_issue_bytes(uint64_t *buf, const uint32_t *offs, const uint8_t *bytes, uint64_t row0,
                                                uint64_t count, uint32_t lane, uint32_t *skew) {
	const uint32_t rows = count - row0 < kVarRows ? static_cast<uint32_t>(count - row0) : kVarRows;
	const uint32_t b0 = offs[0], b1 = offs[var_off_index(rows)];
	const uintptr_t first = reinterpret_cast<uintptr_t>(bytes + b0);
	const uintptr_t base = first & ~static_cast<uintptr_t>(15); // 16-byte aligned start of the copy
	const uint32_t words = static_cast<uint32_t>((reinterpret_cast<uintptr_t>(bytes + b1) - base + 7) >> 3);
	*skew = static_cast<uint32_t>(first - base) - b0;
	if (rows != kVarRows || words > kVarBufWords) {
		return false;
	}
	// 16-byte copies; an odd last word goes alone, so nothing past the 8-byte multiple that holds the last byte is read
	const ulonglong2 *g = reinterpret_cast<const ulonglong2 *>(base);
	ulonglong2 *d = reinterpret_cast<ulonglong2 *>(buf);
	for (uint32_t u = lane; u < (words >> 1); u += 32u) {
		cp_async_16_cg(d + u, g + u);
	}
	if ((words & 1u) && lane == 0) {
		cp_async_8(buf + words - 1, reinterpret_cast<const uint64_t *>(base) + words - 1);
	}
	return true;
}

__global__ void __launch_bounds__(kVarWarps * 32)
    hash_varchar_kernel(const DevCol col, const uint64_t count, uint64_t *__restrict__ out) {
	__shared__ __align__(16) uint64_t s_buf[kVarWarps][2][kVarBufWords + 2];
	__shared__ uint32_t s_off[kVarWarps][3][kVarOffWords];
	const uint32_t lane = threadIdx.x & 31u;
	const uint32_t warp = threadIdx.x >> 5;
	const uint32_t *const goff = reinterpret_cast<const uint32_t *>(col.data);
	const uint8_t *const bytes = reinterpret_cast<const uint8_t *>(col.aux);
	const uint64_t num_steps = (count + kVarRows - 1) / kVarRows;
	const uint64_t stride = static_cast<uint64_t>(gridDim.x) * kVarWarps;
	uint64_t step = static_cast<uint64_t>(blockIdx.x) * kVarWarps + warp;
	if (step >= num_steps) {
		return;
	}
	// prologue: offsets of the first two steps, then the bytes of the first
	var_issue_offsets(s_off[warp][0], goff, step * kVarRows, count, lane);
	if (step + stride < num_steps) {
		var_issue_offsets(s_off[warp][1], goff, (step + stride) * kVarRows, count, lane);
	}
	asm volatile("cp.async.wait_all;" ::: "memory");
	__syncwarp();
	uint32_t skew = 0, skew_next = 0;
	bool staged = var_issue_bytes(s_buf[warp][0], s_off[warp][0], bytes, step * kVarRows, count, lane, &skew);
	bool staged_next = false;
	uint32_t k = 0, par = 0;
	for (; step < num_steps; step += stride) {
		const uint64_t row0 = step * kVarRows;
		const uint32_t rows = count - row0 < kVarRows ? static_cast<uint32_t>(count - row0) : kVarRows;
		const uint32_t k1 = k == 2 ? 0 : k + 1, k2 = k1 == 2 ? 0 : k1 + 1;
		asm volatile("cp.async.wait_all;" ::: "memory");
		__syncwarp(); // this step's bytes and the next step's offsets have landed; last step's buffers are free
		if (step + stride < num_steps) {
			staged_next = var_issue_bytes(s_buf[warp][par ^ 1u], s_off[warp][k1], bytes, (step + stride) * kVarRows, count,
			                              lane, &skew_next);
		}
		if (step + 2 * stride < num_steps) {
			var_issue_offsets(s_off[warp][k2], goff, (step + 2 * stride) * kVarRows, count, lane);
		}
		const uint64_t *const buf = s_buf[warp][par];
		const uint32_t *const offs = s_off[warp][k];
		uint64_t h[4];
#pragma unroll
		for (int i = 0; i < 4; i++) {
			const uint32_t r = lane * 4u + i;
			h[i] = 0;
			if (r < rows) {
				const uint64_t row = row0 + r;
				if (col.validity && !((__ldg(col.validity + (row >> 6)) >> (row & 63)) & 1ULL)) {
					h[i] = kNullHash;
				} else {
					const uint32_t s = offs[var_off_index(r)], e = offs[var_off_index(r + 1)];
					h[i] = staged ? hash_bytes_smem(buf, s + skew, e - s) : hash_bytes(bytes + s, e - s);
				}
			}
		}
		if (rows == kVarRows) {
			ulonglong2 *o = reinterpret_cast<ulonglong2 *>(out + row0 + lane * 4u); // 16-byte aligned: row0 % 128 == 0
			o[0] = make_ulonglong2(h[0], h[1]);
			o[1] = make_ulonglong2(h[2], h[3]);
		} else {
#pragma unroll
			for (int i = 0; i < 4; i++) {
				if (lane * 4u + i < rows) {
					out[row0 + lane * 4u + i] = h[i];
				}
			}
		}
		k = k1;
		par ^= 1u;
		staged = staged_next;
		skew = skew_next;
	}
}

cudaError_t launch_hash_varchar(const DeviceInfo &dev, const DevCol &col, uint64_t count, uint64_t *out,
                                cudaStream_t stream) {
	if (count == 0) {
		return cudaSuccess;
	}
	int per_sm = 0;
	cudaError_t e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, hash_varchar_kernel, kVarWarps * 32, 0);
	if (e != cudaSuccess) {
		return e;
	}
	uint64_t grid = static_cast<uint64_t>(dev.sm_count) * (per_sm > 0 ? per_sm : 1);
	const uint64_t want = ((count + kVarRows - 1) / kVarRows + kVarWarps - 1) / kVarWarps;
	grid = grid < want ? grid : want;
	hash_varchar_kernel<<<static_cast<unsigned>(grid), kVarWarps * 32, 0, stream>>>(col, count, out);
	return cudaGetLastError();
}

cudaError_t launch_hash_entries(const DeviceInfo &dev, const DevCol &col, uint64_t entries, uint64_t *out,
                                cudaStream_t stream) {
	if (entries == 0) {
		return cudaSuccess;
	}
	uint64_t blocks = (entries + 255) / 256;
	const uint64_t cap = static_cast<uint64_t>(dev.sm_count) * 8;
	blocks = blocks < cap ? blocks : cap;
	hash_entries_kernel<<<static_cast<unsigned>(blocks), 256, 0, stream>>>(col, entries, out);
	return cudaGetLastError();
}

} // namespace rpt
