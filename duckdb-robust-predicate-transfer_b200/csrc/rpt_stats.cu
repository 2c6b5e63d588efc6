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
