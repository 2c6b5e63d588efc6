// rpt_kernels.cuh — launcher interface between the C ABI (rpt_capi.cu) and the sm_100a kernels
// (rpt_kernels.cu).  Everything takes device pointers and a stream; nothing synchronises.
#pragma once
#include <cuda_runtime.h>

#include <atomic>
#include <cstdint>

#include "rpt_hash.cuh"

namespace rpt {

struct DeviceInfo {
	int device = -1;
	int sm_count = 0;
	int l2_bytes = 0;
	size_t smem_optin = 0;
};

// cudaFuncAttributeMaxDynamicSharedMemorySize is raised ONCE per kernel instantiation and device, to the most the
// device allows next to the kernel's static shared memory.  Every caller sets the same value, so concurrent
// launches of differently sized filters cannot lower the limit under each other (a per-launch, per-size call could).
template <auto Kern>
inline cudaError_t ensure_max_dyn_smem(const DeviceInfo &dev) {
	static std::atomic<uint64_t> done {0};
	const uint64_t bit = 1ULL << (static_cast<unsigned>(dev.device) & 63u);
	if (done.load(std::memory_order_acquire) & bit) {
		return cudaSuccess;
	}
	cudaFuncAttributes attr;
	cudaError_t e = cudaFuncGetAttributes(&attr, Kern);
	if (e != cudaSuccess) {
		return e;
	}
	const long long room = static_cast<long long>(dev.smem_optin) - static_cast<long long>(attr.sharedSizeBytes);
	e = cudaFuncSetAttribute(Kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(room > 0 ? room : 0));
	if (e == cudaSuccess) {
		done.fetch_or(bit, std::memory_order_release);
	}
	return e;
}

// Largest filter (in 64-bit sectors) that is staged in shared memory by the probe kernel.
constexpr uint64_t kSmemFilterMaxSectors = 1ULL << 14; // 128 KiB
// Filters larger than this are built / probed in sector-range passes of this many sectors, each of which
// stays L2-resident (64 MiB: measured sweet spot on B200's 126 MB, two-die L2; profiles/mem_lab_r1.txt).
constexpr uint64_t kSliceSectors = 1ULL << 23;

// Scratch the probe kernel needs for `count` rows: tile descriptors + ticket counter (bytes).
// prefilter: the probe will be given a folded pre-filter (launch_probe's `pre`), whose partition uses one more region
size_t probe_scratch_bytes(uint64_t count, uint64_t num_sectors, bool prefilter = false);

// Column range (UpdateMinMax, physical_create_bf.cpp:82-176).  Keys are order-preserving uint64 images of the
// values (signed: value ^ 2^63 after sign extension; float/double: sign-magnitude flip); the host decodes them.
// The reduction is fused into the build kernel (launch_build's `mm` argument): the keys are read once.
struct MinMaxState {
	uint64_t min_key; // initialised to ~0
	uint64_t max_key; // initialised to 0
	uint32_t has_value;
	uint32_t has_nan;
};

// Build: filter[h & (S-1)] |= GetMask(h) for every row (InsertHashes, bloom_filter.cpp:43-48,63-70).
// hash_scratch (optional, build_scratch_bytes(cols, count, num_sectors) bytes when that is not 0): a build of a
// generic key source in several sector-range passes hashes the rows once, in its first pass, and runs the other
// passes on the hashes.
size_t build_scratch_bytes(const DevCols &cols, uint64_t count, uint64_t num_sectors);
cudaError_t launch_build(const DeviceInfo &dev, const DevCols &cols, uint64_t count, uint64_t *words,
                         uint64_t num_sectors, cudaStream_t stream, int *launches, MinMaxState *mm = nullptr,
                         uint64_t *hash_scratch = nullptr);

// Probe: ascending row ids of passing rows → sel_out, count → *found_out (LookupHashes,
// bloom_filter.cpp:50-61,72-78).  `scratch` holds probe_scratch_bytes(count) bytes; it is zeroed
// here.  row_offset is added to every emitted row id; translate (optional) maps a row id through
// an incoming selection (BFTableFilter::Filter, bloom_filter.cpp:117-123).
// gate (optional): per-64-MiB-slice events the probe waits for right before the first kernel that reads the slice
// (multi-GPU merge overlapped with the probe).
struct SliceGate {
	cudaEvent_t ev[8]; // S ≤ 2^26 sectors = 8 slices of kSliceSectors
	int n;
};
// pre (optional): the folded pre-filter of `words` (launch_fold below).  Filters of 256 / 512 MiB then take the
// pre-filtered partition: every key meets the L2-resident folded filter first and only its survivors are looked up
// in the real words — same answers (a key the folded filter rejects cannot pass the real one), far fewer records.
struct Prefilter {
	const uint64_t *folded; // `sectors` words: folded[i] = OR of words[i + j·sectors]
	uint64_t sectors;       // kSliceSectors
	uint64_t *stats;        // optional device counters: [0] += rows probed, [1] += rows that survived the folded filter
};
bool prefilter_applicable(uint64_t num_sectors);
cudaError_t launch_fold(const DeviceInfo &dev, const uint64_t *words, uint64_t num_sectors, uint64_t *folded,
                        uint64_t fold_sectors, uint64_t *popcount_out, cudaStream_t stream);
cudaError_t launch_probe(const DeviceInfo &dev, const DevCols &cols, uint64_t count, const uint64_t *words,
                         uint64_t num_sectors, uint32_t *sel_out, uint64_t *found_out, uint32_t row_offset,
                         const uint32_t *translate, void *scratch, cudaStream_t stream, int *launches,
                         const SliceGate *gate = nullptr, const Prefilter *pre = nullptr);

// words[i] |= parts[p*S + i] for p < nparts
cudaError_t launch_or_parts(const DeviceInfo &dev, uint64_t *words, const uint64_t *parts, uint64_t num_sectors,
                            uint64_t nparts, cudaStream_t stream);

// words[off .. off+len) |= srcs[k][off .. off+len) for k < nsrc (peer-memory reduce-scatter-OR leg)
struct PtrList {
	const uint64_t *p[16];
	int n;
};
cudaError_t launch_or_ptrs(const DeviceInfo &dev, uint64_t *words, const PtrList &srcs, uint64_t off, uint64_t len,
                           cudaStream_t stream);

// barrier flags of the library-owned peer merge (rpt_bf_peer_allreduce): store `value` into every listed (peer) word
struct FlagList {
	uint32_t *p[16];
	int n;
};
cudaError_t launch_signal_flags(const FlagList &dst, uint32_t value, cudaStream_t stream);
// poll flags[t] (t < n, t != skip) until each is >= value (fallback of cuStreamWaitValue32)
cudaError_t launch_wait_flags(const uint32_t *flags, int n, int skip, uint32_t value, cudaStream_t stream);

// LookupOne on pre-hashed values (BFTableFilter::FilterValue / CheckStatistics)
cudaError_t launch_lookup_hashes(const uint64_t *words, uint64_t num_sectors, const uint64_t *hashes, uint64_t count,
                                 uint8_t *hits, cudaStream_t stream);

// Zone-map pruning (BFTableFilter::CheckStatistics, duckdb bloom_filter.cpp:138-196): one entry per row group
enum : uint32_t { kStatsHash32 = 0, kStatsHash64 = 1, kStatsHash128 = 2, kStatsSkip = 3 };
enum : uint8_t { kStatsNoPruning = 0, kStatsAlwaysTrue = 1, kStatsAlwaysFalse = 2 }; // = enum rpt_prune
struct StatsRange {
	uint64_t first_lower; // min as a 128-bit two's-complement integer (sign- or zero-extended from its type)
	uint64_t first_upper;
	uint32_t range;       // max - min (≤ 2048): range + 1 consecutive values are looked up
	uint32_t hash_class;  // how the type hashes; kStatsSkip: the host already decided NO_PRUNING_POSSIBLE
};
cudaError_t launch_check_statistics(const uint64_t *words, uint64_t num_sectors, const StatsRange *ranges, uint64_t count,
                                    uint8_t *results, cudaStream_t stream);

// Hashes of the entries of one column (no selection applied): the dictionary of a dictionary vector, hashed once
cudaError_t launch_hash_entries(const DeviceInfo &dev, const DevCol &col, uint64_t entries, uint64_t *out,
                                cudaStream_t stream);

// out[i] = hash of row i of a flat VARCHAR column (offsets + bytes, optional validity), staged through shared memory
cudaError_t launch_hash_varchar(const DeviceInfo &dev, const DevCol &col, uint64_t count, uint64_t *out,
                                cudaStream_t stream);

// Row-ordered bitmap → ordered selection vector (count_segments + scans + compaction), shared with rpt_acero.cu
size_t bitmap_scratch_bytes(uint64_t count);
uint32_t *bitmap_words(void *scratch, uint64_t count);
cudaError_t launch_bitmap_to_sel(const DeviceInfo &dev, uint64_t count, uint32_t *sel_out, uint64_t *found_out,
                                 uint32_t row_offset, void *scratch, cudaStream_t stream, int *launches);

// README-described filter variant (reference README.md:23-32; Arrow Acero BlockedBloomFilter) — rpt_acero.cu.
// Not on the parity path: the reference holds no code for it.
namespace acero {
constexpr int kBitsPerMask = 57, kMinBitsSet = 4, kMaxBitsSet = 5, kLogNumMasks = 10, kNumMasks = 1 << kLogNumMasks;
constexpr int kMaskBytes = (kNumMasks + 64) / 8; // bit vector the 1024 overlapping masks are windows of
constexpr uint64_t kFullMask = (1ULL << kBitsPerMask) - 1;
constexpr int kLogNumBlocksMin = 4; // Fold stops once log_num_blocks <= 4
void generate_mask_bits(uint8_t *out /* kMaskBytes */);
void expand_masks(const uint8_t *bits, uint64_t *out /* kNumMasks */);
cudaError_t launch_build(const DeviceInfo &dev, const DevCols &cols, uint64_t count, uint64_t *blocks,
                         uint64_t num_blocks, const uint64_t *masks, cudaStream_t stream, int *launches);
cudaError_t launch_popcount(const DeviceInfo &dev, const uint64_t *blocks, uint64_t num_blocks, uint64_t *total,
                            cudaStream_t stream);
cudaError_t launch_fold(const DeviceInfo &dev, uint64_t *blocks, uint64_t num_blocks, int num_folds,
                        cudaStream_t stream);
cudaError_t launch_probe(const DeviceInfo &dev, const DevCols &cols, uint64_t count, const uint64_t *blocks,
                         uint64_t num_blocks, const uint64_t *masks, uint32_t *sel_out, uint64_t *found_out,
                         uint32_t row_offset, void *scratch, cudaStream_t stream, int *launches);
} // namespace acero

// SURVEY §8(d) synthetic columns
cudaError_t launch_synth_build(uint64_t *out, uint64_t first, uint64_t count, cudaStream_t stream);
cudaError_t launch_synth_probe(uint64_t *out, uint64_t first, uint64_t count, uint64_t n_build, uint32_t sel_ppm,
                               cudaStream_t stream);

// sets rpt_gpu_last_error() from another translation unit of the library (defined in rpt_capi.cu); returns 1
int report_error(const char *msg);

} // namespace rpt
