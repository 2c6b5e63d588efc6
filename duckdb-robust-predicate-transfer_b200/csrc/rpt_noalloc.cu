// rpt_noalloc.cu — rpt_bf_probe_u64_noalloc: an allocation-free variant of the bitmap probe.
//
// No counterpart in the reference: added for callers that cannot call
// cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).  It is the row-ordered-bitmap probe
// of rpt_kernels.cu (hash → sector test → bitmap → scan → ascending row ids; reference semantics:
// PTBloomFilter::LookupSel, src/bloom_filter.cpp:60-68) restricted to one non-NULL 64-bit integer key column that
// already lies in device memory, with its scratch (bitmap + per-block counts) in a file-scope __device__ array
// instead of the handle's pool allocation.  Results are identical to rpt_bf_probe's on the same inputs.
//
// The scratch is ONE array per device that loads the library (held until exit): every call shares it, so two calls
// that overlap on the GPU (different streams, or different host threads) race.  Nothing in it has to start at zero:
// every word a call reads was written earlier by that same call (mark writes all bitmap words and block counts of its
// blocks, scan rewrites the counts in place), so stale contents of the previous call are never observed.
#include "rpt_hash.cuh"
#include "rpt_kernels.cuh"

#include <cuda_runtime.h>

#include <cstdio>
#include <mutex>

namespace {

constexpr uint64_t kMaxRows = RPT_NOALLOC_MAX_ROWS; // 2^24 rows per call
constexpr int kBlockThreads = 256;
constexpr int kWarps = kBlockThreads / 32;
constexpr int kWordsPerWarp = 8;                            // 8 × 32 consecutive rows per warp
constexpr int kWordsPerBlock = kWarps * kWordsPerWarp;      // 64 bitmap words
constexpr uint64_t kRowsPerBlock = 32ull * kWordsPerBlock;  // 2048 rows
constexpr uint64_t kMaxBlocks = kMaxRows / kRowsPerBlock;   // 8192
constexpr int kScanThreads = 1024;
constexpr int kScanPerThread = static_cast<int>(kMaxBlocks / kScanThreads); // 8

struct Scratch {
	uint32_t bitmap[kMaxRows / 32]; // 2 MiB: bit r%32 of word r/32 = row r passes the filter
	uint32_t block_base[kMaxBlocks]; // 32 KiB: hits per 2048-row block, then their exclusive prefix sums
};

// 2 MiB + 32 KiB of device memory per GPU, for the life of the process
static __device__ __align__(256) unsigned char g_noalloc_scratch[sizeof(Scratch)];

__global__ void __launch_bounds__(kBlockThreads)
    noalloc_mark_kernel(const uint64_t *__restrict__ keys, uint64_t count, const uint64_t *__restrict__ filter,
                        uint64_t sector_mask, Scratch *s) {
	__shared__ uint32_t warp_hits[kWarps];
	const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	const uint64_t word0 = static_cast<uint64_t>(blockIdx.x) * kWordsPerBlock + warp * kWordsPerWarp;
	uint32_t hits = 0;
#pragma unroll
	for (int i = 0; i < kWordsPerWarp; i++) {
		const uint64_t row = (word0 + i) * 32 + lane;
		uint32_t bit = 0;
		if (row < count) {
			const uint64_t h = rpt::murmur64(__ldg(keys + row));
			const uint64_t w = __ldg(filter + (h & sector_mask));
			bit = rpt::sector_test(w, static_cast<uint32_t>(h >> 32));
		}
		const uint32_t word = __ballot_sync(0xffffffffu, bit != 0);
		if (lane == 0) {
			s->bitmap[word0 + i] = word;
		}
		hits += __popc(word);
	}
	if (lane == 0) {
		warp_hits[warp] = hits;
	}
	__syncthreads();
	if (threadIdx.x == 0) {
		uint32_t total = 0;
#pragma unroll
		for (int w = 0; w < kWarps; w++) {
			total += warp_hits[w];
		}
		s->block_base[blockIdx.x] = total;
	}
}

// one CTA: block counts → exclusive prefix sums in place, the grand total to *found_out
__global__ void __launch_bounds__(kScanThreads)
    noalloc_scan_kernel(Scratch *s, uint32_t nblocks, uint64_t *found_out) {
	__shared__ uint32_t warp_sums[kScanThreads / 32];
	const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	uint32_t v[kScanPerThread];
	uint32_t mine = 0;
#pragma unroll
	for (int i = 0; i < kScanPerThread; i++) {
		const uint32_t b = threadIdx.x * kScanPerThread + i;
		v[i] = b < nblocks ? s->block_base[b] : 0u;
		mine += v[i];
	}
	uint32_t incl = mine;
#pragma unroll
	for (int d = 1; d < 32; d <<= 1) {
		const uint32_t up = __shfl_up_sync(0xffffffffu, incl, d);
		if (lane >= d) {
			incl += up;
		}
	}
	if (lane == 31) {
		warp_sums[warp] = incl;
	}
	__syncthreads();
	if (warp == 0) {
		uint32_t ws = warp_sums[lane];
#pragma unroll
		for (int d = 1; d < 32; d <<= 1) {
			const uint32_t up = __shfl_up_sync(0xffffffffu, ws, d);
			if (lane >= d) {
				ws += up;
			}
		}
		warp_sums[lane] = ws; // inclusive over warps
	}
	__syncthreads();
	uint32_t base = (warp ? warp_sums[warp - 1] : 0u) + incl - mine;
#pragma unroll
	for (int i = 0; i < kScanPerThread; i++) {
		const uint32_t b = threadIdx.x * kScanPerThread + i;
		if (b < nblocks) {
			s->block_base[b] = base;
		}
		base += v[i];
	}
	if (threadIdx.x == kScanThreads - 1) {
		*found_out = warp_sums[kScanThreads / 32 - 1];
	}
}

__global__ void __launch_bounds__(kBlockThreads)
    noalloc_compact_kernel(const Scratch *s, uint32_t *__restrict__ sel_out) {
	__shared__ uint32_t words[kWordsPerBlock];
	__shared__ uint32_t offs[kWordsPerBlock];
	const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	const uint64_t word0 = static_cast<uint64_t>(blockIdx.x) * kWordsPerBlock;
	if (threadIdx.x < kWordsPerBlock) {
		words[threadIdx.x] = s->bitmap[word0 + threadIdx.x];
	}
	__syncthreads();
	if (warp == 0) { // exclusive prefix of the 64 popcounts
		const uint32_t a = __popc(words[lane]), b = __popc(words[lane + 32]);
		uint32_t ia = a, ib = b;
#pragma unroll
		for (int d = 1; d < 32; d <<= 1) {
			const uint32_t ua = __shfl_up_sync(0xffffffffu, ia, d), ub = __shfl_up_sync(0xffffffffu, ib, d);
			if (lane >= d) {
				ia += ua;
				ib += ub;
			}
		}
		const uint32_t first_half = __shfl_sync(0xffffffffu, ia, 31);
		offs[lane] = ia - a;
		offs[lane + 32] = first_half + ib - b;
	}
	__syncthreads();
	const uint32_t base = s->block_base[blockIdx.x];
#pragma unroll
	for (int i = 0; i < kWordsPerWarp; i++) {
		const int wi = warp * kWordsPerWarp + i;
		const uint32_t w = words[wi];
		if ((w >> lane) & 1u) {
			const uint32_t pos = base + offs[wi] + __popc(w & ((1u << lane) - 1u));
			sel_out[pos] = static_cast<uint32_t>((word0 + wi) * 32 + lane);
		}
	}
}

// In host code `g_noalloc_scratch` is a host-side stand-in for the symbol, not a device address: the address the
// kernels get comes from cudaGetSymbolAddress, asked once per device with that device current.
constexpr int kMaxDevices = 64;
std::mutex g_mu;
Scratch *g_scratch_of[kMaxDevices] = {};

int fail(const char *fmt, const char *what, cudaError_t e) {
	char buf[384];
	snprintf(buf, sizeof(buf), fmt, what, cudaGetErrorString(e));
	return rpt::report_error(buf);
}

// caller has made `device` current
int scratch_for(int device, Scratch **out) {
	std::lock_guard<std::mutex> lock(g_mu);
	if (!g_scratch_of[device]) {
		void *p = nullptr;
		const cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
		if (e != cudaSuccess || !p) {
			return fail("rpt_bf_probe_u64_noalloc: %s failed: %s", "cudaGetSymbolAddress(g_noalloc_scratch)", e);
		}
		g_scratch_of[device] = static_cast<Scratch *>(p);
	}
	*out = g_scratch_of[device];
	return 0;
}

struct DeviceScope { // makes `device` current for the call, restores the caller's on exit
	int prev = -1;
	bool switched = false;
	cudaError_t enter(int device) {
		cudaError_t e = cudaGetDevice(&prev);
		if (e == cudaSuccess && prev != device) {
			e = cudaSetDevice(device);
			switched = e == cudaSuccess;
		}
		return e;
	}
	~DeviceScope() {
		if (switched) {
			cudaSetDevice(prev);
		}
	}
};

} // namespace

extern "C" int rpt_bf_probe_u64_noalloc(const uint64_t *dev_words, uint64_t num_sectors, const uint64_t *dev_keys,
                                        uint64_t count, uint32_t *dev_sel_out, uint64_t *dev_found_out, int device,
                                        rpt_stream_t stream) {
	if (!dev_words || !dev_found_out || (count && (!dev_keys || !dev_sel_out))) {
		return rpt::report_error("rpt_bf_probe_u64_noalloc: null pointer argument");
	}
	if (num_sectors == 0 || (num_sectors & (num_sectors - 1)) != 0) {
		return rpt::report_error("rpt_bf_probe_u64_noalloc: num_sectors must be a power of two (rpt_bf_num_sectors)");
	}
	if (count > kMaxRows) {
		char buf[256];
		snprintf(buf, sizeof(buf),
		         "rpt_bf_probe_u64_noalloc: count %llu exceeds RPT_NOALLOC_MAX_ROWS (%llu): split the batch or use "
		         "rpt_bf_probe",
		         static_cast<unsigned long long>(count), static_cast<unsigned long long>(kMaxRows));
		return rpt::report_error(buf);
	}
	if (device < 0 || device >= kMaxDevices) {
		return rpt::report_error("rpt_bf_probe_u64_noalloc: device out of range");
	}
	DeviceScope scope;
	cudaError_t e = scope.enter(device);
	if (e != cudaSuccess) {
		return fail("rpt_bf_probe_u64_noalloc: %s failed: %s", "cudaSetDevice", e);
	}
	Scratch *s = nullptr;
	if (scratch_for(device, &s)) {
		return 1;
	}
	cudaStream_t st = static_cast<cudaStream_t>(stream);
	const uint32_t nblocks = static_cast<uint32_t>((count + kRowsPerBlock - 1) / kRowsPerBlock);
	if (nblocks) {
		noalloc_mark_kernel<<<nblocks, kBlockThreads, 0, st>>>(dev_keys, count, dev_words, num_sectors - 1, s);
	}
	noalloc_scan_kernel<<<1, kScanThreads, 0, st>>>(s, nblocks, dev_found_out); // nblocks == 0: *found = 0
	if (nblocks) {
		noalloc_compact_kernel<<<nblocks, kBlockThreads, 0, st>>>(s, dev_sel_out);
	}
	e = cudaGetLastError();
	if (e != cudaSuccess) {
		return fail("rpt_bf_probe_u64_noalloc: %s failed: %s", "kernel launch", e);
	}
	return 0;
}
