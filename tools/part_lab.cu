// part_lab.cu — development microbenchmark (not product): can EVERY large filter be probed as a shared-memory
// filter?  One-pass software-write-combining radix partition of the probe keys into buckets whose filter slice
// (128 KiB = 2^14 sectors) fits shared memory, then one shared-memory lookup kernel per bucket.
//
//   modes (argv[1]):
//     atoms   shared-memory atomicAdd rate on 2048 random counters (the partition's claim primitive)
//     part    partition kernel + lookup kernel, checked against a plain global-memory probe
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -o tools/part_lab tools/part_lab.cu
//   tools/part_lab part [log2_sectors=25] [block=1024] [kpt=8] [rows=1e9] [log2_chunk_rows=32] [density=1] [fake=0]
#include <cuda_runtime.h>

#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#define CK(x)                                                                                                          \
	do {                                                                                                               \
		cudaError_t e_ = (x);                                                                                          \
		if (e_ != cudaSuccess) {                                                                                       \
			printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__);                            \
			exit(1);                                                                                                   \
		}                                                                                                              \
	} while (0)

__host__ __device__ inline uint64_t murmur64(uint64_t x) {
	x ^= x >> 32;
	x *= 0xd6e8feb86659fd93ULL;
	x ^= x >> 32;
	x *= 0xd6e8feb86659fd93ULL;
	x ^= x >> 32;
	return x;
}
__host__ __device__ inline uint64_t splitmix64(uint64_t x) {
	uint64_t z = x + 0x9e3779b97f4a7c15ULL;
	z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ULL;
	z = (z ^ (z >> 27)) * 0x94d049bb133111ebULL;
	return z ^ (z >> 31);
}

__global__ void gen_keys(uint64_t *k, uint64_t n) {
	for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
		k[i] = splitmix64(i);
	}
}
__global__ void mask_keys(uint64_t *k, uint64_t n, uint64_t m) {
	for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
		k[i] &= m;
	}
}
// ~75 % of the bits set: a key passes with probability ≈ 0.75^4 ≈ 0.3
__global__ void fill_filter(uint64_t *f, uint64_t s, int density) {
	for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < s; i += (uint64_t)gridDim.x * blockDim.x) {
		const uint64_t a = splitmix64(i * 77 + 1), b = splitmix64(i * 77 + 2);
		f[i] = density == 0 ? (a | b) : (density == 1 ? a : (a & b)); // 75 / 50 / 25 % of the bits: 32 / 6 / 0.4 % hits
	}
}

constexpr int kSliceLog = 14; // sectors per bucket slice: 128 KiB
constexpr uint64_t kValid = 1ULL << 63;

__device__ __forceinline__ uint32_t test4(uint64_t w, uint32_t p0, uint32_t p1, uint32_t p2, uint32_t p3) {
	return (uint32_t)(w >> p0) & (uint32_t)(w >> p1) & (uint32_t)(w >> p2) & (uint32_t)(w >> p3) & 1u;
}

// reference: plain global lookups into a row-ordered bitmap
__global__ void __launch_bounds__(256) ref_probe(const uint64_t *__restrict__ keys, const uint64_t *__restrict__ filter,
                                                 uint64_t smask, uint64_t n, uint32_t *bitmap) {
	for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
		const uint64_t h = murmur64(keys[i]);
		const uint32_t hi = (uint32_t)(h >> 32);
		const uint64_t w = filter[h & smask];
		const uint32_t hit = test4(w, hi & 63, (hi >> 8) & 63, (hi >> 16) & 63, (hi >> 24) & 63);
		const uint32_t b = __ballot_sync(0xffffffffu, hit);
		if ((threadIdx.x & 31) == 0) {
			bitmap[i >> 5] = b;
		}
	}
}

__global__ void cmp_words(const uint32_t *a, const uint32_t *b, uint64_t n, unsigned long long *diff,
                          unsigned long long *bits) {
	unsigned long long d = 0, c = 0;
	for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
		d += a[i] != b[i];
		c += __popc(a[i]);
	}
	if (d) {
		atomicAdd(diff, d);
	}
	atomicAdd(bits, c);
}

// ---------------------------------------------------------------------------------------------------------
// atoms: every thread does `per_thread` atomicAdd on random counters of a 2048-entry shared array
// ---------------------------------------------------------------------------------------------------------
template <int MODE> // 0: atomicAdd with return, 1: plain ld+st (no atomicity; the smem op floor), 2: red (no return),
                    // 3: __match_any_sync on an 11-bit value, 4: 64-bit random LDS, 5: atomicAdd on 256 counters
__global__ void __launch_bounds__(1024) atoms_kernel(uint32_t per_thread, uint32_t *sink) {
	__shared__ uint64_t ctr64[2048];
	uint32_t *ctr = reinterpret_cast<uint32_t *>(ctr64);
	for (int i = threadIdx.x; i < 2048; i += blockDim.x) {
		ctr64[i] = 0;
	}
	__syncthreads();
	uint64_t x = splitmix64(blockIdx.x * 1024ull + threadIdx.x);
	uint32_t acc = 0;
	for (uint32_t k = 0; k < per_thread; k += 4) {
		x = x * 6364136223846793005ULL + 1442695040888963407ULL;
		const uint64_t r = x >> 16;
#pragma unroll
		for (int j = 0; j < 4; j++) {
			const uint32_t b = (uint32_t)(r >> (11 * j)) & 2047u;
			if (MODE == 0) {
				acc += atomicAdd(&ctr[b], 1u);
			} else if (MODE == 5) {
				acc += atomicAdd(&ctr[b & 255u], 1u);
			} else if (MODE == 1) {
				const uint32_t v = ((volatile uint32_t *)ctr)[b];
				((volatile uint32_t *)ctr)[b] = v + 1;
				acc += v;
			} else if (MODE == 2) {
				asm volatile("red.shared.add.u32 [%0], 1;" ::"r"((uint32_t)__cvta_generic_to_shared(&ctr[b])) : "memory");
			} else if (MODE == 3) {
				acc += __match_any_sync(0xffffffffu, b);
			} else {
				acc += (uint32_t)((volatile uint64_t *)ctr64)[b];
			}
		}
	}
	__syncthreads();
	if (acc == 0x12345678u || ctr[threadIdx.x] == 0xffffffffu) {
		sink[0] = acc;
	}
}

// ---------------------------------------------------------------------------------------------------------
// partition: one CTA per SM owns a contiguous row range; per bucket one 64-byte staging line (8 records) in
// shared memory; a key claims a slot with one shared atomicAdd, the claimer of slot 7 flushes the line to the
// (CTA, bucket) region in global memory and re-opens it.
//   record = valid:1 | row-in-CTA-range:25 | 4 × 6 bit positions | sector-in-slice:14
// ---------------------------------------------------------------------------------------------------------
struct PartArgs {
	const uint64_t *keys;
	uint64_t n;
	uint64_t rows_per_cta;
	uint32_t smask;
	uint32_t nb;        // buckets = (smask + 1) >> kSliceLog
	uint32_t cap_lines; // lines per (CTA, bucket) region
	uint64_t *records;  // [cta][bucket][cap_lines][8]
	uint32_t *counts;   // [cta][bucket] records
	unsigned long long *overflow;
	unsigned long long *retries;
};

__device__ __forceinline__ uint64_t make_record(uint64_t h, uint32_t rowrel) {
	const uint32_t hi = (uint32_t)(h >> 32) & 0x3f3f3f3fu;
	const uint32_t pos = (hi & 0x3fu) | ((hi >> 2) & 0xfc0u) | ((hi >> 4) & 0x3f000u) | ((hi >> 6) & 0xfc0000u);
	const uint32_t lo = ((uint32_t)h & ((1u << kSliceLog) - 1u)) | (pos << kSliceLog); // 14 + 18 of the 24 bits
	const uint32_t up = (pos >> 18) | (rowrel << 6) | 0x80000000u;
	return ((uint64_t)up << 32) | lo;
}

template <int BLOCK, int KPT, bool FAKE = false> // FAKE: no atomics (racy, wrong) — the cost of everything else
__global__ void __launch_bounds__(BLOCK, 1) part_kernel(const PartArgs a) {
	extern __shared__ __align__(128) unsigned char s_dyn[];
	uint64_t *stage = reinterpret_cast<uint64_t *>(s_dyn);                   // nb × 8 records
	uint32_t *ctr = reinterpret_cast<uint32_t *>(s_dyn + (size_t)a.nb * 64); // claims of the open line
	uint32_t *lines = ctr + a.nb;                                            // lines flushed so far
	const uint32_t lane = threadIdx.x & 31u;
	for (uint32_t i = threadIdx.x; i < a.nb * 8; i += BLOCK) {
		stage[i] = 0;
	}
	for (uint32_t i = threadIdx.x; i < a.nb; i += BLOCK) {
		ctr[i] = 0;
		lines[i] = 0;
	}
	__syncthreads();
	const uint64_t row_lo = (uint64_t)blockIdx.x * a.rows_per_cta;
	const uint64_t row_hi = row_lo + a.rows_per_cta < a.n ? row_lo + a.rows_per_cta : a.n;
	uint64_t *const my_records = a.records + (uint64_t)blockIdx.x * a.nb * a.cap_lines * 8;
	uint32_t n_retry = 0, n_over = 0, fake_ctr = threadIdx.x;
	const uint32_t bshift = kSliceLog;

	for (uint64_t base = row_lo; base < row_hi; base += (uint64_t)BLOCK * KPT) {
		// KPT keys per thread: KPT/2 coalesced 16-byte loads, all in flight together
		uint64_t rec[KPT];
		uint32_t bkt[KPT];
		uint32_t pending = 0;
		{
			ulonglong2 k[KPT / 2];
#pragma unroll
			for (int r = 0; r < KPT / 2; r++) {
				const uint64_t row = base + (uint64_t)r * BLOCK * 2 + threadIdx.x * 2;
				k[r] = make_ulonglong2(0, 0);
				if (row + 1 < row_hi) {
					asm volatile("ld.global.nc.L1::no_allocate.v2.u64 {%0, %1}, [%2];"
					             : "=l"(k[r].x), "=l"(k[r].y)
					             : "l"(a.keys + row));
					pending |= 3u << (2 * r);
				} else if (row < row_hi) {
					k[r].x = a.keys[row];
					pending |= 1u << (2 * r);
				}
			}
#pragma unroll
			for (int r = 0; r < KPT / 2; r++) {
				const uint32_t rowrel = (uint32_t)(base - row_lo) + r * BLOCK * 2 + threadIdx.x * 2;
				const uint64_t h0 = murmur64(k[r].x), h1 = murmur64(k[r].y);
				rec[2 * r] = make_record(h0, rowrel);
				rec[2 * r + 1] = make_record(h1, rowrel + 1);
				bkt[2 * r] = ((uint32_t)h0 & a.smask) >> bshift;
				bkt[2 * r + 1] = ((uint32_t)h1 & a.smask) >> bshift;
			}
		}
		// claim rounds: a key whose bucket's line is full (awaiting its flush) stays pending for the next round
		while (true) {
			uint32_t flush = 0;
#pragma unroll
			for (int i = 0; i < KPT; i++) {
				if ((pending >> i) & 1u) {
					uint32_t c;
					if (FAKE) {
						c = (fake_ctr++) & 7u;
					} else {
						c = atomicAdd(&ctr[bkt[i]], 1u);
					}
					if (c < 8u) {
						stage[bkt[i] * 8u + c] = rec[i];
						pending &= ~(1u << i);
						flush |= (c == 7u) ? (1u << i) : 0u;
					}
				}
			}
			__syncwarp();
			// flush the lines this lane completed
			while (flush) {
				const int i = __ffs(flush) - 1;
				flush &= flush - 1;
				uint32_t b = 0;
#pragma unroll
				for (int j = 0; j < KPT; j++) {
					b = (j == i) ? bkt[j] : b;
				}
				ulonglong2 *line = reinterpret_cast<ulonglong2 *>(stage + b * 8u);
				ulonglong2 v[4];
				bool ok;
				do { // the other seven claimers may still be between their atomicAdd and their store
					ok = true;
#pragma unroll
					for (int q = 0; q < 4; q++) {
						asm volatile("ld.volatile.shared.v2.u64 {%0, %1}, [%2];"
						             : "=l"(v[q].x), "=l"(v[q].y)
						             : "r"((uint32_t)__cvta_generic_to_shared(line + q))
						             : "memory");
						ok = ok && (v[q].x & v[q].y & kValid);
					}
				} while (!ok && !FAKE);
#pragma unroll
				for (int q = 0; q < 4; q++) {
					line[q] = make_ulonglong2(0, 0);
				}
				const uint32_t l = FAKE ? (lines[b] % a.cap_lines) : lines[b];
				lines[b] = l + 1;
				__threadfence_block(); // zeroed slots and line count before the line re-opens
				atomicExch(&ctr[b], 0u);
				if (l < a.cap_lines) {
					ulonglong2 *dst = reinterpret_cast<ulonglong2 *>(my_records + ((uint64_t)b * a.cap_lines + l) * 8);
#pragma unroll
					for (int q = 0; q < 4; q++) {
						asm volatile("st.global.cs.v2.u64 [%0], {%1, %2};" ::"l"(dst + q), "l"(v[q].x), "l"(v[q].y)
						             : "memory");
					}
				} else {
					n_over += 8;
				}
			}
			if (!__any_sync(0xffffffffu, pending != 0u)) {
				break;
			}
			n_retry += __popc(pending);
		}
	}
	__syncthreads();
	// partial lines
	for (uint32_t b = threadIdx.x; b < a.nb; b += BLOCK) {
		const uint32_t c = ctr[b] < 8u ? ctr[b] : 8u; // < 8 by construction
		uint32_t l = lines[b];
		if (l > a.cap_lines) {
			l = a.cap_lines;
		}
		uint32_t total = l * 8u;
		if (c && l < a.cap_lines) {
			uint64_t *dst = my_records + ((uint64_t)b * a.cap_lines + l) * 8;
			for (uint32_t j = 0; j < c; j++) {
				dst[j] = stage[b * 8u + j];
			}
			total += c;
		} else if (c) {
			n_over += c;
		}
		a.counts[(uint64_t)blockIdx.x * a.nb + b] = total;
	}
	if (n_over) {
		atomicAdd(a.overflow, (unsigned long long)n_over);
	}
	n_retry = __reduce_add_sync(0xffffffffu, n_retry);
	if (lane == 0 && n_retry) {
		atomicAdd(a.retries, (unsigned long long)n_retry);
	}
}

// ---------------------------------------------------------------------------------------------------------
// lookup: persistent CTAs; per bucket the 128 KiB slice arrives in shared memory through ONE bulk copy, then
// the warps stream the bucket's records region by region and OR hit bits into the row-ordered bitmap.
// ---------------------------------------------------------------------------------------------------------
struct LookArgs {
	const uint64_t *filter;
	uint32_t nb;
	uint32_t nparts; // partition CTAs
	uint32_t cap_lines;
	uint64_t rows_per_cta;
	const uint64_t *records;
	const uint32_t *counts;
	uint32_t *bitmap;
};

__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
	asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"((uint32_t)__cvta_generic_to_shared(bar)), "r"(count)
	             : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
	asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"((uint32_t)__cvta_generic_to_shared(bar)),
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
	             "}" ::"r"((uint32_t)__cvta_generic_to_shared(bar)),
	             "r"(parity)
	             : "memory");
}
__device__ __forceinline__ void bulk_g2s(void *smem_dst, const void *gmem_src, uint32_t bytes, uint64_t *bar) {
	asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
	                 (uint32_t)__cvta_generic_to_shared(smem_dst)),
	             "l"(gmem_src), "r"(bytes), "r"((uint32_t)__cvta_generic_to_shared(bar))
	             : "memory");
}

template <int BLOCK, int U> // U 16-byte loads (2 records each) in flight per lane
__global__ void __launch_bounds__(BLOCK, 1) lookup_kernel(const LookArgs a) {
	extern __shared__ __align__(128) unsigned char s_dyn[];
	uint64_t *slice = reinterpret_cast<uint64_t *>(s_dyn);
	__shared__ __align__(8) uint64_t bar;
	constexpr uint32_t kSliceBytes = (1u << kSliceLog) * 8u;
	const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
	constexpr uint32_t NWARPS = BLOCK / 32;
	if (threadIdx.x == 0) {
		mbar_init(&bar, 1);
		asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
	}
	__syncthreads();
	uint32_t parity = 0;
	for (uint32_t b = blockIdx.x; b < a.nb; b += gridDim.x) {
		if (threadIdx.x == 0) {
			mbar_expect_tx(&bar, kSliceBytes);
			bulk_g2s(slice, a.filter + ((uint64_t)b << kSliceLog), kSliceBytes, &bar);
		}
		mbar_wait(&bar, parity);
		parity ^= 1u;
		for (uint32_t r = warp; r < a.nparts; r += NWARPS) {
			const uint32_t cnt = a.counts[(uint64_t)r * a.nb + b];
			const ulonglong2 *src =
			    reinterpret_cast<const ulonglong2 *>(a.records + ((uint64_t)r * a.nb + b) * a.cap_lines * 8);
			const uint64_t row_base = (uint64_t)r * a.rows_per_cta;
			const uint32_t npairs = (cnt + 1) >> 1;
			for (uint32_t j0 = 0; j0 < npairs; j0 += 32 * U) {
				ulonglong2 v[U];
#pragma unroll
				for (int u = 0; u < U; u++) {
					const uint32_t j = j0 + u * 32 + lane;
					v[u] = make_ulonglong2(0, 0);
					if (j < npairs) {
						asm volatile("ld.global.nc.L1::no_allocate.v2.u64 {%0, %1}, [%2];"
						             : "=l"(v[u].x), "=l"(v[u].y)
						             : "l"(src + j));
						if (2 * j + 1 >= cnt) {
							v[u].y = 0;
						}
					}
				}
				uint64_t w[2 * U];
#pragma unroll
				for (int u = 0; u < U; u++) {
					w[2 * u] = slice[(uint32_t)v[u].x & ((1u << kSliceLog) - 1u)];
					w[2 * u + 1] = slice[(uint32_t)v[u].y & ((1u << kSliceLog) - 1u)];
				}
#pragma unroll
				for (int u = 0; u < 2 * U; u++) {
					const uint64_t rec = (u & 1) ? v[u >> 1].y : v[u >> 1].x;
					const uint32_t lo = (uint32_t)rec, up = (uint32_t)(rec >> 32);
					const uint32_t hit = test4(w[u], (lo >> 14) & 63u, (lo >> 20) & 63u, lo >> 26, up & 63u) & (up >> 31);
					if (hit) {
						const uint64_t row = row_base + ((up >> 6) & 0x1ffffffu);
						atomicOr(a.bitmap + (row >> 5), 1u << (row & 31u));
					}
				}
			}
		}
		__syncthreads(); // every warp is done with the slice before the next bulk copy overwrites it
	}
}


// =========================================================================================================
// v2 — what the ncu capture of v1 said (profiles/part_lab_r2.md): the LSU data pipe is the limit (1.74 wavefronts
// per key, 88 % busy), 1.5 of them spent by the per-lane line flush (4 LDS.128 + 4 STS.128 + 4 STG.128 per line,
// one wavefront per active lane each), and 110 thread-instructions per key.  Shared atomics are NOT the limit
// (9 per clock and SM on 2048 random counters).  v2: the flush is done by groups of SLOTS lanes (one lane per
// record, contiguous LDS.64 / STG.64), right after the key step that completed the line; the bucket word packs
// {claims:16 | line:16}, so the claim that completes a line also learns where the line goes.
//   record = valid:1 | row-in-CTA-range:19 | hash bits 32..61 & 0x3f3f3f3f : 30 | sector-in-slice:14
// =========================================================================================================
struct Part2Args {
	const uint64_t *keys;
	uint64_t n;
	uint64_t rows_per_cta;
	uint32_t smask;
	uint32_t nb;
	uint32_t cap_lines;
	uint64_t *records; // [cta][bucket][cap_lines][SLOTS]
	uint32_t *counts;  // [cta][bucket] records
	const uint64_t *filter;
	uint32_t *bitmap;
	unsigned long long *stats; // [0] records that overflowed their region, [1] failed claims, [2] direct lookups
};

__device__ __forceinline__ uint64_t make_record2(uint64_t h, uint32_t rowrel) {
	const uint32_t hm = (uint32_t)(h >> 32) & 0x3f3f3f3fu;
	const uint32_t lo = ((uint32_t)h & ((1u << kSliceLog) - 1u)) + hm * (1u << kSliceLog);
	const uint32_t up = (hm >> 18) + rowrel * 4096u + 0x80000000u;
	return ((uint64_t)up << 32) | lo;
}
__device__ __forceinline__ uint32_t record2_hit(uint64_t w, uint32_t lo, uint32_t up) {
	return test4(w, (lo >> 14) & 63u, (lo >> 22) & 63u, __funnelshift_r(lo, up, 30) & 63u, (up >> 6) & 63u);
}
__device__ __forceinline__ void direct_lookup2(const Part2Args &a, uint32_t bucket, uint64_t rec, uint64_t row_lo) {
	const uint32_t lo = (uint32_t)rec, up = (uint32_t)(rec >> 32);
	const uint64_t w = a.filter[((uint64_t)bucket << kSliceLog) | (lo & ((1u << kSliceLog) - 1u))];
	if (record2_hit(w, lo, up)) {
		const uint64_t row = row_lo + ((up >> 12) & 0x7ffffu);
		atomicOr(a.bitmap + (row >> 5), 1u << (row & 31u));
	}
}

// forward declarations (defined with the lookup kernel below)
__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count);
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes);
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity);
__device__ __forceinline__ void bulk_g2s(void *smem_dst, const void *gmem_src, uint32_t bytes, uint64_t *bar);

// KPT = key steps per staged buffer: a warp's keys arrive in warp-private buffers of 32·KPT keys through one bulk
// copy each (double buffered, one mbarrier per buffer), so the loop body — one key per lane — exists once.
template <int BLOCK, int KPT, int SLOTS>
__global__ void __launch_bounds__(BLOCK, 1) part2_kernel(const Part2Args a) {
	constexpr int NWARPS = BLOCK / 32;
	constexpr int GROUPS = 32 / SLOTS; // lines flushed per iteration
	constexpr uint32_t SEG = 32u * KPT; // keys per buffer
	extern __shared__ __align__(128) unsigned char s_dyn[];
	uint64_t *stage = reinterpret_cast<uint64_t *>(s_dyn);                          // nb × SLOTS records
	uint32_t *ctr = reinterpret_cast<uint32_t *>(s_dyn + (size_t)a.nb * SLOTS * 8); // {claims:16 | line:16}
	const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
	uint32_t *flist = ctr + a.nb + warp * 32;                                       // warp-private
	uint64_t *kbuf = reinterpret_cast<uint64_t *>(ctr + a.nb + NWARPS * 32) + warp * 2 * SEG;
	__shared__ __align__(8) uint64_t bars[NWARPS][2];
	const uint32_t grp = lane / SLOTS, sub = lane % SLOTS;
	uint32_t lt;
	asm("mov.u32 %0, %%lanemask_lt;" : "=r"(lt));
	for (uint32_t i = threadIdx.x; i < a.nb * SLOTS; i += BLOCK) {
		stage[i] = 0;
	}
	for (uint32_t i = threadIdx.x; i < a.nb; i += BLOCK) {
		ctr[i] = 0;
	}
	if (lane == 0) {
		mbar_init(&bars[warp][0], 1);
		mbar_init(&bars[warp][1], 1);
		asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
	}
	__syncthreads();
	const uint64_t row_lo = (uint64_t)blockIdx.x * a.rows_per_cta;
	const uint64_t row_hi = row_lo + a.rows_per_cta < a.n ? row_lo + a.rows_per_cta : a.n;
	uint64_t *const my_records = a.records + (uint64_t)blockIdx.x * a.nb * a.cap_lines * SLOTS;
	uint32_t n_fail = 0, n_over = 0, n_direct = 0;
	const uint64_t nrows = row_hi > row_lo ? row_hi - row_lo : 0;
	const uint32_t nseg = (uint32_t)((nrows + SEG - 1) / SEG);
	const uint32_t full_segs = (uint32_t)(nrows / SEG);

	auto issue = [&](uint32_t seg, uint32_t buf) { // lane 0: stage segment `seg` into buffer `buf`
		if (seg < full_segs) {
			mbar_expect_tx(&bars[warp][buf], SEG * 8);
			bulk_g2s(kbuf + buf * SEG, a.keys + row_lo + (uint64_t)seg * SEG, SEG * 8, &bars[warp][buf]);
		}
	};
	uint32_t phase = 0; // bit b = parity to wait for on buffer b
	if (lane == 0) {
		issue(warp, 0);
	}
	uint32_t buf = 0;
	for (uint32_t seg = warp; seg < nseg; seg += NWARPS, buf ^= 1u) {
		if (lane == 0) {
			issue(seg + NWARPS, buf ^ 1u);
		}
		const bool staged = seg < full_segs;
		if (staged) {
			mbar_wait(&bars[warp][buf], (phase >> buf) & 1u);
			phase ^= 1u << buf;
		}
#pragma unroll 1
		for (uint32_t st = 0; st < (uint32_t)KPT; st++) {
			const uint32_t rowrel = seg * SEG + st * 32u + lane;
			bool act = rowrel < nrows;
			uint64_t k = 0;
			if (staged) {
				k = kbuf[buf * SEG + st * 32u + lane];
			} else if (act) {
				k = a.keys[row_lo + rowrel];
			}
			const uint64_t h = murmur64(k);
			const uint64_t rec = make_record2(h, rowrel);
			const uint32_t b = ((uint32_t)h & a.smask) >> kSliceLog;
			uint32_t tries = 0;
			while (true) {
				// claim, store, flush the lines this step completed
				uint32_t old = 0xffffu;
				if (act) {
					old = atomicAdd(&ctr[b], 1u);
				}
				const uint32_t c = old & 0xffffu;
				if (c < (uint32_t)SLOTS) {
					stage[b * SLOTS + c] = rec;
				}
				const bool completes = c == (uint32_t)SLOTS - 1u;
				const uint32_t fl = __ballot_sync(0xffffffffu, completes);
				if (fl) { // warp-uniform
					if (completes) {
						flist[__popc(fl & lt)] = b | ((old >> 16) << 12);
					}
					__syncwarp();
					const uint32_t nfl = __popc(fl);
					for (uint32_t it = 0; it < nfl; it += GROUPS) {
						const uint32_t g = it + grp;
						if (g < nfl) {
							const uint32_t e = flist[g];
							const uint32_t fb = e & 4095u, line = e >> 12;
							const uint32_t saddr = (uint32_t)__cvta_generic_to_shared(stage + fb * SLOTS + sub);
							uint64_t v;
							do { // the claimer of this slot may still be between its atomicAdd and its store
								asm volatile("ld.volatile.shared.u64 %0, [%1];" : "=l"(v) : "r"(saddr) : "memory");
							} while (!(v >> 63));
							asm volatile("st.volatile.shared.u64 [%0], %1;" ::"r"(saddr), "l"(0ULL) : "memory");
							__syncwarp(SLOTS == 32 ? 0xffffffffu : (((1u << SLOTS) - 1u) << (grp * SLOTS)));
							if (sub == 0) {
								atomicExch(&ctr[fb], (line + 1u) << 16); // re-opens the line
							}
							if (line < a.cap_lines) {
								uint64_t *dst = my_records + ((uint64_t)fb * a.cap_lines + line) * SLOTS + sub;
								asm volatile("st.global.cs.u64 [%0], %1;" ::"l"(dst), "l"(v) : "memory");
							} else { // skew: the (CTA, bucket) region is full — look the record up on the spot
								direct_lookup2(a, fb, v, row_lo);
								n_over++;
							}
						}
					}
					__syncwarp(); // flist is rewritten by the next step
				}
				act = act && c >= (uint32_t)SLOTS; // failed: the line was full, awaiting its flush
				if (!__any_sync(0xffffffffu, act)) {
					break;
				}
				n_fail += act ? 1u : 0u;
				if (++tries == 3) { // hot bucket = hot sectors: look up on the spot
					if (act) {
						direct_lookup2(a, b, rec, row_lo);
						n_direct++;
					}
					break;
				}
			}
		}
		__syncwarp(); // every lane is done with this buffer before lane 0 refills it (two iterations on)
	}
	__syncthreads();
	// partial lines
	for (uint32_t b = threadIdx.x; b < a.nb; b += BLOCK) {
		const uint32_t c = ctr[b] & 0xffffu; // < SLOTS: every full line has been flushed and re-opened
		uint32_t l = ctr[b] >> 16;
		uint32_t total = (l < a.cap_lines ? l : a.cap_lines) * SLOTS;
		for (uint32_t j = 0; j < c && j < (uint32_t)SLOTS; j++) {
			const uint64_t v = stage[b * SLOTS + j];
			if (l < a.cap_lines) {
				my_records[((uint64_t)b * a.cap_lines + l) * SLOTS + j] = v;
				total++;
			} else {
				direct_lookup2(a, b, v, row_lo);
				n_over++;
			}
		}
		a.counts[(uint64_t)blockIdx.x * a.nb + b] = total;
	}
	n_fail = __reduce_add_sync(0xffffffffu, n_fail);
	n_over = __reduce_add_sync(0xffffffffu, n_over);
	n_direct = __reduce_add_sync(0xffffffffu, n_direct);
	if (lane == 0) {
		if (n_over) {
			atomicAdd(a.stats, (unsigned long long)n_over);
		}
		if (n_fail) {
			atomicAdd(a.stats + 1, (unsigned long long)n_fail);
		}
		if (n_direct) {
			atomicAdd(a.stats + 2, (unsigned long long)n_direct);
		}
	}
}

template <int BLOCK, int U, int SLOTS>
__global__ void __launch_bounds__(BLOCK, 1) lookup2_kernel(const LookArgs a) {
	extern __shared__ __align__(128) unsigned char s_dyn[];
	uint64_t *slice = reinterpret_cast<uint64_t *>(s_dyn);
	__shared__ __align__(8) uint64_t bar;
	constexpr uint32_t kSliceBytes = (1u << kSliceLog) * 8u;
	const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
	constexpr uint32_t NWARPS = BLOCK / 32;
	if (threadIdx.x == 0) {
		mbar_init(&bar, 1);
		asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
	}
	__syncthreads();
	uint32_t parity = 0;
	for (uint32_t b = blockIdx.x; b < a.nb; b += gridDim.x) {
		if (threadIdx.x == 0) {
			mbar_expect_tx(&bar, kSliceBytes);
			bulk_g2s(slice, a.filter + ((uint64_t)b << kSliceLog), kSliceBytes, &bar);
		}
		mbar_wait(&bar, parity);
		parity ^= 1u;
		for (uint32_t r = warp; r < a.nparts; r += NWARPS) {
			const uint32_t cnt = a.counts[(uint64_t)r * a.nb + b];
			const ulonglong2 *src =
			    reinterpret_cast<const ulonglong2 *>(a.records + ((uint64_t)r * a.nb + b) * a.cap_lines * SLOTS);
			const uint64_t row_base = (uint64_t)r * a.rows_per_cta;
			const uint32_t npairs = (cnt + 1) >> 1;
			for (uint32_t j0 = 0; j0 < npairs; j0 += 32 * U) {
				ulonglong2 v[U];
#pragma unroll
				for (int u = 0; u < U; u++) {
					const uint32_t j = j0 + u * 32 + lane;
					v[u] = make_ulonglong2(0, 0);
					if (j < npairs) {
						asm volatile("ld.global.nc.L1::no_allocate.v2.u64 {%0, %1}, [%2];"
						             : "=l"(v[u].x), "=l"(v[u].y)
						             : "l"(src + j));
						if (2 * j + 1 >= cnt) {
							v[u].y = 0;
						}
					}
				}
				uint64_t w[2 * U];
#pragma unroll
				for (int u = 0; u < U; u++) {
					w[2 * u] = slice[(uint32_t)v[u].x & ((1u << kSliceLog) - 1u)];
					w[2 * u + 1] = slice[(uint32_t)v[u].y & ((1u << kSliceLog) - 1u)];
				}
#pragma unroll
				for (int u = 0; u < 2 * U; u++) {
					const uint64_t rec = (u & 1) ? v[u >> 1].y : v[u >> 1].x;
					const uint32_t lo = (uint32_t)rec, up = (uint32_t)(rec >> 32);
					if (record2_hit(w[u], lo, up) & (up >> 31)) {
						const uint64_t row = row_base + ((up >> 12) & 0x7ffffu);
						atomicOr(a.bitmap + (row >> 5), 1u << (row & 31u));
					}
				}
			}
		}
		__syncthreads();
	}
}

static uint64_t rows_per_cta_for(uint64_t n);
struct Part2Run {
	Part2Args pa;
	LookArgs la;
	int block, kpt, slots;
	size_t part_smem;
	uint64_t n_total, chunk_rows;
	int what;
};
template <int BLOCK, int KPT, int SLOTS>
static void launch_part2(const Part2Run *r, const Part2Args &pa, const LookArgs &la, uint64_t cn) {
	if (r->what != 2) {
		CK(cudaFuncSetAttribute(part2_kernel<BLOCK, KPT, SLOTS>, cudaFuncAttributeMaxDynamicSharedMemorySize,
		                        (int)r->part_smem));
		part2_kernel<BLOCK, KPT, SLOTS><<<148, BLOCK, r->part_smem>>>(pa);
	}
	if (r->what != 1) {
		const size_t smem = (1u << kSliceLog) * 8;
		CK(cudaFuncSetAttribute(lookup2_kernel<1024, 4, SLOTS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
		lookup2_kernel<1024, 4, SLOTS><<<148, 1024, smem>>>(la);
	}
}
static void run_part2(void *p) {
	Part2Run *r = (Part2Run *)p;
	for (uint64_t c0 = 0; c0 < r->n_total; c0 += r->chunk_rows) {
		const uint64_t cn = c0 + r->chunk_rows < r->n_total ? r->chunk_rows : r->n_total - c0;
		Part2Args pa = r->pa;
		LookArgs la = r->la;
		pa.keys += c0;
		pa.n = cn;
		pa.rows_per_cta = la.rows_per_cta = rows_per_cta_for(cn);
		pa.bitmap = la.bitmap = r->la.bitmap + c0 / 32;
		// the partition's direct lookups OR into the bitmap too: zero it first
		CK(cudaMemsetAsync(la.bitmap, 0, (cn + 31) / 32 * 4));
		const int key = r->block * 1000 + r->kpt * 10 + (r->slots == 8 ? 0 : (r->slots == 16 ? 1 : 2));
		switch (key) {
		case 1024 * 1000 + 40: launch_part2<1024, 4, 8>(r, pa, la, cn); break;
		case 1024 * 1000 + 20: launch_part2<1024, 2, 8>(r, pa, la, cn); break;
		case 512 * 1000 + 80: launch_part2<512, 8, 8>(r, pa, la, cn); break;
		case 768 * 1000 + 40: launch_part2<768, 4, 8>(r, pa, la, cn); break;
		case 1024 * 1000 + 41: launch_part2<1024, 4, 16>(r, pa, la, cn); break;
		case 1024 * 1000 + 42: launch_part2<1024, 4, 4>(r, pa, la, cn); break;
		default: printf("no such part2 variant\n"); exit(1);
		}
	}
}

// =========================================================================================================
// v3 — v2's ncu capture: 203 thread-instructions per key, issue slots 84 % busy: per-step bookkeeping (flush
// loop, retry loop, one line group per iteration) dominated.  v3 works in phases of KPP keys per lane: all claims
// of a phase first, then ONE flush loop over every line the phase completed (16 bytes per lane: SLOTS/2 lanes per
// line), then a retry round for the claims that found their line full.
//   record = row-in-CTA-range:20 | hm:30 with valid in hm's spare bit 6 | sector-in-slice:14
//   (hm = hash bits 32..61 & 0x3f3f3f3f: bits 6,7,14,15,22,23 of it are always zero)
// =========================================================================================================
__device__ __forceinline__ uint64_t make_record3(uint64_t h, uint32_t rowrel) {
	const uint32_t hm = (uint32_t)(h >> 32) & 0x3f3f3f3fu;
	const uint32_t lo = (((uint32_t)h & ((1u << kSliceLog) - 1u)) | (1u << 20)) + hm * (1u << kSliceLog);
	const uint32_t up = (hm >> 18) + rowrel * 4096u;
	return ((uint64_t)up << 32) | lo;
}
constexpr uint32_t kValid3 = 1u << 20;
__device__ __forceinline__ void direct_lookup3(const Part2Args &a, uint32_t bucket, uint64_t rec, uint64_t row_lo) {
	const uint32_t lo = (uint32_t)rec, up = (uint32_t)(rec >> 32);
	const uint64_t w = a.filter[((uint64_t)bucket << kSliceLog) | (lo & ((1u << kSliceLog) - 1u))];
	if (record2_hit(w, lo, up)) {
		const uint64_t row = row_lo + (up >> 12);
		atomicOr(a.bitmap + (row >> 5), 1u << (row & 31u));
	}
}

template <int BLOCK, int KPP, int SLOTS>
__global__ void __launch_bounds__(BLOCK, 1) part3_kernel(const Part2Args a) {
	constexpr int NWARPS = BLOCK / 32;
	constexpr int LPL = SLOTS / 2;     // lanes per line (16 bytes each)
	constexpr int GROUPS = 32 / LPL;   // lines flushed per iteration
	constexpr uint32_t SEG = 32u * KPP; // keys per phase and warp
	extern __shared__ __align__(128) unsigned char s_dyn[];
	uint64_t *stage = reinterpret_cast<uint64_t *>(s_dyn);                          // nb × SLOTS records
	uint32_t *ctr = reinterpret_cast<uint32_t *>(s_dyn + (size_t)a.nb * SLOTS * 8); // {claims:16 | line:16}
	const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
	uint32_t *flist = ctr + a.nb + warp * SEG;                                      // warp-private, ≤ SEG lines per phase
	uint64_t *kbuf = reinterpret_cast<uint64_t *>(ctr + a.nb + NWARPS * SEG) + warp * 2 * SEG;
	__shared__ __align__(8) uint64_t bars[NWARPS][2];
	const uint32_t grp = lane / LPL, sub = lane % LPL;
	uint32_t lt;
	asm("mov.u32 %0, %%lanemask_lt;" : "=r"(lt));
	for (uint32_t i = threadIdx.x; i < a.nb * SLOTS; i += BLOCK) {
		stage[i] = 0;
	}
	for (uint32_t i = threadIdx.x; i < a.nb; i += BLOCK) {
		ctr[i] = 0;
	}
	if (lane == 0) {
		mbar_init(&bars[warp][0], 1);
		mbar_init(&bars[warp][1], 1);
		asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
	}
	__syncthreads();
	const uint64_t row_lo = (uint64_t)blockIdx.x * a.rows_per_cta;
	const uint64_t row_hi = row_lo + a.rows_per_cta < a.n ? row_lo + a.rows_per_cta : a.n;
	ulonglong2 *const my_records = reinterpret_cast<ulonglong2 *>(a.records + (uint64_t)blockIdx.x * a.nb * a.cap_lines * SLOTS);
	uint32_t n_fail = 0, n_over = 0, n_direct = 0;
	const uint64_t nrows = row_hi > row_lo ? row_hi - row_lo : 0;
	const uint32_t nseg = (uint32_t)((nrows + SEG - 1) / SEG);
	const uint32_t full_segs = (uint32_t)(nrows / SEG);

	auto issue = [&](uint32_t seg, uint32_t buf) { // lane 0: stage segment `seg` into buffer `buf`
		if (seg < full_segs) {
			mbar_expect_tx(&bars[warp][buf], SEG * 8);
			bulk_g2s(kbuf + buf * SEG, a.keys + row_lo + (uint64_t)seg * SEG, SEG * 8, &bars[warp][buf]);
		}
	};
	uint32_t phase = 0;
	if (lane == 0) {
		issue(warp, 0);
	}
	uint32_t buf = 0;
	for (uint32_t seg = warp; seg < nseg; seg += NWARPS, buf ^= 1u) {
		if (lane == 0) {
			issue(seg + NWARPS, buf ^ 1u);
		}
		const bool staged = seg < full_segs;
		if (staged) {
			mbar_wait(&bars[warp][buf], (phase >> buf) & 1u);
			phase ^= 1u << buf;
		}
		uint64_t rec[KPP];
		uint32_t bkt[KPP];
		uint32_t pend = 0;
#pragma unroll
		for (int i = 0; i < KPP; i++) {
			const uint32_t rowrel = seg * SEG + i * 32u + lane;
			const bool act = rowrel < nrows;
			uint64_t k = 0;
			if (staged) {
				k = kbuf[buf * SEG + i * 32u + lane];
			} else if (act) {
				k = a.keys[row_lo + rowrel];
			}
			const uint64_t h = murmur64(k);
			rec[i] = make_record3(h, rowrel);
			bkt[i] = ((uint32_t)h & a.smask) >> kSliceLog;
			pend |= act ? (1u << i) : 0u;
		}
#pragma unroll 1
		for (int tries = 0;; tries++) {
			// claims of this round; flist collects the lines they complete
			uint32_t nfl = 0;
#pragma unroll
			for (int i = 0; i < KPP; i++) {
				const bool want = (pend >> i) & 1u;
				if (tries == 0 || __any_sync(0xffffffffu, want)) {
					uint32_t old = 0xffffu;
					if (want) {
						old = atomicAdd(&ctr[bkt[i]], 1u);
					}
					const uint32_t c = old & 0xffffu;
					if (c < (uint32_t)SLOTS) {
						stage[bkt[i] * SLOTS + c] = rec[i];
						pend &= ~(1u << i);
					}
					const bool completes = c == (uint32_t)SLOTS - 1u;
					const uint32_t fl = __ballot_sync(0xffffffffu, completes);
					if (completes) {
						flist[nfl + __popc(fl & lt)] = bkt[i] | ((old >> 16) << 12);
					}
					nfl += __popc(fl);
				}
			}
			__syncwarp();
			for (uint32_t it = 0; it < nfl; it += GROUPS) {
				const uint32_t g = it + grp;
				if (g < nfl) {
					const uint32_t e = flist[g];
					const uint32_t fb = e & 4095u, line = e >> 12;
					const uint32_t saddr = (uint32_t)__cvta_generic_to_shared(stage + fb * SLOTS + sub * 2);
					ulonglong2 v;
					do { // a claimer of this line may still be between its atomicAdd and its store
						asm volatile("ld.volatile.shared.v2.u64 {%0, %1}, [%2];" : "=l"(v.x), "=l"(v.y) : "r"(saddr) : "memory");
					} while (!((uint32_t)v.x & (uint32_t)v.y & kValid3));
					asm volatile("st.volatile.shared.v2.u64 [%0], {%1, %1};" ::"r"(saddr), "l"(0ULL) : "memory");
					__syncwarp(LPL == 32 ? 0xffffffffu : (((1u << LPL) - 1u) << (grp * LPL)));
					if (sub == 0) {
						atomicExch(&ctr[fb], (line + 1u) << 16); // re-opens the line
					}
					if (line < a.cap_lines) {
						ulonglong2 *dst = my_records + ((uint64_t)fb * a.cap_lines + line) * LPL + sub;
						asm volatile("st.global.cs.v2.u64 [%0], {%1, %2};" ::"l"(dst), "l"(v.x), "l"(v.y) : "memory");
					} else { // skew: the (CTA, bucket) region is full — look the records up on the spot
						direct_lookup3(a, fb, v.x, row_lo);
						direct_lookup3(a, fb, v.y, row_lo);
						n_over += 2;
					}
				}
			}
			__syncwarp(); // flist is rewritten by the next round
			if (!__any_sync(0xffffffffu, pend != 0u)) {
				break;
			}
			n_fail += __popc(pend);
			if (tries == 2) { // hot bucket = hot sectors: look up on the spot
#pragma unroll
				for (int i = 0; i < KPP; i++) {
					if ((pend >> i) & 1u) {
						direct_lookup3(a, bkt[i], rec[i], row_lo);
						n_direct++;
					}
				}
				break;
			}
		}
	}
	__syncthreads();
	// partial lines
	for (uint32_t b = threadIdx.x; b < a.nb; b += BLOCK) {
		const uint32_t c = ctr[b] & 0xffffu; // < SLOTS: every full line has been flushed and re-opened
		uint32_t l = ctr[b] >> 16;
		uint32_t total = (l < a.cap_lines ? l : a.cap_lines) * SLOTS;
		for (uint32_t j = 0; j < c && j < (uint32_t)SLOTS; j++) {
			const uint64_t v = stage[b * SLOTS + j];
			if (l < a.cap_lines) {
				a.records[((uint64_t)blockIdx.x * a.nb + b) * a.cap_lines * SLOTS + (uint64_t)l * SLOTS + j] = v;
				total++;
			} else {
				direct_lookup3(a, b, v, row_lo);
				n_over++;
			}
		}
		a.counts[(uint64_t)blockIdx.x * a.nb + b] = total;
	}
	n_fail = __reduce_add_sync(0xffffffffu, n_fail);
	n_over = __reduce_add_sync(0xffffffffu, n_over);
	n_direct = __reduce_add_sync(0xffffffffu, n_direct);
	if (lane == 0) {
		if (n_over) {
			atomicAdd(a.stats, (unsigned long long)n_over);
		}
		if (n_fail) {
			atomicAdd(a.stats + 1, (unsigned long long)n_fail);
		}
		if (n_direct) {
			atomicAdd(a.stats + 2, (unsigned long long)n_direct);
		}
	}
}

// lookup v3: a warp's work items = (region, round of 32·U record pairs); the loads of item i+1 are issued before
// item i is looked up, and the first item's loads before the slice has arrived.
template <int BLOCK, int U, int SLOTS>
__global__ void __launch_bounds__(BLOCK, 1) lookup3_kernel(const LookArgs a) {
	extern __shared__ __align__(128) unsigned char s_dyn[];
	uint64_t *slice = reinterpret_cast<uint64_t *>(s_dyn);
	__shared__ __align__(8) uint64_t bar;
	constexpr uint32_t kSliceBytes = (1u << kSliceLog) * 8u;
	const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
	constexpr uint32_t NWARPS = BLOCK / 32;
	if (threadIdx.x == 0) {
		mbar_init(&bar, 1);
		asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
	}
	__syncthreads();
	uint32_t parity = 0;
	for (uint32_t b = blockIdx.x; b < a.nb; b += gridDim.x) {
		if (threadIdx.x == 0) {
			mbar_expect_tx(&bar, kSliceBytes);
			bulk_g2s(slice, a.filter + ((uint64_t)b << kSliceLog), kSliceBytes, &bar);
		}
		// this warp's regions: warp, warp + NWARPS, ...; lane k holds the count of the k-th one
		const uint32_t my_r = warp + lane * NWARPS;
		const uint32_t my_cnt = my_r < a.nparts ? a.counts[(uint64_t)my_r * a.nb + b] : 0u;
		uint32_t k = 0, j0 = 0; // next item to load
		uint32_t cnt_k = __shfl_sync(0xffffffffu, my_cnt, 0);
		auto load_item = [&](ulonglong2 (&v)[U], uint32_t &r_out) -> bool {
			while (k < 32u && warp + k * NWARPS < a.nparts && j0 >= ((cnt_k + 1u) >> 1)) { // skip finished / empty regions
				k++;
				j0 = 0;
				cnt_k = __shfl_sync(0xffffffffu, my_cnt, k & 31u);
			}
			const uint32_t r = warp + k * NWARPS;
			if (k >= 32u || r >= a.nparts) {
				return false;
			}
			const ulonglong2 *src =
			    reinterpret_cast<const ulonglong2 *>(a.records + ((uint64_t)r * a.nb + b) * a.cap_lines * SLOTS);
			const uint32_t npairs = (cnt_k + 1u) >> 1;
#pragma unroll
			for (int u = 0; u < U; u++) {
				const uint32_t j = j0 + u * 32 + lane;
				v[u] = make_ulonglong2(0, 0);
				if (j < npairs) {
					asm volatile("ld.global.nc.L1::no_allocate.v2.u64 {%0, %1}, [%2];"
					             : "=l"(v[u].x), "=l"(v[u].y)
					             : "l"(src + j));
					if (2 * j + 1 >= cnt_k) {
						v[u].y = 0;
					}
				}
			}
			r_out = r;
			j0 += 32 * U;
			return true;
		};
		ulonglong2 cur[U], nxt[U];
		uint32_t r_cur = 0, r_nxt = 0;
		bool have = load_item(cur, r_cur);
		mbar_wait(&bar, parity);
		parity ^= 1u;
		while (have) {
			const bool have_nxt = load_item(nxt, r_nxt);
			const uint64_t row_base = (uint64_t)r_cur * a.rows_per_cta;
			uint64_t w[2 * U];
#pragma unroll
			for (int u = 0; u < U; u++) {
				w[2 * u] = slice[(uint32_t)cur[u].x & ((1u << kSliceLog) - 1u)];
				w[2 * u + 1] = slice[(uint32_t)cur[u].y & ((1u << kSliceLog) - 1u)];
			}
#pragma unroll
			for (int u = 0; u < 2 * U; u++) {
				const uint64_t rec = (u & 1) ? cur[u >> 1].y : cur[u >> 1].x;
				const uint32_t lo = (uint32_t)rec, up = (uint32_t)(rec >> 32);
				if (record2_hit(w[u], lo, up) && (lo & kValid3)) {
					const uint64_t row = row_base + (up >> 12);
					atomicOr(a.bitmap + (row >> 5), 1u << (row & 31u));
				}
			}
#pragma unroll
			for (int u = 0; u < U; u++) {
				cur[u] = nxt[u];
			}
			r_cur = r_nxt;
			have = have_nxt;
		}
		__syncthreads();
	}
}

struct Part3Run {
	Part2Args pa;
	LookArgs la;
	int block, kpp, slots;
	size_t part_smem;
	uint64_t n_total, chunk_rows;
	int what;
};
template <int BLOCK, int KPP, int SLOTS>
static void launch_part3(const Part3Run *r, const Part2Args &pa, const LookArgs &la) {
	if (r->what != 2) {
		CK(cudaFuncSetAttribute(part3_kernel<BLOCK, KPP, SLOTS>, cudaFuncAttributeMaxDynamicSharedMemorySize,
		                        (int)r->part_smem));
		part3_kernel<BLOCK, KPP, SLOTS><<<148, BLOCK, r->part_smem>>>(pa);
	}
	if (r->what != 1) {
		const size_t smem = (1u << kSliceLog) * 8;
		CK(cudaFuncSetAttribute(lookup3_kernel<1024, 4, SLOTS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
		lookup3_kernel<1024, 4, SLOTS><<<148, 1024, smem>>>(la);
	}
}
static void run_part3(void *p) {
	Part3Run *r = (Part3Run *)p;
	for (uint64_t c0 = 0; c0 < r->n_total; c0 += r->chunk_rows) {
		const uint64_t cn = c0 + r->chunk_rows < r->n_total ? r->chunk_rows : r->n_total - c0;
		Part2Args pa = r->pa;
		LookArgs la = r->la;
		pa.keys += c0;
		pa.n = cn;
		pa.rows_per_cta = la.rows_per_cta = rows_per_cta_for(cn);
		pa.bitmap = la.bitmap = r->la.bitmap + c0 / 32;
		CK(cudaMemsetAsync(la.bitmap, 0, (cn + 31) / 32 * 4));
		const int key = r->block * 1000 + r->kpp * 10 + (r->slots == 8 ? 0 : 1);
		switch (key) {
		case 1024 * 1000 + 40: launch_part3<1024, 4, 8>(r, pa, la); break;
		case 1024 * 1000 + 20: launch_part3<1024, 2, 8>(r, pa, la); break;
		case 1024 * 1000 + 80: launch_part3<1024, 8, 8>(r, pa, la); break;
		case 768 * 1000 + 40: launch_part3<768, 4, 8>(r, pa, la); break;
		case 1024 * 1000 + 41: launch_part3<1024, 4, 16>(r, pa, la); break;
		default: printf("no such part3 variant\n"); exit(1);
		}
	}
}

// =========================================================================================================
// v4 — v3's source-level ncu page: the claim of one key costs 42 SASS instructions (generic→shared address
// conversions, reconvergence barriers around every predicated shared access) and the retry round re-runs the
// whole claim block.  v4: every shared access is inline PTX on a 32-bit shared address with a predicate (no
// branch), a failed claim is CARRIED into the next phase (one extra claim slot per phase) instead of retried,
// and one launch covers all row chunks (regions per (chunk, CTA, bucket)); the lookup kernel is bucket-outer /
// chunk-inner, so a slice is loaded once and the bitmap window in flight stays a chunk or two wide.
// =========================================================================================================
struct Part4Args {
	const uint64_t *keys;
	uint64_t n;
	uint64_t chunk_rows;   // rows per chunk (multiple of 148·512 is not required; rows_per_cta covers it)
	uint64_t rows_per_cta; // rows of a chunk handled by one CTA (multiple of 512)
	uint32_t nchunks;
	uint32_t smask;
	uint32_t nb;
	uint32_t cap_lines;
	uint64_t *records; // [bucket][chunk·148 + cta][cap_lines][SLOTS]
	uint32_t *counts;  // [bucket][chunk·148 + cta]
	const uint64_t *filter;
	uint32_t *bitmap;
	unsigned long long *stats; // [0] region overflow, [1] carried claims, [2] direct lookups
};

__device__ __forceinline__ uint32_t atoms_add1_pred(uint32_t saddr, bool want) { // returns 0xffff when !want
	uint32_t old = 0xffffu;
	asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.u32 p, %2, 0;\n\t@p atom.shared.add.u32 %0, [%1], 1;\n\t}"
	             : "+r"(old)
	             : "r"(saddr), "r"((uint32_t)want)
	             : "memory");
	return old;
}
__device__ __forceinline__ void sts64_pred(uint32_t saddr, uint32_t lo, uint32_t hi, bool pred) {
	asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.u32 p, %3, 0;\n\t@p st.shared.v2.u32 [%0], {%1, %2};\n\t}" ::"r"(saddr), "r"(lo),
	             "r"(hi), "r"((uint32_t)pred)
	             : "memory");
}
__device__ __forceinline__ void sts32_pred(uint32_t saddr, uint32_t v, bool pred) {
	asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.u32 p, %2, 0;\n\t@p st.shared.u32 [%0], %1;\n\t}" ::"r"(saddr), "r"(v),
	             "r"((uint32_t)pred)
	             : "memory");
}
__device__ __forceinline__ uint32_t lds32(uint32_t saddr) {
	uint32_t v;
	asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(saddr) : "memory");
	return v;
}
__device__ __forceinline__ uint2 lds64(uint32_t saddr) {
	uint2 v;
	asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(saddr) : "memory");
	return v;
}
__device__ __forceinline__ void direct_lookup4(const Part4Args &a, uint32_t bucket, uint32_t lo, uint32_t up,
                                               uint64_t row_lo) {
	const uint64_t w = a.filter[((uint64_t)bucket << kSliceLog) | (lo & ((1u << kSliceLog) - 1u))];
	if (record2_hit(w, lo, up)) {
		const uint64_t row = row_lo + (up >> 12);
		atomicOr(a.bitmap + (row >> 5), 1u << (row & 31u));
	}
}

template <int BLOCK, int KPP, int SLOTS>
__global__ void __launch_bounds__(BLOCK, 1) part4_kernel(const Part4Args a) {
	constexpr int NWARPS = BLOCK / 32;
	constexpr int LPL = SLOTS / 2;      // lanes per line (16 bytes each)
	constexpr int GROUPS = 32 / LPL;    // lines flushed per iteration
	constexpr uint32_t SEG = 32u * KPP; // keys per phase and warp
	constexpr uint32_t LINE_BYTES = SLOTS * 8;
	extern __shared__ __align__(128) unsigned char s_dyn[];
	__shared__ __align__(8) uint64_t bars[NWARPS][2];
	const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
	const uint32_t grp = lane / LPL, sub = lane % LPL;
	// shared layout: [stage nb×LINE_BYTES][ctr nb×4][flist NWARPS×(SEG+32)×4][kbuf NWARPS×2×SEG×8]
	const uint32_t stage_base = (uint32_t)__cvta_generic_to_shared(s_dyn);
	const uint32_t ctr_base = stage_base + a.nb * LINE_BYTES;
	const uint32_t flist_base = ctr_base + a.nb * 4u + warp * (SEG + 32u) * 4u;
	const uint32_t kbuf_base = ctr_base + a.nb * 4u + NWARPS * (SEG + 32u) * 4u + warp * 2u * SEG * 8u;
	unsigned char *const kbuf_gen = s_dyn + (kbuf_base - stage_base);
	uint32_t lt;
	asm("mov.u32 %0, %%lanemask_lt;" : "=r"(lt));
	for (uint32_t i = threadIdx.x; i < a.nb * SLOTS; i += BLOCK) {
		reinterpret_cast<uint64_t *>(s_dyn)[i] = 0;
	}
	for (uint32_t i = threadIdx.x; i < a.nb; i += BLOCK) {
		reinterpret_cast<uint32_t *>(s_dyn + (size_t)a.nb * LINE_BYTES)[i] = 0;
	}
	if (lane == 0) {
		mbar_init(&bars[warp][0], 1);
		mbar_init(&bars[warp][1], 1);
		asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
	}
	__syncthreads();
	const uint32_t nq = a.nchunks * gridDim.x;
	const uint64_t bucket_stride = (uint64_t)nq * a.cap_lines * LINE_BYTES; // bytes between two buckets' region sets
	uint32_t n_carry = 0, n_over = 0, n_direct = 0;
	uint32_t phase = 0; // bit b: parity to wait for on buffer b

	for (uint32_t chunk = 0; chunk < a.nchunks; chunk++) {
		const uint64_t chunk_lo = (uint64_t)chunk * a.chunk_rows;
		const uint64_t chunk_hi = chunk_lo + a.chunk_rows < a.n ? chunk_lo + a.chunk_rows : a.n;
		const uint64_t row_lo = chunk_lo + (uint64_t)blockIdx.x * a.rows_per_cta;
		const uint64_t row_hi = row_lo + a.rows_per_cta < chunk_hi ? row_lo + a.rows_per_cta : chunk_hi;
		const uint64_t nrows = row_hi > row_lo ? row_hi - row_lo : 0;
		const uint32_t nseg = (uint32_t)((nrows + SEG - 1) / SEG);
		const uint32_t full_segs = (uint32_t)(nrows / SEG);
		const uint32_t q = chunk * gridDim.x + blockIdx.x;
		unsigned char *const my_records = reinterpret_cast<unsigned char *>(a.records) + (uint64_t)q * a.cap_lines * LINE_BYTES;

		auto issue = [&](uint32_t seg, uint32_t buf) { // lane 0: stage segment `seg` into buffer `buf`
			if (seg < full_segs) {
				mbar_expect_tx(&bars[warp][buf], SEG * 8);
				bulk_g2s(kbuf_gen + buf * SEG * 8, a.keys + row_lo + (uint64_t)seg * SEG, SEG * 8, &bars[warp][buf]);
			}
		};
		// carried key (a claim that found its line full): claimed again as an extra slot of the next phase
		bool c_valid = false;
		uint32_t c_b = 0, c_lo = 0, c_up = 0;
		uint32_t nfl = 0;
		auto claim = [&](const uint32_t b, const uint32_t lo, const uint32_t up, const bool want) -> bool {
			const uint32_t old = atoms_add1_pred(ctr_base + b * 4u, want);
			const uint32_t c = old & 0xffffu;
			const bool ok = c < (uint32_t)SLOTS;
			sts64_pred(stage_base + b * LINE_BYTES + c * 8u, lo, up, ok);
			const bool completes = c == (uint32_t)SLOTS - 1u;
			const uint32_t fl = __ballot_sync(0xffffffffu, completes);
			sts32_pred(flist_base + (nfl + __popc(fl & lt)) * 4u, b | ((old >> 16) << 12), completes);
			nfl += __popc(fl);
			return want && !ok;
		};

		if (lane == 0) {
			issue(warp, 0);
		}
		uint32_t buf = 0;
		for (uint32_t seg = warp; seg < nseg; seg += NWARPS, buf ^= 1u) {
			if (lane == 0) {
				issue(seg + NWARPS, buf ^ 1u);
			}
			const bool staged = seg < full_segs;
			if (staged) {
				mbar_wait(&bars[warp][buf], (phase >> buf) & 1u);
				phase ^= 1u << buf;
			}
			nfl = 0;
			{ // the carried key first
				const bool failed = claim(c_b, c_lo, c_up, c_valid);
				n_carry += failed ? 1u : 0u;
				c_valid = failed;
			}
#pragma unroll
			for (int i = 0; i < KPP; i++) {
				const uint32_t rowrel = seg * SEG + i * 32u + lane;
				const bool act = rowrel < nrows;
				uint64_t k = 0;
				if (staged) {
					const uint2 kk = lds64(kbuf_base + buf * SEG * 8u + (i * 32u + lane) * 8u);
					k = ((uint64_t)kk.y << 32) | kk.x;
				} else if (act) {
					k = a.keys[row_lo + rowrel];
				}
				const uint64_t h = murmur64(k);
				const uint32_t hm = (uint32_t)(h >> 32) & 0x3f3f3f3fu;
				const uint32_t lo = (((uint32_t)h & ((1u << kSliceLog) - 1u)) | kValid3) + hm * (1u << kSliceLog);
				const uint32_t up = (hm >> 18) + rowrel * 4096u;
				const uint32_t b = ((uint32_t)h & a.smask) >> kSliceLog;
				const bool failed = claim(b, lo, up, act);
				if (failed) {
					if (!c_valid) {
						c_valid = true;
						c_b = b;
						c_lo = lo;
						c_up = up;
					} else { // two failures in one lane and phase: rare — look up on the spot
						direct_lookup4(a, b, lo, up, row_lo);
						n_direct++;
					}
				}
			}
			__syncwarp();
			for (uint32_t it = 0; it < nfl; it += GROUPS) {
				const uint32_t g = it + grp;
				if (g < nfl) {
					const uint32_t e = lds32(flist_base + g * 4u);
					const uint32_t fb = e & 4095u, line = e >> 12;
					const uint32_t saddr = stage_base + fb * LINE_BYTES + sub * 16u;
					uint4 v;
					do { // a claimer of this line may still be between its atomicAdd and its store
						asm volatile("ld.volatile.shared.v4.u32 {%0, %1, %2, %3}, [%4];"
						             : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
						             : "r"(saddr)
						             : "memory");
					} while (!(v.x & v.z & kValid3));
					asm volatile("st.volatile.shared.v4.u32 [%0], {%1, %1, %1, %1};" ::"r"(saddr), "r"(0u) : "memory");
					__syncwarp(LPL == 32 ? 0xffffffffu : (((1u << LPL) - 1u) << (grp * LPL)));
					if (sub == 0) { // re-opens the line
						uint32_t dummy;
						asm volatile("atom.shared.exch.b32 %0, [%1], %2;"
						             : "=r"(dummy)
						             : "r"(ctr_base + fb * 4u), "r"((line + 1u) << 16)
						             : "memory");
					}
					if (line < a.cap_lines) {
						unsigned char *dst = my_records + fb * bucket_stride + line * LINE_BYTES + sub * 16u;
						asm volatile("st.global.cs.v4.u32 [%0], {%1, %2, %3, %4};" ::"l"(dst), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w)
						             : "memory");
					} else { // skew: the (CTA, bucket) region is full — look the records up on the spot
						direct_lookup4(a, fb, v.x, v.y, row_lo);
						direct_lookup4(a, fb, v.z, v.w, row_lo);
						n_over += 2;
					}
				}
			}
			__syncwarp(); // flist is rewritten by the next phase
		}
		if (c_valid) { // the chunk's last carried key
			direct_lookup4(a, c_b, c_lo, c_up, row_lo);
			n_direct++;
		}
		__syncthreads();
		// partial lines of this chunk; the stage and the counters start the next chunk empty
		for (uint32_t b = threadIdx.x; b < a.nb; b += BLOCK) {
			uint32_t *ctr = reinterpret_cast<uint32_t *>(s_dyn + (size_t)a.nb * LINE_BYTES) + b;
			uint64_t *st = reinterpret_cast<uint64_t *>(s_dyn) + b * SLOTS;
			const uint32_t c = *ctr & 0xffffu; // < SLOTS: every full line has been flushed and re-opened
			const uint32_t l = *ctr >> 16;
			uint32_t total = (l < a.cap_lines ? l : a.cap_lines) * SLOTS;
			for (uint32_t j = 0; j < c && j < (uint32_t)SLOTS; j++) {
				const uint64_t v = st[j];
				st[j] = 0;
				if (l < a.cap_lines) {
					*reinterpret_cast<uint64_t *>(my_records + b * bucket_stride + l * LINE_BYTES + j * 8u) = v;
					total++;
				} else {
					direct_lookup4(a, b, (uint32_t)v, (uint32_t)(v >> 32), row_lo);
					n_over++;
				}
			}
			*ctr = 0;
			a.counts[(uint64_t)b * nq + q] = total;
		}
		__syncthreads();
	}
	n_carry = __reduce_add_sync(0xffffffffu, n_carry);
	n_over = __reduce_add_sync(0xffffffffu, n_over);
	n_direct = __reduce_add_sync(0xffffffffu, n_direct);
	if (lane == 0) {
		if (n_over) {
			atomicAdd(a.stats, (unsigned long long)n_over);
		}
		if (n_carry) {
			atomicAdd(a.stats + 1, (unsigned long long)n_carry);
		}
		if (n_direct) {
			atomicAdd(a.stats + 2, (unsigned long long)n_direct);
		}
	}
}

struct Look4Args {
	const uint64_t *filter;
	uint32_t nb;
	uint32_t nparts;  // partition CTAs
	uint32_t nchunks;
	uint32_t cap_lines;
	uint64_t chunk_rows, rows_per_cta;
	const uint64_t *records;
	const uint32_t *counts;
	uint32_t *bitmap;
};

// bucket-outer: the slice arrives once per bucket; a warp's items are the regions q = warp, warp + NWARPS, … of
// the bucket's nchunks·nparts regions (chunk-major, so all warps move through the chunks together and the part of
// the bitmap being written stays a chunk or two wide); item i+1's records are loaded before item i is looked up.
template <int BLOCK, int U, int SLOTS>
__global__ void __launch_bounds__(BLOCK, 1) lookup4_kernel(const Look4Args a) {
	extern __shared__ __align__(128) unsigned char s_dyn[];
	uint64_t *slice = reinterpret_cast<uint64_t *>(s_dyn);
	__shared__ __align__(8) uint64_t bar;
	constexpr uint32_t kSliceBytes = (1u << kSliceLog) * 8u;
	const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
	constexpr uint32_t NWARPS = BLOCK / 32;
	if (threadIdx.x == 0) {
		mbar_init(&bar, 1);
		asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
	}
	__syncthreads();
	const uint32_t nq = a.nchunks * a.nparts;
	const uint32_t nitems = warp < nq ? (nq - warp + NWARPS - 1) / NWARPS : 0; // this warp's regions per bucket
	uint32_t parity = 0;
	for (uint32_t b = blockIdx.x; b < a.nb; b += gridDim.x) {
		if (threadIdx.x == 0) {
			mbar_expect_tx(&bar, kSliceBytes);
			bulk_g2s(slice, a.filter + ((uint64_t)b << kSliceLog), kSliceBytes, &bar);
		}
		const uint32_t *cnts = a.counts + (uint64_t)b * nq;
		const uint64_t *recs = a.records + (uint64_t)b * nq * a.cap_lines * SLOTS;
		uint32_t item = 0, j0 = 0; // next load position: region item, pair offset j0
		uint32_t cnt_reg = 0, cnt_k = 0;
		auto fetch_cnt = [&]() { // counts of items [item & ~31, +32) live in the lanes
			if ((item & 31u) == 0u) {
				const uint32_t it = item + lane;
				cnt_reg = it < nitems ? cnts[warp + it * NWARPS] : 0u;
			}
			cnt_k = __shfl_sync(0xffffffffu, cnt_reg, item & 31u);
		};
		if (nitems) {
			fetch_cnt();
		}
		auto load_item = [&](ulonglong2 (&v)[U], uint32_t &q_out) -> bool {
			while (item < nitems && j0 >= ((cnt_k + 1u) >> 1)) { // region finished (or empty): next one
				item++;
				j0 = 0;
				if (item < nitems) {
					fetch_cnt();
				}
			}
			if (item >= nitems) {
				return false;
			}
			const uint32_t q = warp + item * NWARPS;
			const ulonglong2 *src = reinterpret_cast<const ulonglong2 *>(recs + (uint64_t)q * a.cap_lines * SLOTS);
			const uint32_t npairs = (cnt_k + 1u) >> 1;
#pragma unroll
			for (int u = 0; u < U; u++) {
				const uint32_t j = j0 + u * 32 + lane;
				v[u] = make_ulonglong2(0, 0);
				if (j < npairs) {
					asm volatile("ld.global.nc.L1::no_allocate.v2.u64 {%0, %1}, [%2];"
					             : "=l"(v[u].x), "=l"(v[u].y)
					             : "l"(src + j));
					if (2 * j + 1 >= cnt_k) {
						v[u].y = 0;
					}
				}
			}
			q_out = q;
			j0 += 32 * U;
			return true;
		};
		ulonglong2 cur[U], nxt[U];
		uint32_t q_cur = 0, q_nxt = 0;
		bool have = load_item(cur, q_cur);
		mbar_wait(&bar, parity);
		parity ^= 1u;
		while (have) {
			const bool have_nxt = load_item(nxt, q_nxt);
			const uint32_t chunk = q_cur / a.nparts, part = q_cur - chunk * a.nparts;
			const uint64_t row_base = (uint64_t)chunk * a.chunk_rows + (uint64_t)part * a.rows_per_cta;
			uint64_t w[2 * U];
#pragma unroll
			for (int u = 0; u < U; u++) {
				w[2 * u] = slice[(uint32_t)cur[u].x & ((1u << kSliceLog) - 1u)];
				w[2 * u + 1] = slice[(uint32_t)cur[u].y & ((1u << kSliceLog) - 1u)];
			}
#pragma unroll
			for (int u = 0; u < 2 * U; u++) {
				const uint64_t rec = (u & 1) ? cur[u >> 1].y : cur[u >> 1].x;
				const uint32_t lo = (uint32_t)rec, up = (uint32_t)(rec >> 32);
				if (record2_hit(w[u], lo, up) && (lo & kValid3)) {
					const uint64_t row = row_base + (up >> 12);
					atomicOr(a.bitmap + (row >> 5), 1u << (row & 31u));
				}
			}
#pragma unroll
			for (int u = 0; u < U; u++) {
				cur[u] = nxt[u];
			}
			q_cur = q_nxt;
			have = have_nxt;
		}
		__syncthreads();
	}
}

struct Part4Run {
	Part4Args pa;
	Look4Args la;
	int block, kpp, slots;
	size_t part_smem;
	uint64_t bitmap_bytes;
	int what;
};
template <int BLOCK, int KPP, int SLOTS>
static void launch_part4(const Part4Run *r) {
	if (r->what != 2) {
		CK(cudaMemsetAsync(r->pa.bitmap, 0, r->bitmap_bytes)); // the partition's direct lookups OR into it
		CK(cudaFuncSetAttribute(part4_kernel<BLOCK, KPP, SLOTS>, cudaFuncAttributeMaxDynamicSharedMemorySize,
		                        (int)r->part_smem));
		part4_kernel<BLOCK, KPP, SLOTS><<<148, BLOCK, r->part_smem>>>(r->pa);
	}
	if (r->what != 1) {
		const size_t smem = (1u << kSliceLog) * 8;
		CK(cudaFuncSetAttribute(lookup4_kernel<1024, 4, SLOTS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
		lookup4_kernel<1024, 4, SLOTS><<<148, 1024, smem>>>(r->la);
	}
}
static void run_part4(void *p) {
	Part4Run *r = (Part4Run *)p;
	const int key = r->block * 1000 + r->kpp * 10 + (r->slots == 8 ? 0 : 1);
	switch (key) {
	case 1024 * 1000 + 40: launch_part4<1024, 4, 8>(r); break;
	case 1024 * 1000 + 20: launch_part4<1024, 2, 8>(r); break;
	case 1024 * 1000 + 80: launch_part4<1024, 8, 8>(r); break;
	case 768 * 1000 + 40: launch_part4<768, 4, 8>(r); break;
	case 512 * 1000 + 40: launch_part4<512, 4, 8>(r); break;
	case 1024 * 1000 + 41: launch_part4<1024, 4, 16>(r); break;
	default: printf("no such part4 variant\n"); exit(1);
	}
}

static float time_kernel(int reps, void (*fn)(void *), void *arg) {
	cudaEvent_t e0, e1;
	CK(cudaEventCreate(&e0));
	CK(cudaEventCreate(&e1));
	fn(arg);
	CK(cudaDeviceSynchronize());
	CK(cudaEventRecord(e0));
	for (int i = 0; i < reps; i++) {
		fn(arg);
	}
	CK(cudaEventRecord(e1));
	CK(cudaEventSynchronize(e1));
	float ms = 0;
	CK(cudaEventElapsedTime(&ms, e0, e1));
	return ms / reps;
}

struct AtomsRun {
	int mode;
	int threads;
	uint32_t per_thread;
	uint32_t *sink;
};
static void run_atoms(void *p) {
	AtomsRun *r = (AtomsRun *)p;
	const int threads = r->threads;
	switch (r->mode) {
	case 0: atoms_kernel<0><<<148, threads>>>(r->per_thread, r->sink); break;
	case 1: atoms_kernel<1><<<148, threads>>>(r->per_thread, r->sink); break;
	case 2: atoms_kernel<2><<<148, threads>>>(r->per_thread, r->sink); break;
	case 3: atoms_kernel<3><<<148, threads>>>(r->per_thread, r->sink); break;
	case 4: atoms_kernel<4><<<148, threads>>>(r->per_thread, r->sink); break;
	default: atoms_kernel<5><<<148, threads>>>(r->per_thread, r->sink); break;
	}
}

struct PartRun {
	PartArgs pa;
	LookArgs la;
	int block, kpt;
	bool fake;
	size_t part_smem;
	uint64_t n_total, chunk_rows;
	int what; // 0 both, 1 partition only, 2 lookup only
};
template <int BLOCK, int KPT>
static void launch_part(const PartRun *r, const PartArgs &pa) {
	if (r->fake) {
		CK(cudaFuncSetAttribute(part_kernel<BLOCK, KPT, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)r->part_smem));
		part_kernel<BLOCK, KPT, true><<<148, BLOCK, r->part_smem>>>(pa);
	} else {
		CK(cudaFuncSetAttribute(part_kernel<BLOCK, KPT, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)r->part_smem));
		part_kernel<BLOCK, KPT, false><<<148, BLOCK, r->part_smem>>>(pa);
	}
}
static uint64_t rows_per_cta_for(uint64_t n) {
	return ((n + 147) / 148 + 511) / 512 * 512;
}
static void run_part(void *p) {
	PartRun *r = (PartRun *)p;
	for (uint64_t c0 = 0; c0 < r->n_total; c0 += r->chunk_rows) {
		const uint64_t cn = c0 + r->chunk_rows < r->n_total ? r->chunk_rows : r->n_total - c0;
		PartArgs pa = r->pa;
		LookArgs la = r->la;
		pa.keys += c0;
		pa.n = cn;
		pa.rows_per_cta = la.rows_per_cta = rows_per_cta_for(cn);
		la.bitmap += c0 / 32;
		if (r->what != 2) {
			if (r->block == 1024 && r->kpt == 8) {
				launch_part<1024, 8>(r, pa);
			} else if (r->block == 1024 && r->kpt == 4) {
				launch_part<1024, 4>(r, pa);
			} else if (r->block == 512 && r->kpt == 8) {
				launch_part<512, 8>(r, pa);
			} else if (r->block == 512 && r->kpt == 4) {
				launch_part<512, 4>(r, pa);
			} else {
				launch_part<256, 8>(r, pa);
			}
		}
		if (r->what != 1) {
			CK(cudaMemsetAsync(la.bitmap, 0, (cn + 31) / 32 * 4));
			const size_t smem = (1u << kSliceLog) * 8;
			CK(cudaFuncSetAttribute(lookup_kernel<1024, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
			lookup_kernel<1024, 4><<<148, 1024, smem>>>(la);
		}
	}
}

int main(int argc, char **argv) {
	const char *mode = argc > 1 ? argv[1] : "part";
	if (!strcmp(mode, "atoms")) {
		uint32_t *sink;
		CK(cudaMalloc(&sink, 4));
		const char *names[6] = {"atomicAdd/2048 (return)", "plain ld+st", "red.shared.add", "match_any", "LDS.64 random",
		                        "atomicAdd/256 (return)"};
		for (int threads = 256; threads <= 1024; threads *= 2) {
			for (int m = 0; m < 6; m++) {
				AtomsRun r {m, threads, 4096, sink};
				const float ms = time_kernel(5, run_atoms, &r);
				const double ops = 148.0 * threads * 4096;
				printf("atoms %4d thr/SM %-24s %.3f ms  %.1f Gops/s  %.2f ops/clk/SM (at 1.9 GHz)\n", threads, names[m], ms,
				       ops / ms / 1e6, ops / ms / 1e6 / 148 / 1.9);
			}
		}
		return 0;
	}
	if (!strcmp(mode, "part4")) {
		const int log_sectors = argc > 2 ? atoi(argv[2]) : 25;
		const int block = argc > 3 ? atoi(argv[3]) : 1024;
		const int kpp = argc > 4 ? atoi(argv[4]) : 4;
		const uint64_t n = argc > 5 ? strtoull(argv[5], nullptr, 10) : 1000000000ULL;
		const int chunk_log = argc > 6 ? atoi(argv[6]) : 27;
		const int density = argc > 7 ? atoi(argv[7]) : 1;
		const int slots = argc > 8 ? atoi(argv[8]) : 8;
		const int skew = argc > 9 ? atoi(argv[9]) : 0;
		const uint64_t chunk_rows = 1ULL << chunk_log;
		const uint64_t S = 1ULL << log_sectors;
		const uint32_t nb = (uint32_t)(S >> kSliceLog);
		const uint32_t nparts = 148;
		const uint32_t nchunks = (uint32_t)((n + chunk_rows - 1) / chunk_rows);
		const uint64_t rows_per_cta = rows_per_cta_for(chunk_rows < n ? chunk_rows : n);
		if (rows_per_cta > (1u << 20)) {
			printf("chunk too large for the row field\n");
			return 1;
		}
		const double mean = (double)rows_per_cta / nb;
		const uint32_t cap_lines = (uint32_t)((mean + 6.0 * sqrt(mean) + 2.0 * slots) / slots) + 1;
		const uint64_t nq = (uint64_t)nchunks * nparts;
		printf("part4 n=%llu chunks=%u x 2^%d S=2^%d (%llu MiB) buckets=%u slots=%d rows/cta=%llu mean/bucket=%.1f cap_lines=%u "
		       "scratch=%.2f GB block=%d kpp=%d density=%d skew=%d\n",
		       (unsigned long long)n, nchunks, chunk_log, log_sectors, (unsigned long long)(S * 8 >> 20), nb, slots,
		       (unsigned long long)rows_per_cta, mean, cap_lines, (double)nq * nb * cap_lines * slots * 8 / 1e9, block, kpp,
		       density, skew);
		uint64_t *keys, *filter, *records;
		uint32_t *counts, *bitmap, *bitmap_ref;
		unsigned long long *stats;
		const uint64_t bitmap_words = (n + 31) / 32 + 64;
		CK(cudaMalloc(&keys, n * 8));
		CK(cudaMalloc(&filter, S * 8));
		CK(cudaMalloc(&records, nq * nb * cap_lines * slots * 8));
		CK(cudaMalloc(&counts, nq * nb * 4));
		CK(cudaMalloc(&bitmap, bitmap_words * 4));
		CK(cudaMalloc(&bitmap_ref, bitmap_words * 4));
		CK(cudaMalloc(&stats, 8 * 8));
		CK(cudaMemset(stats, 0, 8 * 8));
		gen_keys<<<148 * 8, 256>>>(keys, n);
		if (skew > 0) {
			mask_keys<<<148 * 8, 256>>>(keys, n, (1ULL << skew) - 1);
		}
		fill_filter<<<148 * 8, 256>>>(filter, S, density);
		CK(cudaDeviceSynchronize());
		Part4Run r;
		r.pa = Part4Args {keys, n, chunk_rows, rows_per_cta, nchunks, (uint32_t)(S - 1), nb, cap_lines, records, counts,
		                  filter, bitmap, stats};
		r.la = Look4Args {filter, nb, nparts, nchunks, cap_lines, chunk_rows, rows_per_cta, records, counts, bitmap};
		r.block = block;
		r.kpp = kpp;
		r.slots = slots;
		r.part_smem = (size_t)nb * slots * 8 + (size_t)nb * 4 + (size_t)(block / 32) * (32 * kpp + 32) * 4 +
		              (size_t)(block / 32) * 2 * 32 * kpp * 8;
		r.bitmap_bytes = bitmap_words * 4;
		unsigned long long hstats[8];
		r.what = 1;
		const float ms_part = time_kernel(3, run_part4, &r);
		CK(cudaMemcpy(hstats, stats, sizeof(hstats), cudaMemcpyDeviceToHost));
		printf("partition  %.3f ms  %.1f Gkeys/s  (16 B/key: %.2f TB/s)  per run: overflow=%.0f carried claims=%.0f (%.3f %%) "
		       "direct=%.0f\n",
		       ms_part, n / ms_part / 1e6, 16.0 * n / ms_part / 1e9, hstats[0] / 4.0, hstats[1] / 4.0,
		       100.0 * hstats[1] / 4.0 / n, hstats[2] / 4.0);
		r.what = 2;
		const float ms_look = time_kernel(3, run_part4, &r);
		printf("lookup     %.3f ms  %.1f Gkeys/s  (8 B/key: %.2f TB/s)\n", ms_look, n / ms_look / 1e6, 8.0 * n / ms_look / 1e9);
		r.what = 0;
		const float ms_both = time_kernel(3, run_part4, &r);
		printf("both       %.3f ms  %.1f Gkeys/s\n", ms_both, n / ms_both / 1e6);
		ref_probe<<<148 * 8, 256>>>(keys, filter, S - 1, n & ~31ULL, bitmap_ref);
		CK(cudaMemset(stats, 0, 8 * 8));
		cmp_words<<<148 * 4, 256>>>(bitmap, bitmap_ref, n / 32, stats + 2, stats + 3);
		CK(cudaDeviceSynchronize());
		CK(cudaMemcpy(hstats, stats, sizeof(hstats), cudaMemcpyDeviceToHost));
		printf("check      %llu differing bitmap words, %llu hits (%.2f %%)  %s\n", hstats[2], hstats[3],
		       100.0 * hstats[3] / n, hstats[2] == 0 ? "OK" : "MISMATCH");
		return hstats[2] == 0 ? 0 : 1;
	}
	if (!strcmp(mode, "part2") || !strcmp(mode, "part3")) {
		const bool v3 = !strcmp(mode, "part3");
		const int log_sectors = argc > 2 ? atoi(argv[2]) : 25;
		const int block = argc > 3 ? atoi(argv[3]) : 1024;
		const int kpt = argc > 4 ? atoi(argv[4]) : 8;
		const uint64_t n = argc > 5 ? strtoull(argv[5], nullptr, 10) : 1000000000ULL;
		const int chunk_log = argc > 6 ? atoi(argv[6]) : 26;
		const int density = argc > 7 ? atoi(argv[7]) : 1;
		const int slots = argc > 8 ? atoi(argv[8]) : 8;
		const int skew = argc > 9 ? atoi(argv[9]) : 0; // keys &= (2^skew - 1) when > 0: few distinct keys
		const uint64_t chunk_rows = 1ULL << chunk_log;
		const uint64_t S = 1ULL << log_sectors;
		const uint32_t nb = (uint32_t)(S >> kSliceLog);
		const uint32_t nparts = 148;
		const uint64_t rows_per_cta = rows_per_cta_for(chunk_rows < n ? chunk_rows : n);
		if (rows_per_cta > (v3 ? (1u << 20) : (1u << 19))) {
			printf("chunk too large for the row field\n");
			return 1;
		}
		const double mean = (double)rows_per_cta / nb;
		const uint32_t cap_lines = (uint32_t)((mean + 6.0 * sqrt(mean) + 2.0 * slots) / slots) + 1;
		printf("part2 n=%llu chunk=2^%d S=2^%d (%llu MiB) buckets=%u slots=%d rows/cta=%llu mean/bucket=%.1f cap_lines=%u "
		       "scratch=%.2f GB block=%d kpt=%d density=%d skew=%d\n",
		       (unsigned long long)n, chunk_log, log_sectors, (unsigned long long)(S * 8 >> 20), nb, slots,
		       (unsigned long long)rows_per_cta, mean, cap_lines, (double)nparts * nb * cap_lines * slots * 8 / 1e9, block,
		       kpt, density, skew);
		uint64_t *keys, *filter, *records;
		uint32_t *counts, *bitmap, *bitmap_ref;
		unsigned long long *stats;
		const uint64_t bitmap_words = (n + 31) / 32 + 64;
		CK(cudaMalloc(&keys, n * 8));
		CK(cudaMalloc(&filter, S * 8));
		CK(cudaMalloc(&records, (uint64_t)nparts * nb * cap_lines * slots * 8));
		CK(cudaMalloc(&counts, (uint64_t)nparts * nb * 4));
		CK(cudaMalloc(&bitmap, bitmap_words * 4));
		CK(cudaMalloc(&bitmap_ref, bitmap_words * 4));
		CK(cudaMalloc(&stats, 8 * 8));
		CK(cudaMemset(stats, 0, 8 * 8));
		gen_keys<<<148 * 8, 256>>>(keys, n);
		if (skew > 0) {
			mask_keys<<<148 * 8, 256>>>(keys, n, (1ULL << skew) - 1);
		}
		fill_filter<<<148 * 8, 256>>>(filter, S, density);
		CK(cudaDeviceSynchronize());
		Part2Run r;
		r.pa = Part2Args {keys, n, rows_per_cta, (uint32_t)(S - 1), nb, cap_lines, records, counts, filter, bitmap, stats};
		r.la = LookArgs {filter, nb, nparts, cap_lines, rows_per_cta, records, counts, bitmap};
		r.block = block;
		r.kpt = kpt;
		r.slots = slots;
		r.part_smem = (size_t)nb * slots * 8 + (size_t)nb * 4 + (block / 32) * 128 + (size_t)(block / 32) * 2 * 32 * kpt * 8;
		r.n_total = n;
		r.chunk_rows = chunk_rows;
		Part3Run r3;
		r3.pa = r.pa; r3.la = r.la; r3.block = block; r3.kpp = kpt; r3.slots = slots; r3.n_total = n; r3.chunk_rows = chunk_rows;
		r3.part_smem = (size_t)nb * slots * 8 + (size_t)nb * 4 + (size_t)(block / 32) * 32 * kpt * 4 +
		               (size_t)(block / 32) * 2 * 32 * kpt * 8;
		void (*fn)(void *) = v3 ? run_part3 : run_part2;
		void *arg = v3 ? (void *)&r3 : (void *)&r;
		unsigned long long hstats[8];
		r.what = r3.what = 1;
		const float ms_part = time_kernel(3, fn, arg);
		CK(cudaMemcpy(hstats, stats, sizeof(hstats), cudaMemcpyDeviceToHost));
		printf("partition  %.3f ms  %.1f Gkeys/s  (16 B/key: %.2f TB/s)  per run: overflow=%.0f failed claims=%.0f (%.3f %%) "
		       "direct=%.0f\n",
		       ms_part, n / ms_part / 1e6, 16.0 * n / ms_part / 1e9, hstats[0] / 4.0, hstats[1] / 4.0,
		       100.0 * hstats[1] / 4.0 / n, hstats[2] / 4.0);
		r.what = r3.what = 0;
		const float ms_both = time_kernel(3, fn, arg);
		printf("both       %.3f ms  %.1f Gkeys/s  (lookup ~ %.3f ms)\n", ms_both, n / ms_both / 1e6, ms_both - ms_part);
		ref_probe<<<148 * 8, 256>>>(keys, filter, S - 1, n & ~31ULL, bitmap_ref);
		CK(cudaMemset(stats, 0, 8 * 8));
		cmp_words<<<148 * 4, 256>>>(bitmap, bitmap_ref, n / 32, stats + 2, stats + 3);
		CK(cudaDeviceSynchronize());
		CK(cudaMemcpy(hstats, stats, sizeof(hstats), cudaMemcpyDeviceToHost));
		printf("check      %llu differing bitmap words, %llu hits (%.2f %%)  %s\n", hstats[2], hstats[3],
		       100.0 * hstats[3] / n, hstats[2] == 0 ? "OK" : "MISMATCH");
		return hstats[2] == 0 ? 0 : 1;
	}
	const int log_sectors = argc > 2 ? atoi(argv[2]) : 25;
	const int block = argc > 3 ? atoi(argv[3]) : 1024;
	const int kpt = argc > 4 ? atoi(argv[4]) : 8;
	const uint64_t n = argc > 5 ? strtoull(argv[5], nullptr, 10) : 1000000000ULL;
	const int chunk_log = argc > 6 ? atoi(argv[6]) : 32;
	const int density = argc > 7 ? atoi(argv[7]) : 1;
	const bool fake = argc > 8 && atoi(argv[8]) != 0;
	const uint64_t chunk_rows = chunk_log >= 32 ? n : (1ULL << chunk_log);
	const uint64_t S = 1ULL << log_sectors;
	const uint32_t nb = (uint32_t)(S >> kSliceLog);
	const uint32_t nparts = 148;
	const uint64_t rows_per_cta = rows_per_cta_for(chunk_rows < n ? chunk_rows : n);
	const double mean = (double)rows_per_cta / nb;
	const uint32_t cap_lines = (uint32_t)((mean + 6.0 * sqrt(mean) + 16.0) / 8.0) + 1;
	printf("n=%llu chunk=%llu S=2^%d (%llu MiB) buckets=%u rows/cta=%llu mean/bucket=%.1f cap_lines=%u scratch=%.2f GB block=%d "
	       "kpt=%d density=%d fake=%d\n",
	       (unsigned long long)n, (unsigned long long)chunk_rows, log_sectors, (unsigned long long)(S * 8 >> 20), nb,
	       (unsigned long long)rows_per_cta, mean, cap_lines, (double)nparts * nb * cap_lines * 64 / 1e9, block, kpt, density,
	       (int)fake);
	uint64_t *keys, *filter, *records;
	uint32_t *counts, *bitmap, *bitmap_ref;
	unsigned long long *stats;
	const uint64_t bitmap_words = (n + 31) / 32 + 64;
	CK(cudaMalloc(&keys, n * 8));
	CK(cudaMalloc(&filter, S * 8));
	CK(cudaMalloc(&records, (uint64_t)nparts * nb * cap_lines * 64));
	CK(cudaMalloc(&counts, (uint64_t)nparts * nb * 4));
	CK(cudaMalloc(&bitmap, bitmap_words * 4));
	CK(cudaMalloc(&bitmap_ref, bitmap_words * 4));
	CK(cudaMalloc(&stats, 4 * 8));
	CK(cudaMemset(stats, 0, 4 * 8));
	gen_keys<<<148 * 8, 256>>>(keys, n);
	fill_filter<<<148 * 8, 256>>>(filter, S, density);
	CK(cudaDeviceSynchronize());

	PartRun r;
	r.pa = PartArgs {keys, n, rows_per_cta, (uint32_t)(S - 1), nb, cap_lines, records, counts, stats, stats + 1};
	r.la = LookArgs {filter, nb, nparts, cap_lines, rows_per_cta, records, counts, bitmap};
	r.block = block;
	r.kpt = kpt;
	r.fake = fake;
	r.part_smem = (size_t)nb * 64 + (size_t)nb * 8;
	r.n_total = n;
	r.chunk_rows = chunk_rows;

	unsigned long long hstats[4];
	if (chunk_rows >= n) { // separate timings only make sense unchunked (the lookup needs its chunk's records)
		r.what = 1;
		const float ms_part = time_kernel(3, run_part, &r);
		CK(cudaMemcpy(hstats, stats, sizeof(hstats), cudaMemcpyDeviceToHost));
		printf("partition  %.3f ms  %.1f Gkeys/s  (16 B/key: %.2f TB/s)  overflow=%llu retries/run=%.0f (%.3f %% of keys)\n",
		       ms_part, n / ms_part / 1e6, 16.0 * n / ms_part / 1e9, hstats[0], hstats[1] / 4.0,
		       100.0 * hstats[1] / 4.0 / n);
		if (fake) {
			return 0;
		}
		r.what = 2;
		const float ms_look = time_kernel(3, run_part, &r);
		printf("lookup     %.3f ms  %.1f Gkeys/s  (8 B/key: %.2f TB/s)\n", ms_look, n / ms_look / 1e6,
		       8.0 * n / ms_look / 1e9);
	}
	r.what = 0;
	const float ms_both = time_kernel(3, run_part, &r);
	printf("both       %.3f ms  %.1f Gkeys/s\n", ms_both, n / ms_both / 1e6);

	// correctness against plain global lookups
	ref_probe<<<148 * 8, 256>>>(keys, filter, S - 1, n & ~31ULL, bitmap_ref);
	CK(cudaMemset(stats, 0, 4 * 8));
	cmp_words<<<148 * 4, 256>>>(bitmap, bitmap_ref, n / 32, stats + 2, stats + 3);
	CK(cudaDeviceSynchronize());
	CK(cudaMemcpy(hstats, stats, sizeof(hstats), cudaMemcpyDeviceToHost));
	printf("check      %llu differing bitmap words, %llu hits (%.2f %%)  %s\n", hstats[2], hstats[3],
	       100.0 * hstats[3] / n, hstats[2] == 0 ? "OK" : "MISMATCH");
	return hstats[2] == 0 ? 0 : 1;
}
