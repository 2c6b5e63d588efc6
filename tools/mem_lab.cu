// mem_lab.cu — development microbenchmark (not product): how B200's L2/HBM treats the bloom filter's access
// pattern (one random 8-byte word per key out of an F-byte array, next to an 8 B/key coalesced stream).
// It exists to choose load instructions / cache policies / L2 fetch granularity for rpt_kernels.cu on evidence.
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o tools/mem_lab tools/mem_lab.cu
//   tools/mem_lab [log2_keys=30]
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <cstdlib>

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
__device__ inline uint64_t mask_of(uint64_t h) {
	const uint32_t hi = (uint32_t)(h >> 32);
	return (1ULL << (hi & 63)) | (1ULL << ((hi >> 8) & 63)) | (1ULL << ((hi >> 16) & 63)) | (1ULL << ((hi >> 24) & 63));
}

__global__ void gen_keys(uint64_t *k, uint64_t n) {
	for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
		k[i] = splitmix64(i);
	}
}
__global__ void fill_filter(uint64_t *f, uint64_t s) {
	for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < s; i += (uint64_t)gridDim.x * blockDim.x) {
		f[i] = splitmix64(i * 77 + 1) & splitmix64(i * 77 + 2);
	}
}

enum { P_NC_EVICT_LAST = 0, P_NC_PLAIN, P_LD_PLAIN, P_NC_NOALLOC, P_NC_EVICT_NORMAL_HINT, P_LD_CG, P_NC_L2_32B_HINTLESS,
       P_LD_CV, P_LD_RELAXED, P_EVICT_FIRST, P_ATOM_OR0, P_LD_NOALLOC_HINT_FRAC };

template <int P>
__device__ __forceinline__ uint64_t ld_filter(const uint64_t *p, uint64_t pol) {
	uint64_t v;
	if (P == P_NC_EVICT_LAST || P == P_NC_EVICT_NORMAL_HINT) {
		asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.u64 %0, [%1], %2;" : "=l"(v) : "l"(p), "l"(pol));
	} else if (P == P_NC_PLAIN) {
		asm volatile("ld.global.nc.u64 %0, [%1];" : "=l"(v) : "l"(p));
	} else if (P == P_LD_PLAIN) {
		asm volatile("ld.global.u64 %0, [%1];" : "=l"(v) : "l"(p));
	} else if (P == P_NC_NOALLOC) {
		asm volatile("ld.global.nc.L1::no_allocate.u64 %0, [%1];" : "=l"(v) : "l"(p));
	} else if (P == P_LD_CG) {
		asm volatile("ld.global.cg.u64 %0, [%1];" : "=l"(v) : "l"(p));
	} else if (P == P_LD_CV) {
		asm volatile("ld.global.cv.u64 %0, [%1];" : "=l"(v) : "l"(p));
	} else if (P == P_LD_RELAXED) {
		asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p));
	} else if (P == P_EVICT_FIRST || P == P_LD_NOALLOC_HINT_FRAC) {
		asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.u64 %0, [%1], %2;" : "=l"(v) : "l"(p), "l"(pol));
	} else if (P == P_ATOM_OR0) {
		asm volatile("atom.global.or.b64 %0, [%1], 0;" : "=l"(v) : "l"(p));
	} else {
		asm volatile("ld.global.nc.L1::no_allocate.L2::64B.u64 %0, [%1];" : "=l"(v) : "l"(p));
	}
	return v;
}

// 16 keys per thread, 256 threads: the probe kernel's memory behaviour without the compaction.
template <int P, int KEYPOL>
__global__ void __launch_bounds__(256) lab_probe(const uint64_t *__restrict__ keys, const uint64_t *__restrict__ filter,
                                                 uint64_t smask, uint64_t n, unsigned long long *hits_out) {
	uint64_t pol_f = 0, pol_k = 0;
	if (P == P_NC_EVICT_LAST) {
		asm("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol_f));
	} else if (P == P_EVICT_FIRST) {
		asm("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol_f));
	} else if (P == P_LD_NOALLOC_HINT_FRAC) {
		asm("createpolicy.fractional.L2::evict_last.L2::evict_first.b64 %0, 0.25;" : "=l"(pol_f));
	} else {
		asm("createpolicy.fractional.L2::evict_normal.b64 %0, 1.0;" : "=l"(pol_f));
	}
	asm("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol_k));
	const uint64_t tile = 256ull * 16;
	uint32_t cnt = 0;
	for (uint64_t base = blockIdx.x * tile; base + tile <= n; base += (uint64_t)gridDim.x * tile) {
		ulonglong2 k[8];
#pragma unroll
		for (int r = 0; r < 8; r++) {
			const uint64_t *p = keys + base + (threadIdx.x >> 5) * 512 + r * 64 + (threadIdx.x & 31) * 2;
			if (KEYPOL == 0) {
				asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.v2.u64 {%0, %1}, [%2], %3;"
				             : "=l"(k[r].x), "=l"(k[r].y)
				             : "l"(p), "l"(pol_k));
			} else {
				asm volatile("ld.global.nc.L1::no_allocate.v2.u64 {%0, %1}, [%2];" : "=l"(k[r].x), "=l"(k[r].y) : "l"(p));
			}
		}
		uint64_t h[16], w[16];
#pragma unroll
		for (int r = 0; r < 8; r++) {
			h[2 * r] = murmur64(k[r].x);
			h[2 * r + 1] = murmur64(k[r].y);
		}
#pragma unroll
		for (int i = 0; i < 16; i++) {
			w[i] = ld_filter<P>(filter + (h[i] & smask), pol_f);
		}
#pragma unroll
		for (int i = 0; i < 16; i++) {
			const uint64_t m = mask_of(h[i]);
			cnt += (w[i] & m) == m;
		}
	}
	cnt = __reduce_add_sync(0xffffffffu, cnt);
	if ((threadIdx.x & 31) == 0 && cnt) {
		atomicAdd(hits_out, (unsigned long long)cnt);
	}
}

// build: atomicOr (no return → RED) with optional cache hint
template <int P>
__global__ void __launch_bounds__(256) lab_build(const uint64_t *__restrict__ keys, unsigned long long *filter,
                                                 uint64_t smask, uint64_t n) {
	uint64_t pol_f;
	asm("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol_f));
	const uint64_t tile = 256ull * 8;
	for (uint64_t base = blockIdx.x * tile; base + tile <= n; base += (uint64_t)gridDim.x * tile) {
		ulonglong2 k[4];
#pragma unroll
		for (int r = 0; r < 4; r++) {
			k[r] = __ldg(reinterpret_cast<const ulonglong2 *>(keys + base + r * 512 + threadIdx.x * 2));
		}
#pragma unroll
		for (int r = 0; r < 4; r++) {
#pragma unroll
			for (int e = 0; e < 2; e++) {
				const uint64_t h = murmur64(e ? k[r].y : k[r].x);
				unsigned long long *p = filter + (h & smask);
				const uint64_t m = mask_of(h);
				if (P == 0) {
					atomicOr(p, (unsigned long long)m);
				} else if (P == 1) {
					asm volatile("red.global.or.b64 [%0], %1;" ::"l"(p), "l"(m) : "memory");
				} else {
					asm volatile("red.global.or.L2::cache_hint.b64 [%0], %1, %2;" ::"l"(p), "l"(m), "l"(pol_f) : "memory");
				}
			}
		}
	}
}


// One pass of the sector-range multi-pass probe: only keys whose sector lies in [lo, hi) touch the filter, so the
// pass's working set is (hi-lo)*8 bytes and can stay L2-resident (evict_last) while the keys stream (evict_first).
template <int WARM>
__global__ void __launch_bounds__(256) lab_probe_slice(const uint64_t *__restrict__ keys,
                                                       const uint64_t *__restrict__ filter, uint64_t smask, uint64_t n,
                                                       uint64_t lo, uint64_t hi, unsigned long long *hits_out) {
	uint64_t pol_f, pol_k;
	asm("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol_f));
	asm("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol_k));
	if (WARM) { // stream the slice in once, coalesced, so that the random phase starts resident
		const ulonglong2 *g = reinterpret_cast<const ulonglong2 *>(filter + lo);
		const uint64_t nvec = (hi - lo) >> 1;
		uint64_t acc = 0;
		for (uint64_t i = blockIdx.x * 256ull + threadIdx.x; i < nvec; i += (uint64_t)gridDim.x * 256) {
			ulonglong2 v;
			asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.v2.u64 {%0, %1}, [%2], %3;"
			             : "=l"(v.x), "=l"(v.y)
			             : "l"(g + i), "l"(pol_f));
			acc ^= v.x ^ v.y;
		}
		if (acc == 0x1234567ull) {
			atomicAdd(hits_out, 1ull);
		}
	}
	const uint64_t tile = 256ull * 16;
	uint32_t cnt = 0;
	for (uint64_t base = blockIdx.x * tile; base + tile <= n; base += (uint64_t)gridDim.x * tile) {
		ulonglong2 k[8];
#pragma unroll
		for (int r = 0; r < 8; r++) {
			const uint64_t *p = keys + base + (threadIdx.x >> 5) * 512 + r * 64 + (threadIdx.x & 31) * 2;
			asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.v2.u64 {%0, %1}, [%2], %3;"
			             : "=l"(k[r].x), "=l"(k[r].y)
			             : "l"(p), "l"(pol_k));
		}
		uint64_t h[16], w[16];
#pragma unroll
		for (int r = 0; r < 8; r++) {
			h[2 * r] = murmur64(k[r].x);
			h[2 * r + 1] = murmur64(k[r].y);
		}
#pragma unroll
		for (int i = 0; i < 16; i++) {
			const uint64_t sct = h[i] & smask;
			w[i] = 0;
			if (sct - lo < hi - lo) {
				w[i] = ld_filter<P_NC_EVICT_LAST>(filter + sct, pol_f);
			}
		}
#pragma unroll
		for (int i = 0; i < 16; i++) {
			const uint64_t m = mask_of(h[i]);
			cnt += (w[i] & m) == m;
		}
	}
	cnt = __reduce_add_sync(0xffffffffu, cnt);
	if ((threadIdx.x & 31) == 0 && cnt) {
		atomicAdd(hits_out, (unsigned long long)cnt);
	}
}

template <int P>
__global__ void __launch_bounds__(256) lab_build_slice(const uint64_t *__restrict__ keys, unsigned long long *filter,
                                                       uint64_t smask, uint64_t n, uint64_t lo, uint64_t hi) {
	uint64_t pol_f, pol_k;
	asm("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol_f));
	asm("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol_k));
	const uint64_t tile = 256ull * 8;
	for (uint64_t base = blockIdx.x * tile; base + tile <= n; base += (uint64_t)gridDim.x * tile) {
		ulonglong2 k[4];
#pragma unroll
		for (int r = 0; r < 4; r++) {
			asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.v2.u64 {%0, %1}, [%2], %3;"
			             : "=l"(k[r].x), "=l"(k[r].y)
			             : "l"(keys + base + r * 512 + threadIdx.x * 2), "l"(pol_k));
		}
#pragma unroll
		for (int r = 0; r < 4; r++) {
#pragma unroll
			for (int e = 0; e < 2; e++) {
				const uint64_t h = murmur64(e ? k[r].y : k[r].x);
				const uint64_t sct = h & smask;
				if (sct - lo < hi - lo) {
					const uint64_t m = mask_of(h);
					if (P == 0) {
						asm volatile("red.global.or.b64 [%0], %1;" ::"l"(filter + sct), "l"(m) : "memory");
					} else {
						asm volatile("red.global.or.L2::cache_hint.b64 [%0], %1, %2;" ::"l"(filter + sct), "l"(m), "l"(pol_f)
						             : "memory");
					}
				}
			}
		}
	}
}

__global__ void __launch_bounds__(256) lab_stream(const uint64_t *__restrict__ keys, uint64_t n,
                                                  unsigned long long *out) {
	uint64_t pol_k;
	asm("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol_k));
	uint64_t acc = 0;
	const uint64_t tile = 256ull * 16;
	for (uint64_t base = blockIdx.x * tile; base + tile <= n; base += (uint64_t)gridDim.x * tile) {
		ulonglong2 k[8];
#pragma unroll
		for (int r = 0; r < 8; r++) {
			const uint64_t *p = keys + base + (threadIdx.x >> 5) * 512 + r * 64 + (threadIdx.x & 31) * 2;
			asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.v2.u64 {%0, %1}, [%2], %3;"
			             : "=l"(k[r].x), "=l"(k[r].y)
			             : "l"(p), "l"(pol_k));
		}
#pragma unroll
		for (int r = 0; r < 8; r++) {
			acc ^= k[r].x ^ k[r].y;
		}
	}
	if (acc == 0x1234567ull) {
		atomicAdd(out, 1ull);
	}
}

static float time_ms(cudaStream_t s, int reps, void (*fn)(void *), void *arg) {
	cudaEvent_t a, b;
	CK(cudaEventCreate(&a));
	CK(cudaEventCreate(&b));
	fn(arg);
	fn(arg);
	CK(cudaStreamSynchronize(s));
	CK(cudaEventRecord(a, s));
	for (int i = 0; i < reps; i++) {
		fn(arg);
	}
	CK(cudaEventRecord(b, s));
	CK(cudaStreamSynchronize(s));
	float ms;
	CK(cudaEventElapsedTime(&ms, a, b));
	return ms / reps;
}

struct Ctx;
static void run_slices(void *p);
static void run_build_slices(void *p);
static void run_stream(void *p);
struct Ctx {
	uint64_t slice_sectors = 0;
	int warm = 0;
	const uint64_t *keys;
	uint64_t *filter;
	uint64_t smask, n;
	unsigned long long *hits;
	cudaStream_t s;
	int grid;
	int variant, keypol;
};

static void run_probe(void *p) {
	Ctx *c = (Ctx *)p;
#define LP(P, K)                                                                                                       \
	lab_probe<P, K><<<c->grid, 256, 0, c->s>>>(c->keys, c->filter, c->smask, c->n, c->hits)
	switch (c->variant * 2 + c->keypol) {
	case 0: LP(0, 0); break;
	case 1: LP(0, 1); break;
	case 2: LP(1, 0); break;
	case 3: LP(1, 1); break;
	case 4: LP(2, 0); break;
	case 5: LP(2, 1); break;
	case 6: LP(3, 0); break;
	case 7: LP(3, 1); break;
	case 8: LP(4, 0); break;
	case 9: LP(4, 1); break;
	case 10: LP(5, 0); break;
	case 11: LP(5, 1); break;
	case 12: LP(6, 0); break;
	case 13: LP(6, 1); break;
	case 14: LP(7, 0); break;
	case 16: LP(8, 0); break;
	case 18: LP(9, 0); break;
	case 20: LP(10, 0); break;
	case 22: LP(11, 0); break;
	}
}
static void run_build(void *p) {
	Ctx *c = (Ctx *)p;
	unsigned long long *f = (unsigned long long *)c->filter;
	if (c->variant == 0) {
		lab_build<0><<<c->grid, 256, 0, c->s>>>(c->keys, f, c->smask, c->n);
	} else if (c->variant == 1) {
		lab_build<1><<<c->grid, 256, 0, c->s>>>(c->keys, f, c->smask, c->n);
	} else {
		lab_build<2><<<c->grid, 256, 0, c->s>>>(c->keys, f, c->smask, c->n);
	}
}

static void run_slices(void *p) {
	Ctx *c = (Ctx *)p;
	const uint64_t S = c->smask + 1;
	for (uint64_t lo = 0; lo < S; lo += c->slice_sectors) {
		const uint64_t hi = lo + c->slice_sectors < S ? lo + c->slice_sectors : S;
		if (c->warm) {
			lab_probe_slice<1><<<c->grid, 256, 0, c->s>>>(c->keys, c->filter, c->smask, c->n, lo, hi, c->hits);
		} else {
			lab_probe_slice<0><<<c->grid, 256, 0, c->s>>>(c->keys, c->filter, c->smask, c->n, lo, hi, c->hits);
		}
	}
}
static void run_build_slices(void *p) {
	Ctx *c = (Ctx *)p;
	const uint64_t S = c->smask + 1;
	unsigned long long *f = (unsigned long long *)c->filter;
	for (uint64_t lo = 0; lo < S; lo += c->slice_sectors) {
		const uint64_t hi = lo + c->slice_sectors < S ? lo + c->slice_sectors : S;
		if (c->variant == 0) {
			lab_build_slice<0><<<c->grid, 256, 0, c->s>>>(c->keys, f, c->smask, c->n, lo, hi);
		} else {
			lab_build_slice<1><<<c->grid, 256, 0, c->s>>>(c->keys, f, c->smask, c->n, lo, hi);
		}
	}
}
static void run_stream(void *p) {
	Ctx *c = (Ctx *)p;
	lab_stream<<<c->grid, 256, 0, c->s>>>(c->keys, c->n, c->hits);
}

int main(int argc, char **argv) {
	const int lg = argc > 1 ? atoi(argv[1]) : 30;
	const uint64_t n = 1ull << lg;
	cudaDeviceProp prop;
	CK(cudaGetDeviceProperties(&prop, 0));
	printf("device %s, %d SMs, L2 %d MiB, persistingL2CacheMaxSize %d MiB, accessPolicyMaxWindowSize %d MiB\n", prop.name,
	       prop.multiProcessorCount, prop.l2CacheSize >> 20, prop.persistingL2CacheMaxSize >> 20,
	       prop.accessPolicyMaxWindowSize >> 20);
	size_t gran = 0;
	CK(cudaDeviceGetLimit(&gran, cudaLimitMaxL2FetchGranularity));
	printf("default cudaLimitMaxL2FetchGranularity = %zu\n", gran);
	cudaStream_t s;
	CK(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
	uint64_t *keys, *filter;
	unsigned long long *hits;
	CK(cudaMalloc(&keys, n * 8));
	CK(cudaMalloc(&filter, 512ull << 20));
	CK(cudaMalloc(&hits, 8));
	gen_keys<<<148 * 8, 256, 0, s>>>(keys, n);
	fill_filter<<<148 * 8, 256, 0, s>>>(filter, (512ull << 20) / 8);
	CK(cudaStreamSynchronize(s));
	const bool quick = argc > 2;
	{
		Ctx c; c.keys = keys; c.filter = filter; c.smask = 0; c.n = n; c.hits = hits; c.s = s; c.grid = 148 * 5; c.variant = 0; c.keypol = 0;
		for (int g = 2; g <= 8; g += 3) {
			c.grid = 148 * g;
			const float ms = time_ms(s, 3, run_stream, &c);
			printf("stream read %llu MiB grid 148x%d: %7.3f ms  %7.1f GB/s\n", (unsigned long long)(n * 8 >> 20), g, ms,
			       n * 8 / ms / 1e6);
		}
		const int fmibs[] = {256, 512};
		const int slices[] = {32, 48, 64, 72, 80, 88, 96, 112, 128};
		for (int fi = 0; fi < (argc <= 3 ? 2 : 0); fi++) {
			for (int si = 0; si < 9; si++) {
				for (int warm = 0; warm < 2; warm++) {
					c.smask = ((uint64_t)fmibs[fi] << 20) / 8 - 1;
					c.slice_sectors = ((uint64_t)slices[si] << 20) / 8;
					c.warm = warm;
					c.grid = 148 * 5;
					const int passes = (fmibs[fi] + slices[si] - 1) / slices[si];
					const float ms = time_ms(s, 2, run_slices, &c);
					printf("sliced probe F=%d MiB slice %3d MiB (%d passes) warm=%d: %7.3f ms  %7.1f Gkeys/s\n", fmibs[fi],
					       slices[si], passes, warm, ms, n / ms / 1e6);
				}
			}
		}
		for (int fi = 0; fi < (argc <= 3 ? 2 : 0); fi++) {
			for (int si = 0; si < 9; si += 2) {
				for (int v = 0; v < 2; v++) {
					c.smask = ((uint64_t)fmibs[fi] << 20) / 8 - 1;
					c.slice_sectors = ((uint64_t)slices[si] << 20) / 8;
					c.n = n / 8;
					c.grid = 148 * 8;
					c.variant = v;
					const int passes = (fmibs[fi] + slices[si] - 1) / slices[si];
					const float ms = time_ms(s, 2, run_build_slices, &c);
					printf("sliced build F=%d MiB slice %3d MiB (%d passes) red%s: %7.3f ms  %7.1f Gkeys/s\n", fmibs[fi],
					       slices[si], passes, v ? "+evict_last" : "", ms, (n / 8) / ms / 1e6);
				}
			}
		}
		if (argc > 3) { // load-flavour scan on the 256 MiB filter
			const char *names[] = {"ld.cv", "ld.relaxed.gpu", "nc+evict_first", "atom.or 0", "nc+frac(evict_last 0.25 / evict_first)"};
			for (int v = 7; v <= 11; v++) {
				Ctx c2; c2.keys = keys; c2.filter = filter; c2.smask = (256ull << 20) / 8 - 1; c2.n = n; c2.hits = hits; c2.s = s; c2.grid = 148 * 5; c2.variant = v; c2.keypol = 0;
				const float ms = time_ms(s, 3, run_probe, &c2);
				printf("probe F= 256 MiB %-40s : %7.3f ms  %7.1f Gkeys/s\n", names[v - 7], ms, n / ms / 1e6);
			}
			return 0;
		}
		if (quick) {
			return 0;
		}
	}
	const char *vn[] = {"nc.noalloc+evict_last", "nc plain", "ld plain", "nc.noalloc", "nc.noalloc+evict_normal hint",
	                    "ld.cg", "nc.noalloc.L2::64B"};
	const size_t grans[] = {0, 32, 64, 128};
	for (int gi = 0; gi < 4; gi++) {
		if (grans[gi]) {
			cudaError_t e = cudaDeviceSetLimit(cudaLimitMaxL2FetchGranularity, grans[gi]);
			CK(cudaDeviceGetLimit(&gran, cudaLimitMaxL2FetchGranularity));
			printf("== set fetch granularity %zu -> %s, now %zu\n", grans[gi], cudaGetErrorString(e), gran);
		}
		for (int fm = 0; fm < 3; fm++) {
			const uint64_t fbytes = fm == 0 ? (16ull << 20) : fm == 1 ? (64ull << 20) : (256ull << 20);
			for (int v = 0; v < 7; v++) {
				for (int kp = 0; kp < 2; kp++) {
					if (gi > 0 && !(v == 0 || v == 3)) {
						continue;
					}
					if (kp == 1 && v != 0 && v != 3) {
						continue;
					}
					Ctx c; c.keys = keys; c.filter = filter; c.smask = fbytes / 8 - 1; c.n = n; c.hits = hits; c.s = s; c.grid = 148 * 5; c.variant = v; c.keypol = kp;
					const float ms = time_ms(s, 3, run_probe, &c);
					printf("probe F=%4llu MiB %-30s keypol=%d : %7.3f ms  %7.1f Gkeys/s\n",
					       (unsigned long long)(fbytes >> 20), vn[v], kp, ms, n / ms / 1e6);
				}
			}
		}
		for (int fm = 0; fm < 2; fm++) {
			const uint64_t fbytes = fm == 0 ? (16ull << 20) : (256ull << 20);
			for (int v = 0; v < 3; v++) {
				const uint64_t nb = n / 8;
				Ctx c; c.keys = keys; c.filter = filter; c.smask = fbytes / 8 - 1; c.n = nb; c.hits = hits; c.s = s; c.grid = 148 * 8; c.variant = v; c.keypol = 0;
				const float ms = time_ms(s, 3, run_build, &c);
				printf("build F=%4llu MiB variant %d (0 atomicOr,1 red,2 red+evict_last) : %7.3f ms  %7.1f Gkeys/s\n",
				       (unsigned long long)(fbytes >> 20), v, ms, nb / ms / 1e6);
			}
		}
	}
	// persisting-L2 access window over the filter (256 MiB case): keep a hitRatio fraction resident
	CK(cudaDeviceSetLimit(cudaLimitMaxL2FetchGranularity, 32));
	if (prop.persistingL2CacheMaxSize > 0) {
		CK(cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, prop.persistingL2CacheMaxSize));
		const float ratios[] = {0.2f, 0.3f, 0.4f, 0.6f, 1.0f};
		for (int ri = 0; ri < 5; ri++) {
			cudaStreamAttrValue attr;
			attr.accessPolicyWindow.base_ptr = filter;
			size_t win = 256ull << 20;
			if (win > (size_t)prop.accessPolicyMaxWindowSize) {
				win = prop.accessPolicyMaxWindowSize;
			}
			attr.accessPolicyWindow.num_bytes = win;
			attr.accessPolicyWindow.hitRatio = ratios[ri];
			attr.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting;
			attr.accessPolicyWindow.missProp = cudaAccessPropertyStreaming;
			CK(cudaStreamSetAttribute(s, cudaStreamAttributeAccessPolicyWindow, &attr));
			Ctx c; c.keys = keys; c.filter = filter; c.smask = (256ull << 20) / 8 - 1; c.n = n; c.hits = hits; c.s = s; c.grid = 148 * 5; c.variant = 3; c.keypol = 1;
			const float ms = time_ms(s, 3, run_probe, &c);
			printf("probe F= 256 MiB window %zu MiB hitRatio %.1f (nc.noalloc, gran 32): %7.3f ms  %7.1f Gkeys/s\n",
			       win >> 20, ratios[ri], ms, n / ms / 1e6);
		}
		CK(cudaCtxResetPersistingL2Cache());
	}
	return 0;
}
