/*
 * rpt_oracle.c — TEST INFRASTRUCTURE, NOT PRODUCT.  See rpt_oracle.h for scope and parity status
 * (PINNED against SURVEY §8(c) known answers and oracle/_ref built from the reference sources).
 * Plain C restatement of the reference algorithm; citations are relative to /root/reference.
 */
#include "rpt_oracle.h"

#include <math.h>
#include <pthread.h>
#include <stdatomic.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#define MURMUR_K 0xd6e8feb86659fd93ULL

/* duckdb/src/include/duckdb/common/types/hash.hpp:36-43 */
uint64_t rpto_murmur64(uint64_t x) {
	x ^= x >> 32;
	x *= MURMUR_K;
	x ^= x >> 32;
	x *= MURMUR_K;
	x ^= x >> 32;
	return x;
}

/* duckdb/src/common/types/hash.cpp:77-112 (HashBytes) and :115-151 (string_t; the inlined
 * branch is an optimisation that yields the same value, asserted at :143-145). */
uint64_t rpto_hash_bytes(const uint8_t *p, uint64_t len) {
	uint64_t h = 0xe17a1465ULL ^ (len * 0xc6a4a7935bd1e995ULL);
	uint64_t rem = len & 7ULL;
	const uint8_t *end = p + (len - rem);
	for (; p != end; p += 8) {
		uint64_t blk;
		memcpy(&blk, p, 8); /* LoadLE on a little-endian host */
		h ^= blk;
		h *= MURMUR_K;
	}
	if (rem) {
		uint64_t hr = 0;
		memcpy(&hr, p, rem);
		h ^= hr;
		h *= MURMUR_K;
	}
	return rpto_murmur64(h);
}

/* duckdb/src/include/duckdb/common/types/interval.hpp:178-193 */
static void interval_normalize(int32_t in_months, int32_t in_days, int64_t in_micros, int64_t *months,
                               int64_t *days, int64_t *micros) {
	const int64_t MICROS_PER_DAY = 86400000000LL;
	const int64_t DAYS_PER_MONTH = 30;
	*micros = in_micros;
	int64_t carry_days = *micros / MICROS_PER_DAY;
	*micros -= carry_days * MICROS_PER_DAY;
	*days = in_days;
	*days += carry_days;
	int64_t carry_months = *days / DAYS_PER_MONTH;
	*days -= carry_months * DAYS_PER_MONTH;
	*months = in_months;
	*months += carry_months;
}

/* hash.hpp:49-61 (generic template: static_cast<uint32_t>), hash.cpp:22-70 (specialisations) */
uint64_t rpto_hash_value(int type, const void *v) {
	switch (type) {
	case RPTO_BOOL:
	case RPTO_INT8: /* vector_hash.cpp:285-287: BOOL is hashed as int8_t */
		return rpto_murmur64((uint64_t)(uint32_t)(int32_t) * (const int8_t *)v);
	case RPTO_UINT8:
		return rpto_murmur64((uint64_t)(uint32_t) * (const uint8_t *)v);
	case RPTO_INT16:
		return rpto_murmur64((uint64_t)(uint32_t)(int32_t) * (const int16_t *)v);
	case RPTO_UINT16:
		return rpto_murmur64((uint64_t)(uint32_t) * (const uint16_t *)v);
	case RPTO_INT32:
		return rpto_murmur64((uint64_t)(uint32_t) * (const int32_t *)v);
	case RPTO_UINT32:
		return rpto_murmur64((uint64_t) * (const uint32_t *)v);
	case RPTO_INT64:
	case RPTO_UINT64:
		return rpto_murmur64(*(const uint64_t *)v);
	case RPTO_INT128:
	case RPTO_UINT128: { /* hash.cpp:22-30 */
		const uint64_t *w = (const uint64_t *)v;
		return rpto_murmur64(w[0]) ^ rpto_murmur64(w[1]);
	}
	case RPTO_FLOAT: { /* hash.cpp:32-50 */
		float f;
		memcpy(&f, v, 4);
		if (f == 0.0f) {
			f = 0.0f;
		} else if (isnan(f)) {
			uint32_t q = 0x7fc00000u; /* std::numeric_limits<float>::quiet_NaN() */
			memcpy(&f, &q, 4);
		}
		uint32_t u;
		memcpy(&u, &f, 4);
		return rpto_murmur64((uint64_t)u);
	}
	case RPTO_DOUBLE: { /* hash.cpp:52-58 */
		double d;
		memcpy(&d, v, 8);
		if (d == 0.0) {
			d = 0.0;
		} else if (isnan(d)) {
			uint64_t q = 0x7ff8000000000000ULL;
			memcpy(&d, &q, 8);
		}
		uint64_t u;
		memcpy(&u, &d, 8);
		return rpto_murmur64(u);
	}
	case RPTO_INTERVAL: { /* hash.cpp:60-65 */
		int32_t m, d;
		int64_t us;
		memcpy(&m, v, 4);
		memcpy(&d, (const char *)v + 4, 4);
		memcpy(&us, (const char *)v + 8, 8);
		int64_t months, days, micros;
		interval_normalize(m, d, us, &months, &days, &micros);
		return rpto_murmur64((uint64_t)days) ^ rpto_murmur64((uint64_t)months) ^ rpto_murmur64((uint64_t)micros);
	}
	default:
		return 0;
	}
}

/* vector_hash.cpp:37-41 */
uint64_t rpto_combine(uint64_t a, uint64_t b) {
	a ^= a >> 32;
	a *= MURMUR_K;
	return a ^ b;
}

/* duckdb/src/common/constants.cpp:30-47; the reference throws when the result wraps to 0 */
uint64_t rpto_next_pow2(uint64_t v) {
	if (v < 1) {
		return 2;
	}
	v--;
	v |= v >> 1;
	v |= v >> 2;
	v |= v >> 4;
	v |= v >> 8;
	v |= v >> 16;
	v |= v >> 32;
	v++;
	return v;
}

/* duckdb/src/planner/filter/bloom_filter.cpp:7-9,17-19 */
uint64_t rpto_num_sectors(uint64_t rows) {
	uint64_t min_bits = rows * 12ULL;
	if (min_bits < 512ULL) {
		min_bits = 512ULL;
	}
	uint64_t s = rpto_next_pow2(min_bits) >> 6;
	const uint64_t cap = 1ULL << 26;
	return s < cap ? s : cap;
}

/* duckdb/src/planner/filter/bloom_filter.cpp:29-41: bytes 4..7 of (h & 0x3F3F..) pick 4 bit positions */
uint64_t rpto_mask(uint64_t h) {
	uint64_t shifts = h & 0x3F3F3F3F3F3F3F3FULL;
	uint64_t mask = 0;
	for (int b = 4; b < 8; b++) {
		mask |= 1ULL << ((shifts >> (8 * b)) & 0xFF);
	}
	return mask;
}

/* src/operators/physical_create_bf.cpp:386-396 */
int rpto_needs_rebuild(uint64_t sized_for, uint64_t actual) {
	if (actual == 0) {
		return 0;
	}
	uint64_t min_bits = sized_for * 12ULL;
	if (min_bits < 512ULL) {
		min_bits = 512ULL;
	}
	uint64_t allocated_bits = rpto_next_pow2(min_bits);
	return actual * 8ULL > allocated_bits;
}

static uint64_t width_of(int type) {
	switch (type) {
	case RPTO_BOOL: case RPTO_INT8: case RPTO_UINT8: return 1;
	case RPTO_INT16: case RPTO_UINT16: return 2;
	case RPTO_INT32: case RPTO_UINT32: case RPTO_FLOAT: return 4;
	case RPTO_INT64: case RPTO_UINT64: case RPTO_DOUBLE: return 8;
	case RPTO_INTERVAL: case RPTO_INT128: case RPTO_UINT128: return 16;
	default: return 0;
	}
}

/* one row of TightLoopHash (vector_hash.cpp:43-61): NULL rows take NULL_HASH, they are not skipped */
static uint64_t hash_row(const rpto_col_t *c, uint64_t i) {
	uint64_t idx = c->sel ? c->sel[i] : i;
	if (c->validity && !((c->validity[idx >> 6] >> (idx & 63)) & 1ULL)) {
		return RPTO_NULL_HASH;
	}
	if (c->type == RPTO_VARCHAR) {
		const uint32_t *off = (const uint32_t *)c->data;
		return rpto_hash_bytes((const uint8_t *)c->aux + off[idx], (uint64_t)(off[idx + 1] - off[idx]));
	}
	return rpto_hash_value(c->type, (const char *)c->data + idx * width_of(c->type));
}

void rpto_hash_column(const rpto_col_t *col, uint64_t count, uint64_t *out, int combine) {
	for (uint64_t i = 0; i < count; i++) {
		uint64_t h = hash_row(col, i);
		out[i] = combine ? rpto_combine(out[i], h) : h;
	}
}

/* src/bloom_filter.cpp:11-24 */
void rpto_hash_columns(const rpto_col_t *cols, int ncols, uint64_t count, uint64_t *out) {
	rpto_hash_column(&cols[0], count, out, 0);
	for (int j = 1; j < ncols; j++) {
		rpto_hash_column(&cols[j], count, out, 1);
	}
}

/* bloom_filter.cpp:43-48 + :63-70 (relaxed fetch_or) */
void rpto_insert_hashes(uint64_t *words, uint64_t S, const uint64_t *h, uint64_t n) {
	const uint64_t bitmask = S - 1;
	for (uint64_t i = 0; i < n; i++) {
		atomic_fetch_or_explicit((_Atomic uint64_t *)&words[h[i] & bitmask], rpto_mask(h[i]), memory_order_relaxed);
	}
}

/* bloom_filter.cpp:50-61 + :72-78: sel[found] is written unconditionally, found advances on a hit */
uint64_t rpto_lookup_hashes(const uint64_t *words, uint64_t S, const uint64_t *h, uint64_t n, uint32_t *sel) {
	const uint64_t bitmask = S - 1;
	uint64_t found = 0;
	for (uint64_t i = 0; i < n; i++) {
		uint64_t m = rpto_mask(h[i]);
		sel[found] = (uint32_t)i;
		found += (words[h[i] & bitmask] & m) == m;
	}
	return found;
}

#define CHUNK 2048 /* STANDARD_VECTOR_SIZE, duckdb/src/include/duckdb/common/vector_size.hpp:16 */

/* hashes rows [base, base+n) of the columns (absolute validity addressing) */
static void hash_rows(const rpto_col_t *cols, int ncols, uint64_t base, uint64_t n, uint64_t *out) {
	for (uint64_t i = 0; i < n; i++) {
		uint64_t h = hash_row(&cols[0], base + i);
		for (int j = 1; j < ncols; j++) {
			h = rpto_combine(h, hash_row(&cols[j], base + i));
		}
		out[i] = h;
	}
}

/* src/bloom_filter.cpp:70-78, chunk by chunk like PhysicalCreateBF::Sink */
void rpto_insert(uint64_t *words, uint64_t S, const rpto_col_t *cols, int ncols, uint64_t count) {
	uint64_t hashes[CHUNK];
	for (uint64_t base = 0; base < count; base += CHUNK) {
		uint64_t n = count - base < CHUNK ? count - base : CHUNK;
		hash_rows(cols, ncols, base, n, hashes);
		rpto_insert_hashes(words, S, hashes, n);
	}
}

/* src/bloom_filter.cpp:60-68, chunk by chunk; chunk-local indices are rebased to the batch */
uint64_t rpto_probe(const uint64_t *words, uint64_t S, const rpto_col_t *cols, int ncols, uint64_t count,
                    uint32_t *sel_out) {
	uint64_t hashes[CHUNK];
	uint32_t sel[CHUNK + 1];
	uint64_t total = 0;
	for (uint64_t base = 0; base < count; base += CHUNK) {
		uint64_t n = count - base < CHUNK ? count - base : CHUNK;
		hash_rows(cols, ncols, base, n, hashes);
		uint64_t f = rpto_lookup_hashes(words, S, hashes, n, sel);
		for (uint64_t k = 0; k < f; k++) {
			sel_out[total + k] = (uint32_t)(base + sel[k]);
		}
		total += f;
	}
	return total;
}

/* BFTableFilter::CheckStatistics / TemplatedCheckStatistics<T> (duckdb/src/planner/filter/bloom_filter.cpp:138-196):
 * zone-map pruning of one row group with statistics [min, max] of an integral physical type.  T-typed
 * subtraction (TrySubtractOperator: the difference has to fit T, so an int8 range of 128..255 is "overflow"),
 * at most 2049 consecutive values are looked up.  Results: 0 = NO_PRUNING_POSSIBLE, 1 = FILTER_ALWAYS_TRUE,
 * 2 = FILTER_ALWAYS_FALSE.  min/max point at one value of the type (16 bytes {lower, upper} for the 128-bit ones). */
static int lookup_one(const uint64_t *words, uint64_t S, uint64_t h) { /* bloom_filter.cpp:72-78 */
	const uint64_t m = rpto_mask(h);
	return (words[h & (S - 1)] & m) == m;
}

int rpto_check_statistics(const uint64_t *words, uint64_t S, int type, const void *min, const void *max,
                          int has_minmax) {
	if (!has_minmax) {
		return 0;
	}
	uint64_t range;
	switch (type) {
#define RPTO_SMALL(TAG, T, WIDE, TMAX)                                                                                  \
	case TAG: {                                                                                                    \
		const T lo = *(const T *)min, hi = *(const T *)max;                                                        \
		if (lo > hi) {                                                                                             \
			return 0;                                                                                              \
		}                                                                                                          \
		const WIDE d = (WIDE)hi - (WIDE)lo;                                                                        \
		if (d > (WIDE)(TMAX) || d > 2048) {                                                                        \
			return 0;                                                                                              \
		}                                                                                                          \
		range = (uint64_t)d;                                                                                       \
		uint64_t hits = 0;                                                                                         \
		T val = lo;                                                                                                \
		for (uint64_t i = 0; i <= range; i++) {                                                                    \
			hits += (uint64_t)lookup_one(words, S, rpto_hash_value(TAG, &val));                                    \
			val = (T)(val + (T)(i < range));                                                                       \
		}                                                                                                          \
		return hits == 0 ? 2 : (hits == range + 1 ? 1 : 0);                                                        \
	}
		RPTO_SMALL(RPTO_UINT8, uint8_t, int64_t, UINT8_MAX)
		RPTO_SMALL(RPTO_UINT16, uint16_t, int64_t, UINT16_MAX)
		RPTO_SMALL(RPTO_UINT32, uint32_t, int64_t, UINT32_MAX)
		RPTO_SMALL(RPTO_INT8, int8_t, int64_t, INT8_MAX)
		RPTO_SMALL(RPTO_INT16, int16_t, int64_t, INT16_MAX)
		RPTO_SMALL(RPTO_INT32, int32_t, int64_t, INT32_MAX)
		RPTO_SMALL(RPTO_UINT64, uint64_t, __int128, UINT64_MAX)
		RPTO_SMALL(RPTO_INT64, int64_t, __int128, INT64_MAX)
#undef RPTO_SMALL
	case RPTO_UINT128:
	case RPTO_INT128: {
		const uint64_t *a = (const uint64_t *)min, *b = (const uint64_t *)max;
		unsigned __int128 lo = ((unsigned __int128)a[1] << 64) | a[0], hi = ((unsigned __int128)b[1] << 64) | b[0];
		unsigned __int128 d;
		if (type == RPTO_INT128) {
			if ((__int128)lo > (__int128)hi) {
				return 0;
			}
			/* hi - lo as hugeint: overflow when the true difference exceeds 2^127 - 1 */
			d = hi - lo;
			if ((d >> 127) != 0) {
				return 0;
			}
		} else {
			if (lo > hi) {
				return 0;
			}
			d = hi - lo;
		}
		if (d > 2048) {
			return 0;
		}
		range = (uint64_t)d;
		uint64_t hits = 0;
		unsigned __int128 val = lo;
		for (uint64_t i = 0; i <= range; i++) {
			uint64_t w[2] = {(uint64_t)val, (uint64_t)(val >> 64)};
			hits += (uint64_t)lookup_one(words, S, rpto_hash_value(type, w));
			val += (unsigned __int128)(i < range);
		}
		return hits == 0 ? 2 : (hits == range + 1 ? 1 : 0);
	}
	default:
		return 0; /* bloom_filter.cpp:193-195: every other physical type */
	}
}

uint64_t rpto_fnv1a_words(const uint64_t *w, uint64_t n) {
	uint64_t h = 0xcbf29ce484222325ULL;
	for (uint64_t i = 0; i < n; i++) {
		h ^= w[i];
		h *= 0x100000001b3ULL;
	}
	return h;
}

/* SURVEY §8(d) */
uint64_t rpto_splitmix64(uint64_t x) {
	uint64_t z = x + 0x9e3779b97f4a7c15ULL;
	z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ULL;
	z = (z ^ (z >> 27)) * 0x94d049bb133111ebULL;
	return z ^ (z >> 31);
}

void rpto_synth_build_keys(uint64_t *out, uint64_t first, uint64_t count) {
	for (uint64_t i = 0; i < count; i++) {
		out[i] = rpto_splitmix64(0xB10F0001ULL + first + i) & 0x7fffffffffffffffULL;
	}
}

void rpto_synth_probe_keys(uint64_t *out, uint64_t first, uint64_t count, uint64_t n_build, uint32_t sel_ppm) {
	for (uint64_t i = 0; i < count; i++) {
		uint64_t j = first + i;
		uint64_t a = rpto_splitmix64(0x9B0BE001ULL + 2 * j);
		uint64_t b = rpto_splitmix64(0x9B0BE001ULL + 2 * j + 1);
		if (a % 1000000ULL < sel_ppm) {
			out[i] = rpto_splitmix64(0xB10F0001ULL + (b % n_build)) & 0x7fffffffffffffffULL;
		} else {
			out[i] = b | 0x8000000000000000ULL;
		}
	}
}

/* ---- threaded CPU baseline ------------------------------------------------------------------ */
typedef struct {
	uint64_t *words;
	uint64_t S;
	const int64_t *keys;
	uint64_t lo, hi;
	uint64_t found;
	int probe;
} bench_arg_t;

static void *bench_worker(void *p) {
	bench_arg_t *a = (bench_arg_t *)p;
	uint64_t hashes[CHUNK];
	uint32_t sel[CHUNK + 1];
	uint64_t found = 0;
	for (uint64_t base = a->lo; base < a->hi; base += CHUNK) {
		uint64_t n = a->hi - base < CHUNK ? a->hi - base : CHUNK;
		for (uint64_t i = 0; i < n; i++) {
			hashes[i] = rpto_murmur64((uint64_t)a->keys[base + i]);
		}
		if (a->probe) {
			found += rpto_lookup_hashes(a->words, a->S, hashes, n, sel);
		} else {
			rpto_insert_hashes(a->words, a->S, hashes, n);
		}
	}
	a->found = found;
	return NULL;
}

static double now_s(void) {
	struct timespec ts;
	clock_gettime(CLOCK_MONOTONIC, &ts);
	return (double)ts.tv_sec + 1e-9 * (double)ts.tv_nsec;
}

static double bench_run(uint64_t *words, uint64_t S, const int64_t *keys, uint64_t n, int threads, int probe,
                        uint64_t *found_out) {
	if (threads < 1) {
		threads = 1;
	}
	pthread_t *tid = (pthread_t *)malloc(sizeof(pthread_t) * (size_t)threads);
	bench_arg_t *args = (bench_arg_t *)malloc(sizeof(bench_arg_t) * (size_t)threads);
	uint64_t chunks = (n + CHUNK - 1) / CHUNK;
	double t0 = now_s();
	for (int t = 0; t < threads; t++) {
		uint64_t c0 = chunks * (uint64_t)t / (uint64_t)threads;
		uint64_t c1 = chunks * (uint64_t)(t + 1) / (uint64_t)threads;
		args[t].words = words;
		args[t].S = S;
		args[t].keys = keys;
		args[t].lo = c0 * CHUNK;
		args[t].hi = c1 * CHUNK < n ? c1 * CHUNK : n;
		args[t].found = 0;
		args[t].probe = probe;
		pthread_create(&tid[t], NULL, bench_worker, &args[t]);
	}
	uint64_t found = 0;
	for (int t = 0; t < threads; t++) {
		pthread_join(tid[t], NULL);
		found += args[t].found;
	}
	double t1 = now_s();
	if (found_out) {
		*found_out = found;
	}
	free(tid);
	free(args);
	return t1 - t0;
}

double rpto_bench_build_i64(uint64_t *words, uint64_t S, const int64_t *keys, uint64_t n, int threads) {
	return bench_run(words, S, keys, n, threads, 0, NULL);
}

double rpto_bench_probe_i64(const uint64_t *words, uint64_t S, const int64_t *keys, uint64_t n, int threads,
                            uint64_t *found_out) {
	return bench_run((uint64_t *)words, S, keys, n, threads, 1, found_out);
}
