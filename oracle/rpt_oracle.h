/*
 * rpt_oracle.h — TEST INFRASTRUCTURE, NOT PRODUCT.
 *
 * CPU restatement (plain C) of the reference's bloom-filter hot path, used only as the checker in
 * tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs.  Nothing in
 * the product (librpt_gpu.so, the package) includes, links or calls this.
 *
 * Parity status: PINNED.  Every function here is checked (tests/test_oracle_cpu.py) against
 *   (1) the known answers recorded from the compiled reference in SURVEY.md §8(c)
 *       (tests/golden/known_answers.json), and
 *   (2) oracle/_ref/librpt_ref.so, which compiles the reference's own sources
 *       (duckdb/src/planner/filter/bloom_filter.cpp, duckdb/src/common/types/hash.cpp,
 *        duckdb/src/common/vector_operations/vector_hash.cpp, duckdb/src/common/constants.cpp)
 *       where they lie — see oracle/ref_harness.cpp and oracle/Makefile — plus the golden vectors
 *       that library generated (tests/golden/ref_vectors.json, made by tests/golden/make_golden.py).
 *
 * All file:line citations are relative to /root/reference.
 */
#ifndef RPT_ORACLE_H
#define RPT_ORACLE_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

/* same numeric values as duckdb::PhysicalType (duckdb/src/include/duckdb/common/types.hpp:65-174) */
enum rpto_type {
	RPTO_BOOL = 1, RPTO_UINT8 = 2, RPTO_INT8 = 3, RPTO_UINT16 = 4, RPTO_INT16 = 5, RPTO_UINT32 = 6,
	RPTO_INT32 = 7, RPTO_UINT64 = 8, RPTO_INT64 = 9, RPTO_FLOAT = 11, RPTO_DOUBLE = 12,
	RPTO_INTERVAL = 21, RPTO_VARCHAR = 200, RPTO_UINT128 = 203, RPTO_INT128 = 204
};

/* pieces of a UnifiedVectorFormat; binary-compatible with rpt_col_t so tests can share buffers */
typedef struct rpto_col {
	const void *data;
	const void *aux;
	const uint32_t *sel;
	const uint64_t *validity;
	uint64_t data_len;
	uint64_t aux_bytes;
	int32_t type;
	int32_t mem;
} rpto_col_t;

#define RPTO_NULL_HASH 0xbf58476d1ce4e5b9ULL /* vector_hash.cpp:17 */

uint64_t rpto_murmur64(uint64_t x);                          /* hash.hpp:36-43 */
uint64_t rpto_hash_bytes(const uint8_t *p, uint64_t len);    /* hash.cpp:77-151 */
uint64_t rpto_hash_value(int type, const void *value);       /* hash.hpp:49-61, hash.cpp:22-70 */
uint64_t rpto_combine(uint64_t a, uint64_t b);               /* vector_hash.cpp:37-41 */
uint64_t rpto_next_pow2(uint64_t v);                         /* constants.cpp:30-47 (0 on overflow) */
uint64_t rpto_num_sectors(uint64_t rows);                    /* bloom_filter.cpp:7-9,17-19 */
uint64_t rpto_mask(uint64_t h);                              /* bloom_filter.cpp:29-41 */
int rpto_needs_rebuild(uint64_t sized_for, uint64_t actual); /* physical_create_bf.cpp:386-396 */

/* VectorOperations::Hash / CombineHash over a column (vector_hash.cpp:43-61, 364-385):
 * combine == 0 writes out[i] = hash(row i); combine != 0 does out[i] = CombineHashScalar(out[i], hash). */
void rpto_hash_column(const rpto_col_t *col, uint64_t count, uint64_t *out, int combine);
/* HashColumns (src/bloom_filter.cpp:11-24) over ncols columns */
void rpto_hash_columns(const rpto_col_t *cols, int ncols, uint64_t count, uint64_t *out);

void rpto_insert_hashes(uint64_t *words, uint64_t S, const uint64_t *h, uint64_t n);     /* bloom_filter.cpp:43-48,63-70 */
uint64_t rpto_lookup_hashes(const uint64_t *words, uint64_t S, const uint64_t *h, uint64_t n,
                            uint32_t *sel);                                              /* bloom_filter.cpp:50-61,72-78 */

/* PTBloomFilter::Insert / LookupSel over whole columns in 2048-row chunks (src/bloom_filter.cpp:60-78) */
void rpto_insert(uint64_t *words, uint64_t S, const rpto_col_t *cols, int ncols, uint64_t count);
uint64_t rpto_probe(const uint64_t *words, uint64_t S, const rpto_col_t *cols, int ncols, uint64_t count,
                    uint32_t *sel_out);

/* BFTableFilter::CheckStatistics (duckdb bloom_filter.cpp:138-196) for one [min, max]: 0 no pruning, 1 always true,
 * 2 always false */
int rpto_check_statistics(const uint64_t *words, uint64_t S, int type, const void *min, const void *max,
                          int has_minmax);

/* FNV-1a-64 over words (xor then multiply per word) — the checksum SURVEY §8(c) froze */
uint64_t rpto_fnv1a_words(const uint64_t *w, uint64_t n);

/* synthetic generators of SURVEY §8(d) */
uint64_t rpto_splitmix64(uint64_t x);
void rpto_synth_build_keys(uint64_t *out, uint64_t first, uint64_t count);
void rpto_synth_probe_keys(uint64_t *out, uint64_t first, uint64_t count, uint64_t n_build, uint32_t sel_ppm);

/* Threaded CPU baseline: T threads insert into ONE shared filter over disjoint row ranges in
 * 2048-row chunks with relaxed atomic fetch_or (the PhysicalCreateBF::Sink pattern,
 * physical_create_bf.cpp:220-227); probe = T threads with private selection buffers
 * (physical_use_bf.cpp:163).  int64 flat keys.  Return seconds; probe returns total found. */
double rpto_bench_build_i64(uint64_t *words, uint64_t S, const int64_t *keys, uint64_t n, int threads);
double rpto_bench_probe_i64(const uint64_t *words, uint64_t S, const int64_t *keys, uint64_t n, int threads,
                            uint64_t *found_out);

#ifdef __cplusplus
}
#endif
#endif
