/*
 * rpt_gpu.h — C ABI of librpt_gpu.so, the B200 (sm_100a) bloom-filter engine behind the
 * `rpt` DuckDB extension's CREATE_BF / USE_BF operators.
 *
 * This is the drop-in boundary: the narrow waist of the reference is the four-method
 * `PTBloomFilter` class (reference src/include/bloom_filter.hpp:22-57, body src/bloom_filter.cpp:27-78)
 * wrapping DuckDB's native `BloomFilter` (duckdb/src/planner/filter/bloom_filter.cpp:7-78).
 * Every entry point below names the reference interface it replaces. The binding a maintainer
 * adds on the reference side is shown in INTEGRATION.md.
 *
 * Conventions
 *   - plain C, no C++ or torch types; every function returns 0 on success, non-zero on error;
 *     the message is in rpt_gpu_last_error() (thread-local).  The reference throws C++
 *     exceptions (OutOfRangeException from NextPowerOfTwo, allocator OOM); the C++ wrapper
 *     re-throws on non-zero status.
 *   - there is NO CPU fallback: with no usable CUDA device every compute entry point fails.
 *   - thread-safety: inserts may run concurrently on one handle (like N sink threads sharing one
 *     filter through relaxed fetch_or, bloom_filter.cpp:63-70); probes may run concurrently with
 *     each other; a probe never overlaps a build on the same handle (the reference guarantees this
 *     by pipeline dependency, physical_use_bf.cpp:200-229).
 *   - filter words are layout-identical to DuckDB's `uint64_t bf[num_sectors]`.
 */
#ifndef RPT_GPU_H
#define RPT_GPU_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define RPT_API __attribute__((visibility("default")))
#else
#define RPT_API
#endif

/* Opaque filter handle: one per build column, owned like the reference's
 * shared_ptr<PTBloomFilter> in PhysicalCreateBF::bloom_filter_map (physical_create_bf.hpp:73). */
typedef struct rpt_bf rpt_bf_t;

/* A cudaStream_t passed as void*.  NULL = the handle's own non-blocking stream. */
typedef void *rpt_stream_t;

/* Physical key types.  Numeric values equal duckdb::PhysicalType
 * (duckdb/src/include/duckdb/common/types.hpp:65-174) so the binding is a cast of
 * `vec.GetType().InternalType()`.  Hash semantics per type: hash.hpp:36-61, hash.cpp:22-151. */
enum rpt_type {
	RPT_BOOL = 1,
	RPT_UINT8 = 2,
	RPT_INT8 = 3,
	RPT_UINT16 = 4,
	RPT_INT16 = 5,
	RPT_UINT32 = 6,
	RPT_INT32 = 7,
	RPT_UINT64 = 8,
	RPT_INT64 = 9,
	RPT_FLOAT = 11,
	RPT_DOUBLE = 12,
	RPT_INTERVAL = 21, /* {int32 months, int32 days, int64 micros}, 16 bytes */
	RPT_VARCHAR = 200, /* data = uint32 offsets[data_len+1], aux = bytes (Arrow utf8 layout) */
	RPT_UINT128 = 203, /* {uint64 lower, uint64 upper} */
	RPT_INT128 = 204,  /* {uint64 lower, int64 upper} */
	RPT_STRING_T = 1200 /* HOST ONLY: data = duckdb::string_t[data_len] (16 B each, string_type.hpp:24);
	                       the library gathers it into the RPT_VARCHAR layout before upload */
};

enum rpt_mem { RPT_MEM_HOST = 0, RPT_MEM_DEVICE = 1 };

#define RPT_MAX_KEY_COLS 8

/* One key column = the pieces of a duckdb::UnifiedVectorFormat
 * (what VectorOperations::Hash iterates over, vector_hash.cpp:43-61):
 * row i of the call reads entry idx = sel ? sel[i] : i of `data`; the row is NULL when
 * `validity` is non-NULL and bit idx of it is clear; NULL rows hash to NULL_HASH
 * (vector_hash.cpp:16-23) and are inserted / probed like any value. */
typedef struct rpt_col {
	const void *data;         /* values; for RPT_VARCHAR the uint32 offsets; for RPT_STRING_T the string_t array */
	const void *aux;          /* RPT_VARCHAR: string bytes; otherwise NULL */
	const uint32_t *sel;      /* optional selection (dictionary / sliced vectors), `count` entries */
	const uint64_t *validity; /* optional validity bitmap over data entries, bit set = valid */
	uint64_t data_len;        /* addressable entries in data/validity; 0 means `count` (flat vector) */
	uint64_t aux_bytes;       /* RPT_VARCHAR: size of the bytes buffer; device buffers must be readable up to
	                             the next multiple of 8 bytes (cudaMalloc'd buffers always are) */
	int32_t type;             /* enum rpt_type */
	int32_t mem;              /* enum rpt_mem: where data/aux/sel/validity live (all in the same space) */
} rpt_col_t;

/* ---- library ---------------------------------------------------------------------------- */
RPT_API const char *rpt_gpu_last_error(void);
RPT_API const char *rpt_gpu_version(void);
RPT_API int rpt_gpu_device_count(int *out);
/* Pinned host memory for full-rate overlapped transfers (any host pointer is accepted by the
 * host-memory entry points; pageable memory is staged through the driver and is slower). */
RPT_API int rpt_host_alloc(void **out, uint64_t bytes);
RPT_API int rpt_host_free(void *ptr);

/* ---- lifetime: PTBloomFilter::Initialize (src/bloom_filter.cpp:27-32) →
 *      BloomFilter::Initialize (duckdb bloom_filter.cpp:14-27) ------------------------------ */
/* est_rows is what the reference passes AFTER its uint32_t truncation (bloom_filter.hpp:25):
 * the caller wraps, the library sizes: S = min(NextPow2(max(512, 12*est_rows)) >> 6, 2^26). */
RPT_API int rpt_bf_create(rpt_bf_t **out, uint64_t est_rows, int device);
RPT_API int rpt_bf_destroy(rpt_bf_t *bf);
/* PTBloomFilter::Initialize called again on a live object (what ReinitializeAndRehash does first,
 * src/bloom_filter.cpp:38-40): re-size for est_rows, zero the words, clear has_data_ and finalized_.
 * When S is unchanged nothing is reallocated and the zero-fill is asynchronous on `stream`. */
RPT_API int rpt_bf_reset(rpt_bf_t *bf, uint64_t est_rows, rpt_stream_t stream);
RPT_API int rpt_bf_num_sectors(const rpt_bf_t *bf, uint64_t *out);    /* BloomFilter::num_sectors */
RPT_API int rpt_bf_sized_for_rows(const rpt_bf_t *bf, uint64_t *out); /* PTBloomFilter::SizedForRows */
RPT_API int rpt_bf_is_empty(const rpt_bf_t *bf, int *out);            /* PTBloomFilter::IsEmpty (!has_data_) */
RPT_API int rpt_bf_is_finalized(const rpt_bf_t *bf, int *out);        /* PTBloomFilter::finalized_ */
RPT_API int rpt_bf_device(const rpt_bf_t *bf, int *out);
/* Sizing helper, no device needed: S for a row count (BloomFilter::Initialize arithmetic). */
RPT_API int rpt_bf_sectors_for_rows(uint64_t rows, uint64_t *out);

/* ---- build: PTBloomFilter::Insert (src/bloom_filter.cpp:70-78) = HashColumns (:11-24) +
 *      BloomFilter::InsertHashes (duckdb bloom_filter.cpp:43-48, InsertOne :63-70) ---------- */
/* Hashes `ncols` key columns row-wise (first column Hash, further columns CombineHash) and ORs
 * GetMask(h) into word h & (S-1).  count == 0 is a no-op that leaves has_data_ unchanged.
 * Device columns: asynchronous on `stream`.  Host columns: chunked H2D pipeline, returns when the
 * last chunk's insert has been enqueued and the host buffers are no longer needed. */
RPT_API int rpt_bf_insert(rpt_bf_t *bf, const rpt_col_t *cols, int ncols, uint64_t count, rpt_stream_t stream);

/* ---- finalize: resize rule of PhysicalCreateBF::Finalize (physical_create_bf.cpp:386-406) +
 *      PTBloomFilter::ReinitializeAndRehash (src/bloom_filter.cpp:34-58) -------------------- */
/* out = 1 iff actual_rows > 0 and actual_rows*8 > NextPow2(max(512, 12*sized_for_rows)). */
RPT_API int rpt_bf_needs_rebuild(const rpt_bf_t *bf, uint64_t actual_rows, int *out);
/* Re-size for `actual_rows` (new S), zero the words, has_data_ = (count > 0), re-insert the given
 * whole-collection key columns (count rows).  Exactly ReinitializeAndRehash. */
RPT_API int rpt_bf_reinitialize_and_rehash(rpt_bf_t *bf, uint64_t actual_rows, const rpt_col_t *cols, int ncols,
                                           uint64_t count, rpt_stream_t stream);
/* Applies the rule (rebuilding from cols when it fires; cols may be NULL when it cannot fire),
 * then sets finalized_.  rebuilt_out (optional) reports whether the rebuild happened. */
RPT_API int rpt_bf_finalize(rpt_bf_t *bf, uint64_t actual_rows, const rpt_col_t *cols, int ncols, uint64_t count,
                            int *rebuilt_out, rpt_stream_t stream);

/* ---- build-side column range: UpdateMinMax / ColumnMinMax (src/operators/physical_create_bf.cpp:82-176), which
 *      PhysicalCreateBF::Sink runs on every chunk next to the insert (:229-239) ----------------------------- */
/* Starts (or restarts) range tracking for a key column of the given physical type: from now on every
 * single-column rpt_bf_insert / rehash of that type also folds the column's valid (non-NULL) rows into a running
 * min / max on the device — inside the build kernel itself (the keys are read once, as Sink reads a chunk once for
 * the insert and for UpdateMinMax; no extra kernel, no host work per chunk).
 * Supported: the fixed-width numeric physical types (BOOL…UINT64, FLOAT, DOUBLE; DATE and TIMESTAMP are INT32 /
 * INT64 physically).  VARCHAR ranges stay on the host. */
RPT_API int rpt_bf_minmax_begin(rpt_bf_t *bf, int type);
RPT_API int rpt_bf_minmax_end(rpt_bf_t *bf);
/* Waits for the handle's work, then: *has_value = any valid row seen (ColumnMinMax::has_value); min_out / max_out
 * receive one value of the column's physical type.  Floating point: the reference compares with C++ `<` / `>`
 * chunk by chunk, which is order-dependent when NaNs (or both zeros) occur; here NaNs are left out of the range
 * and reported through *has_nan (optional) so the caller can widen the range, and -0.0 orders below +0.0. */
RPT_API int rpt_bf_minmax_get(const rpt_bf_t *bf, int *has_value, int *has_nan, void *min_out, void *max_out);

/* ---- probe: PTBloomFilter::LookupSel (src/bloom_filter.cpp:60-68) = HashColumns +
 *      BloomFilter::LookupHashes (duckdb bloom_filter.cpp:50-61, LookupOne :72-78) ---------- */
/* Writes the ascending row indices i in [0,count) whose key passes the filter to sel_out[0..found)
 * and the count to *found_out (LookupHashes' return value).  Entries at and beyond `found` are
 * unspecified, as in the reference.  count must be < 2^32 (sel_t is uint32).
 *   out_mem == RPT_MEM_DEVICE: sel_out (capacity count) and found_out are device pointers; the
 *     call is asynchronous on `stream` (key columns must then be device columns as well).
 *   out_mem == RPT_MEM_HOST:   sel_out / found_out are host pointers; the call returns when they
 *     are filled (key columns may be host or device). */
RPT_API int rpt_bf_probe(const rpt_bf_t *bf, const rpt_col_t *cols, int ncols, uint64_t count, uint32_t *sel_out,
                         uint64_t *found_out, int out_mem, rpt_stream_t stream);
/* Folded pre-filter of large probes.  No counterpart in the reference — an accelerator of LookupHashes
 * (duckdb bloom_filter.cpp:50-61) that never changes its answers.  For filters of 256 / 512 MiB (4 or 8 L2-sized
 * slices) a probe call of >= 2^22 rows first ORs the slices into one 64-MiB "folded" filter (once per content of the
 * words; made again after any insert / reset / merge / import): a key whose four bits are not all set in the folded
 * word cannot have them set in its own word, so every key meets the L2-resident folded filter first and only its
 * survivors are looked up in the real words.  The fold costs one read of the words (0.05 ms at 256 MiB).  The
 * pre-filter is switched off when more than half of the non-member keys would survive it (the fold's bit density:
 * estimated from the rows inserted since the words were zeroed — no host synchronisation — or, for imported / merged
 * words, read from the fold after one stream synchronisation), and, for a while, after a probe whose rows mostly
 * survived it (a probe side rich in members).
 *   mode: -1 library default (environment RPT_PREFILTER, default 1), 0 never, 1 by density, 2 always (tests).
 *   rpt_bf_prefilter_info: *state_out = 0 no fold made for the current words, 1 made and in use, 2 made but too dense;
 *   *pass_rate_out = estimated share of non-member keys that survive the folded filter. */
RPT_API int rpt_bf_set_prefilter(rpt_bf_t *bf, int mode);
RPT_API int rpt_bf_prefilter_info(const rpt_bf_t *bf, int *state_out, double *pass_rate_out);
/* Allocation-free probe.  NO COUNTERPART IN THE REFERENCE: added for callers
 * that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * Same result as rpt_bf_probe with one non-NULL 64-bit integer key column in device memory and device outputs
 * (ascending row ids to dev_sel_out[0..found), found to *dev_found_out), asynchronous on `stream` of `device`; the
 * filter is passed as its words (rpt_bf_device_words) and sector count (rpt_bf_num_sectors, a power of two).
 * Calls no cudaMalloc / cudaFree / allocating routine: the scratch (bitmap + block counts, 2 MiB + 32 KiB) is a
 * static __device__ array that every GPU which loads the library holds until exit.  count > RPT_NOALLOC_MAX_ROWS is
 * rejected with an error (split the batch, or use rpt_bf_probe).  All calls on a device SHARE the array: calls that
 * overlap on the GPU (different streams or threads) race; order them on one stream or with events. */
#define RPT_NOALLOC_MAX_ROWS (1ull << 24)
RPT_API int rpt_bf_probe_u64_noalloc(const uint64_t *dev_words, uint64_t num_sectors, const uint64_t *dev_keys,
                                     uint64_t count, uint32_t *dev_sel_out, uint64_t *dev_found_out, int device,
                                     rpt_stream_t stream);
/* Scan-side form: BFTableFilter::Filter (duckdb bloom_filter.cpp:94-131).  The key column is read
 * through an incoming selection `in_sel` (approved_count entries, may be NULL = identity) and the
 * output is translated back through it: sel_out[k] = in_sel[flat_k] (lines :117-123). */
RPT_API int rpt_bf_filter_sel(const rpt_bf_t *bf, const rpt_col_t *key, const uint32_t *in_sel,
                              uint64_t approved_count, uint32_t *sel_out, uint64_t *found_out, int mem,
                              rpt_stream_t stream);
/* BloomFilter::LookupOne on already-hashed values (BFTableFilter::FilterValue, CheckStatistics
 * :138-169 probe ≤2049 consecutive integers).  hashes/hits are host arrays. */
RPT_API int rpt_bf_lookup_hashes_host(const rpt_bf_t *bf, const uint64_t *hashes, uint64_t count, uint8_t *hits);

/* ---- filter words ------------------------------------------------------------------------ */
/* Zone-map pruning, batched over row groups: BFTableFilter::CheckStatistics / TemplatedCheckStatistics<T>
 * (duckdb/src/planner/filter/bloom_filter.cpp:138-196).  Row group i has statistics [mins[i], maxs[i]] of the
 * column's physical type (`type`; 16-byte {lower, upper} values for the 128-bit types) and has_minmax[i] != 0
 * (NULL = all have).  Like the reference: no statistics, min > max, a difference that does not fit the type or
 * exceeds 2048 → RPT_NO_PRUNING_POSSIBLE; otherwise the max-min+1 consecutive values are hashed as the type
 * hashes and looked up: none hits → RPT_FILTER_ALWAYS_FALSE, all hit → RPT_FILTER_ALWAYS_TRUE.  Types other
 * than the ten integral ones always give RPT_NO_PRUNING_POSSIBLE (:193-195).  Host arrays in, host array out. */
enum rpt_prune { RPT_NO_PRUNING_POSSIBLE = 0, RPT_FILTER_ALWAYS_TRUE = 1, RPT_FILTER_ALWAYS_FALSE = 2 };
RPT_API int rpt_bf_check_statistics(const rpt_bf_t *bf, int type, const void *mins, const void *maxs,
                                    const uint8_t *has_minmax, uint64_t count, uint8_t *results_out);

/* D2H copy of the S words; byte-identical to DuckDB's bf[] so the scan-side BFTableFilter,
 * zone-map CheckStatistics and passthrough USE_BF keep working on the exported copy. */
RPT_API int rpt_bf_export_words(const rpt_bf_t *bf, uint64_t *host_dst, uint64_t capacity_words);
/* H2D replace of the S words (e.g. a filter built by CPU sink threads); marks has_data_. */
RPT_API int rpt_bf_import_words(rpt_bf_t *bf, const uint64_t *host_src, uint64_t num_words);
/* FNV-1a-64 over the S words (xor-then-multiply per 64-bit word, offset 0xcbf29ce484222325, prime 0x100000001b3):
 * SURVEY §8(c)'s recipe, the "fnv1a_words" of tests/golden/ref_vectors.json, and what the ranks compare after a
 * multi-GPU merge (bench.py). */
RPT_API int rpt_bf_checksum(const rpt_bf_t *bf, uint64_t *fnv1a_out);
/* Bumped whenever the words are reallocated (rpt_bf_reset / rehash to another S — or to the same S at a recycled
 * address): peers that mapped the old allocation through rpt_bf_ipc_export must re-open when it changes. */
RPT_API int rpt_bf_words_generation(const rpt_bf_t *bf, uint64_t *out);
/* Raw device pointer to the words (for collective plumbing); valid until destroy / rebuild. */
RPT_API int rpt_bf_device_words(const rpt_bf_t *bf, uint64_t **dev_ptr_out);
RPT_API int rpt_bf_sync(const rpt_bf_t *bf); /* wait for all work enqueued on the handle's own stream */
RPT_API int rpt_bf_set_has_data(rpt_bf_t *bf, int has_data);
/* Number of kernels this handle has launched so far (bench.py's gpu_launches evidence). */
RPT_API int rpt_bf_launch_count(const rpt_bf_t *bf, long long *out);

/* ---- multi-GPU OR-merge: partial filters built from row-range shards are merged
 *      with a bitwise OR; fetch_or is commutative/associative/idempotent (bloom_filter.cpp:69) so
 *      the merged words equal the single-filter build bit for bit ----------------------------- */
/* bf.words |= parts[0..nparts) where parts is nparts*S contiguous device words on bf's device
 * (e.g. the receive buffer of an NCCL all-gather of every rank's partial filter). */
RPT_API int rpt_bf_or_parts(rpt_bf_t *bf, const uint64_t *dev_parts, uint64_t nparts, rpt_stream_t stream);
/* One process per GPU: export a CUDA IPC handle (64 bytes) for this filter's words ... */
RPT_API int rpt_bf_ipc_export(const rpt_bf_t *bf, void *handle64);
/* ... and, after every rank holds all ranks' handles (exchanged by the host plumbing), run the
 * two-phase peer-memory merge.  Phase 0: this rank ORs slice `rank` of every peer's words into its
 * own slice (reads over NVLink).  Phase 1: this rank copies every other rank's reduced slice into
 * its own words.  The caller puts a cross-rank barrier before phase 0, between the phases and
 * after phase 1 (no kernel waits on another rank). */
RPT_API int rpt_bf_peer_open(rpt_bf_t *bf, const void *handles64, int nranks, int rank);
RPT_API int rpt_bf_peer_merge_phase(rpt_bf_t *bf, int phase, rpt_stream_t stream);
RPT_API int rpt_bf_peer_close(rpt_bf_t *bf);
/* The same OR-allreduce as ONE call that owns its barriers (no counterpart in the reference).  Every
 * rank exports a small block of barrier words (rpt_bf_peer_flags_export, once per handle), the handles are
 * exchanged like the filters' (rpt_bf_peer_flags_open), and rpt_bf_peer_allreduce enqueues the whole collective on
 * `stream`: signal "my partial filter is complete" (+ has_local_data) → wait for every peer → reduce-scatter-OR of
 * my slice over NVLink → signal → wait → pull the peers' reduced slices → signal "pulled".  Signals are stores into
 * the peers' words over the IPC mappings, waits are stream memory operations (cuStreamWaitValue32) on the rank's own
 * words: no NCCL call, no host synchronisation, nothing polls.  The last barrier ("everybody has pulled my slice")
 * is not needed by the probes that follow; it is taken by the next rpt_bf_peer_allreduce or by rpt_bf_peer_release,
 * which the caller invokes before it changes the words again (rpt_bf_reset / rpt_bf_insert).  rpt_bf_peer_any_data
 * (synchronises `stream`) tells whether any rank held data — PTBloomFilter::IsEmpty of the merged filter. */
RPT_API int rpt_bf_peer_flags_export(rpt_bf_t *bf, void *handle64);
RPT_API int rpt_bf_peer_flags_open(rpt_bf_t *bf, const void *handles64, int nranks, int rank);
RPT_API int rpt_bf_peer_allreduce(rpt_bf_t *bf, int has_local_data, rpt_stream_t stream);
RPT_API int rpt_bf_peer_release(rpt_bf_t *bf, rpt_stream_t stream);
RPT_API int rpt_bf_peer_any_data(rpt_bf_t *bf, rpt_stream_t stream, int *out);
/* The same two phases restricted to the sectors [first_sector, first_sector + nsectors) (both even): the merge can
 * then proceed slice by slice on a side stream while the probe already runs — see rpt_bf_set_slice_events. */
RPT_API int rpt_bf_peer_merge_range(rpt_bf_t *bf, int phase, uint64_t first_sector, uint64_t nsectors,
                                    rpt_stream_t stream);
/* Sectors per slice of the large-filter probe paths (2^23 = 64 MiB). */
RPT_API int rpt_bf_slice_sectors(uint64_t *out);
/* Overlapping the merge with the probe: events[s] (cudaEvent_t, n ≤ 8) is recorded by the caller once slice s
 * (sectors [s·slice, (s+1)·slice)) holds its merged words.  Until cleared (n = 0), every probe on this handle makes
 * its stream wait for events[s] right before the first kernel that reads slice s — the partition kernel needs slice 0
 * only, the s-th bucket kernel / sector-range pass only slice s — instead of requiring the whole merge up front. */
RPT_API int rpt_bf_set_slice_events(rpt_bf_t *bf, void *const *events, int n);
/* One process driving several GPUs (the DuckDB-extension deployment): OR-allreduce in place over
 * peer access; all filters must have the same S. */
RPT_API int rpt_bf_allreduce_or(rpt_bf_t **per_gpu, int ngpu);

/* ---- synthetic key columns of SURVEY §8(d) / BASELINE.md §4 (benchmark inputs, generated on
 *      the device so that 10^9-row columns need no host RAM) -------------------------------- */
/* B[i] = splitmix64(0xB10F0001 + first + i) & 0x7fff...f */
RPT_API int rpt_synth_build_keys(uint64_t *dev_out, uint64_t first, uint64_t count, int device, rpt_stream_t stream);
/* Row j: member B[r mod n_build] with probability sel_ppm/1e6, else a value with bit 63 set. */
RPT_API int rpt_synth_probe_keys(uint64_t *dev_out, uint64_t first, uint64_t count, uint64_t n_build,
                                 uint32_t sel_ppm, int device, rpt_stream_t stream);

/* ---- README-described filter variant — PARITY UNPINNED, not the reference's code path ---------
 * The reference's README.md:23-32 (and BASELINE.json's north_star) describe an Apache-Arrow-Acero
 * BlockedBloomFilter: one 64-bit block per key, a pre-generated table of 1024 57-bit masks with 4-5
 * bits set, 8 bits per key, and Fold (OR the halves together until more than 25 % of the bits are
 * set).  No such code exists in the reference at this commit: src/bloom_filter.cpp wraps
 * DuckDB's native BloomFilter, which the rpt_bf_* entry points above replace bit for bit.  rpt_abf_*
 * implements the published Arrow algorithm (arrow/acero/bloom_filter.{h,cc}: BloomFilterMasks,
 * BlockedBloomFilter::CreateEmpty/mask/block_id/Insert/Find/Fold/SingleFold) over the same DuckDB key
 * hashes (same rpt_col_t columns, NULL rows hash to NULL_HASH).  It is checked against an independent
 * CPU restatement (oracle/acero_oracle.cpp), never against the reference, and nothing on the rpt_bf_*
 * path calls it.  A maintainer who lands the README's filter in src/bloom_filter.cpp would bind these
 * the way INTEGRATION.md binds rpt_bf_*. */
typedef struct rpt_abf rpt_abf_t;
/* CreateEmpty(num_rows): 2^log_num_blocks 64-bit blocks, log = ceil(log2(max(512, 8*num_rows))) - 6 */
RPT_API int rpt_abf_log_blocks_for_rows(uint64_t rows, int *out);
RPT_API int rpt_abf_create(rpt_abf_t **out, uint64_t num_rows, int device);
RPT_API int rpt_abf_destroy(rpt_abf_t *f);
RPT_API int rpt_abf_log_num_blocks(const rpt_abf_t *f, int *out); /* shrinks with every fold */
/* The 136-byte bit vector whose 57-bit windows are the 1024 masks (BloomFilterMasks::masks_) */
RPT_API int rpt_abf_mask_bits(uint8_t *out136);
/* Insert(hash(row)) for every row.  Device columns: asynchronous on `stream`.  Host columns: returns
 * when the rows are in the filter. */
RPT_API int rpt_abf_insert(rpt_abf_t *f, const rpt_col_t *cols, int ncols, uint64_t count, rpt_stream_t stream);
/* Fold(): synchronises `stream` (the density test needs the popcount on the host).
 * *num_folds_out (optional) = number of halvings performed. */
RPT_API int rpt_abf_fold(rpt_abf_t *f, rpt_stream_t stream, int *num_folds_out);
/* Find(hash(row)) for every row -> ascending row ids of the rows that pass + their number.
 * out_mem = RPT_MEM_DEVICE: sel_out / found_out are device pointers, columns must be device columns,
 * asynchronous on `stream`.  RPT_MEM_HOST: host outputs, returns when they are filled. */
RPT_API int rpt_abf_probe(const rpt_abf_t *f, const rpt_col_t *cols, int ncols, uint64_t count, uint32_t *sel_out,
                          uint64_t *found_out, int out_mem, rpt_stream_t stream);
/* Copies the 2^log_num_blocks blocks to the host (after work on a caller's stream, synchronise it first). */
RPT_API int rpt_abf_export_blocks(const rpt_abf_t *f, uint64_t *host_dst, uint64_t capacity_blocks);
/* blocks |= parts[0..nparts): same-sized partial filters on this device (multi-GPU merge; fold afterwards) */
RPT_API int rpt_abf_or_parts(rpt_abf_t *f, const uint64_t *dev_parts, uint64_t nparts, rpt_stream_t stream);
RPT_API int rpt_abf_device_blocks(const rpt_abf_t *f, uint64_t **dev_ptr_out);
RPT_API int rpt_abf_sync(const rpt_abf_t *f);
RPT_API int rpt_abf_launch_count(const rpt_abf_t *f, long long *out);

#ifdef __cplusplus
}
#endif
#endif /* RPT_GPU_H */
