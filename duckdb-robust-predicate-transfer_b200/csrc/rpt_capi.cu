// rpt_capi.cu — implementation of the C ABI declared in include/rpt_gpu.h.
//
// Host-side responsibilities only: handle lifetime, the sizing / resize arithmetic of the reference
// (duckdb/src/planner/filter/bloom_filter.cpp:7-27, src/operators/physical_create_bf.cpp:386-406),
// staging host key columns through chunked, double-buffered H2D/D2H pipelines, and launching the
// kernels of rpt_kernels.cu.  There is no CPU fallback anywhere in this file: every compute entry
// point ends in a kernel launch or fails.
#include <cuda.h>
#include <cuda_runtime.h>

#include <atomic>
#include <cmath>
#include <memory>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <ctime>
#include <mutex>
#include <new>
#include <stdexcept>
#include <string>
#include <type_traits>
#include <vector>

#include "../../include/rpt_gpu.h"
#include "rpt_kernels.cuh"

using namespace rpt;

// ---------------------------------------------------------------------------------------------
// errors
// ---------------------------------------------------------------------------------------------
static thread_local std::string g_last_error;

static int set_error(const char *fmt, ...) {
	char buf[512];
	va_list ap;
	va_start(ap, fmt);
	vsnprintf(buf, sizeof(buf), fmt, ap);
	va_end(ap);
	g_last_error = buf;
	return 1;
}

// for the library's other translation units (rpt_noalloc.cu): same thread-local message, same status
namespace rpt {
int report_error(const char *msg) {
	return set_error("%s", msg);
}
} // namespace rpt

#define RPT_CUDA(expr)                                                                                                 \
	do {                                                                                                               \
		cudaError_t _e = (expr);                                                                                       \
		if (_e != cudaSuccess) {                                                                                       \
			return set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__);              \
		}                                                                                                              \
	} while (0)

// No C++ exception crosses the C ABI: allocation failures inside a body (std::vector, new) come back as a status.
template <class F>
static int guarded(F &&body) noexcept {
	try {
		return body();
	} catch (const std::bad_alloc &) {
		return set_error("out of host memory");
	} catch (const std::exception &e) {
		return set_error("internal error: %s", e.what());
	} catch (...) {
		return set_error("internal error: unknown exception");
	}
}

#define RPT_TRY(expr)                                                                                                  \
	do {                                                                                                               \
		int _rc = (expr);                                                                                              \
		if (_rc != 0) {                                                                                                \
			return _rc;                                                                                                \
		}                                                                                                              \
	} while (0)

// ---------------------------------------------------------------------------------------------
// reference sizing arithmetic
// ---------------------------------------------------------------------------------------------
// NextPowerOfTwo, duckdb/src/common/constants.cpp:30-47 (returns 0 where the reference throws)
static uint64_t next_pow2(uint64_t v) {
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

// BloomFilter::Initialize, duckdb/src/planner/filter/bloom_filter.cpp:7-9,17-19
static int sectors_for_rows(uint64_t rows, uint64_t *out) {
	uint64_t min_bits = rows * 12ULL; // idx_t arithmetic, wraps like the reference
	if (min_bits < 512ULL) {
		min_bits = 512ULL;
	}
	const uint64_t p2 = next_pow2(min_bits);
	if (p2 == 0) {
		return set_error("Can't find next power of 2 for %llu", (unsigned long long)min_bits); // OutOfRangeException
	}
	const uint64_t s = p2 >> 6;
	*out = s < (1ULL << 26) ? s : (1ULL << 26);
	return 0;
}

// ---------------------------------------------------------------------------------------------
// handle
// ---------------------------------------------------------------------------------------------
// Device memory of one pipeline slot of the host-column calls: the staged key columns and the probe scratch of a chunk
// are carved out of it instead of being allocated from the stream-ordered pool chunk by chunk.  (With pool
// allocations on the upload stream and releases on the compute stream, the first host-column calls after a run of
// device-column calls spent up to a second in cudaMallocFromPoolAsync growing the pool — bench.py's e2e leg swung
// between 161 and 330 ms per step, profiles/bench_r2final.json before this.)  Grow-only: a call that needs more than
// a slot holds falls back to the pool for the excess and the slot is re-sized when the next call starts.
struct SlotArena {
	char *base = nullptr;
	size_t cap = 0;
	size_t want = 0; // the most one chunk has asked for so far
};

struct rpt_bf {
	DeviceInfo dev;
	uint64_t num_sectors = 0;
	uint64_t sized_for_rows = 0;
	uint64_t *words = nullptr;
	uint64_t words_capacity = 0;       // sectors allocated (grow-only: a re-size to fewer sectors reuses the buffer)
	cudaStream_t stream = nullptr;     // handle's own stream
	// host-column pipeline (RPT_MEM_HOST calls): uploads, kernels and downloads run on three streams chained by
	// events, with kPipeSlots chunks in flight; everything here is created once per handle and reused
	static constexpr int kPipeSlots = 3;
	struct PipeSlot {
		uint32_t *dsel = nullptr;   // device selection vector of one chunk (probe)
		uint64_t sel_cap = 0;       // its capacity in rows
		uint64_t *dfound = nullptr; // device survivor count
		uint64_t *hfound = nullptr; // pinned host copy of it
		cudaEvent_t uploaded = nullptr, computed = nullptr, drained = nullptr;
		bool pending = false;
		SlotArena arena;            // staging + scratch of the chunk in this slot
	};
	std::mutex pipe_mu; // host-column calls on one handle take turns (device-column calls never touch the pipeline)
	cudaStream_t h2d = nullptr, compute = nullptr, d2h = nullptr;
	PipeSlot slot[kPipeSlots];
	std::atomic<int> has_data {0};
	std::atomic<int> finalized {0};
	std::atomic<long long> launches {0};
	std::atomic<uint64_t> words_generation {0}; // bumped whenever `words` is reallocated (stale peer mappings)
	// Folded pre-filter of the words (probes of 256 / 512 MiB filters, rpt_kernels.cuh Prefilter): built lazily by
	// the first large probe after a mutation, on that probe's stream; content_version is bumped by every entry point
	// that may change the words (or hands out their address), fold_version is the version the fold was made from.
	std::mutex fold_mu;
	uint64_t *folded = nullptr;        // kSliceSectors words (64 MiB), allocated on first use
	uint64_t *fold_pop_dev = nullptr;  // set bits of the folded filter
	uint64_t *fold_pop_host = nullptr; // pinned copy
	std::atomic<uint64_t> content_version {1};
	uint64_t fold_version = 0;
	double fold_pass = 1.0;            // estimated share of non-member keys that survive the folded filter
	// rows inserted since the words were last zeroed, while inserts are all that happened to them (-1: unknown — an
	// import, a merge, or the address was handed out): lets the host estimate the fold's density WITHOUT waiting
	std::atomic<long long> rows_inserted {0};
	cudaEvent_t fold_ready = nullptr;  // recorded behind the fold kernel; every pre-filtered probe waits for it
	bool fold_pop_pending = false;     // the fold's measured bit count is on its way to fold_pop_host
	std::atomic<int> fold_mode {-1};   // rpt_bf_set_prefilter: -1 = library default (RPT_PREFILTER)
	// feedback: the pre-filtered partition counts the rows it saw and the rows that survived the folded filter
	// (members + its false positives); when too many survive, the pre-filter costs more than it saves and the
	// following probes of this handle go without it for a while (ensure_prefilter)
	uint64_t *fold_stats_dev = nullptr;  // [rows, survivors]
	uint64_t *fold_stats_host = nullptr; // pinned copy
	cudaEvent_t fold_stats_ev = nullptr;
	bool fold_stats_pending = false;     // a probe is counting (or about to); its counters are not on the host yet
	bool fold_stats_recorded = false;    // ... and fold_stats_ev has been recorded behind that probe's kernels
	double fold_alive = -1.0;            // survivors / rows of the last measured probe (< 0: none)
	int fold_holdoff = 0;                // large probes left to run without the pre-filter
	// fused column range (rpt_bf_minmax_*)
	MinMaxState *mm_dev = nullptr;
	int mm_type = 0;
	std::atomic<int> mm_active {0};
	// slice gates of an overlapped multi-GPU merge (rpt_bf_set_slice_events); n == 0: none
	SliceGate gate {{nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr}, 0};
	// peer merge state
	std::vector<uint64_t *> peer_words;
	std::vector<bool> peer_ipc;
	int peer_rank = -1;
	// library-owned collective (rpt_bf_peer_allreduce): 3 phases × 16 ranks of 32-bit barrier words in device memory
	// (phase 0 "my partial filter is complete" + has-data bit, 1 "my slice is reduced", 2 "I have pulled every slice");
	// peers store into my words over their IPC mappings, I wait on my own
	uint32_t *sync_flags = nullptr;
	std::vector<uint32_t *> peer_flags; // [rank] → that rank's flag block (mine included)
	std::vector<bool> peer_flags_ipc;
	uint32_t sync_epoch = 0;
	bool sync_release_pending = false;
};

struct DeviceGuard {
	int prev = -1;
	bool ok = false;
	explicit DeviceGuard(int device) {
		if (cudaGetDevice(&prev) == cudaSuccess && cudaSetDevice(device) == cudaSuccess) {
			ok = true;
		}
	}
	~DeviceGuard() {
		if (prev >= 0) {
			cudaSetDevice(prev);
		}
	}
};

#define RPT_GUARD(bf)                                                                                                  \
	DeviceGuard _guard((bf)->dev.device);                                                                              \
	if (!_guard.ok) {                                                                                                  \
		return set_error("cannot select CUDA device %d: %s", (bf)->dev.device,                                        \
		                 cudaGetErrorString(cudaGetLastError()));                                                      \
	}

static int fill_device_info(int device, DeviceInfo *out) {
	cudaDeviceProp prop;
	RPT_CUDA(cudaGetDeviceProperties(&prop, device));
	if (prop.major < 10) {
		return set_error("device %d is sm_%d%d; librpt_gpu is built for sm_100a (B200) only", device, prop.major,
		                 prop.minor);
	}
	// L2 set-aside for persisting (evict_last) lines: the filter slice of a pass lives there, out of reach of the
	// key stream.  RPT_L2_PERSIST_MIB overrides (0 = leave the device default).
	{
		const char *e = getenv("RPT_L2_PERSIST_MIB");
		const long mib = e ? atol(e) : -1;
		if (mib > 0) {
			size_t want = static_cast<size_t>(mib) << 20;
			if (want > static_cast<size_t>(prop.persistingL2CacheMaxSize)) {
				want = prop.persistingL2CacheMaxSize;
			}
			cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, want);
		}
	}
	cudaGetLastError();
	out->device = device;
	out->sm_count = prop.multiProcessorCount;
	out->l2_bytes = prop.l2CacheSize;
	out->smem_optin = prop.sharedMemPerBlockOptin;
	return 0;
}

// Scratch and staging come from a PRIVATE stream-ordered pool per device (freed blocks stay in the pool across
// synchronisations: release threshold = max) — the host application's default pool is left alone.
static int library_pool(int device, cudaMemPool_t *out) {
	static std::mutex mu;
	static cudaMemPool_t pools[64] = {};
	std::lock_guard<std::mutex> lock(mu);
	if (device < 0 || device >= 64) {
		return set_error("device %d out of range", device);
	}
	if (!pools[device]) {
		cudaMemPoolProps props = {};
		props.allocType = cudaMemAllocationTypePinned;
		props.handleTypes = cudaMemHandleTypeNone;
		props.location.type = cudaMemLocationTypeDevice;
		props.location.id = device;
		RPT_CUDA(cudaMemPoolCreate(&pools[device], &props));
		uint64_t keep = ~0ULL;
		RPT_CUDA(cudaMemPoolSetAttribute(pools[device], cudaMemPoolAttrReleaseThreshold, &keep));
	}
	*out = pools[device];
	return 0;
}

static size_t type_width(int type) {
	switch (type) {
	case RPT_BOOL:
	case RPT_INT8:
	case RPT_UINT8:
		return 1;
	case RPT_INT16:
	case RPT_UINT16:
		return 2;
	case RPT_INT32:
	case RPT_UINT32:
	case RPT_FLOAT:
		return 4;
	case RPT_INT64:
	case RPT_UINT64:
	case RPT_DOUBLE:
		return 8;
	case RPT_INTERVAL:
	case RPT_INT128:
	case RPT_UINT128:
		return 16;
	default:
		return 0;
	}
}

static int validate_cols(const rpt_col_t *cols, int ncols, uint64_t count, int *mem_out) {
	if (!cols || ncols < 1 || ncols > RPT_MAX_KEY_COLS) {
		return set_error("ncols must be in [1, %d]", RPT_MAX_KEY_COLS);
	}
	for (int j = 0; j < ncols; j++) {
		const rpt_col_t &c = cols[j];
		if (c.mem != cols[0].mem) {
			return set_error("all key columns of a call must live in the same memory space");
		}
		const bool known = type_width(c.type) != 0 || c.type == RPT_VARCHAR || c.type == RPT_STRING_T;
		if (!known) {
			return set_error("Invalid type for hash: physical type %d", c.type); // InvalidTypeException
		}
		if (c.type == RPT_STRING_T && c.mem != RPT_MEM_HOST) {
			return set_error("RPT_STRING_T columns are host-only");
		}
		if (count > 0 && !c.data) {
			return set_error("column %d has no data pointer", j);
		}
		if (c.type == RPT_VARCHAR && count > 0 && !c.aux && c.aux_bytes > 0) {
			return set_error("VARCHAR column %d has no bytes buffer", j);
		}
	}
	*mem_out = cols[0].mem;
	return 0;
}

// ---------------------------------------------------------------------------------------------
// staging of key columns: rpt_col_t (host or device) → DevCols (device pointers)
// ---------------------------------------------------------------------------------------------
struct Staged {
	cudaStream_t stream = nullptr;      // allocations and uploads are ordered on this stream
	cudaStream_t free_stream = nullptr; // releases are ordered on this one (the consumer); default = stream
	// set once the consumer stream has been made to wait for the uploads (cudaStreamWaitEvent on `uploaded`):
	// until then a release on free_stream could overtake copies still in flight on `stream`
	bool handed_over = false;
	int device = -1;
	std::vector<void *> dev_allocs;
	SlotArena *arena = nullptr; // host-column pipeline: carve from the slot's memory first (never freed here)
	size_t arena_used = 0, asked = 0;
	~Staged() {
		if (arena && asked > arena->want) {
			arena->want = asked;
		}
		if (free_stream && free_stream != stream && !handed_over && !dev_allocs.empty()) {
			cudaStreamSynchronize(stream); // error path: the uploads finish before their buffers are released
		}
		for (void *p : dev_allocs) {
			cudaFreeAsync(p, free_stream ? free_stream : stream);
		}
	}
	int alloc(void **out, size_t bytes) {
		if (device < 0) {
			RPT_CUDA(cudaGetDevice(&device));
		}
		if (arena) {
			const size_t sz = ((bytes ? bytes : 8) + 255) / 256 * 256;
			asked += sz;
			if (arena_used + sz <= arena->cap) {
				*out = arena->base + arena_used;
				arena_used += sz;
				return 0;
			}
		}
		cudaMemPool_t pool;
		RPT_TRY(library_pool(device, &pool));
		RPT_CUDA(cudaMallocFromPoolAsync(out, bytes ? bytes : 8, pool, stream));
		dev_allocs.push_back(*out);
		return 0;
	}
	int upload(void **out, const void *host, size_t bytes, size_t pad = 0) {
		RPT_TRY(alloc(out, bytes + pad));
		if (bytes) {
			RPT_CUDA(cudaMemcpyAsync(*out, host, bytes, cudaMemcpyHostToDevice, stream));
		}
		return 0;
	}
};

// Rows of a chunked host upload start at multiples of this (validity words and 128-bit key loads
// stay aligned; a multiple of every probe tile size).
static constexpr uint64_t kChunkRows = 1ULL << 24;
// RPT_CHUNK_ROWS overrides the chunk size (rounded down to a multiple of 4096) for experiments.
static uint64_t chunk_rows() {
	static const uint64_t rows = [] {
		const char *e = getenv("RPT_CHUNK_ROWS");
		const uint64_t v = e ? strtoull(e, nullptr, 10) & ~4095ULL : 0;
		return v ? v : kChunkRows;
	}();
	return rows;
}
// RPT_TRACE=1 prints host-side wall times of the host-column pipelines to stderr.
static bool trace_on() {
	static const bool on = getenv("RPT_TRACE") != nullptr;
	return on;
}
static double now_ms() {
	timespec ts;
	clock_gettime(CLOCK_MONOTONIC, &ts);
	return ts.tv_sec * 1e3 + ts.tv_nsec * 1e-6;
}

static void device_cols_view(const rpt_col_t *cols, int ncols, DevCols *out) {
	memset(out, 0, sizeof(*out));
	out->n = ncols;
	for (int j = 0; j < ncols; j++) {
		out->c[j].data = cols[j].data;
		out->c[j].aux = cols[j].aux;
		out->c[j].sel = cols[j].sel;
		out->c[j].validity = cols[j].validity;
		out->c[j].type = cols[j].type;
	}
}

static bool any_sel(const rpt_col_t *cols, int ncols) {
	for (int j = 0; j < ncols; j++) {
		if (cols[j].sel) {
			return true;
		}
	}
	return false;
}

// duckdb::string_t (string_type.hpp:24): uint32 length, then 12 inlined bytes, or a 4-byte prefix
// followed by a pointer.  Gathers entries [e0, e1) into Arrow-style offsets + bytes.  The payload of an entry is
// read only when the entry is valid (validity bit set, or no validity) AND referenced (`used` bit set, or no
// selection): DuckDB leaves the string_t of a NULL row undefined, and the reference never looks at it
// (TightLoopHash: `is_null ? NULL_HASH : Hash(input)`, vector_hash.cpp:43-61).  Skipped entries get length 0.
static void gather_string_t(const void *data, uint64_t e0, uint64_t e1, const uint64_t *validity,
                            const std::vector<uint64_t> *used, std::vector<uint32_t> *offsets,
                            std::vector<uint8_t> *bytes) {
	const uint8_t *base = reinterpret_cast<const uint8_t *>(data);
	auto live = [&](uint64_t e) {
		if (validity && !((validity[e >> 6] >> (e & 63)) & 1ULL)) {
			return false;
		}
		return !used || (((*used)[e >> 6] >> (e & 63)) & 1ULL);
	};
	offsets->resize(e1 - e0 + 1);
	uint64_t total = 0;
	for (uint64_t e = e0; e < e1; e++) {
		(*offsets)[e - e0] = static_cast<uint32_t>(total);
		if (live(e)) {
			uint32_t len;
			memcpy(&len, base + e * 16, 4);
			total += len;
		}
	}
	if (total > 0xffffffffULL) {
		throw std::length_error("string bytes of one call exceed the 32-bit offsets; split the batch");
	}
	(*offsets)[e1 - e0] = static_cast<uint32_t>(total);
	bytes->resize(total);
	for (uint64_t e = e0; e < e1; e++) {
		if (!live(e)) {
			continue;
		}
		uint32_t len;
		memcpy(&len, base + e * 16, 4);
		const uint8_t *src;
		if (len <= 12) {
			src = base + e * 16 + 4;
		} else {
			memcpy(&src, base + e * 16 + 8, sizeof(src));
		}
		if (len) {
			memcpy(bytes->data() + (*offsets)[e - e0], src, len);
		}
	}
}

// Uploads rows [r0, r1) of host columns (r0 % 64 == 0).  When a column carries a selection the
// caller passes the whole call as one chunk (r0 == 0) and the column's data_len entries are
// uploaded once.
static int stage_host_cols(const rpt_col_t *cols, int ncols, uint64_t r0, uint64_t r1, Staged *st, DevCols *out) {
	memset(out, 0, sizeof(*out));
	out->n = ncols;
	const uint64_t n = r1 - r0;
	for (int j = 0; j < ncols; j++) {
		const rpt_col_t &c = cols[j];
		DevCol &d = out->c[j];
		d.type = c.type == RPT_STRING_T ? RPT_VARCHAR : c.type;
		uint64_t e0 = r0, e1 = r1; // data entries to upload
		if (c.sel) {
			if (r0 != 0) {
				return set_error("internal: selection columns are not chunked");
			}
			void *p;
			RPT_TRY(st->upload(&p, c.sel, n * sizeof(uint32_t)));
			d.sel = reinterpret_cast<const uint32_t *>(p);
			if (c.data_len == 0) {
				return set_error("column %d: data_len is required when a selection vector is given", j);
			}
			e0 = 0;
			e1 = c.data_len;
		}
		const uint64_t ne = e1 - e0;
		if (c.validity) {
			void *p;
			RPT_TRY(st->upload(&p, c.validity + (e0 >> 6), ((ne + 63) >> 6) * sizeof(uint64_t)));
			d.validity = reinterpret_cast<const uint64_t *>(p);
		}
		if (c.type == RPT_VARCHAR) {
			const uint32_t *off = reinterpret_cast<const uint32_t *>(c.data);
			void *po, *pb;
			RPT_TRY(st->upload(&po, off + e0, (ne + 1) * sizeof(uint32_t)));
			const uint64_t b0 = ne ? off[e0] : 0, b1 = ne ? off[e1] : 0;
			RPT_TRY(st->upload(&pb, reinterpret_cast<const uint8_t *>(c.aux) + b0, b1 - b0, 8));
			d.data = po;
			d.aux = reinterpret_cast<const uint8_t *>(pb) - b0; // offsets stay absolute
		} else if (c.type == RPT_STRING_T) {
			std::vector<uint32_t> offsets;
			std::vector<uint8_t> bytes;
			std::vector<uint64_t> used; // entries some row of the call references (selection vectors only)
			if (c.sel) {
				used.assign((ne + 63) >> 6, 0);
				for (uint64_t i = 0; i < n; i++) {
					const uint64_t e = c.sel[i];
					if (e >= ne) {
						return set_error("column %d: selection entry %llu is outside data_len %llu", j,
						                 (unsigned long long)e, (unsigned long long)ne);
					}
					used[e >> 6] |= 1ULL << (e & 63);
				}
			}
			// validity is indexed by entry; the chunk's words start at entry e0 (e0 % 64 == 0)
			gather_string_t(c.data, e0, e1, c.validity, c.sel ? &used : nullptr, &offsets, &bytes);
			void *po, *pb;
			RPT_TRY(st->upload(&po, offsets.data(), offsets.size() * sizeof(uint32_t)));
			RPT_TRY(st->upload(&pb, bytes.data(), bytes.size(), 8));
			// pageable sources: cudaMemcpyAsync returns after the buffer was staged, so the vectors may die here
			d.data = po;
			d.aux = pb;
		} else {
			const size_t w = type_width(c.type);
			void *p;
			RPT_TRY(st->upload(&p, reinterpret_cast<const uint8_t *>(c.data) + e0 * w, ne * w));
			d.data = p;
		}
	}
	return 0;
}

// String keys are hashed ahead of the build / probe kernels, which then read a column of 64-bit hashes:
//  * dictionary vectors — a VARCHAR column seen through a selection whose dictionary is at most half as long as
//    the call has rows: the dictionary is hashed once (launch_hash_entries) and the rows gather the hashes; the
//    reference reaches the same values through the dictionary's cached hashes (vector_hash.cpp:484-491);
//  * flat VARCHAR columns of at least kPrehashMinRows rows: launch_hash_varchar stages offsets and bytes through
//    shared memory with coalesced loads instead of one dependent load chain per thread.
// plan_prehash allocates (ordered on the staging stream, before the upload event); run_prehash hashes and swaps
// the column on the stream that consumes it.
static constexpr uint64_t kPrehashMinRows = 1ULL << 15;
struct PrehashPlan {
	uint64_t *hashes = nullptr;
	uint64_t entries = 0;
	bool dictionary = false;
};
static int plan_prehash(const rpt_col_t *cols, int ncols, uint64_t count, const DevCols &dc, Staged *st,
                        PrehashPlan *plan) {
	if (ncols != 1 || dc.c[0].type != RPT_VARCHAR) {
		return 0;
	}
	uint64_t entries;
	if (dc.c[0].sel) {
		entries = cols[0].data_len;
		if (entries == 0 || entries > count / 2) {
			return 0;
		}
		plan->dictionary = true;
	} else {
		if (count < kPrehashMinRows) {
			return 0;
		}
		entries = count;
	}
	void *p;
	RPT_TRY(st->alloc(&p, entries * sizeof(uint64_t)));
	plan->hashes = reinterpret_cast<uint64_t *>(p);
	plan->entries = entries;
	return 0;
}
static int run_prehash(const DeviceInfo &dev, const PrehashPlan &plan, DevCols *dc, cudaStream_t s, int *launches) {
	if (!plan.hashes) {
		return 0;
	}
	if (plan.dictionary) {
		DevCol dict = dc->c[0];
		dict.sel = nullptr;
		RPT_CUDA(launch_hash_entries(dev, dict, plan.entries, plan.hashes, s));
	} else {
		RPT_CUDA(launch_hash_varchar(dev, dc->c[0], plan.entries, plan.hashes, s));
	}
	(*launches)++;
	dc->c[0].data = plan.hashes;
	dc->c[0].aux = nullptr;
	dc->c[0].validity = nullptr;
	dc->c[0].type = kTypePrehashed;
	return 0;
}

static cudaStream_t pick_stream(const rpt_bf *bf, rpt_stream_t s) {
	return s ? reinterpret_cast<cudaStream_t>(s) : bf->stream;
}

// ---------------------------------------------------------------------------------------------
// library
// ---------------------------------------------------------------------------------------------
extern "C" {

const char *rpt_gpu_last_error(void) {
	return g_last_error.c_str();
}

const char *rpt_gpu_version(void) {
	return "rpt_gpu 0.1 (sm_100a)";
}

int rpt_gpu_device_count(int *out) {
	if (!out) {
		return set_error("null argument");
	}
	RPT_CUDA(cudaGetDeviceCount(out));
	return 0;
}

int rpt_host_alloc(void **out, uint64_t bytes) {
	if (!out) {
		return set_error("null argument");
	}
	RPT_CUDA(cudaMallocHost(out, bytes ? bytes : 1));
	return 0;
}

int rpt_host_free(void *ptr) {
	RPT_CUDA(cudaFreeHost(ptr));
	return 0;
}

int rpt_bf_sectors_for_rows(uint64_t rows, uint64_t *out) {
	if (!out) {
		return set_error("null argument");
	}
	return sectors_for_rows(rows, out);
}

// ---------------------------------------------------------------------------------------------
// lifetime
// ---------------------------------------------------------------------------------------------
static int allocate_words(rpt_bf *bf, uint64_t rows) {
	uint64_t s;
	RPT_TRY(sectors_for_rows(rows, &s));
	if (s > bf->words_capacity || !bf->words) {
		// BloomFilter::Initialize re-allocates on every call (bloom_filter.cpp:21); here the buffer only ever grows —
		// cudaFree / cudaMalloc of a 128-512 MiB block cost milliseconds each (measured: a rebuild that re-sizes two
		// filters took 21 ms with them, against 1.1 ms of kernels)
		if (bf->words) {
			RPT_CUDA(cudaFree(bf->words)); // implicit device sync: no kernel still touches the old words
			bf->words = nullptr;
			bf->words_capacity = 0;
		}
		RPT_CUDA(cudaMalloc(&bf->words, s * sizeof(uint64_t)));
		bf->words_capacity = s;
		bf->words_generation++; // IPC / peer mappings of the old allocation are stale, even at an equal address
	} else if (s != bf->num_sectors) {
		RPT_CUDA(cudaDeviceSynchronize()); // kernels on other streams may still read the words with the old S
	}
	bf->num_sectors = s;
	bf->sized_for_rows = rows;
	bf->content_version++;
	bf->rows_inserted.store(0);
	// std::fill_n(bf, num_sectors, 0), bloom_filter.cpp:24
	RPT_CUDA(cudaMemsetAsync(bf->words, 0, s * sizeof(uint64_t), bf->stream));
	RPT_CUDA(cudaStreamSynchronize(bf->stream));
	return 0;
}

int rpt_bf_create(rpt_bf_t **out, uint64_t est_rows, int device) {
	if (!out) {
		return set_error("null argument");
	}
	*out = nullptr;
	int ndev = 0;
	cudaError_t e = cudaGetDeviceCount(&ndev);
	if (e != cudaSuccess || ndev == 0) {
		return set_error("no CUDA device available (%s); librpt_gpu has no CPU fallback",
		                 e == cudaSuccess ? "device count is 0" : cudaGetErrorString(e));
	}
	if (device < 0 || device >= ndev) {
		return set_error("device %d out of range [0, %d)", device, ndev);
	}
	rpt_bf *bf = new (std::nothrow) rpt_bf();
	if (!bf) {
		return set_error("out of host memory");
	}
	int rc = fill_device_info(device, &bf->dev);
	if (rc == 0) {
		DeviceGuard guard(device);
		if (!guard.ok) {
			rc = set_error("cannot select CUDA device %d", device);
		} else {
			auto make = [&]() -> int {
				RPT_CUDA(cudaStreamCreateWithFlags(&bf->stream, cudaStreamNonBlocking));
				RPT_CUDA(cudaStreamCreateWithFlags(&bf->h2d, cudaStreamNonBlocking));
				RPT_CUDA(cudaStreamCreateWithFlags(&bf->compute, cudaStreamNonBlocking));
				RPT_CUDA(cudaStreamCreateWithFlags(&bf->d2h, cudaStreamNonBlocking));
				for (auto &sl : bf->slot) {
					RPT_CUDA(cudaEventCreateWithFlags(&sl.uploaded, cudaEventDisableTiming));
					RPT_CUDA(cudaEventCreateWithFlags(&sl.computed, cudaEventDisableTiming));
					RPT_CUDA(cudaEventCreateWithFlags(&sl.drained, cudaEventDisableTiming));
					RPT_CUDA(cudaMalloc(&sl.dfound, sizeof(uint64_t)));
					RPT_CUDA(cudaMallocHost(&sl.hfound, sizeof(uint64_t)));
				}
				return allocate_words(bf, est_rows);
			};
			rc = make();
		}
	}
	if (rc != 0) {
		rpt_bf_destroy(bf);
		return rc;
	}
	*out = bf;
	return 0;
}

int rpt_bf_destroy(rpt_bf_t *bf) {
	if (!bf) {
		return 0;
	}
	DeviceGuard guard(bf->dev.device);
	rpt_bf_peer_close(bf);
	for (size_t r = 0; r < bf->peer_flags.size(); r++) {
		if (bf->peer_flags_ipc[r] && bf->peer_flags[r]) {
			cudaIpcCloseMemHandle(bf->peer_flags[r]);
		}
	}
	if (bf->sync_flags) {
		cudaFree(bf->sync_flags);
	}
	if (bf->words) {
		cudaFree(bf->words);
	}
	if (bf->mm_dev) {
		cudaFree(bf->mm_dev);
	}
	if (bf->folded) {
		cudaFree(bf->folded);
	}
	if (bf->fold_pop_dev) {
		cudaFree(bf->fold_pop_dev);
	}
	if (bf->fold_pop_host) {
		cudaFreeHost(bf->fold_pop_host);
	}
	if (bf->fold_stats_dev) {
		cudaFree(bf->fold_stats_dev);
	}
	if (bf->fold_stats_host) {
		cudaFreeHost(bf->fold_stats_host);
	}
	if (bf->fold_stats_ev) {
		cudaEventDestroy(bf->fold_stats_ev);
	}
	if (bf->fold_ready) {
		cudaEventDestroy(bf->fold_ready);
	}
	for (cudaStream_t s : {bf->stream, bf->h2d, bf->compute, bf->d2h}) {
		if (s) {
			cudaStreamSynchronize(s);
			cudaStreamDestroy(s);
		}
	}
	for (auto &sl : bf->slot) {
		for (cudaEvent_t ev : {sl.uploaded, sl.computed, sl.drained}) {
			if (ev) {
				cudaEventDestroy(ev);
			}
		}
		if (sl.dsel) {
			cudaFree(sl.dsel);
		}
		if (sl.arena.base) {
			cudaFree(sl.arena.base);
		}
		if (sl.dfound) {
			cudaFree(sl.dfound);
		}
		if (sl.hfound) {
			cudaFreeHost(sl.hfound);
		}
	}
	delete bf;
	return 0;
}

int rpt_bf_reset(rpt_bf_t *bf, uint64_t est_rows, rpt_stream_t stream) {
	if (!bf) {
		return set_error("null filter handle");
	}
	RPT_GUARD(bf);
	uint64_t s;
	RPT_TRY(sectors_for_rows(est_rows, &s));
	if (s > bf->words_capacity || !bf->words) {
		if (stream) {
			RPT_CUDA(cudaStreamSynchronize(reinterpret_cast<cudaStream_t>(stream)));
		}
		RPT_TRY(allocate_words(bf, est_rows));
	} else {
		bf->num_sectors = s; // fits the buffer: stream-ordered like every other use of the words on `stream`
		bf->sized_for_rows = est_rows;
		bf->content_version++;
		bf->rows_inserted.store(0);
		RPT_CUDA(cudaMemsetAsync(bf->words, 0, s * sizeof(uint64_t), pick_stream(bf, stream)));
	}
	bf->has_data.store(0);
	bf->finalized.store(0);
	return 0;
}

#define RPT_GETTER(name, type, expr)                                                                                   \
	int name(const rpt_bf_t *bf, type *out) {                                                                          \
		if (!bf || !out) {                                                                                             \
			return set_error("null argument");                                                                         \
		}                                                                                                              \
		*out = (expr);                                                                                                 \
		return 0;                                                                                                      \
	}
RPT_GETTER(rpt_bf_num_sectors, uint64_t, bf->num_sectors)
RPT_GETTER(rpt_bf_sized_for_rows, uint64_t, bf->sized_for_rows)
RPT_GETTER(rpt_bf_is_empty, int, bf->has_data.load() ? 0 : 1)
RPT_GETTER(rpt_bf_is_finalized, int, bf->finalized.load())
RPT_GETTER(rpt_bf_device, int, bf->dev.device)
// whoever holds the address may write the words (NCCL all-gather targets, tests): the folded pre-filter is stale
RPT_GETTER(rpt_bf_device_words, uint64_t *,
           (const_cast<rpt_bf_t *>(bf)->content_version++, const_cast<rpt_bf_t *>(bf)->rows_inserted.store(-1), bf->words))

int rpt_bf_launch_count(const rpt_bf_t *bf, long long *out) {
	if (!bf || !out) {
		return set_error("null argument");
	}
	*out = bf->launches.load();
	return 0;
}

int rpt_bf_set_has_data(rpt_bf_t *bf, int has_data) {
	if (!bf) {
		return set_error("null argument");
	}
	bf->has_data.store(has_data ? 1 : 0);
	return 0;
}

int rpt_bf_sync(const rpt_bf_t *bf) {
	if (!bf) {
		return set_error("null argument");
	}
	RPT_GUARD(bf);
	RPT_CUDA(cudaStreamSynchronize(bf->stream));
	RPT_CUDA(cudaStreamSynchronize(bf->h2d));
	RPT_CUDA(cudaStreamSynchronize(bf->compute));
	RPT_CUDA(cudaStreamSynchronize(bf->d2h));
	return 0;
}

// ---------------------------------------------------------------------------------------------
// build
// ---------------------------------------------------------------------------------------------
// UpdateMinMax next to the insert (physical_create_bf.cpp:229-239): when range tracking is on for this column's
// physical type, the build kernel itself folds the chunk's valid rows into the range (template flag MM of
// build_kernel) — the keys are read once, no second pass.
static MinMaxState *minmax_target(rpt_bf *bf, const DevCols &dc) {
	if (!bf->mm_active.load() || dc.n != 1 || dc.c[0].type != bf->mm_type) {
		return nullptr;
	}
	return bf->mm_dev;
}

// Start of a host-column call (pipe_mu held, the pipeline idle): slots whose last chunks asked for more device memory
// than they hold are re-sized, so that from the second call of a kind on nothing is allocated chunk by chunk.
static int grow_slot_arenas(rpt_bf *bf) {
	for (auto &sl : bf->slot) {
		SlotArena &a = sl.arena;
		if (a.want > a.cap) {
			if (a.base) {
				RPT_CUDA(cudaFree(a.base)); // implicit device synchronisation
				a.base = nullptr;
				a.cap = 0;
			}
			const size_t cap = (a.want + (a.want >> 4) + 0xfffff) & ~static_cast<size_t>(0xfffff); // +6 %, whole MiB
			RPT_CUDA(cudaMalloc(&a.base, cap));
			a.cap = cap;
		}
	}
	return 0;
}

static int insert_impl(rpt_bf *bf, const rpt_col_t *cols, int ncols, uint64_t count, rpt_stream_t stream_in) {
	int mem = 0;
	RPT_TRY(validate_cols(cols, ncols, count, &mem));
	if (count == 0) {
		return 0; // src/bloom_filter.cpp:71-74: no-op, has_data_ untouched
	}
	bf->has_data.store(1); // src/bloom_filter.cpp:75
	bf->content_version++;
	if (bf->rows_inserted.load() >= 0) {
		bf->rows_inserted += static_cast<long long>(count);
	}
	int launches = 0;
	if (mem == RPT_MEM_DEVICE) {
		DevCols dc;
		device_cols_view(cols, ncols, &dc);
		Staged st;
		st.stream = pick_stream(bf, stream_in);
		PrehashPlan pre;
		RPT_TRY(plan_prehash(cols, ncols, count, dc, &st, &pre));
		MinMaxState *mm = minmax_target(bf, dc);
		RPT_TRY(run_prehash(bf->dev, pre, &dc, st.stream, &launches));
		void *hashes = nullptr;
		if (const size_t hb = build_scratch_bytes(dc, count, bf->num_sectors)) {
			RPT_TRY(st.alloc(&hashes, hb));
		}
		RPT_CUDA(launch_build(bf->dev, dc, count, bf->words, bf->num_sectors, pick_stream(bf, stream_in), &launches, mm,
		                      reinterpret_cast<uint64_t *>(hashes)));
		bf->launches += launches;
		return 0;
	}
	// host columns: chunk c is uploaded on the h2d stream while the insert kernel of chunk c-1 runs on the
	// compute stream (chained by events); at most kPipeSlots chunks are in flight; all chunks OR into the
	// same words
	std::lock_guard<std::mutex> lock(bf->pipe_mu);
	const double t0 = trace_on() ? now_ms() : 0;
	if (stream_in) {
		RPT_CUDA(cudaStreamSynchronize(reinterpret_cast<cudaStream_t>(stream_in)));
	} else {
		RPT_CUDA(cudaStreamSynchronize(bf->stream)); // a zero-fill enqueued by rpt_bf_reset(NULL stream)
	}
	RPT_TRY(grow_slot_arenas(bf));
	const uint64_t chunk = any_sel(cols, ncols) ? count : chunk_rows();
	uint64_t c = 0;
	for (uint64_t r0 = 0; r0 < count; r0 += chunk, c++) {
		const uint64_t r1 = r0 + chunk < count ? r0 + chunk : count;
		rpt_bf::PipeSlot &sl = bf->slot[c % rpt_bf::kPipeSlots];
		if (c >= static_cast<uint64_t>(rpt_bf::kPipeSlots)) {
			RPT_CUDA(cudaEventSynchronize(sl.computed)); // bounds the staging memory in flight
		}
		Staged st;
		st.stream = bf->h2d;
		st.free_stream = bf->compute;
		st.arena = &sl.arena;
		DevCols dc;
		RPT_TRY(stage_host_cols(cols, ncols, r0, r1, &st, &dc));
		PrehashPlan pre;
		RPT_TRY(plan_prehash(cols, ncols, r1 - r0, dc, &st, &pre));
		void *hashes = nullptr; // ordered on the upload stream, like the staged columns: before the `uploaded` event
		if (const size_t hb = build_scratch_bytes(dc, r1 - r0, bf->num_sectors)) {
			RPT_TRY(st.alloc(&hashes, hb));
		}
		RPT_CUDA(cudaEventRecord(sl.uploaded, bf->h2d));
		RPT_CUDA(cudaStreamWaitEvent(bf->compute, sl.uploaded, 0));
		st.handed_over = true;
		MinMaxState *mm = minmax_target(bf, dc); // before run_prehash may retype the column
		RPT_TRY(run_prehash(bf->dev, pre, &dc, bf->compute, &launches));
		RPT_CUDA(launch_build(bf->dev, dc, r1 - r0, bf->words, bf->num_sectors, bf->compute, &launches, mm,
		                      reinterpret_cast<uint64_t *>(hashes)));
		RPT_CUDA(cudaEventRecord(sl.computed, bf->compute));
	}
	bf->launches += launches;
	RPT_CUDA(cudaStreamSynchronize(bf->compute));
	if (trace_on()) {
		fprintf(stderr, "[rpt trace] insert host: %llu rows, %llu chunk(s), %.3f ms\n", (unsigned long long)count,
		        (unsigned long long)c, now_ms() - t0);
	}
	return 0;
}

int rpt_bf_insert(rpt_bf_t *bf, const rpt_col_t *cols, int ncols, uint64_t count, rpt_stream_t stream) {
	if (!bf) {
		return set_error("null filter handle");
	}
	RPT_GUARD(bf);
	return guarded([&] { return insert_impl(bf, cols, ncols, count, stream); });
}

// ---------------------------------------------------------------------------------------------
// finalize
// ---------------------------------------------------------------------------------------------
static bool minmax_type_ok(int type) {
	switch (type) {
	case RPT_BOOL: case RPT_INT8: case RPT_UINT8: case RPT_INT16: case RPT_UINT16: case RPT_INT32: case RPT_UINT32:
	case RPT_INT64: case RPT_UINT64: case RPT_FLOAT: case RPT_DOUBLE:
		return true;
	default:
		return false;
	}
}

int rpt_bf_minmax_begin(rpt_bf_t *bf, int type) {
	if (!bf) {
		return set_error("null filter handle");
	}
	if (!minmax_type_ok(type)) {
		return set_error("min/max is tracked for fixed-width numeric key columns only (physical type %d)", type);
	}
	RPT_GUARD(bf);
	if (!bf->mm_dev) {
		RPT_CUDA(cudaMalloc(&bf->mm_dev, sizeof(MinMaxState)));
	}
	MinMaxState init {~0ULL, 0ULL, 0u, 0u};
	RPT_CUDA(cudaDeviceSynchronize()); // inserts may have been enqueued on caller streams
	RPT_CUDA(cudaMemcpy(bf->mm_dev, &init, sizeof(init), cudaMemcpyHostToDevice));
	bf->mm_type = type;
	bf->mm_active.store(1);
	return 0;
}

int rpt_bf_minmax_end(rpt_bf_t *bf) {
	if (!bf) {
		return set_error("null filter handle");
	}
	bf->mm_active.store(0);
	return 0;
}

int rpt_bf_minmax_get(const rpt_bf_t *bf, int *has_value, int *has_nan, void *min_out, void *max_out) {
	if (!bf || !has_value || !min_out || !max_out) {
		return set_error("null argument");
	}
	if (!bf->mm_dev) {
		return set_error("rpt_bf_minmax_begin was not called on this filter");
	}
	RPT_GUARD(bf);
	RPT_CUDA(cudaDeviceSynchronize()); // inserts may have been enqueued on caller streams
	MinMaxState st;
	RPT_CUDA(cudaMemcpy(&st, bf->mm_dev, sizeof(st), cudaMemcpyDeviceToHost));
	*has_value = st.has_value ? 1 : 0;
	if (has_nan) {
		*has_nan = st.has_nan ? 1 : 0;
	}
	if (!st.has_value) {
		return 0;
	}
	constexpr uint64_t kSign = 1ULL << 63;
	auto decode = [&](uint64_t key, void *out) {
		switch (bf->mm_type) {
		case RPT_BOOL: case RPT_UINT8: { uint8_t v = static_cast<uint8_t>(key); memcpy(out, &v, 1); break; }
		case RPT_INT8: { int8_t v = static_cast<int8_t>(static_cast<int64_t>(key ^ kSign)); memcpy(out, &v, 1); break; }
		case RPT_UINT16: { uint16_t v = static_cast<uint16_t>(key); memcpy(out, &v, 2); break; }
		case RPT_INT16: { int16_t v = static_cast<int16_t>(static_cast<int64_t>(key ^ kSign)); memcpy(out, &v, 2); break; }
		case RPT_UINT32: { uint32_t v = static_cast<uint32_t>(key); memcpy(out, &v, 4); break; }
		case RPT_INT32: { int32_t v = static_cast<int32_t>(static_cast<int64_t>(key ^ kSign)); memcpy(out, &v, 4); break; }
		case RPT_UINT64: { memcpy(out, &key, 8); break; }
		case RPT_INT64: { uint64_t v = key ^ kSign; memcpy(out, &v, 8); break; }
		case RPT_FLOAT: {
			const uint32_t o = static_cast<uint32_t>(key);
			const uint32_t b = (o & 0x80000000u) ? (o & 0x7fffffffu) : ~o;
			memcpy(out, &b, 4);
			break;
		}
		case RPT_DOUBLE: {
			const uint64_t b = (key & kSign) ? (key & ~kSign) : ~key;
			memcpy(out, &b, 8);
			break;
		}
		default:
			break;
		}
	};
	decode(st.min_key, min_out);
	decode(st.max_key, max_out);
	return 0;
}

int rpt_bf_needs_rebuild(const rpt_bf_t *bf, uint64_t actual_rows, int *out) {
	if (!bf || !out) {
		return set_error("null argument");
	}
	*out = 0;
	if (actual_rows == 0) { // physical_create_bf.cpp:387
		return 0;
	}
	uint64_t min_bits = bf->sized_for_rows * 12ULL; // :394
	if (min_bits < 512ULL) {
		min_bits = 512ULL;
	}
	const uint64_t allocated_bits = next_pow2(min_bits); // :395 (not capped at 2^32 sectors' worth)
	if (allocated_bits == 0) {
		return set_error("Can't find next power of 2 for %llu", (unsigned long long)min_bits);
	}
	*out = actual_rows * 8ULL > allocated_bits; // :396
	return 0;
}

int rpt_bf_reinitialize_and_rehash(rpt_bf_t *bf, uint64_t actual_rows, const rpt_col_t *cols, int ncols,
                                   uint64_t count, rpt_stream_t stream) {
	if (!bf) {
		return set_error("null filter handle");
	}
	RPT_GUARD(bf);
	if (stream) {
		RPT_CUDA(cudaStreamSynchronize(reinterpret_cast<cudaStream_t>(stream)));
	}
	RPT_TRY(allocate_words(bf, actual_rows));  // src/bloom_filter.cpp:38-39
	bf->has_data.store(count > 0 ? 1 : 0);     // :40
	if (count == 0) {                          // :42-44
		return 0;
	}
	return guarded([&] { return insert_impl(bf, cols, ncols, count, stream); }); // :46-57
}

int rpt_bf_finalize(rpt_bf_t *bf, uint64_t actual_rows, const rpt_col_t *cols, int ncols, uint64_t count,
                    int *rebuilt_out, rpt_stream_t stream) {
	if (!bf) {
		return set_error("null filter handle");
	}
	int rebuild = 0;
	RPT_TRY(rpt_bf_needs_rebuild(bf, actual_rows, &rebuild));
	if (rebuild) {
		if (!cols) {
			return set_error("filter sized for %llu rows holds %llu: the resize rule fires but no key columns were "
			                 "given to rehash from",
			                 (unsigned long long)bf->sized_for_rows, (unsigned long long)actual_rows);
		}
		RPT_TRY(rpt_bf_reinitialize_and_rehash(bf, actual_rows, cols, ncols, count, stream));
	}
	if (rebuilt_out) {
		*rebuilt_out = rebuild;
	}
	bf->finalized.store(1); // physical_create_bf.cpp:409-413
	return 0;
}

// ---------------------------------------------------------------------------------------------
// probe
// ---------------------------------------------------------------------------------------------
// Folded pre-filter for probes of 256 / 512 MiB filters (rpt_kernels.cuh Prefilter; DESIGN.md §4 regime 3c).  The
// fold is made lazily, by the first probe of ≥ 2^22 rows after a change of the words, on that probe's stream (so it
// is ordered behind the inserts / merges the caller ordered the probe behind); probes on other streams wait for its
// event.  Whether it is used depends on how many non-member keys would survive it (≤ RPT_PREFILTER_MAX_PASS, default
// 0.5), which follows from the fold's bit density: estimated from the number of rows inserted since the words were
// zeroed (no host synchronisation; the fold kernel's own bit count replaces the estimate once it has arrived), or,
// for imported / merged words, read from the fold after one stream synchronisation.  Results never depend on any of
// this — a key the folded filter rejects cannot pass the real one.  RPT_PREFILTER=0 switches it off, =2 forces it.
static int prefilter_mode() {
	static const int mode = [] {
		const char *e = getenv("RPT_PREFILTER");
		return e ? atoi(e) : 1;
	}();
	return mode;
}
// size of the folded filter: one L2-resident slice (64 MiB).  A 32-MiB fold would be fully resident but lets 35 % of
// the non-members through at config 2's density: 8.1 ms against 6.2 ms (profiles/prefilter_r2u.txt)
static uint64_t fold_sectors() {
	return kSliceSectors;
}
static double prefilter_max_pass() {
	static const double v = [] {
		const char *e = getenv("RPT_PREFILTER_MAX_PASS");
		return e ? atof(e) : 0.5;
	}();
	return v;
}

// Share of a probe's rows beyond which the pre-filter stops paying (measured, profiles/prefilter_r2v.txt): at four
// slices the pre-filtered and the plain partition meet at ≈ 30 % survivors (17 %: 6.3 against 7.4 ms; 54 %: 10.4
// against 8.3 ms), at eight slices — where the plain partition writes 7/8 of the keys as records — at ≈ 45 % (36 %:
// 8.5 against 9.4 ms; 66 %: 10.9 against 9.6 ms).  RPT_PREFILTER_MAX_ALIVE overrides both.
static double prefilter_max_alive(uint64_t num_sectors) {
	static const double v = [] {
		const char *e = getenv("RPT_PREFILTER_MAX_ALIVE");
		return e ? atof(e) : -1.0;
	}();
	if (v >= 0) {
		return v;
	}
	return num_sectors > 4 * kSliceSectors ? 0.45 : 0.30;
}

// *measure: the caller must call prefilter_measured(bf, stream) behind the last launch_probe of this call
static int ensure_prefilter(rpt_bf *bf, uint64_t count, cudaStream_t s, Prefilter *out, bool *use, bool *measure,
                            int *launches) {
	*use = false;
	*measure = false;
	const int own = bf->fold_mode.load();
	const int mode = own >= 0 ? own : prefilter_mode();
	if (mode == 0 || bf->gate.n != 0 || !prefilter_applicable(bf->num_sectors) || count <= (1ULL << 18)) {
		return 0;
	}
	const double max_pass = prefilter_max_pass();
	std::lock_guard<std::mutex> lock(bf->fold_mu);
	// feedback of an earlier pre-filtered probe, once its counters have arrived
	if (bf->fold_stats_pending && bf->fold_stats_recorded && cudaEventQuery(bf->fold_stats_ev) == cudaSuccess) {
		bf->fold_stats_pending = false;
		const uint64_t rows = bf->fold_stats_host[0], alive = bf->fold_stats_host[1];
		if (rows >= (1ULL << 20) && alive <= rows) {
			bf->fold_alive = static_cast<double>(alive) / static_cast<double>(rows);
		}
	}
	cudaGetLastError(); // cudaErrorNotReady of the query is not an error
	if (mode == 1 && count >= (1ULL << 22)) {
		if (bf->fold_holdoff > 0) {
			bf->fold_holdoff--;
			return 0;
		}
		if (bf->fold_alive > prefilter_max_alive(bf->num_sectors)) {
			// measured (profiles/prefilter_r2q.txt): with a third of the rows surviving, the pre-filtered partition and
			// its fuller buckets take as long as the plain partition; beyond, longer.  Try again 16 large probes on.
			bf->fold_alive = -1.0;
			bf->fold_holdoff = 15;
			return 0;
		}
	}
	const uint64_t version = bf->content_version.load();
	if (bf->fold_pop_pending && cudaEventQuery(bf->fold_ready) == cudaSuccess) { // the measured density has arrived
		bf->fold_pop_pending = false;
		if (bf->fold_version == version) {
			const double d = static_cast<double>(*bf->fold_pop_host) / (64.0 * static_cast<double>(fold_sectors()));
			bf->fold_pass = 0.91 * d * d * d * d + 0.09 * d * d * d;
		}
	}
	cudaGetLastError();
	if (bf->fold_version != version) {
		if (count < (1ULL << 22) && mode != 2) {
			return 0; // a small probe does not pay for a fold (read S words, write 64 MiB)
		}
		if (!bf->folded) {
			RPT_CUDA(cudaMalloc(&bf->folded, kSliceSectors * sizeof(uint64_t)));
			RPT_CUDA(cudaMalloc(&bf->fold_pop_dev, sizeof(uint64_t)));
			RPT_CUDA(cudaMallocHost(&bf->fold_pop_host, sizeof(uint64_t)));
			RPT_CUDA(cudaMalloc(&bf->fold_stats_dev, 2 * sizeof(uint64_t)));
			RPT_CUDA(cudaMallocHost(&bf->fold_stats_host, 2 * sizeof(uint64_t)));
			RPT_CUDA(cudaEventCreateWithFlags(&bf->fold_stats_ev, cudaEventDisableTiming));
			RPT_CUDA(cudaEventCreateWithFlags(&bf->fold_ready, cudaEventDisableTiming));
		}
		if (bf->fold_pop_pending) { // an earlier fold's count is still in flight into the same pinned word
			RPT_CUDA(cudaEventSynchronize(bf->fold_ready));
			bf->fold_pop_pending = false;
		}
		RPT_CUDA(launch_fold(bf->dev, bf->words, bf->num_sectors, bf->folded, fold_sectors(), bf->fold_pop_dev, s));
		RPT_CUDA(cudaMemcpyAsync(bf->fold_pop_host, bf->fold_pop_dev, sizeof(uint64_t), cudaMemcpyDeviceToHost, s));
		RPT_CUDA(cudaEventRecord(bf->fold_ready, s)); // probes on other streams wait for this event, not the host
		(*launches)++;
		const long long rows = bf->rows_inserted.load();
		double d;
		if (rows >= 0) {
			// only inserts since the words were zeroed: a key sets 4 of a word's 64 bits (with repetition), so a bit of
			// the folded filter is still clear with probability (1 - (1 - (63/64)^4) / FS)^rows.  Duplicate keys make
			// this an over-estimate of the density — the cautious side; the measured count replaces it when it arrives.
			d = 1.0 - exp(-static_cast<double>(rows) * 0.061052 / static_cast<double>(fold_sectors()));
			bf->fold_pop_pending = true;
		} else {
			// imported or merged words: nothing to estimate from — wait for the fold and read its bit count
			RPT_CUDA(cudaStreamSynchronize(s));
			d = static_cast<double>(*bf->fold_pop_host) / (64.0 * static_cast<double>(fold_sectors()));
		}
		// a key sets 4 bit positions, 3 distinct ones with probability ≈ 0.09 (two of its 6-bit fields collide)
		bf->fold_pass = 0.91 * d * d * d * d + 0.09 * d * d * d;
		bf->fold_version = version;
		if (trace_on()) {
			fprintf(stderr, "[rpt trace] fold: %llu -> %llu sectors, density %.4f (%s), estimated pass rate %.4f\n",
			        (unsigned long long)bf->num_sectors, (unsigned long long)fold_sectors(), d,
			        rows >= 0 ? "from the row count" : "measured", bf->fold_pass);
		}
	}
	if (bf->fold_pass <= max_pass || mode == 2) {
		RPT_CUDA(cudaStreamWaitEvent(s, bf->fold_ready, 0)); // a no-op on the stream that made the fold
		out->folded = bf->folded;
		out->sectors = fold_sectors();
		out->stats = nullptr;
		if (!bf->fold_stats_pending && count >= (1ULL << 22)) { // one measurement in flight at a time
			RPT_CUDA(cudaMemsetAsync(bf->fold_stats_dev, 0, 2 * sizeof(uint64_t), s));
			out->stats = bf->fold_stats_dev;
			bf->fold_stats_pending = true; // until prefilter_measured's event has fired
			bf->fold_stats_recorded = false;
			*measure = true;
		}
		*use = true;
	}
	return 0;
}
// behind the last pre-filtered launch of a call that was told to measure: bring the counters to the host
static int prefilter_measured(rpt_bf *bf, cudaStream_t s) {
	std::lock_guard<std::mutex> lock(bf->fold_mu);
	RPT_CUDA(cudaMemcpyAsync(bf->fold_stats_host, bf->fold_stats_dev, 2 * sizeof(uint64_t), cudaMemcpyDeviceToHost, s));
	RPT_CUDA(cudaEventRecord(bf->fold_stats_ev, s));
	bf->fold_stats_recorded = true;
	return 0;
}

int rpt_bf_set_prefilter(rpt_bf_t *bf, int mode) {
	if (!bf) {
		return set_error("null filter handle");
	}
	if (mode < -1 || mode > 2) {
		return set_error("prefilter mode must be -1 (default), 0 (off), 1 (by density) or 2 (always)");
	}
	std::lock_guard<std::mutex> lock(bf->fold_mu);
	bf->fold_mode.store(mode);
	bf->fold_alive = -1.0; // forget the feedback of the probes so far
	bf->fold_holdoff = 0;
	return 0;
}

int rpt_bf_prefilter_info(const rpt_bf_t *bf, int *state_out, double *pass_rate_out) {
	if (!bf || !state_out || !pass_rate_out) {
		return set_error("null argument");
	}
	rpt_bf *m = const_cast<rpt_bf *>(bf);
	std::lock_guard<std::mutex> lock(m->fold_mu);
	const bool made = m->fold_version == m->content_version.load();
	const int own = m->fold_mode.load();
	const int mode = own >= 0 ? own : prefilter_mode();
	*state_out = !made ? 0 : ((m->fold_pass <= prefilter_max_pass() || mode == 2) && mode != 0 ? 1 : 2);
	*pass_rate_out = made ? m->fold_pass : 1.0;
	return 0;
}

// translate_col0: emit in_sel[flat] instead of flat (BFTableFilter::Filter, bloom_filter.cpp:117-123);
// the incoming selection is cols[0].sel.
static int probe_impl(const rpt_bf *bf, const rpt_col_t *cols, int ncols, uint64_t count, uint32_t *sel_out,
                      uint64_t *found_out, int out_mem, rpt_stream_t stream_in, bool translate_col0) {
	int mem = 0;
	RPT_TRY(validate_cols(cols, ncols, count, &mem));
	if (!found_out || (count > 0 && !sel_out)) {
		return set_error("null output argument");
	}
	if (count >= (1ULL << 32)) {
		return set_error("count %llu does not fit the uint32 selection vector; split the batch",
		                 (unsigned long long)count);
	}
	if (out_mem == RPT_MEM_DEVICE && mem != RPT_MEM_DEVICE) {
		return set_error("device outputs require device key columns");
	}
	rpt_bf *mbf = const_cast<rpt_bf *>(bf);
	int launches = 0;

	if (mem == RPT_MEM_DEVICE) {
		cudaStream_t s = pick_stream(bf, stream_in);
		DevCols dc;
		device_cols_view(cols, ncols, &dc);
		const uint32_t *translate = translate_col0 ? dc.c[0].sel : nullptr;
		Staged st;
		st.stream = s;
		Prefilter fold;
		bool use_fold = false, measure = false;
		RPT_TRY(ensure_prefilter(mbf, count, s, &fold, &use_fold, &measure, &launches));
		const Prefilter *foldp = use_fold ? &fold : nullptr;
		void *scratch;
		RPT_TRY(st.alloc(&scratch, probe_scratch_bytes(count, bf->num_sectors, use_fold)));
		PrehashPlan pre;
		RPT_TRY(plan_prehash(cols, ncols, count, dc, &st, &pre));
		RPT_TRY(run_prehash(bf->dev, pre, &dc, s, &launches));
		if (out_mem == RPT_MEM_DEVICE) { // asynchronous (but for the one-off fold of a large filter, ensure_prefilter)
			RPT_CUDA(launch_probe(bf->dev, dc, count, bf->words, bf->num_sectors, sel_out, found_out, 0, translate,
			                      scratch, s, &launches, bf->gate.n ? &bf->gate : nullptr, foldp));
			if (measure) {
				RPT_TRY(prefilter_measured(mbf, s));
			}
			mbf->launches += launches;
			return 0;
		}
		void *dsel, *dfound;
		RPT_TRY(st.alloc(&dsel, count * sizeof(uint32_t)));
		RPT_TRY(st.alloc(&dfound, sizeof(uint64_t)));
		RPT_CUDA(launch_probe(bf->dev, dc, count, bf->words, bf->num_sectors, reinterpret_cast<uint32_t *>(dsel),
		                      reinterpret_cast<uint64_t *>(dfound), 0, translate, scratch, s, &launches,
		                      bf->gate.n ? &bf->gate : nullptr, foldp));
		if (measure) {
			RPT_TRY(prefilter_measured(mbf, s));
		}
		mbf->launches += launches;
		RPT_CUDA(cudaMemcpyAsync(found_out, dfound, sizeof(uint64_t), cudaMemcpyDeviceToHost, s));
		RPT_CUDA(cudaStreamSynchronize(s));
		if (*found_out) {
			RPT_CUDA(cudaMemcpyAsync(sel_out, dsel, *found_out * sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
			RPT_CUDA(cudaStreamSynchronize(s));
		}
		return 0;
	}

	// host columns → host outputs.  Three streams chained by events: chunk c uploads on h2d while chunk c-1 is
	// probed on compute and the survivors of chunk c-2 go out on d2h.  The host only learns a chunk's survivor
	// count two chunks later, so two uploads are always queued behind the one in flight and the H2D copy
	// engine never idles; `total` advances in chunk order, which keeps the selection vector ordered.
	std::lock_guard<std::mutex> lock(mbf->pipe_mu);
	const double t0 = trace_on() ? now_ms() : 0;
	if (stream_in) {
		RPT_CUDA(cudaStreamSynchronize(reinterpret_cast<cudaStream_t>(stream_in)));
	} else {
		RPT_CUDA(cudaStreamSynchronize(bf->stream));
	}
	*found_out = 0;
	if (count == 0) {
		return 0;
	}
	constexpr int kSlots = rpt_bf::kPipeSlots;
	RPT_TRY(grow_slot_arenas(mbf));
	const uint64_t chunk = any_sel(cols, ncols) ? count : chunk_rows();
	const uint64_t cap = chunk < count ? chunk : count;
	for (auto &sl : mbf->slot) { // grow-only device selection buffers, kept for the next call
		if (sl.sel_cap < cap) {
			if (sl.dsel) {
				RPT_CUDA(cudaFree(sl.dsel));
				sl.dsel = nullptr;
				sl.sel_cap = 0;
			}
			RPT_CUDA(cudaMalloc(&sl.dsel, cap * sizeof(uint32_t)));
			sl.sel_cap = cap;
		}
		sl.pending = false;
		if (chunk >= count) {
			break; // a single chunk uses slot 0 only
		}
	}
	Prefilter fold;
	bool use_fold = false, measure = false;
	RPT_TRY(ensure_prefilter(mbf, count, mbf->compute, &fold, &use_fold, &measure, &launches)); // by the call's row count
	const Prefilter *foldp = use_fold ? &fold : nullptr;
	uint64_t total = 0;
	auto drain = [&](uint64_t c) -> int { // chunk c's kernel finished → send its survivors out
		rpt_bf::PipeSlot &sl = mbf->slot[c % kSlots];
		if (!sl.pending) {
			return 0;
		}
		RPT_CUDA(cudaEventSynchronize(sl.computed));
		const uint64_t f = *sl.hfound;
		if (f) {
			RPT_CUDA(cudaMemcpyAsync(sel_out + total, sl.dsel, f * sizeof(uint32_t), cudaMemcpyDeviceToHost, mbf->d2h));
		}
		RPT_CUDA(cudaEventRecord(sl.drained, mbf->d2h));
		total += f;
		sl.pending = false;
		return 0;
	};
	double t_stage = 0, t_alloc = 0, t_launch = 0, t_drain = 0; // host time per part (RPT_TRACE)
	auto run = [&]() -> int {
		uint64_t c = 0;
		for (uint64_t r0 = 0; r0 < count; r0 += chunk, c++) {
			const uint64_t r1 = r0 + chunk < count ? r0 + chunk : count;
			rpt_bf::PipeSlot &sl = mbf->slot[c % kSlots];
			Staged st;
			st.stream = mbf->h2d;
			st.free_stream = mbf->compute;
			st.arena = &sl.arena;
			DevCols dc;
			double t = trace_on() ? now_ms() : 0;
			RPT_TRY(stage_host_cols(cols, ncols, r0, r1, &st, &dc));
			if (trace_on()) {
				t_stage += now_ms() - t;
				t = now_ms();
			}
			void *scratch;
			RPT_TRY(st.alloc(&scratch, probe_scratch_bytes(r1 - r0, bf->num_sectors, use_fold)));
			if (trace_on()) {
				t_alloc += now_ms() - t;
				t = now_ms();
			}
			PrehashPlan pre;
			RPT_TRY(plan_prehash(cols, ncols, r1 - r0, dc, &st, &pre));
			RPT_CUDA(cudaEventRecord(sl.uploaded, mbf->h2d));
			RPT_CUDA(cudaStreamWaitEvent(mbf->compute, sl.uploaded, 0));
			RPT_CUDA(cudaStreamWaitEvent(mbf->compute, sl.drained, 0)); // dsel of chunk c-3 has left the device
			RPT_TRY(run_prehash(bf->dev, pre, &dc, mbf->compute, &launches));
			const uint32_t *translate = translate_col0 ? dc.c[0].sel : nullptr;
			RPT_CUDA(launch_probe(bf->dev, dc, r1 - r0, bf->words, bf->num_sectors, sl.dsel, sl.dfound,
			                      static_cast<uint32_t>(r0), translate, scratch, mbf->compute, &launches,
			                      bf->gate.n ? &bf->gate : nullptr, foldp));
			RPT_CUDA(cudaMemcpyAsync(sl.hfound, sl.dfound, sizeof(uint64_t), cudaMemcpyDeviceToHost, mbf->compute));
			RPT_CUDA(cudaEventRecord(sl.computed, mbf->compute));
			sl.pending = true;
			if (trace_on()) {
				t_launch += now_ms() - t;
				t = now_ms();
			}
			if (c >= static_cast<uint64_t>(kSlots - 1)) {
				RPT_TRY(drain(c - (kSlots - 1))); // frees the slot the next chunk will use
			}
			if (trace_on()) {
				t_drain += now_ms() - t;
			}
		}
		if (measure) {
			RPT_TRY(prefilter_measured(mbf, mbf->compute));
		}
		for (uint64_t d = c >= static_cast<uint64_t>(kSlots - 1) ? c - (kSlots - 1) : 0; d < c; d++) {
			RPT_TRY(drain(d));
		}
		RPT_CUDA(cudaStreamSynchronize(mbf->d2h));
		if (trace_on()) {
			fprintf(stderr, "[rpt trace] probe host: %llu rows, %llu chunk(s), found %llu, %.3f ms (host time: stage + "
			        "upload calls %.1f, scratch alloc %.1f, launches %.1f, waiting for chunks %.1f; pre-filter %d)\n",
			        (unsigned long long)count, (unsigned long long)c, (unsigned long long)total, now_ms() - t0, t_stage,
			        t_alloc, t_launch, t_drain, use_fold ? 1 : 0);
		}
		return 0;
	};
	const int rc = run();
	if (rc != 0) { // leave the pipeline idle so that the handle stays usable
		cudaStreamSynchronize(mbf->h2d);
		cudaStreamSynchronize(mbf->compute);
		cudaStreamSynchronize(mbf->d2h);
		for (auto &sl : mbf->slot) {
			sl.pending = false;
		}
	}
	mbf->launches += launches;
	if (rc == 0) {
		*found_out = total;
	}
	return rc;
}

int rpt_bf_probe(const rpt_bf_t *bf, const rpt_col_t *cols, int ncols, uint64_t count, uint32_t *sel_out,
                 uint64_t *found_out, int out_mem, rpt_stream_t stream) {
	if (!bf) {
		return set_error("null filter handle");
	}
	RPT_GUARD(bf);
	return guarded([&] { return probe_impl(bf, cols, ncols, count, sel_out, found_out, out_mem, stream, false); });
}

int rpt_bf_filter_sel(const rpt_bf_t *bf, const rpt_col_t *key, const uint32_t *in_sel, uint64_t approved_count,
                      uint32_t *sel_out, uint64_t *found_out, int mem, rpt_stream_t stream) {
	if (!bf || !key) {
		return set_error("null argument");
	}
	RPT_GUARD(bf);
	rpt_col_t col = *key;
	if (col.mem != mem) {
		return set_error("key column and selection vectors must live in the same memory space");
	}
	if (in_sel) {
		if (col.sel) {
			return set_error("rpt_bf_filter_sel: the key column must not carry its own selection");
		}
		col.sel = in_sel; // HashInternal slices the keys by `sel` (bloom_filter.cpp:84-92)
	}
	return guarded(
	    [&] { return probe_impl(bf, &col, 1, approved_count, sel_out, found_out, mem, stream, in_sel != nullptr); });
}

static int rpt_bf_lookup_hashes_host_body(const rpt_bf_t *bf, const uint64_t *hashes, uint64_t count, uint8_t *hits) {
	if (!bf || (count && (!hashes || !hits))) {
		return set_error("null argument");
	}
	RPT_GUARD(bf);
	if (count == 0) {
		return 0;
	}
	Staged st;
	st.stream = bf->stream;
	void *dh, *dhits;
	RPT_TRY(st.upload(&dh, hashes, count * sizeof(uint64_t)));
	RPT_TRY(st.alloc(&dhits, count));
	RPT_CUDA(launch_lookup_hashes(bf->words, bf->num_sectors, reinterpret_cast<const uint64_t *>(dh), count,
	                              reinterpret_cast<uint8_t *>(dhits), st.stream));
	const_cast<rpt_bf *>(bf)->launches += 1;
	RPT_CUDA(cudaMemcpyAsync(hits, dhits, count, cudaMemcpyDeviceToHost, st.stream));
	RPT_CUDA(cudaStreamSynchronize(st.stream));
	return 0;
}
int rpt_bf_lookup_hashes_host(const rpt_bf_t *bf, const uint64_t *hashes, uint64_t count, uint8_t *hits) {
	return guarded([&] { return rpt_bf_lookup_hashes_host_body(bf, hashes, count, hits); });
}

// ---------------------------------------------------------------------------------------------
// zone maps (BFTableFilter::CheckStatistics, duckdb bloom_filter.cpp:138-196)
// ---------------------------------------------------------------------------------------------
extern "C++" {
namespace {
// The typed part of TemplatedCheckStatistics<T>: min > max, `max - min` has to be representable in T
// (TrySubtractOperator), at most 2048.  Everything is done on 128-bit two's-complement images of the values.
struct Wide {
	unsigned __int128 v;
	bool is_signed;
};
inline bool wide_less(const Wide &a, const Wide &b) {
	return a.is_signed ? static_cast<__int128>(a.v) < static_cast<__int128>(b.v) : a.v < b.v;
}
template <class T>
inline Wide widen(const void *base, uint64_t i) {
	T x;
	memcpy(&x, static_cast<const char *>(base) + i * sizeof(T), sizeof(T));
	const bool sg = std::is_signed<T>::value;
	return Wide {sg ? static_cast<unsigned __int128>(static_cast<__int128>(x)) : static_cast<unsigned __int128>(x), sg};
}
inline Wide widen128(const void *base, uint64_t i, bool sg) {
	uint64_t w[2];
	memcpy(w, static_cast<const char *>(base) + i * 16, 16);
	return Wide {(static_cast<unsigned __int128>(w[1]) << 64) | w[0], sg};
}
// largest difference T can hold, by type
inline unsigned __int128 max_difference(int type) {
	switch (type) {
	case RPT_UINT8:
		return 0xffu;
	case RPT_INT8:
		return 0x7fu;
	case RPT_UINT16:
		return 0xffffu;
	case RPT_INT16:
		return 0x7fffu;
	case RPT_UINT32:
		return 0xffffffffu;
	case RPT_INT32:
		return 0x7fffffffu;
	case RPT_UINT64:
		return ~0ULL;
	case RPT_INT64:
		return ~0ULL >> 1;
	case RPT_UINT128:
		return ~static_cast<unsigned __int128>(0);
	default: // RPT_INT128
		return ~static_cast<unsigned __int128>(0) >> 1;
	}
}
} // namespace
} // extern "C++"

static int rpt_bf_check_statistics_body(const rpt_bf_t *bf, int type, const void *mins, const void *maxs,
                            const uint8_t *has_minmax, uint64_t count, uint8_t *results_out) {
	if (!bf || (count && (!mins || !maxs || !results_out))) {
		return set_error("null argument");
	}
	RPT_GUARD(bf);
	if (count == 0) {
		return 0;
	}
	uint32_t hash_class;
	switch (type) {
	case RPT_UINT8:
	case RPT_INT8:
	case RPT_UINT16:
	case RPT_INT16:
	case RPT_UINT32:
	case RPT_INT32:
		hash_class = kStatsHash32;
		break;
	case RPT_UINT64:
	case RPT_INT64:
		hash_class = kStatsHash64;
		break;
	case RPT_UINT128:
	case RPT_INT128:
		hash_class = kStatsHash128;
		break;
	default: // bloom_filter.cpp:193-195: no pruning for any other physical type
		memset(results_out, RPT_NO_PRUNING_POSSIBLE, count);
		return 0;
	}
	std::vector<StatsRange> ranges(count);
	const unsigned __int128 limit = max_difference(type);
	for (uint64_t i = 0; i < count; i++) {
		Wide lo, hi;
		switch (type) {
		case RPT_UINT8:
			lo = widen<uint8_t>(mins, i), hi = widen<uint8_t>(maxs, i);
			break;
		case RPT_INT8:
			lo = widen<int8_t>(mins, i), hi = widen<int8_t>(maxs, i);
			break;
		case RPT_UINT16:
			lo = widen<uint16_t>(mins, i), hi = widen<uint16_t>(maxs, i);
			break;
		case RPT_INT16:
			lo = widen<int16_t>(mins, i), hi = widen<int16_t>(maxs, i);
			break;
		case RPT_UINT32:
			lo = widen<uint32_t>(mins, i), hi = widen<uint32_t>(maxs, i);
			break;
		case RPT_INT32:
			lo = widen<int32_t>(mins, i), hi = widen<int32_t>(maxs, i);
			break;
		case RPT_UINT64:
			lo = widen<uint64_t>(mins, i), hi = widen<uint64_t>(maxs, i);
			break;
		case RPT_INT64:
			lo = widen<int64_t>(mins, i), hi = widen<int64_t>(maxs, i);
			break;
		default:
			lo = widen128(mins, i, type == RPT_INT128), hi = widen128(maxs, i, type == RPT_INT128);
			break;
		}
		StatsRange &r = ranges[i];
		r.first_lower = static_cast<uint64_t>(lo.v);
		r.first_upper = static_cast<uint64_t>(lo.v >> 64);
		r.range = 0;
		r.hash_class = kStatsSkip;
		if ((has_minmax && !has_minmax[i]) || wide_less(hi, lo)) {
			continue; // no statistics, or invalid ones
		}
		const unsigned __int128 diff = hi.v - lo.v; // exact: lo <= hi, modulo 2^128
		if (diff > limit || diff > 2048) {
			continue; // the subtraction overflows T, or the range is too wide
		}
		r.range = static_cast<uint32_t>(diff);
		r.hash_class = hash_class;
	}
	Staged st;
	st.stream = bf->stream;
	void *dr, *dres;
	RPT_TRY(st.upload(&dr, ranges.data(), count * sizeof(StatsRange)));
	RPT_TRY(st.alloc(&dres, count));
	RPT_CUDA(launch_check_statistics(bf->words, bf->num_sectors, reinterpret_cast<const StatsRange *>(dr), count,
	                                 reinterpret_cast<uint8_t *>(dres), st.stream));
	const_cast<rpt_bf *>(bf)->launches += 1;
	RPT_CUDA(cudaMemcpyAsync(results_out, dres, count, cudaMemcpyDeviceToHost, st.stream));
	RPT_CUDA(cudaStreamSynchronize(st.stream));
	return 0;
}
int rpt_bf_check_statistics(const rpt_bf_t *bf, int type, const void *mins, const void *maxs,
                            const uint8_t *has_minmax, uint64_t count, uint8_t *results_out) {
	return guarded([&] { return rpt_bf_check_statistics_body(bf, type, mins, maxs, has_minmax, count, results_out); });
}

// ---------------------------------------------------------------------------------------------
// filter words
// ---------------------------------------------------------------------------------------------
int rpt_bf_export_words(const rpt_bf_t *bf, uint64_t *host_dst, uint64_t capacity_words) {
	if (!bf || !host_dst) {
		return set_error("null argument");
	}
	if (capacity_words < bf->num_sectors) {
		return set_error("destination holds %llu words, the filter has %llu", (unsigned long long)capacity_words,
		                 (unsigned long long)bf->num_sectors);
	}
	RPT_GUARD(bf);
	RPT_TRY(rpt_bf_sync(bf));
	RPT_CUDA(cudaMemcpyAsync(host_dst, bf->words, bf->num_sectors * sizeof(uint64_t), cudaMemcpyDeviceToHost,
	                         bf->stream));
	RPT_CUDA(cudaStreamSynchronize(bf->stream));
	return 0;
}

int rpt_bf_import_words(rpt_bf_t *bf, const uint64_t *host_src, uint64_t num_words) {
	if (!bf || !host_src) {
		return set_error("null argument");
	}
	if (num_words != bf->num_sectors) {
		return set_error("source holds %llu words, the filter has %llu", (unsigned long long)num_words,
		                 (unsigned long long)bf->num_sectors);
	}
	RPT_GUARD(bf);
	bf->content_version++;
	bf->rows_inserted.store(-1);
	RPT_CUDA(cudaMemcpyAsync(bf->words, host_src, num_words * sizeof(uint64_t), cudaMemcpyHostToDevice, bf->stream));
	RPT_CUDA(cudaStreamSynchronize(bf->stream));
	bf->has_data.store(1);
	return 0;
}

// FNV-1a-64 over the S words, one xor-then-multiply per WORD (SURVEY §8c's recipe for the known answers): the
// recurrence is sequential, so the words are brought to the host and folded there.  A checksum of the result, not
// part of the build / probe path.
int rpt_bf_checksum(const rpt_bf_t *bf, uint64_t *fnv1a_out) {
	if (!bf || !fnv1a_out) {
		return set_error("null argument");
	}
	RPT_GUARD(bf);
	return guarded([&]() -> int {
		std::vector<uint64_t> host(bf->num_sectors);
		RPT_TRY(rpt_bf_export_words(bf, host.data(), host.size()));
		uint64_t h = 0xcbf29ce484222325ULL;
		for (uint64_t w : host) {
			h = (h ^ w) * 0x100000001b3ULL;
		}
		*fnv1a_out = h;
		return 0;
	});
}

int rpt_bf_words_generation(const rpt_bf_t *bf, uint64_t *out) {
	if (!bf || !out) {
		return set_error("null argument");
	}
	*out = bf->words_generation.load();
	return 0;
}

// ---------------------------------------------------------------------------------------------
// multi-GPU OR-merge
// ---------------------------------------------------------------------------------------------
int rpt_bf_or_parts(rpt_bf_t *bf, const uint64_t *dev_parts, uint64_t nparts, rpt_stream_t stream) {
	if (!bf || (nparts && !dev_parts)) {
		return set_error("null argument");
	}
	RPT_GUARD(bf);
	if (nparts == 0) {
		return 0;
	}
	bf->content_version++;
	bf->rows_inserted.store(-1);
	RPT_CUDA(launch_or_parts(bf->dev, bf->words, dev_parts, bf->num_sectors, nparts, pick_stream(bf, stream)));
	bf->launches += 1;
	return 0;
}

int rpt_bf_ipc_export(const rpt_bf_t *bf, void *handle64) {
	if (!bf || !handle64) {
		return set_error("null argument");
	}
	static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
	RPT_GUARD(bf);
	cudaIpcMemHandle_t h;
	RPT_CUDA(cudaIpcGetMemHandle(&h, bf->words));
	memcpy(handle64, &h, sizeof(h));
	return 0;
}

static int rpt_bf_peer_open_body(rpt_bf_t *bf, const void *handles64, int nranks, int rank) {
	if (!bf || !handles64) {
		return set_error("null argument");
	}
	if (nranks < 1 || nranks > 16 || rank < 0 || rank >= nranks) {
		return set_error("nranks must be in [1,16] and rank in [0,nranks)");
	}
	RPT_GUARD(bf);
	RPT_TRY(rpt_bf_peer_close(bf));
	bf->peer_words.assign(nranks, nullptr);
	bf->peer_ipc.assign(nranks, false);
	bf->peer_rank = rank;
	for (int r = 0; r < nranks; r++) {
		if (r == rank) {
			bf->peer_words[r] = bf->words;
			continue;
		}
		cudaIpcMemHandle_t h;
		memcpy(&h, reinterpret_cast<const uint8_t *>(handles64) + 64 * r, sizeof(h));
		void *p = nullptr;
		RPT_CUDA(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
		bf->peer_words[r] = reinterpret_cast<uint64_t *>(p);
		bf->peer_ipc[r] = true;
	}
	return 0;
}
int rpt_bf_peer_open(rpt_bf_t *bf, const void *handles64, int nranks, int rank) {
	return guarded([&] { return rpt_bf_peer_open_body(bf, handles64, nranks, rank); });
}

// slice r of an S-word filter split N ways (even word boundaries for the 128-bit accesses)
static void peer_slice(uint64_t S, int n, int r, uint64_t *off, uint64_t *len) {
	const uint64_t a = (S * static_cast<uint64_t>(r) / n) & ~1ULL;
	const uint64_t b = r + 1 == n ? S : ((S * static_cast<uint64_t>(r + 1) / n) & ~1ULL);
	*off = a;
	*len = b - a;
}

// One phase of the two-phase merge restricted to the sectors [first, first + nsectors): inside the range rank r owns
// the r-th of N sub-slices.  Phase 0 (reduce-scatter-OR): my sub-slice |= every peer's copy of it, read over NVLink.
// Phase 1 (all-gather): pull every other rank's reduced sub-slice.
static int peer_merge_range(rpt_bf *bf, int phase, uint64_t first, uint64_t nsectors, cudaStream_t s) {
	const int n = static_cast<int>(bf->peer_words.size());
	const int rank = bf->peer_rank;
	bf->content_version++;
	bf->rows_inserted.store(-1);
	if (phase == 0) {
		PtrList srcs;
		srcs.n = 0;
		for (int r = 0; r < n; r++) {
			if (r != rank) {
				srcs.p[srcs.n++] = bf->peer_words[r];
			}
		}
		uint64_t off, len;
		peer_slice(nsectors, n, rank, &off, &len);
		RPT_CUDA(launch_or_ptrs(bf->dev, bf->words, srcs, first + off, len, s));
		bf->launches += 1;
		return 0;
	}
	if (phase == 1) {
		for (int k = 1; k < n; k++) {
			const int r = (rank + k) % n; // staggered so that ranks do not all read the same peer at once
			uint64_t off, len;
			peer_slice(nsectors, n, r, &off, &len);
			if (len) {
				RPT_CUDA(cudaMemcpyAsync(bf->words + first + off, bf->peer_words[r] + first + off, len * sizeof(uint64_t),
				                         cudaMemcpyDeviceToDevice, s));
			}
		}
		return 0;
	}
	return set_error("phase must be 0 or 1");
}

int rpt_bf_peer_merge_phase(rpt_bf_t *bf, int phase, rpt_stream_t stream) {
	if (!bf) {
		return set_error("null filter handle");
	}
	if (bf->peer_rank < 0) {
		return set_error("rpt_bf_peer_open has not been called");
	}
	RPT_GUARD(bf);
	return peer_merge_range(bf, phase, 0, bf->num_sectors, pick_stream(bf, stream));
}

int rpt_bf_peer_merge_range(rpt_bf_t *bf, int phase, uint64_t first_sector, uint64_t nsectors, rpt_stream_t stream) {
	if (!bf) {
		return set_error("null filter handle");
	}
	if (bf->peer_rank < 0) {
		return set_error("rpt_bf_peer_open has not been called");
	}
	if ((first_sector & 1) || (nsectors & 1) || first_sector + nsectors > bf->num_sectors) {
		return set_error("sector range [%llu, +%llu) must be even-aligned and inside the filter",
		                 (unsigned long long)first_sector, (unsigned long long)nsectors);
	}
	RPT_GUARD(bf);
	return peer_merge_range(bf, phase, first_sector, nsectors, pick_stream(bf, stream));
}

int rpt_bf_slice_sectors(uint64_t *out) {
	if (!out) {
		return set_error("null argument");
	}
	*out = kSliceSectors;
	return 0;
}

int rpt_bf_set_slice_events(rpt_bf_t *bf, void *const *events, int n) {
	if (!bf || n < 0 || n > 8 || (n > 0 && !events)) {
		return set_error("rpt_bf_set_slice_events: n must be in [0, 8]");
	}
	for (int i = 0; i < 8; i++) {
		bf->gate.ev[i] = i < n ? reinterpret_cast<cudaEvent_t>(events[i]) : nullptr;
	}
	bf->gate.n = n;
	return 0;
}

int rpt_bf_peer_close(rpt_bf_t *bf) {
	if (!bf) {
		return set_error("null filter handle");
	}
	for (size_t r = 0; r < bf->peer_words.size(); r++) {
		if (bf->peer_ipc[r] && bf->peer_words[r]) {
			cudaIpcCloseMemHandle(bf->peer_words[r]);
		}
	}
	bf->peer_words.clear();
	bf->peer_ipc.clear();
	bf->peer_rank = -1;
	return 0;
}

// ---- the whole OR-allreduce between processes, owned by the library ---------------------------------------------
// Barriers are 32-bit words in device memory: a rank signals by storing the merge's epoch into its slot of every
// peer's flag block (one tiny kernel over the IPC mappings, release at system scope) and waits for its own slots with
// cuStreamWaitValue32 (a stream memory operation: no kernel runs, nothing polls; a one-CTA polling kernel stands in
// where the driver entry point is missing).  Everything is enqueued on the caller's stream; the host never waits.
constexpr int kSyncRanks = 16, kSyncPhases = 3;

int rpt_bf_peer_flags_export(rpt_bf_t *bf, void *handle64) {
	if (!bf || !handle64) {
		return set_error("null argument");
	}
	RPT_GUARD(bf);
	if (!bf->sync_flags) {
		RPT_CUDA(cudaMalloc(&bf->sync_flags, kSyncRanks * kSyncPhases * sizeof(uint32_t)));
		RPT_CUDA(cudaMemset(bf->sync_flags, 0, kSyncRanks * kSyncPhases * sizeof(uint32_t)));
	}
	cudaIpcMemHandle_t h;
	RPT_CUDA(cudaIpcGetMemHandle(&h, bf->sync_flags));
	memcpy(handle64, &h, sizeof(h));
	return 0;
}

static int rpt_bf_peer_flags_open_body(rpt_bf_t *bf, const void *handles64, int nranks, int rank) {
	if (!bf || !handles64) {
		return set_error("null argument");
	}
	if (nranks < 1 || nranks > kSyncRanks || rank < 0 || rank >= nranks) {
		return set_error("nranks must be in [1,16] and rank in [0,nranks)");
	}
	if (!bf->sync_flags) {
		return set_error("rpt_bf_peer_flags_export has not been called on this handle");
	}
	if (!bf->peer_flags.empty()) {
		return set_error("the flag blocks of this handle are already open");
	}
	RPT_GUARD(bf);
	bf->peer_flags.assign(nranks, nullptr);
	bf->peer_flags_ipc.assign(nranks, false);
	for (int r = 0; r < nranks; r++) {
		if (r == rank) {
			bf->peer_flags[r] = bf->sync_flags;
			continue;
		}
		cudaIpcMemHandle_t h;
		memcpy(&h, reinterpret_cast<const uint8_t *>(handles64) + 64 * r, sizeof(h));
		void *p = nullptr;
		RPT_CUDA(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
		bf->peer_flags[r] = reinterpret_cast<uint32_t *>(p);
		bf->peer_flags_ipc[r] = true;
	}
	return 0;
}
int rpt_bf_peer_flags_open(rpt_bf_t *bf, const void *handles64, int nranks, int rank) {
	return guarded([&] { return rpt_bf_peer_flags_open_body(bf, handles64, nranks, rank); });
}

typedef CUresult (*WaitValue32Fn)(CUstream, CUdeviceptr, cuuint32_t, unsigned int);
static WaitValue32Fn wait_value32() {
	static const WaitValue32Fn fn = [] {
		if (getenv("RPT_SYNC_POLL") != nullptr) { // force the polling kernel (tests)
			return static_cast<WaitValue32Fn>(nullptr);
		}
		void *p = nullptr;
		cudaDriverEntryPointQueryResult st;
		if (cudaGetDriverEntryPoint("cuStreamWaitValue32", &p, cudaEnableDefault, &st) != cudaSuccess ||
		    st != cudaDriverEntryPointSuccess) {
			cudaGetLastError();
			p = nullptr;
		}
		return reinterpret_cast<WaitValue32Fn>(p);
	}();
	return fn;
}
// my slot in every peer's block of `phase` := value
static int sync_signal(rpt_bf *bf, int phase, uint32_t value, cudaStream_t s) {
	FlagList dst;
	dst.n = 0;
	const int n = static_cast<int>(bf->peer_flags.size());
	for (int r = 0; r < n; r++) {
		if (r != bf->peer_rank) {
			dst.p[dst.n++] = bf->peer_flags[r] + phase * kSyncRanks + bf->peer_rank;
		}
	}
	RPT_CUDA(launch_signal_flags(dst, value, s));
	bf->launches += 1;
	return 0;
}
// every peer's slot in my block of `phase` >= value (epochs only grow; bit 0 of a phase-0 word is the has-data flag)
static int sync_wait(rpt_bf *bf, int phase, uint32_t value, cudaStream_t s) {
	const int n = static_cast<int>(bf->peer_flags.size());
	uint32_t *mine = bf->sync_flags + phase * kSyncRanks;
	if (WaitValue32Fn wait = wait_value32()) {
		for (int r = 0; r < n; r++) {
			if (r != bf->peer_rank) {
				const CUresult rc = wait(reinterpret_cast<CUstream>(s), reinterpret_cast<CUdeviceptr>(mine + r), value,
				                         CU_STREAM_WAIT_VALUE_GEQ);
				if (rc != CUDA_SUCCESS) {
					return set_error("cuStreamWaitValue32 failed (%d)", static_cast<int>(rc));
				}
			}
		}
		return 0;
	}
	RPT_CUDA(launch_wait_flags(mine, n, bf->peer_rank, value, s));
	bf->launches += 1;
	return 0;
}

static int rpt_bf_peer_allreduce_body(rpt_bf_t *bf, int has_local_data, rpt_stream_t stream) {
	if (!bf) {
		return set_error("null filter handle");
	}
	if (bf->peer_rank < 0 || bf->peer_flags.size() != bf->peer_words.size()) {
		return set_error("rpt_bf_peer_open and rpt_bf_peer_flags_open must both have been called with the same ranks");
	}
	RPT_GUARD(bf);
	cudaStream_t s = pick_stream(bf, stream);
	if (bf->sync_release_pending) { // the peers of the previous merge have pulled my slice before I touch the words again
		RPT_TRY(sync_wait(bf, 2, bf->sync_epoch << 1, s));
		bf->sync_release_pending = false;
	}
	const uint32_t v = ++bf->sync_epoch << 1;
	RPT_TRY(sync_signal(bf, 0, v | (has_local_data ? 1u : 0u), s)); // my partial filter is complete
	RPT_TRY(sync_wait(bf, 0, v, s));
	RPT_TRY(peer_merge_range(bf, 0, 0, bf->num_sectors, s));        // my slice |= every peer's copy of it
	RPT_TRY(sync_signal(bf, 1, v, s));                              // my slice is reduced
	RPT_TRY(sync_wait(bf, 1, v, s));
	RPT_TRY(peer_merge_range(bf, 1, 0, bf->num_sectors, s));        // pull the other ranks' reduced slices
	RPT_TRY(sync_signal(bf, 2, v, s));                              // I have pulled: my peers may overwrite theirs
	bf->sync_release_pending = true; // due before MY next mutation of the words: rpt_bf_peer_release
	if (has_local_data) {
		bf->has_data.store(1);
	}
	return 0;
}
int rpt_bf_peer_allreduce(rpt_bf_t *bf, int has_local_data, rpt_stream_t stream) {
	return guarded([&] { return rpt_bf_peer_allreduce_body(bf, has_local_data, stream); });
}

int rpt_bf_peer_release(rpt_bf_t *bf, rpt_stream_t stream) {
	if (!bf) {
		return set_error("null filter handle");
	}
	if (!bf->sync_release_pending) {
		return 0;
	}
	RPT_GUARD(bf);
	RPT_TRY(sync_wait(bf, 2, bf->sync_epoch << 1, pick_stream(bf, stream)));
	bf->sync_release_pending = false;
	return 0;
}

int rpt_bf_peer_any_data(rpt_bf_t *bf, rpt_stream_t stream, int *out) {
	if (!bf || !out) {
		return set_error("null argument");
	}
	if (bf->peer_flags.empty()) {
		return set_error("rpt_bf_peer_flags_open has not been called");
	}
	RPT_GUARD(bf);
	cudaStream_t s = pick_stream(bf, stream);
	uint32_t host[kSyncRanks];
	RPT_CUDA(cudaMemcpyAsync(host, bf->sync_flags, sizeof(host), cudaMemcpyDeviceToHost, s));
	RPT_CUDA(cudaStreamSynchronize(s));
	int any = bf->has_data.load();
	for (size_t r = 0; r < bf->peer_flags.size(); r++) {
		if (static_cast<int>(r) != bf->peer_rank && (host[r] >> 1) == bf->sync_epoch) {
			any |= static_cast<int>(host[r] & 1u);
		}
	}
	if (any) {
		bf->has_data.store(1);
	}
	*out = any;
	return 0;
}

int rpt_bf_allreduce_or(rpt_bf_t **per_gpu, int ngpu) {
	if (!per_gpu || ngpu < 1 || ngpu > 16) {
		return set_error("ngpu must be in [1,16]");
	}
	for (int i = 0; i < ngpu; i++) {
		if (!per_gpu[i] || per_gpu[i]->num_sectors != per_gpu[0]->num_sectors) {
			return set_error("all partial filters must exist and have the same number of sectors");
		}
	}
	if (ngpu == 1) {
		return 0;
	}
	int prev = 0;
	cudaGetDevice(&prev);
	auto body = [&]() -> int {
		for (int i = 0; i < ngpu; i++) {
			RPT_CUDA(cudaSetDevice(per_gpu[i]->dev.device));
			RPT_TRY(rpt_bf_sync(per_gpu[i]));
			for (int j = 0; j < ngpu; j++) {
				if (per_gpu[j]->dev.device != per_gpu[i]->dev.device) {
					cudaError_t e = cudaDeviceEnablePeerAccess(per_gpu[j]->dev.device, 0);
					if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) {
						return set_error("peer access %d -> %d: %s", per_gpu[i]->dev.device, per_gpu[j]->dev.device,
						                 cudaGetErrorString(e));
					}
					cudaGetLastError();
				}
			}
			RPT_TRY(rpt_bf_peer_close(per_gpu[i])); // IPC mappings opened earlier (rpt_bf_peer_open) are released first
			per_gpu[i]->peer_words.assign(ngpu, nullptr);
			per_gpu[i]->peer_ipc.assign(ngpu, false);
			per_gpu[i]->peer_rank = i;
			for (int j = 0; j < ngpu; j++) {
				per_gpu[i]->peer_words[j] = per_gpu[j]->words;
			}
		}
		for (int phase = 0; phase < 2; phase++) {
			for (int i = 0; i < ngpu; i++) {
				RPT_TRY(rpt_bf_peer_merge_phase(per_gpu[i], phase, nullptr));
			}
			for (int i = 0; i < ngpu; i++) { // every GPU finishes a phase before any starts the next
				RPT_CUDA(cudaSetDevice(per_gpu[i]->dev.device));
				RPT_CUDA(cudaStreamSynchronize(per_gpu[i]->stream));
			}
		}
		int any = 0;
		for (int i = 0; i < ngpu; i++) {
			any |= per_gpu[i]->has_data.load();
		}
		for (int i = 0; i < ngpu; i++) {
			per_gpu[i]->has_data.store(any);
			rpt_bf_peer_close(per_gpu[i]);
		}
		return 0;
	};
	const int rc = guarded(body);
	cudaSetDevice(prev);
	return rc;
}

// ---------------------------------------------------------------------------------------------
// synthetic columns
// ---------------------------------------------------------------------------------------------
int rpt_synth_build_keys(uint64_t *dev_out, uint64_t first, uint64_t count, int device, rpt_stream_t stream) {
	DeviceGuard guard(device);
	if (!guard.ok) {
		return set_error("cannot select CUDA device %d", device);
	}
	RPT_CUDA(launch_synth_build(dev_out, first, count, reinterpret_cast<cudaStream_t>(stream)));
	return 0;
}

int rpt_synth_probe_keys(uint64_t *dev_out, uint64_t first, uint64_t count, uint64_t n_build, uint32_t sel_ppm,
                         int device, rpt_stream_t stream) {
	DeviceGuard guard(device);
	if (!guard.ok) {
		return set_error("cannot select CUDA device %d", device);
	}
	if (n_build == 0) {
		return set_error("n_build must be positive");
	}
	RPT_CUDA(launch_synth_probe(dev_out, first, count, n_build, sel_ppm, reinterpret_cast<cudaStream_t>(stream)));
	return 0;
}

// ---------------------------------------------------------------------------------------------
// README-described filter variant (reference README.md:23-32): Acero-style blocked bloom filter with the
// 1024-entry mask table and Fold.  NOT on the parity path — see rpt_acero.cu and rpt_gpu.h.
// ---------------------------------------------------------------------------------------------
struct rpt_abf {
	DeviceInfo dev;
	int log_num_blocks = 0;
	uint64_t num_blocks = 0;
	uint64_t *blocks = nullptr;  // allocation keeps its CreateEmpty size; Fold only shrinks num_blocks
	uint64_t *masks = nullptr;   // kNumMasks expanded 57-bit masks (8 KiB), staged into shared memory by the kernels
	uint64_t *counter = nullptr; // device scalar (popcount / found)
	cudaStream_t stream = nullptr;
	std::atomic<long long> launches {0};
};

static cudaStream_t abf_stream(const rpt_abf *f, rpt_stream_t s) {
	return s ? reinterpret_cast<cudaStream_t>(s) : f->stream;
}

// BlockedBloomFilter::CreateEmpty: 8 bits per key, at least 512 bits, rounded up to a power of two
static int abf_log_blocks_for_rows(uint64_t rows) {
	uint64_t min_bits = rows > (1ULL << 58) ? (1ULL << 61) : rows * 8;
	if (min_bits < 512) {
		min_bits = 512;
	}
	int log_bits = 0;
	while ((1ULL << log_bits) < min_bits) {
		log_bits++;
	}
	return log_bits - 6;
}

int rpt_abf_log_blocks_for_rows(uint64_t rows, int *out) {
	if (!out) {
		return set_error("null output argument");
	}
	*out = abf_log_blocks_for_rows(rows);
	return 0;
}

int rpt_abf_mask_bits(uint8_t *out136) {
	if (!out136) {
		return set_error("null output argument");
	}
	acero::generate_mask_bits(out136);
	return 0;
}

int rpt_abf_create(rpt_abf_t **out, uint64_t num_rows, int device) {
	if (!out) {
		return set_error("null output argument");
	}
	*out = nullptr;
	const int log_blocks = abf_log_blocks_for_rows(num_rows);
	if (log_blocks > 32) {
		return set_error("a filter for %llu rows would need 2^%d blocks", (unsigned long long)num_rows, log_blocks);
	}
	std::unique_ptr<rpt_abf> f(new rpt_abf());
	RPT_TRY(fill_device_info(device, &f->dev));
	RPT_GUARD(f);
	f->log_num_blocks = log_blocks;
	f->num_blocks = 1ULL << log_blocks;
	RPT_CUDA(cudaStreamCreateWithFlags(&f->stream, cudaStreamNonBlocking));
	RPT_CUDA(cudaMalloc(reinterpret_cast<void **>(&f->blocks), f->num_blocks * sizeof(uint64_t)));
	RPT_CUDA(cudaMalloc(reinterpret_cast<void **>(&f->masks), acero::kNumMasks * sizeof(uint64_t)));
	RPT_CUDA(cudaMalloc(reinterpret_cast<void **>(&f->counter), sizeof(uint64_t)));
	uint8_t bits[acero::kMaskBytes];
	std::vector<uint64_t> masks(acero::kNumMasks);
	acero::generate_mask_bits(bits);
	acero::expand_masks(bits, masks.data());
	RPT_CUDA(cudaMemcpyAsync(f->masks, masks.data(), masks.size() * sizeof(uint64_t), cudaMemcpyHostToDevice, f->stream));
	RPT_CUDA(cudaMemsetAsync(f->blocks, 0, f->num_blocks * sizeof(uint64_t), f->stream));
	RPT_CUDA(cudaStreamSynchronize(f->stream));
	*out = f.release();
	return 0;
}

int rpt_abf_destroy(rpt_abf_t *f) {
	if (!f) {
		return 0;
	}
	{
		RPT_GUARD(f);
		cudaStreamSynchronize(f->stream);
		cudaFree(f->blocks);
		cudaFree(f->masks);
		cudaFree(f->counter);
		cudaStreamDestroy(f->stream);
	}
	delete f;
	return 0;
}

int rpt_abf_log_num_blocks(const rpt_abf_t *f, int *out) {
	if (!f || !out) {
		return set_error("null argument");
	}
	*out = f->log_num_blocks;
	return 0;
}

int rpt_abf_launch_count(const rpt_abf_t *f, long long *out) {
	if (!f || !out) {
		return set_error("null argument");
	}
	*out = f->launches.load();
	return 0;
}

int rpt_abf_sync(const rpt_abf_t *f) {
	if (!f) {
		return set_error("null filter handle");
	}
	RPT_GUARD(f);
	RPT_CUDA(cudaStreamSynchronize(f->stream));
	return 0;
}

int rpt_abf_device_blocks(const rpt_abf_t *f, uint64_t **dev_ptr_out) {
	if (!f || !dev_ptr_out) {
		return set_error("null argument");
	}
	*dev_ptr_out = f->blocks;
	return 0;
}

// Insert for every row.  Device columns: asynchronous on `stream`.  Host columns: uploaded chunk by chunk on the
// handle's stream, returns when the rows are in the filter.
int rpt_abf_insert(rpt_abf_t *f, const rpt_col_t *cols, int ncols, uint64_t count, rpt_stream_t stream) {
	if (!f) {
		return set_error("null filter handle");
	}
	int mem = 0;
	RPT_TRY(validate_cols(cols, ncols, count, &mem));
	if (count == 0) {
		return 0;
	}
	RPT_GUARD(f);
	cudaStream_t s = abf_stream(f, stream);
	if (mem == RPT_MEM_DEVICE) {
		DevCols dc;
		device_cols_view(cols, ncols, &dc);
		int launches = 0;
		RPT_CUDA(acero::launch_build(f->dev, dc, count, f->blocks, f->num_blocks, f->masks, s, &launches));
		f->launches += launches;
		return 0;
	}
	const uint64_t chunk = any_sel(cols, ncols) ? count : chunk_rows();
	for (uint64_t r0 = 0; r0 < count; r0 += chunk) {
		const uint64_t r1 = r0 + chunk < count ? r0 + chunk : count;
		Staged st;
		st.stream = s;
		DevCols dc;
		RPT_TRY(stage_host_cols(cols, ncols, r0, r1, &st, &dc));
		int launches = 0;
		RPT_CUDA(acero::launch_build(f->dev, dc, r1 - r0, f->blocks, f->num_blocks, f->masks, s, &launches));
		f->launches += launches;
	}
	RPT_CUDA(cudaStreamSynchronize(s));
	return 0;
}

// BlockedBloomFilter::Fold: while fewer than a quarter of the bits are set (and more than 2^4 blocks remain), OR
// 2^k equal slices into the first one, k chosen so that the expected density stays under a quarter.  Needs the
// popcount on the host, so it synchronises `stream`.  *num_folds_out = total number of halvings.
int rpt_abf_fold(rpt_abf_t *f, rpt_stream_t stream, int *num_folds_out) {
	if (!f) {
		return set_error("null filter handle");
	}
	RPT_GUARD(f);
	cudaStream_t s = abf_stream(f, stream);
	int total_folds = 0;
	for (;;) {
		if (f->log_num_blocks <= acero::kLogNumBlocksMin) {
			break;
		}
		const uint64_t num_bits = f->num_blocks * 64;
		uint64_t bits_set = 0;
		RPT_CUDA(acero::launch_popcount(f->dev, f->blocks, f->num_blocks, f->counter, s));
		RPT_CUDA(cudaMemcpyAsync(&bits_set, f->counter, sizeof(uint64_t), cudaMemcpyDeviceToHost, s));
		RPT_CUDA(cudaStreamSynchronize(s));
		f->launches += 1;
		if (4 * bits_set >= num_bits) {
			break;
		}
		int num_folds = 1;
		while ((f->log_num_blocks - num_folds) > acero::kLogNumBlocksMin && (4 * bits_set) < (num_bits >> num_folds)) {
			num_folds++;
		}
		RPT_CUDA(acero::launch_fold(f->dev, f->blocks, f->num_blocks, num_folds, s));
		f->launches += 1;
		f->log_num_blocks -= num_folds;
		f->num_blocks = 1ULL << f->log_num_blocks;
		total_folds += num_folds;
	}
	if (num_folds_out) {
		*num_folds_out = total_folds;
	}
	return 0;
}

// Find for every row → ascending row ids of the rows that pass.  out_mem = RPT_MEM_DEVICE: device outputs, device
// columns, asynchronous.  RPT_MEM_HOST: host outputs (columns in either space), returns when sel_out is filled.
int rpt_abf_probe(const rpt_abf_t *f, const rpt_col_t *cols, int ncols, uint64_t count, uint32_t *sel_out,
                  uint64_t *found_out, int out_mem, rpt_stream_t stream) {
	if (!f) {
		return set_error("null filter handle");
	}
	int mem = 0;
	RPT_TRY(validate_cols(cols, ncols, count, &mem));
	if (!found_out || (count > 0 && !sel_out)) {
		return set_error("null output argument");
	}
	if (count >= (1ULL << 32)) {
		return set_error("count %llu does not fit the uint32 selection vector; split the batch",
		                 (unsigned long long)count);
	}
	if (out_mem == RPT_MEM_DEVICE && mem != RPT_MEM_DEVICE) {
		return set_error("device outputs require device key columns");
	}
	RPT_GUARD(f);
	rpt_abf *mf = const_cast<rpt_abf *>(f);
	cudaStream_t s = abf_stream(f, stream);
	int launches = 0;
	if (out_mem == RPT_MEM_DEVICE) {
		DevCols dc;
		device_cols_view(cols, ncols, &dc);
		Staged st;
		st.stream = s;
		void *scratch;
		RPT_TRY(st.alloc(&scratch, bitmap_scratch_bytes(count)));
		RPT_CUDA(acero::launch_probe(f->dev, dc, count, f->blocks, f->num_blocks, f->masks, sel_out, found_out, 0,
		                             scratch, s, &launches));
		mf->launches += launches;
		return 0;
	}
	*found_out = 0;
	const uint64_t chunk = (mem == RPT_MEM_DEVICE || any_sel(cols, ncols)) ? count : chunk_rows();
	uint64_t total = 0;
	for (uint64_t r0 = 0; r0 < count; r0 += chunk) {
		const uint64_t r1 = r0 + chunk < count ? r0 + chunk : count;
		const uint64_t n = r1 - r0;
		Staged st;
		st.stream = s;
		DevCols dc;
		if (mem == RPT_MEM_DEVICE) {
			device_cols_view(cols, ncols, &dc);
		} else {
			RPT_TRY(stage_host_cols(cols, ncols, r0, r1, &st, &dc));
		}
		void *scratch, *dsel, *dfound;
		RPT_TRY(st.alloc(&scratch, bitmap_scratch_bytes(n)));
		RPT_TRY(st.alloc(&dsel, n * sizeof(uint32_t)));
		RPT_TRY(st.alloc(&dfound, sizeof(uint64_t)));
		RPT_CUDA(acero::launch_probe(f->dev, dc, n, f->blocks, f->num_blocks, f->masks,
		                             reinterpret_cast<uint32_t *>(dsel), reinterpret_cast<uint64_t *>(dfound),
		                             static_cast<uint32_t>(r0), scratch, s, &launches));
		uint64_t found = 0;
		RPT_CUDA(cudaMemcpyAsync(&found, dfound, sizeof(uint64_t), cudaMemcpyDeviceToHost, s));
		RPT_CUDA(cudaStreamSynchronize(s));
		if (found) {
			RPT_CUDA(cudaMemcpyAsync(sel_out + total, dsel, found * sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
			RPT_CUDA(cudaStreamSynchronize(s));
		}
		total += found;
	}
	mf->launches += launches;
	*found_out = total;
	return 0;
}

// Multi-GPU merge (SURVEY §8e applied to the variant): blocks |= parts[0 .. nparts), each part a same-sized partial
// filter on this device (e.g. the receive buffer of an all-gather).  Insert is a plain OR, so the merged blocks
// equal a single build; Fold runs after the merge, on every rank alike.
int rpt_abf_or_parts(rpt_abf_t *f, const uint64_t *dev_parts, uint64_t nparts, rpt_stream_t stream) {
	if (!f || (nparts && !dev_parts)) {
		return set_error("null argument");
	}
	RPT_GUARD(f);
	if (nparts == 0) {
		return 0;
	}
	RPT_CUDA(launch_or_parts(f->dev, f->blocks, dev_parts, f->num_blocks, nparts, abf_stream(f, stream)));
	f->launches += 1;
	return 0;
}

int rpt_abf_export_blocks(const rpt_abf_t *f, uint64_t *host_dst, uint64_t capacity_blocks) {
	if (!f || !host_dst) {
		return set_error("null argument");
	}
	if (capacity_blocks < f->num_blocks) {
		return set_error("destination holds %llu blocks, the filter has %llu", (unsigned long long)capacity_blocks,
		                 (unsigned long long)f->num_blocks);
	}
	RPT_GUARD(f);
	RPT_CUDA(cudaStreamSynchronize(f->stream));
	RPT_CUDA(cudaMemcpy(host_dst, f->blocks, f->num_blocks * sizeof(uint64_t), cudaMemcpyDeviceToHost));
	return 0;
}


} // extern "C"
