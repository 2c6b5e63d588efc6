// rpt_gpu.hpp — host-side C++ mirror of the reference's bloom-filter interface over the C ABI (rpt_gpu.h).
//
// Header-only, C++17, no DuckDB and no CUDA headers needed: this is the code a maintainer drops into the rpt
// extension, with DuckDB's DataChunk/Vector replaced by the plain pieces they decompose into
// (rpt_col_t = data / sel / validity of a UnifiedVectorFormat).  Same names, argument meaning and error
// behaviour as the reference:
//
//   rpt::PTBloomFilter   duckdb::PTBloomFilter                 src/include/bloom_filter.hpp:22-57, src/bloom_filter.cpp:27-78
//   rpt::CreateBFBatch   PhysicalCreateBF::Sink / Finalize     src/operators/physical_create_bf.cpp:200-245, 372-413
//   rpt::UseBFBatch      PhysicalUseBF::ExecuteInternal        src/operators/physical_use_bf.cpp:125-195
//
// Errors: the reference throws duckdb exceptions; every non-zero status of the C ABI is re-thrown here as
// rpt::Error carrying rpt_gpu_last_error().  There is no CPU fallback: without a CUDA device Initialize throws.
#pragma once

#include <algorithm>
#include <cstdint>
#include <cstring>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

#include "rpt_gpu.h"

namespace rpt {

struct Error : std::runtime_error {
	using std::runtime_error::runtime_error;
};

inline void Check(int rc) {
	if (rc != 0) {
		throw Error(std::string("rpt_gpu: ") + rpt_gpu_last_error());
	}
}

constexpr uint64_t kStandardVectorSize = 2048; // STANDARD_VECTOR_SIZE: rows per DataChunk

inline size_t TypeWidth(int type) {
	switch (type) {
	case RPT_BOOL: case RPT_INT8: case RPT_UINT8: return 1;
	case RPT_INT16: case RPT_UINT16: return 2;
	case RPT_INT32: case RPT_UINT32: case RPT_FLOAT: return 4;
	case RPT_INT64: case RPT_UINT64: case RPT_DOUBLE: return 8;
	case RPT_INTERVAL: case RPT_INT128: case RPT_UINT128: case RPT_STRING_T: return 16;
	default: return 0;
	}
}

// One key column of one chunk, host memory.  `sel` / `validity` as in UnifiedVectorFormat (nullptr = identity /
// all valid); data_len = entries addressable through sel (dictionary size), 0 for flat vectors.
inline rpt_col_t HostColumn(const void *data, int type, const uint32_t *sel = nullptr,
                            const uint64_t *validity = nullptr, uint64_t data_len = 0) {
	rpt_col_t c {};
	c.data = data;
	c.sel = sel;
	c.validity = validity;
	c.data_len = data_len;
	c.type = type;
	c.mem = RPT_MEM_HOST;
	return c;
}

// Page-locked host buffer (rpt_host_alloc): full-rate, overlapped H2D needs pinned memory.
class PinnedBuffer {
public:
	PinnedBuffer() = default;
	PinnedBuffer(const PinnedBuffer &) = delete;
	PinnedBuffer &operator=(const PinnedBuffer &) = delete;
	~PinnedBuffer() {
		if (ptr_) {
			rpt_host_free(ptr_);
		}
	}
	// grows geometrically, keeps the first `keep` bytes
	void Reserve(size_t bytes, size_t keep) {
		if (bytes <= cap_) {
			return;
		}
		size_t cap = std::max<size_t>(bytes, std::max<size_t>(cap_ * 2, size_t(1) << 20));
		void *p = nullptr;
		Check(rpt_host_alloc(&p, cap));
		if (keep && ptr_) {
			memcpy(p, ptr_, std::min(keep, cap_));
		}
		if (ptr_) {
			rpt_host_free(ptr_);
		}
		ptr_ = p;
		cap_ = cap;
	}
	uint8_t *data() const {
		return static_cast<uint8_t *>(ptr_);
	}

private:
	void *ptr_ = nullptr;
	size_t cap_ = 0;
};

// ------------------------------------------------------------------------------------------------------------
// duckdb::PTBloomFilter (src/include/bloom_filter.hpp:22-57)
// ------------------------------------------------------------------------------------------------------------
class PTBloomFilter {
public:
	explicit PTBloomFilter(int device = 0) : device_(device) {
	}
	PTBloomFilter(const PTBloomFilter &) = delete;
	PTBloomFilter &operator=(const PTBloomFilter &) = delete;
	~PTBloomFilter() {
		if (bf_) {
			rpt_bf_destroy(bf_);
		}
	}

	// src/bloom_filter.cpp:27-32.  est_num_rows is uint32_t in the reference (wraps above 2^32): kept.
	void Initialize(uint32_t est_num_rows) {
		if (bf_) {
			Check(rpt_bf_reset(bf_, est_num_rows, nullptr));
		} else {
			Check(rpt_bf_create(&bf_, est_num_rows, device_));
		}
	}
	// src/bloom_filter.cpp:70-78: count == 0 is a no-op; otherwise has_data_ = true and the hashed rows are OR-ed in.
	void Insert(const rpt_col_t *cols, int ncols, uint64_t count) {
		Check(rpt_bf_insert(bf_, cols, ncols, count, nullptr));
	}
	// src/bloom_filter.cpp:60-68: ascending row ids of the passing rows in sel[0..found), returns found.
	uint64_t LookupSel(const rpt_col_t *cols, int ncols, uint64_t count, uint32_t *sel) const {
		if (count == 0) {
			return 0;
		}
		uint64_t found = 0;
		Check(rpt_bf_probe(bf_, cols, ncols, count, sel, &found, RPT_MEM_HOST, nullptr));
		return found;
	}
	// src/bloom_filter.cpp:34-58
	void ReinitializeAndRehash(uint64_t actual_rows, const rpt_col_t *cols, int ncols, uint64_t count) {
		Check(rpt_bf_reinitialize_and_rehash(bf_, actual_rows, cols, ncols, count, nullptr));
	}
	// resize rule + finalized_ = true of PhysicalCreateBF::Finalize (physical_create_bf.cpp:386-413)
	bool Finalize(uint64_t actual_rows, const rpt_col_t *cols = nullptr, int ncols = 0, uint64_t count = 0) {
		int rebuilt = 0;
		Check(rpt_bf_finalize(bf_, actual_rows, cols, ncols, count, &rebuilt, nullptr));
		return rebuilt != 0;
	}
	bool NeedsRebuild(uint64_t actual_rows) const {
		int out = 0;
		Check(rpt_bf_needs_rebuild(bf_, actual_rows, &out));
		return out != 0;
	}
	uint64_t SizedForRows() const {
		uint64_t v = 0;
		Check(rpt_bf_sized_for_rows(bf_, &v));
		return v;
	}
	bool IsEmpty() const {
		int v = 1;
		Check(rpt_bf_is_empty(bf_, &v));
		return v != 0;
	}
	bool finalized() const { // the reference's public member finalized_
		int v = 0;
		Check(rpt_bf_is_finalized(bf_, &v));
		return v != 0;
	}
	uint64_t NumSectors() const {
		uint64_t v = 0;
		Check(rpt_bf_num_sectors(bf_, &v));
		return v;
	}
	// GetNativeFilter().bf: the S words, byte-identical to DuckDB's array (scan-side BFTableFilter keeps working)
	void ExportWords(uint64_t *dst, uint64_t capacity_words) const {
		Check(rpt_bf_export_words(bf_, dst, capacity_words));
	}
	// Folded pre-filter of large probes (no reference counterpart; rpt_gpu.h): -1 library default, 0 off, 1 by
	// density, 2 always.  Never changes LookupSel's result.
	void SetPrefilter(int mode) {
		Check(rpt_bf_set_prefilter(bf_, mode));
	}
	// state: 0 no fold for the current words, 1 in use, 2 made but too dense
	int PrefilterState(double *pass_rate = nullptr) const {
		int state = 0;
		double rate = 1.0;
		Check(rpt_bf_prefilter_info(bf_, &state, &rate));
		if (pass_rate) {
			*pass_rate = rate;
		}
		return state;
	}
	// BFTableFilter::CheckStatistics (duckdb bloom_filter.cpp:138-196) for `count` row groups at once: mins/maxs are
	// arrays of the column's physical type, results receive enum rpt_prune codes
	void CheckStatistics(int physical_type, const void *mins, const void *maxs, const uint8_t *has_minmax,
	                     uint64_t count, uint8_t *results) const {
		Check(rpt_bf_check_statistics(bf_, physical_type, mins, maxs, has_minmax, count, results));
	}
	rpt_bf_t *handle() const {
		return bf_;
	}

private:
	int device_;
	rpt_bf_t *bf_ = nullptr;
};

// ------------------------------------------------------------------------------------------------------------
// Scan-side mirror of duckdb::BFTableFilter (duckdb/src/planner/filter/bloom_filter.cpp:84-196): the filter a
// forward-pass CREATE_BF pushes into the probe table's scan (physical_create_bf.cpp:321-332).  Same three
// entry points, on the GPU-resident words.
// ------------------------------------------------------------------------------------------------------------
class BFTableFilter {
public:
	explicit BFTableFilter(const PTBloomFilter &filter) : filter_(filter) {
	}
	// Filter (:94-131).  `keys` is the scan's key vector (no selection of its own); `sel` holds the
	// approved_count row ids still alive when sel_is_set, and has room for approved_count entries either way.
	// All rows pass → nothing is touched.  Otherwise sel[0..found) = the surviving row ids (translated back
	// through the incoming selection, :117-123), sel_is_set = true, approved_count = found.  Returns found.
	uint64_t Filter(const rpt_col_t &keys, uint32_t *sel, bool &sel_is_set, uint64_t &approved_count) {
		if (approved_count == 0) {
			return 0;
		}
		if (out_.size() < approved_count) {
			out_.resize(approved_count); // state.bf_sel (:96-100)
		}
		uint64_t found = 0;
		Check(rpt_bf_filter_sel(filter_.handle(), &keys, sel_is_set ? sel : nullptr, approved_count, out_.data(), &found,
		                        RPT_MEM_HOST, nullptr));
		if (found == approved_count) {
			return approved_count; // :112-115
		}
		std::memcpy(sel, out_.data(), found * sizeof(uint32_t));
		sel_is_set = true;
		approved_count = found;
		return found;
	}
	// FilterValue (:133-136) on value.Hash()
	bool FilterValue(uint64_t hash) const {
		uint8_t hit = 0;
		Check(rpt_bf_lookup_hashes_host(filter_.handle(), &hash, 1, &hit));
		return hit != 0;
	}
	// CheckStatistics (:138-196) for one row group
	rpt_prune CheckStatistics(int physical_type, const void *min, const void *max, bool has_minmax = true) const {
		const uint8_t has = has_minmax ? 1 : 0;
		uint8_t res = RPT_NO_PRUNING_POSSIBLE;
		filter_.CheckStatistics(physical_type, min, max, &has, 1, &res);
		return static_cast<rpt_prune>(res);
	}

private:
	const PTBloomFilter &filter_;
	std::vector<uint32_t> out_;
};

// ------------------------------------------------------------------------------------------------------------
// Materialises the key column of successive chunks into one flat pinned column — what
// `local_state.local_data->Append(chunk)` does for the whole chunk in PhysicalCreateBF::Sink
// (physical_create_bf.cpp:214-216), restricted to the key column.  Dictionary / sliced vectors are resolved
// through `sel`; validity is kept as one bitmap; string_t (host pointers) becomes offsets + bytes.
// ------------------------------------------------------------------------------------------------------------
class ColumnStager {
public:
	explicit ColumnStager(int type) : type_(type) {
		if (TypeWidth(type) == 0) {
			throw Error("rpt_gpu: Invalid type for hash: physical type " + std::to_string(type));
		}
	}
	uint64_t Rows() const {
		return rows_;
	}
	void Clear() {
		rows_ = 0;
		str_bytes_ = 0;
		has_validity_ = false;
	}
	void Append(const rpt_col_t &col, uint64_t count) {
		if (count == 0) {
			return;
		}
		if (col.type != type_ || col.mem != RPT_MEM_HOST) {
			throw Error("rpt_gpu: ColumnStager: column type / memory space mismatch");
		}
		if (col.validity && !has_validity_) { // first chunk with NULLs: everything staged so far is valid
			validity_.Reserve(((rows_ + count + 63) / 64 + 1) * 8, 0);
			memset(validity_.data(), 0xff, ((rows_ + 63) / 64) * 8);
			has_validity_ = true;
		}
		if (has_validity_) {
			const size_t old_words = (rows_ + 63) / 64, new_words = (rows_ + count + 63) / 64;
			validity_.Reserve(new_words * 8, old_words * 8);
			uint64_t *v = reinterpret_cast<uint64_t *>(validity_.data());
			for (size_t w = old_words; w < new_words; w++) {
				v[w] = 0;
			}
			if (rows_ & 63) { // clear the unused tail bits of the last old word before OR-ing new rows in
				v[old_words - 1] &= (~0ULL) >> (64 - (rows_ & 63));
			}
			for (uint64_t i = 0; i < count; i++) {
				const uint64_t idx = col.sel ? col.sel[i] : i;
				const bool valid = !col.validity || ((col.validity[idx >> 6] >> (idx & 63)) & 1ULL);
				if (valid) {
					v[(rows_ + i) >> 6] |= 1ULL << ((rows_ + i) & 63);
				}
			}
		}
		if (type_ == RPT_STRING_T) {
			AppendStrings(col, count);
		} else {
			const size_t w = TypeWidth(type_);
			values_.Reserve((rows_ + count) * w, rows_ * w);
			uint8_t *dst = values_.data() + rows_ * w;
			const uint8_t *src = static_cast<const uint8_t *>(col.data);
			if (!col.sel) {
				memcpy(dst, src, count * w);
			} else {
				for (uint64_t i = 0; i < count; i++) {
					memcpy(dst + i * w, src + static_cast<size_t>(col.sel[i]) * w, w);
				}
			}
		}
		rows_ += count;
	}
	// The staged rows as one flat key column (valid until the next Append / Clear).
	rpt_col_t AsColumn() const {
		rpt_col_t c {};
		if (type_ == RPT_STRING_T) {
			c.data = offsets_.data();
			c.aux = bytes_.data();
			c.aux_bytes = str_bytes_;
			c.type = RPT_VARCHAR;
		} else {
			c.data = values_.data();
			c.type = type_;
		}
		c.validity = has_validity_ ? reinterpret_cast<const uint64_t *>(validity_.data()) : nullptr;
		c.mem = RPT_MEM_HOST;
		return c;
	}

private:
	// duckdb::string_t (string_type.hpp:24): uint32 length; ≤ 12 bytes inlined after it, else 4-byte prefix + pointer
	void AppendStrings(const rpt_col_t &col, uint64_t count) {
		offsets_.Reserve((rows_ + count + 1) * 4, (rows_ + 1) * 4);
		uint32_t *off = reinterpret_cast<uint32_t *>(offsets_.data());
		if (rows_ == 0) {
			off[0] = 0;
		}
		const uint8_t *base = static_cast<const uint8_t *>(col.data);
		for (uint64_t i = 0; i < count; i++) {
			const uint8_t *e = base + static_cast<size_t>(col.sel ? col.sel[i] : i) * 16;
			const uint64_t idx = col.sel ? col.sel[i] : i;
			const bool valid = !col.validity || ((col.validity[idx >> 6] >> (idx & 63)) & 1ULL);
			uint32_t len = 0;
			if (valid) { // the payload of a NULL row is undefined: never dereference it
				memcpy(&len, e, 4);
			}
			if (len) {
				const uint8_t *src;
				if (len <= 12) {
					src = e + 4;
				} else {
					memcpy(&src, e + 8, sizeof(src));
				}
				bytes_.Reserve(str_bytes_ + len + 8, str_bytes_);
				memcpy(bytes_.data() + str_bytes_, src, len);
				str_bytes_ += len;
			}
			off[rows_ + i + 1] = static_cast<uint32_t>(str_bytes_);
		}
		bytes_.Reserve(str_bytes_ + 8, str_bytes_);
	}

	int type_;
	uint64_t rows_ = 0;
	size_t str_bytes_ = 0;
	bool has_validity_ = false;
	PinnedBuffer values_, validity_, offsets_, bytes_;
};

// ------------------------------------------------------------------------------------------------------------
// CREATE_BF glue.  The reference inserts every chunk into the filter inside Sink (physical_create_bf.cpp:220-227),
// updates min/max per chunk (:229-239), and at Finalize re-inserts EVERYTHING single-threaded when the filter
// turned out undersized (:386-406).  Here a sink thread only stages the key column (one CreateBFBatch per
// thread = the reference's local sink state); Finalize applies the resize rule first, so the filter is sized
// once, and builds it in one pipelined pass over the staged columns.  The resulting words are bit-identical:
// insertion is an idempotent OR and the final S is the same.
// ------------------------------------------------------------------------------------------------------------
class CreateBFBatch {
public:
	explicit CreateBFBatch(int key_type) : stager_(key_type) {
	}
	// Sink: one chunk's key column
	void Append(const rpt_col_t &col, uint64_t count) {
		stager_.Append(col, count);
	}
	uint64_t Rows() const {
		return stager_.Rows();
	}

	struct Result {
		bool rebuilt = false;   // the :396 rule fired and the filter was re-sized for the actual row count
		bool has_minmax = false; // ColumnMinMax::has_value
		bool has_nan = false;
		uint8_t min_val[8] = {0}, max_val[8] = {0}; // one value of the key's physical type each
	};

	// Finalize over the local batches of all sink threads (gsink.local_data_collections → total_data, :372-376).
	// want_minmax: is_forward_pass && the key type is numeric (:229).
	static Result Finalize(PTBloomFilter &bf, const std::vector<CreateBFBatch *> &batches, bool want_minmax) {
		Result res;
		uint64_t actual_rows = 0;
		for (auto *b : batches) {
			actual_rows += b->Rows();
		}
		int key_type = batches.empty() ? 0 : batches[0]->stager_.AsColumn().type;
		// :386-406 — decided before anything is inserted, so no key is hashed twice
		if (actual_rows > 0 && bf.NeedsRebuild(actual_rows)) {
			bf.ReinitializeAndRehash(actual_rows, nullptr, 0, 0); // re-size + zero only
			res.rebuilt = true;
		}
		// UpdateMinMax's numeric cases (:136-170); VARCHAR ranges stay with the host, other types have none (:171)
		const bool numeric = TypeWidth(key_type) >= 1 && TypeWidth(key_type) <= 8 && key_type != RPT_VARCHAR;
		const bool mm = want_minmax && actual_rows > 0 && numeric;
		if (mm) {
			Check(rpt_bf_minmax_begin(bf.handle(), key_type));
		}
		for (auto *b : batches) {
			if (b->Rows()) {
				const rpt_col_t col = b->stager_.AsColumn();
				bf.Insert(&col, 1, b->Rows());
			}
		}
		if (mm) {
			int has = 0, nan = 0;
			Check(rpt_bf_minmax_get(bf.handle(), &has, &nan, res.min_val, res.max_val));
			Check(rpt_bf_minmax_end(bf.handle()));
			res.has_minmax = has != 0;
			res.has_nan = nan != 0;
		}
		bf.Finalize(actual_rows); // sets finalized_ (:409-413); the rule cannot fire again
		return res;
	}

private:
	ColumnStager stager_;
};

// ------------------------------------------------------------------------------------------------------------
// USE_BF glue (PhysicalUseBF::ExecuteInternal, physical_use_bf.cpp:125-195) for ONE filter: accumulates the
// key column of several chunks, probes them with one call, and hands back per-chunk (sel, count) pairs with
// chunk-relative row ids — what `input.Slice(sel, result_count)` (:172-175) consumes.  The reference's guards are
// kept: an unfinalised filter is skipped (every row passes, :139-142); an empty filter passes nothing (:145-155).
// ------------------------------------------------------------------------------------------------------------
class UseBFBatch {
public:
	UseBFBatch(const PTBloomFilter &bf, int key_type, uint64_t max_rows = uint64_t(1) << 19)
	    : bf_(bf), stager_(key_type), max_rows_(max_rows) {
	}
	// returns true when the batch should be probed before more chunks are added
	bool Append(const rpt_col_t &col, uint64_t count) {
		starts_.push_back(stager_.Rows());
		stager_.Append(col, count);
		return stager_.Rows() >= max_rows_;
	}
	size_t NumChunks() const {
		return starts_.size();
	}
	void Probe() {
		const uint64_t rows = stager_.Rows();
		const size_t nchunks = starts_.size();
		counts_.assign(nchunks, 0);
		first_.assign(nchunks + 1, 0);
		if (sel_.size() < rows) {
			sel_.resize(rows);
		}
		uint64_t found = 0;
		if (!bf_.finalized()) { // "skipped - bloom filter not ready": identity selection
			for (uint64_t i = 0; i < rows; i++) {
				sel_[i] = static_cast<uint32_t>(i);
			}
			found = rows;
		} else if (bf_.IsEmpty() || rows == 0) {
			found = 0; // empty filter means no matches possible
		} else {
			const rpt_col_t col = stager_.AsColumn();
			found = bf_.LookupSel(&col, 1, rows, sel_.data());
		}
		// split the ascending global selection at the chunk starts and make the ids chunk-relative
		size_t c = 0;
		for (uint64_t k = 0; k < found; k++) {
			while (c + 1 < nchunks && sel_[k] >= starts_[c + 1]) {
				first_[++c] = k;
			}
			sel_[k] -= static_cast<uint32_t>(starts_[c]);
			counts_[c]++;
		}
		while (c + 1 < nchunks) {
			first_[++c] = found;
		}
		first_[nchunks] = found;
	}
	// chunk i of the batch: its surviving row ids (relative to the chunk) and their number
	std::pair<const uint32_t *, uint64_t> Chunk(size_t i) const {
		return {sel_.data() + first_[i], counts_[i]};
	}
	void Clear() {
		stager_.Clear();
		starts_.clear();
	}

private:
	const PTBloomFilter &bf_;
	ColumnStager stager_;
	uint64_t max_rows_;
	std::vector<uint64_t> starts_, first_, counts_;
	std::vector<uint32_t> sel_;
};

// ---------------------------------------------------------------------------------------------------------
// README-described filter variant (reference README.md:23-32: Arrow Acero BlockedBloomFilter — 1024-entry mask
// table, 8 bits per key, Fold) over rpt_abf_*.  NOT the reference's code path: src/bloom_filter.cpp wraps DuckDB's
// native BloomFilter (PTBloomFilter above).  Method names follow arrow::acero::BlockedBloomFilter.  Parity
// unpinned (rpt_gpu.h).
// ---------------------------------------------------------------------------------------------------------
class BlockedBloomFilter {
public:
	explicit BlockedBloomFilter(int device = 0) : device_(device) {
	}
	~BlockedBloomFilter() {
		rpt_abf_destroy(handle_);
	}
	BlockedBloomFilter(const BlockedBloomFilter &) = delete;
	BlockedBloomFilter &operator=(const BlockedBloomFilter &) = delete;

	// CreateEmpty(num_rows_to_insert): 8 bits per key, at least 512 bits, power of two
	void CreateEmpty(uint64_t num_rows_to_insert) {
		rpt_abf_destroy(handle_);
		handle_ = nullptr;
		Check(rpt_abf_create(&handle_, num_rows_to_insert, device_));
	}
	// Insert for every row of the key column(s) (DuckDB hashes, NULL rows hash to NULL_HASH)
	void Insert(const rpt_col_t *cols, int ncols, uint64_t count, rpt_stream_t stream = nullptr) {
		Check(rpt_abf_insert(handle_, cols, ncols, count, stream));
	}
	// Fold(): OR the halves together while fewer than a quarter of the bits are set; returns the halvings done
	int Fold(rpt_stream_t stream = nullptr) {
		int folds = 0;
		Check(rpt_abf_fold(handle_, stream, &folds));
		return folds;
	}
	// Find for every row: ascending row ids of the rows that pass
	uint64_t Find(const rpt_col_t *cols, int ncols, uint64_t count, uint32_t *sel_out) const {
		uint64_t found = 0;
		Check(rpt_abf_probe(handle_, cols, ncols, count, sel_out, &found, RPT_MEM_HOST, nullptr));
		return found;
	}
	int log_num_blocks() const {
		int v = 0;
		Check(rpt_abf_log_num_blocks(handle_, &v));
		return v;
	}
	std::vector<uint64_t> Blocks() const {
		std::vector<uint64_t> out(1ULL << log_num_blocks());
		Check(rpt_abf_export_blocks(handle_, out.data(), out.size()));
		return out;
	}
	rpt_abf_t *handle() {
		return handle_;
	}

private:
	int device_;
	rpt_abf_t *handle_ = nullptr;
};

} // namespace rpt
