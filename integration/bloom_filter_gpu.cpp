//===----------------------------------------------------------------------===//
// integration/bloom_filter_gpu.cpp — GPU body of the reference's src/bloom_filter.cpp
//
// The ONLY reference translation unit that is replaced.  PTBloomFilter keeps its signature
// (reference src/include/bloom_filter.hpp:22-57); each method does what the reference's does
// (src/bloom_filter.cpp:27-78), with the hash + insert / lookup running in librpt_gpu.so:
//
//   Initialize             :27-32   rpt_bf_create / rpt_bf_reset (same sizing arithmetic, same uint32 wrap)
//   Insert                 :70-78   the chunk's key column is STAGED (pinned host memory, per filter); the staged
//                                   rows go to the device in one rpt_bf_insert when 4 Mi rows are pending, and at the
//                                   latest before the filter is first read.  Insertion is an idempotent OR, so the
//                                   words equal the reference's chunk-by-chunk inserts bit for bit.
//   ReinitializeAndRehash  :34-58   rpt_bf_reinitialize_and_rehash (re-size + zero), then the collection is streamed
//                                   through the same staging
//   LookupSel              :60-68   rpt_bf_probe on the chunk's key column: ascending row ids, as LookupHashes writes
//   GetNativeFilter                 the device words exported into DuckDB's own BloomFilter::bf (byte-identical
//                                   layout), so PushDynamicFilters / BFTableFilter::Filter / CheckStatistics
//                                   (physical_create_bf.cpp:321-332, duckdb bloom_filter.cpp:94-196) work unchanged
//
// There is no CPU fallback: without a CUDA device rpt_bf_create fails and Initialize throws.
//===----------------------------------------------------------------------===//
#include "bloom_filter.hpp"

#include "duckdb/common/exception.hpp"
#include "duckdb/common/types/column/column_data_collection.hpp"
#include "duckdb/common/types/selection_vector.hpp"
#include "duckdb/common/vector_operations/vector_operations.hpp"

#include "rpt_gpu.h"
#include "rpt_gpu.hpp"

#include <cstdlib>
#include <mutex>

namespace duckdb {

namespace {

void Check(int rc) {
	if (rc != 0) {
		throw InternalException("rpt_gpu: %s", rpt_gpu_last_error());
	}
}

// DuckDB keeps BloomFilter::bf / num_sectors private and offers no accessor.  Explicit template instantiation may
// name private members ([temp.spec]/6), which gives the mirror a standards-conforming way to fill the words without
// touching the DuckDB tree.  (A maintainer who owns the tree adds `uint64_t *Words()` to BloomFilter instead.)
template <class Tag, typename Tag::type Member>
struct PrivateMember {
	friend typename Tag::type Reveal(Tag) {
		return Member;
	}
};
struct BfWordsTag {
	using type = uint64_t *BloomFilter::*;
	friend type Reveal(BfWordsTag);
};
struct BfSectorsTag {
	using type = idx_t BloomFilter::*;
	friend type Reveal(BfSectorsTag);
};
template struct PrivateMember<BfWordsTag, &BloomFilter::bf>;
template struct PrivateMember<BfSectorsTag, &BloomFilter::num_sectors>;

// UnifiedVectorFormat → rpt_col_t (include/rpt_gpu.h): data / selection / validity as they lie in the vector.
// VARCHAR: string_t holds host pointers, the library (or the stager) gathers them (RPT_STRING_T).
// data_len = entries addressable through the selection (max(sel)+1): a constant vector has 1, a dictionary its size.
rpt_col_t ToCol(Vector &vec, idx_t count, UnifiedVectorFormat &fmt) {
	vec.ToUnifiedFormat(count, fmt);
	rpt_col_t c {};
	c.data = fmt.data;
	c.sel = fmt.sel->IsSet() ? reinterpret_cast<const uint32_t *>(fmt.sel->data()) : nullptr;
	c.validity = fmt.validity.AllValid() ? nullptr : reinterpret_cast<const uint64_t *>(fmt.validity.GetData());
	c.data_len = 0;
	if (c.sel) {
		uint64_t top = 0;
		for (idx_t i = 0; i < count; i++) {
			top = c.sel[i] > top ? c.sel[i] : top;
		}
		c.data_len = top + 1;
	}
	const auto pt = vec.GetType().InternalType();
	c.type = pt == PhysicalType::VARCHAR ? static_cast<int32_t>(RPT_STRING_T) : static_cast<int32_t>(pt);
	c.mem = RPT_MEM_HOST;
	return c;
}

int DeviceOrdinal() {
	const char *e = getenv("RPT_GPU_DEVICE");
	return e ? atoi(e) : 0;
}

constexpr idx_t kFlushRows = idx_t(1) << 22; // pending rows that trigger an upload from inside Insert

} // namespace

struct PTBloomFilter::Impl {
	rpt_bf_t *gpu = nullptr;
	std::mutex mu;                             // sink threads share one filter (physical_create_bf.cpp:220-227)
	std::unique_ptr<rpt::ColumnStager> pending; // key column rows not yet on the device
	bool mirror_current = false;               // bf_ holds the device words

	~Impl() {
		if (gpu) {
			rpt_bf_destroy(gpu);
		}
	}
	// caller holds mu
	void Flush() {
		if (pending && pending->Rows() > 0) {
			const rpt_col_t col = pending->AsColumn();
			Check(rpt_bf_insert(gpu, &col, 1, pending->Rows(), nullptr));
			pending->Clear();
			mirror_current = false;
		}
	}
	// caller holds mu
	void Stage(const rpt_col_t &col, idx_t count) {
		if (!pending) {
			pending.reset(new rpt::ColumnStager(col.type));
		}
		pending->Append(col, count);
		if (pending->Rows() >= kFlushRows) {
			Flush();
		}
	}
};

PTBloomFilter::PTBloomFilter() : impl_(new Impl()) {
}

PTBloomFilter::~PTBloomFilter() {
}

void PTBloomFilter::Initialize(ClientContext &context_p, uint32_t est_num_rows) {
	context = &context_p;
	buffer_manager = &BufferManager::GetBufferManager(*context);
	std::lock_guard<std::mutex> lock(impl_->mu);
	if (impl_->gpu) {
		Check(rpt_bf_reset(impl_->gpu, est_num_rows, nullptr));
	} else {
		Check(rpt_bf_create(&impl_->gpu, est_num_rows, DeviceOrdinal()));
	}
	if (impl_->pending) {
		impl_->pending->Clear();
	}
	impl_->mirror_current = false;
	sized_for_rows_ = static_cast<idx_t>(est_num_rows);
	has_data_ = false;
	finalized_ = false;
}

void PTBloomFilter::Insert(DataChunk &chunk, const vector<idx_t> &bound_cols_built) {
	const idx_t count = chunk.size();
	if (count == 0) {
		return;
	}
	has_data_ = true;
	std::lock_guard<std::mutex> lock(impl_->mu);
	if (bound_cols_built.size() == 1) { // what every call site passes
		UnifiedVectorFormat fmt;
		const rpt_col_t col = ToCol(chunk.data[bound_cols_built[0]], count, fmt);
		impl_->Stage(col, count);
		return;
	}
	// HashColumns with several columns (src/bloom_filter.cpp:11-24): hashed row-wise on the device, unstaged
	impl_->Flush();
	if (bound_cols_built.size() > RPT_MAX_KEY_COLS) {
		throw InternalException("rpt_gpu: more than %d key columns", RPT_MAX_KEY_COLS);
	}
	rpt_col_t cols[RPT_MAX_KEY_COLS];
	UnifiedVectorFormat fmt[RPT_MAX_KEY_COLS];
	for (idx_t j = 0; j < bound_cols_built.size(); j++) {
		cols[j] = ToCol(chunk.data[bound_cols_built[j]], count, fmt[j]);
	}
	Check(rpt_bf_insert(impl_->gpu, cols, static_cast<int>(bound_cols_built.size()), count, nullptr));
	impl_->mirror_current = false;
}

void PTBloomFilter::ReinitializeAndRehash(ClientContext &context_p, idx_t actual_rows, ColumnDataCollection &data,
                                          const vector<idx_t> &cols) {
	std::unique_lock<std::mutex> lock(impl_->mu);
	if (impl_->pending) {
		impl_->pending->Clear(); // every row is re-inserted from `data` below
	}
	// re-size for the accurate count and zero (src/bloom_filter.cpp:38-40)
	Check(rpt_bf_reinitialize_and_rehash(impl_->gpu, actual_rows, nullptr, 0, 0, nullptr));
	impl_->mirror_current = false;
	sized_for_rows_ = actual_rows;
	has_data_ = data.Count() > 0;
	lock.unlock();
	if (data.Count() == 0) {
		return;
	}
	DataChunk chunk;
	data.InitializeScanChunk(chunk);
	ColumnDataScanState scan_state;
	data.InitializeScan(scan_state);
	while (data.Scan(scan_state, chunk)) { // :46-57
		if (chunk.size() == 0) {
			continue;
		}
		Insert(chunk, cols);
	}
}

idx_t PTBloomFilter::LookupSel(DataChunk &chunk, SelectionVector &sel, const vector<idx_t> &bound_cols_applied,
                               uint8_t *) const {
	const idx_t count = chunk.size();
	if (count == 0) {
		return 0;
	}
	{
		std::lock_guard<std::mutex> lock(impl_->mu);
		impl_->Flush(); // rows staged by the sink threads reach the device before the first probe
	}
	if (bound_cols_applied.size() > RPT_MAX_KEY_COLS) {
		throw InternalException("rpt_gpu: more than %d key columns", RPT_MAX_KEY_COLS);
	}
	rpt_col_t cols[RPT_MAX_KEY_COLS];
	UnifiedVectorFormat fmt[RPT_MAX_KEY_COLS];
	for (idx_t j = 0; j < bound_cols_applied.size(); j++) {
		cols[j] = ToCol(chunk.data[bound_cols_applied[j]], count, fmt[j]);
	}
	uint64_t found = 0;
	Check(rpt_bf_probe(impl_->gpu, cols, static_cast<int>(bound_cols_applied.size()), count,
	                   reinterpret_cast<uint32_t *>(sel.data()), &found, RPT_MEM_HOST, nullptr));
	return static_cast<idx_t>(found);
}

BloomFilter &PTBloomFilter::GetNativeFilter() {
	std::lock_guard<std::mutex> lock(impl_->mu);
	impl_->Flush();
	if (!impl_->mirror_current) {
		// allocate DuckDB's own buffer for the same row count (same S: same arithmetic), then overwrite its words
		bf_.Initialize(*context, sized_for_rows_);
		uint64_t sectors = 0;
		Check(rpt_bf_num_sectors(impl_->gpu, &sectors));
		if (bf_.*Reveal(BfSectorsTag()) != sectors) {
			throw InternalException("rpt_gpu: host mirror has %llu sectors, the device filter %llu",
			                        static_cast<unsigned long long>(bf_.*Reveal(BfSectorsTag())),
			                        static_cast<unsigned long long>(sectors));
		}
		Check(rpt_bf_export_words(impl_->gpu, bf_.*Reveal(BfWordsTag()), sectors));
		impl_->mirror_current = true;
	}
	return bf_;
}

} // namespace duckdb
