//===----------------------------------------------------------------------===//
// integration/bloom_filter.hpp — drop-in replacement of the reference's src/include/bloom_filter.hpp
//
// Same class, same public surface (reference src/include/bloom_filter.hpp:22-57): the operators
// (src/operators/physical_create_bf.cpp, physical_use_bf.cpp) compile against this header UNCHANGED.
// What changes is behind it: the filter lives on a B200 behind the C ABI of librpt_gpu.so
// (include/rpt_gpu.h), and DuckDB's native BloomFilter member is kept only as the HOST MIRROR that
// GetNativeFilter() hands to the scan-side BFTableFilter (physical_create_bf.cpp:321-332): the GPU-built
// words are byte-identical to DuckDB's bf[] and are exported into it on first use.
//
// Put this directory ahead of the reference's src/include on the include path and compile
// integration/bloom_filter_gpu.cpp instead of src/bloom_filter.cpp (integration/CMakeLists.txt does both).
//===----------------------------------------------------------------------===//
#pragma once

#include "duckdb/planner/filter/bloom_filter.hpp"
#include "duckdb/planner/column_binding.hpp"
#include "duckdb/storage/buffer_manager.hpp"
#include "duckdb/common/types/data_chunk.hpp"

#include <cstdint>
#include <memory>

namespace duckdb {

class ColumnDataCollection;

// wrapper with DataChunk-level operations; the filter itself is device-resident (librpt_gpu.so)
class PTBloomFilter {
public:
	PTBloomFilter();
	~PTBloomFilter();
	PTBloomFilter(const PTBloomFilter &) = delete;
	PTBloomFilter &operator=(const PTBloomFilter &) = delete;

	void Initialize(ClientContext &context_p, uint32_t est_num_rows);

	ClientContext *context = nullptr;
	BufferManager *buffer_manager = nullptr;

	bool finalized_ = false;

public:
	idx_t LookupSel(DataChunk &chunk, SelectionVector &sel, const vector<idx_t> &bound_cols_applied,
	                uint8_t *bit_vector_buf) const;
	void Insert(DataChunk &chunk, const vector<idx_t> &bound_cols_built);

	// re-size the filter for `actual_rows` and re-hash all rows from `data` on `cols`
	void ReinitializeAndRehash(ClientContext &context_p, idx_t actual_rows, ColumnDataCollection &data,
	                           const vector<idx_t> &cols);

	idx_t SizedForRows() const {
		return sized_for_rows_;
	}

	bool IsEmpty() const {
		return !has_data_;
	}

	// the host mirror for scan-side probes: brought up to date with the device filter on demand
	BloomFilter &GetNativeFilter();

private:
	struct Impl; // device handle, staging of the sink threads' key columns, mirror state
	std::unique_ptr<Impl> impl_;
	bool has_data_ = false;
	idx_t sized_for_rows_ = 0;
	BloomFilter bf_;
};

} // namespace duckdb
