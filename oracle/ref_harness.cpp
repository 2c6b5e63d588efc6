/*
 * ref_harness.cpp — TEST INFRASTRUCTURE, NOT PRODUCT.
 *
 * Builds oracle/_ref/librpt_ref.so: the REFERENCE's own hot-path code, compiled from the sources
 * where they lie under /root/reference (nothing is copied into this repository).  The four
 * translation units that hold the arithmetic are textually included below and linked with
 * --gc-sections so that the parts of them that need the rest of DuckDB (BFTableFilter,
 * VectorOperations::Hash's Vector plumbing, serialisation) are discarded; what remains and is
 * exported through the C functions at the bottom is reference code, not a restatement:
 *
 *   duckdb/src/planner/filter/bloom_filter.cpp   GetMask, BloomFilter::InsertOne/LookupOne,
 *                                                InsertHashes, LookupHashes          (:29-78)
 *   duckdb/src/common/types/hash.cpp             Hash<T> specialisations, HashBytes, Hash(string_t)
 *   duckdb/src/common/vector_operations/vector_hash.cpp
 *                                                HashOp::NULL_HASH, CombineHashScalar, TightLoopHash,
 *                                                TightLoopCombineHash                (:16-61,:364-385)
 *   duckdb/src/common/constants.cpp              NextPowerOfTwo                      (:30-47)
 *   duckdb/src/planner/filter/bloom_filter.cpp   TemplatedCheckStatistics<T>          (:138-169), with
 *   duckdb/src/function/scalar/operator/subtract.cpp   TrySubtractOperator::Operation<T> (:98-175) and
 *   duckdb/src/common/types/{hugeint,uhugeint}.cpp     the 128-bit integer arithmetic
 *
 * What cannot be compiled this way (needs the whole engine: BufferManager, DataChunk,
 * ColumnDataCollection, PhysicalOperator): BloomFilter::Initialize's allocation,
 * PTBloomFilter (src/bloom_filter.cpp) and the operators.  Their arithmetic — three lines of
 * sizing and the resize rule — is restated in rpt_oracle.c and pinned by SURVEY §8(c)'s recorded
 * sector counts and filter checksums.
 *
 * The fake `Vector` / `BloomFilter` objects below only carry the fields the reference loops read
 * (Vector::data; BloomFilter::{bf,num_sectors,bitmask}); -fno-access-control lets the harness set them.
 */
/* Zone-map pruning (BFTableFilter::CheckStatistics → TemplatedCheckStatistics<T>, bloom_filter.cpp:138-196): the
 * function template is compiled UNMODIFIED from the reference file included below — the TrySubtractOperator
 * overflow rule, the `range_typed > 2048` cut, the `val += i < range` walk and Hash(val) per type are the
 * reference's.  Only where the range comes from is faked: NumericStats::HasMinMax / GetMin<T> / GetMax<T> normally
 * read a BaseStatistics through duckdb::Value (the whole engine); here they return the harness' current range.
 * The explicit specialisations have to be visible before the template that uses them is instantiated. */
#include "duckdb/storage/statistics/numeric_stats.hpp"
#include "duckdb/common/types/hugeint.hpp"
#include "duckdb/common/types/uhugeint.hpp"
namespace rpt_ref_fake {
struct Range {
	bool has_minmax;
	unsigned char min[16], max[16];
};
static thread_local Range g_range;
} // namespace rpt_ref_fake
namespace duckdb {
#define RPT_REF_MINMAX(T)                                                                                              \
	template <>                                                                                                        \
	T NumericStats::GetMin<T>(const BaseStatistics &) {                                                                \
		T v;                                                                                                           \
		memcpy(&v, rpt_ref_fake::g_range.min, sizeof(T));                                                              \
		return v;                                                                                                      \
	}                                                                                                                  \
	template <>                                                                                                        \
	T NumericStats::GetMax<T>(const BaseStatistics &) {                                                                \
		T v;                                                                                                           \
		memcpy(&v, rpt_ref_fake::g_range.max, sizeof(T));                                                              \
		return v;                                                                                                      \
	}
RPT_REF_MINMAX(uint8_t)
RPT_REF_MINMAX(uint16_t)
RPT_REF_MINMAX(uint32_t)
RPT_REF_MINMAX(uint64_t)
RPT_REF_MINMAX(uhugeint_t)
RPT_REF_MINMAX(int8_t)
RPT_REF_MINMAX(int16_t)
RPT_REF_MINMAX(int32_t)
RPT_REF_MINMAX(int64_t)
RPT_REF_MINMAX(hugeint_t)
#undef RPT_REF_MINMAX
bool NumericStats::HasMinMax(const BaseStatistics &) {
	return rpt_ref_fake::g_range.has_minmax;
}
} // namespace duckdb

#include "/root/reference/duckdb/src/planner/filter/bloom_filter.cpp"
/* TrySubtractOperator::Operation<T> (subtract.cpp:98-175) and the 128-bit integer arithmetic it and the walk use */
#include "/root/reference/duckdb/src/function/scalar/operator/subtract.cpp"
#include "/root/reference/duckdb/src/common/types/hugeint.cpp"
#include "/root/reference/duckdb/src/common/types/uhugeint.cpp"
#include "/root/reference/duckdb/src/common/types/hash.cpp"
#include "/root/reference/duckdb/src/common/vector_operations/vector_hash.cpp"
#include "/root/reference/duckdb/src/common/constants.cpp"

#include <chrono>
#include <cstring>
#include <thread>
#include <vector>

/* Link stubs for the reference's ERROR paths only (NextPowerOfTwo overflow, a type-mismatch check):
 * their real bodies live in duckdb/src/common/exception.cpp, which needs the whole engine.  None of
 * these runs on a valid input; each one just raises. */
namespace duckdb {
Exception::Exception(ExceptionType, const string &message) : std::runtime_error(message) {
}
OutOfRangeException::OutOfRangeException(const string &msg) : Exception(ExceptionType::OUT_OF_RANGE, msg) {
}
InternalException::InternalException(const string &msg) : Exception(ExceptionType::INTERNAL, msg) {
}
string Exception::ConstructMessageRecursive(const string &msg, std::vector<ExceptionFormatValue> &) {
	return msg;
}
template <>
ExceptionFormatValue ExceptionFormatValue::CreateFormatValue(const PhysicalType &) {
	throw std::runtime_error("rpt ref harness: reference error path (PhysicalType)");
}
ExceptionFormatValue::ExceptionFormatValue(int64_t) { // NumericCast's "value out of range" message
	throw std::runtime_error("rpt ref harness: reference error path (numeric cast)");
}
template <>
ExceptionFormatValue ExceptionFormatValue::CreateFormatValue(const hugeint_t &) {
	throw std::runtime_error("rpt ref harness: reference error path (hugeint)");
}
template <>
ExceptionFormatValue ExceptionFormatValue::CreateFormatValue(const uhugeint_t &) {
	throw std::runtime_error("rpt ref harness: reference error path (uhugeint)");
}
template <>
ExceptionFormatValue ExceptionFormatValue::CreateFormatValue(const idx_t &) {
	throw std::runtime_error("rpt ref harness: reference error path (out of range)");
}
} // namespace duckdb

using namespace duckdb;

#define REF_API extern "C" __attribute__((visibility("default")))

namespace {

struct FakeVector {
	alignas(Vector) unsigned char raw[sizeof(Vector)];
	explicit FakeVector(void *data) {
		memset(raw, 0, sizeof(raw));
		get().vector_type = VectorType::FLAT_VECTOR;
		get().type.id_ = LogicalTypeId::UBIGINT; // LogicalType::HASH
		get().type.physical_type_ = PhysicalType::UINT64;
		get().data = reinterpret_cast<data_ptr_t>(data);
	}
	Vector &get() {
		return *reinterpret_cast<Vector *>(raw);
	}
};

struct FakeFilter {
	alignas(BloomFilter) unsigned char raw[sizeof(BloomFilter)];
	FakeFilter(uint64_t *words, uint64_t num_sectors) {
		memset(raw, 0, sizeof(raw));
		get().num_sectors = num_sectors;
		get().bitmask = num_sectors - 1;
		get().initialized = true;
		get().bf = words;
	}
	BloomFilter &get() {
		return *reinterpret_cast<BloomFilter *>(raw);
	}
};

template <class T>
void VectorHashT(const void *data, const uint32_t *sel, const uint64_t *validity, uint64_t data_len, uint64_t count,
                 uint64_t *out, int combine) {
	ValidityMask mask(const_cast<validity_t *>(reinterpret_cast<const validity_t *>(validity)), data_len);
	SelectionVector selv(const_cast<sel_t *>(sel));
	const T *ldata = reinterpret_cast<const T *>(data);
	if (!combine) {
		if (sel) {
			TightLoopHash<false, true, T, false>(ldata, out, nullptr, count, &selv, mask);
		} else {
			TightLoopHash<false, false, T, false>(ldata, out, nullptr, count, &selv, mask);
		}
	} else {
		if (sel) {
			TightLoopCombineHash<false, true, T, false>(ldata, out, nullptr, count, &selv, mask);
		} else {
			TightLoopCombineHash<false, false, T, false>(ldata, out, nullptr, count, &selv, mask);
		}
	}
}

} // namespace

REF_API uint64_t ref_hash_value(int type, const void *v) {
	switch (static_cast<PhysicalType>(type)) {
	case PhysicalType::BOOL:
	case PhysicalType::INT8:
		return Hash<int8_t>(*reinterpret_cast<const int8_t *>(v));
	case PhysicalType::UINT8:
		return Hash<uint8_t>(*reinterpret_cast<const uint8_t *>(v));
	case PhysicalType::INT16:
		return Hash<int16_t>(*reinterpret_cast<const int16_t *>(v));
	case PhysicalType::UINT16:
		return Hash<uint16_t>(*reinterpret_cast<const uint16_t *>(v));
	case PhysicalType::INT32:
		return Hash<int32_t>(*reinterpret_cast<const int32_t *>(v));
	case PhysicalType::UINT32:
		return Hash<uint32_t>(*reinterpret_cast<const uint32_t *>(v));
	case PhysicalType::INT64:
		return Hash<int64_t>(*reinterpret_cast<const int64_t *>(v));
	case PhysicalType::UINT64:
		return Hash<uint64_t>(*reinterpret_cast<const uint64_t *>(v));
	case PhysicalType::INT128:
		return Hash<hugeint_t>(*reinterpret_cast<const hugeint_t *>(v));
	case PhysicalType::UINT128:
		return Hash<uhugeint_t>(*reinterpret_cast<const uhugeint_t *>(v));
	case PhysicalType::FLOAT:
		return Hash<float>(*reinterpret_cast<const float *>(v));
	case PhysicalType::DOUBLE:
		return Hash<double>(*reinterpret_cast<const double *>(v));
	case PhysicalType::INTERVAL:
		return Hash<interval_t>(*reinterpret_cast<const interval_t *>(v));
	default:
		return 0;
	}
}

REF_API uint64_t ref_hash_string(const char *p, uint32_t len) {
	return Hash<string_t>(string_t(p, len));
}

REF_API uint64_t ref_hash_bytes(const char *p, uint64_t len) {
	return Hash(p, static_cast<size_t>(len));
}

REF_API uint64_t ref_null_hash(void) {
	return HashOp::Operation<int64_t>(0, true);
}

REF_API uint64_t ref_combine(uint64_t a, uint64_t b) {
	return CombineHashScalar(a, b);
}

REF_API uint64_t ref_next_pow2(uint64_t v) {
	return NextPowerOfTwo(v);
}

REF_API uint64_t ref_mask(uint64_t h) {
	return GetMask(h);
}

/* TightLoopHash / TightLoopCombineHash over one column; VARCHAR comes as offsets+bytes and is
 * turned into the string_t array the reference loop reads. */
REF_API int ref_vector_hash(int type, const void *data, const void *aux, const uint32_t *sel,
                            const uint64_t *validity, uint64_t data_len, uint64_t count, uint64_t *out, int combine) {
	switch (static_cast<PhysicalType>(type)) {
	case PhysicalType::BOOL:
	case PhysicalType::INT8:
		VectorHashT<int8_t>(data, sel, validity, data_len, count, out, combine);
		return 0;
	case PhysicalType::UINT8:
		VectorHashT<uint8_t>(data, sel, validity, data_len, count, out, combine);
		return 0;
	case PhysicalType::INT16:
		VectorHashT<int16_t>(data, sel, validity, data_len, count, out, combine);
		return 0;
	case PhysicalType::UINT16:
		VectorHashT<uint16_t>(data, sel, validity, data_len, count, out, combine);
		return 0;
	case PhysicalType::INT32:
		VectorHashT<int32_t>(data, sel, validity, data_len, count, out, combine);
		return 0;
	case PhysicalType::UINT32:
		VectorHashT<uint32_t>(data, sel, validity, data_len, count, out, combine);
		return 0;
	case PhysicalType::INT64:
		VectorHashT<int64_t>(data, sel, validity, data_len, count, out, combine);
		return 0;
	case PhysicalType::UINT64:
		VectorHashT<uint64_t>(data, sel, validity, data_len, count, out, combine);
		return 0;
	case PhysicalType::INT128:
		VectorHashT<hugeint_t>(data, sel, validity, data_len, count, out, combine);
		return 0;
	case PhysicalType::UINT128:
		VectorHashT<uhugeint_t>(data, sel, validity, data_len, count, out, combine);
		return 0;
	case PhysicalType::FLOAT:
		VectorHashT<float>(data, sel, validity, data_len, count, out, combine);
		return 0;
	case PhysicalType::DOUBLE:
		VectorHashT<double>(data, sel, validity, data_len, count, out, combine);
		return 0;
	case PhysicalType::INTERVAL:
		VectorHashT<interval_t>(data, sel, validity, data_len, count, out, combine);
		return 0;
	case PhysicalType::VARCHAR: {
		const uint32_t *off = reinterpret_cast<const uint32_t *>(data);
		const char *bytes = reinterpret_cast<const char *>(aux);
		std::vector<string_t> strs(data_len);
		for (uint64_t i = 0; i < data_len; i++) {
			strs[i] = string_t(bytes + off[i], off[i + 1] - off[i]);
		}
		VectorHashT<string_t>(strs.data(), sel, validity, data_len, count, out, combine);
		return 0;
	}
	default:
		return 1;
	}
}

REF_API void ref_insert_hashes(uint64_t *words, uint64_t num_sectors, const uint64_t *hashes, uint64_t n) {
	FakeFilter f(words, num_sectors);
	FakeVector v(const_cast<uint64_t *>(hashes));
	f.get().InsertHashes(v.get(), n);
}

REF_API uint64_t ref_lookup_hashes(const uint64_t *words, uint64_t num_sectors, const uint64_t *hashes, uint64_t n,
                                   uint32_t *sel) {
	FakeFilter f(const_cast<uint64_t *>(words), num_sectors);
	FakeVector v(const_cast<uint64_t *>(hashes));
	SelectionVector selv(sel);
	return f.get().LookupHashes(v.get(), selv, n);
}

/* BFTableFilter::CheckStatistics' type switch (bloom_filter.cpp:171-196) over the reference's own template.
 * min / max: one value of the physical type (16 bytes {lower, upper} for the 128-bit types).
 * Returns FilterPropagateResult: 0 NO_PRUNING_POSSIBLE, 1 FILTER_ALWAYS_TRUE, 2 FILTER_ALWAYS_FALSE. */
REF_API int ref_check_statistics(const uint64_t *words, uint64_t num_sectors, int type, const void *min,
                                 const void *max, int has_minmax) {
	FakeFilter f(const_cast<uint64_t *>(words), num_sectors);
	alignas(16) static unsigned char no_stats[256]; // never read: every accessor is faked above
	const BaseStatistics &stats = *reinterpret_cast<const BaseStatistics *>(no_stats);
	rpt_ref_fake::g_range.has_minmax = has_minmax != 0;
	size_t width = 0;
	switch (static_cast<PhysicalType>(type)) {
	case PhysicalType::UINT8: case PhysicalType::INT8: width = 1; break;
	case PhysicalType::UINT16: case PhysicalType::INT16: width = 2; break;
	case PhysicalType::UINT32: case PhysicalType::INT32: width = 4; break;
	case PhysicalType::UINT64: case PhysicalType::INT64: width = 8; break;
	case PhysicalType::UINT128: case PhysicalType::INT128: width = 16; break;
	default: break;
	}
	memcpy(rpt_ref_fake::g_range.min, min, width);
	memcpy(rpt_ref_fake::g_range.max, max, width);
	const BloomFilter &bf = f.get();
	switch (static_cast<PhysicalType>(type)) { // bloom_filter.cpp:171-196
	case PhysicalType::UINT8:
		return static_cast<int>(TemplatedCheckStatistics<uint8_t>(bf, stats));
	case PhysicalType::UINT16:
		return static_cast<int>(TemplatedCheckStatistics<uint16_t>(bf, stats));
	case PhysicalType::UINT32:
		return static_cast<int>(TemplatedCheckStatistics<uint32_t>(bf, stats));
	case PhysicalType::UINT64:
		return static_cast<int>(TemplatedCheckStatistics<uint64_t>(bf, stats));
	case PhysicalType::UINT128:
		return static_cast<int>(TemplatedCheckStatistics<uhugeint_t>(bf, stats));
	case PhysicalType::INT8:
		return static_cast<int>(TemplatedCheckStatistics<int8_t>(bf, stats));
	case PhysicalType::INT16:
		return static_cast<int>(TemplatedCheckStatistics<int16_t>(bf, stats));
	case PhysicalType::INT32:
		return static_cast<int>(TemplatedCheckStatistics<int32_t>(bf, stats));
	case PhysicalType::INT64:
		return static_cast<int>(TemplatedCheckStatistics<int64_t>(bf, stats));
	case PhysicalType::INT128:
		return static_cast<int>(TemplatedCheckStatistics<hugeint_t>(bf, stats));
	default:
		return static_cast<int>(FilterPropagateResult::NO_PRUNING_POSSIBLE);
	}
}

REF_API int ref_lookup_one(const uint64_t *words, uint64_t num_sectors, uint64_t hash) {
	FakeFilter f(const_cast<uint64_t *>(words), num_sectors);
	return f.get().LookupOne(hash) ? 1 : 0;
}

/* ---- threaded CPU baseline on the reference loops ------------------------------------------------
 * T threads over disjoint 2048-row chunk ranges; per chunk: TightLoopHash<int64_t> into a 2048-entry
 * hash buffer, then InsertHashes into the ONE shared filter (PhysicalCreateBF::Sink pattern,
 * src/operators/physical_create_bf.cpp:220-227 → src/bloom_filter.cpp:70-78) or LookupHashes with a
 * private selection buffer (src/operators/physical_use_bf.cpp:163 → src/bloom_filter.cpp:60-68). */
namespace {
constexpr uint64_t REF_CHUNK = STANDARD_VECTOR_SIZE;

void BenchWorker(uint64_t *words, uint64_t num_sectors, const int64_t *keys, uint64_t lo, uint64_t hi, bool probe,
                 uint64_t *found_out) {
	FakeFilter f(words, num_sectors);
	std::vector<uint64_t> hashes(REF_CHUNK);
	std::vector<sel_t> sel(REF_CHUNK + 1);
	FakeVector hv(hashes.data());
	SelectionVector selv(sel.data());
	SelectionVector no_sel(static_cast<sel_t *>(nullptr));
	ValidityMask all_valid;
	uint64_t found = 0;
	for (uint64_t base = lo; base < hi; base += REF_CHUNK) {
		const uint64_t n = std::min<uint64_t>(REF_CHUNK, hi - base);
		TightLoopHash<false, false, int64_t, false>(keys + base, hashes.data(), nullptr, n, &no_sel, all_valid);
		if (probe) {
			found += f.get().LookupHashes(hv.get(), selv, n);
		} else {
			f.get().InsertHashes(hv.get(), n);
		}
	}
	*found_out = found;
}

double BenchRun(uint64_t *words, uint64_t num_sectors, const int64_t *keys, uint64_t n, int threads, bool probe,
                uint64_t *found_out) {
	threads = std::max(threads, 1);
	const uint64_t chunks = (n + REF_CHUNK - 1) / REF_CHUNK;
	std::vector<std::thread> pool;
	std::vector<uint64_t> found(static_cast<size_t>(threads), 0);
	auto t0 = std::chrono::steady_clock::now();
	for (int t = 0; t < threads; t++) {
		const uint64_t lo = chunks * t / threads * REF_CHUNK;
		const uint64_t hi = std::min<uint64_t>(chunks * (t + 1) / threads * REF_CHUNK, n);
		pool.emplace_back(BenchWorker, words, num_sectors, keys, lo, hi, probe, &found[t]);
	}
	uint64_t total = 0;
	for (int t = 0; t < threads; t++) {
		pool[t].join();
		total += found[t];
	}
	auto t1 = std::chrono::steady_clock::now();
	if (found_out) {
		*found_out = total;
	}
	return std::chrono::duration<double>(t1 - t0).count();
}
} // namespace

REF_API double ref_bench_build_i64(uint64_t *words, uint64_t num_sectors, const int64_t *keys, uint64_t n,
                                   int threads) {
	return BenchRun(words, num_sectors, keys, n, threads, false, nullptr);
}

REF_API double ref_bench_probe_i64(const uint64_t *words, uint64_t num_sectors, const int64_t *keys, uint64_t n,
                                   int threads, uint64_t *found_out) {
	return BenchRun(const_cast<uint64_t *>(words), num_sectors, keys, n, threads, true, found_out);
}
