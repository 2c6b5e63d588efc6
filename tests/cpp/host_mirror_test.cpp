// host_mirror_test.cpp — drives include/rpt_gpu.hpp the way the reference's operators drive PTBloomFilter
// (PhysicalCreateBF::Sink/Finalize, PhysicalUseBF::ExecuteInternal) and checks every result bit for bit
// against the CPU oracle (oracle/rpt_oracle.h — the checker, linked into this TEST only).
//
//   g++ -std=c++17 -O2 -Iinclude -Ioracle tests/cpp/host_mirror_test.cpp -o … -L… -lrpt_gpu -lrpt_oracle
// Needs a CUDA device (the library has no CPU fallback).  Prints "host mirror ok" and exits 0 on success.
#include <csignal>
#include <cstdio>
#include <execinfo.h>
#include <unistd.h>
#include <cstdlib>
#include <string>
#include <vector>

#include "rpt_gpu.hpp"
#include "rpt_oracle.h"

#define REQUIRE(cond)                                                                                                  \
	do {                                                                                                               \
		if (!(cond)) {                                                                                                 \
			fprintf(stderr, "FAILED %s:%d: %s\n", __FILE__, __LINE__, #cond);                                         \
			exit(1);                                                                                                   \
		}                                                                                                              \
	} while (0)

static rpto_col_t AsOracle(const rpt_col_t &c) { // same layout by construction (rpt_oracle.h)
	rpto_col_t o;
	static_assert(sizeof(rpto_col_t) == sizeof(rpt_col_t), "layouts must match");
	memcpy(&o, &c, sizeof(o));
	return o;
}

// One DataChunk's key vector: flat, or a dictionary vector (sel into a small dictionary), with optional NULLs.
struct Chunk {
	std::vector<int64_t> data;
	std::vector<uint32_t> sel;
	std::vector<uint64_t> validity;
	uint64_t count = 0;
	rpt_col_t Col() const {
		return rpt::HostColumn(data.data(), RPT_INT64, sel.empty() ? nullptr : sel.data(),
		                       validity.empty() ? nullptr : validity.data(), sel.empty() ? 0 : data.size());
	}
};

static std::vector<Chunk> MakeChunks(uint64_t rows, uint64_t seed, uint64_t key_space) {
	std::vector<Chunk> out;
	uint64_t x = seed;
	auto next = [&]() { return x = rpto_splitmix64(x); };
	for (uint64_t r = 0, c = 0; r < rows; c++) {
		Chunk ch;
		ch.count = std::min<uint64_t>(rpt::kStandardVectorSize - (c % 7 == 3 ? 5 : 0), rows - r); // some ragged chunks
		if (c % 5 == 2) { // dictionary vector
			ch.data.resize(97);
			for (auto &v : ch.data) {
				v = static_cast<int64_t>(next() % key_space);
			}
			ch.sel.resize(ch.count);
			for (auto &s : ch.sel) {
				s = static_cast<uint32_t>(next() % ch.data.size());
			}
		} else {
			ch.data.resize(ch.count);
			for (auto &v : ch.data) {
				v = static_cast<int64_t>(next() % key_space);
			}
		}
		if (c % 4 == 1) { // NULLs over the data entries
			ch.validity.assign((ch.data.size() + 63) / 64, 0);
			for (size_t i = 0; i < ch.data.size(); i++) {
				if (next() % 10 != 0) {
					ch.validity[i >> 6] |= 1ULL << (i & 63);
				}
			}
		}
		r += ch.count;
		out.push_back(std::move(ch));
	}
	return out;
}

// The reference's flow on the CPU oracle: insert per chunk into a filter sized for est; rebuild for the actual
// row count when the rule fires (physical_create_bf.cpp:386-406).
static std::vector<uint64_t> OracleBuild(const std::vector<Chunk> &chunks, uint32_t est, uint64_t *sized_for) {
	uint64_t rows = 0;
	for (auto &c : chunks) {
		rows += c.count;
	}
	uint64_t size_rows = est;
	if (rows > 0 && rpto_needs_rebuild(est, rows)) {
		size_rows = rows;
	}
	std::vector<uint64_t> words(rpto_num_sectors(size_rows), 0);
	for (auto &c : chunks) {
		const rpto_col_t col = AsOracle(c.Col());
		rpto_insert(words.data(), words.size(), &col, 1, c.count);
	}
	*sized_for = size_rows;
	return words;
}

static void TestCreateAndUse(uint32_t est, uint64_t build_rows, bool expect_rebuild) {
	const auto build = MakeChunks(build_rows, 0xC0FFEE + est, 1 << 20);
	const auto probe = MakeChunks(build_rows * 3 + 17, 0xBEEF + est, 1 << 21);

	// CREATE_BF: three sink threads' local batches, round-robin chunks
	rpt::PTBloomFilter bf(0);
	bf.Initialize(est);
	REQUIRE(bf.IsEmpty() && !bf.finalized() && bf.SizedForRows() == est);
	rpt::CreateBFBatch local[3] = {rpt::CreateBFBatch(RPT_INT64), rpt::CreateBFBatch(RPT_INT64),
	                               rpt::CreateBFBatch(RPT_INT64)};
	for (size_t i = 0; i < build.size(); i++) {
		local[i % 3].Append(build[i].Col(), build[i].count);
	}
	const auto res = rpt::CreateBFBatch::Finalize(bf, {&local[0], &local[1], &local[2]}, /*want_minmax*/ true);
	REQUIRE(res.rebuilt == expect_rebuild);
	REQUIRE(bf.finalized() && !bf.IsEmpty());

	uint64_t sized_for = 0;
	const auto want = OracleBuild(build, est, &sized_for);
	REQUIRE(bf.SizedForRows() == sized_for);
	REQUIRE(bf.NumSectors() == want.size());
	std::vector<uint64_t> got(want.size());
	bf.ExportWords(got.data(), got.size());
	REQUIRE(memcmp(got.data(), want.data(), want.size() * 8) == 0);

	// zone maps (BFTableFilter::CheckStatistics) of row groups around the key space
	{
		std::vector<int64_t> mins, maxs;
		std::vector<uint8_t> has;
		uint64_t x = est;
		for (int g = 0; g < 300; g++) {
			x = rpto_splitmix64(x);
			const int64_t a = static_cast<int64_t>(x % (1 << 21)) - 1000;
			const int64_t w = static_cast<int64_t>((x >> 40) % 2100) - 20; // some inverted, some wider than 2048
			mins.push_back(a);
			maxs.push_back(a + w);
			has.push_back(g % 17 != 0);
		}
		std::vector<uint8_t> res(mins.size(), 0xff);
		bf.CheckStatistics(RPT_INT64, mins.data(), maxs.data(), has.data(), mins.size(), res.data());
		for (size_t g = 0; g < mins.size(); g++) {
			REQUIRE(res[g] == rpto_check_statistics(want.data(), want.size(), RPTO_INT64, &mins[g], &maxs[g], has[g]));
		}
	}

	// min / max of the valid build rows (UpdateMinMax)
	int64_t lo = INT64_MAX, hi = INT64_MIN;
	bool any = false;
	for (auto &c : build) {
		for (uint64_t i = 0; i < c.count; i++) {
			const uint64_t idx = c.sel.empty() ? i : c.sel[i];
			if (!c.validity.empty() && !((c.validity[idx >> 6] >> (idx & 63)) & 1)) {
				continue;
			}
			lo = std::min(lo, c.data[idx]);
			hi = std::max(hi, c.data[idx]);
			any = true;
		}
	}
	int64_t glo, ghi;
	memcpy(&glo, res.min_val, 8);
	memcpy(&ghi, res.max_val, 8);
	REQUIRE(res.has_minmax == any && glo == lo && ghi == hi && !res.has_nan);

	// USE_BF: batches of chunks; per-chunk selection vectors must equal LookupSel chunk by chunk
	rpt::UseBFBatch use(bf, RPT_INT64, 40000);
	std::vector<uint64_t> hashes(rpt::kStandardVectorSize);
	std::vector<uint32_t> sel(rpt::kStandardVectorSize);
	size_t batch_first = 0;
	uint64_t survivors = 0;
	auto flush = [&](size_t end) {
		use.Probe();
		REQUIRE(use.NumChunks() == end - batch_first);
		for (size_t i = batch_first; i < end; i++) {
			const rpto_col_t col = AsOracle(probe[i].Col());
			rpto_hash_columns(&col, 1, probe[i].count, hashes.data());
			const uint64_t f = rpto_lookup_hashes(want.data(), want.size(), hashes.data(), probe[i].count, sel.data());
			const auto got_chunk = use.Chunk(i - batch_first);
			REQUIRE(got_chunk.second == f);
			REQUIRE(memcmp(got_chunk.first, sel.data(), f * 4) == 0);
			survivors += f;
		}
		use.Clear();
		batch_first = end;
	};
	for (size_t i = 0; i < probe.size(); i++) {
		if (use.Append(probe[i].Col(), probe[i].count)) {
			flush(i + 1);
		}
	}
	flush(probe.size());
	REQUIRE(survivors > 0);

	// the plain per-chunk stub gives the same answers
	for (size_t i = 0; i < 3 && i < probe.size(); i++) {
		const rpt_col_t col = probe[i].Col();
		const uint64_t f = bf.LookupSel(&col, 1, probe[i].count, sel.data());
		const rpto_col_t ocol = AsOracle(col);
		std::vector<uint32_t> osel(rpt::kStandardVectorSize);
		REQUIRE(f == rpto_probe(want.data(), want.size(), &ocol, 1, probe[i].count, osel.data()));
		REQUIRE(memcmp(sel.data(), osel.data(), f * 4) == 0);
	}
	// scan-side BFTableFilter: a vector of 2048 keys, first unfiltered, then through a selection left by an earlier
	// table filter; FilterValue and CheckStatistics on single values
	{
		rpt::BFTableFilter scan(bf);
		std::vector<int64_t> vec(rpt::kStandardVectorSize);
		uint64_t x = 77 + est;
		for (size_t i = 0; i < vec.size(); i++) {
			x = rpto_splitmix64(x);
			vec[i] = (i % 3 == 0) ? build[0].data[build[0].sel.empty() ? i % build[0].count : build[0].sel[i % build[0].count]]
			                      : static_cast<int64_t>(x >> 2);
		}
		rpt_col_t keys {};
		keys.data = vec.data();
		keys.type = RPT_INT64;
		keys.mem = RPT_MEM_HOST;
		keys.data_len = vec.size();
		std::vector<uint64_t> hashes(vec.size());
		rpto_col_t okeys = AsOracle(keys);
		rpto_hash_columns(&okeys, 1, vec.size(), hashes.data());
		std::vector<uint32_t> osel(vec.size() + 1), ssel(vec.size());
		const uint64_t of = rpto_lookup_hashes(want.data(), want.size(), hashes.data(), vec.size(), osel.data());
		bool is_set = false;
		uint64_t approved = vec.size();
		REQUIRE(scan.Filter(keys, ssel.data(), is_set, approved) == of);
		REQUIRE(approved == of && (is_set || of == vec.size()));
		REQUIRE(of == vec.size() || memcmp(ssel.data(), osel.data(), of * 4) == 0);
		// now with an incoming selection: every other row
		std::vector<uint32_t> in(vec.size() / 2);
		for (size_t i = 0; i < in.size(); i++) {
			in[i] = static_cast<uint32_t>(2 * i + 1);
		}
		std::vector<uint32_t> expect;
		for (uint32_t r : in) {
			const uint64_t h = hashes[r];
			const uint64_t m = rpto_mask(h);
			if ((want[h & (want.size() - 1)] & m) == m) {
				expect.push_back(r);
			}
		}
		is_set = true;
		approved = in.size();
		const uint64_t f2 = scan.Filter(keys, in.data(), is_set, approved);
		REQUIRE(f2 == expect.size() && approved == f2);
		REQUIRE(f2 == in.size() || memcmp(in.data(), expect.data(), f2 * 4) == 0);
		for (size_t i = 0; i < 64; i++) {
			const uint64_t m = rpto_mask(hashes[i]);
			REQUIRE(scan.FilterValue(hashes[i]) == ((want[hashes[i] & (want.size() - 1)] & m) == m));
		}
		const int64_t lo = vec[0] - 3, hi = vec[0] + 3;
		REQUIRE(scan.CheckStatistics(RPT_INT64, &lo, &hi) ==
		        rpto_check_statistics(want.data(), want.size(), RPTO_INT64, &lo, &hi, 1));
		REQUIRE(scan.CheckStatistics(RPT_INT64, &vec[0], &vec[0]) == RPT_FILTER_ALWAYS_TRUE); // a build key
	}
	printf("  est=%u rows=%llu S=%llu rebuilt=%d survivors=%llu\n", est, (unsigned long long)build_rows,
	       (unsigned long long)want.size(), (int)res.rebuilt, (unsigned long long)survivors);
}

// duckdb::string_t as the extension hands it over: 16 bytes, inlined up to 12 bytes, else prefix + pointer
struct StringT {
	uint8_t raw[16];
};
static StringT MakeStringT(const std::string &s) {
	StringT t;
	memset(&t, 0, sizeof(t));
	const uint32_t len = static_cast<uint32_t>(s.size());
	memcpy(t.raw, &len, 4);
	if (len <= 12) {
		memcpy(t.raw + 4, s.data(), len);
	} else {
		memcpy(t.raw + 4, s.data(), 4);
		const char *p = s.data();
		memcpy(t.raw + 8, &p, sizeof(p));
	}
	return t;
}

static void TestStringsAndGuards() {
	// VARCHAR key column (config 3's second filter): decimal text of the key, 1..19 bytes, some NULLs
	const uint64_t n = 50000;
	std::vector<std::string> pool(n);
	std::vector<StringT> vec(n);
	std::vector<uint64_t> validity((n + 63) / 64, ~0ULL);
	for (uint64_t i = 0; i < n; i++) {
		pool[i] = std::to_string(rpto_splitmix64(i) >> (i % 60));
		vec[i] = MakeStringT(pool[i]);
		if (i % 13 == 5) {
			validity[i >> 6] &= ~(1ULL << (i & 63));
		}
	}
	rpt::PTBloomFilter bf(0);
	bf.Initialize(static_cast<uint32_t>(n));

	// guards of PhysicalUseBF::ExecuteInternal before anything was built
	rpt::UseBFBatch use(bf, RPT_STRING_T);
	use.Append(rpt::HostColumn(vec.data(), RPT_STRING_T), 100);
	use.Probe(); // not finalised: skipped, every row passes
	REQUIRE(use.Chunk(0).second == 100 && use.Chunk(0).first[99] == 99);
	use.Clear();
	bf.Finalize(0); // finalised but empty: nothing passes
	use.Append(rpt::HostColumn(vec.data(), RPT_STRING_T), 100);
	use.Probe();
	REQUIRE(use.Chunk(0).second == 0);
	use.Clear();

	bf.Initialize(static_cast<uint32_t>(n));
	rpt::CreateBFBatch batch(RPT_STRING_T);
	for (uint64_t r = 0; r < n; r += rpt::kStandardVectorSize) {
		const uint64_t cnt = std::min<uint64_t>(rpt::kStandardVectorSize, n - r);
		// validity is addressed by entry index: hand the chunk a shifted copy of its bits
		std::vector<uint64_t> v((cnt + 63) / 64, 0);
		for (uint64_t i = 0; i < cnt; i++) {
			if ((validity[(r + i) >> 6] >> ((r + i) & 63)) & 1) {
				v[i >> 6] |= 1ULL << (i & 63);
			}
		}
		batch.Append(rpt::HostColumn(vec.data() + r, RPT_STRING_T, nullptr, v.data()), cnt);
	}
	const auto res = rpt::CreateBFBatch::Finalize(bf, {&batch}, true);
	REQUIRE(!res.rebuilt && !res.has_minmax); // VARCHAR ranges stay on the host

	// oracle: Arrow layout of the same strings
	std::vector<uint32_t> off(n + 1, 0);
	std::string bytes;
	for (uint64_t i = 0; i < n; i++) {
		const bool valid = (validity[i >> 6] >> (i & 63)) & 1;
		if (valid) {
			bytes += pool[i];
		}
		off[i + 1] = static_cast<uint32_t>(bytes.size());
	}
	bytes.resize(bytes.size() + 8);
	rpto_col_t ocol {};
	ocol.data = off.data();
	ocol.aux = bytes.data();
	ocol.aux_bytes = bytes.size();
	ocol.validity = validity.data();
	ocol.type = RPTO_VARCHAR;
	std::vector<uint64_t> want(rpto_num_sectors(n), 0);
	rpto_insert(want.data(), want.size(), &ocol, 1, n);
	std::vector<uint64_t> got(want.size());
	bf.ExportWords(got.data(), got.size());
	REQUIRE(memcmp(got.data(), want.data(), want.size() * 8) == 0);

	// probe the same strings through a dictionary vector: every non-NULL member passes, NULLs pass too
	// (NULL hashes to NULL_HASH and was inserted)
	std::vector<uint32_t> sel_in(4096);
	for (size_t i = 0; i < sel_in.size(); i++) {
		sel_in[i] = static_cast<uint32_t>((i * 7919) % n);
	}
	std::vector<uint32_t> sel(sel_in.size());
	const rpt_col_t dict = rpt::HostColumn(vec.data(), RPT_STRING_T, sel_in.data(), validity.data(), n);
	REQUIRE(bf.LookupSel(&dict, 1, sel_in.size(), sel.data()) == sel_in.size());

	// error behaviour: a bad type surfaces as an exception carrying the library's message
	bool threw = false;
	try {
		rpt_col_t bad = rpt::HostColumn(vec.data(), 9999);
		bf.Insert(&bad, 1, 1);
	} catch (const rpt::Error &e) {
		threw = std::string(e.what()).find("Invalid type for hash") != std::string::npos;
	}
	REQUIRE(threw);
	printf("  strings: n=%llu S=%llu\n", (unsigned long long)n, (unsigned long long)want.size());
}

static void OnCrash(int sig) { // no debugger on the GPU box: say where
	void *frames[64];
	const int n = backtrace(frames, 64);
	fprintf(stderr, "signal %d, backtrace:\n", sig);
	backtrace_symbols_fd(frames, n, 2);
	_exit(139);
}

int main() {
	signal(SIGSEGV, OnCrash);
	setvbuf(stdout, nullptr, _IONBF, 0);
	try {
		TestCreateAndUse(/*est*/ 300000, /*rows*/ 300000, /*rebuild*/ false);
		TestCreateAndUse(/*est*/ 10000, /*rows*/ 300000, /*rebuild*/ true); // under-estimate: :396 fires
		TestCreateAndUse(/*est*/ 5000000, /*rows*/ 70000, /*rebuild*/ false); // over-estimate: never shrinks
		TestStringsAndGuards();
	} catch (const std::exception &e) {
		fprintf(stderr, "FAILED with exception: %s\n", e.what());
		return 1;
	}
	printf("host mirror ok\n");
	return 0;
}
