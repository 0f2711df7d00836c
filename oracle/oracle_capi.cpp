// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle_roaring.hpp). C entry points for ctypes.
// The `db` arguments are silo::Database* handles created by libsilo_host.so (silo_db_*).
#include <chrono>
#include <cstdlib>
#include <cstring>

#include "../lapis-silo_b200/host/optree_json.hpp"
#include "oracle_engine.hpp"

namespace {
thread_local std::string g_error;
char* dupString(const std::string& s) {
   char* out = static_cast<char*>(malloc(s.size() + 1));
   memcpy(out, s.c_str(), s.size() + 1);
   return out;
}
template <typename F>
int guarded(F&& f) {
   try { f(); return 0; }
   catch (const silo::QueryParseException& e) { g_error = std::string("parse:") + e.what(); return 1; }
   catch (const silo::QueryCompilationException& e) { g_error = std::string("compile:") + e.what(); return 2; }
   catch (const std::exception& e) { g_error = std::string("other:") + e.what(); return 3; }
}
int exportIds(const silo::roaring::Bitmap& bm, uint32_t** ids, size_t* n) {
   auto v = bm.toVector();
   *n = v.size();
   *ids = static_cast<uint32_t*>(malloc(std::max<size_t>(1, v.size()) * sizeof(uint32_t)));
   memcpy(*ids, v.data(), v.size() * sizeof(uint32_t));
   return 0;
}
}  // namespace

extern "C" {

const char* oracle_last_error() { return g_error.c_str(); }
void oracle_free(void* p) { free(p); }

char* oracle_execute_query(const void* db, const char* query_json, int faithful, int threads) {
   char* out = nullptr;
   guarded([&] {
      oracle::Options options{faithful != 0, static_cast<unsigned>(threads), {}};
      auto result = oracle::executeQuery(*static_cast<const silo::Database*>(db), query_json, options);
      out = dupString(silo::toJson(result).dump());
   });
   return out;
}

// Bounded sample for the timed CPU baseline: the query restricted to `partitions`; returns seconds of
// the query restricted to some partitions (the others count as empty filters): what one rank of a sharded database answers
char* oracle_execute_query_subset(const void* db, const char* query_json, const uint32_t* partitions, uint32_t n) {
   char* out = nullptr;
   guarded([&] {
      oracle::Options options{false, 0, std::vector<uint32_t>(partitions, partitions + n)};
      auto result = oracle::executeQuery(*static_cast<const silo::Database*>(db), query_json, options);
      out = dupString(silo::dumpJson(result));
   });
   return out;
}

// wall-clock spent in executeQuery (filter + action), or a negative value on error.
double oracle_time_query_subset(const void* db, const char* query_json, int faithful, int threads, const uint32_t* partitions, uint32_t n, uint64_t* result_rows) {
   double seconds = -1;
   guarded([&] {
      oracle::Options options{faithful != 0, static_cast<unsigned>(threads), std::vector<uint32_t>(partitions, partitions + n)};
      const auto begin = std::chrono::steady_clock::now();
      auto result = oracle::executeQuery(*static_cast<const silo::Database*>(db), query_json, options);
      seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - begin).count();
      if (result_rows) *result_rows = result.query_result.size();
   });
   return seconds;
}

// the same on positions [0, position_limit) only (faithful mode; the result rows are not meaningful): bench.py's bounded CPU sample
double oracle_time_query_sample(const void* db, const char* query_json, int threads, const uint32_t* partitions, uint32_t n, uint32_t position_limit) {
   double seconds = -1;
   guarded([&] {
      oracle::Options options{true, static_cast<unsigned>(threads), std::vector<uint32_t>(partitions, partitions + n), position_limit};
      const auto begin = std::chrono::steady_clock::now();
      oracle::executeQuery(*static_cast<const silo::Database*>(db), query_json, options);
      seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - begin).count();
   });
   return seconds;
}

int oracle_eval_optree(const char* optree_json, uint32_t** ids, size_t* n) {
   return guarded([&] {
      auto holder = silo::opTreeFromJson(optree_json);
      auto result = oracle::evaluate(*holder->root);
      exportIds(result.get(), ids, n);
   });
}

int oracle_filter_ids(const void* db_, const char* filter_json, uint32_t partition, uint32_t** ids, size_t* n) {
   return guarded([&] {
      const auto& db = *static_cast<const silo::Database*>(db_);
      auto expression = silo::filter::parseExpression(silo::json::parse(filter_json));
      auto op = expression->compile(db, *db.partitions.at(partition), silo::filter::AmbiguityMode::NONE);
      auto result = oracle::evaluate(*op);
      exportIds(result.get(), ids, n);
   });
}

// |F_p| only (no id export): lets a test classify partitions as empty / partial / full under a filter
int oracle_filter_cardinality(const void* db_, const char* filter_json, uint32_t partition, uint64_t* cardinality) {
   return guarded([&] {
      const auto& db = *static_cast<const silo::Database*>(db_);
      auto expression = silo::filter::parseExpression(silo::json::parse(filter_json));
      auto op = expression->compile(db, *db.partitions.at(partition), silo::filter::AmbiguityMode::NONE);
      *cardinality = oracle::evaluate(*op).get().cardinality();
   });
}

// Full uint32[Σ][L] count table of one sequence for a Mutations / AminoAcidMutations query.
// Returns 4 if the sequence produced no table (every partition's filter empty): `out` is zeroed.
int oracle_mutation_counts(const void* db_, const char* query_json, const char* sequence_name, int faithful, int threads, uint32_t* out, size_t out_len) {
   int status = 0;
   int rc = guarded([&] {
      const auto& db = *static_cast<const silo::Database*>(db_);
      std::map<std::string, std::vector<uint32_t>> tables;
      oracle::Options options{faithful != 0, static_cast<unsigned>(threads), {}};
      oracle::executeQuery(db, query_json, options, &tables);
      memset(out, 0, out_len * sizeof(uint32_t));
      auto it = tables.find(sequence_name);
      if (it == tables.end()) { status = 4; return; }
      if (it->second.size() != out_len) throw std::runtime_error("count table size mismatch");
      memcpy(out, it->second.data(), out_len * sizeof(uint32_t));
   });
   return rc ? rc : status;
}

// The same restricted to `partitions` (the others count as empty filters): the local table of one rank
// in the world-size-2 tests of the multi-GPU protocol.
int oracle_mutation_counts_subset(const void* db_, const char* query_json, const char* sequence_name, int faithful, int threads,
                                  const uint32_t* partitions, uint32_t n, uint32_t* out, size_t out_len) {
   int status = 0;
   int rc = guarded([&] {
      const auto& db = *static_cast<const silo::Database*>(db_);
      std::map<std::string, std::vector<uint32_t>> tables;
      oracle::Options options{faithful != 0, static_cast<unsigned>(threads), std::vector<uint32_t>(partitions, partitions + n)};
      oracle::executeQuery(db, query_json, options, &tables);
      memset(out, 0, out_len * sizeof(uint32_t));
      auto it = tables.find(sequence_name);
      if (it == tables.end()) { status = 4; return; }
      if (it->second.size() != out_len) throw std::runtime_error("count table size mismatch");
      memcpy(out, it->second.data(), out_len * sizeof(uint32_t));
   });
   return rc ? rc : status;
}

}  // extern "C"
