// C entry points of the GPU query engine for ctypes (tests, bench.py, __graft_entry__) — the same calls
// a C++ host makes through engine.hpp. Every bitmap operation runs in libsilo_b200.so.
#include "capi_common.hpp"
#include "engine.hpp"

using namespace silo;
using engine::DeviceDatabase;

namespace {

int exportIds(const DeviceDatabase::EvaluatedFilters& filters, size_t index, uint32_t** ids, size_t* n) {
   size_t count = 0;
   engine::check(sb_result_export_ids(filters.results.at(index), nullptr, 0, &count));
   *ids = static_cast<uint32_t*>(malloc(std::max<size_t>(1, count) * sizeof(uint32_t)));
   engine::check(sb_result_export_ids(filters.results.at(index), *ids, count, n));
   return 0;
}

// One query in flight across ranks: local filters evaluated, tables reduced by the caller (NCCL), then finished.
struct QueryState {
   const DeviceDatabase* device_db;
   Query query;
   DeviceDatabase::EvaluatedFilters filters;
   std::vector<std::string> names;
   std::map<std::string, std::vector<uint32_t>> tables;   // reduced tables handed back by the caller (host memory)
   std::map<std::string, std::vector<sb_mutation_row>> rows;   // or the rows selected on the device from a reduced device table
};

}  // namespace

extern "C" {

// `partitions` = global partition indices this GPU owns (all when n == 0)
void* silo_engine_create(const void* db_, int device, const uint32_t* partitions, uint32_t n) {
   DeviceDatabase* out = nullptr;
   capi::guarded([&] {
      const auto& db = *static_cast<const Database*>(db_);
      std::vector<uint32_t> owned(partitions, partitions + n);
      if (n == 0) for (uint32_t p = 0; p < db.partitions.size(); ++p) owned.push_back(p);
      out = new DeviceDatabase(db, device, std::move(owned));
   });
   return out;
}
void silo_engine_free(void* engine) { delete static_cast<DeviceDatabase*>(engine); }
void* silo_engine_handle(void* engine) { return static_cast<DeviceDatabase*>(engine)->handle; }

// QueryEngine::executeQuery (query_engine.cpp:30-68) -> {"queryResult": [...]}
char* silo_engine_execute(void* engine, const char* query_json) {
   char* out = nullptr;
   capi::guarded([&] { out = capi::dupString(static_cast<DeviceDatabase*>(engine)->executeQueryJson(query_json)); });
   return out;
}

// Full uint32[sigma][L] table of one sequence; returns 4 (table zeroed) when no partition passes the filter.
int silo_engine_mutation_counts(void* engine_, const char* query_json, const char* sequence_name, uint32_t* out, size_t out_len) {
   int status = 0;
   const int rc = capi::guarded([&] {
      auto& engine = *static_cast<DeviceDatabase*>(engine_);
      Query query = parseQuery(query_json);
      auto filters = engine.evaluateFilter(*query.filter);
      std::vector<uint32_t> counts;
      const bool produced = engine.mutationCounts(filters, sequence_name, query.action.alphabetId(), counts);
      if (counts.size() != out_len) throw std::runtime_error("count table size mismatch");
      memcpy(out, counts.data(), out_len * sizeof(uint32_t));
      if (!produced) status = 4;
   });
   return rc ? rc : status;
}

int silo_engine_filter_ids(void* engine_, const char* filter_json, uint32_t local_partition, uint32_t** ids, size_t* n) {
   return capi::guarded([&] {
      auto& engine = *static_cast<DeviceDatabase*>(engine_);
      auto expression = filter::parseExpression(json::parse(filter_json));
      auto op = expression->compile(engine.db, *engine.db.partitions.at(engine.owned.at(local_partition)), filter::AmbiguityMode::NONE);
      DeviceDatabase::Plan plan;
      engine.lower(*op, local_partition, plan);
      DeviceDatabase::EvaluatedFilters filters;
      filters.handle = engine.handle;
      filters.partitions = {local_partition};
      filters.results = {nullptr};
      engine::check(sb_filter_eval(engine.handle, local_partition, plan.nodes.data(), static_cast<uint32_t>(plan.nodes.size()), &filters.results[0]));
      exportIds(filters, 0, ids, n);
   });
}

// The same filter as CRoaring portable bytes (sb_result_export_portable): what a host-side action of the reference would readSafe
int silo_engine_filter_portable(void* engine_, const char* filter_json, uint32_t local_partition, uint8_t** bytes, size_t* n) {
   return capi::guarded([&] {
      auto& engine = *static_cast<DeviceDatabase*>(engine_);
      auto expression = filter::parseExpression(json::parse(filter_json));
      auto op = expression->compile(engine.db, *engine.db.partitions.at(engine.owned.at(local_partition)), filter::AmbiguityMode::NONE);
      DeviceDatabase::Plan plan;
      engine.lower(*op, local_partition, plan);
      DeviceDatabase::EvaluatedFilters filters;
      filters.handle = engine.handle;
      filters.partitions = {local_partition};
      filters.results = {nullptr};
      engine::check(sb_filter_eval(engine.handle, local_partition, plan.nodes.data(), static_cast<uint32_t>(plan.nodes.size()), &filters.results[0]));
      size_t size = 0;
      engine::check(sb_result_export_portable(filters.results[0], nullptr, 0, &size));
      *bytes = static_cast<uint8_t*>(malloc(std::max<size_t>(1, size)));
      engine::check(sb_result_export_portable(filters.results[0], *bytes, size, n));
   });
}
// Test hook: the ids held by CRoaring portable bytes, read with the host's restated reader (host/roaring.hpp readPortable)
int silo_portable_ids(const uint8_t* bytes, size_t n_bytes, uint32_t** ids, size_t* n) {
   return capi::guarded([&] {
      size_t consumed = 0;
      const Bitmap bitmap = Bitmap::readPortable(bytes, n_bytes, &consumed);
      if (consumed != n_bytes) throw std::runtime_error("portable bytes: " + std::to_string(n_bytes - consumed) + " bytes left over");
      const std::vector<uint32_t> values = bitmap.toVector();
      *ids = static_cast<uint32_t*>(malloc(std::max<size_t>(1, values.size()) * sizeof(uint32_t)));
      memcpy(*ids, values.data(), values.size() * sizeof(uint32_t));
      *n = values.size();
   });
}

// Operator known-answer tests: a literal operator tree evaluated on the GPU.
int silo_engine_eval_optree(int device, const char* optree_json, uint32_t** ids, size_t* n) {
   return capi::guarded([&] {
      static const Database empty_database;
      auto holder = opTreeFromJson(optree_json);
      DeviceDatabase engine(empty_database, device, *holder);
      auto filters = engine.evaluateTrees({holder->root.get()});
      exportIds(filters, 0, ids, n);
   });
}

// ---- multi-GPU protocol: begin (local filters) -> per-sequence device tables -> caller reduces -> finish ----
void* silo_query_begin(void* engine_, const char* query_json) {
   QueryState* state = nullptr;
   capi::guarded([&] {
      auto& engine = *static_cast<DeviceDatabase*>(engine_);
      auto s = std::make_unique<QueryState>();
      s->device_db = &engine;
      s->query = parseQuery(query_json);
      s->query.action.validateOrderByFields();
      if (s->query.action.type != ActionType::AGGREGATED) s->names = engine.mutationSequenceNames(s->query.action);
      else if (!s->query.action.group_by_fields.empty()) throw QueryCompilationException("Aggregated with groupByFields is outside the hot path (SURVEY 8f)");
      s->filters = engine.evaluateFilter(*s->query.filter);
      state = s.release();
   });
   return state;
}
void silo_query_free(void* state) { delete static_cast<QueryState*>(state); }
uint32_t silo_query_sequence_count(void* state) { return static_cast<uint32_t>(static_cast<QueryState*>(state)->names.size()); }
const char* silo_query_sequence_name(void* state, uint32_t i) { return static_cast<QueryState*>(state)->names.at(i).c_str(); }
uint64_t silo_query_table_len(void* state_, uint32_t i) {
   auto* s = static_cast<QueryState*>(state_);
   const int alphabet_id = s->query.action.alphabetId();
   return uint64_t(alphabetById(alphabet_id).count) * s->device_db->db.getReferences(alphabet_id).at(s->names.at(i)).size();
}
int silo_query_local_count(void* state_, uint32_t* count) {
   return capi::guarded([&] { auto* s = static_cast<QueryState*>(state_); *count = s->device_db->aggregatedCount(s->filters); });
}
// local table of sequence i into DEVICE memory (a torch tensor's data_ptr); produced = 0 if all local filters are empty
int silo_query_local_table_device(void* state_, uint32_t i, void* device_counts, uint64_t len, int* produced) {
   return capi::guarded([&] {
      auto* s = static_cast<QueryState*>(state_);
      engine::check(sb_mutations_device(s->device_db->handle, s->names.at(i).c_str(), static_cast<uint32_t>(s->filters.results.size()),
                                        s->filters.partitions.data(), s->filters.results.data(), device_counts, len, produced));
   });
}
int silo_query_set_table(void* state_, uint32_t i, const uint32_t* counts, uint64_t len) {
   return capi::guarded([&] { auto* s = static_cast<QueryState*>(state_); s->tables[s->names.at(i)].assign(counts, counts + len); });
}
// root rank: the reduced table stays in device memory; only the selected rows come back
int silo_query_set_table_device(void* state_, uint32_t i, const void* device_counts) {
   return capi::guarded([&] {
      auto* s = static_cast<QueryState*>(state_);
      const std::string& name = s->names.at(i);
      s->device_db->mutationRowsFromTable(name, s->query.action.alphabetId(), device_counts, s->query.action.min_proportion, s->rows[name]);
   });
}
// root rank: assemble the result from the reduced tables / the reduced count
char* silo_query_finish(void* state_, uint32_t reduced_count) {
   char* out = nullptr;
   capi::guarded([&] {
      auto* s = static_cast<QueryState*>(state_);
      const Action& action = s->query.action;
      QueryResult result = action.executeAndOrder([&]() -> QueryResult {
         if (action.type == ActionType::AGGREGATED) {
            QueryResultEntry entry;
            entry.fields["count"] = static_cast<int32_t>(reduced_count);
            return QueryResult{{entry}};
         }
         QueryResult r;
         const int alphabet_id = action.alphabetId();
         for (const auto& name : s->names) {
            const auto& reference = s->device_db->db.getReferences(alphabet_id).at(name);
            if (auto rit = s->rows.find(name); rit != s->rows.end()) { DeviceDatabase::addRowsToOutput(name, alphabetById(alphabet_id), reference, rit->second, r.query_result); continue; }
            auto it = s->tables.find(name);
            if (it == s->tables.end()) continue;
            addMutationsToOutput(name, alphabetById(alphabet_id), s->device_db->db.getReferences(alphabet_id).at(name), it->second.data(), action.min_proportion, r.query_result);
         }
         return r;
      });
      out = capi::dupString(dumpJson(result));
   });
   return out;
}

// The root rank's last step with a reduced table in HOST memory and no engine: parse, M5, order, serialise.
// (What silo_query_finish does after silo_query_set_table; callable without a device for the gloo tests.)
char* silo_rows_from_table(const void* db_, const char* query_json, const char* sequence_name, const uint32_t* counts, uint64_t len) {
   char* out = nullptr;
   capi::guarded([&] {
      const auto& db = *static_cast<const Database*>(db_);
      Query query = parseQuery(query_json);
      const int alphabet_id = query.action.alphabetId();
      const auto& reference = db.getReferences(alphabet_id).at(sequence_name);
      if (len != uint64_t(alphabetById(alphabet_id).count) * reference.size()) throw std::runtime_error("count table size mismatch");
      QueryResult result = query.action.executeAndOrder([&]() -> QueryResult {
         QueryResult r;
         addMutationsToOutput(sequence_name, alphabetById(alphabet_id), reference, counts, query.action.min_proportion, r.query_result);
         return r;
      });
      out = capi::dupString(dumpJson(result));
   });
   return out;
}

// ---- prepared queries: compile + lower + upload + locate once, replay the whole query as one submission ----
void* silo_prepared_create(void* engine_, const char* query_json, int use_graph) {
   DeviceDatabase::PreparedQuery* out = nullptr;
   capi::guarded([&] { out = static_cast<DeviceDatabase*>(engine_)->prepare(query_json, use_graph != 0).release(); });
   return out;
}
void silo_prepared_free(void* prepared) { delete static_cast<DeviceDatabase::PreparedQuery*>(prepared); }

// One device-resident step: the rows stay in the library's pinned buffer, nothing is formatted. Returns the number of
// result rows over all sequences (Aggregated: the count) through `result_rows`.
int silo_prepared_run_device(void* engine_, void* prepared_, uint64_t* result_rows) {
   return capi::guarded([&] {
      auto& p = *static_cast<DeviceDatabase::PreparedQuery*>(prepared_);
      thread_local DeviceDatabase::FusedResult fused;
      std::lock_guard<std::mutex> lock(p.mutex);
      static_cast<DeviceDatabase*>(engine_)->runPrepared(p, fused);
      if (result_rows) {
         uint64_t n = 0;
         if (p.query.action.type == ActionType::AGGREGATED) n = fused.count();
         else for (const auto& s : fused.sequences) n += s.rows.size();
         *result_rows = n;
      }
   });
}
// A batch of prepared queries from the calling thread (sb_query_run_many): texts[i] receives the response of prepared[i]
// (free each with silo_free). Returns 0 on success.
int silo_prepared_run_many(void* engine_, void* const* prepared, uint32_t n, char** texts) {
   return capi::guarded([&] {
      auto& engine = *static_cast<DeviceDatabase*>(engine_);
      std::vector<DeviceDatabase::PreparedQuery*> queries;
      std::vector<std::unique_lock<std::mutex>> locks;
      for (uint32_t i = 0; i < n; ++i) { queries.push_back(static_cast<DeviceDatabase::PreparedQuery*>(prepared[i])); locks.emplace_back(queries.back()->mutex); }
      std::vector<DeviceDatabase::FusedResult> results;
      engine.runPreparedMany(queries, results);
      for (uint32_t i = 0; i < n; ++i) texts[i] = capi::dupString(engine.responseText(queries[i]->query.action, queries[i]->names, results[i]));
   });
}
// The response text of a prepared query (what silo_engine_execute returns for the same text)
char* silo_prepared_run(void* engine_, void* prepared_) {
   char* out = nullptr;
   capi::guarded([&] {
      auto& engine = *static_cast<DeviceDatabase*>(engine_);
      auto& p = *static_cast<DeviceDatabase::PreparedQuery*>(prepared_);
      DeviceDatabase::FusedResult fused;
      { std::lock_guard<std::mutex> lock(p.mutex); engine.runPrepared(p, fused); }
      out = capi::dupString(engine.responseText(p.query.action, p.names, fused));
   });
   return out;
}
// The uint32 count table of sequence `name_index` of the last run, copied to the host (checksums, parity tests)
int silo_prepared_table(void* prepared_, uint32_t name_index, uint32_t* out, uint64_t len) {
   return capi::guarded([&] {
      auto& p = *static_cast<DeviceDatabase::PreparedQuery*>(prepared_);
      if (!p.handle) throw std::runtime_error("prepared query owns no partition");
      const void* device = nullptr;
      size_t n = 0;
      engine::check(sb_query_table_device(p.handle, name_index, &device, &n));
      if (n != len) throw std::runtime_error("count table size mismatch");
      engine::check(sb_copy_from_device(out, device, n * sizeof(uint32_t)));
   });
}
// Algorithmic bytes of the prepared Mutations query for sequence `name_index`: evaluates the filters through the
// split entry points (costs a filter export: not for timed steps).
int silo_prepared_bytes(void* engine_, void* prepared_, uint32_t name_index, uint64_t* alg_bytes, uint64_t* k1_bytes) {
   return capi::guarded([&] {
      auto& engine = *static_cast<DeviceDatabase*>(engine_);
      auto& p = *static_cast<DeviceDatabase::PreparedQuery*>(prepared_);
      DeviceDatabase::EvaluatedFilters filters;
      filters.handle = engine.handle;
      filters.partitions = p.compiled.partitions;
      filters.results.assign(filters.partitions.size(), nullptr);
      if (filters.partitions.empty()) { *alg_bytes = 0; *k1_bytes = 0; return; }
      engine::check(sb_filter_eval_batch(engine.handle, static_cast<uint32_t>(filters.partitions.size()), filters.partitions.data(), p.compiled.plan_ptrs.data(), p.compiled.plan_sizes.data(), filters.results.data()));
      engine::check(sb_mutations_algorithmic_bytes(engine.handle, p.names.at(name_index).c_str(), static_cast<uint32_t>(filters.results.size()), filters.partitions.data(),
                                                   filters.results.data(), alg_bytes, k1_bytes));
   });
}
float silo_prepared_elapsed_ms(void* prepared_) {
   auto& p = *static_cast<DeviceDatabase::PreparedQuery*>(prepared_);
   float ms = 0;
   if (p.handle) sb_query_elapsed_ms(p.handle, &ms);
   return ms;
}
void silo_engine_set_plan_cache(void* engine_, int enabled) { static_cast<DeviceDatabase*>(engine_)->setPlanCache(enabled != 0); }

// Multi-GPU (one process per GPU, sb_comm_init done): every rank calls with the same query text in the same order; the
// library sums the count tables over the ranks inside the query's one submission. Rank 0 gets the response text, the
// other ranks an empty string. Mutations / AminoAcidMutations only (an Aggregated count is one integer the caller sums).
char* silo_engine_execute_sharded(void* engine_, const char* query_json, int is_root) {
   char* out = nullptr;
   capi::guarded([&] {
      auto& engine = *static_cast<DeviceDatabase*>(engine_);
      Query query = parseQuery(query_json);
      const Action& action = query.action;
      if (action.type == ActionType::AGGREGATED) throw QueryCompilationException("silo_engine_execute_sharded serves Mutations actions");
      action.validateOrderByFields();
      const std::vector<std::string> names = engine.mutationSequenceNames(action);
      DeviceDatabase::CompiledFilter cf;
      engine.compileFilter(*query.filter, cf);
      DeviceDatabase::FusedResult fused;
      engine.runFused(cf, names, action.alphabetId(), action.min_proportion, fused);
      out = capi::dupString(is_root ? engine.responseText(action, names, fused) : std::string());
   });
   return out;
}

// Aggregated (with or without groupByFields) on a sharded database: every rank calls silo_engine_execute_partial, any channel
// carries the texts to one rank (the harness uses torch.distributed.gather_object), silo_merge_partials adds equal tuples up and
// applies orderByFields / limit / offset. `db` supplies the metadata configuration (any rank's Database has it).
char* silo_engine_execute_partial(void* engine_, const char* query_json) {
   char* out = nullptr;
   capi::guarded([&] {
      auto& engine = *static_cast<DeviceDatabase*>(engine_);
      out = capi::dupString(dumpJson(engine.executeAggregatedPartial(parseQuery(query_json))));
   });
   return out;
}
char* silo_merge_partials(const void* db_, const char* query_json, const char* const* partials, uint32_t n) {
   char* out = nullptr;
   capi::guarded([&] {
      const Query query = parseQuery(query_json);
      std::vector<std::string> texts(partials, partials + n);
      out = capi::dupString(dumpJson(engine::mergeAggregatedPartials(query.action, static_cast<const Database*>(db_)->config, texts)));
   });
   return out;
}

// Test hook for the host worker pool (tests/test_host_cpu.py; no device involved): runs n_jobs jobs that add
// their index + 1 to a sum, `rounds` batches in a row, on `n_threads` workers + the caller; job `throw_at` (if >= 0)
// throws. Returns the sum of the last batch, or -1 when the exception came back out of run().
int64_t silo_test_worker_pool(uint32_t n_threads, uint32_t n_jobs, uint32_t rounds, int32_t throw_at) {
   engine::WorkerPool pool(n_threads);
   std::atomic<int64_t> sum{0};
   try {
      for (uint32_t r = 0; r < rounds; ++r) {
         sum = 0;
         pool.run(n_jobs, [&](size_t k) {
            if (throw_at >= 0 && k == static_cast<size_t>(throw_at)) throw std::runtime_error("job failed");
            sum += static_cast<int64_t>(k) + 1;
         });
      }
   } catch (const std::runtime_error&) { return -1; }
   return sum.load();
}

}  // extern "C"
