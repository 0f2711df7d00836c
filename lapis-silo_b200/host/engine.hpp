// Host side of the GPU query path: mirrors QueryEngine::executeQuery
// (src/silo/query_engine/query_engine.cpp:30-68). Everything that touches a bitmap goes through the
// C ABI of libsilo_b200.so (include/silo_b200.h); this file only
//   * ingests a silo::Database into an sb_db (what a loader would do after
//     Database::loadDatabaseState, src/silo/database.cpp:480-560),
//   * lowers a compiled Operator tree to the postfix sb_node plan,
//   * runs M5/M6 of the Mutations action (thresholds, double division, rows) and Aggregated's row.
// There is no CPU evaluation of operators here: sb_* failures surface as exceptions.
#pragma once
#include <charconv>
#include <chrono>
#include <atomic>
#include <condition_variable>
#include <deque>
#include <exception>
#include <functional>
#include <mutex>
#include <thread>
#include <unordered_map>

#include "../../include/silo_b200.h"
#include "group_by.hpp"
#include "optree_json.hpp"
#include "query.hpp"

namespace silo::engine {

// A few host threads that run the per-sequence parts of one query side by side (AminoAcidMutations over all genes: twelve
// independent count tables). Every thread is bound to a lane of the library (streams + scratch of its own), so the
// kernels of different sequences overlap on the GPU. It also shares the per-partition compile + lower of a filter. One batch
// at a time (a second query that arrives meanwhile does its work on its own thread); the calling thread takes part.
class WorkerPool {
 public:
   explicit WorkerPool(unsigned n_threads) {
      for (unsigned k = 0; k < n_threads; ++k) threads_.emplace_back([this] { loop(); });
   }
   ~WorkerPool() {
      { std::lock_guard<std::mutex> lock(mutex_); stop_ = true; }
      wake_.notify_all();
      for (auto& t : threads_) t.join();
   }
   // runs job(0) ... job(n - 1), returns when all are done; rethrows the first exception
   void run(size_t n, const std::function<void(size_t)>& job) {
      std::unique_lock<std::mutex> one_batch(batch_mutex_, std::try_to_lock);
      if (!one_batch.owns_lock()) {      // the pool is busy with another query's batch: this thread does its own work
         for (size_t k = 0; k < n; ++k) job(k);
         return;
      }
      {
         std::lock_guard<std::mutex> lock(mutex_);
         job_ = &job; n_ = n; next_ = 0; left_ = n; error_ = nullptr;
      }
      wake_.notify_all();
      work();
      std::unique_lock<std::mutex> lock(mutex_);
      done_.wait(lock, [this] { return left_ == 0; });
      job_ = nullptr;
      if (error_) std::rethrow_exception(error_);
   }

 private:
   void work() {
      for (;;) {
         size_t index;
         const std::function<void(size_t)>* job;
         {
            std::lock_guard<std::mutex> lock(mutex_);
            if (!job_ || next_ >= n_) return;
            index = next_++;
            job = job_;
         }
         std::exception_ptr error;
         try { (*job)(index); } catch (...) { error = std::current_exception(); }
         std::lock_guard<std::mutex> lock(mutex_);
         if (error && !error_) error_ = error;
         if (--left_ == 0) done_.notify_all();
      }
   }
   void loop() {
      for (;;) {
         {
            std::unique_lock<std::mutex> lock(mutex_);
            wake_.wait(lock, [this] { return stop_ || (job_ && next_ < n_); });
            if (stop_) return;
         }
         work();
      }
   }
   std::vector<std::thread> threads_;
   std::mutex mutex_, batch_mutex_;
   std::condition_variable wake_, done_;
   const std::function<void(size_t)>* job_ = nullptr;
   size_t n_ = 0, next_ = 0, left_ = 0;
   std::exception_ptr error_;
   bool stop_ = false;
};

struct HostTrace {   // SB_TRACE=1: lap times of the host side on stderr (diagnostics only)
   bool on = getenv("SB_TRACE") != nullptr;
   std::chrono::steady_clock::time_point last = std::chrono::steady_clock::now();
   std::string text;
   void lap(const char* what) {
      if (!on) return;
      const auto now = std::chrono::steady_clock::now();
      text += std::string(" ") + what + "=" + std::to_string(std::chrono::duration<double, std::micro>(now - last).count()).substr(0, 7);
      last = now;
   }
   ~HostTrace() { if (on) fprintf(stderr, "[host] executeQuery:%s\n", text.c_str()); }
};

struct DeviceError : std::runtime_error { using std::runtime_error::runtime_error; };
inline void check(int rc) { if (rc != 0) throw DeviceError(std::string("silo_b200: ") + sb_last_error()); }

// Concatenated CRoaring portable serializations (what roaring_serialize.h:23-28 writes per bitmap)
struct PortableBlob {
   std::string bytes;
   std::vector<uint64_t> offsets{0};
   void add(const Bitmap& bm) {
      if (!bm.isEmpty()) bm.writePortable(bytes);
      offsets.push_back(bytes.size());
   }
};

class DeviceDatabase {
  public:
   const Database& db;
   std::vector<uint32_t> owned;   // global partition index of each device-local partition
   sb_db* handle = nullptr;

   DeviceDatabase(const Database& database, int device, std::vector<uint32_t> owned_partitions) : db(database), owned(std::move(owned_partitions)) {
      check(sb_db_create(device, static_cast<uint32_t>(owned.size()), &handle));
      try {
         for (uint32_t local = 0; local < owned.size(); ++local) ingestPartition(local, *db.partitions.at(owned[local]));
         std::sort(store_ranges.begin(), store_ranges.end(), [](const StoreRange& a, const StoreRange& b) { return a.begin < b.begin; });
         check(sb_db_finalize(handle));
         if (!owned.empty()) {
            for (int alphabet_id = 0; alphabet_id < 2; ++alphabet_id) {
               for (const auto& [name, reference] : db.getReferences(alphabet_id)) check(sb_sequence_set_reference(handle, name.c_str(), reference.data(), static_cast<uint32_t>(reference.size())));
            }
         }
      } catch (...) { sb_db_free(handle); handle = nullptr; throw; }
   }
   // Stand-alone operator trees (the reference's operator unit tests): one partition holding the
   // tree's literal missing-symbol stores and columns; literal index bitmaps travel with the plan.
   DeviceDatabase(const Database& empty_database, int device, const OpTreeHolder& holder) : db(empty_database), owned{0} {
      check(sb_db_create(device, 1, &handle));
      try {
         uint32_t counter = 0;
         for (const auto& store : holder.stores) {
            uint32_t length = 1;
            for (const auto& bm : store.missing_symbol_bitmaps) if (!bm.isEmpty()) length = std::max(length, bm.maximum() + 1);
            PortableBlob missing;
            for (const auto& bm : store.missing_symbol_bitmaps) missing.add(bm);
            const std::vector<uint64_t> slot_off(size_t(length) * 16 + 1, 0);
            const std::vector<int8_t> none(length, -1);
            const uint8_t dummy = 0;
            uint32_t store_id = 0;
            check(sb_store_add(handle, 0, ("kat" + std::to_string(counter++)).c_str(), 0, length, static_cast<uint32_t>(store.missing_symbol_bitmaps.size()),
                               &dummy, slot_off.data(), none.data(), none.data(), reinterpret_cast<const uint8_t*>(missing.bytes.data()), missing.offsets.data(), &store_id));
            store_ids[&store] = store_id;
         }
         auto addColumn = [&](int dtype, const void* data, size_t n) {
            if (n == 0) return;
            uint32_t id = 0;
            check(sb_column_add(handle, 0, dtype, data, static_cast<uint32_t>(n), &id));
            column_ids[data] = id;
         };
         for (const auto& c : holder.int_columns) addColumn(0, c.data(), c.size());
         for (const auto& c : holder.float_columns) addColumn(1, c.data(), c.size());
         for (const auto& c : holder.date_columns) addColumn(2, c.data(), c.size());
         check(sb_db_finalize(handle));
      } catch (...) { sb_db_free(handle); handle = nullptr; throw; }
   }
   DeviceDatabase(const DeviceDatabase&) = delete;
   ~DeviceDatabase() { pool_.reset(); cache_.clear(); if (handle) sb_db_free(handle); }

   // ---- plan lowering -------------------------------------------------------------------------------
   struct Plan {
      std::vector<sb_node> nodes;
      std::deque<std::vector<uint32_t>> ranges;
      std::deque<std::vector<sb_predicate>> predicates;
      std::deque<std::string> host_bitmaps;
   };

   void lower(const operators::Operator& op, uint32_t local_partition, Plan& plan) const {
      using namespace operators;
      sb_node node{};
      node.row_count = op.row_count;
      switch (op.type()) {
         case EMPTY: node.op = SB_EMPTY; break;
         case FULL: node.op = SB_FULL; break;
         case INDEX_SCAN: {
            node.op = SB_INDEX_SCAN;
            if (const StoreRange* range = findStore(op.bitmap)) {
               const size_t slot = static_cast<size_t>(op.bitmap - range->begin);
               node.source = 1; node.store_id = range->store_id;
               node.position = static_cast<uint32_t>(slot / range->sigma); node.symbol = static_cast<uint32_t>(slot % range->sigma);
            } else if (auto it = bitmap_ids.find(op.bitmap); it != bitmap_ids.end()) {
               node.source = 0; node.bitmap_id = it->second;
            } else {   // a bitmap the pool does not hold (operator unit tests): ship it like a producer result
               pushHostBitmap(node, *op.bitmap, plan);
            }
            break;
         }
         case BITMAP_PRODUCER: pushHostBitmap(node, op.producer(), plan); break;
         case BITMAP_SELECTION: {
            node.op = SB_BITMAP_SELECTION;
            auto it = store_ids.find(op.store);
            if (it == store_ids.end()) throw DeviceError("BitmapSelection over a store that is not in the device pool");
            node.store_id = it->second; node.position = op.value; node.contains = op.contains ? 1 : 0;
            break;
         }
         case RANGE_SELECTION: {
            node.op = SB_RANGE_SELECTION;
            plan.ranges.emplace_back();
            for (const auto& r : op.ranges) { plan.ranges.back().push_back(r.start); plan.ranges.back().push_back(r.end); }
            node.ranges = plan.ranges.back().data(); node.n_ranges = static_cast<uint32_t>(op.ranges.size());
            break;
         }
         case COMPLEMENT: lower(*op.children[0], local_partition, plan); node.op = SB_COMPLEMENT; node.n_children = 1; break;
         case UNION:
            for (const auto& c : op.children) lower(*c, local_partition, plan);
            node.op = SB_UNION; node.n_children = static_cast<uint32_t>(op.children.size());
            break;
         case INTERSECTION:
         case THRESHOLD:
            for (const auto& c : op.children) lower(*c, local_partition, plan);
            for (const auto& c : op.negated_children) lower(*c, local_partition, plan);
            node.op = op.type() == INTERSECTION ? SB_INTERSECTION : SB_THRESHOLD;
            node.n_children = static_cast<uint32_t>(op.children.size()); node.n_negated = static_cast<uint32_t>(op.negated_children.size());
            node.number_of_matchers = op.number_of_matchers; node.match_exactly = op.match_exactly ? 1 : 0;
            break;
         case SELECTION: lowerSelection(op, local_partition, plan); return;
      }
      plan.nodes.push_back(node);
   }

   // ---- query execution (query_engine.cpp:30-68) ------------------------------------------------------
   struct EvaluatedFilters {
      sb_db* handle = nullptr;
      std::vector<uint32_t> partitions;
      std::vector<sb_result*> results;
      EvaluatedFilters() = default;
      EvaluatedFilters(const EvaluatedFilters&) = delete;
      EvaluatedFilters(EvaluatedFilters&& o) noexcept : handle(o.handle), partitions(std::move(o.partitions)), results(std::move(o.results)) { o.results.clear(); }
      EvaluatedFilters& operator=(EvaluatedFilters&& o) noexcept {
         if (this != &o) { for (auto* r : results) if (r) sb_result_free(r); handle = o.handle; partitions = std::move(o.partitions); results = std::move(o.results); o.results.clear(); }
         return *this;
      }
      ~EvaluatedFilters() { for (auto* r : results) if (r) sb_result_free(r); }
   };

   EvaluatedFilters evaluatePlans(const std::vector<Plan>& plans) const {
      std::vector<const sb_node*> plan_ptrs;
      std::vector<uint32_t> plan_sizes;
      EvaluatedFilters out;
      out.handle = handle;
      for (uint32_t local = 0; local < plans.size(); ++local) {
         plan_ptrs.push_back(plans[local].nodes.data());
         plan_sizes.push_back(static_cast<uint32_t>(plans[local].nodes.size()));
         out.partitions.push_back(local);
      }
      out.results.assign(plans.size(), nullptr);
      if (!plans.empty()) check(sb_filter_eval_batch(handle, static_cast<uint32_t>(plans.size()), out.partitions.data(), plan_ptrs.data(), plan_sizes.data(), out.results.data()));
      return out;
   }
   EvaluatedFilters evaluateTrees(const std::vector<const operators::Operator*>& trees) const {
      std::vector<Plan> plans(trees.size());
      for (uint32_t local = 0; local < trees.size(); ++local) lower(*trees[local], local, plans[local]);
      return evaluatePlans(plans);
   }

   EvaluatedFilters evaluateFilter(const filter::Expression& expression) const {
      // compile + lower per partition (query_engine.cpp:41-46 does it one partition after the other; the partitions are
      // independent, so with more than a few of them the worker pool shares the work); evaluation is one batched call
      const size_t n = owned.size();
      std::vector<operators::OperatorPtr> trees(n);
      std::vector<Plan> plans(n);
      HostTrace trace;
      auto one = [&](size_t local) {
         trees[local] = expression.compile(db, *db.partitions[owned[local]], filter::AmbiguityMode::NONE);
         lower(*trees[local], static_cast<uint32_t>(local), plans[local]);
      };
      if (n >= 8) workers().run(n, one);
      else for (size_t local = 0; local < n; ++local) one(local);
      trace.lap("compile+lower");
      auto out = evaluatePlans(plans);
      trace.lap("eval");
      return out;
   }

   std::vector<std::string> mutationSequenceNames(const Action& action) const {   // mutations.cpp:240-252
      const Alphabet& alphabet = alphabetById(action.alphabetId());
      const auto& references = db.getReferences(action.alphabetId());
      std::vector<std::string> names;
      for (const auto& name : action.sequence_names) {
         CHECK_SILO_QUERY(references.count(name), "Database does not contain the " + std::string(alphabet.name_lower) + " sequence with name: '" + name + "'")
         names.push_back(name);
      }
      if (action.sequence_names.empty()) for (const auto& [name, _] : references) names.push_back(name);
      return names;
   }

   // counts table of one sequence; false when every local filter is empty (no table, mutations.cpp:36-63)
   bool mutationCounts(const EvaluatedFilters& filters, const std::string& name, int alphabet_id, std::vector<uint32_t>& counts) const {
      const auto& reference = db.getReferences(alphabet_id).at(name);
      counts.assign(size_t(alphabetById(alphabet_id).count) * reference.size(), 0);
      int produced = 0;
      check(sb_mutations(handle, name.c_str(), static_cast<uint32_t>(filters.results.size()), filters.partitions.data(), filters.results.data(), counts.data(), counts.size(), &produced));
      return produced != 0;
   }

   // the rows of addMutationsToOutput selected on the device (mutations.cpp:196-215); false as mutationCounts
   bool mutationRows(const EvaluatedFilters& filters, const std::string& name, int alphabet_id, double min_proportion, std::vector<sb_mutation_row>& rows) const {
      const auto& valid = alphabetById(alphabet_id).valid_mutation_symbols;
      int produced = 0;
      size_t n = 0;
      rows.resize(std::max<size_t>(rows.size(), 4096));
      for (;;) {
         check(sb_mutations_rows(handle, name.c_str(), static_cast<uint32_t>(filters.results.size()), filters.partitions.data(), filters.results.data(), min_proportion,
                                 valid.data(), static_cast<uint32_t>(valid.size()), rows.data(), rows.size(), &n, &produced));
         if (n <= rows.size()) break;
         rows.resize(n);
      }
      rows.resize(n);
      return produced != 0;
   }
   // mutationRows for every sequence of the action (mutations.cpp:258-270 loops over them one after the other): the
   // sequences are independent, so more than one runs on the worker pool, each on a lane of the library.
   struct SequenceRows { bool produced = false; std::vector<sb_mutation_row> rows; };
   std::vector<SequenceRows> mutationRowsAll(const EvaluatedFilters& filters, const std::vector<std::string>& names, int alphabet_id, double min_proportion) const {
      std::vector<SequenceRows> out(names.size());
      auto one = [&](size_t k) { out[k].produced = mutationRows(filters, names[k], alphabet_id, min_proportion, out[k].rows); };
      if (names.size() <= 1) { for (size_t k = 0; k < names.size(); ++k) one(k); return out; }
      workers().run(names.size(), one);
      return out;
   }
   void mutationRowsFromTable(const std::string& name, int alphabet_id, const void* device_counts, double min_proportion, std::vector<sb_mutation_row>& rows) const {
      const auto& valid = alphabetById(alphabet_id).valid_mutation_symbols;
      size_t n = 0;
      rows.resize(std::max<size_t>(rows.size(), 4096));
      for (;;) {
         check(sb_mutation_rows_from_table(handle, name.c_str(), device_counts, min_proportion, valid.data(), static_cast<uint32_t>(valid.size()), rows.data(), rows.size(), &n));
         if (n <= rows.size()) break;
         rows.resize(n);
      }
      rows.resize(n);
   }
   // the formatting half of addMutationsToOutput (mutations.cpp:207-214)
   static void addRowsToOutput(const std::string& name, const Alphabet& alphabet, const std::vector<Symbol>& reference, const std::vector<sb_mutation_row>& rows, std::vector<QueryResultEntry>& output) {
      output.reserve(output.size() + rows.size());
      for (const auto& row : rows) {
         QueryResultEntry entry;
         entry.fields["mutation"] = alphabet.symbolToChar(reference[row.position]) + std::to_string(row.position + 1) + alphabet.symbolToChar(Symbol(row.symbol));
         entry.fields["sequenceName"] = name;
         entry.fields["proportion"] = static_cast<double>(row.count) / static_cast<double>(row.total);
         entry.fields["count"] = static_cast<int32_t>(row.count);
         output.push_back(std::move(entry));
      }
   }

   uint32_t aggregatedCount(const EvaluatedFilters& filters) const {
      uint32_t count = 0;
      check(sb_aggregated_count(handle, static_cast<uint32_t>(filters.results.size()), filters.results.data(), &count));
      return count;
   }

   // ---- Aggregated with groupByFields (aggregated.cpp:91-149) ------------------------------------------------------------
   // The group-by columns are coded and uploaded when a query first names them (one pass over the column on the host).
   const GroupDictionary& groupDictionary(const ColumnMetadata& meta) const {
      std::lock_guard<std::mutex> lock(group_mutex_);
      auto it = group_dictionaries_.find(meta.name);
      if (it != group_dictionaries_.end()) return *it->second;
      auto d = std::make_unique<GroupDictionary>(buildGroupDictionary(db, owned, meta));
      for (uint32_t local = 0; local < owned.size(); ++local) {
         uint32_t id = 0;
         check(sb_group_column_add(handle, local, d->codes[local].data(), static_cast<uint32_t>(d->codes[local].size()), &id));
         d->column_ids.push_back(id);
      }
      std::vector<std::vector<uint32_t>>().swap(d->codes);
      return *(group_dictionaries_[meta.name] = std::move(d));
   }
   // This process's share of an Aggregated query (with or without groupByFields): the entries for the partitions it owns, not
   // ordered, not limited -- what mergeAggregatedPartials (host/group_by.hpp) adds up over the processes of a sharded database.
   QueryResult executeAggregatedPartial(const Query& query) const {
      const Action& action = query.action;
      if (action.type != ActionType::AGGREGATED) throw QueryCompilationException("partial results exist for Aggregated actions (the Mutations tables of the ranks are summed on the devices)");
      (void)parseGroupByFields(db.config, action.group_by_fields);
      action.validateOrderByFields();
      if (!action.group_by_fields.empty()) return executeGroupedUnordered(query);
      CompiledFilter cf;
      compileFilter(*query.filter, cf);
      FusedResult fused;
      runFused(cf, {}, 0, 0, fused);
      QueryResultEntry entry;
      entry.fields["count"] = static_cast<int32_t>(fused.count());
      return QueryResult{{entry}};
   }
   QueryResult executeGrouped(const Query& query) const {
      const Action& action = query.action;
      (void)parseGroupByFields(db.config, action.group_by_fields);   // validateOrderByFields starts with it (aggregated.cpp:71-73)
      return action.executeAndOrder([&]() -> QueryResult { return executeGroupedUnordered(query); });
   }
   QueryResult executeGroupedUnordered(const Query& query) const {
      const Action& action = query.action;
      const std::vector<ColumnMetadata> fields = parseGroupByFields(db.config, action.group_by_fields);
      return [&]() -> QueryResult {
         if (owned.empty()) return {};
         CompiledFilter cf;
         compileFilter(*query.filter, cf);
         EvaluatedFilters filters = evaluatePlans(cf.plans);
         std::vector<const GroupDictionary*> dictionaries;
         std::vector<uint32_t> cardinalities, ids(owned.size() * fields.size());
         for (const auto& meta : fields) { dictionaries.push_back(&groupDictionary(meta)); cardinalities.push_back(dictionaries.back()->cardinality); }
         for (size_t local = 0; local < owned.size(); ++local) for (size_t c = 0; c < fields.size(); ++c) ids[local * fields.size() + c] = dictionaries[c]->column_ids[local];
         std::vector<sb_group_row> rows(4096);
         size_t n = 0;
         for (;;) {
            check(sb_aggregated_grouped(handle, static_cast<uint32_t>(filters.results.size()), filters.partitions.data(), filters.results.data(), static_cast<uint32_t>(fields.size()),
                                        ids.data(), cardinalities.data(), rows.data(), rows.size(), &n));
            if (n <= rows.size()) break;
            rows.resize(n);
         }
         return groupedResult(dictionaries, rows.data(), n);
      }();
   }

   // ---- one query = one submission (include/silo_b200.h: sb_query_mutations / sb_query_prepare) -----------------------
   // compile + lower per partition (query_engine.cpp:41-46 does it one partition after the other; the partitions are
   // independent, so with more than a few of them the worker pool shares the work)
   struct CompiledFilter {
      std::vector<operators::OperatorPtr> trees;
      std::vector<Plan> plans;
      std::vector<const sb_node*> plan_ptrs;
      std::vector<uint32_t> plan_sizes, partitions;
   };
   void compileFilter(const filter::Expression& expression, CompiledFilter& out) const {
      const size_t n = owned.size();
      out.trees.resize(n);
      out.plans.assign(n, Plan());
      auto one = [&](size_t local) {
         out.trees[local] = expression.compile(db, *db.partitions[owned[local]], filter::AmbiguityMode::NONE);
         lower(*out.trees[local], static_cast<uint32_t>(local), out.plans[local]);
      };
      if (n >= 8) workers().run(n, one);
      else for (size_t local = 0; local < n; ++local) one(local);
      out.plan_ptrs.clear(); out.plan_sizes.clear(); out.partitions.clear();
      for (uint32_t local = 0; local < n; ++local) {
         out.plan_ptrs.push_back(out.plans[local].nodes.data());
         out.plan_sizes.push_back(static_cast<uint32_t>(out.plans[local].nodes.size()));
         out.partitions.push_back(local);
      }
   }

   struct FusedResult {
      std::vector<SequenceRows> sequences;      // per requested sequence
      std::vector<uint64_t> cardinalities;      // per local partition
      uint32_t count() const { uint32_t c = 0; for (uint64_t v : cardinalities) c += static_cast<uint32_t>(v); return c; }   // aggregated.cpp:58-66 (uint32 accumulation)
   };
   // runs `call(out array, cardinalities)` until every sequence's rows fit
   template <typename Call>
   void runWithRowBuffers(size_t n_sequences, size_t n_partitions, FusedResult& result, Call&& call) const {
      result.sequences.resize(n_sequences);
      result.cardinalities.assign(n_partitions, 0);
      std::vector<sb_sequence_rows> out(n_sequences);
      for (auto& seq : result.sequences) seq.rows.resize(std::max<size_t>(seq.rows.size(), 1024));
      for (;;) {
         for (size_t k = 0; k < n_sequences; ++k) out[k] = sb_sequence_rows{result.sequences[k].rows.data(), result.sequences[k].rows.size(), 0, 0, 0};
         check(call(out.data(), result.cardinalities.data()));
         bool again = false;
         for (size_t k = 0; k < n_sequences; ++k) if (out[k].n_rows > result.sequences[k].rows.size()) { result.sequences[k].rows.resize(out[k].n_rows); again = true; }
         if (!again) break;
      }
      for (size_t k = 0; k < n_sequences; ++k) { result.sequences[k].rows.resize(out[k].n_rows); result.sequences[k].produced = out[k].produced != 0; }
   }
   static std::vector<const char*> cNames(const std::vector<std::string>& names) {
      std::vector<const char*> out;
      for (const auto& n : names) out.push_back(n.c_str());
      return out;
   }
   void runFused(const CompiledFilter& cf, const std::vector<std::string>& names, int alphabet_id, double min_proportion, FusedResult& result) const {
      const auto& valid = alphabetById(alphabet_id).valid_mutation_symbols;
      const std::vector<const char*> c_names = cNames(names);
      if (cf.partitions.empty()) { result.sequences.assign(names.size(), SequenceRows()); result.cardinalities.clear(); return; }
      runWithRowBuffers(names.size(), cf.partitions.size(), result, [&](sb_sequence_rows* out, uint64_t* cards) {
         return sb_query_mutations(handle, static_cast<uint32_t>(cf.partitions.size()), cf.partitions.data(), cf.plan_ptrs.data(), cf.plan_sizes.data(),
                                   static_cast<uint32_t>(names.size()), c_names.data(), min_proportion, valid.data(), static_cast<uint32_t>(valid.size()), out, cards);
      });
   }

   // A query kept on the device for replay (sb_query): what the plan cache holds and what silo_prepared_* hands out.
   struct PreparedQuery {
      Query query;
      std::vector<std::string> names;
      CompiledFilter compiled;
      sb_query* handle = nullptr;
      std::mutex mutex;
      PreparedQuery() = default;
      PreparedQuery(const PreparedQuery&) = delete;
      ~PreparedQuery() { if (handle) sb_query_free(handle); }
   };
   std::unique_ptr<PreparedQuery> prepare(const std::string& query_string, bool use_graph) const {
      auto p = std::make_unique<PreparedQuery>();
      p->query = parseQuery(query_string);
      p->query.action.validateOrderByFields();
      if (p->query.action.type == ActionType::AGGREGATED) {
         if (!p->query.action.group_by_fields.empty()) throw QueryCompilationException("Aggregated with groupByFields cannot be prepared");
      } else {
         p->names = mutationSequenceNames(p->query.action);
      }
      compileFilter(*p->query.filter, p->compiled);
      if (owned.empty()) return p;
      const int alphabet_id = p->query.action.type == ActionType::AGGREGATED ? 0 : p->query.action.alphabetId();
      const auto& valid = alphabetById(alphabet_id).valid_mutation_symbols;
      const std::vector<const char*> c_names = cNames(p->names);
      check(sb_query_prepare(handle, static_cast<uint32_t>(p->compiled.partitions.size()), p->compiled.partitions.data(), p->compiled.plan_ptrs.data(), p->compiled.plan_sizes.data(),
                             static_cast<uint32_t>(p->names.size()), c_names.data(), p->query.action.min_proportion, valid.data(), static_cast<uint32_t>(valid.size()),
                             use_graph ? 1 : 0, &p->handle));
      return p;
   }
   void runPrepared(PreparedQuery& p, FusedResult& result) const {
      if (!p.handle) { result.sequences.assign(p.names.size(), SequenceRows()); result.cardinalities.clear(); return; }
      runWithRowBuffers(p.names.size(), p.compiled.partitions.size(), result, [&](sb_sequence_rows* out, uint64_t* cards) { return sb_query_run(p.handle, out, cards); });
   }

   // Several prepared queries as one batch from this thread (sb_query_run_many): enqueued together, collected together. A query
   // whose rows did not fit its buffers is run again on its own.
   void runPreparedMany(const std::vector<PreparedQuery*>& queries, std::vector<FusedResult>& results) const {
      results.assign(queries.size(), FusedResult());
      std::vector<sb_query*> handles;
      std::vector<std::vector<sb_sequence_rows>> outs(queries.size());
      std::vector<sb_sequence_rows*> out_ptrs;
      std::vector<uint64_t*> card_ptrs;
      std::vector<size_t> batch;      // index into `queries` of every query that has a device handle
      for (size_t i = 0; i < queries.size(); ++i) {
         PreparedQuery& p = *queries[i];
         FusedResult& r = results[i];
         r.sequences.resize(p.names.size());
         r.cardinalities.assign(p.compiled.partitions.size(), 0);
         if (!p.handle) continue;
         for (auto& seq : r.sequences) seq.rows.resize(4096);
         for (size_t k = 0; k < p.names.size(); ++k) outs[i].push_back(sb_sequence_rows{r.sequences[k].rows.data(), r.sequences[k].rows.size(), 0, 0, 0});
         handles.push_back(p.handle); out_ptrs.push_back(outs[i].data()); card_ptrs.push_back(r.cardinalities.data()); batch.push_back(i);
      }
      if (!handles.empty()) check(sb_query_run_many(handles.data(), static_cast<uint32_t>(handles.size()), out_ptrs.data(), card_ptrs.data()));
      for (size_t i : batch) {
         bool fits = true;
         for (size_t k = 0; k < outs[i].size(); ++k) fits = fits && outs[i][k].n_rows <= results[i].sequences[k].rows.size();
         if (!fits) { runPrepared(*queries[i], results[i]); continue; }
         for (size_t k = 0; k < outs[i].size(); ++k) { results[i].sequences[k].rows.resize(outs[i][k].n_rows); results[i].sequences[k].produced = outs[i][k].produced != 0; }
      }
   }

   QueryResult resultFromRows(const Action& action, const std::vector<std::string>& names, const FusedResult& fused) const {
      return action.executeAndOrder([&]() -> QueryResult {
         if (action.type == ActionType::AGGREGATED) {
            QueryResultEntry entry;
            entry.fields["count"] = static_cast<int32_t>(fused.count());   // aggregated.cpp:58-66
            return QueryResult{{entry}};
         }
         const int alphabet_id = action.alphabetId();
         QueryResult result;
         for (size_t k = 0; k < names.size(); ++k) {      // mutations.cpp:258-270
            if (!fused.sequences[k].produced) continue;
            addRowsToOutput(names[k], alphabetById(alphabet_id), db.getReferences(alphabet_id).at(names[k]), fused.sequences[k].rows, result.query_result);
         }
         return result;
      });
   }

   QueryResult executeQuery(const std::string& query_string) const {
      Query query = parseQuery(query_string);
      if (query.action.type == ActionType::AGGREGATED && !query.action.group_by_fields.empty()) return executeGrouped(query);
      std::vector<std::string> names;
      if (query.action.type != ActionType::AGGREGATED) names = mutationSequenceNames(query.action);
      CompiledFilter cf;
      compileFilter(*query.filter, cf);
      FusedResult fused;
      runFused(cf, names, query.action.type == ActionType::AGGREGATED ? 0 : query.action.alphabetId(), query.action.min_proportion, fused);
      return resultFromRows(query.action, names, fused);
   }

   // Mutations rows -> the response text, straight from the selected rows (the same text as dumpJson(resultFromRows(...)):
   // keys in the order of the entry map, doubles through json::Value::writeDouble)
   std::string rowsToJson(const Action& action, const std::vector<std::string>& names, const FusedResult& fused) const {
      const int alphabet_id = action.alphabetId();
      const Alphabet& alphabet = alphabetById(alphabet_id);
      std::string out = "{\"queryResult\":[";
      bool first = true;
      char number[16];
      for (size_t k = 0; k < names.size(); ++k) {
         if (!fused.sequences[k].produced) continue;
         const std::string& name = names[k];
         const std::vector<sb_mutation_row>& rows = fused.sequences[k].rows;
         const auto& reference = db.getReferences(alphabet_id).at(name);
         std::string quoted_name;
         json::Value::writeString(quoted_name, name);
         out.reserve(out.size() + rows.size() * (80 + quoted_name.size()));
         for (const auto& row : rows) {
            if (!first) out += ',';
            first = false;
            out += "{\"count\":";
            out.append(number, std::to_chars(number, number + sizeof number, static_cast<int32_t>(row.count)).ptr);
            out += ",\"mutation\":\"";
            out += alphabet.symbolToChar(reference[row.position]);
            out.append(number, std::to_chars(number, number + sizeof number, row.position + 1).ptr);
            out += alphabet.symbolToChar(Symbol(row.symbol));
            out += "\",\"proportion\":";
            json::Value::writeDouble(out, static_cast<double>(row.count) / static_cast<double>(row.total));
            out += ",\"sequenceName\":";
            out += quoted_name;
            out += '}';
         }
      }
      out += "]}";
      return out;
   }
   std::string responseText(const Action& action, const std::vector<std::string>& names, const FusedResult& fused) const {
      if (action.type == ActionType::AGGREGATED || !action.order_by_fields.empty() || action.limit || action.offset) return dumpJson(resultFromRows(action, names, fused));
      return rowsToJson(action, names, fused);
   }

   // The response body {"queryResult": [...]} of a query. A query text seen before is answered from the plan cache: the
   // compiled + lowered plans stay on the device as a prepared query (from the second sighting on, as a CUDA graph).
   std::string executeQueryJson(const std::string& query_string) const {
      HostTrace trace;
      if (plan_cache_enabled_) {
         std::shared_ptr<PreparedQuery> hit = cacheLookup(query_string);
         if (hit) {
            trace.lap("cache");
            FusedResult fused;
            { std::lock_guard<std::mutex> lock(hit->mutex); runPrepared(*hit, fused); }
            trace.lap("run");
            return responseText(hit->query.action, hit->names, fused);
         }
      }
      Query query = parseQuery(query_string);
      const Action& action = query.action;
      trace.lap("parse");
      if (action.type == ActionType::AGGREGATED && !action.group_by_fields.empty()) return dumpJson(executeGrouped(query));
      action.validateOrderByFields();
      std::vector<std::string> names;
      if (action.type != ActionType::AGGREGATED) names = mutationSequenceNames(action);
      CompiledFilter cf;
      compileFilter(*query.filter, cf);
      trace.lap("compile+lower");
      FusedResult fused;
      runFused(cf, names, action.type == ActionType::AGGREGATED ? 0 : action.alphabetId(), action.min_proportion, fused);
      trace.lap("run");
      return responseText(action, names, fused);
   }
   void setPlanCache(bool enabled) { std::lock_guard<std::mutex> lock(cache_mutex_); plan_cache_enabled_ = enabled; cache_.clear(); cache_order_.clear(); }

  private:
   // Plan cache: query text -> prepared query. The first sighting of a text only counts it; the second prepares it (compile,
   // lower, upload, locate, graph capture), later ones replay. A busy entry (another thread is running it) is a miss.
   struct CacheEntry { uint32_t sightings = 0; std::shared_ptr<PreparedQuery> prepared; };
   static constexpr size_t PLAN_CACHE_ENTRIES = 64;
   mutable std::mutex cache_mutex_;
   mutable std::unordered_map<std::string, CacheEntry> cache_;
   mutable std::deque<std::string> cache_order_;
   mutable bool plan_cache_enabled_ = getenv("SILO_PLAN_CACHE") == nullptr || std::string(getenv("SILO_PLAN_CACHE")) != "0";
   std::shared_ptr<PreparedQuery> cacheLookup(const std::string& text) const {
      bool build = false;
      {
         std::lock_guard<std::mutex> lock(cache_mutex_);
         auto it = cache_.find(text);
         if (it == cache_.end()) {
            if (cache_.size() >= PLAN_CACHE_ENTRIES) { cache_.erase(cache_order_.front()); cache_order_.pop_front(); }
            cache_[text].sightings = 1;
            cache_order_.push_back(text);
            return nullptr;
         }
         if (it->second.prepared) {
            std::unique_lock<std::mutex> busy(it->second.prepared->mutex, std::try_to_lock);
            return busy.owns_lock() ? it->second.prepared : nullptr;
         }
         build = ++it->second.sightings == 2;
      }
      if (!build) return nullptr;
      std::shared_ptr<PreparedQuery> made;
      try { made = prepare(text, /*use_graph=*/true); } catch (...) { return nullptr; }   // e.g. a query the one-shot path rejects with its own message
      std::lock_guard<std::mutex> lock(cache_mutex_);
      auto it = cache_.find(text);
      if (it != cache_.end()) it->second.prepared = made;
      return made;
   }

   WorkerPool& workers() const {
      std::lock_guard<std::mutex> lock(pool_mutex_);
      if (!pool_) pool_ = std::make_unique<WorkerPool>(6);
      return *pool_;
   }
   mutable std::mutex pool_mutex_;
   mutable std::unique_ptr<WorkerPool> pool_;
   mutable std::mutex group_mutex_;
   mutable std::map<std::string, std::unique_ptr<GroupDictionary>> group_dictionaries_;

   struct StoreRange { const Bitmap* begin; const Bitmap* end; uint32_t store_id; uint32_t sigma; };
   std::vector<StoreRange> store_ranges;
   std::unordered_map<const SequenceStorePartition*, uint32_t> store_ids;
   std::unordered_map<const Bitmap*, uint32_t> bitmap_ids;
   std::unordered_map<const void*, uint32_t> column_ids;

   const StoreRange* findStore(const Bitmap* bm) const {
      auto it = std::upper_bound(store_ranges.begin(), store_ranges.end(), bm, [](const Bitmap* b, const StoreRange& r) { return b < r.begin; });
      if (it == store_ranges.begin()) return nullptr;
      --it;
      return bm < it->end ? &*it : nullptr;
   }

   static void pushHostBitmap(sb_node& node, const Bitmap& bm, Plan& plan) {
      node.op = SB_BITMAP_PRODUCER;
      plan.host_bitmaps.emplace_back(bm.isEmpty() ? std::string() : bm.toPortable());
      node.host_bitmap = plan.host_bitmaps.back().data();
      node.host_bitmap_len = plan.host_bitmaps.back().size();
   }

   // Selection (selection.cpp:92-109): numeric / date predicates run on the GPU; SiloString predicates
   // are evaluated on the host and shipped as a bitmap; the parts are intersected.
   void lowerSelection(const operators::Operator& op, uint32_t local_partition, Plan& plan) const {
      using namespace operators;
      uint32_t parts = 0;
      if (!op.children.empty()) { lower(*op.children[0], local_partition, plan); ++parts; }
      std::vector<sb_predicate> device_preds;
      std::vector<const Predicate*> host_preds;
      for (const auto& p : op.predicates) {
         if (p.kind == ColumnKind::STRING) { host_preds.push_back(&p); continue; }
         sb_predicate d{};
         const void* key = p.kind == ColumnKind::INT32 ? static_cast<const void*>(static_cast<const std::vector<int32_t>*>(p.column)->data())
                         : p.kind == ColumnKind::FLOAT64 ? static_cast<const void*>(static_cast<const std::vector<double>*>(p.column)->data())
                                                         : static_cast<const void*>(static_cast<const std::vector<uint32_t>*>(p.column)->data());
         auto it = column_ids.find(key);
         if (it == column_ids.end() && op.row_count == 0) continue;   // empty partition: nothing to compare
         if (it == column_ids.end()) throw DeviceError("Selection over a column that is not in the device pool");
         d.column_id = it->second;
         d.comparator = static_cast<int32_t>(p.comparator);
         d.i32 = p.int_value; d.u32 = p.date_value; d.f64 = p.float_value;
         device_preds.push_back(d);
      }
      if (!device_preds.empty() || (host_preds.empty() && op.children.empty())) {
         sb_node node{};
         node.op = SB_SELECTION; node.row_count = op.row_count;
         plan.predicates.push_back(std::move(device_preds));
         node.predicates = plan.predicates.back().data(); node.n_predicates = static_cast<uint32_t>(plan.predicates.back().size());
         plan.nodes.push_back(node);
         ++parts;
      }
      if (!host_preds.empty()) {
         std::vector<uint32_t> rows;
         for (uint32_t row = 0; row < op.row_count; ++row) {
            bool ok = true;
            for (const auto* p : host_preds) if (!p->match(row)) { ok = false; break; }
            if (ok) rows.push_back(row);
         }
         sb_node node{};
         node.row_count = op.row_count;
         pushHostBitmap(node, Bitmap::fromSorted(rows.data(), rows.size()), plan);
         plan.nodes.push_back(node);
         ++parts;
      }
      if (parts > 1) {
         sb_node node{};
         node.op = SB_INTERSECTION; node.row_count = op.row_count; node.n_children = parts;
         plan.nodes.push_back(node);
      }
   }

   void ingestStore(uint32_t local, const std::string& name, const SequenceStorePartition& store) {
      PortableBlob blob, missing;
      blob.offsets.reserve(store.bitmaps.size() + 1);
      for (const auto& bm : store.bitmaps) blob.add(bm);
      missing.offsets.reserve(store.missing_symbol_bitmaps.size() + 1);
      for (const auto& bm : store.missing_symbol_bitmaps) missing.add(bm);
      uint32_t store_id = 0;
      check(sb_store_add(handle, local, name.c_str(), store.alphabet->id, store.length(), static_cast<uint32_t>(store.missing_symbol_bitmaps.size()),
                         reinterpret_cast<const uint8_t*>(blob.bytes.data()), blob.offsets.data(), store.flipped.data(), store.deleted.data(),
                         reinterpret_cast<const uint8_t*>(missing.bytes.data()), missing.offsets.data(), &store_id));
      store_ids[&store] = store_id;
      if (!store.bitmaps.empty()) store_ranges.push_back({store.bitmaps.data(), store.bitmaps.data() + store.bitmaps.size(), store_id, store.alphabet->count});
   }

   void ingestIndex(uint32_t local, const std::unordered_map<uint32_t, Bitmap>& index) {
      for (const auto& [id, bm] : index) {
         const std::string bytes = bm.isEmpty() ? std::string() : bm.toPortable();
         uint32_t bitmap_id = 0;
         check(sb_bitmap_add_portable(handle, local, bytes.data(), bytes.size(), &bitmap_id));
         bitmap_ids[&bm] = bitmap_id;
      }
   }

   void ingestPartition(uint32_t local, const DatabasePartition& part) {
      for (const auto& [name, store] : part.nuc_sequences) ingestStore(local, name, store);
      for (const auto& [name, store] : part.aa_sequences) ingestStore(local, name, store);
      for (const auto& [name, col] : part.columns.pango_lineage_columns) { ingestIndex(local, col.indexed_values); ingestIndex(local, col.indexed_sublineage_values); }
      for (const auto& [name, col] : part.columns.indexed_string_columns) ingestIndex(local, col.indexed_values);
      auto addColumn = [&](int dtype, const void* data, size_t n) {
         uint32_t id = 0;
         check(sb_column_add(handle, local, dtype, data, static_cast<uint32_t>(n), &id));
         column_ids[data] = id;
      };
      for (const auto& [name, col] : part.columns.int_columns) if (!col.values.empty()) addColumn(0, col.values.data(), col.values.size());
      for (const auto& [name, col] : part.columns.float_columns) if (!col.values.empty()) addColumn(1, col.values.data(), col.values.size());
      for (const auto& [name, col] : part.columns.date_columns) if (!col.values.empty()) addColumn(2, col.values.data(), col.values.size());
   }
};

}  // namespace silo::engine
