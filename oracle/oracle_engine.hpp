// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle_roaring.hpp). CPU restatement of the reference's
// query hot path: Operator::evaluate for all 11 operators, Mutations<SymbolType>::execute,
// Aggregated (count) and QueryEngine::executeQuery. Each function cites the reference lines it
// follows. Parity: PINNED — against the operator known-answer tests ported from
// src/silo/query_engine/operators/*.test.cpp, the 63 hot-path e2e fixtures of
// endToEndTests/test/queries/*.json on testBaseData/exampleDataset, and the size / container
// goldens of src/silo/database.test.cpp:74-111 (tests/test_oracle_*.py).
#pragma once
#include <atomic>
#include <thread>

#include "../lapis-silo_b200/host/query.hpp"
#include "oracle_roaring.hpp"

namespace oracle {

using namespace silo;
using silo::operators::Operator;

// include/silo/query_engine/operator_result.h:9-32: owned-or-borrowed bitmap, copy on first write
struct OperatorResult {
   const Bitmap* borrowed = nullptr;
   Bitmap owned;
   bool is_owned = true;
   OperatorResult() = default;
   explicit OperatorResult(const Bitmap* b) : borrowed(b), is_owned(false) {}
   explicit OperatorResult(Bitmap&& b) : owned(std::move(b)), is_owned(true) {}
   const Bitmap& get() const { return is_owned ? owned : *borrowed; }
   Bitmap& mut() {  // operator_result.cpp:40-54
      if (!is_owned) { owned = *borrowed; is_owned = true; borrowed = nullptr; }
      return owned;
   }
   bool isMutable() const { return is_owned; }
};

OperatorResult evaluate(const Operator& op);

inline OperatorResult evaluateIntersection(const Operator& op) {  // intersection.cpp:76-127
   std::vector<OperatorResult> children, negated;
   for (const auto& c : op.children) children.push_back(evaluate(*c));
   for (const auto& c : op.negated_children) negated.push_back(evaluate(*c));
   std::sort(children.begin(), children.end(), [](const OperatorResult& a, const OperatorResult& b) { return a.get().cardinality() < b.get().cardinality(); });
   std::sort(negated.begin(), negated.end(), [](const OperatorResult& a, const OperatorResult& b) { return a.get().cardinality() > b.get().cardinality(); });
   if (children.size() == 1) {
      OperatorResult result = std::move(children[0]);
      for (auto& n : negated) result.mut() = bitmapAndNot(result.get(), n.get());
      return result;
   }
   OperatorResult result(bitmapAnd(children[0].get(), children[1].get()));  // intersectTwo (:62-74)
   for (size_t i = 2; i < children.size(); ++i) result.owned = bitmapAnd(result.owned, children[i].get());
   for (auto& n : negated) result.owned = bitmapAndNot(result.owned, n.get());
   return result;
}

inline OperatorResult evaluateThreshold(const Operator& op) {  // threshold.cpp:64-138
   const uint32_t dp_table_size = op.match_exactly ? op.number_of_matchers + 1 : op.number_of_matchers;
   std::vector<Bitmap> T(dp_table_size);
   if (op.children.empty()) {
      T[0] = evaluate(*op.negated_children[0]).get();
      T[0].flip(0, op.row_count);
   } else {
      T[0] = evaluate(*op.children[0]).get();
   }
   const int max_table_index = static_cast<int>(dp_table_size - 1);
   const int non_negated_child_count = static_cast<int>(op.children.size());
   const int negated_child_count = static_cast<int>(op.negated_children.size());
   const int n = static_cast<int>(op.number_of_matchers);
   const int k = non_negated_child_count + negated_child_count;
   for (int i = 1; i < non_negated_child_count; ++i) {
      auto bitmap = evaluate(*op.children[i]);
      for (int j = std::min(max_table_index, i); j > std::max(0, n - k + i - 1); --j) {
         T[j] = bitmapOr(T[j], bitmapAnd(T[j - 1], bitmap.get()));
      }
      if (k - i > n - 1) T[0] = bitmapOr(T[0], bitmap.get());
   }
   const int took_first_offset = op.children.empty() ? 1 : 0;
   for (int local_i = took_first_offset; local_i < negated_child_count; ++local_i) {
      auto bitmap = evaluate(*op.negated_children[local_i]);
      const int i = local_i + non_negated_child_count;
      for (int j = std::min(max_table_index, i); j > std::max(0, n - k + i - 1); --j) {
         T[j] = bitmapOr(T[j], bitmapAndNot(T[j - 1], bitmap.get()));
      }
      if (k - i > n - 1) {
         bitmap.mut().flip(0, op.row_count);
         T[0] = bitmapOr(T[0], bitmap.get());
      }
   }
   if (op.match_exactly) {
      T[n - 1] = bitmapAndNot(T[n - 1], T[n]);
      return OperatorResult(std::move(T[n - 1]));
   }
   return OperatorResult(std::move(T.back()));
}

inline OperatorResult evaluate(const Operator& op) {
   switch (op.type()) {
      case operators::EMPTY: return OperatorResult();                         // empty.cpp:24-26
      case operators::FULL: {                                                 // full.cpp:24-28
         OperatorResult r;
         r.owned.addRange(0, op.row_count);
         return r;
      }
      case operators::INDEX_SCAN: return OperatorResult(op.bitmap);           // index_scan.cpp:28-30
      case operators::INTERSECTION: return evaluateIntersection(op);
      case operators::COMPLEMENT: {                                           // complement.cpp:50-54
         auto r = evaluate(*op.children[0]);
         r.mut().flip(0, op.row_count);
         return r;
      }
      case operators::RANGE_SELECTION: {                                      // range_selection.cpp:42-48
         OperatorResult r;
         for (const auto& range : op.ranges) r.owned.addRange(range.start, range.end);
         return r;
      }
      case operators::SELECTION: {                                            // selection.cpp:92-109
         OperatorResult r;
         auto matches = [&](uint32_t row) { for (const auto& p : op.predicates) if (!p.match(row)) return false; return true; };
         if (!op.children.empty()) {
            auto child = evaluate(*op.children[0]);
            child.get().forEach([&](uint32_t row) { if (matches(row)) r.owned.add(row); });
         } else {
            for (uint32_t row = 0; row < op.row_count; ++row) if (matches(row)) r.owned.add(row);
         }
         return r;
      }
      case operators::BITMAP_SELECTION: {                                     // bitmap_selection.cpp:33-52
         OperatorResult r;
         const auto& bitmaps = op.store->missing_symbol_bitmaps;
         for (uint32_t i = 0; i < op.row_count; ++i) if (bitmaps[i].contains(op.value) == op.contains) r.owned.add(i);
         return r;
      }
      case operators::THRESHOLD: return evaluateThreshold(op);
      case operators::UNION: {                                                // union.cpp:35-45
         std::vector<OperatorResult> results;
         for (const auto& c : op.children) results.push_back(evaluate(*c));
         std::vector<const Bitmap*> ptrs;
         for (const auto& r : results) ptrs.push_back(&r.get());
         return OperatorResult(fastUnion(ptrs));
      }
      case operators::BITMAP_PRODUCER: return OperatorResult(op.producer());  // bitmap_producer.cpp:25-27
   }
   throw std::runtime_error("unknown operator type");
}

// ---- Mutations (actions/mutations.cpp) ---------------------------------------------------------

struct FilterAndStore { const Bitmap* filter; const SequenceStorePartition* store; };
struct PrefilteredBitmaps { std::vector<FilterAndStore> bitmaps, full_bitmaps; };

// mutations.cpp:36-63
inline std::map<std::string, PrefilteredBitmaps> preFilterBitmaps(const Database& db, int alphabet_id, std::vector<OperatorResult>& filters) {
   std::map<std::string, PrefilteredBitmaps> out;
   for (size_t i = 0; i < db.partitions.size(); ++i) {
      const DatabasePartition& part = *db.partitions[i];
      OperatorResult& filter = filters[i];
      const uint64_t cardinality = filter.get().cardinality();
      if (cardinality == 0) continue;
      if (cardinality == part.sequence_count) {
         for (const auto& [name, store] : part.getSequenceStores(alphabet_id)) out[name].full_bitmaps.push_back({&filter.get(), &store});
      } else {
         if (filter.isMutable()) filter.owned.runOptimize();
         for (const auto& [name, store] : part.getSequenceStores(alphabet_id)) out[name].bitmaps.push_back({&filter.get(), &store});
      }
   }
   return out;
}

// mutations.cpp:66-98 (faithful: O(|F|) contains() calls per position for the deleted symbol)
inline void addPositionForMixedBitmaps(uint32_t pos, const PrefilteredBitmaps& pre, uint32_t* counts, size_t length, const Alphabet& alphabet) {
   for (const auto& [filter, store] : pre.bitmaps) {
      for (Symbol symbol : alphabet.iteration_order) {
         if (store->isSymbolDeleted(pos, symbol)) {
            counts[symbol * length + pos] += static_cast<uint32_t>(filter->cardinality());
            filter->forEach([&](uint32_t idx) { if (store->missing_symbol_bitmaps[idx].contains(pos)) counts[symbol * length + pos] -= 1; });
            continue;
         }
         const uint32_t symbol_count = static_cast<uint32_t>(
            store->isSymbolFlipped(pos, symbol) ? andnotCardinality(*filter, *store->getBitmap(pos, symbol)) : andCardinality(*filter, *store->getBitmap(pos, symbol)));
         counts[symbol * length + pos] += symbol_count;
         const int8_t deleted = store->deleted[pos];
         if (deleted >= 0 && symbol != static_cast<Symbol>(deleted)) counts[size_t(deleted) * length + pos] -= symbol_count;
      }
   }
}
// mutations.cpp:101-136
inline void addPositionForFullBitmaps(uint32_t pos, const PrefilteredBitmaps& pre, uint32_t* counts, size_t length, const Alphabet& alphabet) {
   for (const auto& [filter, store] : pre.full_bitmaps) {
      for (Symbol symbol : alphabet.iteration_order) {
         if (store->isSymbolDeleted(pos, symbol)) {
            counts[symbol * length + pos] += store->sequence_count;
            for (const Bitmap& n_bitmap : store->missing_symbol_bitmaps) if (n_bitmap.contains(pos)) counts[symbol * length + pos] -= 1;
            continue;
         }
         const uint32_t card = static_cast<uint32_t>(store->getBitmap(pos, symbol)->cardinality());
         const uint32_t symbol_count = store->isSymbolFlipped(pos, symbol) ? store->sequence_count - card : card;
         counts[symbol * length + pos] += symbol_count;
         const int8_t deleted = store->deleted[pos];
         if (deleted >= 0 && symbol != static_cast<Symbol>(deleted)) counts[size_t(deleted) * length + pos] -= card;
      }
   }
}

// "fast mode": same sums, but the per-position missing counts of the deleted symbol come from a
// difference array over the missing runs of the filtered rows instead of
// |F| x L contains() calls. Must equal the faithful mode (tests/test_oracle_synthetic.py).
inline void addStoreFast(const FilterAndStore& fs, bool full, uint32_t* counts, size_t length, const Alphabet& alphabet) {
   const Bitmap& filter = *fs.filter;
   const SequenceStorePartition& store = *fs.store;
   std::vector<int32_t> diff(length + 1, 0);
   auto addRow = [&](uint32_t idx) {
      for (const auto& c : store.missing_symbol_bitmaps[idx].containers) {
         const uint32_t base = static_cast<uint32_t>(c.key) << 16;
         if (c.kind == RUN) {
            for (uint32_t r = 0; r < c.runCount(); ++r) { diff[base + c.data[2 * r]] += 1; diff[base + c.data[2 * r] + c.data[2 * r + 1] + 1] -= 1; }
         } else c.forEach([&](uint32_t v) { diff[v] += 1; diff[v + 1] -= 1; });
      }
   };
   uint32_t filter_card;
   if (full) { filter_card = store.sequence_count; for (uint32_t idx = 0; idx < store.sequence_count; ++idx) addRow(idx); }
   else { filter_card = static_cast<uint32_t>(filter.cardinality()); filter.forEach(addRow); }
   int32_t running = 0;
   for (uint32_t pos = 0; pos < length; ++pos) {
      running += diff[pos];
      for (Symbol symbol : alphabet.iteration_order) {
         if (store.isSymbolDeleted(pos, symbol)) { counts[symbol * length + pos] += filter_card - static_cast<uint32_t>(running); continue; }
         const Bitmap& bm = *store.getBitmap(pos, symbol);
         if (bm.isEmpty() && !store.isSymbolFlipped(pos, symbol)) continue;
         const uint32_t card = static_cast<uint32_t>(full ? bm.cardinality() : andCardinality(filter, bm));
         const uint32_t symbol_count = store.isSymbolFlipped(pos, symbol) ? filter_card - card : card;
         counts[symbol * length + pos] += symbol_count;
         const int8_t deleted = store.deleted[pos];
         if (deleted >= 0 && symbol != static_cast<Symbol>(deleted)) counts[size_t(deleted) * length + pos] -= full ? card : symbol_count;
      }
   }
}

// mutations.cpp:139-164: positions split in chunks of 300 over a thread pool (TBB in the reference)
// `position_limit` (bounded CPU samples of bench.py): faithful mode walks positions [0, limit) only; 0 = all
inline std::vector<uint32_t> calculateMutationsPerPosition(const Alphabet& alphabet, size_t length, const PrefilteredBitmaps& pre, bool faithful, unsigned n_threads,
                                                           uint32_t position_limit = 0) {
   std::vector<uint32_t> counts(size_t(alphabet.count) * length, 0);
   const uint32_t walk = position_limit ? std::min<uint32_t>(position_limit, static_cast<uint32_t>(length)) : static_cast<uint32_t>(length);
   if (n_threads == 0) n_threads = std::max(1u, std::thread::hardware_concurrency());
   if (faithful) {
      constexpr uint32_t POSITIONS_PER_PROCESS = 300;
      std::atomic<uint32_t> next{0};
      auto worker = [&] {
         while (true) {
            const uint32_t begin = next.fetch_add(POSITIONS_PER_PROCESS);
            if (begin >= walk) break;
            const uint32_t end = std::min<uint32_t>(begin + POSITIONS_PER_PROCESS, walk);
            for (uint32_t pos = begin; pos < end; ++pos) {
               addPositionForMixedBitmaps(pos, pre, counts.data(), length, alphabet);
               addPositionForFullBitmaps(pos, pre, counts.data(), length, alphabet);
            }
         }
      };
      std::vector<std::thread> threads;
      for (unsigned t = 1; t < n_threads; ++t) threads.emplace_back(worker);
      worker();
      for (auto& t : threads) t.join();
   } else {
      // fast mode: one private table per worker, stores distributed over workers, summed at the end
      std::vector<std::pair<FilterAndStore, bool>> work;
      for (const auto& fs : pre.bitmaps) work.push_back({fs, false});
      for (const auto& fs : pre.full_bitmaps) work.push_back({fs, true});
      std::atomic<size_t> next{0};
      std::vector<std::vector<uint32_t>> partial(std::min<size_t>(n_threads, std::max<size_t>(1, work.size())));
      auto worker = [&](size_t t) {
         partial[t].assign(counts.size(), 0);
         while (true) {
            const size_t w = next.fetch_add(1);
            if (w >= work.size()) break;
            addStoreFast(work[w].first, work[w].second, partial[t].data(), length, alphabet);
         }
      };
      std::vector<std::thread> threads;
      for (size_t t = 1; t < partial.size(); ++t) threads.emplace_back(worker, t);
      worker(0);
      for (auto& t : threads) t.join();
      for (const auto& p : partial) for (size_t k = 0; k < counts.size(); ++k) counts[k] += p[k];
   }
   return counts;
}

// mutations.cpp:185-232
inline void addMutationsToOutput(const std::string& sequence_name, const Alphabet& alphabet, const std::vector<Symbol>& reference_sequence,
                                 const std::vector<uint32_t>& counts, double min_proportion, std::vector<QueryResultEntry>& output) {
   const size_t sequence_length = reference_sequence.size();
   for (size_t pos = 0; pos < sequence_length; ++pos) {
      uint32_t total = 0;
      for (const Symbol symbol : alphabet.valid_mutation_symbols) total += counts[symbol * sequence_length + pos];
      if (total == 0) continue;
      const auto threshold_count = min_proportion == 0 ? 0 : static_cast<uint32_t>(std::ceil(static_cast<double>(total) * min_proportion) - 1);
      const Symbol symbol_in_reference_genome = reference_sequence.at(pos);
      for (const Symbol symbol : alphabet.valid_mutation_symbols) {
         if (symbol_in_reference_genome != symbol) {
            const uint32_t count = counts[symbol * sequence_length + pos];
            if (count > threshold_count) {
               const double proportion = static_cast<double>(count) / static_cast<double>(total);
               QueryResultEntry entry;
               entry.fields["mutation"] = std::string(1, alphabet.symbolToChar(symbol_in_reference_genome)) + std::to_string(pos + 1) + alphabet.symbolToChar(symbol);
               entry.fields["sequenceName"] = sequence_name;
               entry.fields["proportion"] = proportion;
               entry.fields["count"] = static_cast<int32_t>(count);
               output.push_back(std::move(entry));
            }
         }
      }
   }
}

struct Options {
   bool faithful = true;
   unsigned threads = 0;
   std::vector<uint32_t> partitions;   // bounded samples: evaluate only these partitions (empty = all), the others count as empty filters
   uint32_t position_limit = 0;        // bounded samples: faithful Mutations walks positions [0, limit) only (0 = all)
};

// query_engine.cpp:40-49: serial compile + evaluate per partition
inline std::vector<OperatorResult> evaluateFilters(const Database& db, const filter::Expression& filter_expression,
                                                   std::vector<operators::OperatorPtr>* keep_alive, const std::vector<uint32_t>& subset = {}) {
   std::vector<OperatorResult> partition_filters;
   for (size_t index = 0; index < db.partitions.size(); ++index) {
      const auto& part = db.partitions[index];
      if (!subset.empty() && std::find(subset.begin(), subset.end(), index) == subset.end()) { partition_filters.emplace_back(); continue; }
      auto op = filter_expression.compile(db, *part, filter::AmbiguityMode::NONE);
      partition_filters.push_back(evaluate(*op));
      keep_alive->push_back(std::move(op));  // BitmapSelection / IndexScan results may borrow from the tree's sources
   }
   return partition_filters;
}

// mutations.cpp:235-272; returns per sequence name the full [Σ][L] table through `tables` if given
inline QueryResult executeMutations(const Database& db, const Action& action, std::vector<OperatorResult>& filters, const Options& options,
                                    std::map<std::string, std::vector<uint32_t>>* tables = nullptr) {
   const int alphabet_id = action.alphabetId();
   const Alphabet& alphabet = alphabetById(alphabet_id);
   const auto& references = db.getReferences(alphabet_id);
   std::vector<std::string> names;
   for (const auto& name : action.sequence_names) {
      CHECK_SILO_QUERY(references.count(name), "Database does not contain the " + std::string(alphabet.name_lower) + " sequence with name: '" + name + "'")
      names.push_back(name);
   }
   if (action.sequence_names.empty()) for (const auto& [name, _] : references) names.push_back(name);
   auto pre = preFilterBitmaps(db, alphabet_id, filters);
   QueryResult result;
   for (const auto& name : names) {
      const auto& reference = references.at(name);
      auto it = pre.find(name);
      if (it == pre.end()) continue;
      auto counts = calculateMutationsPerPosition(alphabet, reference.size(), it->second, options.faithful, options.threads, options.position_limit);
      addMutationsToOutput(name, alphabet, reference, counts, action.min_proportion, result.query_result);
      if (tables) (*tables)[name] = std::move(counts);
   }
   return result;
}

// aggregated.cpp:58-66
inline QueryResult aggregateWithoutGrouping(const std::vector<OperatorResult>& filters) {
   uint32_t count = 0;
   for (const auto& f : filters) count += static_cast<uint32_t>(f.get().cardinality());
   QueryResultEntry entry;
   entry.fields["count"] = static_cast<int32_t>(count);
   return QueryResult{{entry}};
}

// ---- Aggregated with groupByFields: aggregated.cpp:24-40 (parseGroupByFields), :91-149 (execute), tuple.cpp:29-155 ----
// A tuple is the byte string of the row's group-by values exactly as assignTupleField lays them out: Date / int32 / double
// raw, the dictionary id for indexed and lineage columns. The reference's String<N> (inline prefix + dictionary id of the
// rest) is one-to-one with the text, which stands for it here behind a length.
struct TupleLayout {
   std::vector<ColumnMetadata> fields;
};
inline std::vector<ColumnMetadata> groupByMetadata(const Database& db, const std::vector<std::string>& group_by_fields) {   // aggregated.cpp:24-40
   std::vector<ColumnMetadata> out;
   for (const auto& field : group_by_fields) {
      bool found = false;
      for (const auto& m : db.config.metadata) if (m.name == field) { out.push_back(m); found = true; break; }
      CHECK_SILO_QUERY(found, "Metadata field '" + field + "' to group by not found")
   }
   return out;
}
inline void appendTupleField(std::string& tuple, const ColumnMetadata& m, const ColumnPartitionGroup& columns, uint32_t row) {   // tuple.cpp:29-76
   auto raw = [&](const void* p, size_t n) { tuple.append(static_cast<const char*>(p), n); };
   auto text = [&](const std::string& v) { const uint32_t n = static_cast<uint32_t>(v.size()); raw(&n, 4); tuple += v; };
   switch (m.type) {
      case ColumnType::DATE: { const uint32_t v = columns.date_columns.at(m.name).values[row]; raw(&v, 4); break; }
      case ColumnType::INT: { const int32_t v = columns.int_columns.at(m.name).values[row]; raw(&v, 4); break; }
      case ColumnType::FLOAT: { const double v = columns.float_columns.at(m.name).values[row]; raw(&v, 8); break; }
      case ColumnType::STRING: text(columns.string_columns.at(m.name).values[row]); break;
      case ColumnType::INDEXED_STRING: { const uint32_t v = columns.indexed_string_columns.at(m.name).value_ids[row]; raw(&v, 4); break; }
      case ColumnType::INDEXED_PANGOLINEAGE: { const uint32_t v = columns.pango_lineage_columns.at(m.name).value_ids[row]; raw(&v, 4); break; }
      case ColumnType::NUC_INSERTION: text(columns.nuc_insertion_columns.at(m.name).values[row]); break;
      case ColumnType::AA_INSERTION: text(columns.aa_insertion_columns.at(m.name).values[row]); break;
   }
}
inline std::map<std::string, FieldValue> tupleFields(const std::string& tuple, const std::vector<ColumnMetadata>& fields, const ColumnPartitionGroup& columns) {   // tuple.cpp:78-155
   std::map<std::string, FieldValue> out;
   size_t at = 0;
   auto take = [&](void* p, size_t n) { memcpy(p, tuple.data() + at, n); at += n; };
   auto text = [&]() { uint32_t n; take(&n, 4); std::string v = tuple.substr(at, n); at += n; return v; };
   auto orNull = [](std::string v) -> FieldValue { if (v.empty()) return std::nullopt; return v; };
   for (const auto& m : fields) {
      switch (m.type) {
         case ColumnType::DATE: {      // common/date.cpp:66-80
            uint32_t v; take(&v, 4);
            if (v == 0) { out[m.name] = std::nullopt; break; }
            char buffer[32];
            snprintf(buffer, sizeof buffer, "%04u-%02u-%02u", v >> 16, (v >> 12) & 0xF, v & 0xFFF);
            out[m.name] = std::string(buffer);
            break;
         }
         case ColumnType::INT: { int32_t v; take(&v, 4); if (v == INT32_MIN) out[m.name] = std::nullopt; else out[m.name] = v; break; }
         case ColumnType::FLOAT: { double v; take(&v, 8); if (std::isnan(v)) out[m.name] = std::nullopt; else out[m.name] = v; break; }
         case ColumnType::STRING: case ColumnType::NUC_INSERTION: case ColumnType::AA_INSERTION: out[m.name] = orNull(text()); break;
         case ColumnType::INDEXED_STRING: { uint32_t v; take(&v, 4); out[m.name] = orNull(columns.indexed_string_columns.at(m.name).lookup->values.at(v)); break; }
         case ColumnType::INDEXED_PANGOLINEAGE: { uint32_t v; take(&v, 4); out[m.name] = orNull(columns.pango_lineage_columns.at(m.name).lookupAliasedValue(v)); break; }
      }
   }
   return out;
}
inline QueryResult aggregateGrouped(const Database& db, const std::vector<std::string>& group_by_fields, const std::vector<OperatorResult>& filters) {   // aggregated.cpp:91-149
   const std::vector<ColumnMetadata> fields = groupByMetadata(db, group_by_fields);
   std::vector<std::unordered_map<std::string, uint32_t>> tuple_maps(db.partitions.size());
   for (size_t p = 0; p < db.partitions.size(); ++p) {      // :111-137, one map per partition
      const ColumnPartitionGroup& columns = db.partitions[p]->columns;
      std::string tuple;
      filters[p].get().forEach([&](uint32_t row) {
         tuple.clear();
         for (const auto& m : fields) appendTupleField(tuple, m, columns, row);
         ++tuple_maps[p][tuple];
      });
   }
   std::unordered_map<std::string, uint32_t> final_map;      // :138-148
   const ColumnPartitionGroup* any_columns = nullptr;
   for (size_t p = 0; p < db.partitions.size(); ++p) {
      for (const auto& [tuple, count] : tuple_maps[p]) final_map[tuple] += count;
      if (!tuple_maps[p].empty()) any_columns = &db.partitions[p]->columns;
   }
   QueryResult result;      // generateResult, :46-56
   for (const auto& [tuple, count] : final_map) {
      QueryResultEntry entry;
      entry.fields = tupleFields(tuple, fields, *any_columns);
      entry.fields["count"] = static_cast<int32_t>(count);
      result.query_result.push_back(std::move(entry));
   }
   // the reference hands the rows out in the order of its hash map; a fixed order (every field, then the count) keeps runs comparable
   std::sort(result.query_result.begin(), result.query_result.end(), [](const QueryResultEntry& a, const QueryResultEntry& b) { return a.fields < b.fields; });
   return result;
}

// query_engine.cpp:30-68
inline QueryResult executeQuery(const Database& db, const std::string& query_string, const Options& options,
                                std::map<std::string, std::vector<uint32_t>>* tables = nullptr) {
   Query query = parseQuery(query_string);
   std::vector<operators::OperatorPtr> trees;
   std::vector<OperatorResult> filters = evaluateFilters(db, *query.filter, &trees, options.partitions);
   if (query.action.type == ActionType::AGGREGATED) (void)groupByMetadata(db, query.action.group_by_fields);   // validateOrderByFields starts with it (aggregated.cpp:71-73)
   return query.action.executeAndOrder([&]() -> QueryResult {
      if (query.action.type == ActionType::AGGREGATED) {
         if (!query.action.group_by_fields.empty()) return aggregateGrouped(db, query.action.group_by_fields, filters);
         return aggregateWithoutGrouping(filters);
      }
      return executeMutations(db, query.action, filters, options, tables);
   });
}

}  // namespace oracle
