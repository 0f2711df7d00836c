// Second half of synth.hpp: turning the row model into a silo::Database.
#pragma once

namespace silo::synth {

inline Bitmap bitmapFromIntervals(const std::vector<Interval>& intervals) {
   Bitmap bm;
   for (const auto& i : intervals) bm.addRange(i.start, i.end);
   bm.runOptimize();
   return bm;
}

// Index of one (partition, sequence model): transposes the rows' events into per-(pos, sym) bitmaps and
// applies the most-numerous-symbol deletion (position.cpp:41-67,102-127).
inline void buildStore(SequenceStorePartition& store, const SequenceModel& model, uint64_t seed, const std::vector<RowMeta>& rows,
                       const std::vector<std::vector<Interval>>& nuc_missing, const std::vector<std::pair<uint32_t, uint32_t>>& gap_ends,
                       std::vector<std::optional<std::string>>* explicit_out) {
   const Alphabet& alphabet = *model.alphabet;
   const uint32_t L = static_cast<uint32_t>(model.reference.size()), sigma = alphabet.count, n = static_cast<uint32_t>(rows.size());
   store.sequence_count = n;
   store.missing_symbol_bitmaps.assign(n, Bitmap());
   std::vector<uint64_t> events;   // (slot << 32) | row, rows ascending
   events.reserve(size_t(n) * (model.is_nucleotide ? 80 : 4));
   std::vector<uint32_t> slot_count(size_t(L) * sigma + 1, 0);
   std::vector<int32_t> missing_diff(L + 1, 0);
   std::vector<std::vector<Interval>> own_missing;
   if (!model.is_nucleotide) own_missing.resize(n);
   std::vector<Event> row_events, scratch;
   for (uint32_t row = 0; row < n; ++row) {
      const std::vector<Interval>* missing = &nuc_missing[row];
      if (!model.is_nucleotide) { own_missing[row] = geneMissing(model, nuc_missing[row]); missing = &own_missing[row]; }
      for (const auto& i : *missing) { missing_diff[i.start] += 1; missing_diff[i.end] -= 1; }
      if (!missing->empty()) store.missing_symbol_bitmaps[row] = bitmapFromIntervals(*missing);
      rowEvents(model, seed, rows[row], *missing, model.is_nucleotide ? gap_ends[row].first : 0, model.is_nucleotide ? gap_ends[row].second : 0, row_events, scratch);
      for (const auto& e : row_events) {
         const uint64_t slot = uint64_t(e.pos) * sigma + e.sym;
         events.push_back((slot << 32) | row);
         ++slot_count[slot];
      }
      if (explicit_out) {
         const bool null_genome = missing->size() == 1 && (*missing)[0].start == 0 && (*missing)[0].end == L && model.is_nucleotide && (rows[row].id % 2 == 0);
         if (null_genome) { explicit_out->push_back(std::nullopt); continue; }   // half of the fully-missing rows become NULL genomes
         std::string text(L, '?');
         for (uint32_t p = 0; p < L; ++p) text[p] = alphabet.symbolToChar(model.reference[p]);
         for (const auto& i : *missing) for (uint32_t p = i.start; p < i.end; ++p) text[p] = alphabet.symbolToChar(alphabet.missing);
         for (const auto& e : row_events) text[e.pos] = alphabet.symbolToChar(e.sym);
         explicit_out->push_back(std::move(text));
      }
   }
   // stable counting sort by slot
   std::vector<uint32_t> slot_begin(slot_count.size(), 0);
   for (size_t s = 1; s < slot_count.size(); ++s) slot_begin[s] = slot_begin[s - 1] + slot_count[s - 1];
   std::vector<uint32_t> sorted_rows(events.size());
   {
      std::vector<uint32_t> cursor(slot_begin.begin(), slot_begin.end());
      for (uint64_t e : events) sorted_rows[cursor[e >> 32]++] = static_cast<uint32_t>(e);
   }
   std::vector<uint64_t>().swap(events);
   // per position: most numerous symbol, in SYMBOLS order with strict > (position.cpp:55-66)
   std::vector<uint32_t> special_positions;   // the deleted symbol is not the reference symbol: its rows must be stored
   int32_t running_missing = 0;
   std::vector<uint32_t> missing_at(L);
   for (uint32_t pos = 0; pos < L; ++pos) {
      running_missing += missing_diff[pos];
      missing_at[pos] = static_cast<uint32_t>(running_missing);
      const Symbol ref = model.reference[pos];
      uint32_t others = 0;
      for (uint32_t s = 0; s < sigma; ++s) others += slot_count[size_t(pos) * sigma + s];
      const uint32_t ref_count = ref == alphabet.missing ? 0 : n - others - missing_at[pos];
      int best = -1;
      uint32_t best_count = 0;
      for (Symbol s : alphabet.iteration_order) {
         const uint32_t c = s == ref ? ref_count : slot_count[size_t(pos) * sigma + s];
         if (c > best_count) { best = s; best_count = c; }
      }
      store.deleted[pos] = static_cast<int8_t>(best);
      if (best >= 0 && static_cast<Symbol>(best) != ref && ref != alphabet.missing && ref_count > 0) special_positions.push_back(pos);
   }
   for (uint32_t pos = 0; pos < L; ++pos) {
      for (uint32_t s = 0; s < sigma; ++s) {
         const size_t slot = size_t(pos) * sigma + s;
         if (slot_count[slot] == 0 || store.deleted[pos] == static_cast<int8_t>(s)) continue;
         Bitmap bm = Bitmap::fromSorted(sorted_rows.data() + slot_begin[slot], slot_count[slot]);
         bm.runOptimize();
         store.bitmaps[slot] = std::move(bm);
      }
   }
   // positions where an alternative symbol won: rows with the reference symbol = all - events - missing
   for (uint32_t pos : special_positions) {
      std::vector<uint8_t> taken(n, 0);
      for (uint32_t s = 0; s < sigma; ++s) {
         const size_t slot = size_t(pos) * sigma + s;
         for (uint32_t k = 0; k < slot_count[slot]; ++k) taken[sorted_rows[slot_begin[slot] + k]] = 1;
      }
      for (uint32_t row = 0; row < n; ++row) {
         const auto& missing = model.is_nucleotide ? nuc_missing[row] : own_missing[row];
         for (const auto& i : missing) if (i.start <= pos && pos < i.end) { taken[row] = 1; break; }
      }
      std::vector<uint32_t> ref_rows;
      for (uint32_t row = 0; row < n; ++row) if (!taken[row]) ref_rows.push_back(row);
      Bitmap bm = Bitmap::fromSorted(ref_rows.data(), ref_rows.size());
      bm.runOptimize();
      store.bitmaps[size_t(pos) * sigma + model.reference[pos]] = std::move(bm);
   }
}

struct Generated {
   std::unique_ptr<Database> db;
   // explicit_sequences: per partition, per sequence name, the rows as text (nullopt = NULL genome)
   std::vector<std::map<std::string, std::vector<std::optional<std::string>>>> explicit_sequences;
   std::vector<std::string> lineage_names;
   std::vector<uint32_t> owned;   // partitions built on this rank (all when world == 1)
   std::vector<std::pair<uint32_t, uint8_t>> clade_mutations;   // (0-based pos, alt) defining substitutions of mid-level lineages (BASELINE config C3)
};

inline Generated generate(const Options& options, const std::map<std::string, std::string>& nuc_refs, const std::map<std::string, std::string>& aa_refs) {
   Generated out;
   out.db = std::make_unique<Database>();
   Database& db = *out.db;
   db.config.instance_name = "synthetic";
   db.config.primary_key = "key";
   db.config.date_to_sort_by = "date";
   db.config.partition_by = "pangoLineage";
   db.config.metadata = {{"date", ColumnType::DATE}, {"pangoLineage", ColumnType::INDEXED_PANGOLINEAGE}, {"region", ColumnType::INDEXED_STRING},
                         {"age", ColumnType::INT}, {"qc", ColumnType::FLOAT}};
   const auto lineages = makeLineages();
   for (const auto& l : lineages) out.lineage_names.push_back(l.name);

   // sequence models
   std::vector<SequenceModel> models;
   {
      uint32_t nuc_offset = 250;
      for (const auto& [name, text] : nuc_refs) {
         std::string t = options.genome_length && options.genome_length < text.size() ? text.substr(0, options.genome_length) : text;
         db.nuc_reference[name] = builder::toSymbols(nucleotideAlphabet(), t, "reference " + name);
         if (name != db.config.default_nucleotide_sequence && nuc_refs.count(db.config.default_nucleotide_sequence)) continue;
         SequenceModel m;
         m.name = name; m.alphabet = &nucleotideAlphabet(); m.reference = db.nuc_reference[name];
         m.substitution_symbols = {nuc::A, nuc::C, nuc::G, nuc::T};
         models.push_back(std::move(m));
      }
      if (options.with_amino_acids) {
         for (const auto& [name, text] : aa_refs) {
            db.aa_reference[name] = builder::toSymbols(aminoAcidAlphabet(), text, "reference " + name);
            SequenceModel m;
            m.name = name; m.alphabet = &aminoAcidAlphabet(); m.reference = db.aa_reference[name]; m.is_nucleotide = false;
            for (Symbol s = aa::A; s <= aa::Y; ++s) m.substitution_symbols.push_back(s);
            m.nuc_offset = nuc_offset;
            nuc_offset += 3 * static_cast<uint32_t>(text.size());
            models.push_back(std::move(m));
         }
      }
   }
   if (models.empty() || !models[0].is_nucleotide) throw std::runtime_error("synthetic generator needs one nucleotide reference sequence");
   const uint32_t L = static_cast<uint32_t>(models[0].reference.size());
   for (auto& model : models) {
      std::vector<std::vector<Event>> node_events;
      for (uint32_t node = 0; node < lineages.size(); ++node) node_events.push_back(nodeEvents(model, options.seed, node));
      for (uint32_t node = 0; node < lineages.size(); ++node) {
         std::vector<int> chain;
         for (int cur = static_cast<int>(node); cur >= 0; cur = lineages[cur].parent) chain.push_back(cur);
         std::vector<Event> merged;
         for (auto it = chain.rbegin(); it != chain.rend(); ++it) merged = overlay(merged, node_events[*it]);
         model.lineage_events.push_back(std::move(merged));
      }
      if (model.is_nucleotide) {
         for (uint32_t node = 1; node <= 19; ++node) for (const auto& e : node_events[node]) if (e.sym != nuc::GAP) out.clade_mutations.push_back({e.pos, e.sym});
      }
   }

   // lineage sizes: Zipf(1) over a seeded permutation, largest share capped at 1/16
   std::vector<uint64_t> weight(lineages.size());
   {
      std::vector<uint32_t> perm(lineages.size());
      for (uint32_t k = 0; k < perm.size(); ++k) perm[k] = k;
      Rng rng(streamSeed(options.seed, 0x5000));
      for (uint32_t k = static_cast<uint32_t>(perm.size()) - 1; k > 0; --k) std::swap(perm[k], perm[rng.below(k + 1)]);
      uint64_t total = 0;
      for (uint32_t r = 0; r < perm.size(); ++r) { weight[perm[r]] = 1000000 / (r + 1); total += weight[perm[r]]; }
      for (int round = 0; round < 8; ++round) {
         uint64_t new_total = 0;
         for (auto& w : weight) { w = std::min<uint64_t>(w, total / 16); new_total += w; }
         total = new_total;
      }
   }
   std::vector<uint64_t> cumulative(weight.size());
   for (size_t k = 0; k < weight.size(); ++k) cumulative[k] = weight[k] + (k ? cumulative[k - 1] : 0);

   // rows -> partitions by the reference rule: sorted keys, extend a group while its count <= N/32
   const uint64_t N = options.n_sequences;
   std::vector<RowMeta> metas(N);
   const unsigned n_threads = options.n_threads ? options.n_threads : std::max(1u, std::thread::hardware_concurrency());
   {
      std::vector<std::thread> threads;
      for (unsigned t = 0; t < n_threads; ++t) threads.emplace_back([&, t] { for (uint64_t id = t; id < N; id += n_threads) metas[id] = rowMeta(options.seed, id, cumulative); });
      for (auto& th : threads) th.join();
   }
   std::vector<uint32_t> order_of_lineage(lineages.size());   // rank of the lineage name in ascending string order
   {
      std::vector<uint32_t> idx(lineages.size());
      for (uint32_t k = 0; k < idx.size(); ++k) idx[k] = k;
      std::sort(idx.begin(), idx.end(), [&](uint32_t a, uint32_t b) { return lineages[a].name < lineages[b].name; });
      for (uint32_t r = 0; r < idx.size(); ++r) order_of_lineage[idx[r]] = r;
   }
   std::vector<uint64_t> lineage_count(lineages.size(), 0);
   for (const auto& m : metas) ++lineage_count[order_of_lineage[m.lineage]];
   std::vector<uint32_t> partition_of_rank(lineages.size(), 0);
   uint32_t n_partitions = 0;
   {
      const double allowed = static_cast<double>(N) / 32.0;
      uint64_t running = 0;
      bool first = true;
      for (uint32_t r = 0; r < lineages.size(); ++r) {
         if (lineage_count[r] == 0) continue;
         if (first || static_cast<double>(running) > allowed) { ++n_partitions; running = 0; first = false; }
         partition_of_rank[r] = n_partitions - 1;
         running += lineage_count[r];
      }
   }
   std::vector<std::vector<RowMeta>> partition_rows(n_partitions);
   for (const auto& m : metas) partition_rows[partition_of_rank[order_of_lineage[m.lineage]]].push_back(m);
   std::vector<RowMeta>().swap(metas);
   if (options.row_slices > 1) {      // row order inside a lineage partition (date, id) first, then consecutive slices become partitions of their own
      {
         std::atomic<uint32_t> next{0};
         std::vector<std::thread> sorters;
         for (unsigned t = 0; t < std::min<unsigned>(n_threads, n_partitions); ++t) sorters.emplace_back([&] {
            for (uint32_t p; (p = next.fetch_add(1)) < n_partitions;) {
               std::sort(partition_rows[p].begin(), partition_rows[p].end(), [](const RowMeta& a, const RowMeta& b) { return a.date != b.date ? a.date < b.date : a.id < b.id; });
            }
         });
         for (auto& th : sorters) th.join();
      }
      std::vector<std::vector<RowMeta>> pieces;
      for (auto& rows : partition_rows) {
         const uint64_t n = rows.size();
         for (uint32_t k = 0; k < options.row_slices; ++k) pieces.emplace_back(rows.begin() + n * k / options.row_slices, rows.begin() + n * (k + 1) / options.row_slices);
         std::vector<RowMeta>().swap(rows);
      }
      partition_rows.swap(pieces);
      n_partitions = static_cast<uint32_t>(partition_rows.size());
   }

   // dictionaries are shared by all partitions: fill them before the parallel phase
   static const char* const REGIONS[6] = {"Africa", "Asia", "Europe", "North America", "Oceania", "South America"};
   db.dictionaries["pangoLineage"] = std::make_unique<Dictionary<std::string>>();
   db.dictionaries["region"] = std::make_unique<Dictionary<std::string>>();
   for (const auto& l : lineages) for (const auto& p : PangoLineageColumnPartition::parentLineages(l.name)) db.dictionaries["pangoLineage"]->getOrCreateId(p);
   for (const char* r : REGIONS) db.dictionaries["region"]->getOrCreateId(r);

   db.partitions.resize(n_partitions);
   // partition -> rank: longest-processing-time-first on row counts, deterministic on every rank
   std::vector<uint32_t> rank_of_partition(n_partitions, 0);
   {
      std::vector<uint32_t> by_size(n_partitions);
      for (uint32_t p = 0; p < n_partitions; ++p) by_size[p] = p;
      std::stable_sort(by_size.begin(), by_size.end(), [&](uint32_t a, uint32_t b) { return partition_rows[a].size() > partition_rows[b].size(); });
      std::vector<uint64_t> load(std::max<uint32_t>(options.world, 1), 0);
      for (uint32_t p : by_size) {
         uint32_t r = static_cast<uint32_t>(std::min_element(load.begin(), load.end()) - load.begin());
         if (options.row_slices > 1 && options.row_slices % std::max<uint32_t>(options.world, 1) == 0) r = (p % options.row_slices) % options.world;   // slice k of every lineage -> rank k mod world
         rank_of_partition[p] = r;
         load[r] += partition_rows[p].size();
      }
      for (uint32_t p = 0; p < n_partitions; ++p) {
         if (rank_of_partition[p] == options.rank) { out.owned.push_back(p); continue; }
         db.partitions[p] = std::make_unique<DatabasePartition>();   // shell: row count only
         db.partitions[p]->sequence_count = static_cast<uint32_t>(partition_rows[p].size());
         std::vector<RowMeta>().swap(partition_rows[p]);
      }
   }
   out.explicit_sequences.resize(options.explicit_sequences ? n_partitions : 0);
   std::atomic<uint32_t> next_partition{0};
   std::vector<std::string> errors(n_threads);
   auto worker = [&](unsigned t) {
      try {
         while (true) {
            const uint32_t p = next_partition.fetch_add(1);
            if (p >= n_partitions) break;
            if (rank_of_partition[p] != options.rank) continue;
            auto& rows = partition_rows[p];
            std::sort(rows.begin(), rows.end(), [](const RowMeta& a, const RowMeta& b) { return a.date != b.date ? a.date < b.date : a.id < b.id; });   // NULL date (0) first
            auto part = std::make_unique<DatabasePartition>();
            const uint32_t n = static_cast<uint32_t>(rows.size());
            part->sequence_count = n;
            part->chunks.push_back({0, n});
            auto& date = part->columns.date_columns["date"];
            date.is_sorted = true;
            auto& lineage_col = part->columns.pango_lineage_columns["pangoLineage"];
            lineage_col.alias = &db.alias; lineage_col.lookup_unaliased = db.dictionaries["pangoLineage"].get();
            auto& region = part->columns.indexed_string_columns["region"];
            region.lookup = db.dictionaries["region"].get();
            auto& age = part->columns.int_columns["age"];
            auto& qc = part->columns.float_columns["qc"];
            std::map<uint32_t, std::vector<uint32_t>> rows_of_lineage, rows_of_prefix, rows_of_region;
            for (uint32_t row = 0; row < n; ++row) {
               const RowMeta& m = rows[row];
               date.values.push_back(m.date);
               age.values.push_back(m.age);
               qc.values.push_back(m.qc);
               const uint32_t lineage_id = *lineage_col.lookup_unaliased->getId(lineages[m.lineage].name);
               lineage_col.value_ids.push_back(lineage_id);
               rows_of_lineage[lineage_id].push_back(row);
               const uint32_t region_id = *region.lookup->getId(REGIONS[m.region]);
               region.value_ids.push_back(region_id);
               rows_of_region[region_id].push_back(row);
            }
            for (const auto& [lineage_id, ids] : rows_of_lineage) {   // pango_lineage_column.cpp:36-55
               lineage_col.indexed_values[lineage_id] = Bitmap::fromSorted(ids.data(), ids.size());
               for (const auto& prefix : PangoLineageColumnPartition::parentLineages(lineage_col.lookup_unaliased->values[lineage_id])) {
                  auto& v = rows_of_prefix[*lineage_col.lookup_unaliased->getId(prefix)];
                  v.insert(v.end(), ids.begin(), ids.end());
               }
            }
            for (auto& [prefix_id, ids] : rows_of_prefix) { std::sort(ids.begin(), ids.end()); lineage_col.indexed_sublineage_values[prefix_id] = Bitmap::fromSorted(ids.data(), ids.size()); }
            for (const auto& [region_id, ids] : rows_of_region) region.indexed_values[region_id] = Bitmap::fromSorted(ids.data(), ids.size());

            std::vector<std::vector<Interval>> nuc_missing(n);
            std::vector<std::pair<uint32_t, uint32_t>> gap_ends(n);
            for (uint32_t row = 0; row < n; ++row) { bool full; nuc_missing[row] = rowMissing(options.seed, rows[row].id, L, full, gap_ends[row].first, gap_ends[row].second); }
            for (const auto& model : models) {
               auto& stores = model.is_nucleotide ? part->nuc_sequences : part->aa_sequences;
               SequenceStorePartition& store = stores[model.name];
               store.init(model.alphabet, model.is_nucleotide ? &db.nuc_reference.at(model.name) : &db.aa_reference.at(model.name));
               buildStore(store, model, options.seed, rows, nuc_missing, gap_ends, options.explicit_sequences ? &out.explicit_sequences[p][model.name] : nullptr);
            }
            db.partitions[p] = std::move(part);
            std::vector<RowMeta>().swap(rows);
         }
      } catch (const std::exception& e) { errors[t] = e.what(); }
   };
   std::vector<std::thread> threads;
   for (unsigned t = 0; t < std::min<unsigned>(n_threads, n_partitions); ++t) threads.emplace_back(worker, t);
   for (auto& th : threads) th.join();
   for (const auto& e : errors) if (!e.empty()) throw std::runtime_error(e);
   return out;
}

}  // namespace silo::synth
