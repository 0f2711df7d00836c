// Mini-builder: dataset directory (database_config.yaml + preprocessing_config.yaml + metadata
// TSV / NDJSON + FASTA files + reference_genomes.json + pangolineage_alias.json) -> Database.
// Restates the parts of the reference's offline preprocessing that define ROW IDS and INDEX
// CONTENT, without DuckDB:
//   * partitioning rule          src/silo/preprocessing/preprocessor.cpp:159-199
//   * row order in a partition   src/silo/config/database_config.cpp:190-198 (NULLS FIRST:
//                                src/silo/preprocessing/preprocessing_database.cpp:41)
//   * column fill / null values  src/silo/storage/column/*.cpp
//   * sequence index fill        src/silo/storage/sequence_store.cpp:101-211, position.cpp:24-127
#pragma once
#include <functional>
#include <algorithm>
#include <thread>

#include "io_util.hpp"
#include "json.hpp"
#include "storage.hpp"

namespace silo::builder {

struct InputRecord {
   std::map<std::string, std::optional<std::string>> metadata;     // nullopt = NULL
   std::map<std::string, std::optional<std::string>> nuc_sequences;
   std::map<std::string, std::optional<std::string>> aa_sequences;
};

inline std::vector<Symbol> toSymbols(const Alphabet& alphabet, const std::string& text, const std::string& what) {
   std::vector<Symbol> out(text.size());
   for (size_t k = 0; k < text.size(); ++k) {
      auto s = alphabet.charToSymbol(text[k]);
      if (!s) throw std::runtime_error("Illegal character " + std::to_string(text[k]) + " contained in " + what);
      out[k] = *s;
   }
   return out;
}

// SequenceStorePartition::fill + optimizeBitmaps, restated without the flipped intermediate state:
// stored[s] = rows carrying s (s != missing); then the most numerous symbol (first in SYMBOLS
// order on ties, position.cpp:41-63) is emptied and recorded as deleted (position.cpp:102-127).
// A null genome contributes to no symbol and is fully missing (sequence_store.cpp:163-170).
inline void fillSequenceStore(SequenceStorePartition& store, const std::vector<const std::optional<std::string>*>& genomes) {
   const Alphabet& alphabet = *store.alphabet;
   const uint32_t length = store.length();
   const uint32_t n = static_cast<uint32_t>(genomes.size());
   store.sequence_count = n;
   store.missing_symbol_bitmaps.assign(n, Bitmap());
   int8_t char_to_symbol[256];
   for (int c = 0; c < 256; ++c) { auto s = alphabet.charToSymbol(static_cast<char>(c)); char_to_symbol[c] = s ? static_cast<int8_t>(*s) : -1; }
   for (uint32_t row = 0; row < n; ++row) {
      if (genomes[row]->has_value() && (*genomes[row])->size() != length) {
         throw std::runtime_error("sequence length " + std::to_string((*genomes[row])->size()) + " does not match the reference length " + std::to_string(length));
      }
   }
   const unsigned n_threads = std::max(1u, std::min(16u, std::thread::hardware_concurrency()));
   std::vector<std::thread> threads;
   std::vector<std::string> errors(n_threads);
   for (unsigned t = 0; t < n_threads; ++t) {
      threads.emplace_back([&, t] {
         try {
            std::vector<std::vector<uint32_t>> ids(alphabet.count);
            for (uint32_t pos = t; pos < length; pos += n_threads) {
               for (auto& v : ids) v.clear();
               for (uint32_t row = 0; row < n; ++row) {
                  if (!genomes[row]->has_value()) continue;
                  const int8_t s = char_to_symbol[static_cast<unsigned char>((**genomes[row])[pos])];
                  if (s < 0) throw std::runtime_error("Illegal character contained in sequence.");
                  if (static_cast<Symbol>(s) != alphabet.missing) ids[s].push_back(row);
               }
               int best = -1;
               size_t best_count = 0;
               for (Symbol s : alphabet.iteration_order) if (ids[s].size() > best_count) { best = s; best_count = ids[s].size(); }
               for (Symbol s = 0; s < alphabet.count; ++s) {
                  if (static_cast<int>(s) == best || ids[s].empty()) continue;
                  Bitmap bm = Bitmap::fromSorted(ids[s].data(), ids[s].size());
                  bm.runOptimize();
                  store.bitmapRef(pos, s) = std::move(bm);
               }
               store.deleted[pos] = static_cast<int8_t>(best);
            }
            // missing-symbol bitmaps, one per row (sequence_store.cpp:153-190)
            std::vector<uint32_t> missing;
            for (uint32_t row = t; row < n; row += n_threads) {
               Bitmap& bm = store.missing_symbol_bitmaps[row];
               if (!genomes[row]->has_value()) { bm.addRange(0, length); bm.runOptimize(); continue; }
               missing.clear();
               const std::string& genome = **genomes[row];
               for (uint32_t pos = 0; pos < length; ++pos) if (char_to_symbol[static_cast<unsigned char>(genome[pos])] == static_cast<int8_t>(alphabet.missing)) missing.push_back(pos);
               if (!missing.empty()) { bm = Bitmap::fromSorted(missing.data(), missing.size()); bm.runOptimize(); }
            }
         } catch (const std::exception& e) { errors[t] = e.what(); }
      });
   }
   for (auto& th : threads) th.join();
   for (const auto& e : errors) if (!e.empty()) throw std::runtime_error(e);
}

inline ColumnType columnTypeFromConfig(const std::string& type, bool generate_index, const std::string& name) {  // database_config.cpp:158-188
   if (type == "string") return generate_index ? ColumnType::INDEXED_STRING : ColumnType::STRING;
   if (type == "date") return ColumnType::DATE;
   if (type == "pango_lineage") return ColumnType::INDEXED_PANGOLINEAGE;
   if (type == "int") return ColumnType::INT;
   if (type == "float") return ColumnType::FLOAT;
   if (type == "insertion") return ColumnType::NUC_INSERTION;
   if (type == "aaInsertion") return ColumnType::AA_INSERTION;
   throw std::runtime_error("unknown metadata type '" + type + "' of column " + name);
}

inline DatabaseConfig parseDatabaseConfig(const std::string& text) {  // config/database_config.cpp:60-130
   DatabaseConfig config;
   std::string cur_name, cur_type;
   bool cur_index = false, cur_index_set = false, in_item = false, has_metadata = false;
   auto flush = [&] {
      if (!in_item) return;
      bool gen = cur_index_set ? cur_index : cur_type == "pango_lineage";
      config.metadata.push_back({cur_name, columnTypeFromConfig(cur_type, gen, cur_name)});
      in_item = false; cur_index_set = false; cur_index = false;
   };
   bool in_metadata = false;      // list items belong to the key above them: only those under `metadata` are columns
   for (const auto& y : io::parseYamlLines(text)) {
      if (!y.list_item && y.indent < 4) { flush(); in_metadata = y.key == "metadata"; }
      if (!in_metadata && (y.list_item || y.indent >= 4)) continue;
      if (y.list_item) { flush(); in_item = true; }
      if (in_item && y.indent >= 4) {
         if (y.key == "name") cur_name = y.value;
         else if (y.key == "type") cur_type = y.value;
         else if (y.key == "generateIndex") { cur_index = y.value == "true"; cur_index_set = true; }
         continue;
      }
      flush();
      if (y.key == "metadata") has_metadata = true;
      if (y.key == "instanceName") config.instance_name = y.value;
      else if (y.key == "primaryKey") config.primary_key = y.value;
      else if (y.key == "dateToSortBy") config.date_to_sort_by = y.value;
      else if (y.key == "partitionBy") config.partition_by = y.value;
      else if (y.key == "defaultNucleotideSequence") config.default_nucleotide_sequence = y.value;
   }
   flush();
   // yaml-cpp's text when schema["metadata"] is dereferenced and absent (database_config.test.cpp:149-160)
   if (!has_metadata) throw std::runtime_error("invalid node; first invalid key: \"metadata\"");
   return config;
}

// preprocessor.cpp:159-199: keys sorted ascending with NULL first; a group is extended while its
// running count is <= sum/32 (double division), otherwise the next key starts a new group.
inline std::vector<std::vector<size_t>> partitionRecords(const std::vector<InputRecord>& records, const std::optional<std::string>& partition_by) {
   std::vector<std::vector<size_t>> partitions;
   if (!partition_by) {
      partitions.emplace_back();
      for (size_t k = 0; k < records.size(); ++k) partitions[0].push_back(k);
      return partitions;
   }
   std::map<std::optional<std::string>, std::vector<size_t>> by_key;  // nullopt < any string
   for (size_t k = 0; k < records.size(); ++k) {
      auto it = records[k].metadata.find(*partition_by);
      by_key[it == records[k].metadata.end() ? std::nullopt : it->second].push_back(k);
   }
   const double allowed = static_cast<double>(records.size()) / 32.0;
   size_t running = 0;
   bool first = true;
   for (auto& [key, rows] : by_key) {
      if (first || static_cast<double>(running) > allowed) { partitions.emplace_back(); running = 0; first = false; }
      partitions.back().insert(partitions.back().end(), rows.begin(), rows.end());
      running += rows.size();
   }
   return partitions;
}

// what turns the aligned sequences of one store into its index: fillSequenceStore, or the device build (host/device_build.hpp)
using StoreFiller = std::function<void(SequenceStorePartition&, const std::vector<const std::optional<std::string>*>&)>;

inline std::unique_ptr<Database> buildDatabase(
   DatabaseConfig config, const std::map<std::string, std::string>& nuc_refs, const std::map<std::string, std::string>& aa_refs,
   PangoLineageAliasLookup alias, const std::vector<InputRecord>& records, const StoreFiller& fill = fillSequenceStore
) {
   auto db = std::make_unique<Database>();
   db->config = std::move(config);
   db->alias = std::move(alias);
   for (const auto& [name, seq] : nuc_refs) db->nuc_reference[name] = toSymbols(nucleotideAlphabet(), seq, "reference " + name);
   for (const auto& [name, seq] : aa_refs) db->aa_reference[name] = toSymbols(aminoAcidAlphabet(), seq, "reference " + name);
   for (const auto& m : db->config.metadata) {
      if (m.type == ColumnType::INDEXED_STRING || m.type == ColumnType::INDEXED_PANGOLINEAGE) db->dictionaries[m.name] = std::make_unique<Dictionary<std::string>>();
      if (m.type == ColumnType::INDEXED_PANGOLINEAGE) db->aliased_dictionaries[m.name] = std::make_unique<Dictionary<std::string>>();
   }
   auto groups = partitionRecords(records, db->config.partition_by);
   const std::string& pk = db->config.primary_key;
   for (auto& rows : groups) {
      // ORDER BY dateToSortBy, primaryKey with NULLS FIRST
      auto sortKey = [&](size_t r) {
         std::optional<uint32_t> date;
         if (db->config.date_to_sort_by) {
            auto it = records[r].metadata.find(*db->config.date_to_sort_by);
            if (it != records[r].metadata.end() && it->second) { uint32_t d = stringToDate(*it->second); if (d != NULL_DATE) date = d; }
         }
         auto it = records[r].metadata.find(pk);
         std::optional<std::string> key = it == records[r].metadata.end() ? std::nullopt : it->second;
         return std::make_pair(date, key);
      };
      std::stable_sort(rows.begin(), rows.end(), [&](size_t a, size_t b) { return sortKey(a) < sortKey(b); });
      auto part = std::make_unique<DatabasePartition>();
      part->sequence_count = static_cast<uint32_t>(rows.size());
      part->chunks.push_back({0, part->sequence_count});  // one chunk per partition (preprocessing_database.cpp:101-103)
      for (const auto& m : db->config.metadata) {
         auto value = [&](size_t r) -> std::optional<std::string> {
            auto it = records[r].metadata.find(m.name);
            return it == records[r].metadata.end() ? std::nullopt : it->second;
         };
         switch (m.type) {
            case ColumnType::STRING: { auto& c = part->columns.string_columns[m.name]; for (size_t r : rows) c.values.push_back(value(r).value_or("")); break; }
            case ColumnType::INDEXED_STRING: {
               auto& c = part->columns.indexed_string_columns[m.name];
               c.lookup = db->dictionaries[m.name].get();
               for (size_t r : rows) c.insert(value(r).value_or(""));
               break;
            }
            case ColumnType::INDEXED_PANGOLINEAGE: {
               auto& c = part->columns.pango_lineage_columns[m.name];
               c.alias = &db->alias; c.lookup_unaliased = db->dictionaries[m.name].get(); c.lookup_aliased = db->aliased_dictionaries[m.name].get();
               for (size_t r : rows) c.insert(value(r).value_or(""));
               break;
            }
            case ColumnType::DATE: {
               auto& c = part->columns.date_columns[m.name];
               c.is_sorted = db->config.date_to_sort_by && *db->config.date_to_sort_by == m.name;  // database.cpp:590-593
               for (size_t r : rows) { auto v = value(r); c.values.push_back(v ? stringToDate(*v) : NULL_DATE); }
               break;
            }
            case ColumnType::INT: {
               auto& c = part->columns.int_columns[m.name];
               for (size_t r : rows) { auto v = value(r); c.values.push_back(v && !v->empty() ? static_cast<int32_t>(std::stol(*v)) : NULL_INT); }
               break;
            }
            case ColumnType::FLOAT: {
               auto& c = part->columns.float_columns[m.name];
               for (size_t r : rows) { auto v = value(r); c.values.push_back(v && !v->empty() ? std::stod(*v) : std::nan("")); }
               break;
            }
            case ColumnType::NUC_INSERTION: { auto& c = part->columns.nuc_insertion_columns[m.name]; for (size_t r : rows) c.values.push_back(value(r).value_or("")); break; }
            case ColumnType::AA_INSERTION: { auto& c = part->columns.aa_insertion_columns[m.name]; for (size_t r : rows) c.values.push_back(value(r).value_or("")); break; }
         }
      }
      static const std::optional<std::string> null_genome;
      for (int alphabet_id = 0; alphabet_id < 2; ++alphabet_id) {
         auto& refs = alphabet_id == 0 ? db->nuc_reference : db->aa_reference;
         auto& stores = alphabet_id == 0 ? part->nuc_sequences : part->aa_sequences;
         for (const auto& [name, ref] : refs) {
            SequenceStorePartition& store = stores[name];
            store.init(&alphabetById(alphabet_id), &ref);
            std::vector<const std::optional<std::string>*> genomes;
            for (size_t r : rows) {
               const auto& seqs = alphabet_id == 0 ? records[r].nuc_sequences : records[r].aa_sequences;
               auto it = seqs.find(name);
               genomes.push_back(it == seqs.end() ? &null_genome : &it->second);
            }
            fill(store, genomes);
         }
      }
      db->partitions.push_back(std::move(part));
   }
   return db;
}

// reference_genomes.json: {"nucleotideSequences":[{name,sequence}], "genes":[{name,sequence}]} (storage/reference_genomes.cpp)
inline void readReferenceGenomes(const std::string& path, std::map<std::string, std::string>& nuc, std::map<std::string, std::string>& aa) {
   json::Value j = json::parse(io::readFile(path));
   for (const auto& e : j["nucleotideSequences"].arr) nuc[e["name"].as_string()] = e["sequence"].as_string();
   for (const auto& e : j["genes"].arr) aa[e["name"].as_string()] = e["sequence"].as_string();
}

// storage/pango_lineage_alias.cpp:79-93
inline PangoLineageAliasLookup readAliasFile(const std::string& path) {
   PangoLineageAliasLookup lookup;
   json::Value j = json::parse(io::readFile(path));
   for (const auto& [key, value] : j.obj) {
      if (value.is_array()) { std::vector<std::string> v; for (const auto& e : value.arr) v.push_back(e.as_string()); lookup.alias_key[key] = v; }
      else if (value.is_string() && !value.s.empty()) lookup.alias_key[key] = {value.s};
   }
   return lookup;
}

// preprocessing/metadata_info.cpp:39-46: every metadata field of the database config has to be in the input (the header of the
// metadata file; for ndjson the keys of the FIRST record's "metadata" object, :129-160)
inline void requireMetadataField(const std::string& name, bool present) {
   if (!present) throw std::runtime_error("The metadata field '" + name + "' which is contained in the database config is not contained in the input.");
}

// preprocessing/sequence_info.cpp:79-151: the sequence names of the FIRST ndjson record and of the reference genomes must be the same
// sets, both ways, for nucleotide sequences and for genes
inline void validateSequenceNames(const json::Value& in_record, const std::map<std::string, std::string>& references, const char* what, const std::string& file) {
   for (const auto& [name, _] : in_record.obj)
      if (!references.count(name)) throw std::runtime_error(std::string("The aligned ") + what + " sequence " + name + " which is contained in the input file " + file + " is not contained in the reference sequences.");
   for (const auto& [name, _] : references)
      if (!in_record.contains(name)) throw std::runtime_error(std::string("The aligned ") + what + " sequence " + name + " which is contained in the reference sequences is not contained in the input file " + file + ".");
}

inline std::string jsonScalarToString(const json::Value& v) {
   if (v.is_string()) return v.s;
   if (v.is_boolean()) return v.b ? "true" : "false";
   return v.dump();
}

// Builds a Database from a directory laid out like the reference's testBaseData/exampleDataset*.
inline std::unique_ptr<Database> buildDatabaseFromDirectory(const std::string& dir, const StoreFiller& fill = fillSequenceStore) {
   DatabaseConfig config = parseDatabaseConfig(io::readFile(dir + "/database_config.yaml"));
   std::string metadata_file, ndjson_file, alias_file = "pangolineage_alias.json", ref_file = "reference_genomes.json";
   for (const auto& y : io::parseYamlLines(io::readFile(dir + "/preprocessing_config.yaml"))) {
      if (y.key == "metadataFilename") metadata_file = y.value;
      else if (y.key == "ndjsonInputFilename") ndjson_file = y.value;
      else if (y.key == "pangoLineageDefinitionFilename") alias_file = y.value;
      else if (y.key == "referenceGenomeFilename") ref_file = y.value;
   }
   std::map<std::string, std::string> nuc_refs, aa_refs;
   readReferenceGenomes(dir + "/" + ref_file, nuc_refs, aa_refs);
   PangoLineageAliasLookup alias;
   if (io::fileExists(dir + "/" + alias_file)) alias = readAliasFile(dir + "/" + alias_file);

   std::vector<InputRecord> records;
   if (!ndjson_file.empty()) {
      std::string text = io::readFile(dir + "/" + ndjson_file);
      if (ndjson_file.size() > 4 && ndjson_file.substr(ndjson_file.size() - 4) == ".zst") text = io::zstdDecompress(text);
      // one record per line; a file whose first line only opens an object is one pretty-printed record (DuckDB's read_json, which the
      // reference reads its input with, detects both layouts: testBaseData/ndjsonFiles/oneline_*.json)
      std::vector<std::pair<size_t, size_t>> documents;      // {offset, length}
      const size_t first_eol = std::min(text.find('\n'), text.size());
      const size_t first_end = text.find_last_not_of(" \t\r", first_eol == 0 ? 0 : first_eol - 1);
      if (first_eol > 0 && first_end != std::string::npos && first_end < first_eol && text.find_first_not_of(" \t") == first_end && text[first_end] == '{') documents.emplace_back(0, text.size());
      else {
         for (size_t pos = 0; pos < text.size();) {
            size_t eol = text.find('\n', pos);
            if (eol == std::string::npos) eol = text.size();
            if (eol > pos) documents.emplace_back(pos, eol - pos);
            pos = eol + 1;
         }
      }
      for (const auto& [offset, length] : documents) {
         {
            json::Value j = json::parse(text.data() + offset, length);
            if (records.empty()) for (const auto& m : config.metadata)      // the two insertion lists live beside "metadata" (metadata_info.cpp:61-95)
               requireMetadataField(m.name, j["metadata"].contains(m.name) || ((m.name == "nucleotideInsertions" || m.name == "aminoAcidInsertions") && j.contains(m.name)));
            if (records.empty()) {
               validateSequenceNames(j["alignedNucleotideSequences"], nuc_refs, "nucleotide", dir + "/" + ndjson_file);
               validateSequenceNames(j["alignedAminoAcidSequences"], aa_refs, "amino acid", dir + "/" + ndjson_file);
            }
            InputRecord rec;
            for (const auto& [k, v] : j["metadata"].obj) rec.metadata[k] = v.is_null() ? std::nullopt : std::optional<std::string>(jsonScalarToString(v));
            for (const auto& [k, v] : j["alignedNucleotideSequences"].obj) rec.nuc_sequences[k] = v.is_null() ? std::nullopt : std::optional<std::string>(v.s);
            for (const auto& [k, v] : j["alignedAminoAcidSequences"].obj) rec.aa_sequences[k] = v.is_null() ? std::nullopt : std::optional<std::string>(v.s);
            // insertion lists become "seq:pos:INS,..." strings (preprocessing/metadata_info.cpp:71)
            for (const char* field : {"nucleotideInsertions", "aminoAcidInsertions"}) {
               if (!j.contains(field)) continue;
               std::string joined;
               for (const auto& [seq, list] : j[field].obj) for (const auto& ins : list.arr) { if (!joined.empty()) joined += ','; joined += seq + ":" + ins.as_string(); }
               rec.metadata[field] = joined;
            }
            records.push_back(std::move(rec));
         }
      }
   } else {
      io::Table table = io::parseTsv(io::readFile(dir + "/" + metadata_file));
      for (const auto& m : config.metadata) requireMetadataField(m.name, table.columnIndex(m.name) >= 0);
      const int pk = table.columnIndex(config.primary_key);
      if (pk < 0) throw std::runtime_error("metadata file has no primary key column " + config.primary_key);
      std::map<std::string, std::map<std::string, std::string>> nuc_fasta, aa_fasta;
      for (const auto& [name, _] : nuc_refs) { bool found; auto text = io::readMaybeCompressed(dir + "/nuc_" + name + ".fasta", found); if (found) nuc_fasta[name] = io::parseFasta(text); }
      for (const auto& [name, _] : aa_refs) { bool found; auto text = io::readMaybeCompressed(dir + "/gene_" + name + ".fasta", found); if (found) aa_fasta[name] = io::parseFasta(text); }
      for (const auto& row : table.rows) {
         InputRecord rec;
         for (size_t c = 0; c < table.header.size(); ++c) rec.metadata[table.header[c]] = row[c].empty() ? std::nullopt : std::optional<std::string>(row[c]);
         const std::string& key = row[pk];
         for (const auto& [name, _] : nuc_refs) {
            auto f = nuc_fasta.find(name);
            if (f == nuc_fasta.end()) { rec.nuc_sequences[name] = std::nullopt; continue; }
            auto it = f->second.find(key);
            rec.nuc_sequences[name] = it == f->second.end() ? std::nullopt : std::optional<std::string>(it->second);
         }
         for (const auto& [name, _] : aa_refs) {
            auto f = aa_fasta.find(name);
            if (f == aa_fasta.end()) { rec.aa_sequences[name] = std::nullopt; continue; }
            auto it = f->second.find(key);
            rec.aa_sequences[name] = it == f->second.end() ? std::nullopt : std::optional<std::string>(it->second);
         }
         records.push_back(std::move(rec));
      }
   }
   return buildDatabase(std::move(config), nuc_refs, aa_refs, std::move(alias), records, fill);
}

}  // namespace silo::builder
