// C entry points for building / inspecting the host Database (used through ctypes by tests,
// bench.py and __graft_entry__; a C++ host links the headers directly).
#include "builder.hpp"
#include "capi_common.hpp"
#include "device_build.hpp"

using namespace silo;

extern "C" {

const char* silo_last_error() { return capi::lastError().c_str(); }
void silo_free(void* p) { free(p); }

void* silo_db_build_from_dir(const char* dir) {
   Database* db = nullptr;
   capi::guarded([&] { db = builder::buildDatabaseFromDirectory(dir).release(); });
   return db;
}
// the same database with every sequence store built on the device (sb_index_build) instead of builder::fillSequenceStore
void* silo_db_build_from_dir_device(const char* dir, int device) {
   Database* db = nullptr;
   capi::guarded([&] { db = builder::buildDatabaseFromDirectory(dir, builder::deviceStoreFiller(device)).release(); });
   return db;
}
// number of differences between the sequence stores of two databases: bitmaps (set and serialized form), flags, missing-symbol
// bitmaps, sequence counts; -1 on error
int64_t silo_db_compare_stores(const void* a_, const void* b_) {
   int64_t differences = -1;
   capi::guarded([&] {
      const auto& a = *static_cast<const Database*>(a_);
      const auto& b = *static_cast<const Database*>(b_);
      if (a.partitions.size() != b.partitions.size()) throw std::runtime_error("different numbers of partitions");
      int64_t diff = 0;
      for (size_t p = 0; p < a.partitions.size(); ++p) {
         for (int alphabet_id = 0; alphabet_id < 2; ++alphabet_id) {
            const auto& stores_b = b.partitions[p]->getSequenceStores(alphabet_id);
            for (const auto& [name, x] : a.partitions[p]->getSequenceStores(alphabet_id)) {
               const SequenceStorePartition& y = stores_b.at(name);
               if (x.sequence_count != y.sequence_count || x.bitmaps.size() != y.bitmaps.size() || x.missing_symbol_bitmaps.size() != y.missing_symbol_bitmaps.size()) { ++diff; continue; }
               for (size_t k = 0; k < x.bitmaps.size(); ++k) if (!(x.bitmaps[k] == y.bitmaps[k]) || x.bitmaps[k].toPortable() != y.bitmaps[k].toPortable()) ++diff;
               for (size_t k = 0; k < x.deleted.size(); ++k) if (x.deleted[k] != y.deleted[k] || x.flipped[k] != y.flipped[k]) ++diff;
               for (size_t k = 0; k < x.missing_symbol_bitmaps.size(); ++k) if (x.missing_symbol_bitmaps[k].toPortable() != y.missing_symbol_bitmaps[k].toPortable()) ++diff;
            }
         }
      }
      differences = diff;
   });
   return differences;
}
// Test hook (tests/test_host_cpu.py): PangoLineageAliasLookup built from a JSON object {"alias": ["value", ...] | "value"},
// applied to one lineage; direction 0 = unalias (pango_lineage_alias.cpp:25-47), 1 = alias (:49-79)
char* silo_alias_apply(const char* alias_json, const char* lineage, int direction) {
   char* out = nullptr;
   capi::guarded([&] {
      PangoLineageAliasLookup lookup;
      const json::Value j = json::parse(alias_json);
      for (const auto& [key, value] : j.obj) {
         if (value.is_array()) { std::vector<std::string> v; for (const auto& e : value.arr) v.push_back(e.as_string()); lookup.alias_key[key] = v; }
         else if (value.is_string() && !value.s.empty()) lookup.alias_key[key] = {value.s};
      }
      out = capi::dupString(direction ? lookup.alias(lineage) : lookup.unalias(lineage));
   });
   return out;
}
// Test hook (tests/test_host_cpu.py): the entries of a FASTA text as the builder reads them, as a JSON array of [key, sequence]
char* silo_parse_fasta(const char* text, size_t n) {
   char* out = nullptr;
   capi::guarded([&] {
      json::Value entries = json::Value::array();
      for (const auto& [key, genome] : io::readFastaEntries(std::string(text, n))) {
         json::Value entry = json::Value::array();
         entry.arr.push_back(json::Value::string(key));
         entry.arr.push_back(json::Value::string(genome));
         entries.arr.push_back(std::move(entry));
      }
      out = capi::dupString(entries.dump());
   });
   return out;
}
// Test hook (tests/test_host_cpu.py): a database_config.yaml text as the builder understands it
char* silo_parse_database_config(const char* text) {
   char* out = nullptr;
   capi::guarded([&] {
      static const char* const names[] = {"STRING", "INDEXED_STRING", "INDEXED_PANGOLINEAGE", "DATE", "INT", "FLOAT", "NUC_INSERTION", "AA_INSERTION"};
      const DatabaseConfig config = builder::parseDatabaseConfig(text);
      json::Value j = json::Value::object();
      j.set("instanceName", json::Value::string(config.instance_name));
      j.set("primaryKey", json::Value::string(config.primary_key));
      j.set("dateToSortBy", config.date_to_sort_by ? json::Value::string(*config.date_to_sort_by) : json::Value::null());
      j.set("partitionBy", config.partition_by ? json::Value::string(*config.partition_by) : json::Value::null());
      json::Value metadata = json::Value::array();
      for (const auto& m : config.metadata) {
         json::Value entry = json::Value::array();
         entry.arr.push_back(json::Value::string(m.name));
         entry.arr.push_back(json::Value::string(names[static_cast<int>(m.type)]));
         metadata.arr.push_back(std::move(entry));
      }
      j.set("metadata", std::move(metadata));
      out = capi::dupString(j.dump());
   });
   return out;
}
void silo_db_free(void* db) { delete static_cast<Database*>(db); }
uint32_t silo_db_partition_count(const void* db) { return static_cast<uint32_t>(static_cast<const Database*>(db)->partitions.size()); }
uint32_t silo_db_partition_rows(const void* db, uint32_t p) { return static_cast<const Database*>(db)->partitions.at(p)->sequence_count; }

// Mirrors Database::getDatabaseInfo / detailedDatabaseInfo (src/silo/database.cpp:122-146,239-305)
char* silo_db_info_json(const void* db_) {
   char* out = nullptr;
   capi::guarded([&] {
      const auto& db = *static_cast<const Database*>(db_);
      uint64_t total_size = 0, n_bitmaps_size = 0, sequence_count = 0;
      json::Value sequences = json::Value::object();
      std::map<std::string, std::array<uint64_t, 8>> per_sequence;  // arr, arr_values, run, run_values, bitset, bitset_values, portable, run_bytes
      for (const auto& part : db.partitions) {
         sequence_count += part->sequence_count;
         for (const auto& [name, store] : part->nuc_sequences) {
            auto& st = per_sequence[name];
            for (const auto& bm : store.bitmaps) {
               total_size += bm.nonPortableSizeInBytes();
               st[6] += bm.portableSizeInBytes();
               for (const auto& c : bm.containers) {
                  if (c.kind == roaring::ARRAY) { st[0]++; st[1] += c.card; }
                  else if (c.kind == roaring::RUN) { st[2]++; st[3] += c.card; st[7] += 2 + 4 * c.runCount(); }
                  else { st[4]++; st[5] += c.card; }
               }
            }
            for (const auto& bm : store.missing_symbol_bitmaps) n_bitmaps_size += bm.nonPortableSizeInBytes();
         }
      }
      for (const auto& [name, st] : per_sequence) {
         json::Value s = json::Value::object();
         s.set("numberOfArrayContainers", json::Value::integer(st[0]));
         s.set("numberOfValuesStoredInArrayContainers", json::Value::integer(st[1]));
         s.set("numberOfRunContainers", json::Value::integer(st[2]));
         s.set("numberOfValuesStoredInRunContainers", json::Value::integer(st[3]));
         s.set("numberOfBitsetContainers", json::Value::integer(st[4]));
         s.set("numberOfValuesStoredInBitsetContainers", json::Value::integer(st[5]));
         s.set("totalBitmapSizeComputed", json::Value::integer(st[6]));
         s.set("totalBitmapSizeArrayContainers", json::Value::integer(2 * st[1]));
         sequences.set(name, std::move(s));
      }
      json::Value info = json::Value::object();
      info.set("sequenceCount", json::Value::integer(sequence_count));
      info.set("totalSize", json::Value::integer(total_size));
      info.set("nBitmapsSize", json::Value::integer(n_bitmaps_size));
      info.set("numberOfPartitions", json::Value::integer(db.partitions.size()));
      json::Value rows = json::Value::array();
      for (const auto& part : db.partitions) rows.arr.push_back(json::Value::integer(part->sequence_count));
      info.set("partitionRows", std::move(rows));
      info.set("sequences", std::move(sequences));
      out = capi::dupString(info.dump());
   });
   return out;
}

}  // extern "C"
