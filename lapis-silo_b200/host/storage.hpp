// In-memory database model of the host side (what the reference keeps in silo::Database).
// Mirrors: include/silo/storage/position.h:22-66, include/silo/storage/sequence_store.h:34-99,
// include/silo/storage/database_partition.h:39-112, include/silo/database.h:36-106 and the
// metadata columns of src/silo/storage/column/*. Only what the query hot path reads is kept.
#pragma once
#include <cmath>
#include <cstdint>
#include <limits>
#include <map>
#include <memory>
#include <optional>
#include <string>
#include <unordered_map>
#include <vector>

#include "roaring.hpp"
#include "symbols.hpp"

namespace silo {

using roaring::Bitmap;

// ---- sequence index ----------------------------------------------------------------------

// One partition of one aligned sequence (nucleotide segment or gene).
// positions[pos] of the reference = bitmaps[pos * Σ + sym] + flipped[pos] + deleted[pos].
struct SequenceStorePartition {
   const Alphabet* alphabet = nullptr;
   const std::vector<Symbol>* reference_sequence = nullptr;
   std::vector<Bitmap> bitmaps;       // [L * Σ]
   std::vector<int8_t> flipped;       // [L], -1 = none  (position.h:36)
   std::vector<int8_t> deleted;       // [L], -1 = none  (position.h:37)
   std::vector<Bitmap> missing_symbol_bitmaps;  // one per row: positions with N / X (sequence_store.h:55)
   uint32_t sequence_count = 0;

   uint32_t length() const { return static_cast<uint32_t>(reference_sequence->size()); }
   const Bitmap* getBitmap(uint32_t pos, Symbol s) const { return &bitmaps[size_t(pos) * alphabet->count + s]; }
   Bitmap& bitmapRef(uint32_t pos, Symbol s) { return bitmaps[size_t(pos) * alphabet->count + s]; }
   bool isSymbolFlipped(uint32_t pos, Symbol s) const { return flipped[pos] == static_cast<int8_t>(s); }
   bool isSymbolDeleted(uint32_t pos, Symbol s) const { return deleted[pos] == static_cast<int8_t>(s); }

   void init(const Alphabet* a, const std::vector<Symbol>* ref) {
      alphabet = a;
      reference_sequence = ref;
      bitmaps.assign(size_t(ref->size()) * a->count, Bitmap());
      flipped.assign(ref->size(), -1);
      deleted.assign(ref->size(), -1);
   }
};

// ---- metadata columns -----------------------------------------------------------------------

constexpr uint32_t NULL_DATE = 0;                                   // common/date.h:11
constexpr int32_t NULL_INT = std::numeric_limits<int32_t>::min();   // int_column.cpp:17

// common/date.cpp:20-64: (year << 16) + (month << 12) + day, 0 for anything unparsable
inline uint32_t stringToDate(const std::string& value) {
   if (value.empty()) return NULL_DATE;
   auto p1 = value.find('-');
   if (p1 == std::string::npos) return NULL_DATE;
   auto p2 = value.find('-', p1 + 1);
   if (p2 == std::string::npos) return NULL_DATE;
   try {
      const uint32_t year = std::stoi(value.substr(0, p1));
      const uint32_t month = std::stoi(value.substr(p1 + 1, p2));
      const uint32_t day = std::stoi(value.substr(p2 + 1));
      if (month > 12 || month == 0) return NULL_DATE;
      if (day > 31 || day == 0) return NULL_DATE;
      return (year << 16) + (month << 12) + day;
   } catch (const std::exception&) {
      return NULL_DATE;
   }
}

template <typename T>
struct Dictionary {  // common/bidirectional_map.h: value <-> dense id, shared by all partitions
   std::vector<T> values;
   std::unordered_map<T, uint32_t> ids;
   uint32_t getOrCreateId(const T& v) {
      auto it = ids.find(v);
      if (it != ids.end()) return it->second;
      uint32_t id = static_cast<uint32_t>(values.size());
      values.push_back(v);
      ids.emplace(v, id);
      return id;
   }
   std::optional<uint32_t> getId(const T& v) const {
      auto it = ids.find(v);
      if (it == ids.end()) return std::nullopt;
      return it->second;
   }
};

struct DateColumnPartition {
   bool is_sorted = false;
   std::vector<uint32_t> values;
};
struct IntColumnPartition { std::vector<int32_t> values; };
struct FloatColumnPartition { std::vector<double> values; };
struct StringColumnPartition { std::vector<std::string> values; };
struct InsertionColumnPartition { std::vector<std::string> values; };  // raw "pos:INS" / "seq:pos:INS" lists

struct IndexedStringColumnPartition {
   Dictionary<std::string>* lookup = nullptr;
   std::vector<uint32_t> value_ids;
   std::unordered_map<uint32_t, Bitmap> indexed_values;
   void insert(const std::string& value) {  // indexed_string_column.cpp:23-33 (null -> "")
      uint32_t id = lookup->getOrCreateId(value);
      indexed_values[id].add(static_cast<uint32_t>(value_ids.size()));
      value_ids.push_back(id);
   }
   const Bitmap* filter(const std::string& value) const {  // indexed_string_column.cpp:14-21
      auto id = lookup->getId(value);
      if (!id) return nullptr;
      auto it = indexed_values.find(*id);
      return it == indexed_values.end() ? nullptr : &it->second;
   }
};

// storage/pango_lineage_alias.cpp:25-47 (only unaliasing is on the query path)
struct PangoLineageAliasLookup {
   std::unordered_map<std::string, std::vector<std::string>> alias_key;
   std::string unalias(const std::string& lineage) const {
      auto dot = lineage.find('.');
      std::string prefix = lineage.substr(0, dot);
      auto it = alias_key.find(prefix);
      if (it == alias_key.end() || it->second.size() != 1) return lineage;
      if (dot == std::string::npos) return it->second[0];
      return it->second[0] + '.' + lineage.substr(dot + 1);
   }
   // storage/pango_lineage_alias.cpp:49-79: the longest prefix of at least three elements that some alias stands for is
   // replaced by that alias (the reference walks its unordered_map; with several aliases of one value the choice is its
   // iteration order, here the smallest alias name)
   std::string alias(const std::string& lineage) const {
      std::vector<std::string> elements;
      for (size_t at = 0;;) {
         const size_t dot = lineage.find('.', at);
         elements.push_back(lineage.substr(at, dot == std::string::npos ? std::string::npos : dot - at));
         if (dot == std::string::npos) break;
         at = dot + 1;
      }
      for (size_t i = elements.size(); i > 3; --i) {
         std::string search;
         for (size_t k = 0; k + 1 < i; ++k) search += (k ? "." : "") + elements[k];
         const std::string* best = nullptr;
         for (const auto& [name, values] : alias_key) if (values.size() == 1 && values[0] == search && (!best || name < *best)) best = &name;
         if (best) {
            std::string value = *best;
            for (size_t k = i - 1; k < elements.size(); ++k) value += "." + elements[k];
            return value;
         }
      }
      return lineage;
   }
};

struct PangoLineageColumnPartition {
   const PangoLineageAliasLookup* alias = nullptr;
   Dictionary<std::string>* lookup_unaliased = nullptr;
   Dictionary<std::string>* lookup_aliased = nullptr;   // filled side by side with lookup_unaliased (pango_lineage_column.cpp:21-33); optional
   std::vector<uint32_t> value_ids;
   std::unordered_map<uint32_t, Bitmap> indexed_values;
   std::unordered_map<uint32_t, Bitmap> indexed_sublineage_values;

   // common/pango_lineage.cpp:25-35: every dot-prefix including the full name
   static std::vector<std::string> parentLineages(const std::string& value) {
      std::vector<std::string> out;
      std::string::size_type pos = 0;
      while (pos != std::string::npos) {
         pos = value.find('.', pos + 1);
         out.push_back(value.substr(0, pos));
      }
      return out;
   }
   void insert(const std::string& raw) {  // pango_lineage_column.cpp:21-55 (null -> "")
      const std::string resolved = alias->unalias(raw);
      const auto parents = parentLineages(resolved);
      for (const auto& p : parents) {
         (void)lookup_unaliased->getOrCreateId(p);
         if (lookup_aliased) (void)lookup_aliased->getOrCreateId(alias->alias(p));
      }
      const uint32_t id = lookup_unaliased->getOrCreateId(resolved);
      if (lookup_aliased) (void)lookup_aliased->getOrCreateId(alias->alias(resolved));
      const uint32_t row = static_cast<uint32_t>(value_ids.size());
      value_ids.push_back(id);
      indexed_values[id].add(row);
      for (const auto& p : parents) indexed_sublineage_values[lookup_unaliased->getOrCreateId(p)].add(row);
   }
   // pango_lineage_column.cpp:82-84: the value Aggregated / Details print for an id -- the id is one of lookup_unaliased and is
   // looked up in lookup_aliased, as the reference does
   std::string lookupAliasedValue(uint32_t id) const {
      if (lookup_aliased) return id < lookup_aliased->values.size() ? lookup_aliased->values[id] : std::string();
      return alias->alias(lookup_unaliased->values.at(id));
   }
   const Bitmap* filter(const std::string& value, bool include_sublineages) const {  // :57-77
      auto id = lookup_unaliased->getId(alias->unalias(value));
      if (!id) return nullptr;
      const auto& map = include_sublineages ? indexed_sublineage_values : indexed_values;
      auto it = map.find(*id);
      return it == map.end() ? nullptr : &it->second;
   }
};

struct ColumnPartitionGroup {  // include/silo/storage/column_group.h:49-76
   std::map<std::string, StringColumnPartition> string_columns;
   std::map<std::string, IndexedStringColumnPartition> indexed_string_columns;
   std::map<std::string, IntColumnPartition> int_columns;
   std::map<std::string, FloatColumnPartition> float_columns;
   std::map<std::string, DateColumnPartition> date_columns;
   std::map<std::string, PangoLineageColumnPartition> pango_lineage_columns;
   std::map<std::string, InsertionColumnPartition> nuc_insertion_columns;
   std::map<std::string, InsertionColumnPartition> aa_insertion_columns;
};

struct PartitionChunk { uint32_t offset = 0; uint32_t size = 0; };  // preprocessing/partition.h

struct DatabasePartition {
   std::map<std::string, SequenceStorePartition> nuc_sequences;
   std::map<std::string, SequenceStorePartition> aa_sequences;
   ColumnPartitionGroup columns;
   std::vector<PartitionChunk> chunks;
   uint32_t sequence_count = 0;
   const std::map<std::string, SequenceStorePartition>& getSequenceStores(int alphabet_id) const {
      return alphabet_id == 0 ? nuc_sequences : aa_sequences;
   }
};

enum class ColumnType { STRING, INDEXED_STRING, INDEXED_PANGOLINEAGE, DATE, INT, FLOAT, NUC_INSERTION, AA_INSERTION };
struct ColumnMetadata { std::string name; ColumnType type; };

struct DatabaseConfig {
   std::string instance_name;
   std::vector<ColumnMetadata> metadata;
   std::string primary_key;
   std::optional<std::string> date_to_sort_by;
   std::optional<std::string> partition_by;
   std::string default_nucleotide_sequence = "main";
};

struct Database {
   DatabaseConfig config;
   // name -> reference sequence (std::map: Mutations iterates stores in name order, mutations.cpp:248-252)
   std::map<std::string, std::vector<Symbol>> nuc_reference;
   std::map<std::string, std::vector<Symbol>> aa_reference;
   std::vector<std::unique_ptr<DatabasePartition>> partitions;
   PangoLineageAliasLookup alias;
   std::map<std::string, std::unique_ptr<Dictionary<std::string>>> dictionaries;  // per indexed / lineage column
   std::map<std::string, std::unique_ptr<Dictionary<std::string>>> aliased_dictionaries;  // per lineage column: lookup_aliased

   const std::map<std::string, std::vector<Symbol>>& getReferences(int alphabet_id) const {
      return alphabet_id == 0 ? nuc_reference : aa_reference;
   }
   uint64_t totalSequenceCount() const {
      uint64_t n = 0;
      for (const auto& p : partitions) n += p->sequence_count;
      return n;
   }
};

}  // namespace silo
