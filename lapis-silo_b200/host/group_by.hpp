// Aggregated with groupByFields, host side (src/silo/query_engine/actions/aggregated.cpp:24-40,91-149 and
// actions/tuple.cpp:29-155). The reference copies the group-by values of every filtered row into a byte tuple and counts
// tuples in hash maps. Here every group-by column is dictionary-coded once -- code = index of the row's value among the
// column's distinct values over all partitions of this database -- the codes live on the device (sb_group_column_add) and a
// tuple is the mixed-radix number of its codes (sb_aggregated_grouped). Two tuples are equal in the reference when their
// bytes are (Tuple::operator==, tuple.cpp:425-427): for a FLOAT column that is equality of the bit pattern, which is what the
// dictionary of a float column is keyed by.
#pragma once
#include <cstring>
#include <set>

#include "../../include/silo_b200.h"
#include "query.hpp"

namespace silo::engine {

// common/date.cpp:66-80
inline FieldValue dateToField(uint32_t date) {
   if (date == 0) return std::nullopt;
   char text[32];
   snprintf(text, sizeof text, "%04u-%02u-%02u", date >> 16, (date >> 12) & 0xFu, date & 0xFFFu);
   return std::string(text);
}
inline FieldValue stringToField(const std::string& value) {   // tuple.cpp:109-116: the empty string is the null of every string-like column
   if (value.empty()) return std::nullopt;
   return value;
}

// aggregated.cpp:24-40
inline std::vector<ColumnMetadata> parseGroupByFields(const DatabaseConfig& config, const std::vector<std::string>& group_by_fields) {
   std::vector<ColumnMetadata> out;
   for (const std::string& field : group_by_fields) {
      const ColumnMetadata* found = nullptr;
      for (const auto& m : config.metadata) if (m.name == field) found = &m;
      CHECK_SILO_QUERY(found != nullptr, "Metadata field '" + field + "' to group by not found")
      out.push_back(*found);
   }
   return out;
}

struct GroupDictionary {
   ColumnMetadata meta;
   uint32_t cardinality = 1;
   std::vector<FieldValue> values;              // code -> the field of the response (tuple.cpp:78-155)
   std::vector<std::vector<uint32_t>> codes;    // per local partition, until uploaded
   std::vector<uint32_t> column_ids;            // per local partition: id from sb_group_column_add
};

template <typename T, typename ToField>
inline void codeByValue(GroupDictionary& d, const std::vector<const std::vector<T>*>& columns, ToField&& to_field) {
   // distinct values in ascending order of their bits / text: the code of a value does not depend on which partition shows it first
   std::set<T> distinct;
   for (const auto* column : columns) distinct.insert(column->begin(), column->end());
   std::vector<T> sorted(distinct.begin(), distinct.end());
   d.cardinality = std::max<uint32_t>(1, static_cast<uint32_t>(sorted.size()));
   d.values.clear();
   for (const T& v : sorted) d.values.push_back(to_field(v));
   if (d.values.empty()) d.values.push_back(std::nullopt);
   for (const auto* column : columns) {
      std::vector<uint32_t> codes(column->size());
      for (size_t row = 0; row < column->size(); ++row) codes[row] = static_cast<uint32_t>(std::lower_bound(sorted.begin(), sorted.end(), (*column)[row]) - sorted.begin());
      d.codes.push_back(std::move(codes));
   }
}

template <typename ColumnT>
inline auto valueColumns(const Database& db, const std::vector<uint32_t>& partitions, std::map<std::string, ColumnT> ColumnPartitionGroup::*member, const std::string& name) {
   std::vector<const decltype(ColumnT::values)*> out;
   for (uint32_t p : partitions) out.push_back(&(db.partitions[p]->columns.*member).at(name).values);
   return out;
}

// The dictionary and the per-partition codes of one metadata column over the given partitions.
inline GroupDictionary buildGroupDictionary(const Database& db, const std::vector<uint32_t>& partitions, const ColumnMetadata& meta) {
   GroupDictionary d;
   d.meta = meta;
   auto columnsOf = [&](auto member) { return valueColumns(db, partitions, member, meta.name); };
   switch (meta.type) {
      case ColumnType::INDEXED_STRING: {
         const Dictionary<std::string>* lookup = nullptr;
         for (uint32_t p : partitions) {
            const auto& col = db.partitions[p]->columns.indexed_string_columns.at(meta.name);
            lookup = col.lookup;
            d.codes.push_back(col.value_ids);
         }
         if (lookup) for (const auto& v : lookup->values) d.values.push_back(stringToField(v));
         break;
      }
      case ColumnType::INDEXED_PANGOLINEAGE: {
         const PangoLineageColumnPartition* any = nullptr;
         for (uint32_t p : partitions) {
            const auto& col = db.partitions[p]->columns.pango_lineage_columns.at(meta.name);
            any = &col;
            d.codes.push_back(col.value_ids);
         }
         if (any) for (uint32_t id = 0; id < any->lookup_unaliased->values.size(); ++id) d.values.push_back(stringToField(any->lookupAliasedValue(id)));
         break;
      }
      case ColumnType::STRING: codeByValue<std::string>(d, columnsOf(&ColumnPartitionGroup::string_columns), stringToField); return d;
      case ColumnType::NUC_INSERTION: codeByValue<std::string>(d, columnsOf(&ColumnPartitionGroup::nuc_insertion_columns), stringToField); return d;
      case ColumnType::AA_INSERTION: codeByValue<std::string>(d, columnsOf(&ColumnPartitionGroup::aa_insertion_columns), stringToField); return d;
      case ColumnType::DATE: codeByValue<uint32_t>(d, columnsOf(&ColumnPartitionGroup::date_columns), dateToField); return d;
      case ColumnType::INT:
         codeByValue<int32_t>(d, columnsOf(&ColumnPartitionGroup::int_columns), [](int32_t v) -> FieldValue { if (v == NULL_INT) return std::nullopt; return v; });
         return d;
      case ColumnType::FLOAT: {      // keyed by the bit pattern (see the head of this file)
         std::vector<std::vector<uint64_t>> bits;
         for (uint32_t p : partitions) {
            const auto& values = db.partitions[p]->columns.float_columns.at(meta.name).values;
            bits.emplace_back(values.size());
            if (!values.empty()) memcpy(bits.back().data(), values.data(), values.size() * sizeof(double));
         }
         std::vector<const std::vector<uint64_t>*> columns;
         for (const auto& b : bits) columns.push_back(&b);
         codeByValue<uint64_t>(d, columns, [](uint64_t b) -> FieldValue { double v; memcpy(&v, &b, sizeof v); if (std::isnan(v)) return std::nullopt; return v; });
         return d;
      }
   }
   d.cardinality = std::max<uint32_t>(1, static_cast<uint32_t>(d.values.size()));
   if (d.values.empty()) d.values.push_back(std::nullopt);
   return d;
}

// {key, count} rows of sb_aggregated_grouped -> result entries (aggregated.cpp:46-56: the tuple's fields + "count")
inline QueryResult groupedResult(const std::vector<const GroupDictionary*>& dictionaries, const sb_group_row* rows, size_t n_rows) {
   QueryResult result;
   result.query_result.reserve(n_rows);
   for (size_t i = 0; i < n_rows; ++i) {
      QueryResultEntry entry;
      uint64_t key = rows[i].key;
      for (size_t c = dictionaries.size(); c-- > 0;) {
         const GroupDictionary& d = *dictionaries[c];
         entry.fields[d.meta.name] = d.values.at(key % d.cardinality);
         key /= d.cardinality;
      }
      entry.fields["count"] = static_cast<int32_t>(rows[i].count);
      result.query_result.push_back(std::move(entry));
   }
   return result;
}

// ---- Aggregated over a database whose partitions live in several processes (one per GPU) ------------------------------------
// Every process answers for ITS partitions without ordering and limits (a partial: the entries of aggregated.cpp:46-56), one
// process merges them the way aggregated.cpp:138-148 merges the partitions' maps -- equal tuples add up -- and then orders.
inline FieldValue fieldFromJson(const json::Value& v) {
   if (v.is_null()) return std::nullopt;
   if (v.is_string()) return v.s;
   if (v.is_number_float()) return v.d;
   return static_cast<int32_t>(v.is_number_unsigned() ? static_cast<int64_t>(v.u) : v.i);
}
inline QueryResult mergeAggregatedPartials(const Action& action, const DatabaseConfig& config, const std::vector<std::string>& partial_responses) {
   if (action.type != ActionType::AGGREGATED) throw QueryCompilationException("only Aggregated results are merged on the host (Mutations tables are summed on the devices)");
   (void)parseGroupByFields(config, action.group_by_fields);
   return action.executeAndOrder([&]() -> QueryResult {
      std::map<std::map<std::string, FieldValue>, uint32_t> merged;      // tuple -> count (uint32 like the reference's map values)
      for (const std::string& text : partial_responses) {
         const json::Value j = json::parse(text);
         for (const auto& entry : j["queryResult"].arr) {
            std::map<std::string, FieldValue> tuple;
            uint32_t count = 0;
            for (const auto& [name, value] : entry.obj) {
               if (name == "count") count = static_cast<uint32_t>(value.is_number_unsigned() ? value.u : static_cast<uint64_t>(value.i));
               else tuple[name] = fieldFromJson(value);
            }
            merged[tuple] += count;
         }
      }
      QueryResult result;
      if (action.group_by_fields.empty() && merged.empty()) merged[{}] = 0;      // aggregateWithoutGrouping always yields one row
      for (auto& [tuple, count] : merged) {
         QueryResultEntry e;
         e.fields = tuple;
         e.fields["count"] = static_cast<int32_t>(count);
         result.query_result.push_back(std::move(e));
      }
      return result;
   });
}

}  // namespace silo::engine
