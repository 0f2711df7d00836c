// Query JSON -> (filter Expression, Action description); QueryResult and its JSON form; ordering.
// Mirrors src/silo/query_engine/query.cpp:13-28, actions/action.cpp:37-187,
// actions/mutations.cpp:276-311, actions/aggregated.cpp:153-157, query_result.cpp:10-25.
#pragma once
#include <map>
#include <optional>
#include <string>
#include <variant>
#include <vector>

#include "expressions.hpp"

namespace silo {

using FieldValue = std::optional<std::variant<std::string, int32_t, double>>;
struct QueryResultEntry { std::map<std::string, FieldValue> fields; };
struct QueryResult { std::vector<QueryResultEntry> query_result; };

inline json::Value toJson(const QueryResult& result) {  // query_result.cpp:10-25
   json::Value arr = json::Value::array();
   for (const auto& entry : result.query_result) {
      json::Value obj = json::Value::object();
      for (const auto& [field, value] : entry.fields) {
         if (!value) obj.set(field, json::Value());
         else if (std::holds_alternative<std::string>(*value)) obj.set(field, json::Value::string(std::get<std::string>(*value)));
         else if (std::holds_alternative<int32_t>(*value)) obj.set(field, json::Value::integer(std::get<int32_t>(*value)));
         else obj.set(field, json::Value::number(std::get<double>(*value)));
      }
      arr.arr.push_back(std::move(obj));
   }
   json::Value out = json::Value::object();
   out.set("queryResult", std::move(arr));
   return out;
}

// The same text as toJson(result).dump(), written straight into the string (no intermediate tree): this is the
// response body of every query.
inline std::string dumpJson(const QueryResult& result) {
   std::string out;
   out.reserve(64 + result.query_result.size() * 96);
   out += "{\"queryResult\":[";
   bool first_entry = true;
   for (const auto& entry : result.query_result) {
      if (!first_entry) out += ',';
      first_entry = false;
      out += '{';
      bool first_field = true;
      for (const auto& [field, value] : entry.fields) {
         if (!first_field) out += ',';
         first_field = false;
         json::Value::writeString(out, field);
         out += ':';
         if (!value) out += "null";
         else if (std::holds_alternative<std::string>(*value)) json::Value::writeString(out, std::get<std::string>(*value));
         else if (std::holds_alternative<int32_t>(*value)) out += std::to_string(std::get<int32_t>(*value));
         else json::Value::writeDouble(out, std::get<double>(*value));
      }
      out += '}';
   }
   out += "]}";
   return out;
}

struct OrderByField { std::string name; bool ascending; };

enum class ActionType { AGGREGATED, MUTATIONS, AA_MUTATIONS };

struct Action {
   ActionType type = ActionType::AGGREGATED;
   std::vector<std::string> sequence_names;  // Mutations
   double min_proportion = 0;                // Mutations
   std::vector<std::string> group_by_fields; // Aggregated
   std::vector<OrderByField> order_by_fields;
   std::optional<uint32_t> limit, offset;

   int alphabetId() const { return type == ActionType::AA_MUTATIONS ? 1 : 0; }

   void validateOrderByFields() const {  // mutations.cpp:167-183, aggregated.cpp:71-88 (the group-by fields exist: parseGroupByFields ran before)
      for (const auto& f : order_by_fields) {
         if (type == ActionType::AGGREGATED) {
            CHECK_SILO_QUERY(f.name == "count" || std::find(group_by_fields.begin(), group_by_fields.end(), f.name) != group_by_fields.end(),
                             "The orderByField '" + f.name + "' cannot be ordered by, as it does not appear in the groupByFields.")
         } else {
            CHECK_SILO_QUERY(f.name == "mutation" || f.name == "proportion" || f.name == "count",
                             "OrderByField " + f.name + " is not contained in the result of this operation.")
         }
      }
   }
   void applySort(QueryResult& result) const {  // action.cpp:37-66
      auto& v = result.query_result;
      auto cmp = [&](const QueryResultEntry& a, const QueryResultEntry& b) {
         for (const auto& f : order_by_fields) {
            if (a.fields.at(f.name) == b.fields.at(f.name)) continue;
            return a.fields.at(f.name) < b.fields.at(f.name) ? f.ascending : !f.ascending;
         }
         return false;
      };
      const size_t end_of_sort = std::min(static_cast<size_t>(limit.value_or(v.size()) + offset.value_or(0UL)), v.size());
      if (!order_by_fields.empty()) {
         if (end_of_sort < v.size()) std::partial_sort(v.begin(), v.begin() + static_cast<int64_t>(end_of_sort), v.end(), cmp);
         else std::sort(v.begin(), v.end(), cmp);
      }
   }
   void applyOffsetAndLimit(QueryResult& result) const {  // action.cpp:68-92
      auto& v = result.query_result;
      size_t end_of_sort = std::min(static_cast<size_t>(limit.value_or(v.size()) + offset.value_or(0UL)), v.size());
      if (offset && *offset >= end_of_sort) { result = {}; return; }
      if (offset && *offset > 0) {
         auto begin = v.begin() + *offset;
         auto end = end_of_sort < v.size() ? v.begin() + static_cast<int64_t>(end_of_sort) : v.end();
         std::copy(begin, end, v.begin());
         end_of_sort -= *offset;
      }
      if (end_of_sort < v.size()) v.resize(end_of_sort);
   }
   // action.cpp:104-117, with `execute` supplied by the caller (GPU engine or CPU oracle)
   template <typename ExecuteFn>
   QueryResult executeAndOrder(ExecuteFn&& execute) const {
      validateOrderByFields();
      QueryResult result = execute();
      if (offset && *offset >= result.query_result.size()) return {};
      applySort(result);
      applyOffsetAndLimit(result);
      return result;
   }
};

struct Query {
   filter::ExpressionPtr filter;
   Action action;
};

inline OrderByField parseOrderByField(const json::Value& j) {  // action.cpp:119-142
   if (j.is_string()) return {j.s, true};
   const std::string message = "The orderByField '" + j.dump() +
      "' must be either a string or an object containing the fields 'field':string and "
      "'order':string, where the value of order is 'ascending' or 'descending'";
   CHECK_SILO_QUERY(j.is_object() && j.contains("field") && j.contains("order") && j["field"].is_string() && j["order"].is_string(), message)
   const std::string order = j["order"].s;
   CHECK_SILO_QUERY(order == "ascending" || order == "descending", message)
   return {j["field"].s, order == "ascending"};
}

inline Action parseAction(const json::Value& j) {  // action.cpp:144-187
   CHECK_SILO_QUERY(j.contains("type"), "The field 'type' is required in any action")
   CHECK_SILO_QUERY(j["type"].is_string(), "The field 'type' in all actions needs to be a string, but is: " + j["type"].dump())
   const std::string type = j["type"].s;
   Action action;
   if (type == "Aggregated") {
      action.type = ActionType::AGGREGATED;
      if (j.contains("groupByFields")) {
         if (!j["groupByFields"].is_array()) throw QueryParseException("The query was not a valid JSON: groupByFields must be an array");
         for (const auto& f : j["groupByFields"].arr) action.group_by_fields.push_back(f.as_string());
      }
   } else if (type == "Mutations" || type == "AminoAcidMutations") {  // mutations.cpp:276-311
      action.type = type == "Mutations" ? ActionType::MUTATIONS : ActionType::AA_MUTATIONS;
      CHECK_SILO_QUERY(!j.contains("sequenceName") || (j["sequenceName"].is_string() || j["sequenceName"].is_array()),
                       "Mutations action can have the field sequenceName of type string or an array of strings, but no other type")
      if (j.contains("sequenceName") && j["sequenceName"].is_array()) {
         for (const auto& c : j["sequenceName"].arr) {
            CHECK_SILO_QUERY(c.is_string(), "The field sequenceName of Mutations action must have type string or an array, if present. Found:" + c.dump())
            action.sequence_names.push_back(c.s);
         }
      } else if (j.contains("sequenceName")) action.sequence_names.push_back(j["sequenceName"].s);
      CHECK_SILO_QUERY(j.contains("minProportion") && j["minProportion"].is_number(),
                       "Mutations action must contain the field minProportion of type number with limits [0.0, 1.0]. Only mutations are returned if the proportion of sequences having this mutation, is at least minProportion")
      action.min_proportion = j["minProportion"].as_double();
      if (action.min_proportion < 0 || action.min_proportion > 1) throw QueryParseException("Invalid proportion: minProportion must be in interval [0.0, 1.0]");
   } else if (type == "Details" || type == "Fasta" || type == "FastaAligned" || type == "Insertions" || type == "AminoAcidInsertions") {
      throw QueryCompilationException("action '" + type + "' is outside the accelerated hot path (SURVEY 8a); use sb_result_export_ids with the reference's host code");
   } else {
      throw QueryParseException(type + " is not a valid action");
   }
   if (j.contains("orderByFields")) {
      if (!j["orderByFields"].is_array()) throw QueryParseException("The query was not a valid JSON: orderByFields must be an array");
      for (const auto& f : j["orderByFields"].arr) action.order_by_fields.push_back(parseOrderByField(f));
   }
   CHECK_SILO_QUERY(!j.contains("limit") || j["limit"].is_number_unsigned(), "If the action contains a limit, it must be a non-negative number")
   CHECK_SILO_QUERY(!j.contains("offset") || j["offset"].is_number_unsigned(), "If the action contains an offset, it must be a non-negative number")
   if (j.contains("limit")) action.limit = static_cast<uint32_t>(j["limit"].u);
   if (j.contains("offset")) action.offset = static_cast<uint32_t>(j["offset"].u);
   return action;
}

inline Query parseQuery(const std::string& query_string) {  // query.cpp:13-28
   json::Value j;
   try {
      j = json::parse(query_string);
   } catch (const json::ParseError& e) {
      throw QueryParseException("The query was not a valid JSON: " + std::string(e.what()));
   }
   if (!j.contains("filterExpression") || !j["filterExpression"].is_object() || !j.contains("action") || !j["action"].is_object()) {
      throw QueryParseException("Query json must contain filterExpression and action.");
   }
   Query q;
   try {
      q.filter = filter::parseExpression(j["filterExpression"]);
      q.action = parseAction(j["action"]);
   } catch (const json::ParseError& e) {
      throw QueryParseException("The query was not a valid JSON: " + std::string(e.what()));
   }
   return q;
}

// mutations.cpp:185-232 (M5): per position total over VALID_MUTATION_SYMBOLS, threshold, emit rows.
// `counts` is the [Σ][L] table (row = enum value of the symbol) produced by the action kernels.
inline void addMutationsToOutput(
   const std::string& sequence_name, const Alphabet& alphabet, const std::vector<Symbol>& reference_sequence,
   const uint32_t* counts, double min_proportion, std::vector<QueryResultEntry>& output
) {
   const size_t length = reference_sequence.size();
   for (size_t pos = 0; pos < length; ++pos) {
      uint32_t total = 0;
      for (Symbol s : alphabet.valid_mutation_symbols) total += counts[size_t(s) * length + pos];
      if (total == 0) continue;
      const uint32_t threshold_count = min_proportion == 0 ? 0 : static_cast<uint32_t>(std::ceil(static_cast<double>(total) * min_proportion) - 1);
      const Symbol ref = reference_sequence[pos];
      for (Symbol s : alphabet.valid_mutation_symbols) {
         if (s == ref) continue;
         const uint32_t count = counts[size_t(s) * length + pos];
         if (count > threshold_count) {
            const double proportion = static_cast<double>(count) / static_cast<double>(total);
            QueryResultEntry entry;
            entry.fields["mutation"] = alphabet.symbolToChar(ref) + std::to_string(pos + 1) + alphabet.symbolToChar(s);
            entry.fields["sequenceName"] = sequence_name;
            entry.fields["proportion"] = proportion;
            entry.fields["count"] = static_cast<int32_t>(count);
            output.push_back(std::move(entry));
         }
      }
   }
}

}  // namespace silo
