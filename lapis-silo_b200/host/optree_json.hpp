// Builds a stand-alone Operator tree from a JSON description with literal bitmaps / columns.
// This is how the reference's operator unit tests construct their inputs
// (e.g. threshold.test.cpp:16-28 wraps roaring::Roaring({..}) literals in IndexScan children);
// tests/ use it to run the same known-answer cases through the GPU path and the oracle.
#pragma once
#include <deque>

#include "json.hpp"
#include "operators.hpp"

namespace silo {

struct OpTreeHolder {
   std::deque<Bitmap> bitmaps;
   std::deque<std::vector<int32_t>> int_columns;
   std::deque<std::vector<double>> float_columns;
   std::deque<std::vector<uint32_t>> date_columns;
   std::deque<std::vector<std::string>> string_columns;
   std::deque<SequenceStorePartition> stores;
   operators::OperatorPtr root;
};

namespace detail {

inline Bitmap bitmapFromJsonIds(const json::Value& ids) {
   Bitmap bm;
   for (const auto& v : ids.arr) bm.add(static_cast<uint32_t>(v.as_int()));
   return bm;
}

inline operators::Comparator parseComparator(const std::string& s) {
   using operators::Comparator;
   if (s == "==") return Comparator::EQUALS;
   if (s == "!=") return Comparator::NOT_EQUALS;
   if (s == "<") return Comparator::LESS;
   if (s == "<=") return Comparator::LESS_OR_EQUALS;
   if (s == ">") return Comparator::HIGHER;
   if (s == ">=") return Comparator::HIGHER_OR_EQUALS;
   throw std::runtime_error("bad comparator " + s);
}

inline operators::OperatorPtr buildOpTree(const json::Value& j, uint32_t inherited_row_count, OpTreeHolder& holder);

inline operators::OperatorPtr buildOpTreeNode(const json::Value& j, uint32_t inherited_row_count, OpTreeHolder& holder) {
   using namespace operators;
   const uint32_t rc = j.contains("row_count") ? static_cast<uint32_t>(j["row_count"].as_int()) : inherited_row_count;
   const std::string op = j["op"].as_string();
   auto children = [&](const char* key) {
      OperatorVector out;
      if (j.contains(key)) for (const auto& c : j[key].arr) out.push_back(buildOpTree(c, rc, holder));
      return out;
   };
   if (op == "empty") return makeEmpty(rc);
   if (op == "full") return makeFull(rc);
   if (op == "index_scan") {
      holder.bitmaps.push_back(bitmapFromJsonIds(j["ids"]));
      if (j.contains("run_optimize") && j["run_optimize"].b) holder.bitmaps.back().runOptimize();
      return makeIndexScan(&holder.bitmaps.back(), rc);
   }
   if (op == "bitmap_producer") {
      Bitmap bm = bitmapFromJsonIds(j["ids"]);
      return makeBitmapProducer([bm]() { return bm; }, rc);
   }
   if (op == "intersection") return makeIntersection(children("children"), children("negated"), rc);
   if (op == "union") return makeUnion(children("children"), rc);
   if (op == "complement") return makeComplement(buildOpTree(j["child"], rc, holder), rc);
   if (op == "threshold") return makeThreshold(children("children"), children("negated"), static_cast<uint32_t>(j["n"].as_int()), j["exact"].b, rc);
   if (op == "range_selection") {
      std::vector<Range> ranges;
      for (const auto& r : j["ranges"].arr) ranges.push_back({static_cast<uint32_t>(r.arr[0].as_int()), static_cast<uint32_t>(r.arr[1].as_int())});
      return makeRangeSelection(std::move(ranges), rc);
   }
   if (op == "bitmap_selection") {
      holder.stores.emplace_back();
      SequenceStorePartition& store = holder.stores.back();
      for (const auto& ids : j["bitmaps"].arr) { store.missing_symbol_bitmaps.push_back(bitmapFromJsonIds(ids)); store.missing_symbol_bitmaps.back().runOptimize(); }
      store.sequence_count = static_cast<uint32_t>(store.missing_symbol_bitmaps.size());
      const uint32_t row_count = j.contains("row_count") ? rc : store.sequence_count;
      return makeBitmapSelection(&store, row_count, j["contains"].b, static_cast<uint32_t>(j["value"].as_int()));
   }
   if (op == "selection") {
      std::vector<Predicate> predicates;
      for (const auto& pj : j["predicates"].arr) {
         Predicate p;
         p.comparator = parseComparator(pj["cmp"].as_string());
         const std::string kind = pj["kind"].as_string();
         if (kind == "int") {
            holder.int_columns.emplace_back();
            for (const auto& v : pj["column"].arr) holder.int_columns.back().push_back(static_cast<int32_t>(v.as_int()));
            p.kind = ColumnKind::INT32; p.column = &holder.int_columns.back(); p.int_value = static_cast<int32_t>(pj["value"].as_int());
         } else if (kind == "float") {
            holder.float_columns.emplace_back();
            for (const auto& v : pj["column"].arr) holder.float_columns.back().push_back(v.is_null() ? std::nan("") : v.as_double());
            p.kind = ColumnKind::FLOAT64; p.column = &holder.float_columns.back(); p.float_value = pj["value"].is_null() ? std::nan("") : pj["value"].as_double();
         } else if (kind == "date") {
            holder.date_columns.emplace_back();
            for (const auto& v : pj["column"].arr) holder.date_columns.back().push_back(static_cast<uint32_t>(v.as_int()));
            p.kind = ColumnKind::DATE; p.column = &holder.date_columns.back(); p.date_value = static_cast<uint32_t>(pj["value"].as_int());
         } else {
            holder.string_columns.emplace_back();
            for (const auto& v : pj["column"].arr) holder.string_columns.back().push_back(v.as_string());
            p.kind = ColumnKind::STRING; p.column = &holder.string_columns.back(); p.string_value = pj["value"].as_string();
         }
         predicates.push_back(std::move(p));
      }
      OperatorPtr child;
      if (j.contains("child")) child = buildOpTree(j["child"], rc, holder);
      return makeSelection(std::move(child), std::move(predicates), rc);
   }
   throw std::runtime_error("unknown op '" + op + "'");
}

// "negate": true applies Operator::negate() to the node (the reference tests call under_test.negate())
inline operators::OperatorPtr buildOpTree(const json::Value& j, uint32_t inherited_row_count, OpTreeHolder& holder) {
   auto op = buildOpTreeNode(j, inherited_row_count, holder);
   if (j.contains("negate") && j["negate"].b) return op->negate();
   return op;
}

}  // namespace detail

inline std::unique_ptr<OpTreeHolder> opTreeFromJson(const std::string& text) {
   auto holder = std::make_unique<OpTreeHolder>();
   json::Value j = json::parse(text);
   holder->root = detail::buildOpTree(j, j.contains("row_count") ? static_cast<uint32_t>(j["row_count"].as_int()) : 0, *holder);
   return holder;
}

}  // namespace silo
