// Physical operator tree (host side). One tagged node type instead of the reference's class
// hierarchy; same 11 operator kinds, same constructor checks, same negate()/copy()/toString().
// Mirrors include/silo/query_engine/operators/operator.h:10-39 and operators/*.cpp.
// evaluate() is NOT here: the product lowers the tree to an sb_plan and runs it on the GPU
// (engine.cpp); the CPU restatement of evaluate() lives in oracle/ (test infrastructure).
#pragma once
#include <functional>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

#include "storage.hpp"

namespace silo {

struct QueryParseException : std::runtime_error { using std::runtime_error::runtime_error; };
struct QueryCompilationException : std::runtime_error { using std::runtime_error::runtime_error; };

#define CHECK_SILO_QUERY(cond, msg) if (!(cond)) { throw ::silo::QueryParseException(msg); }

namespace operators {

enum Type { EMPTY, FULL, INDEX_SCAN, INTERSECTION, COMPLEMENT, RANGE_SELECTION, SELECTION, BITMAP_SELECTION, THRESHOLD, UNION, BITMAP_PRODUCER };
enum class Comparator { EQUALS, LESS, HIGHER, LESS_OR_EQUALS, HIGHER_OR_EQUALS, NOT_EQUALS };
enum class ColumnKind { INT32, FLOAT64, DATE, STRING };

inline std::string displayComparator(Comparator c) {  // selection.cpp:45-61
   switch (c) {
      case Comparator::EQUALS: return "=";
      case Comparator::NOT_EQUALS: return "!=";
      case Comparator::LESS: return "<";
      case Comparator::HIGHER: return ">";
      case Comparator::LESS_OR_EQUALS: return "<=";
      case Comparator::HIGHER_OR_EQUALS: return ">=";
   }
   throw std::runtime_error("found unhandled comparator");
}

// CompareToValueSelection<T> (selection.cpp:134-193), T in {int32_t, double, Date, string}
struct Predicate {
   ColumnKind kind = ColumnKind::INT32;
   const void* column = nullptr;  // std::vector<T>* of the partition's column
   Comparator comparator = Comparator::EQUALS;
   int32_t int_value = 0;
   double float_value = 0;
   uint32_t date_value = 0;
   std::string string_value;

   Predicate negate() const {  // selection.cpp:166-191
      Predicate p = *this;
      switch (comparator) {
         case Comparator::EQUALS: p.comparator = Comparator::NOT_EQUALS; break;
         case Comparator::NOT_EQUALS: p.comparator = Comparator::EQUALS; break;
         case Comparator::LESS: p.comparator = Comparator::HIGHER_OR_EQUALS; break;
         case Comparator::HIGHER_OR_EQUALS: p.comparator = Comparator::LESS; break;
         case Comparator::HIGHER: p.comparator = Comparator::LESS_OR_EQUALS; break;
         case Comparator::LESS_OR_EQUALS: p.comparator = Comparator::HIGHER; break;
      }
      return p;
   }
   template <typename T>
   static bool compare(Comparator c, const T& a, const T& b) {
      switch (c) {
         case Comparator::EQUALS: return a == b;
         case Comparator::NOT_EQUALS: return a != b;
         case Comparator::LESS: return a < b;
         case Comparator::HIGHER_OR_EQUALS: return a >= b;
         case Comparator::HIGHER: return a > b;
         case Comparator::LESS_OR_EQUALS: return a <= b;
      }
      return false;
   }
   bool match(uint32_t row) const {  // selection.cpp:144-164 (host use: string predicates only)
      switch (kind) {
         case ColumnKind::INT32: return compare(comparator, (*static_cast<const std::vector<int32_t>*>(column))[row], int_value);
         case ColumnKind::FLOAT64: return compare(comparator, (*static_cast<const std::vector<double>*>(column))[row], float_value);
         case ColumnKind::DATE: return compare(comparator, (*static_cast<const std::vector<uint32_t>*>(column))[row], date_value);
         case ColumnKind::STRING: return compare(comparator, (*static_cast<const std::vector<std::string>*>(column))[row], string_value);
      }
      return false;
   }
   std::string toString() const {
      switch (kind) {
         case ColumnKind::INT32: return "$int " + displayComparator(comparator) + " " + std::to_string(int_value);
         case ColumnKind::FLOAT64: return "$double " + displayComparator(comparator) + " " + std::to_string(float_value);
         case ColumnKind::DATE: return "$date " + displayComparator(comparator) + " " + std::to_string(date_value);
         case ColumnKind::STRING: return "$string " + displayComparator(comparator) + " " + string_value;
      }
      return "";
   }
};

struct Range { uint32_t start, end; };  // range_selection.h: [start, end)

struct Operator;
using OperatorPtr = std::unique_ptr<Operator>;
using OperatorVector = std::vector<OperatorPtr>;

struct Operator {
   Type op_type = EMPTY;
   uint32_t row_count = 0;
   OperatorVector children;          // Intersection / Threshold: non-negated; Union: all; Complement / Selection: [0]
   OperatorVector negated_children;  // Intersection / Threshold
   // INDEX_SCAN
   const Bitmap* bitmap = nullptr;
   // BITMAP_SELECTION: missing_symbol_bitmaps of `store`, contains / not contains `value`
   const SequenceStorePartition* store = nullptr;
   bool contains = true;
   uint32_t value = 0;
   // RANGE_SELECTION / SELECTION
   std::vector<Range> ranges;
   std::vector<Predicate> predicates;
   // THRESHOLD
   uint32_t number_of_matchers = 0;
   bool match_exactly = false;
   // BITMAP_PRODUCER
   std::function<Bitmap()> producer;

   Type type() const { return op_type; }
   OperatorPtr copy() const;
   OperatorPtr negate() const;
   std::string toString() const;
};

inline OperatorPtr make(Type t, uint32_t row_count) {
   auto op = std::make_unique<Operator>();
   op->op_type = t;
   op->row_count = row_count;
   return op;
}
inline OperatorPtr makeEmpty(uint32_t rc) { return make(EMPTY, rc); }
inline OperatorPtr makeFull(uint32_t rc) { return make(FULL, rc); }
inline OperatorPtr makeIndexScan(const Bitmap* bm, uint32_t rc) { auto op = make(INDEX_SCAN, rc); op->bitmap = bm; return op; }
inline OperatorPtr makeComplement(OperatorPtr child, uint32_t rc) { auto op = make(COMPLEMENT, rc); op->children.push_back(std::move(child)); return op; }
inline OperatorPtr makeUnion(OperatorVector&& children, uint32_t rc) { auto op = make(UNION, rc); op->children = std::move(children); return op; }
inline OperatorPtr makeIntersection(OperatorVector&& children, OperatorVector&& negated, uint32_t rc) {
   auto op = make(INTERSECTION, rc);
   op->children = std::move(children);
   op->negated_children = std::move(negated);
   if (op->children.empty()) {  // intersection.cpp:24-31
      throw QueryCompilationException("Compilation bug: Intersection without non-negated children is not allowed. Should be compiled as a union.");
   }
   if (op->children.size() + op->negated_children.size() < 2) {  // intersection.cpp:33-38
      throw QueryCompilationException("Compilation bug: Intersection needs at least two children.");
   }
   return op;
}
inline OperatorPtr makeThreshold(OperatorVector&& children, OperatorVector&& negated, uint32_t n, bool exact, uint32_t rc) {
   auto op = make(THRESHOLD, rc);
   op->children = std::move(children);
   op->negated_children = std::move(negated);
   op->number_of_matchers = n;
   op->match_exactly = exact;
   if (n >= op->children.size() + op->negated_children.size()) {  // threshold.cpp:28-33
      throw QueryCompilationException("Compilation Error: number_of_matchers must be less than the number of children of a threshold expression");
   }
   if (n == 0) {  // threshold.cpp:34-38
      throw QueryCompilationException("Compilation Error: number_of_matchers must be greater than zero");
   }
   return op;
}
inline OperatorPtr makeRangeSelection(std::vector<Range>&& ranges, uint32_t rc) { auto op = make(RANGE_SELECTION, rc); op->ranges = std::move(ranges); return op; }
inline OperatorPtr makeSelection(OperatorPtr child, std::vector<Predicate>&& predicates, uint32_t rc) {
   auto op = make(SELECTION, rc);
   if (child) op->children.push_back(std::move(child));
   op->predicates = std::move(predicates);
   return op;
}
inline OperatorPtr makeBitmapSelection(const SequenceStorePartition* store, uint32_t rc, bool contains, uint32_t value) {
   auto op = make(BITMAP_SELECTION, rc);
   op->store = store;
   op->contains = contains;
   op->value = value;
   return op;
}
inline OperatorPtr makeBitmapProducer(std::function<Bitmap()> producer, uint32_t rc) { auto op = make(BITMAP_PRODUCER, rc); op->producer = std::move(producer); return op; }

inline OperatorPtr Operator::copy() const {
   auto op = make(op_type, row_count);
   for (const auto& c : children) op->children.push_back(c->copy());
   for (const auto& c : negated_children) op->negated_children.push_back(c->copy());
   op->bitmap = bitmap; op->store = store; op->contains = contains; op->value = value;
   op->ranges = ranges; op->predicates = predicates;
   op->number_of_matchers = number_of_matchers; op->match_exactly = match_exactly;
   op->producer = producer;
   return op;
}

inline OperatorPtr Operator::negate() const {
   switch (op_type) {
      case EMPTY: return makeFull(row_count);                      // empty.cpp:32-34
      case FULL: return makeEmpty(row_count);                      // full.cpp:32-34
      case COMPLEMENT: return children[0]->copy();                 // complement.cpp:60-62
      case RANGE_SELECTION: {                                      // range_selection.cpp:54-67
         std::vector<Range> out;
         uint32_t last_to = 0;
         for (const auto& r : ranges) {
            if (last_to != r.start) out.push_back({last_to, r.start});
            last_to = r.end;
         }
         if (last_to != row_count) out.push_back({last_to, row_count});
         return makeRangeSelection(std::move(out), row_count);
      }
      case BITMAP_SELECTION: {                                     // bitmap_selection.cpp:58-69
         auto op = copy();
         op->contains = !contains;
         return op;
      }
      case SELECTION:                                              // selection.cpp:127-132
         if (children.empty() && predicates.size() == 1) {
            return makeSelection(nullptr, {predicates[0].negate()}, row_count);
         }
         return makeComplement(copy(), row_count);
      default:                                                     // index_scan.cpp:36-38 etc.
         return makeComplement(copy(), row_count);
   }
}

inline std::string Operator::toString() const {
   switch (op_type) {
      case EMPTY: return "Empty";
      case FULL: return "Full";
      case INDEX_SCAN: return "IndexScan(Cardinality: " + std::to_string(bitmap->cardinality()) + ")";
      case COMPLEMENT: return "!" + children[0]->toString();
      case BITMAP_SELECTION: return "BitmapSelection";
      case BITMAP_PRODUCER: return "BitmapProducer";
      case INTERSECTION: {
         std::string res = "(" + children[0]->toString();
         for (size_t i = 1; i < children.size(); ++i) res += " & " + children[i]->toString();
         for (const auto& c : negated_children) res += " &! " + c->toString();
         return res + ")";
      }
      case UNION: {
         std::string res = "(" + (children.empty() ? std::string() : children[0]->toString());
         for (size_t i = 1; i < children.size(); ++i) res += " | " + children[i]->toString();
         return res + ")";
      }
      case THRESHOLD: {
         std::string res = match_exactly ? "=" : ">=";
         for (const auto& c : children) res += ", " + c->toString();
         for (const auto& c : negated_children) res += ", ! " + c->toString();
         return res + ")";
      }
      case RANGE_SELECTION: {
         std::string res = "RangeSelection(";
         for (size_t i = 0; i < ranges.size(); ++i) res += (i ? ", " : "") + std::to_string(ranges[i].start) + "-" + std::to_string(ranges[i].end);
         return res + ")";
      }
      case SELECTION: {
         std::string res = "Select[";
         for (size_t i = 0; i < predicates.size(); ++i) res += (i ? "," : "") + predicates[i].toString();
         return res + "]()";
      }
   }
   return "";
}

}  // namespace operators
}  // namespace silo
