// Filter expression AST: JSON -> Expression -> (per partition) Operator tree.
// Kept host code of the reference, restated from src/silo/query_engine/filter_expressions/*.cpp
// (21 node types; dispatch expression.cpp:49-102). All quirks of SURVEY Appendix B.1 are kept.
#pragma once
#include <algorithm>
#include <cctype>
#include <memory>
#include <optional>
#include <regex>
#include <string>
#include <vector>

#include "json.hpp"
#include "operators.hpp"

namespace silo::filter {

using operators::Operator;
using operators::OperatorPtr;
using operators::OperatorVector;

enum class AmbiguityMode { NONE, UPPER_BOUND, LOWER_BOUND };
inline AmbiguityMode invertMode(AmbiguityMode m) {  // expression.cpp:38-46
   if (m == AmbiguityMode::UPPER_BOUND) return AmbiguityMode::LOWER_BOUND;
   if (m == AmbiguityMode::LOWER_BOUND) return AmbiguityMode::UPPER_BOUND;
   return m;
}

enum class ExprType {
   TRUE_, FALSE_, AND, OR, NOF, NOT, DATE_BETWEEN, NUC_EQUALS, HAS_NUC_MUTATION, AA_EQUALS, HAS_AA_MUTATION,
   PANGO_LINEAGE, STRING_EQUALS, INT_EQUALS, INT_BETWEEN, FLOAT_EQUALS, FLOAT_BETWEEN, MAYBE, EXACT,
   NUC_INSERTION_CONTAINS, AA_INSERTION_CONTAINS
};

struct Expression;
using ExpressionPtr = std::unique_ptr<Expression>;

struct Expression {
   ExprType type = ExprType::TRUE_;
   std::vector<ExpressionPtr> children;  // And / Or / N-Of; Not / Maybe / Exact: [0]
   int number_of_matchers = 0;
   bool match_exactly = false;
   std::optional<std::string> sequence_name;
   uint32_t position = 0;                 // 0-based
   std::optional<Symbol> symbol;          // nullopt = '.' (reference symbol)
   std::string column;
   std::vector<std::string> column_names; // InsertionContains
   std::string value;
   bool include_sublineages = false;
   std::optional<uint32_t> date_from, date_to;
   int32_t int_value = 0;
   std::optional<int32_t> int_from, int_to;
   double float_value = 0;
   std::optional<double> float_from, float_to;

   OperatorPtr compile(const Database& db, const DatabasePartition& part, AmbiguityMode mode) const;
};

ExpressionPtr parseExpression(const json::Value& j);

namespace detail {

inline ExpressionPtr makeSymbolEquals(bool aa, const std::optional<std::string>& seq, uint32_t pos, std::optional<Symbol> sym) {
   auto e = std::make_unique<Expression>();
   e->type = aa ? ExprType::AA_EQUALS : ExprType::NUC_EQUALS;
   e->sequence_name = seq;
   e->position = pos;
   e->symbol = sym;
   return e;
}
inline ExpressionPtr makeNot(ExpressionPtr child) {
   auto e = std::make_unique<Expression>();
   e->type = ExprType::NOT;
   e->children.push_back(std::move(child));
   return e;
}
inline ExpressionPtr makeNary(ExprType t, std::vector<ExpressionPtr>&& children) {
   auto e = std::make_unique<Expression>();
   e->type = t;
   e->children = std::move(children);
   return e;
}

// and.cpp:101-172. Quirk Q7: the reference appends to the vector it is range-iterating (UB); the
// evident intent (also visit the appended operators) is implemented with an index loop.
inline OperatorPtr compileAnd(const Expression& self, const Database& db, const DatabasePartition& part, AmbiguityMode mode) {
   const uint32_t rc = part.sequence_count;
   OperatorVector all;
   for (const auto& c : self.children) all.push_back(c->compile(db, part, mode));
   OperatorVector non_negated, negated;
   std::vector<operators::Predicate> predicates;
   bool found_empty = false;
   for (size_t k = 0; k < all.size(); ++k) {
      OperatorPtr& child = all[k];
      if (child->type() == operators::FULL) continue;
      if (child->type() == operators::EMPTY) { found_empty = true; break; }
      if (child->type() == operators::INTERSECTION) {
         for (auto& c : child->children) non_negated.push_back(std::move(c));
         for (auto& c : child->negated_children) negated.push_back(std::move(c));
      } else if (child->type() == operators::COMPLEMENT) {
         negated.push_back(child->negate());
      } else if (child->type() == operators::SELECTION) {
         for (auto& p : child->predicates) predicates.push_back(p);
         if (!child->children.empty()) {
            OperatorPtr sub = std::move(child->children[0]);
            all.push_back(std::move(sub));  // invalidates `child`; not used afterwards
         }
      } else {
         non_negated.push_back(std::move(child));
      }
   }
   if (found_empty) return operators::makeEmpty(rc);  // and.cpp:124-131 then :174-199 (single child)
   // and.cpp:174-227
   if (non_negated.empty() && negated.empty()) {
      if (predicates.empty()) return operators::makeFull(rc);
      return operators::makeSelection(nullptr, std::move(predicates), rc);
   }
   OperatorPtr index_op;
   if (non_negated.size() == 1 && negated.empty()) index_op = std::move(non_negated[0]);
   else if (negated.size() == 1 && non_negated.empty()) index_op = operators::makeComplement(std::move(negated[0]), rc);
   else if (non_negated.empty()) index_op = operators::makeComplement(operators::makeUnion(std::move(negated), rc), rc);
   else index_op = operators::makeIntersection(std::move(non_negated), std::move(negated), rc);
   if (predicates.empty()) return index_op;
   return operators::makeSelection(std::move(index_op), std::move(predicates), rc);
}

// complement.cpp:22-40
inline OperatorPtr complementFromDeMorgan(OperatorVector&& disjunction, uint32_t rc) {
   OperatorVector non_negated, negated;
   for (auto& c : disjunction) {
      if (c->type() == operators::COMPLEMENT) negated.push_back(c->negate());
      else non_negated.push_back(std::move(c));
   }
   return operators::makeComplement(operators::makeIntersection(std::move(negated), std::move(non_negated), rc), rc);
}

// or.cpp:41-94
inline OperatorPtr compileOr(const Expression& self, const Database& db, const DatabasePartition& part, AmbiguityMode mode) {
   const uint32_t rc = part.sequence_count;
   OperatorVector all;
   for (const auto& c : self.children) all.push_back(c->compile(db, part, mode));
   OperatorVector filtered;
   for (auto& child : all) {
      if (child->type() == operators::EMPTY) continue;
      if (child->type() == operators::FULL) return operators::makeFull(rc);
      if (child->type() == operators::UNION) { for (auto& c : child->children) filtered.push_back(std::move(c)); }
      else filtered.push_back(std::move(child));
   }
   if (filtered.empty()) return operators::makeEmpty(rc);
   if (filtered.size() == 1) return std::move(filtered[0]);
   if (std::any_of(filtered.begin(), filtered.end(), [](const OperatorPtr& c) { return c->type() == operators::COMPLEMENT; })) {
      return complementFromDeMorgan(std::move(filtered), rc);
   }
   return operators::makeUnion(std::move(filtered), rc);
}

// nof.cpp:35-154
inline OperatorPtr nofToOperator(int n, OperatorVector&& non_negated, OperatorVector&& negated, bool exact, uint32_t rc) {
   const int k = static_cast<int>(non_negated.size() + negated.size());
   // handleTrivialCases (nof.cpp:35-88)
   if (n > k) return operators::makeEmpty(rc);
   if (n < 0) return exact ? operators::makeEmpty(rc) : operators::makeFull(rc);
   if (n == 0) {
      if (!exact) return operators::makeFull(rc);
      if (k == 0) return operators::makeFull(rc);
      if (k == 1) {
         if (non_negated.empty()) return std::move(negated[0]);
         return operators::makeComplement(std::move(non_negated[0]), rc);
      }
      if (negated.empty()) return operators::makeComplement(operators::makeUnion(std::move(non_negated), rc), rc);
      return operators::makeIntersection(std::move(negated), std::move(non_negated), rc);
   }
   if (n == 1 && k == 1) {
      if (negated.empty()) return std::move(non_negated[0]);
      return operators::makeComplement(std::move(negated[0]), rc);
   }
   if (n == k) {  // handleAndCase (nof.cpp:90-103)
      if (non_negated.empty()) return operators::makeComplement(operators::makeUnion(std::move(negated), rc), rc);
      return operators::makeIntersection(std::move(non_negated), std::move(negated), rc);
   }
   if (n == 1 && !exact) {  // handleOrCase (nof.cpp:105-118)
      if (negated.empty()) return operators::makeUnion(std::move(non_negated), rc);
      return operators::makeComplement(operators::makeIntersection(std::move(negated), std::move(non_negated), rc), rc);
   }
   return operators::makeThreshold(std::move(non_negated), std::move(negated), static_cast<uint32_t>(n), exact, rc);
}

struct NofChildren { OperatorVector non_negated, negated; int n; };
inline NofChildren nofMapChildren(const Expression& self, const Database& db, const DatabasePartition& part, AmbiguityMode mode) {  // nof.cpp:184-218
   NofChildren out;
   out.n = self.number_of_matchers;
   for (const auto& c : self.children) {
      auto op = c->compile(db, part, mode);
      if (op->type() == operators::EMPTY) continue;
      if (op->type() == operators::FULL) out.n--;
      else if (op->type() == operators::COMPLEMENT) out.negated.push_back(op->negate());
      else out.non_negated.push_back(std::move(op));
   }
   return out;
}
inline OperatorPtr compileNof(const Expression& self, const Database& db, const DatabasePartition& part, AmbiguityMode mode) {  // nof.cpp:220-280
   const uint32_t rc = part.sequence_count;
   auto mapped = nofMapChildren(self, db, part, mode);
   if (mode != AmbiguityMode::NONE && self.match_exactly && self.number_of_matchers < static_cast<int>(self.children.size())) {
      OperatorVector at_least_k, at_least_k_plus_one;
      { auto m = nofMapChildren(self, db, part, mode); at_least_k.push_back(nofToOperator(m.n, std::move(m.non_negated), std::move(m.negated), false, rc)); }
      { auto m = nofMapChildren(self, db, part, mode); at_least_k_plus_one.push_back(nofToOperator(m.n + 1, std::move(m.non_negated), std::move(m.negated), false, rc)); }
      return nofToOperator(2, std::move(at_least_k), std::move(at_least_k_plus_one), false, rc);
   }
   return nofToOperator(mapped.n, std::move(mapped.non_negated), std::move(mapped.negated), self.match_exactly, rc);
}

// nucleotide_symbol_equals.cpp:94-189 and aa_symbol_equals.cpp:41-92
inline OperatorPtr compileSymbolEquals(const Expression& self, bool is_aa, const Database& db, const DatabasePartition& part, AmbiguityMode mode) {
   const uint32_t rc = part.sequence_count;
   const Alphabet& alphabet = is_aa ? aminoAcidAlphabet() : nucleotideAlphabet();
   std::string name;
   const SequenceStorePartition* store = nullptr;
   if (is_aa) {
      name = *self.sequence_name;
      auto it = part.aa_sequences.find(name);
      if (it == part.aa_sequences.end()) throw std::out_of_range("map::at");  // aa_symbol_equals.cpp:46 uses .at()
      store = &it->second;
   } else {
      name = self.sequence_name.value_or(db.config.default_nucleotide_sequence);
      CHECK_SILO_QUERY(db.nuc_reference.count(name), "Database does not contain the nucleotide sequence with name: '" + name + "'")
      store = &part.nuc_sequences.at(name);
   }
   if (self.position >= store->length()) {
      throw QueryParseException(
         std::string(is_aa ? "AminoAcidEquals" : "NucleotideEquals") + " position is out of bounds '" +
         std::to_string(self.position + 1) + "' > '" + std::to_string(store->length()) + "'");
   }
   const Symbol symbol = self.symbol.value_or((*store->reference_sequence)[self.position]);
   if (!is_aa && mode == AmbiguityMode::UPPER_BOUND) {
      std::vector<ExpressionPtr> filters;
      for (Symbol s : ambiguityNucSymbols(symbol)) filters.push_back(makeSymbolEquals(false, name, self.position, s));
      return makeNary(ExprType::OR, std::move(filters))->compile(db, part, AmbiguityMode::NONE);
   }
   if (symbol == alphabet.missing) {
      return operators::makeBitmapSelection(store, static_cast<uint32_t>(store->missing_symbol_bitmaps.size()), true, self.position);
   }
   if (store->isSymbolFlipped(self.position, symbol)) {
      return operators::makeComplement(operators::makeIndexScan(store->getBitmap(self.position, symbol), rc), rc);
   }
   if (store->isSymbolDeleted(self.position, symbol)) {
      // Quirks Q1 / Q4: std::remove without erase -> the vector keeps its length, the tail keeps
      // the old last element. If `symbol` IS the last element of SYMBOLS (amino acid STOP) the
      // expansion still contains ¬Equals(symbol) and the reference recurses without bound; we
      // detect that and raise instead (outside the parity domain).
      std::vector<Symbol> symbols = alphabet.iteration_order;
      (void)std::remove(symbols.begin(), symbols.end(), symbol);
      if (symbols.back() == symbol) {
         throw QueryCompilationException("Equals on the deleted symbol that is last in SYMBOLS recurses without bound in the reference (SURVEY Q4)");
      }
      std::vector<ExpressionPtr> filters;
      for (Symbol s : symbols) filters.push_back(makeNot(makeSymbolEquals(is_aa, name, self.position, s)));
      return makeNary(ExprType::AND, std::move(filters))->compile(db, part, AmbiguityMode::NONE);
   }
   return operators::makeIndexScan(store->getBitmap(self.position, symbol), rc);
}

// has_mutation.cpp:35-78 (quirk Q2) and has_aa_mutation.cpp:33-63 (quirk Q3)
inline OperatorPtr compileHasMutation(const Expression& self, bool is_aa, const Database& db, const DatabasePartition& part, AmbiguityMode mode) {
   std::string name;
   Symbol ref_symbol;
   if (is_aa) {
      name = *self.sequence_name;
      ref_symbol = db.aa_reference.at(name).at(self.position);
   } else {
      name = self.sequence_name.value_or(db.config.default_nucleotide_sequence);
      CHECK_SILO_QUERY(db.nuc_reference.count(name), "Database does not contain the nucleotide sequence with name: '" + name + "'")
      ref_symbol = db.nuc_reference.at(name).at(self.position);
   }
   if (mode == AmbiguityMode::UPPER_BOUND) {
      return makeNot(makeSymbolEquals(is_aa, name, self.position, ref_symbol))->compile(db, part, AmbiguityMode::NONE);
   }
   std::vector<Symbol> symbols;
   if (is_aa) {
      symbols = aminoAcidAlphabet().iteration_order;
      (void)std::remove(symbols.begin(), symbols.end(), Symbol(aa::X));
      (void)std::remove(symbols.begin(), symbols.end(), ref_symbol);
   } else {
      symbols = {nuc::A, nuc::C, nuc::G, nuc::T};
      (void)std::remove(symbols.begin(), symbols.end(), ref_symbol);
   }
   std::vector<ExpressionPtr> filters;
   for (Symbol s : symbols) filters.push_back(makeSymbolEquals(is_aa, name, self.position, s));
   return makeNary(ExprType::OR, std::move(filters))->compile(db, part, AmbiguityMode::NONE);
}

// date_between.cpp:46-99 (quirk Q14)
inline OperatorPtr compileDateBetween(const Expression& self, const DatabasePartition& part) {
   const uint32_t rc = part.sequence_count;
   const auto& column = part.columns.date_columns.at(self.column);
   if (!column.is_sorted) {
      std::vector<operators::Predicate> predicates(2);
      predicates[0].kind = operators::ColumnKind::DATE; predicates[0].column = &column.values;
      predicates[0].comparator = operators::Comparator::HIGHER_OR_EQUALS; predicates[0].date_value = self.date_from.value_or(1);
      predicates[1].kind = operators::ColumnKind::DATE; predicates[1].column = &column.values;
      predicates[1].comparator = operators::Comparator::LESS; predicates[1].date_value = self.date_to.value_or(UINT32_MAX);
      return operators::makeSelection(nullptr, std::move(predicates), rc);
   }
   std::vector<operators::Range> ranges;
   const uint32_t* base = column.values.data();
   for (const auto& chunk : part.chunks) {
      const uint32_t* begin = base + chunk.offset;
      const uint32_t* end = base + chunk.offset + chunk.size;
      const uint32_t* lower = std::lower_bound(begin, end, self.date_from.value_or(1));
      const uint32_t* upper = self.date_to.has_value() ? std::upper_bound(begin, end, *self.date_to) : end;
      ranges.push_back({static_cast<uint32_t>(lower - base), static_cast<uint32_t>(upper - base)});
   }
   return operators::makeRangeSelection(std::move(ranges), rc);
}

// pango_lineage_filter.cpp:37-59 (quirk Q13)
inline OperatorPtr compilePangoLineage(const Expression& self, const DatabasePartition& part) {
   const uint32_t rc = part.sequence_count;
   auto it = part.columns.pango_lineage_columns.find(self.column);
   if (it == part.columns.pango_lineage_columns.end()) return operators::makeEmpty(rc);
   std::string upper = self.value;
   std::transform(upper.begin(), upper.end(), upper.begin(), ::toupper);
   const Bitmap* bitmap = it->second.filter(upper, self.include_sublineages);
   if (!bitmap) return operators::makeEmpty(rc);
   return operators::makeIndexScan(bitmap, rc);
}

// string_equals.cpp:33-65
inline OperatorPtr compileStringEquals(const Expression& self, const DatabasePartition& part) {
   const uint32_t rc = part.sequence_count;
   auto it = part.columns.indexed_string_columns.find(self.column);
   if (it != part.columns.indexed_string_columns.end()) {
      const Bitmap* bitmap = it->second.filter(self.value);
      if (!bitmap || bitmap->isEmpty()) return operators::makeEmpty(rc);
      return operators::makeIndexScan(bitmap, rc);
   }
   auto it2 = part.columns.string_columns.find(self.column);
   if (it2 != part.columns.string_columns.end()) {
      // The reference embeds the value in the column dictionary (embedString) and returns Empty if
      // the string is unknown; comparing std::string values directly selects the same rows.
      operators::Predicate p;
      p.kind = operators::ColumnKind::STRING; p.column = &it2->second.values;
      p.comparator = operators::Comparator::EQUALS; p.string_value = self.value;
      return operators::makeSelection(nullptr, {p}, rc);
   }
   return operators::makeEmpty(rc);
}

inline operators::Predicate intPredicate(const std::vector<int32_t>& col, operators::Comparator c, int32_t v) {
   operators::Predicate p; p.kind = operators::ColumnKind::INT32; p.column = &col; p.comparator = c; p.int_value = v; return p;
}
inline operators::Predicate floatPredicate(const std::vector<double>& col, operators::Comparator c, double v) {
   operators::Predicate p; p.kind = operators::ColumnKind::FLOAT64; p.column = &col; p.comparator = c; p.float_value = v; return p;
}

OperatorPtr compileInsertionContains(const Expression& self, bool is_aa, const Database& db, const DatabasePartition& part);

}  // namespace detail

inline OperatorPtr Expression::compile(const Database& db, const DatabasePartition& part, AmbiguityMode mode) const {
   using namespace detail;
   using operators::Comparator;
   const uint32_t rc = part.sequence_count;
   switch (type) {
      case ExprType::TRUE_: return operators::makeFull(rc);     // true.cpp
      case ExprType::FALSE_: return operators::makeEmpty(rc);   // false.cpp
      case ExprType::AND: return compileAnd(*this, db, part, mode);
      case ExprType::OR: return compileOr(*this, db, part, mode);
      case ExprType::NOF: return compileNof(*this, db, part, mode);
      case ExprType::NOT: return children[0]->compile(db, part, invertMode(mode))->negate();  // negation.cpp:27-34
      case ExprType::MAYBE: return children[0]->compile(db, part, AmbiguityMode::UPPER_BOUND); // maybe.cpp:26-32
      case ExprType::EXACT: return children[0]->compile(db, part, AmbiguityMode::LOWER_BOUND); // exact.cpp:26-32
      case ExprType::NUC_EQUALS: return compileSymbolEquals(*this, false, db, part, mode);
      case ExprType::AA_EQUALS: return compileSymbolEquals(*this, true, db, part, mode);
      case ExprType::HAS_NUC_MUTATION: return compileHasMutation(*this, false, db, part, mode);
      case ExprType::HAS_AA_MUTATION: return compileHasMutation(*this, true, db, part, mode);
      case ExprType::DATE_BETWEEN: return compileDateBetween(*this, part);
      case ExprType::PANGO_LINEAGE: return compilePangoLineage(*this, part);
      case ExprType::STRING_EQUALS: return compileStringEquals(*this, part);
      case ExprType::INT_EQUALS: {  // int_equals.cpp:30-47
         auto it = part.columns.int_columns.find(column);
         if (it == part.columns.int_columns.end()) return operators::makeEmpty(rc);
         return operators::makeSelection(nullptr, {intPredicate(it->second.values, Comparator::EQUALS, int_value)}, rc);
      }
      case ExprType::INT_BETWEEN: {  // int_between.cpp:34-58
         const auto& col = part.columns.int_columns.at(column);
         std::vector<operators::Predicate> predicates;
         predicates.push_back(intPredicate(col.values, Comparator::HIGHER_OR_EQUALS, int_from.value_or(NULL_INT + 1)));
         if (int_to) predicates.push_back(intPredicate(col.values, Comparator::LESS_OR_EQUALS, *int_to));
         return operators::makeSelection(nullptr, std::move(predicates), rc);
      }
      case ExprType::FLOAT_EQUALS: {  // float_equals.cpp:32-49
         auto it = part.columns.float_columns.find(column);
         if (it == part.columns.float_columns.end()) return operators::makeEmpty(rc);
         return operators::makeSelection(nullptr, {floatPredicate(it->second.values, Comparator::EQUALS, float_value)}, rc);
      }
      case ExprType::FLOAT_BETWEEN: {  // float_between.cpp:34-67
         CHECK_SILO_QUERY(part.columns.float_columns.count(column), "The database does not contain the float column '" + column + "'")
         const auto& col = part.columns.float_columns.at(column);
         std::vector<operators::Predicate> predicates;
         if (float_from) predicates.push_back(floatPredicate(col.values, Comparator::HIGHER_OR_EQUALS, *float_from));
         if (float_to) predicates.push_back(floatPredicate(col.values, Comparator::LESS, *float_to));
         if (predicates.empty()) predicates.push_back(floatPredicate(col.values, Comparator::NOT_EQUALS, std::nan("")));
         return operators::makeSelection(nullptr, std::move(predicates), rc);
      }
      case ExprType::NUC_INSERTION_CONTAINS: return compileInsertionContains(*this, false, db, part);
      case ExprType::AA_INSERTION_CONTAINS: return compileInsertionContains(*this, true, db, part);
   }
   throw std::runtime_error("unhandled expression type");
}


namespace detail {

struct ParsedInsertion { std::string sequence_name; uint32_t position; std::string insertion; };
// storage/column/insertion_column.cpp:24-60: "pos:INS" (default sequence) or "seq:pos:INS", comma separated
inline std::vector<ParsedInsertion> parseInsertions(const std::string& value, const std::optional<std::string>& default_sequence) {
   std::vector<ParsedInsertion> out;
   size_t start = 0;
   while (start <= value.size() && !value.empty()) {
      size_t comma = value.find(',', start);
      std::string entry = value.substr(start, comma == std::string::npos ? std::string::npos : comma - start);
      std::vector<std::string> parts;
      size_t ps = 0;
      while (true) {
         size_t colon = entry.find(':', ps);
         std::string part = entry.substr(ps, colon == std::string::npos ? std::string::npos : colon - ps);
         part.erase(std::remove(part.begin(), part.end(), '"'), part.end());
         parts.push_back(part);
         if (colon == std::string::npos) break;
         ps = colon + 1;
      }
      if (parts.size() == 2 && default_sequence) out.push_back({*default_sequence, static_cast<uint32_t>(std::stoul(parts[0])), parts[1]});
      else if (parts.size() == 3) out.push_back({parts[0], static_cast<uint32_t>(std::stoul(parts[1])), parts[2]});
      else throw std::runtime_error("Failed to parse insertion due to invalid format: " + entry);
      if (comma == std::string::npos) break;
      start = comma + 1;
   }
   return out;
}

// insertion_contains.cpp:65-126. The reference's per-column 3-mer index + std::regex_search
// (insertion_index.cpp:225-236,282-291) selects exactly the rows that have an insertion at
// `position` of `sequence` whose string regex_search-matches the pattern; the producer below
// computes that set directly on the host and the device ingests it as a BITMAP_PRODUCER leaf.
inline OperatorPtr compileInsertionContains(const Expression& self, bool is_aa, const Database& db, const DatabasePartition& part) {
   const uint32_t rc = part.sequence_count;
   const auto& columns = is_aa ? part.columns.aa_insertion_columns : part.columns.nuc_insertion_columns;
   for (const auto& name : self.column_names) {
      CHECK_SILO_QUERY(columns.count(name), "The insertion column '" + name + "' does not exist.")
   }
   if (columns.empty()) return operators::makeEmpty(rc);
   std::optional<std::string> default_sequence;
   if (!is_aa) default_sequence = db.config.default_nucleotide_sequence;
   std::string sequence;
   if (self.sequence_name) sequence = *self.sequence_name;
   else {
      CHECK_SILO_QUERY(default_sequence.has_value(), std::string("The database has no default ") + (is_aa ? "amino acid" : "nucleotide") + " sequence name")
      sequence = *default_sequence;
   }
   OperatorVector column_operators;
   for (const auto& [column_name, column] : columns) {
      if (!self.column_names.empty() && std::find(self.column_names.begin(), self.column_names.end(), column_name) == self.column_names.end()) continue;
      // getInsertionIndexes().contains(sequence): some row of this partition has an insertion on it
      bool has_index = false;
      for (const auto& v : column.values) {
         for (const auto& ins : parseInsertions(v, default_sequence)) if (ins.sequence_name == sequence) { has_index = true; break; }
         if (has_index) break;
      }
      if (!has_index) continue;
      const InsertionColumnPartition* col = &column;
      const uint32_t position = self.position;
      const std::string pattern = self.value;
      column_operators.push_back(operators::makeBitmapProducer(
         [col, position, pattern, sequence, default_sequence]() {
            const std::regex re(pattern);
            Bitmap result;
            for (uint32_t row = 0; row < col->values.size(); ++row) {
               for (const auto& ins : parseInsertions(col->values[row], default_sequence)) {
                  if (ins.sequence_name == sequence && ins.position == position && std::regex_search(ins.insertion, re)) { result.add(row); break; }
               }
            }
            return result;
         },
         rc));
   }
   if (column_operators.empty()) return operators::makeEmpty(rc);
   if (column_operators.size() == 1) return std::move(column_operators[0]);
   return operators::makeUnion(std::move(column_operators), rc);
}

inline bool validInsertionPattern(const std::string& value, const Alphabet& alphabet) {  // insertion_contains.cpp:128-146
   size_t k = 0;
   while (k < value.size()) {
      if (value[k] == '.' && k + 1 < value.size() && value[k + 1] == '*') { k += 2; continue; }
      bool ok = false;
      for (char c : alphabet.chars) ok |= c == value[k];
      if (!ok) return false;
      ++k;
   }
   return true;
}

}  // namespace detail

// expression.cpp:49-102 and the from_json of every filter_expressions/*.cpp (messages kept verbatim)
inline ExpressionPtr parseExpression(const json::Value& j) {
   CHECK_SILO_QUERY(j.contains("type"), "The field 'type' is required in any filter expression")
   CHECK_SILO_QUERY(j["type"].is_string(), "The field 'type' in all filter expressions needs to be a string, but is: " + j["type"].dump())
   const std::string type = j["type"].as_string();
   auto e = std::make_unique<Expression>();
   auto parseChildren = [&](const char* what) {
      CHECK_SILO_QUERY(j.contains("children"), std::string("The field 'children' is required in an ") + what + " expression")
      CHECK_SILO_QUERY(j["children"].is_array(), std::string("The field 'children' in an ") + what + " expression needs to be an array")
      for (const auto& c : j["children"].arr) e->children.push_back(parseExpression(c));
   };
   if (type == "True") { e->type = ExprType::TRUE_; }
   else if (type == "False") { e->type = ExprType::FALSE_; }
   else if (type == "And") { e->type = ExprType::AND; parseChildren("And"); }
   else if (type == "Or") { e->type = ExprType::OR; parseChildren("Or"); }
   else if (type == "N-Of") {
      e->type = ExprType::NOF;
      parseChildren("N-Of");
      CHECK_SILO_QUERY(j.contains("numberOfMatchers"), "The field 'numberOfMatchers' is required in an N-Of expression")
      CHECK_SILO_QUERY(j["numberOfMatchers"].is_number_unsigned(), "The field 'numberOfMatchers' in an N-Of expression needs to be an unsigned integer")
      CHECK_SILO_QUERY(j.contains("matchExactly"), "The field 'matchExactly' is required in an N-Of expression")
      CHECK_SILO_QUERY(j["matchExactly"].is_boolean(), "The field 'matchExactly' in an N-Of expression needs to be a boolean")
      e->number_of_matchers = static_cast<int>(static_cast<uint32_t>(j["numberOfMatchers"].u));
      e->match_exactly = j["matchExactly"].b;
   }
   else if (type == "Not" || type == "Maybe" || type == "Exact") {
      e->type = type == "Not" ? ExprType::NOT : type == "Maybe" ? ExprType::MAYBE : ExprType::EXACT;
      CHECK_SILO_QUERY(j.contains("child"), "The field 'child' is required in a " + type + " expression")
      e->children.push_back(parseExpression(j["child"]));
   }
   else if (type == "DateBetween") {
      e->type = ExprType::DATE_BETWEEN;
      CHECK_SILO_QUERY(j.contains("column"), "The field 'column' is required in a DateBetween expression")
      CHECK_SILO_QUERY(j["column"].is_string(), "The field 'column' in a DateBetween expression needs to be a string")
      CHECK_SILO_QUERY(j.contains("from"), "The field 'from' is required in DateBetween expression")
      CHECK_SILO_QUERY(j["from"].is_null() || (j["from"].is_string() && !j["from"].s.empty()), "The field 'from' in a DateBetween expression needs to be a string or null")
      CHECK_SILO_QUERY(j.contains("to"), "The field 'to' is required in a DateBetween expression")
      CHECK_SILO_QUERY(j["to"].is_null() || (j["to"].is_string() && !j["to"].s.empty()), "The field 'to' in a DateBetween expression needs to be a non-empty string or null")
      e->column = j["column"].s;
      if (j["from"].is_string()) e->date_from = stringToDate(j["from"].s);
      if (j["to"].is_string()) e->date_to = stringToDate(j["to"].s);
   }
   else if (type == "NucleotideEquals" || type == "AminoAcidEquals") {
      const bool is_aa = type == "AminoAcidEquals";
      e->type = is_aa ? ExprType::AA_EQUALS : ExprType::NUC_EQUALS;
      if (is_aa) { CHECK_SILO_QUERY(j.contains("sequenceName") && j["sequenceName"].is_string(), "AminoAcidEquals expression requires the string field sequenceName") }
      CHECK_SILO_QUERY(j.is_object() && j.contains("position"), "The field 'position' is required in a " + type + " expression")
      CHECK_SILO_QUERY(j["position"].is_number_unsigned() && static_cast<uint32_t>(j["position"].u) > 0,
                       "The field 'position' in a " + type + " expression needs to be an unsigned integer greater than 0")
      if (is_aa) { CHECK_SILO_QUERY(j.contains("symbol") && j["symbol"].is_string(), "The string field 'symbol' is required in a AminoAcidEquals expression") }
      else {
         CHECK_SILO_QUERY(j.contains("symbol"), "The field 'symbol' is required in a NucleotideEquals expression")
         CHECK_SILO_QUERY(j["symbol"].is_string(), "The field 'symbol' in a NucleotideEquals expression needs to be a string")
      }
      if (j.contains("sequenceName")) e->sequence_name = j["sequenceName"].as_string();
      e->position = static_cast<uint32_t>(j["position"].u) - 1;
      const std::string& sym = j["symbol"].s;
      CHECK_SILO_QUERY(sym.size() == 1, "The string field 'symbol' must be exactly one character long")
      const Alphabet& alphabet = is_aa ? aminoAcidAlphabet() : nucleotideAlphabet();
      e->symbol = alphabet.charToSymbol(sym[0]);  // nucleotide '.' -> GAP (quirk Q5); amino acid '.' -> reference
      CHECK_SILO_QUERY(e->symbol.has_value() || sym[0] == '.',
                       is_aa ? "The string field 'symbol' must be either a valid amino acid or the '.' symbol."
                             : "The string field 'symbol' must be either a valid nucleotide symbol or the '.' symbol.")
   }
   else if (type == "HasNucleotideMutation" || type == "HasAminoAcidMutation") {
      const bool is_aa = type == "HasAminoAcidMutation";
      e->type = is_aa ? ExprType::HAS_AA_MUTATION : ExprType::HAS_NUC_MUTATION;
      CHECK_SILO_QUERY(j.contains("position"), "The field 'position' is required in a " + type + " expression")
      CHECK_SILO_QUERY(j["position"].is_number_unsigned(), "The field 'position' in a " + type + " expression needs to be an unsigned integer")
      if (is_aa) { CHECK_SILO_QUERY(j.contains("sequenceName") && j["sequenceName"].is_string(), "HasAminoAcidMutation expression requires the string field sequenceName") }
      if (j.contains("sequenceName")) e->sequence_name = j["sequenceName"].as_string();
      e->position = static_cast<uint32_t>(j["position"].u) - 1;  // quirk Q8: no > 0 check, wraps
   }
   else if (type == "PangoLineage") {
      e->type = ExprType::PANGO_LINEAGE;
      CHECK_SILO_QUERY(j.contains("column"), "The field 'column' is required in a PangoLineage expression")
      CHECK_SILO_QUERY(j["column"].is_string(), "The field 'column' in a PangoLineage expression needs to be a string")
      CHECK_SILO_QUERY(j.contains("value"), "The field 'value' is required in a PangoLineage expression")
      CHECK_SILO_QUERY(j["value"].is_string(), "The field 'value' in a PangoLineage expression needs to be a string")
      CHECK_SILO_QUERY(j.contains("includeSublineages"), "The field 'includeSublineages' is required in a PangoLineage expression")
      CHECK_SILO_QUERY(j["includeSublineages"].is_boolean(), "The field 'includeSublineages' in a PangoLineage expression needs to be a boolean")
      e->column = j["column"].s; e->value = j["value"].s; e->include_sublineages = j["includeSublineages"].b;
   }
   else if (type == "StringEquals") {
      e->type = ExprType::STRING_EQUALS;
      CHECK_SILO_QUERY(j.contains("column"), "The field 'column' is required in an StringEquals expression")
      CHECK_SILO_QUERY(j["column"].is_string(), "The field 'column' in an StringEquals expression needs to be a string")
      CHECK_SILO_QUERY(j.contains("value"), "The field 'value' is required in an StringEquals expression")
      CHECK_SILO_QUERY(j["value"].is_string() || j["value"].is_null(), "The field 'value' in an StringEquals expression needs to be a string or null")
      e->column = j["column"].s; e->value = j["value"].is_null() ? "" : j["value"].s;
   }
   else if (type == "IntEquals") {
      e->type = ExprType::INT_EQUALS;
      CHECK_SILO_QUERY(j.contains("column"), "The field 'column' is required in an IntEquals expression")
      CHECK_SILO_QUERY(j["column"].is_string(), "The field 'column' in an IntEquals expression must be a string")
      CHECK_SILO_QUERY(j.contains("value"), "The field 'value' is required in an IntEquals expression")
      CHECK_SILO_QUERY(j["value"].is_number_integer() || j["value"].is_null(), "The field 'value' in an IntEquals expression must be an integer or null")
      e->column = j["column"].s; e->int_value = j["value"].is_null() ? NULL_INT : static_cast<int32_t>(j["value"].as_int());
   }
   else if (type == "IntBetween") {
      e->type = ExprType::INT_BETWEEN;
      CHECK_SILO_QUERY(j.contains("column"), "The field 'column' is required in a IntBetween expression")
      CHECK_SILO_QUERY(j["column"].is_string(), "The field 'column' in a IntBetween expression must be a string")
      CHECK_SILO_QUERY(j.contains("from"), "The field 'from' is required in IntBetween expression")
      CHECK_SILO_QUERY(j["from"].is_null() || j["from"].is_number_integer(), "The field 'from' in a IntBetween expression must be an int or null")
      CHECK_SILO_QUERY(j.contains("to"), "The field 'to' is required in a IntBetween expression")
      CHECK_SILO_QUERY(j["to"].is_null() || j["to"].is_number_integer(), "The field 'to' in a IntBetween expression must be an int or null")
      e->column = j["column"].s;
      if (j["from"].is_number_integer()) e->int_from = static_cast<int32_t>(j["from"].as_int());
      if (j["to"].is_number_integer()) e->int_to = static_cast<int32_t>(j["to"].as_int());
   }
   else if (type == "FloatEquals") {
      e->type = ExprType::FLOAT_EQUALS;
      CHECK_SILO_QUERY(j.contains("column"), "The field 'column' is required in an FloatEquals expression")
      CHECK_SILO_QUERY(j["column"].is_string(), "The field 'column' in an FloatEquals expression must be a string")
      CHECK_SILO_QUERY(j.contains("value"), "The field 'value' is required in an FloatEquals expression")
      CHECK_SILO_QUERY(j["value"].is_number_float() || j["value"].is_null(), "The field 'value' in an FloatEquals expression must be a float")
      e->column = j["column"].s; e->float_value = j["value"].is_null() ? std::nan("") : j["value"].d;
   }
   else if (type == "FloatBetween") {
      e->type = ExprType::FLOAT_BETWEEN;
      CHECK_SILO_QUERY(j.contains("column"), "The field 'column' is required in a FloatBetween expression")
      CHECK_SILO_QUERY(j["column"].is_string(), "The field 'column' in a FloatBetween expression must be a string")
      CHECK_SILO_QUERY(j.contains("from"), "The field 'from' is required in FloatBetween expression")
      CHECK_SILO_QUERY(j["from"].is_null() || j["from"].is_number_float(), "The field 'from' in a FloatBetween expression must be a float or null")
      CHECK_SILO_QUERY(j.contains("to"), "The field 'to' is required in a FloatBetween expression")
      CHECK_SILO_QUERY(j["to"].is_null() || j["to"].is_number_float(), "The field 'to' in a FloatBetween expression must be a float or null")
      e->column = j["column"].s;
      if (j["from"].is_number_float()) e->float_from = j["from"].d;
      if (j["to"].is_number_float()) e->float_to = j["to"].d;
   }
   else if (type == "InsertionContains" || type == "AminoAcidInsertionContains") {
      const bool is_aa = type == "AminoAcidInsertionContains";
      e->type = is_aa ? ExprType::AA_INSERTION_CONTAINS : ExprType::NUC_INSERTION_CONTAINS;
      CHECK_SILO_QUERY(!j.contains("column") || (j["column"].is_string() || j["column"].is_array()),
                       "The InsertionsContains filter can have the field column of type string or an array of strings, but no other type")
      if (j.contains("column") && j["column"].is_array()) {
         for (const auto& c : j["column"].arr) {
            CHECK_SILO_QUERY(c.is_string(), "The field column of the InsertionsContains filter must have type string or an array, if present. Found:" + c.dump())
            e->column_names.push_back(c.s);
         }
      } else if (j.contains("column")) e->column_names.push_back(j["column"].s);
      CHECK_SILO_QUERY(j.contains("position"), "The field 'position' is required in an InsertionContains expression")
      CHECK_SILO_QUERY(j["position"].is_number_unsigned() && static_cast<uint32_t>(j["position"].u) > 0,
                       "The field 'position' in an InsertionContains expression needs to be a positive number (> 0)")
      CHECK_SILO_QUERY(!j.contains("sequenceName") || j["sequenceName"].is_string(), "The optional field 'sequenceName' in an InsertionContains expression needs to be a string")
      CHECK_SILO_QUERY(j.contains("value"), "The field 'value' is required in an InsertionContains expression")
      CHECK_SILO_QUERY(j["value"].is_string(), "The field 'value' in an InsertionContains expression needs to be a string")
      if (j.contains("sequenceName")) e->sequence_name = j["sequenceName"].s;
      e->position = static_cast<uint32_t>(j["position"].u);
      e->value = j["value"].s;
      CHECK_SILO_QUERY(!e->value.empty(), "The field 'value' in an InsertionContains expression must not be an empty string")
      const Alphabet& alphabet = is_aa ? aminoAcidAlphabet() : nucleotideAlphabet();
      CHECK_SILO_QUERY(detail::validInsertionPattern(e->value, alphabet),
                       "The field 'value' in the InsertionContains expression does not contain a valid regex pattern: \"" + e->value +
                          "\". It must only consist of " + alphabet.name_lower + " symbols and the regex symbol '.*'.")
   }
   else {
      throw QueryParseException("Unknown object filter type '" + type + "'");
   }
   return e;
}

}  // namespace silo::filter
