// Position<SymbolType> restated over the flat store layout of storage.hpp: one position of a
// SequenceStorePartition = its Sigma bitmaps + the flipped / deleted symbol.
// Follows include/silo/storage/position.h:22-66 and src/silo/storage/position.cpp:24-127
// (addValues, getHighestCardinalitySymbol, flipMostNumerousBitmap, deleteMostNumerousBitmap).
// The hot path only READS positions; these writers exist so that stores in every state the reference
// can serialize (flipped, deleted, neither) can be built for the parity tests and the archive reader.
#pragma once
#include <stdexcept>

#include "storage.hpp"

namespace silo::position {

inline std::optional<Symbol> flippedSymbol(const SequenceStorePartition& st, uint32_t pos) {
   return st.flipped[pos] < 0 ? std::nullopt : std::optional<Symbol>(static_cast<Symbol>(st.flipped[pos]));
}
inline std::optional<Symbol> deletedSymbol(const SequenceStorePartition& st, uint32_t pos) {
   return st.deleted[pos] < 0 ? std::nullopt : std::optional<Symbol>(static_cast<Symbol>(st.deleted[pos]));
}

// Position::fromInitiallyFlipped / fromInitiallyDeleted (position.cpp:6-22)
inline void initFlipped(SequenceStorePartition& st, uint32_t pos, Symbol s) { st.flipped[pos] = static_cast<int8_t>(s); }
inline void initDeleted(SequenceStorePartition& st, uint32_t pos, Symbol s) { st.deleted[pos] = static_cast<int8_t>(s); }

// position.cpp:24-40: values of the deleted symbol are dropped; the flipped symbol's interval is toggled after the add
inline void addValues(SequenceStorePartition& st, uint32_t pos, Symbol symbol, const std::vector<uint32_t>& values, size_t current_offset, size_t interval_size) {
   if (st.isSymbolDeleted(pos, symbol)) return;
   Bitmap& bm = st.bitmapRef(pos, symbol);
   if (!values.empty()) bm.addMany(values.size(), values.data());
   if (st.isSymbolFlipped(pos, symbol)) bm.flip(current_offset, current_offset + interval_size);
}

// position.cpp:42-69: first symbol in SYMBOLS order with the strictly highest count; runOptimize on every bitmap on the way
inline std::optional<Symbol> highestCardinalitySymbol(SequenceStorePartition& st, uint32_t pos, uint32_t sequence_count) {
   if (st.deleted[pos] >= 0) {
      throw std::runtime_error(std::string("Symbol '") + st.alphabet->symbolToChar(static_cast<Symbol>(st.deleted[pos])) +
                               "' is currently deleted. Cannot restore it implicitly and cannot calculate its cardinality as we do not have information about missing symbols");
   }
   std::optional<Symbol> max_symbol;
   uint32_t max_count = 0;
   for (Symbol symbol : st.alphabet->iteration_order) {
      Bitmap& bm = st.bitmapRef(pos, symbol);
      bm.runOptimize();
      const uint32_t card = static_cast<uint32_t>(bm.cardinality());
      const uint32_t count = st.isSymbolFlipped(pos, symbol) ? sequence_count - card : card;
      if (count > max_count) { max_symbol = symbol; max_count = count; }
   }
   return max_symbol;
}

// position.cpp:71-100; returns the newly flipped symbol, nullopt when nothing changed
inline std::optional<Symbol> flipMostNumerousBitmap(SequenceStorePartition& st, uint32_t pos, uint32_t sequence_count) {
   if (st.deleted[pos] >= 0) {
      throw std::runtime_error(std::string("Symbol '") + st.alphabet->symbolToChar(static_cast<Symbol>(st.deleted[pos])) + "' is currently deleted. Cannot restore it implicitly");
   }
   const std::optional<Symbol> max_symbol = highestCardinalitySymbol(st, pos, sequence_count);
   if (max_symbol == flippedSymbol(st, pos)) return std::nullopt;
   if (st.flipped[pos] >= 0) { Bitmap& bm = st.bitmapRef(pos, static_cast<Symbol>(st.flipped[pos])); bm.flip(0, sequence_count); bm.runOptimize(); }
   if (max_symbol) { Bitmap& bm = st.bitmapRef(pos, *max_symbol); bm.flip(0, sequence_count); bm.runOptimize(); }
   st.flipped[pos] = max_symbol ? static_cast<int8_t>(*max_symbol) : int8_t(-1);
   return max_symbol;
}

// position.cpp:102-127; returns the deleted symbol, nullopt when every bitmap is empty
inline std::optional<Symbol> deleteMostNumerousBitmap(SequenceStorePartition& st, uint32_t pos, uint32_t sequence_count) {
   if (st.deleted[pos] >= 0) {
      throw std::runtime_error(std::string("Symbol '") + st.alphabet->symbolToChar(static_cast<Symbol>(st.deleted[pos])) + "' is currently deleted. Cannot restore it implicitly");
   }
   if (st.flipped[pos] >= 0) {
      Bitmap& bm = st.bitmapRef(pos, static_cast<Symbol>(st.flipped[pos]));
      bm.flip(0, sequence_count);
      bm.runOptimize();
      st.flipped[pos] = -1;
   }
   const std::optional<Symbol> max_symbol = highestCardinalitySymbol(st, pos, sequence_count);
   if (!max_symbol) return std::nullopt;
   st.bitmapRef(pos, *max_symbol) = Bitmap();
   st.deleted[pos] = static_cast<int8_t>(*max_symbol);
   return max_symbol;
}

}  // namespace silo::position
