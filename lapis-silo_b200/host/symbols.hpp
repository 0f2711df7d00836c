// Symbol alphabets of the index. Runtime descriptors instead of the reference's class templates.
// Follows include/silo/common/nucleotide_symbols.h:15-84, src/silo/common/nucleotide_symbols.cpp:6-88,
// include/silo/common/aa_symbols.h:15-89, src/silo/common/aa_symbols.cpp (char tables).
// A symbol is its enum value (uint8_t); SymbolMap indexing (symbol_map.h:12-34) is by that value.
// NOTE: AminoAcid iterates SYMBOLS in the order ...,B,Z,X,STOP while the enum values are
// ...,B=21,Z=22,STOP=23,X=24 (aa_symbols.h:17-54); `iteration_order` keeps that difference.
#pragma once
#include <array>
#include <cstdint>
#include <optional>
#include <string>
#include <vector>

namespace silo {

using Symbol = uint8_t;

struct Alphabet {
   int id;                 // 0 nucleotide, 1 amino acid (matches the C ABI)
   uint32_t count;         // 16 / 25
   const char* name;       // "Nucleotide" / "Amino Acid"
   const char* name_lower; // "nucleotide" / "amino acid"
   std::vector<char> chars;               // by enum value
   std::vector<Symbol> iteration_order;   // SymbolType::SYMBOLS
   std::vector<Symbol> valid_mutation_symbols;
   Symbol missing;

   char symbolToChar(Symbol s) const { return chars[s]; }
   std::optional<Symbol> charToSymbol(char c) const {
      if (id == 0) {
         if (c == '.') return Symbol(0);
         if (c == 'U') c = 'T';
      }
      for (uint32_t k = 0; k < count; ++k) if (chars[k] == c) return Symbol(k);
      return std::nullopt;
   }
};

namespace nuc {
enum : Symbol { GAP = 0, A, C, G, T, R, Y, S, W, K, M, B, D, H, V, N, COUNT };
}
namespace aa {
enum : Symbol { GAP = 0, A, C, D, E, F, G, H, I, K, L, M, N, P, Q, R, S, T, V, W, Y, B, Z, STOP, X, COUNT };
}

inline const Alphabet& nucleotideAlphabet() {
   static const Alphabet alphabet = [] {
      Alphabet a;
      a.id = 0; a.count = 16; a.name = "Nucleotide"; a.name_lower = "nucleotide";
      a.chars = {'-', 'A', 'C', 'G', 'T', 'R', 'Y', 'S', 'W', 'K', 'M', 'B', 'D', 'H', 'V', 'N'};
      for (Symbol s = 0; s < 16; ++s) a.iteration_order.push_back(s);
      a.valid_mutation_symbols = {nuc::GAP, nuc::A, nuc::C, nuc::G, nuc::T};
      a.missing = nuc::N;
      return a;
   }();
   return alphabet;
}

inline const Alphabet& aminoAcidAlphabet() {
   static const Alphabet alphabet = [] {
      Alphabet a;
      a.id = 1; a.count = 25; a.name = "Amino Acid"; a.name_lower = "amino acid";
      a.chars = {'-', 'A', 'C', 'D', 'E', 'F', 'G', 'H', 'I', 'K', 'L', 'M', 'N',
                 'P', 'Q', 'R', 'S', 'T', 'V', 'W', 'Y', 'B', 'Z', '*', 'X'};
      for (Symbol s = 0; s <= aa::Z; ++s) a.iteration_order.push_back(s);
      a.iteration_order.push_back(aa::X);
      a.iteration_order.push_back(aa::STOP);
      for (Symbol s = 0; s <= aa::Y; ++s) a.valid_mutation_symbols.push_back(s);
      a.valid_mutation_symbols.push_back(aa::STOP);
      a.missing = aa::X;
      return a;
   }();
   return alphabet;
}

inline const Alphabet& alphabetById(int id) { return id == 0 ? nucleotideAlphabet() : aminoAcidAlphabet(); }

// IUPAC ambiguity expansion used under AmbiguityMode::UPPER_BOUND
// (src/silo/query_engine/filter_expressions/nucleotide_symbol_equals.cpp:28-73).
inline const std::vector<Symbol>& ambiguityNucSymbols(Symbol s) {
   using namespace nuc;
   static const std::array<std::vector<Symbol>, 16> table{{
      {GAP},
      {A, R, M, W, D, H, V, N},
      {C, Y, M, S, B, H, V, N},
      {G, R, K, S, B, D, V, N},
      {T, Y, K, W, B, D, H, N},
      {R}, {Y}, {S}, {W}, {K}, {M}, {B}, {D}, {H}, {V}, {N},
   }};
   return table[s];
}

}  // namespace silo
