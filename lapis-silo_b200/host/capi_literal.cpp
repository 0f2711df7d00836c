// C entry points that build a silo::Database bitmap by bitmap ("literal stores"): the states of a
// SequenceStorePartition that neither the dataset builder nor the synthetic generator produce -- a flipped symbol
// (position.h:36), flipped and deleted at one position, run containers in high keys, arrays of an exact
// cardinality, partitions of an exact row count. The resulting Database* is what silo_engine_create and the oracle take.
// Position writers: host/position.hpp (position.cpp:24-127).
#include "capi_common.hpp"
#include "position.hpp"

using namespace silo;

namespace {

struct LiteralHandle {
   std::unique_ptr<Database> db = std::make_unique<Database>();
};

SequenceStorePartition& storeOf(LiteralHandle& h, uint32_t partition, const char* name) {
   DatabasePartition& part = *h.db->partitions.at(partition);
   if (auto it = part.nuc_sequences.find(name); it != part.nuc_sequences.end()) return it->second;
   if (auto it = part.aa_sequences.find(name); it != part.aa_sequences.end()) return it->second;
   throw std::runtime_error(std::string("literal database has no sequence '") + name + "'");
}
Symbol symbolOf(const SequenceStorePartition& st, char c) {
   auto s = st.alphabet->charToSymbol(c);
   if (!s) throw std::runtime_error(std::string("illegal symbol character '") + c + "'");
   return *s;
}
void checkPosition(const SequenceStorePartition& st, uint32_t pos) {
   if (pos >= st.length()) throw std::runtime_error("position out of range");
}

}  // namespace

extern "C" {

void* silo_literal_create() { return new LiteralHandle(); }
void silo_literal_free(void* h) { delete static_cast<LiteralHandle*>(h); }
void* silo_literal_db(void* h) { return static_cast<LiteralHandle*>(h)->db.get(); }

// a reference sequence (text over the alphabet); call for every sequence before the first partition
int silo_literal_add_sequence(void* h_, int alphabet_id, const char* name, const char* reference_text) {
   return capi::guarded([&] {
      auto& h = *static_cast<LiteralHandle*>(h_);
      if (!h.db->partitions.empty()) throw std::runtime_error("add sequences before partitions");
      const Alphabet& alphabet = alphabetById(alphabet_id);
      std::vector<Symbol> ref;
      for (const char* c = reference_text; *c; ++c) {
         auto s = alphabet.charToSymbol(*c);
         if (!s) throw std::runtime_error("illegal character in reference");
         ref.push_back(*s);
      }
      (alphabet_id == 0 ? h.db->nuc_reference : h.db->aa_reference)[name] = std::move(ref);
   });
}

int silo_literal_add_partition(void* h_, uint32_t sequence_count, uint32_t* index) {
   return capi::guarded([&] {
      auto& h = *static_cast<LiteralHandle*>(h_);
      auto part = std::make_unique<DatabasePartition>();
      part->sequence_count = sequence_count;
      part->chunks.push_back({0, sequence_count});
      for (int alphabet_id = 0; alphabet_id < 2; ++alphabet_id) {
         for (const auto& [name, ref] : h.db->getReferences(alphabet_id)) {
            SequenceStorePartition& st = (alphabet_id == 0 ? part->nuc_sequences : part->aa_sequences)[name];
            st.init(&alphabetById(alphabet_id), &ref);
            st.sequence_count = sequence_count;
            st.missing_symbol_bitmaps.assign(sequence_count, Bitmap());
         }
      }
      *index = static_cast<uint32_t>(h.db->partitions.size());
      h.db->partitions.push_back(std::move(part));
   });
}

// the STORED bitmap of (position, symbol), as is: ascending ids, optionally run-optimised
int silo_literal_set_bitmap(void* h_, uint32_t partition, const char* name, uint32_t pos, char symbol, const uint32_t* ids, uint64_t n, int run_optimize) {
   return capi::guarded([&] {
      auto& st = storeOf(*static_cast<LiteralHandle*>(h_), partition, name);
      checkPosition(st, pos);
      Bitmap bm = Bitmap::fromSorted(ids, n);
      if (run_optimize) bm.runOptimize();
      st.bitmapRef(pos, symbolOf(st, symbol)) = std::move(bm);
   });
}
// symbol < 0 (pass 0 as the character) clears the flag
int silo_literal_set_flags(void* h_, uint32_t partition, const char* name, uint32_t pos, char flipped, char deleted) {
   return capi::guarded([&] {
      auto& st = storeOf(*static_cast<LiteralHandle*>(h_), partition, name);
      checkPosition(st, pos);
      st.flipped[pos] = flipped ? static_cast<int8_t>(symbolOf(st, flipped)) : int8_t(-1);
      st.deleted[pos] = deleted ? static_cast<int8_t>(symbolOf(st, deleted)) : int8_t(-1);
   });
}
// missing-symbol bitmap of a row: [start, end) position ranges
int silo_literal_set_missing(void* h_, uint32_t partition, const char* name, uint32_t row, const uint32_t* ranges, uint32_t n_ranges, int run_optimize) {
   return capi::guarded([&] {
      auto& st = storeOf(*static_cast<LiteralHandle*>(h_), partition, name);
      Bitmap bm;
      for (uint32_t k = 0; k < n_ranges; ++k) bm.addRange(ranges[2 * k], ranges[2 * k + 1]);
      if (run_optimize) bm.runOptimize();
      st.missing_symbol_bitmaps.at(row) = std::move(bm);
   });
}

// the same for rows [0, n_rows): ranges of row r = ranges[2 * range_off[r] .. 2 * range_off[r + 1])
int silo_literal_set_missing_bulk(void* h_, uint32_t partition, const char* name, uint32_t n_rows, const uint64_t* range_off, const uint32_t* ranges, int run_optimize) {
   return capi::guarded([&] {
      auto& st = storeOf(*static_cast<LiteralHandle*>(h_), partition, name);
      if (n_rows > st.missing_symbol_bitmaps.size()) throw std::runtime_error("more rows than the partition holds");
      for (uint32_t row = 0; row < n_rows; ++row) {
         Bitmap bm;
         for (uint64_t k = range_off[row]; k < range_off[row + 1]; ++k) bm.addRange(ranges[2 * k], ranges[2 * k + 1]);
         if (run_optimize) bm.runOptimize();
         st.missing_symbol_bitmaps[row] = std::move(bm);
      }
   });
}

// ---- Position writers (position.cpp:24-127) --------------------------------------------------------
int silo_literal_add_values(void* h_, uint32_t partition, const char* name, uint32_t pos, char symbol, const uint32_t* ids, uint64_t n, uint64_t offset, uint64_t interval) {
   return capi::guarded([&] {
      auto& st = storeOf(*static_cast<LiteralHandle*>(h_), partition, name);
      checkPosition(st, pos);
      position::addValues(st, pos, symbolOf(st, symbol), std::vector<uint32_t>(ids, ids + n), offset, interval);
   });
}
// *result: the symbol character, or 0 for std::nullopt
int silo_literal_flip_most_numerous(void* h_, uint32_t partition, const char* name, uint32_t pos, uint32_t sequence_count, char* result) {
   return capi::guarded([&] {
      auto& st = storeOf(*static_cast<LiteralHandle*>(h_), partition, name);
      checkPosition(st, pos);
      const auto s = position::flipMostNumerousBitmap(st, pos, sequence_count);
      *result = s ? st.alphabet->symbolToChar(*s) : char(0);
   });
}
int silo_literal_delete_most_numerous(void* h_, uint32_t partition, const char* name, uint32_t pos, uint32_t sequence_count, char* result) {
   return capi::guarded([&] {
      auto& st = storeOf(*static_cast<LiteralHandle*>(h_), partition, name);
      checkPosition(st, pos);
      const auto s = position::deleteMostNumerousBitmap(st, pos, sequence_count);
      *result = s ? st.alphabet->symbolToChar(*s) : char(0);
   });
}
// ids of the stored bitmap (malloc'ed; silo_free), and the flags as characters (0 = none)
int silo_literal_get_bitmap(void* h_, uint32_t partition, const char* name, uint32_t pos, char symbol, uint32_t** ids, size_t* n, char* flipped, char* deleted) {
   return capi::guarded([&] {
      auto& st = storeOf(*static_cast<LiteralHandle*>(h_), partition, name);
      checkPosition(st, pos);
      const auto v = st.getBitmap(pos, symbolOf(st, symbol))->toVector();
      *n = v.size();
      *ids = static_cast<uint32_t*>(malloc(std::max<size_t>(1, v.size()) * sizeof(uint32_t)));
      memcpy(*ids, v.data(), v.size() * sizeof(uint32_t));
      *flipped = st.flipped[pos] < 0 ? char(0) : st.alphabet->symbolToChar(static_cast<Symbol>(st.flipped[pos]));
      *deleted = st.deleted[pos] < 0 ? char(0) : st.alphabet->symbolToChar(static_cast<Symbol>(st.deleted[pos]));
   });
}

}  // extern "C"
