// SequenceStorePartition::fill + optimizeBitmaps through the device (sb_index_build, include/silo_b200.h): the O(rows x length)
// pass over the characters runs on the GPU; the host only wraps the row ids of every kept bitmap and the missing runs of
// every row into roaring bitmaps (the same Bitmap::fromSorted + runOptimize the host builder ends with, builder.hpp).
// Mirrors src/silo/storage/sequence_store.cpp:101-211.
#pragma once
#include "../../include/silo_b200.h"
#include "builder.hpp"

namespace silo::builder {

inline StoreFiller deviceStoreFiller(int device) {
   return [device](SequenceStorePartition& store, const std::vector<const std::optional<std::string>*>& genomes) {
      const Alphabet& alphabet = *store.alphabet;
      const uint32_t length = store.length(), n = static_cast<uint32_t>(genomes.size());
      store.sequence_count = n;
      store.missing_symbol_bitmaps.assign(n, Bitmap());
      int8_t char_to_symbol[256];
      for (int c = 0; c < 256; ++c) { auto s = alphabet.charToSymbol(static_cast<char>(c)); char_to_symbol[c] = s ? static_cast<int8_t>(*s) : -1; }
      std::string chars(size_t(n) * length, '\0');
      std::vector<uint8_t> present(n, 0);
      for (uint32_t row = 0; row < n; ++row) {
         if (!genomes[row]->has_value()) continue;
         const std::string& g = **genomes[row];
         if (g.size() != length) throw std::runtime_error("sequence length " + std::to_string(g.size()) + " does not match the reference length " + std::to_string(length));
         memcpy(chars.data() + size_t(row) * length, g.data(), length);
         present[row] = 1;
      }
      sb_index* index = nullptr;
      if (sb_index_build(device, alphabet.count, char_to_symbol, alphabet.iteration_order.data(), alphabet.missing, length, n, chars.data(), present.data(), &index) != 0) {
         throw std::runtime_error(sb_last_error());
      }
      struct Free { sb_index* i; ~Free() { sb_index_free(i); } } free_index{index};
      uint64_t n_ids = 0, n_intervals = 0;
      if (sb_index_sizes(index, &n_ids, &n_intervals) != 0) throw std::runtime_error(sb_last_error());
      std::vector<int8_t> deleted(length);
      std::vector<uint64_t> slot_off(size_t(length) * alphabet.count + 1), missing_off(size_t(n) + 1);
      std::vector<uint32_t> ids(n_ids), intervals(2 * n_intervals);
      if (sb_index_fetch(index, deleted.data(), slot_off.data(), ids.data(), missing_off.data(), intervals.data()) != 0) throw std::runtime_error(sb_last_error());
      for (uint32_t pos = 0; pos < length; ++pos) {
         store.deleted[pos] = deleted[pos];
         for (Symbol s = 0; s < alphabet.count; ++s) {
            const uint64_t b = slot_off[size_t(pos) * alphabet.count + s], e = slot_off[size_t(pos) * alphabet.count + s + 1];
            if (b == e) continue;
            Bitmap bm = Bitmap::fromSorted(ids.data() + b, e - b);
            bm.runOptimize();
            store.bitmapRef(pos, s) = std::move(bm);
         }
      }
      for (uint32_t row = 0; row < n; ++row) {
         if (missing_off[row] == missing_off[row + 1]) continue;
         Bitmap& bm = store.missing_symbol_bitmaps[row];
         for (uint64_t k = missing_off[row]; k < missing_off[row + 1]; ++k) bm.addRange(intervals[2 * k], intervals[2 * k + 1]);
         bm.runOptimize();
      }
   };
}

}  // namespace silo::builder
