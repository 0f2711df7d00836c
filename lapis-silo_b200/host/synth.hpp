// Synthetic SARS-CoV-2-shaped database generator: produces the in-memory index
// directly (no FASTA round trip), deterministic from the seed with integer-only SplitMix64 draws.
// Partitioning and row order follow the reference's rule (preprocessor.cpp:159-199 greedy merge with
// allowed = N/32; ORDER BY date, primaryKey with NULLS FIRST, database_config.cpp:190-198); the index
// content follows SequenceStorePartition::fill + optimizeBitmaps (sequence_store.cpp:101-211): one
// bitmap per (position, symbol) of the non-missing symbols, the most numerous symbol deleted, one
// missing-symbol bitmap per row. `explicit_sequences` additionally materialises every row as a string so
// that tests can rebuild the same database through builder::fillSequenceStore and compare.
#pragma once
#include <atomic>
#include <thread>

#include "builder.hpp"

namespace silo::synth {

struct Rng {   // SplitMix64
   uint64_t state;
   explicit Rng(uint64_t seed) : state(seed) {}
   uint64_t next() {
      uint64_t z = (state += 0x9E3779B97F4A7C15ULL);
      z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
      z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
      return z ^ (z >> 31);
   }
   uint32_t below(uint32_t n) { return static_cast<uint32_t>((next() >> 32) * uint64_t(n) >> 32); }   // [0, n)
   uint32_t range(uint32_t lo, uint32_t hi) { return lo + below(hi - lo + 1); }                       // [lo, hi]
   bool chance(uint32_t numerator, uint32_t denominator) { return below(denominator) < numerator; }
};
inline uint64_t streamSeed(uint64_t seed, uint64_t a, uint64_t b = 0) {
   Rng r(seed ^ (a * 0xD6E8FEB86659FD93ULL) ^ (b * 0xCA5A826395121157ULL));
   r.next();
   return r.next();
}

inline uint64_t nameHash(const std::string& s) {   // FNV-1a
   uint64_t h = 1469598103934665603ULL;
   for (unsigned char c : s) { h ^= c; h *= 1099511628211ULL; }
   return h;
}

struct Options {
   uint64_t n_sequences = 10000;
   uint64_t seed = 0x51100003ULL;
   bool with_amino_acids = false;
   bool explicit_sequences = false;   // tests only (small n)
   uint32_t n_threads = 0;
   uint32_t genome_length = 0;        // 0 = full reference; otherwise truncate (tests)
   uint32_t world = 1, rank = 0;      // multi-GPU: build only the partitions this rank owns (the others stay empty shells)
   uint32_t row_slices = 1;           // every lineage partition is cut into this many DatabasePartitions of consecutive rows (same
                                      // rows, same values: a finer horizontal split, so that N GPUs can share every lineage evenly)
};

struct Event { uint32_t pos; uint8_t sym; };
struct Interval { uint32_t start, end; };   // [start, end)

struct SequenceModel {   // one aligned sequence (nucleotide segment or gene) of the synthetic population
   std::string name;
   const Alphabet* alphabet;
   std::vector<Symbol> reference;
   std::vector<Symbol> substitution_symbols;   // A,C,G,T or the 20 amino acids
   uint32_t nuc_offset = 0;                    // genes: nucleotide coordinate of codon 0
   bool is_nucleotide = true;
   std::vector<std::vector<Event>> lineage_events;   // per lineage: defining events of the whole ancestor chain, sorted by pos
};

struct Lineage { std::string name; int parent; };

inline std::vector<Lineage> makeLineages() {   // 96 pango-like names in a 3-level tree
   std::vector<Lineage> out;
   out.push_back({"B.1", -1});
   for (int k = 1; k <= 19; ++k) out.push_back({"B.1." + std::to_string(k), 0});
   for (int k = 1; k <= 19; ++k) for (int m = 1; m <= 4; ++m) out.push_back({"B.1." + std::to_string(k) + "." + std::to_string(m), k});
   return out;
}

inline uint32_t dayToDate(uint32_t day_index) {   // days since 2020-01-01 -> (year<<16)+(month<<12)+day (common/date.cpp:49-52)
   static const int month_days[12] = {31, 28, 31, 30, 31, 30, 31, 31, 30, 31, 30, 31};
   uint32_t year = 2020;
   while (true) {
      const bool leap = year % 4 == 0;
      const uint32_t len = leap ? 366 : 365;
      if (day_index < len) {
         for (uint32_t m = 0; m < 12; ++m) {
            const uint32_t md = month_days[m] + (m == 1 && leap ? 1 : 0);
            if (day_index < md) return (year << 16) + ((m + 1) << 12) + day_index + 1;
            day_index -= md;
         }
      }
      day_index -= len;
      ++year;
   }
}

// defining events of one lineage node for one sequence model
inline std::vector<Event> nodeEvents(const SequenceModel& model, uint64_t seed, uint32_t node) {
   Rng rng(streamSeed(seed, 0x1000 + node, nameHash(model.name)));
   std::vector<Event> events;
   const uint32_t L = static_cast<uint32_t>(model.reference.size());
   const uint32_t n_subst = model.is_nucleotide ? rng.range(8, 24) : (L >= 60 ? rng.range(0, 2) : rng.below(2) & rng.below(2));
   for (uint32_t k = 0; k < n_subst; ++k) {
      const uint32_t pos = rng.below(L);
      const Symbol ref = model.reference[pos];
      Symbol alt = model.substitution_symbols[rng.below(static_cast<uint32_t>(model.substitution_symbols.size()))];
      if (alt == ref) alt = model.substitution_symbols[(std::find(model.substitution_symbols.begin(), model.substitution_symbols.end(), alt) - model.substitution_symbols.begin() + 1) % model.substitution_symbols.size()];
      if (!model.is_nucleotide && pos + 1 == L) continue;   // keep the terminal stop codon
      events.push_back({pos, alt});
   }
   if (model.is_nucleotide && rng.chance(1, 3)) {   // a defining deletion: gap run of 3-9
      const uint32_t len = rng.range(3, 9), start = rng.below(L - len);
      for (uint32_t k = 0; k < len; ++k) events.push_back({start + k, 0});
   }
   return events;
}

// later (more specific) lists override earlier ones at the same position; result sorted by pos
inline std::vector<Event> overlay(const std::vector<Event>& base, const std::vector<Event>& top) {
   std::map<uint32_t, uint8_t> merged;
   for (const auto& e : base) merged[e.pos] = e.sym;
   for (const auto& e : top) merged[e.pos] = e.sym;
   std::vector<Event> out;
   for (const auto& [pos, sym] : merged) out.push_back({pos, sym});
   return out;
}

struct RowMeta { uint32_t lineage; uint32_t date; uint32_t region; int32_t age; double qc; uint64_t id; };

inline RowMeta rowMeta(uint64_t seed, uint64_t id, const std::vector<uint64_t>& lineage_cumulative) {
   Rng rng(streamSeed(seed, 0x2000, id));
   RowMeta m;
   m.id = id;
   const uint64_t draw = rng.next() % lineage_cumulative.back();
   m.lineage = static_cast<uint32_t>(std::upper_bound(lineage_cumulative.begin(), lineage_cumulative.end(), draw) - lineage_cumulative.begin());
   m.date = rng.chance(1, 200) ? NULL_DATE : dayToDate(rng.below(1461));   // 2020-01-01 .. 2023-12-31
   m.region = rng.below(6);
   m.age = rng.chance(1, 20) ? NULL_INT : static_cast<int32_t>(rng.below(100));
   m.qc = static_cast<double>(rng.below(10001)) / 10000.0;
   return m;
}

// nucleotide-coordinate missing intervals of a row (sorted, merged) + whether the genome is null
inline std::vector<Interval> rowMissing(uint64_t seed, uint64_t id, uint32_t L, bool& fully_missing, uint32_t& gap_prefix, uint32_t& gap_suffix) {
   Rng rng(streamSeed(seed, 0x3000, id));
   std::vector<Interval> iv;
   gap_prefix = gap_suffix = 0;
   fully_missing = rng.chance(1, 500);
   if (fully_missing) return {{0, L}};
   const uint32_t prefix = std::min(rng.range(0, 60), L / 4), suffix = std::min(rng.range(0, 80), L / 4);
   if (rng.chance(3, 4)) { if (prefix) iv.push_back({0, prefix}); } else gap_prefix = prefix;
   if (rng.chance(3, 4)) { if (suffix) iv.push_back({L - suffix, L}); } else gap_suffix = suffix;
   const uint32_t n_runs = __builtin_popcount(rng.below(16));
   for (uint32_t k = 0; k < n_runs; ++k) {
      const uint32_t len = std::min(rng.range(50, 400), L / 8 + 1), start = rng.below(L - len);
      iv.push_back({start, start + len});
   }
   if (rng.chance(1, 100)) { const uint32_t len = L * 3 / 10, start = rng.below(L - len); iv.push_back({start, start + len}); }
   std::sort(iv.begin(), iv.end(), [](const Interval& a, const Interval& b) { return a.start < b.start; });
   std::vector<Interval> merged;
   for (const auto& i : iv) {
      if (!merged.empty() && i.start <= merged.back().end) merged.back().end = std::max(merged.back().end, i.end);
      else merged.push_back(i);
   }
   return merged;
}

// final non-reference, non-missing symbols of one row of one sequence model, sorted by position
inline void rowEvents(const SequenceModel& model, uint64_t seed, const RowMeta& meta, const std::vector<Interval>& missing,
                      uint32_t gap_prefix, uint32_t gap_suffix, std::vector<Event>& out, std::vector<Event>& scratch) {
   const uint32_t L = static_cast<uint32_t>(model.reference.size());
   Rng rng(streamSeed(seed, 0x4000 + nameHash(model.name) % 977, meta.id));
   scratch.clear();
   // private substitutions: Binomial(16, 1/2) for nucleotides, Binomial(4, 1/4) per gene
   const uint32_t n_private = model.is_nucleotide ? __builtin_popcount(rng.below(65536)) : __builtin_popcount(rng.below(16) & rng.below(16));
   for (uint32_t k = 0; k < n_private; ++k) {
      const uint32_t pos = rng.below(L);
      if (!model.is_nucleotide && pos + 1 == L) continue;
      const Symbol ref = model.reference[pos];
      Symbol alt = model.substitution_symbols[rng.below(static_cast<uint32_t>(model.substitution_symbols.size()))];
      if (alt == ref) continue;
      scratch.push_back({pos, alt});
   }
   if (model.is_nucleotide) {
      for (uint32_t t = 0; t < 3; ++t) if (rng.chance(1, 5)) scratch.push_back({rng.below(L), static_cast<uint8_t>(rng.range(nuc::R, nuc::V))});   // IUPAC codes
      for (uint32_t p = 0; p < gap_prefix; ++p) scratch.push_back({p, nuc::GAP});
      for (uint32_t p = 0; p < gap_suffix; ++p) scratch.push_back({L - gap_suffix + p, nuc::GAP});
   }
   std::stable_sort(scratch.begin(), scratch.end(), [](const Event& a, const Event& b) { return a.pos < b.pos; });
   // merge: row-specific events (last one at a position wins) override the lineage's defining events
   const std::vector<Event>& defining = model.lineage_events[meta.lineage];
   out.clear();
   size_t a = 0, b = 0, m = 0;
   auto emit = [&](Event e) {
      while (m < missing.size() && missing[m].end <= e.pos) ++m;
      if (m < missing.size() && missing[m].start <= e.pos) return;   // missing symbol wins
      if (e.sym == model.reference[e.pos]) return;
      out.push_back(e);
   };
   while (a < defining.size() || b < scratch.size()) {
      if (b == scratch.size() || (a < defining.size() && defining[a].pos < scratch[b].pos)) { emit(defining[a++]); continue; }
      const uint32_t pos = scratch[b].pos;
      Event last = scratch[b];
      while (b < scratch.size() && scratch[b].pos == pos) last = scratch[b++];
      if (a < defining.size() && defining[a].pos == pos) ++a;
      emit(last);
   }
}

// X intervals of a gene from the nucleotide missing intervals (a codon with any missing base is X)
inline std::vector<Interval> geneMissing(const SequenceModel& gene, const std::vector<Interval>& nuc_missing) {
   const uint32_t L = static_cast<uint32_t>(gene.reference.size());
   std::vector<Interval> out;
   for (const auto& i : nuc_missing) {
      if (i.end <= gene.nuc_offset) continue;
      const uint32_t s = i.start <= gene.nuc_offset ? 0 : (i.start - gene.nuc_offset) / 3;
      const uint32_t e = std::min<uint32_t>((i.end - 1 - gene.nuc_offset) / 3 + 1, L);
      if (s >= e) continue;
      if (!out.empty() && s <= out.back().end) out.back().end = std::max(out.back().end, e); else out.push_back({s, e});
   }
   return out;
}

}  // namespace silo::synth

#include "synth_build.hpp"
