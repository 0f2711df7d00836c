// ORACLE — TEST INFRASTRUCTURE ONLY. Nothing under lapis-silo_b200/ may include, link or call this.
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs use it.
//
// CPU set algebra over the roaring container model, restating what the reference gets from the
// third-party CRoaring roaring/1.0.0 (conanfile.py:17; source NOT under /root/reference). The
// published algorithm is restated: container-wise merge over keys; array x array by merge or
// galloping, array x bitset by bit probes, bitset x bitset by word ops + popcount, run containers
// by interval arithmetic. All operations are set-semantic, so any correct implementation is
// bit-identical; parity is anchored on the reference's call sites:
//   and_cardinality / andnot_cardinality  mutations.cpp:87-88
//   &= / -= / | / fastunion               intersection.cpp:66-124, threshold.cpp:100-133, union.cpp:44
//   flip / addRange / contains            complement.cpp:52, full.cpp:26, bitmap_selection.cpp:38
#pragma once
#include <cstdint>
#include <vector>

#include "../lapis-silo_b200/host/roaring.hpp"

namespace oracle {

using silo::roaring::ARRAY;
using silo::roaring::Bitmap;
using silo::roaring::BITSET;
using silo::roaring::Container;
using silo::roaring::RUN;

inline uint32_t popcountRange(const uint64_t* w, uint32_t lo, uint32_t hi) {  // bits in [lo, hi]
   uint32_t fw = lo >> 6, lw = hi >> 6;
   uint64_t fm = ~0ULL << (lo & 63), lm = ~0ULL >> (63 - (hi & 63));
   if (fw == lw) return __builtin_popcountll(w[fw] & fm & lm);
   uint32_t c = __builtin_popcountll(w[fw] & fm) + __builtin_popcountll(w[lw] & lm);
   for (uint32_t k = fw + 1; k < lw; ++k) c += __builtin_popcountll(w[k]);
   return c;
}

inline uint32_t arrayArrayAndCard(const uint16_t* a, uint32_t na, const uint16_t* b, uint32_t nb) {
   if (na > nb) { std::swap(a, b); std::swap(na, nb); }
   uint32_t count = 0;
   if (static_cast<uint64_t>(na) * 64 < nb) {  // galloping: small side probes the large side
      uint32_t lo = 0;
      for (uint32_t i = 0; i < na && lo < nb; ++i) {
         const uint16_t v = a[i];
         uint32_t step = 1, hi = lo;
         while (hi < nb && b[hi] < v) { lo = hi + 1; hi += step; step *= 2; }
         if (hi > nb) hi = nb;
         uint32_t l = lo, h = hi < nb ? hi + 1 : nb;
         while (l < h) { uint32_t m = (l + h) / 2; if (b[m] < v) l = m + 1; else h = m; }
         lo = l;
         if (lo < nb && b[lo] == v) { ++count; ++lo; }
      }
      return count;
   }
   uint32_t i = 0, j = 0;
   while (i < na && j < nb) {
      if (a[i] < b[j]) ++i;
      else if (a[i] > b[j]) ++j;
      else { ++count; ++i; ++j; }
   }
   return count;
}

// |A ∩ B| for one container pair of any kinds
inline uint32_t containerAndCard(const Container& a, const Container& b) {
   if (a.kind == BITSET && b.kind == BITSET) {
      const uint64_t *x = a.words(), *y = b.words();
      uint32_t c = 0;
      for (uint32_t k = 0; k < 1024; ++k) c += __builtin_popcountll(x[k] & y[k]);
      return c;
   }
   if (a.kind == ARRAY && b.kind == ARRAY) return arrayArrayAndCard(a.data.data(), a.card, b.data.data(), b.card);
   if (a.kind == ARRAY && b.kind == BITSET) {
      const uint64_t* w = b.words();
      uint32_t c = 0;
      for (uint16_t v : a.data) c += (w[v >> 6] >> (v & 63)) & 1;
      return c;
   }
   if (a.kind == BITSET && b.kind == ARRAY) return containerAndCard(b, a);
   if (a.kind == RUN && b.kind == BITSET) {
      const uint64_t* w = b.words();
      uint32_t c = 0;
      for (uint32_t r = 0; r < a.runCount(); ++r) c += popcountRange(w, a.data[2 * r], a.data[2 * r] + a.data[2 * r + 1]);
      return c;
   }
   if (a.kind == BITSET && b.kind == RUN) return containerAndCard(b, a);
   if (a.kind == RUN && b.kind == ARRAY) {  // walk both
      uint32_t c = 0, r = 0;
      const uint32_t nr = a.runCount();
      for (uint16_t v : b.data) {
         while (r < nr && static_cast<uint32_t>(a.data[2 * r]) + a.data[2 * r + 1] < v) ++r;
         if (r == nr) break;
         if (a.data[2 * r] <= v) ++c;
      }
      return c;
   }
   if (a.kind == ARRAY && b.kind == RUN) return containerAndCard(b, a);
   // run x run: interval intersection
   uint32_t c = 0, i = 0, j = 0;
   const uint32_t ni = a.runCount(), nj = b.runCount();
   while (i < ni && j < nj) {
      uint32_t as = a.data[2 * i], ae = as + a.data[2 * i + 1];
      uint32_t bs = b.data[2 * j], be = bs + b.data[2 * j + 1];
      uint32_t lo = std::max(as, bs), hi = std::min(ae, be);
      if (lo <= hi) c += hi - lo + 1;
      if (ae < be) ++i; else ++j;
   }
   return c;
}

inline uint64_t andCardinality(const Bitmap& a, const Bitmap& b) {
   uint64_t c = 0;
   size_t i = 0, j = 0;
   while (i < a.containers.size() && j < b.containers.size()) {
      const uint16_t ka = a.containers[i].key, kb = b.containers[j].key;
      if (ka < kb) ++i;
      else if (ka > kb) ++j;
      else { c += containerAndCard(a.containers[i], b.containers[j]); ++i; ++j; }
   }
   return c;
}
inline uint64_t andnotCardinality(const Bitmap& a, const Bitmap& b) { return a.cardinality() - andCardinality(a, b); }

// Materialising binary ops. op: 0 = AND, 1 = OR, 2 = ANDNOT (a \ b). Result containers are
// canonical array/bitset (no runs), like CRoaring before runOptimize.
inline void loadWords(const Container& c, uint64_t* w) {
   if (c.kind == BITSET) { memcpy(w, c.words(), 8192); return; }
   memset(w, 0, 8192);
   if (c.kind == ARRAY) { for (uint16_t v : c.data) w[v >> 6] |= 1ULL << (v & 63); }
   else { for (uint32_t r = 0; r < c.runCount(); ++r) Container::setRange(w, c.data[2 * r], static_cast<uint32_t>(c.data[2 * r]) + c.data[2 * r + 1] + 1); }
}
inline Container fromWords(uint16_t key, const uint64_t* w) {
   Container c;
   c.key = key;
   c.kind = BITSET;
   c.data.resize(silo::roaring::BITSET_U16);
   memcpy(c.data.data(), w, 8192);
   c.normalizeFromBitset();
   return c;
}
inline Container canonicalCopy(const Container& c) { Container out = c; out.toArrayOrBitset(); return out; }

inline Container containerOp(const Container& a, const Container& b, int op) {
   if (a.kind == ARRAY && b.kind == ARRAY && (op != 1 || a.card + b.card <= silo::roaring::ARRAY_MAX)) {
      Container out;
      out.key = a.key;
      out.kind = ARRAY;
      const auto &x = a.data, &y = b.data;
      size_t i = 0, j = 0;
      while (i < x.size() && j < y.size()) {
         if (x[i] < y[j]) { if (op != 0) out.data.push_back(x[i]); ++i; }
         else if (x[i] > y[j]) { if (op == 1) out.data.push_back(y[j]); ++j; }
         else { if (op != 2) out.data.push_back(x[i]); ++i; ++j; }
      }
      if (op != 0) for (; i < x.size(); ++i) out.data.push_back(x[i]);
      if (op == 1) for (; j < y.size(); ++j) out.data.push_back(y[j]);
      out.card = static_cast<uint32_t>(out.data.size());
      return out;
   }
   if (a.kind == ARRAY && b.kind != ARRAY && op != 1) {  // probe: result is a subset of the array
      Container out;
      out.key = a.key;
      out.kind = ARRAY;
      for (uint16_t v : a.data) if (b.contains(v) == (op == 0)) out.data.push_back(v);
      out.card = static_cast<uint32_t>(out.data.size());
      return out;
   }
   uint64_t wa[1024], wb[1024];
   loadWords(a, wa);
   loadWords(b, wb);
   if (op == 0) for (int k = 0; k < 1024; ++k) wa[k] &= wb[k];
   else if (op == 1) for (int k = 0; k < 1024; ++k) wa[k] |= wb[k];
   else for (int k = 0; k < 1024; ++k) wa[k] &= ~wb[k];
   return fromWords(a.key, wa);
}

inline Bitmap bitmapOp(const Bitmap& a, const Bitmap& b, int op) {
   Bitmap out;
   size_t i = 0, j = 0;
   while (i < a.containers.size() || j < b.containers.size()) {
      const bool has_a = i < a.containers.size(), has_b = j < b.containers.size();
      if (has_a && (!has_b || a.containers[i].key < b.containers[j].key)) {
         if (op != 0) out.containers.push_back(canonicalCopy(a.containers[i]));
         ++i;
      } else if (has_b && (!has_a || b.containers[j].key < a.containers[i].key)) {
         if (op == 1) out.containers.push_back(canonicalCopy(b.containers[j]));
         ++j;
      } else {
         Container c = containerOp(a.containers[i], b.containers[j], op);
         if (c.card) out.containers.push_back(std::move(c));
         ++i; ++j;
      }
   }
   return out;
}
inline Bitmap bitmapAnd(const Bitmap& a, const Bitmap& b) { return bitmapOp(a, b, 0); }
inline Bitmap bitmapOr(const Bitmap& a, const Bitmap& b) { return bitmapOp(a, b, 1); }
inline Bitmap bitmapAndNot(const Bitmap& a, const Bitmap& b) { return bitmapOp(a, b, 2); }

// Roaring::fastunion (union.cpp:44)
inline Bitmap fastUnion(const std::vector<const Bitmap*>& inputs) {
   Bitmap out;
   for (const Bitmap* b : inputs) out = bitmapOr(out, *b);
   return out;
}

}  // namespace oracle
