// Host-side roaring bitmap model: container model + builder operations + the portable codec.
// Written from scratch against the public Roaring portable format; the
// reference delegates to CRoaring roaring/1.0.0 (conanfile.py:17) and serializes every bitmap with
// Roaring::write(buf, portable=true) (include/silo/roaring/roaring_serialize.h:18-45).
// The product only needs to BUILD, SERIALIZE, PARSE and ENUMERATE bitmaps on the host; all set
// algebra of the query path runs on the GPU (csrc/), never here.
#pragma once
#include <algorithm>
#include <cstdint>
#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>

namespace silo::roaring {

enum ContainerKind : uint8_t { ARRAY = 0, BITSET = 1, RUN = 2 };
constexpr uint32_t ARRAY_MAX = 4096;  // cardinality <= 4096 -> array container
constexpr uint32_t BITSET_U16 = 4096; // 8 KB = 1024 x uint64 = 4096 x uint16

struct Container {
   uint16_t key = 0;
   uint8_t kind = ARRAY;
   uint32_t card = 0;
   // ARRAY: sorted values; BITSET: 4096 uint16 (= 1024 little-endian uint64 words); RUN: (start, len-1) pairs
   std::vector<uint16_t> data;

   uint64_t* words() { return reinterpret_cast<uint64_t*>(data.data()); }
   const uint64_t* words() const { return reinterpret_cast<const uint64_t*>(data.data()); }
   uint32_t runCount() const { return static_cast<uint32_t>(data.size() / 2); }

   bool contains(uint16_t v) const {
      if (kind == BITSET) return (words()[v >> 6] >> (v & 63)) & 1;
      if (kind == ARRAY) return std::binary_search(data.begin(), data.end(), v);
      uint32_t lo = 0, hi = runCount();
      while (lo < hi) {  // last run with start <= v
         uint32_t mid = (lo + hi) / 2;
         if (data[2 * mid] <= v) lo = mid + 1; else hi = mid;
      }
      if (lo == 0) return false;
      uint32_t start = data[2 * (lo - 1)], len1 = data[2 * (lo - 1) + 1];
      return v - start <= len1;
   }

   void toBitset() {
      if (kind == BITSET) return;
      std::vector<uint16_t> bits(BITSET_U16, 0);
      uint64_t* w = reinterpret_cast<uint64_t*>(bits.data());
      if (kind == ARRAY) {
         for (uint16_t v : data) w[v >> 6] |= 1ULL << (v & 63);
      } else {
         for (uint32_t r = 0; r < runCount(); ++r) {
            uint32_t s = data[2 * r], e = s + data[2 * r + 1];
            setRange(w, s, e + 1);
         }
      }
      data.swap(bits);
      kind = BITSET;
   }
   static void setRange(uint64_t* w, uint32_t lo, uint32_t hi) {  // [lo, hi) within 0..65536
      if (lo >= hi) return;
      uint32_t fw = lo >> 6, lw = (hi - 1) >> 6;
      uint64_t fm = ~0ULL << (lo & 63), lm = ~0ULL >> (63 - ((hi - 1) & 63));
      if (fw == lw) { w[fw] |= fm & lm; return; }
      w[fw] |= fm;
      for (uint32_t k = fw + 1; k < lw; ++k) w[k] = ~0ULL;
      w[lw] |= lm;
   }
   static void flipRange(uint64_t* w, uint32_t lo, uint32_t hi) {
      if (lo >= hi) return;
      uint32_t fw = lo >> 6, lw = (hi - 1) >> 6;
      uint64_t fm = ~0ULL << (lo & 63), lm = ~0ULL >> (63 - ((hi - 1) & 63));
      if (fw == lw) { w[fw] ^= fm & lm; return; }
      w[fw] ^= fm;
      for (uint32_t k = fw + 1; k < lw; ++k) w[k] = ~w[k];
      w[lw] ^= lm;
   }
   // Re-derive cardinality from a bitset and pick array/bitset by the roaring size rule.
   void normalizeFromBitset() {
      const uint64_t* w = words();
      uint32_t c = 0;
      for (uint32_t k = 0; k < 1024; ++k) c += __builtin_popcountll(w[k]);
      card = c;
      if (c <= ARRAY_MAX) {
         std::vector<uint16_t> values;
         values.reserve(c);
         for (uint32_t k = 0; k < 1024; ++k) {
            uint64_t x = w[k];
            while (x) { values.push_back(static_cast<uint16_t>(k * 64 + __builtin_ctzll(x))); x &= x - 1; }
         }
         data.swap(values);
         kind = ARRAY;
      }
   }
   void toArrayOrBitset() {  // RUN -> canonical non-run form
      if (kind != RUN) return;
      toBitset();
      normalizeFromBitset();
   }
   uint32_t countRuns() const {
      if (kind == RUN) return runCount();
      if (kind == ARRAY) {
         uint32_t runs = 0;
         for (size_t k = 0; k < data.size(); ++k)
            if (k == 0 || data[k] != data[k - 1] + 1) ++runs;
         return runs;
      }
      const uint64_t* w = words();
      uint32_t runs = 0;
      uint64_t prev_top = 0;
      for (uint32_t k = 0; k < 1024; ++k) {
         uint64_t x = w[k];
         runs += __builtin_popcountll(x & ~((x << 1) | prev_top));
         prev_top = x >> 63;
      }
      return runs;
   }
   // CRoaring runOptimize rule: convert iff the run form is strictly smaller
   // (array: 2 + 4r < 2c + 2 ; bitset: 2 + 4r < 8192). SURVEY Appendix A.
   void runOptimize() {
      if (kind == RUN) {  // convert back if another form became smaller
         uint32_t r = runCount();
         uint32_t run_size = 2 + 4 * r;
         uint32_t other = card <= ARRAY_MAX ? 2 * card + 2 : 8192;
         if (run_size < other) return;
         toArrayOrBitset();
         return;
      }
      uint32_t r = countRuns();
      uint32_t run_size = 2 + 4 * r;
      uint32_t cur = kind == ARRAY ? 2 * card + 2 : 8192;
      if (run_size >= cur) return;
      std::vector<uint16_t> runs;
      runs.reserve(2 * r);
      if (kind == ARRAY) {
         size_t k = 0;
         while (k < data.size()) {
            size_t j = k;
            while (j + 1 < data.size() && data[j + 1] == data[j] + 1) ++j;
            runs.push_back(data[k]);
            runs.push_back(static_cast<uint16_t>(data[j] - data[k]));
            k = j + 1;
         }
      } else {
         const uint64_t* w = words();
         int32_t start = -1;
         for (uint32_t v = 0; v < 65536; ++v) {
            bool bit = (w[v >> 6] >> (v & 63)) & 1;
            if (bit && start < 0) start = static_cast<int32_t>(v);
            if (!bit && start >= 0) { runs.push_back(start); runs.push_back(static_cast<uint16_t>(v - 1 - start)); start = -1; }
            if (!bit && (v & 63) == 0 && w[v >> 6] == 0) v += 63;  // skip empty word
         }
         if (start >= 0) { runs.push_back(start); runs.push_back(static_cast<uint16_t>(65535 - start)); }
      }
      data.swap(runs);
      kind = RUN;
   }
   template <typename F>
   void forEach(F&& f) const {
      uint32_t base = static_cast<uint32_t>(key) << 16;
      if (kind == ARRAY) { for (uint16_t v : data) f(base | v); }
      else if (kind == RUN) {
         for (uint32_t r = 0; r < runCount(); ++r) {
            uint32_t s = data[2 * r], e = s + data[2 * r + 1];
            for (uint32_t v = s; v <= e; ++v) f(base | v);
         }
      } else {
         const uint64_t* w = words();
         for (uint32_t k = 0; k < 1024; ++k) {
            uint64_t x = w[k];
            while (x) { f(base | (k * 64 + __builtin_ctzll(x))); x &= x - 1; }
         }
      }
   }
};

// A set of uint32 values; containers sorted by key, never empty.
class Bitmap {
  public:
   std::vector<Container> containers;

   Bitmap() = default;
   Bitmap(std::initializer_list<uint32_t> values) { for (uint32_t v : values) add(v); }

   bool isEmpty() const { return containers.empty(); }
   uint64_t cardinality() const {
      uint64_t c = 0;
      for (const auto& ct : containers) c += ct.card;
      return c;
   }
   int findContainer(uint16_t key) const {
      int lo = 0, hi = static_cast<int>(containers.size());
      while (lo < hi) { int mid = (lo + hi) / 2; if (containers[mid].key < key) lo = mid + 1; else hi = mid; }
      return lo;
   }
   bool contains(uint32_t v) const {
      uint16_t key = v >> 16;
      int idx = findContainer(key);
      if (idx >= static_cast<int>(containers.size()) || containers[idx].key != key) return false;
      return containers[idx].contains(static_cast<uint16_t>(v));
   }
   Container& getOrCreate(uint16_t key) {
      if (!containers.empty() && containers.back().key == key) return containers.back();
      if (containers.empty() || containers.back().key < key) {
         containers.emplace_back();
         containers.back().key = key;
         return containers.back();
      }
      int idx = findContainer(key);
      if (containers[idx].key == key) return containers[idx];
      Container c; c.key = key;
      return *containers.insert(containers.begin() + idx, std::move(c));
   }
   void add(uint32_t v) {
      Container& c = getOrCreate(v >> 16);
      uint16_t low = static_cast<uint16_t>(v);
      if (c.kind == RUN) c.toArrayOrBitset();
      if (c.kind == ARRAY) {
         if (c.data.empty() || c.data.back() < low) { c.data.push_back(low); ++c.card; }
         else {
            auto it = std::lower_bound(c.data.begin(), c.data.end(), low);
            if (it != c.data.end() && *it == low) return;
            c.data.insert(it, low);
            ++c.card;
         }
         if (c.card > ARRAY_MAX) c.toBitset();
      } else {
         uint64_t& w = c.words()[low >> 6];
         uint64_t m = 1ULL << (low & 63);
         if (!(w & m)) { w |= m; ++c.card; }
      }
   }
   void addMany(size_t n, const uint32_t* values) { for (size_t k = 0; k < n; ++k) add(values[k]); }
   void dropEmpty() {
      containers.erase(std::remove_if(containers.begin(), containers.end(), [](const Container& c) { return c.card == 0; }), containers.end());
   }
   // [lo, hi) as uint64 so that hi = 2^32 is expressible (Roaring::addRange semantics)
   void addRange(uint64_t lo, uint64_t hi) {
      if (lo >= hi) return;
      for (uint64_t key = lo >> 16; key <= ((hi - 1) >> 16); ++key) {
         uint64_t base = key << 16;
         uint32_t s = static_cast<uint32_t>(std::max(lo, base) - base);
         uint32_t e = static_cast<uint32_t>(std::min(hi, base + 65536) - base);  // exclusive
         Container& c = getOrCreate(static_cast<uint16_t>(key));
         if (c.card == 0 && c.kind == ARRAY) {  // fresh container: store as a run directly if smaller
            c.kind = RUN; c.data = {static_cast<uint16_t>(s), static_cast<uint16_t>(e - 1 - s)}; c.card = e - s;
            c.toArrayOrBitset();
            continue;
         }
         c.toBitset();
         Container::setRange(c.words(), s, e);
         c.normalizeFromBitset();
      }
   }
   // Roaring::flip(lo, hi): toggles membership of every value in [lo, hi)
   void flip(uint64_t lo, uint64_t hi) {
      if (lo >= hi) return;
      for (uint64_t key = lo >> 16; key <= ((hi - 1) >> 16); ++key) {
         uint64_t base = key << 16;
         uint32_t s = static_cast<uint32_t>(std::max(lo, base) - base);
         uint32_t e = static_cast<uint32_t>(std::min(hi, base + 65536) - base);
         Container& c = getOrCreate(static_cast<uint16_t>(key));
         c.toBitset();
         Container::flipRange(c.words(), s, e);
         c.normalizeFromBitset();
      }
      dropEmpty();
   }
   void runOptimize() { for (auto& c : containers) c.runOptimize(); }
   void removeRunCompression() { for (auto& c : containers) c.toArrayOrBitset(); }

   template <typename F>
   void forEach(F&& f) const { for (const auto& c : containers) c.forEach(f); }
   std::vector<uint32_t> toVector() const {
      std::vector<uint32_t> out;
      out.reserve(cardinality());
      forEach([&](uint32_t v) { out.push_back(v); });
      return out;
   }
   uint32_t maximum() const {
      uint32_t m = 0;
      if (containers.empty()) return 0;
      containers.back().forEach([&](uint32_t v) { m = v; });
      return m;
   }
   // Build from ascending, duplicate-free ids; the canonical array/bitset choice per container.
   static Bitmap fromSorted(const uint32_t* ids, size_t n) {
      Bitmap bm;
      size_t k = 0;
      while (k < n) {
         uint16_t key = ids[k] >> 16;
         size_t j = k;
         while (j < n && (ids[j] >> 16) == key) ++j;
         Container c; c.key = key; c.card = static_cast<uint32_t>(j - k);
         if (c.card <= ARRAY_MAX) {
            c.kind = ARRAY; c.data.resize(c.card);
            for (size_t t = k; t < j; ++t) c.data[t - k] = static_cast<uint16_t>(ids[t]);
         } else {
            c.kind = BITSET; c.data.assign(BITSET_U16, 0);
            uint64_t* w = c.words();
            for (size_t t = k; t < j; ++t) { uint16_t v = static_cast<uint16_t>(ids[t]); w[v >> 6] |= 1ULL << (v & 63); }
         }
         bm.containers.push_back(std::move(c));
         k = j;
      }
      return bm;
   }
   bool operator==(const Bitmap& other) const { return toVector() == other.toVector(); }

   // ---- CRoaring portable format ----
   static constexpr uint32_t SERIAL_COOKIE_NO_RUN = 12346;
   static constexpr uint32_t SERIAL_COOKIE = 12347;
   static constexpr uint32_t NO_OFFSET_THRESHOLD = 4;

   static uint32_t payloadBytes(const Container& c) {
      if (c.kind == RUN) return 2 + 4 * c.runCount();
      if (c.kind == ARRAY) return 2 * c.card;
      return 8192;
   }
   size_t portableSizeInBytes() const {
      const size_t n = containers.size();
      bool has_run = false;
      for (const auto& c : containers) has_run |= c.kind == RUN;
      size_t size = has_run ? 4 + (n + 7) / 8 : 8;
      size += 4 * n;
      if (!has_run || n >= NO_OFFSET_THRESHOLD) size += 4 * n;
      for (const auto& c : containers) size += payloadBytes(c);
      return size;
   }
   // Roaring::getSizeInBytes(portable=false): min(portable, 4*card + 4) + 1
   size_t nonPortableSizeInBytes() const {
      const size_t portable = portableSizeInBytes();
      const uint64_t as_array = 4 * cardinality() + 4;
      return static_cast<size_t>(std::min<uint64_t>(portable, as_array)) + 1;
   }
   void writePortable(std::string& out) const {
      const uint32_t n = static_cast<uint32_t>(containers.size());
      bool has_run = false;
      for (const auto& c : containers) has_run |= c.kind == RUN;
      auto put32 = [&](uint32_t v) { out.append(reinterpret_cast<const char*>(&v), 4); };
      auto put16 = [&](uint16_t v) { out.append(reinterpret_cast<const char*>(&v), 2); };
      const size_t start = out.size();
      if (has_run) {
         put32(SERIAL_COOKIE | ((n - 1) << 16));
         std::string flags((n + 7) / 8, '\0');
         for (uint32_t k = 0; k < n; ++k) if (containers[k].kind == RUN) flags[k / 8] |= static_cast<char>(1 << (k % 8));
         out += flags;
      } else {
         put32(SERIAL_COOKIE_NO_RUN);
         put32(n);
      }
      for (const auto& c : containers) { put16(c.key); put16(static_cast<uint16_t>(c.card - 1)); }
      if (!has_run || n >= NO_OFFSET_THRESHOLD) {
         uint32_t offset = static_cast<uint32_t>(out.size() - start) + 4 * n;
         for (const auto& c : containers) { put32(offset); offset += payloadBytes(c); }
      }
      for (const auto& c : containers) {
         if (c.kind == RUN) put16(static_cast<uint16_t>(c.runCount()));
         out.append(reinterpret_cast<const char*>(c.data.data()), c.data.size() * 2);
      }
   }
   std::string toPortable() const { std::string out; out.reserve(portableSizeInBytes()); writePortable(out); return out; }

   // Roaring::readSafe equivalent: bounds-checked parse; returns bytes consumed.
   static Bitmap readPortable(const void* buffer, size_t len, size_t* consumed = nullptr) {
      const uint8_t* p = static_cast<const uint8_t*>(buffer);
      size_t pos = 0;
      auto need = [&](size_t k) { if (pos + k > len) throw std::runtime_error("roaring portable: truncated buffer"); };
      auto get32 = [&]() { need(4); uint32_t v; memcpy(&v, p + pos, 4); pos += 4; return v; };
      auto get16 = [&]() { need(2); uint16_t v; memcpy(&v, p + pos, 2); pos += 2; return v; };
      uint32_t cookie = get32();
      uint32_t n;
      bool has_run = false;
      std::vector<uint8_t> flags;
      if ((cookie & 0xFFFF) == SERIAL_COOKIE) {
         has_run = true;
         n = (cookie >> 16) + 1;
         need((n + 7) / 8);
         flags.assign(p + pos, p + pos + (n + 7) / 8);
         pos += (n + 7) / 8;
      } else if (cookie == SERIAL_COOKIE_NO_RUN) {
         n = get32();
      } else {
         throw std::runtime_error("roaring portable: bad cookie");
      }
      if (n > 65536) throw std::runtime_error("roaring portable: too many containers");
      Bitmap bm;
      bm.containers.resize(n);
      for (uint32_t k = 0; k < n; ++k) {
         bm.containers[k].key = get16();
         bm.containers[k].card = static_cast<uint32_t>(get16()) + 1;
         if (k > 0 && bm.containers[k].key <= bm.containers[k - 1].key) throw std::runtime_error("roaring portable: keys not ascending");
      }
      if (!has_run || n >= NO_OFFSET_THRESHOLD) { need(4 * static_cast<size_t>(n)); pos += 4 * static_cast<size_t>(n); }
      for (uint32_t k = 0; k < n; ++k) {
         Container& c = bm.containers[k];
         const bool is_run = has_run && ((flags[k / 8] >> (k % 8)) & 1);
         if (is_run) {
            uint32_t r = get16();
            need(4 * static_cast<size_t>(r));
            c.kind = RUN; c.data.resize(2 * r);
            memcpy(c.data.data(), p + pos, 4 * static_cast<size_t>(r)); pos += 4 * static_cast<size_t>(r);
            uint32_t card = 0;
            for (uint32_t t = 0; t < r; ++t) card += static_cast<uint32_t>(c.data[2 * t + 1]) + 1;
            if (card != c.card) throw std::runtime_error("roaring portable: run cardinality mismatch");
         } else if (c.card <= ARRAY_MAX) {
            need(2 * static_cast<size_t>(c.card));
            c.kind = ARRAY; c.data.resize(c.card);
            memcpy(c.data.data(), p + pos, 2 * static_cast<size_t>(c.card)); pos += 2 * static_cast<size_t>(c.card);
         } else {
            need(8192);
            c.kind = BITSET; c.data.resize(BITSET_U16);
            memcpy(c.data.data(), p + pos, 8192); pos += 8192;
         }
      }
      if (consumed) *consumed = pos;
      return bm;
   }
};

}  // namespace silo::roaring
