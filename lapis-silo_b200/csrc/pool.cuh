// Host side of the HBM container pool: ingest of CRoaring portable bytes, packing into the tiled
// directory + payload pools, upload. Format parsed here: SURVEY Appendix A (the reference writes it
// through include/silo/roaring/roaring_serialize.h:18-45).
#pragma once
#include <algorithm>

#include "common.cuh"
#include "kernels_filter.cuh"

namespace sb {

// A container inside a caller's portable buffer. The payload may sit at ANY byte offset (the run-flag bitset makes it odd for
// many container counts): it is only ever read through memcpy.
struct ParsedContainer {
   uint16_t key; uint8_t kind; uint32_t size; uint32_t card;   // size: array card / #runs / bitset card
   const uint8_t* bytes;
   uint16_t u16(size_t k) const { uint16_t v; memcpy(&v, bytes + 2 * k, 2); return v; }
   uint64_t u64(size_t k) const { uint64_t v; memcpy(&v, bytes + 8 * k, 8); return v; }
};

// Roaring::readSafe equivalent; calls f(ParsedContainer) for every container, ascending keys.
template <typename F>
inline void parsePortable(const uint8_t* p, size_t len, F&& f) {
   if (len == 0) return;
   size_t pos = 0;
   auto need = [&](size_t k) { if (pos + k > len) throw Error("roaring portable: truncated buffer"); };
   auto get32 = [&]() { need(4); uint32_t v; memcpy(&v, p + pos, 4); pos += 4; return v; };
   const uint32_t cookie = get32();
   uint32_t n;
   const uint8_t* run_flags = nullptr;
   if ((cookie & 0xFFFF) == 12347) {
      n = (cookie >> 16) + 1;
      need((n + 7) / 8);
      run_flags = p + pos;
      pos += (n + 7) / 8;
   } else if (cookie == 12346) {
      n = get32();
   } else {
      throw Error("roaring portable: bad cookie");
   }
   SB_REQUIRE(n <= 65536, "roaring portable: too many containers");
   need(4 * size_t(n));
   const uint8_t* header = p + pos;
   pos += 4 * size_t(n);
   if (!run_flags || n >= 4) { need(4 * size_t(n)); pos += 4 * size_t(n); }
   for (uint32_t k = 0; k < n; ++k) {
      uint16_t key, card1;
      memcpy(&key, header + 4 * k, 2);
      memcpy(&card1, header + 4 * k + 2, 2);
      ParsedContainer c;
      c.key = key;
      c.card = uint32_t(card1) + 1;
      const bool is_run = run_flags && ((run_flags[k / 8] >> (k % 8)) & 1);
      if (is_run) {
         need(2);
         uint16_t r;
         memcpy(&r, p + pos, 2);
         pos += 2;
         need(4 * size_t(r));
         SB_REQUIRE(r > 0, "roaring portable: empty run container");
         c.kind = KIND_RUN; c.size = r; c.bytes = p + pos;
         pos += 4 * size_t(r);
      } else if (c.card <= 4096) {
         need(2 * size_t(c.card));
         c.kind = KIND_ARRAY; c.size = c.card; c.bytes = p + pos;
         pos += 2 * size_t(c.card);
      } else {
         need(8192);
         c.kind = KIND_BITSET; c.size = c.card; c.bytes = p + pos;
         pos += 8192;
      }
      f(c);
   }
}

inline uint32_t payloadU16(const ParsedContainer& c) { return c.kind == KIND_ARRAY ? c.size : c.kind == KIND_RUN ? 2 * c.size : 4096; }
inline void appendPayload(std::vector<uint16_t>& dst, const ParsedContainer& c) {
   const size_t at = dst.size(), n16 = payloadU16(c);
   dst.resize(at + n16);
   memcpy(dst.data() + at, c.bytes, n16 * 2);
}

struct TmpContainer { uint32_t slot; uint16_t key; uint8_t kind; uint32_t size; uint64_t pay; };

struct HostStore {   // one SequenceStorePartition as ingested, before packing
   uint32_t partition = 0, seq = 0, n_rows = 0;
   std::vector<TmpContainer> containers;     // ascending (slot, key)
   std::vector<uint16_t> payload;
   std::vector<int8_t> flipped, deleted;
   std::vector<uint32_t> miss_row_off;       // n_rows + 1
   std::vector<uint2> miss_iv;
   std::vector<uint32_t> miss_row_bytes;     // payload(M_i) per row
   uint64_t alg_bytes = 0;                   // sum over containers of payload + 8 (directory)
   uint64_t dir_bytes = 0;                   // 8 per container
   uint64_t miss_alg_bytes = 0;              // sum over rows of payload(M_i) + 4
};

template <typename T>
struct DeviceBuffer {
   T* ptr = nullptr;
   size_t count = 0;
   DeviceBuffer() = default;
   DeviceBuffer(const DeviceBuffer&) = delete;
   DeviceBuffer& operator=(const DeviceBuffer&) = delete;
   ~DeviceBuffer() { if (ptr) cudaFree(ptr); }
   void alloc(size_t n) {
      if (ptr) { cudaFree(ptr); ptr = nullptr; }
      count = n;
      SB_CUDA(cudaMalloc(reinterpret_cast<void**>(&ptr), std::max<size_t>(n, 1) * sizeof(T)));
   }
   void upload(const std::vector<T>& v) {
      alloc(v.size());
      if (!v.empty()) SB_CUDA(cudaMemcpy(ptr, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice));
   }
   size_t bytes() const { return count * sizeof(T); }
};

struct Sequence {   // one logical sequence (store name) over all partitions of this db
   std::string name;
   int alphabet = 0;
   uint32_t length = 0, sigma = 0, tile_positions = TILE_POSITIONS, n_tiles = 0;
   std::vector<int> store_of_partition;   // index into Db::stores or -1
   // device
   DeviceBuffer<uint32_t> postings;
   DeviceBuffer<LargeEntry> large_entries;
   DeviceBuffer<WorkItem> items;
   DeviceBuffer<SegRec> segs;
   DeviceBuffer<SlabRec> slabs;
   DeviceBuffer<RunRec> runs;
   DeviceBuffer<uint32_t> slab_begin, part_rows;
   DeviceBuffer<uint4> large_pool;
   std::vector<uint32_t> h_slab_begin;    // [P + 1]
   std::vector<uint32_t> h_part_rows;     // [P] sequence_count of the partition's store (0 = none)
   DeviceBuffer<int8_t> deleted, flipped;
   DeviceBuffer<uint32_t> miss_row_off;
   DeviceBuffer<uint2> miss_iv;
   DeviceBuffer<uint64_t> miss_row_base, miss_iv_base, miss_tile_base, miss_inline_base, miss_any_base;
   DeviceBuffer<uint32_t> miss_any;
   DeviceBuffer<uint4> miss_inline;
   DeviceBuffer<uint32_t> miss_tile_bits;
   DeviceBuffer<uint32_t> full_tables;   // [P][sigma * length]
   DeviceBuffer<uint8_t> reference;      // [length] symbols of the reference sequence (row selection)
   SeqDev dev{};
   uint32_t n_items = 0, n_slabs = 0, n_runs = 0;
   uint32_t k1_resident[8] = {1, 1, 1, 1, 1, 1, 1, 1};   // CTAs of k1_scan per SM, per compiled variant
   uint64_t n_postings = 0, n_large = 0;
};

struct RegisteredBitmap { uint32_t partition; uint32_t first_container, n_containers; uint32_t max_value; };   // containers in ascending key order

struct Column { uint32_t partition; int dtype; uint32_t n; void* device = nullptr; };
struct GroupColumn { uint32_t partition; uint32_t n; uint32_t* device = nullptr; };   // dictionary codes of a metadata column (group-by)

inline uint64_t itemWeightCap() {   // bytes of postings + large payload per Mutations work item (SB_ITEM_CAP_KB: tuning)
   const char* e = getenv("SB_ITEM_CAP_KB");
   return (e && atoi(e) > 0 ? uint64_t(atoi(e)) : 128) * 1024;
}

// bytes of postings + large payload per RUN of work items (one slab, consecutive tiles): what a CTA of the scan takes at a time.
// Large enough to pay for the 32 KB filter copy, small enough that every resident CTA gets several (SB_RUN_CAP_KB: tuning).
// Items per run: a CTA walks the items of its run one after the other (each costs a few dependent memory round trips whatever it
// holds), so runs of many tiny items serialise a small database on a handful of CTAs -- the 1,000-sequence dataset (1,170 items)
// took 148 us in 20 runs of up to 64 items. Two items per run keep half of the filter copies away and every CTA busy.
constexpr uint32_t RUN_MAX_ITEMS = 2;
inline uint64_t runWeightCap(uint64_t total_weight) {
   const char* e = getenv("SB_RUN_CAP_KB");
   if (e && atoi(e) > 0) return uint64_t(atoi(e)) * 1024;
   (void)total_weight;
   return 96 * 1024;      // measured at 40 M rows: 1 item per run, 96 KB and 256 KB runs take 0.177 / 0.177 / 0.195 ms (long runs cost balance at the end of the scan)
}

// Containers up to this cardinality become postings (SB_FLAT_MAX: tuning). A container that stays in the large pool costs a warp
// two dependent memory round trips (directory entry, payload) whatever it holds: a one-tile gene of the amino-acid stores packs
// hundreds of 20-60-value arrays into one work item, which then took 200-380 us at 10 M rows.
inline uint32_t flatMax() {
   const char* e = getenv("SB_FLAT_MAX");
   const uint32_t v = e && atoi(e) > 0 ? uint32_t(atoi(e)) : FLAT_MAX;
   return std::min<uint32_t>(v, 1024u);
}

inline uint32_t arraySplit() {   // values per large-array entry (SB_ARRAY_SPLIT: tuning; a multiple of 8 keeps the pieces 16-byte aligned)
   const char* e = getenv("SB_ARRAY_SPLIT");
   const uint32_t v = e && atoi(e) >= 64 ? uint32_t(atoi(e)) : 1024u;
   return (v + 7u) & ~7u;
}

// Packs the ingested stores of one sequence into the tiled device layout.
inline void packSequence(Sequence& seq, std::vector<std::unique_ptr<HostStore>>& stores, uint32_t n_partitions, sb_pool_stats& stats) {
   if (const char* e = getenv("SB_TILE_POSITIONS")) {      // tuning: 256 or 512 positions per tile (512 only while the tile's slots fit the posting's 14-bit field)
      const uint32_t v = uint32_t(atoi(e));
      if ((v == 256 || v == 512) && v * seq.sigma < 0x4000) seq.tile_positions = v;
   }
   const uint32_t sigma = seq.sigma, tp = seq.tile_positions, L = seq.length;
   seq.n_tiles = (L + tp - 1) / tp;
   const uint32_t slots_per_tile = tp * sigma;
   SB_REQUIRE(slots_per_tile < 0x4000, "tile does not fit the 14-bit slot field (one value is the dummy slot)");
   SB_REQUIRE(seq.n_tiles <= 0x4000, "sequence longer than 16,384 tiles of positions");
   // slabs: 4 container keys each
   std::vector<SlabRec> slabs;
   seq.h_slab_begin.assign(n_partitions + 1, 0);
   for (uint32_t p = 0; p < n_partitions; ++p) {
      seq.h_slab_begin[p] = uint32_t(slabs.size());
      if (seq.store_of_partition[p] < 0) continue;
      const uint32_t n_rows = stores[seq.store_of_partition[p]]->n_rows;
      const uint32_t n_keys = std::max<uint32_t>(1, uint32_t((uint64_t(n_rows) + 65535) / 65536));
      for (uint32_t k = 0; k < n_keys; k += SLAB_KEYS) slabs.push_back({p, k});
   }
   seq.h_slab_begin[n_partitions] = uint32_t(slabs.size());
   const uint32_t n_slabs = uint32_t(slabs.size());
   seq.n_slabs = n_slabs;

   std::vector<uint32_t> postings;
   std::vector<LargeEntry> large_entries;
   std::vector<WorkItem> items;
   std::vector<uint64_t> item_weight;
   std::vector<SegRec> segs(size_t(seq.n_tiles) * std::max<uint32_t>(n_slabs, 1) + 1);
   std::vector<uint64_t> large_pool;   // 16-byte units = 2 words each
   std::vector<size_t> cursor(n_partitions, 0);
   uint64_t total_containers = 0;
   for (uint32_t p = 0; p < n_partitions; ++p) if (seq.store_of_partition[p] >= 0) total_containers += stores[seq.store_of_partition[p]]->containers.size();
   postings.reserve(total_containers * 4);
   struct Picked { uint32_t slot2; const TmpContainer* c; };
   std::vector<Picked> picked;
   uint64_t hist[16] = {};   // SB_TRACE: array cardinalities 1,2,3,4,<=8,<=16,<=32,<=64,<=256,<=4096 | runs | bitsets
   const uint32_t filler = slots_per_tile << POSTING_ROW_BITS;
   const uint64_t item_cap = itemWeightCap();
   const uint32_t array_split = arraySplit();
   const uint32_t flat_max = flatMax();
   for (uint32_t tile = 0; tile < seq.n_tiles; ++tile) {
      const uint32_t slot_end = (tile + 1) * slots_per_tile;
      for (uint32_t slab_id = 0; slab_id < n_slabs; ++slab_id) {
         while (postings.size() % 4) postings.push_back(filler);
         segs[size_t(tile) * n_slabs + slab_id] = {uint32_t(postings.size()), uint32_t(large_entries.size())};
         const uint32_t p = slabs[slab_id].partition, key0 = slabs[slab_id].key0;
         HostStore& st = *stores[seq.store_of_partition[p]];
         // the (tile, partition) range of the store's containers; pick this slab's keys, order by (slot', key)
         const size_t begin = cursor[p];
         size_t end = begin;
         while (end < st.containers.size() && st.containers[end].slot < slot_end) ++end;
         if (slab_id + 1 == seq.h_slab_begin[p + 1]) cursor[p] = end;   // last slab of the partition consumes the range
         picked.clear();
         for (size_t k = begin; k < end; ++k) {
            const TmpContainer& c = st.containers[k];
            if (c.key < key0 || c.key >= key0 + SLAB_KEYS) continue;
            const uint32_t local = c.slot - tile * slots_per_tile, pl = local / sigma, sym = local - pl * sigma;
            picked.push_back({sym * tp + pl, &c});
         }
         if (picked.empty()) continue;
         std::stable_sort(picked.begin(), picked.end(), [](const Picked& a, const Picked& b) { return a.slot2 < b.slot2; });   // keys stay ascending inside a slot
         const size_t post_begin = postings.size(), large_begin = large_entries.size();
         uint64_t large_bytes = 0;
         for (const Picked& pk : picked) {
            const TmpContainer& c = *pk.c;
            const uint16_t* src = st.payload.data() + c.pay;
            const uint32_t key_local = c.key - key0;
            uint64_t cardinality = c.size;
            if (c.kind == KIND_RUN) { cardinality = 0; for (uint32_t r = 0; r < c.size; ++r) cardinality += uint64_t(src[2 * r + 1]) + 1; }
            if (c.kind == KIND_ARRAY) hist[c.size <= 4 ? c.size - 1 : c.size <= 8 ? 4 : c.size <= 16 ? 5 : c.size <= 32 ? 6 : c.size <= 64 ? 7 : c.size <= 256 ? 8 : 9]++;
            else hist[c.kind == KIND_RUN ? 10 : 11]++;
            if (c.kind == KIND_BITSET) { stats.bitset_containers++; stats.bitset_values += c.size; }
            else if (c.kind == KIND_ARRAY) { stats.array_containers++; stats.array_values += c.size; }
            else { stats.run_containers++; stats.run_count += c.size; stats.run_values += cardinality; }
            const uint32_t head = (pk.slot2 << POSTING_ROW_BITS) | (key_local << 16);
            if (c.kind == KIND_ARRAY && cardinality <= flat_max) {
               for (uint32_t i = 0; i < c.size; ++i) postings.push_back(head | src[i]);
            } else if (c.kind == KIND_RUN && cardinality <= flat_max) {
               for (uint32_t r = 0; r < c.size; ++r) for (uint32_t v = src[2 * r], last = uint32_t(src[2 * r]) + src[2 * r + 1]; v <= last; ++v) postings.push_back(head | v);
            } else {
               // a big array container becomes several entries of at most `array_split` values: an entry is one warp's unit of
               // work in the scan, and a 4,096-value array in one piece is a straggler (set semantics: the pieces are disjoint)
               const uint32_t n16_all = c.kind == KIND_BITSET ? 4096 : c.kind == KIND_ARRAY ? c.size : 2 * c.size;
               const uint32_t piece = c.kind == KIND_ARRAY ? array_split : n16_all;
               for (uint32_t first = 0; first < n16_all; first += piece) {
                  const uint32_t n16 = std::min(piece, n16_all - first);
                  const size_t at = large_pool.size();
                  SB_REQUIRE(at / 2 < 0xFFFFFFFFull, "large pool beyond 64 GB");
                  large_pool.resize(at + size_t((n16 + 7) / 8) * 2, 0);
                  memcpy(large_pool.data() + at, src + first, size_t(n16) * 2);
                  large_entries.push_back({uint32_t(at / 2), c.kind == KIND_ARRAY ? n16 : c.size, pk.slot2 | (uint32_t(c.kind) << 14) | (key_local << 16) | (tile << 18), slab_id});
                  large_bytes += 16 + uint64_t((n16 + 7) / 8) * 16;
               }
            }
         }
         // cut the segment into weight-bounded work items: postings and large entries are split evenly
         const size_t n_post = postings.size() - post_begin, n_large = large_entries.size() - large_begin;
         const uint64_t weight = 4ull * n_post + large_bytes;
         const uint32_t n_cuts = uint32_t(std::max<uint64_t>(1, (weight + item_cap - 1) / item_cap));
         for (uint32_t cut = 0; cut < n_cuts; ++cut) {
            auto postCut = [&](uint32_t c) { return c == n_cuts ? n_post : (n_post * c / n_cuts) & ~size_t(3); };
            const size_t pb = postCut(cut), pe = postCut(cut + 1), lb = n_large * cut / n_cuts, le = n_large * (cut + 1) / n_cuts;
            if (pb == pe && lb == le) continue;
            items.push_back({tile, slab_id, uint32_t(post_begin + pb), uint32_t(post_begin + pe), uint32_t(large_begin + lb), uint32_t(large_begin + le), p, key0});
            item_weight.push_back(4ull * (pe - pb) + (n_large ? large_bytes * (le - lb) / n_large : 0));
         }
         SB_REQUIRE(postings.size() < 0xFFFFFFF0ull, "more than 2^32 postings in one sequence");
      }
   }
   // Items in (slab, tile) order, cut into RUNS: consecutive items of one slab up to a weight cap. A run is what a persistent
   // CTA of the scan takes from the device counter (longest-processing-time hand-out: heaviest run first); inside a run the
   // slab's filter slice stays in shared memory.
   std::vector<RunRec> runs;
   {
      std::vector<uint32_t> order(items.size());
      for (uint32_t k = 0; k < order.size(); ++k) order[k] = k;
      std::stable_sort(order.begin(), order.end(), [&](uint32_t a, uint32_t b) { return items[a].slab != items[b].slab ? items[a].slab < items[b].slab : items[a].tile < items[b].tile; });
      std::vector<WorkItem> sorted(items.size());
      std::vector<uint64_t> sorted_weight(items.size());
      uint64_t total_weight = 0;
      for (uint32_t k = 0; k < order.size(); ++k) { sorted[k] = items[order[k]]; sorted_weight[k] = item_weight[order[k]]; total_weight += sorted_weight[k]; }
      items.swap(sorted);
      item_weight.swap(sorted_weight);
      const uint64_t run_cap = runWeightCap(total_weight);
      std::vector<uint64_t> run_weight;
      for (uint32_t k = 0; k < items.size(); ++k) {
         const bool extend = !runs.empty() && items[runs.back().first_item].slab == items[k].slab && run_weight.back() + item_weight[k] <= run_cap && runs.back().n_items < RUN_MAX_ITEMS;
         if (extend) { runs.back().n_items++; run_weight.back() += item_weight[k]; }
         else { runs.push_back({k, 1, items[k].partition, 0}); run_weight.push_back(item_weight[k]); }
      }
      std::vector<uint32_t> run_order(runs.size());
      for (uint32_t k = 0; k < run_order.size(); ++k) run_order[k] = k;
      std::stable_sort(run_order.begin(), run_order.end(), [&](uint32_t a, uint32_t b) { return run_weight[a] > run_weight[b]; });
      std::vector<RunRec> sorted_runs(runs.size());
      for (uint32_t k = 0; k < run_order.size(); ++k) { sorted_runs[k] = runs[run_order[k]]; sorted_runs[k].weight_kb = uint32_t(run_weight[run_order[k]] >> 10); }
      runs.swap(sorted_runs);
   }
   while (postings.size() % 4) postings.push_back(filler);
   segs.back() = {uint32_t(postings.size()), uint32_t(large_entries.size())};
   for (int k = 0; k < 4; ++k) postings.push_back(filler);
   if (getenv("SB_TRACE")) {
      fprintf(stderr, "[sb] pack '%s': arrays by cardinality 1:%llu 2:%llu 3:%llu 4:%llu <=8:%llu <=16:%llu <=32:%llu <=64:%llu <=256:%llu <=4096:%llu runs:%llu bitsets:%llu postings:%zu large:%zu items:%zu\n", seq.name.c_str(),
              (unsigned long long)hist[0], (unsigned long long)hist[1], (unsigned long long)hist[2], (unsigned long long)hist[3], (unsigned long long)hist[4], (unsigned long long)hist[5], (unsigned long long)hist[6],
              (unsigned long long)hist[7], (unsigned long long)hist[8], (unsigned long long)hist[9], (unsigned long long)hist[10], (unsigned long long)hist[11], postings.size(), large_entries.size(), items.size());
   }

   // flags and missing-symbol intervals
   std::vector<int8_t> deleted(size_t(n_partitions) * L, -1), flipped(size_t(n_partitions) * L, -1);
   std::vector<uint32_t> row_off;
   std::vector<uint2> iv;
   std::vector<uint64_t> row_base(n_partitions, 0), iv_base(n_partitions, 0), inline_base(n_partitions, 0);
   std::vector<uint32_t> inline_words;      // 8 per row
   std::vector<uint32_t> any_words;         // 1 bit per row
   std::vector<uint64_t> any_base(n_partitions, 0);
   const bool narrow = L <= 65535;
   for (uint32_t p = 0; p < n_partitions; ++p) {
      row_base[p] = row_off.size();
      iv_base[p] = iv.size();
      inline_base[p] = inline_words.size() / 8;
      any_base[p] = any_words.size();
      if (seq.store_of_partition[p] < 0) { row_off.push_back(0); continue; }
      HostStore& st = *stores[seq.store_of_partition[p]];
      inline_words.resize(inline_words.size() + size_t(st.n_rows) * 8, 0u);
      any_words.resize(any_words.size() + (size_t(st.n_rows) + 31) / 32 + 1, 0u);
      for (uint32_t row = 0; row < st.n_rows; ++row) {
         uint32_t* rec = inline_words.data() + (inline_base[p] + row) * 8;
         const uint32_t b = st.miss_row_off[row], e = st.miss_row_off[row + 1];
         if (e > b) any_words[any_base[p] + row / 32] |= 1u << (row & 31);
         if (e - b > (narrow ? 8u : 4u)) { for (int k = 0; k < 8; ++k) rec[k] = 0xFFFFFFFFu; continue; }
         for (uint32_t j = b; j < e; ++j) {
            if (narrow) rec[j - b] = st.miss_iv[j].x | (st.miss_iv[j].y << 16);
            else { rec[2 * (j - b)] = st.miss_iv[j].x; rec[2 * (j - b) + 1] = st.miss_iv[j].y; }
         }
      }
      memcpy(deleted.data() + size_t(p) * L, st.deleted.data(), L);
      memcpy(flipped.data() + size_t(p) * L, st.flipped.data(), L);
      row_off.insert(row_off.end(), st.miss_row_off.begin(), st.miss_row_off.end());
      iv.insert(iv.end(), st.miss_iv.begin(), st.miss_iv.end());
      stats.missing_bytes += st.miss_row_off.size() * 4 + st.miss_iv.size() * 8;
      // the packed copy replaces the ingest buffers
      std::vector<TmpContainer>().swap(st.containers);
      std::vector<uint16_t>().swap(st.payload);
      std::vector<uint2>().swap(st.miss_iv);
      std::vector<uint32_t>().swap(st.miss_row_off);
   }
   seq.n_postings = postings.size();
   seq.n_large = large_entries.size();
   seq.n_items = uint32_t(items.size());
   seq.postings.upload(postings);
   seq.large_entries.upload(large_entries);
   seq.items.upload(items);
   seq.n_runs = uint32_t(runs.size());
   seq.runs.upload(runs);
   seq.h_part_rows.assign(n_partitions, 0);
   for (uint32_t p = 0; p < n_partitions; ++p) if (seq.store_of_partition[p] >= 0) seq.h_part_rows[p] = stores[seq.store_of_partition[p]]->n_rows;
   seq.part_rows.upload(seq.h_part_rows);
   seq.segs.upload(segs);
   seq.slabs.upload(slabs);
   seq.slab_begin.upload(seq.h_slab_begin);
   seq.large_pool.alloc(large_pool.size() / 2);
   if (!large_pool.empty()) SB_CUDA(cudaMemcpy(seq.large_pool.ptr, large_pool.data(), large_pool.size() * 8, cudaMemcpyHostToDevice));
   seq.deleted.upload(deleted);
   seq.flipped.upload(flipped);
   seq.miss_row_off.upload(row_off);
   seq.miss_iv.upload(iv);
   seq.miss_row_base.upload(row_base);
   seq.miss_iv_base.upload(iv_base);
   seq.miss_inline_base.upload(inline_base);
   seq.miss_inline.alloc(inline_words.size() / 4);
   if (!inline_words.empty()) SB_CUDA(cudaMemcpy(seq.miss_inline.ptr, inline_words.data(), inline_words.size() * 4, cudaMemcpyHostToDevice));
   seq.miss_any_base.upload(any_base);
   seq.miss_any.upload(any_words);
   stats.missing_bytes += seq.miss_inline.bytes() + seq.miss_any.bytes();
   std::vector<uint32_t>().swap(inline_words);
   {  // coarse index of the missing-symbol intervals: per (partition, tile) the rows that miss some position of the tile.
      // BitmapSelection(position) then searches the intervals of those rows only (a few percent of the partition).
      std::vector<uint64_t> tile_base(n_partitions, 0);
      std::vector<uint32_t> part_rows(n_partitions, 0);
      uint64_t words = 0;
      for (uint32_t p = 0; p < n_partitions; ++p) {
         tile_base[p] = words;
         if (seq.store_of_partition[p] < 0) continue;
         part_rows[p] = stores[seq.store_of_partition[p]]->n_rows;
         words += uint64_t(seq.n_tiles) * ((part_rows[p] + 31) / 32);
      }
      seq.miss_tile_base.upload(tile_base);
      seq.miss_tile_bits.alloc(size_t(std::max<uint64_t>(words, 1)));
      SB_CUDA(cudaMemset(seq.miss_tile_bits.ptr, 0, seq.miss_tile_bits.bytes()));
      for (uint32_t p = 0; p < n_partitions; ++p) {
         if (part_rows[p] == 0) continue;
         build_miss_tiles<<<(part_rows[p] + 255) / 256, 256>>>(seq.miss_row_off.ptr + row_base[p], seq.miss_iv.ptr + iv_base[p], part_rows[p], tp,
                                                             seq.miss_tile_bits.ptr + tile_base[p], (part_rows[p] + 31) / 32);
      }
      SB_CUDA(cudaDeviceSynchronize());
      stats.missing_bytes += seq.miss_tile_bits.bytes();
   }
   stats.directory_bytes += seq.large_entries.bytes() + seq.items.bytes() + seq.runs.bytes() + seq.segs.bytes() + seq.slabs.bytes();
   stats.payload_bytes += seq.postings.bytes() + seq.large_pool.bytes();
   stats.n_work_items += items.size();
   stats.n_chunks += large_entries.size();
   SeqDev& d = seq.dev;
   d.postings = seq.postings.ptr; d.large_entries = seq.large_entries.ptr; d.items = seq.items.ptr; d.segs = seq.segs.ptr; d.slabs = seq.slabs.ptr; d.slab_begin = seq.slab_begin.ptr;
   d.large_pool = seq.large_pool.ptr; d.deleted = seq.deleted.ptr; d.flipped = seq.flipped.ptr;
   d.miss_row_off = seq.miss_row_off.ptr; d.miss_iv = seq.miss_iv.ptr; d.miss_row_base = seq.miss_row_base.ptr; d.miss_iv_base = seq.miss_iv_base.ptr;
   d.miss_tile_bits = seq.miss_tile_bits.ptr; d.miss_tile_base = seq.miss_tile_base.ptr;
   d.miss_inline = seq.miss_inline.ptr; d.miss_inline_base = seq.miss_inline_base.ptr; d.miss_narrow = narrow ? 1u : 0u;
   d.miss_any = seq.miss_any.ptr; d.miss_any_base = seq.miss_any_base.ptr;
   d.part_rows = seq.part_rows.ptr; d.runs = seq.runs.ptr; d.n_runs = seq.n_runs;
   d.length = L; d.sigma = sigma; d.tile_positions = tp; d.n_tiles = seq.n_tiles; d.n_partitions = n_partitions;
   d.n_items = seq.n_items; d.n_slabs = n_slabs;
}

}  // namespace sb
