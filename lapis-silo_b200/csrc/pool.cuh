// Host side of the HBM container pool: ingest of CRoaring portable bytes, packing into the tiled
// directory + payload pools, upload. Format parsed here: SURVEY Appendix A (the reference writes it
// through include/silo/roaring/roaring_serialize.h:18-45).
#pragma once
#include <algorithm>

#include "common.cuh"
#include "kernels_filter.cuh"

namespace sb {

struct ParsedContainer { uint16_t key; uint8_t kind; uint32_t size; uint32_t card; const uint16_t* data; };  // size: array card / #runs / bitset card

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
         c.kind = KIND_RUN; c.size = r; c.data = reinterpret_cast<const uint16_t*>(p + pos);
         pos += 4 * size_t(r);
      } else if (c.card <= 4096) {
         need(2 * size_t(c.card));
         c.kind = KIND_ARRAY; c.size = c.card; c.data = reinterpret_cast<const uint16_t*>(p + pos);
         pos += 2 * size_t(c.card);
      } else {
         need(8192);
         c.kind = KIND_BITSET; c.size = c.card; c.data = reinterpret_cast<const uint16_t*>(p + pos);
         pos += 8192;
      }
      f(c);
   }
}

inline uint32_t payloadU16(const ParsedContainer& c) { return c.kind == KIND_ARRAY ? c.size : c.kind == KIND_RUN ? 2 * c.size : 4096; }

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
   DeviceBuffer<uint64_t> entries;
   DeviceBuffer<StageRec> stages;
   DeviceBuffer<WorkItem> items;
   DeviceBuffer<SegRec> segs;
   DeviceBuffer<SlabRec> slabs;
   DeviceBuffer<uint32_t> slab_begin;
   DeviceBuffer<uint16_t> small_pool;
   DeviceBuffer<uint4> large_pool;
   std::vector<uint32_t> h_slab_begin;    // [P + 1]
   DeviceBuffer<int8_t> deleted, flipped;
   DeviceBuffer<uint32_t> miss_row_off;
   DeviceBuffer<uint2> miss_iv;
   DeviceBuffer<uint64_t> miss_row_base, miss_iv_base;
   DeviceBuffer<uint32_t> full_tables;   // [P][sigma * length]
   DeviceBuffer<uint8_t> reference;      // [length] symbols of the reference sequence (row selection)
   DeviceBuffer<uint32_t> row_area;      // [4] counter + selected rows
   SeqDev dev{};
   uint32_t n_items = 0, n_stages = 0, n_slabs = 0;
   uint64_t n_entries = 0;
};

struct RegisteredBitmap { uint32_t partition; uint32_t first_container, n_containers; uint32_t max_value; };

struct Column { uint32_t partition; int dtype; uint32_t n; void* device = nullptr; };

constexpr uint64_t ITEM_WEIGHT_CAP = 512 * 1024;   // bytes of directory + payload per Mutations work item

// Packs the ingested stores of one sequence into the tiled device layout.
inline void packSequence(Sequence& seq, std::vector<std::unique_ptr<HostStore>>& stores, uint32_t n_partitions, sb_pool_stats& stats) {
   const uint32_t sigma = seq.sigma, tp = seq.tile_positions, L = seq.length;
   seq.n_tiles = (L + tp - 1) / tp;
   const uint32_t slots_per_tile = tp * sigma;
   SB_REQUIRE(slots_per_tile <= 0x4000, "tile does not fit the 14-bit slot field");
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

   std::vector<uint64_t> entries;
   std::vector<StageRec> stage_recs;
   std::vector<WorkItem> items;
   std::vector<uint64_t> item_weight;
   std::vector<SegRec> segs(size_t(seq.n_tiles) * std::max<uint32_t>(n_slabs, 1) + 1);
   std::vector<uint16_t> small_pool;
   std::vector<uint64_t> large_pool;   // 16-byte units = 2 words each
   std::vector<size_t> cursor(n_partitions, 0);
   uint64_t total_containers = 0;
   for (uint32_t p = 0; p < n_partitions; ++p) if (seq.store_of_partition[p] >= 0) total_containers += stores[seq.store_of_partition[p]]->containers.size();
   entries.reserve(total_containers + total_containers / 256 + 16);
   std::vector<const TmpContainer*> picked;
   for (uint32_t tile = 0; tile < seq.n_tiles; ++tile) {
      const uint32_t slot_end = (tile + 1) * slots_per_tile;
      for (uint32_t slab_id = 0; slab_id < n_slabs; ++slab_id) {
         segs[size_t(tile) * n_slabs + slab_id] = {uint32_t(entries.size()), uint32_t(stage_recs.size())};
         const uint32_t p = slabs[slab_id].partition, key0 = slabs[slab_id].key0;
         HostStore& st = *stores[seq.store_of_partition[p]];
         // the (tile, partition) range of the store's containers; pick this slab's keys, keeping (slot, key) order
         const size_t begin = cursor[p];
         size_t end = begin;
         while (end < st.containers.size() && st.containers[end].slot < slot_end) ++end;
         if (slab_id + 1 == seq.h_slab_begin[p + 1]) cursor[p] = end;   // last slab of the partition consumes the range
         picked.clear();
         for (size_t k = begin; k < end; ++k) if (st.containers[k].key >= key0 && st.containers[k].key < key0 + SLAB_KEYS) picked.push_back(&st.containers[k]);
         if (picked.empty()) continue;
         const size_t seg_stage_begin = stage_recs.size();
         std::vector<uint64_t> stage_weight;
         size_t k = 0;
         while (k < picked.size()) {
            // open a stage: entries start even (the previous stage was padded), small payload starts 16-byte aligned
            while (small_pool.size() % 8) small_pool.push_back(0);
            StageRec rec{};
            rec.entry_start = uint32_t(entries.size());
            rec.small_off16 = small_pool.size() / 8;
            rec.large_off16 = large_pool.size() / 2;
            const size_t small_begin = small_pool.size(), large_begin = large_pool.size() / 2;
            uint32_t n_chunks = 0, n_entries = 0;
            while (k < picked.size() && n_chunks < STAGE_CHUNKS && (small_pool.size() - small_begin) + 32 * SMALL_ARRAY_MAX <= STAGE_SMALL_U16) {
               rec.chunk_small[n_chunks] = uint16_t(small_pool.size() - small_begin);
               rec.chunk_large[n_chunks] = uint32_t(large_pool.size() / 2 - large_begin);
               for (uint32_t n = 0; n < 32 && k < picked.size(); ++n, ++k) {
                  const TmpContainer& c = *picked[k];
                  const uint32_t pos = c.slot / sigma, sym = c.slot - pos * sigma;
                  const int8_t del = st.deleted[pos], flip = st.flipped[pos];
                  entries.push_back(EntryBits::pack(c.key, c.slot - tile * slots_per_tile, c.kind, c.kind == KIND_BITSET ? 1 : c.size,
                                                    del < 0 || uint32_t(del) == sym ? 31u : uint32_t(del), flip >= 0 && uint32_t(flip) == sym));
                  ++n_entries;
                  const uint16_t* src = st.payload.data() + c.pay;
                  const uint32_t size_field = c.kind == KIND_BITSET ? 1 : c.size;
                  const uint32_t small_units = EntryBits::smallUnits(c.kind, size_field), large_units = EntryBits::largeUnits(c.kind, size_field);
                  if (small_units) small_pool.insert(small_pool.end(), src, src + small_units);
                  else {
                     const uint32_t n16 = c.kind == KIND_BITSET ? 4096 : c.kind == KIND_ARRAY ? c.size : 2 * c.size;
                     const size_t at = large_pool.size();
                     large_pool.resize(at + size_t(large_units) * 2, 0);
                     memcpy(large_pool.data() + at, src, size_t(n16) * 2);
                  }
                  if (c.kind == KIND_BITSET) { stats.bitset_containers++; stats.bitset_values += c.size; }
                  else if (c.kind == KIND_ARRAY) { stats.array_containers++; stats.array_values += c.size; }
                  else { stats.run_containers++; stats.run_count += c.size; for (uint32_t r = 0; r < c.size; ++r) stats.run_values += uint64_t(src[2 * r + 1]) + 1; }
               }
               ++n_chunks;
            }
            for (uint32_t c = n_chunks; c < STAGE_CHUNKS; ++c) { rec.chunk_small[c] = uint16_t(small_pool.size() - small_begin); rec.chunk_large[c] = uint32_t(large_pool.size() / 2 - large_begin); }
            while (small_pool.size() % 8) small_pool.push_back(0);
            rec.n_entries = n_entries;
            rec.small_bytes = uint32_t((small_pool.size() - small_begin) * 2);
            stage_recs.push_back(rec);
            // keep the next stage's entries 16-byte aligned: a padding entry with the slot of its predecessor
            if (entries.size() % 2) entries.push_back(EntryBits::pack(EntryBits::key(entries.back()), EntryBits::slot(entries.back()), KIND_NULL, 1, 31, 0));
            stage_weight.push_back(128 + 8ull * n_entries + rec.small_bytes + 16ull * (large_pool.size() / 2 - large_begin));
         }
         // cut the segment's stages into weight-bounded work items
         size_t item_begin = seg_stage_begin;
         uint64_t acc = 0;
         for (size_t sidx = seg_stage_begin; sidx < stage_recs.size(); ++sidx) {
            acc += stage_weight[sidx - seg_stage_begin];
            if (acc >= ITEM_WEIGHT_CAP || sidx + 1 == stage_recs.size()) {
               items.push_back({tile, slab_id, uint32_t(item_begin), uint32_t(sidx + 1)});
               item_weight.push_back(acc);
               item_begin = sidx + 1;
               acc = 0;
            }
         }
      }
   }
   {  // heaviest items first: the block scheduler hands out CTAs in index order, so the tail stays short
      std::vector<uint32_t> order(items.size());
      for (uint32_t k = 0; k < order.size(); ++k) order[k] = k;
      std::stable_sort(order.begin(), order.end(), [&](uint32_t a, uint32_t b) { return item_weight[a] > item_weight[b]; });
      std::vector<WorkItem> sorted(items.size());
      for (uint32_t k = 0; k < order.size(); ++k) sorted[k] = items[order[k]];
      items.swap(sorted);
   }
   if (entries.size() % 2) entries.push_back(EntryBits::pack(0, 0, KIND_NULL, 1, 31, 0));
   entries.push_back(EntryBits::pack(0, 0, KIND_NULL, 1, 31, 0));
   entries.push_back(EntryBits::pack(0, 0, KIND_NULL, 1, 31, 0));
   segs.back() = {uint32_t(entries.size() - 2), uint32_t(stage_recs.size())};
   SB_REQUIRE(entries.size() < 0xFFFFFFFFull, "more than 2^32 containers in one sequence");

   // flags and missing-symbol intervals
   std::vector<int8_t> deleted(size_t(n_partitions) * L, -1), flipped(size_t(n_partitions) * L, -1);
   std::vector<uint32_t> row_off;
   std::vector<uint2> iv;
   std::vector<uint64_t> row_base(n_partitions, 0), iv_base(n_partitions, 0);
   for (uint32_t p = 0; p < n_partitions; ++p) {
      row_base[p] = row_off.size();
      iv_base[p] = iv.size();
      if (seq.store_of_partition[p] < 0) { row_off.push_back(0); continue; }
      HostStore& st = *stores[seq.store_of_partition[p]];
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
   while (small_pool.size() % 8) small_pool.push_back(0);
   seq.n_entries = entries.size();
   seq.n_items = uint32_t(items.size());
   seq.n_stages = uint32_t(stage_recs.size());
   seq.entries.upload(entries);
   seq.stages.upload(stage_recs);
   seq.items.upload(items);
   seq.segs.upload(segs);
   seq.slabs.upload(slabs);
   seq.slab_begin.upload(seq.h_slab_begin);
   seq.small_pool.upload(small_pool);
   seq.large_pool.alloc(large_pool.size() / 2);
   if (!large_pool.empty()) SB_CUDA(cudaMemcpy(seq.large_pool.ptr, large_pool.data(), large_pool.size() * 8, cudaMemcpyHostToDevice));
   seq.deleted.upload(deleted);
   seq.flipped.upload(flipped);
   seq.miss_row_off.upload(row_off);
   seq.miss_iv.upload(iv);
   seq.miss_row_base.upload(row_base);
   seq.miss_iv_base.upload(iv_base);
   stats.directory_bytes += seq.entries.bytes() + seq.stages.bytes() + seq.items.bytes() + seq.segs.bytes() + seq.slabs.bytes();
   stats.payload_bytes += seq.small_pool.bytes() + seq.large_pool.bytes();
   stats.n_work_items += items.size();
   stats.n_chunks += stage_recs.size();
   SeqDev& d = seq.dev;
   d.entries = seq.entries.ptr; d.stages = seq.stages.ptr; d.items = seq.items.ptr; d.segs = seq.segs.ptr; d.slabs = seq.slabs.ptr; d.slab_begin = seq.slab_begin.ptr;
   d.small_pool = seq.small_pool.ptr; d.large_pool = seq.large_pool.ptr; d.deleted = seq.deleted.ptr; d.flipped = seq.flipped.ptr;
   d.miss_row_off = seq.miss_row_off.ptr; d.miss_iv = seq.miss_iv.ptr; d.miss_row_base = seq.miss_row_base.ptr; d.miss_iv_base = seq.miss_iv_base.ptr;
   d.length = L; d.sigma = sigma; d.tile_positions = tp; d.n_tiles = seq.n_tiles; d.n_partitions = n_partitions;
   d.n_items = seq.n_items; d.n_stages = seq.n_stages; d.n_slabs = n_slabs;
}

}  // namespace sb
