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
   uint32_t length = 0, sigma = 0, tile_positions = 32, n_tiles = 0;
   std::vector<int> store_of_partition;   // index into Db::stores or -1
   // device
   DeviceBuffer<uint64_t> entries;
   DeviceBuffer<ChunkRec> chunks;
   DeviceBuffer<WorkItem> items;
   DeviceBuffer<SegRec> segs;
   DeviceBuffer<uint16_t> arrays;
   DeviceBuffer<uint4> bitsets;
   DeviceBuffer<int8_t> deleted, flipped;
   DeviceBuffer<uint32_t> miss_row_off;
   DeviceBuffer<uint2> miss_iv;
   DeviceBuffer<uint64_t> miss_row_base, miss_iv_base;
   DeviceBuffer<uint32_t> full_tables;   // [P][sigma * length]
   SeqDev dev{};
   uint32_t n_items = 0, n_chunks = 0;
   uint64_t n_entries = 0;
};

struct RegisteredBitmap { uint32_t partition; uint32_t first_container, n_containers; uint32_t max_value; };

struct Column { uint32_t partition; int dtype; uint32_t n; void* device = nullptr; };

constexpr uint64_t ITEM_WEIGHT_CAP = 256 * 1024;   // bytes of directory + payload per Mutations work item

// Packs the ingested stores of one sequence into the tiled device layout.
inline void packSequence(Sequence& seq, std::vector<std::unique_ptr<HostStore>>& stores, uint32_t n_partitions, sb_pool_stats& stats) {
   const uint32_t sigma = seq.sigma, tp = seq.tile_positions, L = seq.length;
   seq.n_tiles = (L + tp - 1) / tp;
   const uint32_t slots_per_tile = tp * sigma;
   std::vector<uint64_t> entries;
   std::vector<ChunkRec> chunks;
   std::vector<WorkItem> items;
   std::vector<SegRec> segs(size_t(seq.n_tiles) * n_partitions + 1);
   std::vector<uint16_t> arrays;
   std::vector<uint64_t> bitsets;   // 1024 words per container
   std::vector<size_t> cursor(n_partitions, 0);
   uint64_t total_containers = 0;
   for (uint32_t p = 0; p < n_partitions; ++p) if (seq.store_of_partition[p] >= 0) total_containers += stores[seq.store_of_partition[p]]->containers.size();
   entries.reserve(total_containers);
   for (uint32_t tile = 0; tile < seq.n_tiles; ++tile) {
      const uint32_t slot_end = (tile + 1) * slots_per_tile;
      const size_t tile_chunk_begin = chunks.size();
      std::vector<uint64_t> chunk_weight;
      for (uint32_t p = 0; p < n_partitions; ++p) {
         segs[size_t(tile) * n_partitions + p] = {uint32_t(entries.size()), uint32_t(chunks.size())};
         if (seq.store_of_partition[p] < 0) continue;
         HostStore& st = *stores[seq.store_of_partition[p]];
         size_t& cur = cursor[p];
         while (cur < st.containers.size() && st.containers[cur].slot < slot_end) {
            // open a chunk of up to 32 entries; its payload starts 16-byte aligned
            while (arrays.size() % 8) arrays.push_back(0);
            ChunkRec rec{};
            rec.entry_start = uint32_t(entries.size());
            SB_REQUIRE(arrays.size() / 8 <= 0xFFFFFFFFull, "array pool exceeds 64 GB");
            rec.pay_off16 = uint32_t(arrays.size() / 8);
            rec.bitset_idx = uint32_t(bitsets.size() / 1024);
            rec.partition = uint16_t(p);
            uint64_t weight = 16;
            uint32_t n = 0;
            while (n < 32 && cur < st.containers.size() && st.containers[cur].slot < slot_end) {
               const TmpContainer& c = st.containers[cur++];
               const uint32_t pos = c.slot / sigma, sym = c.slot - pos * sigma;
               const int8_t del = st.deleted[pos], flip = st.flipped[pos];
               entries.push_back(EntryBits::pack(c.key, c.slot - tile * slots_per_tile, c.kind, c.kind == KIND_BITSET ? 1 : c.size,
                                                 del < 0 || uint32_t(del) == sym ? 31u : uint32_t(del), flip >= 0 && uint32_t(flip) == sym));
               const uint16_t* src = st.payload.data() + c.pay;
               if (c.kind == KIND_BITSET) {
                  const uint64_t* w = reinterpret_cast<const uint64_t*>(src);
                  bitsets.insert(bitsets.end(), w, w + 1024);
                  weight += 8192 + 8;
                  stats.bitset_containers++; stats.bitset_values += c.size;
               } else if (c.kind == KIND_ARRAY) {
                  arrays.insert(arrays.end(), src, src + c.size);
                  weight += 2 * c.size + 8;
                  stats.array_containers++; stats.array_values += c.size;
               } else {
                  arrays.insert(arrays.end(), src, src + 2 * c.size);
                  weight += 4 * c.size + 8;
                  stats.run_containers++; stats.run_count += c.size;
                  for (uint32_t r = 0; r < c.size; ++r) stats.run_values += uint64_t(src[2 * r + 1]) + 1;
               }
               ++n;
            }
            rec.n_entries = uint8_t(n);
            chunks.push_back(rec);
            chunk_weight.push_back(weight);
         }
      }
      // cut the tile's chunks into weight-bounded work items
      size_t begin = tile_chunk_begin;
      uint64_t acc = 0;
      for (size_t c = tile_chunk_begin; c < chunks.size(); ++c) {
         acc += chunk_weight[c - tile_chunk_begin];
         if (acc >= ITEM_WEIGHT_CAP || c + 1 == chunks.size()) {
            items.push_back({tile, uint32_t(begin), uint32_t(c + 1), 0});
            begin = c + 1;
            acc = 0;
         }
      }
   }
   segs.back() = {uint32_t(entries.size()), uint32_t(chunks.size())};
   SB_REQUIRE(entries.size() < 0xFFFFFFFFull, "more than 2^32 containers in one sequence");
   // the deleted symbol of an entry must not be the entry's own symbol: handled above (31 = none)

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
   while (arrays.size() % 8) arrays.push_back(0);
   seq.n_entries = entries.size();
   seq.n_items = uint32_t(items.size());
   seq.n_chunks = uint32_t(chunks.size());
   seq.entries.upload(entries);
   seq.chunks.upload(chunks);
   seq.items.upload(items);
   seq.segs.upload(segs);
   seq.arrays.upload(arrays);
   seq.bitsets.alloc(bitsets.size() / 2);
   if (!bitsets.empty()) SB_CUDA(cudaMemcpy(seq.bitsets.ptr, bitsets.data(), bitsets.size() * 8, cudaMemcpyHostToDevice));
   seq.deleted.upload(deleted);
   seq.flipped.upload(flipped);
   seq.miss_row_off.upload(row_off);
   seq.miss_iv.upload(iv);
   seq.miss_row_base.upload(row_base);
   seq.miss_iv_base.upload(iv_base);
   stats.directory_bytes += seq.entries.bytes() + seq.chunks.bytes() + seq.items.bytes() + seq.segs.bytes();
   stats.payload_bytes += seq.arrays.bytes() + seq.bitsets.bytes();
   stats.n_work_items += items.size();
   stats.n_chunks += chunks.size();
   SeqDev& d = seq.dev;
   d.entries = seq.entries.ptr; d.chunks = seq.chunks.ptr; d.items = seq.items.ptr; d.segs = seq.segs.ptr;
   d.arrays = seq.arrays.ptr; d.bitsets = seq.bitsets.ptr; d.deleted = seq.deleted.ptr; d.flipped = seq.flipped.ptr;
   d.miss_row_off = seq.miss_row_off.ptr; d.miss_iv = seq.miss_iv.ptr; d.miss_row_base = seq.miss_row_base.ptr; d.miss_iv_base = seq.miss_iv_base.ptr;
   d.length = L; d.sigma = sigma; d.tile_positions = tp; d.n_tiles = seq.n_tiles; d.n_partitions = n_partitions;
   d.n_items = seq.n_items; d.n_chunks = seq.n_chunks;
}

}  // namespace sb
