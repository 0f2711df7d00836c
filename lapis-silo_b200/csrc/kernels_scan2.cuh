// k1_pipe: the container scan with ONE prefetch pipeline for every kind of container (sm_100a).
// k1_scan (kernels_action.cuh) prefetches posting chunks only: a warp that takes a bitset half issues its eight loads and
// waits for them, and the directory entry is one more dependent load in front. Here every unit of work is a 2 KB chunk --
// 128 x 16 bytes of postings, a QUARTER of a bitset container, or an array piece (at most 1,024 values = 2 KB) -- loaded by
// the same routine into the same registers, DEPTH chunks ahead of the one being counted; the directory entries of the
// item's large containers are staged in shared memory by warp 0 while the previous item streams, so that no load of the
// pipeline depends on another load. Everything else (runs of items per slab, the tile table with the flush that zeroes it,
// two barriers per item, flipped / deleted-symbol arithmetic) is k1_scan's.
#pragma once
#include "kernels_action.cuh"

namespace sb {

constexpr uint32_t K1_ENTRY_CACHE = 128;      // directory entries of an item kept in shared memory (more: read from global memory)
constexpr uint32_t K1_CHUNK4 = 128;           // 128-bit words per chunk: 2 KB, four per lane

__host__ __device__ constexpr size_t k1PipeSharedBytes(uint32_t tile_slots) {
   return K1_FILTER_BYTES + size_t(k1TableWords(tile_slots)) * 4 + 16 + 2 * sizeof(K1Next) + 2 * K1_ENTRY_CACHE * sizeof(LargeEntry) + 16;
}

template <int MIN_CTAS, int DEPTH>
__global__ void __launch_bounds__(K1_THREADS, MIN_CTAS)
k1_pipe(const SeqDev s, const ActionPtrs a, uint32_t* __restrict__ counts, int full_mode) {
   extern __shared__ __align__(128) uint8_t k1_smem[];
   uint32_t* filter_s = reinterpret_cast<uint32_t*>(k1_smem);
   uint32_t* tile_counts = reinterpret_cast<uint32_t*>(k1_smem + K1_FILTER_BYTES);
   const uint32_t T = s.tile_positions, tile_slots = T * s.sigma;
   uint64_t* bar = reinterpret_cast<uint64_t*>(tile_counts + k1TableWords(tile_slots));
   uint32_t* unit_counters = reinterpret_cast<uint32_t*>(bar + 1);      // [2]: item i hands out its units through counter i & 1
   K1Next* item_slots = reinterpret_cast<K1Next*>(bar + 2);             // [2]
   LargeEntry* entry_slots = reinterpret_cast<LargeEntry*>(item_slots + 2);   // [2][K1_ENTRY_CACHE]
   constexpr uint32_t N_WARPS = K1_THREADS / 32;
   const uint32_t lane = laneId(), warp = threadIdx.x >> 5;

   // warp 0: the item after the current one (thread 0 walks run -> item, see k1_scan) and its directory entries
   uint32_t run_next = 0, run_end = 0;
   const uint32_t* run_filter = nullptr;
   auto fetchNext = [&](uint32_t slot) {
      if (lane == 0) {
         K1Next n{};
         if (run_next == run_end) {
            const uint32_t k = atomicAdd(&a.head->next_run, 1u);
            if (k < a.head->n_selected) {
               const RunRec run = s.runs[a.selected[k]];
               run_next = run.first_item; run_end = run.first_item + run.n_items;
               run_filter = a.pq[run.partition].filter32;
            }
         }
         if (run_next < run_end) { n.item = s.items[run_next++]; n.filter32 = run_filter; n.valid = 1; }
         item_slots[slot] = n;
      }
      __syncwarp();
      const K1Next& n = item_slots[slot];
      if (!n.valid) return;
      const uint32_t n_entries = min(n.item.large_end - n.item.large_begin, K1_ENTRY_CACHE);
      const uint4* src = reinterpret_cast<const uint4*>(s.large_entries + n.item.large_begin);
      uint4* dst = reinterpret_cast<uint4*>(entry_slots + slot * K1_ENTRY_CACHE);
      for (uint32_t e = lane; e < n_entries; e += 32) dst[e] = __ldg(src + e);
   };
   if (threadIdx.x == 0) {
      mbarInit(bar, 1);
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
      unit_counters[0] = N_WARPS;
   }
   if (warp == 0) fetchNext(0);
   for (uint32_t i = threadIdx.x; i < tile_slots + 32; i += K1_THREADS) tile_counts[i] = 0;
   const uint32_t dummy = tile_slots << POSTING_ROW_BITS;          // a posting of the slot that is never flushed
   const uint32_t counts_s = smemAddr(tile_counts);
   uint32_t parity = 0, turn = 0;
   uint32_t cur_slab = 0xFFFFFFFFu;
   __syncthreads();
   for (;; turn ^= 1u) {
      const K1Next& cur = item_slots[turn];
      if (!cur.valid) break;
      const WorkItem& item = cur.item;
      const LargeEntry* entries = entry_slots + turn * K1_ENTRY_CACHE;
      const bool new_slab = item.slab != cur_slab;
      cur_slab = item.slab;
      uint32_t* next_unit = unit_counters + turn;
      __syncthreads();                                    // the previous flush (and its zeroing) is done; the other slots are free
      if (threadIdx.x == 0) {
         if (new_slab) {
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            mbarExpectTx(bar, K1_FILTER_BYTES);
            bulkLoad(filter_s, cur.filter32 + size_t(item.key0) * 2048u, K1_FILTER_BYTES, bar);
         }
         unit_counters[turn ^ 1u] = N_WARPS;
      }
      constexpr uint32_t FLUSH_MAX = MAX_TILE_POSITIONS / K1_THREADS;
      int flush_del[FLUSH_MAX], flush_flip[FLUSH_MAX];
#pragma unroll
      for (uint32_t f = 0; f < FLUSH_MAX; ++f) {
         const uint32_t pl = threadIdx.x + f * K1_THREADS, pos = item.tile * T + pl;
         flush_del[f] = flush_flip[f] = -1;
         if (pl < T && pos < s.length) {
            flush_del[f] = s.deleted[size_t(item.partition) * s.length + pos];
            flush_flip[f] = s.flipped[size_t(item.partition) * s.length + pos];
         }
      }
      // units: four per large container (quarters of a bitset; an array piece or a run container is its quarter 0, the
      // other three are empty), then the posting chunks
      const uint4* post4 = reinterpret_cast<const uint4*>(s.postings + item.post_begin);
      const uint32_t n4 = (item.post_end - item.post_begin + 3u) >> 2;
      const uint32_t n_large_units = 4u * (item.large_end - item.large_begin);
      const uint32_t n_units = n_large_units + (n4 + K1_CHUNK4 - 1) / K1_CHUNK4;
      auto entryOf = [&](uint32_t unit) -> LargeEntry {
         const uint32_t e = unit >> 2;
         return e < K1_ENTRY_CACHE ? entries[e] : s.large_entries[item.large_begin + e];
      };
      auto loadChunk = [&](uint32_t unit, uint4 (&dst)[4]) {
#pragma unroll
         for (int k = 0; k < 4; ++k) dst[k] = make_uint4(dummy, dummy, dummy, dummy);
         if (unit >= n_units) return;
         if (unit >= n_large_units) {
            const uint32_t base = (unit - n_large_units) * K1_CHUNK4 + lane;
#pragma unroll
            for (int k = 0; k < 4; ++k) if (base + k * 32 < n4) dst[k] = loadStream(post4 + base + k * 32);
            return;
         }
         const LargeEntry w = entryOf(unit);
         const uint32_t quarter = unit & 3u, kind = w.kind();
         const uint4* src = s.large_pool + w.offset16;
         if (kind == KIND_BITSET) {
#pragma unroll
            for (int k = 0; k < 4; ++k) dst[k] = loadStream(src + quarter * K1_CHUNK4 + k * 32 + lane);
         } else if (kind == KIND_ARRAY && quarter == 0) {
            const uint32_t n_vec = (w.size + 7) >> 3;      // at most 128: array pieces hold at most 1,024 values
#pragma unroll
            for (int k = 0; k < 4; ++k) if (k * 32 + lane < n_vec) dst[k] = loadStream(src + k * 32 + lane);
         }
      };
      auto grab = [&]() {
         uint32_t next = 0;
         if (lane == 0) next = atomicAdd(next_unit, 1u);
         return __shfl_sync(0xFFFFFFFFu, next, 0);
      };
      // counts one chunk; `data` was loaded by loadChunk(unit, ...)
      auto countChunk = [&](uint32_t unit, const uint4 (&data)[4]) {
         if (unit >= n_large_units) {      // postings: the filter words first (independent shared-memory loads), then predicated increments
#pragma unroll
            for (int h = 0; h < 2; ++h) {
               const uint32_t p[8] = {data[2 * h].x, data[2 * h].y, data[2 * h].z, data[2 * h].w, data[2 * h + 1].x, data[2 * h + 1].y, data[2 * h + 1].z, data[2 * h + 1].w};
               uint32_t w[8];
#pragma unroll
               for (int t = 0; t < 8; ++t) w[t] = filter_s[(p[t] >> 5) & (SLAB_WORDS32 - 1)];
#pragma unroll
               for (int t = 0; t < 8; ++t) {
                  const uint32_t hit = (w[t] >> (p[t] & 31)) & 1u;      // a posting of the dummy slot carries row 0 of the slab; its slot lies behind the table's last counter
                  asm volatile("{ .reg .pred q; setp.ne.u32 q, %0, 0; @q red.shared.add.u32 [%1], 1; }" ::"r"(hit), "r"(counts_s + ((p[t] >> POSTING_ROW_BITS) << 2)) : "memory");
               }
            }
            return;
         }
         const LargeEntry w = entryOf(unit);
         const uint32_t quarter = unit & 3u, kind = w.kind(), k_local = w.keyLocal();
         int c2 = 0;
         if (kind == KIND_BITSET) {
            const uint4* fk = reinterpret_cast<const uint4*>(filter_s) + k_local * 512u + quarter * K1_CHUNK4;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
               const uint4 y = fk[k * 32 + lane];
               c2 += __popc(data[k].x & y.x) + __popc(data[k].y & y.y) + __popc(data[k].z & y.z) + __popc(data[k].w & y.w);
            }
         } else if (quarter != 0) {
            return;      // uniform over the warp
         } else if (kind == KIND_ARRAY) {
            const uint32_t* fk = filter_s + k_local * 2048u;
#pragma unroll
            for (uint32_t k = 0; k < 4; ++k) {
               const uint32_t wv[4] = {data[k].x, data[k].y, data[k].z, data[k].w};
#pragma unroll
               for (uint32_t t = 0; t < 8; ++t) {
                  const uint32_t val = (wv[t >> 1] >> ((t & 1) * 16)) & 0xFFFFu;
                  c2 += ((fk[val >> 5] >> (val & 31)) & 1u) & uint32_t((k * 32 + lane) * 8 + t < w.size);
               }
            }
            const uint4* src = s.large_pool + w.offset16;      // a piece of more than 1,024 values (SB_ARRAY_SPLIT tuning): the rest without prefetch
            for (uint32_t i0 = K1_CHUNK4 + lane; i0 < ((w.size + 7) >> 3); i0 += 32) {
               const uint4 x = __ldg(src + i0);
               const uint32_t wv[4] = {x.x, x.y, x.z, x.w};
#pragma unroll
               for (uint32_t t = 0; t < 8; ++t) {
                  const uint32_t val = (wv[t >> 1] >> ((t & 1) * 16)) & 0xFFFFu;
                  c2 += ((fk[val >> 5] >> (val & 31)) & 1u) & uint32_t(i0 * 8 + t < w.size);
               }
            }
         } else {      // run container: two rank lookups per run
            const uint16_t* runs = reinterpret_cast<const uint16_t*>(s.large_pool + w.offset16);
            const PartQuery q = a.pq[item.partition];
            for (uint32_t i = lane; i < w.size; i += 32) { const uint32_t start = __ldg(runs + 2 * i), len1 = __ldg(runs + 2 * i + 1); c2 += int(filterRangeCount(q, item.key0 + k_local, start, start + len1)); }
         }
         c2 = warpSum(c2);
         if (lane == 0 && c2 != 0) atomicAdd(&tile_counts[w.slot()], uint32_t(c2));
      };

      // the pipeline: the chunk being counted + DEPTH chunks in flight; a warp's first unit is its warp id
      uint32_t units[DEPTH];
      uint4 ring[DEPTH][4];
      units[0] = warp;
      loadChunk(units[0], ring[0]);
#pragma unroll
      for (int d = 1; d < DEPTH; ++d) { units[d] = grab(); loadChunk(units[d], ring[d]); }
      if (new_slab) { mbarWait(bar, parity); }
      if (warp == 0) fetchNext(turn ^ 1u);
      while (units[0] < n_units) {
         const uint32_t ahead = grab();
         uint4 incoming[4];
         loadChunk(ahead, incoming);
         countChunk(units[0], ring[0]);
#pragma unroll
         for (int d = 0; d + 1 < DEPTH; ++d) {
            units[d] = units[d + 1];
#pragma unroll
            for (int k = 0; k < 4; ++k) ring[d][k] = ring[d + 1][k];
         }
         units[DEPTH - 1] = ahead;
#pragma unroll
         for (int k = 0; k < 4; ++k) ring[DEPTH - 1][k] = incoming[k];
      }
      if (new_slab) parity ^= 1u;
      __syncthreads();
      // flush (and zero) the tile table: see k1_scan
#pragma unroll
      for (uint32_t f = 0; f < FLUSH_MAX; ++f) {
         const uint32_t pl = threadIdx.x + f * K1_THREADS, flush_pos = item.tile * T + pl;
         if (pl >= T || flush_pos >= s.length) continue;
         int acc_deleted = 0;
         for (uint32_t sym0 = 0; sym0 < s.sigma; sym0 += 8) {
            int raw[8];
#pragma unroll
            for (uint32_t q = 0; q < 8; ++q) raw[q] = sym0 + q < s.sigma ? int(tile_counts[(sym0 + q) * T + pl]) : 0;
#pragma unroll
            for (uint32_t q = 0; q < 8; ++q) if (sym0 + q < s.sigma && raw[q] != 0) tile_counts[(sym0 + q) * T + pl] = 0;
#pragma unroll
            for (uint32_t q = 0; q < 8; ++q) {
               const uint32_t sym = sym0 + q;
               if (raw[q] == 0 || int(sym) == flush_del[f]) continue;
               const int add_slot = int(sym) == flush_flip[f] ? -raw[q] : raw[q];
               atomicAdd(counts + size_t(sym) * s.length + flush_pos, uint32_t(add_slot));
               acc_deleted += full_mode ? -raw[q] : -add_slot;
            }
         }
         if (flush_del[f] >= 0 && acc_deleted != 0) atomicAdd(counts + size_t(flush_del[f]) * s.length + flush_pos, uint32_t(acc_deleted));
      }
   }
}

}  // namespace sb
