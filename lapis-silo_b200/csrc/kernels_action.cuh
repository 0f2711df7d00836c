// Action kernels (sm_100a): the Mutations count tables.
//   k1_and_popcount   fused AND-popcount over the container pool; never materialises filter AND bitmap.
//                     Replaces the and_cardinality / andnot_cardinality calls of
//                     src/silo/query_engine/actions/mutations.cpp:85-95 (M2) for all positions at once.
//   k2_missing        per-position count of filtered rows whose missing-symbol bitmap holds the position
//                     (mutations.cpp:74-83, 111-120), as a shared-memory difference array + scan, and the
//                     |F| terms of the deleted / flipped symbols.
//   k3_add_full       adds the precomputed tables of partitions whose filter is full (mutations.cpp:101-136).
// All arithmetic is uint32 with wrap-around, like the reference (mutations.cpp:75,80,94).
#pragma once
#include "common.cuh"

namespace sb {

constexpr int K1_THREADS = 512;    // 16 warps: one chunk of 32 entries per warp per stage
constexpr uint32_t K1_TINY = 8;    // arrays up to this cardinality: one container per lane, fully unrolled probes
constexpr uint32_t K1_STAGE_BYTES = STAGE_CHUNKS * 32 * 8 + STAGE_SMALL_U16 * 2 + 128;   // entries + small payload + StageRec
constexpr uint32_t K1_FILTER_BYTES = SLAB_WORDS32 * 4;

__device__ __forceinline__ uint32_t laneId() { return threadIdx.x & 31; }
__device__ __forceinline__ uint32_t laneMaskLt() { return (1u << laneId()) - 1u; }
__device__ __forceinline__ int warpSum(int v) { return __reduce_add_sync(0xFFFFFFFFu, v); }

// ---- bulk async copy (TMA 1-D) + mbarrier, sm_90+ PTX ------------------------------------------------
__device__ __forceinline__ uint32_t smemAddr(const void* p) { return static_cast<uint32_t>(__cvta_generic_to_shared(p)); }
__device__ __forceinline__ void mbarInit(uint64_t* bar, uint32_t count) {
   asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smemAddr(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbarExpectTx(uint64_t* bar, uint32_t bytes) {
   asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smemAddr(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulkLoad(void* dst_smem, const void* src_global, uint32_t bytes, uint64_t* bar) {
   asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                ::"r"(smemAddr(dst_smem)), "l"(src_global), "r"(bytes), "r"(smemAddr(bar)) : "memory");
}
__device__ __forceinline__ void mbarWait(uint64_t* bar, uint32_t parity) {
   asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "MBAR_WAIT:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra MBAR_DONE;\n"
      "bra MBAR_WAIT;\n"
      "MBAR_DONE:\n"
      "}\n" ::"r"(smemAddr(bar)), "r"(parity) : "memory");
}

// Adds `val` to base[idx] once per distinct idx among the lanes with pred set.
__device__ __forceinline__ void warpAggregatedSharedAdd(uint32_t* base, uint32_t idx, int val, bool pred) {
   const uint32_t mask = __ballot_sync(0xFFFFFFFFu, pred);
   if (!pred) return;
   const uint32_t peers = __match_any_sync(mask, idx);
   const int total = __reduce_add_sync(peers, val);
   if (laneId() == uint32_t(__ffs(peers) - 1) && total != 0) atomicAdd(&base[idx], uint32_t(total));
}

// |F ∩ [key*65536 + a, key*65536 + b]| through the rank tables (a <= b < 65536).
__device__ __forceinline__ uint32_t filterRangeCount(const PartQuery& q, uint32_t key, uint32_t a, uint32_t b) {
   const uint64_t* f64 = reinterpret_cast<const uint64_t*>(q.filter32);
   const uint32_t w0 = key * 1024u + (a >> 6), w1 = key * 1024u + (b >> 6);
   const uint32_t r0 = __ldg(q.rank_block + (w0 >> 8)) + __ldg(q.rank_local + w0);
   const uint32_t r1 = __ldg(q.rank_block + (w1 >> 8)) + __ldg(q.rank_local + w1);
   const uint64_t x0 = __ldg(f64 + w0), x1 = __ldg(f64 + w1);
   const uint64_t below_a = (a & 63) ? (x0 & (~0ULL >> (64 - (a & 63)))) : 0ULL;
   const uint64_t upto_b = x1 & (~0ULL >> (63 - (b & 63)));
   return r1 - r0 + __popcll(upto_b) - __popcll(below_a);
}

// Shared memory of one CTA: [filter slab 32 KB][stage buffer 0][stage buffer 1][tile counts][3 mbarriers].
// A stage buffer holds the stage's entries (4 KB), its small payload (12 KB) and its StageRec (128 B).
struct K1Stage {
   uint64_t entries[STAGE_CHUNKS * 32];
   uint16_t small[STAGE_SMALL_U16];
   StageRec rec;
};
static_assert(sizeof(K1Stage) == K1_STAGE_BYTES, "stage buffer layout");
// Large containers (bitsets, big arrays / runs) of the current stage, queued by the lanes that own their
// entries and then spread over ALL warps of the CTA (a chunk of 32 bitsets must not serialise on one warp).
struct K1Large { uint32_t offset16; uint32_t meta; uint32_t target; };   // meta: kind | key_local << 2 | size << 4 ; target: slot | deleted << 14 | flipped << 19 | key << 20 (runs)
constexpr uint32_t K1_LARGE_BYTES = STAGE_CHUNKS * 32 * sizeof(K1Large) + 16;

__device__ __forceinline__ void k1IssueStage(const SeqDev& s, uint32_t stage, K1Stage* buf, uint64_t* bar) {
   // one thread: read the issue fields of the stage record, then three bulk copies completing on `bar`
   const uint4 head = __ldg(reinterpret_cast<const uint4*>(s.stages + stage));        // entry_start, n_entries, small_off16 (lo, hi)
   const uint32_t small_bytes = __ldg(reinterpret_cast<const uint32_t*>(s.stages + stage) + 4);
   const uint32_t entry_bytes = ((head.y + 1u) & ~1u) * 8u;
   const uint64_t small_off16 = uint64_t(head.z) | (uint64_t(head.w) << 32);
   mbarExpectTx(bar, entry_bytes + small_bytes + 128u);
   bulkLoad(buf->entries, s.entries + head.x, entry_bytes, bar);
   if (small_bytes) bulkLoad(buf->small, reinterpret_cast<const uint4*>(s.small_pool) + small_off16, small_bytes, bar);
   bulkLoad(&buf->rec, s.stages + stage, 128u, bar);
}

__global__ void __launch_bounds__(K1_THREADS)
k1_and_popcount(const SeqDev s, const PartQuery* __restrict__ pq, uint32_t* __restrict__ counts, int full_mode) {
   extern __shared__ __align__(128) uint8_t k1_smem[];
   const WorkItem item = s.items[blockIdx.x];
   const SlabRec slab = s.slabs[item.slab];
   const PartQuery q = pq[slab.partition];
   if (q.state != PART_PARTIAL) return;

   uint32_t* filter_s = reinterpret_cast<uint32_t*>(k1_smem);
   K1Stage* stage_s = reinterpret_cast<K1Stage*>(k1_smem + K1_FILTER_BYTES);
   uint32_t* tile_counts = reinterpret_cast<uint32_t*>(k1_smem + K1_FILTER_BYTES + 2 * K1_STAGE_BYTES);
   const uint32_t tile_slots = s.tile_positions * s.sigma;
   uint64_t* bars = reinterpret_cast<uint64_t*>(tile_counts + tile_slots + (tile_slots & 1));   // [0] filter, [1], [2] stage buffers
   uint32_t* n_large_s = reinterpret_cast<uint32_t*>(bars + 3);                                  // [0] queue length (16-byte slot)
   K1Large* large_list = reinterpret_cast<K1Large*>(n_large_s + 4);

   if (threadIdx.x == 0) {
      mbarInit(&bars[0], 1); mbarInit(&bars[1], 1); mbarInit(&bars[2], 1);
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
   }
   __syncthreads();
   if (threadIdx.x == 0) {
      mbarExpectTx(&bars[0], K1_FILTER_BYTES);
      bulkLoad(filter_s, q.filter32 + size_t(slab.key0) * 2048u, K1_FILTER_BYTES, &bars[0]);
      k1IssueStage(s, item.stage_begin, &stage_s[0], &bars[1]);
   }
   for (uint32_t i = threadIdx.x; i < tile_slots; i += blockDim.x) tile_counts[i] = 0;
   if (threadIdx.x == 0) { n_large_s[0] = 0; n_large_s[1] = 0; }
   __syncthreads();
   mbarWait(&bars[0], 0);

   const uint32_t lane = laneId(), warp = threadIdx.x >> 5;
   const uint4* filter128 = reinterpret_cast<const uint4*>(filter_s);
   for (uint32_t st = item.stage_begin; st < item.stage_end; ++st) {
      const uint32_t it = st - item.stage_begin, b = it & 1;
      if (threadIdx.x == 0) {
         n_large_s[b ^ 1] = 0;   // the queue counter of the next stage; its last users left at the previous barrier
         if (st + 1 < item.stage_end) k1IssueStage(s, st + 1, &stage_s[b ^ 1], &bars[1 + (b ^ 1)]);
      }
      mbarWait(&bars[1 + b], (it >> 1) & 1);
      const K1Stage& stage = stage_s[b];
      const uint32_t n_entries = stage.rec.n_entries;
      const uint32_t c = warp;                      // chunk of this warp
      if (c * 32 < n_entries) {
         const uint32_t idx = c * 32 + lane;
         const bool active = idx < n_entries;
         const uint64_t e = active ? stage.entries[idx] : (uint64_t(KIND_NULL) << 30);
         const uint32_t kind = EntryBits::kind(e), size = EntryBits::size(e), key_local = (EntryBits::key(e) - slab.key0) & (SLAB_KEYS - 1);
         const uint32_t small_units = EntryBits::smallUnits(kind, size), large_units = EntryBits::largeUnits(kind, size);
         // warp prefix sums of both payload streams (packed: small in the low half, large in the high half)
         uint64_t incl = uint64_t(small_units) | (uint64_t(large_units) << 32);
         const uint64_t own = incl;
#pragma unroll
         for (int d = 1; d < 32; d <<= 1) { const uint64_t t = __shfl_up_sync(0xFFFFFFFFu, incl, d); if (lane >= uint32_t(d)) incl += t; }
         const uint64_t excl = incl - own;
         const uint16_t* small = stage.small + stage.rec.chunk_small[c] + uint32_t(excl);
         const uint32_t* f = filter_s + key_local * 2048u;
         int cnt = 0;

         // (1) tiny arrays: one container per lane, predicated unrolled probes of the shared-memory filter
         if (kind == KIND_ARRAY && size <= K1_TINY) {
#pragma unroll
            for (uint32_t i = 0; i < K1_TINY; ++i) {
               if (i < size) { const uint32_t v = small[i]; cnt += (f[v >> 5] >> (v & 31)) & 1u; }
            }
         } else if (kind == KIND_ARRAY && size <= SMALL_ARRAY_MAX) {   // (2) small arrays: per-lane loop
            for (uint32_t i = 0; i < size; ++i) { const uint32_t v = small[i]; cnt += (f[v >> 5] >> (v & 31)) & 1u; }
         } else if (kind == KIND_RUN && size <= SMALL_RUN_MAX) {        // (3) small run containers: rank lookups per run
            for (uint32_t i = 0; i < size; ++i) { const uint32_t start = small[2 * i], len1 = small[2 * i + 1]; cnt += int(filterRangeCount(q, EntryBits::key(e), start, start + len1)); }
         }
         // (4) large containers are queued for the whole CTA
         if (large_units != 0) {
            const uint32_t at = atomicAdd(&n_large_s[b], 1u);
            large_list[at] = {stage.rec.chunk_large[c] + uint32_t(excl >> 32), kind | (key_local << 2) | (size << 4),
                              EntryBits::slot(e) | (EntryBits::deleted(e) << 14) | (EntryBits::flipped(e) << 19)};
         }

         // accumulate: symbol slot, and the deleted symbol of the same (partition, position) (mutations.cpp:90-95)
         const uint32_t slot = EntryBits::slot(e), deleted = EntryBits::deleted(e);
         const int add_slot = EntryBits::flipped(e) ? -cnt : cnt;
         const int add_deleted = full_mode ? -cnt : -add_slot;
         warpAggregatedSharedAdd(tile_counts, slot, add_slot, cnt != 0);
         warpAggregatedSharedAdd(tile_counts, (slot / s.sigma) * s.sigma + deleted, add_deleted, cnt != 0 && deleted != 31u);
      }
      __syncthreads();
      // phase B: the queued large containers, one per warp at a time, 128-bit loads straight from the large pool
      const uint32_t n_large = n_large_s[b];
      if (n_large) {
         const uint4* large_base = s.large_pool + stage.rec.large_off16;
         for (uint32_t j = warp; j < n_large; j += blockDim.x >> 5) {
            const K1Large w = large_list[j];
            const uint4* src = large_base + w.offset16;
            const uint32_t k_kind = w.meta & 3u, k_local = (w.meta >> 2) & 3u, k_size = w.meta >> 4;
            int c2 = 0;
            if (k_kind == KIND_BITSET) {
               const uint4* fk = filter128 + k_local * 512u;
#pragma unroll
               for (uint32_t half = 0; half < 2; ++half) {
                  uint4 x[8];
#pragma unroll
                  for (uint32_t t = 0; t < 8; ++t) x[t] = __ldg(src + (half * 8 + t) * 32 + lane);
#pragma unroll
                  for (uint32_t t = 0; t < 8; ++t) {
                     const uint4 y = fk[(half * 8 + t) * 32 + lane];
                     c2 += __popc(x[t].x & y.x) + __popc(x[t].y & y.y) + __popc(x[t].z & y.z) + __popc(x[t].w & y.w);
                  }
               }
            } else if (k_kind == KIND_ARRAY) {
               const uint32_t* fk = filter_s + k_local * 2048u;
               for (uint32_t i = lane; i * 8 < k_size; i += 32) {
                  const uint4 x = __ldg(src + i);
                  const uint32_t wv[4] = {x.x, x.y, x.z, x.w};
#pragma unroll
                  for (uint32_t t = 0; t < 8; ++t) {
                     const uint32_t v = (wv[t >> 1] >> ((t & 1) * 16)) & 0xFFFFu;
                     if (i * 8 + t < k_size) c2 += (fk[v >> 5] >> (v & 31)) & 1u;
                  }
               }
            } else {   // large run container: two rank lookups per run
               const uint16_t* runs = reinterpret_cast<const uint16_t*>(src);
               for (uint32_t i = lane; i < k_size; i += 32) { const uint32_t start = __ldg(runs + 2 * i), len1 = __ldg(runs + 2 * i + 1); c2 += int(filterRangeCount(q, slab.key0 + k_local, start, start + len1)); }
            }
            c2 = warpSum(c2);
            if (lane == 0 && c2 != 0) {
               const uint32_t slot = w.target & 0x3FFFu, deleted = (w.target >> 14) & 31u;
               const int add_slot = ((w.target >> 19) & 1u) ? -c2 : c2;
               atomicAdd(&tile_counts[slot], uint32_t(add_slot));
               if (deleted != 31u) atomicAdd(&tile_counts[(slot / s.sigma) * s.sigma + deleted], uint32_t(full_mode ? -c2 : -add_slot));
            }
         }
      }
      __syncthreads();   // every warp is done with buffer b (and the queue) before it is refilled
   }
   // flush: consecutive threads -> consecutive positions of one symbol row of counts[sym][pos]
   const uint32_t tp = s.tile_positions;
   for (uint32_t i = threadIdx.x; i < tile_slots; i += blockDim.x) {
      const uint32_t sym = i / tp, pl = i - sym * tp;
      const uint32_t v = tile_counts[pl * s.sigma + sym];
      const uint32_t pos = item.tile * tp + pl;
      if (v != 0 && pos < s.length) atomicAdd(counts + size_t(sym) * s.length + pos, v);
   }
}

// ---- K2 ---------------------------------------------------------------------------------------------
constexpr int K2_THREADS = 1024;
constexpr uint32_t K2_MAX_WINDOW = 49152;   // positions per shared-memory difference array (192 KB)

// grid: (row blocks, listed partitions, position windows). `plist` = partitions to process; a block owns
// `rows_per_block` rows of one partition (the host sizes it so that the grid is about one wave).
__global__ void __launch_bounds__(K2_THREADS)
k2_missing(const SeqDev s, const PartQuery* __restrict__ pq, const uint32_t* __restrict__ plist, uint32_t* __restrict__ counts,
           uint32_t window, uint32_t rows_per_block, int full_mode) {
   extern __shared__ int diff[];   // [window + 1]
   __shared__ int warp_sums[32];
   __shared__ uint32_t n_filtered_s;
   const uint32_t partition = plist[blockIdx.y];
   const PartQuery q = pq[partition];
   const uint32_t row0 = blockIdx.x * rows_per_block;
   if (row0 >= q.n_rows) return;
   const uint32_t row1 = min(q.n_rows, row0 + rows_per_block);
   const uint32_t w0 = blockIdx.z * window, w1 = min(s.length, w0 + window), width = w1 - w0;
   for (uint32_t i = threadIdx.x; i <= window; i += blockDim.x) diff[i] = 0;
   if (threadIdx.x == 0) n_filtered_s = 0;
   __syncthreads();

   const uint32_t* row_off = s.miss_row_off + s.miss_row_base[partition];
   const uint2* iv = s.miss_iv + s.miss_iv_base[partition];
   uint32_t local_filtered = 0;
   for (uint32_t base = row0 + threadIdx.x; base < row1; base += 4 * K2_THREADS) {   // 4 rows in flight per thread
      uint32_t b[4], e[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) {
         const uint32_t row = base + k * K2_THREADS;
         const bool in = row < row1 && ((__ldg(q.filter32 + (row >> 5)) >> (row & 31)) & 1u);
         b[k] = e[k] = 0;
         if (in) { ++local_filtered; b[k] = __ldg(row_off + row); e[k] = __ldg(row_off + row + 1); }
      }
#pragma unroll
      for (int k = 0; k < 4; ++k) {
         for (uint32_t j = b[k]; j < e[k]; ++j) {
            const uint2 r = __ldg(iv + j);
            const uint32_t a = max(r.x, w0), z = min(r.y, w1);
            if (a < z) { atomicAdd(&diff[a - w0], 1); atomicAdd(&diff[z - w0], -1); }
         }
      }
   }
   local_filtered = uint32_t(warpSum(int(local_filtered)));
   if (laneId() == 0 && local_filtered) atomicAdd(&n_filtered_s, local_filtered);
   __syncthreads();
   const uint32_t n_filtered = n_filtered_s;
   if (n_filtered == 0) return;

   // inclusive scan of diff in place: every thread owns `per` consecutive elements
   const uint32_t per = (width + K2_THREADS - 1) / K2_THREADS;
   const uint32_t begin = min(threadIdx.x * per, width), end = min(begin + per, width);
   int sum = 0;
   for (uint32_t i = begin; i < end; ++i) sum += diff[i];
   const uint32_t warp = threadIdx.x >> 5, lane = laneId();
   int incl = sum;
#pragma unroll
   for (int d = 1; d < 32; d <<= 1) { const int t = __shfl_up_sync(0xFFFFFFFFu, incl, d); if (lane >= uint32_t(d)) incl += t; }
   if (lane == 31) warp_sums[warp] = incl;
   __syncthreads();
   if (warp == 0) {
      int ws = warp_sums[lane];
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) { const int t = __shfl_up_sync(0xFFFFFFFFu, ws, d); if (lane >= uint32_t(d)) ws += t; }
      warp_sums[lane] = ws;
   }
   __syncthreads();
   int running = incl - sum + (warp ? warp_sums[warp - 1] : 0);
   for (uint32_t i = begin; i < end; ++i) { running += diff[i]; diff[i] = running; }
   __syncthreads();

   const int8_t* deleted = s.deleted + size_t(partition) * s.length;
   const int8_t* flipped = s.flipped + size_t(partition) * s.length;
   for (uint32_t i = threadIdx.x; i < width; i += blockDim.x) {   // coalesced over positions
      const uint32_t pos = w0 + i;
      const int d = deleted[pos], f = flipped[pos];
      if (d >= 0) {
         uint32_t add = n_filtered - uint32_t(diff[i]);          // |F| - #{i in F : pos in M_i}
         if (f >= 0 && !full_mode) add -= n_filtered;              // deleted -= |F| part of the flipped symbol's count
         if (add) atomicAdd(counts + size_t(d) * s.length + pos, add);
      }
      if (f >= 0) atomicAdd(counts + size_t(f) * s.length + pos, n_filtered);
   }
}

// ---- row selection (mutations.cpp:196-215) --------------------------------------------------------------
struct ValidSymbols { uint8_t symbol[32]; uint32_t n; };

__global__ void k4_select_rows(const uint32_t* __restrict__ counts, uint32_t length, const uint8_t* __restrict__ reference, ValidSymbols valid,
                               double min_proportion, sb_mutation_row* __restrict__ rows, uint32_t capacity, uint32_t* __restrict__ n_rows) {
   const uint32_t pos = blockIdx.x * blockDim.x + threadIdx.x;
   if (pos >= length) return;
   uint32_t total = 0;
   for (uint32_t k = 0; k < valid.n; ++k) total += counts[size_t(valid.symbol[k]) * length + pos];
   if (total == 0) return;
   const uint32_t threshold = min_proportion == 0 ? 0u : static_cast<uint32_t>(ceil(__dmul_rn(static_cast<double>(total), min_proportion)) - 1);
   const uint32_t ref = reference[pos];
   for (uint32_t k = 0; k < valid.n; ++k) {
      const uint32_t sym = valid.symbol[k];
      if (sym == ref) continue;
      const uint32_t count = counts[size_t(sym) * length + pos];
      if (count > threshold) {
         const uint32_t at = atomicAdd(n_rows, 1u);
         if (at < capacity) rows[at] = {pos, k, count, total};   // symbol field carries the index into `valid` until sorted on the host
      }
   }
}

// counts[i] += sum over listed full partitions of their precomputed table
__global__ void k3_add_full(const uint32_t* __restrict__ full_tables, size_t table_len, const uint32_t* __restrict__ plist, uint32_t n_list,
                            uint32_t* __restrict__ counts) {
   const size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
   if (i >= table_len) return;
   uint32_t acc = 0;
   for (uint32_t k = 0; k < n_list; ++k) acc += __ldg(full_tables + size_t(plist[k]) * table_len + i);
   counts[i] += acc;
}

}  // namespace sb
