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

// tuning knobs of the container scan; the defaults are the measured best (tools/k1_experiment.py builds variants with -D)
#ifndef SB_K1_THREADS
#define SB_K1_THREADS 256
#endif
#ifndef SB_K1_UNROLL
#define SB_K1_UNROLL 4
#endif
#ifndef SB_K1_PROBE
#define SB_K1_PROBE 0
#endif
#ifndef SB_K1_MIN_CTAS
#define SB_K1_MIN_CTAS 4
#endif
constexpr int K1_THREADS = SB_K1_THREADS;   // 8 warps; up to 4 CTAs per SM (48-57 KB of shared memory each)
constexpr int K1_UNROLL = SB_K1_UNROLL;     // 128-bit posting loads in flight per thread
#ifdef SB_K1_TIMELINE
// diagnostics build only: per-CTA time stamps {smid, start, filter ready, warp 0 out of units, after the barrier, end, units, weight}
__device__ unsigned long long* g_k1_timeline = nullptr;
__device__ __forceinline__ unsigned long long globalTimerNs() { unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); return t; }
__device__ __forceinline__ uint32_t smId() { uint32_t v; asm volatile("mov.u32 %0, %%smid;" : "=r"(v)); return v; }
#define K1_STAMP(k) do { if (g_k1_timeline && threadIdx.x == 0) g_k1_timeline[size_t(blockIdx.x) * 8 + (k)] = globalTimerNs(); } while (0)
#else
#define K1_STAMP(k) do { } while (0)
#endif
constexpr uint32_t K1_FILTER_BYTES = SLAB_WORDS32 * 4;
// count table of a tile in shared memory: tile_slots counters + 32 dummy slots (one per lane: where the misses go)
__host__ __device__ constexpr uint32_t k1TableWords(uint32_t tile_slots) { return (tile_slots + 32 + 1) & ~1u; }

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

// 128-bit load of data that is read once per query
__device__ __forceinline__ uint4 loadStream(const uint4* p) {
#ifdef SB_K1_LDG_NOALLOC
   uint4 v;
   asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
   return v;
#else
   return __ldg(p);
#endif
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

// Block -> work item: a per-query list of the work items of the partial partitions, heaviest first over all
// of them (longest-processing-time order: the block scheduler hands out CTAs in index order). The filter
// pointers of the first K1_INLINE_PARTITIONS partitions travel in kernel-parameter (constant) memory so that
// the bulk copy of the filter slab is one dependent load closer.
constexpr uint32_t K1_INLINE_PARTITIONS = 64;
struct K1Map {
   const uint32_t* item_list;                              // k1_stream: run table (kernels_stream.cuh)
   const WorkItem* query_items;                            // k1_and_popcount: [grid] the query's work items, heaviest first (a copy: one dependent load less per CTA)
   const uint32_t* inline_filter[K1_INLINE_PARTITIONS];    // filter32 of partition p (p < K1_INLINE_PARTITIONS)
};

// One CTA per work item. Shared memory: [filter slab 32 KB][tile counts (tile_slots + 1 words)][mbarrier].
// tile_counts[sym * T + pos_in_tile] accumulates the RAW |F AND B[pos][sym]|; the flipped / deleted-symbol
// arithmetic (mutations.cpp:85-95, 122-133) is linear in it and is applied once per slot in the flush.
__global__ void __launch_bounds__(K1_THREADS, SB_K1_MIN_CTAS)
k1_and_popcount(const SeqDev s, const PartQuery* __restrict__ pq, const K1Map map, uint32_t* __restrict__ counts, int full_mode) {
   extern __shared__ __align__(128) uint8_t k1_smem[];
   K1_STAMP(1);
   const WorkItem item = map.query_items[blockIdx.x];
   const SlabRec slab{item.partition, item.key0};
   const uint32_t* filter32 = slab.partition < K1_INLINE_PARTITIONS ? map.inline_filter[slab.partition] : pq[slab.partition].filter32;

   uint32_t* filter_s = reinterpret_cast<uint32_t*>(k1_smem);
   uint32_t* tile_counts = reinterpret_cast<uint32_t*>(k1_smem + K1_FILTER_BYTES);
   const uint32_t T = s.tile_positions, tile_slots = T * s.sigma;
   uint64_t* bar = reinterpret_cast<uint64_t*>(tile_counts + k1TableWords(tile_slots));
   uint32_t* next_unit = reinterpret_cast<uint32_t*>(bar + 1);

   if (threadIdx.x == 0) {
      mbarInit(bar, 1);
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
      mbarExpectTx(bar, K1_FILTER_BYTES);
      bulkLoad(filter_s, filter32 + size_t(slab.key0) * 2048u, K1_FILTER_BYTES, bar);
   }
   for (uint32_t i = threadIdx.x; i < tile_slots + 32; i += K1_THREADS) tile_counts[i] = 0;

   // flags of this thread's flush position, fetched now so that the flush does not wait for them
   const uint32_t flush_pos = item.tile * T + threadIdx.x;
   int flush_del = -1, flush_flip = -1;
   if (threadIdx.x < T && flush_pos < s.length) {
      flush_del = s.deleted[size_t(slab.partition) * s.length + flush_pos];
      flush_flip = s.flipped[size_t(slab.partition) * s.length + flush_pos];
   }
   // Work units of the item, handed out to the warps through a shared counter (a warp that is busy with a bitset
   // simply takes fewer): first the large containers -- a bitset as two halves of 4 KB, anything else whole --
   // then the postings in chunks of 128 x 16 bytes. The first unit of a warp is its warp id; the loads of a posting
   // chunk are issued one unit ahead (and before waiting for the filter: they do not depend on it).
   constexpr uint32_t CHUNK4 = 32 * K1_UNROLL;    // 128-bit words per chunk
   constexpr uint32_t N_WARPS = K1_THREADS / 32;
   const uint32_t lane = laneId(), warp = threadIdx.x >> 5;
   const uint4* post4 = reinterpret_cast<const uint4*>(s.postings + item.post_begin);
   const uint32_t n4 = (item.post_end - item.post_begin + 3u) >> 2;
   const uint32_t n_large_units = 2u * (item.large_end - item.large_begin);
   const uint32_t n_units = n_large_units + (n4 + CHUNK4 - 1) / CHUNK4;
   const uint32_t dummy = tile_slots << POSTING_ROW_BITS;          // a posting of the slot that is never flushed
   const uint4 none = make_uint4(dummy, dummy, dummy, dummy);
   uint4 v[K1_UNROLL];
   auto loadChunk = [&](uint32_t unit, uint4 (&dst)[K1_UNROLL]) {
      if (unit < n_large_units || unit >= n_units) return;
      const uint32_t base = (unit - n_large_units) * CHUNK4 + lane;
#pragma unroll
      for (int k = 0; k < K1_UNROLL; ++k) dst[k] = base + k * 32 < n4 ? loadStream(post4 + base + k * 32) : none;
   };
   auto grab = [&]() {
      uint32_t next = 0;
      if (lane == 0) next = atomicAdd(next_unit, 1u);
      return __shfl_sync(0xFFFFFFFFu, next, 0);
   };
   uint32_t unit = warp;
   loadChunk(unit, v);
   if (threadIdx.x == 0) *next_unit = N_WARPS;
   __syncthreads();      // counts zeroed, mbarrier and unit counter initialised
   mbarWait(bar, 0);
   K1_STAMP(2);

   // Eight postings at a time: the eight filter words are fetched first (independent shared-memory loads), then the hits
   // are counted with unconditional shared-memory increments -- no data-dependent branch, one load round trip per eight postings.
   auto probe8 = [&](const uint4& a, const uint4& b) {
      const uint32_t p[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
      uint32_t w[8];
#pragma unroll
      for (int t = 0; t < 8; ++t) w[t] = filter_s[(p[t] >> 5) & (SLAB_WORDS32 - 1)];
#pragma unroll
      for (int t = 0; t < 8; ++t) {
         const bool hit = (w[t] >> (p[t] & 31)) & 1u;
#if SB_K1_PROBE == 0      // a miss increments the lane's dummy slot: unconditional, so the compiler emits no branch
         atomicAdd(&tile_counts[hit ? (p[t] >> POSTING_ROW_BITS) : tile_slots + lane], 1u);
#elif SB_K1_PROBE == 1    // one dummy slot for all lanes
         atomicAdd(&tile_counts[hit ? (p[t] >> POSTING_ROW_BITS) : tile_slots], 1u);
#else                     // branch around the increment
         if (hit) atomicAdd(&tile_counts[p[t] >> POSTING_ROW_BITS], 1u);
#endif
      }
   };
   while (unit < n_units) {
      if (unit < n_large_units) {
         const LargeEntry w = s.large_entries[item.large_begin + (unit >> 1)];
         const uint32_t half = unit & 1u;
         const uint4* src = s.large_pool + w.offset16;
         const uint32_t k_kind = w.kind(), k_local = w.keyLocal(), k_size = w.size;
         int c2 = 0;
         if (k_kind == KIND_BITSET) {
            const uint4* fk = reinterpret_cast<const uint4*>(filter_s) + k_local * 512u;
            uint4 x[8];
#pragma unroll
            for (uint32_t t = 0; t < 8; ++t) x[t] = __ldg(src + (half * 8 + t) * 32 + lane);
#pragma unroll
            for (uint32_t t = 0; t < 8; ++t) {
               const uint4 y = fk[(half * 8 + t) * 32 + lane];
               c2 += __popc(x[t].x & y.x) + __popc(x[t].y & y.y) + __popc(x[t].z & y.z) + __popc(x[t].w & y.w);
            }
         } else if (half == 0 && k_kind == KIND_ARRAY) {
            const uint32_t* fk = filter_s + k_local * 2048u;
            const uint32_t n_vec = (k_size + 7) >> 3;
            for (uint32_t i0 = lane; i0 < n_vec; i0 += 128) {      // four 128-bit loads per lane in flight
               uint4 x[4];
#pragma unroll
               for (uint32_t k = 0; k < 4; ++k) x[k] = i0 + k * 32 < n_vec ? __ldg(src + i0 + k * 32) : make_uint4(0, 0, 0, 0);
#pragma unroll
               for (uint32_t k = 0; k < 4; ++k) {
                  const uint32_t wv[4] = {x[k].x, x[k].y, x[k].z, x[k].w};
#pragma unroll
                  for (uint32_t t = 0; t < 8; ++t) {      // the padding of the last vector is zero: probed like a value, masked out
                     const uint32_t val = (wv[t >> 1] >> ((t & 1) * 16)) & 0xFFFFu;
                     c2 += ((fk[val >> 5] >> (val & 31)) & 1u) & uint32_t((i0 + k * 32) * 8 + t < k_size);
                  }
               }
            }
         } else if (half == 0) {   // long run container: two rank lookups per run
            const uint16_t* runs = reinterpret_cast<const uint16_t*>(src);
            const PartQuery q = pq[slab.partition];
            for (uint32_t i = lane; i < k_size; i += 32) { const uint32_t start = __ldg(runs + 2 * i), len1 = __ldg(runs + 2 * i + 1); c2 += int(filterRangeCount(q, slab.key0 + k_local, start, start + len1)); }
         }
         c2 = warpSum(c2);
         if (lane == 0 && c2 != 0) atomicAdd(&tile_counts[w.slot()], uint32_t(c2));
         unit = grab();
         loadChunk(unit, v);
      } else {
         const uint32_t next = grab();
         uint4 nx[K1_UNROLL];
         loadChunk(next, nx);
#pragma unroll
         for (int k = 0; k < K1_UNROLL; k += 2) probe8(v[k], v[k + 1]);   // loads past the item's end carry postings of the dummy slot
#pragma unroll
         for (int k = 0; k < K1_UNROLL; ++k) v[k] = nx[k];
         unit = next;
      }
   }
   K1_STAMP(3);
   __syncthreads();
   K1_STAMP(4);
   // (3) flush: thread = position of the tile; symbol rows of counts[sym][pos] get coalesced REDs
   static_assert(TILE_POSITIONS <= K1_THREADS, "one flush position per thread");
   if (threadIdx.x < T && flush_pos < s.length) {
      const uint32_t pl = threadIdx.x;
      int acc_deleted = 0;
      for (uint32_t sym0 = 0; sym0 < s.sigma; sym0 += 8) {      // eight counters in flight, then their REDs
         int raw[8];
#pragma unroll
         for (uint32_t q = 0; q < 8; ++q) raw[q] = sym0 + q < s.sigma ? int(tile_counts[(sym0 + q) * T + pl]) : 0;
#pragma unroll
         for (uint32_t q = 0; q < 8; ++q) {
            const uint32_t sym = sym0 + q;
            if (raw[q] == 0 || int(sym) == flush_del) continue;
            const int add_slot = int(sym) == flush_flip ? -raw[q] : raw[q];
            atomicAdd(counts + size_t(sym) * s.length + flush_pos, uint32_t(add_slot));
            acc_deleted += full_mode ? -raw[q] : -add_slot;
         }
      }
      if (flush_del >= 0 && acc_deleted != 0) atomicAdd(counts + size_t(flush_del) * s.length + flush_pos, uint32_t(acc_deleted));
   }
#ifdef SB_K1_TIMELINE
   if (g_k1_timeline && threadIdx.x == 0) {
      g_k1_timeline[size_t(blockIdx.x) * 8 + 0] = smId();
      g_k1_timeline[size_t(blockIdx.x) * 8 + 5] = globalTimerNs();
      g_k1_timeline[size_t(blockIdx.x) * 8 + 6] = (uint64_t(n_large_units) << 32) | n_units;
      g_k1_timeline[size_t(blockIdx.x) * 8 + 7] = 16ull * n4;
   }
#endif
}

// ---- K2 ---------------------------------------------------------------------------------------------
constexpr int K2_THREADS = 1024;
constexpr uint32_t K2_MAX_WINDOW = 49152;   // positions per shared-memory difference array (192 KB)

// Row blocks of a partition: the work of a block is its filtered rows (three dependent scattered loads each) far more than
// its rows, and the filter's density differs from partition to partition, so a partition gets blocks in proportion to
// cost = rows / 16 + filtered rows out of a budget of about one wave (host and kernel evaluate the same formula).
__host__ __device__ inline uint64_t k2Cost(uint32_t n_rows, uint32_t cardinality) { return uint64_t(n_rows) / 16 + cardinality; }
__host__ __device__ inline uint32_t k2Blocks(uint32_t n_rows, uint32_t cardinality, uint64_t total_cost, uint32_t budget) {
   const uint32_t share = 1u + uint32_t(k2Cost(n_rows, cardinality) * budget / (total_cost ? total_cost : 1));
   const uint32_t cap = (n_rows + 4095u) / 4096u;      // no block under 4,096 rows: every block pays the scan + flush over the positions
   return share < cap ? share : (cap ? cap : 1u);
}

// grid: (row blocks, listed partitions, position windows). `plist` = partitions to process; a block owns
// an equal share of the rows of one partition (k2Blocks of them; the grid is about one wave).
__global__ void __launch_bounds__(K2_THREADS)
k2_missing(const SeqDev s, const PartQuery* __restrict__ pq, const uint32_t* __restrict__ plist, uint32_t* __restrict__ acc,
           uint32_t window, uint64_t total_cost, uint32_t budget) {
   extern __shared__ int diff[];   // [window + 1]
   __shared__ int warp_sums[32];
   __shared__ uint32_t n_filtered_s;
   const uint32_t partition = plist[blockIdx.y];
   const PartQuery q = pq[partition];
   const uint32_t n_blocks = k2Blocks(q.n_rows, q.cardinality, total_cost, budget);
   if (blockIdx.x >= n_blocks) return;
   const uint32_t rows_per_block = ((q.n_rows + n_blocks - 1) / n_blocks + 31u) & ~31u;
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

   // hand the block's per-position counts to the per-partition accumulator: only positions that some filtered row
   // misses carry a value (the ends of the genome and scattered stretches), so this is far fewer atomics than positions
   uint32_t* mine = acc + size_t(blockIdx.y) * (s.length + 1);
   for (uint32_t i = threadIdx.x; i < width; i += blockDim.x) {
      const int v = diff[i];
      if (v != 0) atomicAdd(mine + w0 + i, uint32_t(v));
   }
   if (threadIdx.x == 0 && blockIdx.z == 0) atomicAdd(mine + s.length, n_filtered);
}

// acc[j][pos] = #{i in F_p : pos in M_i}, acc[j][L] = |F_p| for the j-th listed partition p  ->  the |F| terms of the
// deleted / flipped symbols (mutations.cpp:74-83, 111-133). grid: (ceil(L / 256), listed partitions)
__global__ void k2_finish(const SeqDev s, const uint32_t* __restrict__ plist, const uint32_t* __restrict__ acc, uint32_t* __restrict__ counts, int full_mode) {
   const uint32_t pos = blockIdx.x * blockDim.x + threadIdx.x;
   if (pos >= s.length) return;
   const uint32_t* mine = acc + size_t(blockIdx.y) * (s.length + 1);
   const uint32_t n_filtered = mine[s.length];
   if (n_filtered == 0) return;
   const uint32_t partition = plist[blockIdx.y];
   const int d = s.deleted[size_t(partition) * s.length + pos], f = s.flipped[size_t(partition) * s.length + pos];
   if (d >= 0) {
      uint32_t add = n_filtered - mine[pos];                    // |F| - #{i in F : pos in M_i}
      if (f >= 0 && !full_mode) add -= n_filtered;              // deleted -= |F| part of the flipped symbol's count
      if (add) atomicAdd(counts + size_t(d) * s.length + pos, add);
   }
   if (f >= 0) atomicAdd(counts + size_t(f) * s.length + pos, n_filtered);
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
