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

constexpr int K1_THREADS = 256;
constexpr uint32_t K1_TINY = 8;   // arrays up to this cardinality are handled one per lane

__device__ __forceinline__ uint32_t laneId() { return threadIdx.x & 31; }
__device__ __forceinline__ uint32_t laneMaskLt() { return (1u << laneId()) - 1u; }

__device__ __forceinline__ int warpSum(int v) { return __reduce_add_sync(0xFFFFFFFFu, v); }

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

__global__ void __launch_bounds__(K1_THREADS)
k1_and_popcount(const SeqDev s, const PartQuery* __restrict__ pq, uint32_t* __restrict__ counts, int full_mode) {
   extern __shared__ uint32_t tile_counts[];
   const WorkItem item = s.items[blockIdx.x];
   const uint32_t tile_slots = s.tile_positions * s.sigma;
   for (uint32_t i = threadIdx.x; i < tile_slots; i += blockDim.x) tile_counts[i] = 0;
   __syncthreads();

   const uint32_t lane = laneId();
   const uint32_t warp = threadIdx.x >> 5, n_warps = blockDim.x >> 5;
   for (uint32_t c = item.chunk_begin + warp; c < item.chunk_end; c += n_warps) {
      const uint4 raw = __ldg(reinterpret_cast<const uint4*>(s.chunks + c));
      const uint32_t partition = raw.w & 0xFFFF, n_entries = (raw.w >> 16) & 0xFF;
      const PartQuery q = pq[partition];
      if (q.state != PART_PARTIAL) continue;
      const bool active = lane < n_entries;
      const uint64_t e = active ? __ldg(s.entries + raw.x + lane) : 0ULL;
      const uint32_t kind = EntryBits::kind(e), size = EntryBits::size(e), key = EntryBits::key(e);
      const uint32_t units = !active ? 0u : (kind == KIND_ARRAY ? size : (kind == KIND_RUN ? 2u * size : 0u));
      uint32_t incl = units;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) { const uint32_t t = __shfl_up_sync(0xFFFFFFFFu, incl, d); if (lane >= uint32_t(d)) incl += t; }
      const uint64_t off = uint64_t(raw.y) * 8ULL + (incl - units);
      const uint32_t bitset_mask = __ballot_sync(0xFFFFFFFFu, active && kind == KIND_BITSET);
      const uint32_t bitset_idx = raw.z + __popc(bitset_mask & laneMaskLt());
      int cnt = 0;

      // (1) tiny arrays: one container per lane, one filter-bit probe per value
      if (active && kind == KIND_ARRAY && size <= K1_TINY) {
         const uint16_t* p = s.arrays + off;
         const uint32_t* f = q.filter32 + key * 2048u;
#pragma unroll 4
         for (uint32_t i = 0; i < size; ++i) { const uint32_t v = __ldg(p + i); cnt += (__ldg(f + (v >> 5)) >> (v & 31)) & 1u; }
      }
      // (2) larger arrays: the warp strides over one container
      uint32_t m = __ballot_sync(0xFFFFFFFFu, active && kind == KIND_ARRAY && size > K1_TINY);
      while (m) {
         const int l = __ffs(m) - 1; m &= m - 1;
         const uint64_t o = __shfl_sync(0xFFFFFFFFu, off, l);
         const uint32_t sz = __shfl_sync(0xFFFFFFFFu, size, l), k = __shfl_sync(0xFFFFFFFFu, key, l);
         const uint16_t* p = s.arrays + o;
         const uint32_t* f = q.filter32 + k * 2048u;
         int c2 = 0;
         for (uint32_t i = lane; i < sz; i += 32) { const uint32_t v = __ldg(p + i); c2 += (__ldg(f + (v >> 5)) >> (v & 31)) & 1u; }
         c2 = warpSum(c2);
         if (int(lane) == l) cnt = c2;
      }
      // (3) bitset containers: 128-bit loads of container and filter, AND + popcount
      m = bitset_mask;
      while (m) {
         const int l = __ffs(m) - 1; m &= m - 1;
         const uint32_t bi = __shfl_sync(0xFFFFFFFFu, bitset_idx, l), k = __shfl_sync(0xFFFFFFFFu, key, l);
         const uint4* b = s.bitsets + size_t(bi) * 512u;
         const uint4* f = reinterpret_cast<const uint4*>(q.filter32 + k * 2048u);
         int c2 = 0;
#pragma unroll 4
         for (uint32_t i = lane; i < 512; i += 32) {
            const uint4 x = __ldg(b + i), y = __ldg(f + i);
            c2 += __popc(x.x & y.x) + __popc(x.y & y.y) + __popc(x.z & y.z) + __popc(x.w & y.w);
         }
         c2 = warpSum(c2);
         if (int(lane) == l) cnt = c2;
      }
      // (4) run containers: two rank lookups per run
      m = __ballot_sync(0xFFFFFFFFu, active && kind == KIND_RUN);
      while (m) {
         const int l = __ffs(m) - 1; m &= m - 1;
         const uint64_t o = __shfl_sync(0xFFFFFFFFu, off, l);
         const uint32_t n_runs = __shfl_sync(0xFFFFFFFFu, size, l), k = __shfl_sync(0xFFFFFFFFu, key, l);
         const uint16_t* p = s.arrays + o;
         int c2 = 0;
         for (uint32_t i = lane; i < n_runs; i += 32) {
            const uint32_t start = __ldg(p + 2 * i), len1 = __ldg(p + 2 * i + 1);
            c2 += int(filterRangeCount(q, k, start, start + len1));
         }
         c2 = warpSum(c2);
         if (int(lane) == l) cnt = c2;
      }

      // accumulate: symbol slot, and the deleted symbol of the same (partition, position) (mutations.cpp:90-95)
      const uint32_t slot = EntryBits::slot(e), deleted = EntryBits::deleted(e);
      const int add_slot = EntryBits::flipped(e) ? -cnt : cnt;
      const int add_deleted = full_mode ? -cnt : -add_slot;
      warpAggregatedSharedAdd(tile_counts, slot, add_slot, active && cnt != 0);
      warpAggregatedSharedAdd(tile_counts, (slot / s.sigma) * s.sigma + deleted, add_deleted, active && cnt != 0 && deleted != 31u);
   }
   __syncthreads();
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
constexpr uint32_t K2_ROWS_PER_BLOCK = 32768;
constexpr uint32_t K2_MAX_WINDOW = 49152;   // positions per shared-memory difference array (192 KB)

// grid: (row blocks, listed partitions, position windows). `plist` = partitions to process.
__global__ void __launch_bounds__(K2_THREADS)
k2_missing(const SeqDev s, const PartQuery* __restrict__ pq, const uint32_t* __restrict__ plist, uint32_t* __restrict__ counts,
           uint32_t window, int full_mode) {
   extern __shared__ int diff[];   // [window + 1]
   __shared__ int warp_sums[32];
   __shared__ int carry_s;
   __shared__ uint32_t n_filtered_s;
   const uint32_t partition = plist[blockIdx.y];
   const PartQuery q = pq[partition];
   const uint32_t row0 = blockIdx.x * K2_ROWS_PER_BLOCK;
   if (row0 >= q.n_rows) return;
   const uint32_t row1 = min(q.n_rows, row0 + K2_ROWS_PER_BLOCK);
   const uint32_t w0 = blockIdx.z * window, w1 = min(s.length, w0 + window);
   for (uint32_t i = threadIdx.x; i <= window; i += blockDim.x) diff[i] = 0;
   if (threadIdx.x == 0) { n_filtered_s = 0; carry_s = 0; }
   __syncthreads();

   const uint32_t* row_off = s.miss_row_off + s.miss_row_base[partition];
   const uint2* iv = s.miss_iv + s.miss_iv_base[partition];
   uint32_t local_filtered = 0;
   for (uint32_t row = row0 + threadIdx.x; row < row1; row += blockDim.x) {
      if (!((__ldg(q.filter32 + (row >> 5)) >> (row & 31)) & 1u)) continue;
      ++local_filtered;
      const uint32_t b = __ldg(row_off + row), e = __ldg(row_off + row + 1);
      for (uint32_t k = b; k < e; ++k) {
         const uint2 r = __ldg(iv + k);
         const uint32_t a = max(r.x, w0), z = min(r.y, w1);
         if (a < z) { atomicAdd(&diff[a - w0], 1); atomicAdd(&diff[z - w0], -1); }
      }
   }
   local_filtered = uint32_t(warpSum(int(local_filtered)));
   if (laneId() == 0 && local_filtered) atomicAdd(&n_filtered_s, local_filtered);
   __syncthreads();
   const uint32_t n_filtered = n_filtered_s;
   if (n_filtered == 0) return;

   const int8_t* deleted = s.deleted + size_t(partition) * s.length;
   const int8_t* flipped = s.flipped + size_t(partition) * s.length;
   const uint32_t warp = threadIdx.x >> 5, lane = laneId();
   for (uint32_t base = 0; base < w1 - w0; base += blockDim.x) {   // blocked inclusive scan of diff
      const uint32_t i = base + threadIdx.x;
      int v = i < w1 - w0 ? diff[i] : 0;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) { const int t = __shfl_up_sync(0xFFFFFFFFu, v, d); if (lane >= uint32_t(d)) v += t; }
      if (lane == 31) warp_sums[warp] = v;
      __syncthreads();
      if (warp == 0) {
         int ws = warp_sums[lane];
#pragma unroll
         for (int d = 1; d < 32; d <<= 1) { const int t = __shfl_up_sync(0xFFFFFFFFu, ws, d); if (lane >= uint32_t(d)) ws += t; }
         warp_sums[lane] = ws;
      }
      __syncthreads();
      const int missing = v + (warp ? warp_sums[warp - 1] : 0) + carry_s;
      if (i < w1 - w0) {
         const uint32_t pos = w0 + i;
         const int d = deleted[pos], f = flipped[pos];
         if (d >= 0) {
            uint32_t add = n_filtered - uint32_t(missing);          // |F| - #{i in F : pos in M_i}
            if (f >= 0 && !full_mode) add -= n_filtered;              // deleted -= |F| part of the flipped symbol's count
            if (add) atomicAdd(counts + size_t(d) * s.length + pos, add);
         }
         if (f >= 0) atomicAdd(counts + size_t(f) * s.length + pos, n_filtered);
      }
      __syncthreads();
      if (threadIdx.x == 0) carry_s += warp_sums[31];
      __syncthreads();
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
