// Index build kernels (sm_100a): SequenceStorePartition<S>::fill + optimizeBitmaps on the device.
//   The reference (src/silo/storage/sequence_store.cpp:101-211, position.cpp:24-127) walks the aligned sequences of a partition
//   position by position, adds every row to the bitmap of its symbol, collects the positions of the missing symbol (N / X)
//   per row, and then drops the bitmap of the most numerous symbol of each position. Here, from the row-major characters
//   genomes[n_rows][length] in device memory:
//   ib_histogram      per (position, symbol) the number of rows          -> the symbol to delete, the size of every bitmap
//   ib_pick           most numerous symbol in the alphabet's iteration order (the first among equals), bitmap sizes
//   ib_scan           exclusive prefix sum (one CTA)                     -> where each bitmap's row ids start
//   ib_emit           row ids of every kept (position, symbol), ascending: a CTA owns 32 positions and walks the rows in
//                     order, a warp appends the rows of a 32-row group with one match / popc per distinct symbol
//   ib_missing_*      per row the maximal runs of the missing symbol as [start, end) intervals (count, scan, emit)
// Everything is integer work; the characters are read three times (histogram, ids, missing runs), 1 byte per lane.
#pragma once
#include "common.cuh"

namespace sb {

constexpr int IB_THREADS = 256;
constexpr uint32_t IB_ROWS_PER_BLOCK = 2048;      // rows a histogram CTA walks
constexpr uint32_t IB_MAX_SIGMA = 32;

struct BuildAlphabet {
   int8_t lut[256];                 // character -> symbol, -1 = illegal
   uint8_t order[IB_MAX_SIGMA];     // iteration order of the symbols (SymbolType::SYMBOLS)
   uint32_t sigma, missing;
};

__device__ __forceinline__ uint32_t ibLane() { return threadIdx.x & 31; }

// grid: (position tiles of 256, row blocks). Thread t owns position tile * 256 + t: no atomics inside the CTA.
__global__ void __launch_bounds__(IB_THREADS)
ib_histogram(const uint8_t* __restrict__ genomes, const uint8_t* __restrict__ present, uint32_t n_rows, uint32_t length, const BuildAlphabet alphabet,
             uint32_t* __restrict__ hist, unsigned long long* __restrict__ first_illegal) {
   extern __shared__ uint32_t counts_s[];      // [256][sigma]
   __shared__ int8_t lut_s[256];
   lut_s[threadIdx.x] = alphabet.lut[threadIdx.x];
   for (uint32_t i = threadIdx.x; i < IB_THREADS * alphabet.sigma; i += IB_THREADS) counts_s[i] = 0;
   __syncthreads();
   const uint32_t pos = blockIdx.x * IB_THREADS + threadIdx.x;
   const uint32_t row0 = blockIdx.y * IB_ROWS_PER_BLOCK, row1 = min(n_rows, row0 + IB_ROWS_PER_BLOCK);
   if (pos < length) {
      uint32_t* mine = counts_s + threadIdx.x * alphabet.sigma;
      for (uint32_t row = row0; row < row1; ++row) {
         if (!present[row]) continue;      // uniform over the CTA
         const int sym = lut_s[genomes[size_t(row) * length + pos]];
         if (sym < 0) atomicMin(first_illegal, (unsigned long long)(row) * length + pos);
         else mine[sym]++;
      }
   }
   __syncthreads();
   for (uint32_t i = threadIdx.x; i < IB_THREADS * alphabet.sigma; i += IB_THREADS) {
      const uint32_t p = blockIdx.x * IB_THREADS + i / alphabet.sigma;
      if (p < length && counts_s[i]) atomicAdd(hist + size_t(p) * alphabet.sigma + i % alphabet.sigma, counts_s[i]);
   }
}

// thread per position: deleted[pos] = the most numerous symbol (sequence_store.cpp:193-211 via position.cpp:103-127: strictly
// more rows than every symbol before it in iteration order; -1 when no row carries a symbol), sizes[slot] = rows of the bitmap
// that is kept (0 for the deleted and the missing symbol)
__global__ void ib_pick(const uint32_t* __restrict__ hist, uint32_t length, const BuildAlphabet alphabet, int8_t* __restrict__ deleted, unsigned long long* __restrict__ sizes) {
   const uint32_t pos = blockIdx.x * blockDim.x + threadIdx.x;
   if (pos >= length) return;
   int best = -1;
   uint32_t best_count = 0;
   for (uint32_t k = 0; k < alphabet.sigma; ++k) {
      const uint32_t sym = alphabet.order[k], c = hist[size_t(pos) * alphabet.sigma + sym];
      if (sym != alphabet.missing && c > best_count) { best = int(sym); best_count = c; }
   }
   deleted[pos] = int8_t(best);
   for (uint32_t sym = 0; sym < alphabet.sigma; ++sym)
      sizes[size_t(pos) * alphabet.sigma + sym] = (int(sym) == best || sym == alphabet.missing) ? 0ull : hist[size_t(pos) * alphabet.sigma + sym];
}

// exclusive prefix sum of n values in place, out[n] = total. One CTA of 1,024 threads (n is at most a few million).
__global__ void __launch_bounds__(1024)
ib_scan(unsigned long long* __restrict__ values, size_t n) {
   __shared__ unsigned long long warp_sums[32];
   __shared__ unsigned long long carry_s;
   if (threadIdx.x == 0) carry_s = 0;
   __syncthreads();
   const uint32_t lane = ibLane(), warp = threadIdx.x >> 5;
   for (size_t base = 0; base < n; base += 1024) {
      const size_t i = base + threadIdx.x;
      const unsigned long long v = i < n ? values[i] : 0ull;
      unsigned long long incl = v;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) { const unsigned long long t = __shfl_up_sync(0xFFFFFFFFu, incl, d); if (lane >= uint32_t(d)) incl += t; }
      if (lane == 31) warp_sums[warp] = incl;
      __syncthreads();
      if (warp == 0) {
         unsigned long long ws = warp_sums[lane];
#pragma unroll
         for (int d = 1; d < 32; d <<= 1) { const unsigned long long t = __shfl_up_sync(0xFFFFFFFFu, ws, d); if (lane >= uint32_t(d)) ws += t; }
         warp_sums[lane] = ws;
      }
      __syncthreads();
      const unsigned long long carry = carry_s, before = warp ? warp_sums[warp - 1] : 0ull;
      if (i < n) values[i] = carry + before + incl - v;
      __syncthreads();
      if (threadIdx.x == 1023) carry_s = carry + before + incl;
      __syncthreads();
   }
   if (threadIdx.x == 0) values[n] = carry_s;
}

// grid: position groups of 32. The CTA walks all rows in chunks of 256: a chunk's [256 rows][32 positions] characters go
// through shared memory (a row's 32 bytes are one sector), then warp w appends the rows of positions 4w .. 4w+3 -- it alone
// writes those positions' bitmaps, so their fill levels are plain shared-memory counters and the ids come out ascending.
__global__ void __launch_bounds__(IB_THREADS)
ib_emit(const uint8_t* __restrict__ genomes, const uint8_t* __restrict__ present, uint32_t n_rows, uint32_t length, const BuildAlphabet alphabet,
        const int8_t* __restrict__ deleted, const unsigned long long* __restrict__ slot_off, uint32_t* __restrict__ ids) {
   __shared__ int8_t lut_s[256];
   __shared__ int8_t tile_s[256][33];                      // symbol of (row of the chunk, position of the group); -1 = nothing to emit
   __shared__ uint32_t fill_s[32][IB_MAX_SIGMA];           // rows written so far per (position of the group, symbol)
   lut_s[threadIdx.x] = alphabet.lut[threadIdx.x];
   for (uint32_t i = threadIdx.x; i < 32 * IB_MAX_SIGMA; i += IB_THREADS) fill_s[i / IB_MAX_SIGMA][i % IB_MAX_SIGMA] = 0;
   const uint32_t lane = ibLane(), warp = threadIdx.x >> 5;
   const uint32_t pos0 = blockIdx.x * 32, my_pos = pos0 + lane;
   const int my_deleted = my_pos < length ? deleted[my_pos] : -1;
   __syncthreads();
   for (uint32_t chunk = 0; chunk < n_rows; chunk += 256) {
      for (uint32_t r = warp; r < 256; r += IB_THREADS / 32) {      // lane = position: one 32-byte read per row
         const uint32_t row = chunk + r;
         int sym = -1;
         if (row < n_rows && my_pos < length && present[row]) {
            sym = lut_s[genomes[size_t(row) * length + my_pos]];
            if (sym == my_deleted || sym == int(alphabet.missing)) sym = -1;
         }
         tile_s[r][lane] = int8_t(sym);
      }
      __syncthreads();
      for (uint32_t q = warp * 4; q < warp * 4 + 4; ++q) {
         const uint32_t pos = pos0 + q;
         if (pos >= length) break;
         for (uint32_t group = 0; group < 256; group += 32) {
            const int sym = tile_s[group + lane][q];
            const uint32_t active = __ballot_sync(0xFFFFFFFFu, sym >= 0);
            if (active == 0) continue;      // the usual case: every row carries the position's most numerous symbol
            if (sym >= 0) {
               const uint32_t peers = __match_any_sync(active, sym);
               const uint32_t leader = uint32_t(__ffs(int(peers))) - 1u, rank = __popc(peers & ((1u << lane) - 1u));
               uint32_t base = 0;
               if (lane == leader) { base = fill_s[q][sym]; fill_s[q][sym] = base + __popc(peers); }
               base = __shfl_sync(peers, base, int(leader));
               ids[slot_off[size_t(pos) * alphabet.sigma + uint32_t(sym)] + base + rank] = chunk + group + lane;
            }
            __syncwarp();
         }
      }
      __syncthreads();
   }
}

// warp per row: maximal runs of the missing symbol. COUNT pass: n_intervals[row]; EMIT pass: intervals at missing_off[row].
template <bool EMIT>
__global__ void __launch_bounds__(IB_THREADS)
ib_missing(const uint8_t* __restrict__ genomes, const uint8_t* __restrict__ present, uint32_t n_rows, uint32_t length, const BuildAlphabet alphabet,
           unsigned long long* __restrict__ counts_or_off, uint2* __restrict__ intervals) {
   __shared__ int8_t lut_s[256];
   lut_s[threadIdx.x] = alphabet.lut[threadIdx.x];
   __syncthreads();
   const uint32_t row = (blockIdx.x * IB_THREADS + threadIdx.x) >> 5, lane = ibLane();
   if (row >= n_rows) return;
   if (!present[row]) {      // a null genome misses everything (sequence_store.cpp:170-176)
      if (lane == 0) { if (EMIT) intervals[counts_or_off[row]] = make_uint2(0, length); else counts_or_off[row] = 1; }
      return;
   }
   const uint8_t* genome = genomes + size_t(row) * length;
   uint32_t n = 0;            // intervals closed so far (EMIT: starts written so far)
   uint32_t n_ends = 0;
   uint32_t carry = 0;        // was the position before this group missing?
   const unsigned long long out = EMIT ? counts_or_off[row] : 0ull;
   for (uint32_t base = 0; base < length; base += 32) {
      const uint32_t pos = base + lane;
      const bool miss = pos < length && lut_s[genome[pos]] == int(alphabet.missing);
      const uint32_t m = __ballot_sync(0xFFFFFFFFu, miss);
      const uint32_t before = (m << 1) | carry;
      const uint32_t starts = m & ~before, ends = ~m & before;      // a run starts where the bit before is clear; it ends at the first clear bit behind it
      if (EMIT) {
         if (starts >> lane & 1u) intervals[out + n + __popc(starts & ((1u << lane) - 1u))].x = pos;
         if (ends >> lane & 1u) intervals[out + n_ends + __popc(ends & ((1u << lane) - 1u))].y = pos;
      }
      n += __popc(starts);
      n_ends += __popc(ends);
      carry = m >> 31;
   }
   if (carry) {               // the last run reaches the end of the sequence
      if (EMIT && lane == 0) intervals[out + n_ends].y = length;
      ++n_ends;
   }
   if (!EMIT && lane == 0) counts_or_off[row] = n;
}

}  // namespace sb
