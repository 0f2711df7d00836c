// Filter kernels (sm_100a). A compiled Operator tree is evaluated in two steps:
//  (1) every leaf that is not a pure range is densified into a zeroed scratch bitset:
//        leaf_position_bitmap  IndexScan of a stored (position, symbol) bitmap   index_scan.cpp:28-30
//        leaf_registered       IndexScan of a registered index bitmap / uploaded BitmapProducer result
//        leaf_bitmap_selection BitmapSelection over missing_symbol_bitmaps       bitmap_selection.cpp:33-52
//        leaf_selection        CompareToValueSelection predicates                selection.cpp:92-109,144-164
//  (2) interpret_plans runs the postfix program word-by-word (64 rows per thread) with a register /
//      local-memory stack: Intersection (intersection.cpp:76-127), Union (union.cpp:35-45), Complement
//      (complement.cpp:50-54), Threshold (threshold.cpp:64-138) as a thermometer-coded bit-sliced
//      counter with the reference's band pruning, Full / Empty / RangeSelection as computed masks.
//      It also emits the per-word rank tables the run-container path of the Mutations scan needs.
#pragma once
#include "common.cuh"

namespace sb {

enum : uint32_t {
   P_ZERO = 0, P_RANGES = 1, P_LEAF = 2, P_NOT = 3, P_AND = 4, P_OR = 5, P_ANDNOT = 6,
   P_THR_INIT = 7, P_THR_ADD = 8, P_THR_FIN = 9, P_END = 10
};
constexpr int INTERP_THREADS = 256;   // == words per rank block
constexpr int INTERP_STACK = 48;

struct PlanDesc {
   uint32_t prog_off, prog_len;   // into the program stream
   uint32_t n_words;              // 64-bit words that can hold a set bit (max(row_count, leaf universe))
   uint32_t n_words_padded;       // multiple of 4096 (whole 4-key slabs of 262,144 rows), zero beyond n_words
   uint64_t out_off;              // 64-bit word offset of the result in the result slab
   uint64_t rank_local_off;       // uint32 offset
   uint64_t rank_block_off;       // uint32 offset
   uint64_t* cardinality;         // device counter
   uint32_t leaf_off, n_leaves;   // the plan's slice of the leaf table (P_LEAF operands index it)
};
struct LeafRef { uint64_t off; uint32_t words; uint32_t pad; };   // dense scratch bitset of a leaf: word offset, length

struct PosLeafJob { uint32_t seq, partition, slot, slab; uint64_t dst_off; uint32_t dst_words, pad2; };   // one per (leaf, slab)
struct RegContainer { uint64_t pay_off; uint32_t size; uint16_t key; uint8_t kind; uint8_t pad; };  // size: array card / #runs
struct RegLeafJob { uint32_t container; uint32_t dst_words; uint64_t dst_off; };
struct BselJob { uint32_t seq, partition, value, contains, n_rows, pad; uint64_t dst_off; };
struct PredDev { const void* column; int32_t dtype, comparator; int32_t i32; uint32_t u32; double f64; };
struct SelJob { uint32_t pred_begin, n_preds, n_rows, pad; uint64_t dst_off; };

__device__ __forceinline__ void setBit(uint64_t* dst, uint32_t n_words, uint32_t bit) {
   if ((bit >> 6) < n_words) atomicOr(reinterpret_cast<unsigned long long*>(dst + (bit >> 6)), 1ULL << (bit & 63));
}
// sets [a, b] (inclusive) with word-granular atomics; the warp's lanes split the words
__device__ __forceinline__ void setRangeWarp(uint64_t* dst, uint32_t n_words, uint32_t a, uint32_t b, uint32_t lane) {
   const uint32_t wa = a >> 6, wb = b >> 6;
   for (uint32_t w = wa + lane; w <= wb; w += 32) {
      if (w >= n_words) break;
      uint64_t mask = ~0ULL;
      if (w == wa) mask &= ~0ULL << (a & 63);
      if (w == wb) mask &= ~0ULL >> (63 - (b & 63));
      atomicOr(reinterpret_cast<unsigned long long*>(dst + w), mask);
   }
}

__device__ __forceinline__ void scatterContainer(uint64_t* dst, uint32_t n_words, uint32_t key, uint32_t kind, uint32_t size,
                                                 const uint16_t* pay, const uint4* bitset, uint32_t lane) {
   const uint32_t base = key << 16;
   if (kind == KIND_ARRAY) {
      if ((reinterpret_cast<uintptr_t>(pay) & 15u) == 0) {   // 8 values per load, 4 loads per lane in flight
         const uint4* pay4 = reinterpret_cast<const uint4*>(pay);
         const uint32_t n4 = (size + 7) / 8;
         for (uint32_t i0 = lane; i0 < n4; i0 += 128) {
            uint4 x[4];
#pragma unroll
            for (uint32_t k = 0; k < 4; ++k) x[k] = i0 + k * 32 < n4 ? __ldg(pay4 + i0 + k * 32) : make_uint4(0, 0, 0, 0);
#pragma unroll
            for (uint32_t k = 0; k < 4; ++k) {
               const uint32_t wv[4] = {x[k].x, x[k].y, x[k].z, x[k].w};
#pragma unroll
               for (uint32_t t = 0; t < 8; ++t) {
                  if ((i0 + k * 32) * 8 + t < size) setBit(dst, n_words, base | ((wv[t >> 1] >> ((t & 1) * 16)) & 0xFFFFu));
               }
            }
         }
      } else {
         for (uint32_t i = lane; i < size; i += 32) setBit(dst, n_words, base | __ldg(pay + i));
      }
   } else if (kind == KIND_RUN) {
      for (uint32_t r = 0; r < size; ++r) {
         const uint32_t start = __ldg(pay + 2 * r), len1 = __ldg(pay + 2 * r + 1);
         setRangeWarp(dst, n_words, base + start, base + start + len1, lane);
      }
   } else {
      // a bitmap has one container per key and every leaf owns its scratch region: plain 128-bit copy
      if ((key + 1u) * 1024u <= n_words) {
         uint4* out = reinterpret_cast<uint4*>(dst + size_t(key) * 1024u);
#pragma unroll
         for (uint32_t half = 0; half < 2; ++half) {
            uint4 x[8];
#pragma unroll
            for (uint32_t t = 0; t < 8; ++t) x[t] = __ldg(bitset + (half * 8 + t) * 32 + lane);
#pragma unroll
            for (uint32_t t = 0; t < 8; ++t) out[(half * 8 + t) * 32 + lane] = x[t];
         }
      } else {
         const uint64_t* src = reinterpret_cast<const uint64_t*>(bitset);
         for (uint32_t i = lane; i < 1024; i += 32) {
            const uint32_t w = key * 1024u + i;
            if (w < n_words) dst[w] = __ldg(src + i);
         }
      }
   }
}

// one warp per (leaf, slab): finds the postings and large containers of slot' in the sorted segment and scatters them
__global__ void leaf_position_bitmap(const SeqDev* __restrict__ seqs, const PosLeafJob* __restrict__ jobs, uint32_t n_jobs, uint64_t* __restrict__ scratch) {
   const uint32_t job_id = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = laneId();
   if (job_id >= n_jobs) return;
   const PosLeafJob job = jobs[job_id];
   const SeqDev& s = seqs[job.seq];
   const uint32_t pos = job.slot / s.sigma, sym = job.slot - pos * s.sigma;
   const uint32_t tile = pos / s.tile_positions, slot2 = sym * s.tile_positions + (pos - tile * s.tile_positions);
   uint64_t* dst = scratch + job.dst_off;
   const SegRec seg = s.segs[size_t(tile) * s.n_slabs + job.slab];
   const SegRec next = s.segs[size_t(tile) * s.n_slabs + job.slab + 1];
   const uint32_t key0 = s.slabs[job.slab].key0;
   {  // postings: warp-wide 32-ary lower bound of slot2 << 18 (3 dependent loads for a 32 K segment), then the run of equal slots
      const uint32_t target = slot2 << POSTING_ROW_BITS;
      uint32_t lo = seg.post_start, hi = next.post_start;   // answer in [lo, hi]
      while (hi - lo > 32) {
         const uint32_t step = (hi - lo + 31) / 32;           // probes at lo + k*step - 1 ... cover the range
         const uint32_t at = min(lo + (lane + 1) * step, hi) - 1;
         const bool below = __ldg(s.postings + at) < target;   // monotone in lane
         const uint32_t n_below = __popc(__ballot_sync(0xFFFFFFFFu, below));
         const uint32_t new_lo = n_below == 0 ? lo : min(lo + n_below * step, hi);
         const uint32_t new_hi = n_below == 32 ? hi : min(lo + (n_below + 1) * step, hi) - 1;
         lo = new_lo; hi = max(new_hi, new_lo);
      }
      {
         const bool below = lo + lane < hi && __ldg(s.postings + lo + lane) < target;
         lo += __popc(__ballot_sync(0xFFFFFFFFu, below));
      }
      for (uint32_t e = lo + lane; e < next.post_start; e += 32) {
         const uint32_t p = __ldg(s.postings + e);
         if ((p >> POSTING_ROW_BITS) != slot2) break;
         setBit(dst, job.dst_words, (key0 << 16) + (p & POSTING_ROW_MASK));
      }
   }
   {  // large containers of the slot (few per segment: plain binary search)
      uint32_t lo = seg.large_start, hi = next.large_start;
      while (lo < hi) { const uint32_t mid = (lo + hi) >> 1; if (s.large_entries[mid].slot() < slot2) lo = mid + 1; else hi = mid; }
      for (uint32_t e = lo; e < next.large_start; ++e) {
         const LargeEntry w = s.large_entries[e];
         if (w.slot() != slot2) break;
         const uint4* pay = s.large_pool + w.offset16;
         scatterContainer(dst, job.dst_words, key0 + w.keyLocal(), w.kind(), w.size, reinterpret_cast<const uint16_t*>(pay), pay, lane);
      }
   }
}

// one warp per container of a registered / uploaded bitmap
__global__ void leaf_registered(const RegContainer* __restrict__ containers, const uint16_t* __restrict__ pool,
                                const RegLeafJob* __restrict__ jobs, uint32_t n_jobs, uint64_t* __restrict__ scratch) {
   const uint32_t job_id = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = laneId();
   if (job_id >= n_jobs) return;
   const RegLeafJob job = jobs[job_id];
   const RegContainer c = containers[job.container];
   const uint16_t* pay = pool + c.pay_off;
   scatterContainer(scratch + job.dst_off, job.dst_words, c.key, c.kind, c.size, pay, reinterpret_cast<const uint4*>(pay), lane);
}

// Builds the coarse index of the missing-symbol intervals of one partition (load time): bit (tile, row) is set when
// the row misses some position of the tile. Thread per row.
__global__ void build_miss_tiles(const uint32_t* __restrict__ row_off, const uint2* __restrict__ iv, uint32_t n_rows, uint32_t tile_positions,
                                 uint32_t* __restrict__ bits, uint32_t words_per_tile) {
   const uint32_t row = blockIdx.x * blockDim.x + threadIdx.x;
   if (row >= n_rows) return;
   for (uint32_t j = row_off[row]; j < row_off[row + 1]; ++j) {
      const uint2 r = iv[j];
      if (r.x >= r.y) continue;
      for (uint32_t t = r.x / tile_positions; t <= (r.y - 1) / tile_positions; ++t) atomicOr(bits + size_t(t) * words_per_tile + (row >> 5), 1u << (row & 31));
   }
}

// does the row's missing-symbol bitmap contain `value`?  grid.y = job. A warp owns 1,024 consecutive rows (32 words of the
// result). The coarse index (build_miss_tiles) names the rows that miss anything in value's tile -- a few percent; the
// warp writes them into a list in shared memory and its lanes share them evenly, four at a time per lane, so that the
// dependent loads of the interval search (sorted, disjoint intervals: binary search) are few rounds with many loads in flight.
constexpr uint32_t BSEL_ROWS = 32;     // rows per thread
__global__ void __launch_bounds__(256)
leaf_bitmap_selection(const SeqDev* __restrict__ seqs, const BselJob* __restrict__ jobs, uint64_t* __restrict__ scratch) {
   __shared__ uint32_t cand_rows[8][1024];
   __shared__ uint32_t found_words[8][32];
   const BselJob job = jobs[blockIdx.y];
   const uint32_t lane = laneId(), warp = threadIdx.x >> 5;
   const uint32_t word = blockIdx.x * 256 + threadIdx.x, n_words = (job.n_rows + 31) / 32;
   if (blockIdx.x * 256 + warp * 32 >= n_words) return;        // whole warp beyond the partition
   const SeqDev& s = seqs[job.seq];
   const uint32_t tile = job.value / s.tile_positions;
   uint32_t candidates = 0;
   if (word < n_words && tile < s.n_tiles) candidates = __ldg(s.miss_tile_bits + s.miss_tile_base[job.partition] + size_t(tile) * n_words + word);
   // candidate rows of the warp -> list
   const uint32_t mine = __popc(candidates);
   uint32_t before = mine;
#pragma unroll
   for (int d = 1; d < 32; d <<= 1) { const uint32_t t = __shfl_up_sync(0xFFFFFFFFu, before, d); if (lane >= uint32_t(d)) before += t; }
   const uint32_t total = __shfl_sync(0xFFFFFFFFu, before, 31);
   before -= mine;
   found_words[warp][lane] = 0;
   for (uint32_t c = candidates, k = before; c; c &= c - 1u) cand_rows[warp][k++] = word * 32 + uint32_t(__ffs(int(c))) - 1u;
   __syncwarp();
   if (total) {
      const uint32_t* row_off = s.miss_row_off + s.miss_row_base[job.partition];
      const uint2* iv = s.miss_iv + s.miss_iv_base[job.partition];
      // four candidates per lane at a time, their searches walked in lockstep: four independent chains of dependent loads
      for (uint32_t k0 = lane; k0 < total; k0 += 128) {
         uint32_t row[4], lo[4], hi[4];
#pragma unroll
         for (uint32_t q = 0; q < 4; ++q) {
            lo[q] = hi[q] = 0; row[q] = 0;
            if (k0 + q * 32 < total) { row[q] = cand_rows[warp][k0 + q * 32]; lo[q] = __ldg(row_off + row[q]); hi[q] = __ldg(row_off + row[q] + 1); }
         }
         bool any = true;
         while (any) {
            any = false;
            uint2 r[4];
            uint32_t mid[4];
#pragma unroll
            for (uint32_t q = 0; q < 4; ++q) { mid[q] = (lo[q] + hi[q]) >> 1; if (lo[q] < hi[q]) r[q] = __ldg(iv + mid[q]); }
#pragma unroll
            for (uint32_t q = 0; q < 4; ++q) {
               if (lo[q] < hi[q]) {
                  if (job.value < r[q].x) hi[q] = mid[q];
                  else if (job.value >= r[q].y) lo[q] = mid[q] + 1;
                  else { atomicOr(&found_words[warp][(row[q] >> 5) & 31u], 1u << (row[q] & 31)); hi[q] = lo[q]; }
                  any |= lo[q] < hi[q];
               }
            }
         }
      }
      __syncwarp();
   }
   if (word >= n_words) return;
   const uint32_t found = found_words[warp][lane];
   const uint32_t valid = (word + 1) * 32 <= job.n_rows ? ~0u : (~0u >> (32 - (job.n_rows & 31)));
   const uint32_t bits = (job.contains ? found : ~found) & valid;
   if (bits) reinterpret_cast<uint32_t*>(scratch + job.dst_off)[word] = bits;
}

template <typename T>
__device__ __forceinline__ bool compareValues(int32_t comparator, T a, T b) {
   switch (comparator) {
      case SB_EQUALS: return a == b;
      case SB_NOT_EQUALS: return a != b;
      case SB_LESS: return a < b;
      case SB_HIGHER_OR_EQUALS: return a >= b;
      case SB_HIGHER: return a > b;
      default: return a <= b;
   }
}

// thread per row: conjunction of column predicates.  grid.y = job
__global__ void leaf_selection(const SelJob* __restrict__ jobs, const PredDev* __restrict__ preds, uint64_t* __restrict__ scratch) {
   const SelJob job = jobs[blockIdx.y];
   const uint32_t row = blockIdx.x * blockDim.x + threadIdx.x;
   bool hit = row < job.n_rows;
   for (uint32_t k = 0; hit && k < job.n_preds; ++k) {
      const PredDev p = preds[job.pred_begin + k];
      if (p.dtype == 0) hit = compareValues<int32_t>(p.comparator, __ldg(static_cast<const int32_t*>(p.column) + row), p.i32);
      else if (p.dtype == 1) hit = compareValues<double>(p.comparator, __ldg(static_cast<const double*>(p.column) + row), p.f64);
      else hit = compareValues<uint32_t>(p.comparator, __ldg(static_cast<const uint32_t*>(p.column) + row), p.u32);
   }
   const uint32_t bits = __ballot_sync(0xFFFFFFFFu, hit);
   if (laneId() == 0 && bits) reinterpret_cast<uint32_t*>(scratch + job.dst_off)[row >> 5] = bits;
}

// mask of the bits of word `w` that lie in [a, b)
__device__ __forceinline__ uint64_t rangeMask(uint32_t w, uint32_t a, uint32_t b) {
   const uint64_t lo = uint64_t(w) << 6, hi = lo + 64;
   if (b <= lo || a >= hi || a >= b) return 0;
   uint64_t m = ~0ULL;
   if (a > lo) m &= ~0ULL << (a - lo);
   if (b < hi) m &= ~0ULL >> (hi - b);
   return m;
}

// grid: (ceil(max n_words_padded / 256), n_plans); each thread owns one 64-bit word of one plan.
// The interpreter is a serial fetch-decode chain per thread, so memory latency is taken out of it: the program and
// the plan's leaf table are staged in shared memory, and the thread's words of the leaves -- consumed strictly in
// program order -- are fetched eight at a time into a register window (eight independent loads in flight).
// The top of the evaluation stack lives in a register; deeper entries in local memory.
constexpr uint32_t INTERP_PROG_SMEM = 2048;   // program words staged in shared memory (longer programs are read from global)
constexpr uint32_t INTERP_LEAF_SMEM = 128;    // leaf table entries staged in shared memory
__global__ void __launch_bounds__(INTERP_THREADS)
interpret_plans(const PlanDesc* __restrict__ plans, const uint32_t* __restrict__ programs, const LeafRef* __restrict__ leaf_table,
                const uint64_t* __restrict__ scratch, uint64_t* __restrict__ out, uint32_t* __restrict__ ranks) {
   __shared__ uint32_t prog_s[INTERP_PROG_SMEM];
   __shared__ LeafRef leaf_s[INTERP_LEAF_SMEM];
   __shared__ int warp_sums[INTERP_THREADS / 32];
   const PlanDesc plan = plans[blockIdx.y];
   const uint32_t w = blockIdx.x * INTERP_THREADS + threadIdx.x;
   if (blockIdx.x * INTERP_THREADS >= plan.n_words_padded) return;
   uint64_t result = 0;
   if (blockIdx.x * INTERP_THREADS < plan.n_words) {
      const uint32_t* prog = programs + plan.prog_off;
      const LeafRef* leaves = leaf_table + plan.leaf_off;
      if (plan.prog_len <= INTERP_PROG_SMEM) {
         for (uint32_t i = threadIdx.x; i < plan.prog_len; i += INTERP_THREADS) prog_s[i] = prog[i];
         prog = prog_s;
      }
      if (plan.n_leaves <= INTERP_LEAF_SMEM) {
         for (uint32_t i = threadIdx.x; i < plan.n_leaves; i += INTERP_THREADS) leaf_s[i] = leaves[i];
         leaves = leaf_s;
      }
      __syncthreads();
      uint64_t win[8];      // words of leaves [8 * batch, 8 * batch + 8)
#pragma unroll
      for (int k = 0; k < 8; ++k) win[k] = 0;
      uint32_t next_leaf = 0;
      uint64_t stack[INTERP_STACK];
      uint64_t tos = 0;   // top of stack; stack[0 .. sp-2] below it
      int sp = 0;
      const uint32_t* pc = prog;
      const uint32_t* end = pc + plan.prog_len;
#define SB_PUSH(value) do { if (sp > 0) stack[sp - 1] = tos; tos = (value); ++sp; } while (0)
      while (pc < end) {
         const uint32_t op = *pc++;
         switch (op) {
            case P_ZERO: SB_PUSH(0ULL); break;
            case P_RANGES: {
               const uint32_t n = *pc++;
               uint64_t m = 0;
               for (uint32_t k = 0; k < n; ++k) { m |= rangeMask(w, pc[0], pc[1]); pc += 2; }
               SB_PUSH(m);
               break;
            }
            case P_LEAF: {   // operand = index in the leaf table = number of leaves consumed so far
               ++pc;
               const uint32_t k = next_leaf & 7u;
               if (k == 0) {
#pragma unroll
                  for (uint32_t q = 0; q < 8; ++q) {
                     win[q] = 0;
                     if (next_leaf + q < plan.n_leaves) { const LeafRef r = leaves[next_leaf + q]; if (w < r.words) win[q] = __ldg(scratch + r.off + w); }
                  }
               }
               uint64_t v;
               switch (k) {
                  case 0: v = win[0]; break; case 1: v = win[1]; break; case 2: v = win[2]; break; case 3: v = win[3]; break;
                  case 4: v = win[4]; break; case 5: v = win[5]; break; case 6: v = win[6]; break; default: v = win[7]; break;
               }
               ++next_leaf;
               SB_PUSH(v);
               break;
            }
            case P_NOT: { const uint32_t rc = *pc++; tos ^= rangeMask(w, 0, rc); break; }   // flip(0, row_count)
            case P_AND: --sp; tos = stack[sp - 1] & tos; break;
            case P_OR: --sp; tos = stack[sp - 1] | tos; break;
            case P_ANDNOT: --sp; tos = stack[sp - 1] & ~tos; break;
            case P_THR_INIT: { const uint32_t d = *pc++; for (uint32_t j = 0; j < d; ++j) SB_PUSH(0ULL); break; }
            case P_THR_ADD: {   // threshold.cpp:92-129, one child; T[j] = rows matching >= j+1 children so far
               const int d = int(pc[0]), i = int(pc[1]), n = int(pc[2]), k = int(pc[3]);
               const uint32_t negated = pc[4], rc = pc[5];
               pc += 6;
               const uint64_t c = tos;
               --sp;
               uint64_t* t = stack + sp - d;     // the d planes are all in memory while a child is evaluated
               const int lo = max(0, n - k + i - 1);
               for (int j = min(d - 1, i); j > lo; --j) t[j] |= negated ? (t[j - 1] & ~c) : (t[j - 1] & c);
               if (k - i > n - 1) t[0] |= negated ? (c ^ rangeMask(w, 0, rc)) : c;
               tos = t[d - 1];
               break;
            }
            case P_THR_FIN: {
               const int d = int(pc[0]), n = int(pc[1]);
               const uint32_t exact = pc[2];
               pc += 3;
               stack[sp - 1] = tos;
               const uint64_t* t = stack + sp - d;
               const uint64_t r = exact ? (t[n - 1] & ~t[n]) : t[d - 1];
               sp -= d;
               ++sp;
               tos = r;
               break;
            }
            default: pc = end; break;
         }
      }
#undef SB_PUSH
      if (w < plan.n_words) result = tos;
   }
   out[plan.out_off + w] = result;
   // rank tables: exclusive popcount prefix inside this 256-word block, and the block total
   const int pc_w = __popcll(result);
   const uint32_t lane = laneId(), warp = threadIdx.x >> 5;
   int incl = pc_w;
#pragma unroll
   for (int d = 1; d < 32; d <<= 1) { const int t = __shfl_up_sync(0xFFFFFFFFu, incl, d); if (lane >= uint32_t(d)) incl += t; }
   __syncthreads();
   if (lane == 31) warp_sums[warp] = incl;
   __syncthreads();
   int before = 0, total = 0;
#pragma unroll
   for (int k = 0; k < INTERP_THREADS / 32; ++k) { const int v = warp_sums[k]; if (k < int(warp)) before += v; total += v; }
   ranks[plan.rank_local_off + w] = uint32_t(before + incl - pc_w);
   if (threadIdx.x == 0) {
      ranks[plan.rank_block_off + blockIdx.x] = uint32_t(total);
      if (total) atomicAdd(reinterpret_cast<unsigned long long*>(plan.cardinality), static_cast<unsigned long long>(total));
   }
}

// one warp per plan: block totals -> exclusive prefix
__global__ void rank_block_prefix(const PlanDesc* __restrict__ plans, uint32_t n_plans, uint32_t* __restrict__ ranks) {
   const uint32_t plan_id = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = laneId();
   if (plan_id >= n_plans) return;
   const PlanDesc plan = plans[plan_id];
   uint32_t* blocks = ranks + plan.rank_block_off;
   const uint32_t n_blocks = plan.n_words_padded / INTERP_THREADS;
   uint32_t carry = 0;
   for (uint32_t base = 0; base < n_blocks; base += 32) {
      const uint32_t i = base + lane;
      const uint32_t v = i < n_blocks ? blocks[i] : 0;
      uint32_t incl = v;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) { const uint32_t t = __shfl_up_sync(0xFFFFFFFFu, incl, d); if (lane >= uint32_t(d)) incl += t; }
      if (i < n_blocks) blocks[i] = carry + incl - v;
      carry += __shfl_sync(0xFFFFFFFFu, incl, 31);
   }
}

}  // namespace sb
