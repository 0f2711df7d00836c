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
};

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
      for (uint32_t i = lane; i < size; i += 32) setBit(dst, n_words, base | __ldg(pay + i));
   } else if (kind == KIND_RUN) {
      for (uint32_t r = 0; r < size; ++r) {
         const uint32_t start = __ldg(pay + 2 * r), len1 = __ldg(pay + 2 * r + 1);
         setRangeWarp(dst, n_words, base + start, base + start + len1, lane);
      }
   } else {
      // a bitmap has one container per key and every leaf owns its scratch region: plain 128-bit copy
      if ((key + 1u) * 1024u <= n_words) {
         uint4* out = reinterpret_cast<uint4*>(dst + size_t(key) * 1024u);
#pragma unroll 4
         for (uint32_t i = lane; i < 512; i += 32) out[i] = __ldg(bitset + i);
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
   {  // postings: lower bound of slot2 << 18, then the run of equal slots, lanes strided
      const uint32_t target = slot2 << POSTING_ROW_BITS;
      uint32_t lo = seg.post_start, hi = next.post_start;
      while (lo < hi) { const uint32_t mid = (lo + hi) >> 1; if (__ldg(s.postings + mid) < target) lo = mid + 1; else hi = mid; }
      for (uint32_t e = lo + lane; e < next.post_start; e += 32) {
         const uint32_t p = __ldg(s.postings + e);
         if ((p >> POSTING_ROW_BITS) != slot2) break;
         setBit(dst, job.dst_words, (key0 << 16) + (p & POSTING_ROW_MASK));
      }
   }
   {  // large containers of the slot
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

// thread per row: does the row's missing-symbol bitmap contain `value`?  grid.y = job
__global__ void leaf_bitmap_selection(const SeqDev* __restrict__ seqs, const BselJob* __restrict__ jobs, uint64_t* __restrict__ scratch) {
   const BselJob job = jobs[blockIdx.y];
   const uint32_t row = blockIdx.x * blockDim.x + threadIdx.x;
   const SeqDev& s = seqs[job.seq];
   bool hit = false;
   if (row < job.n_rows) {
      const uint32_t* row_off = s.miss_row_off + s.miss_row_base[job.partition];
      const uint2* iv = s.miss_iv + s.miss_iv_base[job.partition];
      uint32_t lo = __ldg(row_off + row), hi = __ldg(row_off + row + 1);
      bool contains = false;
      while (lo < hi) {   // intervals are sorted and disjoint
         const uint32_t mid = (lo + hi) >> 1;
         const uint2 r = __ldg(iv + mid);
         if (job.value < r.x) hi = mid; else if (job.value >= r.y) lo = mid + 1; else { contains = true; break; }
      }
      hit = contains == (job.contains != 0);
   }
   const uint32_t bits = __ballot_sync(0xFFFFFFFFu, hit);
   if (laneId() == 0 && bits) reinterpret_cast<uint32_t*>(scratch + job.dst_off)[row >> 5] = bits;
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

// grid: (ceil(max n_words_padded / 256), n_plans); each thread owns one 64-bit word of one plan
__global__ void __launch_bounds__(INTERP_THREADS)
interpret_plans(const PlanDesc* __restrict__ plans, const uint32_t* __restrict__ programs, const uint64_t* __restrict__ scratch,
                uint64_t* __restrict__ out, uint32_t* __restrict__ ranks) {
   const PlanDesc plan = plans[blockIdx.y];
   const uint32_t w = blockIdx.x * INTERP_THREADS + threadIdx.x;
   if (blockIdx.x * INTERP_THREADS >= plan.n_words_padded) return;
   uint64_t result = 0;
   if (w < plan.n_words) {
      uint64_t stack[INTERP_STACK];
      int sp = 0;
      const uint32_t* pc = programs + plan.prog_off;
      const uint32_t* end = pc + plan.prog_len;
      while (pc < end) {
         const uint32_t op = *pc++;
         switch (op) {
            case P_ZERO: stack[sp++] = 0; break;
            case P_RANGES: {
               const uint32_t n = *pc++;
               uint64_t m = 0;
               for (uint32_t k = 0; k < n; ++k) { m |= rangeMask(w, pc[0], pc[1]); pc += 2; }
               stack[sp++] = m;
               break;
            }
            case P_LEAF: {
               const uint64_t off = uint64_t(pc[0]) | (uint64_t(pc[1]) << 32);
               const uint32_t words = pc[2];
               pc += 3;
               stack[sp++] = w < words ? __ldg(scratch + off + w) : 0ULL;
               break;
            }
            case P_NOT: { const uint32_t rc = *pc++; stack[sp - 1] ^= rangeMask(w, 0, rc); break; }   // flip(0, row_count)
            case P_AND: --sp; stack[sp - 1] &= stack[sp]; break;
            case P_OR: --sp; stack[sp - 1] |= stack[sp]; break;
            case P_ANDNOT: --sp; stack[sp - 1] &= ~stack[sp]; break;
            case P_THR_INIT: { const uint32_t d = *pc++; for (uint32_t j = 0; j < d; ++j) stack[sp++] = 0; break; }
            case P_THR_ADD: {   // threshold.cpp:92-129, one child; T[j] = rows matching >= j+1 children so far
               const int d = int(pc[0]), i = int(pc[1]), n = int(pc[2]), k = int(pc[3]);
               const uint32_t negated = pc[4], rc = pc[5];
               pc += 6;
               const uint64_t c = stack[--sp];
               uint64_t* t = stack + sp - d;
               const int lo = max(0, n - k + i - 1);
               for (int j = min(d - 1, i); j > lo; --j) t[j] |= negated ? (t[j - 1] & ~c) : (t[j - 1] & c);
               if (k - i > n - 1) t[0] |= negated ? (c ^ rangeMask(w, 0, rc)) : c;
               break;
            }
            case P_THR_FIN: {
               const int d = int(pc[0]), n = int(pc[1]);
               const uint32_t exact = pc[2];
               pc += 3;
               const uint64_t* t = stack + sp - d;
               const uint64_t r = exact ? (t[n - 1] & ~t[n]) : t[d - 1];
               sp -= d;
               stack[sp++] = r;
               break;
            }
            default: pc = end; break;
         }
      }
      result = stack[0];
   }
   out[plan.out_off + w] = result;
   // rank tables: exclusive popcount prefix inside this 256-word block, and the block total
   __shared__ int warp_sums[INTERP_THREADS / 32];
   const int pc_w = __popcll(result);
   const uint32_t lane = laneId(), warp = threadIdx.x >> 5;
   int incl = pc_w;
#pragma unroll
   for (int d = 1; d < 32; d <<= 1) { const int t = __shfl_up_sync(0xFFFFFFFFu, incl, d); if (lane >= uint32_t(d)) incl += t; }
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
