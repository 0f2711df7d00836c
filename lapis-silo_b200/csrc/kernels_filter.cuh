// Filter kernels (sm_100a). A compiled Operator tree (postfix program) is evaluated per partition in ONE pass over its
// leaves -- no leaf of a roaring bitmap is ever densified into global memory:
//   locate_leaves          where the containers of an IndexScan'ed (position, symbol) bitmap lie in the posting layout
//                          (index_scan.cpp:28-30 borrows the stored bitmap; here: a warp search per (leaf, slab))
//   leaf_bitmap_selection  BitmapSelection over missing_symbol_bitmaps   bitmap_selection.cpp:33-52   (row-wise: dense scratch)
//   leaf_selection         CompareToValueSelection predicates            selection.cpp:92-109,144-164 (row-wise: dense scratch)
//   eval_plans             one CTA per (plan, 65,536-row container key): every thread owns four 64-bit words of the key and
//                          all threads run the same postfix program. A leaf's container of that key is read straight from the
//                          pool -- a bitset with coalesced 64-bit loads, an array / run / posting list through an 8 KB shared-
//                          memory staging bitset -- and combined on an evaluation stack in shared memory (top of stack in
//                          registers): Intersection (intersection.cpp:76-127), Union (union.cpp:35-45), Complement
//                          (complement.cpp:50-54), Threshold (threshold.cpp:64-138) as a thermometer-coded bit-sliced counter
//                          with the reference's band pruning, Full / Empty / RangeSelection as computed masks.
//                          Emits the result bitset and the rank tables the run-container path of the Mutations scan needs.
//   finish_filters         per plan: block totals -> exclusive prefix, cardinality
#pragma once
#include "common.cuh"

namespace sb {

enum : uint32_t {
   P_ZERO = 0, P_RANGES = 1, P_LEAF = 2, P_NOT = 3, P_AND = 4, P_OR = 5, P_ANDNOT = 6,
   P_THR_INIT = 7, P_THR_ADD = 8, P_THR_FIN = 9, P_END = 10
};
constexpr int EVAL_THREADS = 256;
constexpr uint32_t EVAL_WORDS = 1024;                 // 64-bit words per CTA: one container key (65,536 rows)
constexpr int EVAL_PER_THREAD = EVAL_WORDS / EVAL_THREADS;
constexpr uint32_t RANK_BLOCK_WORDS = 256;            // words per rank block
constexpr uint32_t EVAL_MAX_SMEM_LEVELS = 24;         // evaluation stack levels kept in shared memory (8 KB each); deeper trees use global memory
constexpr uint32_t EVAL_MAX_LEVELS = 4096;

struct PlanDesc {
   uint32_t prog_off, prog_len;   // into the program stream
   uint32_t n_words;              // 64-bit words that can hold a set bit (max(row_count, leaf universe))
   uint32_t n_words_padded;       // multiple of 4096 (whole 4-key slabs of 262,144 rows), zero beyond n_words
   uint64_t out_off;              // 64-bit word offset of the result in the result area
   uint64_t rank_local_off;       // uint32 offset
   uint64_t rank_block_off;       // uint32 offset
   uint64_t* cardinality;         // device counter
   uint32_t leaf_off, n_leaves;   // the plan's slice of the leaf table (P_LEAF operands index it)
};

// A leaf of a plan: where its set bits come from.
enum : uint32_t { LEAF_DENSE = 0, LEAF_POS = 1, LEAF_REG = 2, LEAF_UP = 3 };
struct LeafDesc {
   uint32_t kind;
   uint32_t a;       // DENSE: words of the scratch bitset | POS: sequence index | REG / UP: first container record
   uint32_t b;       // POS: slabs of the partition        | REG / UP: number of container records
   uint32_t c;       // POS: slot' (symbol * tile_positions + position in tile)
   uint64_t off;     // DENSE: word offset in the scratch area | POS: index of the leaf's first LeafLoc (one per slab)
};
static_assert(sizeof(LeafDesc) == 24, "LeafDesc must be 24 bytes");
struct LeafLoc { uint32_t post_begin, post_end, large_begin, large_end; };   // of one (leaf, slab): its postings and large entries

struct LocJob { uint32_t seq, tile, slot2, slab; };                          // one per (leaf, slab); result at locs[job index]
struct RegContainer { uint64_t pay_off; uint32_t size; uint16_t key; uint8_t kind; uint8_t pad; };  // size: array card / #runs
struct BselJob { uint32_t seq, partition, value, contains, n_rows, pad; uint64_t dst_off; };
struct PredDev { const void* column; int32_t dtype, comparator; int32_t i32; uint32_t u32; double f64; };
struct SelJob { uint32_t pred_begin, n_preds, n_rows, pad; uint64_t dst_off; };

__device__ __forceinline__ uint32_t laneIdF() { return threadIdx.x & 31; }

// one warp per (leaf, slab): the postings and large entries of slot' in the sorted (tile, slab) segment
__global__ void locate_leaves(const SeqDev* __restrict__ seqs, const LocJob* __restrict__ jobs, uint32_t n_jobs, LeafLoc* __restrict__ locs) {
   const uint32_t job_id = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = laneIdF();
   if (job_id >= n_jobs) return;
   const LocJob job = jobs[job_id];
   const SeqDev& s = seqs[job.seq];
   const SegRec seg = s.segs[size_t(job.tile) * s.n_slabs + job.slab];
   const SegRec next = s.segs[size_t(job.tile) * s.n_slabs + job.slab + 1];
   LeafLoc loc;
   {  // postings: warp-wide 32-ary lower bound of slot' << 18 (3 dependent loads for a 32 K segment), then the run of equal slots
      const uint32_t target = job.slot2 << POSTING_ROW_BITS;
      uint32_t lo = seg.post_start, hi = next.post_start;   // answer in [lo, hi]
      while (hi - lo > 32) {
         const uint32_t step = (hi - lo + 31) / 32;
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
      // the postings of the slot in this slab (at most SLAB_KEYS small containers)
      uint32_t n = 0;
      for (uint32_t base = lo; base < next.post_start; base += 32) {
         const bool same = base + lane < next.post_start && (__ldg(s.postings + base + lane) >> POSTING_ROW_BITS) == job.slot2;
         const uint32_t m = __ballot_sync(0xFFFFFFFFu, same);
         n += __popc(m);
         if (m != 0xFFFFFFFFu) break;
      }
      loc.post_begin = lo; loc.post_end = lo + n;
   }
   {  // large containers of the slot (few per segment: plain binary search)
      uint32_t lo = seg.large_start, hi = next.large_start;
      while (lo < hi) { const uint32_t mid = (lo + hi) >> 1; if (s.large_entries[mid].slot() < job.slot2) lo = mid + 1; else hi = mid; }
      uint32_t e = lo;
      while (e < next.large_start && s.large_entries[e].slot() == job.slot2) ++e;
      loc.large_begin = lo; loc.large_end = e;
   }
   if (lane == 0) locs[job_id] = loc;
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
// Every 32-bit word of the leaf's scratch bitset (an even number of them) is written: the scratch area is never zeroed.
constexpr uint32_t BSEL_ROWS = 32;     // rows per thread
__global__ void __launch_bounds__(256, 6)
leaf_bitmap_selection(const SeqDev* __restrict__ seqs, const BselJob* __restrict__ jobs, uint64_t* __restrict__ scratch) {
   __shared__ uint32_t cand_rows[8][1024];
   __shared__ uint32_t found_words[8][32];
   const BselJob job = jobs[blockIdx.y];
   const uint32_t lane = laneIdF(), warp = threadIdx.x >> 5;
   const uint32_t word = blockIdx.x * 256 + threadIdx.x, n_words = (job.n_rows + 31) / 32, n_words_out = (n_words + 1u) & ~1u;
   if (blockIdx.x * 256 + warp * 32 >= n_words_out) return;        // whole warp beyond the partition
   const SeqDev& s = seqs[job.seq];
   const uint32_t tile = job.value / s.tile_positions;
   // the coarse index is laid out with the STORE's row count; a BitmapSelection over fewer rows reads the same words
   const uint32_t index_words = (s.part_rows[job.partition] + 31) / 32;
   uint32_t candidates = 0;
   if (word < n_words && tile < s.n_tiles) candidates = __ldg(s.miss_tile_bits + s.miss_tile_base[job.partition] + size_t(tile) * index_words + word);
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
      const uint4* records = s.miss_inline + s.miss_inline_base[job.partition] * 2;
      // two candidates per lane at a time (a warp of 1,024 rows rarely has more than 64): their inline interval records are four independent 128-bit loads; only a row with
      // more intervals than a record holds (all words set) is searched in the CSR arrays
      for (uint32_t k0 = lane; k0 < total; k0 += 64) {
         uint32_t row[2];
         uint4 rec[2][2];
#pragma unroll
         for (uint32_t q = 0; q < 2; ++q) {
            row[q] = 0xFFFFFFFFu;
            rec[q][0] = rec[q][1] = make_uint4(0, 0, 0, 0);
            if (k0 + q * 32 < total) { row[q] = cand_rows[warp][k0 + q * 32]; rec[q][0] = __ldg(records + size_t(row[q]) * 2); rec[q][1] = __ldg(records + size_t(row[q]) * 2 + 1); }
         }
#pragma unroll
         for (uint32_t q = 0; q < 2; ++q) {
            if (row[q] == 0xFFFFFFFFu) continue;
            const uint32_t w[8] = {rec[q][0].x, rec[q][0].y, rec[q][0].z, rec[q][0].w, rec[q][1].x, rec[q][1].y, rec[q][1].z, rec[q][1].w};
            bool found = false;
            if (w[0] == 0xFFFFFFFFu && w[1] == 0xFFFFFFFFu) {      // sorted, disjoint intervals: binary search
               uint32_t lo = __ldg(row_off + row[q]), hi = __ldg(row_off + row[q] + 1);
               while (lo < hi) {
                  const uint32_t mid = (lo + hi) >> 1;
                  const uint2 r = __ldg(iv + mid);
                  if (job.value < r.x) hi = mid;
                  else if (job.value >= r.y) lo = mid + 1;
                  else { found = true; break; }
               }
            } else if (s.miss_narrow) {
#pragma unroll
               for (int t = 0; t < 8; ++t) found |= job.value >= (w[t] & 0xFFFFu) && job.value < (w[t] >> 16);
            } else {
#pragma unroll
               for (int t = 0; t < 4; ++t) found |= job.value >= w[2 * t] && job.value < w[2 * t + 1];
            }
            if (found) atomicOr(&found_words[warp][(row[q] >> 5) & 31u], 1u << (row[q] & 31));
         }
      }
      __syncwarp();
   }
   if (word >= n_words_out) return;
   uint32_t bits = 0;
   if (word < n_words) {
      const uint32_t found = found_words[warp][lane];
      const uint32_t valid = (word + 1) * 32 <= job.n_rows ? ~0u : (~0u >> (32 - (job.n_rows & 31)));
      bits = (job.contains ? found : ~found) & valid;
   }
   reinterpret_cast<uint32_t*>(scratch + job.dst_off)[word] = bits;
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

// thread per row: conjunction of column predicates.  grid.y = job; every 32-bit word of the scratch bitset is written
__global__ void leaf_selection(const SelJob* __restrict__ jobs, const PredDev* __restrict__ preds, uint64_t* __restrict__ scratch) {
   const SelJob job = jobs[blockIdx.y];
   const uint32_t row = blockIdx.x * blockDim.x + threadIdx.x;
   const uint32_t n_words_out = (((job.n_rows + 31) / 32) + 1u) & ~1u;
   if ((row >> 5) >= n_words_out) return;
   bool hit = row < job.n_rows;
   for (uint32_t k = 0; hit && k < job.n_preds; ++k) {
      const PredDev p = preds[job.pred_begin + k];
      if (p.dtype == 0) hit = compareValues<int32_t>(p.comparator, __ldg(static_cast<const int32_t*>(p.column) + row), p.i32);
      else if (p.dtype == 1) hit = compareValues<double>(p.comparator, __ldg(static_cast<const double*>(p.column) + row), p.f64);
      else hit = compareValues<uint32_t>(p.comparator, __ldg(static_cast<const uint32_t*>(p.column) + row), p.u32);
   }
   const uint32_t bits = __ballot_sync(0xFFFFFFFFu, hit);
   if (laneIdF() == 0) reinterpret_cast<uint32_t*>(scratch + job.dst_off)[row >> 5] = bits;
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

struct EvalArgs {
   const PlanDesc* plans;
   const uint32_t* programs;
   const LeafDesc* leaves;
   const LeafLoc* locs;
   const SeqDev* seqs;
   const RegContainer* reg_containers;   // registered index bitmaps of the database
   const uint16_t* reg_pool;
   const RegContainer* up_containers;    // bitmaps that came with the plan (BitmapProducer results, host-evaluated selections)
   const uint16_t* up_pool;
   const uint64_t* scratch;              // dense bitsets of the row-wise leaves
   uint64_t* out;
   uint32_t* ranks;
   uint64_t* deep_stack;                 // evaluation stacks in global memory when the batch needs more levels than fit in shared memory
   uint32_t levels;                      // stack levels below the top of stack that any plan of the batch needs
};

// sets bit v of the staging bitset (65,536 bits as 2,048 32-bit words)
__device__ __forceinline__ void stageBit(uint32_t* staging, uint32_t v) { atomicOr(staging + (v >> 5), 1u << (v & 31)); }
// sets [a, b] (inclusive, < 65,536); `part` of `parts` cooperating threads takes every parts-th word
__device__ __forceinline__ void stageRange(uint32_t* staging, uint32_t a, uint32_t b, uint32_t part, uint32_t parts) {
   const uint32_t wa = a >> 5, wb = b >> 5;
   for (uint32_t w = wa + part; w <= wb; w += parts) {
      uint32_t mask = ~0u;
      if (w == wa) mask &= ~0u << (a & 31);
      if (w == wb) mask &= ~0u >> (31 - (b & 31));
      atomicOr(staging + w, mask);
   }
}

// One roaring container -> staging (arrays, runs). All threads of the CTA call it with the same arguments.
__device__ __forceinline__ void stageContainer(uint32_t* staging, uint32_t kind, uint32_t size, const uint16_t* pay) {
   if (kind == KIND_ARRAY) {
      if ((reinterpret_cast<uintptr_t>(pay) & 15u) == 0) {   // 8 values per load
         const uint4* pay4 = reinterpret_cast<const uint4*>(pay);
         const uint32_t n4 = (size + 7) / 8;
         for (uint32_t i = threadIdx.x; i < n4; i += EVAL_THREADS) {
            const uint4 x = __ldg(pay4 + i);
            const uint32_t wv[4] = {x.x, x.y, x.z, x.w};
#pragma unroll
            for (uint32_t t = 0; t < 8; ++t) if (i * 8 + t < size) stageBit(staging, (wv[t >> 1] >> ((t & 1) * 16)) & 0xFFFFu);
         }
      } else {
         for (uint32_t i = threadIdx.x; i < size; i += EVAL_THREADS) stageBit(staging, __ldg(pay + i));
      }
   } else {   // KIND_RUN: few long runs are split over the threads word by word, many runs go one per thread
      if (size <= 8) {
         for (uint32_t r = 0; r < size; ++r) { const uint32_t start = __ldg(pay + 2 * r), len1 = __ldg(pay + 2 * r + 1); stageRange(staging, start, start + len1, threadIdx.x, EVAL_THREADS); }
      } else {
         for (uint32_t r = threadIdx.x; r < size; r += EVAL_THREADS) { const uint32_t start = __ldg(pay + 2 * r), len1 = __ldg(pay + 2 * r + 1); stageRange(staging, start, start + len1, 0, 1); }
      }
   }
}

// What a leaf contributes to ONE container key, resolved before the program runs: the descriptor, location and directory loads
// of a leaf are a chain of dependent loads, and a plan has many leaves -- one thread per leaf walks the chains of a whole
// window of leaves side by side instead of the CTA walking them one leaf after the other.
enum : uint32_t { RES_NONE = 0, RES_DIRECT = 1, RES_CONTAINER = 2, RES_POS = 3 };
struct LeafRes {
   const void* ptr;          // DIRECT: 64-bit words of this key as they lie (bitset container, dense scratch) | CONTAINER: payload
   uint32_t mode;
   uint32_t size;            // DIRECT: readable words (<= 1,024) | CONTAINER: cardinality / number of runs | POS: sequence index
   uint32_t e_begin, e_end;  // POS: large entries (arrays, runs) of this key | CONTAINER: e_begin = kind
   uint32_t post_begin, post_end;   // POS: postings of the slot in the slab (all of its keys)
};
static_assert(sizeof(LeafRes) == 32, "LeafRes must be 32 bytes");
constexpr uint32_t EVAL_RES_WINDOW = 64;
constexpr uint32_t EVAL_PROG_SMEM = 768;      // program words kept in shared memory (a 10-leaf N-Of is about 120)

__device__ __forceinline__ LeafRes resolveLeaf(const EvalArgs& a, const LeafDesc& leaf, uint32_t chunk) {
   LeafRes r{nullptr, RES_NONE, 0, 0, 0, 0, 0};
   const uint32_t w_base = chunk * EVAL_WORDS;
   if (leaf.kind == LEAF_DENSE) {
      if (leaf.a > w_base) { r.ptr = a.scratch + leaf.off + w_base; r.mode = RES_DIRECT; r.size = min(EVAL_WORDS, leaf.a - w_base); }
   } else if (leaf.kind == LEAF_POS) {
      const uint32_t slab_local = chunk / SLAB_KEYS, key_local = chunk % SLAB_KEYS;
      if (slab_local < leaf.b) {
         const SeqDev& s = a.seqs[leaf.a];
         const LeafLoc loc = a.locs[leaf.off + slab_local];
         uint32_t e_begin = loc.large_end, e_end = loc.large_end;
         for (uint32_t e = loc.large_begin; e < loc.large_end; ++e) {      // sorted by key inside the slot
            const LargeEntry le = s.large_entries[e];
            if (le.keyLocal() < key_local) continue;
            if (le.keyLocal() > key_local) break;
            if (le.kind() == KIND_BITSET) { r.ptr = s.large_pool + le.offset16; r.mode = RES_DIRECT; r.size = EVAL_WORDS; break; }
            if (e_begin == loc.large_end) e_begin = e;
            e_end = e + 1;
         }
         if (r.mode == RES_NONE && (e_begin < e_end || loc.post_begin < loc.post_end)) {
            r.mode = RES_POS; r.size = leaf.a; r.e_begin = e_begin; r.e_end = e_end; r.post_begin = loc.post_begin; r.post_end = loc.post_end;
         }
      }
   } else {
      const RegContainer* containers = (leaf.kind == LEAF_REG ? a.reg_containers : a.up_containers) + leaf.a;
      const uint16_t* pool = leaf.kind == LEAF_REG ? a.reg_pool : a.up_pool;
      uint32_t lo = 0, hi = leaf.b;       // containers are sorted by key
      while (lo < hi) { const uint32_t mid = (lo + hi) >> 1; if (containers[mid].key < chunk) lo = mid + 1; else hi = mid; }
      if (lo < leaf.b && containers[lo].key == chunk) {
         const RegContainer c = containers[lo];
         r.ptr = pool + c.pay_off;
         if (c.kind == KIND_BITSET) { r.mode = RES_DIRECT; r.size = EVAL_WORDS; }
         else { r.mode = RES_CONTAINER; r.size = c.size; r.e_begin = c.kind; }
      }
   }
   return r;
}

// grid: (ceil(max n_words_padded / 1024), n_plans). Thread t owns words t, t + 256, t + 512, t + 768 of the key.
// Shared memory: [levels][1024] uint64 evaluation stack (GLOBAL_STACK: in global memory instead) + 8 KB staging bitset.
template <bool GLOBAL_STACK>
__global__ void __launch_bounds__(EVAL_THREADS)
eval_plans(const EvalArgs a) {
   extern __shared__ __align__(16) uint64_t eval_smem[];
   __shared__ int warp_sums[EVAL_PER_THREAD][EVAL_THREADS / 32];
   __shared__ LeafRes res[EVAL_RES_WINDOW];
   __shared__ uint32_t prog_s[EVAL_PROG_SMEM];
   const PlanDesc plan = a.plans[blockIdx.y];
   const uint32_t chunk = blockIdx.x, w_base = chunk * EVAL_WORDS, tid = threadIdx.x;
   if (w_base >= plan.n_words_padded) return;
   uint64_t tos[EVAL_PER_THREAD];
#pragma unroll
   for (int j = 0; j < EVAL_PER_THREAD; ++j) tos[j] = 0;
   if (w_base < plan.n_words) {
      uint64_t* stk = GLOBAL_STACK ? a.deep_stack + (size_t(blockIdx.y) * gridDim.x + chunk) * size_t(a.levels) * EVAL_WORDS : eval_smem;
      uint32_t* staging = reinterpret_cast<uint32_t*>(GLOBAL_STACK ? eval_smem : eval_smem + size_t(a.levels) * EVAL_WORDS);
      uint64_t* staging64 = reinterpret_cast<uint64_t*>(staging);
#pragma unroll
      for (int j = 0; j < EVAL_PER_THREAD; ++j) staging64[j * EVAL_THREADS + tid] = 0;     // invariant: all zero between leaves
      __syncthreads();
      // the program: in shared memory when it fits (every operand fetch from global memory is a dependent load in front of the op)
      const uint32_t* pc = a.programs + plan.prog_off;
      if (plan.prog_len <= EVAL_PROG_SMEM) {
         for (uint32_t i = tid; i < plan.prog_len; i += EVAL_THREADS) prog_s[i] = __ldg(pc + i);
         pc = prog_s;
         __syncthreads();
      }
      const uint32_t* end = pc + plan.prog_len;
      uint64_t ahead[EVAL_PER_THREAD];      // the words of the NEXT leaf when they can be read in place: issued while this leaf's operator runs
      uint32_t ahead_leaf = 0xFFFFFFFFu;
      const LeafDesc* leaves = a.leaves + plan.leaf_off;
      int sp = 0;      // entries on the stack: levels 0 .. sp-2 in `stk`, the top in `tos`
      uint32_t res_begin = 0, res_end = 0;      // leaves [res_begin, res_end) of the plan are resolved in `res` (leaves are numbered in program order)
#define SB_LEVEL(l, j) stk[size_t(l) * EVAL_WORDS + (j) * EVAL_THREADS + tid]
#define SB_SPILL() do { if (sp > 0) { _Pragma("unroll") for (int j = 0; j < EVAL_PER_THREAD; ++j) SB_LEVEL(sp - 1, j) = tos[j]; } ++sp; } while (0)
      while (pc < end) {
         const uint32_t op = (*pc++);
         switch (op) {
            case P_ZERO:
               SB_SPILL();
#pragma unroll
               for (int j = 0; j < EVAL_PER_THREAD; ++j) tos[j] = 0;
               break;
            case P_RANGES: {
               const uint32_t n = (*pc++);
               SB_SPILL();
#pragma unroll
               for (int j = 0; j < EVAL_PER_THREAD; ++j) tos[j] = 0;
               for (uint32_t k = 0; k < n; ++k) {
                  const uint32_t r0 = pc[0], r1 = pc[1];
                  pc += 2;
#pragma unroll
                  for (int j = 0; j < EVAL_PER_THREAD; ++j) tos[j] |= rangeMask(w_base + j * EVAL_THREADS + tid, r0, r1);
               }
               break;
            }
            case P_LEAF: {
               const uint32_t li = (*pc++);
               if (li >= res_end || li < res_begin) {      // uniform over the CTA: resolve the next window of leaves, one thread per leaf
                  __syncthreads();
                  if (tid < EVAL_RES_WINDOW && li + tid < plan.n_leaves) res[tid] = resolveLeaf(a, leaves[li + tid], chunk);
                  res_begin = li; res_end = li + EVAL_RES_WINDOW;
                  __syncthreads();
               }
               const LeafRes r = res[li - res_begin];
               SB_SPILL();
#pragma unroll
               for (int j = 0; j < EVAL_PER_THREAD; ++j) tos[j] = 0;
               if (r.mode == RES_DIRECT) {
                  if (ahead_leaf == li) {
#pragma unroll
                     for (int j = 0; j < EVAL_PER_THREAD; ++j) tos[j] = ahead[j];
                  } else {
                     const uint64_t* direct = static_cast<const uint64_t*>(r.ptr);
#pragma unroll
                     for (int j = 0; j < EVAL_PER_THREAD; ++j) if (uint32_t(j * EVAL_THREADS + tid) < r.size) tos[j] = __ldg(direct + j * EVAL_THREADS + tid);
                  }
               } else if (r.mode != RES_NONE) {      // uniform over the CTA
                  if (r.mode == RES_CONTAINER) {
                     stageContainer(staging, r.e_begin, r.size, static_cast<const uint16_t*>(r.ptr));
                  } else {
                     const SeqDev& s = a.seqs[r.size];
                     const uint32_t key_local = chunk % SLAB_KEYS;
                     for (uint32_t e = r.e_begin; e < r.e_end; ++e) {
                        const LargeEntry le = s.large_entries[e];
                        stageContainer(staging, le.kind(), le.size, reinterpret_cast<const uint16_t*>(s.large_pool + le.offset16));
                     }
                     for (uint32_t i = r.post_begin + tid; i < r.post_end; i += EVAL_THREADS) {
                        const uint32_t p = __ldg(s.postings + i);
                        if (((p >> 16) & (SLAB_KEYS - 1)) == key_local) stageBit(staging, p & 0xFFFFu);
                     }
                  }
                  __syncthreads();
#pragma unroll
                  for (int j = 0; j < EVAL_PER_THREAD; ++j) { tos[j] = staging64[j * EVAL_THREADS + tid]; staging64[j * EVAL_THREADS + tid] = 0; }
                  __syncthreads();
               }
               if (li + 1 < res_end && li + 1 < plan.n_leaves && res[li + 1 - res_begin].mode == RES_DIRECT) {      // leaves are numbered in program order
                  const LeafRes nx = res[li + 1 - res_begin];
                  const uint64_t* direct = static_cast<const uint64_t*>(nx.ptr);
#pragma unroll
                  for (int j = 0; j < EVAL_PER_THREAD; ++j) ahead[j] = uint32_t(j * EVAL_THREADS + tid) < nx.size ? __ldg(direct + j * EVAL_THREADS + tid) : 0ull;
                  ahead_leaf = li + 1;
               }
               break;
            }
            case P_NOT: {   // flip(0, row_count)
               const uint32_t rc = (*pc++);
#pragma unroll
               for (int j = 0; j < EVAL_PER_THREAD; ++j) tos[j] ^= rangeMask(w_base + j * EVAL_THREADS + tid, 0, rc);
               break;
            }
            case P_AND:
               --sp;
#pragma unroll
               for (int j = 0; j < EVAL_PER_THREAD; ++j) tos[j] = SB_LEVEL(sp - 1, j) & tos[j];
               break;
            case P_OR:
               --sp;
#pragma unroll
               for (int j = 0; j < EVAL_PER_THREAD; ++j) tos[j] = SB_LEVEL(sp - 1, j) | tos[j];
               break;
            case P_ANDNOT:
               --sp;
#pragma unroll
               for (int j = 0; j < EVAL_PER_THREAD; ++j) tos[j] = SB_LEVEL(sp - 1, j) & ~tos[j];
               break;
            case P_THR_INIT: {
               const uint32_t d = (*pc++);
               for (uint32_t k = 0; k < d; ++k) {
                  SB_SPILL();
#pragma unroll
                  for (int j = 0; j < EVAL_PER_THREAD; ++j) tos[j] = 0;
               }
               break;
            }
            case P_THR_ADD: {   // threshold.cpp:92-129, one child; plane j = rows matching >= j + 1 children so far
               const int d = int(pc[0]), i = int(pc[1]), n = int(pc[2]), k = int(pc[3]);
               const uint32_t negated = pc[4], rc = pc[5];
               pc += 6;
               --sp;                              // the child is in `tos`; the d planes are levels sp-d .. sp-1, all in `stk`
               const int base = sp - d;
               const int lo = max(0, n - k + i - 1);
#pragma unroll
               for (int j = 0; j < EVAL_PER_THREAD; ++j) {
                  const uint64_t c = tos[j];
                  for (int q = min(d - 1, i); q > lo; --q) SB_LEVEL(base + q, j) |= negated ? (SB_LEVEL(base + q - 1, j) & ~c) : (SB_LEVEL(base + q - 1, j) & c);
                  if (k - i > n - 1) SB_LEVEL(base, j) |= negated ? (c ^ rangeMask(w_base + j * EVAL_THREADS + tid, 0, rc)) : c;
                  tos[j] = SB_LEVEL(base + d - 1, j);
               }
               break;
            }
            case P_THR_FIN: {
               const int d = int(pc[0]), n = int(pc[1]);
               const uint32_t exact = pc[2];
               pc += 3;
               const int base = sp - d;           // top plane in `tos`, the others in levels base .. sp-2
#pragma unroll
               for (int j = 0; j < EVAL_PER_THREAD; ++j) {
                  if (exact) tos[j] = (n - 1 == d - 1 ? tos[j] : SB_LEVEL(base + n - 1, j)) & ~(n == d - 1 ? tos[j] : SB_LEVEL(base + n, j));
               }
               sp = base + 1;
               break;
            }
            default: pc = end; break;
         }
      }
#undef SB_LEVEL
#undef SB_SPILL
#pragma unroll
      for (int j = 0; j < EVAL_PER_THREAD; ++j) if (w_base + j * EVAL_THREADS + tid >= plan.n_words) tos[j] = 0;
   }
   // result, rank tables: exclusive popcount prefix inside every 256-word block (block j of the key = words j*256 + t), block totals
   const uint32_t lane = laneIdF(), warp = tid >> 5;
   int pc_w[EVAL_PER_THREAD], incl[EVAL_PER_THREAD];
#pragma unroll
   for (int j = 0; j < EVAL_PER_THREAD; ++j) {
      a.out[plan.out_off + w_base + j * EVAL_THREADS + tid] = tos[j];
      pc_w[j] = __popcll(tos[j]);
      incl[j] = pc_w[j];
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) { const int t = __shfl_up_sync(0xFFFFFFFFu, incl[j], d); if (lane >= uint32_t(d)) incl[j] += t; }
      if (lane == 31) warp_sums[j][warp] = incl[j];
   }
   __syncthreads();
#pragma unroll
   for (int j = 0; j < EVAL_PER_THREAD; ++j) {
      int before = 0, total = 0;
#pragma unroll
      for (int k = 0; k < EVAL_THREADS / 32; ++k) { const int v = warp_sums[j][k]; if (k < int(warp)) before += v; total += v; }
      a.ranks[plan.rank_local_off + w_base + j * EVAL_THREADS + tid] = uint32_t(before + incl[j] - pc_w[j]);
      if (tid == 0) a.ranks[plan.rank_block_off + chunk * EVAL_PER_THREAD + j] = uint32_t(total);
   }
}

// one warp per plan: block totals -> exclusive prefix; the grand total is the cardinality
__global__ void finish_filters(const PlanDesc* __restrict__ plans, uint32_t n_plans, uint32_t* __restrict__ ranks) {
   const uint32_t plan_id = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = laneIdF();
   if (plan_id >= n_plans) return;
   const PlanDesc plan = plans[plan_id];
   uint32_t* blocks = ranks + plan.rank_block_off;
   const uint32_t n_blocks = plan.n_words_padded / RANK_BLOCK_WORDS;
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
   if (lane == 0) *plan.cardinality = carry;
}

}  // namespace sb
