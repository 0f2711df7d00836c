// Action kernels (sm_100a): the Mutations count tables.
//   k1_scan           fused AND-popcount over the container pool; never materialises filter AND bitmap.
//                     Replaces the and_cardinality / andnot_cardinality calls of
//                     src/silo/query_engine/actions/mutations.cpp:85-95 (M2) for all positions at once.
//   k2_missing        per-position count of filtered rows whose missing-symbol bitmap holds the position
//                     (mutations.cpp:74-83, 111-120), as a shared-memory difference array + scan, and the
//                     |F| terms of the deleted / flipped symbols.
//   k3_init_table     the table starts as the sum of the precomputed tables of the partitions whose filter is full
//                     (mutations.cpp:101-136).
//   classify_partitions / compact_runs    Mutations::preFilterBitmaps (mutations.cpp:36-63) on the device: a query is one
//                     submission, the host learns the cardinalities only with the result.
// All arithmetic is uint32 with wrap-around, like the reference (mutations.cpp:75,80,94).
#pragma once
#include "common.cuh"

namespace sb {

constexpr int K1_THREADS = 256;   // 8 warps per CTA of the container scan
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
__device__ __forceinline__ uint4 loadStream(const uint4* p) { return __ldg(p); }

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

// Row blocks of a partition: the work of a block is its filtered rows (dependent scattered loads each) far more than
// its rows, and the filter's density differs from partition to partition, so a partition gets blocks in proportion to
// cost = rows / 16 + filtered rows out of a budget of about one wave.
__host__ __device__ inline uint64_t k2Cost(uint32_t n_rows, uint32_t cardinality) { return uint64_t(n_rows) / 16 + cardinality; }
__host__ __device__ inline uint32_t k2Blocks(uint32_t n_rows, uint32_t cardinality, uint64_t total_cost, uint32_t budget) {
   const uint32_t share = 1u + uint32_t(k2Cost(n_rows, cardinality) * budget / (total_cost ? total_cost : 1));
   const uint32_t cap = (n_rows + 4095u) / 4096u;      // no block under 4,096 rows: every block pays the scan + flush over the positions
   return share < cap ? share : (cap ? cap : 1u);
}

// ---- classification (Mutations::preFilterBitmaps, mutations.cpp:36-63) on the device ----------------------------------
// One CTA per requested sequence: cardinality 0 -> skip, == sequence_count -> full, else partial. Leaves PartQuery[P], the
// lists of partial and full partitions, the cost total the missing-symbol kernel balances its blocks with, and zeroed
// hand-out counters. `force_partial`: the load-time pass that precomputes the full-filter tables.
struct ActionPtrs {            // device arrays of one (query, sequence)
   PartQuery* pq;              // [P]
   uint32_t* partial;          // [P]
   uint32_t* full;             // [P]
   ActionHead* head;
   uint32_t* selected;         // [n_runs] indices into SeqDev::runs of the partial partitions' runs, heaviest first
};
struct ClassifyArgs {
   uint32_t seq_index, pad;    // into the database's SeqDev array
   ActionPtrs out;
};
constexpr uint32_t CLASSIFY_MAX_SEQS = 32;      // sequences per launch (AminoAcidMutations over all genes: 12)
struct ClassifyBatch { ClassifyArgs seq[CLASSIFY_MAX_SEQS]; };

struct ClassifySeq { const uint32_t* part_rows; uint32_t n_partitions; };
__global__ void __launch_bounds__(256)
classify_partitions(const ClassifyBatch batch, const SeqDev* __restrict__ seqs, const PartInput* __restrict__ in, uint32_t n_in, int force_partial) {
   struct { ClassifySeq seq; ActionPtrs out; } a;
   a.seq.part_rows = seqs[batch.seq[blockIdx.x].seq_index].part_rows;
   a.seq.n_partitions = seqs[batch.seq[blockIdx.x].seq_index].n_partitions;
   a.out = batch.seq[blockIdx.x].out;
   const uint32_t P = a.seq.n_partitions;
   for (uint32_t p = threadIdx.x; p < P; p += blockDim.x) a.out.pq[p] = PartQuery{nullptr, nullptr, nullptr, PART_EMPTY, 0, 0, 0};
   __syncthreads();
   for (uint32_t i = threadIdx.x; i < n_in; i += blockDim.x) {
      const PartInput pi = in[i];
      const uint32_t rows = a.seq.part_rows[pi.partition];
      const uint64_t card = *pi.cardinality;
      if (rows == 0 || card == 0) continue;
      PartQuery q;
      q.filter32 = pi.filter32; q.rank_local = pi.rank_local; q.rank_block = pi.rank_block;
      q.cardinality = uint32_t(card); q.n_rows = rows; q.pad = 0;
      q.state = (!force_partial && card == rows) ? PART_FULL : PART_PARTIAL;
      a.out.pq[pi.partition] = q;
   }
   __syncthreads();
   if (threadIdx.x == 0) {      // ascending partition order, like the host loop it replaces; P is a few dozen
      uint32_t n_partial = 0, n_full = 0;
      uint64_t cost = 0;
      for (uint32_t p = 0; p < P; ++p) {
         const PartQuery q = a.out.pq[p];
         if (q.state == PART_PARTIAL) { a.out.partial[n_partial++] = p; cost += k2Cost(q.n_rows, q.cardinality); }
         else if (q.state == PART_FULL) a.out.full[n_full++] = p;
      }
      ActionHead h{};
      h.n_partial = n_partial; h.n_full = n_full; h.produced = (n_partial + n_full) ? 1u : 0u; h.k2_total_cost = cost;
      *a.out.head = h;
   }
}

// indices of the runs of the partial partitions, in the order of SeqDev::runs (heaviest first; blocks may land out of order)
__global__ void __launch_bounds__(256)
compact_runs(const SeqDev s, const ActionPtrs a) {
   __shared__ uint32_t warp_base[8];
   __shared__ uint32_t block_base;
   const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x, lane = laneId(), warp = threadIdx.x >> 5;
   const bool take = i < s.n_runs && a.pq[s.runs[i].partition].state == PART_PARTIAL;
   const uint32_t mask = __ballot_sync(0xFFFFFFFFu, take);
   if (lane == 0) warp_base[warp] = __popc(mask);
   __syncthreads();
   if (threadIdx.x == 0) {
      uint32_t total = 0;
      for (int k = 0; k < 8; ++k) { const uint32_t v = warp_base[k]; warp_base[k] = total; total += v; }
      block_base = total ? atomicAdd(&a.head->n_selected, total) : 0u;
   }
   __syncthreads();
   if (take) a.selected[block_base + warp_base[warp] + __popc(mask & laneMaskLt())] = i;
}

// counts[i] = sum over the full partitions of their precomputed table (mutations.cpp:101-136); zero when there is none.
// First kernel of the action: it initialises the whole table, the "produced" word in front of it and the row counter.
__global__ void k3_init_table(const uint32_t* __restrict__ full_tables, size_t table_len, const ActionPtrs a, uint32_t* __restrict__ counts,
                              uint32_t* __restrict__ flag, uint32_t* __restrict__ row_counter) {
   const size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
   if (i >= table_len) return;
   const uint32_t n_full = a.head->n_full;
   if (i == 0) {
      if (flag) *flag = a.head->produced;
      if (row_counter) *row_counter = 0;
   }
   uint32_t acc = 0;
   for (uint32_t k = 0; k < n_full; ++k) acc += __ldg(full_tables + size_t(a.full[k]) * table_len + i);
   counts[i] = acc;
}

// ---- K1: the container scan -----------------------------------------------------------------------------------------
// Persistent CTAs (4 per SM) take the selected RUNS of work items from a device counter, heaviest first; the items of a run
// belong to one slab (ascending tiles), so the slab's 32 KB filter slice is copied once per run. Shared memory:
// [filter slab 32 KB][tile counts (tile_slots + 32 words)][mbarrier][two unit counters][next item].
// tile_counts[sym * T + pos_in_tile] accumulates the RAW |F AND B[pos][sym]|; the flipped / deleted-symbol arithmetic
// (mutations.cpp:85-95, 122-133) is linear in it and is applied once per slot in the flush, which also leaves the table
// zeroed for the next item. Two CTA barriers per item: before its first increment (the previous flush is done) and before
// its flush (all units are counted).
struct K1Next { WorkItem item; const uint32_t* filter32; uint32_t valid, pad; };

// FLAGS (compiled side by side): 1 = a hit increments its slot with a predicated reduction (a miss costs no shared-memory
// operation) instead of every posting incrementing either its slot or the lane's dummy slot.
template <int UNROLL, int MIN_CTAS, int FLAGS>
__global__ void __launch_bounds__(K1_THREADS, MIN_CTAS)
k1_scan(const SeqDev s, const ActionPtrs a, uint32_t* __restrict__ counts, int full_mode) {
   constexpr int K1_UNROLL = UNROLL;
   extern __shared__ __align__(128) uint8_t k1_smem[];
   uint32_t* filter_s = reinterpret_cast<uint32_t*>(k1_smem);
   uint32_t* tile_counts = reinterpret_cast<uint32_t*>(k1_smem + K1_FILTER_BYTES);
   const uint32_t T = s.tile_positions, tile_slots = T * s.sigma;
   uint64_t* bar = reinterpret_cast<uint64_t*>(tile_counts + k1TableWords(tile_slots));
   uint32_t* unit_counters = reinterpret_cast<uint32_t*>(bar + 1);      // [2]: item i hands out its units through counter i & 1
   K1Next* item_slots = reinterpret_cast<K1Next*>(bar + 2);      // [2]: the current item and the one thread 0 fetches meanwhile (read in place: no registers held)
   constexpr uint32_t CHUNK4 = 32 * K1_UNROLL;    // 128-bit words per chunk
   constexpr uint32_t N_WARPS = K1_THREADS / 32;

   // thread 0 only: the item after the current one -- the next item of the run, or the first item of the next run from the
   // device counter (index, run, item and filter pointer are dependent loads: fetched while the other warps stream)
   uint32_t run_next = 0, run_end = 0;
   const uint32_t* run_filter = nullptr;
   auto fetchNext = [&](K1Next* into) {
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
      *into = n;
   };
   if (threadIdx.x == 0) {
      mbarInit(bar, 1);
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
      unit_counters[0] = N_WARPS;
      fetchNext(item_slots);
   }
   for (uint32_t i = threadIdx.x; i < tile_slots + 32; i += K1_THREADS) tile_counts[i] = 0;
   const uint32_t lane = laneId(), warp = threadIdx.x >> 5;
   const uint32_t dummy = tile_slots << POSTING_ROW_BITS;          // a posting of the slot that is never flushed
   const uint4 none = make_uint4(dummy, dummy, dummy, dummy);
   uint32_t parity = 0, turn = 0;
   uint32_t cur_slab = 0xFFFFFFFFu;
   __syncthreads();
   for (;; turn ^= 1u) {
      const K1Next& cur = item_slots[turn];
      if (!cur.valid) break;
      const WorkItem& item = cur.item;
      const bool new_slab = item.slab != cur_slab;        // the items of a run keep the filter slice
      cur_slab = item.slab;
      uint32_t* next_unit = unit_counters + turn;
      // Work units of the item, handed out to the warps through a shared counter (a warp that is busy with a bitset
      // simply takes fewer): first the large containers -- a bitset as two halves of 4 KB, anything else whole --
      // then the postings in chunks of 128 x 16 bytes. The first unit of a warp is its warp id; the loads of a posting
      // chunk are issued one unit ahead (and before waiting for the filter: they do not depend on it).
      const uint4* post4 = reinterpret_cast<const uint4*>(s.postings + item.post_begin);
      const uint32_t n4 = (item.post_end - item.post_begin + 3u) >> 2;
      const uint32_t n_large_units = 2u * (item.large_end - item.large_begin);
      const uint32_t n_units = n_large_units + (n4 + CHUNK4 - 1) / CHUNK4;
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
      loadChunk(unit, v);      // issued before the barrier: the loads travel while the other warps finish the previous item's flush
      __syncthreads();                                    // the previous flush (and its zeroing) is done; nobody reads the other item slot any more
      if (threadIdx.x == 0) {
         if (new_slab) {
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            mbarExpectTx(bar, K1_FILTER_BYTES);
            bulkLoad(filter_s, cur.filter32 + size_t(item.key0) * 2048u, K1_FILTER_BYTES, bar);
         }
         unit_counters[turn ^ 1u] = N_WARPS;              // the next item's counter: nobody touches it before the next barrier
      }

      // flags of this thread's flush positions (one or two: a tile holds 256 or 512), fetched now so that the flush does not wait for them
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
      if (new_slab) { mbarWait(bar, parity); }
      // thread 0 fetches the CTA's next item while the other warps stream
      if (threadIdx.x == 0) fetchNext(item_slots + (turn ^ 1u));

      // Eight postings at a time: the eight filter words are fetched first (independent shared-memory loads: one round trip per
      // eight postings), then the hits are counted -- FLAGS & 1: a reduction under a predicate; otherwise an unconditional
      // increment of the posting's slot or, for a miss, of the lane's dummy slot.
      const uint32_t counts_s = smemAddr(tile_counts);
      auto probe8 = [&](const uint4& p0, const uint4& p1) {
         const uint32_t p[8] = {p0.x, p0.y, p0.z, p0.w, p1.x, p1.y, p1.z, p1.w};
         uint32_t w[8];
#pragma unroll
         for (int t = 0; t < 8; ++t) w[t] = filter_s[(p[t] >> 5) & (SLAB_WORDS32 - 1)];
#pragma unroll
         for (int t = 0; t < 8; ++t) {
            const uint32_t hit = (w[t] >> (p[t] & 31)) & 1u;
            if (FLAGS & 1) {
               // (ptxas wraps the predicated reduction in a branch either way; with few hits per warp it still beats the unconditional form)
               asm volatile("{ .reg .pred q; setp.ne.u32 q, %0, 0; @q red.shared.add.u32 [%1], 1; }" ::"r"(hit), "r"(counts_s + ((p[t] >> POSTING_ROW_BITS) << 2)) : "memory");
            } else {
               atomicAdd(&tile_counts[hit ? (p[t] >> POSTING_ROW_BITS) : tile_slots + lane], 1u);   // a miss increments the lane's dummy slot: no branch
            }
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
               const PartQuery q = a.pq[item.partition];
               for (uint32_t i = lane; i < k_size; i += 32) { const uint32_t start = __ldg(runs + 2 * i), len1 = __ldg(runs + 2 * i + 1); c2 += int(filterRangeCount(q, item.key0 + k_local, start, start + len1)); }
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
      if (new_slab) parity ^= 1u;
      __syncthreads();
      // flush: thread = position of the tile; symbol rows of counts[sym][pos] get coalesced REDs
      static_assert(MAX_TILE_POSITIONS % K1_THREADS == 0, "whole flush positions per thread");
#pragma unroll
      for (uint32_t f = 0; f < FLUSH_MAX; ++f) {
         const uint32_t pl = threadIdx.x + f * K1_THREADS, flush_pos = item.tile * T + pl;
         if (pl >= T || flush_pos >= s.length) continue;
         int acc_deleted = 0;
         for (uint32_t sym0 = 0; sym0 < s.sigma; sym0 += 8) {      // eight counters in flight, then their REDs
            int raw[8];
#pragma unroll
            for (uint32_t q = 0; q < 8; ++q) raw[q] = sym0 + q < s.sigma ? int(tile_counts[(sym0 + q) * T + pl]) : 0;
#pragma unroll
            for (uint32_t q = 0; q < 8; ++q) if (sym0 + q < s.sigma && raw[q] != 0) tile_counts[(sym0 + q) * T + pl] = 0;      // the next item starts from zero
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

// ---- K2: rows that miss a position -------------------------------------------------------------------------------------
constexpr int K2_THREADS = 1024;
constexpr uint32_t K2_MAX_WINDOW = 49152;   // positions per shared-memory difference array (192 KB)

// grid: (about one wave of row blocks, position windows). Block b walks the list of partial partitions to find the partition
// and the share of its rows it owns. Per block: shared-memory difference array over the positions fed by the missing
// intervals of the block's FILTERED rows (a thread takes a 32-row word of the filter and walks its set bits, four rows in
// flight), in-place scan, then the block's share of the deleted / flipped-symbol terms goes straight into the table
// (mutations.cpp:74-83, 111-133): counts[deleted][pos] -= #{block's filtered rows missing pos}, and the first
// block of a partition adds the |F| terms.
__global__ void __launch_bounds__(K2_THREADS)
k2_missing(const SeqDev s, const ActionPtrs a, uint32_t* __restrict__ counts, uint32_t window, uint32_t budget, int full_mode) {
   extern __shared__ int diff[];   // [window + 1]
   __shared__ int warp_sums[32];
   __shared__ uint32_t found[3];
   const ActionHead head = *a.head;
   if (threadIdx.x == 0) {
      uint32_t b = blockIdx.x, hit = 0xFFFFFFFFu, blocks = 0;
      for (uint32_t j = 0; j < head.n_partial; ++j) {
         const PartQuery q = a.pq[a.partial[j]];
         blocks = k2Blocks(q.n_rows, q.cardinality, head.k2_total_cost, budget);
         if (b < blocks) { hit = a.partial[j]; break; }
         b -= blocks;
      }
      found[0] = hit; found[1] = b; found[2] = blocks;
   }
   __syncthreads();
   const uint32_t partition = found[0];
   if (partition == 0xFFFFFFFFu) return;
   const PartQuery q = a.pq[partition];
   const uint32_t n_blocks = found[2];
   const uint32_t rows_per_block = ((q.n_rows + n_blocks - 1) / n_blocks + 31u) & ~31u;
   const uint32_t row0 = found[1] * rows_per_block;
   if (row0 >= q.n_rows) return;
   const uint32_t row1 = min(q.n_rows, row0 + rows_per_block);
   const uint32_t w0 = blockIdx.y * window, w1 = min(s.length, w0 + window), width = w1 - w0;
   for (uint32_t i = threadIdx.x; i <= window; i += blockDim.x) diff[i] = 0;
   __syncthreads();

   const uint32_t* row_off = s.miss_row_off + s.miss_row_base[partition];
   const uint2* iv = s.miss_iv + s.miss_iv_base[partition];
   bool any_row = false;
   // A thread takes a 32-row word of the filter at a time and walks its set bits four rows at a time: the rows' inline interval
   // records are eight independent 128-bit loads, nothing depends on another load (the CSR arrays are read only for the few
   // rows with more intervals than a record holds).
   const uint32_t lane = laneId(), warp = threadIdx.x >> 5;
   const uint32_t word_begin = row0 >> 5, word_end = (row1 + 31) >> 5;
   int at_window_start = 0;      // intervals that begin at the window's first position (leading runs: nearly every row) are counted in a register
   const uint4* records = s.miss_inline + s.miss_inline_base[partition] * 2;
   const uint32_t* any_missing = s.miss_any + s.miss_any_base[partition];
   auto apply = [&](uint32_t start, uint32_t end) {      // an empty slot is (0, 0): nothing happens
      const uint32_t lo = max(start, w0), hi = min(end, w1);
      if (lo < hi) {
         if (lo == w0) ++at_window_start; else atomicAdd(&diff[lo - w0], 1);
         if (hi < w1) atomicAdd(&diff[hi - w0], -1);      // diff[width] is never read
      }
   };
   for (uint32_t word = word_begin + threadIdx.x; word < word_end; word += K2_THREADS) {
      uint32_t bits = __ldg(q.filter32 + word) & __ldg(any_missing + word);      // filtered rows that miss anything at all (a short gene: few)
      if ((word + 1) * 32 > row1) bits &= ~0u >> (32 - (row1 & 31));      // row1 is the partition's end here: row blocks are multiples of 32 rows
      if (bits) any_row = true;
      while (bits) {
         uint4 rec[4][2];
         uint32_t rows[4];
#pragma unroll
         for (int k = 0; k < 4; ++k) {
            rows[k] = 0xFFFFFFFFu;
            rec[k][0] = rec[k][1] = make_uint4(0, 0, 0, 0);
            if (bits) {
               rows[k] = word * 32 + uint32_t(__ffs(int(bits))) - 1u;
               bits &= bits - 1u;
               rec[k][0] = __ldg(records + size_t(rows[k]) * 2);
               rec[k][1] = __ldg(records + size_t(rows[k]) * 2 + 1);
            }
         }
#pragma unroll
         for (int k = 0; k < 4; ++k) {
            const uint32_t w[8] = {rec[k][0].x, rec[k][0].y, rec[k][0].z, rec[k][0].w, rec[k][1].x, rec[k][1].y, rec[k][1].z, rec[k][1].w};
            if (w[0] == 0xFFFFFFFFu && w[1] == 0xFFFFFFFFu) {      // more intervals than a record holds: the CSR arrays
               for (uint32_t j = __ldg(row_off + rows[k]), e = __ldg(row_off + rows[k] + 1); j < e; ++j) { const uint2 r = __ldg(iv + j); apply(r.x, r.y); }
            } else if (s.miss_narrow) {
#pragma unroll
               for (int t = 0; t < 8; ++t) apply(w[t] & 0xFFFFu, w[t] >> 16);
            } else {
#pragma unroll
               for (int t = 0; t < 4; ++t) apply(w[2 * t], w[2 * t + 1]);
            }
         }
      }
   }
   at_window_start = warpSum(at_window_start);
   if (lane == 0 && at_window_start) atomicAdd(&diff[0], at_window_start);
   const bool first_block = found[1] == 0;
   if (!__syncthreads_or(any_row) && !first_block) return;

   // inclusive scan of diff in place: every thread owns `per` consecutive elements
   const uint32_t per = (width + K2_THREADS - 1) / K2_THREADS;
   const uint32_t begin = min(threadIdx.x * per, width), end = min(begin + per, width);
   int sum = 0;
   for (uint32_t i = begin; i < end; ++i) sum += diff[i];
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

   // the block's share of  deleted += |F| - #{i in F : pos in M_i}  and of the |F| terms of a flipped symbol
   const uint32_t n_filtered = q.cardinality;
   const int8_t* deleted = s.deleted + size_t(partition) * s.length;
   const int8_t* flipped = s.flipped + size_t(partition) * s.length;
   // four positions per thread at a time: their flag loads are issued together (one round trip instead of four in a row)
   for (uint32_t i0 = threadIdx.x; i0 < width; i0 += 4 * blockDim.x) {
      int d[4], f[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) {
         const uint32_t i = i0 + k * blockDim.x;
         d[k] = f[k] = -1;
         if (i < width && (first_block || diff[i] != 0)) { d[k] = deleted[w0 + i]; if (first_block) f[k] = flipped[w0 + i]; }
      }
#pragma unroll
      for (int k = 0; k < 4; ++k) {
         const uint32_t i = i0 + k * blockDim.x;
         if (i >= width) continue;
         const uint32_t pos = w0 + i;
         const uint32_t miss = uint32_t(diff[i]);
         if (miss == 0 && !first_block) continue;
         uint32_t add = 0u - miss;
         if (first_block) {
            add += n_filtered;
            if (f[k] >= 0) {
               if (!full_mode) add -= n_filtered;               // the flipped symbol's |F| part of the deleted symbol's count
               atomicAdd(counts + size_t(f[k]) * s.length + pos, n_filtered);
            }
         }
         if (d[k] >= 0 && add) atomicAdd(counts + size_t(d[k]) * s.length + pos, add);
      }
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

}  // namespace sb
