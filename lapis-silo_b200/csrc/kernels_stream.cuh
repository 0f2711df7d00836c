// k1_stream (sm_100a): the container scan of the Mutations action as ONE persistent CTA per SM.
//
// k1_and_popcount (kernels_action.cuh) launches one CTA per work item; every CTA pays a start-up (item fetch, 32 KB
// filter copy), a block barrier and a flush during which its warps do not stream: measured with the per-CTA timeline
// (tools/k1_experiment.py) those phases are a third of the CTA time, and a read-only kernel with the same loads is
// 1.4x faster (tools/membench.cu). Here a CTA owns a cost-balanced share of the query's work items, cut on the host into
// RUNS (consecutive items of one slab). Per run the filter slab is copied into shared memory once; after that the 32
// warps never meet at a barrier:
//   * units (posting chunks, bitset halves, array pieces) of all items of the run form one ticket sequence; a warp takes
//     the next ticket from a shared counter and has the postings of its next unit in flight while it probes the current;
//   * item j counts into table j mod n_tables (as many tables as fit in shared memory, at most 8); a warp that finishes a
//     unit bumps the item's `done` counter, and the warp that completes the item puts its table up for flushing;
//   * the flush of a table is T / 32 PARTS of 32 positions that any warp takes (compare-and-swap on the table's state word)
//     before it takes its next unit: read the part's counters into registers, zero them, and -- once the part is handed
//     back, so that the table can go to item j + n_tables as early as possible -- apply the flipped / deleted-symbol
//     arithmetic of mutations.cpp:85-95, 122-133 and emit coalesced REDs from the registers.
// Same arithmetic, same layouts, same results as k1_and_popcount (tests run both). Opt-in (SB_K1=stream): measured 1.6x
// slower than k1_and_popcount so far (DESIGN.md 4 has the numbers and the reasons).
#pragma once
#include "kernels_action.cuh"

namespace sb {

constexpr int KS_THREADS = 1024;
constexpr int KS_WARPS = KS_THREADS / 32;
constexpr uint32_t KS_MAX_TABLES = 8;    // count tables in rotation: as many as fit (ksTables)
constexpr uint32_t KS_MAX_ITEMS = 256;   // items per run (the host cuts longer runs)
#ifndef SB_KS_ITEM_COST
#define SB_KS_ITEM_COST 6
#endif
constexpr uint32_t KS_ITEM_COST = SB_KS_ITEM_COST;   // cost of flushing an item's table, in units (for the cost-balanced shares of the CTAs)

struct KsRun { uint32_t item_begin, item_end; };     // items [begin, end) of one slab, at most KS_MAX_ITEMS
struct KsMap {
   const uint32_t* run_off;                              // [grid + 1] runs of CTA b: runs[run_off[b] .. run_off[b + 1])
   const KsRun* runs;
   const uint32_t* inline_filter[K1_INLINE_PARTITIONS];  // filter32 of partition p (p < K1_INLINE_PARTITIONS)
};

struct KsItem {          // what the unit loop needs of a WorkItem, in shared memory
   uint32_t post_begin, n4, large_begin, n_large_units, tile, unit_base, n_units, pad;
};
static_assert(sizeof(KsItem) == 32, "KsItem must be 32 bytes");

constexpr size_t KS_SMEM_BUDGET = 200 * 1024;     // of the 227 KB a CTA may have; the rest stays L1
inline size_t ksFixedBytes() { return K1_FILTER_BYTES + size_t(KS_MAX_ITEMS) * sizeof(KsItem) + 256; }
inline uint32_t ksTables(uint32_t tile_slots) {
   const size_t fit = (KS_SMEM_BUDGET - ksFixedBytes()) / (size_t(k1TableWords(tile_slots)) * 4);
   return uint32_t(std::max<size_t>(2, std::min<size_t>(KS_MAX_TABLES, fit)));
}
inline size_t ksSharedBytes(uint32_t tile_slots) { return ksFixedBytes() + size_t(ksTables(tile_slots)) * k1TableWords(tile_slots) * 4; }


#ifdef SB_K1_TIMELINE
// diagnostics build: per-CTA sums over the warps of the clock cycles spent {0 smid, 1 start ns, 2 end ns, 3 prologue, 4 waiting for a table,
// 5 large units, 6 posting units, 7 flushing} (row of 8 per CTA in g_k1_timeline)
#define KS_CLOCK(var) const long long var = clock64()
#define KS_ACCOUNT(k, from, to) acct[k] += (to) - (from)
#else
#define KS_CLOCK(var) do { } while (0)
#define KS_ACCOUNT(k, from, to) do { } while (0)
#endif

constexpr uint32_t KS_PENDING = 0x80000000u;    // state word of a table: flush parts taken so far | KS_PENDING while parts are left

// Shared-memory control block of a CTA (behind the item descriptors).
struct KsControl {
   uint64_t bar;                          // mbarrier of the filter copy
   uint32_t ticket;                       // next unit of the run
   uint32_t flushed;                      // items of the run whose table went back into rotation
   uint32_t done[KS_MAX_TABLES];          // finished units of the item that owns the table
   uint32_t owner[KS_MAX_TABLES];         // item (index in the run) that may count into the table
   uint32_t fstate[KS_MAX_TABLES];        // flush state word
   uint32_t fdone[KS_MAX_TABLES];         // finished flush parts
   uint32_t warp_sums[8];                 // scan scratch of the prologue
   uint32_t run_partition, run_key0;
};
static_assert(sizeof(KsControl) <= 256, "ksFixedBytes reserves 256 bytes for the control block");

template <uint32_t SIGMA>
__global__ void __launch_bounds__(KS_THREADS, 1)
k1_stream(const SeqDev s, const PartQuery* __restrict__ pq, const KsMap map, uint32_t* __restrict__ counts, int full_mode, uint32_t n_tables) {
   extern __shared__ __align__(128) uint8_t ks_smem[];
   const uint32_t T = s.tile_positions, tile_slots = T * SIGMA, table_words = k1TableWords(tile_slots), n_parts = T / 32;
   uint32_t* filter_s = reinterpret_cast<uint32_t*>(ks_smem);
   uint32_t* tables = reinterpret_cast<uint32_t*>(ks_smem + K1_FILTER_BYTES);
   KsItem* items_s = reinterpret_cast<KsItem*>(tables + size_t(n_tables) * table_words);
   KsControl* ctl = reinterpret_cast<KsControl*>(items_s + KS_MAX_ITEMS);
   volatile uint32_t* owner = ctl->owner;
   volatile uint32_t* fstate = ctl->fstate;
   volatile uint32_t* flushed = &ctl->flushed;

   const uint32_t lane = laneId(), warp = threadIdx.x >> 5;
   const uint32_t run0 = __ldg(map.run_off + blockIdx.x), run1 = __ldg(map.run_off + blockIdx.x + 1);
   if (run0 == run1) return;
#ifdef SB_K1_TIMELINE
   long long acct[8] = {};
   if (g_k1_timeline && threadIdx.x == 0) { g_k1_timeline[size_t(blockIdx.x) * 8 + 0] = smId(); g_k1_timeline[size_t(blockIdx.x) * 8 + 1] = globalTimerNs(); }
#endif
   if (threadIdx.x == 0) {
      mbarInit(&ctl->bar, 1);
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
   }
   for (uint32_t i = threadIdx.x; i < n_tables * table_words; i += KS_THREADS) tables[i] = 0;

   constexpr uint32_t CHUNK4 = 32 * K1_UNROLL;    // 128-bit words per posting chunk
   const uint32_t dummy = tile_slots << POSTING_ROW_BITS;
   const uint4 none = make_uint4(dummy, dummy, dummy, dummy);

   for (uint32_t run = run0; run < run1; ++run) {
      KS_CLOCK(c_run);
      const KsRun r = map.runs[run];
      const uint32_t n_items = r.item_end - r.item_begin;
      // ---- run prologue: item descriptors + unit prefix into shared memory, filter slab by bulk copy ----
      uint32_t my_units = 0;
      WorkItem w{};
      if (threadIdx.x < n_items) {
         w = s.items[r.item_begin + threadIdx.x];
         const uint32_t n4 = (w.post_end - w.post_begin + 3u) >> 2;
         my_units = 2u * (w.large_end - w.large_begin) + (n4 + CHUNK4 - 1) / CHUNK4;
      }
      uint32_t incl = my_units;
      if (warp < KS_MAX_ITEMS / 32) {
#pragma unroll
         for (int d = 1; d < 32; d <<= 1) { const uint32_t t = __shfl_up_sync(0xFFFFFFFFu, incl, d); if (lane >= uint32_t(d)) incl += t; }
      }
      __syncthreads();          // every warp is out of the previous run: its tables are flushed and zero, its control words are free
      if (warp < KS_MAX_ITEMS / 32 && lane == 31) ctl->warp_sums[warp] = incl;
      if (threadIdx.x == 0) { ctl->run_partition = w.partition; ctl->run_key0 = w.key0; ctl->ticket = KS_WARPS; ctl->flushed = 0; }
      if (threadIdx.x < KS_MAX_TABLES) { ctl->done[threadIdx.x] = 0; ctl->owner[threadIdx.x] = threadIdx.x; ctl->fstate[threadIdx.x] = 0; ctl->fdone[threadIdx.x] = 0; }
      __syncthreads();
      if (threadIdx.x == 0) {
         const uint32_t* filter32 = w.partition < K1_INLINE_PARTITIONS ? map.inline_filter[w.partition] : pq[w.partition].filter32;
         mbarExpectTx(&ctl->bar, K1_FILTER_BYTES);
         bulkLoad(filter_s, filter32 + size_t(w.key0) * 2048u, K1_FILTER_BYTES, &ctl->bar);
      }
      if (threadIdx.x < n_items) {
         uint32_t before = 0;
         for (uint32_t k = 0; k < warp; ++k) before += ctl->warp_sums[k];
         KsItem it;
         it.post_begin = w.post_begin; it.n4 = (w.post_end - w.post_begin + 3u) >> 2; it.large_begin = w.large_begin;
         it.n_large_units = 2u * (w.large_end - w.large_begin); it.tile = w.tile; it.unit_base = before + incl - my_units; it.n_units = my_units; it.pad = 0;
         items_s[threadIdx.x] = it;
      }
      __syncthreads();
      const uint32_t partition = ctl->run_partition, key0 = ctl->run_key0;
      const uint32_t total_units = items_s[n_items - 1].unit_base + items_s[n_items - 1].n_units;
      const int8_t* del_row = s.deleted + size_t(partition) * s.length;
      const int8_t* flip_row = s.flipped + size_t(partition) * s.length;
      mbarWait(&ctl->bar, (run - run0) & 1u);
      KS_CLOCK(c_ready);
      KS_ACCOUNT(3, c_run, c_ready);

      // ---- flush helper: takes one part (32 positions) of a table that is up for flushing; false if there is none ----
      auto helpFlush = [&]() -> bool {
         const uint32_t st = lane < n_tables ? fstate[lane] : 0u;
         const uint32_t cand = __ballot_sync(0xFFFFFFFFu, (st & KS_PENDING) != 0u);
         if (cand == 0u) return false;
         const uint32_t tb = uint32_t(__ffs(int(cand))) - 1u;
         const uint32_t old = __shfl_sync(0xFFFFFFFFu, st, int(tb));
         const uint32_t part = old & 0xFFu;
         const uint32_t taken = part + 1u == n_parts ? n_parts : old + 1u;      // the last part clears KS_PENDING
         uint32_t got = 0;
         if (lane == 0) got = atomicCAS(&ctl->fstate[tb], old, taken) == old;
         if (!__shfl_sync(0xFFFFFFFFu, got, 0)) return true;                     // another warp was faster: look again
         const uint32_t item = owner[tb];        // stable while the table is being flushed
         uint32_t* table = tables + size_t(tb) * table_words;
         const uint32_t pl = part * 32u + lane, pos = items_s[item].tile * T + pl;
         int del = -1, flip = -1;
         if (pos < s.length) { del = del_row[pos]; flip = flip_row[pos]; }
         __threadfence_block();
         int raw[SIGMA];
#pragma unroll
         for (uint32_t sym = 0; sym < SIGMA; ++sym) raw[sym] = int(table[sym * T + pl]);
#pragma unroll
         for (uint32_t sym = 0; sym < SIGMA; ++sym) table[sym * T + pl] = 0;
         if (part == 0) table[tile_slots + lane] = 0;                            // the dummy slots
         __threadfence_block();
         if (lane == 0) {
            if (atomicAdd(&ctl->fdone[tb], 1u) + 1u == n_parts) {                // last part: the table goes to item + n_tables
               ctl->fdone[tb] = 0; ctl->done[tb] = 0; fstate[tb] = 0;
               __threadfence_block();
               owner[tb] = item + n_tables;
               atomicAdd(&ctl->flushed, 1u);
            }
         }
         if (pos < s.length) {     // from here on only registers and global REDs (mutations.cpp:85-95, 122-133)
            int acc_deleted = 0;
#pragma unroll
            for (uint32_t sym = 0; sym < SIGMA; ++sym) {
               if (raw[sym] == 0 || int(sym) == del) continue;
               const int add_slot = int(sym) == flip ? -raw[sym] : raw[sym];
               atomicAdd(counts + size_t(sym) * s.length + pos, uint32_t(add_slot));
               acc_deleted += full_mode ? -raw[sym] : -add_slot;
            }
            if (del >= 0 && acc_deleted != 0) atomicAdd(counts + size_t(del) * s.length + pos, uint32_t(acc_deleted));
         }
         return true;
      };

      // ---- unit loop ----
      uint4 v[K1_UNROLL];
      auto locate = [&](uint32_t t, uint32_t& cur) {     // tickets of a warp ascend: the cursor only moves forward
         while (cur + 1 < n_items && t >= items_s[cur + 1].unit_base) ++cur;
      };
      auto loadChunk = [&](uint32_t t, uint32_t cur, uint4 (&dst)[K1_UNROLL]) {
         if (t >= total_units) return;
         const KsItem& it = items_s[cur];
         const uint32_t u = t - it.unit_base;
         if (u < it.n_large_units) return;
         const uint4* post4 = reinterpret_cast<const uint4*>(s.postings + it.post_begin);
         const uint32_t base = (u - it.n_large_units) * CHUNK4 + lane, n4 = it.n4;
#pragma unroll
         for (int k = 0; k < K1_UNROLL; ++k) dst[k] = base + k * 32 < n4 ? loadStream(post4 + base + k * 32) : none;
      };
      auto grab = [&]() {
         uint32_t next = 0;
         if (lane == 0) next = atomicAdd(&ctl->ticket, 1u);
         return __shfl_sync(0xFFFFFFFFu, next, 0);
      };
      uint32_t t = warp, cur = 0;
      if (t < total_units) locate(t, cur);
      loadChunk(t, cur, v);
      for (;;) {
         KS_CLOCK(c_top);
         while (helpFlush()) { }                        // tables first: a table that waits for its flush stalls the items behind it
         if (t >= total_units) {
            KS_CLOCK(c_idle);
            KS_ACCOUNT(7, c_top, c_idle);
            if (*flushed == n_items) break;
            __nanosleep(64);
            continue;
         }
         const KsItem it = items_s[cur];
         const uint32_t u = t - it.unit_base, tb = cur % n_tables;
         uint32_t* table = tables + size_t(tb) * table_words;
         KS_CLOCK(c_wait);
         KS_ACCOUNT(7, c_top, c_wait);
         while (owner[tb] != cur) { if (!helpFlush()) __nanosleep(32); }     // the table still belongs to item cur - n_tables
         uint32_t next, next_cur = cur;
         KS_CLOCK(c_unit);
         KS_ACCOUNT(4, c_wait, c_unit);
         if (u < it.n_large_units) {
            const LargeEntry e = s.large_entries[it.large_begin + (u >> 1)];
            const uint32_t half = u & 1u;
            const uint4* src = s.large_pool + e.offset16;
            const uint32_t k_kind = e.kind(), k_local = e.keyLocal(), k_size = e.size;
            int c2 = 0;
            if (k_kind == KIND_BITSET) {
               const uint4* fk = reinterpret_cast<const uint4*>(filter_s) + k_local * 512u;
               uint4 x[8];
#pragma unroll
               for (uint32_t q = 0; q < 8; ++q) x[q] = loadStream(src + (half * 8 + q) * 32 + lane);
#pragma unroll
               for (uint32_t q = 0; q < 8; ++q) {
                  const uint4 y = fk[(half * 8 + q) * 32 + lane];
                  c2 += __popc(x[q].x & y.x) + __popc(x[q].y & y.y) + __popc(x[q].z & y.z) + __popc(x[q].w & y.w);
               }
            } else if (half == 0 && k_kind == KIND_ARRAY) {
               const uint32_t* fk = filter_s + k_local * 2048u;
               const uint32_t n_vec = (k_size + 7) >> 3;
               for (uint32_t i0 = lane; i0 < n_vec; i0 += 128) {
                  uint4 x[4];
#pragma unroll
                  for (uint32_t k = 0; k < 4; ++k) x[k] = i0 + k * 32 < n_vec ? loadStream(src + i0 + k * 32) : make_uint4(0, 0, 0, 0);
#pragma unroll
                  for (uint32_t k = 0; k < 4; ++k) {
                     const uint32_t wv[4] = {x[k].x, x[k].y, x[k].z, x[k].w};
#pragma unroll
                     for (uint32_t q = 0; q < 8; ++q) {
                        const uint32_t val = (wv[q >> 1] >> ((q & 1) * 16)) & 0xFFFFu;
                        c2 += ((fk[val >> 5] >> (val & 31)) & 1u) & uint32_t((i0 + k * 32) * 8 + q < k_size);
                     }
                  }
               }
            } else if (half == 0) {   // long run container: two rank lookups per run
               const uint16_t* runs16 = reinterpret_cast<const uint16_t*>(src);
               const PartQuery q = pq[partition];
               for (uint32_t i = lane; i < k_size; i += 32) { const uint32_t start = __ldg(runs16 + 2 * i), len1 = __ldg(runs16 + 2 * i + 1); c2 += int(filterRangeCount(q, key0 + k_local, start, start + len1)); }
            }
            c2 = warpSum(c2);
            if (lane == 0 && c2 != 0) atomicAdd(&table[e.slot()], uint32_t(c2));
            next = grab();
            if (next < total_units) locate(next, next_cur);
            loadChunk(next, next_cur, v);
         } else {
            next = grab();
            if (next < total_units) locate(next, next_cur);
            uint4 nx[K1_UNROLL];
            loadChunk(next, next_cur, nx);
#pragma unroll
            for (int k = 0; k < K1_UNROLL; k += 2) {
               const uint32_t p[8] = {v[k].x, v[k].y, v[k].z, v[k].w, v[k + 1].x, v[k + 1].y, v[k + 1].z, v[k + 1].w};
               uint32_t f[8];
#pragma unroll
               for (int q = 0; q < 8; ++q) f[q] = filter_s[(p[q] >> 5) & (SLAB_WORDS32 - 1)];
#pragma unroll
               for (int q = 0; q < 8; ++q) {
                  const bool hit = (f[q] >> (p[q] & 31)) & 1u;
                  atomicAdd(&table[hit ? (p[q] >> POSTING_ROW_BITS) : tile_slots + lane], 1u);
               }
            }
#pragma unroll
            for (int k = 0; k < K1_UNROLL; ++k) v[k] = nx[k];
         }
         // ---- unit finished: the warp that completes the item puts its table up for flushing ----
         KS_CLOCK(c_done);
         KS_ACCOUNT(u < it.n_large_units ? 5 : 6, c_unit, c_done);
#ifndef SB_KS_NOFENCE
         __threadfence_block();
#else
         __syncwarp();
#endif
         if (lane == 0 && atomicAdd(&ctl->done[tb], 1u) + 1u == it.n_units) fstate[tb] = KS_PENDING;
         t = next;
         cur = next_cur;
      }
   }
#ifdef SB_K1_TIMELINE
   if (g_k1_timeline) {
      if (lane == 0) for (int k = 3; k < 8; ++k) atomicAdd(g_k1_timeline + size_t(blockIdx.x) * 8 + k, static_cast<unsigned long long>(acct[k]));
      __syncthreads();
      if (threadIdx.x == 0) g_k1_timeline[size_t(blockIdx.x) * 8 + 2] = globalTimerNs();
   }
#endif
}

}  // namespace sb
