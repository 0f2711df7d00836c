// k1_scan: the fused AND-popcount of the Mutations action (mutations.cpp:85-95, M2) as a persistent,
// warp-specialised pipeline -- one CTA per SM that owns all of its shared memory:
//
//   producer warp   walks the CTA's work items (longest first). Per item it bulk-copies the 32 KB filter slab
//                   (cp.async.bulk + mbarrier), then publishes the item's payload as STAGES of <= 2 KB into a
//                   32-slot shared-memory ring, again with bulk copies: postings (512 per stage), quarters of
//                   bitset containers, array containers up to 1024 values. HBM requests are up to 32 stages
//                   deep, independent of how the consumer warps are scheduled; no register holds data in flight.
//   consumer warps  take stages in ticket order (shared counter), wait on the stage's full barrier, probe the
//                   shared-memory filter (postings, arrays) or AND-popcount against it (bitsets), add into the
//                   item's shared-memory count table and release the slot. Nothing a consumer touches is in
//                   global memory except long run containers and arrays > 2048 values (read in place).
//   The filter slab and the count table are double-buffered: warps drift into the next item while the last
//   stages of the previous one are still being processed. The warp that completes an item's last stage flushes
//   the item's counts (applying the flipped / deleted-symbol arithmetic once per slot) and frees the buffers.
// There is no CTA-wide barrier after the prologue.
#pragma once
#include "kernels_action.cuh"

namespace sb {

constexpr uint32_t KS_THREADS = 1024;                 // 31 consumer warps + 1 producer warp
constexpr uint32_t KS_CONSUMER_WARPS = KS_THREADS / 32 - 1;
// Ring: a consumer waits for "phase (ticket / KS_SLOTS) of slot ticket % KS_SLOTS" (see the note in ksConsume).
constexpr uint32_t KS_SLOTS = 32;                     // ring slots
constexpr uint32_t KS_SLOT_BYTES = 2048;
constexpr uint32_t KS_SLOT_U4 = KS_SLOT_BYTES / 16;   // 128 x 16 bytes
constexpr uint32_t KS_BATCH = 16;                     // stages the producer warp publishes per round (half the ring: refill overlaps consumption)
enum : uint32_t { KS_POSTINGS = 0, KS_BITSET_PART = 1, KS_ARRAY = 2, KS_DIRECT = 3, KS_NOP = 4, KS_EXIT = 5 };

struct KsHeader {            // 40 bytes, written by the producer before it publishes the stage
   uint32_t kind;
   uint32_t buf;             // filter / count buffer of the stage's item (item sequence number & 1)
   uint32_t item_seq;        // sequence number of the item inside this CTA
   uint32_t n16;             // valid 16-byte units in the slot
   uint32_t slot;            // large containers: slot' (sym * T + pos_in_tile)
   uint32_t key_local;       // large containers: key inside the slab;  bitset: | quarter << 8
   uint32_t size;            // array: cardinality; run: number of runs
   uint32_t offset16;        // KS_DIRECT: payload offset in the large pool
   uint32_t stage;           // number of the stage in this CTA's sequence (= the consumer's ticket)
   uint32_t pad;
};
struct KsItemInfo { uint32_t tile, partition, key0, n_stages; };

// shared memory: [filter 2 x 32 KB][ring 16 x 4 KB][counts 2 x stride][headers][item info][barriers][counters]
__host__ __device__ inline uint32_t ksCountStride(uint32_t tile_slots) { return (tile_slots + 4u) & ~3u; }   // words, includes the dummy slot
inline size_t ksSharedBytes(uint32_t tile_slots) {
   return 2 * K1_FILTER_BYTES + KS_SLOTS * KS_SLOT_BYTES + 2 * size_t(ksCountStride(tile_slots)) * 4 + KS_SLOTS * sizeof(KsHeader) + 2 * sizeof(KsItemInfo) +
          (2 * KS_SLOTS + 4) * 8 + 16 + 32 * 4;
}

__device__ __forceinline__ void mbarArrive(uint64_t* bar) {
   asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smemAddr(bar)) : "memory");
}

// mbarrier wait with a watchdog: a pipeline bug must show up as an error, not as a hung GPU
__device__ __forceinline__ void ksWait(uint64_t* bar, uint32_t parity, int tag, uint32_t a, uint32_t b, const uint32_t* dbg = nullptr) {
   uint32_t done = 0;
   for (uint32_t spins = 0; !done; ++spins) {
      asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}\n" : "=r"(done) : "r"(smemAddr(bar)), "r"(parity) : "memory");
      if (!done && spins > (1u << 22)) {
         printf("k1_scan stuck: tag %d block %u warp %u lane %u parity %u a %u b %u\n", tag, blockIdx.x, threadIdx.x >> 5, threadIdx.x & 31, parity, a, b);
         if (dbg) { for (int k = 0; k < 31; ++k) printf("  block %u consumer %d: ticket %u state %u\n", blockIdx.x, k, dbg[k] >> 4, dbg[k] & 15u); printf("  block %u next_ticket %u done0 %u done1 %u\n", blockIdx.x, dbg[-4], dbg[-3], dbg[-2]); }
         __trap();
      }
   }
}

#ifndef KS_DEBUG
#define KS_DEBUG 0
#endif
#define KS_CHECK(cond, tag, a, b) do { if (KS_DEBUG && !(cond)) { printf("k1_scan check %d failed: block %u warp %u lane %u a %u b %u\n", tag, blockIdx.x, threadIdx.x >> 5, threadIdx.x & 31, uint32_t(a), uint32_t(b)); __trap(); } } while (0)

struct KsShared {
   uint32_t* filter;       // [2][SLAB_WORDS32]
   uint4* ring;            // [KS_SLOTS][KS_SLOT_U4]
   uint32_t* counts;       // [2][stride]
   KsHeader* headers;      // [KS_SLOTS]
   KsItemInfo* info;       // [2]
   uint64_t* full;         // [KS_SLOTS]   producer -> consumers (expect_tx of the slot's bulk copy)
   uint64_t* empty;        // [KS_SLOTS]   consumer -> producer
   uint64_t* filter_full;  // [2]
   uint64_t* buf_free;     // [2]          item finished (counts flushed and zeroed) -> producer
   uint32_t* next_ticket;  // consumer ticket counter
   uint32_t* stages_done;  // [2]
   uint32_t* wstate;       // [32] debug: ticket << 4 | state of every consumer warp
};

__device__ __forceinline__ KsShared ksCarve(uint8_t* base, uint32_t tile_slots) {
   KsShared sh;
   const uint32_t stride = ksCountStride(tile_slots);
   sh.filter = reinterpret_cast<uint32_t*>(base);
   sh.ring = reinterpret_cast<uint4*>(base + 2 * K1_FILTER_BYTES);
   sh.counts = reinterpret_cast<uint32_t*>(base + 2 * K1_FILTER_BYTES + KS_SLOTS * KS_SLOT_BYTES);
   sh.headers = reinterpret_cast<KsHeader*>(sh.counts + 2 * stride);
   sh.info = reinterpret_cast<KsItemInfo*>(sh.headers + KS_SLOTS);
   sh.full = reinterpret_cast<uint64_t*>(sh.info + 2);
   sh.empty = sh.full + KS_SLOTS;
   sh.filter_full = sh.empty + KS_SLOTS;
   sh.buf_free = sh.filter_full + 2;
   sh.next_ticket = reinterpret_cast<uint32_t*>(sh.buf_free + 2);
   sh.stages_done = sh.next_ticket + 1;
   sh.wstate = sh.stages_done + 3;
   return sh;
}

// ---- producer ------------------------------------------------------------------------------------------
// One warp; every lane publishes its own stages (lane-parallel: a single lane issuing one stage at a time was the
// bottleneck of the whole pipeline). Stage numbers are handed out with warp prefix sums, so the ring order is fixed
// no matter in which order the lanes get their slots.
__device__ __forceinline__ void ksPublish(const KsShared& sh, uint32_t n, const KsHeader& hdr, const void* src, uint32_t bytes) {
   const uint32_t slot = n % KS_SLOTS;
   if (n >= KS_SLOTS) ksWait(&sh.empty[slot], ((n / KS_SLOTS) - 1u) & 1u, 1, n, slot, sh.wstate);
   sh.headers[slot] = hdr;
   sh.headers[slot].stage = n;
   if (src != nullptr && bytes != 0) {
      mbarExpectTx(&sh.full[slot], bytes);   // arrives and arms the transaction count; releases the header
      bulkLoad(sh.ring + size_t(slot) * KS_SLOT_U4, src, bytes, &sh.full[slot]);
   } else {
      mbarArrive(&sh.full[slot]);
   }
}

__device__ __forceinline__ uint32_t warpExclusiveSum(uint32_t v, uint32_t& total) {
   uint32_t incl = v;
#pragma unroll
   for (int d = 1; d < 32; d <<= 1) { const uint32_t t = __shfl_up_sync(0xFFFFFFFFu, incl, d); if (laneId() >= uint32_t(d)) incl += t; }
   total = __shfl_sync(0xFFFFFFFFu, incl, 31);
   return incl - v;
}

__device__ __forceinline__ void ksProduce(const SeqDev& s, const PartQuery* __restrict__ pq, const K1Map& map, uint32_t n_items, const KsShared& sh) {
   const uint32_t lane = laneId();
   uint32_t n_published = 0;   // warp-uniform
   uint32_t seq = 0;
   WorkItem next_item{};
   if (blockIdx.x < n_items) next_item = s.items[__ldg(map.item_list + blockIdx.x)];
   for (uint32_t t = blockIdx.x; t < n_items; t += gridDim.x, ++seq) {
      const WorkItem item = next_item;
      if (t + gridDim.x < n_items) next_item = s.items[__ldg(map.item_list + t + gridDim.x)];   // in flight while this item is published
      const uint32_t b = seq & 1u;
      const uint32_t n_post4 = (item.post_end - item.post_begin + 3u) >> 2;
      const uint32_t n_large = item.large_end - item.large_begin;
      if (lane == 0) {
         if (seq >= 2) ksWait(&sh.buf_free[b], ((seq >> 1) - 1u) & 1u, 2, seq, t);   // item seq-2 flushed, its filter and counts are free
         sh.info[b] = {item.tile, item.partition, item.key0, 0xFFFFFFFFu};   // the stage count follows with the item's last stage
         const uint32_t* filter32 = item.partition < K1_INLINE_PARTITIONS ? map.inline_filter[item.partition] : pq[item.partition].filter32;
         mbarExpectTx(&sh.filter_full[b], K1_FILTER_BYTES);
         bulkLoad(sh.filter + size_t(b) * SLAB_WORDS32, filter32 + size_t(item.key0) * 2048u, K1_FILTER_BYTES, &sh.filter_full[b]);
      }
      __syncwarp();
      const uint32_t first_stage = n_published;
      // large containers first (the longest stages): a bitset is four stages, anything else one. Eight containers per
      // round, one (container, part) per lane: the <= 32 stages of a round occupy distinct slots and wait only for
      // stages of earlier rounds (lanes run in lockstep: a lane must never wait for a stage of its own round).
      constexpr uint32_t PARTS = 8192 / KS_SLOT_BYTES;
      for (uint32_t e0 = 0; e0 < n_large; e0 += KS_BATCH / PARTS) {
         const uint32_t e = e0 + lane / PARTS, part = lane % PARTS;
         LargeEntry w{};
         const bool have = lane < KS_BATCH && e < n_large;
         if (have) w = s.large_entries[item.large_begin + e];
         const bool mine = have && (part == 0 || w.kind() == KIND_BITSET);
         const uint32_t active = __ballot_sync(0xFFFFFFFFu, mine);
         if (mine) {
            const uint32_t n = n_published + __popc(active & ((1u << lane) - 1u));
            KsHeader hdr{KS_NOP, b, seq, 0, w.slot(), w.keyLocal(), w.size, w.offset16};
            const uint4* src = s.large_pool + w.offset16;
            if (w.kind() == KIND_BITSET) {
               hdr.kind = KS_BITSET_PART; hdr.n16 = KS_SLOT_U4; hdr.key_local = w.keyLocal() | (part << 8);
               ksPublish(sh, n, hdr, src + part * KS_SLOT_U4, KS_SLOT_BYTES);
            } else if (w.kind() == KIND_ARRAY && w.size <= KS_SLOT_BYTES / 2) {
               hdr.kind = KS_ARRAY; hdr.n16 = (w.size + 7u) / 8u;
               ksPublish(sh, n, hdr, src, hdr.n16 * 16u);
            } else {
               hdr.kind = KS_DIRECT | (w.kind() << 8);
               ksPublish(sh, n, hdr, nullptr, 0);
            }
         }
         n_published += __popc(active);
         __syncwarp();
      }
      const uint4* post4 = reinterpret_cast<const uint4*>(s.postings + item.post_begin);
      const uint32_t n_chunks = (n_post4 + KS_SLOT_U4 - 1) / KS_SLOT_U4;
      for (uint32_t c0 = 0; c0 < n_chunks; c0 += KS_BATCH) {
         const uint32_t c = c0 + lane;
         if (lane < KS_BATCH && c < n_chunks) {
            const uint32_t n16 = min(KS_SLOT_U4, n_post4 - c * KS_SLOT_U4);
            KsHeader hdr{KS_POSTINGS, b, seq, n16, 0, 0, n_published + lane, item.post_begin / 4 + c * KS_SLOT_U4};
            ksPublish(sh, n_published + lane, hdr, post4 + size_t(c) * KS_SLOT_U4, n16 * 16u);
         }
         n_published += min(KS_BATCH, n_chunks - c0);
         __syncwarp();
      }
      // end of item: the stage count becomes known to the consumers with this last (empty) stage
      if (lane == 0) {
         sh.info[b].n_stages = n_published - first_stage + 1u;
         KsHeader end_hdr{KS_NOP, b, seq, 0, 0, 0, 0, 0};
         ksPublish(sh, n_published, end_hdr, nullptr, 0);
      }
      ++n_published;
      __syncwarp();
   }
   if (lane < KS_CONSUMER_WARPS) {
      KsHeader hdr{KS_EXIT, 0, 0, 0, 0, 0, 0, 0};
      ksPublish(sh, n_published + lane, hdr, nullptr, 0);
   }
}

// ---- consumers -----------------------------------------------------------------------------------------
// Flush of one finished item by one warp: lane = position of the tile; symbol rows of counts[sym][pos] get
// coalesced REDs; the table is zeroed on the way (it is reused two items later).
__device__ __forceinline__ void ksFlush(const SeqDev& s, const KsItemInfo info, uint32_t* tile_counts, uint32_t* __restrict__ counts, int full_mode) {
   constexpr uint32_t T = TILE_POSITIONS, ROUNDS = TILE_POSITIONS / 32;
   const uint32_t lane = laneId();
   int del[ROUNDS], flip[ROUNDS];
#pragma unroll
   for (uint32_t r = 0; r < ROUNDS; ++r) {   // all flag loads first: one memory round trip instead of eight
      const uint32_t pos = info.tile * T + r * 32 + lane;
      del[r] = flip[r] = -1;
      if (pos < s.length) { del[r] = s.deleted[size_t(info.partition) * s.length + pos]; flip[r] = s.flipped[size_t(info.partition) * s.length + pos]; }
   }
#pragma unroll
   for (uint32_t r = 0; r < ROUNDS; ++r) {
      const uint32_t pl = r * 32 + lane, pos = info.tile * T + pl;
      const bool live = pos < s.length;
      int acc_deleted = 0;
      for (uint32_t sym = 0; sym < s.sigma; ++sym) {
         const int raw = int(tile_counts[sym * T + pl]);
         if (raw == 0) continue;
         tile_counts[sym * T + pl] = 0;
         if (!live || int(sym) == del[r]) continue;
         const int add_slot = int(sym) == flip[r] ? -raw : raw;
         atomicAdd(counts + size_t(sym) * s.length + pos, uint32_t(add_slot));
         acc_deleted += full_mode ? -raw : -add_slot;
      }
      if (del[r] >= 0 && acc_deleted != 0) atomicAdd(counts + size_t(del[r]) * s.length + pos, uint32_t(acc_deleted));
   }
   if (lane == 0) tile_counts[T * s.sigma] = 0;   // the dummy slot of the filler postings
}

__device__ __forceinline__ void ksConsume(const SeqDev& s, const PartQuery* __restrict__ pq, const KsShared& sh, uint32_t* __restrict__ counts, int full_mode) {
   const uint32_t lane = laneId();
   const uint32_t stride = ksCountStride(s.tile_positions * s.sigma);
   for (;;) {
      uint32_t ticket = 0;
      if (lane == 0) ticket = atomicAdd(sh.next_ticket, 1u);
      ticket = __shfl_sync(0xFFFFFFFFu, ticket, 0);
      const uint32_t slot = ticket % KS_SLOTS;
      if (KS_DEBUG && lane == 0) sh.wstate[threadIdx.x >> 5] = (ticket << 4) | 1u;
      // The parity test of an mbarrier tells the current phase from the previous one only, and a warp can hold a ticket
      // a whole round (or more) ahead of the stage that currently occupies the slot. So it first waits until the
      // producer has written THIS stage's header -- which it does only after the previous occupant was released, i.e.
      // when the slot's full barrier is in this stage's phase -- and only then waits on the barrier.
      if (lane == 0) {
         const volatile uint32_t* stage_word = &sh.headers[slot].stage;
         for (uint32_t spins = 0; *stage_word != ticket; ++spins) {
            __nanosleep(20);
            if (spins > (1u << 24)) { printf("k1_scan stuck: header of ticket %u block %u warp %u sees %u\n", ticket, blockIdx.x, threadIdx.x >> 5, *stage_word); __trap(); }
         }
      }
      __syncwarp();
      ksWait(&sh.full[slot], (ticket / KS_SLOTS) & 1u, 3, ticket, slot);
      const KsHeader hdr = sh.headers[slot];
      const uint32_t kind = hdr.kind & 0xFFu;
      if (KS_DEBUG && lane == 0) sh.wstate[threadIdx.x >> 5] = (ticket << 4) | (2u + kind);
      if (kind == KS_EXIT) { __syncwarp(); if (lane == 0) mbarArrive(&sh.empty[slot]); break; }
      const uint32_t b = hdr.buf;
      KS_CHECK(b < 2 && kind <= KS_EXIT && hdr.slot < s.tile_positions * s.sigma && (hdr.key_local & 0xFFu) < 4, 12, hdr.kind, hdr.slot);
      const uint32_t* filter_s = sh.filter + size_t(b) * SLAB_WORDS32;
      uint32_t* tile_counts = sh.counts + size_t(b) * stride;
      const uint4* data = sh.ring + size_t(slot) * KS_SLOT_U4;
      if (kind != KS_NOP) ksWait(&sh.filter_full[b], (hdr.item_seq >> 1) & 1u, 4, hdr.item_seq, ticket);   // (an already completed phase returns at once)
      if (kind == KS_POSTINGS) {
         auto probe = [&](uint32_t p) {
            KS_CHECK((p >> POSTING_ROW_BITS) <= s.tile_positions * s.sigma, 10, p, ticket);
            if ((filter_s[(p >> 5) & (SLAB_WORDS32 - 1)] >> (p & 31)) & 1u) atomicAdd(&tile_counts[p >> POSTING_ROW_BITS], 1u);
         };
         KS_CHECK(hdr.n16 <= KS_SLOT_U4, 11, hdr.n16, ticket);
         KS_CHECK(hdr.size == ticket, 13, hdr.size, ticket);
         if (KS_DEBUG && lane < hdr.n16) { const uint4 want = __ldg(reinterpret_cast<const uint4*>(s.postings) + hdr.offset16 + lane), got = data[lane]; KS_CHECK(want.x == got.x && want.w == got.w, 14, got.x, want.x); }
         uint4 v[4];
#pragma unroll
         for (uint32_t k = 0; k < 4; ++k) { const uint32_t i = k * 32 + lane; if (i < hdr.n16) v[k] = data[i]; }
#pragma unroll
         for (uint32_t k = 0; k < 4; ++k) {
            if (k * 32 + lane < hdr.n16) { probe(v[k].x); probe(v[k].y); probe(v[k].z); probe(v[k].w); }
         }
      } else if (kind == KS_BITSET_PART) {
         const uint32_t k_local = hdr.key_local & 0xFFu, part = hdr.key_local >> 8;
         const uint4* fk = reinterpret_cast<const uint4*>(filter_s) + k_local * 512u + part * KS_SLOT_U4;
         int c2 = 0;
#pragma unroll
         for (uint32_t t = 0; t < KS_SLOT_U4 / 32; ++t) {
            const uint4 x = data[t * 32 + lane], y = fk[t * 32 + lane];
            c2 += __popc(x.x & y.x) + __popc(x.y & y.y) + __popc(x.z & y.z) + __popc(x.w & y.w);
         }
         c2 = warpSum(c2);
         if (lane == 0 && c2 != 0) atomicAdd(&tile_counts[hdr.slot], uint32_t(c2));
      } else if (kind == KS_ARRAY) {
         const uint32_t* fk = filter_s + (hdr.key_local & 0xFFu) * 2048u;
         int c2 = 0;
         for (uint32_t i = lane; i < hdr.n16; i += 32) {
            const uint4 x = data[i];
            const uint32_t wv[4] = {x.x, x.y, x.z, x.w};
#pragma unroll
            for (uint32_t t = 0; t < 8; ++t) {
               const uint32_t val = (wv[t >> 1] >> ((t & 1) * 16)) & 0xFFFFu;
               if (i * 8 + t < hdr.size) c2 += (fk[val >> 5] >> (val & 31)) & 1u;
            }
         }
         c2 = warpSum(c2);
         if (lane == 0 && c2 != 0) atomicAdd(&tile_counts[hdr.slot], uint32_t(c2));
      } else if (kind == KS_DIRECT) {   // arrays beyond 2048 values and long run containers, read in place
         const uint32_t c_kind = hdr.kind >> 8, k_local = hdr.key_local & 0xFFu;
         const uint4* src = s.large_pool + hdr.offset16;
         int c2 = 0;
         if (c_kind == KIND_ARRAY) {
            const uint32_t* fk = filter_s + k_local * 2048u;
            for (uint32_t i = lane; i * 8 < hdr.size; i += 32) {
               const uint4 x = __ldg(src + i);
               const uint32_t wv[4] = {x.x, x.y, x.z, x.w};
#pragma unroll
               for (uint32_t t = 0; t < 8; ++t) {
                  const uint32_t val = (wv[t >> 1] >> ((t & 1) * 16)) & 0xFFFFu;
                  if (i * 8 + t < hdr.size) c2 += (fk[val >> 5] >> (val & 31)) & 1u;
               }
            }
         } else {   // two rank lookups per run
            const KsItemInfo info = sh.info[b];
            const uint16_t* runs = reinterpret_cast<const uint16_t*>(src);
            const PartQuery q = pq[info.partition];
            for (uint32_t i = lane; i < hdr.size; i += 32) { const uint32_t start = __ldg(runs + 2 * i), len1 = __ldg(runs + 2 * i + 1); c2 += int(filterRangeCount(q, info.key0 + k_local, start, start + len1)); }
         }
         c2 = warpSum(c2);
         if (lane == 0 && c2 != 0) atomicAdd(&tile_counts[hdr.slot], uint32_t(c2));
      }
      // release the slot; the warp that completes the item's last stage flushes it
      __syncwarp();
      uint32_t last = 0;
      if (lane == 0) {
         mbarArrive(&sh.empty[slot]);
         __threadfence_block();
         const uint32_t done = atomicAdd(&sh.stages_done[b], 1u) + 1u;
         __threadfence_block();
         last = done == sh.info[b].n_stages;
      }
      last = __shfl_sync(0xFFFFFFFFu, last, 0);
      if (last) {
         if (KS_DEBUG && lane == 0) sh.wstate[threadIdx.x >> 5] = (ticket << 4) | 9u;
         ksFlush(s, sh.info[b], tile_counts, counts, full_mode);
         __syncwarp();
         if (lane == 0) { sh.stages_done[b] = 0; __threadfence_block(); mbarArrive(&sh.buf_free[b]); }
      }
   }
}

__global__ void __launch_bounds__(KS_THREADS, 1)
k1_scan(const SeqDev s, const PartQuery* __restrict__ pq, const K1Map map, uint32_t n_items, uint32_t* __restrict__ counts, int full_mode) {
   extern __shared__ __align__(128) uint8_t ks_smem[];
   const uint32_t tile_slots = s.tile_positions * s.sigma;
   const KsShared sh = ksCarve(ks_smem, tile_slots);
   if (threadIdx.x == 0) {
      for (uint32_t k = 0; k < KS_SLOTS; ++k) { mbarInit(&sh.full[k], 1); mbarInit(&sh.empty[k], 1); }
      for (uint32_t k = 0; k < 2; ++k) { mbarInit(&sh.filter_full[k], 1); mbarInit(&sh.buf_free[k], 1); }
      *sh.next_ticket = 0;
      for (uint32_t k = 0; k < KS_SLOTS; ++k) sh.headers[k].stage = 0xFFFFFFFFu;
      sh.stages_done[0] = 0; sh.stages_done[1] = 0;
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
   }
   for (uint32_t i = threadIdx.x; i < 2 * ksCountStride(tile_slots); i += KS_THREADS) sh.counts[i] = 0;
   __syncthreads();
   if (threadIdx.x >= KS_THREADS - 32) ksProduce(s, pq, map, n_items, sh);
   else ksConsume(s, pq, sh, counts, full_mode);
}

}  // namespace sb
