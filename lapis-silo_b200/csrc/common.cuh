// Shared declarations of libsilo_b200.so: error slot, CUDA checks, device-side record layouts.
#pragma once
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <cstring>
#include <map>
#include <memory>
#include <mutex>
#include <stdexcept>
#include <string>
#include <vector>

#include "../../include/silo_b200.h"

namespace sb {

inline std::string& lastError() { thread_local std::string e; return e; }

struct Error : std::runtime_error { using std::runtime_error::runtime_error; };

#define SB_CUDA(call)                                                                                   \
   do {                                                                                                 \
      cudaError_t err__ = (call);                                                                       \
      if (err__ != cudaSuccess) {                                                                       \
         throw ::sb::Error(std::string(#call) + " failed: " + cudaGetErrorString(err__) + " (" + __FILE__ + ":" + std::to_string(__LINE__) + ")"); \
      }                                                                                                 \
   } while (0)

#define SB_REQUIRE(cond, msg) do { if (!(cond)) throw ::sb::Error(msg); } while (0)

template <typename F>
int guarded(F&& f) {
   try { f(); return 0; }
   catch (const std::exception& e) { lastError() = e.what(); return 1; }
   catch (...) { lastError() = "unknown error"; return 1; }
}

// ---- container directory -------------------------------------------------------------------------
// One 8-byte entry per stored roaring container of a sequence, ordered by
// (position tile, partition, slot = pos*Sigma+sym, key). Payload location is NOT in the entry: a warp
// owns one chunk of <= 32 consecutive entries and derives the offsets with a prefix sum of `size`.
enum : uint32_t { KIND_ARRAY = 0, KIND_BITSET = 1, KIND_RUN = 2 };

struct EntryBits {
   // bits  0..15 key (high 16 bits of the row ids)      16..29 slot within the tile
   //      30..31 kind                                   32..47 size-1 (array: cardinality, run: #runs; bitset: unused)
   //      48..52 deleted symbol of (partition, position), 31 = none      53 slot is the flipped symbol
   static __host__ __device__ inline uint64_t pack(uint32_t key, uint32_t slot_local, uint32_t kind, uint32_t size, uint32_t deleted, uint32_t flipped) {
      return uint64_t(key) | (uint64_t(slot_local) << 16) | (uint64_t(kind) << 30) | (uint64_t((size - 1) & 0xFFFF) << 32) |
             (uint64_t(deleted & 31) << 48) | (uint64_t(flipped & 1) << 53);
   }
   static __host__ __device__ inline uint32_t key(uint64_t e) { return uint32_t(e) & 0xFFFF; }
   static __host__ __device__ inline uint32_t slot(uint64_t e) { return (uint32_t(e) >> 16) & 0x3FFF; }
   static __host__ __device__ inline uint32_t kind(uint64_t e) { return uint32_t(e) >> 30; }
   static __host__ __device__ inline uint32_t size(uint64_t e) { return (uint32_t(e >> 32) & 0xFFFF) + 1; }
   static __host__ __device__ inline uint32_t deleted(uint64_t e) { return uint32_t(e >> 48) & 31; }
   static __host__ __device__ inline uint32_t flipped(uint64_t e) { return uint32_t(e >> 53) & 1; }
};

struct ChunkRec {          // 16 bytes per <= 32 entries
   uint32_t entry_start;   // index into Seq::entries
   uint32_t pay_off16;     // payload of the chunk's first array/run container, in 16-byte units of the array pool
   uint32_t bitset_idx;    // index of the chunk's first bitset container in the bitset pool (8 KB each)
   uint16_t partition;     // db-local partition index
   uint8_t n_entries;      // 1..32
   uint8_t pad;
};

struct WorkItem {          // one CTA of the Mutations scan: a weight-bounded run of chunks of one tile
   uint32_t tile;
   uint32_t chunk_begin;
   uint32_t chunk_end;
   uint32_t pad;
};

struct SegRec {            // (tile, partition) segment: where its entries and chunks start
   uint32_t entry_start;
   uint32_t chunk_base;
};

// Per-partition, per-query state the action kernels read.
enum : uint32_t { PART_EMPTY = 0, PART_PARTIAL = 1, PART_FULL = 2 };
struct PartQuery {
   const uint32_t* filter32;   // dense filter bitset (padded to whole 65,536-row blocks)
   const uint32_t* rank_local; // per 64-bit word: popcount of the preceding words of its 256-word block
   const uint32_t* rank_block; // per 256-word block: popcount of all preceding blocks
   uint32_t state;
   uint32_t cardinality;
   uint32_t n_rows;
   uint32_t pad;
};

// Device view of one logical sequence (all partitions of one store name on this GPU).
struct SeqDev {
   const uint64_t* entries;
   const ChunkRec* chunks;
   const WorkItem* items;
   const SegRec* segs;          // [n_tiles * P + 1]
   const uint16_t* arrays;      // array + run payloads
   const uint4* bitsets;        // 512 uint4 per bitset container
   const int8_t* deleted;       // [P][L]
   const int8_t* flipped;       // [P][L]
   const uint32_t* miss_row_off;  // concatenated per partition: n_rows+1 offsets (relative to miss_iv_base[p])
   const uint2* miss_iv;          // [start, end) position intervals
   const uint64_t* miss_row_base; // [P] index of the partition's first row_off entry
   const uint64_t* miss_iv_base;  // [P]
   uint32_t length, sigma, tile_positions, n_tiles, n_partitions, n_items, n_chunks, pad;
};

}  // namespace sb
