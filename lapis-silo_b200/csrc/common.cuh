// Shared declarations of libsilo_b200.so: error slot, CUDA checks, device-side record layouts.
#pragma once
#include <cuda_runtime.h>

#include <chrono>
#include <cstdint>
#include <cstdlib>
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

// SB_TRACE=1: host-side lap times of the entry points on stderr (diagnostics only)
struct Trace {
   const char* name;
   bool on;
   std::chrono::steady_clock::time_point last;
   std::string text;
   explicit Trace(const char* n) : name(n), on(getenv("SB_TRACE") != nullptr), last(std::chrono::steady_clock::now()) {}
   void lap(const char* what) {
      if (!on) return;
      const auto now = std::chrono::steady_clock::now();
      text += std::string(" ") + what + "=" + std::to_string(std::chrono::duration<double, std::micro>(now - last).count()).substr(0, 7);
      last = now;
   }
   ~Trace() { if (on) fprintf(stderr, "[sb] %s:%s\n", name, text.c_str()); }
};

template <typename F>
int guarded(F&& f) {
   try { f(); return 0; }
   catch (const std::exception& e) { lastError() = e.what(); return 1; }
   catch (...) { lastError() = "unknown error"; return 1; }
}

// ---- container directory -------------------------------------------------------------------------
// Rows of a partition are cut into SLABS of 4 container keys (262,144 rows, a 32 KB slice of the dense
// filter bitset that the scan keeps in shared memory). One 8-byte entry per stored roaring container of a
// sequence, ordered by (position tile, slab, slot = pos*Sigma+sym, key). Entries are grouped into chunks of
// <= 32 (one warp) and STAGES of <= 16 chunks: the unit the scan streams into shared memory with bulk
// async copies. Payload location is not in the entry: it follows from the stage record's per-chunk offsets
// plus a warp prefix sum of the sizes.
enum : uint32_t { KIND_ARRAY = 0, KIND_BITSET = 1, KIND_RUN = 2, KIND_NULL = 3 };
constexpr uint32_t TILE_POSITIONS = 256;     // positions per tile (shared-memory count table = 256 * Sigma words)
constexpr uint32_t SLAB_KEYS = 4;            // container keys per slab
constexpr uint32_t SLAB_WORDS32 = SLAB_KEYS * 2048;   // 32-bit filter words per slab (32 KB)
constexpr uint32_t SMALL_ARRAY_MAX = 32;     // arrays up to this cardinality live in the staged (small) pool
constexpr uint32_t SMALL_RUN_MAX = 16;       // run containers up to this many runs live in the small pool
constexpr uint32_t STAGE_CHUNKS = 16;        // chunks per stage (512 entries = 4 KB of directory)
constexpr uint32_t STAGE_SMALL_U16 = 6144;   // small-payload capacity of a stage in uint16 units (12 KB)

struct EntryBits {
   // bits  0..15 key (high 16 bits of the row ids)      16..29 slot within the tile
   //      30..31 kind (3 = padding entry)               32..47 size-1 (array: cardinality, run: #runs; bitset: unused)
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
   // payload in the small pool (uint16 units) / in the large pool (16-byte units); exactly one is non-zero
   static __host__ __device__ inline uint32_t smallUnits(uint32_t kind, uint32_t size) {
      return kind == KIND_ARRAY ? (size <= SMALL_ARRAY_MAX ? size : 0u) : (kind == KIND_RUN && size <= SMALL_RUN_MAX ? 2u * size : 0u);
   }
   static __host__ __device__ inline uint32_t largeUnits(uint32_t kind, uint32_t size) {
      return kind == KIND_BITSET ? 512u : (kind == KIND_ARRAY ? (size > SMALL_ARRAY_MAX ? (size + 7u) / 8u : 0u)
                                                              : (kind == KIND_RUN && size > SMALL_RUN_MAX ? (size + 3u) / 4u : 0u));
   }
};

struct StageRec {            // 128 bytes per <= 512 entries; copied to shared memory with the stage
   uint32_t entry_start;     // index into entries (even)
   uint32_t n_entries;       // real entries (a padding entry follows when odd)
   uint64_t small_off16;     // small pool offset of the stage's payload, in 16-byte units
   uint32_t small_bytes;     // staged payload bytes (multiple of 16)
   uint32_t pad;
   uint64_t large_off16;     // large pool offset of the stage's first large container, in 16-byte units
   uint16_t chunk_small[STAGE_CHUNKS];   // uint16-unit offset of each chunk's payload inside the staged payload
   uint32_t chunk_large[STAGE_CHUNKS];   // 16-byte-unit offset of each chunk's large containers from large_off16
};
static_assert(sizeof(StageRec) == 128, "StageRec must be 128 bytes");

struct WorkItem {            // one CTA of the Mutations scan: a weight-bounded run of stages of one (tile, slab)
   uint32_t tile;
   uint32_t slab;
   uint32_t stage_begin;
   uint32_t stage_end;
};

struct SlabRec {
   uint32_t partition;       // db-local partition index
   uint32_t key0;            // first container key of the slab
};

struct SegRec {              // (tile, slab) segment: where its entries and stages start
   uint32_t entry_start;
   uint32_t stage_start;
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
   const StageRec* stages;
   const WorkItem* items;
   const SegRec* segs;          // [n_tiles * n_slabs + 1]
   const SlabRec* slabs;        // [n_slabs]
   const uint32_t* slab_begin;  // [P + 1] first slab of each partition
   const uint16_t* small_pool;  // small arrays / runs, staged through shared memory
   const uint4* large_pool;     // bitsets (512 x 16 B), large arrays and runs, 16-byte aligned each
   const int8_t* deleted;       // [P][L]
   const int8_t* flipped;       // [P][L]
   const uint32_t* miss_row_off;  // concatenated per partition: n_rows+1 offsets (relative to miss_iv_base[p])
   const uint2* miss_iv;          // [start, end) position intervals
   const uint64_t* miss_row_base; // [P] index of the partition's first row_off entry
   const uint64_t* miss_iv_base;  // [P]
   uint32_t length, sigma, tile_positions, n_tiles, n_partitions, n_items, n_stages, n_slabs;
};

}  // namespace sb
