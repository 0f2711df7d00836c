// Shared declarations of libsilo_b200.so: error slot, CUDA checks, device-side record layouts.
#pragma once
#include <cuda_runtime.h>

#include <chrono>
#include <condition_variable>
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

// ---- posting layout ---------------------------------------------------------------------------------
// Rows of a partition are cut into SLABS of 4 container keys (262,144 rows, a 32 KB slice of the dense
// filter bitset that the scan keeps in shared memory). The stored roaring containers of a sequence are
// grouped by (position tile, slab) SEGMENTS. Inside a segment
//   * containers (array or run) of cardinality <= FLAT_MAX -- 99.9 % of them hold at most 16 values in real data: a handful of
//     rows carrying a private mutation -- are flattened into 4-byte POSTINGS
//         slot':14 << 18 | row_in_slab:18          slot' = symbol * tile_positions + position_in_tile
//     sorted ascending; the scan streams them with 128-bit loads, one probe of the shared-memory filter each;
//   * everything else (bitsets, big arrays, long run containers) keeps its roaring payload in the large pool
//     (16-byte aligned) behind a 16-byte LargeEntry, sorted by (slot', key).
// A segment's postings start 16-byte aligned; the gap is filled with postings of the dummy slot
// `tile_slots` (never flushed), so the scan needs no tail handling.
enum : uint32_t { KIND_ARRAY = 0, KIND_BITSET = 1, KIND_RUN = 2, KIND_NULL = 3 };
constexpr uint32_t TILE_POSITIONS = 256;     // positions per tile (shared-memory count table = 256 * Sigma words); SB_TILE_POSITIONS: tuning
constexpr uint32_t MAX_TILE_POSITIONS = 512;
constexpr uint32_t SLAB_KEYS = 4;            // container keys per slab
constexpr uint32_t SLAB_WORDS32 = SLAB_KEYS * 2048;   // 32-bit filter words per slab (32 KB)
constexpr uint32_t FLAT_MAX = 256;           // containers up to this cardinality become postings (pool.cuh flatMax)
constexpr uint32_t POSTING_ROW_BITS = 18;
constexpr uint32_t POSTING_ROW_MASK = (1u << POSTING_ROW_BITS) - 1u;

struct LargeEntry {
   uint32_t offset16;        // payload offset in the large pool, 16-byte units
   uint32_t size;            // array: cardinality, run: number of runs, bitset: cardinality (unused by the scan)
   uint32_t where;           // slot':14 | kind:2 << 14 | key_local:2 << 16 | tile:14 << 18
   uint32_t slab;            // the (tile, slab) segment it belongs to: a unit of the large-container scan finds its partition by itself
   __host__ __device__ inline uint32_t slot() const { return where & 0x3FFFu; }
   __host__ __device__ inline uint32_t kind() const { return (where >> 14) & 3u; }
   __host__ __device__ inline uint32_t keyLocal() const { return (where >> 16) & 3u; }
   __host__ __device__ inline uint32_t tile() const { return where >> 18; }
};
static_assert(sizeof(LargeEntry) == 16, "LargeEntry must be 16 bytes");

struct WorkItem {            // one unit of the Mutations scan: a weight-bounded part of one (tile, slab) segment
   uint32_t tile;
   uint32_t slab;
   uint32_t post_begin, post_end;     // postings [begin, end), begin a multiple of 4
   uint32_t large_begin, large_end;   // large entries
   uint32_t partition, key0;          // of the slab (saves the scan a dependent load)
};
static_assert(sizeof(WorkItem) == 32, "WorkItem must be 32 bytes");

// What a CTA of the scan takes from the hand-out counter: consecutive work items (ascending tile) of ONE slab, so that the
// 32 KB filter slice of the slab is copied into shared memory once per run instead of once per item.
struct RunRec {
   uint32_t first_item, n_items;      // items [first_item, first_item + n_items) of SeqDev::items
   uint32_t partition, weight_kb;
};
static_assert(sizeof(RunRec) == 16, "RunRec must be 16 bytes");

struct SlabRec {
   uint32_t partition;       // db-local partition index
   uint32_t key0;            // first container key of the slab
};

struct SegRec {              // (tile, slab) segment: where its postings and large entries start
   uint32_t post_start;
   uint32_t large_start;
};

// Per-partition, per-query state the action kernels read. Written on the DEVICE (classify_partitions): a query is one
// submission, the host never sees the filter cardinalities before the count kernels are enqueued.
enum : uint32_t { PART_EMPTY = 0, PART_PARTIAL = 1, PART_FULL = 2 };
struct PartQuery {
   const uint32_t* filter32;   // dense filter bitset (padded to whole 4-key slabs)
   const uint32_t* rank_local; // per 64-bit word: popcount of the preceding words of its 256-word block
   const uint32_t* rank_block; // per 256-word block: popcount of all preceding blocks
   uint32_t state;
   uint32_t cardinality;
   uint32_t n_rows;
   uint32_t pad;
};
// What the host knows of a partition's filter result when it enqueues the action: where things are, not what they hold.
struct PartInput {
   const uint32_t* filter32;
   const uint32_t* rank_local;
   const uint32_t* rank_block;
   const uint64_t* cardinality;   // device counter written by finish_filters
   uint32_t partition;
   uint32_t pad;
};
// Per-(query, sequence) header the classification leaves for the count kernels.
struct ActionHead {
   uint32_t n_partial, n_full, produced, n_selected;   // n_selected: runs of work items of the partial partitions (compact_runs)
   uint32_t next_run, pad0;                            // hand-out counter of the container scan
   uint64_t k2_total_cost;
};

// Device view of one logical sequence (all partitions of one store name on this GPU).
struct SeqDev {
   const uint32_t* postings;
   const LargeEntry* large_entries;
   const WorkItem* items;       // ordered by (slab, tile): the items of a run are consecutive
   const RunRec* runs;          // heaviest first over all partitions (the scan hands them out in this order)
   const SegRec* segs;          // [n_tiles * n_slabs + 1]
   const SlabRec* slabs;        // [n_slabs]
   const uint32_t* slab_begin;  // [P + 1] first slab of each partition
   const uint4* large_pool;     // bitsets (512 x 16 B), large arrays and runs, 16-byte aligned each
   const int8_t* deleted;       // [P][L]
   const int8_t* flipped;       // [P][L]
   const uint32_t* miss_row_off;  // concatenated per partition: n_rows+1 offsets (relative to miss_iv_base[p])
   const uint2* miss_iv;          // [start, end) position intervals
   const uint64_t* miss_row_base; // [P] index of the partition's first row_off entry
   const uint64_t* miss_iv_base;  // [P]
   // One 32-byte record per row with the row's first intervals inline, so that a kernel that needs the missing positions of a
   // row issues ONE load instead of walking row_off -> intervals (dependent loads, two or three scattered sectors):
   //   length <= 65,535 (miss_narrow): eight words start | end << 16, unused ones 0; otherwise four (start, end) word pairs.
   //   A row with more intervals has every word 0xFFFFFFFF: its intervals are read from the CSR arrays above.
   const uint4* miss_inline;         // [2 per row]
   const uint64_t* miss_inline_base; // [P] index of the partition's first row
   uint32_t miss_narrow, pad_inline;
   const uint32_t* miss_any;         // per partition one bit per row: the row misses at least one position (a short gene: a few percent of the rows)
   const uint64_t* miss_any_base;    // [P] word offset of the partition's bits
   const uint32_t* miss_tile_bits;   // per partition [n_tiles][ceil(n_rows / 32)]: bit = the row misses some position of the tile
   const uint64_t* miss_tile_base;   // [P] word offset of the partition's block
   const uint32_t* part_rows;        // [P] sequence_count of the partition's store, 0 = the partition has no store of this sequence
   uint32_t length, sigma, tile_positions, n_tiles, n_partitions, n_items, n_slabs, n_runs;
};

}  // namespace sb
