// Group-by kernels (sm_100a): Aggregated with groupByFields.
//   The reference (src/silo/query_engine/actions/aggregated.cpp:100-149) walks the filter bitmap of every partition row by
//   row, builds the tuple of the row's group-by values (actions/tuple.cpp:29-76) and counts tuples in a hash map per
//   partition, then merges the maps. Here the group-by columns are dictionary-coded at ingest (a tuple is a mixed-radix key
//   over the columns' codes), so counting is a filtered histogram:
//   group_count_dense    key space up to GROUP_DENSE_MAX: one uint32 bin per key; key spaces up to GROUP_SMEM_BINS are
//                        counted in a shared-memory copy per CTA first (a few hot bins -- lineages, countries -- would
//                        otherwise serialise in L2)
//   group_count_hash     larger key spaces: open addressing over 64-bit keys (atomicCAS claim + atomicAdd count)
//   group_compact_*      non-empty bins -> {key, count} rows
// Counts are uint32 like the reference's map values (aggregated.cpp:104).
#pragma once
#include "common.cuh"

namespace sb {

constexpr uint32_t GROUP_MAX_COLUMNS = 16;
constexpr uint64_t GROUP_DENSE_MAX = 1ull << 24;      // bins of the dense histogram (64 MB)
constexpr uint32_t GROUP_SMEM_BINS = 8192;            // bins counted in shared memory first (32 KB)
constexpr uint64_t GROUP_EMPTY_KEY = ~0ull;

struct GroupJob {            // one partition of the query
   const uint32_t* filter32; // dense filter bitset
   uint32_t n_rows;          // rows of the partition (bits beyond are ignored: the columns end there)
   uint32_t pad;
   const uint32_t* codes[GROUP_MAX_COLUMNS];
};
struct GroupShape { uint32_t n_columns; uint32_t cardinality[GROUP_MAX_COLUMNS]; };

__device__ __forceinline__ uint64_t groupKey(const GroupJob& job, const GroupShape& shape, uint32_t row) {
   uint64_t key = 0;
   for (uint32_t c = 0; c < shape.n_columns; ++c) key = key * shape.cardinality[c] + __ldg(job.codes[c] + row);
   return key;
}

// grid: (row blocks, partitions); a thread takes 32-row words of the filter, block-strided
template <bool SMEM>
__global__ void __launch_bounds__(256)
group_count_dense(const GroupJob* __restrict__ jobs, const GroupShape shape, uint32_t n_bins, uint32_t* __restrict__ bins) {
   extern __shared__ uint32_t local_bins[];
   const GroupJob job = jobs[blockIdx.y];
   if (SMEM) {
      for (uint32_t i = threadIdx.x; i < n_bins; i += blockDim.x) local_bins[i] = 0;
      __syncthreads();
   }
   const uint32_t n_words = (job.n_rows + 31) / 32;
   for (uint32_t word = blockIdx.x * blockDim.x + threadIdx.x; word < n_words; word += gridDim.x * blockDim.x) {
      uint32_t bits = __ldg(job.filter32 + word);
      if ((word + 1) * 32 > job.n_rows) bits &= ~0u >> (32 - (job.n_rows & 31));
      while (bits) {
         const uint32_t row = word * 32 + uint32_t(__ffs(int(bits))) - 1u;
         bits &= bits - 1u;
         const uint32_t key = uint32_t(groupKey(job, shape, row));
         if (SMEM) atomicAdd(&local_bins[key], 1u);
         else atomicAdd(bins + key, 1u);
      }
   }
   if (SMEM) {
      __syncthreads();
      for (uint32_t i = threadIdx.x; i < n_bins; i += blockDim.x) { const uint32_t v = local_bins[i]; if (v) atomicAdd(bins + i, v); }
   }
}

__device__ __forceinline__ uint64_t groupHash(uint64_t key) {      // splitmix64 finaliser
   key ^= key >> 30; key *= 0xbf58476d1ce4e5b9ull;
   key ^= key >> 27; key *= 0x94d049bb133111ebull;
   return key ^ (key >> 31);
}

__global__ void __launch_bounds__(256)
group_count_hash(const GroupJob* __restrict__ jobs, const GroupShape shape, uint64_t capacity_mask, unsigned long long* __restrict__ keys, uint32_t* __restrict__ counts) {
   const GroupJob job = jobs[blockIdx.y];
   const uint32_t n_words = (job.n_rows + 31) / 32;
   for (uint32_t word = blockIdx.x * blockDim.x + threadIdx.x; word < n_words; word += gridDim.x * blockDim.x) {
      uint32_t bits = __ldg(job.filter32 + word);
      if ((word + 1) * 32 > job.n_rows) bits &= ~0u >> (32 - (job.n_rows & 31));
      while (bits) {
         const uint32_t row = word * 32 + uint32_t(__ffs(int(bits))) - 1u;
         bits &= bits - 1u;
         const uint64_t key = groupKey(job, shape, row);
         uint64_t at = groupHash(key) & capacity_mask;
         for (;;) {      // the table holds at least twice as many slots as there are filtered rows: a free slot always exists
            const unsigned long long seen = atomicCAS(keys + at, GROUP_EMPTY_KEY, key);
            if (seen == GROUP_EMPTY_KEY || seen == key) { atomicAdd(counts + at, 1u); break; }
            at = (at + 1) & capacity_mask;
         }
      }
   }
}

struct GroupRow { uint64_t key; uint32_t count, pad; };
static_assert(sizeof(GroupRow) == sizeof(sb_group_row), "GroupRow is the ABI's sb_group_row");

__global__ void group_compact_dense(const uint32_t* __restrict__ bins, uint64_t n_bins, GroupRow* __restrict__ rows, uint32_t* __restrict__ n_rows) {
   const uint64_t i = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;
   if (i >= n_bins) return;
   const uint32_t v = bins[i];
   if (v) rows[atomicAdd(n_rows, 1u)] = GroupRow{i, v, 0};
}
__global__ void group_compact_hash(const unsigned long long* __restrict__ keys, const uint32_t* __restrict__ counts, uint64_t capacity, GroupRow* __restrict__ rows, uint32_t* __restrict__ n_rows) {
   const uint64_t i = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;
   if (i >= capacity) return;
   const unsigned long long k = keys[i];
   if (k != GROUP_EMPTY_KEY) rows[atomicAdd(n_rows, 1u)] = GroupRow{k, counts[i], 0};
}

}  // namespace sb
