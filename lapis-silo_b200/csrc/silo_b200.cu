// libsilo_b200.so — C ABI (include/silo_b200.h) over the sm_100a kernels. One translation unit:
// this file holds the database object and the ingest entry points, pipeline.inc the lowering of plans and the launch
// sequences of a query, api_filter.inc the filter evaluation (replaces Operator::evaluate), api_action.inc finalize + the
// Mutations / Aggregated actions on evaluated filters, api_query.inc whole queries as one submission (and as CUDA graphs).
#include <atomic>
#include <shared_mutex>
#include <dlfcn.h>
#include <nccl.h>
#include "common.cuh"
#include "kernels_action.cuh"
#include "kernels_filter.cuh"
#include "kernels_group.cuh"
#include "kernels_build.cuh"
#include "pool.cuh"

using namespace sb;

namespace sb {

struct PinnedBuffer {   // growable page-locked staging buffer
   uint8_t* ptr = nullptr;
   size_t capacity = 0;
   ~PinnedBuffer() { if (ptr) cudaFreeHost(ptr); }
   uint8_t* ensure(size_t n) {
      if (n > capacity) {
         if (ptr) cudaFreeHost(ptr);
         capacity = std::max<size_t>(n * 2, 1 << 16);
         SB_CUDA(cudaMallocHost(reinterpret_cast<void**>(&ptr), capacity));
      }
      return ptr;
   }
};
// growable device buffer, allocated and released in stream order (no device-wide synchronisation on the query path)
struct GrowDevice {
   uint8_t* ptr = nullptr;
   size_t capacity = 0;
   GrowDevice() = default;
   GrowDevice(const GrowDevice&) = delete;
   GrowDevice& operator=(const GrowDevice&) = delete;
   ~GrowDevice() { if (ptr) cudaFree(ptr); }
   uint8_t* ensure(size_t n, cudaStream_t stream) {
      if (n > capacity) {
         if (ptr) SB_CUDA(cudaFreeAsync(ptr, stream));
         ptr = nullptr;
         capacity = std::max<size_t>(n + n / 2, 1 << 16);
         SB_CUDA(cudaMallocAsync(reinterpret_cast<void**>(&ptr), capacity, stream));
      }
      return ptr;
   }
};

// device memory of one sb_filter_eval_batch call; shared by its results
struct Slab {
   int device;
   cudaStream_t stream;
   uint8_t* ptr = nullptr;
   size_t bytes = 0;
   ~Slab() { if (ptr) { cudaSetDevice(device); cudaFreeAsync(ptr, stream); } }
};

// Everything one query in flight needs besides the index: the lowered plans, the filter results, the scratch of the row-wise
// leaves, the per-sequence action state, the count tables, the selected rows, and the pinned images the copies go through.
// A lane owns one (one-shot queries); a prepared query owns its own (fixed addresses: its launches are a CUDA graph).
// A query over several sequences (AminoAcidMutations over all genes: twelve count tables) runs the sequences' count kernels side
// by side: sequence k goes to branch k mod n -- a stream pair of its own that forks from the query's stream after the
// classification and joins it before the query ends (parallel branches of the CUDA graph for a prepared query).
struct SeqBranch {
   cudaStream_t main = nullptr, side = nullptr;
   cudaEvent_t fork2 = nullptr, join2 = nullptr, done = nullptr;
};
constexpr size_t SEQ_BRANCHES = 4;
struct QueryBuffers {
   GrowDevice blob, results, scratch, locs, deep_stack, action, tables, rows;
   PinnedBuffer staging, out_host;
   std::vector<SeqBranch> branches;
   cudaEvent_t ev_branches = nullptr;
   void ensureBranches(size_t n) {      // never inside a stream capture
      if (!ev_branches) SB_CUDA(cudaEventCreateWithFlags(&ev_branches, cudaEventDisableTiming));
      while (branches.size() < std::min(n, SEQ_BRANCHES)) {
         SeqBranch b;
         SB_CUDA(cudaStreamCreateWithFlags(&b.main, cudaStreamNonBlocking));
         SB_CUDA(cudaStreamCreateWithFlags(&b.side, cudaStreamNonBlocking));
         for (cudaEvent_t* e : {&b.fork2, &b.join2, &b.done}) SB_CUDA(cudaEventCreateWithFlags(e, cudaEventDisableTiming));
         branches.push_back(b);
      }
   }
   ~QueryBuffers() {
      for (auto& b : branches) { cudaEventDestroy(b.fork2); cudaEventDestroy(b.join2); cudaEventDestroy(b.done); cudaStreamDestroy(b.side); cudaStreamDestroy(b.main); }
      if (ev_branches) cudaEventDestroy(ev_branches);
   }
};

}  // namespace sb

struct sb_result {
   sb_db* db;
   std::shared_ptr<Slab> slab;
   uint32_t partition, n_words, n_words_padded;
   const uint64_t* words;
   const uint32_t* rank_local;
   const uint32_t* rank_block;
   const uint64_t* d_cardinality;   // the device counter finish_filters wrote (classification reads it on the device)
   uint64_t cardinality;
};

// One query in flight = one lane: its own streams, events and buffers. A calling thread is bound to a lane
// (round robin); calls of different threads overlap on the GPU, calls on one lane are serialised by its mutex.
struct sb_lane {
   std::mutex mutex;
   cudaStream_t stream = nullptr, stream2 = nullptr;   // stream2: kernels independent of the main chain (fork / join through events)
   cudaEvent_t ev_fork = nullptr, ev_join = nullptr, ev_fork2 = nullptr, ev_join2 = nullptr, ev_begin = nullptr, ev_k1_begin = nullptr, ev_k1_end = nullptr, ev_end = nullptr;
   bool timing_valid = false, k1_ran = false;
   uint32_t last_launches = 0;
   QueryBuffers buffers;
   void create() {
      SB_CUDA(cudaStreamCreateWithFlags(&stream, cudaStreamNonBlocking));
      SB_CUDA(cudaStreamCreateWithFlags(&stream2, cudaStreamNonBlocking));
      for (cudaEvent_t* e : {&ev_fork, &ev_join, &ev_fork2, &ev_join2}) SB_CUDA(cudaEventCreateWithFlags(e, cudaEventDisableTiming));
      for (cudaEvent_t* e : {&ev_begin, &ev_k1_begin, &ev_k1_end, &ev_end}) SB_CUDA(cudaEventCreate(e));
   }
   ~sb_lane() {
      for (cudaEvent_t e : {ev_fork, ev_join, ev_fork2, ev_join2, ev_begin, ev_k1_begin, ev_k1_end, ev_end}) if (e) cudaEventDestroy(e);
      if (stream2) cudaStreamDestroy(stream2);
      if (stream) cudaStreamDestroy(stream);
   }
};
constexpr uint32_t SB_LANES = 8;

namespace sb {
// NCCL entry points, resolved at run time (the process may already hold torch's copy of libnccl.so.2: dlopen returns that one)
struct NcclApi {
   ncclResult_t (*getUniqueId)(ncclUniqueId*) = nullptr;
   ncclResult_t (*commInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
   ncclResult_t (*reduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, int, ncclComm_t, cudaStream_t) = nullptr;
   ncclResult_t (*commDestroy)(ncclComm_t) = nullptr;
   const char* (*getErrorString)(ncclResult_t) = nullptr;
   static NcclApi& get() {
      static NcclApi api = [] {
         NcclApi a;
         // SB_NCCL_LIB names another copy. A process that also imports PyTorch must import it first: the loader keeps one
         // libnccl.so.2 per process, and torch needs its own (newer) one.
         const char* path = getenv("SB_NCCL_LIB") ? getenv("SB_NCCL_LIB") : "libnccl.so.2";
         void* lib = dlopen(path, RTLD_NOW | RTLD_LOCAL);
         if (!lib) throw Error(std::string(path) + " cannot be loaded: " + dlerror());
         auto sym = [&](const char* name) { void* p = dlsym(lib, name); if (!p) throw Error(std::string("libnccl.so.2 lacks ") + name); return p; };
         a.getUniqueId = reinterpret_cast<decltype(a.getUniqueId)>(sym("ncclGetUniqueId"));
         a.commInitRank = reinterpret_cast<decltype(a.commInitRank)>(sym("ncclCommInitRank"));
         a.reduce = reinterpret_cast<decltype(a.reduce)>(sym("ncclReduce"));
         a.commDestroy = reinterpret_cast<decltype(a.commDestroy)>(sym("ncclCommDestroy"));
         a.getErrorString = reinterpret_cast<decltype(a.getErrorString)>(sym("ncclGetErrorString"));
         return a;
      }();
      return api;
   }
};
#define SB_NCCL(call) do { ncclResult_t r__ = (call); if (r__ != ncclSuccess) throw ::sb::Error(std::string(#call) + " failed: " + ::sb::NcclApi::get().getErrorString(r__)); } while (0)
}  // namespace sb

struct sb_db {
   int device = 0, sm_count = 148;
   ncclComm_t comm = nullptr;          // sb_comm_init: count tables are summed over the ranks (root 0)
   int comm_rank = 0, comm_world = 1;
   std::mutex comm_mutex;
   uint32_t n_partitions = 0;
   bool finalized = false;
   std::mutex mutex;
   std::vector<std::unique_ptr<sb_lane>> lanes;
   std::atomic<uint32_t> next_lane{0};
   sb_lane* last_timed = nullptr;
   std::mutex k1_mutex;
   cudaEvent_t ev_timer_begin = nullptr, ev_timer_end = nullptr;
   std::atomic<uint64_t> launch_count{0}, h2d_bytes{0}, d2h_bytes{0};
   uint64_t k1_launches = 0;
   double k1_total_ms = 0;
   bool overlap = true;
   uint32_t k1_ctas = getenv("SB_K1_CTAS") ? uint32_t(atoi(getenv("SB_K1_CTAS"))) : 0;   // tuning / tests: persistent CTAs of the container scan (0 = as many as are resident at once)
   uint32_t k1_pad_bytes = getenv("SB_K1_PAD_KB") ? uint32_t(atoi(getenv("SB_K1_PAD_KB"))) * 1024u : 0;   // experiments: extra dynamic shared memory per CTA (fewer CTAs per SM, more L1)
   uint32_t k1_variant = getenv("SB_K1_VARIANT") ? uint32_t(atoi(getenv("SB_K1_VARIANT"))) : 0;
   // multi-rank: NCCL matches the collectives of a communicator by issue order, so a query that reduces (one ncclReduce per
   // requested sequence) holds this lock from its first launch to its last: the reduces of two host threads never interleave.
   // The ORDER of the queries must still be the same on every rank -- that is the caller's contract (include/silo_b200.h).
   std::mutex collective_mutex;

   std::vector<std::unique_ptr<HostStore>> stores;
   std::vector<std::unique_ptr<Sequence>> seqs;
   std::map<std::string, int> seq_by_name;
   DeviceBuffer<SeqDev> d_seqs;

   std::vector<RegContainer> reg_containers;
   std::vector<uint16_t> reg_pool;
   std::vector<RegisteredBitmap> reg_bitmaps;
   DeviceBuffer<RegContainer> d_reg_containers;
   DeviceBuffer<uint16_t> d_reg_pool;
   std::vector<Column> columns;
   std::vector<GroupColumn> group_columns;      // may grow after finalize (a column is coded when a query first groups by it)
   std::shared_mutex group_mutex;

   sb_pool_stats stats{};

   ~sb_db() {
      cudaSetDevice(device);
      if (comm) { sb::NcclApi::get().commDestroy(comm); comm = nullptr; }
      for (auto& c : columns) if (c.device) cudaFree(c.device);
      for (auto& c : group_columns) if (c.device) cudaFree(c.device);
      if (ev_timer_begin) { cudaEventDestroy(ev_timer_begin); cudaEventDestroy(ev_timer_end); }
      lanes.clear();
   }
};

namespace sb {
inline std::unique_lock<std::mutex> collectiveLock(sb_db* db) {      // held only while a communicator is active
   std::unique_lock<std::mutex> lock(db->collective_mutex, std::defer_lock);
   if (db->comm) lock.lock();
   return lock;
}
inline void useDevice(const sb_db* db) { SB_CUDA(cudaSetDevice(db->device)); }
// A lane for the calling thread, locked: the thread's home lane (threads are numbered in the order in which they first call in)
// if it is free, else the next free one -- a ninth request thread does not queue behind the first while other lanes idle.
struct LaneHold { sb_lane* lane; std::unique_lock<std::mutex> lock; };
inline sb_lane*& lastLaneOfThread() { thread_local sb_lane* last = nullptr; return last; }
inline LaneHold holdLane(sb_db* db) {
   static std::atomic<uint32_t> next_thread{0};
   thread_local uint32_t my_index = next_thread.fetch_add(1);
   const size_t n = db->lanes.size();
   for (size_t k = 0; k < n; ++k) {
      sb_lane* ln = db->lanes[(my_index + k) % n].get();
      std::unique_lock<std::mutex> lock(ln->mutex, std::try_to_lock);
      if (lock.owns_lock()) { lastLaneOfThread() = ln; return {ln, std::move(lock)}; }
   }
   sb_lane* ln = db->lanes[my_index % n].get();
   lastLaneOfThread() = ln;
   return {ln, std::unique_lock<std::mutex>(ln->mutex)};
}
}

extern "C" {

const char* sb_last_error(void) { return lastError().c_str(); }

int sb_device_count(int* n) {
   return guarded([&] { SB_CUDA(cudaGetDeviceCount(n)); });
}

int sb_db_create(int device, uint32_t n_partitions, sb_db** out) {
   return guarded([&] {
      int count = 0;
      SB_CUDA(cudaGetDeviceCount(&count));
      SB_REQUIRE(count > 0, "no CUDA device: libsilo_b200 has no CPU fallback");
      SB_REQUIRE(device >= 0 && device < count, "invalid device index");
      SB_REQUIRE(n_partitions > 0 && n_partitions < 65536, "n_partitions must be in [1, 65535]");
      SB_CUDA(cudaSetDevice(device));
      auto db = std::make_unique<sb_db>();
      db->device = device;
      SB_CUDA(cudaDeviceGetAttribute(&db->sm_count, cudaDevAttrMultiProcessorCount, device));
      db->n_partitions = n_partitions;
      for (uint32_t k = 0; k < SB_LANES; ++k) { db->lanes.push_back(std::make_unique<sb_lane>()); db->lanes.back()->create(); }
      {  // keep freed result slabs in the stream-ordered pool: a query must not pay for cudaMalloc
         cudaMemPool_t pool;
         SB_CUDA(cudaDeviceGetDefaultMemPool(&pool, device));
         uint64_t threshold = ~0ULL;
         SB_CUDA(cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &threshold));
      }
      SB_CUDA(cudaEventCreate(&db->ev_timer_begin));
      SB_CUDA(cudaEventCreate(&db->ev_timer_end));
      db->stats.n_partitions = n_partitions;
      *out = db.release();
   });
}

int sb_comm_unique_id(void* id_128_bytes) {
   return guarded([&] {
      static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is 128 bytes");
      ncclUniqueId id;
      SB_NCCL(NcclApi::get().getUniqueId(&id));
      memcpy(id_128_bytes, &id, sizeof id);
   });
}

int sb_comm_init(sb_db* db, const void* id_128_bytes, int rank, int world_size) {
   return guarded([&] {
      SB_REQUIRE(db && !db->comm, "sb_comm_init called twice");
      SB_REQUIRE(world_size >= 1 && rank >= 0 && rank < world_size, "rank must be in [0, world_size)");
      useDevice(db);
      ncclUniqueId id;
      memcpy(&id, id_128_bytes, sizeof id);
      SB_NCCL(NcclApi::get().commInitRank(&db->comm, world_size, id, rank));
      db->comm_rank = rank;
      db->comm_world = world_size;
   });
}

int sb_comm_free(sb_db* db) {
   return guarded([&] {
      if (!db || !db->comm) return;
      useDevice(db);
      cudaDeviceSynchronize();
      SB_NCCL(NcclApi::get().commDestroy(db->comm));
      db->comm = nullptr;
      db->comm_world = 1;
   });
}

int sb_db_free(sb_db* db) {
   return guarded([&] {
      if (!db) return;
      cudaSetDevice(db->device);
      cudaDeviceSynchronize();
      delete db;
   });
}

int sb_bitmap_add_portable(sb_db* db, uint32_t partition, const void* bytes, size_t len, uint32_t* bitmap_id) {
   return guarded([&] {
      std::lock_guard<std::mutex> lock(db->mutex);
      SB_REQUIRE(!db->finalized, "sb_bitmap_add_portable after sb_db_finalize");
      SB_REQUIRE(partition < db->n_partitions, "partition out of range");
      RegisteredBitmap bm{partition, uint32_t(db->reg_containers.size()), 0, 0};
      parsePortable(static_cast<const uint8_t*>(bytes), len, [&](const ParsedContainer& c) {
         while (db->reg_pool.size() % 8) db->reg_pool.push_back(0);
         RegContainer rc{};
         rc.pay_off = db->reg_pool.size();
         rc.size = c.kind == KIND_BITSET ? c.card : c.size;
         rc.key = c.key;
         rc.kind = c.kind;
         appendPayload(db->reg_pool, c);
         db->reg_containers.push_back(rc);
         bm.n_containers++;
         uint32_t last = 0;
         if (c.kind == KIND_ARRAY) last = c.u16(c.size - 1);
         else if (c.kind == KIND_RUN) last = uint32_t(c.u16(2 * (c.size - 1))) + c.u16(2 * (c.size - 1) + 1);
         else { for (int k = 1023; k >= 0; --k) { const uint64_t x = c.u64(k); if (x) { last = uint32_t(k) * 64 + 63 - __builtin_clzll(x); break; } } }
         bm.max_value = (uint32_t(c.key) << 16) | last;
      });
      *bitmap_id = uint32_t(db->reg_bitmaps.size());
      db->reg_bitmaps.push_back(bm);
   });
}

int sb_store_add(sb_db* db, uint32_t partition, const char* name, int alphabet, uint32_t length, uint32_t n_rows,
                 const uint8_t* blob, const uint64_t* slot_off, const int8_t* flipped, const int8_t* deleted,
                 const uint8_t* missing_blob, const uint64_t* missing_off, uint32_t* store_id) {
   return guarded([&] {
      std::lock_guard<std::mutex> lock(db->mutex);
      SB_REQUIRE(!db->finalized, "sb_store_add after sb_db_finalize");
      SB_REQUIRE(partition < db->n_partitions, "partition out of range");
      SB_REQUIRE(alphabet == 0 || alphabet == 1, "alphabet must be 0 (nucleotide) or 1 (amino acid)");
      SB_REQUIRE(length > 0, "sequence length must be positive");
      const uint32_t sigma = alphabet == 0 ? 16 : 25;
      int seq_index;
      auto it = db->seq_by_name.find(name);
      if (it == db->seq_by_name.end()) {
         auto seq = std::make_unique<Sequence>();
         seq->name = name; seq->alphabet = alphabet; seq->length = length; seq->sigma = sigma;
         seq->store_of_partition.assign(db->n_partitions, -1);
         seq_index = int(db->seqs.size());
         db->seq_by_name[name] = seq_index;
         db->seqs.push_back(std::move(seq));
      } else {
         seq_index = it->second;
         Sequence& seq = *db->seqs[seq_index];
         SB_REQUIRE(seq.alphabet == alphabet && seq.length == length, std::string("store '") + name + "' registered with a different alphabet or length");
      }
      Sequence& seq = *db->seqs[seq_index];
      SB_REQUIRE(seq.store_of_partition[partition] < 0, std::string("store '") + name + "' registered twice for one partition");
      auto st = std::make_unique<HostStore>();
      st->partition = partition; st->seq = uint32_t(seq_index); st->n_rows = n_rows;
      st->flipped.assign(flipped, flipped + length);
      st->deleted.assign(deleted, deleted + length);
      for (uint32_t pos = 0; pos < length; ++pos) {
         SB_REQUIRE(st->flipped[pos] >= -1 && st->flipped[pos] < int(sigma) && st->deleted[pos] >= -1 && st->deleted[pos] < int(sigma), "flipped / deleted symbol out of range");
      }
      const size_t n_slots = size_t(length) * sigma;
      for (size_t slot = 0; slot < n_slots; ++slot) {
         const uint64_t b = slot_off[slot], e = slot_off[slot + 1];
         if (e == b) continue;
         const uint32_t pos = uint32_t(slot / sigma), sym = uint32_t(slot - size_t(pos) * sigma);
         // the reference never reads the bitmap of the deleted symbol (mutations.cpp:74-83)
         const bool is_deleted = st->deleted[pos] == int(sym);
         parsePortable(blob + b, e - b, [&](const ParsedContainer& c) {
            SB_REQUIRE(uint64_t(c.key) * 65536 < std::max<uint64_t>(n_rows, 1), "a position bitmap holds row ids beyond sequence_count");
            if (is_deleted) return;
            TmpContainer t{uint32_t(slot), c.key, c.kind, c.kind == KIND_BITSET ? c.card : c.size, st->payload.size()};
            appendPayload(st->payload, c);
            st->containers.push_back(t);
            st->alg_bytes += (c.kind == KIND_BITSET ? 8192 : c.kind == KIND_ARRAY ? 2 * c.size : 4 * c.size) + 8;
            st->dir_bytes += 8;
         });
      }
      st->miss_row_off.resize(size_t(n_rows) + 1, 0);
      st->miss_row_bytes.resize(n_rows, 0);
      for (uint32_t row = 0; row < n_rows; ++row) {
         st->miss_row_off[row] = uint32_t(st->miss_iv.size());
         const uint64_t b = missing_off[row], e = missing_off[row + 1];
         uint32_t bytes = 0;
         parsePortable(missing_blob + b, e - b, [&](const ParsedContainer& c) {
            const uint32_t base = uint32_t(c.key) << 16;
            auto push = [&](uint32_t start, uint32_t end) {   // [start, end), merged with an adjacent predecessor
               if (st->miss_iv.size() > st->miss_row_off[row] && st->miss_iv.back().y == start) st->miss_iv.back().y = end;
               else st->miss_iv.push_back(make_uint2(start, end));
            };
            if (c.kind == KIND_RUN) {
               for (uint32_t r = 0; r < c.size; ++r) push(base + c.u16(2 * r), base + c.u16(2 * r) + c.u16(2 * r + 1) + 1);
               bytes += 4 * c.size;
            } else if (c.kind == KIND_ARRAY) {
               for (uint32_t k = 0; k < c.size; ++k) push(base + c.u16(k), base + c.u16(k) + 1);
               bytes += 2 * c.size;
            } else {
               for (uint32_t k = 0; k < 1024; ++k) { uint64_t x = c.u64(k); while (x) { const uint32_t bit = __builtin_ctzll(x); push(base + k * 64 + bit, base + k * 64 + bit + 1); x &= x - 1; } }
               bytes += 8192;
            }
         });
         st->miss_row_bytes[row] = bytes;
         st->miss_alg_bytes += bytes + 4;
      }
      st->miss_row_off[n_rows] = uint32_t(st->miss_iv.size());
      seq.store_of_partition[partition] = int(db->stores.size());
      *store_id = uint32_t(db->stores.size());
      db->stores.push_back(std::move(st));
      db->stats.n_stores++;
   });
}

int sb_column_add(sb_db* db, uint32_t partition, int dtype, const void* values, uint32_t n, uint32_t* column_id) {
   return guarded([&] {
      std::lock_guard<std::mutex> lock(db->mutex);
      SB_REQUIRE(partition < db->n_partitions, "partition out of range");
      SB_REQUIRE(dtype >= 0 && dtype <= 2, "dtype must be 0 (int32), 1 (float64) or 2 (uint32 date)");
      useDevice(db);
      Column c{partition, dtype, n, nullptr};
      const size_t bytes = size_t(n) * (dtype == 1 ? 8 : 4);
      SB_CUDA(cudaMalloc(&c.device, std::max<size_t>(bytes, 16)));
      if (bytes) SB_CUDA(cudaMemcpy(c.device, values, bytes, cudaMemcpyHostToDevice));
      *column_id = uint32_t(db->columns.size());
      db->columns.push_back(c);
      db->stats.total_device_bytes += bytes;
   });
}

}  // extern "C"

#include "pipeline.inc"
#include "api_filter.inc"
#include "api_action.inc"
#include "api_query.inc"
#include "api_group.inc"
#include "api_build.inc"

// ---- row selection without any allocation ------------------------------------------------------------------------------
// For callers that must not reach cudaMalloc (graph capture around them, tight serving loops, harnesses that forbid it):
// sb_mutation_rows_from_table with the scratch of k4_select_rows -- the row counter and the row area -- in a file-scope
// device array instead of the lane's growable buffer. Every GPU that loads the library holds the array (4 MiB) until exit.
// All calls share it and it keeps the last call's contents: the counter is cleared on the stream before every launch, the
// row area needs no clearing (only the first `counter` rows are read back). Two calls that overlap on one GPU would race
// on it; inside one process the mutex below serialises them.
namespace sb {
constexpr size_t STATIC_SCRATCH_BYTES = size_t(4) << 20;
constexpr size_t STATIC_HEADER_BYTES = 256;      // [0] row counter; the rows start 256-byte aligned behind it
constexpr size_t STATIC_MAX_ROWS = (STATIC_SCRATCH_BYTES - STATIC_HEADER_BYTES) / sizeof(sb_mutation_row);
static_assert(STATIC_MAX_ROWS == SB_STATIC_MAX_ROWS, "include/silo_b200.h states the limit");
static __device__ __align__(256) uint8_t g_static_scratch[STATIC_SCRATCH_BYTES];

// In host code g_static_scratch is only the symbol's host-side stand-in: the device address comes from
// cudaGetSymbolAddress, asked once per device with that device current.
constexpr int STATIC_MAX_DEVICES = 64;
static std::mutex g_static_mutex;
static uint8_t* g_static_ptr[STATIC_MAX_DEVICES] = {};
inline uint8_t* staticScratchOfCurrentDevice(int device) {      // g_static_mutex held
   SB_REQUIRE(device >= 0 && device < STATIC_MAX_DEVICES, "device index beyond the static scratch table");
   if (!g_static_ptr[device]) {
      void* p = nullptr;
      SB_CUDA(cudaGetSymbolAddress(&p, g_static_scratch));
      SB_REQUIRE(p != nullptr, "cudaGetSymbolAddress returned no address for the static scratch");
      g_static_ptr[device] = static_cast<uint8_t*>(p);
   }
   return g_static_ptr[device];
}
}  // namespace sb

extern "C" int sb_mutation_rows_from_table_static(sb_db* db, const char* store_name, const void* device_counts, double min_proportion,
                                                  const uint8_t* valid, uint32_t n_valid, sb_mutation_row* rows, size_t capacity, size_t* n_rows) {
   return guarded([&] {
      SB_REQUIRE(db && db->finalized, "sb_mutation_rows_from_table_static before sb_db_finalize");
      SB_REQUIRE(device_counts && n_rows && (rows || capacity == 0), "null argument");
      const LaneHold hold = holdLane(db);
      sb_lane* ln = hold.lane;
      useDevice(db);
      Sequence& seq = findSequence(db, store_name);
      const RowSelection rs = makeRowSelection(seq, min_proportion, valid, n_valid);
      const size_t max_rows = size_t(seq.length) * n_valid;      // what k4_select_rows can emit
      SB_REQUIRE(max_rows <= STATIC_MAX_ROWS, "sb_mutation_rows_from_table_static: length * n_valid = " + std::to_string(max_rows) + " rows exceed the static scratch (" +
                                                 std::to_string(STATIC_MAX_ROWS) + " rows); use sb_mutation_rows_from_table");
      std::lock_guard<std::mutex> scratch_lock(g_static_mutex);
      uint8_t* scratch = staticScratchOfCurrentDevice(db->device);
      uint32_t* d_counter = reinterpret_cast<uint32_t*>(scratch);
      sb_mutation_row* d_rows = reinterpret_cast<sb_mutation_row*>(scratch + STATIC_HEADER_BYTES);
      SB_CUDA(cudaMemsetAsync(d_counter, 0, 16, ln->stream));
      k4_select_rows<<<(seq.length + 255) / 256, 256, 0, ln->stream>>>(static_cast<const uint32_t*>(device_counts), seq.length, seq.reference.ptr, rs.valid, rs.min_proportion,
                                                                       d_rows, uint32_t(max_rows), d_counter);
      SB_CUDA(cudaGetLastError());
      db->launch_count += 1;
      uint32_t n = 0;      // results go straight into the caller's memory: no staging buffer of the library is grown
      SB_CUDA(cudaMemcpyAsync(&n, d_counter, 4, cudaMemcpyDeviceToHost, ln->stream));
      SB_CUDA(cudaStreamSynchronize(ln->stream));
      *n_rows = n;
      db->d2h_bytes += 4;
      if (n == 0 || n > capacity) return;
      SB_CUDA(cudaMemcpyAsync(rows, d_rows, size_t(n) * sizeof(sb_mutation_row), cudaMemcpyDeviceToHost, ln->stream));
      SB_CUDA(cudaStreamSynchronize(ln->stream));
      db->d2h_bytes += size_t(n) * sizeof(sb_mutation_row);
      std::sort(rows, rows + n, [](const sb_mutation_row& a, const sb_mutation_row& b) { return a.position != b.position ? a.position < b.position : a.symbol < b.symbol; });
      for (size_t i = 0; i < n; ++i) rows[i].symbol = valid[rows[i].symbol];
   });
}
