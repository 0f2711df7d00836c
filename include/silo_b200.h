/* silo_b200.h — C ABI of the B200 query hot path (libsilo_b200.so).
 *
 * The reference (pflanze/LAPIS-SILO) has no plugin / FFI layer; its hot path sits behind two C++
 * virtuals that are called from exactly one place:
 *     Operator::evaluate()          include/silo/query_engine/operators/operator.h:32
 *                                   (called at src/silo/query_engine/query_engine.cpp:48)
 *     Action::execute(db, filters)  include/silo/query_engine/actions/action.h:37-40
 *                                   (called through executeAndOrder, query_engine.cpp:60)
 * This header is what a binding for those two call sites links against. Plain pointers and sizes
 * only; no C++ or torch types; every function returns 0 on success, non-zero on failure with the
 * text available from sb_last_error() (thread-local). A failure never aborts the process: the host
 * turns it into an exception -> HTTP 500 (src/silo_api/query_handler.cpp:47-73).
 *
 * Threading: create / finalize / free are exclusive (the reference swaps databases under a unique lock,
 * src/silo_api/database_mutex.cpp:15-18); query entry points may be called concurrently (see below).
 * There is no CPU fallback: every entry point needs a CUDA device and fails loudly without one.
 */
#ifndef SILO_B200_H
#define SILO_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* Threading: after sb_db_finalize every query entry point may be called from several host threads at once, like the
 * reference's request handlers on its immutable Database. A calling thread is bound to one of 8 lanes (own CUDA streams
 * and scratch buffers); calls on different lanes overlap on the GPU, calls that share a lane are serialised. Every call
 * is synchronous: results are complete when it returns. */
typedef struct sb_db sb_db;         /* device-resident index of the partitions owned by this GPU   */
typedef struct sb_result sb_result; /* filter result of one partition: dense device bitset + |F|   */

/* ---- lifecycle ------------------------------------------------------------------------------ */
const char* sb_last_error(void);
int sb_device_count(int* n);
/* One sb_db per process+GPU. `n_partitions` = number of DatabasePartitions held by THIS db
 * (include/silo/database.h:41). Replaces the in-memory Database for the hot path; built where
 * Database::loadDatabaseState finishes (src/silo/database.cpp:480-560). */
int sb_db_create(int device, uint32_t n_partitions, sb_db** out);
int sb_db_free(sb_db* db);

/* ---- ingest (before sb_db_finalize) ------------------------------------------------------------ */
/* A roaring bitmap exactly as roaring::Roaring::write(buf, portable=true) emits it
 * (include/silo/roaring/roaring_serialize.h:23-28). len == 0 means the empty bitmap.
 * Used for the metadata index bitmaps that IndexScan leaves borrow:
 * PangoLineageColumnPartition::indexed_values / indexed_sublineage_values
 * (include/silo/storage/column/pango_lineage_column.h:37-39) and
 * IndexedStringColumnPartition::indexed_values (indexed_string_column.h:34-36). */
int sb_bitmap_add_portable(sb_db* db, uint32_t partition, const void* bytes, size_t len, uint32_t* bitmap_id);

/* One SequenceStorePartition<S> (include/silo/storage/sequence_store.h:34-88):
 *   alphabet   0 nucleotide (16 symbols), 1 amino acid (25 symbols); symbol = enum value
 *   length     number of positions L;  n_rows = sequence_count
 *   blob/slot_off   concatenated portable bitmaps; bitmap of (pos, sym) = bytes
 *                   [slot_off[pos*S+sym], slot_off[pos*S+sym+1])   (Position::bitmaps, position.h:35)
 *   flipped / deleted   per position symbol or -1 (position.h:36-37); both are honoured
 *   missing_blob/missing_off   per row: portable bitmap over POSITIONS (sequence_store.h:55)
 * Stores with the same `name` in different partitions form one logical sequence. */
int sb_store_add(sb_db* db, uint32_t partition, const char* name, int alphabet, uint32_t length, uint32_t n_rows,
                 const uint8_t* blob, const uint64_t* slot_off, const int8_t* flipped, const int8_t* deleted,
                 const uint8_t* missing_blob, const uint64_t* missing_off, uint32_t* store_id);

/* A metadata column vector borrowed by Selection predicates (operators/selection.h:33).
 * dtype: 0 int32 (IntColumnPartition), 1 float64 (FloatColumnPartition), 2 uint32 date (DateColumnPartition). */
int sb_column_add(sb_db* db, uint32_t partition, int dtype, const void* values, uint32_t n, uint32_t* column_id);

/* Packs everything into the HBM container pool and precomputes the per-partition full-filter count
 * tables (the K3 path of mutations.cpp:101-136). No ingest call is valid afterwards. */
int sb_db_finalize(sb_db* db);

/* ---- filter plans ------------------------------------------------------------------------------ */
/* Postfix (children before parent) encoding of a compiled Operator tree; same 11 tags as
 * operators::Type (include/silo/query_engine/operators/operator.h:10-22). */
enum sb_op {
   SB_EMPTY = 0, SB_FULL = 1, SB_INDEX_SCAN = 2, SB_INTERSECTION = 3, SB_COMPLEMENT = 4, SB_RANGE_SELECTION = 5,
   SB_SELECTION = 6, SB_BITMAP_SELECTION = 7, SB_THRESHOLD = 8, SB_UNION = 9, SB_BITMAP_PRODUCER = 10
};
enum sb_comparator { SB_EQUALS = 0, SB_LESS = 1, SB_HIGHER = 2, SB_LESS_OR_EQUALS = 3, SB_HIGHER_OR_EQUALS = 4, SB_NOT_EQUALS = 5 };

typedef struct sb_predicate {   /* CompareToValueSelection<T> (selection.cpp:134-193) */
   uint32_t column_id;          /* from sb_column_add                                  */
   int32_t comparator;          /* enum sb_comparator                                  */
   int32_t i32;                 /* value for dtype 0                                   */
   uint32_t u32;                /* value for dtype 2                                   */
   double f64;                  /* value for dtype 1                                   */
} sb_predicate;

typedef struct sb_node {
   int32_t op;                  /* enum sb_op                                                              */
   uint32_t row_count;          /* every operator carries it (complement.cpp:52, full.cpp:26)              */
   uint32_t n_children;         /* INTERSECTION / THRESHOLD: non-negated; UNION: all; COMPLEMENT: 1;
                                   SELECTION: 0 or 1. Children precede the node: first the non-negated
                                   ones in order, then the negated ones.                                   */
   uint32_t n_negated;          /* INTERSECTION / THRESHOLD                                                */
   uint32_t number_of_matchers; /* THRESHOLD (threshold.cpp:28-38: 0 < n < k)                              */
   uint32_t match_exactly;      /* THRESHOLD                                                               */
   /* INDEX_SCAN: either a registered bitmap (source 0, bitmap_id) or a stored position bitmap
      (source 1: store_id, position, symbol)                                                               */
   uint32_t source;
   uint32_t bitmap_id;
   uint32_t store_id;           /* also BITMAP_SELECTION: the store whose missing_symbol_bitmaps are probed */
   uint32_t position;           /* also BITMAP_SELECTION: `value`                                          */
   uint32_t symbol;
   uint32_t contains;           /* BITMAP_SELECTION: 1 CONTAINS, 0 NOT_CONTAINS (bitmap_selection.h:14)     */
   const uint32_t* ranges;      /* RANGE_SELECTION: n_ranges pairs [start, end)                            */
   uint32_t n_ranges;
   const sb_predicate* predicates; /* SELECTION                                                            */
   uint32_t n_predicates;
   const void* host_bitmap;     /* BITMAP_PRODUCER (and host-evaluated string Selections): portable bytes  */
   size_t host_bitmap_len;
} sb_node;

/* ---- filter evaluation: replaces Operator::evaluate() (query_engine.cpp:48) -------------------- */
/* Evaluates plan[i] (n_nodes[i] nodes) on partition partitions[i] for i < n; one batch = one query.
 * results[i] receives an owned handle. */
int sb_filter_eval_batch(sb_db* db, uint32_t n, const uint32_t* partitions, const sb_node* const* plans,
                         const uint32_t* n_nodes, sb_result** results);
int sb_filter_eval(sb_db* db, uint32_t partition, const sb_node* plan, uint32_t n_nodes, sb_result** result);
int sb_result_cardinality(const sb_result* r, uint64_t* cardinality);
/* Ascending row ids (what iterating the reference's OperatorResult yields). Call with out == NULL
 * to get the count. Needed by the host-side actions that stay outside the hot path. */
int sb_result_export_ids(const sb_result* r, uint32_t* out, size_t capacity, size_t* n);
/* The same row set as CRoaring portable bytes, i.e. what roaring::Roaring::readSafe (roaring_serialize.h:44) turns back into the
 * bitmap an OperatorResult holds (operator_result.h:9-32) -- for host-side actions that want a roaring, not an id list. Array
 * containers up to 4,096 values, bitset containers above, no run containers (the reader does not care; the caller may
 * runOptimize). Call with out == NULL to get the size. */
int sb_result_export_portable(const sb_result* r, void* out, size_t capacity, size_t* n_bytes);
int sb_result_free(sb_result* r);

/* ---- actions: replace Action::execute (query_engine.cpp:60) ------------------------------------ */
/* Mutations<S>::calculateMutationsPerPosition (mutations.cpp:139-164) for the sequence `store_name`:
 * counts[sym * length + pos], uint32 with wrap-around, summed over the n partitions given
 * (results[i] belongs to partitions[i]). `counts` is a HOST buffer of n_symbols*length entries.
 * Returns 0 and sets *produced = 0 (counts untouched) when every filter is empty, mirroring
 * preFilterBitmaps (mutations.cpp:36-63) which then yields no entry for the sequence. */
int sb_mutations(sb_db* db, const char* store_name, uint32_t n, const uint32_t* partitions,
                 sb_result* const* results, uint32_t* counts, size_t counts_len, int* produced);
/* ---- multi-GPU: partitions are sharded over processes (one per GPU); the only exchange of the path is the sum of the
 * count tables (the reference sums over partitions in mutations.cpp:71,108). The library issues it itself:
 * one ncclReduce (uint32, sum, root rank 0) on the stream of the count kernels, so the exchange costs one launch.
 * NCCL is loaded at run time (dlopen libnccl.so.2, or the file SB_NCCL_LIB names); single-GPU use never touches it. A
 * process that also uses PyTorch imports torch first, so that the one libnccl.so.2 of the process is torch's.
 *   sb_comm_unique_id   rank 0: a fresh ncclUniqueId (128 bytes) for the caller to hand to every rank
 *   sb_comm_init        every rank: joins the communicator on the db's device. From then on sb_mutations_device leaves the
 *                       SUM over all ranks in rank 0's buffer (other ranks: their local table). Calls that reduce
 *                       (sb_mutations_device, sb_query_mutations, sb_query_run) are then COLLECTIVE: every rank must issue them
 *                       for the same query in the same order. Inside a rank the library runs one collective call at a time
 *                       (the reduces of two host threads never interleave); agreeing on the order across ranks is the
 *                       caller's job -- with several request threads, e.g. a sequence number handed out by rank 0.
 *   sb_comm_free        leaves the communicator (sb_db_free does it too). */
int sb_comm_unique_id(void* id_128_bytes);
int sb_comm_init(sb_db* db, const void* id_128_bytes, int rank, int world_size);
int sb_comm_free(sb_db* db);

/* Same, but leaves the table in device memory (for the reduce over NVLink before the copy; see sb_comm_init). */
int sb_mutations_device(sb_db* db, const char* store_name, uint32_t n, const uint32_t* partitions,
                        sb_result* const* results, void* device_counts, size_t counts_len, int* produced);
/* The selection step of addMutationsToOutput (mutations.cpp:196-215) on the device, so that only the
 * emitted rows cross PCIe instead of the whole table: for every position, total = sum of the counts of the
 * `valid` symbols (SymbolType::VALID_MUTATION_SYMBOLS); skipped when 0; threshold = min_proportion == 0 ? 0 :
 * uint32(ceil(double(total) * min_proportion) - 1) -- the same IEEE double multiply and ceil --; a row for
 * every valid symbol != reference[position] with count > threshold. Rows come back sorted by (position,
 * order of `valid`). The proportion double(count) / double(total) stays with the caller. */
typedef struct sb_mutation_row { uint32_t position, symbol, count, total; } sb_mutation_row;
/* reference symbols (enum values) of a sequence, one byte per position; needed by the row selection */
int sb_sequence_set_reference(sb_db* db, const char* store_name, const uint8_t* symbols, uint32_t length);
/* rows of a count table that already is in device memory (e.g. after an NCCL reduce) */
int sb_mutation_rows_from_table(sb_db* db, const char* store_name, const void* device_counts, double min_proportion,
                                const uint8_t* valid, uint32_t n_valid, sb_mutation_row* rows, size_t capacity, size_t* n_rows);
/* sb_mutation_rows_from_table for callers that cannot reach cudaMalloc (code under CUDA graph capture, tight serving
 * loops, harnesses that forbid allocation). No counterpart in the reference.
 * Same arguments, same rows in the same order; the kernel's scratch (row counter + row area) is a file-scope device array
 * of 4 MiB that every GPU holds from the moment it loads the library until exit, so the call allocates nothing -- no
 * cudaMalloc / cudaFree / cudaMallocHost, no growth of a library buffer; the rows are copied straight into `rows`.
 * Limit: length * n_valid <= SB_STATIC_MAX_ROWS (nucleotides: 29,903 * 5 = 149,515), otherwise the call fails with a
 * message that says so. All calls share the one array: calls that overlap on a GPU would race on it; within a process
 * the library serialises them with a mutex. */
#define SB_STATIC_MAX_ROWS 262128u
int sb_mutation_rows_from_table_static(sb_db* db, const char* store_name, const void* device_counts, double min_proportion,
                                       const uint8_t* valid, uint32_t n_valid, sb_mutation_row* rows, size_t capacity, size_t* n_rows);
/* sb_mutations + the selection in one call with one synchronisation; *n_rows may exceed capacity (then call again) */
int sb_mutations_rows(sb_db* db, const char* store_name, uint32_t n, const uint32_t* partitions, sb_result* const* results,
                      double min_proportion, const uint8_t* valid, uint32_t n_valid,
                      sb_mutation_row* rows, size_t capacity, size_t* n_rows, int* produced);

/* ---- a whole query as ONE submission --------------------------------------------------------------
 * QueryEngine::executeQuery (query_engine.cpp:30-68) evaluates the partitions' filters (:40-49) and then runs the action
 * (:60). These entry points take both call sites together: the plans of all listed partitions, the sequences whose
 * Mutations rows are wanted (n_sequences == 0: filters only, e.g. for Aggregated) and the parameters of the row selection
 * (see sb_mutation_rows_from_table). The library enqueues filter evaluation, Mutations::preFilterBitmaps (on the device),
 * the count kernels and the row selection without reading anything back in between and synchronises once.
 * out[k] belongs to sequence_names[k]: the caller provides rows / capacity; n_rows may exceed capacity (then call again
 * with a larger buffer); produced == 0 mirrors preFilterBitmaps yielding no entry for the sequence (mutations.cpp:36-63).
 * cardinalities (optional) receives |F_p| per listed partition.
 * With sb_comm_init active every rank calls with its own partitions, in the same order on all ranks (see sb_comm_init); the
 * tables are summed over the ranks before the selection and only rank 0 receives rows. */
typedef struct sb_sequence_rows { sb_mutation_row* rows; size_t capacity; size_t n_rows; int produced; int pad; } sb_sequence_rows;
int sb_query_mutations(sb_db* db, uint32_t n, const uint32_t* partitions, const sb_node* const* plans, const uint32_t* n_nodes,
                       uint32_t n_sequences, const char* const* sequence_names, double min_proportion, const uint8_t* valid, uint32_t n_valid,
                       sb_sequence_rows* out, uint64_t* cardinalities);
/* The same query kept for replay: plans lowered and resident on the device, leaf bitmaps located once (a compiled Operator
 * tree holds its bitmap pointers the same way, index_scan.h:18), buffers of its own. use_graph != 0 captures the launches
 * in a CUDA graph: sb_query_run is then one graph launch + one synchronisation. One run at a time per sb_query; different
 * sb_query objects run concurrently. */
typedef struct sb_query sb_query;
int sb_query_prepare(sb_db* db, uint32_t n, const uint32_t* partitions, const sb_node* const* plans, const uint32_t* n_nodes,
                     uint32_t n_sequences, const char* const* sequence_names, double min_proportion, const uint8_t* valid, uint32_t n_valid,
                     int use_graph, sb_query** out);
int sb_query_run(sb_query* q, sb_sequence_rows* out, uint64_t* cardinalities);
/* A batch of prepared queries from one host thread (SURVEY 8f "batched filter evaluation"): every query is enqueued on its own
 * stream before the first is waited for. outs[i] (and cardinalities[i], may be null) belong to queries[i]; a query appears once. */
int sb_query_run_many(sb_query* const* queries, uint32_t n, sb_sequence_rows* const* outs, uint64_t* const* cardinalities);
/* the uint32[n_symbols][length] table of sequence `index` of the last run, in device memory (valid until the next run) */
int sb_query_table_device(sb_query* q, uint32_t index, const void** device_counts, size_t* counts_len);
/* CUDA-event time of the last sb_query_run, recorded on the query's own stream: first kernel to last copy */
int sb_query_elapsed_ms(sb_query* q, float* ms);
int sb_query_free(sb_query* q);
/* synchronous copy out of device memory the library handed out (sb_query_table_device) */
int sb_copy_from_device(void* host, const void* device, size_t bytes);

/* aggregateWithoutGrouping (aggregated.cpp:58-66): sum of cardinalities, uint32 wrap-around. */
int sb_aggregated_count(sb_db* db, uint32_t n, sb_result* const* results, uint32_t* count);

/* Aggregated with groupByFields (Aggregated::execute, aggregated.cpp:91-149; tuples: actions/tuple.cpp:29-76). The host
 * dictionary-codes every group-by column once (code = index of the row's value among the column's distinct values, the same
 * dictionary for all partitions) and registers the codes per partition:
 *   sb_group_column_add   uint32 codes[n] of one column of one partition; allowed after sb_db_finalize (a column is coded
 *                         when a query first groups by it)
 *   sb_aggregated_grouped counts the filtered rows of the n listed partitions per tuple. group_column_ids[i * n_columns + c]
 *                         is column c of partitions[i]; cardinalities[c] bounds the codes of column c. A tuple comes back as
 *                         the mixed-radix key ((code_0 * cardinalities[1] + code_1) * cardinalities[2] + code_2) ...; rows
 *                         are sorted by key (the reference's order is that of its hash map and is fixed by orderByFields);
 *                         *n_rows may exceed capacity (then call again). count is uint32 like the reference's map values. */
typedef struct sb_group_row { uint64_t key; uint32_t count; uint32_t pad; } sb_group_row;
int sb_group_column_add(sb_db* db, uint32_t partition, const uint32_t* codes, uint32_t n, uint32_t* group_column_id);
int sb_aggregated_grouped(sb_db* db, uint32_t n, const uint32_t* partitions, sb_result* const* results, uint32_t n_columns,
                          const uint32_t* group_column_ids, const uint32_t* cardinalities, sb_group_row* rows, size_t capacity, size_t* n_rows);

/* ---- index build: SequenceStorePartition<S>::fill + optimizeBitmaps on the device ---------------------------------
 * (src/silo/storage/sequence_store.cpp:101-211, position.cpp:24-127). genomes = n_rows * length characters, row-major: the
 * aligned sequences of one store of one partition as the FASTA reader yields them; present[row] == 0 is a null genome
 * (sequence_store.cpp:170-176: the row misses every position). char_to_symbol[256] maps a character to its symbol (enum
 * value) or -1 (then the call fails with the reference's "Illegal character contained in sequence."), iteration_order lists
 * the symbols as SymbolType::SYMBOLS does (it breaks ties for the most numerous symbol), missing_symbol is N / X.
 * sb_index_fetch hands back what the host wraps into roaring bitmaps:
 *   deleted[length]                      most numerous symbol per position (its bitmap is not kept), -1 if no row has a symbol
 *   slot_off[length * sigma + 1], ids    ascending row ids of bitmap (pos, sym) = ids[slot_off[pos*sigma+sym] .. slot_off[pos*sigma+sym+1])
 *   missing_off[n_rows + 1], missing_intervals   per row the maximal runs of the missing symbol as (start, end) pairs, end exclusive */
typedef struct sb_index sb_index;
int sb_index_build(int device, uint32_t sigma, const int8_t* char_to_symbol, const uint8_t* iteration_order, uint32_t missing_symbol, uint32_t length, uint32_t n_rows,
                   const char* genomes, const uint8_t* present, sb_index** out);
int sb_index_sizes(const sb_index* index, uint64_t* n_ids, uint64_t* n_missing_intervals);
int sb_index_fetch(const sb_index* index, int8_t* deleted, uint64_t* slot_off, uint32_t* ids, uint64_t* missing_off, uint32_t* missing_intervals);
int sb_index_free(sb_index* index);

/* ---- introspection (mirrors /info?details=true, database.cpp:239-305) -------------------------- */
typedef struct sb_pool_stats {
   uint64_t array_containers, array_values, run_containers, run_values, run_count, bitset_containers, bitset_values;
   uint64_t directory_bytes, payload_bytes, missing_bytes, full_table_bytes, total_device_bytes;
   uint64_t n_stores, n_partitions, n_work_items, n_chunks;
} sb_pool_stats;
int sb_pool_get_stats(const sb_db* db, sb_pool_stats* out);
/* Algorithmic bytes of sb_mutations for these results: payload+directory of every stored
 * container of the partial partitions, their filters, missing bitmaps of filtered rows, the table.
 * k1_bytes = the share the fused AND-popcount kernel is accountable for (containers + directory + filters). */
int sb_mutations_algorithmic_bytes(sb_db* db, const char* store_name, uint32_t n, const uint32_t* partitions,
                                   sb_result* const* results, uint64_t* bytes, uint64_t* k1_bytes);
/* CUDA-event time of the kernels of the last sb_mutations* call on this db, in milliseconds
 * (k1 = fused AND-popcount scan, total = first kernel to last kernel). */
int sb_last_timing(const sb_db* db, float* k1_ms, float* total_ms, uint32_t* n_launches);

/* Measurement helpers: CUDA events recorded on the db's own stream (torch.cuda.Event would only see
 * torch's stream), and the number of kernels launched by this db so far. */
int sb_timer_begin(sb_db* db);
int sb_timer_end(sb_db* db, float* elapsed_ms);
int sb_launch_count(const sb_db* db, uint64_t* launches);
/* Bytes moved by cudaMemcpyAsync on the db's stream so far (plans / PartQuery up, cardinalities / tables down). */
int sb_transfer_bytes(const sb_db* db, uint64_t* h2d, uint64_t* d2h);
/* By default independent kernels of a query overlap on a second stream (the row-wise leaves beside the leaf location;
 * the missing-symbol kernel beside the container scan). enabled = 0 serialises them on one stream, which is
 * how bench.py times the container scan alone for its roofline line. */
int sb_set_overlap(sb_db* db, int enabled);
/* Tuning: which compiled shape of the container scan runs (0 = default); tools/ measure them side by side in one process. */
int sb_debug_set_k1_variant(sb_db* db, uint32_t variant);
/* Accumulated CUDA-event time of the container scan (k1_scan) since the last call with reset != 0, and its launch count
 * (one-shot entry points only: a graph replay records no events). */
int sb_k1_accumulated(sb_db* db, int reset, double* total_ms, uint64_t* launches);

#ifdef __cplusplus
}
#endif
#endif /* SILO_B200_H */
