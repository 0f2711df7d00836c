// The calls INTEGRATION.md shows a maintainer, as code that has to compile against include/silo_b200.h and link against
// libsilo_b200.so (tests/test_host_cpu.py::test_integration_snippets_compile_and_link). Nothing here runs: main() returns before the
// first call. The reference's types are stood in for by minimal structs with the members the snippets touch.
#include <cstdint>
#include <cstdio>
#include <stdexcept>
#include <string>
#include <vector>

extern "C" {
#include "silo_b200.h"
}

static void check(int rc) {
   if (rc != 0) throw std::runtime_error(sb_last_error());
}

struct StoreBytes {   // what addStore concatenates: bitmap.write(ptr, true) of all L * sigma bitmaps + the missing-symbol bitmaps
   std::vector<uint8_t> blob, missing_blob;
   std::vector<uint64_t> slot_off, missing_off;
   std::vector<int8_t> flipped, deleted;
   uint32_t length = 0, n_rows = 0;
};

// upload after Database::build / loadDatabaseState
sb_db* uploadToDevice(int device, uint32_t n_partitions, const std::vector<StoreBytes>& nuc_main, const std::vector<std::vector<int32_t>>& age,
                      const std::vector<std::vector<uint8_t>>& lineage_bitmap, const std::vector<uint8_t>& reference) {
   sb_db* db = nullptr;
   check(sb_db_create(device, n_partitions, &db));
   for (uint32_t p = 0; p < n_partitions; ++p) {
      const StoreBytes& s = nuc_main[p];
      uint32_t store_id = 0, column_id = 0, bitmap_id = 0;
      check(sb_store_add(db, p, "main", /*alphabet=*/0, s.length, s.n_rows, s.blob.data(), s.slot_off.data(), s.flipped.data(), s.deleted.data(),
                         s.missing_blob.data(), s.missing_off.data(), &store_id));
      check(sb_column_add(db, p, /*dtype int32=*/0, age[p].data(), static_cast<uint32_t>(age[p].size()), &column_id));
      check(sb_bitmap_add_portable(db, p, lineage_bitmap[p].data(), lineage_bitmap[p].size(), &bitmap_id));
   }
   check(sb_sequence_set_reference(db, "main", reference.data(), static_cast<uint32_t>(reference.size())));
   check(sb_db_finalize(db));
   return db;
}

// Operator::evaluate() of all partitions in one call, then what host-side actions take from a result
void evaluateFilters(sb_db* db, const std::vector<std::vector<sb_node>>& plans, std::vector<sb_result*>& partition_filters) {
   std::vector<uint32_t> partition_ids, plan_sizes;
   std::vector<const sb_node*> plan_ptrs;
   for (size_t p = 0; p < plans.size(); ++p) {
      partition_ids.push_back(static_cast<uint32_t>(p));
      plan_ptrs.push_back(plans[p].data());
      plan_sizes.push_back(static_cast<uint32_t>(plans[p].size()));
   }
   partition_filters.assign(plans.size(), nullptr);
   check(sb_filter_eval_batch(db, static_cast<uint32_t>(plans.size()), partition_ids.data(), plan_ptrs.data(), plan_sizes.data(), partition_filters.data()));
   for (sb_result* result : partition_filters) {
      uint64_t cardinality = 0;
      check(sb_result_cardinality(result, &cardinality));
      size_t n_ids = 0, size = 0;
      check(sb_result_export_ids(result, nullptr, 0, &n_ids));
      std::vector<uint32_t> ids(n_ids);
      check(sb_result_export_ids(result, ids.data(), ids.size(), &n_ids));
      check(sb_result_export_portable(result, nullptr, 0, &size));
      std::vector<char> bytes(size);
      check(sb_result_export_portable(result, bytes.data(), size, &size));      // roaring::Roaring::readSafe(bytes.data(), size)
   }
}

// Action::execute for Aggregated and Mutations on evaluated filters
void actions(sb_db* db, const std::vector<sb_result*>& filters, const std::vector<uint32_t>& partition_ids, double min_proportion,
             const uint8_t* valid_symbols, uint32_t n_valid, uint32_t sigma, uint32_t length) {
   uint32_t count = 0;
   check(sb_aggregated_count(db, static_cast<uint32_t>(filters.size()), filters.data(), &count));
   std::vector<sb_mutation_row> rows(4096);
   size_t n = 0;
   int produced = 0;
   check(sb_mutations_rows(db, "main", static_cast<uint32_t>(filters.size()), partition_ids.data(), filters.data(), min_proportion, valid_symbols, n_valid,
                           rows.data(), rows.size(), &n, &produced));
   std::vector<uint32_t> table(size_t(sigma) * length);
   check(sb_mutations(db, "main", static_cast<uint32_t>(filters.size()), partition_ids.data(), filters.data(), table.data(), table.size(), &produced));
   for (sb_result* result : filters) check(sb_result_free(result));
}

// both call sites in one submission; a prepared query; a batch of prepared queries
void oneSubmission(sb_db* db, const std::vector<uint32_t>& partition_ids, const std::vector<const sb_node*>& plan_ptrs, const std::vector<uint32_t>& plan_sizes,
                   const std::string& sequence_name, double min_proportion, const uint8_t* valid_symbols, uint32_t n_valid) {
   std::vector<sb_mutation_row> buffer(4096);
   sb_sequence_rows out{buffer.data(), buffer.size(), 0, 0, 0};
   const char* names[] = {sequence_name.c_str()};
   check(sb_query_mutations(db, static_cast<uint32_t>(partition_ids.size()), partition_ids.data(), plan_ptrs.data(), plan_sizes.data(), 1, names, min_proportion,
                            valid_symbols, n_valid, &out, /*cardinalities=*/nullptr));
   std::vector<uint64_t> cardinalities(partition_ids.size());
   check(sb_query_mutations(db, static_cast<uint32_t>(partition_ids.size()), partition_ids.data(), plan_ptrs.data(), plan_sizes.data(), 0, nullptr, 0.0,
                            nullptr, 0, nullptr, cardinalities.data()));      // filters only: Aggregated = sum of the cardinalities
   sb_query* q = nullptr;
   check(sb_query_prepare(db, static_cast<uint32_t>(partition_ids.size()), partition_ids.data(), plan_ptrs.data(), plan_sizes.data(), 1, names, min_proportion,
                          valid_symbols, n_valid, /*use_graph=*/1, &q));
   check(sb_query_run(q, &out, nullptr));
   sb_query* batch[] = {q};
   sb_sequence_rows* outs[] = {&out};
   uint64_t* no_cardinalities[] = {nullptr};
   check(sb_query_run_many(batch, 1, outs, no_cardinalities));
   const void* device_table = nullptr;
   size_t table_len = 0;
   float ms = 0;
   check(sb_query_table_device(q, 0, &device_table, &table_len));
   std::vector<uint32_t> table(table_len);
   check(sb_copy_from_device(table.data(), device_table, table_len * sizeof(uint32_t)));
   check(sb_query_elapsed_ms(q, &ms));
   check(sb_query_free(q));
}

// Aggregated with groupByFields; the index build
void groupedAndBuild(sb_db* db, const std::vector<sb_result*>& filters, const std::vector<uint32_t>& partition_ids, const std::vector<std::vector<uint32_t>>& codes,
                     uint32_t dictionary_size, const int8_t* char_to_symbol, const uint8_t* iteration_order, const std::string& genomes, const std::vector<uint8_t>& present, uint32_t length) {
   std::vector<uint32_t> column_ids(partition_ids.size());
   for (size_t p = 0; p < partition_ids.size(); ++p) check(sb_group_column_add(db, partition_ids[p], codes[p].data(), static_cast<uint32_t>(codes[p].size()), &column_ids[p]));
   const uint32_t cardinalities[] = {dictionary_size};
   std::vector<sb_group_row> rows(4096);
   size_t n = 0;
   check(sb_aggregated_grouped(db, static_cast<uint32_t>(filters.size()), partition_ids.data(), filters.data(), 1, column_ids.data(), cardinalities, rows.data(), rows.size(), &n));
   sb_index* index = nullptr;
   check(sb_index_build(/*device=*/0, /*sigma=*/16, char_to_symbol, iteration_order, /*missing_symbol=*/15, length, static_cast<uint32_t>(present.size()), genomes.data(), present.data(), &index));
   uint64_t n_ids = 0, n_intervals = 0;
   check(sb_index_sizes(index, &n_ids, &n_intervals));
   std::vector<int8_t> deleted(length);
   std::vector<uint64_t> slot_off(size_t(length) * 16 + 1), missing_off(present.size() + 1);
   std::vector<uint32_t> ids(n_ids), intervals(2 * n_intervals);
   check(sb_index_fetch(index, deleted.data(), slot_off.data(), ids.data(), missing_off.data(), intervals.data()));
   check(sb_index_free(index));
}

// one process per GPU; the library sums the count tables (ncclReduce to rank 0) once a communicator is set
void multiGpu(sb_db* db, int rank, int world, void (*hand_to_every_rank)(void* id_128_bytes), const std::vector<sb_result*>& filters,
              const std::vector<uint32_t>& partition_ids, void* device_table, size_t table_len, const uint8_t* valid_symbols, uint32_t n_valid) {
   unsigned char id[128] = {0};
   if (rank == 0) check(sb_comm_unique_id(id));
   hand_to_every_rank(id);
   check(sb_comm_init(db, id, rank, world));
   int produced = 0;
   check(sb_mutations_device(db, "main", static_cast<uint32_t>(filters.size()), partition_ids.data(), filters.data(), device_table, table_len, &produced));
   if (rank == 0) {
      std::vector<sb_mutation_row> rows(4096);
      size_t n = 0;
      check(sb_mutation_rows_from_table(db, "main", device_table, 0.05, valid_symbols, n_valid, rows.data(), rows.size(), &n));
      check(sb_mutation_rows_from_table_static(db, "main", device_table, 0.05, valid_symbols, n_valid, rows.data(), rows.size(), &n));
   }
   check(sb_comm_free(db));
   check(sb_db_free(db));
}

int main(int argc, char**) {
   int devices = 0;
   if (argc < 1000) { std::printf("links\n"); return 0; }      // never goes on: the functions above only have to compile and link
   check(sb_device_count(&devices));
   return devices;
}
