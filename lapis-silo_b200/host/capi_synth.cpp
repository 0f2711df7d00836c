// C entry points of the synthetic dataset generator for ctypes.
#include "capi_common.hpp"
#include "device_build.hpp"
#include "synth.hpp"

using namespace silo;

namespace {
struct SynthHandle {
   synth::Generated generated;
};
}

extern "C" {

// reference_genomes_json: path of a reference_genomes.json (testBaseData layout). Returns a handle
// owning the Database; silo_synth_db() yields the Database* usable with silo_engine_create / oracle_*.
void* silo_synth_generate(const char* reference_genomes_json, uint64_t n_sequences, uint64_t seed, int with_amino_acids,
                          int explicit_sequences, uint32_t genome_length, uint32_t n_threads, uint32_t world, uint32_t rank, uint32_t row_slices) {
   SynthHandle* handle = nullptr;
   capi::guarded([&] {
      std::map<std::string, std::string> nuc, aa;
      builder::readReferenceGenomes(reference_genomes_json, nuc, aa);
      if (nuc.count("main")) { std::string main = nuc["main"]; nuc.clear(); nuc["main"] = main; }
      synth::Options options;
      options.n_sequences = n_sequences; options.seed = seed; options.with_amino_acids = with_amino_acids != 0;
      options.explicit_sequences = explicit_sequences != 0; options.genome_length = genome_length; options.n_threads = n_threads; options.world = world ? world : 1; options.rank = rank; options.row_slices = row_slices ? row_slices : 1;
      auto h = std::make_unique<SynthHandle>();
      h->generated = synth::generate(options, nuc, aa);
      handle = h.release();
   });
   return handle;
}
uint32_t silo_synth_owned_count(void* handle) { return static_cast<uint32_t>(static_cast<SynthHandle*>(handle)->generated.owned.size()); }
uint32_t silo_synth_owned(void* handle, uint32_t i) { return static_cast<SynthHandle*>(handle)->generated.owned.at(i); }
void silo_synth_free(void* handle) { delete static_cast<SynthHandle*>(handle); }
void* silo_synth_db(void* handle) { return static_cast<SynthHandle*>(handle)->generated.db.get(); }
uint32_t silo_synth_clade_mutation_count(void* handle) { return static_cast<uint32_t>(static_cast<SynthHandle*>(handle)->generated.clade_mutations.size()); }
void silo_synth_clade_mutation(void* handle, uint32_t i, uint32_t* position, char* symbol) {
   const auto& m = static_cast<SynthHandle*>(handle)->generated.clade_mutations.at(i);
   *position = m.first;
   *symbol = nucleotideAlphabet().symbolToChar(m.second);
}
const char* silo_synth_lineage_name(void* handle, uint32_t i) { return static_cast<SynthHandle*>(handle)->generated.lineage_names.at(i).c_str(); }

// Rebuilds the database from the explicit row strings through the reference-shaped builder
// (builder::fillSequenceStore) and compares every bitmap and flag with the generated index.
// Returns the number of differences (0 = identical), or -1 on error.
// device < 0: builder::fillSequenceStore on the host; otherwise the device build (sb_index_build) on that GPU.
int64_t silo_synth_verify_against_builder_on(void* handle_, int device);
int64_t silo_synth_verify_against_builder(void* handle_) { return silo_synth_verify_against_builder_on(handle_, -1); }
int64_t silo_synth_verify_against_builder_on(void* handle_, int device) {
   int64_t differences = -1;
   capi::guarded([&] {
      const builder::StoreFiller fill = device < 0 ? builder::StoreFiller(builder::fillSequenceStore) : builder::deviceStoreFiller(device);
      auto& g = static_cast<SynthHandle*>(handle_)->generated;
      if (g.explicit_sequences.empty()) throw std::runtime_error("generated without explicit_sequences");
      int64_t diff = 0;
      for (size_t p = 0; p < g.db->partitions.size(); ++p) {
         for (int alphabet_id = 0; alphabet_id < 2; ++alphabet_id) {
            for (const auto& [name, store] : g.db->partitions[p]->getSequenceStores(alphabet_id)) {
               SequenceStorePartition rebuilt;
               rebuilt.init(store.alphabet, store.reference_sequence);
               std::vector<const std::optional<std::string>*> genomes;
               for (const auto& s : g.explicit_sequences[p].at(name)) genomes.push_back(&s);
               fill(rebuilt, genomes);
               if (rebuilt.sequence_count != store.sequence_count) ++diff;
               for (size_t k = 0; k < store.bitmaps.size(); ++k) if (!(rebuilt.bitmaps[k] == store.bitmaps[k]) || rebuilt.bitmaps[k].toPortable() != store.bitmaps[k].toPortable()) ++diff;
               for (size_t k = 0; k < store.deleted.size(); ++k) if (rebuilt.deleted[k] != store.deleted[k] || rebuilt.flipped[k] != store.flipped[k]) ++diff;
               for (size_t k = 0; k < store.missing_symbol_bitmaps.size(); ++k) if (rebuilt.missing_symbol_bitmaps[k].toPortable() != store.missing_symbol_bitmaps[k].toPortable()) ++diff;
            }
         }
      }
      differences = diff;
   });
   return differences;
}

}  // extern "C"
