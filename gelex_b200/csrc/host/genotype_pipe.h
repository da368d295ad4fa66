// Whole-matrix standardised genotype export: the sink next to the GRM for the same decode + code-policy stages.
//   GenotypePipe    include/gelex/data/genotype/genotype_pipe.h:38-119, src/data/genotype/genotype_pipe.cpp:28-93
//                   -> <prefix>.bmat (column-major FP64 n x m, no header: src/data/io/binary_matrix_writer.cpp:34-52) and
//                      <prefix>.snpstats (int64 {n, m, #mono}, int64 mono indices, f64 means[m], f64 stddevs[m]:
//                      src/data/io/snp_stats_writer.cpp:50-91), the pair GenotypeMap (genotype_mmap.cpp:27-106) maps back in
//   GenotypeLoader  include/gelex/data/genotype/genotype_loader.h:38-113, src/data/genotype/genotype_loader.cpp:51-95
//                   -> the same matrix in host memory
// Same constructor checks and exception types/messages; the chunk body (decode, statistics, standardise) runs on the GPU through
// ggrm_genotype_chunk - there is no CPU path.
#pragma once
#include <cstdint>
#include <filesystem>
#include <memory>
#include <string>
#include <vector>

#include "grm_engine.h"
#include "plink_meta.h"

namespace gelex::b200 {

struct GenotypeStats {
    int64_t num_samples = 0, num_variants = 0;
    std::vector<int64_t> monomorphic_indices;  // ascending (the reference's loader appends in OpenMP order and sorts on read)
    std::vector<double> means, stddevs, freqs;
    bool is_monomorphic(int64_t snp) const;
};

// Shared input side: FAM order + mapped BED (GRM::GRM / BedPipe construction, bed_pipe.cpp:69-98).
class GenotypeSource {
   public:
    explicit GenotypeSource(const std::filesystem::path& bed_path);
    int64_t num_samples() const { return static_cast<int64_t>(order_.ids.size()); }
    int64_t num_raw_samples() const { return n_raw_; }
    int64_t num_variants() const { return m_; }
    const SampleOrder& order() const { return order_; }
    const BedMapping& bed() const { return *bed_; }

   private:
    SampleOrder order_;
    int64_t n_raw_ = 0, m_ = 0;
    std::unique_ptr<BedMapping> bed_;
};

class GenotypePipe {
   public:
    // throws FileExistsException("<prefix>: existing files") when either output is already there (genotype_pipe.cpp:64-75)
    GenotypePipe(const std::filesystem::path& bed_path, const std::filesystem::path& output_prefix, int device = 0);
    // additive == true: GeneticEffectType::Add, false: Dom.  Writes both files and returns what .snpstats holds.
    GenotypeStats process(bool additive, GenotypeProcessMethod method, size_t chunk_size = 10000);
    int64_t num_samples() const { return src_.num_samples(); }
    int64_t num_variants() const { return src_.num_variants(); }
    const std::filesystem::path& matrix_path() const { return matrix_path_; }
    const std::filesystem::path& stats_path() const { return stats_path_; }

   private:
    GenotypeSource src_;
    std::filesystem::path matrix_path_, stats_path_;
    int device_;
};

class GenotypeLoader {
   public:
    explicit GenotypeLoader(const std::filesystem::path& bed_path, int device = 0);
    // column-major n x m into `out` (caller-owned, n*m doubles; may be pinned memory from ggrm_host_alloc)
    GenotypeStats process(bool additive, GenotypeProcessMethod method, double* out, size_t chunk_size = 10000);
    int64_t num_samples() const { return src_.num_samples(); }
    int64_t num_variants() const { return src_.num_variants(); }

   private:
    GenotypeSource src_;
    int device_;
};

// SnpStatsWriter::write (snp_stats_writer.cpp:50-91) incl. its argument checks.
void write_snp_stats(const std::filesystem::path& path, const GenotypeStats& stats);

}  // namespace gelex::b200

// C entry points (tests / bindings), defined in capi.cpp.  dims3 = {n, m, #monomorphic}; mean/stddev/freq hold m doubles; mono receives
// at most mono_cap ascending SNP indices; any of them may be NULL.  Return 0 or a negative / GGRM_ERR_* status with "Type: message" in err.
extern "C" int gelex_b200_genotype_export(const char* bfile, const char* out_prefix, int method, int additive, int chunk_size, int device,
                                          int64_t* dims3, double* mean, double* stddev, double* freq, int64_t* mono, int64_t mono_cap,
                                          char* err, int err_len);
extern "C" int gelex_b200_genotype_load(const char* bfile, int method, int additive, int chunk_size, int device, double* out, int64_t out_cap,
                                        int64_t* dims3, double* mean, double* stddev, double* freq, int64_t* mono, int64_t mono_cap, char* err,
                                        int err_len);
