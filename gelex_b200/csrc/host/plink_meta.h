// Host-side PLINK metadata for the GRM path: FAM ids + row order, BIM chromosome runs, memory-mapped BED payload.
// Behaviour (accepted inputs, row order, error types and messages) follows the reference files cited per function;
// the implementation is this repo's own.
#pragma once
#include <cstdint>
#include <filesystem>
#include <fstream>
#include <string>
#include <utility>
#include <vector>

namespace gelex::b200 {

constexpr char kSampleIdSeparator = '\x1F';  // include/gelex/types/sample_id.h:30

// "<prefix>" or "<prefix>.bed" -> existing .bed path (src/data/genotype/bed_path.cpp:25-39; FileNotFoundException).
std::filesystem::path format_bed_path(const std::string& bfile);

// open_file<std::ifstream>(path, std::ios::in) of the reference (include/gelex/io/parser.h:48-93): directory, missing, unreadable and
// EMPTY files are rejected here, with the reference's exception types and messages.
std::ifstream open_text(const std::filesystem::path& path);

// Number of lines, a trailing line without '\n' counts (src/io/parser.cpp:30-61).
size_t count_total_lines(const std::filesystem::path& path);

// FAM ids as FID<US>IID in file order (src/data/loader/fam_loader.cpp:33-82, src/io/parser.cpp:63-82,105-116):
// delimiter = tab if the first line contains a tab, else a single space; empty lines skipped.
std::vector<std::string> load_fam_ids(const std::filesystem::path& fam_path);

struct SampleOrder {
    std::vector<std::string> ids;     // sorted + de-duplicated canonical ids = GRM row order (sample_manager.cpp:37-40)
    std::vector<int32_t> raw_to_out;  // FAM line -> GRM row (sample_projection.cpp:32-55)
    bool dense = false;               // identity mapping (the reference's memcpy decode path)
};
SampleOrder make_sample_order(const std::vector<std::string>& raw_ids);

// BIM: 6 columns, empty tokens skipped, chromosome kept as a string (src/data/loader/bim_loader.cpp:37-91).
std::vector<std::string> load_bim_chromosomes(const std::filesystem::path& bim_path);

struct ChrRun {
    std::string name;
    int64_t begin, end;
};
// Maximal runs of consecutive equal chromosome names; a chromosome that re-appears gives a second run with the same
// name (src/pipeline/grm_work_plan.cpp:54-81).
std::vector<ChrRun> chromosome_runs(const std::vector<std::string>& chroms);

// Read-only mapping of a SNP-major BED file (src/data/bed_pipe/mmap_reader.cpp:43-119): magic 6C 1B 01, size >=
// 3 + m*bpv, otherwise FileFormatException; payload() points at SNP 0.
class BedMapping {
   public:
    BedMapping(const std::filesystem::path& bed_path, int64_t num_snps, int64_t bytes_per_variant);
    ~BedMapping();
    BedMapping(const BedMapping&) = delete;
    BedMapping& operator=(const BedMapping&) = delete;
    const uint8_t* chunk_ptr(int64_t snp_begin, int64_t snp_count) const;  // nullptr when out of range
    int64_t bytes_per_variant() const { return bpv_; }

   private:
    void* base_ = nullptr;
    size_t size_ = 0;
    int64_t bpv_ = 0;
    size_t payload_snps_ = 0;
};

// FID<US>IID -> ("FID", "IID") (include/gelex/types/sample_id.h:47-71).
std::pair<std::string, std::string> split_sample_id(const std::string& id);

}  // namespace gelex::b200
