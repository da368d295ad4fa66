// B200 GrmEngine: same Config in, same files + events out as the reference's gelex::GrmEngine
// (include/gelex/pipeline/grm_engine.h:30-51, src/pipeline/grm_engine.cpp:52-124), with the chunk loop of
// include/gelex/data/grm/grm.h:94-137 replaced by the C ABI of include/ggrm.h.
//
// Differences that are improvements, not behaviour changes:
//   * AD mode decodes once and runs two SYRKs per slab (the reference re-reads the BED per work item, grm_work_plan.cpp:95-103)
//   * --loco produces every per-chromosome partial in ONE pass; with write_whole the whole-genome matrix is the device-side
//     sum of the partials (the reference needs a second invocation)
//   * .bin files are written from row blocks (or, in 2D-tile mode, 2048 x 2048 super-tiles) fetched into page-locked buffers and
//     pwrite()n at their offsets by a thread pool - several GPUs write disjoint byte ranges of one file at once - instead of
//     one 4-byte ostream::write per element (grm_bin_writer.cpp:54-61); bytes are identical.
#pragma once
#include <cstdint>
#include <filesystem>
#include <functional>
#include <string>
#include <variant>
#include <vector>

#include "../../../include/ggrm.h"

namespace gelex::b200 {

enum class GrmType : uint8_t { A, D, AD };  // include/gelex/types/freq_effect.h

enum class GenotypeProcessMethod : uint8_t {  // include/gelex/data/genotype/genotype_processor.h:38-48
    StandardizeHWE = 1, CenterHWE, OrthStandardizeHWE, OrthCenterHWE, Standardize, Center, OrthStandardize, OrthCenter
};
// "SH" / "StandardizeHWE" / ... case-insensitive (apps/cli/cli_helper.h:41-82); throws InvalidInputException.
GenotypeProcessMethod parse_genotype_process_method(const std::string& value);
const char* method_name(GenotypeProcessMethod m);

// Events, include/gelex/infra/logging/grm_event.h:28-67 (same names and field meanings).
struct GrmConfigLoadedEvent { std::string method; GrmType mode; bool do_loco; };
struct GrmDataLoadedEvent { size_t num_samples; size_t num_snps; };
struct GrmComputeStartedEvent { size_t total_snps; };
struct GrmProgressEvent { size_t current; size_t total; bool done; };
struct GrmFilesWrittenEvent { size_t num_files; std::string output_dir; std::string file_pattern; };
using GrmEvent = std::variant<GrmConfigLoadedEvent, GrmDataLoadedEvent, GrmComputeStartedEvent, GrmProgressEvent, GrmFilesWrittenEvent>;
using GrmObserver = std::function<void(const GrmEvent&)>;

struct GrmWorkItem {  // src/pipeline/grm_work_plan.h:34-39
    std::vector<std::pair<int64_t, int64_t>> ranges;
    bool is_additive;
    std::string output_name;
    int group;  // index into the chromosome runs, -1 = whole genome
};

// Multi-GPU decomposition.  Auto = SNP shards (chromosome shards with --loco) below 100 000 samples, 2D output
// tiles from there on.
enum class Partition : uint8_t { Auto, SnpShard, Tile };

// How blocks reach the .bin files (grm_engine.cpp, BinFile): Auto = memcpy into a shared mapping of the pre-sized file for
// scattered segments (2D-tile rows) and on tmpfs, pwrite otherwise; Pwrite / Mmap force one of them.  A mapping is only used
// when the output file system had room for every output at the start of the run (a full disk must fail with
// FileWriteException, not SIGBUS).  GELEX_B200_WRITER=auto|pwrite|mmap overrides Auto.
enum class WriterMode : uint8_t { Auto, Pwrite, Mmap };

class GrmEngine {
   public:
    struct Config {
        // --- the reference's fields (grm_engine.h:33-43) ---
        std::filesystem::path bed_path;
        GrmType mode = GrmType::A;
        GenotypeProcessMethod method = GenotypeProcessMethod::OrthStandardizeHWE;
        bool do_loco = false;
        std::string out_prefix = "grm";
        int chunk_size = 10000;  // host push granularity = progress-event granularity (grm.h:111-132)
        // --- opt-in additions (SURVEY §5 "config / flag system") ---
        int precision = GGRM_PREC_FP16;  // GGRM_PREC_FP16 | GGRM_PREC_BF16 | GGRM_PREC_FP64
        std::vector<int> gpus = {0};     // device ordinals; >1 = SNP-sharded (or chromosome-sharded with --loco)
        bool write_whole = false;        // with do_loco: also write <out>.<add|dom>.{bin,id} from the sum of partials
        bool skip_all_missing = false;   // GGRM_ALLMISS_SKIP instead of the reference's NaN poisoning
        Partition partition = Partition::Auto;
        int writer_threads = 4;          // writer threads per GPU draining the page-locked read-out ring
        WriterMode writer = WriterMode::Auto;
        // 2D-tile partition only: upload every host chunk of the BED once (to GPU chunk % N) and let the other GPUs take it over
        // NVLink peer memory, instead of every GPU reading the whole file over its own PCIe link (grm_engine.cpp, BedRelay).
        // GELEX_B200_BED_RELAY=1 turns it on as well.  Off by default until it has been measured on an 8-GPU box.
        bool bed_relay = false;
        // Non-LOCO runs whose outputs are mapped: create the .bin up front and populate its pages (MADV_POPULATE_WRITE, a few host
        // threads) while the GPUs accumulate, so that the read-out copies into allocated pages (grm_engine.cpp, Prefaulter).
        // LOCO runs prepare the per-chromosome files two ahead of the writer the same way (OutputPreparer).  On by default
        // (measured on the GPU box, profiles/r02a_cli_writers.jsonl); GELEX_B200_PREFAULT=0/1 overrides.
        bool prefault = true;
        // GCTA binary format behind a flag (docs/source/data_formats.rst:188-208): instead of <out>.add.{bin,id} write
        // <out>.grm.bin (A_ij = G_ij / N_ij), <out>.grm.N.bin (N_ij = SNPs called in both samples) and <out>.grm.id.  Mode A, no LOCO.
        bool gcta = false;
        // ... with GCTA's default diagonal in .grm.bin (its --make-grm-alg 0: A_jj = 1 + (1/N) sum (x^2 - (1 + 2p) x + 2p^2) / (2p(1 - p)))
        // instead of sum z^2 / N.  Needs method StandardizeHWE, the one whose off-diagonal values are GCTA's.
        bool gcta_diag = false;
    };

    explicit GrmEngine(Config config);
    void compute(const GrmObserver& observer = {});

    // What the last compute() did (for the CLI's closing line and for tests).
    struct Summary {
        size_t num_samples = 0, num_snps = 0, num_files = 0;
        double wall_s = 0, device_pass_ms = 0, syrk_ms = 0, write_s = 0;
        std::vector<std::string> files;
        int precision_used = GGRM_PREC_FP16;  // Config::precision, or BF16 where FP16 cannot hold the code values (see compute())
        bool outputs_mapped = false;  // the run was allowed to write through shared mappings (WriterMode, free-space check)
        // Host-side phases of the run in seconds since compute() began, (name, begin, end); slowest GPU thread per phase.
        // "metadata" FAM / BIM / BED mapping, "device setup" context + buffers, "push" first SNP pushed .. pass enqueued,
        // "read-out" waiting for / fetching / writing the matrices still pending when the pass is enqueued, "teardown".
        struct Phase { std::string name; double begin, end; };
        std::vector<Phase> phases;
    };
    const Summary& summary() const { return summary_; }

    // Work plan exactly as GrmNormalPlan / GrmLocoPlan build it (src/pipeline/grm_work_plan.cpp:83-147).
    static std::vector<GrmWorkItem> build_plan(const std::vector<std::string>& chroms, GrmType mode, bool do_loco,
                                               std::string* output_pattern_suffix, size_t* num_groups);

    // How many of `offered` GPUs a one-shot run should use: the k that minimises 0.3 k (CUDA context creation, which the driver
    // serialises across the devices of a process: 2.3 s for eight) + T1 / k (T1 = n(n+1) m kinds / rate, rate = 1.3e15 flop/s for
    // 16-bit operands, 3.0e13 for FP64) + 0.4 s for k > 1 (reduce-scatter and piece-grain read-out); a run that writes more than
    // 32 GB (n_outputs matrices of 2 n^2 bytes: a LOCO run) takes at least two - the second PCIe link and writer pool are what it
    // is short of - and never fewer than hold the accumulators.  C1 / C2 / C3 -> 1, C4 -> 2, C5 -> 8; all measured
    // (profiles/r02_summary.md).  The CLI applies it to --gpus (an upper bound) unless --gpus-exact; Config::gpus is always taken as is.
    static int gpus_worth_using(int64_t n, int64_t m, int kinds, int precision, int offered, int n_outputs = 1);

    // Row ranges the GPUs of an SNP-shard run reduce and write (cuts on 2 048-row super-rows, balanced on triangle area).
    static std::vector<std::pair<int64_t, int64_t>> row_slices_for(int64_t n, int parts);
    // 2 048-row strips of the band-pipelined SNP-shard run, dealt by LPT: (rank, band, row range), rank-major, rows ascending.
    struct OwnedStrip { int rank, band; int64_t row_begin, row_end; };
    static std::vector<OwnedStrip> strip_owners_for(int64_t n, int world, const std::vector<int64_t>& band_row_end);
    // ... and the finer dealing the run actually uses: column pieces of the strips (whole 2 048 x 2 048 super-tile columns, ~64 MB
    // each), every band spread over all ranks: (rank, band, super-row, super-column range), rank-major, (band, super-row, column) order.
    struct OwnedPiece { int rank, band; int64_t super_row, super_col_begin, super_col_end; };
    static std::vector<OwnedPiece> piece_owners_for(int64_t n, int world, const std::vector<int64_t>& band_row_end);

    // CPU-only check of the writer pool (no GPU involved): writes the packed lower triangle of an n x n matrix with entry
    // (i, j) = float((31 i + j) mod 65521) + 0.5 through the same jobs the engine submits - whole-row blocks of `parts` row
    // slices (tiles = false), or 2 048 x 2 048 super-tile row segments dealt round-robin to `parts` concurrent pools (tiles = true) -
    // and returns the seconds spent.  tests/test_host_engine.py compares the file with numpy.
    // prefault_threads > 0 (and a mapped mode): the file's pages are populated first, as Config::prefault does during the pass;
    // that phase is timed separately (prefault_seconds) and the returned seconds are the writers' alone.
    static double writer_selftest(const std::filesystem::path& bin_path, int64_t n, bool tiles, int parts, int writer_threads, WriterMode mode,
                                  bool* mapped, int prefault_threads = 0, double* prefault_seconds = nullptr);

    // CPU-only check of the BED relay's protocol (ChunkRelay in grm_engine.cpp: `world` uploader threads filling `slots` slots
    // each, `world` consumers taking every chunk in order) with host buffers in place of device staging slots.  Returns the number
    // of (consumer, chunk) pairs that did not read the chunk's own bytes (0 = no slot was refilled early); fail_at_chunk >= 0
    // injects an upload failure, which must surface as that FileFormatException and not as a hang.
    static int64_t relay_selftest(int n_chunks, int world, int slots, int chunk_bytes, int fail_at_chunk = -1);

   private:
    Config config_;
    Summary summary_;
};

// Writers with the reference's byte layouts (grm_bin_writer.cpp:34-69, grm_id_writer.cpp:41-54).
void write_id_file(const std::filesystem::path& path, const std::vector<std::string>& sorted_ids);

}  // namespace gelex::b200

// C entry point for tests / bindings: runs GrmEngine::compute; returns 0 or the ggrm_status-like negative code,
// error text in err (NUL-terminated, truncated to err_len).  mode: 0 = A, 1 = D, 2 = AD.
// _ex adds: partition 0 = auto, 1 = SNP shards, 2 = 2D tiles; writer_threads >= 1.
extern "C" int gelex_b200_grm_run_ex(const char* bfile, const char* out_prefix, int method, int mode, int do_loco, int chunk_size,
                                     int precision, int n_gpus, int write_whole, int partition, int writer_threads, char* err, int err_len);
extern "C" int gelex_b200_grm_run(const char* bfile, const char* out_prefix, int method, int mode, int do_loco, int chunk_size,
                                  int precision, int n_gpus, int write_whole, char* err, int err_len);
// gelex_b200_grm_run_ex (one partition rule: auto) with the observer's events recorded as (type, a, b) triples; see capi.cpp.
extern "C" int gelex_b200_grm_run_events(const char* bfile, const char* out_prefix, int method, int mode, int do_loco, int chunk_size,
                                         int precision, int n_gpus, int64_t* events3, int64_t cap, int64_t* n_events, char* err, int err_len);
// Config::gcta run (GCTA binary format out); see capi.cpp.
extern "C" int gelex_b200_grm_run_gcta(const char* bfile, const char* out_prefix, int method, int chunk_size, int precision, int n_gpus,
                                       int writer_threads, char* err, int err_len);
// _ex adds: gcta_diag (Config::gcta_diag) and partition (1 = SNP shards, 2 = 2D tiles; 0 = auto).
extern "C" int gelex_b200_grm_run_gcta_ex(const char* bfile, const char* out_prefix, int method, int chunk_size, int precision, int n_gpus,
                                          int writer_threads, int gcta_diag, int partition, char* err, int err_len);
// strip_owners_for through C (CPU tests): out4 receives (rank, band, row_begin, row_end) per strip, up to cap strips; returns the strip count.
extern "C" int gelex_b200_strip_owners(int64_t n, int world, const int64_t* band_row_end, int n_bands, int64_t* out4, int cap);
// GrmEngine::gpus_worth_using through C (CPU tests).
extern "C" int gelex_b200_gpus_worth_using(int64_t n, int64_t m, int kinds, int precision, int offered, int n_outputs);
// piece_owners_for through C: out5 receives (rank, band, super_row, super_col_begin, super_col_end) per piece; returns the piece count.
extern "C" int gelex_b200_piece_owners(int64_t n, int world, const int64_t* band_row_end, int n_bands, int64_t* out5, int cap);
// GrmEngine::writer_selftest through C (CPU tests): writer 0 = auto, 1 = pwrite, 2 = mmap.
extern "C" int gelex_b200_writer_selftest(const char* bin_path, int64_t n, int tiles, int parts, int writer_threads, int writer,
                                          int prefault_threads, double* seconds_out, double* prefault_seconds_out, int* mapped_out, char* err,
                                          int err_len);
// GrmEngine::relay_selftest through C (CPU tests): returns the mismatch count (>= 0) or a negative status with err set.
extern "C" int64_t gelex_b200_relay_selftest(int n_chunks, int world, int slots, int chunk_bytes, int fail_at_chunk, char* err, int err_len);
