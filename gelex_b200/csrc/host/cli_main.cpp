// `gelex-b200 grm ...` — drop-in for `gelex grm` (apps/cli/grm/grm_args.cpp:31-62): identical flags and defaults, the
// same [Config] / [Dataset Summary] / progress / [File Summary] report (apps/cli/grm/grm_reporter.cpp:30-94), messages also
// written to <out>.log, exit code 0 / 1 (apps/main.cpp:54-98).  Opt-in additions: --gpus, --precision, --with-whole,
// --skip-all-missing, --partition, --writer-threads, --writer, --prefault, --bed-relay, --gcta, --gcta-diag.
#include <unistd.h>

#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <fstream>
#include <iostream>
#include <string>
#include <thread>
#include <vector>

#include "exceptions.h"
#include "grm_engine.h"
#include "plink_meta.h"

using namespace gelex::b200;

namespace {

struct Logger {
    std::ofstream file;
    void info(const std::string& s) {
        std::cout << s << "\n";
        if (file.is_open()) file << s << "\n";
    }
    void error(const std::string& s) {
        std::cerr << "[error] " << s << "\n";
        if (file.is_open()) file << "[error] " << s << "\n";
    }
};

const char* kUsage =
    "Usage: gelex-b200 grm [--help] --bfile <BFILE> [--out <OUT>] [--geno-method <STR>] [--chunk-size <SIZE>]\n"
    "                      [--threads <N>] [--add] [--dom] [--loco] [--gpus <N>] [--gpus-exact] [--precision <fp16|bf16|fp64>]\n"
    "                      [--partition <auto|snp|tile>] [--writer-threads <N>] [--writer <auto|pwrite|mmap>]\n"
    "                      [--no-prefault] [--bed-relay] [--with-whole] [--skip-all-missing] [--gcta [--gcta-diag]]\n\n"
    "Compute genomic relationship matrix (GRM) from PLINK binary files and output in GCTA format\n\n"
    "Data Files:\n"
    "  -b, --bfile <BFILE>        PLINK binary file prefix (.bed/.bim/.fam) [required]\n"
    "  -o, --out <OUT>            Output file prefix [default: \"grm\"]\n\n"
    "GRM Options:\n"
    "  --geno-method, --gm <STR>  Genotype method: StandardizeHWE(SH), CenterHWE(CH), OrthStandardizeHWE(OSH), OrthCenterHWE(OCH),\n"
    "                             Standardize(S), Center(C), OrthStandardize(OS), OrthCenter(OC) [default: \"OSH\"]\n"
    "  -c, --chunk-size <SIZE>    Chunk size for memory-efficient computation [default: 10000]\n"
    "  -t, --threads <N>          Number of threads (-1 for all cores); host staging only on the B200 engine\n"
    "  --add                      Compute additive GRM\n"
    "  --dom                      Compute dominance GRM\n"
    "  --loco                     Compute GRM for each chromosome\n\n"
    "B200 engine options:\n"
    "  --gpus <N>                 Use up to N GPUs of this node: as many as the problem size pays for [default: 1]\n"
    "  --gpus-exact               Use exactly the N GPUs of --gpus\n"
    "  --precision <P>            fp16 (default) | bf16 : 16-bit operands, FP32 accumulate on tcgen05; fp64 : DMMA\n"
    "  --partition <auto|snp|tile> Multi-GPU split: SNP shards + one reduce-scatter, or 2D output tiles (auto: tile from 100k samples)\n"
    "  --writer-threads <N>       writer threads per GPU draining the read-out ring [default: 4]\n"
    "  --writer <auto|pwrite|mmap> How blocks reach the .bin: pwrite, or memcpy into a shared mapping of the pre-sized file\n"
    "                             (auto: mapping for scattered tile rows and on tmpfs, when the file system has room) [default: auto]\n"
    "  --no-prefault              Do not create the .bin files ahead of the writers and populate their pages meanwhile (default: do)\n"
    "  --bed-relay                2D-tile partition: upload each BED chunk once and share it between the GPUs over NVLink\n"
    "  --with-whole               With --loco also write the whole-genome matrix (sum of the partials, same pass)\n"
    "  --skip-all-missing         Treat an all-missing SNP as contributing 0 (the reference lets it poison the GRM with NaN)\n"
    "  --gcta                     GCTA binary format: <out>.grm.bin (G_ij / N_ij), <out>.grm.N.bin, <out>.grm.id (with --add, no --loco)\n"
    "  --gcta-diag                with --gcta and --method SH: GCTA's default diagonal, 1 + (1/N) sum (x^2 - (1+2p)x + 2p^2) / (2p(1-p))\n";

std::string mode_name(GrmType m) { return m == GrmType::A ? "A" : (m == GrmType::D ? "D" : "AD"); }

int run_grm(int argc, char** argv, Logger& log) {
    std::string bfile, out = "grm", gm = "OSH", precision = "fp16", partition = "auto", writer = "auto";
    int chunk = 10000, gpus = 1, writer_threads = 4;
    int threads = static_cast<int>(std::thread::hardware_concurrency() / 2);
    bool add = false, dom = false, loco = false, with_whole = false, skip_missing = false, gcta = false, gcta_diag = false, bed_relay = false, prefault = true, gpus_exact = false;
    auto need = [&](int& i) -> std::string {
        if (i + 1 >= argc) throw ArgumentValidationException(std::string("Too few arguments for '") + argv[i] + "'.");
        return argv[++i];
    };
    auto to_int = [](const std::string& s, const char* flag) {
        try {
            size_t pos = 0;
            const int v = std::stoi(s, &pos);
            if (pos != s.size()) throw std::invalid_argument(s);
            return v;
        } catch (const std::exception&) {
            throw ArgumentValidationException(std::string("pattern '") + s + "' not identified as integer for " + flag);
        }
    };
    for (int i = 2; i < argc; ++i) {
        const std::string a = argv[i];
        if (a == "-h" || a == "--help") { std::cout << kUsage; return 0; }
        else if (a == "-b" || a == "--bfile") bfile = need(i);
        else if (a == "-o" || a == "--out") out = need(i);
        else if (a == "--geno-method" || a == "--gm") gm = need(i);
        else if (a == "-c" || a == "--chunk-size") chunk = to_int(need(i), "--chunk-size");
        else if (a == "-t" || a == "--threads") threads = to_int(need(i), "--threads");
        else if (a == "--add") add = true;
        else if (a == "--dom") dom = true;
        else if (a == "--loco") loco = true;
        else if (a == "--gpus") gpus = to_int(need(i), "--gpus");
        else if (a == "--gpus-exact") gpus_exact = true;
        else if (a == "--precision") precision = need(i);
        else if (a == "--partition") partition = need(i);
        else if (a == "--writer-threads") writer_threads = to_int(need(i), "--writer-threads");
        else if (a == "--writer") writer = need(i);
        else if (a == "--with-whole") with_whole = true;
        else if (a == "--bed-relay") bed_relay = true;
        else if (a == "--prefault") prefault = true;
        else if (a == "--no-prefault") prefault = false;
        else if (a == "--skip-all-missing") skip_missing = true;
        else if (a == "--gcta") gcta = true;
        else if (a == "--gcta-diag") gcta_diag = true;
        else throw ArgumentValidationException("Unknown argument: " + a);
    }
    (void)threads;
    if (bfile.empty()) throw ArgumentValidationException("-b: required.");
    log.file.open(out + ".log", std::ios::out | std::ios::trunc);  // logging::initialize("<out>.log"), apps/main.cpp:60
    GrmEngine::Config cfg;
    cfg.mode = (add && dom) ? GrmType::AD : (dom ? GrmType::D : GrmType::A);  // grm_config.cpp:31-41 (neither flag -> A)
    if (chunk <= 0) throw ArgumentValidationException("chunk_size must be positive");
    cfg.bed_path = format_bed_path(bfile);
    cfg.method = parse_genotype_process_method(gm);
    cfg.do_loco = loco;
    cfg.out_prefix = out;
    cfg.chunk_size = chunk;
    if (precision == "fp16") cfg.precision = GGRM_PREC_FP16;
    else if (precision == "bf16") cfg.precision = GGRM_PREC_BF16;
    else if (precision == "fp64") cfg.precision = GGRM_PREC_FP64;
    else throw ArgumentValidationException("--precision must be fp16, bf16 or fp64");
    if (gpus < 1) throw ArgumentValidationException("--gpus must be >= 1");
    // --gpus is an upper bound: a one-shot process pays ~0.3 s of context creation per GPU and ~0.4 s for going multi-GPU at all, which a small problem never earns back
    // (GrmEngine::gpus_worth_using).  Metadata errors are left to the engine, which reports them in the reference's order.
    const int gpus_offered = gpus;
    if (gpus > 1 && !gpus_exact) {
        try {
            std::filesystem::path fam = cfg.bed_path, bim = cfg.bed_path;
            fam.replace_extension(".fam");
            bim.replace_extension(".bim");
            const auto n = static_cast<int64_t>(count_total_lines(fam)), m = static_cast<int64_t>(count_total_lines(bim));
            const int kinds = cfg.mode == GrmType::AD ? 2 : 1;
            gpus = GrmEngine::gpus_worth_using(n, m, kinds, cfg.precision, gpus, kinds * (loco ? 22 : 1));  // (22: an estimate; the BIM is parsed later)
        } catch (const std::exception&) {
        }
    }
    cfg.gpus.clear();
    for (int g = 0; g < gpus; ++g) cfg.gpus.push_back(g);
    if (partition == "auto") cfg.partition = Partition::Auto;
    else if (partition == "snp") cfg.partition = Partition::SnpShard;
    else if (partition == "tile") cfg.partition = Partition::Tile;
    else throw ArgumentValidationException("--partition must be auto, snp or tile");
    if (writer_threads < 1) throw ArgumentValidationException("--writer-threads must be >= 1");
    cfg.writer_threads = writer_threads;
    if (writer == "auto") cfg.writer = WriterMode::Auto;
    else if (writer == "pwrite") cfg.writer = WriterMode::Pwrite;
    else if (writer == "mmap") cfg.writer = WriterMode::Mmap;
    else throw ArgumentValidationException("--writer must be auto, pwrite or mmap");
    cfg.write_whole = with_whole;
    cfg.skip_all_missing = skip_missing;
    cfg.gcta = gcta;
    cfg.gcta_diag = gcta_diag;
    cfg.bed_relay = bed_relay;
    cfg.prefault = prefault;

    // CUDA initialises every visible device when the first context is created (several hundred ms each on an 8-GPU box): a run
    // that uses the first `gpus` devices hides the others, unless the caller has already chosen what is visible
    if (std::getenv("CUDA_VISIBLE_DEVICES") == nullptr) {
        std::string vis;
        for (int g = 0; g < gpus; ++g) vis += (g ? "," : "") + std::to_string(g);
        ::setenv("CUDA_VISIBLE_DEVICES", vis.c_str(), 0);
    }

    log.info("gelex-b200 | GRM Computation (B200 tcgen05 engine)");
    log.info("");
    log.info("[Config]");
    log.info(std::string("  Method      : ") + method_name(cfg.method));
    log.info("  Mode        : " + mode_name(cfg.mode));
    log.info(std::string("  LOCO        : ") + (loco ? "yes" : "no"));
    log.info("  Engine      : " + std::to_string(gpus) + " GPU(s)" +
             (gpus < gpus_offered ? " of the " + std::to_string(gpus_offered) + " offered (more would cost more start-up than they save; --gpus-exact overrides)" : std::string()) +
             ", " + precision + " operands");
    log.info("");

    const bool tty = ::isatty(fileno(stderr)) != 0;
    size_t global_total = 0, base = 0, last_pct = 1000;
    GrmEngine engine(cfg);
    engine.compute([&](const GrmEvent& ev) {
        if (const auto* d = std::get_if<GrmDataLoadedEvent>(&ev)) {
            log.info("[Dataset Summary]");
            log.info("  Samples    : " + std::to_string(d->num_samples) + " samples");
            log.info("  SNPs       : " + std::to_string(d->num_snps) + " markers");
            log.info("");
        } else if (const auto* s = std::get_if<GrmComputeStartedEvent>(&ev)) {
            global_total = s->total_snps;
        } else if (const auto* p = std::get_if<GrmProgressEvent>(&ev)) {
            if (p->done) {
                if (tty) std::fprintf(stderr, "\r100.0%% (%zu/%zu)\n", global_total, global_total);
                return;
            }
            const size_t progress = base + p->current;  // grm_reporter.cpp:73-85
            if (p->current == p->total) base += p->total;
            const size_t pct = global_total ? progress * 1000 / global_total : 0;
            if (tty && pct != last_pct) {
                std::fprintf(stderr, "\r%5.1f%% (%zu/%zu)", pct / 10.0, progress, global_total);
                last_pct = pct;
            }
        } else if (const auto* f = std::get_if<GrmFilesWrittenEvent>(&ev)) {
            log.info("[File Summary]");
            log.info("  Num Files : " + std::to_string(f->num_files));
            log.info("  Output Dir  : " + f->output_dir);
            log.info("  Pattern     : " + f->file_pattern);
        }
    });
    const auto& sm = engine.summary();
    char line[256];
    std::snprintf(line, sizeof(line), "  Device      : accumulate %.1f ms (SYRK %.1f ms), read-out + file write %.2f s, engine wall %.2f s", sm.device_pass_ms,
                  sm.syrk_ms, sm.write_s, sm.wall_s);
    log.info(line);
    std::string ph = "  Host phases :";
    for (const auto& p : sm.phases) {
        std::snprintf(line, sizeof(line), " %s %.2f-%.2f s |", p.name.c_str(), p.begin, p.end);
        ph += line;
    }
    if (!sm.phases.empty()) {
        ph.pop_back();
        log.info(ph);
    }
    return 0;
}

}  // namespace

int main(int argc, char** argv) {
    Logger log;
    if (argc < 2 || std::string(argv[1]) == "-h" || std::string(argv[1]) == "--help") {
        std::cout << "Usage: gelex-b200 grm [options]\n\nSubcommands:\n  grm   Compute genomic relationship matrix (the B200-native hot path)\n\n"
                     "The other gelex subcommands (fit, assoc, predict, simulate, post) are not part of this engine.\n";
        return argc < 2 ? 1 : 0;
    }
    if (std::string(argv[1]) != "grm") {
        std::cerr << "[error] unknown subcommand '" << argv[1] << "' (only `grm` is provided by gelex-b200)\n";
        return 1;
    }
    try {
        const auto t0 = std::chrono::steady_clock::now();
        const int rc = run_grm(argc, argv, log);
        const double el = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
        char line[128];
        std::snprintf(line, sizeof(line), "Done in %.2f s", el);
        log.info("");
        log.info(line);
        return rc;
    } catch (const std::exception& e) {
        log.error(e.what());
        return 1;
    } catch (...) {
        log.error("unknown exception");
        return 1;
    }
}
