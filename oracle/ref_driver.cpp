// TEST INFRASTRUCTURE ONLY.  Driver around the UNMODIFIED reference sources of the `gelex grm`
// path (compiled from /root/reference by oracle/Makefile into oracle/_ref/gelex_ref).
// Nothing in the product (gelex_b200/, include/) links, loads or executes this program; it is
// used by tests (golden-vector generation, parity checks) and by bench.py's CPU-baseline legs.
//
// Sub-commands
//   grm     <bfile> <out> <method 1..8> <A|D|AD> <loco 0|1> <chunk> <threads>
//           -> GrmEngine::compute (src/pipeline/grm_engine.cpp:52-124), prints one JSON line with wall s
//   stages  <bfile> <method> <A|D> <chunk> <threads>
//           -> the chunk loop of grm.h:111-132 re-timed stage by stage (decode / standardise / dsyrk)
//   dense   <bfile> <method> <A|D> <chunk> <threads> <out.f64>
//           -> GRM::compute<GT> (include/gelex/data/grm/grm.h:94-137), the FP64 result BEFORE the float32 cast of the writer:
//              dumps n (int64), denominator (double), then the packed lower triangle row-major as doubles
//   chunk   <bfile> <start> <end> <method> <A|D> <out.bin>
//           -> BedPipe::load_chunk + process_matrix; dumps n, c, raw doubles, freqs, processed doubles
//   process <method> <A|D> v0 v1 ...   (values 0,1,2 or nan) -> prints processed column, maf, mean, sd
//   loco    <whole_prefix> <chr_prefix> <out.bin>
//           -> LocoGRMLoader arithmetic (src/data/grm/loco_grm_loader.cpp:62-76) dumped as doubles
//   load    <prefix> <out.bin>  -> GrmLoader::load (normalised, src/data/grm/grm_loader.cpp:125-165)
//   pipe    <bfile> <out_prefix> <method> <A|D> <chunk>
//           -> GenotypePipe(bed, SampleManager::create_finalized(bed), out_prefix).process<GT>(method, chunk)
//              (include/gelex/data/genotype/genotype_pipe.h:55-89): writes <out_prefix>.bmat and <out_prefix>.snpstats
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <iostream>
#include <string>
#include <vector>

#include <omp.h>
#include <Eigen/Core>

#include "ref_shim/cblas.h"

#include "gelex/data/genotype/bed_pipe.h"
#include "gelex/data/genotype/genotype_pipe.h"
#include "gelex/infra/logger.h"
#include "gelex/data/genotype/genotype_processor.h"
#include "gelex/data/genotype/sample_manager.h"
#include "gelex/data/grm/grm.h"
#include "gelex/data/grm/grm_loader.h"
#include "gelex/data/grm/loco_grm_loader.h"
#include "gelex/pipeline/grm_engine.h"

extern "C" void scipy_openblas_set_num_threads(int) __attribute__((weak));
extern "C" char* scipy_openblas_get_config(void) __attribute__((weak));
extern "C" void openblas_set_num_threads(int) __attribute__((weak));
extern "C" char* openblas_get_config(void) __attribute__((weak));

// gelex::logging::get() without spdlog (oracle/ref_shim/spdlog/logger.h): the genotype sinks log an error line before they throw.
namespace gelex::logging
{
std::shared_ptr<spdlog::logger>& get()
{
    static std::shared_ptr<spdlog::logger> l = std::make_shared<spdlog::logger>();
    return l;
}
}  // namespace gelex::logging

namespace
{
using Clock = std::chrono::steady_clock;
double secs(Clock::time_point a, Clock::time_point b)
{
    return std::chrono::duration<double>(b - a).count();
}

void set_threads(int t)
{
    if (t <= 0) { return; }
    omp_set_num_threads(t);
    Eigen::setNbThreads(t);
    if (scipy_openblas_set_num_threads) { scipy_openblas_set_num_threads(t); }
    if (openblas_set_num_threads) { openblas_set_num_threads(t); }
}

std::string blas_name()
{
    if (scipy_openblas_get_config) { return scipy_openblas_get_config(); }
    if (openblas_get_config) { return openblas_get_config(); }
    return "unknown";
}

gelex::GenotypeProcessMethod method_of(const char* s)
{
    int v = std::atoi(s);
    if (v < 1 || v > 8) { throw std::runtime_error("method must be 1..8"); }
    return static_cast<gelex::GenotypeProcessMethod>(v);
}

void write_doubles(std::ofstream& f, const double* p, size_t n)
{
    f.write(reinterpret_cast<const char*>(p), static_cast<std::streamsize>(n * sizeof(double)));
}

int cmd_grm(int argc, char** argv)
{
    if (argc < 9) { return 2; }
    gelex::GrmEngine::Config cfg;
    cfg.bed_path = std::string(argv[2]) + ".bed";
    cfg.out_prefix = argv[3];
    cfg.method = method_of(argv[4]);
    std::string mode = argv[5];
    cfg.mode = mode == "AD" ? gelex::freq::GrmType::AD
                            : (mode == "D" ? gelex::freq::GrmType::D : gelex::freq::GrmType::A);
    cfg.do_loco = std::atoi(argv[6]) != 0;
    cfg.chunk_size = std::atoi(argv[7]);
    int threads = std::atoi(argv[8]);
    set_threads(threads);
    size_t n = 0, m = 0, total_snps = 0, num_files = 0, progress_events = 0, progress_sum = 0;
    std::string file_pattern;
    auto t0 = Clock::now();
    gelex::GrmEngine engine(cfg);
    engine.compute(
        [&](const gelex::GrmEvent& ev)
        {
            if (const auto* d = std::get_if<gelex::GrmDataLoadedEvent>(&ev))
            {
                n = d->num_samples;
                m = d->num_snps;
            }
            else if (const auto* s = std::get_if<gelex::GrmComputeStartedEvent>(&ev)) { total_snps = s->total_snps; }
            else if (const auto* p = std::get_if<gelex::GrmProgressEvent>(&ev))
            {
                if (!p->done)
                {
                    ++progress_events;
                    progress_sum += p->current;
                }
            }
            else if (const auto* f = std::get_if<gelex::GrmFilesWrittenEvent>(&ev))
            {
                num_files = f->num_files;
                file_pattern = f->file_pattern;
            }
        });
    auto t1 = Clock::now();
    // events as the CLI reporter receives them (grm_event.h:28-67): totals, the files-written summary, and the per-chunk progress
    // stream reduced to its count and the sum of `current`
    std::printf(
        "{\"cmd\":\"grm\",\"n\":%zu,\"m\":%zu,\"wall_s\":%.6f,\"threads\":%d,\"blas\":\"%s\",\"total_snps\":%zu,\"num_files\":%zu,"
        "\"file_pattern\":\"%s\",\"progress_events\":%zu,\"progress_sum\":%zu}\n",
        n, m, secs(t0, t1), omp_get_max_threads(), blas_name().c_str(), total_snps, num_files, file_pattern.c_str(), progress_events,
        progress_sum);
    return 0;
}

template <gelex::GeneticEffectType GT>
void stages_impl(gelex::BedPipe& bed, gelex::GenotypeProcessMethod method, Eigen::Index chunk,
                 double& t_dec, double& t_std, double& t_syrk, Eigen::MatrixXd& grm)
{
    const Eigen::Index n = bed.num_samples();
    const Eigen::Index m = bed.num_snps();
    grm = Eigen::MatrixXd::Zero(n, n);
    for (Eigen::Index s = 0; s < m; s += chunk)
    {
        const Eigen::Index e = std::min(s + chunk, m);
        auto a = Clock::now();
        Eigen::MatrixXd g = bed.load_chunk(s, e);
        auto b = Clock::now();
        gelex::process_matrix<GT>(method, g);
        auto c = Clock::now();
        cblas_dsyrk(CblasColMajor, CblasLower, CblasNoTrans, static_cast<int>(n),
                    static_cast<int>(e - s), 1.0, g.data(), static_cast<int>(n), 1.0, grm.data(),
                    static_cast<int>(n));
        auto d = Clock::now();
        t_dec += secs(a, b);
        t_std += secs(b, c);
        t_syrk += secs(c, d);
    }
}

int cmd_stages(int argc, char** argv)
{
    if (argc < 7) { return 2; }
    std::string bed_path = std::string(argv[2]) + ".bed";
    auto method = method_of(argv[3]);
    bool add = std::string(argv[4]) != "D";
    Eigen::Index chunk = std::atol(argv[5]);
    set_threads(std::atoi(argv[6]));
    auto sm = gelex::SampleManager::create_finalized(bed_path);
    gelex::BedPipe bed(bed_path, sm);
    double t_dec = 0, t_std = 0, t_syrk = 0;
    Eigen::MatrixXd grm;
    auto t0 = Clock::now();
    if (add) { stages_impl<gelex::GeneticEffectType::Add>(bed, method, chunk, t_dec, t_std, t_syrk, grm); }
    else { stages_impl<gelex::GeneticEffectType::Dom>(bed, method, chunk, t_dec, t_std, t_syrk, grm); }
    auto t1 = Clock::now();
    const double n = static_cast<double>(bed.num_samples());
    const double m = static_cast<double>(bed.num_snps());
    std::printf(
        "{\"cmd\":\"stages\",\"n\":%.0f,\"m\":%.0f,\"wall_s\":%.6f,\"decode_s\":%.6f,\"standardise_s\":%.6f,"
        "\"dsyrk_s\":%.6f,\"flops\":%.6e,\"trace_over_n\":%.10e,\"threads\":%d,\"blas\":\"%s\"}\n",
        n, m, secs(t0, t1), t_dec, t_std, t_syrk, n * (n + 1.0) * m, grm.trace() / n,
        omp_get_max_threads(), blas_name().c_str());
    return 0;
}

int cmd_dense(int argc, char** argv)
{
    if (argc < 8) { return 2; }
    std::string bed_path = std::string(argv[2]) + ".bed";
    auto method = method_of(argv[3]);
    bool add = std::string(argv[4]) != "D";
    Eigen::Index chunk = std::atol(argv[5]);
    set_threads(std::atoi(argv[6]));
    gelex::GRM grm(bed_path);
    auto t0 = Clock::now();
    gelex::GrmResult res = add ? grm.compute<gelex::GeneticEffectType::Add>(method, chunk)
                               : grm.compute<gelex::GeneticEffectType::Dom>(method, chunk);
    auto t1 = Clock::now();
    std::ofstream f(argv[7], std::ios::binary | std::ios::trunc);
    const int64_t n = res.grm.rows();
    f.write(reinterpret_cast<const char*>(&n), sizeof(n));
    f.write(reinterpret_cast<const char*>(&res.denominator), sizeof(double));
    std::vector<double> row(static_cast<size_t>(n));
    for (int64_t i = 0; i < n; ++i)
    {
        for (int64_t j = 0; j <= i; ++j) { row[static_cast<size_t>(j)] = res.grm(i, j); }
        write_doubles(f, row.data(), static_cast<size_t>(i + 1));
    }
    std::printf("{\"cmd\":\"dense\",\"n\":%lld,\"wall_s\":%.6f,\"denominator\":%.17g}\n", static_cast<long long>(n), secs(t0, t1),
                res.denominator);
    return f.good() ? 0 : 1;
}

int cmd_chunk(int argc, char** argv)
{
    if (argc < 8) { return 2; }
    std::string bed_path = std::string(argv[2]) + ".bed";
    Eigen::Index start = std::atol(argv[3]);
    Eigen::Index end = std::atol(argv[4]);
    auto method = method_of(argv[5]);
    bool add = std::string(argv[6]) != "D";
    auto sm = gelex::SampleManager::create_finalized(bed_path);
    gelex::BedPipe bed(bed_path, sm);
    Eigen::MatrixXd raw = bed.load_chunk(start, end);
    Eigen::MatrixXd z = raw;
    Eigen::VectorXd freqs(z.cols());
    if (add) { gelex::process_matrix<gelex::GeneticEffectType::Add>(method, z, &freqs); }
    else { gelex::process_matrix<gelex::GeneticEffectType::Dom>(method, z, &freqs); }
    std::ofstream f(argv[7], std::ios::binary | std::ios::trunc);
    int64_t hdr[2] = {raw.rows(), raw.cols()};
    f.write(reinterpret_cast<const char*>(hdr), sizeof(hdr));
    write_doubles(f, raw.data(), static_cast<size_t>(raw.size()));
    write_doubles(f, freqs.data(), static_cast<size_t>(freqs.size()));
    write_doubles(f, z.data(), static_cast<size_t>(z.size()));
    return f.good() ? 0 : 1;
}

int cmd_process(int argc, char** argv)
{
    if (argc < 5) { return 2; }
    auto method = method_of(argv[2]);
    bool add = std::string(argv[3]) != "D";
    Eigen::VectorXd v(argc - 4);
    for (int i = 4; i < argc; ++i) { v(i - 4) = std::strtod(argv[i], nullptr); }
    gelex::LocusStatistic st = add ? gelex::get_genotype_process_method<gelex::GeneticEffectType::Add>(method)(v)
                                   : gelex::get_genotype_process_method<gelex::GeneticEffectType::Dom>(method)(v);
    std::printf("{\"maf\":%.17g,\"mean\":%.17g,\"stddev\":%.17g,\"mono\":%d,\"z\":[", st.maf, st.mean,
                st.stddev, st.is_monomorphic ? 1 : 0);
    for (Eigen::Index i = 0; i < v.size(); ++i)
    {
        if (std::isnan(v(i))) { std::printf("%s\"nan\"", i ? "," : ""); }
        else { std::printf("%s%.17g", i ? "," : "", v(i)); }
    }
    std::printf("]}\n");
    return 0;
}

std::unordered_map<std::string, Eigen::Index> identity_map(const std::string& prefix)
{
    std::unordered_map<std::string, Eigen::Index> m;
    std::ifstream f(prefix + ".id");
    std::string line;
    Eigen::Index i = 0;
    while (std::getline(f, line))
    {
        auto tab = line.find('\t');
        m.emplace(line.substr(0, tab) + '\x1F' + line.substr(tab + 1), i++);
    }
    return m;
}

int cmd_loco(int argc, char** argv)
{
    if (argc < 5) { return 2; }
    auto ids = identity_map(argv[2]);
    gelex::LocoGRMLoader loader(argv[2], ids);
    Eigen::MatrixXd out;
    loader.load_loco_grm(argv[3], ids, out);
    std::ofstream f(argv[4], std::ios::binary | std::ios::trunc);
    int64_t hdr[2] = {out.rows(), out.cols()};
    f.write(reinterpret_cast<const char*>(hdr), sizeof(hdr));
    write_doubles(f, out.data(), static_cast<size_t>(out.size()));
    return f.good() ? 0 : 1;
}

int cmd_pipe(int argc, char** argv)
{
    if (argc < 7) { return 2; }
    std::string bed_path = std::string(argv[2]) + ".bed";
    auto method = method_of(argv[4]);
    bool add = std::string(argv[5]) != "D";
    size_t chunk = static_cast<size_t>(std::atol(argv[6]));
    // process() ends in finalize(), which maps the files it has just written through GenotypeMap while the two writers are still
    // open: with outputs smaller than the stream buffers nothing has reached the file yet and the unmodified reference throws
    // "Failed to read scalar: rows" there.  Both files are complete once the pipe is destroyed (the writers flush), so that
    // exception is reported, not fatal; constructor errors (existing outputs, bad inputs) propagate as they are.
    std::string finalize_error;
    {
        gelex::GenotypePipe pipe(bed_path, gelex::SampleManager::create_finalized(bed_path), argv[3]);
        try
        {
            if (add) { pipe.process<gelex::GeneticEffectType::Add>(method, chunk); }
            else { pipe.process<gelex::GeneticEffectType::Dom>(method, chunk); }
        }
        catch (const std::exception& e)
        {
            finalize_error = e.what();
        }
    }
    std::printf("{\"cmd\":\"pipe\",\"finalize_error\":\"%s\"}\n", finalize_error.c_str());
    return 0;
}

int cmd_load(int argc, char** argv)
{
    if (argc < 4) { return 2; }
    gelex::detail::GrmLoader loader(argv[2]);
    Eigen::MatrixXd out = loader.load();
    std::ofstream f(argv[3], std::ios::binary | std::ios::trunc);
    int64_t hdr[2] = {out.rows(), out.cols()};
    f.write(reinterpret_cast<const char*>(hdr), sizeof(hdr));
    write_doubles(f, out.data(), static_cast<size_t>(out.size()));
    return f.good() ? 0 : 1;
}
}  // namespace

int main(int argc, char** argv)
{
    if (argc < 2)
    {
        std::fprintf(stderr, "usage: gelex_ref grm|stages|dense|chunk|process|loco|load|pipe ...\n");
        return 2;
    }
    try
    {
        std::string c = argv[1];
        if (c == "grm") { return cmd_grm(argc, argv); }
        if (c == "stages") { return cmd_stages(argc, argv); }
        if (c == "dense") { return cmd_dense(argc, argv); }
        if (c == "chunk") { return cmd_chunk(argc, argv); }
        if (c == "process") { return cmd_process(argc, argv); }
        if (c == "loco") { return cmd_loco(argc, argv); }
        if (c == "load") { return cmd_load(argc, argv); }
        if (c == "pipe") { return cmd_pipe(argc, argv); }
        std::fprintf(stderr, "unknown command %s\n", c.c_str());
        return 2;
    }
    catch (const std::exception& e)
    {
        std::fprintf(stderr, "error: %s\n", e.what());
        return 1;
    }
}
