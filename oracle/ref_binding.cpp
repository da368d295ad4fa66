// TEST INFRASTRUCTURE ONLY.  The reference-side binding of INTEGRATION.md §2, compiled against the UNMODIFIED reference:
// gelex's own SampleManager, BedPipe (projection + mmap reader), GrmNormalPlan / GrmLocoPlan, GrmBinWriter and GrmIdWriter do
// everything they do in `gelex grm`; only the chunk loop of GRM::compute<GT> (include/gelex/data/grm/grm.h:94-137:
// load_chunk -> process_matrix -> update_grm) is replaced by the C ABI of include/ggrm.h.  Built by oracle/Makefile into
// oracle/_ref/gelex_ref_b200 (needs /root/reference to build, a B200 to run); tests/test_host_engine.py compares its files with
// those of the plain reference.  Nothing in the product links or runs this program.
//
//   gelex_ref_b200 <bfile> <out_prefix> <method 1..8> <A|D|AD> <loco 0|1> <chunk> <fp16|bf16|fp64>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <filesystem>
#include <format>
#include <memory>
#include <string>
#include <vector>

#include <Eigen/Core>

#include "../include/ggrm.h"

#include "gelex/data/genotype/sample_manager.h"
#include "gelex/exception.h"
#include "gelex/types/freq_effect.h"

// BedPipe keeps its projection and its mmap reader private (bed_pipe.h:66-70).  A maintainer adds three one-line accessors
// (INTEGRATION.md §2); this test program reaches the same members without touching the reference's sources.
#include "data/bed_pipe/mmap_reader.h"
#include "data/bed_pipe/sample_projection.h"
#include "data/bed_pipe/variant_decoder.h"
#define private public
#include "gelex/data/genotype/bed_pipe.h"
#undef private

#include "gelex/data/grm/grm.h"
#include "gelex/data/grm/grm_bin_writer.h"
#include "gelex/data/grm/grm_id_writer.h"
#include "pipeline/grm_work_plan.h"

namespace
{

[[noreturn]] void raise(int status, const ggrm_t* h)  // ggrm_status -> GelexException subclass (include/ggrm.h:27-38)
{
    const std::string msg = ggrm_last_error(h);
    switch (status)
    {
        case GGRM_ERR_INVALID_ARGUMENT: throw gelex::ArgumentValidationException(msg);
        case GGRM_ERR_COLUMN_RANGE: throw gelex::ColumnRangeException(msg);
        case GGRM_ERR_INVALID_OPERATION: throw gelex::InvalidOperationException(msg);
        case GGRM_ERR_OUT_OF_MEMORY: throw std::bad_alloc();
        case GGRM_ERR_FILE_FORMAT: throw gelex::FileFormatException(msg);
        default: throw gelex::GelexException(msg);
    }
}

// GRM::compute<GT> with the chunk loop on the GPU: same arguments, same GrmResult.
gelex::GrmResult compute_b200(
    const gelex::BedPipe& bed,
    gelex::GenotypeProcessMethod method,
    bool additive,
    const std::vector<std::pair<Eigen::Index, Eigen::Index>>& ranges,
    Eigen::Index chunk_size,
    int precision)
{
    const auto& proj = *bed.projection_;
    const auto& reader = *bed.bed_reader_;
    const std::vector<int32_t> raw_to_out(proj.mapping().begin(), proj.mapping().end());  // Eigen::Index -> the ABI's int32
    const Eigen::Index n = bed.num_samples();

    ggrm_config_t cfg{};
    cfg.abi_version = GGRM_ABI_VERSION;
    cfg.n_raw = static_cast<int64_t>(raw_to_out.size());
    cfg.n_out = n;
    cfg.raw_to_out = proj.is_dense() ? nullptr : raw_to_out.data();
    cfg.m_total = bed.num_snps();
    cfg.method = static_cast<int>(method);  // enum values are identical (genotype_processor.h:38-48)
    cfg.want_add = additive ? 1 : 0;
    cfg.want_dom = additive ? 0 : 1;
    cfg.precision = precision;
    cfg.n_groups = 1;
    cfg.tile_world = 1;
    const int64_t gb = ranges.front().first, ge = ranges.back().second;
    cfg.group_begin = &gb;
    cfg.group_end = &ge;

    ggrm_t* h = nullptr;
    if (int st = ggrm_create(&cfg, &h); st != GGRM_OK) { raise(st, nullptr); }
    std::unique_ptr<ggrm_t, void (*)(ggrm_t*)> guard(h, ggrm_destroy);

    for (auto [rs, re] : ranges)
    {
        for (Eigen::Index s = rs; s < re; s += chunk_size)
        {
            const Eigen::Index c = std::min(chunk_size, re - s);
            // the bytes BedMmapReader::chunk_ptr(s, c) points at (mmap_reader.cpp:96-119); the caller keeps ownership
            if (int st = ggrm_push(h, reader.chunk_ptr(s, c), s, c, reader.bytes_per_variant()); st != GGRM_OK) { raise(st, h); }
        }
    }
    if (int st = ggrm_finish(h); st != GGRM_OK) { raise(st, h); }

    std::vector<double> rows(static_cast<size_t>(n) * static_cast<size_t>(n + 1) / 2);
    double trace = 0;
    if (int st = ggrm_fetch_f64(h, additive ? GGRM_KIND_ADD : GGRM_KIND_DOM, GGRM_GROUP_WHOLE, 0, n, rows.data(), &trace); st != GGRM_OK)
    {
        raise(st, h);
    }
    Eigen::MatrixXd G = Eigen::MatrixXd::Zero(n, n);
    size_t k = 0;
    for (Eigen::Index i = 0; i < n; ++i)
    {
        for (Eigen::Index j = 0; j <= i; ++j) { G(i, j) = rows[k++]; }
    }
    return {std::move(G), trace / static_cast<double>(n)};
}

}  // namespace

int main(int argc, char** argv)
{
    if (argc < 8)
    {
        std::fprintf(stderr, "usage: gelex_ref_b200 <bfile> <out_prefix> <method 1..8> <A|D|AD> <loco 0|1> <chunk> <fp16|bf16|fp64>\n");
        return 2;
    }
    try
    {
        const std::filesystem::path bed_path = std::string(argv[1]) + ".bed";
        const std::string out_prefix = argv[2];
        const auto method = static_cast<gelex::GenotypeProcessMethod>(std::atoi(argv[3]));
        const std::string mode_s = argv[4];
        const auto mode = mode_s == "AD" ? gelex::freq::GrmType::AD : (mode_s == "D" ? gelex::freq::GrmType::D : gelex::freq::GrmType::A);
        const bool loco = std::atoi(argv[5]) != 0;
        const Eigen::Index chunk = std::atol(argv[6]);
        const std::string prec_s = argv[7];
        const int precision = prec_s == "fp64" ? GGRM_PREC_FP64 : (prec_s == "bf16" ? GGRM_PREC_BF16 : GGRM_PREC_FP16);

        // GRM::GRM (grm.cpp:36-40)
        auto sample_manager = gelex::SampleManager::create_finalized(bed_path);
        gelex::BedPipe bed(bed_path, sample_manager);
        const auto& sample_ids = sample_manager->common_ids();

        auto bim_path = bed_path;
        bim_path.replace_extension(".bim");
        auto run_plan = [&](const auto& plan)  // GrmEngine::compute, grm_engine.cpp:79-113, minus the events
        {
            for (const auto& item : plan.items())
            {
                const gelex::GrmResult result = compute_b200(bed, method, item.is_additive, item.ranges, chunk, precision);
                const std::string path = std::format("{}.{}", out_prefix, item.output_name);
                gelex::GrmBinWriter(path + ".bin").write(result.grm);
                gelex::GrmIdWriter(path + ".id").write(sample_ids);
            }
            std::printf("{\"cmd\":\"grm_b200\",\"items\":%zu,\"n\":%zu}\n", plan.items().size(), sample_ids.size());
        };
        if (loco) { run_plan(gelex::GrmLocoPlan(bim_path, mode)); }
        else { run_plan(gelex::GrmNormalPlan(bim_path, mode)); }
        return 0;
    }
    catch (const std::exception& e)
    {
        std::fprintf(stderr, "error: %s\n", e.what());
        return 1;
    }
}
