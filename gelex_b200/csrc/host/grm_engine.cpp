#include "grm_engine.h"

#include <algorithm>
#include <cctype>
#include <chrono>
#include <cstdio>
#include <cstring>
#include <exception>
#include <fstream>
#include <mutex>
#include <numeric>
#include <thread>
#include <unordered_map>

#include "exceptions.h"
#include "plink_meta.h"

namespace fs = std::filesystem;

namespace gelex::b200 {

namespace {

using Clock = std::chrono::steady_clock;
double since(Clock::time_point t0) { return std::chrono::duration<double>(Clock::now() - t0).count(); }

void check(ggrm_t* h, int status) {
    if (status != GGRM_OK) throw_status(status, ggrm_last_error(h));
}

struct HandleGuard {
    ggrm_t* h = nullptr;
    ~HandleGuard() { if (h != nullptr) ggrm_destroy(h); }
};

// .bin: row-major lower triangle incl. diagonal as float32, no header (grm_bin_writer.cpp:54-61).  Rows are fetched
// in file order in blocks of <= 64 MiB and written with one fwrite per block.
void write_bin_file(ggrm_t* h, int kind, int group, int64_t n, const fs::path& path, double* write_s) {
    std::FILE* f = std::fopen(path.c_str(), "wb");
    if (f == nullptr) throw FileOpenException(path.string() + ": failed to open file");
    const int64_t max_elems = int64_t{16} << 20;
    std::vector<float> buf(static_cast<size_t>(std::max<int64_t>(n, std::min<int64_t>(max_elems, n * (n + 1) / 2))));
    int64_t r = 0;
    while (r < n) {
        int64_t cnt = 1;
        while (r + cnt < n && (r + cnt + 1) * (r + cnt + 2) / 2 - r * (r + 1) / 2 <= static_cast<int64_t>(buf.size())) ++cnt;
        const int64_t elems = (r + cnt) * (r + cnt + 1) / 2 - r * (r + 1) / 2;
        const int st = ggrm_fetch(h, kind, group, r, cnt, buf.data(), nullptr);
        if (st != GGRM_OK) {
            std::fclose(f);
            throw_status(st, ggrm_last_error(h));
        }
        const auto t0 = Clock::now();
        const size_t w = std::fwrite(buf.data(), sizeof(float), static_cast<size_t>(elems), f);
        if (write_s != nullptr) *write_s += since(t0);
        if (w != static_cast<size_t>(elems)) {
            std::fclose(f);
            throw FileWriteException(path.string() + ": failed to write GRM data to binary file");
        }
        r += cnt;
    }
    if (std::fclose(f) != 0) throw FileWriteException(path.string() + ": failed to write GRM data to binary file");
}

std::string lower(std::string s) {
    std::transform(s.begin(), s.end(), s.begin(), [](unsigned char c) { return static_cast<char>(std::tolower(c)); });
    return s;
}

}  // namespace

GenotypeProcessMethod parse_genotype_process_method(const std::string& value) {
    static const std::unordered_map<std::string, GenotypeProcessMethod> kMap = {
        {"standardizehwe", GenotypeProcessMethod::StandardizeHWE}, {"sh", GenotypeProcessMethod::StandardizeHWE},
        {"centerhwe", GenotypeProcessMethod::CenterHWE}, {"ch", GenotypeProcessMethod::CenterHWE},
        {"orthstandardizehwe", GenotypeProcessMethod::OrthStandardizeHWE}, {"osh", GenotypeProcessMethod::OrthStandardizeHWE},
        {"orthcenterhwe", GenotypeProcessMethod::OrthCenterHWE}, {"och", GenotypeProcessMethod::OrthCenterHWE},
        {"standardize", GenotypeProcessMethod::Standardize}, {"s", GenotypeProcessMethod::Standardize},
        {"center", GenotypeProcessMethod::Center}, {"c", GenotypeProcessMethod::Center},
        {"orthstandardize", GenotypeProcessMethod::OrthStandardize}, {"os", GenotypeProcessMethod::OrthStandardize},
        {"orthcenter", GenotypeProcessMethod::OrthCenter}, {"oc", GenotypeProcessMethod::OrthCenter},
    };
    const auto it = kMap.find(lower(value));
    if (it == kMap.end())
        throw InvalidInputException("Invalid genotype process method: \"" + value +
                                    "\". Valid: StandardizeHWE(SH), CenterHWE(CH), OrthStandardizeHWE(OSH), OrthCenterHWE(OCH),"
                                    " Standardize(S), Center(C), OrthStandardize(OS), OrthCenter(OC)");
    return it->second;
}

const char* method_name(GenotypeProcessMethod m) {
    switch (m) {
        case GenotypeProcessMethod::StandardizeHWE: return "StandardizeHWE";
        case GenotypeProcessMethod::CenterHWE: return "CenterHWE";
        case GenotypeProcessMethod::OrthStandardizeHWE: return "OrthStandardizeHWE";
        case GenotypeProcessMethod::OrthCenterHWE: return "OrthCenterHWE";
        case GenotypeProcessMethod::Standardize: return "Standardize";
        case GenotypeProcessMethod::Center: return "Center";
        case GenotypeProcessMethod::OrthStandardize: return "OrthStandardize";
        case GenotypeProcessMethod::OrthCenter: return "OrthCenter";
    }
    return "Unknown Genotype Process Method";
}

void write_id_file(const fs::path& path, const std::vector<std::string>& sorted_ids) {
    std::ofstream f(path, std::ios::out | std::ios::trunc);
    if (!f.is_open()) throw FileOpenException(path.string() + ": failed to open file");
    for (const auto& id : sorted_ids) {
        const auto [fid, iid] = split_sample_id(id);
        f << fid << '\t' << iid << '\n';
    }
    f.flush();
    if (!f.good()) throw FileWriteException(path.string() + ": failed to write ID data");
}

std::vector<GrmWorkItem> GrmEngine::build_plan(const std::vector<std::string>& chroms, GrmType mode, bool do_loco,
                                               std::string* output_pattern_suffix, size_t* num_groups) {
    std::vector<std::pair<std::string, bool>> tasks;  // build_tasks, grm_work_plan.cpp:38-50
    if (mode != GrmType::D) tasks.emplace_back("add", true);
    if (mode != GrmType::A) tasks.emplace_back("dom", false);
    const std::string task_pattern = tasks.size() == 1 ? tasks[0].first : std::string("{add|dom}");
    std::vector<GrmWorkItem> items;
    const int64_t m = static_cast<int64_t>(chroms.size());
    if (!do_loco) {
        for (const auto& t : tasks) items.push_back({{{0, m}}, t.second, t.first, -1});
        if (num_groups != nullptr) *num_groups = 0;
        if (output_pattern_suffix != nullptr) *output_pattern_suffix = "." + task_pattern + ".{bin|id}";
    } else {
        const auto runs = chromosome_runs(chroms);
        for (size_t g = 0; g < runs.size(); ++g)
            for (const auto& t : tasks)
                items.push_back({{{runs[g].begin, runs[g].end}}, t.second, t.first + ".chr" + runs[g].name, static_cast<int>(g)});
        if (num_groups != nullptr) *num_groups = runs.size();
        if (output_pattern_suffix != nullptr)
            *output_pattern_suffix = "." + task_pattern + ".chr{1.." + std::to_string(runs.size()) + "}.{bin|id}";
    }
    return items;
}

GrmEngine::GrmEngine(Config config) : config_(std::move(config)) {}

namespace {

struct RunContext {
    const GrmEngine::Config* cfg;
    const SampleOrder* order;
    const BedMapping* bed;
    int64_t n_raw, n_out, m, bpv;
    std::vector<ChrRun> groups;  // one pseudo-group [0, m) when not LOCO
    bool loco;
    bool want_add, want_dom;
    // progress (reference: per work item, counted in SNPs; grm.h:124-130)
    const GrmObserver* observer;
    std::mutex* mu;
    std::vector<std::string>* files;
    double* write_s;
};

ggrm_t* create_handle(const RunContext& rc, int device, const std::vector<int>& my_groups, bool keep_whole) {
    ggrm_config_t c{};
    c.abi_version = GGRM_ABI_VERSION;
    c.device = device;
    c.n_raw = rc.n_raw;
    c.n_out = rc.n_out;
    c.raw_to_out = rc.order->dense ? nullptr : rc.order->raw_to_out.data();
    c.m_total = rc.m;
    c.method = static_cast<int>(rc.cfg->method);
    c.want_add = rc.want_add;
    c.want_dom = rc.want_dom;
    c.precision = rc.cfg->precision;
    std::vector<int64_t> gb, ge;
    for (int g : my_groups) {
        gb.push_back(rc.groups[static_cast<size_t>(g)].begin);
        ge.push_back(rc.groups[static_cast<size_t>(g)].end);
    }
    c.n_groups = static_cast<int>(my_groups.size());
    c.group_begin = gb.data();
    c.group_end = ge.data();
    c.keep_whole = keep_whole ? 1 : 0;
    c.all_missing_policy = rc.cfg->skip_all_missing ? GGRM_ALLMISS_SKIP : GGRM_ALLMISS_REFERENCE_NAN;
    c.tile_rank = 0;
    c.tile_world = 1;
    ggrm_t* h = nullptr;
    const int st = ggrm_create(&c, &h);
    if (st != GGRM_OK) throw_status(st, ggrm_last_error(nullptr));
    return h;
}

void emit_progress(const RunContext& rc, size_t current, size_t total) {
    if (rc.observer == nullptr || !*rc.observer) return;
    std::lock_guard<std::mutex> lk(*rc.mu);
    const int kinds = (rc.want_add ? 1 : 0) + (rc.want_dom ? 1 : 0);
    for (int k = 0; k < kinds; ++k) (*rc.observer)(GrmProgressEvent{current, total, false});
}

// Push SNPs [begin, end) in host chunks of cfg.chunk_size (the reference's load_chunk granularity).
void push_range(const RunContext& rc, ggrm_t* h, int64_t begin, int64_t end, int64_t item_total, int64_t* item_done) {
    for (int64_t s = begin; s < end; s += rc.cfg->chunk_size) {
        const int64_t c = std::min<int64_t>(rc.cfg->chunk_size, end - s);
        if (s < 0 || s + c > rc.m || c <= 0)  // validate_chunk_range, bed_pipe.cpp:38-52
            throw ColumnRangeException("invalid chunk range: [" + std::to_string(s) + ", " + std::to_string(s + c) + "). Total SNPs: " + std::to_string(rc.m));
        const uint8_t* p = rc.bed->chunk_ptr(s, c);
        if (p == nullptr) throw FileFormatException("BedPipe::load_chunk: mapped BED payload is truncated");
        check(h, ggrm_push(h, p, s, c, rc.bpv));
        *item_done += c;
        emit_progress(rc, static_cast<size_t>(*item_done), static_cast<size_t>(item_total));
    }
}

void write_outputs(const RunContext& rc, ggrm_t* h, int local_group, const std::string& suffix /* "" or ".chrNAME" */) {
    for (int kind = 0; kind < 2; ++kind) {
        if (!(kind == 0 ? rc.want_add : rc.want_dom)) continue;
        const std::string base = rc.cfg->out_prefix + "." + (kind == 0 ? "add" : "dom") + suffix;
        double ws = 0;
        write_bin_file(h, kind, local_group, rc.n_out, base + ".bin", &ws);
        write_id_file(base + ".id", rc.order->ids);
        std::lock_guard<std::mutex> lk(*rc.mu);
        rc.files->push_back(base + ".bin");
        rc.files->push_back(base + ".id");
        *rc.write_s += ws;
    }
}

// One GPU, a list of groups (chromosome runs) in SNP order.  Group g is written while the GPU already works on g+1.
void run_groups_on_device(const RunContext& rc, int device, const std::vector<int>& my_groups, bool keep_whole, ggrm_t** keep_handle,
                          double* pass_ms, double* syrk_ms) {
    HandleGuard hg;
    hg.h = create_handle(rc, device, my_groups, keep_whole);
    int pending = -1;  // local index of a group that is complete on the device but not yet written
    for (size_t li = 0; li < my_groups.size(); ++li) {
        const ChrRun& run = rc.groups[static_cast<size_t>(my_groups[li])];
        int64_t done = 0;
        // first chunk of this group makes the previous group fetchable
        const int64_t first_end = std::min<int64_t>(run.end, run.begin + rc.cfg->chunk_size);
        push_range(rc, hg.h, run.begin, first_end, run.end - run.begin, &done);
        if (pending >= 0) {
            write_outputs(rc, hg.h, pending, ".chr" + rc.groups[static_cast<size_t>(my_groups[static_cast<size_t>(pending)])].name);
            for (int kind = 0; kind < 2; ++kind)
                if (kind == 0 ? rc.want_add : rc.want_dom) check(hg.h, ggrm_release_group(hg.h, kind, pending));
            pending = -1;
        }
        push_range(rc, hg.h, first_end, run.end, run.end - run.begin, &done);
        if (rc.loco) pending = static_cast<int>(li);
    }
    check(hg.h, ggrm_finish(hg.h));
    if (rc.loco) {
        if (pending >= 0) write_outputs(rc, hg.h, pending, ".chr" + rc.groups[static_cast<size_t>(my_groups[static_cast<size_t>(pending)])].name);
    }
    double t[8] = {0};
    ggrm_get_timers(hg.h, t);
    {
        std::lock_guard<std::mutex> lk(*rc.mu);
        *pass_ms = std::max(*pass_ms, t[7]);
        *syrk_ms = std::max(*syrk_ms, t[3]);
    }
    if (keep_handle != nullptr) {
        *keep_handle = hg.h;
        hg.h = nullptr;
    }
}

// One GPU, one contiguous SNP shard of the whole-genome matrix (multi-GPU SNP-shard mode); the handle is kept for the reduce.
void run_shard_on_device(const RunContext& rc, int device, int64_t begin, int64_t end, ggrm_t** keep_handle, double* pass_ms, double* syrk_ms) {
    HandleGuard hg;
    hg.h = create_handle(rc, device, {0}, false);
    int64_t done = 0;
    push_range(rc, hg.h, begin, end, end - begin, &done);
    check(hg.h, ggrm_finish(hg.h));
    double t[8] = {0};
    ggrm_get_timers(hg.h, t);
    {
        std::lock_guard<std::mutex> lk(*rc.mu);
        *pass_ms = std::max(*pass_ms, t[7]);
        *syrk_ms = std::max(*syrk_ms, t[3]);
    }
    *keep_handle = hg.h;
    hg.h = nullptr;
}

// Sum the whole-genome accumulators of all handles into handles[0] over NVLink peer memory (no NCCL: one process owns
// every GPU, so peers' device pointers are directly addressable once peer access is enabled).
void reduce_onto_first(const RunContext& rc, const std::vector<ggrm_t*>& handles, const std::vector<int>& devices) {
    if (handles.size() < 2) return;
    for (size_t p = 1; p < devices.size(); ++p) {
        const int st = ggrm_enable_peer_access(devices[0], devices[p]);
        if (st != GGRM_OK) throw_status(st, ggrm_last_error(nullptr));
    }
    for (int kind = 0; kind < 2; ++kind) {
        if (!(kind == 0 ? rc.want_add : rc.want_dom)) continue;
        std::vector<const void*> peers;
        int64_t n_elems = 0;
        for (size_t p = 0; p < handles.size(); ++p) {
            void* ptr = nullptr;
            int64_t ne = 0;
            int32_t eb = 0;
            check(handles[p], ggrm_accum_buffer(handles[p], kind, GGRM_GROUP_WHOLE, &ptr, &ne, &eb));
            if (p == 0) n_elems = ne;
            else peers.push_back(ptr);
        }
        for (size_t off = 0; off < peers.size(); off += 7) {  // ggrm_reduce_peer takes up to 7 peers per call
            const int cnt = static_cast<int>(std::min<size_t>(7, peers.size() - off));
            check(handles[0], ggrm_reduce_peer(handles[0], kind, GGRM_GROUP_WHOLE, peers.data() + off, cnt, 0, n_elems));
        }
    }
}

void run_multi_gpu(const RunContext& rc, double* pass_ms, double* syrk_ms) {
    const std::vector<int>& gpus = rc.cfg->gpus;
    const size_t ng = gpus.size();
    std::vector<ggrm_t*> handles(ng, nullptr);
    std::vector<std::exception_ptr> errors(ng);
    std::vector<std::thread> threads;
    struct Cleanup {
        std::vector<ggrm_t*>* hs;
        ~Cleanup() { for (ggrm_t* h : *hs) if (h != nullptr) ggrm_destroy(h); }
    } cleanup{&handles};
    const bool want_whole = !rc.loco || rc.cfg->write_whole;
    std::vector<std::vector<int>> mine(ng);  // outlives the worker threads (joined below), which read it by reference
    if (rc.loco) {
        // chromosome runs are dealt to GPUs by LPT on their SNP counts: every partial is produced by ONE GPU, no reduce
        std::vector<size_t> order(rc.groups.size());
        std::iota(order.begin(), order.end(), size_t{0});
        std::stable_sort(order.begin(), order.end(), [&](size_t a, size_t b) {
            return rc.groups[a].end - rc.groups[a].begin > rc.groups[b].end - rc.groups[b].begin;
        });
        std::vector<int64_t> load(ng, 0);
        for (size_t g : order) {
            const size_t r = static_cast<size_t>(std::min_element(load.begin(), load.end()) - load.begin());
            mine[r].push_back(static_cast<int>(g));
            load[r] += rc.groups[g].end - rc.groups[g].begin;
        }
        for (size_t r = 0; r < ng; ++r) {
            std::sort(mine[r].begin(), mine[r].end());
            threads.emplace_back([&, r] {
                try {
                    if (!mine[r].empty()) run_groups_on_device(rc, gpus[r], mine[r], want_whole, want_whole ? &handles[r] : nullptr, pass_ms, syrk_ms);
                } catch (...) {
                    errors[r] = std::current_exception();
                }
            });
        }
    } else {
        for (size_t r = 0; r < ng; ++r) {
            const int64_t b = std::min<int64_t>(rc.m, (rc.m * static_cast<int64_t>(r) / static_cast<int64_t>(ng) + 63) / 64 * 64);
            const int64_t e = r + 1 == ng ? rc.m : std::min<int64_t>(rc.m, (rc.m * static_cast<int64_t>(r + 1) / static_cast<int64_t>(ng) + 63) / 64 * 64);
            threads.emplace_back([&, r, b, e] {
                try {
                    run_shard_on_device(rc, gpus[r], b, e, &handles[r], pass_ms, syrk_ms);
                } catch (...) {
                    errors[r] = std::current_exception();
                }
            });
        }
    }
    for (auto& t : threads) t.join();
    for (auto& e : errors) if (e) std::rethrow_exception(e);
    if (want_whole) {
        std::vector<ggrm_t*> live;
        std::vector<int> devs;
        for (size_t r = 0; r < ng; ++r)
            if (handles[r] != nullptr) {
                live.push_back(handles[r]);
                devs.push_back(gpus[r]);
            }
        if (!live.empty()) {
            reduce_onto_first(rc, live, devs);
            write_outputs(rc, live[0], GGRM_GROUP_WHOLE, "");
        }
    }
}

}  // namespace

void GrmEngine::compute(const GrmObserver& observer) {
    const auto t_start = Clock::now();
    summary_ = Summary{};
    if (config_.chunk_size <= 0) throw ArgumentValidationException("chunk_size must be positive");
    if (config_.gpus.empty()) throw ArgumentValidationException("at least one GPU is required (there is no CPU fallback)");
    fs::path fam = config_.bed_path, bim = config_.bed_path;
    fam.replace_extension(".fam");
    bim.replace_extension(".bim");

    // GRM::GRM(bed_path): SampleManager::create_finalized + BedPipe (grm.cpp:36-40, bed_pipe.cpp:69-98)
    const std::vector<std::string> raw_ids = load_fam_ids(fam);
    const SampleOrder order = make_sample_order(raw_ids);
    const int64_t m = static_cast<int64_t>(count_total_lines(bim));
    const int64_t n_raw = static_cast<int64_t>(raw_ids.size());
    const int64_t bpv = (n_raw + 3) / 4;
    const BedMapping bed(config_.bed_path, m, bpv);
    if (observer) observer(GrmDataLoadedEvent{order.ids.size(), static_cast<size_t>(m)});

    const std::vector<std::string> chroms = load_bim_chromosomes(bim);
    std::string pattern_suffix;
    size_t n_groups = 0;
    const std::vector<GrmWorkItem> items = build_plan(chroms, config_.mode, config_.do_loco, &pattern_suffix, &n_groups);
    size_t total_work = 0;
    for (const auto& it : items)
        for (const auto& r : it.ranges) total_work += static_cast<size_t>(r.second - r.first);
    if (observer) observer(GrmComputeStartedEvent{total_work});

    std::mutex mu;
    std::vector<std::string> files;
    double write_s = 0, pass_ms = 0, syrk_ms = 0;
    RunContext rc{};
    rc.cfg = &config_;
    rc.order = &order;
    rc.bed = &bed;
    rc.n_raw = n_raw;
    rc.n_out = static_cast<int64_t>(order.ids.size());
    rc.m = m;
    rc.bpv = bpv;
    rc.loco = config_.do_loco;
    rc.groups = config_.do_loco ? chromosome_runs(chroms) : std::vector<ChrRun>{{"", 0, static_cast<int64_t>(chroms.size())}};
    rc.want_add = config_.mode != GrmType::D;
    rc.want_dom = config_.mode != GrmType::A;
    rc.observer = &observer;
    rc.mu = &mu;
    rc.files = &files;
    rc.write_s = &write_s;

    const int n_gpus = static_cast<int>(config_.gpus.size());
    if (items.empty() || rc.groups.empty() || chroms.empty()) {
        // nothing to accumulate (empty BIM is rejected earlier as "is empty")
    } else if (n_gpus == 1) {
        std::vector<int> all(rc.groups.size());
        std::iota(all.begin(), all.end(), 0);
        ggrm_t* h = nullptr;
        const bool need_handle = !config_.do_loco || config_.write_whole;
        run_groups_on_device(rc, config_.gpus[0], all, config_.do_loco && config_.write_whole, need_handle ? &h : nullptr, &pass_ms, &syrk_ms);
        HandleGuard hg;
        hg.h = h;
        if (need_handle) write_outputs(rc, h, GGRM_GROUP_WHOLE, "");
    } else {
        run_multi_gpu(rc, &pass_ms, &syrk_ms);
    }

    if (observer) {
        observer(GrmProgressEvent{total_work, total_work, true});
        observer(GrmFilesWrittenEvent{items.size() * 2, fs::absolute(fs::path(config_.out_prefix)).parent_path().string(),
                                      config_.out_prefix + pattern_suffix});
    }
    summary_.num_samples = order.ids.size();
    summary_.num_snps = static_cast<size_t>(m);
    summary_.files = files;
    summary_.num_files = files.size();
    summary_.write_s = write_s;
    summary_.device_pass_ms = pass_ms;
    summary_.syrk_ms = syrk_ms;
    summary_.wall_s = since(t_start);
}

}  // namespace gelex::b200
