// C entry point over the C++ host engine (for ctypes tests and for embedding); never throws.
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "exceptions.h"
#include "genotype_pipe.h"
#include "grm_engine.h"
#include "grm_loader.h"
#include "plink_meta.h"

using namespace gelex::b200;

static void set_err(char* err, int err_len, const std::string& m) {
    if (err == nullptr || err_len <= 0) return;
    std::strncpy(err, m.c_str(), static_cast<size_t>(err_len - 1));
    err[err_len - 1] = '\0';
}

// Current exception -> (status, "Type: message"); call inside a catch (...) block.
static int translate_exception(char* err, int err_len) {
    try {
        throw;
    } catch (const FileNotFoundException& e) { set_err(err, err_len, std::string("FileNotFoundException: ") + e.what()); return -10;
    } catch (const FileOpenException& e) { set_err(err, err_len, std::string("FileOpenException: ") + e.what()); return -11;
    } catch (const FileWriteException& e) { set_err(err, err_len, std::string("FileWriteException: ") + e.what()); return -12;
    } catch (const FileExistsException& e) { set_err(err, err_len, std::string("FileExistsException: ") + e.what()); return -13;
    } catch (const FileFormatException& e) { set_err(err, err_len, std::string("FileFormatException: ") + e.what()); return GGRM_ERR_FILE_FORMAT;
    } catch (const ColumnRangeException& e) { set_err(err, err_len, std::string("ColumnRangeException: ") + e.what()); return GGRM_ERR_COLUMN_RANGE;
    } catch (const ArgumentValidationException& e) { set_err(err, err_len, std::string("ArgumentValidationException: ") + e.what()); return GGRM_ERR_INVALID_ARGUMENT;
    } catch (const InvalidInputException& e) { set_err(err, err_len, std::string("InvalidInputException: ") + e.what()); return GGRM_ERR_INVALID_ARGUMENT;
    } catch (const InvalidOperationException& e) { set_err(err, err_len, std::string("InvalidOperationException: ") + e.what()); return GGRM_ERR_INVALID_OPERATION;
    } catch (const GelexException& e) { set_err(err, err_len, std::string("GelexException: ") + e.what()); return GGRM_ERR_CUDA;
    } catch (const std::bad_alloc&) { set_err(err, err_len, "std::bad_alloc"); return GGRM_ERR_OUT_OF_MEMORY;
    } catch (const std::exception& e) { set_err(err, err_len, std::string("std::exception: ") + e.what()); return -99; }
}

extern "C" int gelex_b200_grm_run(const char* bfile, const char* out_prefix, int method, int mode, int do_loco, int chunk_size,
                                  int precision, int n_gpus, int write_whole, char* err, int err_len) {
    return gelex_b200_grm_run_ex(bfile, out_prefix, method, mode, do_loco, chunk_size, precision, n_gpus, write_whole, 0, 4, err, err_len);
}

extern "C" int gelex_b200_grm_run_ex(const char* bfile, const char* out_prefix, int method, int mode, int do_loco, int chunk_size,
                                     int precision, int n_gpus, int write_whole, int partition, int writer_threads, char* err, int err_len) {
    try {
        if (bfile == nullptr || out_prefix == nullptr) throw ArgumentValidationException("null argument");
        if (method < 1 || method > 8) throw InvalidInputException("Invalid genotype process method.");
        GrmEngine::Config cfg;
        cfg.bed_path = format_bed_path(bfile);
        cfg.mode = mode == 2 ? GrmType::AD : (mode == 1 ? GrmType::D : GrmType::A);
        cfg.method = static_cast<GenotypeProcessMethod>(method);
        cfg.do_loco = do_loco != 0;
        cfg.out_prefix = out_prefix;
        cfg.chunk_size = chunk_size;
        cfg.precision = precision;
        cfg.write_whole = write_whole != 0;
        if (partition < 0 || partition > 2) throw ArgumentValidationException("partition must be 0 (auto), 1 (snp) or 2 (tile)");
        cfg.partition = static_cast<Partition>(partition);
        cfg.writer_threads = writer_threads;
        cfg.gpus.clear();
        // Test aid: GELEX_B200_OVERSUBSCRIBE=1 maps rank i to device i % device_count, so the multi-GPU decompositions (several
        // handles, peer reduce, concurrent writers) can be exercised on a one-GPU box.  Never set by the CLI.
        const char* over = std::getenv("GELEX_B200_OVERSUBSCRIBE");
        const int devs = ggrm_device_count();
        for (int i = 0; i < (n_gpus > 0 ? n_gpus : 1); ++i)
            cfg.gpus.push_back(over != nullptr && over[0] == '1' && devs > 0 ? i % devs : i);
        GrmEngine engine(cfg);
        engine.compute();
        set_err(err, err_len, "");
        return 0;
    } catch (...) {
        return translate_exception(err, err_len);
    }
}

// gelex_b200_grm_run_ex with the observer's events recorded: events3 receives up to cap triples (type, a, b) in the order the observer
// saw them - 0 GrmDataLoadedEvent (samples, snps), 1 GrmComputeStartedEvent (total_snps, 0), 2 GrmProgressEvent (current, total),
// 3 GrmProgressEvent with done (current, total), 4 GrmFilesWrittenEvent (num_files, 0) - and *n_events their number.
extern "C" int gelex_b200_grm_run_events(const char* bfile, const char* out_prefix, int method, int mode, int do_loco, int chunk_size,
                                         int precision, int n_gpus, int64_t* events3, int64_t cap, int64_t* n_events, char* err, int err_len) {
    try {
        if (bfile == nullptr || out_prefix == nullptr || n_events == nullptr) throw ArgumentValidationException("null argument");
        if (method < 1 || method > 8) throw InvalidInputException("Invalid genotype process method.");
        GrmEngine::Config cfg;
        cfg.bed_path = format_bed_path(bfile);
        cfg.mode = mode == 2 ? GrmType::AD : (mode == 1 ? GrmType::D : GrmType::A);
        cfg.method = static_cast<GenotypeProcessMethod>(method);
        cfg.do_loco = do_loco != 0;
        cfg.out_prefix = out_prefix;
        cfg.chunk_size = chunk_size;
        cfg.precision = precision;
        cfg.gpus.clear();
        const char* over = std::getenv("GELEX_B200_OVERSUBSCRIBE");
        const int devs = ggrm_device_count();
        for (int i = 0; i < (n_gpus > 0 ? n_gpus : 1); ++i) cfg.gpus.push_back(over != nullptr && over[0] == '1' && devs > 0 ? i % devs : i);
        int64_t count = 0;
        auto put = [&](int64_t type, size_t a, size_t b) {
            if (events3 != nullptr && count < cap) {
                events3[3 * count] = type;
                events3[3 * count + 1] = static_cast<int64_t>(a);
                events3[3 * count + 2] = static_cast<int64_t>(b);
            }
            ++count;
        };
        GrmEngine engine(cfg);
        engine.compute([&](const GrmEvent& ev) {
            if (const auto* e = std::get_if<GrmDataLoadedEvent>(&ev)) put(0, e->num_samples, e->num_snps);
            else if (const auto* e = std::get_if<GrmComputeStartedEvent>(&ev)) put(1, e->total_snps, 0);
            else if (const auto* e = std::get_if<GrmProgressEvent>(&ev)) put(e->done ? 3 : 2, e->current, e->total);
            else if (const auto* e = std::get_if<GrmFilesWrittenEvent>(&ev)) put(4, e->num_files, 0);
        });
        *n_events = count;
        set_err(err, err_len, "");
        return 0;
    } catch (...) {
        return translate_exception(err, err_len);
    }
}

// Metadata-only entry (no GPU needed): sample order, projection and chromosome runs as the engine will use them.
// ids_out: n_out lines "FID\tIID\n" (the .id payload) written into buf (truncated to buf_len); returns n_out or < 0.
extern "C" int64_t gelex_b200_plan(const char* bfile, int mode, int do_loco, int32_t* raw_to_out, int64_t raw_cap, int64_t* n_raw_out,
                                   int64_t* m_out, int* dense_out, char* names_buf, int names_len, int64_t* group_begin,
                                   int64_t* group_end, int group_cap, int* n_items_out, char* err, int err_len) {
    try {
        std::filesystem::path bed = format_bed_path(bfile), fam = bed, bim = bed;
        fam.replace_extension(".fam");
        bim.replace_extension(".bim");
        const auto raw = load_fam_ids(fam);
        const auto order = make_sample_order(raw);
        const auto chroms = load_bim_chromosomes(bim);
        if (n_raw_out) *n_raw_out = static_cast<int64_t>(raw.size());
        if (m_out) *m_out = static_cast<int64_t>(count_total_lines(bim));
        if (dense_out) *dense_out = order.dense ? 1 : 0;
        if (raw_to_out)
            for (int64_t i = 0; i < static_cast<int64_t>(raw.size()) && i < raw_cap; ++i) raw_to_out[i] = order.raw_to_out[static_cast<size_t>(i)];
        std::string suffix;
        size_t ng = 0;
        const auto items = GrmEngine::build_plan(chroms, mode == 2 ? GrmType::AD : (mode == 1 ? GrmType::D : GrmType::A), do_loco != 0, &suffix, &ng);
        std::string names;
        int gi = 0;
        for (const auto& it : items) {
            names += it.output_name + "\n";
            if (gi < group_cap && group_begin && group_end) {
                group_begin[gi] = it.ranges[0].first;
                group_end[gi] = it.ranges[0].second;
            }
            ++gi;
        }
        names += suffix + "\n";
        if (names_buf && names_len > 0) {
            std::strncpy(names_buf, names.c_str(), static_cast<size_t>(names_len - 1));
            names_buf[names_len - 1] = '\0';
        }
        if (n_items_out) *n_items_out = static_cast<int>(items.size());
        set_err(err, err_len, "");
        return static_cast<int64_t>(order.ids.size());
    } catch (...) {
        return translate_exception(err, err_len);  // "Type: message", status < 0
    }
}

// GrmLoader / LocoGRMLoader through C (tests, bindings).  ids = n_ids canonical sample ids "FID\x1FIID" separated by '\n' with
// their target rows in `targets` (n_ids == 0: no id_map, the file's own order).  chr_prefix NULL/"" -> GrmLoader::load (normalize != 0)
// or load_unnormalized; otherwise LocoGRMLoader(prefix, id_map).load_loco_grm(chr_prefix, id_map).  Returns the matrix size (the
// column-major matrix is written to out when out_cap >= size^2; call with out == NULL to query) or a negative status.
extern "C" int64_t gelex_b200_grm_load(const char* prefix, const char* chr_prefix, int normalize, const char* ids, const int64_t* targets,
                                       int64_t n_ids, double* out, int64_t out_cap, char* err, int err_len) {
    try {
        if (prefix == nullptr) throw ArgumentValidationException("null argument");
        IdMap id_map;
        if (n_ids > 0) {
            if (ids == nullptr || targets == nullptr) throw ArgumentValidationException("null id list");
            const std::string all(ids);
            size_t pos = 0;
            for (int64_t k = 0; k < n_ids; ++k) {
                const size_t nl = all.find('\n', pos);
                id_map[all.substr(pos, nl == std::string::npos ? std::string::npos : nl - pos)] = targets[k];
                pos = nl == std::string::npos ? all.size() : nl + 1;
            }
        }
        DenseMatrix m;
        if (chr_prefix != nullptr && chr_prefix[0] != '\0') {
            if (n_ids == 0) {  // every sample of the whole-genome file, in file order
                const GrmLoader probe(prefix);
                for (int64_t i = 0; i < probe.num_samples(); ++i) id_map[probe.sample_ids()[static_cast<size_t>(i)]] = i;
            }
            const LocoGRMLoader loco(prefix, id_map);
            m = loco.load_loco_grm(chr_prefix, id_map);
        } else {
            const GrmLoader loader(prefix);
            if (n_ids > 0) m = normalize ? loader.load(id_map) : loader.load_unnormalized(id_map);
            else m = normalize ? loader.load() : loader.load_unnormalized();
        }
        if (out != nullptr && out_cap >= m.n * m.n) std::memcpy(out, m.data.data(), m.data.size() * sizeof(double));
        set_err(err, err_len, "");
        return m.n;
    } catch (const FileNotFoundException& e) { set_err(err, err_len, std::string("FileNotFoundException: ") + e.what()); return -10;
    } catch (const FileOpenException& e) { set_err(err, err_len, std::string("FileOpenException: ") + e.what()); return -11;
    } catch (const FileFormatException& e) { set_err(err, err_len, std::string("FileFormatException: ") + e.what()); return GGRM_ERR_FILE_FORMAT;
    } catch (const InvalidInputException& e) { set_err(err, err_len, std::string("InvalidInputException: ") + e.what()); return GGRM_ERR_INVALID_ARGUMENT;
    } catch (const GelexException& e) { set_err(err, err_len, std::string("GelexException: ") + e.what()); return GGRM_ERR_CUDA;
    } catch (const std::bad_alloc&) { set_err(err, err_len, "std::bad_alloc"); return GGRM_ERR_OUT_OF_MEMORY;
    } catch (const std::exception& e) { set_err(err, err_len, std::string("std::exception: ") + e.what()); return -99; }
}

// Host-side sharding rule of the engine, for CPU tests: writes `parts` (begin, end) pairs; returns parts or -1.
extern "C" int gelex_b200_row_slices(int64_t n, int parts, int64_t* begin_end) {
    if (n <= 0 || parts <= 0 || begin_end == nullptr) return -1;
    const auto sl = GrmEngine::row_slices_for(n, parts);
    for (size_t r = 0; r < sl.size(); ++r) {
        begin_end[2 * r] = sl[r].first;
        begin_end[2 * r + 1] = sl[r].second;
    }
    return static_cast<int>(sl.size());
}

static void copy_stats(const GenotypeStats& st, int64_t* dims3, double* mean, double* stddev, double* freq, int64_t* mono, int64_t mono_cap) {
    if (dims3 != nullptr) {
        dims3[0] = st.num_samples;
        dims3[1] = st.num_variants;
        dims3[2] = static_cast<int64_t>(st.monomorphic_indices.size());
    }
    const size_t m = static_cast<size_t>(st.num_variants);
    if (mean != nullptr) std::memcpy(mean, st.means.data(), m * sizeof(double));
    if (stddev != nullptr) std::memcpy(stddev, st.stddevs.data(), m * sizeof(double));
    if (freq != nullptr) std::memcpy(freq, st.freqs.data(), m * sizeof(double));
    if (mono != nullptr)
        for (size_t k = 0; k < st.monomorphic_indices.size() && static_cast<int64_t>(k) < mono_cap; ++k) mono[k] = st.monomorphic_indices[k];
}

// GenotypePipe(bed, prefix).process<GT>(method, chunk) through C: writes <prefix>.bmat + <prefix>.snpstats.  dims3 = {n, m, #mono};
// mean/stddev/freq (m doubles each) and mono (mono_cap indices) may be NULL.  Returns 0 or a negative / GGRM_ERR_* status.
extern "C" int gelex_b200_genotype_export(const char* bfile, const char* out_prefix, int method, int additive, int chunk_size, int device,
                                          int64_t* dims3, double* mean, double* stddev, double* freq, int64_t* mono, int64_t mono_cap,
                                          char* err, int err_len) {
    try {
        if (bfile == nullptr || out_prefix == nullptr) throw ArgumentValidationException("null argument");
        if (method < 1 || method > 8) throw InvalidInputException("Invalid genotype process method.");
        if (chunk_size <= 0) throw ArgumentValidationException("chunk_size must be positive");
        GenotypePipe pipe(format_bed_path(bfile), out_prefix, device);
        const GenotypeStats st = pipe.process(additive != 0, static_cast<GenotypeProcessMethod>(method), static_cast<size_t>(chunk_size));
        copy_stats(st, dims3, mean, stddev, freq, mono, mono_cap);
        set_err(err, err_len, "");
        return 0;
    } catch (...) {
        return translate_exception(err, err_len);
    }
}

// GenotypeLoader(bed).process<GT>(method, chunk) through C.  Call with out == NULL to query dims3 = {n, m, 0} (no GPU work); otherwise
// out must hold n*m doubles (column-major).
extern "C" int gelex_b200_genotype_load(const char* bfile, int method, int additive, int chunk_size, int device, double* out, int64_t out_cap,
                                        int64_t* dims3, double* mean, double* stddev, double* freq, int64_t* mono, int64_t mono_cap, char* err,
                                        int err_len) {
    try {
        if (bfile == nullptr) throw ArgumentValidationException("null argument");
        if (method < 1 || method > 8) throw InvalidInputException("Invalid genotype process method.");
        if (chunk_size <= 0) throw ArgumentValidationException("chunk_size must be positive");
        GenotypeLoader loader(format_bed_path(bfile), device);
        if (out == nullptr) {
            if (dims3 != nullptr) {
                dims3[0] = loader.num_samples();
                dims3[1] = loader.num_variants();
                dims3[2] = 0;
            }
            set_err(err, err_len, "");
            return 0;
        }
        if (out_cap < loader.num_samples() * loader.num_variants()) throw ArgumentValidationException("output buffer too small");
        const GenotypeStats st = loader.process(additive != 0, static_cast<GenotypeProcessMethod>(method), out, static_cast<size_t>(chunk_size));
        copy_stats(st, dims3, mean, stddev, freq, mono, mono_cap);
        set_err(err, err_len, "");
        return 0;
    } catch (...) {
        return translate_exception(err, err_len);
    }
}

// GrmEngine with Config::gcta: <out>.grm.bin (A = G / N), <out>.grm.N.bin, <out>.grm.id.  Additive, whole genome; SNP shards when n_gpus > 1.
extern "C" int gelex_b200_grm_run_gcta_ex(const char* bfile, const char* out_prefix, int method, int chunk_size, int precision, int n_gpus,
                                          int writer_threads, int gcta_diag, int partition, char* err, int err_len) {
    try {
        if (bfile == nullptr || out_prefix == nullptr) throw ArgumentValidationException("null argument");
        if (method < 1 || method > 8) throw InvalidInputException("Invalid genotype process method.");
        if (partition < 0 || partition > 2) throw ArgumentValidationException("partition must be 0 (auto), 1 (snp) or 2 (tile)");
        GrmEngine::Config cfg;
        cfg.bed_path = format_bed_path(bfile);
        cfg.mode = GrmType::A;
        cfg.method = static_cast<GenotypeProcessMethod>(method);
        cfg.out_prefix = out_prefix;
        cfg.chunk_size = chunk_size;
        cfg.precision = precision;
        cfg.partition = partition == 2 ? Partition::Tile : (partition == 1 ? Partition::SnpShard : Partition::Auto);
        cfg.writer_threads = writer_threads;
        cfg.gcta = true;
        cfg.gcta_diag = gcta_diag != 0;
        cfg.gpus.clear();
        const char* over = std::getenv("GELEX_B200_OVERSUBSCRIBE");  // test aid, see gelex_b200_grm_run_ex
        const int devs = ggrm_device_count();
        for (int i = 0; i < (n_gpus > 0 ? n_gpus : 1); ++i) cfg.gpus.push_back(over != nullptr && over[0] == '1' && devs > 0 ? i % devs : i);
        GrmEngine engine(cfg);
        engine.compute();
        set_err(err, err_len, "");
        return 0;
    } catch (...) {
        return translate_exception(err, err_len);
    }
}

extern "C" int gelex_b200_grm_run_gcta(const char* bfile, const char* out_prefix, int method, int chunk_size, int precision, int n_gpus,
                                       int writer_threads, char* err, int err_len) {
    return gelex_b200_grm_run_gcta_ex(bfile, out_prefix, method, chunk_size, precision, n_gpus, writer_threads, 0, 1, err, err_len);
}

extern "C" int gelex_b200_strip_owners(int64_t n, int world, const int64_t* band_row_end, int n_bands, int64_t* out4, int cap) {
    if (n <= 0 || world <= 0 || band_row_end == nullptr || n_bands <= 0 || out4 == nullptr) return -1;
    const auto strips = GrmEngine::strip_owners_for(n, world, std::vector<int64_t>(band_row_end, band_row_end + n_bands));
    for (size_t i = 0; i < strips.size() && static_cast<int>(i) < cap; ++i) {
        out4[4 * i] = strips[i].rank;
        out4[4 * i + 1] = strips[i].band;
        out4[4 * i + 2] = strips[i].row_begin;
        out4[4 * i + 3] = strips[i].row_end;
    }
    return static_cast<int>(strips.size());
}

extern "C" int gelex_b200_gpus_worth_using(int64_t n, int64_t m, int kinds, int precision, int offered, int n_outputs) {
    return GrmEngine::gpus_worth_using(n, m, kinds, precision, offered, n_outputs);
}

extern "C" int gelex_b200_piece_owners(int64_t n, int world, const int64_t* band_row_end, int n_bands, int64_t* out5, int cap) {
    if (n <= 0 || world <= 0 || band_row_end == nullptr || n_bands <= 0 || out5 == nullptr) return -1;
    const auto pieces = GrmEngine::piece_owners_for(n, world, std::vector<int64_t>(band_row_end, band_row_end + n_bands));
    for (size_t i = 0; i < pieces.size() && static_cast<int>(i) < cap; ++i) {
        out5[5 * i] = pieces[i].rank;
        out5[5 * i + 1] = pieces[i].band;
        out5[5 * i + 2] = pieces[i].super_row;
        out5[5 * i + 3] = pieces[i].super_col_begin;
        out5[5 * i + 4] = pieces[i].super_col_end;
    }
    return static_cast<int>(pieces.size());
}

// GrmEngine::writer_selftest through C (CPU tests, no GPU): writer 0 = auto, 1 = pwrite, 2 = mmap; seconds_out / mapped_out optional.
extern "C" int gelex_b200_writer_selftest(const char* bin_path, int64_t n, int tiles, int parts, int writer_threads, int writer,
                                          int prefault_threads, double* seconds_out, double* prefault_seconds_out, int* mapped_out, char* err,
                                          int err_len) {
    try {
        if (bin_path == nullptr) throw ArgumentValidationException("null argument");
        if (writer < 0 || writer > 2) throw ArgumentValidationException("writer must be 0 (auto), 1 (pwrite) or 2 (mmap)");
        bool mapped = false;
        double pf = 0.0;
        const double s = GrmEngine::writer_selftest(bin_path, n, tiles != 0, parts, writer_threads, static_cast<WriterMode>(writer), &mapped,
                                                    prefault_threads, &pf);
        if (prefault_seconds_out != nullptr) *prefault_seconds_out = pf;
        if (seconds_out != nullptr) *seconds_out = s;
        if (mapped_out != nullptr) *mapped_out = mapped ? 1 : 0;
        set_err(err, err_len, "");
        return 0;
    } catch (...) {
        return translate_exception(err, err_len);
    }
}

extern "C" int64_t gelex_b200_relay_selftest(int n_chunks, int world, int slots, int chunk_bytes, int fail_at_chunk, char* err, int err_len) {
    try {
        const int64_t bad = GrmEngine::relay_selftest(n_chunks, world, slots, chunk_bytes, fail_at_chunk);
        set_err(err, err_len, "");
        return bad;
    } catch (...) {
        return translate_exception(err, err_len);
    }
}

// write_id_file (GrmIdWriter::write, grm_id_writer.cpp:41-54) through C: ids = n_ids canonical "FID<US>IID" strings, each NUL-terminated,
// back to back.  For the reference's id-writer test cases (tests/test_grm_id_writer.cpp).
extern "C" int gelex_b200_write_id_file(const char* path, const char* ids, int64_t n_ids, char* err, int err_len) {
    try {
        if (path == nullptr || (ids == nullptr && n_ids > 0) || n_ids < 0) throw ArgumentValidationException("null argument");
        std::vector<std::string> v;
        const char* p = ids;
        for (int64_t i = 0; i < n_ids; ++i) {
            v.emplace_back(p);
            p += v.back().size() + 1;
        }
        write_id_file(path, v);
        set_err(err, err_len, "");
        return 0;
    } catch (...) {
        return translate_exception(err, err_len);
    }
}

// format_bed_path (src/data/genotype/bed_path.cpp:26-40) through C: out receives the .bed path.
extern "C" int gelex_b200_format_bed_path(const char* bfile, char* out, int out_len, char* err, int err_len) {
    try {
        if (bfile == nullptr || out == nullptr || out_len <= 0) throw ArgumentValidationException("null argument");
        const std::string r = format_bed_path(bfile).string();
        if (static_cast<int>(r.size()) >= out_len) throw ArgumentValidationException("output buffer too small");
        std::memcpy(out, r.c_str(), r.size() + 1);
        set_err(err, err_len, "");
        return 0;
    } catch (...) {
        return translate_exception(err, err_len);
    }
}
