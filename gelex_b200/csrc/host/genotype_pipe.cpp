// See genotype_pipe.h.  The per-chunk arithmetic is ggrm_genotype_chunk (CUDA); this file is file handling and bookkeeping.
#include "genotype_pipe.h"

#include <fcntl.h>
#include <unistd.h>

#include <algorithm>
#include <array>
#include <cerrno>
#include <cstring>
#include <fstream>
#include <future>

#include "exceptions.h"

namespace fs = std::filesystem;

namespace gelex::b200 {

bool GenotypeStats::is_monomorphic(int64_t snp) const {
    return std::binary_search(monomorphic_indices.begin(), monomorphic_indices.end(), snp);
}

GenotypeSource::GenotypeSource(const fs::path& bed_path) {
    fs::path fam = bed_path, bim = bed_path;
    fam.replace_extension(".fam");
    bim.replace_extension(".bim");
    const std::vector<std::string> raw_ids = load_fam_ids(fam);
    order_ = make_sample_order(raw_ids);
    n_raw_ = static_cast<int64_t>(raw_ids.size());
    m_ = static_cast<int64_t>(count_total_lines(bim));
    bed_ = std::make_unique<BedMapping>(bed_path, m_, (n_raw_ + 3) / 4);
}

namespace {

struct PinnedBuffer {
    double* p = nullptr;
    explicit PinnedBuffer(size_t doubles) {
        void* q = nullptr;
        const int st = ggrm_host_alloc(doubles * sizeof(double), &q);
        if (st != GGRM_OK) throw_status(st, ggrm_last_error(nullptr));
        p = static_cast<double*>(q);
    }
    ~PinnedBuffer() {
        if (p != nullptr) ggrm_host_free(p);
    }
    PinnedBuffer(const PinnedBuffer&) = delete;
    PinnedBuffer& operator=(const PinnedBuffer&) = delete;
};

// One chunk through the GPU; per-SNP outputs land at their global positions in `st`.
void run_chunk(const GenotypeSource& src, int device, bool additive, GenotypeProcessMethod method, int64_t begin, int64_t count, double* out,
               GenotypeStats& st, std::vector<uint8_t>& mono_scratch) {
    const uint8_t* payload = src.bed().chunk_ptr(begin, count);
    if (payload == nullptr) throw ColumnRangeException("SNP range out of bounds of the BED payload");
    mono_scratch.assign(static_cast<size_t>(count), 0);
    const int rc = ggrm_genotype_chunk(device, payload, count, src.bed().bytes_per_variant(), src.num_raw_samples(), src.num_samples(),
                                       src.order().dense ? nullptr : src.order().raw_to_out.data(), static_cast<int>(method),
                                       additive ? GGRM_KIND_ADD : GGRM_KIND_DOM, out, st.freqs.data() + begin, st.means.data() + begin,
                                       st.stddevs.data() + begin, mono_scratch.data());
    if (rc != GGRM_OK) throw_status(rc, ggrm_last_error(nullptr));
    for (int64_t j = 0; j < count; ++j)
        if (mono_scratch[static_cast<size_t>(j)] != 0) st.monomorphic_indices.push_back(begin + j);
}

GenotypeStats empty_stats(const GenotypeSource& src) {
    GenotypeStats st;
    st.num_samples = src.num_samples();
    st.num_variants = src.num_variants();
    st.means.assign(static_cast<size_t>(st.num_variants), 0.0);
    st.stddevs.assign(static_cast<size_t>(st.num_variants), 0.0);
    st.freqs.assign(static_cast<size_t>(st.num_variants), 0.0);
    return st;
}

void write_all(int fd, const void* data, size_t bytes, const fs::path& path, const char* what) {
    const char* p = static_cast<const char*>(data);
    while (bytes > 0) {
        const ssize_t w = ::write(fd, p, std::min<size_t>(bytes, size_t{1} << 30));
        if (w < 0) {
            if (errno == EINTR) continue;
            throw FileWriteException(path.string() + ": " + what);
        }
        p += w;
        bytes -= static_cast<size_t>(w);
    }
}

struct Fd {
    int fd = -1;
    Fd(const fs::path& path) : fd(::open(path.c_str(), O_WRONLY | O_CREAT | O_TRUNC, 0644)) {
        if (fd < 0) throw FileOpenException(path.string() + ": failed to open file");
    }
    ~Fd() {
        if (fd >= 0) ::close(fd);
    }
};

}  // namespace

void write_snp_stats(const fs::path& path, const GenotypeStats& st) {
    if (st.means.size() != st.stddevs.size())
        throw ArgumentValidationException("means (" + std::to_string(st.means.size()) + ") and stddevs (" + std::to_string(st.stddevs.size()) +
                                          ") must have the same length.");
    if (st.means.empty()) throw ArgumentValidationException("means and stddevs cannot be empty");
    const int64_t m = static_cast<int64_t>(st.means.size());
    if (!st.monomorphic_indices.empty()) {
        const int64_t last = st.monomorphic_indices.back();
        if (last >= m || last < 0)
            throw ArgumentValidationException("Monomorphic SNP index " + std::to_string(last) + " is out of range [0, " + std::to_string(m) + ").");
    }
    Fd f(path);
    const std::array<int64_t, 3> header = {st.num_samples, m, static_cast<int64_t>(st.monomorphic_indices.size())};
    write_all(f.fd, header.data(), sizeof(header), path, "failed to write header");
    if (!st.monomorphic_indices.empty())
        write_all(f.fd, st.monomorphic_indices.data(), st.monomorphic_indices.size() * sizeof(int64_t), path, "failed to write monomorphic snp indices");
    write_all(f.fd, st.means.data(), st.means.size() * sizeof(double), path, "failed to write means");
    write_all(f.fd, st.stddevs.data(), st.stddevs.size() * sizeof(double), path, "failed to write stddevs");
}

GenotypePipe::GenotypePipe(const fs::path& bed_path, const fs::path& output_prefix, int device) : src_(bed_path), device_(device) {
    matrix_path_ = output_prefix;
    matrix_path_ += ".bmat";
    stats_path_ = output_prefix;
    stats_path_ += ".snpstats";
    if (fs::exists(matrix_path_) || fs::exists(stats_path_)) throw FileExistsException(output_prefix.string() + ": existing files");
}

GenotypeStats GenotypePipe::process(bool additive, GenotypeProcessMethod method, size_t chunk_size) {
    if (chunk_size == 0 || chunk_size > 65535) throw ArgumentValidationException("chunk_size must be in [1, 65535]");
    GenotypeStats st = empty_stats(src_);
    const int64_t n = st.num_samples, m = st.num_variants, c = static_cast<int64_t>(chunk_size);
    Fd out(matrix_path_);
    // two pinned chunk buffers: the file write of chunk i runs on a helper thread under the GPU pass of chunk i + 1
    PinnedBuffer bufs[2] = {PinnedBuffer(static_cast<size_t>(n * std::min(c, m))), PinnedBuffer(static_cast<size_t>(n * std::min(c, m)))};
    std::future<void> pending[2];
    std::vector<uint8_t> mono;
    int b = 0;
    try {
        for (int64_t begin = 0; begin < m; begin += c, b ^= 1) {
            const int64_t count = std::min(c, m - begin);
            if (pending[b].valid()) pending[b].get();
            run_chunk(src_, device_, additive, method, begin, count, bufs[b].p, st, mono);
            const double* data = bufs[b].p;
            const size_t bytes = static_cast<size_t>(n * count) * sizeof(double);
            if (pending[b ^ 1].valid()) pending[b ^ 1].get();  // keep the file in chunk order: one writer at a time
            pending[b] = std::async(std::launch::async, [fd = out.fd, data, bytes, this] { write_all(fd, data, bytes, matrix_path_, "failed to write matrix data to binary file"); });
        }
        for (auto& p : pending)
            if (p.valid()) p.get();
    } catch (...) {
        for (auto& p : pending)
            if (p.valid()) p.wait();
        throw;
    }
    write_snp_stats(stats_path_, st);
    return st;
}

GenotypeLoader::GenotypeLoader(const fs::path& bed_path, int device) : src_(bed_path), device_(device) {}

GenotypeStats GenotypeLoader::process(bool additive, GenotypeProcessMethod method, double* out, size_t chunk_size) {
    if (out == nullptr) throw ArgumentValidationException("null output matrix");
    if (chunk_size == 0 || chunk_size > 65535) throw ArgumentValidationException("chunk_size must be in [1, 65535]");
    GenotypeStats st = empty_stats(src_);
    const int64_t n = st.num_samples, m = st.num_variants, c = static_cast<int64_t>(chunk_size);
    std::vector<uint8_t> mono;
    for (int64_t begin = 0; begin < m; begin += c)
        run_chunk(src_, device_, additive, method, begin, std::min(c, m - begin), out + begin * n, st, mono);
    return st;
}

}  // namespace gelex::b200
