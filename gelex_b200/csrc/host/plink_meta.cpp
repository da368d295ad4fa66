#include "plink_meta.h"

#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include <algorithm>
#include <charconv>
#include <fstream>
#include <unordered_map>

#include "exceptions.h"

namespace fs = std::filesystem;

namespace gelex::b200 {

// Same checks, in the same order, with the same messages as the reference's open_file (include/gelex/io/parser.h:48-93).
std::ifstream open_text(const fs::path& path) {
    if (fs::is_directory(path)) throw FileOpenException(path.string() + ": is a directory, not a regular file");
    std::ifstream f(path, std::ios::in);
    if (!f.is_open()) {
        if (!fs::exists(path)) throw FileNotFoundException(path.string() + ": not found");
        throw FileOpenException(path.string() + ": failed to open file");
    }
    if (fs::is_regular_file(path)) {
        std::error_code ec;
        if (fs::file_size(path, ec) == 0 && !ec) throw FileFormatException(path.string() + ": is empty");
    }
    return f;
}

namespace {

char detect_delimiter(std::ifstream& f) {
    std::string probe;
    std::getline(f, probe);
    const bool tab = !probe.empty() && probe.find('\t') != std::string::npos;
    f.clear();
    f.seekg(0);
    return tab ? '\t' : ' ';
}

// First two delimiter-separated fields -> FID<US>IID (src/io/parser.cpp:63-82, include/gelex/types/sample_id.h:32-45).
std::string parse_id(const std::string& line, char delim) {
    const size_t p1 = line.find(delim);
    if (p1 == std::string::npos) throw FileFormatException("failed to parse FID and IID (missing delimiter)");
    const size_t p2 = line.find(delim, p1 + 1);
    const std::string fid = line.substr(0, p1);
    const std::string iid = line.substr(p1 + 1, p2 == std::string::npos ? std::string::npos : p2 - p1 - 1);
    if (fid.empty()) throw ArgumentValidationException("FID cannot be empty");
    if (iid.empty()) throw ArgumentValidationException("IID cannot be empty");
    return fid + kSampleIdSeparator + iid;
}

}  // namespace

fs::path format_bed_path(const std::string& bfile) {
    fs::path bed(bfile);
    if (bed.extension() != ".bed") bed += ".bed";
    if (!fs::exists(bed)) throw FileNotFoundException(bed.string() + ": file not found");
    return bed;
}

size_t count_total_lines(const fs::path& path) {
    if (fs::is_directory(path)) throw FileOpenException(path.string() + ": is a directory, not a regular file");
    std::ifstream f(path, std::ios::in | std::ios::binary);
    if (!f.is_open()) {
        if (!fs::exists(path)) throw FileNotFoundException(path.string() + ": not found");
        throw FileOpenException(path.string() + ": failed to open file");
    }
    std::error_code ec;
    if (fs::is_regular_file(path) && fs::file_size(path, ec) == 0 && !ec) throw FileFormatException(path.string() + ": is empty");
    std::vector<char> buf(128 * 1024);
    size_t lines = 0;
    char last = '\n';
    while (f) {
        f.read(buf.data(), static_cast<std::streamsize>(buf.size()));
        const std::streamsize got = f.gcount();
        if (got == 0) break;
        lines += static_cast<size_t>(std::count(buf.begin(), buf.begin() + got, '\n'));
        last = buf[static_cast<size_t>(got) - 1];
    }
    if (last != '\n') ++lines;
    return lines;
}

std::vector<std::string> load_fam_ids(const fs::path& fam_path) {
    try {
        std::ifstream f = open_text(fam_path);
        const char delim = detect_delimiter(f);
        std::vector<std::string> ids;
        ids.reserve(1024);
        std::string line;
        int n_line = 0;
        while (std::getline(f, line)) {
            ++n_line;
            if (line.empty()) continue;
            try {
                ids.emplace_back(parse_id(line, delim));
            } catch (const GelexException& e) {
                throw DataParseException(std::to_string(n_line) + ": " + e.what());
            }
        }
        return ids;
    } catch (const GelexException& e) {
        throw FileFormatException(fam_path.string() + ":" + e.what());  // fam_loader.cpp:35-45
    }
}

SampleOrder make_sample_order(const std::vector<std::string>& raw_ids) {
    SampleOrder so;
    so.ids = raw_ids;
    std::sort(so.ids.begin(), so.ids.end());
    so.ids.erase(std::unique(so.ids.begin(), so.ids.end()), so.ids.end());
    std::unordered_map<std::string, int32_t> index;
    index.reserve(so.ids.size() * 2);
    for (size_t i = 0; i < so.ids.size(); ++i) index.emplace(so.ids[i], static_cast<int32_t>(i));
    so.raw_to_out.resize(raw_ids.size());
    bool sequential = true;
    for (size_t i = 0; i < raw_ids.size(); ++i) {
        const int32_t t = index.at(raw_ids[i]);
        so.raw_to_out[i] = t;
        if (t != static_cast<int32_t>(i)) sequential = false;
    }
    so.dense = sequential && raw_ids.size() == so.ids.size();
    return so;
}

std::vector<std::string> load_bim_chromosomes(const fs::path& bim_path) {
    std::ifstream f = open_text(bim_path);
    const char delim = detect_delimiter(f);
    std::vector<std::string> chroms;
    try {
        std::string line;
        int n_line = 0;
        while (std::getline(f, line)) {
            ++n_line;
            std::string cols[6];
            int count = 0;
            size_t pos = 0;
            while (pos <= line.size() && count < 6) {
                size_t next = line.find(delim, pos);
                if (next == std::string::npos) next = line.size();
                if (next > pos) cols[count++] = line.substr(pos, next - pos);  // empty tokens are skipped
                pos = next + 1;
            }
            if (count != 6)
                throw InconsistentColumnCountException(std::to_string(n_line) + ": has " + std::to_string(count) + " columns, expected 6");
            int bp = 0;
            const char* first = cols[3].data();
            const char* last = first + cols[3].size();
            auto [ptr, ec] = std::from_chars(first, last, bp);
            if (ec != std::errc() || ptr != last)
                throw DataParseException(std::to_string(n_line) + ": failed to parse '" + cols[3] + "' as number");
            chroms.emplace_back(std::move(cols[0]));
        }
    } catch (const GelexException& e) {
        throw FileFormatException(bim_path.string() + ":" + e.what());  // bim_loader.cpp:41-49
    }
    return chroms;
}

std::vector<ChrRun> chromosome_runs(const std::vector<std::string>& chroms) {
    std::vector<ChrRun> runs;
    const int64_t m = static_cast<int64_t>(chroms.size());
    int64_t start = 0;
    for (int64_t i = 1; i <= m; ++i) {
        if (i == m || chroms[static_cast<size_t>(i)] != chroms[static_cast<size_t>(start)]) {
            if (m > 0) runs.push_back({chroms[static_cast<size_t>(start)], start, i});
            start = i;
        }
    }
    return runs;
}

BedMapping::BedMapping(const fs::path& bed_path, int64_t num_snps, int64_t bytes_per_variant) : bpv_(bytes_per_variant) {
    if (bpv_ <= 0) throw FileFormatException(bed_path.string() + ": invalid bytes per variant");
    const int fd = ::open(bed_path.c_str(), O_RDONLY);
    if (fd < 0) throw FileOpenException(bed_path.string() + ": failed to mmap bed file");
    struct stat st {};
    if (::fstat(fd, &st) != 0) {
        ::close(fd);
        throw FileOpenException(bed_path.string() + ": failed to mmap bed file");
    }
    size_ = static_cast<size_t>(st.st_size);
    if (size_ <= 3) {
        ::close(fd);
        throw FileFormatException(bed_path.string() + ": bed file too short");
    }
    base_ = ::mmap(nullptr, size_, PROT_READ, MAP_PRIVATE, fd, 0);
    ::close(fd);
    if (base_ == MAP_FAILED) {
        base_ = nullptr;
        throw FileOpenException(bed_path.string() + ": failed to mmap bed file");
    }
    // a constructor that throws never runs the destructor: release the mapping first
    auto reject = [this](const std::string& msg) {
        ::munmap(base_, size_);
        base_ = nullptr;
        throw FileFormatException(msg);
    };
    const auto* b = static_cast<const uint8_t*>(base_);
    if (!(b[0] == 0x6C && b[1] == 0x1B && b[2] == 0x01)) reject(bed_path.string() + ": invalid BED magic number");
    const size_t expected = 3 + static_cast<size_t>(num_snps) * static_cast<size_t>(bpv_);
    if (size_ < expected)
        reject(bed_path.string() + ": bed file truncated. Expected " + std::to_string(expected) + " bytes, got " + std::to_string(size_));
    payload_snps_ = (size_ - 3) / static_cast<size_t>(bpv_);
    ::madvise(base_, size_, MADV_SEQUENTIAL);
}

BedMapping::~BedMapping() {
    if (base_ != nullptr) ::munmap(base_, size_);
}

const uint8_t* BedMapping::chunk_ptr(int64_t snp_begin, int64_t snp_count) const {
    if (snp_begin < 0 || snp_count < 0) return nullptr;
    const size_t s = static_cast<size_t>(snp_begin), c = static_cast<size_t>(snp_count);
    if (s > payload_snps_ || c > payload_snps_ - s) return nullptr;
    return static_cast<const uint8_t*>(base_) + 3 + s * static_cast<size_t>(bpv_);
}

std::pair<std::string, std::string> split_sample_id(const std::string& id) {
    if (id.empty()) throw ArgumentValidationException("sample ID cannot be empty");
    const size_t p = id.find(kSampleIdSeparator);
    if (p == std::string::npos) throw ArgumentValidationException("sample ID is not in the canonical FID<US>IID format");
    std::string fid = id.substr(0, p), iid = id.substr(p + 1);
    if (fid.empty() || iid.empty()) throw ArgumentValidationException("sample ID must contain non-empty FID and IID");
    return {fid, iid};
}

}  // namespace gelex::b200
