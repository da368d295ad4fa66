#include "grm_loader.h"

#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include <algorithm>
#include <fstream>

#include "../../../include/ggrm.h"
#include "exceptions.h"
#include "plink_meta.h"

#include <format>

namespace fs = std::filesystem;

namespace gelex::b200 {

namespace {

LoadedGrmType type_of(const std::string& stem) {
    if (stem.find("add") != std::string::npos) return LoadedGrmType::A;
    if (stem.find("dom") != std::string::npos) return LoadedGrmType::D;
    return LoadedGrmType::Unknown;
}

}  // namespace

GrmLoader::GrmLoader(const fs::path& prefix, int device)
    : bin_path_(prefix.string() + ".bin"), id_path_(prefix.string() + ".id"), type_(type_of(prefix.string())), device_(device) {
    // load_sample_ids, grm_loader.cpp:58-95: "FID\tIID" per line, empty lines skipped
    {
        std::ifstream f = open_text(id_path_);
        std::string line;
        while (std::getline(f, line)) {
            if (line.empty()) continue;
            const size_t tab = line.find('\t');
            if (tab == std::string::npos) throw FileFormatException(id_path_.string() + ": invalid ID line '" + line + "' (expected FID and IID)");
            // make_sample_id, include/gelex/types/sample_id.h:32-45
            if (tab == 0) throw ArgumentValidationException("FID cannot be empty");
            if (tab + 1 == line.size()) throw ArgumentValidationException("IID cannot be empty");
            sample_ids_.push_back(line.substr(0, tab) + kSampleIdSeparator + line.substr(tab + 1));
        }
        if (sample_ids_.empty()) throw FileFormatException(id_path_.string() + ": no sample IDs found");
        num_samples_ = static_cast<int64_t>(sample_ids_.size());
    }
    // init_mmap, grm_loader.cpp:97-123: the payload is exactly n(n+1)/2 float32
    const int fd = ::open(bin_path_.c_str(), O_RDONLY);
    if (fd < 0) throw FileOpenException(bin_path_.string() + ": failed to mmap file");
    struct stat st {};
    if (::fstat(fd, &st) != 0) {
        ::close(fd);
        throw FileOpenException(bin_path_.string() + ": failed to mmap file");
    }
    map_size_ = static_cast<size_t>(st.st_size);
    if (map_size_ == 0) {  // mio cannot map an empty file: the reference reports that before it compares sizes
        ::close(fd);
        throw FileOpenException(bin_path_.string() + ": failed to mmap file");
    }
    const size_t expected = static_cast<size_t>(num_samples_) * (static_cast<size_t>(num_samples_) + 1) / 2 * sizeof(float);
    if (map_size_ != expected) {
        ::close(fd);
        throw FileFormatException(bin_path_.string() + ": file size mismatch. Expected " + std::to_string(expected) + " bytes (" +
                                  std::to_string(num_samples_) + " samples), got " + std::to_string(map_size_) + " bytes");
    }
    map_base_ = ::mmap(nullptr, map_size_, PROT_READ, MAP_PRIVATE, fd, 0);
    ::close(fd);
    if (map_base_ == MAP_FAILED) {
        map_base_ = nullptr;
        throw FileOpenException(bin_path_.string() + ": failed to mmap file");
    }
    data_ = static_cast<const float*>(map_base_);
}

GrmLoader::~GrmLoader() {
    if (map_base_ != nullptr) ::munmap(map_base_, map_size_);
}

std::vector<int64_t> GrmLoader::target_map(const IdMap& id_map, int64_t* out_size) const {
    std::unordered_map<std::string, int64_t> file_index;
    file_index.reserve(sample_ids_.size());
    for (int64_t i = 0; i < num_samples_; ++i) file_index[sample_ids_[static_cast<size_t>(i)]] = i;  // a repeated id keeps its last row
    int64_t max_target = 0;
    for (const auto& [id, target] : id_map) {
        if (file_index.find(id) == file_index.end()) throw InvalidInputException(bin_path_.string() + ": sample ID '" + id + "' not found in GRM file");
        if (target < 0) throw InvalidInputException(bin_path_.string() + ": negative target index for sample ID '" + id + "'");
        max_target = std::max(max_target, target);
    }
    const int64_t n_target = id_map.empty() ? 0 : max_target + 1;
    std::vector<int64_t> map(static_cast<size_t>(n_target), -1);
    for (const auto& [id, target] : id_map) map[static_cast<size_t>(target)] = file_index[id];
    if (out_size != nullptr) *out_size = n_target;
    return map;
}

double GrmLoader::filtered_trace(const std::vector<int64_t>& target_to_file) const {
    double tr = 0.0;
    for (const int64_t f : target_to_file)
        tr += f >= 0 ? static_cast<double>(data_[static_cast<size_t>(f) * (static_cast<size_t>(f) + 1) / 2 + static_cast<size_t>(f)]) : 0.0;
    return tr;
}

DenseMatrix GrmLoader::expand(const std::vector<int64_t>* target_to_file, int64_t n_target, double denominator) const {
    DenseMatrix out;
    out.n = n_target;
    if (n_target == 0) return out;
    out.data.resize(static_cast<size_t>(n_target) * static_cast<size_t>(n_target));
    const int st = ggrm_bin_to_dense(device_, data_, num_samples_, target_to_file != nullptr ? target_to_file->data() : nullptr, nullptr, 0, nullptr,
                                     n_target, denominator, out.data.data());
    if (st != GGRM_OK) throw_status(st, ggrm_last_error(nullptr));
    return out;
}

DenseMatrix GrmLoader::load_unnormalized() const { return expand(nullptr, num_samples_, 1.0); }

DenseMatrix GrmLoader::load() const {
    double tr = 0.0;
    for (int64_t i = 0; i < num_samples_; ++i) tr += static_cast<double>(data_[static_cast<size_t>(i) * (static_cast<size_t>(i) + 1) / 2 + static_cast<size_t>(i)]);
    return expand(nullptr, num_samples_, tr / static_cast<double>(num_samples_));
}

DenseMatrix GrmLoader::load_unnormalized(const IdMap& id_map) const {
    int64_t n_target = 0;
    const std::vector<int64_t> map = target_map(id_map, &n_target);
    return expand(&map, n_target, 1.0);
}

DenseMatrix GrmLoader::load(const IdMap& id_map) const {
    int64_t n_target = 0;
    const std::vector<int64_t> map = target_map(id_map, &n_target);
    if (n_target == 0) return DenseMatrix{};
    return expand(&map, n_target, filtered_trace(map) / static_cast<double>(n_target));
}

LocoGRMLoader::LocoGRMLoader(const fs::path& whole_grm_prefix, const IdMap& id_map, int device) : whole_(whole_grm_prefix, device), device_(device) {
    whole_map_ = whole_.target_map(id_map, &n_target_);
    k_whole_ = n_target_ > 0 ? whole_.filtered_trace(whole_map_) / static_cast<double>(n_target_) : 0.0;
}

DenseMatrix LocoGRMLoader::load_loco_grm(const fs::path& chr_grm_prefix, const IdMap& id_map) const {
    const fs::path bin_path = chr_grm_prefix.string() + ".bin";
    if (!fs::exists(bin_path)) throw InvalidInputException("LOCO error: GRM file not found: " + bin_path.string());
    const GrmLoader chr(chr_grm_prefix, device_);
    int64_t n_chr = 0;
    const std::vector<int64_t> chr_map = chr.target_map(id_map, &n_chr);
    const double k_i = n_chr > 0 ? chr.filtered_trace(chr_map) / static_cast<double>(n_chr) : 0.0;
    const double k_loco = k_whole_ - k_i;
    if (k_loco <= 0)
        throw InvalidInputException(std::format("LOCO error: Chromosome GRM denominator ({}) is greater than or equal to Whole GRM denominator ({})", k_i,
                                                k_whole_));  // "{}" of a double: shortest round-trip digits, as the reference prints them
    // the two files are aligned through the SAME id_map (each by its own .id list), as in the reference
    if (n_chr != n_target_) throw InvalidInputException("LOCO error: " + bin_path.string() + " maps to a different matrix size than the whole-genome GRM");
    DenseMatrix out;
    out.n = n_target_;
    if (n_target_ == 0) return out;
    out.data.resize(static_cast<size_t>(n_target_) * static_cast<size_t>(n_target_));
    const int st = ggrm_bin_to_dense(device_, whole_.packed(), whole_.num_samples(), whole_map_.data(), chr.packed(), chr.num_samples(), chr_map.data(),
                                     n_target_, k_loco, out.data.data());
    if (st != GGRM_OK) throw_status(st, ggrm_last_error(nullptr));
    return out;
}

}  // namespace gelex::b200
