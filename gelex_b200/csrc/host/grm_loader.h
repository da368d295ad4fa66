// Readers of the .bin/.id pair with the semantics of the reference's consumer side:
//   GrmLoader      src/data/grm/grm_loader.cpp:47-243      (include/gelex/data/grm/grm_loader.h)
//   LocoGRMLoader  src/data/grm/loco_grm_loader.cpp:27-98  (include/gelex/data/grm/loco_grm_loader.h)
// Same constructor arguments, same validation and exception types/messages, same values; the O(n^2) expansion of the
// packed float32 triangle into the dense symmetric FP64 matrix (filter + reorder by id_map, subtraction, division)
// runs on the GPU (ggrm_bin_to_dense), the O(n) denominators are summed on the host in the reference's order.
#pragma once
#include <cstdint>
#include <filesystem>
#include <string>
#include <unordered_map>
#include <vector>

namespace gelex::b200 {

struct DenseMatrix {  // column-major n x n, the layout of Eigen::MatrixXd
    int64_t n = 0;
    std::vector<double> data;
    double operator()(int64_t i, int64_t j) const { return data[static_cast<size_t>(j * n + i)]; }
};

using IdMap = std::unordered_map<std::string, int64_t>;  // canonical sample id (FID<US>IID) -> target row

enum class LoadedGrmType : uint8_t { A, D, Unknown };  // get_type(), grm_loader.cpp:31-42

class GrmLoader {
   public:
    explicit GrmLoader(const std::filesystem::path& prefix, int device = 0);
    ~GrmLoader();
    GrmLoader(const GrmLoader&) = delete;
    GrmLoader& operator=(const GrmLoader&) = delete;

    DenseMatrix load() const;                               // G / (trace / n)
    DenseMatrix load_unnormalized() const;
    DenseMatrix load(const IdMap& id_map) const;            // filtered + reordered, then / (trace / rows)
    DenseMatrix load_unnormalized(const IdMap& id_map) const;

    const std::vector<std::string>& sample_ids() const { return sample_ids_; }
    int64_t num_samples() const { return num_samples_; }
    LoadedGrmType type() const { return type_; }
    const float* packed() const { return data_; }

    // target row -> file row (-1 = none) and the matrix size max_target + 1; throws InvalidInputException for an id that is
    // not in the file (grm_loader.cpp:196-205)
    std::vector<int64_t> target_map(const IdMap& id_map, int64_t* out_size) const;
    // trace of the filtered matrix, summed in target order like Eigen's trace()
    double filtered_trace(const std::vector<int64_t>& target_to_file) const;

   private:
    DenseMatrix expand(const std::vector<int64_t>* target_to_file, int64_t n_target, double denominator) const;
    std::filesystem::path bin_path_, id_path_;
    LoadedGrmType type_;
    int device_;
    std::vector<std::string> sample_ids_;
    int64_t num_samples_ = 0;
    void* map_base_ = nullptr;
    size_t map_size_ = 0;
    const float* data_ = nullptr;
};

class LocoGRMLoader {
   public:
    LocoGRMLoader(const std::filesystem::path& whole_grm_prefix, const IdMap& id_map, int device = 0);
    // (G_whole - G_chr) / (k_whole - k_chr); InvalidInputException when the file is missing or k_loco <= 0
    DenseMatrix load_loco_grm(const std::filesystem::path& chr_grm_prefix, const IdMap& id_map) const;
    int64_t num_samples() const { return n_target_; }
    double k_whole() const { return k_whole_; }

   private:
    GrmLoader whole_;
    int device_;
    std::vector<int64_t> whole_map_;
    int64_t n_target_ = 0;
    double k_whole_ = 0;
};

}  // namespace gelex::b200
