// Part of ggrm_api.cu: stateless test hooks (bit-exact layers) and the synthetic BED generator entry points.

namespace ggrm {

struct ScratchFree {
    std::vector<void*> ptrs;
    ~ScratchFree() { for (void* p : ptrs) if (p != nullptr) cudaFree(p); }
    template <typename T>
    T* alloc(size_t n) {
        void* p = nullptr;
        GGRM_CUDA(cudaMalloc(&p, std::max<size_t>(n, 1) * sizeof(T)));
        ptrs.push_back(p);
        return static_cast<T*>(p);
    }
};

static void decode_common(int device, const uint8_t* payload, int64_t snp_count, int64_t bpv, int64_t n_raw, int64_t n_out,
                          const int32_t* raw_to_out, int method, int kind, int policy, bool standardise, double* out, double* freq,
                          double* mean = nullptr, double* stddev = nullptr, uint8_t* monomorphic = nullptr) {
    check_device(device, nullptr);
    if (payload == nullptr || out == nullptr) throw CudaError{GGRM_ERR_INVALID_ARGUMENT, "null argument"};
    if (n_raw <= 0 || n_out <= 0 || snp_count <= 0 || snp_count > 65535) throw CudaError{GGRM_ERR_INVALID_ARGUMENT, "test hook limits: 0 < snp_count <= 65535"};
    if (bpv != (n_raw + 3) / 4) throw CudaError{GGRM_ERR_INVALID_ARGUMENT, "bytes_per_variant != ceil(n_raw/4)"};
    if (raw_to_out == nullptr && n_out != n_raw) throw CudaError{GGRM_ERR_INVALID_ARGUMENT, "identity projection requires n_out == n_raw"};
    if (standardise && (method < 1 || method > 8)) throw CudaError{GGRM_ERR_INVALID_ARGUMENT, "Invalid genotype process method."};
    const int64_t n_pad = round_up(n_out, kTile), bpv_pad = n_pad / 4;
    ScratchFree sf;
    uint8_t* d_raw = sf.alloc<uint8_t>(static_cast<size_t>(snp_count * bpv));
    uint8_t* d_dense = sf.alloc<uint8_t>(static_cast<size_t>(snp_count * bpv_pad));
    double* d_out = sf.alloc<double>(static_cast<size_t>(snp_count * n_out));
    int32_t* d_inv = nullptr;
    if (raw_to_out != nullptr) {
        std::vector<int32_t> inv(static_cast<size_t>(n_out), -1);
        for (int64_t i = 0; i < n_raw; ++i) {
            const int32_t t = raw_to_out[i];
            if (t < -1 || t >= n_out) throw CudaError{GGRM_ERR_INVALID_ARGUMENT, "raw_to_out entry out of range"};
            if (t >= 0) inv[static_cast<size_t>(t)] = static_cast<int32_t>(i);
        }
        d_inv = sf.alloc<int32_t>(inv.size());
        GGRM_CUDA(cudaMemcpy(d_inv, inv.data(), inv.size() * sizeof(int32_t), cudaMemcpyHostToDevice));
    }
    GGRM_CUDA(cudaMemcpy(d_raw, payload, static_cast<size_t>(snp_count * bpv), cudaMemcpyHostToDevice));
    repack_kernel<<<repack_grid(snp_count, bpv_pad), 256>>>(d_raw, bpv, d_dense, bpv_pad, n_out, d_inv, snp_count);
    GGRM_CUDA(cudaGetLastError());
    SnpTable* d_tab = nullptr;
    double* d_freq = nullptr;
    SnpMoments* d_mom = nullptr;
    const bool want_moments = mean != nullptr || stddev != nullptr || monomorphic != nullptr;
    if (standardise) {
        d_tab = sf.alloc<SnpTable>(static_cast<size_t>(snp_count));
        d_freq = sf.alloc<double>(static_cast<size_t>(snp_count));
        if (want_moments) d_mom = sf.alloc<SnpMoments>(static_cast<size_t>(snp_count));
        StatsOut so{};
        so.freq = d_freq;
        so.tab64[kind] = d_tab;
        so.moments[kind] = d_mom;
        count_stats_kernel<<<static_cast<unsigned>((snp_count + 7) / 8), 256>>>(d_dense, bpv_pad, n_out, n_pad, snp_count, method, policy, 0, so, 0);
        GGRM_CUDA(cudaGetLastError());
    }
    const dim3 grid(static_cast<unsigned>((n_out + 255) / 256), static_cast<unsigned>(snp_count));
    decode_colmajor_kernel<<<grid, 256>>>(d_dense, bpv_pad, n_out, snp_count, d_tab, d_out);
    GGRM_CUDA(cudaGetLastError());
    GGRM_CUDA(cudaMemcpy(out, d_out, static_cast<size_t>(snp_count * n_out) * sizeof(double), cudaMemcpyDeviceToHost));
    if (standardise && freq != nullptr) GGRM_CUDA(cudaMemcpy(freq, d_freq, static_cast<size_t>(snp_count) * sizeof(double), cudaMemcpyDeviceToHost));
    if (d_mom != nullptr) {
        std::vector<SnpMoments> mo(static_cast<size_t>(snp_count));
        GGRM_CUDA(cudaMemcpy(mo.data(), d_mom, mo.size() * sizeof(SnpMoments), cudaMemcpyDeviceToHost));
        for (size_t j = 0; j < mo.size(); ++j) {
            if (mean != nullptr) mean[j] = mo[j].mean;
            if (stddev != nullptr) stddev[j] = mo[j].stddev;
            if (monomorphic != nullptr) monomorphic[j] = static_cast<uint8_t>(mo[j].monomorphic);
        }
    }
}

}  // namespace ggrm

#define GGRM_STATELESS_BEGIN try {
#define GGRM_STATELESS_END                    \
    }                                         \
    catch (const CudaError& e) {              \
        g_create_error = e.msg;               \
        return e.status;                      \
    }                                         \
    catch (const std::bad_alloc&) {           \
        g_create_error = "host allocation failed"; \
        return GGRM_ERR_OUT_OF_MEMORY;        \
    }                                         \
    catch (const std::exception& e) {         \
        g_create_error = e.what();            \
        return GGRM_ERR_INVALID_OPERATION;    \
    }                                         \
    return GGRM_OK;

extern "C" int ggrm_decode_chunk(int device, const uint8_t* bed_payload, int64_t snp_count, int64_t bytes_per_variant, int64_t n_raw,
                                 int64_t n_out, const int32_t* raw_to_out, double* out_colmajor) {
    GGRM_STATELESS_BEGIN
    decode_common(device, bed_payload, snp_count, bytes_per_variant, n_raw, n_out, raw_to_out, 0, 0, 0, false, out_colmajor, nullptr);
    GGRM_STATELESS_END
}

extern "C" int ggrm_standardise_chunk(int device, const uint8_t* bed_payload, int64_t snp_count, int64_t bytes_per_variant, int64_t n_raw,
                                      int64_t n_out, const int32_t* raw_to_out, int method, int kind, int all_missing_policy,
                                      double* out_colmajor, double* freq) {
    GGRM_STATELESS_BEGIN
    if (kind != GGRM_KIND_ADD && kind != GGRM_KIND_DOM) throw CudaError{GGRM_ERR_INVALID_ARGUMENT, "kind must be 0 (add) or 1 (dom)"};
    decode_common(device, bed_payload, snp_count, bytes_per_variant, n_raw, n_out, raw_to_out, method, kind, all_missing_policy, true,
                  out_colmajor, freq);
    GGRM_STATELESS_END
}

// dispatch_assoc_chunk_by_method (src/pipeline/assoc_detail.cpp:28-47): the GWAS scan's chunk pre-processing is the same decode +
// code policy, restricted to the centre-family methods, with the per-SNP A1 frequencies handed to the result writer.
extern "C" int ggrm_assoc_chunk(int device, const uint8_t* bed_payload, int64_t snp_count, int64_t bytes_per_variant, int64_t n_raw,
                                int64_t n_out, const int32_t* raw_to_out, int method, int kind, double* out_colmajor, double* freq) {
    GGRM_STATELESS_BEGIN
    if (method != GGRM_CENTER_HWE && method != GGRM_ORTH_CENTER_HWE && method != GGRM_CENTER && method != GGRM_ORTH_CENTER)
        throw CudaError{GGRM_ERR_INVALID_ARGUMENT,
                        "assoc --geno-method supports only center-family methods: 2 (center-hwe), 4 (orth-center-hwe), 6 (center), 8 (orth-center)"};
    if (kind != GGRM_KIND_ADD && kind != GGRM_KIND_DOM) throw CudaError{GGRM_ERR_INVALID_ARGUMENT, "kind must be 0 (add) or 1 (dom)"};
    decode_common(device, bed_payload, snp_count, bytes_per_variant, n_raw, n_out, raw_to_out, method, kind, GGRM_ALLMISS_REFERENCE_NAN, true,
                  out_colmajor, freq);
    GGRM_STATELESS_END
}

// GenotypeLoader::process_chunk / GenotypePipe::process_chunk (src/data/genotype/genotype_loader.cpp:51-79, genotype_pipe.cpp:28-49):
// the standardised chunk itself is the product, with the LocusStatistic of every SNP beside it.
extern "C" int ggrm_genotype_chunk(int device, const uint8_t* bed_payload, int64_t snp_count, int64_t bytes_per_variant, int64_t n_raw,
                                   int64_t n_out, const int32_t* raw_to_out, int method, int kind, double* out_colmajor, double* freq,
                                   double* mean, double* stddev, uint8_t* is_monomorphic) {
    GGRM_STATELESS_BEGIN
    if (kind != GGRM_KIND_ADD && kind != GGRM_KIND_DOM) throw CudaError{GGRM_ERR_INVALID_ARGUMENT, "kind must be 0 (add) or 1 (dom)"};
    decode_common(device, bed_payload, snp_count, bytes_per_variant, n_raw, n_out, raw_to_out, method, kind, GGRM_ALLMISS_REFERENCE_NAN, true,
                  out_colmajor, freq, mean, stddev, is_monomorphic);
    GGRM_STATELESS_END
}

extern "C" int ggrm_enable_peer_access(int device, int peer_device) {
    GGRM_STATELESS_BEGIN
    check_device(device, nullptr);
    if (device == peer_device) return GGRM_OK;
    int can = 0;
    GGRM_CUDA(cudaDeviceCanAccessPeer(&can, device, peer_device));
    if (!can) throw CudaError{GGRM_ERR_CUDA, "CUDA: device " + std::to_string(device) + " cannot access peer " + std::to_string(peer_device)};
    const cudaError_t e = cudaDeviceEnablePeerAccess(peer_device, 0);
    if (e == cudaErrorPeerAccessAlreadyEnabled) cudaGetLastError();
    else GGRM_CUDA(e);
    GGRM_STATELESS_END
}

extern "C" int ggrm_ipc_open(int device, const void* handle64, void** d_ptr) {
    GGRM_STATELESS_BEGIN
    if (handle64 == nullptr || d_ptr == nullptr) throw CudaError{GGRM_ERR_INVALID_ARGUMENT, "null argument"};
    check_device(device, nullptr);
    cudaIpcMemHandle_t ipc;
    std::memcpy(&ipc, handle64, sizeof(ipc));
    GGRM_CUDA(cudaIpcOpenMemHandle(d_ptr, ipc, cudaIpcMemLazyEnablePeerAccess));
    GGRM_STATELESS_END
}

extern "C" int ggrm_ipc_close(int device, void* d_ptr) {
    GGRM_STATELESS_BEGIN
    check_device(device, nullptr);
    if (d_ptr != nullptr) GGRM_CUDA(cudaIpcCloseMemHandle(d_ptr));
    GGRM_STATELESS_END
}

extern "C" int ggrm_bin_to_dense(int device, const float* whole_packed, int64_t n_whole, const int64_t* target_to_whole, const float* chr_packed,
                                 int64_t n_chr, const int64_t* target_to_chr, int64_t n_target, double denominator, double* out_colmajor) {
    GGRM_STATELESS_BEGIN
    check_device(device, nullptr);
    if (whole_packed == nullptr || out_colmajor == nullptr) throw CudaError{GGRM_ERR_INVALID_ARGUMENT, "null argument"};
    const int64_t lim = int64_t{1} << 30;
    if (n_whole <= 0 || n_target <= 0 || n_whole > lim || n_target > lim || (chr_packed != nullptr && (n_chr <= 0 || n_chr > lim)))
        throw CudaError{GGRM_ERR_INVALID_ARGUMENT, "matrix sizes must be positive"};
    auto check_map = [&](const int64_t* map, int64_t n_file) {
        if (map == nullptr) {
            if (n_target != n_file) throw CudaError{GGRM_ERR_INVALID_ARGUMENT, "identity map requires n_target == n_file"};
            return;
        }
        for (int64_t t = 0; t < n_target; ++t)
            if (map[t] < -1 || map[t] >= n_file) throw CudaError{GGRM_ERR_INVALID_ARGUMENT, "target map entry out of range"};
    };
    check_map(target_to_whole, n_whole);
    if (chr_packed != nullptr) check_map(target_to_chr, n_chr);
    ScratchFree sf;
    auto upload_packed = [&](const float* src, int64_t n) {
        const size_t packed = static_cast<size_t>(n) * static_cast<size_t>(n + 1) / 2;
        float* d = sf.alloc<float>(packed);
        GGRM_CUDA(cudaMemcpy(d, src, packed * sizeof(float), cudaMemcpyHostToDevice));
        return d;
    };
    auto upload_map = [&](const int64_t* map) -> int64_t* {
        if (map == nullptr) return nullptr;
        int64_t* d = sf.alloc<int64_t>(static_cast<size_t>(n_target));
        GGRM_CUDA(cudaMemcpy(d, map, static_cast<size_t>(n_target) * sizeof(int64_t), cudaMemcpyHostToDevice));
        return d;
    };
    const float* d_w = upload_packed(whole_packed, n_whole);
    const float* d_c = chr_packed != nullptr ? upload_packed(chr_packed, n_chr) : nullptr;
    const int64_t* d_mw = upload_map(target_to_whole);
    const int64_t* d_mc = chr_packed != nullptr ? upload_map(target_to_chr) : nullptr;
    double* d_out = sf.alloc<double>(static_cast<size_t>(n_target) * static_cast<size_t>(n_target));
    const unsigned nb = static_cast<unsigned>((n_target + 31) / 32);
    bin_to_dense_kernel<<<dim3(nb, nb), 256>>>(d_w, d_c, d_mw, d_mc, n_target, denominator, d_out);
    GGRM_CUDA(cudaGetLastError());
    GGRM_CUDA(cudaMemcpy(out_colmajor, d_out, static_cast<size_t>(n_target) * static_cast<size_t>(n_target) * sizeof(double), cudaMemcpyDeviceToHost));
    GGRM_STATELESS_END
}

extern "C" int ggrm_host_alloc(int64_t bytes, void** out) {
    GGRM_STATELESS_BEGIN
    if (out == nullptr || bytes <= 0) throw CudaError{GGRM_ERR_INVALID_ARGUMENT, "invalid pinned allocation request"};
    *out = nullptr;
    const cudaError_t e = cudaHostAlloc(out, static_cast<size_t>(bytes), cudaHostAllocPortable);
    if (e == cudaErrorNoDevice || e == cudaErrorInsufficientDriver) {  // a box without a GPU: say so the way ggrm_create does
        cudaGetLastError();
        throw CudaError{GGRM_ERR_NO_DEVICE, "no CUDA device visible; the B200 GRM engine has no CPU fallback"};
    }
    GGRM_CUDA(e);
    GGRM_STATELESS_END
}

extern "C" int ggrm_host_free(void* p) {
    GGRM_STATELESS_BEGIN
    if (p != nullptr) GGRM_CUDA(cudaFreeHost(p));
    GGRM_STATELESS_END
}

extern "C" int ggrm_host_register(void* p, int64_t bytes) {
    if (p == nullptr || bytes <= 0) {
        g_create_error = "null argument";
        return GGRM_ERR_INVALID_ARGUMENT;
    }
    const cudaError_t e = cudaHostRegister(p, static_cast<size_t>(bytes), cudaHostRegisterPortable);
    if (e != cudaSuccess) {
        cudaGetLastError();
        g_create_error = std::string("CUDA: ") + cudaGetErrorString(e) + " at cudaHostRegister";
        return GGRM_ERR_CUDA;
    }
    return GGRM_OK;
}

extern "C" int ggrm_host_unregister(void* p) {
    if (p == nullptr) return GGRM_OK;
    const cudaError_t e = cudaHostUnregister(p);
    if (e != cudaSuccess) {
        cudaGetLastError();
        g_create_error = std::string("CUDA: ") + cudaGetErrorString(e) + " at cudaHostUnregister";
        return GGRM_ERR_CUDA;
    }
    return GGRM_OK;
}

static uint32_t miss_threshold(double missing_rate) {
    if (!(missing_rate > 0.0)) return 0u;
    if (missing_rate >= 1.0) return 1u << 24;
    return static_cast<uint32_t>(missing_rate * 16777216.0);
}

extern "C" int ggrm_synth_bed_host(uint8_t* out, int64_t n, int64_t snp_begin, int64_t snp_count, uint64_t seed, double missing_rate) {
    if (out == nullptr || n <= 0 || snp_begin < 0 || snp_count < 0) {
        g_create_error = "invalid argument";
        return GGRM_ERR_INVALID_ARGUMENT;
    }
    const int64_t bpv = (n + 3) / 4;
    const uint32_t thr = miss_threshold(missing_rate);
    for (int64_t s = 0; s < snp_count; ++s) {
        const ggrm_snp_thresholds t = ggrm_synth_thresholds(seed, snp_begin + s);
        for (int64_t b = 0; b < bpv; ++b) out[s * bpv + b] = ggrm_synth_byte(seed, snp_begin + s, b, n, t, thr);
    }
    return GGRM_OK;
}

extern "C" int ggrm_synth_bed_device(int device, uint8_t* d_out, int64_t n, int64_t snp_begin, int64_t snp_count, uint64_t seed,
                                     double missing_rate) {
    GGRM_STATELESS_BEGIN
    check_device(device, nullptr);
    if (d_out == nullptr || n <= 0 || snp_begin < 0 || snp_count < 0) throw CudaError{GGRM_ERR_INVALID_ARGUMENT, "invalid argument"};
    if (snp_count > 0) {
        const int64_t bpv = (n + 3) / 4;
        synth_bed_kernel<<<static_cast<unsigned>(snp_count), 256>>>(d_out, n, bpv, snp_begin, snp_count, seed, miss_threshold(missing_rate));
        GGRM_CUDA(cudaGetLastError());
        GGRM_CUDA(cudaDeviceSynchronize());
    }
    GGRM_STATELESS_END
}
