// Part of ggrm_api.cu (same translation unit): ggrm_snp_stats_noalloc, an entry point that allocates nothing.
// No counterpart in the reference; added for callers that cannot call cudaMalloc (code
// under CUDA graph capture, tight serving loops, harnesses that forbid it).
//
// It is the K1a + K1b/K2 pair of the path - repack_kernel then count_stats_kernel (prep_kernels.cuh) - whose intermediate, the
// dense 2-bit chunk in GRM row order padded to 256 samples with "missing", every other entry point keeps in a cudaMalloc'ed
// buffer (Handle::d_dense, or the ScratchFree buffers of the stateless hooks).  Here that scratch is a file-scope
// `static __device__` array of GGRM_NOALLOC_SCRATCH_BYTES: every GPU that loads the library holds it until the process exits.
//
// g_noalloc_scratch below is, in HOST code, only a stand-in for the symbol - not a device address.  The address a kernel can use
// comes from cudaGetSymbolAddress, asked once per device with that device current (noalloc_scratch_for), and is what the launches
// are given.  All calls on one device share the array: it keeps the previous call's bytes (repack_kernel overwrites every byte
// count_stats_kernel then reads, so nothing has to start at zero), and two calls that overlap on the GPU - different streams,
// different host threads - race on it.

namespace ggrm {

static __device__ __align__(256) uint8_t g_noalloc_scratch[GGRM_NOALLOC_SCRATCH_BYTES];

constexpr int kNoallocMaxDevices = 64;
static std::mutex g_noalloc_mu;
static uint8_t* g_noalloc_addr[kNoallocMaxDevices] = {};  // device address of g_noalloc_scratch per device ordinal, nullptr = not asked yet

// Device address of the scratch array on `device`; makes `device` current.  First call per device: sm_100 check + symbol look-up.
static uint8_t* noalloc_scratch_for(int device) {
    if (device < 0 || device >= kNoallocMaxDevices) throw CudaError{GGRM_ERR_INVALID_ARGUMENT, "device ordinal " + std::to_string(device) + " out of range"};
    std::lock_guard<std::mutex> lk(g_noalloc_mu);
    if (g_noalloc_addr[device] == nullptr) {
        check_device(device, nullptr);  // cudaSetDevice(device) included; NO_DEVICE without an sm_100 GPU
        void* p = nullptr;
        GGRM_CUDA(cudaGetSymbolAddress(&p, g_noalloc_scratch));
        if (p == nullptr) throw CudaError{GGRM_ERR_CUDA, "CUDA: cudaGetSymbolAddress returned no address for the scratch array"};
        g_noalloc_addr[device] = static_cast<uint8_t*>(p);
    } else {
        GGRM_CUDA(cudaSetDevice(device));
    }
    return g_noalloc_addr[device];
}

}  // namespace ggrm

extern "C" int ggrm_snp_stats_noalloc(int device, const uint8_t* d_bed_payload, int64_t snp_count, int64_t bytes_per_variant, int64_t n_raw,
                                      int32_t* d_counts4, double* d_freq, void* stream) {
    GGRM_STATELESS_BEGIN
    if (d_bed_payload == nullptr || (d_counts4 == nullptr && d_freq == nullptr)) throw CudaError{GGRM_ERR_INVALID_ARGUMENT, "null argument"};
    if (n_raw <= 0 || snp_count <= 0) throw CudaError{GGRM_ERR_INVALID_ARGUMENT, "n_raw and snp_count must be positive"};
    if (bytes_per_variant != (n_raw + 3) / 4)
        throw CudaError{GGRM_ERR_INVALID_ARGUMENT, "bytes_per_variant " + std::to_string(bytes_per_variant) + " != ceil(n_raw/4) = " + std::to_string((n_raw + 3) / 4)};
    if ((reinterpret_cast<uintptr_t>(d_counts4) & 15u) != 0) throw CudaError{GGRM_ERR_INVALID_ARGUMENT, "d_counts4 must be 16-byte aligned"};
    if (n_raw > (int64_t{1} << 30)) throw CudaError{GGRM_ERR_INVALID_ARGUMENT, "n_raw too large"};
    const int64_t n_pad = round_up(n_raw, kTile), bpv_pad = n_pad / 4;
    const int64_t need = snp_count * bpv_pad;
    if (snp_count > int64_t{GGRM_NOALLOC_SCRATCH_BYTES} / bpv_pad)
        throw CudaError{GGRM_ERR_INVALID_ARGUMENT,
                        "ggrm_snp_stats_noalloc: " + std::to_string(snp_count) + " SNPs x " + std::to_string(bpv_pad) + " padded bytes = " + std::to_string(need) +
                            " bytes of scratch exceed the static device array of " + std::to_string(int64_t{GGRM_NOALLOC_SCRATCH_BYTES}) +
                            " bytes (GGRM_NOALLOC_SCRATCH_BYTES): at most " + std::to_string(int64_t{GGRM_NOALLOC_SCRATCH_BYTES} / bpv_pad) +
                            " SNPs per call at this sample count"};
    uint8_t* scratch = noalloc_scratch_for(device);
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    repack_kernel<<<repack_grid(snp_count, bpv_pad), 256, 0, s>>>(d_bed_payload, bytes_per_variant, scratch, bpv_pad, n_raw, nullptr, snp_count);
    GGRM_CUDA(cudaGetLastError());
    StatsOut so{};  // no value tables: counts and frequency only
    so.counts = reinterpret_cast<int4*>(d_counts4);
    so.freq = d_freq;
    count_stats_kernel<<<static_cast<unsigned>((snp_count + 7) / 8), 256, 0, s>>>(scratch, bpv_pad, n_raw, n_pad, snp_count, GGRM_STANDARDIZE_HWE,
                                                                                  GGRM_ALLMISS_REFERENCE_NAN, 0, so, 0);
    GGRM_CUDA(cudaGetLastError());
    GGRM_STATELESS_END
}
