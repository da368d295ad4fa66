// C-ABI implementation (include/ggrm.h): handle life cycle, chunk streaming, kernel launches, read-out.
// Host-side orchestration only; all arithmetic happens in the kernels of prep_kernels.cuh, syrk_tc.cuh,
// syrk_f64.cuh and pack_kernels.cuh.  There is deliberately no CPU fallback anywhere in this file.
#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <new>
#include <numeric>

#include "ggrm_internal.h"
#include "pack_kernels.cuh"

namespace ggrm {

static thread_local std::string g_create_error;

static inline int64_t round_up(int64_t x, int64_t m) { return (x + m - 1) / m * m; }

// ---------------------------------------------------------------------------------------------------------------
// TMA descriptor for a K-major 16-bit operand matrix Z[n_pad][k_cap]: box = 64 elements (128 B) x 128 rows,
// 128-byte swizzle — the smem image tcgen05 descriptors (make_smem_desc_sw128) expect.
void encode_tensor_map(CUtensorMap* map, void* base, int64_t k_cap, int64_t n_pad, bool fp16) {  // k_cap = columns of the matrix
    typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                 const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                 CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
    // resolved once; a function-local static is initialised under the compiler's guard, so handles created concurrently from
    // several host threads (the multi-GPU engine does that) do not race on it, and a failed lookup is retried by the next call
    static const EncodeFn encode = [] {
        void* fn = nullptr;
        cudaDriverEntryPointQueryResult qres;
        GGRM_CUDA(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres));
        if (fn == nullptr || qres != cudaDriverEntryPointSuccess) {
            throw CudaError{GGRM_ERR_CUDA, "CUDA: cuTensorMapEncodeTiled entry point not available"};
        }
        return reinterpret_cast<EncodeFn>(fn);
    }();
    const cuuint64_t dims[2] = {static_cast<cuuint64_t>(k_cap), static_cast<cuuint64_t>(n_pad)};
    const cuuint64_t strides[1] = {static_cast<cuuint64_t>(k_cap) * 2};
    const cuuint32_t box[2] = {64, 128};
    const cuuint32_t estr[2] = {1, 1};
    const CUresult r = encode(map, fp16 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, base,
                              dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                              CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        throw CudaError{GGRM_ERR_CUDA, "CUDA: cuTensorMapEncodeTiled failed with code " + std::to_string(static_cast<int>(r))};
    }
}

// ---------------------------------------------------------------------------------------------------------------
// Tile ownership + persistent work lists.  The lower triangle of 256x256 tiles is cut into 8x8-tile super-tiles
// (2048x2048 entries); super-tiles are the unit of ownership (2D-tile multi-GPU mode, LPT over their tile counts)
// and of scheduling order (CTAs that run at the same time work inside one or two super-tiles and share Z panels in L2).
void build_ownership_and_items(Handle& h) {
    const int64_t T = h.ntiles_side;
    constexpr int64_t ST = 8;
    const int64_t nst = (T + ST - 1) / ST;
    struct Super { int64_t si, sj, cost; int owner; };
    std::vector<Super> supers;
    for (int64_t si = 0; si < nst; ++si) {
        for (int64_t sj = 0; sj <= si; ++sj) {
            int64_t cost = 0;
            for (int64_t I = si * ST; I < std::min(T, (si + 1) * ST); ++I)
                for (int64_t J = sj * ST; J < std::min(T, (sj + 1) * ST); ++J) cost += (J <= I);
            supers.push_back({si, sj, cost, 0});
        }
    }
    const int world = std::max(1, h.cfg.tile_world);
    if (world > 1) {
        std::vector<size_t> order(supers.size());
        std::iota(order.begin(), order.end(), size_t{0});
        std::stable_sort(order.begin(), order.end(), [&](size_t a, size_t b) { return supers[a].cost > supers[b].cost; });
        std::vector<int64_t> load(world, 0);
        for (size_t idx : order) {
            const int r = static_cast<int>(std::min_element(load.begin(), load.end()) - load.begin());
            supers[idx].owner = r;
            load[r] += supers[idx].cost;
        }
    }
    h.slot_of_host.assign(static_cast<size_t>(T * (T + 1) / 2), -1);
    h.owned_supers.clear();
    struct Owned { int64_t I, J; int32_t slot; };
    std::vector<Owned> owned;
    int32_t slot = 0;
    for (const Super& s : supers) {
        if (s.owner != h.cfg.tile_rank) continue;
        h.owned_supers.push_back(s.si);
        h.owned_supers.push_back(s.sj);
        for (int64_t I = s.si * ST; I < std::min(T, (s.si + 1) * ST); ++I) {
            for (int64_t J = s.sj * ST; J < std::min(T, (s.sj + 1) * ST); ++J) {
                if (J > I) continue;
                h.slot_of_host[static_cast<size_t>(I * (I + 1) / 2 + J)] = slot;
                owned.push_back({I, J, slot});
                ++slot;
            }
        }
    }
    // Scheduling order (storage slots stay in super-tile order).  strip_h > 0: horizontal strips of strip_h tile rows,
    // walked column by column (snake: odd strips right-to-left), so that the ~sm_count/2 tiles in flight at any time
    // share strip_h row panels and ~sm_count/(2*strip_h) column panels, and consecutive waves keep the row panels in L2.
    if (h.strip_h > 0 && !h.fp64) {
        const int64_t sh = h.strip_h;
        std::stable_sort(owned.begin(), owned.end(), [sh](const Owned& a, const Owned& b) {
            const int64_t sa = a.I / sh, sb = b.I / sh;
            if (sa != sb) return sa < sb;
            if (a.J != b.J) return (sa & 1) ? a.J > b.J : a.J < b.J;
            return a.I < b.I;
        });
    }
    // Row bands for the final slab (see Handle::band_*): close a band at a strip boundary once it holds its share of tiles.
    h.band_item_end.clear();
    h.band_row_end.clear();
    if (h.strip_h > 0 && !h.fp64 && !owned.empty()) {
        int nb = 8;
        if (const char* e = std::getenv("GGRM_FETCH_BANDS")) nb = std::max(1, std::min(32, std::atoi(e)));
        const int64_t total = static_cast<int64_t>(owned.size());
        int64_t min_tiles = 2048;  // ~28 waves of 74 CTA pairs per band: the extra tail wave costs ~2 % of the last slab only
        if (const char* e = std::getenv("GGRM_BAND_MIN_TILES")) min_tiles = std::max<int64_t>(1, std::atoll(e));
        const int bands = static_cast<int>(std::max<int64_t>(1, std::min<int64_t>(nb, total / min_tiles)));
        const int items_per_tile = h.tc_mode == 2 ? 1 : 2;
        int64_t prev_strip_end = 0;  // tile index where the strip before the current one ended
        for (int64_t i = 0; i < total; ++i) {
            const bool strip_ends = i + 1 == total || owned[static_cast<size_t>(i + 1)].I / h.strip_h != owned[static_cast<size_t>(i)].I / h.strip_h;
            if (!strip_ends) continue;
            const int64_t b = static_cast<int64_t>(h.band_item_end.size());
            const bool last_strip = i + 1 == total;
            // the LAST band is what a pipelined read-out cannot hide: when it would hold several strips, the final strip becomes a
            // band of its own (168 MB instead of 560 MB at n = 50 000)
            if (last_strip && bands > 1 && prev_strip_end > 0 && (h.band_item_end.empty() || h.band_item_end.back() < prev_strip_end * items_per_tile)) {
                h.band_item_end.push_back(static_cast<int>(prev_strip_end * items_per_tile));
                h.band_row_end.push_back(std::min<int64_t>(h.n_out, (owned[static_cast<size_t>(prev_strip_end - 1)].I / h.strip_h + 1) * h.strip_h * kTile));
            }
            if (last_strip || (b + 1 < bands && i + 1 >= (b + 1) * total / bands)) {
                h.band_item_end.push_back(static_cast<int>((i + 1) * items_per_tile));
                h.band_row_end.push_back(last_strip ? h.n_out : std::min<int64_t>(h.n_out, (owned[static_cast<size_t>(i)].I / h.strip_h + 1) * h.strip_h * kTile));
            }
            prev_strip_end = i + 1;
        }
        for (size_t b = h.band_done.size(); b < h.band_item_end.size(); ++b) {
            cudaEvent_t ev = nullptr;
            GGRM_CUDA(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
            h.band_done.push_back(ev);
        }
    }
    std::vector<TcItem> tc;
    std::vector<F64Item> f64;
    for (const Owned& o : owned) {
        const int64_t I = o.I, J = o.J;
        const int64_t off = static_cast<int64_t>(o.slot) * 65536;
        if (h.fp64) {
            for (int bi = 0; bi < 4; ++bi)
                for (int bj = 0; bj < 4; ++bj) {
                    if (I == J && bj > bi) continue;
                    f64.push_back({static_cast<int32_t>(I * 256 + bi * 64), static_cast<int32_t>(J * 256 + bj * 64),
                                   off + static_cast<int64_t>(bi) * 64 * 256 + bj * 64});
                }
        } else if (h.tc_mode == 2) {
            tc.push_back({static_cast<int32_t>(I * 256), static_cast<int32_t>(J * 256), off, 0, 0});
        } else {
            tc.push_back({static_cast<int32_t>(I * 256), static_cast<int32_t>(J * 256), off, 0, 0});
            tc.push_back({static_cast<int32_t>(I * 256 + 128), static_cast<int32_t>(J * 256), off + kHalfTileElems, 0, 0});
        }
    }
    h.owned_tiles = slot;
    h.accum_elems = static_cast<int64_t>(slot) * 65536;
    // Split-K list (Handle::ksplit): as many K parts as whole waves of CTA pairs the tiles leave empty, at most 8; part-major, so
    // that the parts of one tile run in the same wave and share its operand panels in L2.
    h.ksplit = 1;
    std::vector<TcItem> split;
    if (!h.fp64 && h.tc_mode == 2 && slot > 0) {
        // from the tile count of the WHOLE triangle, not of this handle's share: every 2D-tile rank then cuts K exactly like the
        // single-GPU run does, and a tile's bits do not depend on who owns it
        int want = static_cast<int>(std::min<int64_t>(8, (h.sm_count / 2) / std::max<int64_t>(1, T * (T + 1) / 2)));
        // Few-tile problems are launch-bound (measured at C1: every extra slab costs a ~40 us gap, profiles/r02_summary.md section 3), so
        // they take slabs of up to 65 536 SNPs - C1 is ONE launch - and enough K parts that one TMEM accumulation chain stays within
        // 16 384 SNPs (the FP32 truncation bias of a chain grows with its length, see create_impl)
        if (want > 1) want = static_cast<int>(std::min<int64_t>(8, std::max<int64_t>(want, (std::min<int64_t>(round_up(h.cfg.m_total, 64), 65536) + 16383) / 16384)));
        if (const char* e = std::getenv("GGRM_KSPLIT")) want = std::max(1, std::min(8, std::atoi(e)));
        const bool keeps_whole = h.cfg.n_groups > 1 && h.cfg.keep_whole;  // the in-epilogue fold into `whole` assumes one copy
        if (want > 1 && !keeps_whole && h.band_item_end.size() <= 1) {
            h.ksplit = want;
            for (const TcItem& t : tc)
                for (int p = 0; p < want; ++p) split.push_back({t.a_row0, t.b_row0, t.out_off + static_cast<int64_t>(p) * h.accum_elems, p, 0});
        }
    }
    GGRM_CUDA(cudaMalloc(&h.d_slot_of, std::max<size_t>(1, h.slot_of_host.size()) * sizeof(int32_t)));
    GGRM_CUDA(cudaMemcpy(h.d_slot_of, h.slot_of_host.data(), h.slot_of_host.size() * sizeof(int32_t), cudaMemcpyHostToDevice));
    h.num_items_tc = static_cast<int>(tc.size());
    h.num_items_f64 = static_cast<int>(f64.size());
    if (!tc.empty()) {
        GGRM_CUDA(cudaMalloc(&h.d_items_tc, tc.size() * sizeof(TcItem)));
        GGRM_CUDA(cudaMemcpy(h.d_items_tc, tc.data(), tc.size() * sizeof(TcItem), cudaMemcpyHostToDevice));
    }
    h.num_items_split = static_cast<int>(split.size());
    if (!split.empty()) {
        GGRM_CUDA(cudaMalloc(&h.d_items_split, split.size() * sizeof(TcItem)));
        GGRM_CUDA(cudaMemcpy(h.d_items_split, split.data(), split.size() * sizeof(TcItem), cudaMemcpyHostToDevice));
    }
    if (!f64.empty()) {
        GGRM_CUDA(cudaMalloc(&h.d_items_f64, f64.size() * sizeof(F64Item)));
        GGRM_CUDA(cudaMemcpy(h.d_items_f64, f64.data(), f64.size() * sizeof(F64Item), cudaMemcpyHostToDevice));
    }
}

// ---- device allocations, optionally between guard bands (Handle::guard) ----------------------------------------------------
constexpr size_t kGuardBytes = size_t{64} << 10;

static void dev_alloc(Handle& h, void** p, size_t bytes, const char* name) {
    bytes = std::max<size_t>(bytes, 1);
    if (!h.guard) {
        GGRM_CUDA(cudaMalloc(p, bytes));
        return;
    }
    const size_t tail = kGuardBytes + (256 - bytes % 256) % 256;  // the back band starts at the first byte past the buffer
    char* base = nullptr;
    GGRM_CUDA(cudaMalloc(&base, kGuardBytes + bytes + tail));
    GGRM_CUDA(cudaMemset(base, 0xA5, kGuardBytes));
    GGRM_CUDA(cudaMemset(base + kGuardBytes + bytes, 0xA5, tail));
    h.guards.push_back({base, bytes, tail, name});
    *p = base + kGuardBytes;
}
template <typename T>
static void dev_alloc(Handle& h, T** p, size_t bytes, const char* name) { dev_alloc(h, reinterpret_cast<void**>(p), bytes, name); }

static void dev_free(Handle& h, void* p) {
    if (p == nullptr) return;
    for (auto& g : h.guards)
        if (g.base + kGuardBytes == p) {
            cudaFree(g.base);
            g.base = nullptr;
            return;
        }
    cudaFree(p);
}

// Called with the device idle.  Throws on the first band that no longer reads 0xA5 everywhere.
static void check_guards(Handle& h) {
    if (!h.guard) return;
    std::vector<unsigned char> host;
    for (const auto& g : h.guards) {
        if (g.base == nullptr) continue;
        for (int side = 0; side < 2; ++side) {
            const size_t len = side == 0 ? kGuardBytes : g.tail;
            const char* src = side == 0 ? g.base : g.base + kGuardBytes + g.bytes;
            host.resize(len);
            GGRM_CUDA(cudaMemcpy(host.data(), src, len, cudaMemcpyDeviceToHost));
            for (size_t i = 0; i < len; ++i)
                if (host[i] != 0xA5) {
                    const long long off = side == 0 ? static_cast<long long>(i) - static_cast<long long>(kGuardBytes) : static_cast<long long>(g.bytes + i);
                    throw CudaError{GGRM_ERR_CUDA, std::string("guard band of device buffer '") + g.name + "' (" + std::to_string(g.bytes) +
                                                       " bytes) overwritten at byte offset " + std::to_string(off)};
                }
        }
    }
}

static void check_device(int device, int* sm_count) {
    int count = 0;
    const cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count <= 0) {
        throw CudaError{GGRM_ERR_NO_DEVICE, "no CUDA device visible; the B200 GRM engine has no CPU fallback"};
    }
    if (device < 0 || device >= count) {
        throw CudaError{GGRM_ERR_INVALID_ARGUMENT, "device ordinal " + std::to_string(device) + " out of range"};
    }
    cudaDeviceProp prop;
    GGRM_CUDA(cudaGetDeviceProperties(&prop, device));
    if (prop.major != 10) {
        throw CudaError{GGRM_ERR_NO_DEVICE, std::string("device '") + prop.name + "' is sm_" + std::to_string(prop.major) +
                                                std::to_string(prop.minor) + "; this library is built for sm_100a only"};
    }
    if (sm_count != nullptr) *sm_count = prop.multiProcessorCount;
    GGRM_CUDA(cudaSetDevice(device));
}

static void free_handle(Handle* h) {
    if (h == nullptr) return;
    cudaSetDevice(h->cfg.device);
    cudaDeviceSynchronize();
    auto fr = [h](void* p) { dev_free(*h, p); };
    fr(h->d_slot_of); fr(h->d_items_tc); fr(h->d_items_split); fr(h->d_items_f64); fr(h->d_out_to_raw);
    fr(h->d_raw[0]); fr(h->d_raw[1]); fr(h->d_dense); fr(h->d_counts); fr(h->d_freq);
    for (int k = 0; k < 2; ++k) {
        fr(h->d_tab16[k]); fr(h->d_tab64[k]); fr(h->whole[k]);
        for (int b = 0; b < Handle::kMaxZBuf; ++b) fr(h->d_Z[k][b]);
        for (void* p : h->pool[k]) fr(p);
    }
    fr(h->d_stage); fr(h->d_trace); fr(h->d_fault); fr(h->d_wave_ctr); fr(h->d_range_flag);
    fr(h->d_gcta_w); fr(h->d_gcta_lanes); fr(h->d_gcta_diag); fr(h->d_gcta_total);
    for (int i = 0; i < Handle::kRingSlots; ++i) {
        if (h->ring[i] != nullptr) cudaFreeHost(h->ring[i]);
        if (h->ring_free[i] != nullptr) cudaEventDestroy(h->ring_free[i]);
    }
    for (int i = 0; i < 2; ++i)
        if (h->raw_free[i] != nullptr) cudaEventDestroy(h->raw_free[i]);
    for (int i = 0; i < Handle::kMaxZBuf; ++i) {
        if (h->prep_done[i] != nullptr) cudaEventDestroy(h->prep_done[i]);
        if (h->z_free[i] != nullptr) cudaEventDestroy(h->z_free[i]);
    }
    if (h->h2d_done != nullptr) cudaEventDestroy(h->h2d_done);
    if (h->pass_begin != nullptr) cudaEventDestroy(h->pass_begin);
    if (h->pass_end != nullptr) cudaEventDestroy(h->pass_end);
    for (auto& g : h->groups) if (g.done != nullptr) cudaEventDestroy(g.done);
    for (cudaEvent_t ev : h->band_done) cudaEventDestroy(ev);
    for (auto& t : h->timers) { cudaEventDestroy(t.a); cudaEventDestroy(t.b); }
    if (h->s_copy != nullptr) cudaStreamDestroy(h->s_copy);
    if (h->s_prep != nullptr) cudaStreamDestroy(h->s_prep);
    if (h->s_comp != nullptr) cudaStreamDestroy(h->s_comp);
    if (h->s_fetch != nullptr) cudaStreamDestroy(h->s_fetch);
    delete h;
}

static void create_impl(const ggrm_config_t* cfg, Handle* h) {
    if (cfg->abi_version != GGRM_ABI_VERSION) throw CudaError{GGRM_ERR_INVALID_ARGUMENT, "abi_version mismatch"};
    if (cfg->n_raw <= 0 || cfg->n_out <= 0) throw CudaError{GGRM_ERR_INVALID_ARGUMENT, "n_raw and n_out must be positive"};
    if (cfg->raw_to_out == nullptr && cfg->n_out != cfg->n_raw)
        throw CudaError{GGRM_ERR_INVALID_ARGUMENT, "identity projection requires n_out == n_raw"};
    if (cfg->method < 1 || cfg->method > 8) throw CudaError{GGRM_ERR_INVALID_ARGUMENT, "Invalid genotype process method."};
    if (!cfg->want_add && !cfg->want_dom) throw CudaError{GGRM_ERR_INVALID_ARGUMENT, "neither additive nor dominance requested"};
    if (cfg->precision < 0 || cfg->precision > 2) throw CudaError{GGRM_ERR_INVALID_ARGUMENT, "invalid precision"};
    if (cfg->n_groups < 1) throw CudaError{GGRM_ERR_INVALID_ARGUMENT, "n_groups must be >= 1"};
    if (cfg->n_groups > 1 && (cfg->group_begin == nullptr || cfg->group_end == nullptr))
        throw CudaError{GGRM_ERR_INVALID_ARGUMENT, "group ranges required when n_groups > 1"};
    if (cfg->m_total <= 0) throw CudaError{GGRM_ERR_INVALID_ARGUMENT, "m_total must be positive"};
    if (cfg->tile_world < 1 || cfg->tile_rank < 0 || cfg->tile_rank >= cfg->tile_world)
        throw CudaError{GGRM_ERR_INVALID_ARGUMENT, "invalid tile partition"};
    if (cfg->n_out > (int64_t{1} << 30)) throw CudaError{GGRM_ERR_INVALID_ARGUMENT, "n_out too large"};
    if (cfg->second_matrix != GGRM_SECOND_DOMINANCE && cfg->second_matrix != GGRM_SECOND_NONMISSING && cfg->second_matrix != GGRM_SECOND_NONMISSING_GCTA_DIAG)
        throw CudaError{GGRM_ERR_INVALID_ARGUMENT, "invalid second_matrix"};
    if (cfg->second_matrix == GGRM_SECOND_NONMISSING_GCTA_DIAG && (cfg->method != GGRM_STANDARDIZE_HWE || cfg->n_groups != 1 || !cfg->want_add || !cfg->want_dom))
        throw CudaError{GGRM_ERR_INVALID_ARGUMENT, "GCTA's diagonal is defined for the standardize-HWE additive matrix of one group: needs method SH, n_groups = 1, want_add and want_dom"};
    if (cfg->second_matrix != GGRM_SECOND_DOMINANCE && cfg->precision != GGRM_PREC_FP64 && cfg->m_total >= (int64_t{1} << 24))
        throw CudaError{GGRM_ERR_INVALID_ARGUMENT, "non-missing pair counts are exact in FP32 only below 2^24 SNPs: use the FP64 precision"};

    h->cfg = *cfg;
    const bool dbg_time = std::getenv("GGRM_DEBUG_TIMING") != nullptr;
    const auto t_create0 = std::chrono::steady_clock::now();
    auto lap = [&](const char* what) {
        if (dbg_time) std::fprintf(stderr, "[ggrm] create: %-28s at %.3f s\n", what, std::chrono::duration<double>(std::chrono::steady_clock::now() - t_create0).count());
    };
    if (const char* e = std::getenv("GGRM_GUARD")) h->guard = std::atoi(e) != 0;
    h->cfg.raw_to_out = nullptr;
    h->cfg.group_begin = h->cfg.group_end = nullptr;
    check_device(cfg->device, &h->sm_count);
    lap("device checked");
    h->n_raw = cfg->n_raw;
    h->n_out = cfg->n_out;
    h->n_pad = round_up(cfg->n_out, kTile);
    h->bpv_raw = (cfg->n_raw + 3) / 4;
    h->bpv_pad = h->n_pad / 4;
    h->ntiles_side = h->n_pad / kTile;
    h->fp64 = cfg->precision == GGRM_PREC_FP64;
    h->elem_bytes = h->fp64 ? 8 : 4;
    h->tc_mode = 2;  // CTA-pair UMMA 256x256x16 (measured ~20 % faster than the single-CTA 128x256 form)
    if (const char* e = std::getenv("GGRM_CACHE_HINTS")) h->cache_hints = std::atoi(e) & 7;
    if (const char* e = std::getenv("GGRM_TC_MODE")) h->tc_mode = std::atoi(e) == 1 ? 1 : 2;
    if (const char* e = std::getenv("GGRM_STRIP_H")) h->strip_h = std::max(0, std::atoi(e));
    if (const char* e = std::getenv("GGRM_FIRST_SLAB")) h->first_slab_snps = std::max<int64_t>(0, std::atoll(e));
    if (const char* e = std::getenv("GGRM_FIRST_SLAB_SPLIT")) h->first_slab_with_split = std::atoi(e) != 0;

    // groups
    h->groups.resize(static_cast<size_t>(cfg->n_groups));
    for (int g = 0; g < cfg->n_groups; ++g) {
        h->groups[g].begin = cfg->n_groups == 1 && cfg->group_begin == nullptr ? 0 : cfg->group_begin[g];
        h->groups[g].end = cfg->n_groups == 1 && cfg->group_end == nullptr ? cfg->m_total : cfg->group_end[g];
        if (h->groups[g].begin < 0 || h->groups[g].end > cfg->m_total || h->groups[g].begin > h->groups[g].end ||
            (g > 0 && h->groups[g].begin < h->groups[g - 1].end))
            throw CudaError{GGRM_ERR_COLUMN_RANGE, "invalid group range " + std::to_string(g)};
        GGRM_CUDA(cudaEventCreateWithFlags(&h->groups[g].done, cudaEventDisableTiming));
    }

    // projection: invert raw->out, later raw sample wins (variant_decoder.cpp:79-126)
    if (cfg->raw_to_out != nullptr) {
        std::vector<int32_t> inv(static_cast<size_t>(cfg->n_out), -1);
        bool identity = cfg->n_out == cfg->n_raw;
        for (int64_t i = 0; i < cfg->n_raw; ++i) {
            const int32_t t = cfg->raw_to_out[i];
            if (t < -1 || t >= cfg->n_out) throw CudaError{GGRM_ERR_INVALID_ARGUMENT, "raw_to_out entry out of range"};
            if (t >= 0) inv[static_cast<size_t>(t)] = static_cast<int32_t>(i);
            if (t != i) identity = false;
        }
        if (!identity) {
            GGRM_CUDA(cudaMalloc(&h->d_out_to_raw, inv.size() * sizeof(int32_t)));
            GGRM_CUDA(cudaMemcpy(h->d_out_to_raw, inv.data(), inv.size() * sizeof(int32_t), cudaMemcpyHostToDevice));
        }
    }

    GGRM_CUDA(cudaStreamCreateWithFlags(&h->s_copy, cudaStreamNonBlocking));
    GGRM_CUDA(cudaStreamCreateWithFlags(&h->s_comp, cudaStreamNonBlocking));
    {
        // K1-K3 blocks are small; a higher priority lets them slot in beside the persistent SYRK CTAs as soon as they are ready
        int prio_lo = 0, prio_hi = 0;
        GGRM_CUDA(cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi));
        GGRM_CUDA(cudaStreamCreateWithPriority(&h->s_prep, cudaStreamNonBlocking, prio_hi));
    }
    if (const char* e = std::getenv("GGRM_OVERLAP_PREP")) h->overlap_prep = std::atoi(e) != 0;
    GGRM_CUDA(cudaStreamCreateWithFlags(&h->s_fetch, cudaStreamNonBlocking));
    GGRM_CUDA(cudaEventCreateWithFlags(&h->h2d_done, cudaEventDisableTiming));
    for (int i = 0; i < 2; ++i) GGRM_CUDA(cudaEventCreateWithFlags(&h->raw_free[i], cudaEventDisableTiming));
    for (int i = 0; i < Handle::kMaxZBuf; ++i) {
        GGRM_CUDA(cudaEventCreateWithFlags(&h->prep_done[i], cudaEventDisableTiming));
        GGRM_CUDA(cudaEventCreateWithFlags(&h->z_free[i], cudaEventDisableTiming));
    }

    // 16-bit operands and the 1e-4 bound on G/(trace/n).  The rounding error of an output matrix accumulated over m SNPs measures
    // c / sqrt(m) (independent per-SNP rounding of the four table values): c ~ 0.018 for BF16 (8 mantissa bits; 4.0e-4 at
    // m = 2 048, 2.3e-5 at C2's 500 000) and ~ 0.0023 for FP16 (11 bits; 1.1e-5 at C1's 50 000).  The scaled orthogonal dominance
    // codes (OrthStandardizeHWE / OrthStandardize: homozygote = -q/p, heavy-tailed) measure 3.5x that (FP16: 1.0e-4 at m = 6 000).
    // Below the SNP count where a plain 16-bit product keeps the bound with a ~3x margin, operands are split into hi + lo (3 MMAs
    // per K block, ~2^-17 relative error per product), so both 16-bit precisions meet 1e-4 at every size.  What counts is the
    // SHORTEST group: each chromosome partial of a LOCO run is an output matrix of its own (C4's smallest has 9 800 SNPs - the
    // reason the engine's default is FP16, whose threshold those partials clear).  GGRM_OPERAND_SPLIT=0/1 forces it off / on.
    {
        int64_t m_short = cfg->m_total;
        for (const auto& gr : h->groups)
            if (gr.end > gr.begin) m_short = std::min(m_short, gr.end - gr.begin);
        const bool heavy = cfg->want_dom && cfg->second_matrix == GGRM_SECOND_DOMINANCE &&
                           (cfg->method == GGRM_ORTH_STANDARDIZE_HWE || cfg->method == GGRM_ORTH_STANDARDIZE);
        int64_t below = 0;
        if (cfg->precision == GGRM_PREC_BF16) below = heavy ? (int64_t{1} << 22) : (int64_t{1} << 18);
        if (cfg->precision == GGRM_PREC_FP16) below = heavy ? (int64_t{1} << 16) : (int64_t{1} << 13);
        h->op_split = m_short < below;
        const char* e = std::getenv("GGRM_OPERAND_SPLIT");
        if (e == nullptr) e = std::getenv("GGRM_BF16_SPLIT");
        if (e != nullptr) h->op_split = cfg->precision != GGRM_PREC_FP64 && std::atoi(e) != 0;
    }

    build_ownership_and_items(*h);

    // ---- memory plan -------------------------------------------------------------------------------------------
    lap("streams + work lists");
    size_t free_b = 0, total_b = 0;
    GGRM_CUDA(cudaMemGetInfo(&free_b, &total_b));
    const int kinds = (cfg->want_add ? 1 : 0) + (cfg->want_dom ? 1 : 0);
    const int64_t accum_bytes = std::max<int64_t>(h->accum_elems, 4) * h->elem_bytes;
    const bool want_whole = cfg->n_groups > 1 && cfg->keep_whole;
    const int64_t z_elem = h->fp64 ? 8 : (h->op_split ? 4 : 2);  // bytes per sample per SNP of one operand slab (hi + lo when split)
    const int64_t z_cols_per_k = h->op_split ? 2 : 1;
    const int64_t per_snp = 2 * h->n_pad * z_elem * kinds + 2 * h->bpv_raw + h->bpv_pad + 128;
    int64_t avail = static_cast<int64_t>(free_b) - (int64_t{1} << 30) - (want_whole ? accum_bytes * kinds : 0);
    const int64_t min_chunk = std::min<int64_t>(round_up(cfg->m_total, 64), 4096) * per_snp;
    int pool_n = 1;
    if (cfg->n_groups > 1) {
        int64_t fit = (avail - min_chunk) / (accum_bytes * h->ksplit * kinds);
        if (cfg->max_resident_groups > 0) fit = std::min<int64_t>(fit, cfg->max_resident_groups);
        pool_n = static_cast<int>(std::max<int64_t>(1, std::min<int64_t>(fit, cfg->n_groups)));
    }
    avail -= static_cast<int64_t>(pool_n) * accum_bytes * h->ksplit * kinds;
    int64_t k_cap;
    if (cfg->chunk_snps > 0) {
        k_cap = round_up(cfg->chunk_snps, 64);
    } else {
        // Measured on B200 (profiles/r01_summary.md, "wave pacing"): with the producers paced wave by wave the Z traffic per
        // SNP no longer depends on the slab length (tiles in flight share panels through L2 whatever K is), while the FP32
        // accumulator read-modify-write (2 x 5 GB at n = 50k) is paid once per slab - so long slabs win on speed (+1.5 % from
        // 16k to 32k).  Accuracy caps it: the tensor core truncates when it adds into the FP32 TMEM accumulator, which biases
        // large positive sums (the diagonal, ~1 per SNP) by about -5.4e-9 x K relative: 16k SNPs keeps that under 1e-4
        // (tools/precision_probe.py; the bias is uniform, so G/(trace/n) is unaffected either way).
        // split-K launches (few tiles) accumulate only 1 / ksplit of a slab per TMEM chain, so the slab may be ksplit times longer
        // for the same bias - and a small-n problem is launch-bound: fewer, longer slabs
        int64_t target = std::min<int64_t>(int64_t{16384} * h->ksplit, 65536);
        if (const char* e = std::getenv("GGRM_CHUNK_SNPS")) target = std::max<int64_t>(64, std::atoll(e));
        const int64_t budget = std::min<int64_t>(int64_t{24} << 30, avail / 2);
        k_cap = std::max<int64_t>(1024, std::min<int64_t>(target, budget / per_snp / 1024 * 1024));
    }
    k_cap = std::min<int64_t>(k_cap, round_up(cfg->m_total, 64));
    // Extra operand slabs for the band-major tail (Handle::tail_slabs): as many of the wanted 4 as fit in a quarter of what is left
    {
        int want = h->fp64 ? 0 : 4;
        if (const char* e = std::getenv("GGRM_TAIL_SLABS")) want = std::max(0, std::min(Handle::kMaxZBuf - 2, std::atoi(e)));
        if (h->fp64) want = 0;
        const int64_t slab_bytes = h->n_pad * k_cap * z_elem * kinds;
        const int64_t spare = (avail - k_cap * per_snp) / 4;
        h->tail_slabs = static_cast<int>(std::max<int64_t>(0, std::min<int64_t>(want, slab_bytes > 0 ? spare / slab_bytes : 0)));
        h->n_zbuf = 2 + h->tail_slabs;
    }
    if (k_cap * per_snp > avail)
        throw CudaError{GGRM_ERR_OUT_OF_MEMORY, "not enough device memory for accumulators plus one SNP chunk (need " +
                                                    std::to_string((k_cap * per_snp) >> 20) + " MiB more)"};
    h->k_cap = k_cap;

    for (int kind = 0; kind < 2; ++kind) {
        if (!(kind == 0 ? cfg->want_add : cfg->want_dom)) continue;
        for (int p = 0; p < pool_n; ++p) {
            void* buf = nullptr;
            dev_alloc(*h, &buf, static_cast<size_t>(accum_bytes * h->ksplit), kind == 0 ? "accumulator (kind 0)" : "accumulator (kind 1)");
            h->pool[kind].push_back(buf);
            h->pool_owner[kind].push_back(-1);
        }
        if (want_whole) {
            dev_alloc(*h, &h->whole[kind], static_cast<size_t>(accum_bytes), "whole-genome accumulator");
            GGRM_CUDA(cudaMemsetAsync(h->whole[kind], 0, static_cast<size_t>(accum_bytes), h->s_comp));
        }
        for (int b = 0; b < h->n_zbuf; ++b) dev_alloc(*h, &h->d_Z[kind][b], static_cast<size_t>(h->n_pad * k_cap * z_elem), "operand slab Z");
        if (h->fp64) {
            dev_alloc(*h, &h->d_tab64[kind], static_cast<size_t>(k_cap) * sizeof(SnpTable), "value tables (f64)");
        } else {
            dev_alloc(*h, &h->d_tab16[kind], static_cast<size_t>(k_cap * z_cols_per_k) * sizeof(uint2), "value tables (16-bit)");
            for (int b = 0; b < h->n_zbuf; ++b)
                encode_tensor_map(&h->tmap[kind][b], h->d_Z[kind][b], k_cap * z_cols_per_k, h->n_pad, cfg->precision == GGRM_PREC_FP16);
        }
    }
    for (int i = 0; i < 2; ++i) dev_alloc(*h, &h->d_raw[i], static_cast<size_t>(k_cap * h->bpv_raw), "raw BED chunk");
    dev_alloc(*h, &h->d_dense, static_cast<size_t>(k_cap * h->bpv_pad), "dense 2-bit chunk");
    if (cfg->keep_stats) {
        dev_alloc(*h, &h->d_counts, static_cast<size_t>(cfg->m_total) * sizeof(int4), "per-SNP counts");
        dev_alloc(*h, &h->d_freq, static_cast<size_t>(cfg->m_total) * sizeof(double), "per-SNP frequencies");
        GGRM_CUDA(cudaMemsetAsync(h->d_counts, 0, static_cast<size_t>(cfg->m_total) * sizeof(int4), h->s_comp));
        GGRM_CUDA(cudaMemsetAsync(h->d_freq, 0, static_cast<size_t>(cfg->m_total) * sizeof(double), h->s_comp));
    }
    // one launch for K1-K3 when the whole pass is one slab of few samples (nothing to overlap the prep with; see prep_fused_kernel)
    h->fused_prep = !h->fp64 && h->n_pad <= 4096 && round_up(cfg->m_total, 64) <= k_cap;
    if (const char* e = std::getenv("GGRM_FUSED_PREP")) h->fused_prep = std::atoi(e) != 0 && !h->fp64 && h->n_pad <= 4096;
    if (const char* e = std::getenv("GGRM_DIRECT_SRC")) h->direct_src_ok = std::atoi(e) != 0;
    if (h->fused_prep)
        GGRM_CUDA(cudaFuncSetAttribute(prep_fused_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(prep_fused_smem_bytes(h->bpv_pad))));
    if (cfg->second_matrix == GGRM_SECOND_NONMISSING_GCTA_DIAG) {
        dev_alloc(*h, &h->d_gcta_w, static_cast<size_t>(k_cap) * sizeof(double4), "GCTA diagonal terms of a chunk");
        dev_alloc(*h, &h->d_gcta_lanes, static_cast<size_t>(kGctaLanes * h->n_pad) * sizeof(double), "GCTA diagonal lane sums");
        GGRM_CUDA(cudaMemsetAsync(h->d_gcta_lanes, 0, static_cast<size_t>(kGctaLanes * h->n_pad) * sizeof(double), h->s_comp));
        dev_alloc(*h, &h->d_gcta_diag, static_cast<size_t>(h->n_pad) * sizeof(double), "GCTA diagonal sums");
        dev_alloc(*h, &h->d_gcta_total, static_cast<size_t>(h->n_pad) * sizeof(double), "GCTA diagonal sums over all shards");
        GGRM_CUDA(cudaMemsetAsync(h->d_gcta_diag, 0, static_cast<size_t>(h->n_pad) * sizeof(double), h->s_comp));
        GGRM_CUDA(cudaMemsetAsync(h->d_gcta_total, 0, static_cast<size_t>(h->n_pad) * sizeof(double), h->s_comp));
    }
    lap("device buffers allocated");
    GGRM_CUDA(cudaMalloc(&h->d_trace, sizeof(double)));
    GGRM_CUDA(cudaMalloc(&h->d_wave_ctr, sizeof(unsigned long long)));
    GGRM_CUDA(cudaMemsetAsync(h->d_wave_ctr, 0, sizeof(unsigned long long), h->s_comp));
    if (const char* e = std::getenv("GGRM_WAVE_SYNC")) h->sync_every = std::max(0, std::atoi(e));
    GGRM_CUDA(cudaMalloc(&h->d_fault, sizeof(int)));
    GGRM_CUDA(cudaMemsetAsync(h->d_fault, 0, sizeof(int), h->s_comp));
    GGRM_CUDA(cudaMalloc(&h->d_range_flag, sizeof(unsigned long long)));
    GGRM_CUDA(cudaMemsetAsync(h->d_range_flag, 0, sizeof(unsigned long long), h->s_comp));

    if (!h->fp64) {
        GGRM_CUDA(cudaFuncSetAttribute(syrk_tc_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, TcConfig<1>::kSmemBytes));
        GGRM_CUDA(cudaFuncSetAttribute(syrk_tc_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, TcConfig<2>::kSmemBytes));
        // Ask for the largest smem carve-out (228 KB): the ring needs ~194 KB, and the ~30 KB left over is what lets K1-K3
        // blocks of the next slab (<= 6.6 KB each) become resident beside the persistent SYRK CTA.  With the default
        // (smallest fitting carve-out, 196 KB) they were starved until the SYRK kernel exited.
        GGRM_CUDA(cudaFuncSetAttribute(syrk_tc_kernel<1>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
        GGRM_CUDA(cudaFuncSetAttribute(syrk_tc_kernel<2>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
    }
    GGRM_CUDA(cudaStreamSynchronize(h->s_comp));
    lap("kernel attributes + sync");
}

}  // namespace ggrm

using namespace ggrm;

#define GGRM_API_BEGIN(h)                          \
    if ((h) == nullptr) return GGRM_ERR_INVALID_ARGUMENT; \
    try {                                          \
        GGRM_CUDA(cudaSetDevice((h)->cfg.device));
#define GGRM_API_END(h)                            \
    }                                              \
    catch (const CudaError& e) {                   \
        (h)->last_error = e.msg;                   \
        return e.status;                           \
    }                                              \
    catch (const std::bad_alloc&) {                \
        (h)->last_error = "host allocation failed"; \
        return GGRM_ERR_OUT_OF_MEMORY;             \
    }                                              \
    catch (const std::exception& e) {              \
        (h)->last_error = e.what();                \
        return GGRM_ERR_INVALID_OPERATION;         \
    }                                              \
    return GGRM_OK;

extern "C" int ggrm_create(const ggrm_config_t* cfg, ggrm_t** out) {
    if (cfg == nullptr || out == nullptr) {
        g_create_error = "null argument";
        return GGRM_ERR_INVALID_ARGUMENT;
    }
    *out = nullptr;
    Handle* h = nullptr;
    try {
        h = new Handle();
        create_impl(cfg, h);
    } catch (const CudaError& e) {
        g_create_error = e.msg;
        free_handle(h);
        return e.status;
    } catch (const std::bad_alloc&) {
        g_create_error = "host allocation failed";
        free_handle(h);
        return GGRM_ERR_OUT_OF_MEMORY;
    } catch (const std::exception& e) {  // nothing may unwind through the C ABI
        g_create_error = e.what();
        free_handle(h);
        return GGRM_ERR_INVALID_OPERATION;
    }
    *out = reinterpret_cast<ggrm_t*>(h);
    return GGRM_OK;
}

extern "C" void ggrm_destroy(ggrm_t* h) { free_handle(reinterpret_cast<Handle*>(h)); }

extern "C" const char* ggrm_last_error(const ggrm_t* h) {
    if (h == nullptr) return g_create_error.c_str();
    return reinterpret_cast<const Handle*>(h)->last_error.c_str();
}

extern "C" int ggrm_device_count(void) {
    int c = 0;
    if (cudaGetDeviceCount(&c) != cudaSuccess) return 0;
    return c;
}

extern "C" int ggrm_warm_device(int device) {
    try {
        check_device(device, nullptr);
        GGRM_CUDA(cudaFree(nullptr));  // primary context
        cudaFuncAttributes fa{};       // module load (lazy loading would otherwise pay it at the first launch)
        GGRM_CUDA(cudaFuncGetAttributes(&fa, syrk_tc_kernel<2>));
        GGRM_CUDA(cudaFuncGetAttributes(&fa, expand16_kernel));
        GGRM_CUDA(cudaFuncGetAttributes(&fa, count_stats_kernel));
    } catch (const CudaError& e) {
        g_create_error = e.msg;
        return e.status;
    } catch (const std::exception& e) {
        g_create_error = e.what();
        return GGRM_ERR_INVALID_OPERATION;
    }
    return GGRM_OK;
}

extern "C" const char* ggrm_version(void) { return "gelex-grm-b200 0.1 (sm_100a; tcgen05 SYRK)"; }

#include "ggrm_api_stream.inl"
#include "ggrm_api_fetch.inl"
#include "ggrm_api_hooks.inl"
#include "ggrm_api_stage.inl"
#include "ggrm_api_noalloc.inl"
