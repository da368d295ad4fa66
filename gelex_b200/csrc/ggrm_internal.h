// Internal state of a ggrm handle (one CUDA device).  Not part of the ABI.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>

#include <cstdint>
#include <deque>
#include <string>
#include <vector>

#include "../../include/ggrm.h"
#include "prep_kernels.cuh"
#include "syrk_f64.cuh"
#include "syrk_tc.cuh"

namespace ggrm {

struct CudaError {
    int status;
    std::string msg;
};

#define GGRM_CUDA(expr)                                                                                              \
    do {                                                                                                             \
        cudaError_t _e = (expr);                                                                                     \
        if (_e != cudaSuccess) {                                                                                     \
            throw ::ggrm::CudaError{_e == cudaErrorMemoryAllocation ? GGRM_ERR_OUT_OF_MEMORY : GGRM_ERR_CUDA,        \
                                    std::string("CUDA: ") + cudaGetErrorString(_e) + " at " #expr};                  \
        }                                                                                                            \
    } while (0)

struct GroupState {
    int64_t begin = 0, end = 0;
    int buffer[2] = {-1, -1};     // accumulator pool slot per kind (-1 = none yet / released)
    bool started = false;
    bool touched[2] = {false, false};  // a SYRK has written this group's accumulator of that kind (first slab stores, later slabs add)
    bool split_touched[2] = {false, false};  // ... and its split-K copies 1 .. ksplit-1 (Handle::ksplit); summed into copy 0 when the group completes
    bool complete = false;        // all SNPs accumulated (+ folded into whole); `done` event recorded
    bool folded = false;
    cudaEvent_t done = nullptr;
};

struct Timer {
    cudaEvent_t a, b;
    int phase;
};

struct Handle {
    ggrm_config_t cfg{};
    std::string last_error;
    int sm_count = 148;
    // geometry
    int64_t n_raw = 0, n_out = 0, n_pad = 0, bpv_raw = 0, bpv_pad = 0;
    int64_t ntiles_side = 0;              // n_pad / 256
    int64_t k_cap = 0;                    // SNP capacity of one device chunk (multiple of 64)
    bool fp64 = false;
    int elem_bytes = 4;
    int cache_hints = 3;                  // bit 0: Z panels evict_last, bit 1: accumulator stream evict_first (GGRM_CACHE_HINTS)
    int strip_h = 8;                      // SYRK item order: 0 = by 8x8 super-tile, h>0 = strips of h tile rows (GGRM_STRIP_H)
    int tc_mode = 1;                      // 1 = single-CTA UMMA 128x256, 2 = CTA-pair UMMA 256x256
    // ownership and work lists
    std::vector<int32_t> slot_of_host;    // [ntiles_side*(ntiles_side+1)/2] -> slot or -1
    int64_t owned_tiles = 0;
    std::vector<int64_t> owned_supers;    // (super-row, super-col) pairs of the 2048x2048 super-tiles this handle owns, storage order
    int32_t* d_slot_of = nullptr;
    TcItem* d_items_tc = nullptr;
    int num_items_tc = 0;
    // Split-K work list for few-tile problems (fewer tiles than CTA pairs, e.g. config C1: 36 tiles for 74 pairs): every tile
    // appears ksplit times, part p accumulating K blocks [p nkb / ksplit, (p + 1) nkb / ksplit) into copy p of the accumulator
    // (copy p at element offset p * accum_elems; pool buffers hold ksplit copies).  Copies are summed in a fixed order when the
    // group completes, so results stay deterministic.  ksplit == 1: not used.
    TcItem* d_items_split = nullptr;
    int num_items_split = 0;
    int ksplit = 1;
    // 16-bit operands on short inputs: operands are kept as hi = op16(z) and lo = op16(z - hi) (Z slabs and value tables hold
    // both, lo at K offset k_cap) and every K block is accumulated as hi hi^T + hi lo^T + lo hi^T (syrk_tc.cuh, lo_blk).  Decided
    // in create_impl from the precision, the code policy and the SHORTEST group (every LOCO partial is an output of its own).
    bool op_split = false;
    F64Item* d_items_f64 = nullptr;
    int num_items_f64 = 0;
    // Row bands of the tcgen05 work list (whole strips of strip_h tile rows, ~equal tile counts).  The LAST slab of a pass is
    // launched band by band with an event after each, so the read-out of the first rows (pack + D2H on s_fetch) overlaps the
    // SYRK of the later bands (ggrm_finish_async + ggrm_fetch*).
    std::vector<int> band_item_end;       // exclusive item index per band
    std::vector<int64_t> band_row_end;    // exclusive GRM row per band
    std::vector<cudaEvent_t> band_done;
    // ggrm_set_band_flags: flags[b] = value is stored from a CUDA host callback once band b of the pass's final accumulation is
    // complete (multi-process SNP shards pipeline their reduce-scatter + read-out band by band on these)
    int32_t* band_flags = nullptr;
    int32_t band_flag_value = 0;
    bool band_flags_posted = false;
    struct FlagMsg { int32_t* addr; int32_t value; };
    std::deque<FlagMsg> flag_msgs;        // payloads of in-flight callbacks (stable addresses; cleared at reset)
    int banded_group = -1;                // group whose partial was finished band by band in the final flush (-1: none)
    bool banded_whole = false;            // ... and `whole` too (its last fold happened inside those launches)
    // projection
    int32_t* d_out_to_raw = nullptr;      // nullptr = identity
    // streams
    cudaStream_t s_copy = nullptr, s_prep = nullptr, s_comp = nullptr, s_fetch = nullptr;  // H2D | K1-K3 | K4,K6 | K5+D2H
    // chunk buffers
    uint8_t* d_raw[2] = {nullptr, nullptr};
    cudaEvent_t raw_free[2] = {nullptr, nullptr};
    cudaEvent_t h2d_done = nullptr;
    uint8_t* d_dense = nullptr;
    int4* d_counts = nullptr;             // [m_total] when keep_stats
    double* d_freq = nullptr;
    uint2* d_tab16[2] = {nullptr, nullptr};
    SnpTable* d_tab64[2] = {nullptr, nullptr};
    // operand slabs are double buffered so that K1-K3 of slab i+1 (s_prep) overlap the SYRK of slab i (s_comp)
    // plus `tail_slabs` more buffers: the last slabs of a pass stay resident so that they can be accumulated row band by row band
    // (all deferred slabs of band 0, then of band 1, ...) and the read-out of the first bands overlaps ~tail_slabs slabs of SYRK
    static constexpr int kMaxZBuf = 8;
    void* d_Z[2][kMaxZBuf] = {};                                   // [kind][z_buf]
    CUtensorMap tmap[2][kMaxZBuf];
    cudaEvent_t prep_done[kMaxZBuf] = {};                          // recorded on s_prep after the expands into Z[.][b]
    cudaEvent_t z_free[kMaxZBuf] = {};                             // recorded on s_comp after the SYRKs that read Z[.][b]
    int n_zbuf = 2;
    int tail_slabs = 0;                                            // GGRM_TAIL_SLABS (default 4 when memory allows; tcgen05 path only)
    struct DeferredSlab { int zb; int64_t count; int64_t first_snp; };
    std::vector<DeferredSlab> deferred;                            // prepared (Z resident) but not yet accumulated
    int deferred_group = -1;
    bool cur_trailing = false;                                     // the slab being filled is one of the last tail_slabs of the pass
    bool first_slab_done = false;                                  // this pass has started its first slab (see push_impl)
    int64_t slabs_started = 0;                                     // slabs begun in this pass (bounds the parked tail, see push_impl)
    bool first_slab_with_split = false;                            // GGRM_FIRST_SLAB_SPLIT=1: short first slab for few-tile (split-K) problems too
    int64_t first_slab_snps = 4096;                                // GGRM_FIRST_SLAB (0 = off)
    int64_t push_end_hint = -1;                                    // ggrm_set_push_end: no SNP >= this will be pushed (-1: the last group's end)
    int z_buf = 0;
    bool overlap_prep = true;                                      // GGRM_OVERLAP_PREP=0: run K1-K3 on s_comp (serial)
    const uint8_t* direct_src = nullptr;                           // fused prep: the slab being filled IS this device payload (one push = the whole pass): read in place
    bool direct_src_ok = true;                                     // GGRM_DIRECT_SRC=0: always copy
    bool fused_prep = false;                                       // K1-K3 as one launch (prep_fused_kernel): single-slab passes of few samples; GGRM_FUSED_PREP=0/1
    // pinned staging ring (pageable callers)
    static constexpr int kRingSlots = 4;
    uint8_t* ring[kRingSlots] = {nullptr, nullptr, nullptr, nullptr};
    cudaEvent_t ring_free[kRingSlots] = {nullptr, nullptr, nullptr, nullptr};
    int64_t ring_bytes = 0;
    int ring_next = 0;
    // accumulators
    int64_t accum_elems = 0;              // owned_tiles * 65536
    std::vector<void*> pool[2];           // per kind
    std::vector<int> pool_owner[2];       // group index or -1
    void* whole[2] = {nullptr, nullptr};  // n_groups>1 && keep_whole
    std::vector<GroupState> groups;
    // current chunk
    int cur_buf = 0;
    int64_t cur_count = 0, cur_first_snp = -1;
    int64_t cur_cap = 0;                  // SNP capacity of the slab being filled: the rest of its group cut into equal slabs <= k_cap
    int cur_group = -1;
    int64_t last_snp_pushed = -1;
    bool finished = false;
    // fetch staging
    void* d_stage = nullptr;
    int64_t stage_bytes = 0;

    // Debug aid (GGRM_GUARD=1): every device buffer the kernels write sits between two 64 KiB guard bands filled with 0xA5,
    // verified at ggrm_finish - an out-of-bounds store next to a buffer becomes an error instead of silent corruption.
    struct GuardRec { char* base; size_t bytes; size_t tail; const char* name; };
    bool guard = false;
    std::vector<GuardRec> guards;
    double* d_trace = nullptr;
    unsigned long long* d_wave_ctr = nullptr;  // SYRK wave pacing counter (monotonic; wave_base = value all earlier launches add up to)
    unsigned long long wave_base = 0;
    int sync_every = 1;                   // waves between pacing points, 0 = off (GGRM_WAVE_SYNC)
    int* d_fault = nullptr;               // mbarrier-timeout location written by a stuck SYRK before it traps
    // GGRM_SECOND_NONMISSING_GCTA_DIAG: per-SNP diagonal terms of the chunk being prepared, this handle's per-sample sums, and the
    // sum over all SNP shards (ggrm_gcta_diag_reduce_async; the read-out uses it once it has been produced in this pass)
    double4* d_gcta_w = nullptr;          // [k_cap]
    double* d_gcta_lanes = nullptr;       // [16][n_pad] running sums by SNP index mod 16 (gcta_diag_kernel)
    double* d_gcta_diag = nullptr;        // [n_pad] their sum, refreshed after every chunk
    double* d_gcta_total = nullptr;       // [n_pad]
    bool gcta_total_valid = false;
    unsigned long long* d_range_flag = nullptr;  // FP16 operands: 1 + index of a SNP whose standardised value overflows binary16 (0 = none)
    // timers
    std::vector<Timer> timers;
    double phase_ms[7] = {0, 0, 0, 0, 0, 0, 0};
    cudaEvent_t pass_begin = nullptr, pass_end = nullptr;  // first push enqueued -> last accumulation done
    bool pass_open = false;
    double pass_ms = 0;
    int64_t syrk_launches = 0, kernel_launches = 0;
};

// implemented in ggrm_api.cu
void build_ownership_and_items(Handle& h);
void encode_tensor_map(CUtensorMap* map, void* base, int64_t k_cap, int64_t n_pad, bool fp16);

}  // namespace ggrm
