// K4: lower-triangle SYRK  G += Z Z^T  on the 5th-generation tensor cores (tcgen05 + TMEM), replacing cblas_dsyrk
// (src/data/grm/grm.cpp:51-62).  16-bit operands (BF16 or FP16), FP32 accumulation in TMEM.
//
//   Z      [n_pad][k_cap] 16-bit, K-major (row = sample, SNPs contiguous), read by TMA with 128-byte swizzle
//   G      tile-packed FP32 accumulator: every owned 256x256 tile is two 128-row halves, each stored as
//          [64 float4-columns][128 rows][4] so that an epilogue warp (one TMEM lane = one row per thread) writes
//          512 contiguous bytes per instruction
//   items  persistent work list, ordered by 2048x2048 super-tiles so that CTAs running concurrently share Z panels
//          in L2 (each panel K-slab leaves HBM about once per super-tile, ~1000 FLOP per HBM byte)
//
// Warp roles (256 threads): warp 0 TMA producer, warp 1 MMA issuer, warp 2 TMEM allocator, warps 4-7 epilogue.
// Pipelines: smem ring full/empty (TMA <-> MMA), TMEM double buffer full/empty (MMA <-> epilogue).
//
// kCtaGroup = 1 : one CTA computes a 128x256 half tile  (UMMA 128x256x16)
// kCtaGroup = 2 : a CTA pair computes a 256x256 tile     (UMMA 256x256x16, each CTA stages its own 128 A rows and half
//                 of the B rows, halving smem/L2 operand traffic per FLOP)
#pragma once
#include <cstdint>
#include <cuda.h>
#include <cuda_runtime.h>

#include "ptx_sm100.cuh"

namespace ggrm {

struct TcItem {
    int32_t a_row0;   // first GRM row of the item (128 rows for kCtaGroup 1, 256 for kCtaGroup 2)
    int32_t b_row0;   // first GRM column (= Z row) of the 256-column block
    int64_t out_off;  // element offset of the item's first half tile in the accumulator (of copy `kpart` in split-K launches)
    int32_t kpart;    // split-K: this item covers K blocks [kpart * nkb / ksplit, (kpart + 1) * nkb / ksplit) of the slab
    int32_t pad_;
};

// K range of one work item.  ksplit == 1: the whole slab.  Few-tile problems (n <= ~2 300: fewer tiles than CTA pairs) are cut along
// K instead, every part accumulating into its own copy of the tile (summed once per group by axpy_accum_kernel), so that one launch
// still fills the machine - config C1 has 36 tiles for 74 CTA pairs.
__device__ __forceinline__ void tc_k_range(const TcItem& item, int num_k_blocks, int ksplit, int* kb0, int* kb1) {
    *kb0 = static_cast<int>(static_cast<int64_t>(item.kpart) * num_k_blocks / ksplit);
    *kb1 = static_cast<int>(static_cast<int64_t>(item.kpart + 1) * num_k_blocks / ksplit);
}

constexpr int kTcBlockK = 64;                 // 64 x 16-bit = 128 B = one swizzle row
constexpr int kTcABytes = 128 * 128;          // 128 rows x 128 B
constexpr int kHalfTileElems = 128 * 256;     // one 128-row half of a storage tile

template <int kCtaGroup>
struct TcConfig {
    static constexpr int kBRows = kCtaGroup == 1 ? 256 : 128;  // B rows staged by one CTA
    static constexpr int kBBytes = kBRows * 128;
    static constexpr int kStageBytes = kTcABytes + kBBytes;
    static constexpr int kStages = kCtaGroup == 1 ? 4 : 6;
    static constexpr int kSmemBytes = kStages * kStageBytes + 1024 /*align*/ + 256 /*barriers*/;
    static constexpr uint32_t kUmmaM = 128 * kCtaGroup;
    static constexpr uint32_t kUmmaN = 256;
};

template <int kCtaGroup>
__global__ void __launch_bounds__(256, 1)
syrk_tc_kernel(const __grid_constant__ CUtensorMap tmap_z, float* __restrict__ accum, const TcItem* __restrict__ items,
               int num_items, int num_k_blocks, int ksplit, int lo_blk, int k_rotate, int accumulate, int accumulate_split, uint32_t ab_format,
               int cache_hints, int* fault_flag, unsigned long long* wave_ctr, unsigned long long wave_base, int sync_every,
               float* __restrict__ fold) {
    // lo_blk > 0: BF16 hi + lo operands (GGRM_PREC_BF16 on short inputs).  The operand matrix holds hi = bf16(z) in K blocks
    // [0, lo_blk) and lo = bf16(z - hi) in [lo_blk, 2 lo_blk); every K block is accumulated three times - hi hi^T, hi lo^T,
    // lo hi^T - which restores ~16 mantissa bits for 3x the tensor time (lo lo^T ~ 2^-18 relative is dropped).
    const int nsub = lo_blk > 0 ? 3 : 1;
    using Cfg = TcConfig<kCtaGroup>;
    constexpr int kStages = Cfg::kStages;
    extern __shared__ uint8_t smem_raw[];
    // SWIZZLE_128B operand tiles need 1024-byte alignment.
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* smem_a = smem;
    uint8_t* smem_b = smem + kStages * kTcABytes;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kStages * Cfg::kStageBytes);
    uint64_t* full_bar = bars;
    uint64_t* empty_bar = bars + kStages;
    uint64_t* tmem_full_bar = bars + 2 * kStages;
    uint64_t* tmem_empty_bar = bars + 2 * kStages + 2;
    uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(bars + 2 * kStages + 4);

    const int warp = threadIdx.x >> 5;
    const uint32_t cta_rank = kCtaGroup == 2 ? ptx::cluster_ctarank() : 0u;
    const bool leader = cta_rank == 0;
    const int cluster_id = blockIdx.x / kCtaGroup;
    const int num_clusters = gridDim.x / kCtaGroup;

    if (warp == 0 && ptx::elect_one()) ptx::prefetch_tensormap(&tmap_z);
    if (warp == 1 && ptx::elect_one()) {
        for (int s = 0; s < kStages; ++s) {
            ptx::mbar_init(&full_bar[s], kCtaGroup);  // leader: own arrive(+expect_tx) and, in pair mode, the peer's arrive
            ptx::mbar_init(&empty_bar[s], 1);         // one tcgen05.commit
        }
        for (int s = 0; s < 2; ++s) {
            ptx::mbar_init(&tmem_full_bar[s], 1);                   // one tcgen05.commit
            ptx::mbar_init(&tmem_empty_bar[s], 128 * kCtaGroup);    // every epilogue thread of the pair
        }
        ptx::fence_mbar_init();
    }
    if constexpr (kCtaGroup == 2) ptx::cluster_sync_all();  // peer barriers must exist before anyone signals them
    if (warp == 2) ptx::tmem_alloc<kCtaGroup>(tmem_ptr_smem, 512);
    ptx::tc_fence_before();
    if constexpr (kCtaGroup == 2) ptx::cluster_sync_all(); else __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr_smem;

    if (warp == 0) {
        // ===================================================== TMA producer
        uint32_t stage = 0, phase = 0;
        // cache_hints bit 0: keep Z panels in L2 (evict_last) — they are re-read by every CTA of the super-tile
        // cache_hints bit 2: only the row (A) panels are kept; column (B) panels change every wave and are marked evict_first
        const uint64_t z_hint = (cache_hints & 1) ? ptx::kCacheEvictLast : ptx::kCacheEvictNormal;
        const uint64_t b_hint = (cache_hints & 4) ? ptx::kCacheEvictFirst : z_hint;
        // Wave pacing (sync_every > 0): every sync_every-th wave, each CTA's producer checks in on a global counter and
        // waits (bounded) until all CTAs have — the tiles in flight then stay within a few K blocks of each other, so a Z
        // panel needed by several CTAs is fetched from HBM once and hit in L2 by the rest.  Purely a timing hint: on
        // timeout the producer just proceeds.
        int wave = 0;
        unsigned long long syncs_done = 0;
        for (int it = cluster_id; it < num_items; it += num_clusters, ++wave) {
            if (sync_every > 0 && wave > 0 && wave % sync_every == 0) {
                ++syncs_done;
                if (ptx::elect_one()) {
                    atomicAdd(wave_ctr, 1ull);
                    const unsigned long long target = wave_base + syncs_done * gridDim.x;
                    const uint64_t t0 = ptx::global_timer_ns();
                    while (ptx::ld_relaxed_gpu_u64(wave_ctr) < target && ptx::global_timer_ns() - t0 < 200000ull) __nanosleep(100);
                }
                __syncwarp();
            }
            const TcItem item = items[it];
            const int32_t a_row = item.a_row0 + static_cast<int32_t>(cta_rank) * 128;
            const int32_t b_row = item.b_row0 + static_cast<int32_t>(cta_rank) * Cfg::kBRows;
            int kb0, kb1;
            tc_k_range(item, num_k_blocks, ksplit, &kb0, &kb1);
            // k_rotate != 0 (operand slab resident in L2): a tile starts its K range at a block derived from its own coordinates and
            // wraps around - the sum over K blocks is the same set in another order that depends on the TILE only, so a tile gives
            // the same bits whichever handle owns it (2D-tile ranks reproduce the single-GPU file byte for byte)
            const int klen = kb1 - kb0;
            const unsigned tile_key = static_cast<unsigned>(item.a_row0 >> 7) * 7u + static_cast<unsigned>(item.b_row0 >> 8) * static_cast<unsigned>(k_rotate);
            const int rot = k_rotate != 0 && klen > 1 ? static_cast<int>(tile_key % static_cast<unsigned>(klen)) : 0;
            for (int kk = 0; kk < klen; ++kk) {
                const int kb = kb0 + (kk + rot < klen ? kk + rot : kk + rot - klen);
                for (int sub = 0; sub < nsub; ++sub) {
                    ptx::mbar_wait(&empty_bar[stage], phase ^ 1, fault_flag, 1);
                    if (ptx::elect_one()) {
                        const int32_t kc_a = (kb + (sub == 2 ? lo_blk : 0)) * kTcBlockK;  // hi, hi, lo
                        const int32_t kc_b = (kb + (sub == 1 ? lo_blk : 0)) * kTcBlockK;  // hi, lo, hi
                        uint8_t* sa = smem_a + stage * kTcABytes;
                        uint8_t* sb = smem_b + stage * Cfg::kBBytes;
                        if constexpr (kCtaGroup == 1) {
                            ptx::mbar_arrive_expect_tx(&full_bar[stage], Cfg::kStageBytes);
                            ptx::tma_load_2d(&tmap_z, &full_bar[stage], sa, kc_a, a_row, z_hint);
                            ptx::tma_load_2d(&tmap_z, &full_bar[stage], sb, kc_b, b_row, b_hint);
                            ptx::tma_load_2d(&tmap_z, &full_bar[stage], sb + kTcABytes, kc_b, b_row + 128, b_hint);
                        } else {
                            // both CTAs' bytes are counted on the leader's barrier
                            if (leader) ptx::mbar_arrive_expect_tx(&full_bar[stage], 2 * Cfg::kStageBytes);
                            else ptx::mbar_arrive_cluster(&full_bar[stage], 0);
                            ptx::tma_load_2d_pair(&tmap_z, &full_bar[stage], sa, kc_a, a_row, z_hint);
                            ptx::tma_load_2d_pair(&tmap_z, &full_bar[stage], sb, kc_b, b_row, b_hint);
                        }
                    }
                    __syncwarp();
                    if (++stage == kStages) { stage = 0; phase ^= 1; }
                }
            }
        }
        if (sync_every > 0) {  // check in for the sync points of waves this CTA had no tile in
            const int waves = (num_items + num_clusters - 1) / num_clusters;
            const unsigned long long total = waves > 1 ? static_cast<unsigned long long>((waves - 1) / sync_every) : 0ull;
            if (total > syncs_done && ptx::elect_one()) atomicAdd(wave_ctr, total - syncs_done);
        }
    } else if (warp == 1 && leader) {
        // ===================================================== MMA issuer (leader CTA only)
        const uint32_t idesc = ptx::make_idesc_f16(Cfg::kUmmaM, Cfg::kUmmaN, ab_format);
        uint32_t stage = 0, phase = 0, local_it = 0;
        for (int it = cluster_id; it < num_items; it += num_clusters, ++local_it) {
            const uint32_t as = local_it & 1, aphase = (local_it >> 1) & 1;
            ptx::mbar_wait(&tmem_empty_bar[as], aphase ^ 1, fault_flag, 2);
            ptx::tc_fence_after();
            const uint32_t tmem_d = tmem_base + as * Cfg::kUmmaN;
            int kb0, kb1;
            tc_k_range(items[it], num_k_blocks, ksplit, &kb0, &kb1);
            const int n_steps = (kb1 - kb0) * nsub;  // >= 1: the host never launches a split with more parts than K blocks
            for (int kb = 0; kb < n_steps; ++kb) {
                ptx::mbar_wait(&full_bar[stage], phase, fault_flag, 3);
                ptx::tc_fence_after();
                if (ptx::elect_one()) {
                    const uint64_t da = ptx::make_smem_desc_sw128(ptx::smem_u32(smem_a + stage * kTcABytes));
                    const uint64_t db = ptx::make_smem_desc_sw128(ptx::smem_u32(smem_b + stage * Cfg::kBBytes));
#pragma unroll
                    for (int k = 0; k < kTcBlockK / 16; ++k) {
                        // advance 16 elements = 32 bytes inside the 128-byte swizzle row: +2 in the (addr >> 4) field
                        ptx::umma_f16<kCtaGroup>(tmem_d, da + 2 * k, db + 2 * k, idesc, (kb | k) != 0 ? 1u : 0u);
                    }
                    // frees the smem slot (both CTAs in pair mode) once the MMAs above have read it
                    if constexpr (kCtaGroup == 1) ptx::umma_commit(&empty_bar[stage]);
                    else ptx::umma_commit_pair(&empty_bar[stage], 0x3);
                    if (kb == n_steps - 1) {
                        if constexpr (kCtaGroup == 1) ptx::umma_commit(&tmem_full_bar[as]);
                        else ptx::umma_commit_pair(&tmem_full_bar[as], 0x3);
                    }
                }
                __syncwarp();
                if (++stage == kStages) { stage = 0; phase ^= 1; }
            }
        }
    } else if (warp >= 4) {
        // ===================================================== epilogue: TMEM -> registers -> G (+=)
        const uint32_t quad = warp & 3;  // TMEM lanes 32*quad .. 32*quad+31 are the ones this warp may read
        const uint32_t lane = ptx::lane_id();
        const uint32_t row = quad * 32 + lane;
        uint32_t local_it = 0;
        const uint64_t stream_policy = ptx::l2_policy_evict_first();
        const bool stream = (cache_hints & 2) != 0;  // bit 1: accumulator read-modify-write bypasses L1 and is first to leave L2
        for (int it = cluster_id; it < num_items; it += num_clusters, ++local_it) {
            const TcItem item = items[it];
            const uint32_t as = local_it & 1, aphase = (local_it >> 1) & 1;
            ptx::mbar_wait(&tmem_full_bar[as], aphase, fault_flag, 4);
            ptx::tc_fence_after();
            const int64_t tile_off = item.out_off + static_cast<int64_t>(cta_rank) * kHalfTileElems + row * 4;
            float* out = accum + tile_off;
            // LOCO: on the last slab of a chromosome the finished partial is also added into the whole-genome accumulator here,
            // while the tile is in registers (replaces a separate 3-access whole += partial pass over 5 GB per chromosome)
            float* out2 = fold != nullptr ? fold + tile_off : nullptr;
            const int acc_item = item.kpart == 0 ? accumulate : accumulate_split;
#pragma unroll 1
            for (int chunk = 0; chunk < 8; ++chunk) {
                uint32_t v[32];
                ptx::tmem_ld_32x32(tmem_base + ((quad * 32u) << 16) + as * Cfg::kUmmaN + chunk * 32, v);
                ptx::tmem_ld_wait();
#pragma unroll
                for (int q = 0; q < 8; ++q) {
                    float4* dst = reinterpret_cast<float4*>(out + (chunk * 8 + q) * (128 * 4));
                    float4 r = make_float4(__uint_as_float(v[4 * q]), __uint_as_float(v[4 * q + 1]),
                                           __uint_as_float(v[4 * q + 2]), __uint_as_float(v[4 * q + 3]));
                    if (acc_item) {
                        const float4 o = stream ? ptx::ld_stream_f4(dst, stream_policy) : *dst;
                        r.x += o.x; r.y += o.y; r.z += o.z; r.w += o.w;
                    }
                    if (stream) ptx::st_stream_f4(dst, r, stream_policy); else *dst = r;
                    if (out2 != nullptr) {
                        float4* dst2 = reinterpret_cast<float4*>(out2 + (chunk * 8 + q) * (128 * 4));
                        float4 w = stream ? ptx::ld_stream_f4(dst2, stream_policy) : *dst2;
                        w.x += r.x; w.y += r.y; w.z += r.z; w.w += r.w;
                        if (stream) ptx::st_stream_f4(dst2, w, stream_policy); else *dst2 = w;
                    }
                }
            }
            ptx::tc_fence_before();
            if constexpr (kCtaGroup == 1) ptx::mbar_arrive(&tmem_empty_bar[as]);
            else ptx::mbar_arrive_cluster(&tmem_empty_bar[as], 0);
        }
    }

    ptx::tc_fence_before();
    if constexpr (kCtaGroup == 2) ptx::cluster_sync_all(); else __syncthreads();
    if (warp == 2) ptx::tmem_dealloc<kCtaGroup>(tmem_base, 512);
}

}  // namespace ggrm
