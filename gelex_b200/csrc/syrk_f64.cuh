// FP64 SYRK for the close-agreement mode (north star: |G - G_ref| <= 1e-10 on G/(trace/n)).
// G_tile += Zd_I * Zd_J^T with FP64 tensor-core DMMA (mma.sync.aligned.m8n8k4.f64), replacing cblas_dsyrk
// (src/data/grm/grm.cpp:51-62) for the lower-triangle blocks this handle owns.
//   Zd : [n_pad][k_pad] doubles, K-major (one row = one sample)
//   G  : tile-packed, each owned 256x256 tile stored row-major (doubles)
// Block = 64x64 output sub-block, 4 warps (2x2), warp tile 32x32 = 4x4 DMMA tiles, K step 16 with cp.async double
// buffering.  Arithmetic intensity is low by tcgen05 standards but the FP64 pipe (~37 TFLOP/s) is the bound here.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace ggrm {

struct F64Item {
    int32_t row0;     // first global row of the 64-row A block
    int32_t col0;     // first global row (= GRM column) of the 64-row B block
    int64_t out_off;  // element offset of (row0, col0) inside the tile-packed accumulator (row pitch 256)
};

__device__ __forceinline__ void cp_async16(void* smem, const void* gmem) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(static_cast<uint32_t>(__cvta_generic_to_shared(smem))),
                 "l"(gmem)
                 : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
    asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ void dmma_8x8x4(double& c0, double& c1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                 : "+d"(c0), "+d"(c1)
                 : "d"(a), "d"(b));
}

constexpr int kF64Pitch = 20;  // doubles per smem row (16 data + 4 pad): conflict-free half-warp fragment loads

__global__ void __launch_bounds__(128) syrk_f64_kernel(const double* __restrict__ Z, int64_t k_pad, int64_t k_len,
                                                       double* __restrict__ G, const F64Item* __restrict__ items,
                                                       int num_items, int accumulate) {
    __shared__ __align__(16) double sA[2][64][kF64Pitch];
    __shared__ __align__(16) double sB[2][64][kF64Pitch];
    for (int item_idx = blockIdx.x; item_idx < num_items; item_idx += gridDim.x) {
        const F64Item it = items[item_idx];
        const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
        const int wr = (warp >> 1) * 32, wc = (warp & 1) * 32;
        const bool diag = it.row0 == it.col0;
        double acc[4][4][2];
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) acc[i][j][0] = acc[i][j][1] = 0.0;

        const int ksteps = static_cast<int>((k_len + 15) / 16);
        auto load_stage = [&](int stage, int ks) {
            // 64 rows x 16 doubles = 512 x 16-byte pieces per operand; 128 threads -> 4 pieces each
#pragma unroll
            for (int p = 0; p < 4; ++p) {
                const int piece = p * 128 + tid;
                const int r = piece >> 3, c = (piece & 7) * 2;
                cp_async16(&sA[stage][r][c], Z + static_cast<int64_t>(it.row0 + r) * k_pad + ks * 16 + c);
                if (!diag) cp_async16(&sB[stage][r][c], Z + static_cast<int64_t>(it.col0 + r) * k_pad + ks * 16 + c);
            }
            cp_async_commit();
        };
        load_stage(0, 0);
        for (int ks = 0; ks < ksteps; ++ks) {
            const int cur = ks & 1;
            if (ks + 1 < ksteps) {
                load_stage(cur ^ 1, ks + 1);
                cp_async_wait<1>();
            } else {
                cp_async_wait<0>();
            }
            __syncthreads();
            const double(*pa)[kF64Pitch] = sA[cur];
            const double(*pb)[kF64Pitch] = diag ? sA[cur] : sB[cur];
#pragma unroll
            for (int kk = 0; kk < 4; ++kk) {
                double a[4], b[4];
#pragma unroll
                for (int i = 0; i < 4; ++i) a[i] = pa[wr + i * 8 + (lane >> 2)][kk * 4 + (lane & 3)];
#pragma unroll
                for (int j = 0; j < 4; ++j) b[j] = pb[wc + j * 8 + (lane >> 2)][kk * 4 + (lane & 3)];
#pragma unroll
                for (int i = 0; i < 4; ++i)
#pragma unroll
                    for (int j = 0; j < 4; ++j) dmma_8x8x4(acc[i][j][0], acc[i][j][1], a[i], b[j]);
            }
            __syncthreads();
        }
        // C fragment: lane holds (row = lane/4, cols 2*(lane%4) + {0,1}) of each 8x8 tile
#pragma unroll
        for (int i = 0; i < 4; ++i) {
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int r = wr + i * 8 + (lane >> 2);
                const int c = wc + j * 8 + 2 * (lane & 3);
                double2* dst = reinterpret_cast<double2*>(G + it.out_off + static_cast<int64_t>(r) * 256 + c);
                double2 v = make_double2(acc[i][j][0], acc[i][j][1]);
                if (accumulate) {
                    const double2 old = *dst;
                    v.x += old.x;
                    v.y += old.y;
                }
                *dst = v;
            }
        }
    }
}

}  // namespace ggrm
