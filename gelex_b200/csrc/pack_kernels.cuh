// K5: read-out of the tile-packed accumulators.
//   pack_rows     : accumulator tiles -> packed row-major lower-triangle rows, the byte order GrmBinWriter emits
//                   (src/data/grm/grm_bin_writer.cpp:54-61), FP32 or FP64, plus the FP64 trace of those rows
//                   (GrmResult::denominator, include/gelex/data/grm/grm.h:134).  Optional affine form
//                   out = (A - B) * scale serves the consumer-side arithmetic of grm_loader.cpp:125-165
//                   (B = null, scale = n/trace) and loco_grm_loader.cpp:62-76 (A = whole, B = chromosome).
//   axpy_accum    : whole += partial (LOCO: whole = sum of chromosome partials, so the whole-genome GRM costs no
//                   extra SYRK pass).
//   reduce_peers  : SNP-shard reduce over NVLink peer memory — this rank sums one slice of every peer's accumulator
//                   into its own with plain P2P loads (reduce-scatter step; no NCCL on this path).
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace ggrm {

// Element offset inside one 256x256 storage tile.
//   FP32 (tcgen05 path): two 128-row halves, each [64 float4-columns][128 rows][4]  (see syrk_tc.cuh epilogue)
//   FP64 (DMMA path)   : plain row-major
template <typename T>
__device__ __forceinline__ int64_t tile_elem_offset(int r, int c);
template <>
__device__ __forceinline__ int64_t tile_elem_offset<float>(int r, int c) {
    return static_cast<int64_t>(r >> 7) * 32768 + (static_cast<int64_t>(c >> 2) * 128 + (r & 127)) * 4 + (c & 3);
}
template <>
__device__ __forceinline__ int64_t tile_elem_offset<double>(int r, int c) {
    return static_cast<int64_t>(r) * 256 + c;
}

template <typename TAcc, typename TOut>
__global__ void __launch_bounds__(256) pack_rows_kernel(const TAcc* __restrict__ A, const TAcc* __restrict__ B,
                                                        const int32_t* __restrict__ slot_of, int64_t row_begin,
                                                        int64_t row_count, double scale, TOut* __restrict__ out,
                                                        double* __restrict__ trace) {
    const int64_t i = row_begin + blockIdx.x;
    if (blockIdx.x >= row_count) return;
    const int64_t base = i * (i + 1) / 2 - row_begin * (row_begin + 1) / 2;
    const int64_t I = i >> 8;
    const int r = static_cast<int>(i & 255);
    for (int64_t j = threadIdx.x; j <= i; j += blockDim.x) {
        const int64_t J = j >> 8;
        const int32_t slot = slot_of[I * (I + 1) / 2 + J];
        double v = 0.0;
        if (slot >= 0) {
            const int64_t off = static_cast<int64_t>(slot) * 65536 + tile_elem_offset<TAcc>(r, static_cast<int>(j & 255));
            v = static_cast<double>(A[off]);
            if (B != nullptr) v -= static_cast<double>(B[off]);
            v *= scale;
        }
        out[base + j] = static_cast<TOut>(v);
        if (j == i && trace != nullptr) atomicAdd(trace, v);
    }
}

template <typename T>
__global__ void __launch_bounds__(256) axpy_accum_kernel(T* __restrict__ dst, const T* __restrict__ src, int64_t n) {
    const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x * 4;
    for (int64_t i = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) * 4; i < n; i += stride) {
        if constexpr (sizeof(T) == 4) {
            float4 a = *reinterpret_cast<const float4*>(dst + i);
            const float4 b = *reinterpret_cast<const float4*>(src + i);
            a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w;
            *reinterpret_cast<float4*>(dst + i) = a;
        } else {
            double2 a0 = *reinterpret_cast<const double2*>(dst + i), a1 = *reinterpret_cast<const double2*>(dst + i + 2);
            const double2 b0 = *reinterpret_cast<const double2*>(src + i), b1 = *reinterpret_cast<const double2*>(src + i + 2);
            a0.x += b0.x; a0.y += b0.y; a1.x += b1.x; a1.y += b1.y;
            *reinterpret_cast<double2*>(dst + i) = a0;
            *reinterpret_cast<double2*>(dst + i + 2) = a1;
        }
    }
}

constexpr int kMaxPeers = 7;
struct PeerPtrs {
    const void* p[kMaxPeers];
};

template <typename T>
__global__ void __launch_bounds__(256) reduce_peers_kernel(T* __restrict__ mine, PeerPtrs peers, int n_peers,
                                                           int64_t begin, int64_t end) {
    const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x * 4;
    for (int64_t i = begin + (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) * 4; i < end; i += stride) {
        if constexpr (sizeof(T) == 4) {
            float4 a = *reinterpret_cast<const float4*>(mine + i);
#pragma unroll
            for (int p = 0; p < kMaxPeers; ++p) {
                if (p < n_peers) {
                    const float4 b = *reinterpret_cast<const float4*>(static_cast<const T*>(peers.p[p]) + i);
                    a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w;
                }
            }
            *reinterpret_cast<float4*>(mine + i) = a;
        } else {
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                double2 a = *reinterpret_cast<const double2*>(mine + i + 2 * h);
#pragma unroll
                for (int p = 0; p < kMaxPeers; ++p) {
                    if (p < n_peers) {
                        const double2 b = *reinterpret_cast<const double2*>(static_cast<const T*>(peers.p[p]) + i + 2 * h);
                        a.x += b.x; a.y += b.y;
                    }
                }
                *reinterpret_cast<double2*>(mine + i + 2 * h) = a;
            }
        }
    }
}

}  // namespace ggrm
