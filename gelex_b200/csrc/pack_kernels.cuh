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
                                                        const TAcc* __restrict__ D,  // entrywise divisor (GCTA N matrix) or nullptr
                                                        const int32_t* __restrict__ slot_of, int64_t row_begin,
                                                        int64_t row_count, double scale, TOut* __restrict__ out,
                                                        double* __restrict__ trace,
                                                        const double* __restrict__ gdiag = nullptr) {  // GCTA diagonal sums (with D): A_ii = 1 + gdiag[i] / D_ii
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
            if (D != nullptr) {
                const double d = static_cast<double>(D[off]);
                if (gdiag != nullptr && j == i) v = d > 0.0 ? 1.0 + __ddiv_rn(gdiag[i], d) : 0.0;
                else v = d > 0.0 ? __ddiv_rn(v, d) : 0.0;
            }
        }
        out[base + j] = static_cast<TOut>(v);
        if (j == i && trace != nullptr) atomicAdd(trace, v);
    }
}

// FP32 accumulators (the tcgen05 path): the same read-out with one WARP per row and eight consecutive rows per block.  In the
// tile layout ([float4-column][row][4]) a row's float4s are 2 KB apart, so a block that packs a single row uses 16 bytes of
// every 32-byte sector it pulls (measured: 0.35 of the HBM peak, 1.19x the useful bytes from DRAM).  Rows r and r + 1 share the
// sector and eight rows share the 128-byte line: with the eight rows in one block the line is fetched once and served to the
// other seven warps from L1.  A lane reads one float4 (four columns) per step and writes 16 contiguous bytes; a warp writes 512
// contiguous bytes of its row; blockIdx.y walks the row in chunks of kPackTilesPerBlock tile columns so that a read-out of a few
// hundred rows still fills the machine.  Arithmetic is the per-element FP64 form of pack_rows_kernel, so results are bit-identical.
constexpr int kPackTilesPerBlock = 8;
template <typename TOut>
__global__ void __launch_bounds__(256) pack_rows8_kernel(const float* __restrict__ A, const float* __restrict__ B,
                                                         const float* __restrict__ D, const int32_t* __restrict__ slot_of,
                                                         int64_t row_begin, int64_t row_count, double scale,
                                                         TOut* __restrict__ out, double* __restrict__ trace,
                                                         const double* __restrict__ gdiag = nullptr) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int64_t local = static_cast<int64_t>(blockIdx.x) * 8 + warp;
    if (local >= row_count) return;
    const int64_t i = row_begin + local;
    TOut* dst_row = out + (i * (i + 1) / 2 - row_begin * (row_begin + 1) / 2);
    const int64_t I = i >> 8;
    const int r = static_cast<int>(i & 255);
    const int64_t row_off = static_cast<int64_t>(r >> 7) * 32768 + static_cast<int64_t>(r & 127) * 4;  // + float4-column * 512
    const int64_t J_begin = static_cast<int64_t>(blockIdx.y) * kPackTilesPerBlock;
    const int64_t J_end = J_begin + kPackTilesPerBlock < I + 1 ? J_begin + kPackTilesPerBlock : I + 1;
    for (int64_t J = J_begin; J < J_end; ++J) {
        const int32_t slot = slot_of[I * (I + 1) / 2 + J];
        const int64_t tile = static_cast<int64_t>(slot) * 65536 + row_off;
#pragma unroll
        for (int k = 0; k < 2; ++k) {
            const int c4 = lane + 32 * k;
            const int64_t j0 = J * 256 + 4 * c4;
            if (j0 > i) continue;
            float4 a = make_float4(0.f, 0.f, 0.f, 0.f), b = a, d = a;
            if (slot >= 0) {
                a = __ldg(reinterpret_cast<const float4*>(A + tile + c4 * 512));
                if (B != nullptr) b = __ldg(reinterpret_cast<const float4*>(B + tile + c4 * 512));
                if (D != nullptr) d = __ldg(reinterpret_cast<const float4*>(D + tile + c4 * 512));
            }
            const float av[4] = {a.x, a.y, a.z, a.w}, bv[4] = {b.x, b.y, b.z, b.w}, dv[4] = {d.x, d.y, d.z, d.w};
            TOut o[4];
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                double v = 0.0;
                if (slot >= 0) {
                    v = static_cast<double>(av[e]);
                    if (B != nullptr) v -= static_cast<double>(bv[e]);
                    v *= scale;
                    if (D != nullptr) {
                        const double dd = static_cast<double>(dv[e]);
                        if (gdiag != nullptr && j0 + e == i) v = dd > 0.0 ? 1.0 + __ddiv_rn(gdiag[i], dd) : 0.0;
                        else v = dd > 0.0 ? __ddiv_rn(v, dd) : 0.0;
                    }
                }
                o[e] = static_cast<TOut>(v);
                if (j0 + e == i && trace != nullptr) atomicAdd(trace, v);
            }
            TOut* dst = dst_row + j0;
            if (j0 + 3 <= i && (reinterpret_cast<uintptr_t>(dst) & (4 * sizeof(TOut) - 1)) == 0) {
                if constexpr (sizeof(TOut) == 4) {
                    *reinterpret_cast<float4*>(dst) = make_float4(o[0], o[1], o[2], o[3]);
                } else {
                    *reinterpret_cast<double2*>(dst) = make_double2(o[0], o[1]);
                    *reinterpret_cast<double2*>(dst + 2) = make_double2(o[2], o[3]);
                }
            } else {
#pragma unroll
                for (int e = 0; e < 4; ++e)
                    if (j0 + e <= i) dst[e] = o[e];
            }
        }
    }
}

// Rectangular read-out for the 2D-tile multi-GPU mode: rows [row_begin, +row_count) x columns [col_begin, +col_count) of
// the lower triangle as a dense row-major [row_count][col_count] block (entries above the diagonal or in tiles this
// handle does not own read as 0), so a rank can ship exactly the super-tiles it owns to their offsets in the .bin file.
// D / gdiag: the GCTA ratio form of pack_rows_kernel (same FP64 arithmetic, so both partitions write the same bytes).
template <typename TAcc, typename TOut>
__global__ void __launch_bounds__(256) pack_block_kernel(const TAcc* __restrict__ A, const int32_t* __restrict__ slot_of,
                                                         int64_t row_begin, int64_t row_count, int64_t col_begin,
                                                         int64_t col_count, TOut* __restrict__ out,
                                                         const TAcc* __restrict__ D = nullptr, const double* __restrict__ gdiag = nullptr) {
    if (blockIdx.x >= row_count) return;
    const int64_t i = row_begin + blockIdx.x;
    const int64_t I = i >> 8;
    const int r = static_cast<int>(i & 255);
    TOut* dst = out + static_cast<int64_t>(blockIdx.x) * col_count;
    for (int64_t jl = threadIdx.x; jl < col_count; jl += blockDim.x) {
        const int64_t j = col_begin + jl;
        TOut v = TOut(0);
        if (j <= i) {
            const int32_t slot = slot_of[I * (I + 1) / 2 + (j >> 8)];
            if (slot >= 0) {
                const int64_t off = static_cast<int64_t>(slot) * 65536 + tile_elem_offset<TAcc>(r, static_cast<int>(j & 255));
                if (D == nullptr) {
                    v = static_cast<TOut>(A[off]);
                } else {
                    const double d = static_cast<double>(D[off]);
                    double w = static_cast<double>(A[off]) * 1.0;
                    if (gdiag != nullptr && j == i) w = d > 0.0 ? 1.0 + __ddiv_rn(gdiag[i], d) : 0.0;
                    else w = d > 0.0 ? __ddiv_rn(w, d) : 0.0;
                    v = static_cast<TOut>(w);
                }
            }
        }
        dst[jl] = v;
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

// Split-K launches (few-tile problems, Handle::ksplit): copy 0 += copy 1 + ... + copy (parts - 1), summed in that fixed order, in
// ONE pass (parts reads + one write per element; separate axpy launches cost ~8 us each, which shows at C1's 0.3 ms pass).
__global__ void __launch_bounds__(256) sum_copies_kernel(float* __restrict__ base, int64_t copy_elems, int parts) {
    const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x * 4;
    for (int64_t i = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) * 4; i < copy_elems; i += stride) {
        float4 a = *reinterpret_cast<const float4*>(base + i);
        for (int p = 1; p < parts; ++p) {
            const float4 b = *reinterpret_cast<const float4*>(base + static_cast<int64_t>(p) * copy_elems + i);
            a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w;
        }
        *reinterpret_cast<float4*>(base + i) = a;
    }
}

// Consumer side of the .bin format: packed float32 lower rows of one or two files already on the device ->
// the dense symmetric FP64 matrix GrmLoader / LocoGRMLoader hand to the solvers (src/data/grm/grm_loader.cpp:125-243,
// loco_grm_loader.cpp:62-76):  out(i, j) = (W(fw(i), fw(j)) - C(fc(i), fc(j))) / denom, fw / fc = target row -> row of that file
// (-1: the target has no sample there and contributes 0), column-major.  One 32 x 32 tile per block, lower block
// triangle only: the tile is gathered once with its contiguous file index on threadIdx.x and written twice (as is and
// transposed through shared memory), so both the reads and the 8-byte writes are coalesced for the identity map.
__global__ void __launch_bounds__(256) bin_to_dense_kernel(const float* __restrict__ W, const float* __restrict__ C,
                                                           const int64_t* __restrict__ map_w, const int64_t* __restrict__ map_c,
                                                           int64_t n_target, double denom, double* __restrict__ out) {
    const int64_t bi = blockIdx.y, bj = blockIdx.x;
    if (bj > bi) return;
    __shared__ double tile[32][33];
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;  // 32 x 8
    const int64_t tj = bj * 32 + tx;
    const int64_t wj = tj < n_target ? (map_w != nullptr ? map_w[tj] : tj) : -1;
    const int64_t cj = C != nullptr && tj < n_target ? (map_c != nullptr ? map_c[tj] : tj) : -1;
#pragma unroll
    for (int r = ty; r < 32; r += 8) {
        const int64_t ti = bi * 32 + r;
        double v = 0.0;
        if (ti < n_target) {
            const int64_t wi = map_w != nullptr ? map_w[ti] : ti;
            if (wi >= 0 && wj >= 0) {
                const int64_t a = wi > wj ? wi : wj, b = wi > wj ? wj : wi;
                v = static_cast<double>(W[a * (a + 1) / 2 + b]);
            }
            if (C != nullptr) {
                const int64_t ci = map_c != nullptr ? map_c[ti] : ti;
                if (ci >= 0 && cj >= 0) {
                    const int64_t a = ci > cj ? ci : cj, b = ci > cj ? cj : ci;
                    v -= static_cast<double>(C[a * (a + 1) / 2 + b]);
                }
            }
            v = __ddiv_rn(v, denom);
        }
        tile[r][tx] = v;  // tile[row i][col j]
    }
    __syncthreads();
#pragma unroll
    for (int r = ty; r < 32; r += 8) {
        // out(i, j) at out[j * n + i]: column j = bj*32 + r, rows i = bi*32 + tx contiguous
        const int64_t j = bj * 32 + r, i = bi * 32 + tx;
        if (j < n_target && i < n_target) out[j * n_target + i] = tile[tx][r];
        if (bi != bj) {  // mirror: out(j, i) at out[i * n + j]: column i = bi*32 + r, rows j = bj*32 + tx contiguous
            const int64_t i2 = bi * 32 + r, j2 = bj * 32 + tx;
            if (i2 < n_target && j2 < n_target) out[i2 * n_target + j2] = tile[r][tx];
        }
    }
}

constexpr int kMaxPeers = 7;
struct PeerPtrs {
    const void* p[kMaxPeers];
};

// 16-byte streaming load (peer memory over NVLink, or local HBM that will not be re-read): bypass L1.
__device__ __forceinline__ uint4 ld_stream_u4(const void* p) {
    uint4 v;
    asm volatile("ld.global.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
    return v;
}

// mine[begin, end) += sum_p peers[p][begin, end).  NVLink P2P loads have microseconds of latency, so every thread keeps
// kChunks x (1 + n_peers) independent 16-byte loads in flight (all issued before the first add); a block sweeps
// kChunks x 4 KB contiguous pieces so each warp-level request is a full 512-byte line group.
template <typename T>
__global__ void __launch_bounds__(256) reduce_peers_kernel(T* __restrict__ mine, PeerPtrs peers, int n_peers,
                                                           int64_t begin, int64_t end) {
    constexpr int kVec = 16 / static_cast<int>(sizeof(T));   // elements per 16-byte access
    constexpr int kChunks = 4;
    const int64_t block_span = static_cast<int64_t>(blockDim.x) * kVec * kChunks;
    for (int64_t base = begin + static_cast<int64_t>(blockIdx.x) * block_span; base < end; base += static_cast<int64_t>(gridDim.x) * block_span) {
        uint4 a[kChunks];
        uint4 b[kChunks];
        int64_t idx[kChunks];
#pragma unroll
        for (int u = 0; u < kChunks; ++u) {
            idx[u] = base + (static_cast<int64_t>(u) * blockDim.x + threadIdx.x) * kVec;
            if (idx[u] < end) a[u] = ld_stream_u4(mine + idx[u]);
        }
        for (int p = 0; p < n_peers; ++p) {
            const T* src = static_cast<const T*>(peers.p[p]);
#pragma unroll
            for (int u = 0; u < kChunks; ++u)
                if (idx[u] < end) b[u] = ld_stream_u4(src + idx[u]);
#pragma unroll
            for (int u = 0; u < kChunks; ++u) {
                if (idx[u] >= end) continue;
                if constexpr (sizeof(T) == 4) {
                    a[u].x = __float_as_uint(__uint_as_float(a[u].x) + __uint_as_float(b[u].x));
                    a[u].y = __float_as_uint(__uint_as_float(a[u].y) + __uint_as_float(b[u].y));
                    a[u].z = __float_as_uint(__uint_as_float(a[u].z) + __uint_as_float(b[u].z));
                    a[u].w = __float_as_uint(__uint_as_float(a[u].w) + __uint_as_float(b[u].w));
                } else {
                    const double a0 = __hiloint2double(static_cast<int>(a[u].y), static_cast<int>(a[u].x)) +
                                      __hiloint2double(static_cast<int>(b[u].y), static_cast<int>(b[u].x));
                    const double a1 = __hiloint2double(static_cast<int>(a[u].w), static_cast<int>(a[u].z)) +
                                      __hiloint2double(static_cast<int>(b[u].w), static_cast<int>(b[u].z));
                    a[u].x = static_cast<uint32_t>(__double2loint(a0)); a[u].y = static_cast<uint32_t>(__double2hiint(a0));
                    a[u].z = static_cast<uint32_t>(__double2loint(a1)); a[u].w = static_cast<uint32_t>(__double2hiint(a1));
                }
            }
        }
#pragma unroll
        for (int u = 0; u < kChunks; ++u)
            if (idx[u] < end) *reinterpret_cast<uint4*>(mine + idx[u]) = a[u];
    }
}

}  // namespace ggrm
