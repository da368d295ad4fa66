// K1-K3 of the GRM path: sample projection (repack), genotype counts + allele frequency + code-policy value
// tables, and expansion of packed 2-bit genotypes into K-major 16-bit (or FP64) operand panels.
// HBM-bound byte/integer work: coalesced 16-byte loads, warp-shuffle reductions, smem-staged transposes.
//
// Reference semantics reproduced here (all file:line under /root/reference):
//   decode LUT           src/data/decode_lut.h:40-41          00->2.0  01->NaN  10->1.0  11->0.0
//   sample projection    src/data/bed_pipe/variant_decoder.cpp:79-126 (scatter, later raw sample wins)
//   per-SNP process()    include/gelex/internal/genotype_processor/genotype_processor.h:35-58
//   encode policies      include/gelex/internal/genotype_processor/encode_policy.h:32-81
//   statistic policies   include/gelex/internal/genotype_processor/statistic_policy.h:43-175
#pragma once
#include <cstdint>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>

#include "../../include/ggrm.h"

namespace ggrm {

constexpr int kTile = 256;            // GRM storage tile edge; n is padded to a multiple of this
constexpr uint8_t kAllMissingByte = 0x55;  // four BED "01" codes

// ------------------------------------------------------------------------------------------------------------------
// K1a: repack.  raw [c][bpv_raw] -> dense [c][bpv_pad] in GRM row order; rows >= n_out are BED 01 (missing) so that
// every later kernel can treat padding exactly like a missing call (contributes 0 after mean imputation).
// out_to_raw == nullptr means identity (the reference's dense path, sample_projection.cpp:51-54).
// A thread produces 4 output bytes (16 samples) per iteration; threads walk the chunk's output words (snp, word) in one flat
// grid-stride sequence, so short rows (n = 2 000: 128 words) fill their blocks as well as long ones.
// One output word (16 samples, GRM rows 16 w .. 16 w + 15) of one SNP row.
__device__ __forceinline__ uint32_t repack_word(const uint8_t* __restrict__ src, int64_t w, int64_t n_out, const int32_t* __restrict__ out_to_raw,
                                                bool words_aligned) {
    uint32_t out = 0;
    const int64_t s0 = w * 16;
    if (out_to_raw == nullptr && words_aligned && s0 + 16 <= n_out) {
        out = __ldg(reinterpret_cast<const uint32_t*>(src) + w);  // identity projection, whole word of real samples
    } else if (out_to_raw == nullptr) {
#pragma unroll
        for (int b = 0; b < 4; ++b) {
            const int64_t byte = w * 4 + b;
            const int64_t first = byte * 4;
            uint32_t v = kAllMissingByte;
            if (first < n_out) {
                v = __ldg(src + byte);
                const int64_t valid = n_out - first;  // samples of this byte that exist
                if (valid < 4) {
                    const uint32_t keep = (1u << (2 * valid)) - 1u;
                    v = (v & keep) | (kAllMissingByte & ~keep);
                }
            }
            out |= v << (8 * b);
        }
    } else {
#pragma unroll
        for (int k = 0; k < 16; ++k) {
            const int64_t s = s0 + k;
            uint32_t code = 1u;
            if (s < n_out) {
                const int32_t r = __ldg(out_to_raw + s);
                if (r >= 0) code = (__ldg(src + (r >> 2)) >> (2 * (r & 3))) & 3u;
            }
            out |= code << (2 * k);
        }
    }
    return out;
}

__global__ void __launch_bounds__(256) repack_kernel(const uint8_t* __restrict__ raw, int64_t bpv_raw,
                                                     uint8_t* __restrict__ dense, int64_t bpv_pad, int64_t n_out,
                                                     const int32_t* __restrict__ out_to_raw, int64_t snp_count) {
    const int64_t words = bpv_pad / 4;
    const bool words_aligned = ((bpv_raw & 3) == 0) && ((reinterpret_cast<uintptr_t>(raw) & 3) == 0);
    const int64_t total = snp_count * words;
    // (snp, word) of the thread's first output word by one division, then stepped: a 64-bit division per word cost more than the copy
    const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
    const int64_t step_snp = stride / words, step_w = stride - step_snp * words;
    int64_t idx = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    int64_t snp = idx / words, w = idx - snp * words;
    for (; idx < total; idx += stride) {
        reinterpret_cast<uint32_t*>(dense + snp * bpv_pad)[w] = repack_word(raw + snp * bpv_raw, w, n_out, out_to_raw, words_aligned);
        snp += step_snp;
        w += step_w;
        if (w >= words) {
            w -= words;
            ++snp;
        }
    }
}

// blocks for a chunk of snp_count rows: enough 256-thread blocks to cover every output word once, capped at one resident wave
// (148 SMs x 8 blocks; grid-stride beyond, which also spreads the two index divisions of a thread over many words)
inline unsigned repack_grid(int64_t snp_count, int64_t bpv_pad) {
    const int64_t total = snp_count * (bpv_pad / 4);
    const int64_t blocks = (total + 255) / 256;
    return static_cast<unsigned>(blocks < 1 ? 1 : (blocks > 148 * 8 ? 148 * 8 : blocks));
}

// ------------------------------------------------------------------------------------------------------------------
// Code-policy value table of one SNP: z[code] for BED code 0..3 (00=g2, 01=missing, 10=g1, 11=g0), in FP64 with the
// reference's operation order and no FMA contraction (explicit _rn intrinsics), so that `freq` is bit-identical
// (integer sum / (2*valid), one IEEE division) and every other quantity agrees to the last ulp or two.
struct SnpTable {
    double z[4];
};
// LocusStatistic (include/gelex/types/genetic_effect_type.h:20-26) minus maf: what GenotypeLoader / GenotypePipe keep per SNP.
struct SnpMoments {
    double mean, stddev;
    int monomorphic;
};

__device__ __forceinline__ void compute_snp_table(int method, int kind, int n0, int n1, int n2, int nmiss,
                                                  int all_missing_policy, double* freq_out, SnpTable* tab,
                                                  SnpMoments* mom = nullptr) {
    const double qnan = __longlong_as_double(0x7ff8000000000000ll);
    const int validi = n0 + n1 + n2;
    (void)nmiss;
    if (validi <= 0) {
        // genotype_processor.h:38-42: the reference returns early and leaves the column NaN.
        const double v = (all_missing_policy == GGRM_ALLMISS_SKIP) ? 0.0 : qnan;
        tab->z[0] = tab->z[1] = tab->z[2] = tab->z[3] = v;
        *freq_out = 0.0;  // LocusStatistic{} default
        if (mom != nullptr) *mom = SnpMoments{0.0, 0.0, 0};
        return;
    }
    const double valid = static_cast<double>(validi);
    const double sum = static_cast<double>(n1 + 2 * n2);  // exact integer
    const double maf = __ddiv_rn(sum, __dmul_rn(2.0, valid));
    *freq_out = maf;

    const bool orth = (method == GGRM_ORTH_STANDARDIZE_HWE || method == GGRM_ORTH_CENTER_HWE ||
                       method == GGRM_ORTH_STANDARDIZE || method == GGRM_ORTH_CENTER);
    const bool hwe = (method <= GGRM_ORTH_CENTER_HWE);
    const bool scale = (method == GGRM_STANDARDIZE_HWE || method == GGRM_ORTH_STANDARDIZE_HWE ||
                        method == GGRM_STANDARDIZE || method == GGRM_ORTH_STANDARDIZE);
    // encode (encode_policy.h): Add no-op; Dom raw 2->0; Dom orth 1->2p, 2->4p-2
    double e0 = 0.0, e1 = 1.0, e2 = 2.0;
    if (kind == GGRM_KIND_DOM) {
        if (orth) {
            e2 = __dsub_rn(__dmul_rn(4.0, maf), 2.0);
            e1 = __dmul_rn(2.0, maf);
        } else {
            e2 = 0.0;
        }
    }
    double mean, sd;
    const double one_m = __dsub_rn(1.0, maf);
    if (hwe && !orth) {  // HWEPolicy, statistic_policy.h:81-127
        if (kind == GGRM_KIND_ADD) {
            mean = __dmul_rn(maf, 2.0);
            sd = __dsqrt_rn(__dmul_rn(__dmul_rn(2.0, maf), one_m));
        } else {
            mean = __dmul_rn(__dmul_rn(2.0, one_m), maf);
            const double p_sq = __dmul_rn(maf, maf);
            const double q_sq = __dmul_rn(one_m, one_m);
            sd = __dsqrt_rn(__dmul_rn(__dmul_rn(__dmul_rn(2.0, maf), one_m), __dadd_rn(p_sq, q_sq)));
        }
    } else if (hwe && orth) {  // OrthHWEPolicy, statistic_policy.h:129-175
        const double dsd = __dmul_rn(__dmul_rn(2.0, maf), one_m);
        if (kind == GGRM_KIND_ADD) {
            mean = __dmul_rn(maf, 2.0);
            sd = __dsqrt_rn(dsd);
        } else {
            mean = __dmul_rn(__dmul_rn(2.0, maf), maf);
            sd = dsd;
        }
    } else {  // SamplePolicy, statistic_policy.h:43-79
        if (kind == GGRM_KIND_ADD) {
            mean = __dmul_rn(maf, 2.0);
        } else {
            const double s = __dadd_rn(__dmul_rn(static_cast<double>(n1), e1), __dmul_rn(static_cast<double>(n2), e2));
            mean = __ddiv_rn(s, valid);
        }
        const double c0 = __dsub_rn(e0, mean), c1 = __dsub_rn(e1, mean), c2 = __dsub_rn(e2, mean);
        double ss = __dmul_rn(static_cast<double>(n0), __dmul_rn(c0, c0));
        ss = __dadd_rn(ss, __dmul_rn(static_cast<double>(n1), __dmul_rn(c1, c1)));
        ss = __dadd_rn(ss, __dmul_rn(static_cast<double>(n2), __dmul_rn(c2, c2)));
        sd = __dsqrt_rn(__ddiv_rn(ss, __dsub_rn(valid, 1.0)));
    }
    double z0 = __dsub_rn(e0, mean), z1 = __dsub_rn(e1, mean), z2 = __dsub_rn(e2, mean);
    double zm = __dsub_rn(mean, mean);  // missing -> mean -> centred (0, or NaN if mean is not finite)
    const bool mono = sd < 1e-10;       // MONOMORPHIC_TOL; NaN compares false exactly like the reference
    if (mom != nullptr) *mom = SnpMoments{mean, sd, mono ? 1 : 0};
    if (scale && !mono) {
        z0 = __ddiv_rn(z0, sd);
        z1 = __ddiv_rn(z1, sd);
        z2 = __ddiv_rn(z2, sd);
        zm = __ddiv_rn(zm, sd);
    }
    tab->z[0] = z2;  // BED 00
    tab->z[1] = zm;  // BED 01
    tab->z[2] = z1;  // BED 10
    tab->z[3] = z0;  // BED 11
}

__device__ __forceinline__ uint16_t to_op16(double v, int fp16) {
    const float f = __double2float_rn(v);
    if (fp16) return __half_as_ushort(__float2half_rn(f));
    return __bfloat16_as_ushort(__float2bfloat16_rn(f));
}

// lo half of a hi + lo 16-bit operand: op16(z - hi), hi = op16(z).  NaN / inf (all-missing SNP under the reference's NaN policy,
// FP16 overflow) keep lo = 0 so that hi alone carries the non-finite value.
__device__ __forceinline__ uint16_t op16_residual(double v, int fp16) {
    const uint16_t hb = to_op16(v, fp16);
    const float hi = fp16 ? __half2float(__ushort_as_half(hb)) : __bfloat162float(__ushort_as_bfloat16(hb));
    const double r = v - static_cast<double>(hi);
    if (!(fabs(r) <= 3.0e38)) return 0;
    return to_op16(r, fp16);
}

// ------------------------------------------------------------------------------------------------------------------
// K1b+K2: counts + stats.  One warp per SNP: lanes stride over the dense row in 16-byte words (64 samples each),
// classify the four 2-bit codes with bit tricks + popc, butterfly-reduce, lane 0 derives frequency and tables.
//   counts[snp] = {n0, n1, n2, nmiss}     freq[snp]     tab16[kind][snp] = 4 x 16-bit     tab64[kind][snp] = 4 x f64
struct StatsOut {
    int4* counts;       // may be nullptr
    double* freq;       // may be nullptr
    uint2* tab16[2];    // [kind]; nullptr when that kind / precision is not wanted
    uint2* tab16_lo[2]; // [kind]; hi + lo operands: the table of lo = op16(z - hi) (nullptr: single 16-bit operand)
    unsigned long long* range_flag;  // FP16 operands: receives 1 + SNP index when a value that occurs in the data overflows binary16
    SnpTable* tab64[2];
    SnpMoments* moments[2];  // [kind]; only the genotype-export hook asks for them (with tab64 of the same kind)
    int second_nonmissing;   // kind 1 = the 0/1 non-missing indicator instead of the dominance policy (GGRM_SECOND_NONMISSING)
    double4* gcta_w;         // [snp] GCTA diagonal terms by BED code {x = 2, missing, x = 1, x = 0}; nullptr unless GGRM_SECOND_NONMISSING_GCTA_DIAG
};

// Genotype class counts of one dense row over the lanes of a warp (every lane returns the totals).
__device__ __forceinline__ void count_row_classes(const uint4* row, int64_t nvec, int lane, int* miss, int* het, int* hom0) {
    int c_miss = 0, c_het = 0, c_hom0 = 0;
    for (int64_t v = lane; v < nvec; v += 32) {
        const uint4 w = row[v];
        const uint32_t ws[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const uint32_t lo = ws[k] & 0x55555555u;
            const uint32_t hi = (ws[k] >> 1) & 0x55555555u;
            c_miss += __popc(lo & ~hi);  // 01
            c_het += __popc(hi & ~lo);   // 10
            c_hom0 += __popc(hi & lo);   // 11
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        c_miss += __shfl_xor_sync(0xffffffffu, c_miss, o);
        c_het += __shfl_xor_sync(0xffffffffu, c_het, o);
        c_hom0 += __shfl_xor_sync(0xffffffffu, c_hom0, o);
    }
    *miss = c_miss;
    *het = c_het;
    *hom0 = c_hom0;
}

// Everything one SNP leaves behind (counts, frequency, value tables, GCTA terms), from its class counts.  One thread.
__device__ __forceinline__ void emit_snp_outputs(int64_t snp, int c_miss, int c_het, int c_hom0, int64_t n_out, int64_t n_pad, int method,
                                                 int all_missing_policy, int fp16, const StatsOut& out, int64_t out_offset,
                                                 uint2* tabs = nullptr) {  // tabs[kind * 2 + (lo ? 1 : 0)]: the 16-bit tables as written (zeros where none)
    if (tabs != nullptr) tabs[0] = tabs[1] = tabs[2] = tabs[3] = make_uint2(0u, 0u);
    const int total = static_cast<int>(n_pad);
    const int n2 = total - c_miss - c_het - c_hom0;            // 00
    const int nmiss = c_miss - static_cast<int>(n_pad - n_out);  // padding rows were encoded as missing
    const int n1 = c_het, n0 = c_hom0;
    const int64_t o = out_offset + snp;
    if (out.counts != nullptr) out.counts[o] = make_int4(n0, n1, n2, nmiss);
    double freq = 0.0;
#pragma unroll
    for (int kind = 0; kind < 2; ++kind) {
        if (out.tab16[kind] == nullptr && out.tab64[kind] == nullptr) continue;
        SnpTable t;
        SnpMoments mo;
        if (kind == GGRM_KIND_DOM && out.second_nonmissing) {
            compute_snp_table(method, GGRM_KIND_ADD, n0, n1, n2, nmiss, all_missing_policy, &freq, &t, &mo);  // frequency only
            t.z[0] = t.z[2] = t.z[3] = 1.0;  // called genotype (BED 00 / 10 / 11)
            t.z[1] = 0.0;                    // BED 01 = missing (padding rows are encoded as missing too)
        } else {
            compute_snp_table(method, kind, n0, n1, n2, nmiss, all_missing_policy, &freq, &t, &mo);
        }
        if (out.moments[kind] != nullptr) out.moments[kind][snp] = mo;
        if (out.tab64[kind] != nullptr) out.tab64[kind][snp] = t;
        if (out.tab16[kind] != nullptr) {
            const uint32_t a = to_op16(t.z[0], fp16) | (static_cast<uint32_t>(to_op16(t.z[1], fp16)) << 16);
            const uint32_t b = to_op16(t.z[2], fp16) | (static_cast<uint32_t>(to_op16(t.z[3], fp16)) << 16);
            out.tab16[kind][snp] = make_uint2(a, b);
            if (tabs != nullptr) tabs[kind * 2] = make_uint2(a, b);
            if (out.tab16_lo[kind] != nullptr) {  // lo = op16(z - float(hi)); exact subtraction in FP64, one rounding
                const uint32_t la = op16_residual(t.z[0], fp16) | (static_cast<uint32_t>(op16_residual(t.z[1], fp16)) << 16);
                const uint32_t lb = op16_residual(t.z[2], fp16) | (static_cast<uint32_t>(op16_residual(t.z[3], fp16)) << 16);
                out.tab16_lo[kind][snp] = make_uint2(la, lb);
                if (tabs != nullptr) tabs[kind * 2 + 1] = make_uint2(la, lb);
            }
            if (fp16 && out.range_flag != nullptr) {
                // a value counts only if some sample carries that genotype (z[1], the imputed missing call, is 0)
                const bool over = (n2 > 0 && fabs(t.z[0]) > 65504.0) || (n1 > 0 && fabs(t.z[2]) > 65504.0) || (n0 > 0 && fabs(t.z[3]) > 65504.0);
                if (over) atomicMax(out.range_flag, static_cast<unsigned long long>(o) + 1ull);
            }
        }
    }
    if (out.tab16[0] == nullptr && out.tab64[0] == nullptr && out.tab16[1] == nullptr && out.tab64[1] == nullptr) {
        SnpTable t;
        compute_snp_table(method, GGRM_KIND_ADD, n0, n1, n2, nmiss, all_missing_policy, &freq, &t);
    }
    if (out.freq != nullptr) out.freq[o] = freq;
    if (out.gcta_w != nullptr) {
        // GCTA's default diagonal term of one SNP: (x^2 - (1 + 2p) x + 2p^2) / (2p(1 - p)), x = A1 count, p = A1 frequency;
        // 0 for a missing call and for a SNP without variation (which contributes 0 to the off-diagonal sums as well)
        const int validi = n0 + n1 + n2;
        const double p = validi > 0 ? static_cast<double>(n1 + 2 * n2) / (2.0 * static_cast<double>(validi)) : 0.0;
        const double den = 2.0 * p * (1.0 - p);
        double4 w = make_double4(0.0, 0.0, 0.0, 0.0);
        if (den > 0.0) {
            w.x = (4.0 - 2.0 * (1.0 + 2.0 * p) + 2.0 * p * p) / den;  // BED 00: x = 2
            w.z = (1.0 - (1.0 + 2.0 * p) + 2.0 * p * p) / den;        // BED 10: x = 1
            w.w = (2.0 * p * p) / den;                                // BED 11: x = 0
        }
        out.gcta_w[snp] = w;
    }
}

__global__ void __launch_bounds__(256) count_stats_kernel(const uint8_t* __restrict__ dense, int64_t bpv_pad,
                                                          int64_t n_out, int64_t n_pad, int64_t snp_count, int method,
                                                          int all_missing_policy, int fp16, StatsOut out,
                                                          int64_t out_offset) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int64_t snp = static_cast<int64_t>(blockIdx.x) * (blockDim.x >> 5) + warp;
    if (snp >= snp_count) return;
    int c_miss, c_het, c_hom0;
    count_row_classes(reinterpret_cast<const uint4*>(dense + snp * bpv_pad), bpv_pad / 16, lane, &c_miss, &c_het, &c_hom0);
    if (lane == 0) emit_snp_outputs(snp, c_miss, c_het, c_hom0, n_out, n_pad, method, all_missing_policy, fp16, out, out_offset);
}

// GCTA diagonal sums.  lanes[L][sample] = sum over the SNPs pushed so far whose global index is L mod 16 of
// gcta_w[snp][code(sample, snp)], each lane added one SNP after the other in SNP order, and diag[sample] = lanes[0] + ... +
// lanes[15] in that order: the result does not depend on how the SNPs were cut into chunks or pushes.  One block per 128
// samples (32 bytes = one sector of every dense row): thread (x, y) owns half-word x of those bytes (8 samples, 8 FP64
// accumulators in registers) for the SNPs u = y, y + 16, ... of the chunk.  The chunk is walked 256 SNPs at a time: their tables
// go to shared memory (a term is then ONE indexed 8-byte load - entry [1], "missing", is 0), and a thread issues its 16 row
// loads of the batch before it uses the first.  Reads the dense chunk once more (n/4 bytes per SNP) and the lanes twice.
constexpr int kGctaLanes = 16;
constexpr int kGctaSamplesPerBlock = 128;
__global__ void __launch_bounds__(256) gcta_diag_kernel(const uint8_t* __restrict__ dense, int64_t bpv_pad, const double4* __restrict__ gcta_w,
                                                        int64_t snp_count, int64_t first_snp, int64_t n_pad, double* __restrict__ lanes,
                                                        double* __restrict__ diag) {
    __shared__ __align__(32) double s_w[256 * 4];
    const int x = threadIdx.x & 15, y = threadIdx.x >> 4;
    const int64_t b0 = static_cast<int64_t>(blockIdx.x) * (kGctaSamplesPerBlock / 4);  // byte offset of the block's samples inside a dense row
    const int64_t sample0 = b0 * 4 + x * 8;                                          // first of this thread's 8 samples
    const int lane = static_cast<int>((first_snp + y) & (kGctaLanes - 1));
    double* mine = lanes + static_cast<int64_t>(lane) * n_pad + sample0;
    double acc[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) acc[k] = mine[k];
    const uint16_t* col = reinterpret_cast<const uint16_t*>(dense + b0) + x;
    const int64_t pitch = bpv_pad / 2;
    for (int64_t s0 = 0; s0 < snp_count; s0 += 256) {
        const int cnt = static_cast<int>(snp_count - s0 < 256 ? snp_count - s0 : 256);
        __syncthreads();
        {
            double4 w = make_double4(0.0, 0.0, 0.0, 0.0);
            if (static_cast<int>(threadIdx.x) < cnt) {
                const double2* src = reinterpret_cast<const double2*>(gcta_w + s0 + threadIdx.x);
                const double2 w01 = __ldg(src), w23 = __ldg(src + 1);
                w = make_double4(w01.x, 0.0, w23.x, w23.y);
            }
            reinterpret_cast<double4*>(s_w)[threadIdx.x] = w;
        }
        uint32_t word[16];
#pragma unroll
        for (int i = 0; i < 16; ++i) {  // SNPs past the end of the chunk read as "missing everywhere"
            const int ul = y + 16 * i;
            word[i] = ul < cnt ? static_cast<uint32_t>(__ldg(col + (s0 + ul) * pitch)) : 0x5555u;
        }
        __syncthreads();
#pragma unroll
        for (int i = 0; i < 16; ++i) {
            const double* t = s_w + (y + 16 * i) * 4;
#pragma unroll
            for (int k = 0; k < 8; ++k) acc[k] += t[(word[i] >> (2 * k)) & 3u];
        }
    }
#pragma unroll
    for (int k = 0; k < 8; ++k) mine[k] = acc[k];
    __syncthreads();  // the block's lane sums are in global memory and visible to the block
    if (threadIdx.x < kGctaSamplesPerBlock) {
        const int64_t sample = b0 * 4 + threadIdx.x;
        double total = 0.0;
#pragma unroll
        for (int l = 0; l < kGctaLanes; ++l) total += lanes[static_cast<int64_t>(l) * n_pad + sample];
        diag[sample] = total;
    }
}

// ------------------------------------------------------------------------------------------------------------------
// K3: expand to 16-bit operands.  dense [c][bpv_pad] + tab16[c] -> Z [n_pad][k_pad] (K-major: one GRM row = one
// sample, SNPs contiguous), the layout the TMA/tcgen05 SYRK consumes.  Tile = 64 SNPs x 256 samples:
//   load : 64 rows x 64 bytes, one 16-byte load per thread, transposed into smem as [byte][snp]
//   emit : thread (sample quad, octet) reads 8 contiguous smem bytes (8 SNPs of four samples), looks the 32 values up in
//          the per-SNP tables with byte permutes and writes four 16-byte stores; 8 lanes cover one sample's 128 contiguous bytes.
// SNPs in [snp_count, k_pad) are written as 0 so the SYRK can run whole 64-wide K blocks.
constexpr int kExpandPitch = 64 + 4;  // row pitch of the transposed tile [byte][snp]: 17 words, so that the word stores of stage and the reads of emit spread over the banks

// load half of a tile.  Thread (word column wc = tid & 15, SNP quad sq = tid >> 4) holds word wc (4 bytes = 16 samples) of the SNP rows
// 4 sq .. 4 sq + 3, transposes the 4 x 4 bytes in registers and stores one word per byte column: s_t[4 wc + b][4 sq .. 4 sq + 3].
__device__ __forceinline__ void expand16_stage(const uint32_t w0, const uint32_t w1, const uint32_t w2, const uint32_t w3, int tid,
                                               uint8_t (*s_t)[kExpandPitch]) {
    const int wc = tid & 15, sq = tid >> 4;
    const uint32_t t0 = __byte_perm(w0, w1, 0x5140u), t1 = __byte_perm(w2, w3, 0x5140u);  // {a.b0, b.b0, a.b1, b.b1}
    const uint32_t t2 = __byte_perm(w0, w1, 0x7362u), t3 = __byte_perm(w2, w3, 0x7362u);  // {a.b2, b.b2, a.b3, b.b3}
    *reinterpret_cast<uint32_t*>(&s_t[4 * wc + 0][4 * sq]) = __byte_perm(t0, t1, 0x5410u);
    *reinterpret_cast<uint32_t*>(&s_t[4 * wc + 1][4 * sq]) = __byte_perm(t0, t1, 0x7632u);
    *reinterpret_cast<uint32_t*>(&s_t[4 * wc + 2][4 * sq]) = __byte_perm(t2, t3, 0x5410u);
    *reinterpret_cast<uint32_t*>(&s_t[4 * wc + 3][4 * sq]) = __byte_perm(t2, t3, 0x7632u);
}

// emit half: the staged tile (64 SNPs x 256 samples) through the 64 tables of s_tab into Z rows 4 b0 .. 4 b0 + 255, columns k0 .. k0 + 63
__device__ __forceinline__ void expand16_emit(const uint8_t (*s_t)[kExpandPitch], const uint2* s_tab, uint16_t* __restrict__ Z, int64_t k_pad,
                                              int64_t k0, int64_t b0, int tid) {
    // A SNP's table is 8 bytes {z(00), z(01) | z(10), z(11)}: one PRMT with selector nibbles (2c, 2c+1, 2c', 2c'+1) looks up the
    // values of two samples' codes c, c' at once; a second PRMT pairs the values of two neighbouring SNPs of one sample.
    const int octet = tid & 7;
    uint2 tab[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) tab[u] = s_tab[octet * 8 + u];
#pragma unroll
    for (int it = 0; it < 2; ++it) {
        const int quad = it * 32 + (tid >> 3);  // byte column of the tile = samples 4*quad .. 4*quad + 3
        const uint2 bytes = make_uint2(*reinterpret_cast<const uint32_t*>(&s_t[quad][octet * 8]),       // that quad's byte at 8 consecutive SNPs
                                       *reinterpret_cast<const uint32_t*>(&s_t[quad][octet * 8 + 4]));
        uint32_t x01[8], x23[8];  // per SNP: {value of sample 0, sample 1}, {sample 2, sample 3}
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const uint32_t w = h == 0 ? bytes.x : bytes.y;
            const uint32_t c0 = w & 0x03030303u, c1 = (w >> 2) & 0x03030303u, c2 = (w >> 4) & 0x03030303u, c3 = (w >> 6) & 0x03030303u;
#pragma unroll
            for (int v = 0; v < 4; ++v) {
                const uint32_t gather = static_cast<uint32_t>(v | ((4 + v) << 4));  // low 16 bits <- {byte v of a, byte v of b}
                const uint32_t sel01 = __byte_perm(c0, c1, gather) * 0x22u + 0x1010u;  // PRMT reads the low 16 bits only
                const uint32_t sel23 = __byte_perm(c2, c3, gather) * 0x22u + 0x1010u;
                x01[h * 4 + v] = __byte_perm(tab[h * 4 + v].x, tab[h * 4 + v].y, sel01);
                x23[h * 4 + v] = __byte_perm(tab[h * 4 + v].x, tab[h * 4 + v].y, sel23);
            }
        }
        uint32_t o[4][4];
#pragma unroll
        for (int p = 0; p < 4; ++p) {
            o[0][p] = __byte_perm(x01[2 * p], x01[2 * p + 1], 0x5410u);
            o[1][p] = __byte_perm(x01[2 * p], x01[2 * p + 1], 0x7632u);
            o[2][p] = __byte_perm(x23[2 * p], x23[2 * p + 1], 0x5410u);
            o[3][p] = __byte_perm(x23[2 * p], x23[2 * p + 1], 0x7632u);
        }
        const int64_t row = (b0 + quad) * 4;
#pragma unroll
        for (int j = 0; j < 4; ++j)
            *reinterpret_cast<uint4*>(Z + (row + j) * k_pad + k0 + octet * 8) = make_uint4(o[j][0], o[j][1], o[j][2], o[j][3]);
    }
}

__global__ void __launch_bounds__(256) expand16_kernel(const uint8_t* __restrict__ dense, int64_t bpv_pad,
                                                       const uint2* __restrict__ tab16, int64_t snp_count,
                                                       uint16_t* __restrict__ Z, int64_t k_pad) {
    __shared__ __align__(16) uint8_t s_t[64][kExpandPitch];  // [byte][snp]
    __shared__ uint2 s_tab[64];
    const int64_t k0 = static_cast<int64_t>(blockIdx.x) * 64;     // first SNP of the tile
    const int64_t b0 = static_cast<int64_t>(blockIdx.y) * 64;     // first byte (sample/4) of the tile
    const int tid = threadIdx.x;
    {
        uint32_t w[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {  // SNP rows past the chunk read as missing everywhere (and their tables are zero)
            const int64_t snp = k0 + 4 * (tid >> 4) + i;
            w[i] = snp < snp_count ? __ldg(reinterpret_cast<const uint32_t*>(dense + snp * bpv_pad + b0) + (tid & 15)) : 0x55555555u;
        }
        expand16_stage(w[0], w[1], w[2], w[3], tid, s_t);
        if (tid < 64) s_tab[tid] = (k0 + tid < snp_count) ? __ldg(tab16 + k0 + tid) : make_uint2(0u, 0u);
    }
    __syncthreads();
    expand16_emit(s_t, s_tab, Z, k_pad, k0, b0, tid);
}

// ------------------------------------------------------------------------------------------------------------------
// K1-K3 in ONE launch for problems that are a single operand slab of few samples (n_pad <= 4096, every SNP of the pass in one
// chunk; C1 of BASELINE.json): nothing overlaps the prep of the only slab, so its three kernels and the gaps between them are
// the pass's serial front.  One block takes 64 SNPs through all three steps with the repacked rows held in shared memory: repack
// (also written to `dense`, which other kernels read), class counts with a warp per row and the per-SNP outputs on 8 lanes side by
// side, then the 16-bit expansion of every 256-sample tile - the transposed tile is staged once and emitted through each
// wanted table (kind, hi / lo).  Same device functions as the separate kernels on the same bytes: bit-identical outputs.
constexpr int kFusedRowPad = 16;  // shared-memory row pitch = bpv_pad + 16: rows stay 16-byte aligned for the counting loads, and rows 4 apart (the two
                                  // SNP quads a warp stages) start 16 banks apart
inline size_t prep_fused_smem_bytes(int64_t bpv_pad) { return static_cast<size_t>(64 * (bpv_pad + kFusedRowPad)); }
__global__ void __launch_bounds__(256, 4) prep_fused_kernel(const uint8_t* __restrict__ raw, int64_t bpv_raw, uint8_t* __restrict__ dense, int64_t bpv_pad,
                                                         int64_t n_out, int64_t n_pad, const int32_t* __restrict__ out_to_raw, int64_t snp_count, int method,
                                                         int all_missing_policy, int fp16, StatsOut out, int64_t out_offset, uint16_t* __restrict__ Z0,
                                                         uint16_t* __restrict__ Z1, int64_t z_pitch, int64_t lo_offset) {
    extern __shared__ __align__(16) uint8_t s_d[];  // [64][bpv_pad + kFusedRowPad]
    __shared__ __align__(16) uint8_t s_t[2][64][kExpandPitch];  // double buffered: one barrier per tile
    __shared__ uint2 s_tab[4][64];                  // [kind * 2 + (lo ? 1 : 0)]
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int64_t k0 = static_cast<int64_t>(blockIdx.x) * 64;
    const int64_t pitch = bpv_pad + kFusedRowPad;
    const int64_t words = bpv_pad / 4;
    {   // K1a: warp w repacks rows 8 w .. 8 w + 7, a lane every 32nd word of a row
        const bool words_aligned = ((bpv_raw & 3) == 0) && ((reinterpret_cast<uintptr_t>(raw) & 3) == 0);
        const int nwords = static_cast<int>(words);
        // identity projection: the words made of real samples only are plain copies - independent loads, unrolled so that a lane
        // has up to 32 of them in flight; everything else (last word of a row, padding, projections) goes through repack_word
        const int ncopy = out_to_raw == nullptr && words_aligned ? static_cast<int>(n_out / 16 < words ? n_out / 16 : words) : 0;
        for (int w0 = 0; w0 < ncopy; w0 += 128) {  // 128 words of each of the warp's 8 rows per round: 32 loads issued, then 64 stores
            uint32_t v[8][4];
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const int64_t snp = k0 + warp * 8 + j;
                const uint32_t* src = reinterpret_cast<const uint32_t*>(raw + snp * bpv_raw);
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const int w = w0 + lane + 32 * i;
                    v[j][i] = snp < snp_count && w < ncopy ? __ldg(src + w) : 0u;
                }
            }
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const int r = warp * 8 + j;
                const int64_t snp = k0 + r;
                uint32_t* dst_s = reinterpret_cast<uint32_t*>(s_d + r * pitch);
                uint32_t* dst_g = reinterpret_cast<uint32_t*>(dense + snp * bpv_pad);
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const int w = w0 + lane + 32 * i;
                    if (snp < snp_count && w < ncopy) {
                        dst_s[w] = v[j][i];
                        dst_g[w] = v[j][i];
                    }
                }
            }
        }
#pragma unroll 1
        for (int j = 0; j < 8; ++j) {
            const int r = warp * 8 + j;
            const int64_t snp = k0 + r;
            uint32_t* dst_s = reinterpret_cast<uint32_t*>(s_d + r * pitch);
            if (snp < snp_count) {
                uint32_t* dst_g = reinterpret_cast<uint32_t*>(dense + snp * bpv_pad);
#pragma unroll 1
                for (int w = ncopy + lane; w < nwords; w += 32) {
                    const uint32_t v = repack_word(raw + snp * bpv_raw, w, n_out, out_to_raw, words_aligned);
                    dst_s[w] = v;
                    dst_g[w] = v;
                }
            } else {  // rows past the chunk: missing everywhere (their tables are zero as well)
                for (int w = lane; w < nwords; w += 32) dst_s[w] = 0x55555555u;
            }
        }
    }
    __syncthreads();
    {   // K1b + K2: warp w counts rows 8 w .. 8 w + 7, then lane j derives everything for row 8 w + j
        int my_miss = 0, my_het = 0, my_hom0 = 0;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            int a, b, c;
            count_row_classes(reinterpret_cast<const uint4*>(s_d + (warp * 8 + j) * pitch), bpv_pad / 16, lane, &a, &b, &c);
            if (lane == j) {
                my_miss = a;
                my_het = b;
                my_hom0 = c;
            }
        }
        const int64_t snp = k0 + warp * 8 + lane;
        if (lane < 8) {  // the tables go to the expansion below through shared memory (and to global memory for everybody else)
            uint2 tabs[4] = {make_uint2(0u, 0u), make_uint2(0u, 0u), make_uint2(0u, 0u), make_uint2(0u, 0u)};
            if (snp < snp_count) emit_snp_outputs(snp, my_miss, my_het, my_hom0, n_out, n_pad, method, all_missing_policy, fp16, out, out_offset, tabs);
#pragma unroll
            for (int t = 0; t < 4; ++t) s_tab[t][warp * 8 + lane] = tabs[t];
        }
    }
    // K3
    const uint8_t* my_rows = s_d + 4 * (tid >> 4) * pitch + 4 * (tid & 15);  // word (tid & 15) of rows 4 (tid >> 4) ..
    int buf = 0;
    for (int64_t b0 = 0; b0 < bpv_pad; b0 += 64, buf ^= 1) {
        expand16_stage(*reinterpret_cast<const uint32_t*>(my_rows + b0), *reinterpret_cast<const uint32_t*>(my_rows + pitch + b0),
                       *reinterpret_cast<const uint32_t*>(my_rows + 2 * pitch + b0), *reinterpret_cast<const uint32_t*>(my_rows + 3 * pitch + b0), tid, s_t[buf]);
        __syncthreads();  // tile staged (first pass: s_tab complete); the other buffer is free again once every warp has passed this barrier
#pragma unroll 1
        for (int t = 0; t < 4; ++t) {  // (kind, hi / lo): one copy of the emit code, the staged tile is reused
            uint16_t* base = t < 2 ? Z0 : Z1;
            if (base == nullptr || ((t & 1) != 0 && lo_offset == 0)) continue;
            expand16_emit(s_t[buf], s_tab[t], base + ((t & 1) != 0 ? lo_offset : 0), z_pitch, k0, b0, tid);
        }
    }
}

// FP64 twin for the DMMA path: Zd [n_pad][k_pad] doubles.  Same tile; a thread writes 8 doubles (64 B).
__global__ void __launch_bounds__(256) expand64_kernel(const uint8_t* __restrict__ dense, int64_t bpv_pad,
                                                       const SnpTable* __restrict__ tab64, int64_t snp_count,
                                                       double* __restrict__ Z, int64_t k_pad) {
    __shared__ __align__(16) uint8_t s_t[64][64 + 16];
    __shared__ SnpTable s_tab[64];
    const int64_t k0 = static_cast<int64_t>(blockIdx.x) * 64;
    const int64_t b0 = static_cast<int64_t>(blockIdx.y) * 64;
    const int tid = threadIdx.x;
    {
        const int s = tid >> 2, part = tid & 3;
        const int64_t snp = k0 + s;
        uint4 w = make_uint4(0x55555555u, 0x55555555u, 0x55555555u, 0x55555555u);
        if (snp < snp_count) w = __ldg(reinterpret_cast<const uint4*>(dense + snp * bpv_pad + b0 + part * 16));
        const uint32_t ws[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
        for (int q = 0; q < 16; ++q) s_t[part * 16 + q][s] = static_cast<uint8_t>(ws[q >> 2] >> (8 * (q & 3)));
        if (tid < 64) {
            SnpTable t;
            t.z[0] = t.z[1] = t.z[2] = t.z[3] = 0.0;
            if (k0 + tid < snp_count) t = tab64[k0 + tid];
            s_tab[tid] = t;
        }
    }
    __syncthreads();
    const int octet = tid & 7;
#pragma unroll
    for (int it = 0; it < 8; ++it) {
        const int sample = it * 32 + (tid >> 3);
        const uint2 bytes = *reinterpret_cast<const uint2*>(&s_t[sample >> 2][octet * 8]);
        const int shift = 2 * (sample & 3);
        const int64_t row = b0 * 4 + sample;
        double* dst = Z + row * k_pad + k0 + octet * 8;
#pragma unroll
        for (int u = 0; u < 8; u += 2) {
            const uint32_t c0 = ((((u < 4 ? bytes.x : bytes.y) >> (8 * (u & 3))) & 0xFFu) >> shift) & 3u;
            const uint32_t c1 = ((((u + 1 < 4 ? bytes.x : bytes.y) >> (8 * ((u + 1) & 3))) & 0xFFu) >> shift) & 3u;
            *reinterpret_cast<double2*>(dst + u) = make_double2(s_tab[octet * 8 + u].z[c0], s_tab[octet * 8 + u + 1].z[c1]);
        }
    }
}

// ------------------------------------------------------------------------------------------------------------------
// Test hook kernels (bit-exact layer): decode to the reference's FP64 column-major chunk, optionally standardised.
// out[snp * n_out + row]; mode 0 = raw decode (2/1/0/NaN), mode 1 = table lookup (missing -> z[1]).
__global__ void __launch_bounds__(256) decode_colmajor_kernel(const uint8_t* __restrict__ dense, int64_t bpv_pad,
                                                              int64_t n_out, int64_t snp_count,
                                                              const SnpTable* __restrict__ tab64,
                                                              double* __restrict__ out) {
    const int64_t snp = blockIdx.y;
    const int64_t row = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (snp >= snp_count || row >= n_out) return;
    const uint32_t code = (__ldg(dense + snp * bpv_pad + (row >> 2)) >> (2 * (row & 3))) & 3u;
    double v;
    if (tab64 == nullptr) {
        const double qnan = __longlong_as_double(0x7ff8000000000000ll);
        v = code == 0u ? 2.0 : (code == 1u ? qnan : (code == 2u ? 1.0 : 0.0));
    } else {
        v = tab64[snp].z[code];
    }
    out[snp * n_out + row] = v;
}

// Synthetic payload generator (see synth.h): one thread per output byte.
}  // namespace ggrm

#include "synth.h"

namespace ggrm {
__global__ void __launch_bounds__(256) synth_bed_kernel(uint8_t* __restrict__ out, int64_t n, int64_t bpv,
                                                        int64_t snp_begin, int64_t snp_count, uint64_t seed,
                                                        uint32_t miss_thr24) {
    const int64_t snp_local = blockIdx.x;  // one block per SNP row (grid.x may be up to 2^31-1)
    if (snp_local >= snp_count) return;
    const int64_t snp = snp_begin + snp_local;
    const ggrm_snp_thresholds t = ggrm_synth_thresholds(seed, snp);
    for (int64_t b = threadIdx.x; b < bpv; b += blockDim.x) {
        out[snp_local * bpv + b] = ggrm_synth_byte(seed, snp, b, n, t, miss_thr24);
    }
}
}  // namespace ggrm
