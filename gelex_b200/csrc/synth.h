// Synthetic PLINK BED payload generator shared by host and device code.
// Model = tests/bed_fixture.cpp:138-177 of the reference (per-SNP allele frequency, HWE genotypes, i.i.d. missing calls),
// but with a counter-based integer RNG so any box regenerates identical bytes, in parallel, on CPU or GPU.
//   p_j   ~ U(0.05, 0.5)   as 32-bit fixed point  P = 0.05*2^32 + h32 * 0.45
//   g_ij  : u32 < p^2 -> A1A1 (BED 00, value 2) ; u32 < p^2+2pq -> het (BED 10) ; else A2A2 (BED 11)
//   miss  : low 24 bits of the hash < missing_rate * 2^24 -> BED 01
// Sample i occupies bits 2*(i%4) of byte i/4 (tests/bed_fixture.cpp:266-282); padding bits are 0.
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define GGRM_HD __host__ __device__ __forceinline__
#else
#define GGRM_HD static inline
#endif

GGRM_HD uint64_t ggrm_splitmix64(uint64_t x) {
    x += 0x9E3779B97F4A7C15ull;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
    return x ^ (x >> 31);
}

struct ggrm_snp_thresholds {
    uint32_t t_hom2;  // u < t_hom2          -> genotype 2 (A1A1)
    uint32_t t_het;   // u < t_het (>= hom2) -> genotype 1
};

GGRM_HD ggrm_snp_thresholds ggrm_synth_thresholds(uint64_t seed, int64_t snp) {
    uint64_t h = ggrm_splitmix64(seed ^ (0xD1B54A32D192ED03ull * (uint64_t)(snp + 1)));
    uint64_t h32 = h >> 32;
    uint64_t P = 214748365ull + ((h32 * 1932735283ull) >> 32);  // p in [0.05, 0.5) as Q32
    uint64_t Q = 4294967296ull - P;
    uint64_t t1 = (P * P) >> 32;
    uint64_t t2 = t1 + ((2ull * P * Q) >> 32);
    ggrm_snp_thresholds t;
    t.t_hom2 = (uint32_t)t1;
    t.t_het = (uint32_t)(t2 > 0xFFFFFFFFull ? 0xFFFFFFFFull : t2);
    return t;
}

// BED 2-bit code of sample i at SNP snp.
GGRM_HD uint32_t ggrm_synth_code(uint64_t seed, int64_t snp, int64_t i, ggrm_snp_thresholds t, uint32_t miss_thr24) {
    uint64_t h = ggrm_splitmix64((seed + 0x632BE59BD9B4E019ull * (uint64_t)(snp + 1)) ^ (0x9E3779B97F4A7C15ull * (uint64_t)(i + 1)));
    uint32_t u = (uint32_t)(h >> 32);
    uint32_t m = (uint32_t)(h & 0xFFFFFFu);
    if (m < miss_thr24) return 1u;      // 01 missing
    if (u < t.t_hom2) return 0u;        // 00 -> 2
    if (u < t.t_het) return 2u;         // 10 -> 1
    return 3u;                          // 11 -> 0
}

GGRM_HD uint8_t ggrm_synth_byte(uint64_t seed, int64_t snp, int64_t byte_idx, int64_t n, ggrm_snp_thresholds t,
                                uint32_t miss_thr24) {
    uint32_t b = 0;
    for (int k = 0; k < 4; ++k) {
        int64_t i = byte_idx * 4 + k;
        if (i < n) b |= ggrm_synth_code(seed, snp, i, t, miss_thr24) << (2 * k);
    }
    return (uint8_t)b;
}
