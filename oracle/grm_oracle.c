/* TEST INFRASTRUCTURE ONLY — CPU restatement ("oracle") of the reference's `gelex grm` arithmetic in plain C.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may load this library; the product
 * (gelex_b200/, include/ggrm.h) never links, loads or executes it and has no CPU fallback.
 *
 * Parity is PINNED: tests/test_oracle_golden.py checks every function here against
 *   (a) the known-answer vectors of the reference's own tests (tests/test_genotype_processor.cpp:71-270,
 *       tests/test_grm_code_policy.cpp:30-450, tests/test_grm.cpp:271-394, tests/test_grm_bin_writer.cpp:142-314), and
 *   (b) outputs of the unmodified reference compiled from /root/reference (oracle/_ref/gelex_ref), committed as
 *       fixtures under tests/golden/ together with the generating script (tests/golden/make_golden.py).
 *
 * Every function cites the reference file:line it restates (paths relative to /root/reference).
 * Compiled with -ffp-contract=off so FP64 expressions round exactly like the reference's non-FMA x86-64 build.
 */
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#define ORACLE_API __attribute__((visibility("default")))

/* src/data/decode_lut.h:40-41 — BED code -> genotype value (A1 allele count); 01 is missing (NaN). */
static double decode_code(unsigned code) {
    switch (code & 3u) {
        case 0u: return 2.0;
        case 1u: return NAN;
        case 2u: return 1.0;
        default: return 0.0;
    }
}

/* src/data/genotype/bed_pipe.cpp:121-158 + src/data/bed_pipe/variant_decoder.cpp:51-126.
 * out is column-major n_out x snp_count.  raw_to_out == NULL is the dense path (identity, n_out == n_raw);
 * otherwise the sparse path: pre-fill NaN (bed_pipe.cpp:133-136), scatter raw sample i to row raw_to_out[i]
 * unless it is -1; when several raw samples map to one row the later one overwrites the earlier. */
ORACLE_API void oracle_decode_chunk(const uint8_t* payload, int64_t snp_count, int64_t bpv, int64_t n_raw, int64_t n_out,
                                    const int32_t* raw_to_out, double* out) {
    for (int64_t j = 0; j < snp_count; ++j) {
        const uint8_t* src = payload + j * bpv;
        double* col = out + j * n_out;
        if (raw_to_out == NULL) {
            for (int64_t i = 0; i < n_raw && i < n_out; ++i) col[i] = decode_code(src[i >> 2] >> (2 * (i & 3)));
        } else {
            for (int64_t i = 0; i < n_out; ++i) col[i] = NAN;
            for (int64_t i = 0; i < n_raw; ++i) {
                const int32_t t = raw_to_out[i];
                if (t != -1) col[t] = decode_code(src[i >> 2] >> (2 * (i & 3)));
            }
        }
    }
}

typedef struct oracle_locus_stat {
    double maf, mean, stddev;
    int32_t is_monomorphic;
    int32_t valid;
} oracle_locus_stat;

/* include/gelex/internal/genotype_processor/genotype_processor.h:35-58 with
 * encode_policy.h:32-81 and statistic_policy.h:43-175.  method = GenotypeProcessMethod 1..8
 * (include/gelex/data/genotype/genotype_processor.h:38-48), kind 0 = Add, 1 = Dom.  In place. */
ORACLE_API void oracle_process_column(double* locus, int64_t n, int method, int kind, oracle_locus_stat* st) {
    memset(st, 0, sizeof(*st));
    int64_t n_nan = 0;
    for (int64_t i = 0; i < n; ++i) n_nan += isnan(locus[i]) ? 1 : 0;
    const double valid = (double)(n - n_nan);
    st->valid = (int32_t)(n - n_nan);
    if (valid <= 0) return; /* :38-42 column stays NaN */
    double sum = 0.0;
    for (int64_t i = 0; i < n; ++i) sum += isnan(locus[i]) ? 0.0 : locus[i];
    st->maf = sum / (2.0 * valid); /* :44 */
    const double maf = st->maf;
    const int orth = (method == 3 || method == 4 || method == 7 || method == 8);
    const int hwe = method <= 4;
    const int scale = (method == 1 || method == 3 || method == 5 || method == 7);
    /* Encode (encode_policy.h:35-52 raw, :59-79 orthogonal); NaN compares false and stays NaN */
    if (kind == 1) {
        for (int64_t i = 0; i < n; ++i) {
            const double e = locus[i];
            if (orth) {
                if (e == 2.0) locus[i] = (4 * maf) - 2;
                else if (e == 1.0) locus[i] = 2 * maf;
            } else if (e == 2.0) {
                locus[i] = 0;
            }
        }
    }
    /* Statistic policy: mean */
    if (kind == 0) {
        st->mean = maf * 2.0; /* statistic_policy.h:56-59, 95-98, 144-147 */
    } else if (!hwe) {
        double s = 0.0; /* :61-65 */
        for (int64_t i = 0; i < n; ++i) s += isnan(locus[i]) ? 0.0 : locus[i];
        st->mean = s / valid;
    } else if (!orth) {
        st->mean = 2 * (1 - maf) * maf; /* :100-102 */
    } else {
        st->mean = 2.0 * maf * maf; /* :149-151 */
    }
    /* impute + centre (:66-68, 103-105, 152-154) */
    for (int64_t i = 0; i < n; ++i) locus[i] = (isnan(locus[i]) ? st->mean : locus[i]) - st->mean;
    /* stddev */
    if (!hwe) {
        double ss = 0.0; /* :70-78 */
        for (int64_t i = 0; i < n; ++i) ss += locus[i] * locus[i];
        st->stddev = sqrt(ss / (valid - 1.0));
    } else if (!orth) {
        if (kind == 0) {
            st->stddev = sqrt(2.0 * maf * (1.0 - maf)); /* :111-115 */
        } else {
            const double q = 1 - maf, p_sq = maf * maf, q_sq = q * q; /* :116-122 */
            st->stddev = sqrt(2.0 * maf * q * (p_sq + q_sq));
        }
    } else {
        const double dsd = 2.0 * maf * (1.0 - maf); /* :159-168 */
        st->stddev = kind == 0 ? sqrt(dsd) : dsd;
    }
    if (st->stddev < 1e-10) st->is_monomorphic = 1; /* MONOMORPHIC_TOL :41 */
    if (st->is_monomorphic) return;                 /* genotype_processor.h:49-52 */
    if (scale)
        for (int64_t i = 0; i < n; ++i) locus[i] /= st->stddev; /* :53-56 */
}

/* include/gelex/data/genotype/genotype_processor.h:142-158 process_matrix: every column, optional freqs. */
ORACLE_API void oracle_process_matrix(double* chunk, int64_t n, int64_t c, int method, int kind, double* freqs) {
#pragma omp parallel for schedule(static)
    for (int64_t j = 0; j < c; ++j) {
        oracle_locus_stat st;
        oracle_process_column(chunk + j * n, n, method, kind, &st);
        if (freqs != NULL) freqs[j] = st.maf;
    }
}

/* src/data/grm/grm.cpp:42-63 — cblas_dsyrk(ColMajor, Lower, NoTrans, n, c, 1.0, Z, n, 1.0, G, n):
 * only entries (i >= j) of the column-major n x n matrix are touched; the upper triangle stays as it was. */
ORACLE_API void oracle_syrk_lower(const double* Z, int64_t n, int64_t c, double* G) {
#pragma omp parallel for schedule(dynamic, 4)
    for (int64_t j = 0; j < n; ++j) {
        for (int64_t k = 0; k < c; ++k) {
            const double zjk = Z[k * n + j];
            const double* zk = Z + k * n;
            double* gj = G + j * n;
            for (int64_t i = j; i < n; ++i) gj[i] += zk[i] * zjk;
        }
    }
}

/* include/gelex/data/grm/grm.h:94-137 GRM::compute: zero G, for each range, for each chunk: decode -> process -> syrk;
 * denominator = trace / n.  G is column-major n_out x n_out (caller-allocated), lower triangle only. */
ORACLE_API int oracle_grm_compute(const uint8_t* payload, int64_t m_total, int64_t bpv, int64_t n_raw, int64_t n_out,
                                  const int32_t* raw_to_out, int method, int kind, const int64_t* range_begin,
                                  const int64_t* range_end, int n_ranges, int64_t chunk_size, double* G, double* denominator,
                                  double* freqs_out) {
    if (chunk_size <= 0) return -1;
    memset(G, 0, (size_t)(n_out * n_out) * sizeof(double));
    double* chunk = (double*)malloc((size_t)(n_out * chunk_size) * sizeof(double));
    if (chunk == NULL) return -2;
    for (int r = 0; r < n_ranges; ++r) {
        if (range_begin[r] < 0 || range_end[r] > m_total || range_begin[r] >= range_end[r]) { /* bed_pipe.cpp:38-52 */
            free(chunk);
            return -3;
        }
        for (int64_t s = range_begin[r]; s < range_end[r]; s += chunk_size) {
            const int64_t e = s + chunk_size < range_end[r] ? s + chunk_size : range_end[r];
            oracle_decode_chunk(payload + s * bpv, e - s, bpv, n_raw, n_out, raw_to_out, chunk);
            oracle_process_matrix(chunk, n_out, e - s, method, kind, freqs_out != NULL ? freqs_out + s : NULL);
            oracle_syrk_lower(chunk, n_out, e - s, G);
        }
    }
    double tr = 0.0;
    for (int64_t i = 0; i < n_out; ++i) tr += G[i * n_out + i];
    if (denominator != NULL) *denominator = tr / (double)n_out;
    free(chunk);
    return 0;
}

/* src/data/grm/grm_bin_writer.cpp:34-69 — row-major lower triangle incl. diagonal, each value cast to float32,
 * no header; n == 0 writes nothing.  Here: into a caller buffer of n(n+1)/2 floats. */
ORACLE_API void oracle_pack_lower_f32(const double* G, int64_t n, float* out) {
    int64_t o = 0;
    for (int64_t i = 0; i < n; ++i)
        for (int64_t j = 0; j <= i; ++j) out[o++] = (float)G[j * n + i]; /* grm(i, j), column-major */
}

ORACLE_API int oracle_write_bin(const double* G, int64_t n, const char* path) {
    FILE* f = fopen(path, "wb");
    if (f == NULL) return -1;
    for (int64_t i = 0; i < n; ++i)
        for (int64_t j = 0; j <= i; ++j) {
            const float v = (float)G[j * n + i];
            if (fwrite(&v, sizeof(float), 1, f) != 1) { fclose(f); return -2; }
        }
    return fclose(f) == 0 ? 0 : -2;
}

/* src/data/grm/grm_loader.cpp:125-165 — read the packed float32 lower triangle back into a full symmetric FP64 matrix
 * and normalise by trace/n.  packed has n(n+1)/2 floats; out is n x n (symmetric, so layout-agnostic). */
ORACLE_API void oracle_load_normalized(const float* packed, int64_t n, double* out) {
    int64_t o = 0;
    double tr = 0.0;
    for (int64_t i = 0; i < n; ++i)
        for (int64_t j = 0; j <= i; ++j) {
            const double v = (double)packed[o++];
            out[i * n + j] = v;
            out[j * n + i] = v;
            if (i == j) tr += v;
        }
    const double k = tr / (double)n;
    for (int64_t i = 0; i < n * n; ++i) out[i] /= k;
}

/* src/data/grm/loco_grm_loader.cpp:62-76 — (G_whole - G_chr) / (k_whole - k_chr), k = trace/n of the un-normalised
 * matrices as read from the float32 files. */
ORACLE_API void oracle_loco(const float* whole_packed, const float* chr_packed, int64_t n, double* out) {
    int64_t o = 0;
    double tw = 0.0, tc = 0.0;
    for (int64_t i = 0; i < n; ++i)
        for (int64_t j = 0; j <= i; ++j) {
            const double w = (double)whole_packed[o], c = (double)chr_packed[o];
            ++o;
            out[i * n + j] = w - c;
            out[j * n + i] = w - c;
            if (i == j) { tw += w; tc += c; }
        }
    const double k = tw / (double)n - tc / (double)n;
    for (int64_t i = 0; i < n * n; ++i) out[i] /= k;
}

/* Genotype counts the GPU path derives its tables from (n0, n1, n2, nmiss over the projected column). */
ORACLE_API void oracle_counts(const double* col, int64_t n, int32_t* out4) {
    out4[0] = out4[1] = out4[2] = out4[3] = 0;
    for (int64_t i = 0; i < n; ++i) {
        if (isnan(col[i])) out4[3]++;
        else if (col[i] == 0.0) out4[0]++;
        else if (col[i] == 1.0) out4[1]++;
        else out4[2]++;
    }
}

/* Checker for the full-size configs (C2-C4), where the n x n FP64 matrix of oracle_grm_compute (20 GB at n = 50 000) and its
 * n^2 m flops are out of reach on the host: the same per-column arithmetic (oracle_decode_chunk + oracle_process_column, i.e.
 * bed_pipe.cpp:121-158 and genotype_processor.h:35-58) for `c` SNPs in parallel over columns, leaving the standardised chunk Z
 * (column-major n_out x c) for the caller - tests form selected ROWS of Z Z^T from it with an FP64 GEMM - and adding every
 * sample's z^2 into diag_accum[n_out] (the diagonal of Z Z^T, whose sum is the trace in GrmResult::denominator, grm.h:134). */
ORACLE_API void oracle_standardise_chunk(const uint8_t* payload, int64_t c, int64_t bpv, int64_t n_raw, int64_t n_out,
                                         const int32_t* raw_to_out, int method, int kind, double* Z, double* diag_accum,
                                         double* freqs) {
#pragma omp parallel
    {
        double* local = diag_accum != NULL ? (double*)calloc((size_t)n_out, sizeof(double)) : NULL;
#pragma omp for schedule(static)
        for (int64_t j = 0; j < c; ++j) {
            double* col = Z + j * n_out;
            oracle_locus_stat st;
            oracle_decode_chunk(payload + j * bpv, 1, bpv, n_raw, n_out, raw_to_out, col);
            oracle_process_column(col, n_out, method, kind, &st);
            if (freqs != NULL) freqs[j] = st.maf;
            if (local != NULL)
                for (int64_t i = 0; i < n_out; ++i) local[i] += col[i] * col[i];
        }
        if (local != NULL) {
#pragma omp critical
            for (int64_t i = 0; i < n_out; ++i) diag_accum[i] += local[i];
            free(local);
        }
    }
}
