/* ggrm.h — C ABI of the B200-native GRM engine (libggrm_b200.so).
 *
 * This is the drop-in boundary for the `gelex grm` hot path.  The reference (r1cheu/gelex) has no FFI layer;
 * the narrowest seam it has is `GrmEngine::compute` (include/gelex/pipeline/grm_engine.h:30-51) whose inner loop
 * (include/gelex/data/grm/grm.h:94-137) calls, per chunk of SNPs,
 *     BedPipe::load_chunk        (src/data/genotype/bed_pipe.cpp:107-158)   2-bit decode + sample projection
 *     process_matrix<GT>         (include/gelex/data/genotype/genotype_processor.h:142-158)   code policy
 *     GRM::update_grm/cblas_dsyrk (src/data/grm/grm.cpp:42-63)              G_lower += Z Z^T
 * and then writes float32 row-major lower-triangle rows (src/data/grm/grm_bin_writer.cpp:54-61).
 * The entry points below replace exactly that: packed BED bytes in, packed lower-triangle float32 rows out.
 * Plain pointers and sizes only; no C++ or torch types cross this boundary; nothing throws across it.
 *
 * Threading: a handle is not thread-safe; distinct handles (one per GPU) may be driven from distinct threads
 * or processes.  All functions return GGRM_OK (0) or a negative ggrm_status; ggrm_last_error gives the text.
 */
#ifndef GGRM_H_
#define GGRM_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GGRM_ABI_VERSION 1

/* Status codes.  The C++ host shim re-throws each as the GelexException subclass named on the right
 * (include/gelex/exception.h:26-117) so CLI behaviour (log + exit 1, apps/main.cpp:73-98) is unchanged. */
typedef enum ggrm_status {
    GGRM_OK = 0,
    GGRM_ERR_INVALID_ARGUMENT = -1, /* ArgumentValidationException / InvalidInputException */
    GGRM_ERR_COLUMN_RANGE = -2,     /* ColumnRangeException (bed_pipe.cpp:38-52)            */
    GGRM_ERR_INVALID_OPERATION = -3,/* InvalidOperationException (call order)               */
    GGRM_ERR_OUT_OF_MEMORY = -4,    /* std::bad_alloc                                       */
    GGRM_ERR_CUDA = -5,             /* GelexException("CUDA: ...")                          */
    GGRM_ERR_NO_DEVICE = -6,        /* GelexException: no sm_100 device — there is NO CPU fallback */
    GGRM_ERR_FILE_FORMAT = -7       /* FileFormatException (truncated payload)              */
} ggrm_status;

/* GenotypeProcessMethod, include/gelex/data/genotype/genotype_processor.h:38-48 (same numeric values). */
typedef enum ggrm_method {
    GGRM_STANDARDIZE_HWE = 1,      /* SH  : "Yang" additive / "Zeng" [0,1,0] dominance */
    GGRM_CENTER_HWE = 2,           /* CH  */
    GGRM_ORTH_STANDARDIZE_HWE = 3, /* OSH : CLI default (apps/cli/grm/grm_args.cpp:42-48) */
    GGRM_ORTH_CENTER_HWE = 4,      /* OCH : "Vitezica" with --add --dom */
    GGRM_STANDARDIZE = 5,          /* S   */
    GGRM_CENTER = 6,               /* C   */
    GGRM_ORTH_STANDARDIZE = 7,     /* OS  */
    GGRM_ORTH_CENTER = 8           /* OC  */
} ggrm_method;

typedef enum ggrm_precision {
    GGRM_PREC_BF16 = 0, /* BF16 operands, FP32 accumulate in TMEM (tcgen05); tolerance 1e-4 on G/(trace/n) */
    GGRM_PREC_FP64 = 1, /* FP64 operands and accumulate (DMMA mma.sync.m8n8k4.f64); tolerance 1e-10       */
    GGRM_PREC_FP16 = 2  /* FP16 operands (11-bit mantissa), same tcgen05 pipe and throughput as BF16      */
} ggrm_precision;

enum { GGRM_KIND_ADD = 0, GGRM_KIND_DOM = 1 };
enum { GGRM_GROUP_WHOLE = -1 };

/* What an all-missing SNP contributes.  The reference leaves the column NaN
 * (include/gelex/internal/genotype_processor/genotype_processor.h:38-42) and dsyrk then poisons the whole GRM. */
enum { GGRM_ALLMISS_REFERENCE_NAN = 0, GGRM_ALLMISS_SKIP = 1 };
/* What the second accumulator (the `want_dom` slot, kind 1) holds: the dominance matrix of `method`, or the number of SNPs
 * non-missing in BOTH samples, N = M M^T with M = the 0/1 non-missing indicator (exact: integers < 2^24 in FP32) - the
 * `.grm.N.bin` matrix of the GCTA format. */
enum { GGRM_SECOND_DOMINANCE = 0, GGRM_SECOND_NONMISSING = 1,
       /* as GGRM_SECOND_NONMISSING, and ggrm_fetch_gcta(which = 0) returns GCTA's DEFAULT diagonal (--make-grm-alg 0),
        *   A_jj = 1 + (1 / N_jj) sum_i (x_ij^2 - (1 + 2 p_i) x_ij + 2 p_i^2) / (2 p_i (1 - p_i))     over the SNPs called in sample j,
        * instead of sum z^2 / N: one more per-sample FP64 sum per SNP chunk (x = A1 count, p = A1 frequency; monomorphic SNPs skipped).
        * Needs method GGRM_STANDARDIZE_HWE, whose off-diagonal sum (x_ij - 2p)(x_ik - 2p) / 2pq is GCTA's. */
       GGRM_SECOND_NONMISSING_GCTA_DIAG = 2 };

typedef struct ggrm_handle ggrm_t;

typedef struct ggrm_config {
    int32_t abi_version;       /* GGRM_ABI_VERSION */
    int32_t device;            /* CUDA device ordinal this handle owns */
    int64_t n_raw;             /* samples per SNP row in the BED = FAM lines (metadata.cpp:39-43) */
    int64_t n_out;             /* GRM rows = sorted, de-duplicated sample IDs (sample_manager.cpp:37-40) */
    const int32_t* raw_to_out; /* [n_raw] raw sample -> GRM row, -1 = dropped, later raw wins on duplicates
                                  (sample_projection.cpp:32-55, variant_decoder.cpp:79-126); NULL = identity */
    int64_t m_total;           /* number of SNPs in the BED (upper bound of any snp index pushed) */
    int32_t method;            /* ggrm_method */
    int32_t want_add;          /* accumulate the additive matrix  (GrmType A / AD, grm_work_plan.cpp:38-50) */
    int32_t want_dom;          /* accumulate the dominance matrix (GrmType D / AD); both = one decode, two SYRKs */
    int32_t precision;         /* ggrm_precision */
    int32_t n_groups;          /* >=1.  Groups are chromosome runs (grm_work_plan.cpp:54-81); 1 = no LOCO */
    const int64_t* group_begin;/* [n_groups] first SNP index of each group; NULL iff n_groups==1 (=> [0,m_total)) */
    const int64_t* group_end;  /* [n_groups] one past the last SNP index */
    int32_t keep_whole;        /* n_groups>1: also keep whole = sum of group partials (LOCO by subtraction) */
    int32_t keep_stats;        /* retain per-SNP counts + allele frequency for ggrm_snp_stats */
    int32_t all_missing_policy;/* GGRM_ALLMISS_* */
    int32_t tile_rank;         /* 2D-tile partition: this handle owns every tile whose LPT slot % tile_world == tile_rank */
    int32_t tile_world;        /* 1 = own all tiles (single GPU or SNP-shard mode) */
    int64_t chunk_snps;        /* device K-chunk (SNPs expanded to 16-bit operands at once); 0 = auto */
    int32_t max_resident_groups; /* accumulator pool size per kind; 0 = as many as fit / n_groups */
    int32_t second_matrix;     /* GGRM_SECOND_*; 0 in every reference-compatible run */
} ggrm_config_t;

/* Create / destroy.  Fails with GGRM_ERR_NO_DEVICE when no sm_100 GPU is present (no CPU fallback). */
int ggrm_create(const ggrm_config_t* cfg, ggrm_t** out);
void ggrm_destroy(ggrm_t* h);
const char* ggrm_last_error(const ggrm_t* h); /* h may be NULL: error of the last failed ggrm_create on this thread */

/* Feed packed BED payload: `snp_count` SNP rows of `bytes_per_variant` = ceil(n_raw/4) bytes, the bytes
 * BedMmapReader::chunk_ptr(snp_begin, snp_count) points at (mmap_reader.cpp:96-119).  SNP indices must be pushed in
 * increasing order; ranges may skip SNPs (SNP-sharded multi-GPU) but never overlap.  The caller keeps ownership:
 * pageable memory is staged through the library's pinned ring (may block while the ring is full); memory that is
 * already pinned (cudaHostRegister/cudaHostAlloc) is copied from directly and must stay valid until ggrm_finish. */
int ggrm_push(ggrm_t* h, const uint8_t* bed_payload, int64_t snp_begin, int64_t snp_count, int64_t bytes_per_variant);
/* Same, payload already resident in device memory of cfg.device (used by bench.py's HBM-resident leg).  The payload must stay
 * valid until ggrm_wait_copies or ggrm_finish has returned: it is normally copied on the copy stream, and a pass that is ONE slab
 * of few samples handed over in ONE call (the fused K1-K3 launch, n_pad <= 4096) reads it in place instead of copying it. */
int ggrm_push_device(ggrm_t* h, const uint8_t* d_bed_payload, int64_t snp_begin, int64_t snp_count,
                     int64_t bytes_per_variant);
/* Drain: every pushed SNP is accumulated, every group is complete.  Blocks until the device is idle. */
int ggrm_finish(ggrm_t* h);
/* Same, but returns as soon as the rest of the pass is ENQUEUED.  The last slab is accumulated in row bands, so a read-out
 * that follows (ggrm_fetch / ggrm_fetch_f64 / ggrm_fetch_block, which wait on the device for exactly the rows they return)
 * starts with the first rows while the later bands are still on the tensor cores (the last few slabs of a pass are kept
 * resident and accumulated band-major for this, so the overlap window is several slabs long): the D2H of a 5 GB matrix overlaps the
 * tail of the accumulation instead of following it.  After it only ggrm_fetch*, ggrm_finish (blocks, reports device faults),
 * ggrm_reset, ggrm_get_timers and ggrm_destroy may be called; no further ggrm_push. */
int ggrm_finish_async(ggrm_t* h);
/* Optional hint for callers that push only part of a group (SNP shards): no SNP index >= snp_end will be pushed in this
 * pass (-1 clears it).  Lets the library cut the shard into equal slabs and recognise its last slabs, which are accumulated
 * row band by row band at finish (see ggrm_finish_async).  Purely a scheduling hint: results do not depend on it. */
int ggrm_set_push_end(ggrm_t* h, int64_t snp_end);
/* Re-arm the handle for another pass (e.g. the dominance pass of a second invocation, or a benchmark step): all
 * accumulators read as zero again, buffers / work lists / streams are reused. */
int ggrm_reset(ggrm_t* h);

/* Bit-exact test hook: per-SNP genotype counts after projection (n0 = #g==0, n1, n2 = #A1A1, nmiss) and the A1
 * allele frequency `maf` of genotype_processor.h:44.  Requires keep_stats.  Any output pointer may be NULL. */
int ggrm_snp_stats(ggrm_t* h, int64_t snp_begin, int64_t snp_count, int32_t* n0, int32_t* n1, int32_t* n2,
                   int32_t* nmiss, double* freq);

/* Read rows [row_begin, row_begin+row_count) of matrix (kind, group) as packed row-major lower-triangle float32 —
 * exactly the .bin byte order of grm_bin_writer.cpp:54-61: row i contributes i+1 values (i,0..i).
 * out_lower_rows (host, caller-owned) must hold sum_{i}(i+1) floats.  trace_out (may be NULL) receives the FP64
 * sum of the diagonal entries in the requested rows (GrmResult::denominator = trace/n, grm.h:134).
 * group = GGRM_GROUP_WHOLE for the whole-genome matrix (n_groups==1, or keep_whole).  A group may be fetched as soon
 * as a SNP of a later group has been pushed (or after ggrm_finish).  In tile mode (tile_world>1) entries of tiles
 * this handle does not own are returned as 0 so per-rank results can be summed or written at their own offsets. */
int ggrm_fetch(ggrm_t* h, int kind, int group, int64_t row_begin, int64_t row_count, float* out_lower_rows,
               double* trace_out);
/* GCTA-format read-out (docs/source/data_formats.rst:188-208 names the format, the reference writes only its own
 * .bin): needs want_add, want_dom and second_matrix = GGRM_SECOND_NONMISSING.  which = 0: rows of A_ij = G_ij / N_ij (0 where
 * N_ij = 0), G the additive matrix of `method` (StandardizeHWE gives GCTA's sum of (x_i - 2p)(x_j - 2p) / 2pq over the SNPs
 * called in both samples; the diagonal uses the same sum, not GCTA's inbreeding-based variant); which = 1: rows of N_ij.
 * Same row-major packed lower-triangle float32 layout as ggrm_fetch = the byte order of .grm.bin / .grm.N.bin. */
int ggrm_fetch_gcta(ggrm_t* h, int which, int group, int64_t row_begin, int64_t row_count, float* out_lower_rows);
/* The same read-out as a rectangular block (layout and ownership rules of ggrm_fetch_block): what the 2D-tile writer stores
 * for the GCTA files.  Every rank of a 2D-tile run streams all SNPs, so its diagonal sums are complete without an exchange. */
int ggrm_fetch_block_gcta(ggrm_t* h, int which, int group, int64_t row_begin, int64_t row_count, int64_t col_begin,
                          int64_t col_count, float* out, int64_t ld);
/* SNP shards with GGRM_SECOND_NONMISSING_GCTA_DIAG: every handle holds the per-sample diagonal sums of ITS SNPs.
 * ggrm_gcta_diag_address gives the device address of that vector (n_out doubles, stable for the handle's life, final once
 * the handle's pass is - e.g. when any of its band flags is up); ggrm_gcta_diag_reduce_async adds the peers' vectors to this
 * handle's OWN sums into a separate total (on the read-out stream, so peers may do the same at the same time), which the
 * GCTA read-outs that follow use. */
int ggrm_gcta_diag_address(ggrm_t* h, void** d_ptr);
int ggrm_gcta_diag_reduce_async(ggrm_t* h, const void* const* peers, int n_peers);

/* 2D-tile mode read-out: the dense block rows [row_begin, +row_count) x columns [col_begin, +col_count) of the lower
 * triangle as float32 into `out` (host, caller-owned, row-major with leading dimension `ld` >= col_count elements).
 * Entries above the diagonal, and entries of tiles this handle does not own, are written as 0.  The writer calls it once
 * per owned super-tile (ggrm_owned_supertiles) and stores row i of the block at file offset 4*(i(i+1)/2 + col_begin),
 * min(col_count, i + 1 - col_begin) values - which is how "each GPU's own tiles go straight to the writer". */
int ggrm_fetch_block(ggrm_t* h, int kind, int group, int64_t row_begin, int64_t row_count, int64_t col_begin,
                     int64_t col_count, float* out, int64_t ld);
/* The 2048 x 2048 super-tiles (units of tile ownership, cfg.tile_rank / tile_world) this handle owns, as (super-row,
 * super-column) index pairs, super-column <= super-row; entry rows/cols = index * 2048.  *count receives the number
 * owned; at most `capacity` pairs are stored (either array may be NULL to query the count only). */
int ggrm_owned_supertiles(ggrm_t* h, int64_t* super_rows, int64_t* super_cols, int64_t capacity, int64_t* count);
/* Same rows in FP64 (only meaningful for GGRM_PREC_FP64; FP32 accumulators are widened). */
int ggrm_fetch_f64(ggrm_t* h, int kind, int group, int64_t row_begin, int64_t row_count, double* out_lower_rows,
                   double* trace_out);
/* Return the accumulator of a fetched group to the pool (needed only when n_groups > max_resident_groups). */
int ggrm_release_group(ggrm_t* h, int kind, int group);

/* Multi-GPU SNP-shard mode: device pointer + element count of the tile-packed accumulator of (kind, group) so the
 * host plumbing (NCCL / torch.distributed, or ggrm_reduce_peer) can sum partial GRMs across ranks in place.
 * Element type is float (BF16/FP16 precision) or double (FP64). */
int ggrm_accum_buffer(ggrm_t* h, int kind, int group, void** d_ptr, int64_t* n_elems, int32_t* elem_bytes);
/* Sum-reduce over NVLink peer memory without NCCL: `peers` are the accum buffers of the other ranks (device pointers
 * mapped into this process, e.g. cudaIpcOpenMemHandle); this rank reduces the slice [slice_begin, slice_end) of
 * elements into its own buffer with one kernel that reads all peers (reduce-scatter by slices).  */
int ggrm_reduce_peer(ggrm_t* h, int kind, int group, const void* const* peers, int n_peers, int64_t slice_begin,
                     int64_t slice_end);

/* Band-pipelined reduce-scatter (SNP shards, one process per GPU).  The final accumulation of a pass completes row band by row
 * band (ggrm_finish_async); ggrm_band_layout gives the bands (band b = rows [row_end[b-1], row_end[b]), whole 2 048-row
 * super-rows, the same on every rank of equal n_out).  With ggrm_set_band_flags, `flags[b] = value` is stored by a CUDA host
 * callback the moment band b is final on this device - put `flags` in memory every rank can see (POSIX shared memory) and the
 * owner of band b knows when all its peers' rows are ready without a collective.  It then calls ggrm_reduce_rows_async (sums
 * the peers' rows into its own on the read-out stream, behind its own band event, without blocking) followed by ggrm_fetch of
 * those rows - while every GPU is still accumulating the later bands.  `flags` must stay valid until the pass is finished
 * (ggrm_finish / ggrm_reset); NULL disables. */
int ggrm_band_layout(ggrm_t* h, int64_t* row_end, int capacity, int* n_bands);
/* Device address the accumulator of `kind` has for the whole pass on a handle with ONE group (SNP-shard mode), available
 * before anything is pushed - so peers can be wired up ahead of a single pass (ggrm_accum_buffer answers only once the group
 * is complete).  The contents are defined only after this handle's own band flags / ggrm_finish. */
int ggrm_accum_address(ggrm_t* h, int kind, void** d_ptr);
int ggrm_set_band_flags(ggrm_t* h, int32_t* flags, int32_t value);
int ggrm_reduce_rows_async(ggrm_t* h, int kind, int group, const void* const* peers, int n_peers, int64_t row_begin,
                           int64_t row_end);

/* The same at a finer grain, for the bands that finish last: sums the peers' 2 048 x 2 048 super-tiles
 * [super_col_begin, super_col_end) of super-row `super_row` (rows super_row * 2048 ..; super_col <= super_row; the super-tiles of
 * a super-row are contiguous in the accumulator) into this handle's, on the read-out stream, behind the band event of those rows.
 * A strip of the LAST band is hundreds of MB at n = 50 000; cut along its columns and dealt to all ranks, its reduce + read-out
 * (ggrm_fetch_block of the same rows x columns) is spread over every GPU's NVLink and PCIe link instead of one.  Needs
 * tile_world == 1. */
int ggrm_reduce_block_async(ggrm_t* h, int kind, int group, const void* const* peers, int n_peers, int64_t super_row,
                            int64_t super_col_begin, int64_t super_col_end);

/* Multi-PROCESS form of the same (one process per GPU, torch.distributed-style launch): export the accumulator of
 * (kind, group) as a CUDA IPC handle (64 bytes, cudaIpcMemHandle_t), ship it to the other ranks over any host channel,
 * open it there (maps the peer's memory over NVLink, enabling peer access lazily) and pass the mapped pointers to
 * ggrm_reduce_peer.  Row ranges whose tiles are contiguous in the accumulator (whole 8-tile super-rows, 2048 rows) are
 * given by ggrm_row_slice, so rank r reduces and then fetches/writes only its own rows of the .bin file. */
int ggrm_ipc_export(ggrm_t* h, int kind, int group, void* handle64);
int ggrm_ipc_open(int device, const void* handle64, void** d_ptr);
int ggrm_ipc_close(int device, void* d_ptr);
/* Element range [*elem_begin, *elem_end) of the tile-packed accumulator that holds every tile of GRM rows
 * [row_begin, row_end); both must be multiples of 2048 or equal to n_out (tile_world == 1); row_begin == row_end gives an
 * empty range. */
int ggrm_row_slice(ggrm_t* h, int64_t row_begin, int64_t row_end, int64_t* elem_begin, int64_t* elem_end);

/* Next-row helpers on data already resident: normalised and leave-one-chromosome-out matrices,
 * G/(trace/n) (grm_loader.cpp:125-165) and (G_whole - G_chr)/(k_whole - k_chr) (loco_grm_loader.cpp:62-76),
 * returned as packed lower rows in FP64. */
int ggrm_fetch_normalized(ggrm_t* h, int kind, int group, int64_t row_begin, int64_t row_count, double* out_lower_rows);
int ggrm_fetch_loco(ggrm_t* h, int kind, int group, int64_t row_begin, int64_t row_count, double* out_lower_rows);

/* The same consumer-side arithmetic from FILES (the reader half of SURVEY 8f-1): `whole_packed` (n_whole samples) and
 * optionally `chr_packed` (n_chr samples) are the float32 payloads of .bin files (host pointers, e.g. the mmap).  Writes
 * the dense symmetric column-major FP64 matrix
 *     out(i, j) = (W(fw(i), fw(j)) - C(fc(i), fc(j))) / denominator,   i, j < n_target,
 * fw = target_to_whole, fc = target_to_chr (NULL = identity; -1 = the target has no sample in that file and contributes
 * 0): GrmLoader::load_unnormalized(id_map) (grm_loader.cpp:171-231) with denominator 1, GrmLoader::load (:125-131) with
 * trace/n, LocoGRMLoader::load_loco_grm (loco_grm_loader.cpp:46-77) with the chromosome file and k_whole - k_chr.  The
 * caller computes the denominators from the file diagonals (n values) so their summation order is the reference's. */
int ggrm_bin_to_dense(int device, const float* whole_packed, int64_t n_whole, const int64_t* target_to_whole,
                      const float* chr_packed, int64_t n_chr, const int64_t* target_to_chr, int64_t n_target,
                      double denominator, double* out_colmajor);

/* Phase timers (CUDA events on the handle's streams), milliseconds accumulated since create:
 * [0] H2D copies  [1] repack+count+stats  [2] expand  [3] SYRK  [4] pack+D2H  [5] SYRK launches  [6] all kernels launched
 * [7] passes: first push enqueued -> last accumulation finished (events on the copy / compute streams) */
int ggrm_get_timers(ggrm_t* h, double* out8);

/* ---- Stateless test hooks (CUDA kernels; bit-exact layers of SURVEY §8c) ------------------------------------ */
/* Decode `snp_count` packed SNP rows into a column-major n_out x snp_count FP64 matrix on the host, with the same
 * values BedPipe::load_chunk produces: 2.0 / 1.0 / 0.0 / NaN (decode_lut.h:40-41), rows never written stay NaN. */
int ggrm_decode_chunk(int device, const uint8_t* bed_payload, int64_t snp_count, int64_t bytes_per_variant,
                      int64_t n_raw, int64_t n_out, const int32_t* raw_to_out, double* out_colmajor);
/* Decode + code policy (process_matrix<GT>) in FP64: out is column-major n_out x snp_count, freq[snp_count]. */
int ggrm_standardise_chunk(int device, const uint8_t* bed_payload, int64_t snp_count, int64_t bytes_per_variant,
                           int64_t n_raw, int64_t n_out, const int32_t* raw_to_out, int method, int kind,
                           int all_missing_policy, double* out_colmajor, double* freq);

/* Next consumer of the same two stages: the GWAS scan's chunk pre-processing,
 * dispatch_assoc_chunk_by_method (src/pipeline/assoc_detail.cpp:28-47) = load_chunk + process_matrix<GT>(method, Z, &freqs)
 * restricted to the centre-family methods 2, 4, 6, 8 (anything else: GGRM_ERR_INVALID_ARGUMENT with the reference's
 * message).  kind = 0 additive model, 1 dominance model.  Output as ggrm_standardise_chunk. */
int ggrm_assoc_chunk(int device, const uint8_t* bed_payload, int64_t snp_count, int64_t bytes_per_variant, int64_t n_raw,
                     int64_t n_out, const int32_t* raw_to_out, int method, int kind, double* out_colmajor, double* freq);

/* Whole-matrix standardised genotype export: the chunk body of GenotypeLoader::process
 * (include/gelex/data/genotype/genotype_loader.h:52-80, src/data/genotype/genotype_loader.cpp:51-79) and GenotypePipe::process
 * (genotype_pipe.h:55-89, genotype_pipe.cpp:28-49) = load_chunk + process_matrix<GT> for any of the 8 methods, the
 * standardised column-major n_out x snp_count chunk being the product, with every SNP's LocusStatistic
 * (include/gelex/types/genetic_effect_type.h:20-26) beside it: freq = maf, mean, stddev, is_monomorphic (stddev < 1e-10; an
 * all-missing SNP returns the default {0, 0, 0, false} and a NaN column, genotype_processor.h:38-42).  Any of the four
 * per-SNP outputs may be NULL. */
int ggrm_genotype_chunk(int device, const uint8_t* bed_payload, int64_t snp_count, int64_t bytes_per_variant, int64_t n_raw,
                        int64_t n_out, const int32_t* raw_to_out, int method, int kind, double* out_colmajor, double* freq,
                        double* mean, double* stddev, uint8_t* is_monomorphic);

/* ---- Synthetic PLINK BED payload: p_j~U(0.05,0.5), HWE genotypes, i.i.d. missing -------------------
 * Counter-based integer RNG, bit-identical on host and device.  Writes snp_count rows of ceil(n/4) bytes for SNPs
 * [snp_begin, snp_begin+snp_count); padding bits of the last byte are 0. */
int ggrm_synth_bed_host(uint8_t* out, int64_t n, int64_t snp_begin, int64_t snp_count, uint64_t seed,
                        double missing_rate);
int ggrm_synth_bed_device(int device, uint8_t* d_out, int64_t n, int64_t snp_begin, int64_t snp_count, uint64_t seed,
                          double missing_rate);

/* Single-process multi-GPU host (the C++ GrmEngine): let kernels on `device` read device memory of `peer_device`
 * over NVLink (cudaDeviceEnablePeerAccess; idempotent), so ggrm_reduce_peer can take the peers' ggrm_accum_buffer
 * pointers directly. */
int ggrm_enable_peer_access(int device, int peer_device);

/* Page-locked host memory for callers that want ggrm_push / ggrm_fetch* to copy without staging (the host engine's
 * read-out ring).  Plain cudaHostAlloc / cudaFreeHost, exposed so that a host written against this header needs no CUDA
 * headers of its own. */
int ggrm_host_alloc(int64_t bytes, void** out);
int ggrm_host_free(void* p);
/* Page-lock memory the caller already owns (cudaHostRegister, portable) - e.g. a populated shared mapping of an output file on
 * tmpfs - so that ggrm_fetch* can DMA rows straight into it instead of through a staging buffer and a host copy
 * (grm_bin_writer.cpp:54-61 is the step this replaces).  Fails (GGRM_ERR_CUDA) for memory the driver cannot pin, such as the
 * mapping of a file on a disk file system; the caller then keeps staging.  Unregister before unmapping. */
int ggrm_host_register(void* p, int64_t bytes);
int ggrm_host_unregister(void* p);

/* ---- One host read of the BED feeding several handles (2D-tile multi-GPU mode, SURVEY 8e input side) ------------------
 * Every GPU of a 2D-tile run accumulates every SNP.  Rather than each GPU pulling the whole file over its own PCIe link
 * (what BedPipe::load_chunk per GPU would amount to, bed_pipe.cpp:107-158), the host uploads chunk j ONCE, into a staging
 * slot on GPU j % N, and every handle takes it from there with ggrm_push_device (a peer copy over NVLink when the slot is
 * on another GPU; ggrm_enable_peer_access first).
 *   ggrm_stage_create   n_slots device buffers of slot_bytes on `device`
 *   ggrm_stage_upload   host bytes -> slot; pageable sources are staged through two pinned 32 MiB bounce buffers; returns
 *                       when the bytes are in device memory
 *   ggrm_stage_slot     device address of a slot
 *   ggrm_wait_copies    (on a handle) every copy enqueued by earlier ggrm_push* calls has read its source, so the slot
 *                       may be overwritten
 * Errors of the stage calls are reported through ggrm_last_error(NULL) on the calling thread. */
typedef struct ggrm_stage ggrm_stage_t;
int ggrm_stage_create(int device, int64_t slot_bytes, int n_slots, ggrm_stage_t** out);
void ggrm_stage_destroy(ggrm_stage_t* s);
int ggrm_stage_upload(ggrm_stage_t* s, int slot, const uint8_t* host_src, int64_t bytes);
int ggrm_stage_slot(ggrm_stage_t* s, int slot, const uint8_t** d_ptr);
int ggrm_wait_copies(ggrm_t* h);

/* ---- An entry point that allocates nothing ------------------------------------------------------------------------------
 * No counterpart in the reference: added for callers that cannot call cudaMalloc (code
 * under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * Per-SNP genotype counts and A1 allele frequency - the values of ggrm_snp_stats, bit for bit - for `snp_count` packed BED rows
 * that are ALREADY in device memory of `device`, identity projection (n_out == n_raw), written to DEVICE memory:
 *   d_counts4  [snp_count][4] int32 = {n0, n1, n2 (= #A1A1), nmiss} per SNP, 16-byte aligned; may be NULL
 *   d_freq     [snp_count] double; may be NULL (not both)
 *   stream     a cudaStream_t passed as a plain pointer, NULL = the default stream.  The two kernels are only ENQUEUED: the
 *              caller synchronises the stream before reading the outputs.
 * It calls no cudaMalloc / cudaFree or any other allocating routine.  The one intermediate of the computation (the chunk
 * repacked to 256-sample padding) lives in a static device array of GGRM_NOALLOC_SCRATCH_BYTES that every GPU loading this
 * library holds until the process exits; a call that needs more - snp_count * (round_up(n_raw, 256) / 4) bytes - is rejected
 * with GGRM_ERR_INVALID_ARGUMENT and a message naming the limit (ggrm_last_error(NULL)): split the chunk.  At 50 000 samples
 * that is 668 SNPs per call.  All calls on one device share that array: calls that overlap on the GPU (different streams or
 * host threads) race on it - issue them on one stream, or serialise them.  Nothing in it has to be cleared between calls. */
#define GGRM_NOALLOC_SCRATCH_BYTES (8 * 1024 * 1024)
int ggrm_snp_stats_noalloc(int device, const uint8_t* d_bed_payload, int64_t snp_count, int64_t bytes_per_variant, int64_t n_raw,
                           int32_t* d_counts4, double* d_freq, void* stream);

/* Start-up overlap for hosts: creates the CUDA context of `device` and loads this library's kernels (0.3-0.5 s on a cold
 * process) so that a host can do it on a helper thread while it parses FAM / BIM metadata, as the C++ GrmEngine does
 * (the reference has nothing comparable: its start-up is the BedPipe / SampleManager construction, grm.cpp:36-40).
 * Idempotent; errors are reported like those of ggrm_create (ggrm_last_error(NULL) on the calling thread). */
int ggrm_warm_device(int device);

/* Library / device facts for the harness. */
int ggrm_device_count(void);
const char* ggrm_version(void);

#ifdef __cplusplus
}
#endif
#endif /* GGRM_H_ */
