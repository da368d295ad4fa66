// Part of ggrm_api.cu: read-out (packed lower rows, trace, normalised / LOCO forms), accumulator access for the
// multi-GPU reduce, peer reduce.

#include <type_traits>

namespace ggrm {

static inline int64_t tri(int64_t i) { return i * (i + 1) / 2; }

// Resolve (kind, group) to a device accumulator; makes sure the data is final and returns the event to wait on
// (nullptr when the device is already idle for it).
static void* resolve_accum(Handle& h, int kind, int group, cudaEvent_t* wait_on, bool* banded = nullptr) {
    *wait_on = nullptr;
    if (banded != nullptr) *banded = false;
    if (kind != GGRM_KIND_ADD && kind != GGRM_KIND_DOM) throw CudaError{GGRM_ERR_INVALID_ARGUMENT, "kind must be 0 (add) or 1 (dom)"};
    if (h.pool[kind].empty()) throw CudaError{GGRM_ERR_INVALID_OPERATION, kind == 0 ? "additive matrix was not requested" : "dominance matrix was not requested"};
    const int ng = static_cast<int>(h.groups.size());
    if (group == GGRM_GROUP_WHOLE && ng > 1) {
        if (h.whole[kind] == nullptr) throw CudaError{GGRM_ERR_INVALID_OPERATION, "whole-genome matrix requires keep_whole when n_groups > 1"};
        if (!h.finished) throw CudaError{GGRM_ERR_INVALID_OPERATION, "whole-genome matrix is available after ggrm_finish"};
        if (h.banded_whole) {
            if (banded != nullptr) *banded = true;               // caller waits band by band
            else *wait_on = h.band_done.back();
        } else {
            for (auto it = h.groups.rbegin(); it != h.groups.rend(); ++it)
                if (it->complete) { *wait_on = it->done; break; }  // the last fold into `whole` precedes it on the compute stream
        }
        return h.whole[kind];
    }
    const int g = group == GGRM_GROUP_WHOLE ? 0 : group;
    if (g < 0 || g >= ng) throw CudaError{GGRM_ERR_INVALID_ARGUMENT, "group index out of range"};
    GroupState& gs = h.groups[g];
    if (!gs.started) {
        if (!h.finished && !(h.cur_group > g)) throw CudaError{GGRM_ERR_INVALID_OPERATION, "group is not complete yet: push a later group or call ggrm_finish"};
        start_group(h, g, true);  // no SNP of this group was pushed on this handle: an all-zero partial
        gs.folded = true;
        complete_group(h, g);
    }
    if (!gs.complete) throw CudaError{GGRM_ERR_INVALID_OPERATION, "group is not complete yet: push a later group or call ggrm_finish"};
    if (gs.buffer[kind] < 0) throw CudaError{GGRM_ERR_INVALID_OPERATION, "group accumulator was released"};
    *wait_on = gs.done;
    if (banded != nullptr && g == h.banded_group) *banded = true;  // rows become final band by band (Handle::band_done)
    return h.pool[kind][gs.buffer[kind]];
}

template <typename TAcc>
__global__ void __launch_bounds__(256) trace_kernel(const TAcc* __restrict__ A, const int32_t* __restrict__ slot_of, int64_t n,
                                                    double* __restrict__ trace) {
    double local = 0.0;
    for (int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
        const int64_t I = i >> 8;
        const int32_t slot = slot_of[I * (I + 1) / 2 + I];
        if (slot >= 0) local += static_cast<double>(A[static_cast<int64_t>(slot) * 65536 + tile_elem_offset<TAcc>(static_cast<int>(i & 255), static_cast<int>(i & 255))]);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) local += __shfl_xor_sync(0xffffffffu, local, o);
    if ((threadIdx.x & 31) == 0) atomicAdd(trace, local);
}

// GCTA diagonal sums the ratio read-outs use: the all-shard total once ggrm_gcta_diag_reduce_async has produced it in this
// pass, else this handle's own sums (complete when the handle saw every SNP: one GPU, or a 2D-tile rank); nullptr without the option.
static const double* gcta_diag_for_readout(Handle& h) {
    if (h.d_gcta_diag == nullptr) return nullptr;
    return h.gcta_total_valid ? h.d_gcta_total : h.d_gcta_diag;
}

// total[i] = own[i] + sum_p peers[p][i], in peer order: every rank passes its peers in rank order around itself, so the
// totals of two ranks may differ in the last bit; each rank only ever writes the diagonal entries of rows it owns.
__global__ void __launch_bounds__(256) gcta_diag_total_kernel(const double* __restrict__ own, PeerPtrs peers, int n_peers, int64_t n, double* __restrict__ total) {
    const int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= n) return;
    double v = own[i];
    for (int p = 0; p < n_peers; ++p) v += static_cast<const double*>(peers.p[p])[i];
    total[i] = v;
}

template <typename TAcc>
static double full_trace(Handle& h, const TAcc* A) {
    GGRM_CUDA(cudaMemsetAsync(h.d_trace, 0, sizeof(double), h.s_fetch));
    trace_kernel<TAcc><<<std::max(1, std::min<int>(static_cast<int>((h.n_out + 255) / 256), h.sm_count * 4)), 256, 0, h.s_fetch>>>(A, h.d_slot_of, h.n_out, h.d_trace);
    GGRM_CUDA(cudaGetLastError());
    ++h.kernel_launches;
    double t = 0.0;
    GGRM_CUDA(cudaMemcpyAsync(&t, h.d_trace, sizeof(double), cudaMemcpyDeviceToHost, h.s_fetch));
    GGRM_CUDA(cudaStreamSynchronize(h.s_fetch));
    return t;
}

// mode 0: raw accumulator   mode 1: G/(trace/n)   mode 2: (G_whole - G_group)/(k_whole - k_group)
// mode 3: G_add / N entrywise, N = the non-missing-pair counts in the second accumulator (GCTA .grm.bin values; kind must be add)
template <typename TAcc, typename TOut>
static void fetch_typed(Handle& h, int kind, int group, int64_t row_begin, int64_t row_count, TOut* out, double* trace_out, int mode) {
    if (row_begin < 0 || row_count < 0 || row_begin + row_count > h.n_out)
        throw CudaError{GGRM_ERR_COLUMN_RANGE, "row range outside [0, n_out)"};
    if (out == nullptr && row_count > 0) throw CudaError{GGRM_ERR_INVALID_ARGUMENT, "null output buffer"};
    if (mode != 0 && h.cfg.tile_world > 1) throw CudaError{GGRM_ERR_INVALID_OPERATION, "normalised / LOCO read-out needs the full trace: not available in 2D-tile mode"};
    cudaEvent_t ev = nullptr;
    bool banded = false;
    const TAcc* A = static_cast<const TAcc*>(resolve_accum(h, kind, group, &ev, mode == 0 ? &banded : nullptr));
    if (ev != nullptr && !banded) GGRM_CUDA(cudaStreamWaitEvent(h.s_fetch, ev, 0));
    int band_waited = -1;
    const TAcc* B = nullptr;
    const TAcc* D = nullptr;
    const double* gdiag = nullptr;
    double scale = 1.0;
    if (mode == 3) {
        cudaEvent_t evn = nullptr;
        D = static_cast<const TAcc*>(resolve_accum(h, GGRM_KIND_DOM, group, &evn));
        if (evn != nullptr) GGRM_CUDA(cudaStreamWaitEvent(h.s_fetch, evn, 0));
        gdiag = gcta_diag_for_readout(h);
    }
    if (mode == 1) {
        const double tr = full_trace<TAcc>(h, A);
        scale = static_cast<double>(h.n_out) / tr;
    } else if (mode == 2) {
        cudaEvent_t ev2 = nullptr;
        const TAcc* W = static_cast<const TAcc*>(resolve_accum(h, kind, GGRM_GROUP_WHOLE, &ev2));
        if (ev2 != nullptr) GGRM_CUDA(cudaStreamWaitEvent(h.s_fetch, ev2, 0));
        const double kw = full_trace<TAcc>(h, W) / static_cast<double>(h.n_out);
        const double kc = full_trace<TAcc>(h, A) / static_cast<double>(h.n_out);
        B = A;
        A = W;
        scale = 1.0 / (kw - kc);
    }
    if (h.d_stage == nullptr) {
        const int64_t want = std::min<int64_t>(int64_t{256} << 20, std::max<int64_t>(tri(h.n_out) * 8, h.n_out * 8 + 1024));
        h.stage_bytes = std::max<int64_t>(want, h.n_out * 8);
        dev_alloc(h, &h.d_stage, static_cast<size_t>(h.stage_bytes), "read-out staging buffer");
    }
    const int64_t stage_elems = h.stage_bytes / static_cast<int64_t>(sizeof(TOut));
    const size_t tfetch = phase_begin(h, h.s_fetch, 4);
    GGRM_CUDA(cudaMemsetAsync(h.d_trace, 0, sizeof(double), h.s_fetch));
    int64_t r = row_begin;
    const int64_t r_end = row_begin + row_count;
    int64_t written = 0;
    while (r < r_end) {
        int64_t cnt = 1;
        int64_t limit = r_end;
        if (banded) {  // rows of band b are final once band_done[b] has fired: never pack across a band boundary
            int b = std::max(band_waited, 0);
            while (h.band_row_end[static_cast<size_t>(b)] <= r) ++b;
            limit = std::min(limit, h.band_row_end[static_cast<size_t>(b)]);
            if (b > band_waited) {
                GGRM_CUDA(cudaStreamWaitEvent(h.s_fetch, h.band_done[static_cast<size_t>(b)], 0));
                band_waited = b;
            }
        }
        // largest cnt with tri(r+cnt) - tri(r) <= stage_elems (at least one row; a row always fits by construction)
        {
            int64_t lo = 1, hi = limit - r;
            while (lo < hi) {
                const int64_t mid = (lo + hi + 1) / 2;
                if (tri(r + mid) - tri(r) <= stage_elems) lo = mid; else hi = mid - 1;
            }
            cnt = lo;
        }
        const int64_t elems = tri(r + cnt) - tri(r);
        if constexpr (std::is_same<TAcc, float>::value)  // one warp per row, eight consecutive rows per block (see pack_rows8_kernel)
            pack_rows8_kernel<TOut><<<dim3(static_cast<unsigned>((cnt + 7) / 8), static_cast<unsigned>(((r + cnt - 1) / kTile) / kPackTilesPerBlock + 1)), 256, 0, h.s_fetch>>>(
                A, B, D, h.d_slot_of, r, cnt, scale, static_cast<TOut*>(h.d_stage), h.d_trace, gdiag);
        else
            pack_rows_kernel<TAcc, TOut><<<static_cast<unsigned>(cnt), 256, 0, h.s_fetch>>>(A, B, D, h.d_slot_of, r, cnt, scale, static_cast<TOut*>(h.d_stage), h.d_trace, gdiag);
        GGRM_CUDA(cudaGetLastError());
        ++h.kernel_launches;
        GGRM_CUDA(cudaMemcpyAsync(out + written, h.d_stage, static_cast<size_t>(elems) * sizeof(TOut), cudaMemcpyDeviceToHost, h.s_fetch));
        written += elems;
        r += cnt;
    }
    double tr = 0.0;
    GGRM_CUDA(cudaMemcpyAsync(&tr, h.d_trace, sizeof(double), cudaMemcpyDeviceToHost, h.s_fetch));
    phase_end(h, h.s_fetch, tfetch);
    GGRM_CUDA(cudaStreamSynchronize(h.s_fetch));
    if (trace_out != nullptr) *trace_out = tr;
    if (h.finished) collect_timers(h);
    if (h.guard && h.finished) {  // debug mode only; skipped while a ggrm_finish_async tail is still running
        if (cudaStreamQuery(h.s_comp) == cudaSuccess) check_guards(h);
        else cudaGetLastError();
    }
}

// Rectangular block of the lower triangle -> host buffer with leading dimension ld (elements).  Used by the 2D-tile
// writer: one call per owned 2048x2048 super-tile, rows then pwrite()n at their .bin offsets.
template <typename TAcc>
static void fetch_block_typed(Handle& h, int kind, int group, int64_t row_begin, int64_t row_count, int64_t col_begin, int64_t col_count,
                              float* out, int64_t ld, bool gcta_ratio = false) {
    if (row_begin < 0 || row_count < 0 || row_begin + row_count > h.n_out || col_begin < 0 || col_count < 0 || col_begin + col_count > h.n_out)
        throw CudaError{GGRM_ERR_COLUMN_RANGE, "block outside [0, n_out) x [0, n_out)"};
    if (row_count == 0 || col_count == 0) return;
    if (out == nullptr) throw CudaError{GGRM_ERR_INVALID_ARGUMENT, "null output buffer"};
    if (ld < col_count) throw CudaError{GGRM_ERR_INVALID_ARGUMENT, "leading dimension smaller than col_count"};
    cudaEvent_t ev = nullptr;
    bool banded = false;
    const TAcc* A = static_cast<const TAcc*>(resolve_accum(h, kind, group, &ev, &banded));
    if (banded) {
        size_t b = 0;
        while (h.band_row_end[b] < row_begin + row_count) ++b;
        ev = h.band_done[b];
    }
    if (ev != nullptr) GGRM_CUDA(cudaStreamWaitEvent(h.s_fetch, ev, 0));
    const TAcc* D = nullptr;
    const double* gdiag = nullptr;
    if (gcta_ratio) {  // the pair counts of the same rows: same band, on the same stream
        cudaEvent_t evn = nullptr;
        bool bandedn = false;
        D = static_cast<const TAcc*>(resolve_accum(h, GGRM_KIND_DOM, group, &evn, &bandedn));
        if (bandedn) evn = ev;
        if (evn != nullptr && evn != ev) GGRM_CUDA(cudaStreamWaitEvent(h.s_fetch, evn, 0));
        gdiag = gcta_diag_for_readout(h);
    }
    if (h.d_stage == nullptr) {
        const int64_t want = std::min<int64_t>(int64_t{256} << 20, std::max<int64_t>(tri(h.n_out) * 8, h.n_out * 8 + 1024));
        h.stage_bytes = std::max<int64_t>(want, h.n_out * 8);
        dev_alloc(h, &h.d_stage, static_cast<size_t>(h.stage_bytes), "read-out staging buffer");
    }
    const int64_t rows_per_pass = std::max<int64_t>(1, h.stage_bytes / (col_count * static_cast<int64_t>(sizeof(float))));
    const size_t tfetch = phase_begin(h, h.s_fetch, 4);
    for (int64_t r = 0; r < row_count; r += rows_per_pass) {
        const int64_t cnt = std::min(rows_per_pass, row_count - r);
        pack_block_kernel<TAcc, float><<<static_cast<unsigned>(cnt), 256, 0, h.s_fetch>>>(A, h.d_slot_of, row_begin + r, cnt, col_begin, col_count,
                                                                                         static_cast<float*>(h.d_stage), D, gdiag);
        GGRM_CUDA(cudaGetLastError());
        ++h.kernel_launches;
        GGRM_CUDA(cudaMemcpy2DAsync(out + r * ld, static_cast<size_t>(ld) * sizeof(float), h.d_stage, static_cast<size_t>(col_count) * sizeof(float),
                                    static_cast<size_t>(col_count) * sizeof(float), static_cast<size_t>(cnt), cudaMemcpyDeviceToHost, h.s_fetch));
    }
    phase_end(h, h.s_fetch, tfetch);
    GGRM_CUDA(cudaStreamSynchronize(h.s_fetch));
    if (h.finished) collect_timers(h);
}

template <typename TOut>
static void fetch_dispatch(Handle& h, int kind, int group, int64_t row_begin, int64_t row_count, TOut* out, double* trace_out, int mode) {
    if (h.fp64) fetch_typed<double, TOut>(h, kind, group, row_begin, row_count, out, trace_out, mode);
    else fetch_typed<float, TOut>(h, kind, group, row_begin, row_count, out, trace_out, mode);
}

}  // namespace ggrm

extern "C" int ggrm_fetch(ggrm_t* hh, int kind, int group, int64_t row_begin, int64_t row_count, float* out_lower_rows, double* trace_out) {
    Handle* h = reinterpret_cast<Handle*>(hh);
    GGRM_API_BEGIN(h)
    fetch_dispatch<float>(*h, kind, group, row_begin, row_count, out_lower_rows, trace_out, 0);
    GGRM_API_END(h)
}

extern "C" int ggrm_fetch_f64(ggrm_t* hh, int kind, int group, int64_t row_begin, int64_t row_count, double* out_lower_rows, double* trace_out) {
    Handle* h = reinterpret_cast<Handle*>(hh);
    GGRM_API_BEGIN(h)
    fetch_dispatch<double>(*h, kind, group, row_begin, row_count, out_lower_rows, trace_out, 0);
    GGRM_API_END(h)
}

extern "C" int ggrm_fetch_normalized(ggrm_t* hh, int kind, int group, int64_t row_begin, int64_t row_count, double* out_lower_rows) {
    Handle* h = reinterpret_cast<Handle*>(hh);
    GGRM_API_BEGIN(h)
    fetch_dispatch<double>(*h, kind, group, row_begin, row_count, out_lower_rows, nullptr, 1);
    GGRM_API_END(h)
}

extern "C" int ggrm_fetch_loco(ggrm_t* hh, int kind, int group, int64_t row_begin, int64_t row_count, double* out_lower_rows) {
    Handle* h = reinterpret_cast<Handle*>(hh);
    GGRM_API_BEGIN(h)
    if (h->groups.size() < 2) throw CudaError{GGRM_ERR_INVALID_OPERATION, "LOCO read-out needs n_groups > 1 and keep_whole"};
    fetch_dispatch<double>(*h, kind, group, row_begin, row_count, out_lower_rows, nullptr, 2);
    GGRM_API_END(h)
}

extern "C" int ggrm_fetch_gcta(ggrm_t* hh, int which, int group, int64_t row_begin, int64_t row_count, float* out_lower_rows) {
    Handle* h = reinterpret_cast<Handle*>(hh);
    GGRM_API_BEGIN(h)
    if (h->cfg.second_matrix == GGRM_SECOND_DOMINANCE || !h->cfg.want_add || !h->cfg.want_dom)
        throw CudaError{GGRM_ERR_INVALID_OPERATION, "GCTA read-out needs want_add, want_dom and second_matrix = GGRM_SECOND_NONMISSING"};
    if (which != 0 && which != 1) throw CudaError{GGRM_ERR_INVALID_ARGUMENT, "which must be 0 (A = G/N) or 1 (N)"};
    if (which == 0) fetch_dispatch<float>(*h, GGRM_KIND_ADD, group, row_begin, row_count, out_lower_rows, nullptr, 3);
    else fetch_dispatch<float>(*h, GGRM_KIND_DOM, group, row_begin, row_count, out_lower_rows, nullptr, 0);
    GGRM_API_END(h)
}

extern "C" int ggrm_fetch_block_gcta(ggrm_t* hh, int which, int group, int64_t row_begin, int64_t row_count, int64_t col_begin, int64_t col_count,
                                     float* out, int64_t ld) {
    Handle* h = reinterpret_cast<Handle*>(hh);
    GGRM_API_BEGIN(h)
    if (h->cfg.second_matrix == GGRM_SECOND_DOMINANCE || !h->cfg.want_add || !h->cfg.want_dom)
        throw CudaError{GGRM_ERR_INVALID_OPERATION, "GCTA read-out needs want_add, want_dom and second_matrix = GGRM_SECOND_NONMISSING"};
    if (which != 0 && which != 1) throw CudaError{GGRM_ERR_INVALID_ARGUMENT, "which must be 0 (A = G/N) or 1 (N)"};
    const int kind = which == 0 ? GGRM_KIND_ADD : GGRM_KIND_DOM;
    if (h->fp64) fetch_block_typed<double>(*h, kind, group, row_begin, row_count, col_begin, col_count, out, ld, which == 0);
    else fetch_block_typed<float>(*h, kind, group, row_begin, row_count, col_begin, col_count, out, ld, which == 0);
    GGRM_API_END(h)
}

extern "C" int ggrm_gcta_diag_address(ggrm_t* hh, void** d_ptr) {
    Handle* h = reinterpret_cast<Handle*>(hh);
    GGRM_API_BEGIN(h)
    if (h->d_gcta_diag == nullptr) throw CudaError{GGRM_ERR_INVALID_OPERATION, "the handle was not created with second_matrix = GGRM_SECOND_NONMISSING_GCTA_DIAG"};
    if (d_ptr == nullptr) throw CudaError{GGRM_ERR_INVALID_ARGUMENT, "null argument"};
    *d_ptr = h->d_gcta_diag;
    GGRM_API_END(h)
}

extern "C" int ggrm_gcta_diag_reduce_async(ggrm_t* hh, const void* const* peers, int n_peers) {
    Handle* h = reinterpret_cast<Handle*>(hh);
    GGRM_API_BEGIN(h)
    if (h->d_gcta_diag == nullptr) throw CudaError{GGRM_ERR_INVALID_OPERATION, "the handle was not created with second_matrix = GGRM_SECOND_NONMISSING_GCTA_DIAG"};
    if (n_peers < 0 || n_peers > kMaxPeers || (n_peers > 0 && peers == nullptr)) throw CudaError{GGRM_ERR_INVALID_ARGUMENT, "n_peers must be 0..7"};
    // this handle's own sums are final once its pass is: behind the group's completion (band 0 of a banded pass follows every
    // prepared slab of the pass, see accumulate_slabs)
    cudaEvent_t ev = nullptr;
    bool banded = false;
    resolve_accum(*h, GGRM_KIND_ADD, GGRM_GROUP_WHOLE, &ev, &banded);
    if (banded) ev = h->band_done.front();
    if (ev != nullptr) GGRM_CUDA(cudaStreamWaitEvent(h->s_fetch, ev, 0));
    PeerPtrs pp{};
    for (int i = 0; i < n_peers; ++i) pp.p[i] = peers[i];
    gcta_diag_total_kernel<<<static_cast<unsigned>((h->n_out + 255) / 256), 256, 0, h->s_fetch>>>(h->d_gcta_diag, pp, n_peers, h->n_out, h->d_gcta_total);
    GGRM_CUDA(cudaGetLastError());
    ++h->kernel_launches;
    h->gcta_total_valid = true;
    GGRM_API_END(h)
}

extern "C" int ggrm_fetch_block(ggrm_t* hh, int kind, int group, int64_t row_begin, int64_t row_count, int64_t col_begin, int64_t col_count,
                                float* out, int64_t ld) {
    Handle* h = reinterpret_cast<Handle*>(hh);
    GGRM_API_BEGIN(h)
    if (h->fp64) fetch_block_typed<double>(*h, kind, group, row_begin, row_count, col_begin, col_count, out, ld);
    else fetch_block_typed<float>(*h, kind, group, row_begin, row_count, col_begin, col_count, out, ld);
    GGRM_API_END(h)
}

extern "C" int ggrm_owned_supertiles(ggrm_t* hh, int64_t* super_rows, int64_t* super_cols, int64_t capacity, int64_t* count) {
    Handle* h = reinterpret_cast<Handle*>(hh);
    GGRM_API_BEGIN(h)
    const int64_t n = static_cast<int64_t>(h->owned_supers.size() / 2);
    if (count != nullptr) *count = n;
    if (super_rows != nullptr && super_cols != nullptr) {
        for (int64_t i = 0; i < std::min(n, capacity); ++i) {
            super_rows[i] = h->owned_supers[static_cast<size_t>(2 * i)];
            super_cols[i] = h->owned_supers[static_cast<size_t>(2 * i + 1)];
        }
    }
    GGRM_API_END(h)
}

extern "C" int ggrm_release_group(ggrm_t* hh, int kind, int group) {
    Handle* h = reinterpret_cast<Handle*>(hh);
    GGRM_API_BEGIN(h)
    if (kind < 0 || kind > 1 || group < 0 || group >= static_cast<int>(h->groups.size())) throw CudaError{GGRM_ERR_INVALID_ARGUMENT, "invalid kind/group"};
    GroupState& gs = h->groups[group];
    if (!gs.complete) throw CudaError{GGRM_ERR_INVALID_OPERATION, "cannot release a group that is still accumulating"};
    if (gs.buffer[kind] >= 0) {
        GGRM_CUDA(cudaStreamSynchronize(h->s_fetch));
        GGRM_CUDA(cudaEventSynchronize(gs.done));  // the fold into `whole` must have consumed it
        h->pool_owner[kind][gs.buffer[kind]] = -1;
        gs.buffer[kind] = -1;
    }
    GGRM_API_END(h)
}

extern "C" int ggrm_accum_buffer(ggrm_t* hh, int kind, int group, void** d_ptr, int64_t* n_elems, int32_t* elem_bytes) {
    Handle* h = reinterpret_cast<Handle*>(hh);
    GGRM_API_BEGIN(h)
    cudaEvent_t ev = nullptr;
    void* p = resolve_accum(*h, kind, group, &ev);
    if (ev != nullptr) GGRM_CUDA(cudaEventSynchronize(ev));
    if (d_ptr != nullptr) *d_ptr = p;
    if (n_elems != nullptr) *n_elems = h->accum_elems;
    if (elem_bytes != nullptr) *elem_bytes = h->elem_bytes;
    GGRM_API_END(h)
}

extern "C" int ggrm_reduce_peer(ggrm_t* hh, int kind, int group, const void* const* peers, int n_peers, int64_t slice_begin, int64_t slice_end) {
    Handle* h = reinterpret_cast<Handle*>(hh);
    GGRM_API_BEGIN(h)
    if (n_peers < 0 || n_peers > kMaxPeers) throw CudaError{GGRM_ERR_INVALID_ARGUMENT, "n_peers must be 0..7"};
    if (slice_begin < 0 || slice_end > h->accum_elems || slice_begin > slice_end || (slice_begin & 3) || (slice_end & 3))
        throw CudaError{GGRM_ERR_INVALID_ARGUMENT, "slice must be 4-element aligned and inside the accumulator"};
    cudaEvent_t ev = nullptr;
    void* mine = resolve_accum(*h, kind, group, &ev);
    if (ev != nullptr) GGRM_CUDA(cudaStreamWaitEvent(h->s_comp, ev, 0));
    PeerPtrs pp{};
    for (int i = 0; i < n_peers; ++i) pp.p[i] = peers[i];
    const int64_t len = slice_end - slice_begin;
    if (len > 0 && n_peers > 0) {
        const int64_t span = 256 * (16 / h->elem_bytes) * 4;  // elements one block covers per sweep (see reduce_peers_kernel)
        const int blocks = static_cast<int>(std::min<int64_t>((len + span - 1) / span, h->sm_count * 8));
        if (h->fp64) reduce_peers_kernel<double><<<blocks, 256, 0, h->s_comp>>>(static_cast<double*>(mine), pp, n_peers, slice_begin, slice_end);
        else reduce_peers_kernel<float><<<blocks, 256, 0, h->s_comp>>>(static_cast<float*>(mine), pp, n_peers, slice_begin, slice_end);
        GGRM_CUDA(cudaGetLastError());
        ++h->kernel_launches;
    }
    GGRM_CUDA(cudaStreamSynchronize(h->s_comp));
    GGRM_API_END(h)
}

// Band-pipelined form of the reduce-scatter: sums the peers' rows [row_begin, row_end) (whole super-rows) into this handle's on
// the READ-OUT stream, behind the band events that make this handle's own rows final, and returns without waiting.  A
// ggrm_fetch of those rows that follows is ordered after it.  The caller guarantees the peers' rows are final (their band flags).
extern "C" int ggrm_reduce_rows_async(ggrm_t* hh, int kind, int group, const void* const* peers, int n_peers, int64_t row_begin, int64_t row_end) {
    Handle* h = reinterpret_cast<Handle*>(hh);
    GGRM_API_BEGIN(h)
    if (n_peers < 0 || n_peers > kMaxPeers) throw CudaError{GGRM_ERR_INVALID_ARGUMENT, "n_peers must be 0..7"};
    if (h->cfg.tile_world != 1) throw CudaError{GGRM_ERR_INVALID_OPERATION, "row slices are defined for handles that own every tile"};
    if (row_begin < 0 || row_end < row_begin || row_end > h->n_out || ((row_begin % 2048) != 0 && row_begin != h->n_out) ||
        ((row_end % 2048) != 0 && row_end != h->n_out))
        throw CudaError{GGRM_ERR_INVALID_ARGUMENT, "row range must be cut on multiples of 2048 rows (one super-row of tiles)"};
    cudaEvent_t ev = nullptr;
    bool banded = false;
    void* mine = resolve_accum(*h, kind, group, &ev, &banded);
    if (banded && row_end > row_begin) {
        size_t b = 0;
        while (h->band_row_end[b] < row_end) ++b;
        ev = h->band_done[b];
    }
    if (ev != nullptr) GGRM_CUDA(cudaStreamWaitEvent(h->s_fetch, ev, 0));
    const int64_t A = row_begin == h->n_out ? h->ntiles_side : row_begin / kTile;
    const int64_t B = row_end == h->n_out ? h->ntiles_side : row_end / kTile;
    const int64_t slice_begin = A * (A + 1) / 2 * 65536, slice_end = B * (B + 1) / 2 * 65536;
    PeerPtrs pp{};
    for (int i = 0; i < n_peers; ++i) pp.p[i] = peers[i];
    const int64_t len = slice_end - slice_begin;
    if (len > 0 && n_peers > 0) {
        const int64_t span = 256 * (16 / h->elem_bytes) * 4;
        const int blocks = static_cast<int>(std::min<int64_t>((len + span - 1) / span, h->sm_count * 8));
        if (h->fp64) reduce_peers_kernel<double><<<blocks, 256, 0, h->s_fetch>>>(static_cast<double*>(mine), pp, n_peers, slice_begin, slice_end);
        else reduce_peers_kernel<float><<<blocks, 256, 0, h->s_fetch>>>(static_cast<float*>(mine), pp, n_peers, slice_begin, slice_end);
        GGRM_CUDA(cudaGetLastError());
        ++h->kernel_launches;
    }
    GGRM_API_END(h)
}

extern "C" int ggrm_reduce_block_async(ggrm_t* hh, int kind, int group, const void* const* peers, int n_peers, int64_t super_row,
                                       int64_t super_col_begin, int64_t super_col_end) {
    Handle* h = reinterpret_cast<Handle*>(hh);
    GGRM_API_BEGIN(h)
    if (n_peers < 0 || n_peers > kMaxPeers) throw CudaError{GGRM_ERR_INVALID_ARGUMENT, "n_peers must be 0..7"};
    if (h->cfg.tile_world != 1) throw CudaError{GGRM_ERR_INVALID_OPERATION, "super-tile ranges are defined for handles that own every tile"};
    const int64_t T = h->ntiles_side, nsr = (T + 7) / 8;
    if (super_row < 0 || super_row >= nsr || super_col_begin < 0 || super_col_end < super_col_begin || super_col_end > super_row + 1)
        throw CudaError{GGRM_ERR_INVALID_ARGUMENT, "super-tile range outside the lower triangle"};
    // storage slots: super-row by super-row, inside one super-tile by super-tile (build_ownership_and_items); a super-tile left of
    // the diagonal holds Ri x 8 tiles, the diagonal one Ri (Ri + 1) / 2 (Ri = tile rows of this super-row: 8, or fewer in the last)
    const int64_t A = super_row * 8, Ri = std::min<int64_t>(T, A + 8) - A;
    auto slot_at = [&](int64_t sc) { return A * (A + 1) / 2 + (sc <= super_row ? sc * Ri * 8 : super_row * Ri * 8 + Ri * (Ri + 1) / 2); };
    const int64_t slice_begin = slot_at(super_col_begin) * 65536, slice_end = slot_at(super_col_end) * 65536;
    cudaEvent_t ev = nullptr;
    bool banded = false;
    void* mine = resolve_accum(*h, kind, group, &ev, &banded);
    if (banded) {
        const int64_t row_end = std::min<int64_t>(h->n_out, (super_row + 1) * 2048);
        size_t b = 0;
        while (h->band_row_end[b] < row_end) ++b;
        ev = h->band_done[b];
    }
    if (ev != nullptr) GGRM_CUDA(cudaStreamWaitEvent(h->s_fetch, ev, 0));
    PeerPtrs pp{};
    for (int i = 0; i < n_peers; ++i) pp.p[i] = peers[i];
    const int64_t len = slice_end - slice_begin;
    if (len > 0 && n_peers > 0) {
        const int64_t span = 256 * (16 / h->elem_bytes) * 4;
        const int blocks = static_cast<int>(std::min<int64_t>((len + span - 1) / span, h->sm_count * 8));
        if (h->fp64) reduce_peers_kernel<double><<<blocks, 256, 0, h->s_fetch>>>(static_cast<double*>(mine), pp, n_peers, slice_begin, slice_end);
        else reduce_peers_kernel<float><<<blocks, 256, 0, h->s_fetch>>>(static_cast<float*>(mine), pp, n_peers, slice_begin, slice_end);
        GGRM_CUDA(cudaGetLastError());
        ++h->kernel_launches;
    }
    GGRM_API_END(h)
}

extern "C" int ggrm_ipc_export(ggrm_t* hh, int kind, int group, void* handle64) {
    Handle* h = reinterpret_cast<Handle*>(hh);
    GGRM_API_BEGIN(h)
    if (h->guard) throw CudaError{GGRM_ERR_INVALID_OPERATION, "GGRM_GUARD offsets device buffers inside their allocations: IPC export is not available"};
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "ggrm.h documents a 64-byte handle");
    if (handle64 == nullptr) throw CudaError{GGRM_ERR_INVALID_ARGUMENT, "null handle buffer"};
    cudaEvent_t ev = nullptr;
    void* p = resolve_accum(*h, kind, group, &ev);
    cudaIpcMemHandle_t ipc;
    GGRM_CUDA(cudaIpcGetMemHandle(&ipc, p));
    std::memcpy(handle64, &ipc, sizeof(ipc));
    GGRM_API_END(h)
}

extern "C" int ggrm_row_slice(ggrm_t* hh, int64_t row_begin, int64_t row_end, int64_t* elem_begin, int64_t* elem_end) {
    Handle* h = reinterpret_cast<Handle*>(hh);
    GGRM_API_BEGIN(h)
    if (h->cfg.tile_world != 1) throw CudaError{GGRM_ERR_INVALID_OPERATION, "row slices are defined for handles that own every tile"};
    if (row_begin < 0 || row_end < row_begin || row_end > h->n_out || ((row_begin % 2048) != 0 && row_begin != h->n_out) ||
        ((row_end % 2048) != 0 && row_end != h->n_out))
        throw CudaError{GGRM_ERR_INVALID_ARGUMENT, "row slice must be cut on multiples of 2048 rows (one super-row of tiles)"};
    // storage slots are numbered super-row by super-row (build_ownership_and_items), so the tiles of tile rows [A, B) are the
    // contiguous slots [A(A+1)/2, B(B+1)/2)
    const int64_t A = row_begin == h->n_out ? h->ntiles_side : row_begin / kTile;
    const int64_t B = row_end == h->n_out ? h->ntiles_side : row_end / kTile;
    if (elem_begin != nullptr) *elem_begin = A * (A + 1) / 2 * 65536;
    if (elem_end != nullptr) *elem_end = B * (B + 1) / 2 * 65536;
    GGRM_API_END(h)
}
