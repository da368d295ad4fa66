// Part of ggrm_api.cu: chunk streaming (push / flush / finish), per-SNP statistics read-back, timers.

namespace ggrm {

static size_t phase_begin(Handle& h, cudaStream_t s, int phase) {
    Timer t;
    GGRM_CUDA(cudaEventCreate(&t.a));
    GGRM_CUDA(cudaEventCreate(&t.b));
    t.phase = phase;
    GGRM_CUDA(cudaEventRecord(t.a, s));
    h.timers.push_back(t);
    return h.timers.size() - 1;
}
static void phase_end(Handle& h, cudaStream_t s, size_t idx) { GGRM_CUDA(cudaEventRecord(h.timers[idx].b, s)); }

static void collect_timers(Handle& h) {
    if (std::getenv("GGRM_DEBUG_GAPS") != nullptr) {  // developer aid: idle time on s_comp between consecutive SYRK phases
        const Timer* prev = nullptr;
        double gap_sum = 0;
        int n = 0;
        for (auto& t : h.timers) {
            if (t.phase != 3) continue;
            if (prev != nullptr) {
                float ms = 0.f;
                if (cudaEventElapsedTime(&ms, prev->b, t.a) == cudaSuccess) { gap_sum += ms; ++n; }
            }
            prev = &t;
        }
        std::fprintf(stderr, "[ggrm] SYRK gaps: %d gaps, total %.3f ms\n", n, gap_sum);
        if (h.pass_begin != nullptr && h.timers.size() <= 16) {  // short passes: every timed phase against the start of the pass
            static const char* names[] = {"copy", "prep", "expand", "syrk", "fetch", "5", "6"};
            for (auto& t : h.timers) {
                float a = 0.f, b = 0.f;
                if (cudaEventElapsedTime(&a, h.pass_begin, t.a) == cudaSuccess && cudaEventElapsedTime(&b, h.pass_begin, t.b) == cudaSuccess)
                    std::fprintf(stderr, "[ggrm]   %-7s %8.1f .. %8.1f us\n", names[t.phase < 7 ? t.phase : 6], a * 1e3, b * 1e3);
            }
            float e = 0.f;
            if (h.pass_end != nullptr && cudaEventElapsedTime(&e, h.pass_begin, h.pass_end) == cudaSuccess) std::fprintf(stderr, "[ggrm]   pass end %8.1f us\n", e * 1e3);
            cudaGetLastError();
        }
    }
    std::vector<Timer> pending;  // still running (ggrm_finish_async + partial read-out): collected by a later call
    for (auto& t : h.timers) {
        if (cudaEventQuery(t.b) == cudaErrorNotReady) {
            pending.push_back(t);
            continue;
        }
        float ms = 0.f;
        if (cudaEventElapsedTime(&ms, t.a, t.b) == cudaSuccess) h.phase_ms[t.phase] += ms;
        cudaEventDestroy(t.a);
        cudaEventDestroy(t.b);
    }
    cudaGetLastError();  // an unrecorded timer event must not poison later error checks
    h.timers.swap(pending);
}

static int group_of(const Handle& h, int64_t snp) {
    int lo = 0, hi = static_cast<int>(h.groups.size()) - 1;
    while (lo <= hi) {
        const int mid = (lo + hi) / 2;
        if (snp < h.groups[mid].begin) hi = mid - 1;
        else if (snp >= h.groups[mid].end) lo = mid + 1;
        else return mid;
    }
    return -1;
}

// zero = the accumulator must read as zero before the first SYRK.  The tcgen05 path does not need that: its first slab
// stores every element of every owned tile (accumulate = 0), which saves a 5 GB memset and a 5 GB read per group at n = 50k.
static void start_group(Handle& h, int g, bool zero) {
    GroupState& gs = h.groups[g];
    if (gs.started) return;
    for (int kind = 0; kind < 2; ++kind) {
        if (h.pool[kind].empty()) continue;
        int slot = -1;
        for (size_t p = 0; p < h.pool_owner[kind].size(); ++p)
            if (h.pool_owner[kind][p] < 0) { slot = static_cast<int>(p); break; }
        if (slot < 0)
            throw CudaError{GGRM_ERR_OUT_OF_MEMORY,
                            "accumulator pool exhausted: fetch and ggrm_release_group completed groups, or raise max_resident_groups"};
        h.pool_owner[kind][slot] = g;
        gs.buffer[kind] = slot;
        gs.touched[kind] = false;
        gs.split_touched[kind] = false;
        if (zero) {
            GGRM_CUDA(cudaMemsetAsync(h.pool[kind][slot], 0, static_cast<size_t>(std::max<int64_t>(h.accum_elems, 4)) * h.elem_bytes, h.s_comp));
            gs.touched[kind] = true;
        }
    }
    gs.started = true;
}

static void accumulate_slabs(Handle& h, int g, bool banded);

static void complete_group(Handle& h, int g) {
    GroupState& gs = h.groups[g];
    if (gs.complete) return;
    if (!h.deferred.empty() && h.deferred_group == g) accumulate_slabs(h, g, false);
    // split-K launches (Handle::ksplit): copies 1 .. ksplit-1 of the tiles are summed into copy 0, in a fixed order
    for (int kind = 0; kind < 2 && gs.started; ++kind) {
        if (!gs.split_touched[kind] || gs.buffer[kind] < 0) continue;
        float* base = static_cast<float*>(h.pool[kind][gs.buffer[kind]]);
        const int blocks = static_cast<int>(std::min<int64_t>((h.accum_elems / 4 + 255) / 256, h.sm_count * 8));
        sum_copies_kernel<<<blocks, 256, 0, h.s_comp>>>(base, h.accum_elems, h.ksplit);
        GGRM_CUDA(cudaGetLastError());
        ++h.kernel_launches;
        gs.split_touched[kind] = false;
    }
    if (gs.started && !gs.folded) {
        for (int kind = 0; kind < 2; ++kind) {
            if (h.whole[kind] == nullptr || gs.buffer[kind] < 0 || h.accum_elems == 0) continue;
            const int blocks = static_cast<int>(std::min<int64_t>((h.accum_elems / 4 + 255) / 256, h.sm_count * 8));
            if (h.fp64)
                axpy_accum_kernel<double><<<blocks, 256, 0, h.s_comp>>>(static_cast<double*>(h.whole[kind]),
                                                                         static_cast<const double*>(h.pool[kind][gs.buffer[kind]]), h.accum_elems);
            else
                axpy_accum_kernel<float><<<blocks, 256, 0, h.s_comp>>>(static_cast<float*>(h.whole[kind]),
                                                                        static_cast<const float*>(h.pool[kind][gs.buffer[kind]]), h.accum_elems);
            GGRM_CUDA(cudaGetLastError());
            ++h.kernel_launches;
        }
        gs.folded = true;
    }
    GGRM_CUDA(cudaEventRecord(gs.done, h.s_comp));
    gs.complete = true;
}

// items [item_begin, item_begin + item_count) of the tcgen05 work list (the whole list except in the banded final slab)
static void launch_syrk(Handle& h, int kind, int g, int64_t c, int zb, float* fold, int accumulate, int item_begin, int item_count) {
    GroupState& gs = h.groups[g];
    void* accum = h.pool[kind][gs.buffer[kind]];
    if (h.fp64) {
        if (h.num_items_f64 == 0) return;
        const int grid = std::min(h.num_items_f64, h.sm_count * 6);
        syrk_f64_kernel<<<grid, 128, 0, h.s_comp>>>(static_cast<const double*>(h.d_Z[kind][zb]), h.k_cap, c, static_cast<double*>(accum),
                                                    h.d_items_f64, h.num_items_f64, accumulate);
        GGRM_CUDA(cudaGetLastError());
    } else {
        if (item_count <= 0) return;
        const TcItem* items = h.d_items_tc + item_begin;
        const int nkb = static_cast<int>(round_up(c, 64) / 64);
        const uint32_t fmt = h.cfg.precision == GGRM_PREC_FP16 ? 0u : 1u;
        const int lo_blk = h.op_split ? static_cast<int>(h.k_cap / 64) : 0;
        // operand slab small enough to live in L2 (few-tile problems): every CTA pair walks its K range from a different starting
        // block, so that the pairs sharing a panel do not all ask one L2 slice for the same line at the same moment
        const int k_rotate = h.n_pad * round_up(c, 64) * 2 * (h.op_split ? 2 : 1) <= (int64_t{48} << 20) ? 13 : 0;
        // few tiles: cut this slab along K as well (whole list only; never combined with bands or the in-epilogue fold)
        int ksplit = 1, accumulate_split = 0;
        if (h.ksplit > 1 && nkb >= h.ksplit && item_begin == 0 && item_count == h.num_items_tc && fold == nullptr) {
            ksplit = h.ksplit;
            items = h.d_items_split;
            item_count = h.num_items_split;
            accumulate_split = gs.split_touched[kind] ? 1 : 0;
            gs.split_touched[kind] = true;
        }
        int grid_ctas = 0, n_clusters = 1;
        if (h.tc_mode == 2) {
            const int clusters = std::min(item_count, h.sm_count / 2);
            cudaLaunchConfig_t lc{};
            lc.gridDim = dim3(static_cast<unsigned>(2 * clusters));
            lc.blockDim = dim3(256);
            lc.dynamicSmemBytes = TcConfig<2>::kSmemBytes;
            lc.stream = h.s_comp;
            cudaLaunchAttribute attr[1];
            attr[0].id = cudaLaunchAttributeClusterDimension;
            attr[0].val.clusterDim.x = 2;
            attr[0].val.clusterDim.y = 1;
            attr[0].val.clusterDim.z = 1;
            lc.attrs = attr;
            lc.numAttrs = 1;
            GGRM_CUDA(cudaLaunchKernelEx(&lc, syrk_tc_kernel<2>, h.tmap[kind][zb], static_cast<float*>(accum),
                                         items, item_count, nkb, ksplit, lo_blk, k_rotate, accumulate, accumulate_split, fmt, h.cache_hints, h.d_fault,
                                         h.d_wave_ctr, h.wave_base, h.sync_every, fold));
            grid_ctas = 2 * clusters;
            n_clusters = clusters;
        } else {
            const int grid = std::min(item_count, h.sm_count);
            syrk_tc_kernel<1><<<grid, 256, TcConfig<1>::kSmemBytes, h.s_comp>>>(h.tmap[kind][zb], static_cast<float*>(accum), items,
                                                                                item_count, nkb, 1, lo_blk, k_rotate, accumulate, 0, fmt, h.cache_hints, h.d_fault,
                                                                                h.d_wave_ctr, h.wave_base, h.sync_every, fold);
            GGRM_CUDA(cudaGetLastError());
            grid_ctas = grid;
            n_clusters = grid;
        }
        if (h.sync_every > 0) {  // every CTA checks in once per pacing point (see syrk_tc.cuh)
            const int waves = (item_count + n_clusters - 1) / n_clusters;
            if (waves > 1) h.wave_base += static_cast<unsigned long long>((waves - 1) / h.sync_every) * static_cast<unsigned long long>(grid_ctas);
        }
    }
    ++h.syrk_launches;
    ++h.kernel_launches;
}


static void CUDART_CB band_flag_callback(void* p) {
    const auto* m = static_cast<const Handle::FlagMsg*>(p);
    __atomic_store_n(m->addr, m->value, __ATOMIC_RELEASE);
}

// Enqueue "flags[band] = value" behind everything already on the compute stream.
static void post_band_flag(Handle& h, size_t band) {
    if (h.band_flags == nullptr) return;
    h.flag_msgs.push_back({h.band_flags + band, h.band_flag_value});
    GGRM_CUDA(cudaLaunchHostFunc(h.s_comp, band_flag_callback, &h.flag_msgs.back()));
}

// K4 over the parked slabs of group g (Handle::deferred; normally exactly one).  banded: row band by row band - every slab of
// band 0, event, every slab of band 1, event, ... - so that rows become final progressively (ggrm_finish_async + ggrm_fetch*).
static void accumulate_slabs(Handle& h, int g, bool banded) {
    if (h.deferred.empty()) return;
    GroupState& gs = h.groups[g];
    for (const auto& sl : h.deferred) GGRM_CUDA(cudaStreamWaitEvent(h.s_comp, h.prep_done[sl.zb], 0));
    const size_t ti = phase_begin(h, h.s_comp, 3);
    // last slab of a chromosome whose partial also feeds `whole`: fold inside that slab's SYRK epilogue (tcgen05 path)
    const auto& last = h.deferred.back();
    const bool fold_here = !h.fp64 && !gs.folded && last.first_snp + last.count == gs.end;
    int accumulate[2];
    for (int kind = 0; kind < 2; ++kind) {  // first slab of a group stores, later slabs add (see start_group)
        accumulate[kind] = gs.touched[kind] ? 1 : 0;
        gs.touched[kind] = true;
    }
    const bool final_pass = banded;  // called from finish: completion flags are posted per band (or once, when there is one band)
    banded = banded && !h.fp64 && h.band_item_end.size() > 1;
    const size_t nbands = banded ? h.band_item_end.size() : 1;
    int item_begin = 0;
    for (size_t b = 0; b < nbands; ++b) {
        const int item_end = banded ? h.band_item_end[b] : h.num_items_tc;
        for (size_t si = 0; si < h.deferred.size(); ++si) {
            const auto& sl = h.deferred[si];
            const bool is_last = si + 1 == h.deferred.size();
            for (int kind = 0; kind < 2; ++kind) {
                if (h.d_Z[kind][sl.zb] == nullptr) continue;
                launch_syrk(h, kind, g, sl.count, sl.zb, fold_here && is_last ? static_cast<float*>(h.whole[kind]) : nullptr,
                            si == 0 ? accumulate[kind] : 1, item_begin, item_end - item_begin);
            }
        }
        if (banded) GGRM_CUDA(cudaEventRecord(h.band_done[b], h.s_comp));
        if (final_pass && g + 1 == static_cast<int>(h.groups.size())) post_band_flag(h, b);
        item_begin = item_end;
    }
    if (final_pass && g + 1 == static_cast<int>(h.groups.size())) h.band_flags_posted = true;
    if (banded) {
        h.banded_group = g;
        h.banded_whole = fold_here;
    }
    if (fold_here) gs.folded = true;
    phase_end(h, h.s_comp, ti);
    for (const auto& sl : h.deferred) GGRM_CUDA(cudaEventRecord(h.z_free[sl.zb], h.s_comp));
    h.deferred.clear();
}

// K1-K3 of a slab run on s_prep, beside the previous slab's SYRK (which owns s_comp); the serial mode (GGRM_OVERLAP_PREP=0) and
// the fused single-slab launch - nothing to run beside, and one event hop less in front of the only SYRK launch - use s_comp
static cudaStream_t prep_stream(const Handle& h) { return h.overlap_prep && !h.fused_prep ? h.s_prep : h.s_comp; }

static void flush_chunk(Handle& h, bool final_flush = false) {
    if (h.cur_count == 0) return;
    const int b = h.cur_buf;
    const int zb = h.z_buf;
    const int64_t c = h.cur_count;
    const int g = h.cur_group;
    cudaStream_t sp = prep_stream(h);
    GGRM_CUDA(cudaEventRecord(h.h2d_done, h.s_copy));
    GGRM_CUDA(cudaStreamWaitEvent(sp, h.h2d_done, 0));
    GGRM_CUDA(cudaStreamWaitEvent(sp, h.z_free[zb], 0));  // the SYRKs of two slabs ago have finished reading Z[.][zb]
    start_group(h, g, h.fp64);

    StatsOut so{};
    so.counts = h.d_counts;
    so.freq = h.d_freq;
    so.second_nonmissing = h.cfg.second_matrix != GGRM_SECOND_DOMINANCE ? 1 : 0;
    so.gcta_w = h.d_gcta_w;
    so.range_flag = h.cfg.precision == GGRM_PREC_FP16 ? h.d_range_flag : nullptr;
    for (int kind = 0; kind < 2; ++kind) {
        so.tab16[kind] = h.d_tab16[kind];
        so.tab16_lo[kind] = h.op_split && h.d_tab16[kind] != nullptr ? h.d_tab16[kind] + h.k_cap : nullptr;
        so.tab64[kind] = h.d_tab64[kind];
    }
    const int64_t z_pitch = h.op_split ? 2 * h.k_cap : h.k_cap;  // hi in columns [0, k_cap), lo in [k_cap, 2 k_cap)
    size_t ti = phase_begin(h, sp, 1);
    if (h.fused_prep) {
        prep_fused_kernel<<<static_cast<unsigned>((c + 63) / 64), 256, prep_fused_smem_bytes(h.bpv_pad), sp>>>(
            h.direct_src != nullptr ? h.direct_src : h.d_raw[b], h.bpv_raw, h.d_dense, h.bpv_pad, h.n_out, h.n_pad, h.d_out_to_raw, c, h.cfg.method, h.cfg.all_missing_policy,
            h.cfg.precision == GGRM_PREC_FP16 ? 1 : 0, so, h.cur_first_snp, static_cast<uint16_t*>(h.d_Z[0][zb]), static_cast<uint16_t*>(h.d_Z[1][zb]), z_pitch,
            h.op_split ? h.k_cap : 0);
        GGRM_CUDA(cudaGetLastError());
        GGRM_CUDA(cudaEventRecord(h.raw_free[b], sp));
        ++h.kernel_launches;
        h.direct_src = nullptr;
    } else {
        repack_kernel<<<repack_grid(c, h.bpv_pad), 256, 0, sp>>>(h.d_raw[b], h.bpv_raw, h.d_dense, h.bpv_pad, h.n_out, h.d_out_to_raw, c);
        GGRM_CUDA(cudaGetLastError());
        GGRM_CUDA(cudaEventRecord(h.raw_free[b], sp));
        count_stats_kernel<<<static_cast<unsigned>((c + 7) / 8), 256, 0, sp>>>(
            h.d_dense, h.bpv_pad, h.n_out, h.n_pad, c, h.cfg.method, h.cfg.all_missing_policy,
            h.cfg.precision == GGRM_PREC_FP16 ? 1 : 0, so, h.cur_first_snp);
        GGRM_CUDA(cudaGetLastError());
        h.kernel_launches += 2;
    }
    if (h.d_gcta_w != nullptr) {  // GGRM_SECOND_NONMISSING_GCTA_DIAG: this chunk's share of the per-sample diagonal sums
        gcta_diag_kernel<<<static_cast<unsigned>(h.n_pad / kGctaSamplesPerBlock), 256, 0, sp>>>(h.d_dense, h.bpv_pad, h.d_gcta_w, c, h.cur_first_snp, h.n_pad, h.d_gcta_lanes,
                                                                                 h.d_gcta_diag);
        GGRM_CUDA(cudaGetLastError());
        ++h.kernel_launches;
    }
    phase_end(h, sp, ti);

    const dim3 egrid(static_cast<unsigned>((c + 63) / 64), static_cast<unsigned>(h.bpv_pad / 64));
    if (!h.fused_prep) ti = phase_begin(h, sp, 2);  // (the fused launch did K3 already: no empty phase, its two event records cost ~5 us of a 0.25 ms pass)
    for (int kind = 0; kind < 2 && !h.fused_prep; ++kind) {
        if (h.d_Z[kind][zb] == nullptr) continue;
        if (h.fp64)
            expand64_kernel<<<egrid, 256, 0, sp>>>(h.d_dense, h.bpv_pad, h.d_tab64[kind], c, static_cast<double*>(h.d_Z[kind][zb]), h.k_cap);
        else {
            expand16_kernel<<<egrid, 256, 0, sp>>>(h.d_dense, h.bpv_pad, h.d_tab16[kind], c, static_cast<uint16_t*>(h.d_Z[kind][zb]), z_pitch);
            if (h.op_split) {
                GGRM_CUDA(cudaGetLastError());
                ++h.kernel_launches;
                expand16_kernel<<<egrid, 256, 0, sp>>>(h.d_dense, h.bpv_pad, h.d_tab16[kind] + h.k_cap, c,
                                                       static_cast<uint16_t*>(h.d_Z[kind][zb]) + h.k_cap, z_pitch);
            }
        }
        GGRM_CUDA(cudaGetLastError());
        ++h.kernel_launches;
    }
    if (!h.fused_prep) phase_end(h, sp, ti);
    GGRM_CUDA(cudaEventRecord(h.prep_done[zb], sp));

    // K4.  A slab among the last tail_slabs of the pass is only parked here (its operands stay resident in their ring buffer);
    // ggrm_finish / ggrm_finish_async accumulates all parked slabs band-major so the read-out can start after the first band.
    const Handle::DeferredSlab slab{zb, c, h.cur_first_snp};
    if (final_flush) {
        if (!h.deferred.empty() && h.deferred_group != g) accumulate_slabs(h, h.deferred_group, true);
        h.deferred_group = g;
        h.deferred.push_back(slab);
        accumulate_slabs(h, g, true);
    } else if (h.cur_trailing && (h.deferred.empty() || h.deferred_group == g)) {
        // never park more slabs than there are spare operand buffers: pushes that skip SNPs inside the tail cut it into more
        // than tail_slabs pieces, and the ring would otherwise wrap onto a parked slab that has not been accumulated yet
        if (static_cast<int>(h.deferred.size()) + 1 > h.n_zbuf - 2) accumulate_slabs(h, h.deferred_group, false);
        h.deferred_group = g;
        h.deferred.push_back(slab);
    } else {
        if (!h.deferred.empty()) accumulate_slabs(h, h.deferred_group, false);
        h.deferred_group = g;
        h.deferred.push_back(slab);
        accumulate_slabs(h, g, false);
    }
    h.z_buf = (h.z_buf + 1) % h.n_zbuf;
    h.cur_buf ^= 1;
    h.cur_count = 0;
    h.cur_first_snp = -1;
    GGRM_CUDA(cudaStreamWaitEvent(h.s_copy, h.raw_free[h.cur_buf], 0));
}

// true when p is memory of this handle's own device (a peer's memory is copied: the copy engine does not need peer access)
static bool device_pointer_is_local(const Handle& h, const uint8_t* p) {
    cudaPointerAttributes attr{};
    if (cudaPointerGetAttributes(&attr, p) != cudaSuccess) {
        cudaGetLastError();
        return false;
    }
    return attr.type == cudaMemoryTypeDevice && attr.device == h.cfg.device;
}

// The slab being filled was registered in place (Handle::direct_src) and something else wants the chunk buffer or the source
// back: copy it now, as the ordinary path would have.
static void materialize_direct_src(Handle& h) {
    if (h.direct_src == nullptr) return;
    GGRM_CUDA(cudaMemcpyAsync(h.d_raw[h.cur_buf], h.direct_src, static_cast<size_t>(h.cur_count * h.bpv_raw), cudaMemcpyDeviceToDevice, h.s_copy));
    h.direct_src = nullptr;
}

static void copy_rows(Handle& h, const uint8_t* src, bool src_is_device, bool src_is_pinned, int64_t rows, uint8_t* dst) {
    const int64_t bytes = rows * h.bpv_raw;
    if (src_is_device) {
        GGRM_CUDA(cudaMemcpyAsync(dst, src, static_cast<size_t>(bytes), cudaMemcpyDeviceToDevice, h.s_copy));
        return;
    }
    if (src_is_pinned) {
        GGRM_CUDA(cudaMemcpyAsync(dst, src, static_cast<size_t>(bytes), cudaMemcpyHostToDevice, h.s_copy));
        return;
    }
    if (h.ring[0] == nullptr) {
        // 4 x 8 MiB: page-locking host memory costs ~0.5 ms per MiB and the driver serialises it across the handles of a multi-GPU
        // run; the pushing thread's memcpy (~8 GB/s) is the pace setter, so a deeper ring buys nothing
        h.ring_bytes = std::max<int64_t>(int64_t{8} << 20, h.bpv_raw);
        for (int i = 0; i < Handle::kRingSlots; ++i) {
            GGRM_CUDA(cudaHostAlloc(reinterpret_cast<void**>(&h.ring[i]), static_cast<size_t>(h.ring_bytes), cudaHostAllocDefault));
            GGRM_CUDA(cudaEventCreateWithFlags(&h.ring_free[i], cudaEventDisableTiming));
        }
    }
    int64_t done = 0;
    while (done < bytes) {
        const int slot = h.ring_next;
        h.ring_next = (h.ring_next + 1) % Handle::kRingSlots;
        GGRM_CUDA(cudaEventSynchronize(h.ring_free[slot]));  // blocks while the ring is full
        const int64_t n = std::min<int64_t>(h.ring_bytes, bytes - done);
        std::memcpy(h.ring[slot], src + done, static_cast<size_t>(n));
        GGRM_CUDA(cudaMemcpyAsync(dst + done, h.ring[slot], static_cast<size_t>(n), cudaMemcpyHostToDevice, h.s_copy));
        GGRM_CUDA(cudaEventRecord(h.ring_free[slot], h.s_copy));
        done += n;
    }
}

static void push_impl(Handle& h, const uint8_t* payload, bool is_device, int64_t snp_begin, int64_t snp_count, int64_t bpv) {
    if (h.finished) throw CudaError{GGRM_ERR_INVALID_OPERATION, "ggrm_push after ggrm_finish"};
    if (payload == nullptr && snp_count > 0) throw CudaError{GGRM_ERR_INVALID_ARGUMENT, "null payload"};
    if (bpv != h.bpv_raw)
        throw CudaError{GGRM_ERR_INVALID_ARGUMENT, "bytes_per_variant " + std::to_string(bpv) + " != ceil(n_raw/4) = " + std::to_string(h.bpv_raw)};
    if (snp_begin < 0 || snp_count < 0 || snp_begin + snp_count > h.cfg.m_total)
        throw CudaError{GGRM_ERR_COLUMN_RANGE, "invalid chunk range: [" + std::to_string(snp_begin) + ", " + std::to_string(snp_begin + snp_count) +
                                                   "). Total SNPs: " + std::to_string(h.cfg.m_total)};
    if (snp_count == 0) return;
    if (snp_begin <= h.last_snp_pushed) throw CudaError{GGRM_ERR_INVALID_OPERATION, "SNP ranges must be pushed in increasing order without overlap"};
    bool pinned = false;
    if (!is_device) {
        cudaPointerAttributes attr{};
        if (cudaPointerGetAttributes(&attr, payload) == cudaSuccess) pinned = attr.type == cudaMemoryTypeHost;
        else cudaGetLastError();
    }
    if (!h.pass_open) {
        if (h.pass_begin == nullptr) {
            GGRM_CUDA(cudaEventCreate(&h.pass_begin));
            GGRM_CUDA(cudaEventCreate(&h.pass_end));
        }
        GGRM_CUDA(cudaEventRecord(h.pass_begin, h.s_copy));
        h.pass_open = true;
    }
    size_t tcopy = SIZE_MAX;  // the copy phase is timed from its first copy (a payload read in place enqueues none)
    const int64_t end = snp_begin + snp_count;
    int64_t s = snp_begin;
    while (s < end) {
        const int g = group_of(h, s);
        if (g < 0) {  // SNP outside every group: not part of any requested matrix
            int64_t next = end;
            for (const auto& gr : h.groups) if (gr.begin > s) { next = std::min(next, gr.begin); break; }
            s = next;
            continue;
        }
        const int64_t run_end = std::min(end, h.groups[g].end);
        while (s < run_end) {
            if (h.cur_count > 0 && (g != h.cur_group || s != h.cur_first_snp + h.cur_count || h.cur_count == h.cur_cap)) flush_chunk(h);
            if (h.cur_count == 0) {
                for (int og = 0; og < g; ++og)
                    if (h.groups[og].started && !h.groups[og].complete) complete_group(h, og);
                h.cur_group = g;
                h.cur_first_snp = s;
                // equal slabs: the rest of the group in ceil(rest / k_cap) pieces (every launch pays the same 2 x accumulator
                // read-modify-write, so a short tail slab would be HBM-bound)
                int64_t seg_end = h.groups[g].end;
                if (h.push_end_hint > s) seg_end = std::min(seg_end, h.push_end_hint);  // SNP shard: the caller stops before the group does
                const int64_t rest = seg_end - s;
                int64_t pieces = (rest + h.k_cap - 1) / h.k_cap;
                h.cur_cap = std::min<int64_t>(h.k_cap, round_up((rest + pieces - 1) / pieces, 64));
                // The first slab of a pass is short: the tensor cores wait for its upload and its K1-K3 (8.5 + 0.5 ms for a
                // 15 600-SNP slab at 23 GB/s per rank on the 8-GPU box), so they get 4 096 SNPs to start on while the first
                // full slab is still arriving.  It stores instead of accumulating, so it costs one launch and no extra accumulator read.
                if (!h.first_slab_done && !h.fp64 && (h.ksplit == 1 || h.first_slab_with_split) && h.first_slab_snps > 0 && rest > h.k_cap) {  // (few-tile problems are launch-bound: no extra slab unless asked, GGRM_FIRST_SLAB_SPLIT)
                    h.cur_cap = std::min<int64_t>(h.cur_cap, round_up(h.first_slab_snps, 64));
                    pieces = 1 + (rest - h.cur_cap + h.k_cap - 1) / h.k_cap;
                }
                h.first_slab_done = true;
                ++h.slabs_started;
                // the last tail_slabs slabs of the pass are parked and accumulated band-major at finish
                // ... but never more than half of the pass: parked slabs are not accumulated before the last push, and a short pass
                // (an SNP shard of 8 GPUs is 5 slabs) would leave the tensor cores idle for its whole upload - 34 ms of H2D in front
                // of 112 ms of SYRK at N = 8
                const int64_t pass_slabs = h.slabs_started - 1 + pieces;
                const int64_t tail = std::min<int64_t>(h.tail_slabs, std::max<int64_t>(1, pass_slabs / 2));
                h.cur_trailing = h.tail_slabs > 0 && g + 1 == static_cast<int>(h.groups.size()) && pieces <= tail;
            }
            const int64_t take = std::min(run_end - s, h.cur_cap - h.cur_count);
            const uint8_t* src = payload + (s - snp_begin) * h.bpv_raw;
            // one device-resident push that is the whole single-slab pass: the fused prep launch reads it where it lies (no 2 x 25 MB
            // copy in front of a 0.25 ms pass at C1); anything else is copied into the chunk buffer as before
            if (is_device && h.fused_prep && h.direct_src_ok && h.cur_count == 0 && s == h.groups[g].begin && take == h.groups[g].end - s &&
                device_pointer_is_local(h, src)) {
                h.direct_src = src;
            } else {
                if (tcopy == SIZE_MAX) tcopy = phase_begin(h, h.s_copy, 0);
                materialize_direct_src(h);
                copy_rows(h, src, is_device, pinned, take, h.d_raw[h.cur_buf] + h.cur_count * h.bpv_raw);
            }
            h.cur_count += take;
            s += take;
        }
    }
    if (tcopy != SIZE_MAX) phase_end(h, h.s_copy, tcopy);
    h.last_snp_pushed = end - 1;
}

// Enqueue the rest of the pass (last slab band by band, folds, completion events) without waiting for the device.
static void finish_enqueue(Handle& h) {
    if (h.finished) return;
    flush_chunk(h, true);
    if (!h.deferred.empty()) accumulate_slabs(h, h.deferred_group, true);  // the pass ended exactly on a parked slab's boundary
    for (size_t g = 0; g < h.groups.size(); ++g)
        if (h.groups[g].started) complete_group(h, static_cast<int>(g));
    h.finished = true;
    if (h.pass_open) GGRM_CUDA(cudaEventRecord(h.pass_end, h.s_comp));
    if (h.band_flags != nullptr && !h.band_flags_posted) {  // nothing was accumulated at finish: every band is final already
        const size_t nb = std::max<size_t>(1, h.fp64 ? 1 : h.band_row_end.size());
        for (size_t b = 0; b < nb; ++b) post_band_flag(h, b);
        h.band_flags_posted = true;
    }
}

// Wait for the device, surface faults, collect timers.
static void finish_sync(Handle& h) {
    cudaError_t e1 = cudaStreamSynchronize(h.s_copy);
    const cudaError_t e1b = cudaStreamSynchronize(h.s_prep);
    if (e1 == cudaSuccess) e1 = e1b;
    const cudaError_t e1c = cudaStreamSynchronize(h.s_fetch);  // ggrm_reduce_rows_async runs there
    if (e1 == cudaSuccess) e1 = e1c;
    const cudaError_t e2 = cudaStreamSynchronize(h.s_comp);
    if (e1 != cudaSuccess || e2 != cudaSuccess) {
        int where = 0;
        cudaMemcpy(&where, h.d_fault, sizeof(int), cudaMemcpyDeviceToHost);
        throw CudaError{GGRM_ERR_CUDA, std::string("CUDA: device fault during GRM accumulation: ") +
                                           cudaGetErrorString(e2 != cudaSuccess ? e2 : e1) + " (pipeline wait id " + std::to_string(where) + ")"};
    }
    if (h.guard) {
        const char* e = std::getenv("GGRM_GUARD_SELFTEST");  // proves the detector: one byte past the dense chunk buffer
        if (e != nullptr && std::atoi(e) != 0) cudaMemset(reinterpret_cast<char*>(h.d_dense) + h.k_cap * h.bpv_pad, 0, 1);
    }
    check_guards(h);
    collect_timers(h);
    if (h.cfg.precision == GGRM_PREC_FP16) {
        unsigned long long bad = 0;
        GGRM_CUDA(cudaMemcpy(&bad, h.d_range_flag, sizeof(bad), cudaMemcpyDeviceToHost));
        if (bad != 0) {
            GGRM_CUDA(cudaMemset(h.d_range_flag, 0, sizeof(bad)));
            throw CudaError{GGRM_ERR_INVALID_OPERATION,
                            "FP16 operand overflow: a standardised genotype value of SNP " + std::to_string(bad - 1) +
                                " exceeds the binary16 range (65504); the matrix would hold inf/NaN where the FP64 reference is finite. "
                                "Use GGRM_PREC_BF16 (FP32 exponent range) or GGRM_PREC_FP64 for this input"};
        }
    }
    if (h.pass_open) {
        float ms = 0.f;
        if (cudaEventElapsedTime(&ms, h.pass_begin, h.pass_end) == cudaSuccess) h.pass_ms += ms;
        cudaGetLastError();
        h.pass_open = false;
    }
}

static void finish_impl(Handle& h) {
    finish_enqueue(h);
    finish_sync(h);
}

// Re-arm a finished handle for another pass over (the same or other) SNPs: accumulators are zeroed lazily when their
// group restarts, statistics are overwritten, buffers and work lists are reused.  Timers keep accumulating.
static void reset_impl(Handle& h) {
    GGRM_CUDA(cudaStreamSynchronize(h.s_fetch));
    if (!h.finished) h.pass_open = false;  // abandoned pass: no end event to measure against
    finish_sync(h);  // also closes a pass that was finished with ggrm_finish_async and read out without a blocking finish
    h.banded_group = -1;
    h.banded_whole = false;
    h.flag_msgs.clear();  // every callback of the finished pass has run (streams are idle)
    h.band_flags_posted = false;
    h.deferred.clear();  // (a finished pass has none; an abandoned one may)
    h.cur_trailing = false;
    for (auto& g : h.groups) {
        g.buffer[0] = g.buffer[1] = -1;
        g.started = g.complete = g.folded = false;
        g.touched[0] = g.touched[1] = false;
        g.split_touched[0] = g.split_touched[1] = false;
    }
    for (int kind = 0; kind < 2; ++kind) {
        std::fill(h.pool_owner[kind].begin(), h.pool_owner[kind].end(), -1);
        if (h.whole[kind] != nullptr)
            GGRM_CUDA(cudaMemsetAsync(h.whole[kind], 0, static_cast<size_t>(std::max<int64_t>(h.accum_elems, 4)) * h.elem_bytes, h.s_comp));
    }
    if (h.d_gcta_diag != nullptr) {  // the streams are idle (finish_sync); the next pass's first chunk adds to it on the prep stream
        cudaStream_t sp = prep_stream(h);
        GGRM_CUDA(cudaMemsetAsync(h.d_gcta_lanes, 0, static_cast<size_t>(kGctaLanes * h.n_pad) * sizeof(double), sp));
        GGRM_CUDA(cudaMemsetAsync(h.d_gcta_diag, 0, static_cast<size_t>(h.n_pad) * sizeof(double), sp));
        h.gcta_total_valid = false;
    }
    h.direct_src = nullptr;
    h.cur_count = 0;
    h.cur_first_snp = -1;
    h.cur_group = -1;
    h.last_snp_pushed = -1;
    h.finished = false;
    h.pass_open = false;
    h.first_slab_done = false;
    h.slabs_started = 0;
}

}  // namespace ggrm

extern "C" int ggrm_push(ggrm_t* hh, const uint8_t* bed_payload, int64_t snp_begin, int64_t snp_count, int64_t bytes_per_variant) {
    Handle* h = reinterpret_cast<Handle*>(hh);
    GGRM_API_BEGIN(h)
    push_impl(*h, bed_payload, false, snp_begin, snp_count, bytes_per_variant);
    GGRM_API_END(h)
}

extern "C" int ggrm_push_device(ggrm_t* hh, const uint8_t* d_bed_payload, int64_t snp_begin, int64_t snp_count, int64_t bytes_per_variant) {
    Handle* h = reinterpret_cast<Handle*>(hh);
    GGRM_API_BEGIN(h)
    push_impl(*h, d_bed_payload, true, snp_begin, snp_count, bytes_per_variant);
    GGRM_API_END(h)
}

extern "C" int ggrm_reset(ggrm_t* hh) {
    Handle* h = reinterpret_cast<Handle*>(hh);
    GGRM_API_BEGIN(h)
    reset_impl(*h);
    GGRM_API_END(h)
}

extern "C" int ggrm_finish(ggrm_t* hh) {
    Handle* h = reinterpret_cast<Handle*>(hh);
    GGRM_API_BEGIN(h)
    finish_impl(*h);
    GGRM_API_END(h)
}

extern "C" int ggrm_set_push_end(ggrm_t* hh, int64_t snp_end) {
    Handle* h = reinterpret_cast<Handle*>(hh);
    GGRM_API_BEGIN(h)
    if (snp_end < -1 || snp_end > h->cfg.m_total) throw CudaError{GGRM_ERR_COLUMN_RANGE, "push end outside [0, m_total]"};
    h->push_end_hint = snp_end;
    GGRM_API_END(h)
}

extern "C" int ggrm_set_band_flags(ggrm_t* hh, int32_t* flags, int32_t value) {
    Handle* h = reinterpret_cast<Handle*>(hh);
    GGRM_API_BEGIN(h)
    if (h->finished) throw CudaError{GGRM_ERR_INVALID_OPERATION, "ggrm_set_band_flags must precede ggrm_finish of the pass"};
    h->band_flags = flags;
    h->band_flag_value = value;
    GGRM_API_END(h)
}

extern "C" int ggrm_band_layout(ggrm_t* hh, int64_t* row_end, int capacity, int* n_bands) {
    Handle* h = reinterpret_cast<Handle*>(hh);
    GGRM_API_BEGIN(h)
    const bool banded = !h->fp64 && h->band_row_end.size() > 1;
    const int nb = banded ? static_cast<int>(h->band_row_end.size()) : 1;
    if (n_bands != nullptr) *n_bands = nb;
    if (row_end != nullptr)
        for (int b = 0; b < std::min(nb, capacity); ++b) row_end[b] = banded ? h->band_row_end[static_cast<size_t>(b)] : h->n_out;
    GGRM_API_END(h)
}

extern "C" int ggrm_accum_address(ggrm_t* hh, int kind, void** d_ptr) {
    Handle* h = reinterpret_cast<Handle*>(hh);
    GGRM_API_BEGIN(h)
    if (kind != GGRM_KIND_ADD && kind != GGRM_KIND_DOM) throw CudaError{GGRM_ERR_INVALID_ARGUMENT, "kind must be 0 (add) or 1 (dom)"};
    if (d_ptr == nullptr) throw CudaError{GGRM_ERR_INVALID_ARGUMENT, "null argument"};
    if (h->groups.size() != 1) throw CudaError{GGRM_ERR_INVALID_OPERATION, "ggrm_accum_address is defined for handles with one group"};
    if (h->pool[kind].empty()) throw CudaError{GGRM_ERR_INVALID_OPERATION, kind == 0 ? "additive matrix was not requested" : "dominance matrix was not requested"};
    const int slot = h->groups[0].started ? h->groups[0].buffer[kind] : 0;  // start_group takes the first free slot: 0 on a fresh / reset handle
    if (slot < 0) throw CudaError{GGRM_ERR_INVALID_OPERATION, "group accumulator was released"};
    *d_ptr = h->pool[kind][static_cast<size_t>(slot)];
    GGRM_API_END(h)
}

extern "C" int ggrm_finish_async(ggrm_t* hh) {
    Handle* h = reinterpret_cast<Handle*>(hh);
    GGRM_API_BEGIN(h)
    finish_enqueue(*h);
    GGRM_API_END(h)
}

extern "C" int ggrm_snp_stats(ggrm_t* hh, int64_t snp_begin, int64_t snp_count, int32_t* n0, int32_t* n1, int32_t* n2, int32_t* nmiss, double* freq) {
    Handle* h = reinterpret_cast<Handle*>(hh);
    GGRM_API_BEGIN(h)
    if (!h->cfg.keep_stats) throw CudaError{GGRM_ERR_INVALID_OPERATION, "ggrm_snp_stats requires keep_stats"};
    if (snp_begin < 0 || snp_count < 0 || snp_begin + snp_count > h->cfg.m_total)
        throw CudaError{GGRM_ERR_COLUMN_RANGE, "invalid SNP range for ggrm_snp_stats"};
    if (!h->finished) flush_chunk(*h);
    GGRM_CUDA(cudaStreamSynchronize(h->s_prep));
    GGRM_CUDA(cudaStreamSynchronize(h->s_comp));
    std::vector<int4> c(static_cast<size_t>(snp_count));
    GGRM_CUDA(cudaMemcpy(c.data(), h->d_counts + snp_begin, c.size() * sizeof(int4), cudaMemcpyDeviceToHost));
    for (int64_t i = 0; i < snp_count; ++i) {
        if (n0 != nullptr) n0[i] = c[i].x;
        if (n1 != nullptr) n1[i] = c[i].y;
        if (n2 != nullptr) n2[i] = c[i].z;
        if (nmiss != nullptr) nmiss[i] = c[i].w;
    }
    if (freq != nullptr) GGRM_CUDA(cudaMemcpy(freq, h->d_freq + snp_begin, static_cast<size_t>(snp_count) * sizeof(double), cudaMemcpyDeviceToHost));
    GGRM_API_END(h)
}

extern "C" int ggrm_get_timers(ggrm_t* hh, double* out8) {
    Handle* h = reinterpret_cast<Handle*>(hh);
    GGRM_API_BEGIN(h)
    if (out8 == nullptr) throw CudaError{GGRM_ERR_INVALID_ARGUMENT, "null output"};
    for (int i = 0; i < 5; ++i) out8[i] = h->phase_ms[i];
    out8[5] = static_cast<double>(h->syrk_launches);
    out8[6] = static_cast<double>(h->kernel_launches);
    out8[7] = h->pass_ms;
    GGRM_API_END(h)
}
