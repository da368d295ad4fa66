"""Timings of the SURVEY 8f rows (the callers either side of the GRM path) through their C-ABI entry points, host buffers in and out,
with the CPU restatement (oracle, one thread) timed beside each on a bounded sample.  One JSON line per row.

    python tools/next_rows_probe.py [--quick]
"""
import ctypes as C
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import gelex_b200 as g  # noqa: E402
import oracle_lib as o  # noqa: E402  (checker / CPU baseline only)


def pinned(nbytes):
    lib = g._lib.load()
    p = C.c_void_p()
    assert lib.ggrm_host_alloc(nbytes, C.byref(p)) == 0
    return p


def best_of(fn, reps=3):
    ts = []
    for _ in range(reps):
        t0 = time.perf_counter()
        fn()
        ts.append(time.perf_counter() - t0)
    return min(ts)


def row_8f4_genotype_chunk(n, c):
    """ggrm_genotype_chunk (also the body of ggrm_assoc_chunk, 8f-2): packed chunk in, standardised FP64 n x c + LocusStatistic out."""
    lib = g._lib.load()
    bpv = (n + 3) // 4
    bed = g.synth_bed_host(n, 0, c, 5, 0.01)
    out_p = pinned(n * c * 8)
    freq, mean, sd = (np.empty(c) for _ in range(3))
    mono = np.empty(c, dtype=np.uint8)
    ptr = lambda a: a.ctypes.data_as(C.c_void_p)  # noqa: E731

    def run():
        st = lib.ggrm_genotype_chunk(0, ptr(bed), c, bpv, n, n, None, 3, 0, out_p, ptr(freq), ptr(mean), ptr(sd), ptr(mono))
        assert st == 0, lib.ggrm_last_error(None)

    run()
    t = best_of(run)
    # CPU restatement on a bounded sample of the same chunk (first c_cpu SNPs), one thread
    c_cpu = min(c, 400)
    raw = o.decode_chunk(bed[:c_cpu], n)
    t0 = time.perf_counter()
    raw2 = o.decode_chunk(bed[:c_cpu], n)
    zr, fr = o.process_matrix(raw2, 3, 0)
    t_cpu = time.perf_counter() - t0
    z = np.ctypeslib.as_array(C.cast(out_p, C.POINTER(C.c_double)), shape=(c, n))[:c_cpu].T
    ok = bool(np.array_equal(freq[:c_cpu], fr) and np.allclose(z, zr, rtol=1e-12, atol=1e-13, equal_nan=True))
    lib.ggrm_host_free(out_p)
    del raw
    return {"row": "8f-4 / 8f-2 ggrm_genotype_chunk (decode + OSH policy + LocusStatistic -> FP64 n x c on the host, pinned)", "n": n, "chunk_snps": c,
            "wall_ms": t * 1e3, "snps_per_s": c / t, "out_bytes": n * c * 8, "out_GBps": n * c * 8 / t / 1e9,
            "note": "bound by the D2H of the FP64 chunk over PCIe (one cudaMemcpy) plus one cudaMalloc/cudaFree set per call",
            "cpu_port": {"snps": c_cpu, "wall_ms": t_cpu * 1e3, "snps_per_s": c_cpu / t_cpu, "cores": 1}, "speedup_vs_cpu_port": (c / t) / (c_cpu / t_cpu),
            "matches_cpu_port": ok}


def row_8f1_bin_to_dense(n):
    """ggrm_bin_to_dense: packed float32 triangle (the .bin payload) -> dense symmetric FP64 / denominator (GrmLoader::load)."""
    lib = g._lib.load()
    rng = np.random.default_rng(1)
    packed = rng.standard_normal(g.tri(n), dtype=np.float32)
    diag = np.cumsum(np.arange(1, n + 1, dtype=np.int64)) - 1
    packed[diag] = np.abs(packed[diag]) + 1.0
    den = float(packed[diag].astype(np.float64).sum() / n)
    out_p = pinned(n * n * 8)
    ptr = lambda a: a.ctypes.data_as(C.c_void_p)  # noqa: E731
    lib.ggrm_bin_to_dense.argtypes = [C.c_int, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_int64, C.c_double, C.c_void_p]

    def run():
        st = lib.ggrm_bin_to_dense(0, ptr(packed), n, None, None, 0, None, n, den, out_p)
        assert st == 0, lib.ggrm_last_error(None)

    run()
    t = best_of(run)
    n_cpu = min(n, 6000)
    pk = np.ascontiguousarray(packed[: g.tri(n_cpu)])
    t0 = time.perf_counter()
    ref = o.load_normalized(pk, n_cpu)
    t_cpu = time.perf_counter() - t0
    # same arithmetic, different denominator (the sample's own trace): compare un-normalised entries
    got = np.ctypeslib.as_array(C.cast(out_p, C.POINTER(C.c_double)), shape=(n, n))
    den_cpu = float(pk[diag[:n_cpu]].astype(np.float64).sum() / n_cpu)
    ok = bool(np.allclose(got[:n_cpu, :n_cpu] * den, np.asarray(ref).reshape(n_cpu, n_cpu) * den_cpu, rtol=1e-12, atol=1e-12))
    lib.ggrm_host_free(out_p)
    b = g.tri(n) * 4 + n * n * 8
    return {"row": "8f-1 ggrm_bin_to_dense (GrmLoader::load: packed f32 triangle -> dense symmetric FP64 / (trace/n))", "n": n, "wall_ms": t * 1e3,
            "host_bytes_moved": b, "GBps": b / t / 1e9, "note": "H2D of the packed file (pageable) + kernel + D2H of n^2 doubles (pinned)",
            "cpu_port": {"n": n_cpu, "wall_ms": t_cpu * 1e3, "entries_per_s": n_cpu * n_cpu / t_cpu, "cores": 1},
            "speedup_vs_cpu_port": (n * n / t) / (n_cpu * n_cpu / t_cpu), "matches_cpu_port": ok}


def row_8f3_gcta(n, m):
    """GCTA mode: G and N = M M^T in one pass (two SYRKs per slab), then the G/N read-out; BED resident in HBM."""
    import torch

    bpv = (n + 3) // 4
    bed = torch.empty(m * bpv, dtype=torch.uint8, device="cuda")
    for s in range(0, m, 32768):
        g.synth_bed_device(bed.data_ptr() + s * bpv, n, s, min(32768, m - s), 20261106, 0.01)
    torch.cuda.synchronize()
    res = {}
    for tag, kw in (("plain", {}), ("gcta", {"want_dom": True, "second_matrix": 1})):
        acc = g.Accumulator(n, m, method="SH", precision=g.PREC_BF16, **kw)
        best = None
        prev = acc.timers()
        for p in range(2):
            if p:
                acc.reset()
            acc.push_device(bed.data_ptr(), 0, m)
            acc.finish()
            t = acc.timers()
            d = t["pass_ms"] - prev["pass_ms"]
            prev = t
            best = d if best is None else min(best, d)
        res[tag + "_pass_ms"] = best
        if tag == "gcta":
            lib = g._lib.load()
            buf = pinned(g.tri(n) * 4)
            for which in (0, 1):
                t0 = time.perf_counter()
                st = lib.ggrm_fetch_gcta(acc.h, which, g.GROUP_WHOLE, 0, n, buf)
                assert st == 0
                res["fetch_%s_ms" % ("ratio" if which == 0 else "counts")] = (time.perf_counter() - t0) * 1e3
                arr = np.ctypeslib.as_array(C.cast(buf, C.POINTER(C.c_float)), shape=(g.tri(n),))
                diag = np.cumsum(np.arange(1, n + 1, dtype=np.int64)) - 1
                if which == 0:
                    res["diag_mean"] = float(arr[diag].astype(np.float64).mean())       # ~1 for StandardizeHWE
                else:
                    res["counts_diag_mean_over_m"] = float(arr[diag].astype(np.float64).mean() / m)   # ~0.99
                    res["counts_are_integers"] = bool(np.all(arr[:: max(1, arr.size // 1000003)] == np.round(arr[:: max(1, arr.size // 1000003)])))
            lib.ggrm_host_free(buf)
        acc.close()
    fl = n * (n + 1) * m
    return {"row": "8f-3 GCTA mode (G and the non-missing pair counts N in one pass, G/N read-out)", "n": n, "m": m, **res,
            "plain_tflops": fl / res["plain_pass_ms"] / 1e9, "gcta_tflops_both_matrices": 2 * fl / res["gcta_pass_ms"] / 1e9}


def main():
    quick = "--quick" in sys.argv
    print(json.dumps(row_8f4_genotype_chunk(50_000 if not quick else 5000, 10_000 if not quick else 1000)), flush=True)
    print(json.dumps(row_8f1_bin_to_dense(20_000 if not quick else 3000)), flush=True)
    print(json.dumps(row_8f3_gcta(*((50_000, 500_000) if not quick else (4000, 20_000)))), flush=True)


if __name__ == "__main__":
    main()
