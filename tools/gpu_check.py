"""Developer bring-up check run under gpurun: exercises every CUDA stage against the CPU oracle and prints max errors.
Not a test (tests/ holds those); this prints diagnostics without stopping at the first failure."""
import os
import sys
import time
import traceback

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import gelex_b200 as g  # noqa: E402
import oracle_lib as o  # noqa: E402


def normalised_err(packed, G_ref):
    n = G_ref.shape[0]
    idx = np.tril_indices(n)
    ref = G_ref[idx]
    k_ref = np.trace(G_ref) / n
    diag = np.cumsum(np.arange(1, n + 1)) - 1
    k = packed[diag].astype(np.float64).sum() / n
    return float(np.max(np.abs(packed.astype(np.float64) / k - ref / k_ref))), k, k_ref


def stage(name, fn):
    t = time.time()
    try:
        r = fn()
        print(f"[ok ] {name}: {r}  ({time.time() - t:.2f}s)", flush=True)
        return True
    except Exception as e:  # noqa: BLE001
        print(f"[ERR] {name}: {type(e).__name__}: {e}", flush=True)
        traceback.print_exc()
        return False


def main():
    only = sys.argv[1:] or ["decode", "std", "fp64", "tc1", "tc2"]
    n, m = 1000, 3000
    bed = g.synth_bed_host(n, 0, m, 7, 0.02)
    rng = np.random.default_rng(1)

    if "decode" in only:
        def dec():
            a = g.decode_chunk(bed[:500], n)
            b = o.decode_chunk(bed[:500], n)
            assert np.array_equal(np.isnan(a), np.isnan(b)) and np.array_equal(np.nan_to_num(a, nan=-1), np.nan_to_num(b, nan=-1))
            perm = rng.permutation(n).astype(np.int32)
            perm[::17] = -1
            keep = np.sort(perm[perm >= 0])
            remap = {v: i for i, v in enumerate(keep)}
            r2o = np.array([remap[v] if v >= 0 else -1 for v in perm], dtype=np.int32)
            n_out = len(keep)
            a = g.decode_chunk(bed[:300], n, n_out, r2o)
            b = o.decode_chunk(bed[:300], n, n_out, r2o)
            assert np.array_equal(np.nan_to_num(a, nan=-1), np.nan_to_num(b, nan=-1))
            return "dense+sparse bit-exact"
        stage("decode", dec)

    if "std" in only:
        def std():
            worst = 0.0
            for method in range(1, 9):
                for kind in (0, 1):
                    z, f = g.standardise_chunk(bed[:400], n, method, kind)
                    zr, fr = o.process_matrix(o.decode_chunk(bed[:400], n), method, kind)
                    assert np.array_equal(f, fr), f"freq mismatch method {method} kind {kind}"
                    err = np.max(np.abs(z - zr) / np.maximum(1.0, np.abs(zr)))
                    worst = max(worst, err)
            return f"freq bit-exact, max rel z err {worst:.3e}"
        stage("standardise 8 methods x add/dom", std)

    G_ref = {}

    def ref(method, kind):
        key = (method, kind)
        if key not in G_ref:
            G_ref[key] = o.grm_compute(bed, n, method, kind, chunk=1000)
        return G_ref[key]

    def run(precision, method=1, kinds=(0,), chunk=0, groups=None, label=""):
        acc = g.Accumulator(n, m, method=method, want_add=0 in kinds, want_dom=1 in kinds, precision=precision,
                            keep_stats=True, chunk_snps=chunk, groups=groups)
        try:
            for s in range(0, m, 700):
                e = min(m, s + 700)
                acc.push(bed[s:e], s)
            acc.finish()
            out = {}
            for kind in kinds:
                packed, tr = acc.fetch(kind, -1, dtype=np.float64 if precision == g.PREC_FP64 else np.float32)
                G, den, fr = ref(method, kind)
                err, k, k_ref = normalised_err(packed, G)
                out[kind] = (err, tr / n, den)
            n0, n1, n2, nm, f = acc.snp_stats(0, m)
            assert np.array_equal(f, ref(method, kinds[0])[2]), "freq not bit-exact"
            return out, acc.timers()
        finally:
            acc.close()

    if "fp64" in only:
        stage("fp64 SH add", lambda: run(g.PREC_FP64))
        stage("fp64 OCH add+dom", lambda: run(g.PREC_FP64, method=4, kinds=(0, 1)))
        stage("fp64 chunk=512", lambda: run(g.PREC_FP64, chunk=512))
    for mode in ("1", "2"):
        if ("tc" + mode) in only:
            os.environ["GGRM_TC_MODE"] = mode
            stage(f"bf16 tc_mode={mode} SH add", lambda: run(g.PREC_BF16))
            stage(f"fp16 tc_mode={mode} SH add", lambda: run(g.PREC_FP16))
            stage(f"bf16 tc_mode={mode} chunk=640 OCH add+dom", lambda: run(g.PREC_BF16, method=4, kinds=(0, 1), chunk=640))


if __name__ == "__main__":
    main()
