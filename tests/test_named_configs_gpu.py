"""Parity on the NAMED configs of BASELINE.json, GPU against the reference (run with `-m gpu` on a B200).

  C1 (2 000 x 50 000, Yang additive): the whole matrix against the UNMODIFIED reference binary (oracle/_ref/gelex_ref dense =
      GRM::compute<Add>, include/gelex/data/grm/grm.h:94-137) on G/(trace/n): FP64 path <= 1e-10, FP16 <= 1e-4, BF16 <= 1e-4 -
      the north star's bounds, nothing scaled.
  C2 (50 000 x 500 000), C3 (20 000 x 300 000, OCH add + dom), C4 (50 000 x 600 000, 22 chromosome partials + whole) at FULL
      shape: 48 rows x all columns recomputed in FP64 on the host from the oracle's standardised Z (oracle_standardise_chunk =
      the oracle's decode + process_column, then an FP64 GEMM), the trace from the full FP64 diagonal.  C4 adds the 16-bit
      leave-one-chromosome-out read-out (G_w - G_c)/(k_w - k_c) of loco_grm_loader.cpp:62-76.
The host side of a full-shape check costs 1-2 minutes of CPU per config (n x m standardised doubles pass through the cache).
"""
import json
import subprocess

import numpy as np
import pytest

import gelex_b200 as g
import oracle_lib as o
import plink_io

pytestmark = pytest.mark.gpu

TOL16 = 1e-4   # north star: "max abs error <= 1e-4 on the BF16/FP32 path"
TOL64 = 1e-10  # "<= 1e-10 on the FP64 path"
CHR_MB = [248, 242, 198, 190, 182, 171, 159, 145, 138, 134, 135, 133, 114, 107, 102, 90, 83, 80, 59, 64, 47, 51]


def chromosome_runs(m):
    """22 contiguous runs with lengths proportional to the human autosomes (same rule as tools/config_probe.py)."""
    cuts, acc = [0], 0
    for mb in CHR_MB:
        acc += mb
        cuts.append(round(m * acc / sum(CHR_MB)))
    cuts[-1] = m
    return [(cuts[i], cuts[i + 1]) for i in range(22)]


def device_payload(n, m, seed):
    import torch

    bpv = (n + 3) // 4
    d = torch.empty(m * bpv, dtype=torch.uint8, device="cuda")
    for s in range(0, m, 65536):
        g.synth_bed_device(d.data_ptr() + s * bpv, n, s, min(65536, m - s), seed, 0.01)
    torch.cuda.synchronize()
    return d


def spot_rows(n, k, seed):
    rng = np.random.default_rng(seed)
    rows = set(rng.integers(0, n, size=k // 2).tolist()) | set(rng.integers(n - n // 8, n, size=k // 2 - 2).tolist()) | {0, n - 1}
    return np.array(sorted(rows), dtype=np.int64)


def test_c1_matches_reference_binary(tmp_path):
    if not o.REF_BIN.exists():
        pytest.skip("oracle/_ref/gelex_ref not built")
    n, m = 2000, 50000
    bed = g.synth_bed_host(n, 0, m, 20261105, 0.01)
    chroms = [str(c + 1) for c, (b, e) in enumerate(chromosome_runs(m)) for _ in range(e - b)]
    plink_io.write_plink(tmp_path / "c1", bed, [("F", f"s{i:07d}") for i in range(n)], chroms)
    r = subprocess.run([str(o.REF_BIN), "dense", str(tmp_path / "c1"), "1", "A", "10000", "0", str(tmp_path / "ref.f64")],
                       capture_output=True, text=True, check=True)
    info = json.loads(r.stdout.strip().splitlines()[-1])
    raw = np.fromfile(tmp_path / "ref.f64", dtype=np.float64)
    assert int(raw[:1].view(np.int64)[0]) == n and raw.size == 2 + g.tri(n)
    den, ref = raw[1], raw[2:]
    assert den == pytest.approx(info["denominator"], rel=1e-15)
    diag = np.cumsum(np.arange(1, n + 1)) - 1
    assert ref[diag].sum() / n == pytest.approx(den, rel=1e-12)
    ref_n = ref / den
    for prec, tol, name in ((g.PREC_FP64, TOL64, "fp64"), (g.PREC_FP16, TOL16, "fp16"), (g.PREC_BF16, TOL16, "bf16")):
        with g.Accumulator(n, m, method=1, precision=prec) as acc:
            for s in range(0, m, 10000):
                acc.push(bed[s:s + 10000], s)
            acc.finish()
            packed, tr = acc.fetch(g.KIND_ADD, g.GROUP_WHOLE, dtype=np.float64)
        err = float(np.max(np.abs(packed / (tr / n) - ref_n)))
        print(f"C1 vs reference binary, {name}: max |G/k - ref| = {err:.3e} (bound {tol:g})")
        assert err <= tol, (name, err)
        if prec == g.PREC_FP64:  # the float32 file view: what GrmBinWriter would have written (grm_bin_writer.cpp:54-61), to one ulp
            ref32 = ref.astype(np.float32)
            assert np.all(np.abs(packed.astype(np.float32) - ref32) <= np.spacing(np.abs(ref32)))


def check_rows(acc, kind, group, rows, R, D, n, tol, label, loco_of=None):
    """GPU rows (normalised, or the LOCO read-out of `group` when loco_of = (R_whole, D_whole)) against the FP64 host rows."""
    worst = 0.0
    if loco_of is None:
        k = D.sum() / n
        for a, r in enumerate(rows):
            got = acc.fetch_normalized(kind, group, int(r), 1)
            worst = max(worst, float(np.max(np.abs(got - R[a, : r + 1] / k))))
    else:
        Rw, Dw = loco_of
        k = Dw.sum() / n - D.sum() / n
        for a, r in enumerate(rows):
            got = acc.fetch_loco(kind, group, int(r), 1)
            worst = max(worst, float(np.max(np.abs(got - (Rw[a, : r + 1] - R[a, : r + 1]) / k))))
    print(f"{label}: max abs err over {len(rows)} rows x all columns = {worst:.3e} (bound {tol:g})")
    assert worst <= tol, (label, worst)
    return worst


def test_c2_full_shape_rows_match_the_oracle():
    n, m = 50_000, 500_000
    d_bed = device_payload(n, m, 20261104 + 2)
    bed = d_bed.cpu().numpy().reshape(m, -1)
    rows = spot_rows(n, 48, 2)
    (R,), (D,) = o.row_spot(bed, n, 1, g.KIND_ADD, rows, chunk=2048)
    for precision, name in ((g.PREC_BF16, "bf16"), (g.PREC_FP16, "fp16")):
        with g.Accumulator(n, m, method=1, precision=precision, keep_stats=True) as acc:
            acc.push_device(d_bed.data_ptr(), 0, m)
            acc.finish()
            check_rows(acc, g.KIND_ADD, g.GROUP_WHOLE, rows, R, D, n, TOL16, f"C2 Yang additive 50000 x 500000 {name}")
            # raw diagonal entries (the .bin payload): FP32 accumulation in TMEM truncates, a uniform relative bias below 1.5e-4
            # that cancels in G/k; fetching one row returns its diagonal entry as the trace of that row
            for r in rows[::6]:
                _, grr = acc.fetch(g.KIND_ADD, g.GROUP_WHOLE, row_begin=int(r), row_count=1)
                assert abs(grr / D[r] - 1.0) < 1.5e-4, (r, grr, D[r])
            # allele frequencies of the full-size run: bit-exact against the oracle's (spot SNPs)
            for s in (0, 123_456, m - 64):
                z = np.empty((64, n))
                fr = np.empty(64)
                o.lib().oracle_standardise_chunk(o._p(bed[s:s + 64]), 64, (n + 3) // 4, n, n, None, 1, 0, o._p(z), None, o._p(fr))
                assert np.array_equal(acc.snp_stats(s, 64)[4], fr)


def test_c3_full_shape_add_and_dom_rows_match_the_oracle():
    n, m = 20_000, 300_000
    d_bed = device_payload(n, m, 20261104 + 3)
    bed = d_bed.cpu().numpy().reshape(m, -1)
    rows = spot_rows(n, 48, 3)
    host = {kind: o.row_spot(bed, n, 4, kind, rows, chunk=4096) for kind in (g.KIND_ADD, g.KIND_DOM)}
    for precision, name in ((g.PREC_FP16, "fp16"), (g.PREC_BF16, "bf16")):
        with g.Accumulator(n, m, method="OCH", want_add=True, want_dom=True, precision=precision) as acc:
            acc.push_device(d_bed.data_ptr(), 0, m)
            acc.finish()
            for kind in (g.KIND_ADD, g.KIND_DOM):
                (R,), (D,) = host[kind]
                check_rows(acc, kind, g.GROUP_WHOLE, rows, R, D, n, TOL16, f"C3 OCH {'add' if kind == 0 else 'dom'} 20000 x 300000 {name}")


def test_c4_full_shape_loco_readout_16bit():
    # 22 chromosome partials + whole in one pass.  FP16 operands are what the engine and `bench.py --config c4` run: every partial
    # (the shortest has 9 800 SNPs) clears FP16's 8 192-SNP threshold unsplit.  BF16 on the same input splits its operands (hi + lo)
    # because those partials are far below BF16's 2^18 - checked as well, on fewer partials (it costs 3x the tensor time).
    n, m = 50_000, 600_000
    groups = chromosome_runs(m)
    d_bed = device_payload(n, m, 20261104 + 4)
    bed = d_bed.cpu().numpy().reshape(m, -1)
    rows = spot_rows(n, 32, 4)
    R, D = o.row_spot(bed, n, 1, g.KIND_ADD, rows, groups=groups, chunk=2048)
    Rw, Dw = sum(R), sum(D)
    for precision, name, partials in ((g.PREC_FP16, "fp16", (0, 7, 20, 21)), (g.PREC_BF16, "bf16 hi+lo", (21,))):
        with g.Accumulator(n, m, method=1, precision=precision, groups=groups, keep_whole=True) as acc:
            acc.push_device(d_bed.data_ptr(), 0, m)
            acc.finish()
            check_rows(acc, g.KIND_ADD, g.GROUP_WHOLE, rows, Rw, Dw, n, TOL16, f"C4 {name} whole = sum of 22 partials")
            for c in partials:  # largest, middle, smallest chromosomes
                check_rows(acc, g.KIND_ADD, c, rows, R[c], D[c], n, TOL16, f"C4 {name} partial chr{c + 1} normalised")
                check_rows(acc, g.KIND_ADD, c, rows, R[c], D[c], n, TOL16, f"C4 {name} LOCO read-out without chr{c + 1}", loco_of=(Rw, Dw))
            # the whole-genome matrix is the device-side sum of the 22 partials: raw rows add up to FP32 rounding
            for r in rows[::8]:
                whole, _ = acc.fetch(g.KIND_ADD, g.GROUP_WHOLE, row_begin=int(r), row_count=1)
                parts = sum(acc.fetch(g.KIND_ADD, c, row_begin=int(r), row_count=1)[0].astype(np.float64) for c in range(22))
                assert np.max(np.abs(whole - parts)) <= 22 * 2.0 ** -23 * np.max(np.abs(parts)) + 1e-3, r
