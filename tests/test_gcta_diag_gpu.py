"""SURVEY 8f-3 leftovers: GCTA's default diagonal (second_matrix = GGRM_SECOND_NONMISSING_GCTA_DIAG) and the GCTA ratio read-out
as rectangular blocks (what the 2D-tile writer stores), through the C ABI.

There is no reference code for this option (the reference's docs name the GCTA format, docs/source/data_formats.rst:188-208,
its code never writes it): the checker is a numpy evaluation of the formula GCTA publishes for --make-grm (Yang et al. 2011,
AJHG 88:76, eq. 3-4), so parity of the DIAGONAL is "unpinned".  Off-diagonal values and pair counts are compared byte for byte
with the existing GGRM_SECOND_NONMISSING read-out, which is pinned (test_gpu_parity.py::test_gcta_ratio_and_pair_counts)."""
import numpy as np
import pytest

import gelex_b200 as g
from gelex_b200 import _lib
import plink_io

pytestmark = pytest.mark.gpu


def gcta_diag_numpy(geno):
    """A_jj = 1 + (1 / N_jj) sum_i (x_ij^2 - (1 + 2 p_i) x_ij + 2 p_i^2) / (2 p_i (1 - p_i)) over the SNPs called in sample j;
    geno (n, m) of 0/1/2/NaN counts of A1, p_i the A1 frequency among called samples; SNPs without variation are skipped in the sum
    (their standardised values are 0 everywhere) but still count in N, like in the off-diagonal entries."""
    called = ~np.isnan(geno)
    x = np.where(called, geno, 0.0)
    cnt = called.sum(axis=0)
    p = np.divide(x.sum(axis=0), 2.0 * cnt, out=np.zeros(geno.shape[1]), where=cnt > 0)
    den = 2.0 * p * (1.0 - p)
    ok = den > 0
    term = np.zeros_like(x)
    term[:, ok] = (x[:, ok] ** 2 - (1.0 + 2.0 * p[ok]) * x[:, ok] + 2.0 * p[ok] ** 2) / den[ok]
    term[~called] = 0.0
    nj = called.sum(axis=1).astype(np.float64)
    return np.where(nj > 0, 1.0 + np.divide(term.sum(axis=1), nj, out=np.zeros_like(nj), where=nj > 0), 0.0)


def make_case(n, m, seed, miss, all_missing=True):
    geno = plink_io.decode_payload(g.synth_bed_host(n, 0, m, seed, miss), n)
    geno[:, 7] = 2.0                 # no variation: skipped by both sums
    if all_missing:                  # an all-missing SNP: needs GGRM_ALLMISS_SKIP (the reference's policy turns the whole matrix into NaN)
        geno[:, 11] = np.nan
    geno[3, :] = np.nan              # a sample without any call: N_33 = 0 -> A_33 = 0
    return geno, plink_io.encode_matrix(geno)


def diag_index(n):
    i = np.arange(n, dtype=np.int64)
    return i * (i + 1) // 2 + i


def run(bed, n, m, prec, second, push=333, **kw):
    acc = g.Accumulator(n, m, method="SH", want_add=True, want_dom=True, precision=prec, second_matrix=second,
                        all_missing_policy=_lib.ALLMISS_SKIP, **kw)
    for s in range(0, m, push):
        acc.push(bed[s:s + push], s)
    acc.finish()
    return acc


@pytest.mark.parametrize("prec", [g.PREC_FP64, g.PREC_FP16, g.PREC_BF16])
def test_gcta_default_diagonal_matches_published_formula(prec):
    n, m = 515, 900
    geno, bed = make_case(n, m, 31, 0.08)
    want = gcta_diag_numpy(geno)
    with run(bed, n, m, prec, _lib.SECOND_NONMISSING, chunk_snps=256) as acc:
        base_a, base_n = acc.fetch_gcta(0), acc.fetch_gcta(1)
    with run(bed, n, m, prec, _lib.SECOND_NONMISSING_GCTA_DIAG, chunk_snps=256) as acc:
        got_a, got_n = acc.fetch_gcta(0), acc.fetch_gcta(1)
        part = acc.fetch_gcta(0, row_begin=200, row_count=57)
        blk = acc.fetch_block_gcta(0, g.GROUP_WHOLE, 100, 57, 31, 400)   # a block crossing the diagonal
        blk_n = acc.fetch_block_gcta(1, g.GROUP_WHOLE, 100, 57, 31, 400)
        acc.reset()                                                    # a second pass starts its sums from zero
        for s in range(0, m, 450):
            acc.push(bed[s:s + 450], s)
        acc.finish()
        again = acc.fetch_gcta(0)
    d = diag_index(n)
    off = np.ones(got_a.size, dtype=bool)
    off[d] = False
    assert np.array_equal(got_n, base_n)
    assert np.array_equal(got_a[off], base_a[off])                     # only the diagonal differs from the plain ratio
    # the diagonal sums are FP64 in every precision (they never pass through the 16-bit operands): float32(formula) to one ulp
    ref32 = want.astype(np.float32)
    assert np.all(np.abs(got_a[d] - ref32) <= np.spacing(np.abs(ref32))), np.max(np.abs(got_a[d] - ref32))
    assert got_a[d[3]] == 0.0 and np.all(np.isfinite(got_a))
    assert np.array_equal(part, got_a[g.tri(200):g.tri(257)])
    assert np.array_equal(again[d], got_a[d])                          # a sample's terms are added in SNP order whatever the chunking
    dense = np.zeros((n, n), dtype=np.float32)
    dense[np.tril_indices(n)] = got_a
    assert np.array_equal(blk, dense[100:157, 31:431])
    dn = np.zeros((n, n), dtype=np.float32)
    dn[np.tril_indices(n)] = got_n
    assert np.array_equal(blk_n, dn[100:157, 31:431])


def test_gcta_diagonal_with_dropped_and_reordered_samples():
    # the sums are taken over GRM rows (after the FAM -> sample-order projection), not over BED columns
    n_raw, m = 700, 400
    rng = np.random.default_rng(5)
    raw_to_out = rng.permutation(n_raw).astype(np.int32)
    raw_to_out[raw_to_out >= 650] = -1                                 # 50 samples dropped
    n_out = 650
    geno_raw = plink_io.decode_payload(g.synth_bed_host(n_raw, 0, m, 77, 0.05), n_raw)
    bed = plink_io.encode_matrix(geno_raw)
    geno = np.full((n_out, m), np.nan)
    keep = raw_to_out >= 0
    geno[raw_to_out[keep]] = geno_raw[keep]
    want = gcta_diag_numpy(geno).astype(np.float32)
    with g.Accumulator(n_raw, m, method="SH", want_add=True, want_dom=True, precision=g.PREC_FP16, second_matrix=_lib.SECOND_NONMISSING_GCTA_DIAG,
                       raw_to_out=raw_to_out, n_out=n_out) as acc:
        acc.push(bed, 0)
        acc.finish()
        got = acc.fetch_gcta(0)[diag_index(n_out)]
    assert np.all(np.abs(got - want) <= np.spacing(np.abs(want)))


@pytest.mark.parametrize("second", [_lib.SECOND_NONMISSING, _lib.SECOND_NONMISSING_GCTA_DIAG])
def test_gcta_blocks_of_tile_ranks_reproduce_the_packed_rows(second):
    # 2D tiles: every rank streams all SNPs, so its diagonal sums are complete; the owned super-tiles of all ranks, read as GCTA
    # ratio blocks, assemble to the bytes one handle returns as packed rows
    n, m = 4300, 500   # 3 super-rows of 2048
    geno, bed = make_case(n, m, 83, 0.03)
    with run(bed, n, m, g.PREC_FP16, second) as one:
        want_a, want_n = one.fetch_gcta(0), one.fetch_gcta(1)
    got_a = np.zeros((n, n), dtype=np.float32)
    got_n = np.zeros((n, n), dtype=np.float32)
    for r in range(3):
        with run(bed, n, m, g.PREC_FP16, second, tile_rank=r, tile_world=3) as acc:
            for si, sj in acc.owned_supertiles():
                r0, c0 = si * 2048, sj * 2048
                rows, cols = min(2048, n - r0), min(2048, n - c0)
                got_a[r0:r0 + rows, c0:c0 + cols] += acc.fetch_block_gcta(0, g.GROUP_WHOLE, r0, rows, c0, cols)
                got_n[r0:r0 + rows, c0:c0 + cols] += acc.fetch_block_gcta(1, g.GROUP_WHOLE, r0, rows, c0, cols)
    assert np.array_equal(got_n[np.tril_indices(n)], want_n)
    assert np.array_equal(got_a[np.tril_indices(n)], want_a)
    assert not np.triu(got_a, 1).any()


def test_gcta_diagonal_of_snp_shards_is_reduced_over_peers():
    # SNP shards: each handle holds the diagonal sums of its own SNPs; ggrm_gcta_diag_reduce_async adds the peers' (here two
    # handles on one device, whose addresses are each other's "peer memory") before the read-out
    n, m = 2301, 1200
    geno, bed = make_case(n, m, 41, 0.05)
    want = gcta_diag_numpy(geno).astype(np.float32)
    kw = dict(method="SH", want_add=True, want_dom=True, precision=g.PREC_FP64, second_matrix=_lib.SECOND_NONMISSING_GCTA_DIAG,
              all_missing_policy=_lib.ALLMISS_SKIP)
    with g.Accumulator(n, m, **kw) as a, g.Accumulator(n, m, **kw) as b, run(bed, n, m, g.PREC_FP64, _lib.SECOND_NONMISSING_GCTA_DIAG) as one:
        a.push(bed[:500], 0)
        b.push(bed[500:], 500)
        a.finish()
        b.finish()
        for kind in (0, 1):
            pb, nb, _ = b.accum_buffer(kind, g.GROUP_WHOLE)
            a.reduce_peer(kind, g.GROUP_WHOLE, [pb], 0, nb)
        before = a.fetch_gcta(0)[diag_index(n)]        # own sums only: not the answer yet
        a.gcta_diag_reduce_async([b.gcta_diag_address()])
        got = a.fetch_gcta(0)
        ref = one.fetch_gcta(0)
    d = diag_index(n)
    assert np.max(np.abs(before - want)) > 1e-3
    assert np.all(np.abs(got[d] - want) <= np.spacing(np.abs(want)))
    assert np.all(np.abs(got - ref) <= 2 * np.spacing(np.abs(ref)) + 1e-12)   # FP64 sums in another order, rounded to float32


def test_gcta_diag_option_is_validated():
    with pytest.raises(g.GrmError, match="needs method SH"):
        g.Accumulator(100, 50, method="OSH", want_add=True, want_dom=True, second_matrix=_lib.SECOND_NONMISSING_GCTA_DIAG)
    with pytest.raises(g.GrmError, match="invalid second_matrix"):
        g.Accumulator(100, 50, method="SH", want_add=True, want_dom=True, second_matrix=3)
    with g.Accumulator(100, 50, method="SH", want_add=True, want_dom=True, second_matrix=_lib.SECOND_NONMISSING) as acc:
        with pytest.raises(g.GrmError, match="GCTA_DIAG"):
            acc.gcta_diag_address()


def test_engine_gcta_diag_files_one_gpu_snp_shards_and_tiles(tmp_path, monkeypatch):
    # GrmEngine / CLI level: --gcta --gcta-diag on one GPU, on two SNP shards (the diagonal sums are reduced over peer memory beside
    # the two matrices) and on 2D tiles (GCTA files were refused there before); FAM order differs from the sample order
    import subprocess

    from gelex_b200 import engine

    monkeypatch.setenv("GELEX_B200_OVERSUBSCRIBE", "1")
    n, m = 4301, 900   # three super-rows
    rng = np.random.default_rng(3)
    perm = rng.permutation(n)
    fam = [(f"fam{i % 7}", f"s{perm[i]:05d}") for i in range(n)]
    ids, r2o = plink_io.sample_order(fam)
    geno_raw, bed = make_case(n, m, 41, 0.05, all_missing=False)
    plink_io.write_plink(tmp_path / "d", bed, fam, ["1"] * m)
    geno = np.empty_like(geno_raw)
    geno[np.asarray(r2o)] = geno_raw
    want = gcta_diag_numpy(geno).astype(np.float32)
    d = diag_index(n)
    engine.grm_run_gcta(tmp_path / "d", tmp_path / "p1", method="SH", precision=g.PREC_FP16, chunk_size=400)                      # sum z^2 / N
    engine.grm_run_gcta(tmp_path / "d", tmp_path / "g1", method="SH", precision=g.PREC_FP16, chunk_size=400, gcta_diag=True)
    plain, got = plink_io.read_bin(tmp_path / "p1.grm.bin", n), plink_io.read_bin(tmp_path / "g1.grm.bin", n)
    off = np.ones(got.size, dtype=bool)
    off[d] = False
    assert np.array_equal(got[off], plain[off])
    assert np.all(np.abs(got[d] - want) <= np.spacing(np.abs(want)))
    assert (tmp_path / "g1.grm.N.bin").read_bytes() == (tmp_path / "p1.grm.N.bin").read_bytes()
    # 2D tiles: same bytes as one GPU, with and without the diagonal option
    engine.grm_run_gcta(tmp_path / "d", tmp_path / "t1", method="SH", precision=g.PREC_FP16, chunk_size=400, n_gpus=3, gcta_diag=True, partition="tile")
    engine.grm_run_gcta(tmp_path / "d", tmp_path / "t0", method="SH", precision=g.PREC_FP16, chunk_size=400, n_gpus=2, partition="tile")
    for a, b in (("t1", "g1"), ("t0", "p1")):
        for suffix in (".grm.bin", ".grm.N.bin", ".grm.id"):
            assert (tmp_path / (a + suffix)).read_bytes() == (tmp_path / (b + suffix)).read_bytes(), (a, suffix)
    # SNP shards, band-pipelined (16-bit) and blocking (FP64) reduce
    engine.grm_run_gcta(tmp_path / "d", tmp_path / "s1", method="SH", precision=g.PREC_FP16, chunk_size=400, n_gpus=2, gcta_diag=True)
    s1 = plink_io.read_bin(tmp_path / "s1.grm.bin", n)
    assert np.all(np.abs(s1[d] - want) <= np.spacing(np.abs(want)))
    assert (tmp_path / "s1.grm.N.bin").read_bytes() == (tmp_path / "g1.grm.N.bin").read_bytes()
    assert np.max(np.abs(s1[off].astype(np.float64) - got[off])) <= 1e-5      # FP32 partial sums added in another order
    engine.grm_run_gcta(tmp_path / "d", tmp_path / "s2", method="SH", precision=g.PREC_FP64, chunk_size=400, n_gpus=3, gcta_diag=True)
    s2 = plink_io.read_bin(tmp_path / "s2.grm.bin", n)
    assert np.all(np.abs(s2[d] - want) <= np.spacing(np.abs(want)))
    # CLI flag and its validation
    r = subprocess.run([str(engine.CLI_PATH), "grm", "-b", str(tmp_path / "d"), "--add", "--gcta", "--gcta-diag", "--geno-method", "SH", "--precision", "fp16", "-c", "400",
                        "-o", str(tmp_path / "c1")], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    assert (tmp_path / "c1.grm.bin").read_bytes() == (tmp_path / "g1.grm.bin").read_bytes()
    r = subprocess.run([str(engine.CLI_PATH), "grm", "-b", str(tmp_path / "d"), "--add", "--gcta", "--gcta-diag", "--geno-method", "OSH", "-o", str(tmp_path / "c2")],
                       capture_output=True, text=True)
    assert r.returncode == 1 and "--gcta-diag is GCTA's diagonal" in (r.stdout + r.stderr)
