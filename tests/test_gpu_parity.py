"""GPU parity tests (run with `-m gpu` on a B200).  Everything goes through the C ABI (include/ggrm.h) and is checked
against the CPU oracle (oracle/grm_oracle.c, itself pinned to the reference) and the committed golden fixtures.

Tolerances (north star): decode, counts, allele frequencies bit-exact; GRM entries on G/(trace/n):
  FP64 path <= 1e-10;  16-bit/FP32 path <= 1e-4, at every size tested here - nothing is scaled.  FP16 operands carry 11
  mantissa bits; BF16 (8 bits) is split into hi + lo below 2^18 SNPs (ggrm_api.cu, Handle::bf16_split), which is what keeps it
  inside the bound on short inputs.  tests/test_named_configs_gpu.py holds the named-shape checks against the reference.
"""
import os

import numpy as np
import pytest

import gelex_b200 as g
import oracle_lib as o
import plink_io
from test_oracle_golden import load_dense, load_tiny, same, close

pytestmark = pytest.mark.gpu
TOL64 = 1e-10


TOL16 = 1e-4


def norm_packed(packed):
    packed = np.asarray(packed, dtype=np.float64)
    n = int((np.sqrt(8 * packed.size + 1) - 1) / 2)
    diag = np.cumsum(np.arange(1, n + 1)) - 1
    return packed / (packed[diag].sum() / n)


def norm_ref(G):
    n = G.shape[0]
    return G[np.tril_indices(n)] / (np.trace(G) / n)


def run_gpu(payload, n_raw, m, method, kinds, precision, chunk=0, push=997, **kw):
    acc = g.Accumulator(n_raw, m, method=method, want_add=0 in kinds, want_dom=1 in kinds, precision=precision,
                        chunk_snps=chunk, keep_stats=True, **kw)
    for s in range(0, m, push):
        acc.push(payload[s:min(m, s + push)], s)
    acc.finish()
    return acc


# ------------------------------------------------------------------------------------------------ bit-exact layers
def test_decode_bit_exact_dense_and_sparse():
    rng = np.random.default_rng(0)
    for n in (1000, 1001, 1003, 257):  # n % 4 != 0 exercises the tail byte
        bed = g.synth_bed_host(n, 0, 200, 11, 0.03)
        assert same(g.decode_chunk(bed, n), o.decode_chunk(bed, n))
        perm = rng.permutation(n)
        fam = [(f"fam{(i % 5) + 1}", f"s{perm[i]}") for i in range(n)]
        fam[7] = fam[3]  # duplicated ID: later raw sample wins
        ids, r2o = plink_io.sample_order(fam)
        assert same(g.decode_chunk(bed, n, len(ids), r2o), o.decode_chunk(bed, n, len(ids), r2o))


def test_decode_with_dropped_samples():
    n = 500
    bed = g.synth_bed_host(n, 0, 64, 5, 0.05)
    keep = np.arange(n) % 3 != 0
    r2o = np.full(n, -1, dtype=np.int32)
    r2o[keep] = np.random.default_rng(1).permutation(int(keep.sum())).astype(np.int32)
    n_out = int(keep.sum())
    assert same(g.decode_chunk(bed, n, n_out, r2o), o.decode_chunk(bed, n, n_out, r2o))


def test_decode_matches_reference_golden(golden_dir):
    payload, n_raw, ids, r2o, _ = load_tiny(golden_dir)
    gold = np.load(golden_dir / "tiny_chunks.npz")
    assert same(g.decode_chunk(payload, n_raw, len(ids), r2o), gold["raw"])
    dpayload, dn, _ = load_dense(golden_dir)
    assert same(g.decode_chunk(dpayload, dn), np.load(golden_dir / "dense_chunks.npz")["raw"])


@pytest.mark.parametrize("method", range(1, 9))
@pytest.mark.parametrize("kind", ["A", "D"])
def test_standardise_matches_reference_golden(golden_dir, method, kind):
    payload, n_raw, ids, r2o, _ = load_tiny(golden_dir)
    gold = np.load(golden_dir / "tiny_chunks.npz")
    z, freq = g.standardise_chunk(payload, n_raw, method, 0 if kind == "A" else 1, n_out=len(ids), raw_to_out=r2o)
    assert np.array_equal(freq, gold[f"freq_{method}_{kind}"])  # bit-exact allele frequencies
    assert close(z, gold[f"z_{method}_{kind}"], rtol=1e-12, atol=1e-13)


def test_standardise_vs_oracle_random():
    n = 1537
    bed = g.synth_bed_host(n, 0, 300, 3, 0.04)
    raw = o.decode_chunk(bed, n)
    for method in range(1, 9):
        for kind in (0, 1):
            z, f = g.standardise_chunk(bed, n, method, kind)
            zr, fr = o.process_matrix(raw, method, kind)
            assert np.array_equal(f, fr)
            assert close(z, zr, rtol=1e-12, atol=1e-13)


def test_assoc_chunk_preprocessing_center_family_only(golden_dir):
    # the GWAS scan's chunk pre-processing (assoc_detail.cpp:28-47) reuses the decode + policy kernels; only the
    # centre-family methods are accepted, with the reference's message otherwise
    n = 1201
    fam = [(f"f{i % 3}", f"s{(i * 7919) % n}") for i in range(n)]  # shuffled IDs: the projection is part of the stage
    ids, r2o = plink_io.sample_order(fam)
    bed = g.synth_bed_host(n, 0, 150, 13, 0.03)
    raw = o.decode_chunk(bed, n, len(ids), r2o)
    for method in (2, 4, 6, 8):
        for kind in (0, 1):
            z, f = g.assoc_chunk(bed, n, method, kind, n_out=len(ids), raw_to_out=r2o)
            zr, fr = o.process_matrix(raw, method, kind)
            assert np.array_equal(f, fr)  # A1 frequency written to the .gwas.tsv: bit-exact
            assert close(z, zr, rtol=1e-12, atol=1e-13)
    d = np.load(golden_dir / "tiny_chunks.npz")  # outputs of the unmodified reference's process_matrix
    bedt, n_raw_t, ids_t, r2o_t, _ = load_tiny(golden_dir)
    z, f = g.assoc_chunk(bedt, n_raw_t, 4, 1, n_out=len(ids_t), raw_to_out=r2o_t)
    assert np.array_equal(f, d["freq_4_D"]) and close(z, d["z_4_D"], rtol=1e-12, atol=1e-13)
    for method in (1, 3, 5, 7):
        with pytest.raises(g.GrmError) as e:
            g.assoc_chunk(bed, n, method, 0, n_out=len(ids), raw_to_out=r2o)
        assert e.value.status == -1 and "supports only center-family methods" in e.value.message


def test_genotype_chunk_locus_statistics(golden_dir):
    # the standardised chunk as the product (GenotypeLoader / GenotypePipe::process_chunk) with every SNP's
    # LocusStatistic.  (1) the columns of the reference's own unit tests, outputs of the unmodified reference
    import json

    vectors = json.load(open(golden_dir / "process_vectors.json"))
    for v in vectors:
        col = np.array([float("nan") if x == "nan" else float(x) for x in v["input"]], dtype=np.float64)
        ref = np.array([float("nan") if x == "nan" else x for x in v["z"]], dtype=np.float64)
        z, st = g.genotype_chunk(plink_io.encode_matrix(col[:, None]), col.size, v["method"], 0 if v["kind"] == "A" else 1)
        assert st["freq"][0] == v["maf"], v
        assert close(z[:, 0], ref, rtol=1e-12, atol=1e-13), v
        assert bool(st["mono"][0]) == bool(v["mono"]), v
        assert abs(st["mean"][0] - v["mean"]) <= 1e-14 * abs(v["mean"]), v
        assert abs(st["stddev"][0] - v["stddev"]) <= 1e-12 * abs(v["stddev"]) + 1e-300, v
    # (2) random SNPs through a shuffled projection + monomorphic / all-missing / single-valid columns, against the oracle
    n = 777
    fam = [(f"f{i % 4}", f"s{(i * 5011) % n}") for i in range(n)]
    ids, r2o = plink_io.sample_order(fam)
    bed = g.synth_bed_host(n, 0, 96, 17, 0.05)
    special = np.full((n, 4), 2.0)
    special[:, 1] = np.nan           # all missing: default LocusStatistic, NaN column
    special[:, 2] = 0.0              # monomorphic
    special[5:, 3] = np.nan          # five valid calls only
    special[:5, 3] = [0, 1, 2, 1, 1]
    bed = np.concatenate([bed, plink_io.encode_matrix(special)])
    raw = o.decode_chunk(bed, n, len(ids), r2o)
    for method in range(1, 9):
        for kind in (0, 1):
            z, st = g.genotype_chunk(bed, n, method, kind, n_out=len(ids), raw_to_out=r2o)
            zr, fr = o.process_matrix(raw, method, kind)
            assert np.array_equal(st["freq"], fr)
            assert close(z, zr, rtol=1e-12, atol=1e-13)
            for j in range(raw.shape[1]):
                _, ls = o.process_column(raw[:, j], method, kind)
                assert bool(st["mono"][j]) == bool(ls.is_monomorphic), (method, kind, j)
                assert close(st["mean"][j], ls.mean, rtol=1e-13, atol=0), (method, kind, j)  # sample-policy means are sums over the column
                assert close(st["stddev"][j], ls.stddev, rtol=1e-12, atol=1e-300), (method, kind, j)
    with pytest.raises(g.GrmError) as e:
        g.genotype_chunk(bed, n, 9, 0, n_out=len(ids), raw_to_out=r2o)
    assert "Invalid genotype process method." in e.value.message


def test_all_missing_snp_policy():
    n = 64
    geno = np.random.default_rng(2).integers(0, 3, size=(n, 6)).astype(np.float64)
    geno[:, 2] = np.nan
    bed = plink_io.encode_matrix(geno)
    z, f = g.standardise_chunk(bed, n, 1, 0)  # reference behaviour: the column stays NaN
    zr, fr = o.process_matrix(o.decode_chunk(bed, n), 1, 0)
    assert same(np.isnan(z), np.isnan(zr)) and np.all(np.isnan(z[:, 2])) and f[2] == fr[2] == 0.0
    z2, _ = g.standardise_chunk(bed, n, 1, 0, all_missing_policy=g._lib.ALLMISS_SKIP)
    assert np.all(z2[:, 2] == 0.0)


def test_snp_counts_and_freq_bit_exact():
    n, m = 2001, 1500
    bed = g.synth_bed_host(n, 0, m, 21, 0.02)
    acc = run_gpu(bed, n, m, 1, (0,), g.PREC_FP16)
    n0, n1, n2, nm, f = acc.snp_stats(0, m)
    acc.close()
    raw = o.decode_chunk(bed, n)
    assert np.array_equal(nm, np.isnan(raw).sum(axis=0))
    assert np.array_equal(n0, (raw == 0).sum(axis=0)) and np.array_equal(n1, (raw == 1).sum(axis=0))
    assert np.array_equal(n2, (raw == 2).sum(axis=0))
    _, fr = o.process_matrix(raw, 1, 0)
    assert np.array_equal(f, fr)


# ------------------------------------------------------------------------------------------------ GRM entries
@pytest.mark.parametrize("method,kinds", [(1, (0,)), (3, (0, 1)), (4, (0, 1)), (5, (1,)), (8, (0, 1)), (2, (1,))])
def test_grm_fp64_vs_oracle(method, kinds):
    n, m = 777, 2100
    bed = g.synth_bed_host(n, 0, m, 31, 0.02)
    acc = run_gpu(bed, n, m, method, kinds, g.PREC_FP64, chunk=640)
    for kind in kinds:
        packed, tr = acc.fetch(kind, g.GROUP_WHOLE, dtype=np.float64)
        G, den, _ = o.grm_compute(bed, n, method, kind, chunk=500)
        assert abs(tr / n - den) <= 1e-9 * abs(den)
        assert np.max(np.abs(norm_packed(packed) - norm_ref(G))) <= TOL64
        # on-disk float32 view: identical after rounding or within 1 float32 ulp
        f32, _ = acc.fetch(kind, g.GROUP_WHOLE, dtype=np.float32)
        ref32 = o.pack_lower_f32(G)
        assert np.all(np.abs(f32 - ref32) <= np.spacing(np.abs(ref32)))
    acc.close()


@pytest.mark.parametrize("precision", [g.PREC_BF16, g.PREC_FP16])
@pytest.mark.parametrize("method,kinds", [(1, (0,)), (3, (0, 1)), (4, (0, 1)), (6, (0, 1))])
def test_grm_16bit_vs_oracle(precision, method, kinds):
    n, m = 1100, 6000
    bed = g.synth_bed_host(n, 0, m, 41, 0.01)
    acc = run_gpu(bed, n, m, method, kinds, precision, chunk=2048)
    for kind in kinds:
        packed, tr = acc.fetch(kind, g.GROUP_WHOLE)
        G, den, _ = o.grm_compute(bed, n, method, kind, chunk=2000)
        err = np.max(np.abs(norm_packed(packed) - norm_ref(G)))
        assert err <= TOL16, err
        assert abs(tr / n - den) <= 2e-3 * abs(den)
    acc.close()


def test_chunk_size_and_push_size_invariance_fp64():
    # reference tests/test_grm.cpp:187-225 (chunk 1 / 7 / m / m+100 within 1e-10) restated for device chunks and pushes
    n, m = 300, 700
    bed = g.synth_bed_host(n, 0, m, 51, 0.02)
    base = None
    for chunk, push in ((64, 1), (448, 7), (0, m), (1024, 333)):
        acc = run_gpu(bed, n, m, 3, (0,), g.PREC_FP64, chunk=chunk, push=push)
        packed, _ = acc.fetch(0, g.GROUP_WHOLE, dtype=np.float64)
        acc.close()
        if base is None:
            base = norm_packed(packed)
        assert np.max(np.abs(norm_packed(packed) - base)) <= TOL64


def test_tc_single_cta_and_pair_kernels_agree(monkeypatch):
    n, m = 1300, 4000
    bed = g.synth_bed_host(n, 0, m, 61, 0.01)
    outs = []
    for mode in ("1", "2"):
        monkeypatch.setenv("GGRM_TC_MODE", mode)
        acc = run_gpu(bed, n, m, 1, (0,), g.PREC_FP16, chunk=1024)
        outs.append(acc.fetch(0, g.GROUP_WHOLE)[0])
        acc.close()
    # same operands; the pair kernel may walk K in another order (K rotation / split-K on few-tile problems), so the two agree to
    # FP32 rounding of a 4 000-term accumulation, not bit for bit (measured 1.1e-6)
    assert np.max(np.abs(norm_packed(outs[0]) - norm_packed(outs[1]))) <= 3e-6


def test_projection_row_order_and_golden_bin(golden_dir):
    # shuffled FAM with a duplicated ID: rows must come out in sorted-ID order
    payload, n_raw, ids, r2o, chroms = load_tiny(golden_dir)
    n, m = len(ids), len(chroms)
    for tag, method, kind in (("SH.add", 1, 0), ("OSH.dom", 3, 1), ("OCH.add", 4, 0), ("OCH.dom", 4, 1)):
        acc = run_gpu(payload, n_raw, m, method, (kind,), g.PREC_FP64, n_out=n, raw_to_out=r2o, push=13)
        got, _ = acc.fetch(kind, g.GROUP_WHOLE, dtype=np.float32)
        acc.close()
        ref = plink_io.read_bin(golden_dir / f"tiny_{tag}.bin", n)
        assert np.all(np.abs(got - ref) <= np.spacing(np.abs(ref)) + 1e-30), tag


def test_reference_nan_poisoning_is_reproduced(golden_dir):
    # tiny SNP 11 has one valid call: Standardize (sample sd, valid-1 = 0) yields NaN and poisons the matrix in the
    # reference (tiny_S.*.bin are all-NaN); the default policy reproduces that bit for bit.
    payload, n_raw, ids, r2o, chroms = load_tiny(golden_dir)
    n, m = len(ids), len(chroms)
    acc = run_gpu(payload, n_raw, m, 5, (0, 1), g.PREC_FP64, n_out=n, raw_to_out=r2o)
    for kind, tag in ((0, "S.add"), (1, "S.dom")):
        got, _ = acc.fetch(kind, g.GROUP_WHOLE, dtype=np.float32)
        ref = plink_io.read_bin(golden_dir / f"tiny_{tag}.bin", n)
        assert np.array_equal(np.isnan(got), np.isnan(ref))
    acc.close()


def test_loco_groups_whole_and_subtraction(golden_dir):
    payload, n_raw, ids, r2o, chroms = load_tiny(golden_dir)
    n, m = len(ids), len(chroms)
    runs = plink_io.chrom_runs(chroms)
    groups = [(b, e) for _, b, e in runs]
    acc = run_gpu(payload, n_raw, m, 4, (0, 1), g.PREC_FP64, n_out=n, raw_to_out=r2o, groups=groups, keep_whole=True, push=10)
    gold = np.load(golden_dir / "tiny_loaders.npz")
    for kind, kname in ((0, "add"), (1, "dom")):
        whole, _ = acc.fetch(kind, g.GROUP_WHOLE, dtype=np.float32)
        ref_w = plink_io.read_bin(golden_dir / f"tiny_OCH.{kname}.bin", n)
        assert np.all(np.abs(whole - ref_w) <= 2 * np.spacing(np.abs(ref_w)) + 1e-6)  # whole = sum of partials
        for gi, (name, _, _) in enumerate(runs):
            part, _ = acc.fetch(kind, gi, dtype=np.float32)
            ref = plink_io.read_bin(golden_dir / f"tiny_OCH.{kname}.chr{name}.bin", n)
            assert np.all(np.abs(part - ref) <= np.spacing(np.abs(ref)) + 1e-30), (kname, name)
    # consumer-side arithmetic on device: (G_w - G_c)/(k_w - k_c) as LocoGRMLoader computes it
    for gi, (name, _, _) in enumerate(runs):
        loco = acc.fetch_loco(0, gi)
        ref = gold[f"OCH_add_loco_chr{name}"][np.tril_indices(n)]
        assert np.allclose(loco, ref, rtol=1e-5, atol=1e-6)  # the reference works from float32 files
    acc.close()
    acc = run_gpu(payload, n_raw, m, 1, (0,), g.PREC_FP64, n_out=n, raw_to_out=r2o)
    nrm = acc.fetch_normalized(0, g.GROUP_WHOLE)
    assert np.allclose(nrm, gold["SH_add_normalized"][np.tril_indices(n)], rtol=1e-6, atol=1e-7)
    acc.close()


def test_group_pool_release_and_row_range_fetch():
    n, m = 520, 900
    bed = g.synth_bed_host(n, 0, m, 71, 0.01)
    groups = [(0, 200), (200, 450), (450, 900)]
    acc = g.Accumulator(n, m, method=1, precision=g.PREC_FP64, groups=groups, keep_whole=True, max_resident_groups=1)
    fetched = []
    for gi, (b, e) in enumerate(groups):
        acc.push(bed[b:e], b)
        if gi > 0:  # previous group is complete as soon as a later group has been pushed
            fetched.append(acc.fetch(0, gi - 1, dtype=np.float64)[0])
            acc.release_group(0, gi - 1)
    acc.finish()
    fetched.append(acc.fetch(0, 2, dtype=np.float64)[0])
    whole, _ = acc.fetch(0, g.GROUP_WHOLE, dtype=np.float64)
    assert np.allclose(whole, sum(fetched), rtol=1e-12, atol=1e-9)
    for gi, (b, e) in enumerate(groups):
        G, _, _ = o.grm_compute(bed, n, 1, 0, ranges=[(b, e)])
        assert np.allclose(fetched[gi], G[np.tril_indices(n)], rtol=1e-10, atol=1e-9)
    # row-range fetch returns exactly the corresponding slice of the packed file
    part, tr = acc.fetch(0, g.GROUP_WHOLE, row_begin=100, row_count=57, dtype=np.float64)
    assert np.array_equal(part, whole[g.tri(100):g.tri(157)])
    diag = np.cumsum(np.arange(1, n + 1)) - 1
    assert abs(tr - whole[diag[100:157]].sum()) <= 1e-9 * abs(tr)
    acc.close()


def test_tile_partition_and_snp_shard_recombine():
    # multi-GPU decompositions emulated on one device: tile ranks own disjoint tiles (sum = whole);
    # SNP shards hold partial sums (sum = whole); peer reduce kernel sums accumulators in place.
    n, m = 2300, 1200   # 9 x 9 tiles of 256 -> 3 super-tiles
    bed = g.synth_bed_host(n, 0, m, 81, 0.01)
    full = run_gpu(bed, n, m, 1, (0,), g.PREC_FP64)
    whole, _ = full.fetch(0, g.GROUP_WHOLE, dtype=np.float64)
    parts = []
    for r in range(3):
        acc = run_gpu(bed, n, m, 1, (0,), g.PREC_FP64, tile_rank=r, tile_world=3)
        parts.append(acc.fetch(0, g.GROUP_WHOLE, dtype=np.float64)[0])
        acc.close()
    assert all(np.count_nonzero(p) > 0 for p in parts)
    assert np.array_equal(sum(parts), whole)  # every entry is owned by exactly one rank
    a = g.Accumulator(n, m, method=1, precision=g.PREC_FP64)
    b = g.Accumulator(n, m, method=1, precision=g.PREC_FP64)
    a.push(bed[:500], 0)
    b.push(bed[500:], 500)
    a.finish()
    b.finish()
    pa, na, _ = a.accum_buffer(0, g.GROUP_WHOLE)
    pb, nb, _ = b.accum_buffer(0, g.GROUP_WHOLE)
    assert na == nb
    a.reduce_peer(0, g.GROUP_WHOLE, [pb], 0, na)
    summed, _ = a.fetch(0, g.GROUP_WHOLE, dtype=np.float64)
    assert np.allclose(summed, whole, rtol=1e-12, atol=1e-9)
    for x in (a, b, full):
        x.close()


def test_fetch_block_and_owned_supertiles():
    # rectangular read-out used by the 2D-tile writer: blocks agree with the packed rows, zeros above the diagonal and in
    # tiles of other ranks; the owned super-tile lists of all ranks partition the triangle
    n, m = 4300, 700   # 3 super-rows of 2048
    bed = g.synth_bed_host(n, 0, m, 83, 0.01)
    full = run_gpu(bed, n, m, 1, (0,), g.PREC_FP64)
    whole, _ = full.fetch(0, g.GROUP_WHOLE, dtype=np.float64)
    dense = np.zeros((n, n))
    dense[np.tril_indices(n)] = whole
    assert full.owned_supertiles() == [(0, 0), (1, 0), (1, 1), (2, 0), (2, 1), (2, 2)]
    blk = full.fetch_block(0, g.GROUP_WHOLE, 2048, 2048, 0, 2048)
    assert np.array_equal(blk, dense[2048:4096, 0:2048].astype(np.float32))
    blk = full.fetch_block(0, g.GROUP_WHOLE, 4096, n - 4096, 4096, n - 4096)  # ragged diagonal super-tile
    assert np.array_equal(blk, dense[4096:, 4096:].astype(np.float32))
    blk = full.fetch_block(0, g.GROUP_WHOLE, 100, 57, 31, 400)  # arbitrary block crossing the diagonal
    assert np.array_equal(blk, dense[100:157, 31:431].astype(np.float32))
    full.close()
    seen = []
    acc16 = np.zeros((n, n), dtype=np.float32)
    for r in range(2):
        acc = run_gpu(bed, n, m, 1, (0,), g.PREC_FP16, tile_rank=r, tile_world=2)
        mine = acc.owned_supertiles()
        assert mine, "every rank owns something"
        seen += mine
        for si, sj in mine:
            r0, c0 = si * 2048, sj * 2048
            rows, cols = min(2048, n - r0), min(2048, n - c0)
            acc16[r0:r0 + rows, c0:c0 + cols] += acc.fetch_block(0, g.GROUP_WHOLE, r0, rows, c0, cols)
        # a block of a super-tile this rank does NOT own reads as zeros
        other = next(t for t in [(0, 0), (1, 0), (1, 1), (2, 0), (2, 1), (2, 2)] if t not in mine)
        z = acc.fetch_block(0, g.GROUP_WHOLE, other[0] * 2048, min(2048, n - other[0] * 2048), other[1] * 2048, min(2048, n - other[1] * 2048))
        assert not z.any()
        acc.close()
    assert sorted(seen) == [(0, 0), (1, 0), (1, 1), (2, 0), (2, 1), (2, 2)]
    one = run_gpu(bed, n, m, 1, (0,), g.PREC_FP16)
    p16, _ = one.fetch(0, g.GROUP_WHOLE)
    one.close()
    assert np.array_equal(acc16[np.tril_indices(n)], p16)  # tile ranks reproduce the single-GPU FP32 values bit for bit


def test_finish_async_banded_readout_matches_blocking(monkeypatch):
    # ggrm_finish_async: the last slab is accumulated in row bands and every read-out waits only for the bands it needs.
    # Same kernels, same order per tile -> identical bits to the blocking path, for full and piecewise reads, LOCO partials
    # and the whole-genome sum folded inside the last slab.
    monkeypatch.setenv("GGRM_BAND_MIN_TILES", "4")  # n = 4300 has 153 tiles in 3 strips: force one band per strip
    n, m = 4300, 1500
    bed = g.synth_bed_host(n, 0, m, 91, 0.01)
    groups = [(0, 400), (400, 1500)]

    def build(async_finish, **kw):
        # 128-SNP slabs: 12 per pass (4 + 9 with the LOCO split) - the first ones accumulate as they arrive, the last four stay
        # resident and are accumulated band-major at finish
        acc = g.Accumulator(n, m, method=3, want_add=True, want_dom=True, precision=g.PREC_FP16, chunk_snps=128, **kw)
        for s in range(0, m, 300):
            acc.push(bed[s:min(m, s + 300)], s)
        (acc.finish_async if async_finish else acc.finish)()
        return acc

    for kw in ({}, {"groups": groups, "keep_whole": True}):
        ref = build(False, **kw)
        acc = build(True, **kw)
        targets = [(k, g.GROUP_WHOLE) for k in (0, 1)] + ([(0, 1), (1, 0)] if kw else [])
        for kind, group in targets:
            want, tr_want = ref.fetch(kind, group)
            got, tr = acc.fetch(kind, group)
            assert np.array_equal(got, want) and tr == tr_want, (kw, kind, group)
        # piecewise, out of order, straddling the 2048-row band boundaries; rectangular blocks
        want, _ = ref.fetch(0, g.GROUP_WHOLE)
        acc2 = build(True, **kw)
        for r0, cnt in ((4000, 300), (2000, 100), (0, 2048), (2048, 1952)):
            part, _ = acc2.fetch(0, g.GROUP_WHOLE, r0, cnt)
            assert np.array_equal(part, want[g.tri(r0):g.tri(r0 + cnt)])
        blk = acc2.fetch_block(0, g.GROUP_WHOLE, 4096, n - 4096, 2048, 2048)
        dense = np.zeros((n, n), dtype=np.float32)
        dense[np.tril_indices(n)] = want
        assert np.array_equal(blk, dense[4096:, 2048:4096])
        acc2.finish()  # blocking finish after the async one is allowed and reports device faults
        t = acc2.timers()
        assert t["pass_ms"] > 0 and t["syrk_launches"] >= 3
        acc2.reset()   # and the handle can be re-armed
        acc2.push(bed[:300], 0)
        acc2.finish()
        for a in (ref, acc, acc2):
            a.close()
    # parking the last slabs is scheduling only: with it switched off the bits are the same
    monkeypatch.setenv("GGRM_TAIL_SLABS", "0")
    plain = build(False)
    monkeypatch.delenv("GGRM_TAIL_SLABS")
    parked = build(True)
    for kind in (0, 1):
        assert np.array_equal(plain.fetch(kind, g.GROUP_WHOLE)[0], parked.fetch(kind, g.GROUP_WHOLE)[0])
    # an SNP shard that announces where it stops (ggrm_set_push_end) parks its own last slabs; same values as without the hint
    hinted = g.Accumulator(n, m, method=3, want_add=True, want_dom=False, precision=g.PREC_FP16, chunk_snps=128)
    hinted.set_push_end(900)
    unhinted = g.Accumulator(n, m, method=3, want_add=True, want_dom=False, precision=g.PREC_FP16, chunk_snps=128)
    for acc in (hinted, unhinted):
        for s0 in range(0, 900, 300):
            acc.push(bed[s0:s0 + 300], s0)
    hinted.finish_async()
    unhinted.finish()
    assert np.array_equal(hinted.fetch(0, g.GROUP_WHOLE)[0], unhinted.fetch(0, g.GROUP_WHOLE)[0])
    for a in (plain, parked, hinted, unhinted):
        a.close()


@pytest.mark.parametrize("n,m", [(1, 5), (2, 1), (7, 1), (3, 64), (5, 65), (255, 3), (257, 130)])
def test_minimal_and_ragged_shapes_vs_oracle(n, m):
    # the reference's own edge cases (tests/test_bed_pipe.cpp:289-327: one sample, one SNP) plus sizes around the padding
    # units of this implementation (4 samples per byte, 64-SNP operand blocks, 256-row tiles); center family so that a
    # monomorphic or one-sample SNP stays finite (CenterHWE keeps sd out of the values)
    bed = g.synth_bed_host(n, 0, m, 1000 + 17 * n + m, 0.1)
    for method, kind in ((2, 0), (4, 1), (6, 0)):
        G, den, freqs = o.grm_compute(bed, n, method, kind)
        want = G[np.tril_indices(n)]
        for prec, tol in ((g.PREC_FP64, 1e-12), (g.PREC_FP16, 2e-3)):
            acc = run_gpu(bed, n, m, method, (kind,), prec, push=max(1, m // 2))
            got, tr = acc.fetch(kind, g.GROUP_WHOLE, dtype=np.float64)
            f = acc.snp_stats(0, m)[4]
            acc.close()
            assert got.shape == want.shape
            assert np.array_equal(f, freqs)
            assert same(np.isnan(got), np.isnan(want))
            ok = ~np.isnan(want)
            scale = max(1.0, float(np.max(np.abs(want[ok]))) if ok.any() else 1.0)
            assert np.max(np.abs(got[ok] - want[ok]), initial=0.0) <= tol * scale, (n, m, method, kind, prec)
            if np.isfinite(den):
                assert abs(tr / n - den) <= tol * max(1.0, abs(den))


def test_error_codes_follow_reference_semantics():
    acc = g.Accumulator(100, 50, method=1, precision=g.PREC_FP64)
    bed = g.synth_bed_host(100, 0, 50, 1, 0.0)
    with pytest.raises(g.GrmError) as e:
        acc.push(bed[:10], 45)  # bed_pipe.cpp:38-52 ColumnRangeException
    assert e.value.status == -2 and "invalid chunk range" in e.value.message
    acc.push(bed[:10], 0)
    with pytest.raises(g.GrmError) as e:
        acc.push(bed[:10], 5)   # overlapping range
    assert e.value.status == -3
    with pytest.raises(g.GrmError) as e:
        acc.fetch(1, g.GROUP_WHOLE)  # dominance was not requested
    assert e.value.status == -3
    acc.close()
    with pytest.raises(g.GrmError) as e:
        g.Accumulator(100, 50, method=9)
    assert e.value.status == -1 and "Invalid genotype process method" in e.value.message


def test_reset_allows_a_second_pass():
    n, m = 400, 600
    bed = g.synth_bed_host(n, 0, m, 91, 0.01)
    acc = run_gpu(bed, n, m, 1, (0,), g.PREC_FP16)
    first, _ = acc.fetch(0, g.GROUP_WHOLE)
    acc.reset()
    acc.push(bed, 0)
    acc.finish()
    second, _ = acc.fetch(0, g.GROUP_WHOLE)
    assert np.array_equal(first, second)
    acc.close()


def test_c1_shape_tolerance_16bit():
    # Config C1 (2 000 x 50 000, Yang additive = StandardizeHWE + Add): the 1e-4 bound at a named shape.
    # The CPU oracle would need minutes at this size, so the FP64 tensor-core path (itself within 1e-10 of the oracle
    # above) is the comparison; entries are checked on G/(trace/n).
    n, m = 2000, 50000
    bed = g.synth_bed_host(n, 0, m, 20261105, 0.01)
    ref = run_gpu(bed, n, m, 1, (0,), g.PREC_FP64, push=10000)
    r, _ = ref.fetch(0, g.GROUP_WHOLE, dtype=np.float64)
    ref.close()
    rn = norm_packed(r)
    for prec, bound in ((g.PREC_FP16, 1e-4), (g.PREC_BF16, 1e-4)):
        acc = run_gpu(bed, n, m, 1, (0,), prec, push=10000)
        p, _ = acc.fetch(0, g.GROUP_WHOLE)
        acc.close()
        err = np.max(np.abs(norm_packed(p) - rn))
        assert err <= bound, (prec, err)


def test_c2_snp_count_tolerance_16bit_device_resident():
    # The 1e-4 bound at the SNP count of the headline config C2 (m = 500 000).  The rounding error of 16-bit operands
    # depends on m (it averages out as 1/sqrt(m)), not on n, so n is kept small enough (4 096) for the FP64 DMMA path —
    # itself within 1e-10 of the oracle above — to serve as the comparison.  Payload is generated on the device.
    import torch

    n, m = 4096, 500_000
    bpv = n // 4
    d_bed = torch.empty(m * bpv, dtype=torch.uint8, device="cuda")
    for s in range(0, m, 65536):
        g.synth_bed_device(d_bed.data_ptr() + s * bpv, n, s, min(65536, m - s), 20261107, 0.01)
    torch.cuda.synchronize()

    def build(prec):
        with g.Accumulator(n, m, method=1, precision=prec) as acc:
            acc.push_device(d_bed.data_ptr(), 0, m)
            acc.finish()
            p, tr = acc.fetch(0, g.GROUP_WHOLE, dtype=np.float64)
        return norm_packed(p), tr

    ref, tr_ref = build(g.PREC_FP64)
    assert abs(tr_ref / n / (m * 0.99) - 1.0) < 0.01  # StandardizeHWE: diag mean ~ number of non-missing SNPs
    for prec in (g.PREC_FP16, g.PREC_BF16):
        got, tr = build(prec)
        err = float(np.max(np.abs(got - ref)))
        assert err <= 1e-4, (prec, err)
        # raw (un-normalised) entries: the FP32 TMEM accumulation truncates, a uniform relative bias that grows with the slab
        # length (-8.6e-5 at the default 16 384-SNP slabs) and cancels in G/(trace/n)
        assert abs(tr / tr_ref - 1.0) < 1.5e-4, (prec, tr / tr_ref - 1.0)


def test_gcta_ratio_and_pair_counts():
    # the second accumulator holds N_ij = SNPs called in both samples (exact integers), ggrm_fetch_gcta returns
    # A_ij = G_ij / N_ij and N in the .grm.bin / .grm.N.bin byte order
    n, m = 515, 900
    geno = plink_io.decode_payload(g.synth_bed_host(n, 0, m, 31, 0.08), n)   # (n, m) 0/1/2/NaN
    geno[0, : m // 2] = np.nan      # samples 0 and 1 share no called SNP: N_10 = 0 -> A_10 = 0
    geno[1, m // 2:] = np.nan
    bed = plink_io.encode_matrix(geno)
    called = (~np.isnan(geno)).astype(np.float64)
    N = (called @ called.T)[np.tril_indices(n)]
    G, _, _ = o.grm_compute(bed, n, 1, 0, chunk=256)
    A = np.divide(G[np.tril_indices(n)], N, out=np.zeros_like(N), where=N > 0)
    assert N[1] == 0.0                                                       # packed index of (1, 0)
    for prec, tol in ((g.PREC_FP64, None), (g.PREC_FP16, 2e-3), (g.PREC_BF16, 8e-3)):
        with g.Accumulator(n, m, method="SH", want_add=True, want_dom=True, precision=prec, chunk_snps=256, second_matrix=1) as acc:
            for s in range(0, m, 333):
                acc.push(bed[s:s + 333], s)
            acc.finish()
            got_n = acc.fetch_gcta(1)
            got_a = acc.fetch_gcta(0)
            part = acc.fetch_gcta(0, row_begin=200, row_count=57)
        assert np.array_equal(got_n.astype(np.float64), N), prec               # counts: bit-exact in every precision
        assert got_a[1] == 0.0 and np.all(np.isfinite(got_a))
        assert np.array_equal(part, got_a[g.tri(200):g.tri(257)])
        if tol is None:
            ref32 = A.astype(np.float32)
            assert np.all(np.abs(got_a - ref32) <= np.spacing(np.abs(ref32)) + 1e-12)   # float32(G64 / N) to one ulp
        else:
            assert np.max(np.abs(got_a.astype(np.float64) - A)) <= tol, prec
    with g.Accumulator(n, m, method="SH", want_add=True, want_dom=True, precision=g.PREC_FP64) as acc:
        acc.push(bed, 0)
        acc.finish()
        with pytest.raises(g.GrmError) as e:
            acc.fetch_gcta(0)
        assert "second_matrix" in e.value.message


def test_guard_bands_stay_intact_over_ragged_shapes(monkeypatch):
    # compute-sanitizer is not available on this pool, so the library carries its own bounds check: with GGRM_GUARD=1 every
    # device buffer the kernels write sits between 64 KiB guard bands that ggrm_finish / ggrm_fetch verify.  Shapes are chosen
    # so that every kernel works on ragged edges (n % 256, n % 4, m % 64, last slab shorter than the others, projection with
    # dropped samples, LOCO groups with the fused fold, the band-major tail, the second matrix of the GCTA mode).
    monkeypatch.setenv("GGRM_GUARD", "1")
    monkeypatch.setenv("GGRM_BAND_MIN_TILES", "1")
    rng = np.random.default_rng(9)
    for n, m, chunk in ((257, 130, 64), (1001, 777, 200), (2307, 1000, 192), (515, 65, 0)):
        bed = g.synth_bed_host(n, 0, m, 100 + n, 0.03)
        r2o = rng.permutation(n).astype(np.int32)
        r2o[::7] = -1
        keep = r2o >= 0
        r2o[keep] = np.argsort(np.argsort(r2o[keep])).astype(np.int32)
        n_out = int(keep.sum())
        cuts = sorted({0, m // 3, m // 3 + 1, (2 * m) // 3, m})
        groups = [(cuts[i], cuts[i + 1]) for i in range(len(cuts) - 1)]
        for prec in (g.PREC_BF16, g.PREC_FP16, g.PREC_FP64):
            for kw in ({"want_dom": True}, {"groups": groups, "keep_whole": True}, {"want_dom": True, "second_matrix": 1},
                       {"raw_to_out": r2o, "n_out": n_out, "keep_stats": True}):
                with g.Accumulator(n, m, method="OSH", precision=prec, chunk_snps=chunk, **kw) as acc:
                    for s in range(0, m, 97):
                        acc.push(bed[s:s + 97], s)
                    acc.finish_async()
                    acc.fetch(g.KIND_ADD, g.GROUP_WHOLE, dtype=np.float64 if prec == g.PREC_FP64 else np.float32)
                    acc.finish()                                  # verifies every guard band; raises GrmError on a stray store
                    if kw.get("second_matrix"):
                        acc.fetch_gcta(0)
                    acc.reset()
                    acc.push(bed, 0)
                    acc.finish()
    # the detector itself: a deliberate one-byte store past a buffer is reported
    monkeypatch.setenv("GGRM_GUARD_SELFTEST", "1")
    with g.Accumulator(300, 64, method="SH", precision=g.PREC_BF16) as acc:
        acc.push(g.synth_bed_host(300, 0, 64, 1, 0.0), 0)
        with pytest.raises(g.GrmError) as e:
            acc.finish()
        assert "guard band of device buffer 'dense 2-bit chunk'" in e.value.message and "overwritten at byte offset" in e.value.message


@pytest.mark.parametrize("seed", [1, 2, 3])
def test_fp64_path_agrees_with_oracle_on_random_small_inputs(seed):
    # the generator of tests/test_reference_differential.py::test_oracle_agrees_with_the_reference_on_random_small_inputs (where the
    # oracle is compared with the unmodified reference on the same cases): tiny ragged shapes, heavy missingness, duplicate and
    # unsorted ids, monomorphic / all-heterozygous / all-missing SNPs, every policy, ragged device slabs
    rng = np.random.default_rng(seed)
    for k in range(24):
        n_raw, m = int(rng.integers(1, 24)), int(rng.integers(1, 40))
        p = rng.uniform(0.02, 0.98, size=m)
        geno = rng.binomial(2, p, size=(n_raw, m)).astype(np.float64)
        geno[rng.random((n_raw, m)) < rng.choice([0.0, 0.05, 0.4])] = np.nan
        for j in range(m):
            q = rng.random()
            if q < 0.06:
                geno[:, j] = rng.integers(0, 3)
            elif q < 0.09:
                geno[:, j] = 1.0
            elif q < 0.11:
                geno[:, j] = np.nan
        fam = [(f"f{int(rng.integers(0, 3))}", f"s{int(rng.integers(0, max(2, n_raw)))}" if rng.random() < 0.3 else f"s{i:03d}") for i in range(n_raw)]
        fam = [fam[i] for i in rng.permutation(n_raw)]
        rng.choice(["\t", " "])  # (keeps the stream in step with the CPU test's generator)
        payload = plink_io.encode_matrix(geno)
        ids, r2o = plink_io.sample_order(fam)
        method, kind, chunk = int(rng.integers(1, 9)), int(rng.integers(0, 2)), int(rng.choice([1, 3, 7, 10000]))
        G, _, freqs = o.grm_compute(payload, n_raw, method, kind, chunk=chunk, n_out=len(ids), raw_to_out=r2o)
        want = o.pack_lower_f64(G)
        acc = run_gpu(payload, n_raw, m, method, (kind,), g.PREC_FP64, chunk=64, push=max(1, chunk if chunk < 100 else 13), n_out=len(ids), raw_to_out=r2o)
        got, _ = acc.fetch(kind, g.GROUP_WHOLE, dtype=np.float64)
        f = acc.snp_stats(0, m)[4]
        acc.close()
        ctx = f"seed {seed} case {k}: n_raw={n_raw} n={len(ids)} m={m} method={method} kind={kind}"
        assert np.array_equal(f, freqs), ctx  # allele frequencies: bit-exact
        assert np.array_equal(np.isnan(got), np.isnan(want)), ctx
        ok = ~np.isnan(want)
        if ok.any():
            assert np.max(np.abs(got[ok] - want[ok])) <= 1e-10 * max(1.0, float(np.max(np.abs(want[ok])))), ctx


def test_stage_slots_feed_two_handles_like_host_pushes():
    # ggrm_stage_* + ggrm_push_device + ggrm_wait_copies (the 2D-tile BED relay at the ABI level): every chunk is uploaded once into one
    # of two staging slots, two handles take it from there, and a slot is only refilled after both have waited for their copies -
    # bit-identical to two handles fed by ggrm_push from host memory
    from gelex_b200._lib import Stage

    n, m, chunk = 1500, 1200, 100
    bed = g.synth_bed_host(n, 0, m, 47, 0.02)
    kw = dict(method=1, precision=g.PREC_FP64, chunk_snps=256)
    ref = g.Accumulator(n, m, **kw)
    a, b = g.Accumulator(n, m, **kw), g.Accumulator(n, m, **kw)
    with Stage(chunk * bed.shape[1], n_slots=2) as st:
        for j, s in enumerate(range(0, m, chunk)):
            ref.push(bed[s:s + chunk], s)
            st.upload(j % 2, bed[s:s + chunk])
            for acc in (a, b):
                acc.push_device(st.slot(j % 2), s, min(chunk, m - s))
            for acc in (a, b):
                acc.wait_copies()
        for acc in (ref, a, b):
            acc.finish()
    want, _ = ref.fetch(g.KIND_ADD, g.GROUP_WHOLE, dtype=np.float64)
    for acc in (a, b):
        got, _ = acc.fetch(g.KIND_ADD, g.GROUP_WHOLE, dtype=np.float64)
        assert np.array_equal(got, want)
    with pytest.raises(g.GrmError):
        Stage(0)
    for acc in (ref, a, b):
        acc.close()


# ------------------------------------------------------------------------------------------------ round-2 additions
def test_loco_and_normalised_readout_16bit_vs_oracle():
    # The consumer-side read-outs on the 16-BIT path: G/(trace/n) of every chromosome partial and of the folded
    # whole, and (G_w - G_c)/(k_w - k_c) (loco_grm_loader.cpp:62-76).  Groups of very different lengths, so the partials are
    # accumulated in slabs of different lengths (the FP32 truncation bias of a slab grows with its length) and the whole mixes them.
    n, m = 700, 12000
    bed = g.synth_bed_host(n, 0, m, 83, 0.02)
    groups = [(0, 5000), (5000, 5300), (5300, 9000), (9000, 9064), (9064, 12000)]
    tl = np.tril_indices(n)
    parts = [o.grm_compute(bed, n, 1, 0, ranges=[gr], chunk=1000)[0] for gr in groups]
    whole = sum(parts)
    k = lambda G: np.trace(G) / n  # noqa: E731
    for prec in (g.PREC_FP16, g.PREC_BF16):
        with g.Accumulator(n, m, method=1, precision=prec, groups=groups, keep_whole=True, chunk_snps=2048) as acc:
            for s in range(0, m, 1234):
                acc.push(bed[s:s + 1234], s)
            acc.finish()
            assert np.max(np.abs(acc.fetch_normalized(0, g.GROUP_WHOLE) - (whole / k(whole))[tl])) <= TOL16
            for gi, P in enumerate(parts):
                if groups[gi][1] - groups[gi][0] >= 2000:  # a 64- or 300-SNP partial alone is noise-dominated at any precision budget
                    assert np.max(np.abs(acc.fetch_normalized(0, gi) - (P / k(P))[tl])) <= TOL16, (prec, gi)
                ref = ((whole - P) / (k(whole) - k(P)))[tl]
                assert np.max(np.abs(acc.fetch_loco(0, gi) - ref)) <= TOL16, (prec, gi)


def test_gapped_pushes_across_the_tail_match_the_oracle():
    # ggrm.h allows pushes that skip SNPs.  Gaps inside the last tail_slabs * k_cap SNPs cut the tail into more parked slabs than
    # there are spare operand buffers; the ring must accumulate parked slabs before it wraps onto them (round-1 advisor finding).
    n, m, k = 600, 6000, 256
    bed = g.synth_bed_host(n, 0, m, 91, 0.01)
    ranges = [(0, 3000)] + [(s, min(m, s + 100)) for s in range(3200, 6000, 130)]  # ~22 gapped pieces, the last ones inside the parked tail
    sel = np.concatenate([np.arange(b, e) for b, e in ranges])
    G, den, _ = o.grm_compute(bed, n, 1, 0, ranges=ranges, chunk=500)
    ref = norm_ref(G)
    for prec, tol in ((g.PREC_FP64, TOL64), (g.PREC_FP16, TOL16)):
        with g.Accumulator(n, m, method=1, precision=prec, chunk_snps=k) as acc:
            for b, e in ranges:
                acc.push(bed[b:e], b)
            acc.finish()
            packed, tr = acc.fetch(0, g.GROUP_WHOLE, dtype=np.float64)
        assert np.max(np.abs(norm_packed(packed) - ref)) <= tol, prec
    assert sel.size == sum(e - b for b, e in ranges)


def test_split_k_few_tiles_agrees_with_unsplit_and_oracle(monkeypatch):
    # n <= ~2 300 leaves CTA pairs idle (C1: 36 tiles for 74 pairs), so slabs are also cut along K, each part into its own copy of
    # the tile, summed in a fixed order when the group completes (Handle::ksplit).  Same operands, other summation order: equal to
    # the unsplit launch up to FP32 rounding, deterministic from run to run, and inside the bound against the oracle rows.
    n, m = 2000, 20000
    bed = g.synth_bed_host(n, 0, m, 97, 0.01)
    rows = np.array([0, 5, 777, 1500, 1999])
    (R,), (D,) = o.row_spot(bed, n, 1, 0, rows, chunk=2048)
    outs = {}
    for ks in ("1", "auto", "auto", "3"):
        if ks == "auto":
            monkeypatch.delenv("GGRM_KSPLIT", raising=False)
        else:
            monkeypatch.setenv("GGRM_KSPLIT", ks)
        acc = run_gpu(bed, n, m, 1, (0,), g.PREC_FP16, push=5000)
        outs.setdefault(ks, []).append(acc.fetch(0, g.GROUP_WHOLE)[0])
        acc.close()
    assert np.array_equal(outs["auto"][0], outs["auto"][1])
    base = norm_packed(outs["1"][0])
    for ks in ("auto", "3"):  # other chain lengths, other truncation pattern of the FP32 TMEM accumulation: agreement far inside the bound
        diff = float(np.max(np.abs(norm_packed(outs[ks][0]) - base)))
        print(f"split-K {ks} vs unsplit: max abs diff on G/k = {diff:.3e}")
        assert diff <= 1e-5, (ks, diff)
    kd = D.sum() / n
    for a, r in enumerate(rows):
        got = norm_packed(outs["auto"][0])[g.tri(r):g.tri(r + 1)]
        assert np.max(np.abs(got - R[a, : r + 1] / kd)) <= TOL16


@pytest.mark.parametrize("precision,method,kinds,m,n", [
    (g.PREC_FP16, 1, (0,), 8192, 1024),          # FP16, additive: first unsplit size
    (g.PREC_FP16, 3, (0, 1), 65536, 512),        # FP16, OrthStandardizeHWE add + dom (heavy-tailed dominance codes): first unsplit size
    (g.PREC_BF16, 1, (0,), 1 << 18, 512),        # BF16, additive: first unsplit size
    (g.PREC_BF16, 3, (0, 1), 1 << 22, 256),      # BF16, OrthStandardizeHWE add + dom: first unsplit size
], ids=["fp16-add-8k", "fp16-osh-dom-64k", "bf16-add-256k", "bf16-osh-dom-4M"])
def test_first_unsplit_sizes_keep_the_bound(precision, method, kinds, m, n):
    # create_impl splits 16-bit operands into hi + lo below a SNP count that depends on the precision and on the code policy.
    # Right AT each threshold the plain 16-bit product runs: the bound must hold there with room to spare (the margin is the point
    # of the thresholds), against every row of the FP64 host matrix.
    bed = g.synth_bed_host(n, 0, m, 101 + m % 97, 0.01)
    rows = np.arange(n)
    with g.Accumulator(n, m, method=method, want_add=0 in kinds, want_dom=1 in kinds, precision=precision) as acc:
        for s in range(0, m, 65536):
            acc.push(bed[s:s + 65536], s)
        acc.finish()
        for kind in kinds:
            (R,), (D,) = o.row_spot(bed, n, method, kind, rows, chunk=8192)
            got = acc.fetch_normalized(kind, g.GROUP_WHOLE)
            err = float(np.max(np.abs(got - (R / (D.sum() / n))[np.tril_indices(n)])))
            print(f"unsplit precision={precision} method={method} kind={kind} m={m}: max abs err {err:.3e}  (err * sqrt(m) = {err * m ** 0.5:.4f})")
            assert err <= 0.6 * TOL16, (precision, method, kind, m, err)


def test_bf16_hi_lo_split_keeps_the_bound_on_short_inputs(monkeypatch):
    # GGRM_PREC_BF16 below 2^18 SNPs runs hi + lo operands (3 MMAs per K block): inside 1e-4 where plain BF16 is not
    n, m = 600, 2048
    bed = g.synth_bed_host(n, 0, m, 20261104, 0.01)
    G, den, _ = o.grm_compute(bed, n, 1, 0, chunk=1024)
    ref = norm_ref(G)
    err = {}
    for split in ("0", "1"):
        monkeypatch.setenv("GGRM_BF16_SPLIT", split)
        acc = run_gpu(bed, n, m, 1, (0,), g.PREC_BF16, chunk=512)
        err[split] = float(np.max(np.abs(norm_packed(acc.fetch(0, g.GROUP_WHOLE)[0]) - ref)))
        acc.close()
    monkeypatch.delenv("GGRM_BF16_SPLIT")
    acc = run_gpu(bed, n, m, 1, (0,), g.PREC_BF16, chunk=512)
    auto = float(np.max(np.abs(norm_packed(acc.fetch(0, g.GROUP_WHOLE)[0]) - ref)))
    acc.close()
    assert err["1"] <= 2e-5 and auto == err["1"], (err, auto)
    assert err["0"] > 1e-4  # plain BF16 at this size: the reason the split exists (round 1 measured 4.0e-4)


def test_fp16_overflow_is_reported_and_bf16_stays_finite():
    # OrthStandardizeHWE dominance codes a homozygote as -q/p: with more than 65 504 samples a SNP whose rare allele appears once,
    # as a homozygote, gives |z| ~ n, outside binary16.  FP16 operands must fail loudly instead of writing inf/NaN rows; BF16
    # (FP32 exponent range) stays finite and agrees with the FP64 host rows (round-1 advisor finding).
    n, m = 66_000, 192
    geno_bed = g.synth_bed_host(n, 0, m, 5, 0.0).copy()
    geno_bed[7] = 0xFF               # SNP 7: every sample homozygous A2 (BED 11) ...
    geno_bed[7, 100] = 0xFC          # ... except sample 400: A1A1 (BED 00), the only carrier
    rows = np.array([0, 400, 401, n - 1])
    (R,), (D,) = o.row_spot(geno_bed, n, 3, 1, rows, chunk=64)
    assert np.max(np.abs(R)) > 65504.0 ** 2 * 0.9
    with g.Accumulator(n, m, method=3, want_add=False, want_dom=True, precision=g.PREC_FP16) as acc:
        acc.push(geno_bed, 0)
        with pytest.raises(g.GrmError, match="FP16 operand overflow"):
            acc.finish()
    with g.Accumulator(n, m, method=3, want_add=False, want_dom=True, precision=g.PREC_BF16) as acc:
        acc.push(geno_bed, 0)
        acc.finish()
        for a, r in enumerate(rows):
            got, _ = acc.fetch(g.KIND_DOM, g.GROUP_WHOLE, row_begin=int(r), row_count=1)
            assert np.all(np.isfinite(got))
            ref = R[a, : r + 1]
            assert np.max(np.abs(got - ref)) <= 1e-4 * np.max(np.abs(ref)) + 1e-3, r


@pytest.mark.gpu
@pytest.mark.parametrize("precision", [g.PREC_FP16, g.PREC_BF16])
def test_fused_prep_launch_is_bit_identical_to_the_three_kernels(monkeypatch, precision):
    # single-slab passes of few samples run K1-K3 as ONE launch (prep_fused_kernel); the separate kernels stay the path of every
    # other shape.  Same device functions on the same bytes: operands, statistics and therefore the accumulators are bit-identical.
    rng = np.random.default_rng(17)
    cases = [
        dict(n=515, m=900, kw=dict(method="OSH", want_add=True, want_dom=True)),                       # ragged n, two kinds
        dict(n=2000, m=3000, kw=dict(method="SH", want_add=True, want_dom=True, second_matrix=2)),     # GCTA counts + diagonal terms
        dict(n=3100, m=130, kw=dict(method="OC", want_add=False, want_dom=True), drop=200),            # projection with dropped samples, m % 64 != 0
        dict(n=4096, m=257, kw=dict(method="S", want_add=True, want_dom=False)),                       # the largest sample count the launch takes
    ]
    for case in cases:
        n, m, kw = case["n"], case["m"], dict(case["kw"])
        bed = g.synth_bed_host(n, 0, m, 1234 + n, 0.03)
        n_out = n
        if "drop" in case:
            r2o = rng.permutation(n).astype(np.int32)
            n_out = n - case["drop"]
            r2o[r2o >= n_out] = -1
            kw.update(raw_to_out=r2o, n_out=n_out)
        got = {}
        for fused in ("0", "1"):
            monkeypatch.setenv("GGRM_FUSED_PREP", fused)
            with g.Accumulator(n, m, precision=precision, keep_stats=True, **kw) as acc:
                acc.push(bed[: m // 3], 0)
                acc.push(bed[m // 3:], m // 3)
                acc.finish()
                out = [acc.fetch(k, g.GROUP_WHOLE)[0] for k in (0, 1) if (kw.get("want_add", True), kw.get("want_dom", False))[k]]
                out += list(acc.snp_stats(0, m))
                if kw.get("second_matrix") == 2:
                    out.append(acc.fetch_gcta(0))
                launches = acc.timers()["kernel_launches"]
            got[fused] = (out, launches)
        assert got["1"][1] < got["0"][1], "the fused launch replaces at least two"
        for a, b in zip(got["0"][0], got["1"][0]):
            assert np.array_equal(a, b, equal_nan=True), (n, m)
    monkeypatch.delenv("GGRM_FUSED_PREP")
    # more samples than the launch holds in shared memory, or more than one slab: the separate kernels, whatever the environment says
    with g.Accumulator(4100, 100, method="SH", precision=precision) as acc:
        acc.push(g.synth_bed_host(4100, 0, 100, 5, 0.0), 0)
        acc.finish()
        assert acc.timers()["kernel_launches"] >= 4


@pytest.mark.gpu
def test_device_payload_read_in_place_matches_the_copied_one(monkeypatch):
    # one ggrm_push_device call that is the whole single-slab pass is read where it lies by the fused prep launch (no copy into the
    # chunk buffer); ggrm_wait_copies hands the memory back early by making the copy after all.  Same bits every way.
    import torch

    n, m = 1000, 4000
    bed = g.synth_bed_host(n, 0, m, 99, 0.02)
    d_bed = torch.from_numpy(bed).cuda()
    torch.cuda.synchronize()

    def run(mode):
        with g.Accumulator(n, m, method="SH", precision=g.PREC_FP16, keep_stats=True) as acc:
            if mode == "two pushes":        # not the whole pass in one call: copied
                acc.push_device(d_bed.data_ptr(), 0, m // 2)
                acc.push_device(d_bed.data_ptr() + (m // 2) * bed.shape[1], m // 2, m - m // 2)
            elif mode == "wait_copies":     # registered in place, then claimed back before the pass runs
                scratch = d_bed.clone()
                torch.cuda.synchronize()
                acc.push_device(scratch.data_ptr(), 0, m)
                acc.wait_copies()
                scratch.fill_(0x55)          # the caller's memory is its own again: all missing now
                torch.cuda.synchronize()
            else:
                acc.push_device(d_bed.data_ptr(), 0, m)
            acc.finish()
            out = acc.fetch(0, g.GROUP_WHOLE)[0]
            launches = acc.timers()["kernel_launches"]
            # a second pass over the same handle
            acc.reset()
            acc.push_device(d_bed.data_ptr(), 0, m)
            acc.finish()
            again = acc.fetch(0, g.GROUP_WHOLE)[0]
        assert np.array_equal(out, again)
        return out, launches

    monkeypatch.setenv("GGRM_DIRECT_SRC", "0")
    copied, _ = run("one push")
    monkeypatch.delenv("GGRM_DIRECT_SRC")
    assert np.all(np.isfinite(copied)) and copied.any()
    for mode in ("one push", "two pushes", "wait_copies"):
        got, _ = run(mode)
        assert np.array_equal(got, copied), mode
    # LOCO groups inside one slab: every group's push range is registered and flushed on its own
    groups = [(0, 1500), (1500, 4000)]
    res = []
    for direct in ("0", "1"):
        monkeypatch.setenv("GGRM_DIRECT_SRC", direct)
        with g.Accumulator(n, m, method="SH", precision=g.PREC_FP16, groups=groups, keep_whole=True) as acc:
            acc.push_device(d_bed.data_ptr(), 0, m)
            acc.finish()
            res.append([acc.fetch(0, gi)[0] for gi in (0, 1, g.GROUP_WHOLE)])
    for a, b in zip(*res):
        assert np.array_equal(a, b)
