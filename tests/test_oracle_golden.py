"""CPU tests (no GPU): pin the oracle (oracle/grm_oracle.c) against
  (a) the known-answer vectors of the reference's own unit tests, and
  (b) the golden fixtures produced by the unmodified reference (tests/golden/make_golden.py).
"""
import json
import math

import numpy as np
import pytest

import oracle_lib as o
import plink_io

TOL = 1e-10  # the reference's own tolerance (tests/test_genotype_processor.cpp: WithinRel(1e-10))
KIND = {"A": 0, "D": 1}


def load_tiny(golden_dir):
    fam = [tuple(l.split()[:2]) for l in open(golden_dir / "tiny.fam")]
    chroms = [l.split()[0] for l in open(golden_dir / "tiny.bim")]
    raw = np.fromfile(golden_dir / "tiny.bed", dtype=np.uint8)
    assert bytes(raw[:3]) == plink_io.BED_MAGIC
    n_raw = len(fam)
    payload = raw[3:].reshape(len(chroms), (n_raw + 3) // 4)
    ids, r2o = plink_io.sample_order(fam)
    return payload, n_raw, ids, r2o, chroms


def load_dense(golden_dir):
    fam = [tuple(l.split()[:2]) for l in open(golden_dir / "dense.fam")]
    chroms = [l.split()[0] for l in open(golden_dir / "dense.bim")]
    raw = np.fromfile(golden_dir / "dense.bed", dtype=np.uint8)
    payload = raw[3:].reshape(len(chroms), (len(fam) + 3) // 4)
    return payload, len(fam), chroms


def same(a, b):
    """NaN-aware exact equality (reference tests/bed_fixture.cpp:31-59 are_matrices_equal)."""
    a, b = np.asarray(a), np.asarray(b)
    return a.shape == b.shape and np.array_equal(np.isnan(a), np.isnan(b)) and np.array_equal(np.nan_to_num(a), np.nan_to_num(b))


def close(a, b, rtol=TOL, atol=1e-12):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    if a.shape != b.shape or not np.array_equal(np.isnan(a), np.isnan(b)):
        return False
    m = ~np.isnan(a)
    fin = m & np.isfinite(a)
    return np.array_equal(a[m & ~fin], b[m & ~fin]) and np.allclose(a[fin], b[fin], rtol=rtol, atol=atol)


# ---------------------------------------------------------------- (a) reference unit-test vectors
def test_standardize_add_known_answer():
    # tests/test_genotype_processor.cpp:75-89
    z, st = o.process_column([0, 1, 2, 1, 0], 5, 0)
    assert st.mean == pytest.approx(0.8, rel=TOL)
    assert st.stddev == pytest.approx(0.8366600265340756, rel=TOL)
    assert z[0] == pytest.approx(-0.9561828874675147, rel=TOL)
    assert z[1] == pytest.approx(0.23904572186687866, rel=TOL)
    assert z[2] == pytest.approx(1.434274331201319, rel=TOL)


def test_standardize_dom_known_answer():
    # tests/test_genotype_processor.cpp:144-153  Dom raw {0,1,2,1,0,2} -> mean 1/3, sd 0.5163977794943222
    z, st = o.process_column([0, 1, 2, 1, 0, 2], 5, 1)
    assert st.mean == pytest.approx(1.0 / 3.0, rel=TOL)
    assert st.stddev == pytest.approx(0.5163977794943222, rel=TOL)


def test_hwe_known_answers():
    # tests/test_genotype_processor.cpp:205-270 (maf = 0.4 for {0,1,2,1,0})
    _, st = o.process_column([0, 1, 2, 1, 0], 1, 0)
    assert st.maf == pytest.approx(0.4, rel=TOL) and st.stddev == pytest.approx(math.sqrt(2 * 0.4 * 0.6), rel=TOL)
    _, st = o.process_column([0, 1, 2, 1, 0], 1, 1)
    assert st.mean == pytest.approx(0.48, rel=TOL) and st.stddev == pytest.approx(math.sqrt(0.48 * 0.52), rel=TOL)
    _, st = o.process_column([0, 1, 2, 1, 0], 3, 1)
    assert st.mean == pytest.approx(0.32, rel=TOL) and st.stddev == pytest.approx(0.48, rel=TOL)


def test_center_family_known_answers():
    # tests/test_grm_code_policy.cpp:34-44, 121-135, 182-200, 243-264
    z, _ = o.process_column([0, 1, 2, 1, 0], 6, 0)
    assert np.allclose(z, [-0.8, 0.2, 1.2, 0.2, -0.8], rtol=TOL)
    z, _ = o.process_column([0, 1, 2, 1, 0], 6, 1)
    assert np.allclose(z, [-0.4, 0.6, -0.4, 0.6, -0.4], rtol=TOL)
    z, _ = o.process_column([0, 1, 2, 1, 0], 8, 1)
    assert np.allclose(z, [-0.24, 0.56, -0.64, 0.56, -0.24], rtol=TOL)
    z, _ = o.process_column([0, 1, 2, 1, 0], 4, 1)
    assert np.allclose(z, [-0.32, 0.48, -0.72, 0.48, -0.32], rtol=TOL)


def test_monomorphic_is_centred_not_scaled():
    # tests/test_genotype_processor.cpp:92-102,156-165
    for method in (5, 7):  # sample statistics: any constant column is monomorphic
        z, st = o.process_column([1, 1, 1, 1, 1], method, 0)
        assert st.is_monomorphic == 1 and np.all(z == 0.0)
    for method in (1, 3, 5, 7):  # HWE statistics: p = 0 gives sd = 0
        z, st = o.process_column([0, 0, 0, 0], method, 0)
        assert st.is_monomorphic == 1 and np.all(z == 0.0)


def test_hand_computed_grm_4x3():
    # tests/test_grm.cpp:271-394: G = Z Z^T of the processed matrix, lower triangle only, tol 1e-8
    geno = np.array([[0, 1, 2], [1, 1, 0], [2, 0, 1], [1, 2, 1]], dtype=np.float64)
    payload = plink_io.encode_matrix(geno)
    G, den, _ = o.grm_compute(payload, 4, 6, 0, chunk=2)
    Z = geno - geno.mean(axis=0)  # Center Add: mean = 2*maf = column mean when nothing is missing
    ref = Z @ Z.T
    assert np.allclose(np.tril(G), np.tril(ref), atol=1e-8)
    assert np.all(np.triu(G, 1) == 0.0)  # upper triangle untouched (tests/test_grm.cpp:100,155)
    assert den == pytest.approx(np.trace(ref) / 4, rel=1e-12)


def test_chunk_size_invariance():
    # tests/test_grm.cpp:187-225: chunk 1 / 7 / m / m+100 agree within 1e-10
    rng = np.random.default_rng(3)
    geno = rng.integers(0, 3, size=(23, 40)).astype(np.float64)
    payload = plink_io.encode_matrix(geno)
    base, _, _ = o.grm_compute(payload, 23, 3, 0, chunk=40)
    for c in (1, 7, 140):
        G, _, _ = o.grm_compute(payload, 23, 3, 0, chunk=c)
        assert np.allclose(G, base, rtol=0, atol=1e-10)


def test_bin_layout_known_answer():
    # tests/test_grm_bin_writer.cpp:142-181: 3x3 -> (0,0),(1,0),(1,1),(2,0),(2,1),(2,2), cast to float32
    G = np.array([[1.0, 0, 0], [2.0, 3.0, 0], [4.0, 5.0, 6.5]])
    assert np.array_equal(o.pack_lower_f32(G), np.array([1, 2, 3, 4, 5, 6.5], dtype=np.float32))
    assert o.pack_lower_f32(np.zeros((0, 0))).size == 0  # :89-113 empty matrix -> empty file


# ---------------------------------------------------------------- (b) golden fixtures from the unmodified reference
def test_decode_matches_reference(golden_dir):
    payload, n_raw, ids, r2o, _ = load_tiny(golden_dir)
    gold = np.load(golden_dir / "tiny_chunks.npz")
    assert same(o.decode_chunk(payload, n_raw, len(ids), r2o), gold["raw"])          # sparse path, duplicate ID
    assert same(o.decode_chunk(payload[17:61], n_raw, len(ids), r2o), gold["raw_17_61"])
    dpayload, dn, _ = load_dense(golden_dir)
    assert same(o.decode_chunk(dpayload, dn), np.load(golden_dir / "dense_chunks.npz")["raw"])  # dense path


@pytest.mark.parametrize("method", range(1, 9))
@pytest.mark.parametrize("kind", ["A", "D"])
def test_process_matrix_matches_reference(golden_dir, method, kind):
    payload, n_raw, ids, r2o, _ = load_tiny(golden_dir)
    gold = np.load(golden_dir / "tiny_chunks.npz")
    z, freqs = o.process_matrix(o.decode_chunk(payload, n_raw, len(ids), r2o), method, KIND[kind])
    assert np.array_equal(freqs, gold[f"freq_{method}_{kind}"])  # allele frequencies: bit-exact
    assert close(z, gold[f"z_{method}_{kind}"])


def test_process_vectors_match_reference(golden_dir):
    vectors = json.load(open(golden_dir / "process_vectors.json"))
    assert len(vectors) == 6 * 16
    for v in vectors:
        col = [float("nan") if x == "nan" else float(x) for x in v["input"]]
        z, st = o.process_column(col, v["method"], KIND[v["kind"]])
        ref = np.array([float("nan") if x == "nan" else x for x in v["z"]], dtype=np.float64)
        assert st.maf == v["maf"], v
        assert close(z, ref), v
        assert st.is_monomorphic == v["mono"], v
        # LocusStatistic moments (what GenotypeLoader / GenotypePipe keep per SNP)
        assert st.mean == v["mean"] or abs(st.mean - v["mean"]) <= 1e-15 * abs(v["mean"]), v
        assert st.stddev == v["stddev"] or abs(st.stddev - v["stddev"]) <= 4e-16 * abs(v["stddev"]) + 1e-300, v


@pytest.mark.parametrize("tag,method,kind,chunk", [("SH.add", 1, 0, 25), ("OSH.add", 3, 0, 10000), ("OSH.dom", 3, 1, 10000),
                                                    ("OCH.add", 4, 0, 7), ("OCH.dom", 4, 1, 7), ("S.add", 5, 0, 96), ("S.dom", 5, 1, 96)])
def test_grm_bin_matches_reference(golden_dir, tag, method, kind, chunk):
    payload, n_raw, ids, r2o, _ = load_tiny(golden_dir)
    n = len(ids)
    G, _, _ = o.grm_compute(payload, n_raw, method, kind, chunk=chunk, n_out=n, raw_to_out=r2o)
    ref = plink_io.read_bin(golden_dir / f"tiny_{tag}.bin", n)
    got = o.pack_lower_f32(G)
    assert np.array_equal(np.isnan(got), np.isnan(ref))
    m = ~np.isnan(ref)
    # float32 file values: identical after rounding, or within 1 float32 ulp (BLAS summation order differs)
    assert np.all(np.abs(got[m] - ref[m]) <= np.spacing(np.abs(ref[m])) + 1e-30)


def test_loco_partials_match_reference(golden_dir):
    payload, n_raw, ids, r2o, chroms = load_tiny(golden_dir)
    n = len(ids)
    for name, b, e in plink_io.chrom_runs(chroms):
        G, _, _ = o.grm_compute(payload, n_raw, 4, 0, ranges=[(b, e)], chunk=7, n_out=n, raw_to_out=r2o)
        ref = plink_io.read_bin(golden_dir / f"tiny_OCH.add.chr{name}.bin", n)
        got = o.pack_lower_f32(G)
        assert np.all(np.abs(got - ref) <= np.spacing(np.abs(ref)) + 1e-30)


def test_id_file_matches_reference(golden_dir):
    _, _, ids, _, _ = load_tiny(golden_dir)
    text = "".join(s.replace("\x1f", "\t") + "\n" for s in ids)
    assert text == open(golden_dir / "tiny_SH.add.id").read()
    assert len(ids) == 36  # 37 FAM lines, one duplicated ID


def test_loaders_match_reference(golden_dir):
    n = 36
    gold = np.load(golden_dir / "tiny_loaders.npz")
    whole = plink_io.read_bin(golden_dir / "tiny_SH.add.bin", n)
    assert np.allclose(o.load_normalized(whole, n), gold["SH_add_normalized"], rtol=1e-14, atol=0)
    w = plink_io.read_bin(golden_dir / "tiny_OCH.add.bin", n)
    for c in ("1", "2", "X", "7"):
        part = plink_io.read_bin(golden_dir / f"tiny_OCH.add.chr{c}.bin", n)
        assert np.allclose(o.loco(w, part, n), gold[f"OCH_add_loco_chr{c}"], rtol=1e-13, atol=1e-15)


def test_whole_equals_sum_of_chromosome_partials(golden_dir):
    # the identity the GPU engine relies on to get the whole-genome matrix for free
    n = 36
    w = plink_io.read_bin(golden_dir / "tiny_OCH.add.bin", n).astype(np.float64)
    s = sum(plink_io.read_bin(golden_dir / f"tiny_OCH.add.chr{c}.bin", n).astype(np.float64) for c in ("1", "2", "X", "7"))
    assert np.allclose(w, s, rtol=1e-5, atol=1e-5)


PIPE_CASES = [("OSH", 3, True, 25), ("SH", 1, False, 7), ("OC", 8, False, 40), ("S", 5, True, 33)]


@pytest.mark.parametrize("tag,code,additive,chunk", PIPE_CASES)
def test_genotype_pipe_files_of_the_reference(golden_dir, tag, code, additive, chunk):
    # .bmat / .snpstats written by the unmodified reference's GenotypePipe::process<GT> (tests/golden/make_golden.py pipe): pins the
    # layout this repo's GenotypePipe mirror writes - .bmat = column-major FP64 without header, .snpstats =
    # {int64 n, int64 m, int64 #mono} + int64 indices + m means + m stddevs - and the oracle's LocusStatistic beside it.
    from gelex_b200 import engine

    payload, n_raw, ids, r2o, _ = load_tiny(golden_dir)
    n, m = len(ids), payload.shape[0]
    kind = 0 if additive else 1
    base = golden_dir / f"tiny_pipe_{tag}.{'add' if additive else 'dom'}"
    bmat = np.fromfile(str(base) + ".bmat", dtype=np.float64)
    assert bmat.size == n * m
    gold = np.load(golden_dir / "tiny_chunks.npz")
    z = bmat.reshape(m, n).T
    assert same(z, gold[f"z_{code}_{'A' if additive else 'D'}"])  # the file holds process_matrix's output, bit for bit, whatever the chunking
    meta = engine.read_snp_stats(str(base) + ".snpstats")
    assert (meta["num_samples"], meta["num_variants"]) == (n, m)
    raw = o.decode_chunk(payload, n_raw, n, r2o)
    mono = []
    for j in range(m):
        _, ls = o.process_column(raw[:, j], code, kind)
        assert close(meta["mean"][j], ls.mean, rtol=1e-15, atol=0) and close(meta["stddev"][j], ls.stddev, rtol=1e-15, atol=1e-300), j  # sample policies: Eigen sums the centred column in another order
        if ls.is_monomorphic:
            mono.append(j)
    assert meta["monomorphic_indices"].tolist() == mono
