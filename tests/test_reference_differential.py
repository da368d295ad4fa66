"""Differential test of the host side against the UNMODIFIED reference (oracle/_ref/gelex_ref, built by oracle/Makefile from
/root/reference; skipped where that binary does not exist): random small PLINK triples with the quirks real files have - tab / space /
repeated-space / mixed FAM delimiters, duplicate and unsorted sample IDs, short and long lines, blank and whitespace-only lines, '#'
lines, non-ASCII ids, CRLF, missing final newline,
chromosomes that re-appear, non-numeric positions, wrong BED magic, truncated or over-long payloads, empty files - go through
`gelex grm` of the reference and through this repo's GrmEngine front half (everything up to the first device call; there is no
GPU here, so a run that passes validation ends in the engine's "no CUDA device" error).

Checked per case: the same inputs are accepted or rejected; a rejection carries the reference's exception message verbatim
(so validation happens in the reference's ORDER: FAM, BIM line count, BED, BIM records - grm.cpp:36-40, bed_pipe.cpp:69-98,
grm_engine.cpp:52-70); an accepted case plans the files the reference wrote (names incl. `chr<NAME>` of re-appearing chromosomes),
the same matrix size n(n+1)/2 (sort + de-duplicate of the IDs), and the same `.id` rows in the same order."""
import json
import os
import random
import shutil
import subprocess
from pathlib import Path

import numpy as np
import pytest

from gelex_b200 import engine

REF = Path(__file__).resolve().parent.parent / "oracle" / "_ref" / "gelex_ref"
pytestmark = pytest.mark.skipif(not REF.exists(), reason="oracle/_ref/gelex_ref not built (needs /root/reference)")


def make_case(rng, d: Path):
    n, m = rng.randint(1, 9), rng.randint(1, 14)
    fam_delim = rng.choice(["\t", " ", " ", "  "])
    lines = []
    for _ in range(n):
        cols = [rng.choice(["f1", "f2", "F", "fam_a", "a.b", "10", "9"]), rng.choice(["i1", "i2", "s01", "s1", "x_y", "2", "11"]), "0", "0",
                str(rng.randint(0, 2)), "-9"]
        q = rng.random()
        if q < 0.04:
            cols = cols[:rng.randint(0, 1)]
        elif q < 0.08:
            cols = cols[:rng.randint(2, 5)]
        elif q < 0.12:
            cols += ["extra"]
        line = fam_delim.join(cols)
        if rng.random() < 0.05:
            line = " " + line
        if rng.random() < 0.05:
            line += fam_delim
        if rng.random() < 0.04:
            line = line.replace(fam_delim, "\t" if fam_delim != "\t" else " ", 1)  # the other delimiter inside a line
        if rng.random() < 0.03:
            line = "   "
        if rng.random() < 0.03:
            line = "# " + line
        if rng.random() < 0.03:
            line = line.replace("s", "\u00e9", 1)
        lines.append(line)
    if rng.random() < 0.06:
        lines.insert(rng.randint(0, len(lines)), "")
    eol = "\r\n" if rng.random() < 0.06 else "\n"
    fam = eol.join(lines) + (eol if rng.random() < 0.85 else "")
    if rng.random() < 0.03:
        fam = ""
    bim_delim = rng.choice(["\t", "\t", " "])
    c, bl = rng.choice(["1", "2", "X", "chr1", "23"]), []
    for j in range(m):
        if rng.random() < 0.3:
            c = rng.choice(["1", "2", "3", "X", "chr1", "23", "MT"])
        cols = [c, f"rs{j}", "0", str(100 + j), "A", "G"]
        q = rng.random()
        if q < 0.03:
            cols = cols[:rng.randint(1, 5)]
        elif q < 0.05:
            cols += ["x"]
        elif q < 0.07:
            cols[3] = "notint"
        bl.append(bim_delim.join(cols))
    if rng.random() < 0.04:
        bl.insert(rng.randint(0, len(bl)), "")
    beol = "\r\n" if rng.random() < 0.08 else "\n"
    if rng.random() < 0.05 and bl:
        bl[rng.randrange(len(bl))] = "  "
    bim = beol.join(bl) + (beol if rng.random() < 0.9 else "")
    if rng.random() < 0.03:
        bim = ""
    payload = bytes(rng.randrange(256) for _ in range(m * ((n + 3) // 4)))
    magic = bytes([0x6C, 0x1B, 0x01])
    q = rng.random()
    if q < 0.03:
        magic = bytes([0x6C, 0x1B, 0x00])  # sample-major
    elif q < 0.05:
        magic = bytes([0x6C, 0x1A, 0x01])
    elif q < 0.08:
        payload = payload[:-1]
    elif q < 0.10:
        payload += b"\0\0"
    (d / "t.fam").write_text(fam, newline="")
    (d / "t.bim").write_text(bim, newline="")
    (d / "t.bed").write_bytes(magic + payload)
    return fam


@pytest.mark.parametrize("seed", [1, 2, 3])
def test_front_half_agrees_with_the_reference_on_quirky_inputs(tmp_path, seed):
    rng = random.Random(seed)
    accepted = rejected = 0
    for k in range(90):
        d = tmp_path / f"c{k}"
        d.mkdir()
        fam = make_case(rng, d)
        loco, mode = rng.random() < 0.5, rng.choice(["A", "D", "AD"])
        chunk = rng.choice([5, 5, 5, 5, 5, 5, 0, -3])  # the engine (unlike the CLI) meets a chunk size <= 0 at its first load_chunk
        r = subprocess.run([str(REF), "grm", str(d / "t"), str(d / "ref"), "1", mode, "1" if loco else "0", str(chunk), "1"], capture_output=True)
        ref_msg = r.stderr.decode().rstrip("\n").replace("error: ", "", 1)  # bytes: a '\r' inside a message must survive
        ref_files = sorted(p.name for p in d.iterdir() if p.name.startswith("ref."))
        try:
            engine.grm_run(d / "t", d / "mine", method="SH", mode=mode, loco=loco, chunk_size=chunk)
            mine_err = None  # (a GPU is present and the run went through)
        except engine.EngineError as e:
            mine_err = None if "no CUDA device" in e.message else e.message
        ctx = f"seed {seed} case {k}: fam={fam!r} bim={(d / 't.bim').read_text()!r} loco={loco} mode={mode} chunk={chunk}"
        assert (r.returncode != 0) == (mine_err is not None), f"{ctx}\nreference: rc={r.returncode} {ref_msg!r}\nhere: {mine_err!r}"
        if r.returncode != 0:
            rejected += 1
            assert mine_err.split(": ", 1)[1] == ref_msg, ctx  # "<ExceptionName>: <reference message>"
            shutil.rmtree(d)
            continue
        accepted += 1
        p = engine.plan(d / "t", mode=mode, loco=loco)
        assert sorted({f"ref.{name}.{ext}" for name, _, _ in p["items"] for ext in ("bin", "id")}) == ref_files, ctx
        # the events the reference's reporter received (grm_event.h:28-67) against what this engine's plan emits: dataset summary,
        # total work, files-written summary, and the per-chunk progress stream (one event per work item and chunk, `current` counted
        # inside the item - grm.h:124-130; the one-pass engine emits the same events, add and dom interleaved)
        ev = json.loads(r.stdout.decode().strip().splitlines()[-1])
        assert (ev["n"], ev["m"]) == (p["n_out"], p["m"]), ctx
        assert ev["total_snps"] == sum(e - b for _, b, e in p["items"]), ctx
        assert ev["num_files"] == 2 * len(p["items"]), ctx
        assert ev["file_pattern"] == str(d / "ref") + p["pattern_suffix"], ctx
        currents = [min(s + chunk, e) - b for _, b, e in p["items"] for s in range(b, e, chunk)]
        assert (ev["progress_events"], ev["progress_sum"]) == (len(currents), sum(currents)), ctx
        for f in ref_files:
            if f.endswith(".bin"):
                assert (d / f).stat().st_size == p["n_out"] * (p["n_out"] + 1) // 2 * 4, ctx
        # .id rows: row t = the FAM line i with raw_to_out[i] == t
        delim = "\t" if "\t" in fam.split("\n")[0] else " "
        raw = [ln for ln in fam.split("\n") if ln != ""]
        assert len(raw) == p["n_raw"], ctx
        rows = {}
        for i, ln in enumerate(raw):
            # parse_id (src/io/parser.cpp:63-82): FID = up to the first delimiter, IID = up to the next one; nothing is trimmed
            # (the '\r' of a CRLF file stays in a last column, "a.b\t\r" is FID "a.b" with IID "\r")
            p1 = ln.find(delim)
            p2 = ln.find(delim, p1 + 1)
            if p["raw_to_out"][i] >= 0:
                rows[int(p["raw_to_out"][i])] = ln[:p1] + "\t" + (ln[p1 + 1:p2] if p2 >= 0 else ln[p1 + 1:]) + "\n"
        ref_id = (d / [f for f in ref_files if f.endswith(".id")][0]).read_bytes().decode()
        assert "".join(rows[t] for t in range(p["n_out"])) == ref_id, ctx
        shutil.rmtree(d)
    assert accepted >= 8 and rejected >= 10  # the generator exercises both sides


def _reference(cmd):
    r = subprocess.run([str(REF)] + [str(c) for c in cmd], capture_output=True)  # bytes: '\r' inside a message must survive
    return r.returncode, r.stderr.decode().rstrip("\n").replace("error: ", "", 1)


def _here(call):
    """None when validation passed (the call then needs a GPU, or ran on one), else '<ExceptionName>: message'."""
    try:
        call()
        return None
    except engine.EngineError as e:
        # past validation the first device call fails here: ggrm_create reports "no CUDA device ... no CPU fallback", a bare
        # pinned allocation reports the runtime's own "CUDA: ..." error
        return None if "no CUDA device" in e.message or e.message.startswith("GelexException: CUDA:") else e.message


@pytest.mark.parametrize("seed", [1, 2])
def test_grm_loader_rejects_what_the_reference_rejects(tmp_path, seed):
    # GrmLoader(prefix) on damaged .id / .bin pairs (grm_loader.cpp:48-123): lines without a tab, empty FID or IID, three columns,
    # blank lines, CRLF, empty and missing files, payloads one sample too long / short or cut mid-float
    rng = random.Random(seed)
    rejected = 0
    for k in range(150):
        d = tmp_path / f"l{k}"
        d.mkdir()
        n = rng.randint(0, 6)
        lines = []
        for i in range(n):
            cols = [rng.choice(["F1", "F2", "a b", "F3"] * 3 + [""]), rng.choice(["s1", "s2", "s3", "x"]) + str(i if rng.random() < 0.8 else 0)]
            q = rng.random()
            if q < 0.06:
                cols = cols[:1]
            elif q < 0.12:
                cols += ["z"]
            lines.append(("\t" if rng.random() < 0.9 else " ").join(cols))
        if rng.random() < 0.08:
            lines.insert(rng.randint(0, len(lines)), "")
        ids = "\n".join(lines) + ("\n" if rng.random() < 0.9 else "")
        if rng.random() < 0.05 and lines:
            ids = ids.replace("\n", "\r\n")
        nn = n + 1 if rng.random() < 0.1 else (max(0, n - 1) if rng.random() < 0.1 else n)
        b = np.array([rng.uniform(0.5, 2.0) for _ in range(nn * (nn + 1) // 2)], dtype=np.float32).tobytes()
        q = rng.random()
        if q < 0.06:
            b = b[:-1]
        elif q < 0.1:
            b += b"\0\0\0"
        if rng.random() < 0.95:
            (d / "g.bin").write_bytes(b)
        if rng.random() < 0.95:
            (d / "g.id").write_text(ids, newline="")
        rc, ref_msg = _reference(["load", d / "g", d / "ref.out"])
        mine = _here(lambda: engine.load_grm(d / "g"))
        ctx = f"seed {seed} case {k}: ids={ids!r} bin bytes={len(b)}"
        assert (rc != 0) == (mine is not None), f"{ctx}\nreference: {ref_msg!r}\nhere: {mine!r}"
        if rc != 0:
            rejected += 1
            assert mine.split(": ", 1)[1] == ref_msg, ctx
        shutil.rmtree(d)
    assert rejected >= 30


@pytest.mark.parametrize("seed", [1, 2])
def test_loco_loader_rejects_what_the_reference_rejects(tmp_path, seed):
    # LocoGRMLoader(whole, id_map).load_loco_grm(chr, id_map) (loco_grm_loader.cpp:27-77): chromosome file with other / fewer / more /
    # reordered samples, damaged payloads, missing files, and k_chr >= k_whole - the denominators in that message are the
    # reference's own digits, so trace / n is computed the same way on the host
    rng = random.Random(seed)
    rejected = 0

    def write(prefix, ids, scale, n_bin=None, corrupt=False):
        n = len(ids) if n_bin is None else n_bin
        v = [(scale if i == j else 0.1 * scale) * rng.uniform(0.9, 1.1) for i in range(n) for j in range(i + 1)]
        b = np.array(v, dtype=np.float32).tobytes()
        Path(str(prefix) + ".bin").write_bytes(b[:-2] if corrupt else b)
        Path(str(prefix) + ".id").write_text("".join(f"{f}\t{i}\n" for f, i in ids))

    for k in range(150):
        d = tmp_path / f"o{k}"
        d.mkdir()
        ids = [(rng.choice(["F1", "F2"]), f"s{i}") for i in range(rng.randint(1, 6))]
        cids = list(ids)
        q = rng.random()
        if q < 0.15:
            rng.shuffle(cids)
        elif q < 0.3:
            cids = cids[:-1] if len(cids) > 1 else cids
        elif q < 0.4:
            cids = cids + [("F9", "new")]
        elif q < 0.5:
            cids[0] = ("F9", "other")
        if rng.random() < 0.05:
            cids = cids + [cids[0]]
        write(d / "w", ids, 10.0, corrupt=rng.random() < 0.05)
        write(d / "c", cids, rng.choice([1.0, 1.0, 1.0, 10.0, 12.0]), n_bin=(len(cids) + 1 if rng.random() < 0.05 else None), corrupt=rng.random() < 0.05)
        if rng.random() < 0.04:
            os.unlink(d / "c.bin")
        if rng.random() < 0.04:
            os.unlink(d / "c.id")
        rc, ref_msg = _reference(["loco", d / "w", d / "c", d / "ref.out"])
        mine = _here(lambda: engine.load_grm(d / "w", chr_prefix=d / "c", id_map={fi: t for t, fi in enumerate(ids)}))
        ctx = f"seed {seed} case {k}: ids={ids} chr ids={cids}"
        assert (rc != 0) == (mine is not None), f"{ctx}\nreference: {ref_msg!r}\nhere: {mine!r}"
        if rc != 0:
            rejected += 1
            assert mine.split(": ", 1)[1] == ref_msg, ctx
        shutil.rmtree(d)
    assert rejected >= 30


@pytest.mark.parametrize("seed", [1, 2, 3])
def test_oracle_agrees_with_the_reference_on_random_small_inputs(tmp_path, seed):
    # The CPU oracle (oracle/grm_oracle.c, what the GPU parity tests compare against) beside `gelex grm` of the unmodified reference:
    # random small matrices with the cases the reference's own GRM tests leave out (missing data through GRM::compute,
    # n not a multiple of 4, duplicate and unsorted FAM ids) plus monomorphic, all-heterozygous and all-missing SNPs (the last one
    # poisons the matrix with NaN in the reference, and must here too), every code policy, additive and dominance, ragged chunks.
    import oracle_lib as o
    import plink_io

    rng = np.random.default_rng(seed)
    for k in range(24):
        n_raw, m = int(rng.integers(1, 24)), int(rng.integers(1, 40))
        p = rng.uniform(0.02, 0.98, size=m)
        geno = rng.binomial(2, p, size=(n_raw, m)).astype(np.float64)
        geno[rng.random((n_raw, m)) < rng.choice([0.0, 0.05, 0.4])] = np.nan
        for j in range(m):
            q = rng.random()
            if q < 0.06:
                geno[:, j] = rng.integers(0, 3)  # monomorphic
            elif q < 0.09:
                geno[:, j] = 1.0                 # all heterozygous
            elif q < 0.11:
                geno[:, j] = np.nan              # all missing
        fam = [(f"f{int(rng.integers(0, 3))}", f"s{int(rng.integers(0, max(2, n_raw)))}" if rng.random() < 0.3 else f"s{i:03d}") for i in range(n_raw)]
        order = rng.permutation(n_raw)
        fam = [fam[i] for i in order]
        payload = plink_io.encode_matrix(geno)
        plink_io.write_plink(tmp_path / f"t{k}", payload, fam, ["1"] * m, fam_delim=str(rng.choice(["\t", " "])))
        ids, r2o = plink_io.sample_order(fam)
        method, kind, chunk = int(rng.integers(1, 9)), int(rng.integers(0, 2)), int(rng.choice([1, 3, 7, 10000]))
        rc, msg = _reference(["grm", tmp_path / f"t{k}", tmp_path / f"r{k}", method, "AD"[kind], 0, chunk, 1])
        assert rc == 0, msg
        ref = plink_io.read_bin(tmp_path / f"r{k}.{('add', 'dom')[kind]}.bin", len(ids))
        G, _, _ = o.grm_compute(payload, n_raw, method, kind, chunk=chunk, n_out=len(ids), raw_to_out=r2o)
        got = o.pack_lower_f32(G)
        ctx = f"seed {seed} case {k}: n_raw={n_raw} n={len(ids)} m={m} method={method} kind={kind} chunk={chunk}"
        assert np.array_equal(np.isnan(got), np.isnan(ref)), ctx
        ok = ~np.isnan(ref)
        # float32 file values: identical after rounding, or one float32 ulp apart (BLAS summation order); entries that cancel to zero
        # carry the FP64 rounding noise of the sum (|x| ~ 1e-17 against 0 exactly), hence the floor relative to the largest entry
        floor = 1e-14 * float(np.max(np.abs(ref[ok]))) if ok.any() else 0.0
        assert np.all(np.abs(got[ok].astype(np.float64) - ref[ok]) <= np.spacing(np.abs(ref[ok])) + floor), ctx
        assert (tmp_path / f"r{k}.{('add', 'dom')[kind]}.id").read_text() == "".join(s.replace("\x1f", "\t") + "\n" for s in ids), ctx


@pytest.mark.parametrize("seed", [1, 2])
def test_genotype_pipe_front_half_agrees_with_the_reference(tmp_path, seed):
    # GenotypePipe(bed, SampleManager::create_finalized(bed), out_prefix) (genotype_pipe.cpp:51-82): the same quirky triples, plus
    # outputs that already exist - accepted / rejected alike, the reference's message, in the reference's order (inputs first, then
    # the existing-files check)
    rng = random.Random(1000 + seed)
    accepted = rejected = 0
    for k in range(60):
        d = tmp_path / f"p{k}"
        d.mkdir()
        fam = make_case(rng, d)
        q = rng.random()
        if q < 0.15:
            (d / "out.bmat").write_bytes(b"")
        elif q < 0.3:
            (d / "out.snpstats").write_bytes(b"x")
        before = {p.name for p in d.iterdir()}
        rc, ref_msg = _reference(["pipe", d / "t", d / "out", 3, "A", 5])
        for p in list(d.iterdir()):  # same prefix for both (it is part of the message): drop what the reference wrote
            if p.name not in before:
                p.unlink()
        mine = _here(lambda: engine.genotype_export(d / "t", d / "out", method="OSH", additive=True, chunk_size=5))
        ctx = f"seed {seed} case {k}: fam={fam!r} bim={(d / 't.bim').read_text()!r}"
        assert (rc != 0) == (mine is not None), f"{ctx}\nreference: {ref_msg!r}\nhere: {mine!r}"
        if rc != 0:
            rejected += 1
            assert mine.split(": ", 1)[1] == ref_msg, ctx
        else:
            accepted += 1
        shutil.rmtree(d)
    assert accepted >= 5 and rejected >= 10


@pytest.mark.parametrize("seed", [1, 2])
def test_oracle_loco_partials_agree_with_the_reference(tmp_path, seed):
    # `gelex grm --loco` of the reference on random small inputs whose chromosomes re-appear (a later run of the same name
    # overwrites the earlier file, grm_work_plan.cpp:54-81 + grm_engine.cpp:88-110) beside the oracle run over the same SNP
    # ranges; AD mode writes both matrices per run
    import oracle_lib as o
    import plink_io

    rng = np.random.default_rng(100 + seed)
    for k in range(10):
        n_raw, m = int(rng.integers(2, 20)), int(rng.integers(3, 40))
        geno = rng.binomial(2, rng.uniform(0.05, 0.95, size=m), size=(n_raw, m)).astype(np.float64)
        geno[rng.random((n_raw, m)) < 0.1] = np.nan
        chroms, c = [], "1"
        for _ in range(m):
            if rng.random() < 0.25:
                c = str(rng.choice(["1", "2", "X", "7"]))
            chroms.append(c)
        fam = [("F", f"s{i:03d}") for i in rng.permutation(n_raw)]
        payload = plink_io.encode_matrix(geno)
        plink_io.write_plink(tmp_path / f"t{k}", payload, fam, chroms)
        ids, r2o = plink_io.sample_order(fam)
        method, chunk = int(rng.integers(1, 9)), int(rng.choice([2, 5, 10000]))
        rc, msg = _reference(["grm", tmp_path / f"t{k}", tmp_path / f"r{k}", method, "AD", 1, chunk, 1])
        assert rc == 0, msg
        last = {}
        for name, b, e in plink_io.chrom_runs(chroms):
            last[name] = (b, e)  # the file that survives is the last run of that name
        written = sorted(p.name for p in tmp_path.iterdir() if p.name.startswith(f"r{k}.") and p.name.endswith(".bin"))
        assert written == sorted(f"r{k}.{kd}.chr{name}.bin" for name in last for kd in ("add", "dom"))
        for name, (b, e) in last.items():
            for kind, kd in ((0, "add"), (1, "dom")):
                ref = plink_io.read_bin(tmp_path / f"r{k}.{kd}.chr{name}.bin", len(ids))
                G, _, _ = o.grm_compute(payload, n_raw, method, kind, ranges=[(b, e)], chunk=chunk, n_out=len(ids), raw_to_out=r2o)
                got = o.pack_lower_f32(G)
                ctx = f"seed {seed} case {k} chr {name} [{b},{e}) {kd}: n={len(ids)} m={m} method={method}"
                assert np.array_equal(np.isnan(got), np.isnan(ref)), ctx
                ok = ~np.isnan(ref)
                floor = 1e-14 * float(np.max(np.abs(ref[ok]))) if ok.any() else 0.0
                assert np.all(np.abs(got[ok].astype(np.float64) - ref[ok]) <= np.spacing(np.abs(ref[ok])) + floor), ctx


def test_reference_side_binding_builds_and_reaches_the_c_abi(tmp_path):
    # oracle/ref_binding.cpp = INTEGRATION.md section 2 compiled against the unmodified reference headers and linked with its own
    # loaders / planners / writers plus libggrm_b200.so.  Without a GPU the reference's front half still runs (and rejects what it
    # rejects) and the first ABI call fails loudly; with one, tests/test_host_engine.py compares its files with the reference's.
    exe = REF.parent / "gelex_ref_b200"
    if not exe.exists():
        pytest.skip("oracle/_ref/gelex_ref_b200 not built")
    golden = Path(__file__).resolve().parent / "golden"
    r = subprocess.run([str(exe), str(golden / "tiny"), str(tmp_path / "o"), "1", "A", "0", "25", "fp64"], capture_output=True, text=True)
    if r.returncode == 0:  # a GPU is present
        assert (tmp_path / "o.add.id").read_bytes() == (golden / "tiny_SH.add.id").read_bytes()
    else:
        assert "no CUDA device visible; the B200 GRM engine has no CPU fallback" in r.stderr
    r = subprocess.run([str(exe), str(tmp_path / "missing"), str(tmp_path / "o"), "1", "A", "0", "25", "fp64"], capture_output=True, text=True)
    assert r.returncode == 1 and "not found" in r.stderr  # the reference's own FileNotFoundException, before any ABI call
