"""Host-side logic of the C++ engine (FAM/BIM parsing, row order, work plan, file validation, error behaviour).
CPU-only parts run everywhere; the end-to-end file parity tests need a B200 and are marked gpu."""
import os
import shutil
import subprocess
from pathlib import Path

import numpy as np
import pytest

import gelex_b200 as g
from gelex_b200 import engine
import plink_io


def test_plan_matches_reference_row_order_and_items(golden_dir):
    p = engine.plan(golden_dir / "tiny", mode="AD", loco=True)
    fam = [tuple(l.split()[:2]) for l in open(golden_dir / "tiny.fam")]
    ids, r2o = plink_io.sample_order(fam)
    assert p["n_raw"] == 37 and p["n_out"] == 36 and p["m"] == 96 and not p["dense"]
    assert np.array_equal(p["raw_to_out"], r2o)
    # GrmLocoPlan: for each chromosome run, add then dom; names add.chr<NAME> (grm_work_plan.cpp:122-147)
    assert [i[0] for i in p["items"]] == ["add.chr1", "dom.chr1", "add.chr2", "dom.chr2", "add.chrX", "dom.chrX", "add.chr7", "dom.chr7"]
    assert [i[1:] for i in p["items"][::2]] == [(0, 30), (30, 55), (55, 72), (72, 96)]
    assert p["pattern_suffix"] == ".{add|dom}.chr{1..4}.{bin|id}"
    q = engine.plan(golden_dir / "dense", mode="D", loco=False)
    assert q["dense"] and q["items"] == [("dom", 0, 50)] and q["pattern_suffix"] == ".dom.{bin|id}"


def test_fam_delimiter_detection_and_errors(tmp_path):
    payload = plink_io.encode_matrix(np.zeros((3, 2)))
    plink_io.write_plink(tmp_path / "t", payload, [("f b", "i1"), ("f a", "i2"), ("f a", "i0")], ["1", "1"], fam_delim="\t")
    p = engine.plan(tmp_path / "t")  # tab-delimited: FID may contain a space (parser.cpp:105-116)
    assert p["raw_to_out"].tolist() == [2, 1, 0]
    (tmp_path / "bad.fam").write_text("onlyonefield\n")
    (tmp_path / "bad.bim").write_text("1\ts\t0\t1\tA\tG\n")
    (tmp_path / "bad.bed").write_bytes(bytes([0x6C, 0x1B, 0x01, 0]))
    with pytest.raises(engine.EngineError) as e:
        engine.plan(tmp_path / "bad")
    assert "1: failed to parse FID and IID (missing delimiter)" in e.value.message
    (tmp_path / "bad.fam").write_text("f\ti\t0\t0\t1\t-9\n")
    (tmp_path / "bad.bim").write_text("1\ts\t0\t1\tA\n")
    with pytest.raises(engine.EngineError) as e:
        engine.plan(tmp_path / "bad")
    assert "1: has 5 columns, expected 6" in e.value.message  # bim_loader.cpp:69-74
    with pytest.raises(engine.EngineError) as e:
        engine.plan(tmp_path / "missing")
    assert "file not found" in e.value.message  # bed_path.cpp:33-37


def test_bim_and_fam_cases_of_the_reference_loader_tests(tmp_path):
    # reference tests/test_bim_loader.cpp:110-152 (column count, non-numeric / empty position), :74-107 (space-delimited),
    # tests/test_fam_loader.cpp:90-104 (missing IID), :139-165 (single line, space or tab delimited)
    bed = bytes([0x6C, 0x1B, 0x01, 0])

    def triple(name, fam, bim):
        (tmp_path / f"{name}.fam").write_text(fam)
        (tmp_path / f"{name}.bim").write_text(bim)
        (tmp_path / f"{name}.bed").write_bytes(bed[:3] + bytes(max(1, bim.count("\n") + (0 if bim.endswith("\n") else 1))))
        return tmp_path / name

    for bim, msg in (("1\trs12345\t0\t1000\tA\tG\n1\trs67890\t0.001\t2000\tC", "2: has 5 columns, expected 6"),
                     ("1\trs12345\t0\tinvalid\tA\tG", "failed to parse 'invalid' as number"),
                     ("1\trs12345\t0\t\tA\tG", "has 5 columns, expected 6")):
        with pytest.raises(engine.EngineError) as e:
            engine.plan(triple("b", "1 sample1 0 0 1 2.5\n", bim))
        assert e.value.message.startswith("FileFormatException") and e.value.message.endswith(msg), e.value.message
    with pytest.raises(engine.EngineError) as e:
        engine.plan(triple("f", "1\n2 sample2\n", "1\ts\t0\t1\tA\tG\n"))
    assert e.value.message.startswith("FileFormatException") and e.value.message.endswith("failed to parse FID and IID (missing delimiter)")
    for fam in ("1 sample1 0 0 1 2.5\n", "1\tsample1\t0\t0\t1\t2.5\n"):
        p = engine.plan(triple("one", fam, "1 rs1 0 1000 A G\n1 rs2 0.001 2000 C T\n2 rs3 0.002 3000 T A\n"), loco=True)  # space-delimited BIM
        assert p["n_raw"] == 1 and p["n_out"] == 1 and p["m"] == 3 and p["dense"]
        assert p["items"] == [("add.chr1", 0, 2), ("add.chr2", 2, 3)]


def test_id_writer_cases_of_the_reference_tests(tmp_path):
    US = "\x1f"
    p = tmp_path / "t.grm.id"
    engine.write_id_file(p, [])  # test_grm_id_writer.cpp:118-133: empty list -> the file exists and is empty
    assert p.exists() and p.stat().st_size == 0
    engine.write_id_file(p, [f"FAM1{US}IND1"])  # :139-150
    assert p.read_bytes() == b"FAM1\tIND1\n"
    engine.write_id_file(p, [f"FAM1{US}IND1", f"FAM2{US}IND2", f"FAM3{US}IND3"])  # :152-168 (and: an existing file is truncated, :100-115)
    assert p.read_bytes() == b"FAM1\tIND1\nFAM2\tIND2\nFAM3\tIND3\n"
    engine.write_id_file(p, [f"FAM_A{US}IND_1", f"F.1-x y{US}I 2"])  # :170-186, :390-406: underscores, dots, dashes, spaces pass through
    assert p.read_bytes() == b"FAM_A\tIND_1\nF.1-x y\tI 2\n"
    for ids, msg in (([ "FAM1_IND1"], "canonical FID<US>IID format"), (["NOSPLIT"], "canonical FID<US>IID format"), ([""], "cannot be empty"),
                     ([f"{US}IND1"], "non-empty FID and IID"), ([f"FAM1{US}"], "non-empty FID and IID")):  # :188-253
        with pytest.raises(engine.EngineError, match="ArgumentValidationException: .*" + msg):
            engine.write_id_file(p, ids)
    ids = [f"FAM{i}{US}IND{i}" for i in range(10000)]  # :339-365
    engine.write_id_file(p, ids)
    lines = p.read_text().split("\n")
    assert lines[-1] == "" and len(lines) == 10001 and lines[0] == "FAM0\tIND0" and lines[9999] == "FAM9999\tIND9999"
    with pytest.raises(engine.EngineError, match="FileOpenException"):
        engine.write_id_file(tmp_path / "missing_dir" / "x.id", ids[:1])


def test_format_bed_path_cases_of_the_reference_tests(tmp_path):
    bed = tmp_path / "cohort.bed"  # test_bed_path.cpp:37-78
    bed.write_bytes(b"")
    assert engine.format_bed_path(bed) == str(bed)
    assert engine.format_bed_path(tmp_path / "cohort") == str(bed)
    tricky = tmp_path / "my.bedsets"
    tricky.mkdir()
    (tricky / "cohort1.bed").write_bytes(b"")
    assert engine.format_bed_path(tricky / "cohort1") == str(tricky / "cohort1.bed")
    with pytest.raises(engine.EngineError, match=r"FileNotFoundException: .*missing_prefix\.bed: file not found$"):
        engine.format_bed_path(tmp_path / "missing_prefix")


def test_bed_validation_errors_precede_any_gpu_work(tmp_path):
    # reference tests/test_bed_pipe.cpp:154-257: bad magic, truncated file -> FileFormatException
    payload = plink_io.encode_matrix(np.zeros((8, 5)))
    plink_io.write_plink(tmp_path / "ok", payload, [("F", f"s{i}") for i in range(8)], ["1"] * 5)
    raw = (tmp_path / "ok.bed").read_bytes()
    for name, content, msg in (("magic", bytes([0x6C, 0x1B, 0x00]) + raw[3:], "invalid BED magic number"),
                               ("trunc", raw[:-3], "bed file truncated. Expected 13 bytes, got 10"),
                               ("short", raw[:3], "bed file too short")):
        for ext in ("bim", "fam"):
            shutil.copy(tmp_path / f"ok.{ext}", tmp_path / f"{name}.{ext}")
        (tmp_path / f"{name}.bed").write_bytes(content)
        with pytest.raises(engine.EngineError) as e:
            engine.grm_run(tmp_path / name, tmp_path / "out")
        assert e.value.message.startswith("FileFormatException") and msg in e.value.message
    (tmp_path / "nobim.bed").write_bytes(raw)
    shutil.copy(tmp_path / "ok.fam", tmp_path / "nobim.fam")
    with pytest.raises(engine.EngineError) as e:
        engine.grm_run(tmp_path / "nobim", tmp_path / "out")
    assert "nobim.bim: not found" in e.value.message


def test_engine_fails_loudly_without_gpu(golden_dir, tmp_path):
    if g.device_count() > 0:
        pytest.skip("a GPU is present")
    with pytest.raises(engine.EngineError) as e:
        engine.grm_run(golden_dir / "tiny", tmp_path / "o")
    assert "no CPU fallback" in e.value.message
    r = subprocess.run([str(engine.CLI_PATH), "grm", "-b", str(golden_dir / "tiny"), "-o", str(tmp_path / "o")], capture_output=True, text=True)
    assert r.returncode == 1 and "no CPU fallback" in r.stderr  # apps/main.cpp:73-98: log the error, exit code 1


# ------------------------------------------------------------------------------------------------ end to end (GPU)
def assert_bin_close(path, ref_path, n, ulps=1):
    got, ref = plink_io.read_bin(path, n), plink_io.read_bin(ref_path, n)
    assert np.array_equal(np.isnan(got), np.isnan(ref))
    m = ~np.isnan(ref)
    assert np.all(np.abs(got[m] - ref[m]) <= ulps * np.spacing(np.abs(ref[m])) + 1e-30), path


@pytest.mark.gpu
def test_engine_files_match_reference_outputs(golden_dir, tmp_path):
    n = 36
    engine.grm_run(golden_dir / "tiny", tmp_path / "SH", method="SH", mode="A", chunk_size=25, precision=g.PREC_FP64)
    assert_bin_close(tmp_path / "SH.add.bin", golden_dir / "tiny_SH.add.bin", n)
    assert (tmp_path / "SH.add.id").read_bytes() == (golden_dir / "tiny_SH.add.id").read_bytes()
    engine.grm_run(golden_dir / "tiny", tmp_path / "OSH", method="OSH", mode="AD", precision=g.PREC_FP64)  # one decode, two SYRKs
    for k in ("add", "dom"):
        assert_bin_close(tmp_path / f"OSH.{k}.bin", golden_dir / f"tiny_OSH.{k}.bin", n)
        assert (tmp_path / f"OSH.{k}.id").read_bytes() == (golden_dir / f"tiny_OSH.{k}.id").read_bytes()
    engine.grm_run(golden_dir / "dense", tmp_path / "dSH", method=1, mode="A", precision=g.PREC_FP64)  # dense (identity) projection
    assert_bin_close(tmp_path / "dSH.add.bin", golden_dir / "dense_SH.add.bin", 40)


@pytest.mark.gpu
def test_engine_loco_one_pass_matches_reference_two_invocations(golden_dir, tmp_path):
    n = 36
    engine.grm_run(golden_dir / "tiny", tmp_path / "OCH", method="OCH", mode="AD", loco=True, chunk_size=7, precision=g.PREC_FP64, write_whole=True)
    names = sorted(p.name for p in tmp_path.iterdir())
    expect = sorted(f"OCH.{k}{c}.{e}" for k in ("add", "dom") for c in ("", ".chr1", ".chr2", ".chrX", ".chr7") for e in ("bin", "id"))
    assert names == expect
    for k in ("add", "dom"):
        for c in ("1", "2", "X", "7"):
            assert_bin_close(tmp_path / f"OCH.{k}.chr{c}.bin", golden_dir / f"tiny_OCH.{k}.chr{c}.bin", n)
        assert_bin_close(tmp_path / f"OCH.{k}.bin", golden_dir / f"tiny_OCH.{k}.bin", n, ulps=4)  # whole = sum of 4 partials
    # without --with-whole the file set is exactly the reference's (no whole-genome files)
    out2 = tmp_path / "ref_like"
    out2.mkdir()
    engine.grm_run(golden_dir / "tiny", out2 / "g", method="OCH", mode="A", loco=True, precision=g.PREC_FP64)
    assert sorted(p.name for p in out2.iterdir()) == sorted(f"g.add.chr{c}.{e}" for c in ("1", "2", "X", "7") for e in ("bin", "id"))


@pytest.mark.gpu
@pytest.mark.parametrize("n_gpus", [2, 3])
def test_engine_multi_gpu_matches_single_gpu(tmp_path, monkeypatch, n_gpus):
    # On a box with fewer GPUs the ranks share devices (GELEX_B200_OVERSUBSCRIBE, capi.cpp): same handles, same peer reduce,
    # same concurrent writers - only the NVLink hop is missing.
    monkeypatch.setenv("GELEX_B200_OVERSUBSCRIBE", "1")
    n, m = 4500, 3000  # 3 super-rows of 2048: uneven row slices, 6 super-tiles, ragged last one
    payload = g.synth_bed_host(n, 0, m, 17, 0.01)
    bounds = [round(m * i / 5) for i in range(6)]
    chroms = [str(c + 1) for c in range(5) for _ in range(bounds[c + 1] - bounds[c])]
    plink_io.write_plink(tmp_path / "d", payload, [("F", f"s{i:05d}") for i in range(n)], chroms)
    # SNP shards: partial GRMs reduce-scattered by row slices over peer memory (ggrm_reduce_peer), every GPU writes its own rows
    engine.grm_run(tmp_path / "d", tmp_path / "g1", method="SH", precision=g.PREC_FP64, n_gpus=1)
    engine.grm_run(tmp_path / "d", tmp_path / "g2", method="SH", precision=g.PREC_FP64, n_gpus=n_gpus, partition="snp")
    assert_bin_close(tmp_path / "g2.add.bin", tmp_path / "g1.add.bin", n, ulps=2)
    assert (tmp_path / "g2.add.id").read_bytes() == (tmp_path / "g1.add.id").read_bytes()
    # 2D output tiles: every GPU streams all SNPs, owns disjoint super-tiles and writes them at their own file offsets; the
    # per-entry arithmetic is the single-GPU one, so the bytes are identical
    engine.grm_run(tmp_path / "d", tmp_path / "g3", method="SH", precision=g.PREC_FP64, n_gpus=n_gpus, partition="tile")
    assert (tmp_path / "g3.add.bin").read_bytes() == (tmp_path / "g1.add.bin").read_bytes()
    assert (tmp_path / "g3.add.id").read_bytes() == (tmp_path / "g1.add.id").read_bytes()
    # LOCO: chromosome runs dealt to GPUs, whole = cross-GPU sum of per-GPU wholes
    engine.grm_run(tmp_path / "d", tmp_path / "l1", method="OCH", mode="AD", loco=True, precision=g.PREC_FP64, n_gpus=1, write_whole=True)
    engine.grm_run(tmp_path / "d", tmp_path / "l2", method="OCH", mode="AD", loco=True, precision=g.PREC_FP64, n_gpus=n_gpus, write_whole=True, partition="snp")
    engine.grm_run(tmp_path / "d", tmp_path / "l3", method="OCH", mode="AD", loco=True, precision=g.PREC_FP64, n_gpus=n_gpus, write_whole=True, partition="tile")
    for k in ("add", "dom"):
        for c in ("", ".chr1", ".chr2", ".chr3", ".chr4", ".chr5"):
            assert_bin_close(tmp_path / f"l2.{k}{c}.bin", tmp_path / f"l1.{k}{c}.bin", n, ulps=2 if c == "" else 0)
            assert (tmp_path / f"l3.{k}{c}.bin").read_bytes() == (tmp_path / f"l1.{k}{c}.bin").read_bytes()
    assert sorted(p.name for p in tmp_path.iterdir() if p.name.startswith("l3.")) == sorted(p.name.replace("l1.", "l3.") for p in tmp_path.iterdir() if p.name.startswith("l1."))


@pytest.mark.gpu
def test_engine_16bit_tile_mode_large_rows(tmp_path, monkeypatch):
    # tcgen05 path through the tile writer and the row-slice writer (FP32 accumulators): bytes identical to the single-GPU file in
    # tile mode; SNP shards differ only by FP32 summation order
    monkeypatch.setenv("GELEX_B200_OVERSUBSCRIBE", "1")
    n, m = 4100, 2048
    payload = g.synth_bed_host(n, 0, m, 23, 0.01)
    plink_io.write_plink(tmp_path / "d", payload, [("F", f"s{i:05d}") for i in range(n)], ["1"] * m)
    engine.grm_run(tmp_path / "d", tmp_path / "a1", method="SH", precision=g.PREC_FP16, n_gpus=1)
    engine.grm_run(tmp_path / "d", tmp_path / "a2", method="SH", precision=g.PREC_FP16, n_gpus=2, partition="tile")
    engine.grm_run(tmp_path / "d", tmp_path / "a3", method="SH", precision=g.PREC_FP16, n_gpus=2, partition="snp", writer_threads=1)
    assert (tmp_path / "a2.add.bin").read_bytes() == (tmp_path / "a1.add.bin").read_bytes()
    a1, a3 = plink_io.read_bin(tmp_path / "a1.add.bin", n).astype(np.float64), plink_io.read_bin(tmp_path / "a3.add.bin", n).astype(np.float64)
    assert np.max(np.abs(a3 - a1)) <= 1e-4 * np.max(np.abs(a1))


@pytest.mark.parametrize("writer", ["pwrite", "mmap", "auto"])
@pytest.mark.parametrize("tiles,parts", [(False, 1), (False, 3), (True, 1), (True, 3)])
def test_writer_pool_bytes_need_no_gpu(tmp_path, writer, tiles, parts):
    # The .bin writer of the engine (BinFile + WriterPool, grm_engine.cpp) on a synthetic matrix: whole-row blocks of `parts` row
    # slices (single GPU / SNP-shard read-out) and super-tile row segments dealt to `parts` concurrent pools (2D-tile read-out),
    # through pwrite and through the shared mapping: the reference's layout (grm_bin_writer.cpp:54-61: row-major lower triangle,
    # float32, no header), byte for byte.  n = 4500: three super-rows, ragged last super-tile; n = 1 and 5: a single block.
    for n in (1, 5, 4500):
        path = tmp_path / f"w{n}.bin"
        _, mapped = engine.writer_selftest(path, n, tiles=tiles, parts=parts, writer_threads=2, writer=writer)
        assert mapped == (writer != "pwrite")
        got = np.fromfile(path, dtype=np.float32)
        assert path.stat().st_size == n * (n + 1) // 2 * 4  # exactly n(n+1)/2 floats (test_grm_bin_writer.cpp:40-43)
        assert np.array_equal(got, engine.writer_selftest_expected(n))


@pytest.mark.parametrize("tiles,parts", [(False, 1), (False, 3), (True, 3)])
def test_writer_pool_after_prefault_needs_no_gpu(tmp_path, tiles, parts):
    # Config::prefault: the pages of the pre-sized file are populated (MADV_POPULATE_WRITE on a shared mapping, contents untouched)
    # and the writers then copy through THAT mapping (SharedMappings) - same bytes; with pwrite forced nothing is mapped or populated
    for n in (5, 4500):
        for writer in ("mmap", "auto", "pwrite"):
            path = tmp_path / f"p{n}.bin"
            _, mapped, pf_s = engine.writer_selftest(path, n, tiles=tiles, parts=parts, writer_threads=2, writer=writer, prefault_threads=3,
                                                     with_prefault_seconds=True)
            assert mapped == (writer != "pwrite") and (pf_s > 0) == (writer != "pwrite")
            assert np.array_equal(np.fromfile(path, dtype=np.float32), engine.writer_selftest_expected(n))


def test_writer_mode_env_and_validation_need_no_gpu(tmp_path, monkeypatch):
    monkeypatch.setenv("GELEX_B200_WRITER", "pwrite")  # the environment overrides Auto only
    assert engine.writer_selftest(tmp_path / "a.bin", 300, writer="auto")[1] is False
    assert engine.writer_selftest(tmp_path / "a.bin", 300, writer="mmap")[1] is True
    monkeypatch.setenv("GELEX_B200_WRITER", "floppy")
    with pytest.raises(engine.EngineError, match="ArgumentValidationException: GELEX_B200_WRITER must be auto, pwrite or mmap"):
        engine.writer_selftest(tmp_path / "a.bin", 300, writer="auto")
    monkeypatch.delenv("GELEX_B200_WRITER")
    with pytest.raises(engine.EngineError, match="FileOpenException"):
        engine.writer_selftest(tmp_path / "no_such_dir" / "a.bin", 300)
    with pytest.raises(engine.EngineError, match="ArgumentValidationException"):
        engine.writer_selftest(tmp_path / "a.bin", 0)


@pytest.mark.gpu
@pytest.mark.parametrize("partition", ["snp", "tile"])
def test_engine_writer_modes_give_identical_files(tmp_path, monkeypatch, partition):
    # same run through pwrite and through the shared mapping: identical bytes (two ranks on one device when the box has one GPU)
    monkeypatch.setenv("GELEX_B200_OVERSUBSCRIBE", "1")
    n, m = 4100, 1024
    payload = g.synth_bed_host(n, 0, m, 29, 0.01)
    plink_io.write_plink(tmp_path / "d", payload, [("F", f"s{i:05d}") for i in range(n)], ["1"] * m)
    for w in ("pwrite", "mmap"):
        monkeypatch.setenv("GELEX_B200_WRITER", w)
        engine.grm_run(tmp_path / "d", tmp_path / w, method="SH", precision=g.PREC_FP64, n_gpus=2, partition=partition)
    assert (tmp_path / "mmap.add.bin").read_bytes() == (tmp_path / "pwrite.add.bin").read_bytes()
    assert (tmp_path / "mmap.add.id").read_bytes() == (tmp_path / "pwrite.add.id").read_bytes()


@pytest.mark.parametrize("n_chunks,world,slots", [(0, 2, 2), (1, 3, 2), (37, 1, 1), (200, 2, 1), (333, 3, 2), (600, 8, 2)])
def test_bed_relay_protocol_needs_no_gpu(n_chunks, world, slots):
    # 2D-tile BED relay (grm_engine.cpp, ChunkRelay): `world` uploaders fill `slots` slots each, `world` consumers take every chunk
    # in order; a slot may only be refilled once every consumer has read it.  Host buffers stand in for the device staging slots;
    # uploaders and consumers are slowed unevenly.  0 = every consumer read every chunk's own bytes.
    assert engine.relay_selftest(n_chunks, world, slots, chunk_bytes=8192) == 0


def test_bed_relay_upload_failure_surfaces_instead_of_hanging():
    for at in (0, 5, 99):
        with pytest.raises(engine.EngineError, match="FileFormatException: relay_selftest: injected upload failure"):
            engine.relay_selftest(100, 4, 2, chunk_bytes=512, fail_at_chunk=at)
    with pytest.raises(engine.EngineError, match="ArgumentValidationException"):
        engine.relay_selftest(10, 0, 2)


@pytest.mark.gpu
@pytest.mark.parametrize("loco", [False, True])
def test_engine_tile_mode_bed_relay_matches_direct_push(tmp_path, monkeypatch, loco):
    # same 2D-tile run with every GPU reading the whole BED itself, and with one upload per chunk shared through staging slots
    # (ggrm_stage_* + ggrm_push_device + ggrm_wait_copies): identical bytes
    monkeypatch.setenv("GELEX_B200_OVERSUBSCRIBE", "1")
    n, m = 4100, 3000
    payload = g.synth_bed_host(n, 0, m, 31, 0.01)
    bounds = [round(m * i / 3) for i in range(4)]
    chroms = [str(c + 1) for c in range(3) for _ in range(bounds[c + 1] - bounds[c])]
    plink_io.write_plink(tmp_path / "d", payload, [("F", f"s{i:05d}") for i in range(n)], chroms)
    kw = dict(method="SH", precision=g.PREC_FP64, n_gpus=3, partition="tile", loco=loco, write_whole=loco, chunk_size=250)
    engine.grm_run(tmp_path / "d", tmp_path / "direct", **kw)
    monkeypatch.setenv("GELEX_B200_BED_RELAY", "1")
    engine.grm_run(tmp_path / "d", tmp_path / "relay", **kw)
    names = sorted(p.name for p in tmp_path.iterdir() if p.name.startswith("direct."))
    assert len(names) == (8 if loco else 2)
    for name in names:
        assert (tmp_path / name.replace("direct.", "relay.")).read_bytes() == (tmp_path / name).read_bytes(), name


@pytest.mark.gpu
@pytest.mark.parametrize("n_gpus", [1, 3])
def test_engine_loco_reappearing_chromosome_last_run_wins(tmp_path, monkeypatch, n_gpus):
    # BIM order 1,2,1,X: the reference writes add.chr1 twice, the second run over the first (pinned against the reference binary in
    # tests/test_reference_differential.py::test_oracle_loco_partials_agree_with_the_reference); here the first is never written
    import oracle_lib as o

    monkeypatch.setenv("GELEX_B200_OVERSUBSCRIBE", "1")
    n, m = 300, 400
    payload = g.synth_bed_host(n, 0, m, 37, 0.02)
    chroms = ["1"] * 100 + ["2"] * 120 + ["1"] * 80 + ["X"] * 100
    plink_io.write_plink(tmp_path / "d", payload, [("F", f"s{i:05d}") for i in range(n)], chroms)
    engine.grm_run(tmp_path / "d", tmp_path / "g", method="SH", loco=True, precision=g.PREC_FP64, n_gpus=n_gpus, write_whole=True, chunk_size=64,
                   partition="snp")
    assert sorted(p.name for p in tmp_path.iterdir() if p.name.startswith("g.") and p.name.endswith(".bin")) == \
        ["g.add.bin", "g.add.chr1.bin", "g.add.chr2.bin", "g.add.chrX.bin"]
    for name, (b, e) in {"1": (220, 300), "2": (100, 220), "X": (300, 400), None: (0, 400)}.items():
        G, _, _ = o.grm_compute(payload, n, 1, 0, ranges=[(b, e)], chunk=64)
        ref = o.pack_lower_f32(G)
        got = plink_io.read_bin(tmp_path / ("g.add.bin" if name is None else f"g.add.chr{name}.bin"), n)
        assert np.all(np.abs(got - ref) <= 2 * np.spacing(np.abs(ref)) + 1e-12), name


@pytest.mark.gpu
@pytest.mark.parametrize("n_gpus,partition", [(1, "auto"), (2, "snp"), (2, "tile")])
def test_engine_prefault_gives_identical_files(tmp_path, monkeypatch, n_gpus, partition):
    # Config::prefault (here through GELEX_B200_PREFAULT): the .bin exists from the start of the pass and its pages are populated while the
    # GPU accumulates; AD mode = two files, two prefaulters.  Same bytes as the plain run.
    monkeypatch.setenv("GELEX_B200_OVERSUBSCRIBE", "1")
    n, m = 4100, 1024
    payload = g.synth_bed_host(n, 0, m, 43, 0.01)
    plink_io.write_plink(tmp_path / "d", payload, [("F", f"s{i:05d}") for i in range(n)], ["1"] * m)
    kw = dict(method="OCH", mode="AD", precision=g.PREC_FP64, n_gpus=n_gpus, partition=partition)
    engine.grm_run(tmp_path / "d", tmp_path / "plain", **kw)
    monkeypatch.setenv("GELEX_B200_PREFAULT", "1")
    monkeypatch.setenv("GELEX_B200_WRITER", "mmap")
    engine.grm_run(tmp_path / "d", tmp_path / "pf", **kw)
    for k in ("add", "dom"):
        assert (tmp_path / f"pf.{k}.bin").read_bytes() == (tmp_path / f"plain.{k}.bin").read_bytes(), k
        assert (tmp_path / f"pf.{k}.id").read_bytes() == (tmp_path / f"plain.{k}.id").read_bytes(), k


REF_BINDING = Path(__file__).resolve().parent.parent / "oracle" / "_ref" / "gelex_ref_b200"


@pytest.mark.gpu
@pytest.mark.skipif(not REF_BINDING.exists(), reason="oracle/_ref/gelex_ref_b200 (the reference-side binding program) is not built")
def test_reference_host_code_over_the_c_abi_writes_the_reference_files(golden_dir, tmp_path):
    # INTEGRATION.md section 2 as a program (oracle/ref_binding.cpp): the UNMODIFIED reference's SampleManager, BedPipe projection + mmap,
    # work plans and GrmBinWriter / GrmIdWriter, with only the chunk loop of GRM::compute<GT> replaced by ggrm_create / push / finish /
    # fetch_f64.  Its files against the plain reference's (tests/golden): .id identical, .bin within one float32 ulp (FP64 mode).
    r = subprocess.run([str(REF_BINDING), str(golden_dir / "tiny"), str(tmp_path / "sh"), "1", "A", "0", "25", "fp64"], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    r = subprocess.run([str(REF_BINDING), str(golden_dir / "tiny"), str(tmp_path / "och"), "4", "AD", "1", "7", "fp64"], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    pairs = [("sh.add", "tiny_SH.add")] + [(f"och.{k}.chr{c}", f"tiny_OCH.{k}.chr{c}") for k in ("add", "dom") for c in ("1", "2", "X", "7")]
    for mine, gold in pairs:
        assert (tmp_path / f"{mine}.id").read_bytes() == (golden_dir / f"{gold}.id").read_bytes(), mine
        got, ref = plink_io.read_bin(tmp_path / f"{mine}.bin", 36), plink_io.read_bin(golden_dir / f"{gold}.bin", 36)
        assert np.all(np.abs(got.astype(np.float64) - ref) <= np.spacing(np.abs(ref)) + 1e-12), mine


@pytest.mark.gpu
def test_cli_is_a_drop_in(golden_dir, tmp_path):
    r = subprocess.run([str(engine.CLI_PATH), "grm", "-b", str(golden_dir / "tiny"), "--gm", "SH", "--add", "-c", "25", "--precision", "fp64",
                        "-o", str(tmp_path / "cli")], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    assert "Samples    : 36 samples" in r.stdout and "Pattern     : " + str(tmp_path / "cli") + ".add.{bin|id}" in r.stdout
    assert (tmp_path / "cli.log").exists()
    assert_bin_close(tmp_path / "cli.add.bin", golden_dir / "tiny_SH.add.bin", 36)
    # default flags: OSH, additive, fp16 operands -> within the 16-bit tolerance of the reference file (normalised)
    r = subprocess.run([str(engine.CLI_PATH), "grm", "-b", str(golden_dir / "tiny"), "-o", str(tmp_path / "d")], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    got, ref = plink_io.read_bin(tmp_path / "d.add.bin", 36).astype(np.float64), plink_io.read_bin(golden_dir / "tiny_OSH.add.bin", 36).astype(np.float64)
    diag = np.cumsum(np.arange(1, 37)) - 1
    assert np.max(np.abs(got / (got[diag].sum() / 36) - ref / (ref[diag].sum() / 36))) < 5e-3  # m = 96 only


# ------------------------------------------------------------------------------------------------ consumer side
def test_grm_loader_validation_errors_need_no_gpu(golden_dir, tmp_path):
    # GrmLoader::GrmLoader (grm_loader.cpp:47-123): id parsing and the exact-size check happen before any device work
    import shutil

    shutil.copy(golden_dir / "tiny_SH.add.id", tmp_path / "x.add.id")
    (tmp_path / "x.add.bin").write_bytes((golden_dir / "tiny_SH.add.bin").read_bytes()[:-4])
    with pytest.raises(engine.EngineError, match="FileFormatException.*file size mismatch. Expected 2664 bytes \\(36 samples\\), got 2660 bytes"):
        engine.load_grm(tmp_path / "x.add")
    (tmp_path / "y.id").write_text("F1 s1\n")
    (tmp_path / "y.bin").write_bytes(b"\0" * 4)
    with pytest.raises(engine.EngineError, match="FileFormatException.*invalid ID line 'F1 s1' \\(expected FID and IID\\)"):
        engine.load_grm(tmp_path / "y")
    (tmp_path / "z.id").write_text("\n\n")
    (tmp_path / "z.bin").write_bytes(b"")
    with pytest.raises(engine.EngineError, match="FileFormatException.*no sample IDs found"):
        engine.load_grm(tmp_path / "z")
    with pytest.raises(engine.EngineError, match="InvalidInputException.*sample ID 'nope.*' not found in GRM file"):
        engine.load_grm(golden_dir / "tiny_SH.add", id_map={("nope", "x"): 0})
    with pytest.raises(engine.EngineError, match="InvalidInputException: LOCO error: GRM file not found"):
        engine.load_grm(golden_dir / "tiny_OCH.add", chr_prefix=tmp_path / "missing.chr9")


@pytest.mark.gpu
def test_grm_loaders_match_reference_outputs(golden_dir):
    # The fixtures are the dense matrices the UNMODIFIED reference's GrmLoader::load and LocoGRMLoader::load_loco_grm returned for the
    # golden .bin files (tests/golden/make_golden.py).  Same float32 inputs, same FP64 operations in the same order -> same bits.
    gold = np.load(golden_dir / "tiny_loaders.npz")
    got = engine.load_grm(golden_dir / "tiny_SH.add")
    assert np.array_equal(got, gold["SH_add_normalized"])
    for c in ("1", "2", "X", "7"):
        got = engine.load_grm(golden_dir / "tiny_OCH.add", chr_prefix=str(golden_dir / "tiny_OCH.add") + f".chr{c}")
        assert np.array_equal(got, gold[f"OCH_add_loco_chr{c}"]), c
    # filter + reorder (GrmLoader::load_unnormalized(id_map), grm_loader.cpp:171-231) against a numpy statement of the same rule
    n = 36
    packed = plink_io.read_bin(golden_dir / "tiny_SH.add.bin", n).astype(np.float64)
    full = np.zeros((n, n))
    full[np.tril_indices(n)] = packed
    full = full + np.tril(full, -1).T
    ids = [tuple(ln.split("\t")) for ln in (golden_dir / "tiny_SH.add.id").read_text().split("\n") if ln]
    rng = np.random.default_rng(5)
    pick = rng.permutation(n)[:20]
    id_map = {ids[f]: t for t, f in zip([0, 2, 3, 5, 6, 7, 8, 9, 10, 11, 12, 13, 14, 15, 16, 17, 18, 19, 21, 22], pick)}  # rows 1, 4, 20 stay empty
    want = np.zeros((23, 23))
    for (fa, ta) in zip(pick, id_map.values()):
        for (fb, tb) in zip(pick, id_map.values()):
            want[ta, tb] = full[fa, fb]
    got = engine.load_grm(golden_dir / "tiny_SH.add", normalize=False, id_map=id_map)
    assert np.array_equal(got, want)
    got = engine.load_grm(golden_dir / "tiny_SH.add", normalize=True, id_map=id_map)
    tr = 0.0
    for t in range(23):
        tr += want[t, t]
    assert np.array_equal(got, want / (tr / 23))


@pytest.mark.gpu
def test_grm_loader_large_roundtrip_from_engine_files(tmp_path):
    # files written by the engine (LOCO + whole) -> loaders -> the same matrices the resident accumulators give (ggrm_fetch_loco),
    # at a size with many 32 x 32 loader tiles and a ragged edge
    n, m = 1111, 900
    payload = g.synth_bed_host(n, 0, m, 29, 0.01)
    chroms = ["1"] * 300 + ["2"] * 600
    plink_io.write_plink(tmp_path / "d", payload, [("F", f"s{i:05d}") for i in range(n)], chroms)
    engine.grm_run(tmp_path / "d", tmp_path / "o", method="OSH", mode="A", loco=True, precision=g.PREC_FP64, write_whole=True)
    w = plink_io.read_bin(tmp_path / "o.add.bin", n).astype(np.float64)
    for ci, c in enumerate(("1", "2")):
        p = plink_io.read_bin(tmp_path / f"o.add.chr{c}.bin", n).astype(np.float64)
        diag = np.cumsum(np.arange(1, n + 1)) - 1
        kw = kc = 0.0
        for i in range(n):
            kw += w[diag[i]]
            kc += p[diag[i]]
        want = np.zeros((n, n))
        want[np.tril_indices(n)] = (w - p) / (kw / n - kc / n)
        want = want + np.tril(want, -1).T
        got = engine.load_grm(tmp_path / "o.add", chr_prefix=tmp_path / f"o.add.chr{c}")
        assert np.array_equal(got, want), c
        assert np.array_equal(got, got.T)


def _write_grm(prefix, matrix, ids):
    """GrmBinWriter + GrmIdWriter byte layouts (grm_bin_writer.cpp:54-61, grm_id_writer.cpp:41-54) for loader fixtures."""
    m = np.asarray(matrix, dtype=np.float64)
    n = m.shape[0]
    m[np.tril_indices(n)].astype(np.float32).tofile(str(prefix) + ".bin")
    with open(str(prefix) + ".id", "w") as f:
        for fid, iid in ids:
            f.write(f"{fid}\t{iid}\n")


def _f32(v):
    return np.float64(np.float32(v))


@pytest.mark.gpu
def test_grm_loader_known_answers_of_the_reference_tests(tmp_path):
    # The hand-written vectors of the reference's own tests/test_grm_loader.cpp, restated against this loader.
    ids3 = [(f"FAM{i}", f"IND{i}") for i in range(3)]
    # :290-325  3 x 3, trace/n = 1: every element is float(original)
    m3 = np.array([[1.0, 0.5, 0.3], [0.5, 1.0, 0.4], [0.3, 0.4, 1.0]])
    _write_grm(tmp_path / "a", m3, ids3)
    got = engine.load_grm(tmp_path / "a")
    assert got.shape == (3, 3) and all(got[i, j] == _f32(m3[i, j]) for i in range(3) for j in range(3))
    # :423-470  subset: samples 1 and 3 of 4 -> rows 0 and 1
    m4 = np.array([[1.0, 0.1, 0.2, 0.3], [0.1, 2.0, 0.4, 0.5], [0.2, 0.4, 3.0, 0.6], [0.3, 0.5, 0.6, 4.0]])
    ids4 = [(f"FAM{i}", f"IND{i}") for i in range(4)]
    _write_grm(tmp_path / "b", m4, ids4)
    got = engine.load_grm(tmp_path / "b", normalize=False, id_map={("FAM1", "IND1"): 0, ("FAM3", "IND3"): 1})
    assert got.shape == (2, 2)
    assert got[0, 0] == _f32(2.0) and got[1, 1] == _f32(4.0) and got[0, 1] == _f32(0.5) and got[1, 0] == _f32(0.5)
    # :474-520  reversed order
    got = engine.load_grm(tmp_path / "b", normalize=False, id_map={ids4[i]: 3 - i for i in range(4)})
    assert all(got[3 - i, 3 - j] == _f32(m4[i, j]) for i in range(4) for j in range(4))
    # :590-638  sparse target indices: 0 -> 0, 2 -> 5 gives a 6 x 6 matrix, unmapped rows zero
    m3b = np.array([[1.0, 0.1, 0.2], [0.1, 2.0, 0.3], [0.2, 0.3, 3.0]])
    _write_grm(tmp_path / "c", m3b, ids3)
    got = engine.load_grm(tmp_path / "c", normalize=False, id_map={("FAM0", "IND0"): 0, ("FAM2", "IND2"): 5})
    assert got.shape == (6, 6)
    assert got[0, 0] == _f32(1.0) and got[5, 5] == _f32(3.0) and got[0, 5] == _f32(0.2) and got[5, 0] == _f32(0.2)
    assert all(got[k, k] == 0.0 for k in (1, 2, 3, 4)) and np.count_nonzero(got) == 4
    # :735-777  inf and NaN survive
    _write_grm(tmp_path / "d", np.array([[np.inf, 0.5], [0.5, 1.0]]), ids3[:2])
    got = engine.load_grm(tmp_path / "d", normalize=False)
    assert np.isinf(got[0, 0]) and got[0, 0] > 0
    _write_grm(tmp_path / "e", np.array([[np.nan, 0.5], [0.5, 1.0]]), ids3[:2])
    assert np.isnan(engine.load_grm(tmp_path / "e", normalize=False)[0, 0])
    # :780-812  trace-based normalisation: trace 12, n 3 -> denominator 4
    _write_grm(tmp_path / "f", np.array([[2.0, 1.0, 0.6], [1.0, 4.0, 0.8], [0.6, 0.8, 6.0]]), ids3)
    got = engine.load_grm(tmp_path / "f")
    assert got[0, 0] == _f32(2.0 / 4.0) and got[1, 1] == _f32(4.0 / 4.0) and got[2, 2] == _f32(6.0 / 4.0)
    assert got[1, 0] == _f32(1.0) / 4.0 and got[2, 1] == got[1, 2] == _f32(0.8) / 4.0
    # :523-544  empty id_map -> empty matrix is an API property of the C++ class (IdMap{}); through this binding an empty map means
    # "no filter", so it is covered in the C++ sources only.


@pytest.mark.gpu
def test_loco_grm_loader_known_answers_of_the_reference_tests(tmp_path):
    # tests/test_loco_grm_loader.cpp:60-110 (basic) and :112-173 (filtered + reordered), tolerance 1e-5 relative as there;
    # plus chromosome files whose .id order differs from the whole-genome file (each loader maps by ID)
    rng = np.random.default_rng(12)
    n = 3
    ids = [("F1", "I1"), ("F1", "I2"), ("F1", "I3")]
    xw, xi = rng.uniform(-1, 1, (n, 10)), rng.uniform(-1, 1, (n, 3))
    gw, gi = xw @ xw.T, xi @ xi.T
    _write_grm(tmp_path / "whole", gw, ids)
    _write_grm(tmp_path / "chr1", gi, ids)
    got = engine.load_grm(tmp_path / "whole", chr_prefix=tmp_path / "chr1", id_map={ids[i]: i for i in range(n)})
    want = (gw - gi) / (np.trace(gw) / n - np.trace(gi) / n)
    assert np.allclose(got, want, rtol=1e-5, atol=0)
    got = engine.load_grm(tmp_path / "whole", chr_prefix=tmp_path / "chr1", id_map={("F1", "I3"): 0, ("F1", "I1"): 1})
    sel = [2, 0]
    gws, gis = gw[np.ix_(sel, sel)], gi[np.ix_(sel, sel)]
    want = (gws - gis) / (np.trace(gws) / 2 - np.trace(gis) / 2)
    assert got.shape == (2, 2) and np.allclose(got, want, rtol=1e-5, atol=0)
    perm = [2, 0, 1]  # chromosome file lists the samples in another order
    _write_grm(tmp_path / "chr2", gi[np.ix_(perm, perm)], [ids[p] for p in perm])
    got2 = engine.load_grm(tmp_path / "whole", chr_prefix=tmp_path / "chr2", id_map={("F1", "I3"): 0, ("F1", "I1"): 1})
    assert np.array_equal(got2, got)
    # k_loco <= 0 (loco_grm_loader.cpp:62-70)
    _write_grm(tmp_path / "big", gw * 2, ids)
    with pytest.raises(engine.EngineError, match="LOCO error: Chromosome GRM denominator .* is greater than or equal to Whole GRM denominator"):
        engine.load_grm(tmp_path / "whole", chr_prefix=tmp_path / "big")


# ------------------------------------------------------------------------------------------------ standardised genotype export
def test_genotype_export_validation_needs_no_gpu(golden_dir, tmp_path):
    # GenotypePipe::GenotypePipe (genotype_pipe.cpp:51-82): inputs are opened and existing outputs refused before any device work
    assert engine._genotype_dims(golden_dir / "tiny") == (36, 96)   # sorted + de-duplicated samples, BIM lines
    (tmp_path / "have.bmat").write_bytes(b"")
    with pytest.raises(engine.EngineError, match="FileExistsException: .*have: existing files"):
        engine.genotype_export(golden_dir / "tiny", tmp_path / "have")
    (tmp_path / "have.bmat").unlink()
    (tmp_path / "have.snpstats").write_bytes(b"")
    with pytest.raises(engine.EngineError, match="FileExistsException"):
        engine.genotype_export(golden_dir / "tiny", tmp_path / "have")
    with pytest.raises(engine.EngineError, match="FileNotFoundException"):
        engine.genotype_export(tmp_path / "nothing", tmp_path / "out")
    h = engine.load_host()
    import ctypes as C

    err = C.create_string_buffer(512)
    st = h.gelex_b200_genotype_export(str(golden_dir / "tiny").encode(), str(tmp_path / "o").encode(), 9, 1, 100, 0, None, None, None, None, None, 0,
                                      err, len(err))
    assert st == -1 and b"Invalid genotype process method." in err.value
    assert not (tmp_path / "o.bmat").exists()


@pytest.mark.gpu
@pytest.mark.parametrize("method,additive,chunk", [("OSH", True, 10000), ("SH", False, 7), ("OC", False, 40), ("S", True, 33)])
def test_genotype_export_files_match_reference_layout(golden_dir, tmp_path, method, additive, chunk):
    # tiny fixture (37 FAM lines, one duplicated id, shuffled): the values are the unmodified reference's process_matrix outputs
    # (tiny_chunks.npz); .bmat = column-major FP64 without header, .snpstats = {n, m, #mono} + indices + means + stddevs
    import oracle_lib as o
    from test_oracle_golden import close, load_tiny

    payload, n_raw, ids, r2o, _ = load_tiny(golden_dir)
    n, m = len(ids), payload.shape[0]
    code, kind = g.METHODS[method], 0 if additive else 1
    st = engine.genotype_export(golden_dir / "tiny", tmp_path / "t", method=method, additive=additive, chunk_size=chunk)
    gold = np.load(golden_dir / "tiny_chunks.npz")
    bmat = np.fromfile(tmp_path / "t.bmat", dtype=np.float64)
    assert bmat.size == n * m                                      # GenotypeMap's exact-size check (genotype_mmap.cpp:44-53)
    z = bmat.reshape(m, n).T
    assert close(z, gold[f"z_{code}_{'A' if additive else 'D'}"], rtol=1e-12, atol=1e-13)
    assert np.array_equal(st["freq"], gold[f"freq_{code}_{'A' if additive else 'D'}"])
    meta = engine.read_snp_stats(tmp_path / "t.snpstats")
    assert (meta["num_samples"], meta["num_variants"]) == (n, m)
    raw = o.decode_chunk(payload, n_raw, n, r2o)
    mono = []
    for j in range(m):
        _, ls = o.process_column(raw[:, j], code, kind)
        # (a column with one valid call has stddev 0/0 = NaN under the sample policies, in the reference too)
        assert close(meta["mean"][j], ls.mean, rtol=1e-13, atol=0) and close(meta["stddev"][j], ls.stddev, rtol=1e-12, atol=1e-300), j
        if ls.is_monomorphic:
            mono.append(j)
    assert meta["monomorphic_indices"].tolist() == mono == st["monomorphic_indices"].tolist()
    assert np.array_equal(meta["mean"], st["mean"], equal_nan=True) and np.array_equal(meta["stddev"], st["stddev"], equal_nan=True)
    # GenotypeLoader: the same matrix in memory, bit for bit, whatever the chunking
    mat, st2 = engine.genotype_load(golden_dir / "tiny", method=method, additive=additive, chunk_size=5)
    assert mat.shape == (n, m) and np.array_equal(np.nan_to_num(mat, nan=-9.0), np.nan_to_num(z, nan=-9.0))
    assert np.array_equal(st2["mean"], st["mean"], equal_nan=True) and st2["monomorphic_indices"].tolist() == mono


@pytest.mark.gpu
@pytest.mark.parametrize("method,additive,chunk", [("OSH", True, 25), ("SH", False, 7), ("OC", False, 40), ("S", True, 33)])
def test_genotype_export_files_against_the_reference_pipe_files(golden_dir, tmp_path, method, additive, chunk):
    # our .bmat / .snpstats beside the files the unmodified reference's GenotypePipe wrote for the same input (tests/golden/tiny_pipe_*):
    # identical sizes, header and monomorphic indices; values to 2 ulp (FP64 on the device, explicit _rn arithmetic)
    engine.genotype_export(golden_dir / "tiny", tmp_path / "t", method=method, additive=additive, chunk_size=chunk)
    base = golden_dir / f"tiny_pipe_{method}.{'add' if additive else 'dom'}"
    ref, got = np.fromfile(str(base) + ".bmat", dtype=np.float64), np.fromfile(tmp_path / "t.bmat", dtype=np.float64)
    assert got.shape == ref.shape and np.array_equal(np.isnan(got), np.isnan(ref))
    ok = ~np.isnan(ref)
    assert np.all(np.abs(got[ok] - ref[ok]) <= 2 * np.spacing(np.abs(ref[ok])) + 1e-15)
    rm, gm = engine.read_snp_stats(str(base) + ".snpstats"), engine.read_snp_stats(tmp_path / "t.snpstats")
    assert (tmp_path / "t.snpstats").stat().st_size == Path(str(base) + ".snpstats").stat().st_size
    assert (gm["num_samples"], gm["num_variants"]) == (rm["num_samples"], rm["num_variants"])
    assert gm["monomorphic_indices"].tolist() == rm["monomorphic_indices"].tolist()
    # mean / stddev: 2 ulp where they are closed forms of the allele frequency (HWE policies, additive sample mean).  The sample
    # policy's dominance mean and every sample stddev are SUMS over the column in the reference (statistic_policy.h:61-65, 70-78):
    # Eigen adds n rounded terms in its own (vectorised) order, the device evaluates n1*e1 + n2*e2 and the three-term sum of
    # squares from the integer counts, so the two differ by the reference's own summation rounding - a few ulp on the tiny fixture (measured
    # 4), bounded here by 16.
    sums = method in ("S", "C", "OS", "OC")
    ulps = {"mean": 16 if (sums and not additive) else 2, "stddev": 16 if sums else 2}
    for key in ("mean", "stddev"):
        assert np.array_equal(np.isnan(gm[key]), np.isnan(rm[key]))
        ok = ~np.isnan(rm[key])
        assert np.all(np.abs(gm[key][ok] - rm[key][ok]) <= ulps[key] * np.spacing(np.abs(rm[key][ok])) + 1e-300), key


@pytest.mark.gpu
def test_genotype_export_larger_dense_with_monomorphic_snps(tmp_path):
    import oracle_lib as o
    from test_oracle_golden import close

    n, m = 1003, 700
    bed = g.synth_bed_host(n, 0, m, 23, 0.02).copy()
    bed[[10, 400, 699]] = 0xFF                # code 11 everywhere: genotype 0 for every sample -> monomorphic
    bed[[10, 400, 699], -1] = 0x3F            # padding bits of the tail byte stay 0 (n % 4 == 3)
    plink_io.write_plink(tmp_path / "d", bed, [("F", f"s{i:05d}") for i in range(n)], ["1"] * m)
    st = engine.genotype_export(tmp_path / "d", tmp_path / "d_out", method="OSH", additive=True, chunk_size=256)
    assert st["monomorphic_indices"].tolist() == [10, 400, 699]
    z = np.fromfile(tmp_path / "d_out.bmat", dtype=np.float64).reshape(m, n).T
    zr, fr = o.process_matrix(o.decode_chunk(bed, n), 3, 0)
    assert np.array_equal(st["freq"], fr) and close(z, zr, rtol=1e-12, atol=1e-13)
    meta = engine.read_snp_stats(tmp_path / "d_out.snpstats")
    assert meta["monomorphic_indices"].tolist() == [10, 400, 699] and meta["num_samples"] == n and meta["num_variants"] == m
    assert np.all(z[:, 10] == 0.0)            # centred, not scaled (genotype_processor.h:49-52)


# ------------------------------------------------------------------------------------------------ GCTA binary format behind a flag
def test_gcta_flag_validation_needs_no_gpu(golden_dir, tmp_path):
    r = subprocess.run([str(engine.CLI_PATH), "grm", "-b", str(golden_dir / "tiny"), "--add", "--dom", "--gcta", "-o", str(tmp_path / "x")],
                       capture_output=True, text=True)
    assert r.returncode == 1 and "--gcta writes the additive matrix of the whole genome" in (r.stdout + r.stderr)
    r = subprocess.run([str(engine.CLI_PATH), "grm", "-b", str(golden_dir / "tiny"), "--loco", "--gcta", "-o", str(tmp_path / "x")],
                       capture_output=True, text=True)
    assert r.returncode == 1 and not list(tmp_path.glob("x.grm*"))
    for extra in (["--gcta", "--gcta-diag", "--geno-method", "OSH"], ["--gcta-diag"]):   # GCTA's diagonal: with --gcta and StandardizeHWE only
        r = subprocess.run([str(engine.CLI_PATH), "grm", "-b", str(golden_dir / "tiny"), "--add", *extra, "-o", str(tmp_path / "x")],
                           capture_output=True, text=True)
        assert r.returncode == 1 and "--gcta-diag is GCTA's diagonal" in (r.stdout + r.stderr)


@pytest.mark.gpu
def test_engine_gcta_files(tmp_path, monkeypatch):
    import oracle_lib as o

    monkeypatch.setenv("GELEX_B200_OVERSUBSCRIBE", "1")
    n, m = 2301, 1500   # two super-rows: the two-GPU run reduces and writes two row slices
    rng = np.random.default_rng(3)
    perm = rng.permutation(n)
    fam = [(f"fam{i % 7}", f"s{perm[i]:05d}") for i in range(n)]
    ids, r2o = plink_io.sample_order(fam)
    bed = g.synth_bed_host(n, 0, m, 41, 0.05)
    plink_io.write_plink(tmp_path / "d", bed, fam, ["1"] * m)
    engine.grm_run_gcta(tmp_path / "d", tmp_path / "g1", method="SH", precision=g.PREC_FP64, chunk_size=400)
    assert sorted(p.name for p in tmp_path.glob("g1.*")) == ["g1.grm.N.bin", "g1.grm.bin", "g1.grm.id"]
    raw = o.decode_chunk(bed, n, len(ids), r2o)
    called = (~np.isnan(raw)).astype(np.float64)
    N = (called @ called.T)[np.tril_indices(n)]
    G, _, _ = o.grm_compute(bed, n, 1, 0, chunk=500, n_out=len(ids), raw_to_out=r2o)
    ref = (G[np.tril_indices(n)] / N).astype(np.float32)
    got_n, got_a = plink_io.read_bin(tmp_path / "g1.grm.N.bin", n), plink_io.read_bin(tmp_path / "g1.grm.bin", n)
    assert np.array_equal(got_n.astype(np.float64), N)
    assert np.all(np.abs(got_a - ref) <= np.spacing(np.abs(ref)) + 1e-12)
    assert (tmp_path / "g1.grm.id").read_text() == "".join(f"{i.split(chr(31))[0]}\t{i.split(chr(31))[1]}\n" for i in ids)
    # two SNP shards: both matrices reduce-scattered over peer memory, each GPU divides and writes its own rows
    engine.grm_run_gcta(tmp_path / "d", tmp_path / "g2", method="SH", precision=g.PREC_FP64, chunk_size=400, n_gpus=2)
    assert (tmp_path / "g2.grm.N.bin").read_bytes() == (tmp_path / "g1.grm.N.bin").read_bytes()
    assert_bin_close(tmp_path / "g2.grm.bin", tmp_path / "g1.grm.bin", n, ulps=2)
    assert (tmp_path / "g2.grm.id").read_bytes() == (tmp_path / "g1.grm.id").read_bytes()
    # 16-bit operands: counts still exact
    engine.grm_run_gcta(tmp_path / "d", tmp_path / "g3", method="SH", precision=g.PREC_BF16)
    assert (tmp_path / "g3.grm.N.bin").read_bytes() == (tmp_path / "g1.grm.N.bin").read_bytes()
    a3 = plink_io.read_bin(tmp_path / "g3.grm.bin", n).astype(np.float64)
    assert np.max(np.abs(a3 - ref)) <= 8e-3


@pytest.mark.gpu
def test_engine_direct_dma_into_the_output_mapping_matches_the_staged_writer(tmp_path, monkeypatch):
    # With GELEX_B200_DIRECT_DMA=1 (opt-in: measured slower end to end, the page-locking holds the driver lock) the populated mapping
    # of an output file on tmpfs is page-locked and whole-row read-outs DMA straight into it (no staging buffer, no host copy);
    # otherwise rows go through the writer pool.  Same bytes either way:
    # whole-genome file, every chromosome file of a LOCO run and its whole (one GPU; two ranks' row slices of the LOCO whole).
    import shutil
    import tempfile

    if not Path("/dev/shm").is_dir():
        pytest.skip("no tmpfs at /dev/shm")
    shm = Path(tempfile.mkdtemp(prefix="ggrm_dma_", dir="/dev/shm"))
    try:
        monkeypatch.setenv("GELEX_B200_OVERSUBSCRIBE", "1")
        n, m = 4500, 3000
        payload = g.synth_bed_host(n, 0, m, 29, 0.01)
        bounds = [round(m * i / 5) for i in range(6)]
        chroms = [str(c + 1) for c in range(5) for _ in range(bounds[c + 1] - bounds[c])]
        plink_io.write_plink(shm / "d", payload, [("F", f"s{i:05d}") for i in range(n)], chroms)
        runs = {}
        for tag, dma in (("dma", "1"), ("staged", "0")):
            monkeypatch.setenv("GELEX_B200_DIRECT_DMA", dma)
            engine.grm_run(shm / "d", shm / f"{tag}_w", method="SH", precision=g.PREC_FP16)
            engine.grm_run(shm / "d", shm / f"{tag}_l", method="OCH", mode="AD", loco=True, precision=g.PREC_FP16, write_whole=True)
            engine.grm_run(shm / "d", shm / f"{tag}_l2", method="SH", loco=True, precision=g.PREC_FP16, write_whole=True, n_gpus=2, partition="snp")
            runs[tag] = sorted(p.name[len(tag):] for p in shm.glob(f"{tag}_*.bin"))
        assert runs["dma"] == runs["staged"] and len(runs["dma"]) == 1 + 12 + 6
        for suffix in runs["dma"]:
            assert (shm / f"dma{suffix}").read_bytes() == (shm / f"staged{suffix}").read_bytes(), suffix
    finally:
        shutil.rmtree(shm, ignore_errors=True)


@pytest.mark.gpu
def test_observer_sees_the_events_of_the_reference_in_its_order(tmp_path, monkeypatch):
    # The reference's observer gets, per work item in plan order (per chromosome group: add, then dom), one (current, total) event per
    # chunk with `current` counted inside the item (grm.h:124-130, grm_engine.cpp:78-110), bracketed by the dataset / total-work events
    # and the done + files-written events.  The engine computes add and dom of a range in ONE pass: the dom item's events are replayed
    # when the pass over the range completes, so the sequence is the reference's.
    n, m, chunk = 300, 1000, 128
    bed = g.synth_bed_host(n, 0, m, 5, 0.01)
    chroms = ["1"] * 400 + ["2"] * 250 + ["7"] * 350
    plink_io.write_plink(tmp_path / "d", bed, [("F", f"s{i:04d}") for i in range(n)], chroms)
    for mode in ("A", "D", "AD"):
        for loco in (False, True):
            p = engine.plan(tmp_path / "d", mode=mode, loco=loco)
            total = sum(e - b for _, b, e in p["items"])
            want = [("data_loaded", n, m), ("compute_started", total, 0)]
            for _, b, e in p["items"]:   # plan order = the reference's item order
                want += [("progress", min(s + chunk, e) - b, e - b) for s in range(b, e, chunk)]
            want += [("progress_done", total, total), ("files_written", 2 * len(p["items"]), 0)]
            got = engine.grm_run_events(tmp_path / "d", tmp_path / f"o_{mode}_{int(loco)}", method="SH", mode=mode, loco=loco, chunk_size=chunk)
            assert got == want, (mode, loco)
    # SNP shards on two GPUs: chunks are cut inside every shard, so `current` advances in other steps; the two items still come one
    # after the other, each monotone and ending at (m, m)
    monkeypatch.setenv("GELEX_B200_OVERSUBSCRIBE", "1")
    got = engine.grm_run_events(tmp_path / "d", tmp_path / "o2", method="SH", mode="AD", loco=False, chunk_size=chunk, n_gpus=2)
    prog = [(a, b) for name, a, b in got if name == "progress"]
    half = len(prog) // 2
    assert len(prog) % 2 == 0 and prog[:half] == prog[half:]
    cur = [a for a, _ in prog[:half]]
    assert cur == sorted(cur) and cur[-1] == m and all(b == m for _, b in prog)
