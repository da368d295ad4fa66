"""ctypes access to the C++ host engine (libggrm_host.so): the GrmEngine mirror of the reference
(include/gelex/pipeline/grm_engine.h:30-51) and its metadata planner.  Marshalling only."""
from __future__ import annotations

import ctypes as C
from pathlib import Path

import numpy as np

from . import _lib

HOST_LIB_PATH = Path(__file__).resolve().parent / "libggrm_host.so"
CLI_PATH = Path(__file__).resolve().parent / "gelex-b200"
MODES = {"A": 0, "D": 1, "AD": 2}
_host = None


class EngineError(RuntimeError):
    def __init__(self, status, message):
        super().__init__(message)
        self.status = status
        self.message = message


def load_host():
    global _host
    if _host is None:
        _lib.load()  # dependency (resolved through rpath $ORIGIN as well)
        if not HOST_LIB_PATH.exists():
            raise ImportError(f"{HOST_LIB_PATH} not built; run __graft_entry__.build()")
        h = C.CDLL(str(HOST_LIB_PATH))
        h.gelex_b200_grm_run.argtypes = [C.c_char_p, C.c_char_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_char_p, C.c_int]
        h.gelex_b200_grm_run_ex.argtypes = [C.c_char_p, C.c_char_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                            C.c_char_p, C.c_int]
        h.gelex_b200_plan.argtypes = [C.c_char_p, C.c_int, C.c_int, C.c_void_p, C.c_int64, C.POINTER(C.c_int64), C.POINTER(C.c_int64),
                                      C.POINTER(C.c_int), C.c_char_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.POINTER(C.c_int), C.c_char_p, C.c_int]
        h.gelex_b200_plan.restype = C.c_int64
        h.gelex_b200_grm_load.argtypes = [C.c_char_p, C.c_char_p, C.c_int, C.c_char_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_int64, C.c_char_p, C.c_int]
        h.gelex_b200_grm_load.restype = C.c_int64
        h.gelex_b200_row_slices.argtypes = [C.c_int64, C.c_int, C.c_void_p]
        h.gelex_b200_strip_owners.argtypes = [C.c_int64, C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_int]
        h.gelex_b200_gpus_worth_using.argtypes = [C.c_int64, C.c_int64, C.c_int, C.c_int, C.c_int, C.c_int]
        h.gelex_b200_piece_owners.argtypes = [C.c_int64, C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_int]
        h.gelex_b200_grm_run_events.argtypes = [C.c_char_p, C.c_char_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_int64,
                                                C.POINTER(C.c_int64), C.c_char_p, C.c_int]
        h.gelex_b200_grm_run_gcta.argtypes = [C.c_char_p, C.c_char_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_char_p, C.c_int]
        h.gelex_b200_grm_run_gcta_ex.argtypes = [C.c_char_p, C.c_char_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_char_p, C.c_int]
        vp = C.c_void_p
        h.gelex_b200_genotype_export.argtypes = [C.c_char_p, C.c_char_p, C.c_int, C.c_int, C.c_int, C.c_int, vp, vp, vp, vp, vp, C.c_int64,
                                                 C.c_char_p, C.c_int]
        h.gelex_b200_genotype_load.argtypes = [C.c_char_p, C.c_int, C.c_int, C.c_int, C.c_int, vp, C.c_int64, vp, vp, vp, vp, vp, C.c_int64,
                                               C.c_char_p, C.c_int]
        h.gelex_b200_writer_selftest.argtypes = [C.c_char_p, C.c_int64, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_double),
                                                 C.POINTER(C.c_double), C.POINTER(C.c_int), C.c_char_p, C.c_int]
        h.gelex_b200_write_id_file.argtypes = [C.c_char_p, C.c_char_p, C.c_int64, C.c_char_p, C.c_int]
        h.gelex_b200_format_bed_path.argtypes = [C.c_char_p, C.c_char_p, C.c_int, C.c_char_p, C.c_int]
        h.gelex_b200_relay_selftest.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_char_p, C.c_int]
        h.gelex_b200_relay_selftest.restype = C.c_int64
        _host = h
    return _host


WRITERS = {"auto": 0, "pwrite": 1, "mmap": 2}


def writer_selftest(bin_path, n, tiles=False, parts=1, writer_threads=4, writer="auto", prefault_threads=0, with_prefault_seconds=False):
    """The engine's .bin writer pool on a synthetic n x n matrix, no GPU involved (GrmEngine::writer_selftest): entry (i, j) of
    the packed lower triangle is float32((31 i + j) mod 65521 + 0.5).  tiles=False: `parts` row-slice writers of whole-row
    blocks (SNP-shard read-out); tiles=True: 2048 x 2048 super-tiles dealt round-robin to `parts` pools as row segments (2D-tile
    read-out).  prefault_threads > 0: the file's pages are populated first (what Config::prefault does during the pass), timed on its own.
    Returns (writer seconds, outputs_mapped), plus the prefault seconds with with_prefault_seconds."""
    h = load_host()
    err = C.create_string_buffer(2048)
    sec, pf, mapped = C.c_double(0), C.c_double(0), C.c_int(0)
    st = h.gelex_b200_writer_selftest(str(bin_path).encode(), int(n), int(bool(tiles)), int(parts), int(writer_threads), WRITERS[writer],
                                      int(prefault_threads), C.byref(sec), C.byref(pf), C.byref(mapped), err, len(err))
    if st != 0:
        raise EngineError(st, err.value.decode())
    if with_prefault_seconds:
        return sec.value, bool(mapped.value), pf.value
    return sec.value, bool(mapped.value)


def writer_selftest_expected(n):
    """Packed lower triangle writer_selftest must produce."""
    i, j = np.tril_indices(n)
    return (((31 * i.astype(np.int64) + j) % 65521).astype(np.float32) + np.float32(0.5))


PARTITIONS = {"auto": 0, "snp": 1, "tile": 2}


def grm_run(bfile, out_prefix, method="OSH", mode="A", loco=False, chunk_size=10000, precision=_lib.PREC_FP16, n_gpus=1, write_whole=False,
            partition="auto", writer_threads=4):
    """GrmEngine::compute through the C entry point; raises EngineError("<ExceptionName>: message")."""
    h = load_host()
    m = _lib.METHODS[method] if isinstance(method, str) else int(method)
    err = C.create_string_buffer(2048)
    st = h.gelex_b200_grm_run_ex(str(bfile).encode(), str(out_prefix).encode(), m, MODES[mode], int(loco), int(chunk_size), int(precision),
                                 int(n_gpus), int(write_whole), PARTITIONS[partition], int(writer_threads), err, len(err))
    if st != 0:
        raise EngineError(st, err.value.decode())


EVENT_NAMES = {0: "data_loaded", 1: "compute_started", 2: "progress", 3: "progress_done", 4: "files_written"}


def grm_run_events(bfile, out_prefix, method="OSH", mode="A", loco=False, chunk_size=10000, precision=_lib.PREC_FP16, n_gpus=1, max_events=1 << 16):
    """grm_run with the observer's events recorded: [(name, a, b)] in the order the observer saw them (see gelex_b200_grm_run_events)."""
    h = load_host()
    m = _lib.METHODS[method] if isinstance(method, str) else int(method)
    err = C.create_string_buffer(2048)
    ev = np.zeros((max_events, 3), dtype=np.int64)
    n = C.c_int64()
    st = h.gelex_b200_grm_run_events(str(bfile).encode(), str(out_prefix).encode(), m, MODES[mode], int(loco), int(chunk_size), int(precision),
                                     int(n_gpus), ev.ctypes.data_as(C.c_void_p), max_events, C.byref(n), err, len(err))
    if st != 0:
        raise EngineError(st, err.value.decode())
    assert n.value <= max_events, "event buffer too small"
    return [(EVENT_NAMES[int(t)], int(a), int(b)) for t, a, b in ev[: n.value]]


def plan(bfile, mode="A", loco=False, max_samples=1 << 22, max_items=4096):
    """Metadata the engine derives from .fam/.bim (no GPU needed): row order projection, SNP count, work items."""
    h = load_host()
    r2o = np.full(max_samples, -2, dtype=np.int32)
    n_raw, m, dense, n_items = C.c_int64(), C.c_int64(), C.c_int(), C.c_int()
    names = C.create_string_buffer(1 << 20)
    gb = np.zeros(max_items, dtype=np.int64)
    ge = np.zeros(max_items, dtype=np.int64)
    err = C.create_string_buffer(2048)
    n_out = h.gelex_b200_plan(str(bfile).encode(), MODES[mode], int(loco), r2o.ctypes.data_as(C.c_void_p), max_samples, C.byref(n_raw), C.byref(m),
                              C.byref(dense), names, len(names), gb.ctypes.data_as(C.c_void_p), ge.ctypes.data_as(C.c_void_p), max_items,
                              C.byref(n_items), err, len(err))
    if n_out < 0:
        raise EngineError(int(n_out), err.value.decode())
    lines = names.value.decode().split("\n")
    k = n_items.value
    return {"n_out": int(n_out), "n_raw": n_raw.value, "m": m.value, "dense": bool(dense.value), "raw_to_out": r2o[:n_raw.value].copy(),
            "items": [(lines[i], int(gb[i]), int(ge[i])) for i in range(k)], "pattern_suffix": lines[k]}


def load_grm(prefix, chr_prefix=None, normalize=True, id_map=None):
    """GrmLoader::load / load_unnormalized (optionally filtered + reordered by id_map = {(fid, iid): target_row}) or, with chr_prefix,
    LocoGRMLoader::load_loco_grm: dense symmetric FP64 matrix as the reference's consumers receive it.  The expansion runs on the GPU."""
    h = load_host()
    ids, targets, n_ids = None, None, 0
    if id_map:
        keys = list(id_map.keys())
        ids = "\n".join(f"{fid}\x1f{iid}" for fid, iid in keys).encode()
        targets = np.array([id_map[k] for k in keys], dtype=np.int64)
        n_ids = len(keys)
    err = C.create_string_buffer(2048)
    cp = str(chr_prefix).encode() if chr_prefix else None
    tp = targets.ctypes.data_as(C.c_void_p) if targets is not None else None
    if n_ids:
        n = int(targets.max()) + 1
    else:
        id_path = Path(str(prefix) + ".id")
        n = sum(1 for ln in id_path.read_text().split("\n") if ln) if id_path.exists() else 0
    out = np.empty((n, n), dtype=np.float64, order="F")
    got = h.gelex_b200_grm_load(str(prefix).encode(), cp, int(normalize), ids, tp, n_ids, out.ctypes.data_as(C.c_void_p), n * n, err, len(err))
    if got < 0:
        raise EngineError(int(got), err.value.decode())
    assert got == n, (got, n)
    return out


def row_slices(n, parts):
    """The C++ engine's row-slice rule (same contract as sharding.row_slices; compared in tests/test_sharding.py)."""
    h = load_host()
    out = np.zeros(2 * parts, dtype=np.int64)
    got = h.gelex_b200_row_slices(int(n), int(parts), out.ctypes.data_as(C.c_void_p))
    if got != parts:
        raise EngineError(got, "row_slices failed")
    return [(int(out[2 * r]), int(out[2 * r + 1])) for r in range(parts)]


def _genotype_dims(bfile):
    h = load_host()
    dims = np.zeros(3, dtype=np.int64)
    err = C.create_string_buffer(2048)
    st = h.gelex_b200_genotype_load(str(bfile).encode(), 1, 1, 1, 0, None, 0, dims.ctypes.data_as(C.c_void_p), None, None, None, None, 0, err, len(err))
    if st != 0:
        raise EngineError(int(st), err.value.decode())
    return int(dims[0]), int(dims[1])


def _stats_dict(dims, mean, sd, freq, mono):
    return {"num_samples": int(dims[0]), "num_variants": int(dims[1]), "mean": mean, "stddev": sd, "freq": freq,
            "monomorphic_indices": mono[: int(dims[2])].copy()}


def genotype_export(bfile, out_prefix, method="OSH", additive=True, chunk_size=10000, device=0):
    """GenotypePipe(bed, prefix).process<GT>(method, chunk) (genotype_pipe.h:55-89): writes <prefix>.bmat (column-major FP64 n x m) and
    <prefix>.snpstats; returns the per-SNP statistics the .snpstats file holds (+ A1 frequencies)."""
    h = load_host()
    n, m = _genotype_dims(bfile)
    mth = _lib.METHODS[method] if isinstance(method, str) else int(method)
    dims = np.zeros(3, dtype=np.int64)
    mean, sd, freq = (np.empty(m, dtype=np.float64) for _ in range(3))
    mono = np.empty(max(1, m), dtype=np.int64)
    err = C.create_string_buffer(2048)
    p = lambda a: a.ctypes.data_as(C.c_void_p)  # noqa: E731
    st = h.gelex_b200_genotype_export(str(bfile).encode(), str(out_prefix).encode(), mth, int(additive), int(chunk_size), int(device), p(dims),
                                      p(mean), p(sd), p(freq), p(mono), m, err, len(err))
    if st != 0:
        raise EngineError(int(st), err.value.decode())
    return _stats_dict(dims, mean, sd, freq, mono)


def genotype_load(bfile, method="OSH", additive=True, chunk_size=10000, device=0):
    """GenotypeLoader(bed).process<GT>(method, chunk) (genotype_loader.h:52-80): the standardised (n, m) matrix (Fortran order, as
    Eigen holds it) + statistics."""
    h = load_host()
    n, m = _genotype_dims(bfile)
    mth = _lib.METHODS[method] if isinstance(method, str) else int(method)
    out = np.empty((n, m), dtype=np.float64, order="F")
    dims = np.zeros(3, dtype=np.int64)
    mean, sd, freq = (np.empty(m, dtype=np.float64) for _ in range(3))
    mono = np.empty(max(1, m), dtype=np.int64)
    err = C.create_string_buffer(2048)
    p = lambda a: a.ctypes.data_as(C.c_void_p)  # noqa: E731
    st = h.gelex_b200_genotype_load(str(bfile).encode(), mth, int(additive), int(chunk_size), int(device), p(out), n * m, p(dims), p(mean), p(sd),
                                    p(freq), p(mono), m, err, len(err))
    if st != 0:
        raise EngineError(int(st), err.value.decode())
    return out, _stats_dict(dims, mean, sd, freq, mono)


def read_snp_stats(path):
    """The .snpstats layout (snp_stats_writer.cpp:50-91) as GenotypeMap::load_metadata reads it (genotype_mmap.cpp:64-106)."""
    raw = Path(path).read_bytes()
    n, m, k = (int(v) for v in np.frombuffer(raw, dtype=np.int64, count=3))
    off = 24
    mono = np.frombuffer(raw, dtype=np.int64, count=k, offset=off).copy()
    off += 8 * k
    mean = np.frombuffer(raw, dtype=np.float64, count=m, offset=off).copy()
    sd = np.frombuffer(raw, dtype=np.float64, count=m, offset=off + 8 * m).copy()
    assert len(raw) == off + 16 * m, "trailing bytes in .snpstats"
    return {"num_samples": n, "num_variants": m, "monomorphic_indices": mono, "mean": mean, "stddev": sd}


def grm_run_gcta(bfile, out_prefix, method="SH", chunk_size=10000, precision=_lib.PREC_FP16, n_gpus=1, writer_threads=4,
                 gcta_diag=False, partition="snp"):
    """GrmEngine with Config::gcta: writes <out>.grm.bin (A_ij = G_ij / N_ij), <out>.grm.N.bin (N_ij = SNPs called in both
    samples) and <out>.grm.id - the GCTA binary format the reference's docs name (docs/source/data_formats.rst:188-208).
    gcta_diag: GCTA's default diagonal instead of sum z^2 / N (method SH only); partition: "snp" | "tile" | "auto" when n_gpus > 1."""
    h = load_host()
    m = _lib.METHODS[method] if isinstance(method, str) else int(method)
    err = C.create_string_buffer(2048)
    st = h.gelex_b200_grm_run_gcta_ex(str(bfile).encode(), str(out_prefix).encode(), m, int(chunk_size), int(precision), int(n_gpus),
                                      int(writer_threads), int(bool(gcta_diag)), {"auto": 0, "snp": 1, "tile": 2}[partition], err, len(err))
    if st != 0:
        raise EngineError(int(st), err.value.decode())


def strip_owners(n, world, bands):
    """The C++ engine's strip dealing for the band-pipelined SNP-shard run (same contract as sharding.strip_owners: per rank a list of
    (band, row_begin, row_end)); `bands` = [(row_begin, row_end)] as ggrm_band_layout gives them."""
    h = load_host()
    ends = np.array([e for _, e in bands], dtype=np.int64)
    cap = (int(n) + 2047) // 2048 + 1
    out = np.zeros(4 * cap, dtype=np.int64)
    cnt = h.gelex_b200_strip_owners(int(n), int(world), ends.ctypes.data_as(C.c_void_p), len(ends), out.ctypes.data_as(C.c_void_p), cap)
    if cnt < 0 or cnt > cap:
        raise EngineError(cnt, "strip_owners failed")
    per = [[] for _ in range(world)]
    for i in range(cnt):
        r, b, lo, hi = (int(v) for v in out[4 * i:4 * i + 4])
        per[r].append((b, lo, hi))
    return per


def gpus_worth_using(n, m, kinds=1, precision=_lib.PREC_FP16, offered=8, n_outputs=1):
    """How many of `offered` GPUs the CLI uses for an n x m problem writing n_outputs matrices (GrmEngine::gpus_worth_using)."""
    return int(load_host().gelex_b200_gpus_worth_using(int(n), int(m), int(kinds), int(precision), int(offered), int(n_outputs)))


def piece_owners(n, world, bands):
    """The C++ engine's dealing of strip column pieces (same contract as sharding.piece_owners: per rank a list of
    (band, super_row, super_col_begin, super_col_end))."""
    h = load_host()
    ends = np.array([e for _, e in bands], dtype=np.int64)
    ns = (int(n) + 2047) // 2048
    cap = ns * (ns + 1) // 2 + 1
    out = np.zeros(5 * cap, dtype=np.int64)
    cnt = h.gelex_b200_piece_owners(int(n), int(world), ends.ctypes.data_as(C.c_void_p), len(ends), out.ctypes.data_as(C.c_void_p), cap)
    if cnt < 0 or cnt > cap:
        raise EngineError(cnt, "piece_owners failed")
    per = [[] for _ in range(world)]
    for i in range(cnt):
        r, b, si, c0, c1 = (int(v) for v in out[5 * i:5 * i + 5])
        per[r].append((b, si, c0, c1))
    return per


def relay_selftest(n_chunks, world, slots, chunk_bytes=4096, fail_at_chunk=-1):
    """Protocol of the 2D-tile BED relay on host buffers (GrmEngine::relay_selftest): number of (consumer, chunk) pairs that read
    anything but the chunk's own bytes."""
    h = load_host()
    err = C.create_string_buffer(2048)
    r = h.gelex_b200_relay_selftest(int(n_chunks), int(world), int(slots), int(chunk_bytes), int(fail_at_chunk), err, len(err))
    if r < 0:
        raise EngineError(int(r), err.value.decode())
    return int(r)


def write_id_file(path, ids):
    """GrmIdWriter::write mirror: canonical 'FID\x1fIID' strings -> 'FID\tIID\n' lines."""
    h = load_host()
    err = C.create_string_buffer(2048)
    blob = b"".join(i.encode() + b"\0" for i in ids)
    st = h.gelex_b200_write_id_file(str(path).encode(), blob, len(ids), err, len(err))
    if st != 0:
        raise EngineError(st, err.value.decode())


def format_bed_path(bfile):
    h = load_host()
    err, out = C.create_string_buffer(2048), C.create_string_buffer(4096)
    st = h.gelex_b200_format_bed_path(str(bfile).encode(), out, len(out), err, len(err))
    if st != 0:
        raise EngineError(st, err.value.decode())
    return out.value.decode()
