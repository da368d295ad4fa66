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
        _host = h
    return _host


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
