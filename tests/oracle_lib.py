"""ctypes loader of oracle/liboracle.so (the CPU restatement; test infrastructure only)."""
import ctypes as C
import subprocess
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
ORACLE_SO = ROOT / "oracle" / "liboracle.so"
REF_BIN = ROOT / "oracle" / "_ref" / "gelex_ref"


class LocusStat(C.Structure):
    _fields_ = [("maf", C.c_double), ("mean", C.c_double), ("stddev", C.c_double), ("is_monomorphic", C.c_int32), ("valid", C.c_int32)]


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not ORACLE_SO.exists():
            subprocess.check_call(["make", "-C", str(ROOT / "oracle"), str(ORACLE_SO)])
        l = C.CDLL(str(ORACLE_SO))
        vp, i64 = C.c_void_p, C.c_int64
        l.oracle_decode_chunk.argtypes = [vp, i64, i64, i64, i64, vp, vp]
        l.oracle_decode_chunk.restype = None
        l.oracle_process_column.argtypes = [vp, i64, C.c_int, C.c_int, C.POINTER(LocusStat)]
        l.oracle_process_column.restype = None
        l.oracle_process_matrix.argtypes = [vp, i64, i64, C.c_int, C.c_int, vp]
        l.oracle_process_matrix.restype = None
        l.oracle_syrk_lower.argtypes = [vp, i64, i64, vp]
        l.oracle_syrk_lower.restype = None
        l.oracle_grm_compute.argtypes = [vp, i64, i64, i64, i64, vp, C.c_int, C.c_int, vp, vp, C.c_int, i64, vp, C.POINTER(C.c_double), vp]
        l.oracle_pack_lower_f32.argtypes = [vp, i64, vp]
        l.oracle_pack_lower_f32.restype = None
        l.oracle_write_bin.argtypes = [vp, i64, C.c_char_p]
        l.oracle_load_normalized.argtypes = [vp, i64, vp]
        l.oracle_load_normalized.restype = None
        l.oracle_loco.argtypes = [vp, vp, i64, vp]
        l.oracle_loco.restype = None
        l.oracle_counts.argtypes = [vp, i64, vp]
        l.oracle_counts.restype = None
        l.oracle_standardise_chunk.argtypes = [vp, i64, i64, i64, i64, vp, C.c_int, C.c_int, vp, vp, vp]
        l.oracle_standardise_chunk.restype = None
        _lib = l
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def decode_chunk(payload, n_raw, n_out=None, raw_to_out=None):
    n_out = n_raw if n_out is None else n_out
    bpv = (n_raw + 3) // 4
    payload = np.ascontiguousarray(payload, dtype=np.uint8)
    c = payload.size // bpv
    out = np.empty((c, n_out), dtype=np.float64)
    r2o = None if raw_to_out is None else np.ascontiguousarray(raw_to_out, dtype=np.int32)
    lib().oracle_decode_chunk(_p(payload), c, bpv, n_raw, n_out, _p(r2o), _p(out))
    return out.T


def process_column(col, method, kind):
    v = np.array(col, dtype=np.float64)
    st = LocusStat()
    lib().oracle_process_column(_p(v), v.size, method, kind, C.byref(st))
    return v, st


def process_matrix(chunk_nc, method, kind):
    """chunk_nc: (n, c) array; returns processed (n, c) and freqs."""
    z = np.array(np.asarray(chunk_nc, dtype=np.float64).T, order="C", copy=True)  # (c, n) C-order == column-major n x c; never alias the input
    freqs = np.empty(z.shape[0], dtype=np.float64)
    lib().oracle_process_matrix(_p(z), z.shape[1], z.shape[0], method, kind, _p(freqs))
    return z.T, freqs


def grm_compute(payload, n_raw, method, kind, ranges=None, chunk=10000, n_out=None, raw_to_out=None):
    """Returns (G lower-triangle as (n,n) array with G[i,j] valid for i>=j, denominator, freqs)."""
    n_out = n_raw if n_out is None else n_out
    bpv = (n_raw + 3) // 4
    payload = np.ascontiguousarray(payload, dtype=np.uint8)
    m = payload.size // bpv
    if ranges is None:
        ranges = [(0, m)]
    rb = np.array([r[0] for r in ranges], dtype=np.int64)
    re = np.array([r[1] for r in ranges], dtype=np.int64)
    G = np.empty((n_out, n_out), dtype=np.float64)  # column-major: G[j, i] in numpy index == grm(i, j)
    den = C.c_double()
    freqs = np.zeros(m, dtype=np.float64)
    r2o = None if raw_to_out is None else np.ascontiguousarray(raw_to_out, dtype=np.int32)
    st = lib().oracle_grm_compute(_p(payload), m, bpv, n_raw, n_out, _p(r2o), method, kind, _p(rb), _p(re), len(ranges), chunk,
                                  _p(G), C.byref(den), _p(freqs))
    if st != 0:
        raise RuntimeError(f"oracle_grm_compute failed: {st}")
    return G.T, den.value, freqs


def pack_lower_f32(G_rowview):
    """G_rowview[i, j] = grm(i, j) for i >= j.  Returns packed float32 rows (the .bin payload)."""
    n = G_rowview.shape[0]
    Gc = np.ascontiguousarray(G_rowview.T)  # column-major storage of grm
    out = np.empty(n * (n + 1) // 2, dtype=np.float32)
    lib().oracle_pack_lower_f32(_p(Gc), n, _p(out))
    return out


def pack_lower_f64(G_rowview):
    n = G_rowview.shape[0]
    idx = np.tril_indices(n)
    return np.ascontiguousarray(G_rowview[idx])


def load_normalized(packed_f32, n):
    out = np.empty((n, n), dtype=np.float64)
    packed = np.ascontiguousarray(packed_f32, dtype=np.float32)
    lib().oracle_load_normalized(_p(packed), n, _p(out))
    return out


def loco(whole_packed, chr_packed, n):
    out = np.empty((n, n), dtype=np.float64)
    w = np.ascontiguousarray(whole_packed, dtype=np.float32)
    c = np.ascontiguousarray(chr_packed, dtype=np.float32)
    lib().oracle_loco(_p(w), _p(c), n, _p(out))
    return out


def row_spot(payload, n_raw, method, kind, rows, groups=None, chunk=4096, n_out=None, raw_to_out=None):
    """Checker for full-size configs: selected ROWS of the FP64 matrix Z Z^T and its full diagonal, per SNP group, without
    ever forming the n x n matrix.  The standardised chunk comes from the oracle's own per-column arithmetic
    (oracle_standardise_chunk = oracle_decode_chunk + oracle_process_column); the rows are an FP64 GEMM of it (numpy / BLAS).
    groups: list of (begin, end) SNP ranges (default one group = all SNPs).
    Returns (R, D): R[g] is (len(rows), n_out) with R[g][k, j] = G_g(rows[k], j); D[g] is the diagonal of G_g (n_out)."""
    n_out = n_raw if n_out is None else n_out
    bpv = (n_raw + 3) // 4
    payload = np.ascontiguousarray(payload, dtype=np.uint8).reshape(-1, bpv)
    m = payload.shape[0]
    groups = [(0, m)] if groups is None else list(groups)
    rows = np.asarray(rows, dtype=np.int64)
    r2o = None if raw_to_out is None else np.ascontiguousarray(raw_to_out, dtype=np.int32)
    R = [np.zeros((rows.size, n_out), dtype=np.float64) for _ in groups]
    D = [np.zeros(n_out, dtype=np.float64) for _ in groups]
    buf = np.empty((min(chunk, m), n_out), dtype=np.float64)  # (c, n) C-order == column-major n x c
    for gi, (b, e) in enumerate(groups):
        for s in range(b, e, chunk):
            c = min(chunk, e - s)
            z = buf[:c]
            lib().oracle_standardise_chunk(_p(payload[s:s + c]), c, bpv, n_raw, n_out, _p(r2o), method, kind, _p(z), _p(D[gi]), None)
            R[gi] += z[:, rows].T @ z
    return R, D
