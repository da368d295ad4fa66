"""ctypes binding of the C ABI in include/ggrm.h (libggrm_b200.so, built in-tree by gelex_b200/csrc/Makefile).

The library is the product; this module only marshals numpy buffers across the boundary.  There is no fallback:
if the shared object is missing, import fails loudly; if no sm_100 GPU is present, ggrm_create fails with
GGRM_ERR_NO_DEVICE and GrmError is raised.
"""
from __future__ import annotations

import ctypes as C
from pathlib import Path

import numpy as np

_HERE = Path(__file__).resolve().parent
LIB_PATH = _HERE / "libggrm_b200.so"

ABI_VERSION = 1
KIND_ADD, KIND_DOM = 0, 1
GROUP_WHOLE = -1
PREC_BF16, PREC_FP64, PREC_FP16 = 0, 1, 2
SECOND_DOMINANCE, SECOND_NONMISSING, SECOND_NONMISSING_GCTA_DIAG = 0, 1, 2
ALLMISS_REFERENCE_NAN, ALLMISS_SKIP = 0, 1

# GenotypeProcessMethod (reference include/gelex/data/genotype/genotype_processor.h:38-48)
METHODS = {
    "StandardizeHWE": 1, "SH": 1,
    "CenterHWE": 2, "CH": 2,
    "OrthStandardizeHWE": 3, "OSH": 3,
    "OrthCenterHWE": 4, "OCH": 4,
    "Standardize": 5, "S": 5,
    "Center": 6, "C": 6,
    "OrthStandardize": 7, "OS": 7,
    "OrthCenter": 8, "OC": 8,
}

STATUS_NAMES = {
    0: "GGRM_OK", -1: "GGRM_ERR_INVALID_ARGUMENT", -2: "GGRM_ERR_COLUMN_RANGE", -3: "GGRM_ERR_INVALID_OPERATION",
    -4: "GGRM_ERR_OUT_OF_MEMORY", -5: "GGRM_ERR_CUDA", -6: "GGRM_ERR_NO_DEVICE", -7: "GGRM_ERR_FILE_FORMAT",
}


class GrmError(RuntimeError):
    def __init__(self, status: int, message: str):
        super().__init__(f"{STATUS_NAMES.get(status, status)}: {message}")
        self.status = status
        self.message = message


class Config(C.Structure):
    _fields_ = [
        ("abi_version", C.c_int32), ("device", C.c_int32),
        ("n_raw", C.c_int64), ("n_out", C.c_int64),
        ("raw_to_out", C.POINTER(C.c_int32)),
        ("m_total", C.c_int64),
        ("method", C.c_int32), ("want_add", C.c_int32), ("want_dom", C.c_int32), ("precision", C.c_int32),
        ("n_groups", C.c_int32),
        ("group_begin", C.POINTER(C.c_int64)), ("group_end", C.POINTER(C.c_int64)),
        ("keep_whole", C.c_int32), ("keep_stats", C.c_int32), ("all_missing_policy", C.c_int32),
        ("tile_rank", C.c_int32), ("tile_world", C.c_int32),
        ("chunk_snps", C.c_int64),
        ("max_resident_groups", C.c_int32), ("second_matrix", C.c_int32),
    ]


EXPORTS = [
    "ggrm_create", "ggrm_destroy", "ggrm_last_error", "ggrm_push", "ggrm_push_device", "ggrm_finish", "ggrm_finish_async", "ggrm_set_push_end", "ggrm_reset", "ggrm_snp_stats",
    "ggrm_fetch", "ggrm_fetch_f64", "ggrm_release_group", "ggrm_accum_buffer", "ggrm_reduce_peer",
    "ggrm_fetch_normalized", "ggrm_fetch_loco", "ggrm_get_timers", "ggrm_decode_chunk", "ggrm_standardise_chunk",
    "ggrm_synth_bed_host", "ggrm_synth_bed_device", "ggrm_device_count", "ggrm_version", "ggrm_enable_peer_access",
    "ggrm_ipc_export", "ggrm_ipc_open", "ggrm_ipc_close", "ggrm_row_slice",
    "ggrm_fetch_block", "ggrm_owned_supertiles", "ggrm_host_alloc", "ggrm_host_free", "ggrm_bin_to_dense", "ggrm_assoc_chunk",
    "ggrm_band_layout", "ggrm_set_band_flags", "ggrm_reduce_rows_async", "ggrm_genotype_chunk", "ggrm_fetch_gcta", "ggrm_accum_address",
    "ggrm_stage_create", "ggrm_stage_destroy", "ggrm_stage_upload", "ggrm_stage_slot", "ggrm_wait_copies",
    "ggrm_warm_device", "ggrm_reduce_block_async", "ggrm_host_register", "ggrm_host_unregister",
    "ggrm_snp_stats_noalloc", "ggrm_fetch_block_gcta", "ggrm_gcta_diag_address", "ggrm_gcta_diag_reduce_async",
]

_lib = None


def load() -> C.CDLL:
    """Load libggrm_b200.so (fails loudly if it has not been built: run `python -c 'import __graft_entry__ as g; g.build()'`)."""
    global _lib
    if _lib is not None:
        return _lib
    if not LIB_PATH.exists():
        raise ImportError(f"{LIB_PATH} not built; run __graft_entry__.build() (nvcc, sm_100a). There is no CPU fallback.")
    lib = C.CDLL(str(LIB_PATH))
    vp, i32, i64, u64, dbl = C.c_void_p, C.c_int32, C.c_int64, C.c_uint64, C.c_double
    lib.ggrm_create.argtypes = [C.POINTER(Config), C.POINTER(vp)]
    lib.ggrm_destroy.argtypes = [vp]
    lib.ggrm_destroy.restype = None
    lib.ggrm_last_error.argtypes = [vp]
    lib.ggrm_last_error.restype = C.c_char_p
    lib.ggrm_push.argtypes = [vp, vp, i64, i64, i64]
    lib.ggrm_push_device.argtypes = [vp, vp, i64, i64, i64]
    lib.ggrm_finish.argtypes = [vp]
    lib.ggrm_finish_async.argtypes = [vp]
    lib.ggrm_set_push_end.argtypes = [vp, i64]
    lib.ggrm_reset.argtypes = [vp]
    lib.ggrm_snp_stats.argtypes = [vp, i64, i64, vp, vp, vp, vp, vp]
    lib.ggrm_fetch.argtypes = [vp, C.c_int, C.c_int, i64, i64, vp, C.POINTER(dbl)]
    lib.ggrm_fetch_f64.argtypes = [vp, C.c_int, C.c_int, i64, i64, vp, C.POINTER(dbl)]
    lib.ggrm_release_group.argtypes = [vp, C.c_int, C.c_int]
    lib.ggrm_accum_buffer.argtypes = [vp, C.c_int, C.c_int, C.POINTER(vp), C.POINTER(i64), C.POINTER(i32)]
    lib.ggrm_reduce_peer.argtypes = [vp, C.c_int, C.c_int, C.POINTER(vp), C.c_int, i64, i64]
    lib.ggrm_fetch_normalized.argtypes = [vp, C.c_int, C.c_int, i64, i64, vp]
    lib.ggrm_fetch_loco.argtypes = [vp, C.c_int, C.c_int, i64, i64, vp]
    lib.ggrm_get_timers.argtypes = [vp, C.POINTER(dbl)]
    lib.ggrm_ipc_export.argtypes = [vp, C.c_int, C.c_int, vp]
    lib.ggrm_ipc_open.argtypes = [C.c_int, vp, C.POINTER(vp)]
    lib.ggrm_ipc_close.argtypes = [C.c_int, vp]
    lib.ggrm_row_slice.argtypes = [vp, i64, i64, C.POINTER(i64), C.POINTER(i64)]
    lib.ggrm_band_layout.argtypes = [vp, C.POINTER(i64), C.c_int, C.POINTER(C.c_int)]
    lib.ggrm_set_band_flags.argtypes = [vp, vp, i32]
    lib.ggrm_reduce_rows_async.argtypes = [vp, C.c_int, C.c_int, vp, C.c_int, i64, i64]
    lib.ggrm_fetch_block.argtypes = [vp, C.c_int, C.c_int, i64, i64, i64, i64, vp, i64]
    lib.ggrm_owned_supertiles.argtypes = [vp, C.POINTER(i64), C.POINTER(i64), i64, C.POINTER(i64)]
    lib.ggrm_bin_to_dense.argtypes = [C.c_int, vp, i64, vp, vp, i64, vp, i64, dbl, vp]
    lib.ggrm_host_alloc.argtypes = [i64, C.POINTER(vp)]
    lib.ggrm_host_free.argtypes = [vp]
    lib.ggrm_decode_chunk.argtypes = [C.c_int, vp, i64, i64, i64, i64, vp, vp]
    lib.ggrm_standardise_chunk.argtypes = [C.c_int, vp, i64, i64, i64, i64, vp, C.c_int, C.c_int, C.c_int, vp, vp]
    lib.ggrm_assoc_chunk.argtypes = [C.c_int, vp, i64, i64, i64, i64, vp, C.c_int, C.c_int, vp, vp]
    lib.ggrm_accum_address.argtypes = [vp, C.c_int, C.POINTER(vp)]
    lib.ggrm_fetch_gcta.argtypes = [vp, C.c_int, C.c_int, i64, i64, vp]
    lib.ggrm_genotype_chunk.argtypes = [C.c_int, vp, i64, i64, i64, i64, vp, C.c_int, C.c_int, vp, vp, vp, vp, vp]
    lib.ggrm_synth_bed_host.argtypes = [vp, i64, i64, i64, u64, dbl]
    lib.ggrm_synth_bed_device.argtypes = [C.c_int, vp, i64, i64, i64, u64, dbl]
    lib.ggrm_stage_create.argtypes = [C.c_int, i64, C.c_int, C.POINTER(vp)]
    lib.ggrm_stage_destroy.argtypes = [vp]
    lib.ggrm_stage_destroy.restype = None
    lib.ggrm_stage_upload.argtypes = [vp, C.c_int, vp, i64]
    lib.ggrm_stage_slot.argtypes = [vp, C.c_int, C.POINTER(vp)]
    lib.ggrm_wait_copies.argtypes = [vp]
    lib.ggrm_warm_device.argtypes = [C.c_int]
    lib.ggrm_host_register.argtypes = [vp, i64]
    lib.ggrm_host_unregister.argtypes = [vp]
    lib.ggrm_snp_stats_noalloc.argtypes = [C.c_int, vp, i64, i64, i64, vp, vp, vp]
    lib.ggrm_fetch_block_gcta.argtypes = [vp, C.c_int, C.c_int, i64, i64, i64, i64, vp, i64]
    lib.ggrm_gcta_diag_address.argtypes = [vp, C.POINTER(vp)]
    lib.ggrm_gcta_diag_reduce_async.argtypes = [vp, C.POINTER(vp), C.c_int]
    lib.ggrm_reduce_block_async.argtypes = [vp, C.c_int, C.c_int, C.POINTER(vp), C.c_int, i64, i64, i64]
    lib.ggrm_device_count.argtypes = []
    lib.ggrm_version.argtypes = []
    lib.ggrm_version.restype = C.c_char_p
    _lib = lib
    return lib


def _ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def tri(n: int) -> int:
    return n * (n + 1) // 2


class Accumulator:
    """One ggrm handle = one GPU's GRM accumulator(s).  Thin, explicit mirror of the C ABI."""

    def __init__(self, n_raw, m_total, method=3, want_add=True, want_dom=False, precision=PREC_BF16, raw_to_out=None,
                 n_out=None, groups=None, keep_whole=True, keep_stats=False, all_missing_policy=ALLMISS_REFERENCE_NAN,
                 device=0, chunk_snps=0, tile_rank=0, tile_world=1, max_resident_groups=0, second_matrix=0):
        self.lib = load()
        self.n_raw = int(n_raw)
        self.n_out = int(n_out if n_out is not None else n_raw)
        self.m_total = int(m_total)
        self.bpv = (self.n_raw + 3) // 4
        cfg = Config()
        cfg.abi_version = ABI_VERSION
        cfg.device = device
        cfg.n_raw, cfg.n_out, cfg.m_total = self.n_raw, self.n_out, self.m_total
        self._r2o = None
        if raw_to_out is not None:
            self._r2o = np.ascontiguousarray(raw_to_out, dtype=np.int32)
            cfg.raw_to_out = self._r2o.ctypes.data_as(C.POINTER(C.c_int32))
        cfg.method = METHODS[method] if isinstance(method, str) else int(method)
        cfg.want_add, cfg.want_dom = int(bool(want_add)), int(bool(want_dom))
        cfg.precision = precision
        self.groups = None
        if groups is not None and len(groups) > 0:
            self.groups = [(int(b), int(e)) for b, e in groups]
            self._gb = np.array([g[0] for g in self.groups], dtype=np.int64)
            self._ge = np.array([g[1] for g in self.groups], dtype=np.int64)
            cfg.n_groups = len(self.groups)
            cfg.group_begin = self._gb.ctypes.data_as(C.POINTER(C.c_int64))
            cfg.group_end = self._ge.ctypes.data_as(C.POINTER(C.c_int64))
        else:
            cfg.n_groups = 1
        cfg.keep_whole, cfg.keep_stats = int(bool(keep_whole)), int(bool(keep_stats))
        cfg.all_missing_policy = all_missing_policy
        cfg.tile_rank, cfg.tile_world = tile_rank, tile_world
        cfg.chunk_snps = chunk_snps
        cfg.max_resident_groups = max_resident_groups
        cfg.second_matrix = second_matrix  # 1: the dom slot holds the non-missing pair counts (GCTA .grm.N.bin); 2: + GCTA's default diagonal
        self.precision = precision
        h = C.c_void_p()
        st = self.lib.ggrm_create(C.byref(cfg), C.byref(h))
        if st != 0:
            raise GrmError(st, self.lib.ggrm_last_error(None).decode())
        self.h = h

    def _check(self, st):
        if st != 0:
            raise GrmError(st, self.lib.ggrm_last_error(self.h).decode())

    def close(self):
        if getattr(self, "h", None):
            self.lib.ggrm_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def push(self, payload: np.ndarray, snp_begin: int, snp_count: int | None = None):
        payload = np.ascontiguousarray(payload, dtype=np.uint8)
        if snp_count is None:
            snp_count = payload.size // self.bpv
        self._check(self.lib.ggrm_push(self.h, _ptr(payload), snp_begin, snp_count, self.bpv))

    def push_ptr(self, host_ptr: int, snp_begin: int, snp_count: int):
        self._check(self.lib.ggrm_push(self.h, C.c_void_p(host_ptr), snp_begin, snp_count, self.bpv))

    def push_device(self, dev_ptr: int, snp_begin: int, snp_count: int):
        self._check(self.lib.ggrm_push_device(self.h, C.c_void_p(dev_ptr), snp_begin, snp_count, self.bpv))

    def wait_copies(self):
        """Every copy enqueued by earlier push calls has read its source (a staging slot may be overwritten)."""
        self._check(self.lib.ggrm_wait_copies(self.h))

    def finish(self):
        self._check(self.lib.ggrm_finish(self.h))

    def finish_async(self):
        self._check(self.lib.ggrm_finish_async(self.h))

    def set_push_end(self, snp_end: int):
        self._check(self.lib.ggrm_set_push_end(self.h, snp_end))

    def reset(self):
        self._check(self.lib.ggrm_reset(self.h))

    def snp_stats(self, snp_begin: int, snp_count: int):
        n0, n1, n2, nm = (np.empty(snp_count, dtype=np.int32) for _ in range(4))
        fr = np.empty(snp_count, dtype=np.float64)
        self._check(self.lib.ggrm_snp_stats(self.h, snp_begin, snp_count, _ptr(n0), _ptr(n1), _ptr(n2), _ptr(nm), _ptr(fr)))
        return n0, n1, n2, nm, fr

    def fetch(self, kind=KIND_ADD, group=GROUP_WHOLE, row_begin=0, row_count=None, dtype=np.float32, out=None):
        if row_count is None:
            row_count = self.n_out - row_begin
        size = tri(row_begin + row_count) - tri(row_begin)
        if out is None:
            out = np.empty(size, dtype=dtype)
        assert out.size >= size and out.dtype == dtype
        tr = C.c_double(0.0)
        fn = self.lib.ggrm_fetch if dtype == np.float32 else self.lib.ggrm_fetch_f64
        self._check(fn(self.h, kind, group, row_begin, row_count, _ptr(out), C.byref(tr)))
        return out, tr.value

    def fetch_normalized(self, kind=KIND_ADD, group=GROUP_WHOLE, row_begin=0, row_count=None):
        row_count = self.n_out - row_begin if row_count is None else row_count
        out = np.empty(tri(row_begin + row_count) - tri(row_begin), dtype=np.float64)
        self._check(self.lib.ggrm_fetch_normalized(self.h, kind, group, row_begin, row_count, _ptr(out)))
        return out

    def fetch_loco(self, kind, group, row_begin=0, row_count=None):
        row_count = self.n_out - row_begin if row_count is None else row_count
        out = np.empty(tri(row_begin + row_count) - tri(row_begin), dtype=np.float64)
        self._check(self.lib.ggrm_fetch_loco(self.h, kind, group, row_begin, row_count, _ptr(out)))
        return out

    def release_group(self, kind, group):
        self._check(self.lib.ggrm_release_group(self.h, kind, group))

    def accum_buffer(self, kind=KIND_ADD, group=GROUP_WHOLE):
        p, n, eb = C.c_void_p(), C.c_int64(), C.c_int32()
        self._check(self.lib.ggrm_accum_buffer(self.h, kind, group, C.byref(p), C.byref(n), C.byref(eb)))
        return p.value, n.value, eb.value

    def reduce_peer(self, kind, group, peer_ptrs, slice_begin, slice_end):
        arr = (C.c_void_p * max(1, len(peer_ptrs)))(*peer_ptrs)
        self._check(self.lib.ggrm_reduce_peer(self.h, kind, group, arr, len(peer_ptrs), slice_begin, slice_end))

    def accum_address(self, kind=KIND_ADD):
        p = C.c_void_p()
        self._check(self.lib.ggrm_accum_address(self.h, kind, C.byref(p)))
        return p.value

    def band_layout(self):
        """[(row_begin, row_end)] of the bands the final accumulation completes in (see ggrm_band_layout)."""
        ends = (C.c_int64 * 64)()
        nb = C.c_int()
        self._check(self.lib.ggrm_band_layout(self.h, ends, 64, C.byref(nb)))
        out, prev = [], 0
        for b in range(nb.value):
            out.append((prev, ends[b]))
            prev = ends[b]
        return out

    def set_band_flags(self, flags_addr: int, value: int):
        self._check(self.lib.ggrm_set_band_flags(self.h, C.c_void_p(flags_addr), value))

    def reduce_rows_async(self, kind, group, peer_ptrs, row_begin, row_end):
        arr = (C.c_void_p * max(1, len(peer_ptrs)))(*peer_ptrs)
        self._check(self.lib.ggrm_reduce_rows_async(self.h, kind, group, arr, len(peer_ptrs), row_begin, row_end))

    def reduce_block_async(self, kind, group, peer_ptrs, super_row, super_col_begin, super_col_end):
        arr = (C.c_void_p * max(1, len(peer_ptrs)))(*peer_ptrs)
        self._check(self.lib.ggrm_reduce_block_async(self.h, kind, group, arr, len(peer_ptrs), super_row, super_col_begin, super_col_end))

    def ipc_export(self, kind=KIND_ADD, group=GROUP_WHOLE) -> bytes:
        buf = C.create_string_buffer(64)
        self._check(self.lib.ggrm_ipc_export(self.h, kind, group, buf))
        return buf.raw

    def row_slice(self, row_begin: int, row_end: int):
        b, e = C.c_int64(), C.c_int64()
        self._check(self.lib.ggrm_row_slice(self.h, row_begin, row_end, C.byref(b), C.byref(e)))
        return b.value, e.value

    def fetch_gcta(self, which, group=GROUP_WHOLE, row_begin=0, row_count=None):
        """which = 0: A = G_add / N (GCTA .grm.bin values); 1: N (.grm.N.bin).  Needs second_matrix=1, want_add and want_dom."""
        if row_count is None:
            row_count = self.n_out - row_begin
        out = np.empty(tri(row_begin + row_count) - tri(row_begin), dtype=np.float32)
        self._check(self.lib.ggrm_fetch_gcta(self.h, which, group, row_begin, row_count, _ptr(out)))
        return out

    def fetch_block_gcta(self, which, group, row_begin, row_count, col_begin, col_count):
        """fetch_gcta as a rectangular block (the 2D-tile writer's unit)."""
        out = np.empty((row_count, col_count), dtype=np.float32)
        self._check(self.lib.ggrm_fetch_block_gcta(self.h, which, group, row_begin, row_count, col_begin, col_count, _ptr(out), col_count))
        return out

    def gcta_diag_address(self) -> int:
        p = C.c_void_p()
        self._check(self.lib.ggrm_gcta_diag_address(self.h, C.byref(p)))
        return p.value

    def gcta_diag_reduce_async(self, peer_ptrs):
        """SNP shards with second_matrix=2: total diagonal sums = own + the peers' (gcta_diag_address of their handles)."""
        arr = (C.c_void_p * max(1, len(peer_ptrs)))(*peer_ptrs)
        self._check(self.lib.ggrm_gcta_diag_reduce_async(self.h, arr, len(peer_ptrs)))

    def fetch_block(self, kind, group, row_begin, row_count, col_begin, col_count):
        out = np.empty((row_count, col_count), dtype=np.float32)
        self._check(self.lib.ggrm_fetch_block(self.h, kind, group, row_begin, row_count, col_begin, col_count, _ptr(out), col_count))
        return out

    def owned_supertiles(self):
        cnt = C.c_int64()
        self._check(self.lib.ggrm_owned_supertiles(self.h, None, None, 0, C.byref(cnt)))
        r = (C.c_int64 * max(1, cnt.value))()
        c = (C.c_int64 * max(1, cnt.value))()
        self._check(self.lib.ggrm_owned_supertiles(self.h, r, c, cnt.value, C.byref(cnt)))
        return [(r[i], c[i]) for i in range(cnt.value)]

    def timers(self):
        out = (C.c_double * 8)()
        self._check(self.lib.ggrm_get_timers(self.h, out))
        keys = ["h2d_ms", "prep_ms", "expand_ms", "syrk_ms", "fetch_ms", "syrk_launches", "kernel_launches", "pass_ms"]
        return dict(zip(keys, list(out)))


def decode_chunk(payload, n_raw, n_out=None, raw_to_out=None, device=0):
    lib = load()
    n_out = n_raw if n_out is None else n_out
    bpv = (n_raw + 3) // 4
    payload = np.ascontiguousarray(payload, dtype=np.uint8)
    c = payload.size // bpv
    out = np.empty((c, n_out), dtype=np.float64)  # column-major n_out x c == C-order (c, n_out)
    r2o = None if raw_to_out is None else np.ascontiguousarray(raw_to_out, dtype=np.int32)
    st = lib.ggrm_decode_chunk(device, _ptr(payload), c, bpv, n_raw, n_out, _ptr(r2o), _ptr(out))
    if st != 0:
        raise GrmError(st, lib.ggrm_last_error(None).decode())
    return out.T  # (n_out, c) view


def standardise_chunk(payload, n_raw, method, kind, n_out=None, raw_to_out=None, all_missing_policy=0, device=0):
    lib = load()
    n_out = n_raw if n_out is None else n_out
    bpv = (n_raw + 3) // 4
    payload = np.ascontiguousarray(payload, dtype=np.uint8)
    c = payload.size // bpv
    out = np.empty((c, n_out), dtype=np.float64)
    freq = np.empty(c, dtype=np.float64)
    r2o = None if raw_to_out is None else np.ascontiguousarray(raw_to_out, dtype=np.int32)
    m = METHODS[method] if isinstance(method, str) else int(method)
    st = lib.ggrm_standardise_chunk(device, _ptr(payload), c, bpv, n_raw, n_out, _ptr(r2o), m, kind, all_missing_policy,
                                    _ptr(out), _ptr(freq))
    if st != 0:
        raise GrmError(st, lib.ggrm_last_error(None).decode())
    return out.T, freq


def assoc_chunk(payload, n_raw, method, kind, n_out=None, raw_to_out=None, device=0):
    """dispatch_assoc_chunk_by_method (assoc_detail.cpp:28-47): decode + centre-family code policy + per-SNP A1 frequency."""
    lib = load()
    n_out = n_raw if n_out is None else n_out
    bpv = (n_raw + 3) // 4
    payload = np.ascontiguousarray(payload, dtype=np.uint8)
    c = payload.size // bpv
    out = np.empty((c, n_out), dtype=np.float64)
    freq = np.empty(c, dtype=np.float64)
    r2o = None if raw_to_out is None else np.ascontiguousarray(raw_to_out, dtype=np.int32)
    m = METHODS[method] if isinstance(method, str) else int(method)
    st = lib.ggrm_assoc_chunk(device, _ptr(payload), c, bpv, n_raw, n_out, _ptr(r2o), m, kind, _ptr(out), _ptr(freq))
    if st != 0:
        raise GrmError(st, lib.ggrm_last_error(None).decode())
    return out.T, freq


def genotype_chunk(payload, n_raw, method, kind, n_out=None, raw_to_out=None, device=0):
    """GenotypeLoader/GenotypePipe::process_chunk (genotype_loader.cpp:51-79): standardised (n_out, c) chunk + per-SNP LocusStatistic
    as a dict {freq, mean, stddev, mono}."""
    lib = load()
    n_out = n_raw if n_out is None else n_out
    bpv = (n_raw + 3) // 4
    payload = np.ascontiguousarray(payload, dtype=np.uint8)
    c = payload.size // bpv
    out = np.empty((c, n_out), dtype=np.float64)
    freq, mean, sd = (np.empty(c, dtype=np.float64) for _ in range(3))
    mono = np.empty(c, dtype=np.uint8)
    r2o = None if raw_to_out is None else np.ascontiguousarray(raw_to_out, dtype=np.int32)
    m = METHODS[method] if isinstance(method, str) else int(method)
    st = lib.ggrm_genotype_chunk(device, _ptr(payload), c, bpv, n_raw, n_out, _ptr(r2o), m, kind, _ptr(out), _ptr(freq), _ptr(mean), _ptr(sd),
                                 _ptr(mono))
    if st != 0:
        raise GrmError(st, lib.ggrm_last_error(None).decode())
    return out.T, {"freq": freq, "mean": mean, "stddev": sd, "mono": mono.astype(bool)}


NOALLOC_SCRATCH_BYTES = 8 * 1024 * 1024  # GGRM_NOALLOC_SCRATCH_BYTES


def snp_stats_noalloc(dev_bed_ptr: int, snp_count: int, n_raw: int, dev_counts4_ptr: int | None, dev_freq_ptr: int | None, device=0, stream=None):
    """ggrm_snp_stats_noalloc: device pointers in and out, kernels only enqueued on `stream` (None = default); allocates nothing."""
    lib = load()
    st = lib.ggrm_snp_stats_noalloc(device, C.c_void_p(dev_bed_ptr), snp_count, (n_raw + 3) // 4, n_raw,
                                    C.c_void_p(dev_counts4_ptr) if dev_counts4_ptr else None, C.c_void_p(dev_freq_ptr) if dev_freq_ptr else None,
                                    C.c_void_p(stream) if stream else None)
    if st != 0:
        raise GrmError(st, lib.ggrm_last_error(None).decode())


def synth_bed_host(n, snp_begin, snp_count, seed, missing_rate=0.01):
    lib = load()
    out = np.empty((snp_count, (n + 3) // 4), dtype=np.uint8)
    st = lib.ggrm_synth_bed_host(_ptr(out), n, snp_begin, snp_count, seed, missing_rate)
    if st != 0:
        raise GrmError(st, "synth_bed_host failed")
    return out


def synth_bed_device(dev_ptr, n, snp_begin, snp_count, seed, missing_rate=0.01, device=0):
    lib = load()
    st = lib.ggrm_synth_bed_device(device, C.c_void_p(dev_ptr), n, snp_begin, snp_count, seed, missing_rate)
    if st != 0:
        raise GrmError(st, lib.ggrm_last_error(None).decode())


def ipc_open(handle: bytes, device=0) -> int:
    lib = load()
    p = C.c_void_p()
    st = lib.ggrm_ipc_open(device, C.c_char_p(handle), C.byref(p))
    if st != 0:
        raise GrmError(st, lib.ggrm_last_error(None).decode())
    return p.value


def ipc_close(ptr: int, device=0):
    lib = load()
    st = lib.ggrm_ipc_close(device, C.c_void_p(ptr))
    if st != 0:
        raise GrmError(st, lib.ggrm_last_error(None).decode())


def device_count() -> int:
    return load().ggrm_device_count()


class Stage:
    """Device staging slots (ggrm_stage_*): one host upload of a BED chunk that several handles take with push_device."""

    def __init__(self, slot_bytes: int, n_slots: int = 2, device: int = 0):
        self.lib = load()
        self.h = C.c_void_p()
        st = self.lib.ggrm_stage_create(device, int(slot_bytes), int(n_slots), C.byref(self.h))
        if st != 0:
            raise GrmError(st, self.lib.ggrm_last_error(None).decode())

    def upload(self, slot: int, payload: np.ndarray):
        payload = np.ascontiguousarray(payload, dtype=np.uint8)
        st = self.lib.ggrm_stage_upload(self.h, int(slot), _ptr(payload), payload.size)
        if st != 0:
            raise GrmError(st, self.lib.ggrm_last_error(None).decode())

    def slot(self, slot: int) -> int:
        p = C.c_void_p()
        st = self.lib.ggrm_stage_slot(self.h, int(slot), C.byref(p))
        if st != 0:
            raise GrmError(st, self.lib.ggrm_last_error(None).decode())
        return p.value

    def close(self):
        if getattr(self, "h", None):
            self.lib.ggrm_stage_destroy(self.h)
            self.h = None

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()
