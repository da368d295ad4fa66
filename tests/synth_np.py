"""numpy restatement of gelex_b200/csrc/synth.h (the synthetic PLINK BED payload generator of SURVEY 8d), for harness code
that must not load the product library: bench.py's reference arm generates its PLINK triple with this, so that arm's process
maps none of the repo's shared objects.  Byte-equal to ggrm_synth_bed_host (tests/test_abi.py::test_numpy_synth_matches_library).
Test / benchmark infrastructure, not product code."""
from concurrent.futures import ThreadPoolExecutor

import numpy as np

_M64 = np.uint64(0xFFFFFFFFFFFFFFFF)


def _splitmix64(x):
    """x: uint64 array (modified copy); arithmetic wraps modulo 2^64 like the C code."""
    x = x + np.uint64(0x9E3779B97F4A7C15)
    x = (x ^ (x >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
    x = (x ^ (x >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
    return x ^ (x >> np.uint64(31))


def _thresholds(seed, snps):
    """ggrm_synth_thresholds for an int64 array of SNP indices -> (t_hom2, t_het) as uint64 arrays (values < 2^32)."""
    s1 = (snps + 1).astype(np.uint64)
    h = _splitmix64(np.uint64(seed) ^ (np.uint64(0xD1B54A32D192ED03) * s1))
    h32 = h >> np.uint64(32)
    P = np.uint64(214748365) + ((h32 * np.uint64(1932735283)) >> np.uint64(32))
    Q = np.uint64(4294967296) - P
    t1 = (P * P) >> np.uint64(32)
    # 2*P*Q < 2^64 needs care: P < 2^31, Q <= 2^32 - 2^27.7 -> 2PQ < 2^64 holds (P*Q < 2^63)
    t2 = t1 + ((np.uint64(2) * P * Q) >> np.uint64(32))
    t2 = np.minimum(t2, np.uint64(0xFFFFFFFF))
    return t1, t2


def _rows(n, snp_begin, snp_count, seed, miss_thr24):
    bpv = (n + 3) // 4
    n4 = bpv * 4
    snps = np.arange(snp_begin, snp_begin + snp_count, dtype=np.int64)
    t1, t2 = _thresholds(seed, snps)
    with np.errstate(over="ignore"):
        a = (np.uint64(seed) + np.uint64(0x632BE59BD9B4E019) * (snps + 1).astype(np.uint64))[:, None]
        b = (np.uint64(0x9E3779B97F4A7C15) * np.arange(1, n4 + 1, dtype=np.uint64))[None, :]
        h = _splitmix64(a ^ b)
    u = h >> np.uint64(32)
    mlow = h & np.uint64(0xFFFFFF)
    code = np.full(h.shape, 3, dtype=np.uint8)
    code[u < t2[:, None]] = 2
    code[u < t1[:, None]] = 0
    code[mlow < np.uint64(miss_thr24)] = 1
    code[:, n:] = 0  # padding bits of the last byte
    c = code.reshape(snp_count, bpv, 4)
    return (c[:, :, 0] | (c[:, :, 1] << 2) | (c[:, :, 2] << 4) | (c[:, :, 3] << 6)).astype(np.uint8)


def synth_bed(n, snp_begin, snp_count, seed, missing_rate, threads=8, block=256):
    """(snp_count, ceil(n/4)) uint8, the bytes ggrm_synth_bed_host writes for the same arguments."""
    # miss_threshold() of ggrm_api_hooks.inl
    miss_thr24 = 0 if not missing_rate > 0.0 else (1 << 24 if missing_rate >= 1.0 else int(missing_rate * 16777216.0))
    out = np.empty((snp_count, (n + 3) // 4), dtype=np.uint8)
    starts = list(range(0, snp_count, block))

    def work(s):
        c = min(block, snp_count - s)
        with np.errstate(over="ignore"):
            out[s:s + c] = _rows(n, snp_begin + s, c, seed, miss_thr24)

    if threads <= 1 or len(starts) == 1:
        for s in starts:
            work(s)
    else:
        with ThreadPoolExecutor(max_workers=threads) as ex:
            list(ex.map(work, starts))
    return out
