"""Tiny PLINK triple writer/reader helpers for tests and benchmarks (not product code)."""
from pathlib import Path

import numpy as np

BED_MAGIC = bytes([0x6C, 0x1B, 0x01])


def encode_matrix(geno: np.ndarray) -> np.ndarray:
    """geno: (n, m) float array of 0/1/2/NaN (A1 allele counts) -> packed (m, ceil(n/4)) uint8.
    Encoding per reference tests/bed_fixture.cpp:72-95: NaN->01, 0->11, 1->10, 2->00; sample i in bits 2*(i%4) of byte i/4."""
    n, m = geno.shape
    codes = np.full((n, m), 1, dtype=np.uint8)
    codes[geno == 0] = 3
    codes[geno == 1] = 2
    codes[geno == 2] = 0
    bpv = (n + 3) // 4
    padded = np.zeros((bpv * 4, m), dtype=np.uint8)
    padded[:n] = codes
    p = padded.reshape(bpv, 4, m)
    packed = p[:, 0] | (p[:, 1] << 2) | (p[:, 2] << 4) | (p[:, 3] << 6)
    return np.ascontiguousarray(packed.T)


def decode_payload(payload: np.ndarray, n: int) -> np.ndarray:
    """packed (m, bpv) -> (n, m) float64 with NaN for missing (pure numpy; for building expectations in tests)."""
    lut = np.array([2.0, np.nan, 1.0, 0.0])
    m = payload.shape[0]
    codes = np.stack([(payload >> (2 * k)) & 3 for k in range(4)], axis=2).reshape(m, -1)[:, :n]
    return lut[codes].T


def write_plink(prefix, payload: np.ndarray, fam_ids, chroms, fam_delim="\t", bim_delim="\t"):
    """fam_ids: list of (fid, iid); chroms: list of chromosome names, one per SNP."""
    prefix = Path(prefix)
    prefix.parent.mkdir(parents=True, exist_ok=True)
    with open(str(prefix) + ".bed", "wb") as f:
        f.write(BED_MAGIC)
        f.write(np.ascontiguousarray(payload, dtype=np.uint8).tobytes())
    with open(str(prefix) + ".fam", "w") as f:
        for fid, iid in fam_ids:
            f.write(fam_delim.join([fid, iid, "0", "0", "1", "-9"]) + "\n")
    with open(str(prefix) + ".bim", "w") as f:
        for j, c in enumerate(chroms):
            f.write(bim_delim.join([str(c), f"snp{j + 1}", "0", str(1000 + j), "A", "G"]) + "\n")


def sample_order(fam_ids):
    """Reference row order: sort + unique of FID<US>IID (sample_manager.cpp:37-40). Returns (sorted_ids, raw_to_out)."""
    canon = [fid + "\x1f" + iid for fid, iid in fam_ids]
    uniq = sorted(set(canon))
    idx = {s: i for i, s in enumerate(uniq)}
    raw_to_out = np.array([idx[c] for c in canon], dtype=np.int32)
    return uniq, raw_to_out


def chrom_runs(chroms):
    """Maximal runs of equal consecutive chromosome names (grm_work_plan.cpp:54-81) -> [(name, begin, end)]."""
    runs = []
    start = 0
    for i in range(1, len(chroms) + 1):
        if i == len(chroms) or chroms[i] != chroms[start]:
            runs.append((str(chroms[start]), start, i))
            start = i
    return runs


def read_bin(path, n):
    a = np.fromfile(path, dtype=np.float32)
    assert a.size == n * (n + 1) // 2, (a.size, n)
    return a
