"""Generates the committed golden fixtures in this directory by RUNNING THE UNMODIFIED REFERENCE
(oracle/_ref/gelex_ref, built from /root/reference by oracle/Makefile).  Run in the build container only:

    make -C oracle && python tests/golden/make_golden.py

Fixtures (all small):
  tiny.bed/.bim/.fam      37 raw samples (shuffled IDs, one duplicated ID), 96 SNPs on 4 chromosome runs, 5 % missing,
                          one monomorphic SNP, one all-but-one-missing SNP
  tiny_chunks.npz         BedPipe::load_chunk output + process_matrix output + freqs for 8 methods x {Add, Dom}
  tiny_<tag>.bin/.id      GrmEngine outputs: SH add, OSH add+dom, OCH add+dom --loco
  tiny_loaders.npz        GrmLoader::load and LocoGRMLoader::load_loco_grm outputs
  dense.bed/.bim/.fam     40 samples already in sorted order (dense decode path), 50 SNPs; dense_chunks.npz, dense_SH_add.bin
  process_vectors.json    per-column process() outputs for the columns used in the reference's own unit tests
  tiny_pipe_<tag>.bmat/.snpstats   GenotypePipe::process<GT> outputs on the tiny fixture (`python tests/golden/make_golden.py pipe`
                          regenerates only these): OSH add chunk 25, SH dom chunk 7, OC dom chunk 40, S add chunk 33
"""
import json
import subprocess
import sys
from pathlib import Path

import numpy as np

HERE = Path(__file__).resolve().parent
ROOT = HERE.parent.parent
sys.path.insert(0, str(ROOT / "tests"))
import plink_io  # noqa: E402

REF = ROOT / "oracle" / "_ref" / "gelex_ref"


def run(*args):
    out = subprocess.run([str(REF), *map(str, args)], check=True, capture_output=True, text=True)
    return out.stdout


def read_dump(path):
    raw = np.fromfile(path, dtype=np.uint8)
    n, c = np.frombuffer(raw[:16].tobytes(), dtype=np.int64)
    body = np.frombuffer(raw[16:].tobytes(), dtype=np.float64)
    return int(n), int(c), body


def chunk_dump(prefix, start, end, method, kind, tmp):
    run("chunk", prefix, start, end, method, kind, tmp)
    n, c, body = read_dump(tmp)
    raw = body[: n * c].reshape(c, n).T
    freqs = body[n * c: n * c + c]
    z = body[n * c + c:].reshape(c, n).T
    return raw, freqs, z


def make_tiny():
    rng = np.random.default_rng(20261104)
    n, m = 37, 96
    p = rng.uniform(0.05, 0.5, size=m)
    geno = rng.binomial(2, p, size=(n, m)).astype(np.float64)
    geno[rng.random((n, m)) < 0.05] = np.nan
    geno[:, 10] = 0.0                 # monomorphic
    geno[:, 11] = np.nan
    geno[5, 11] = 1.0                 # a single valid call
    geno[:, 12] = 1.0                 # all heterozygous
    payload = plink_io.encode_matrix(geno)
    ids = [(f"fam{(i % 5) + 1}", f"sample{i + 1}") for i in range(n)]   # reference tests/bed_fixture.cpp:353-371
    ids[20] = ids[3]                  # duplicated ID: later raw sample wins
    chroms = ["1"] * 30 + ["2"] * 25 + ["X"] * 17 + ["7"] * 24
    plink_io.write_plink(HERE / "tiny", payload, ids, chroms, fam_delim=" ", bim_delim="\t")
    return n, m


def make_dense():
    rng = np.random.default_rng(7)
    n, m = 40, 50
    p = rng.uniform(0.1, 0.5, size=m)
    geno = rng.binomial(2, p, size=(n, m)).astype(np.float64)
    geno[rng.random((n, m)) < 0.02] = np.nan
    payload = plink_io.encode_matrix(geno)
    ids = [("F", f"s{i:04d}") for i in range(n)]
    chroms = ["1"] * 20 + ["2"] * 30
    plink_io.write_plink(HERE / "dense", payload, ids, chroms)
    return n, m


def main():
    if not REF.exists():
        raise SystemExit("oracle/_ref/gelex_ref missing: run `make -C oracle` in a container that has /root/reference")
    tmp = HERE / "_tmp.bin"
    n, m = make_tiny()
    chunks = {}
    for method in range(1, 9):
        for kind in ("A", "D"):
            raw, freqs, z = chunk_dump(HERE / "tiny", 0, m, method, kind, tmp)
            chunks[f"z_{method}_{kind}"] = z
            chunks[f"freq_{method}_{kind}"] = freqs
    chunks["raw"] = raw
    raw2, _, _ = chunk_dump(HERE / "tiny", 17, 61, 1, "A", tmp)
    chunks["raw_17_61"] = raw2
    np.savez_compressed(HERE / "tiny_chunks.npz", **chunks)

    run("grm", HERE / "tiny", HERE / "tiny_SH", 1, "A", 0, 25, 2)
    run("grm", HERE / "tiny", HERE / "tiny_OSH", 3, "AD", 0, 10000, 2)
    run("grm", HERE / "tiny", HERE / "tiny_OCH", 4, "AD", 0, 7, 2)
    run("grm", HERE / "tiny", HERE / "tiny_OCH", 4, "AD", 1, 7, 2)
    run("grm", HERE / "tiny", HERE / "tiny_S", 5, "AD", 0, 96, 2)
    loaders = {}
    run("load", HERE / "tiny_SH.add", tmp)
    nn, _, body = read_dump(tmp)
    loaders["SH_add_normalized"] = body.reshape(nn, nn)
    for c in ("1", "2", "X", "7"):
        run("loco", HERE / "tiny_OCH.add", str(HERE / "tiny_OCH.add") + f".chr{c}", tmp)
        nn, _, body = read_dump(tmp)
        loaders[f"OCH_add_loco_chr{c}"] = body.reshape(nn, nn)
    np.savez_compressed(HERE / "tiny_loaders.npz", **loaders)

    n2, m2 = make_dense()
    dchunks = {}
    for method, kind in ((1, "A"), (3, "D"), (6, "D")):
        raw, freqs, z = chunk_dump(HERE / "dense", 0, m2, method, kind, tmp)
        dchunks[f"z_{method}_{kind}"] = z
        dchunks[f"freq_{method}_{kind}"] = freqs
    dchunks["raw"] = raw
    np.savez_compressed(HERE / "dense_chunks.npz", **dchunks)
    run("grm", HERE / "dense", HERE / "dense_SH", 1, "A", 0, 10000, 2)

    # Columns of the reference's own unit tests (tests/test_genotype_processor.cpp, tests/test_grm_code_policy.cpp)
    cols = {
        "c01210": [0, 1, 2, 1, 0],
        "c012102": [0, 1, 2, 1, 0, 2],
        "mono1": [1, 1, 1, 1, 1],
        "mono0": [0, 0, 0, 0],
        "with_nan": [0, "nan", 2, 1, 1, "nan", 0, 2],
        "hom2": [2, 2, 2, 2],
    }
    vectors = []
    for name, col in cols.items():
        for method in range(1, 9):
            for kind in ("A", "D"):
                out = json.loads(run("process", method, kind, *col))
                vectors.append({"column": name, "input": col, "method": method, "kind": kind, **out})
    with open(HERE / "process_vectors.json", "w") as f:
        json.dump(vectors, f)
    tmp.unlink(missing_ok=True)
    print("golden fixtures written to", HERE)


PIPE_CASES = [("OSH", 3, "A", 25), ("SH", 1, "D", 7), ("OC", 8, "D", 40), ("S", 5, "A", 33)]


def make_pipe():
    """GenotypePipe outputs of the unmodified reference (ref_driver `pipe`).  Its process() ends by mapping the files it has just
    written while the writers are still open, which throws on outputs this small ("Failed to read scalar: rows"); the files are
    complete once the pipe is destroyed, and the driver reports that exception instead of failing."""
    for tag, method, kind, chunk in PIPE_CASES:
        prefix = HERE / f"tiny_pipe_{tag}.{'add' if kind == 'A' else 'dom'}"
        for ext in (".bmat", ".snpstats"):
            Path(str(prefix) + ext).unlink(missing_ok=True)
        info = json.loads(run("pipe", HERE / "tiny", prefix, method, kind, chunk).strip().splitlines()[-1])
        print(prefix.name, info)


if __name__ == "__main__":
    if sys.argv[1:] == ["pipe"]:
        make_pipe()
    else:
        main()
        make_pipe()
