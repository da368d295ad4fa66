"""CPU-only measurement of the engine's .bin writer pool (GrmEngine::writer_selftest: the BinFile + WriterPool code the engine uses,
fed from host memory instead of ggrm_fetch*), pwrite against the shared mapping, and the mapping after a prefault (Config::prefault: the populate phase, which the
engine overlaps with the GPU pass, is timed separately).  No GPU involved; the numbers are those of the
host it runs on (say which in the output).

    python tools/writer_probe.py [--n 36000] [--dirs /dev/shm /tmp] > profiles/r01_writer_cpu.jsonl
"""
import argparse
import json
import os
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from gelex_b200 import engine  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=36000)
    ap.add_argument("--dirs", nargs="+", default=["/dev/shm", "/tmp"])
    ap.add_argument("--reps", type=int, default=2)
    a = ap.parse_args()
    gb = a.n * (a.n + 1) / 2 * 4 / 1e9
    for d in a.dirs:
        path = Path(d) / f"gelex_b200_writer_probe_{os.getpid()}.bin"
        for tiles, parts, wt in ((False, 1, 4), (False, 8, 4), (True, 1, 4), (True, 8, 4)):
            for w, pft in (("pwrite", 0), ("mmap", 0), ("auto", 0), ("auto", 4)):
                best = None
                for _ in range(a.reps):
                    s, mapped, pf = engine.writer_selftest(path, a.n, tiles=tiles, parts=parts, writer_threads=wt, writer=w, prefault_threads=pft,
                                                           with_prefault_seconds=True)
                    path.unlink()
                    if best is None or s < best[0]:
                        best = (s, pf)
                print(json.dumps({"dir": d, "n": a.n, "file_gb": round(gb, 3), "jobs": "super-tile row segments" if tiles else "whole-row blocks",
                                  "pools": parts, "writer_threads": wt, "writer": w, "mapped": mapped, "prefault_threads": pft,
                                  "prefault_seconds": round(best[1], 3), "seconds": round(best[0], 3), "gb_per_s": round(gb / best[0], 2),
                                  "host_cpus": os.cpu_count()}), flush=True)


if __name__ == "__main__":
    main()
