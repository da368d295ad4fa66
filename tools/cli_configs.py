"""End-to-end timing of the named configs through the drop-in CLI (`gelex-b200 grm`), real PLINK files in, real .bin/.id
files out: t_wall (process start -> files closed) next to the device pass time the CLI reports.

    python tools/cli_configs.py C1 C3 C2 [--gpus N] [--dir /dev/shm/ggrm] [--precision bf16] [--partition auto]

Files live on tmpfs (/dev/shm) unless --dir says otherwise; the BED payload is generated on the GPU (same counter-based
generator as bench.py) and written once.  With --loco configs (C4) finished per-chromosome files are deleted as soon as the
next one appears (115 GB of output does not fit beside the inputs), after their size has been checked.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time
from pathlib import Path

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))
import gelex_b200 as g  # noqa: E402
from config_probe import CONFIGS, chromosome_runs  # noqa: E402
from gelex_b200.engine import CLI_PATH  # noqa: E402


def write_triple(prefix: Path, n, m, seed, loco):
    bpv = (n + 3) // 4
    t0 = time.time()
    with open(str(prefix) + ".bed", "wb") as f:
        f.write(bytes([0x6C, 0x1B, 0x01]))
        step = max(1, (1 << 30) // bpv)
        d = torch.empty(step * bpv, dtype=torch.uint8, device="cuda")
        for s in range(0, m, step):
            c = min(step, m - s)
            g.synth_bed_device(d.data_ptr(), n, s, c, seed, 0.01)
            f.write(d[: c * bpv].cpu().numpy().tobytes())
        del d
    with open(str(prefix) + ".fam", "w") as f:
        f.write("".join(f"F\ts{i:07d}\t0\t0\t1\t-9\n" for i in range(n)))
    runs = chromosome_runs(m)
    with open(str(prefix) + ".bim", "w") as f:
        for ci, (b, e) in enumerate(runs):
            f.write("".join(f"{ci + 1}\tsnp{j + 1}\t0\t{1000 + j}\tA\tG\n" for j in range(b, e)))
    torch.cuda.empty_cache()
    return time.time() - t0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("configs", nargs="+")
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--dir", default="/dev/shm/ggrm")
    ap.add_argument("--precision", default="bf16")
    ap.add_argument("--partition", default="auto")
    ap.add_argument("--writer-threads", type=int, default=4)
    args = ap.parse_args()
    d = Path(args.dir)
    d.mkdir(parents=True, exist_ok=True)
    for name in args.configs:
        c = CONFIGS[name]
        n, m = c["n"], c["m"]
        prefix = d / name
        gen_s = write_triple(prefix, n, m, 20261104 + int(name[1]), c["loco"])
        out = d / (name + "_out")
        cmd = [str(CLI_PATH), "grm", "-b", str(prefix), "-o", str(out), "--gm", c["method"], "--precision", args.precision, "--gpus", str(args.gpus),
               "--partition", args.partition, "--writer-threads", str(args.writer_threads)]
        if c["add"]:
            cmd.append("--add")
        if c["dom"]:
            cmd.append("--dom")
        if c["loco"]:
            cmd += ["--loco", "--with-whole"]
        expect_bytes = n * (n + 1) // 2 * 4
        seen, stop = {}, threading.Event()

        def reaper():  # LOCO only: keep the newest 1 + gpus .bin files, delete older finished ones after recording their size
            while not stop.is_set():
                bins = sorted(d.glob(name + "_out.*.chr*.bin"), key=lambda p: p.stat().st_mtime)
                for p in bins[:-(1 + args.gpus)]:  # every GPU may have one file in progress (those have the newest mtimes)
                    seen[p.name] = p.stat().st_size
                    p.unlink()
                time.sleep(0.2)

        th = threading.Thread(target=reaper, daemon=True)
        if c["loco"]:
            th.start()
        t0 = time.time()
        r = subprocess.run(cmd, capture_output=True, text=True)
        wall = time.time() - t0
        stop.set()
        for p in d.glob(name + "_out.*.bin"):
            seen[p.name] = p.stat().st_size
        dev_line = [ln for ln in r.stdout.splitlines() if "Device" in ln]
        kinds = int(c["add"]) + int(c["dom"])
        flops = n * (n + 1) * m * kinds
        rec = {"config": name, "n": n, "m": m, "gpus": args.gpus, "precision": args.precision, "partition": args.partition, "rc": r.returncode,
               "cli_wall_s": wall, "effective_tflops_wall": flops / wall / 1e12, "bed_bytes": m * ((n + 3) // 4),
               "out_files": len(seen), "out_bytes": int(sum(seen.values())), "all_sizes_ok": all(v == expect_bytes for v in seen.values()),
               "device_line": dev_line[0].strip() if dev_line else None, "input_gen_s": gen_s, "dir": str(d),
               "writer": os.environ.get("GELEX_B200_WRITER", "auto"), "bed_relay": os.environ.get("GELEX_B200_BED_RELAY", "0"), "prefault": os.environ.get("GELEX_B200_PREFAULT", "0")}
        if r.returncode != 0:
            rec["stderr"] = r.stderr[-800:]
            rec["stdout"] = r.stdout[-800:]
        ph = [ln for ln in r.stdout.splitlines() if "Host phases" in ln]
        rec["host_phases"] = ph[0].strip() if ph else None
        if os.environ.get("GGRM_DEBUG_TIMING"):
            rec["create_timing"] = [ln.strip() for ln in r.stderr.splitlines() if "[ggrm] create" in ln][:40]
        # spot check: diagonal mean of the whole-genome additive file ~ non-missing SNP count (StandardizeHWE)
        whole = d / (name + "_out.add.bin")
        if whole.exists() and c["method"] == "SH":
            mm = np.memmap(whole, dtype=np.float32, mode="r")
            idx = np.cumsum(np.arange(1, n + 1, dtype=np.int64)) - 1
            rec["diag_mean_over_m"] = float(np.mean(mm[idx[:: max(1, n // 4096)]].astype(np.float64)) / m)
            del mm
        print(json.dumps(rec), flush=True)
        for p in list(d.glob(name + "*")):
            p.unlink()


if __name__ == "__main__":
    main()
