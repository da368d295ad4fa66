"""Times the named configs of BASELINE.json (C1-C5) on ONE GPU with the packed BED resident in HBM, through the C ABI
(ggrm_push_device -> ggrm_finish), and prints one JSON line per config.  Run under gpurun:

    python tools/config_probe.py C1 C2 C3 C4 C5 [--precision bf16|fp16] [--passes 2]

FLOPs are SURVEY 8d's algorithmic figure n(n+1)m per output matrix (LOCO partials sum to the whole: counted once).
"""
import argparse
import json
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import gelex_b200 as g  # noqa: E402

# GRCh38 autosome lengths in Mb: LOCO run lengths are proportional to these (SURVEY 8d "run lengths ∝ human chromosome sizes")
CHR_MB = [248, 242, 198, 190, 182, 171, 159, 145, 138, 134, 135, 133, 114, 107, 102, 90, 83, 80, 59, 64, 47, 51]

try:  # the driver-written measured peaks (bf16 cuBLAS, sustained); fallback = the value this pool measured on 2026-11-04
    SUSTAINED = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["bf16_tflops_sustained"])
except Exception:  # noqa: BLE001
    SUSTAINED = 1385.8

CONFIGS = {
    "C1": dict(n=2_000, m=50_000, method="SH", add=True, dom=False, loco=False),
    "C2": dict(n=50_000, m=500_000, method="SH", add=True, dom=False, loco=False),
    "C3": dict(n=20_000, m=300_000, method="OCH", add=True, dom=True, loco=False),
    "C4": dict(n=50_000, m=600_000, method="SH", add=True, dom=False, loco=True),
    "C5": dict(n=200_000, m=600_000, method="SH", add=True, dom=False, loco=False),
}


def chromosome_runs(m):
    tot = sum(CHR_MB)
    cuts = [0]
    acc = 0
    for mb in CHR_MB:
        acc += mb
        cuts.append(round(m * acc / tot))
    cuts[-1] = m
    return [(cuts[i], cuts[i + 1]) for i in range(22)]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("configs", nargs="+")
    ap.add_argument("--precision", default="bf16", choices=["bf16", "fp16"])
    ap.add_argument("--passes", type=int, default=2)
    args = ap.parse_args()
    prec = g.PREC_BF16 if args.precision == "bf16" else g.PREC_FP16
    for name in args.configs:
        c = CONFIGS[name]
        n, m = c["n"], c["m"]
        bpv = (n + 3) // 4
        seed = 20261104 + int(name[1])
        bed = torch.empty(m * bpv, dtype=torch.uint8, device="cuda")
        for s in range(0, m, 65536):
            g.synth_bed_device(bed.data_ptr() + s * bpv, n, s, min(65536, m - s), seed, 0.01)
        torch.cuda.synchronize()
        groups = chromosome_runs(m) if c["loco"] else None
        acc = g.Accumulator(n, m, method=c["method"], want_add=c["add"], want_dom=c["dom"], precision=prec, groups=groups, keep_whole=True)
        kinds = int(c["add"]) + int(c["dom"])
        flops = n * (n + 1) * m * kinds
        best = None
        prev = acc.timers()
        for p in range(args.passes):
            if p:
                acc.reset()
            torch.cuda.synchronize()
            t0 = time.time()
            acc.push_device(bed.data_ptr(), 0, m)
            acc.finish()
            wall = time.time() - t0
            t = acc.timers()
            d = {k: t[k] - prev[k] for k in t}
            prev = t
            if best is None or d["pass_ms"] < best["pass_ms"]:
                best = dict(d, wall_s=wall)
        # sanity: StandardizeHWE diagonal mean ~ number of non-missing SNPs; every LOCO partial's trace sums to the whole's
        rows = min(n, 2048)
        _, tr = acc.fetch(g.KIND_ADD, g.GROUP_WHOLE, 0, rows)
        check = {"diag_mean_first_rows_over_m": tr / rows / m}
        if c["loco"]:
            parts = sum(acc.fetch(g.KIND_ADD, gi, 0, rows)[1] for gi in range(22))
            check["loco_partials_trace_over_whole"] = parts / tr
        print(json.dumps({"config": name, "n": n, "m": m, "method": c["method"], "kinds": kinds, "loco_groups": 22 if c["loco"] else 0,
                          "precision": args.precision, "device_pass_ms": best["pass_ms"], "wall_s": best["wall_s"],
                          "effective_tflops": flops / best["pass_ms"] / 1e9, "syrk_ms": best["syrk_ms"], "prep_ms": best["prep_ms"],
                          "expand_ms": best["expand_ms"], "syrk_launches": int(best["syrk_launches"]),
                          "frac_of_sustained": flops / best["pass_ms"] / 1e9 / SUSTAINED, "sustained_peak_tflops": SUSTAINED, "check": check}), flush=True)
        acc.close()
        del bed
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
