"""FP64 close-agreement mode vs the FP64 rate of this GPU ("FP64 mode is reported against the FP64 tensor peak
measured the same way" as MEASURED_PEAKS.json: torch.matmul, 2*N^3 flops, best of several)."""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import gelex_b200 as g  # noqa: E402


def dgemm_peak(nn=8192, reps=5):
    a = torch.randn(nn, nn, dtype=torch.float64, device="cuda")
    b = torch.randn(nn, nn, dtype=torch.float64, device="cuda")
    torch.matmul(a, b)
    torch.cuda.synchronize()
    best = 1e9
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        torch.matmul(a, b)
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1) * 1e-3)
    return 2 * nn ** 3 / best / 1e12


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
    m = int(sys.argv[2]) if len(sys.argv) > 2 else 32768
    peak = dgemm_peak()
    bpv = (n + 3) // 4
    bed = torch.empty(m * bpv, dtype=torch.uint8, device="cuda")
    for s in range(0, m, 32768):
        g.synth_bed_device(bed.data_ptr() + s * bpv, n, s, min(32768, m - s), 5, 0.01)
    torch.cuda.synchronize()
    acc = g.Accumulator(n, m, method="SH", precision=g.PREC_FP64)
    prev = acc.timers()
    out = []
    for p in range(3):
        if p:
            acc.reset()
        acc.push_device(bed.data_ptr(), 0, m)
        acc.finish()
        t = acc.timers()
        d = {k: t[k] - prev[k] for k in t}
        prev = t
        out.append(d)
    best = min(out, key=lambda d: d["pass_ms"])
    fl = n * (n + 1) * m
    print(json.dumps({"n": n, "m": m, "dgemm_tflops_torch_8192": peak, "fp64_pass_ms": best["pass_ms"], "fp64_syrk_ms": best["syrk_ms"],
                      "fp64_effective_tflops": fl / best["pass_ms"] / 1e9, "fp64_syrk_tflops": fl / best["syrk_ms"] / 1e9,
                      "frac_of_dgemm": fl / best["syrk_ms"] / 1e9 / peak, "expand_ms": best["expand_ms"], "prep_ms": best["prep_ms"]}))
    acc.close()


if __name__ == "__main__":
    main()
