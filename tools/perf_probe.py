"""Developer perf probe (run under gpurun): times full passes of the device-resident path and prints TFLOP/s.
usage: python tools/perf_probe.py n m [precision 0|1|2] [chunk_snps] [passes]"""
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import gelex_b200 as g  # noqa: E402


def main():
    n = int(sys.argv[1])
    m = int(sys.argv[2])
    prec = int(sys.argv[3]) if len(sys.argv) > 3 else 2
    chunk = int(sys.argv[4]) if len(sys.argv) > 4 else 0
    passes = int(sys.argv[5]) if len(sys.argv) > 5 else 2
    bpv = (n + 3) // 4
    bed = torch.empty(m * bpv, dtype=torch.uint8, device="cuda")
    t0 = time.time()
    step = 65536
    for s in range(0, m, step):
        c = min(step, m - s)
        g.synth_bed_device(bed.data_ptr() + s * bpv, n, s, c, 20261104, 0.01)
    torch.cuda.synchronize()
    print(f"synth {m}x{n}: {time.time() - t0:.2f}s", flush=True)
    acc = g.Accumulator(n, m, method=1, precision=prec, chunk_snps=chunk)
    flops = n * (n + 1) * m
    prev = acc.timers()
    for p in range(passes):
        if p:
            acc.reset()
        t0 = time.time()
        acc.push_device(bed.data_ptr(), 0, m)
        acc.finish()
        wall = time.time() - t0
        t = acc.timers()
        d = {k: t[k] - prev[k] for k in t}
        prev = t
        print(f"pass {p}: wall {wall:.3f}s pass_ms {d['pass_ms']:.1f} -> {flops / d['pass_ms'] / 1e9:.1f} TFLOP/s | syrk_ms {d['syrk_ms']:.1f} "
              f"({flops / max(d['syrk_ms'], 1e-9) / 1e9:.1f} TFLOP/s, {int(d['syrk_launches'])} launches) prep {d['prep_ms']:.2f} expand {d['expand_ms']:.2f} "
              f"h2d/d2d {d['h2d_ms']:.2f}", flush=True)
    t0 = time.time()
    out, tr = acc.fetch(0, -1, 0, min(n, 4096))
    print(f"fetch 4096 rows {time.time() - t0:.3f}s trace/n part {tr / min(n, 4096):.3f}", flush=True)
    acc.close()


if __name__ == "__main__":
    main()
