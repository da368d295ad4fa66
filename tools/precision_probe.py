"""Developer probe: 16-bit vs FP64 error at m = 500 000 (n = 4096) for several slab lengths."""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import gelex_b200 as g

def norm(p):
    n = int((np.sqrt(8 * p.size + 1) - 1) / 2)
    d = np.cumsum(np.arange(1, n + 1)) - 1
    return p / (p[d].sum() / n), d

n, m = 4096, 500_000
bpv = n // 4
d_bed = torch.empty(m * bpv, dtype=torch.uint8, device="cuda")
for s in range(0, m, 65536):
    g.synth_bed_device(d_bed.data_ptr() + s * bpv, n, s, min(65536, m - s), 20261107, 0.01)
torch.cuda.synchronize()

def build(prec, chunk):
    with g.Accumulator(n, m, method=1, precision=prec, chunk_snps=chunk) as acc:
        acc.push_device(d_bed.data_ptr(), 0, m)
        acc.finish()
        p, tr = acc.fetch(0, g.GROUP_WHOLE, dtype=np.float64)
    return p, tr

ref, tr_ref = build(g.PREC_FP64, 8192)
rn, diag = norm(ref)
for chunk in (4096, 8192, 16384, 32768, 65536):
    for name, prec in (("fp16", g.PREC_FP16), ("bf16", g.PREC_BF16)):
        p, tr = build(prec, chunk)
        pn, _ = norm(p)
        e = np.abs(pn - rn)
        off = np.ones(e.size, bool); off[diag] = False
        print(f"K={chunk:6d} {name}: max|norm err| {e.max():.3e} (diag {e[diag].max():.3e}, offdiag {e[off].max():.3e})  trace rel {tr / tr_ref - 1:+.3e}  "
              f"raw diag rel err max {np.max(np.abs(p[diag] / ref[diag] - 1)):.3e}", flush=True)
