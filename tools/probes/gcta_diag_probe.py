"""Cost of the GCTA default-diagonal option (second_matrix = 2) against the plain pair-count mode (1): device pass and prep phase
of one GPU at a given shape, BED resident in HBM (generated on the device), pushed 65 536 SNPs at a time.  One JSON line per mode."""
import json
import sys

import torch

sys.path.insert(0, ".")
import gelex_b200 as g  # noqa: E402

n, m = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (20000, 30000)
bpv = (n + 3) // 4
bed = torch.empty(m * bpv, dtype=torch.uint8, device="cuda")
for s in range(0, m, 65536):
    g.synth_bed_device(bed.data_ptr() + s * bpv, n, s, min(65536, m - s), 20261104 + 9, 0.01)
torch.cuda.synchronize()


def one_pass(acc):
    for s in range(0, m, 65536):
        acc.push_device(bed.data_ptr() + s * bpv, s, min(65536, m - s))
    acc.finish()


for second in (1, 2, 1, 2):
    with g.Accumulator(n, m, method="SH", want_add=True, want_dom=True, precision=g.PREC_FP16, second_matrix=second) as acc:
        one_pass(acc)
        t0 = acc.timers()
        acc.reset()
        one_pass(acc)
        t1 = acc.timers()
        d = float(acc.fetch_gcta(0, row_begin=n - 1, row_count=1)[-1])
    print(json.dumps({"n": n, "m": m, "second_matrix": second, "pass_ms": round(t1["pass_ms"] - t0["pass_ms"], 3),
                      "prep_ms": round(t1["prep_ms"] - t0["prep_ms"], 3), "syrk_ms": round(t1["syrk_ms"] - t0["syrk_ms"], 3), "last_diag": d}), flush=True)
