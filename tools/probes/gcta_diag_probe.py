"""Cost of the GCTA default-diagonal option (second_matrix = 2) against the plain pair-count mode (1): device pass and prep phase
of one GPU at a given shape, BED pushed from pinned host memory.  One JSON line per mode."""
import json
import sys

import numpy as np

sys.path.insert(0, ".")
import gelex_b200 as g  # noqa: E402

n, m = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (20000, 30000)
bed = g.synth_bed_host(n, 0, m, 20261104 + 9, 0.01)
for second in (1, 2, 1, 2):
    with g.Accumulator(n, m, method="SH", want_add=True, want_dom=True, precision=g.PREC_FP16, second_matrix=second) as acc:
        for rep in range(2):
            if rep:
                acc.reset()
            for s in range(0, m, 8192):
                acc.push(bed[s:s + 8192], s)
            acc.finish()
        t0 = acc.timers()
        acc.reset()
        for s in range(0, m, 8192):
            acc.push(bed[s:s + 8192], s)
        acc.finish()
        t1 = acc.timers()
        d = float(acc.fetch_gcta(0, row_begin=n - 1, row_count=1)[-1])
    print(json.dumps({"n": n, "m": m, "second_matrix": second, "pass_ms": round(t1["pass_ms"] - t0["pass_ms"], 3),
                      "prep_ms": round(t1["prep_ms"] - t0["prep_ms"], 3), "syrk_ms": round(t1["syrk_ms"] - t0["syrk_ms"], 3), "last_diag": d}))
