"""BASELINE.json's headline shape (C2: 50 000 samples x 500 000 SNPs) at FULL size, checked through properties that do not need
a CPU answer (the oracle would need hours here):

  1. diagonal: StandardizeHWE makes every sample's diagonal ~ the number of non-missing SNPs (SURVEY 8c smoke property)
  2. linearity in SNPs: G(all SNPs) = G(first half) + G(second half) - the SNP-shard decomposition, summed by the peer-reduce
     kernel - and the banded ggrm_finish_async read-out returns the same bits as a blocking finish
  3. centring: every SNP column of Z sums to ~0 over samples, so every full row of the symmetric matrix sums to ~0
  4. agreement of the two 16-bit operand types with each other on G/(trace/n) (each is within 1e-4 of FP64 at this SNP count:
     tests/test_gpu_parity.py::test_c2_snp_count_tolerance_16bit_device_resident)
"""
import numpy as np
import pytest

import gelex_b200 as g

pytestmark = pytest.mark.gpu

N, M, SEED = 50_000, 500_000, 20261106


def _norm(rows, tr, n):
    return rows.astype(np.float64) / (tr / n)


def test_c2_full_size_properties():
    import torch

    if torch.cuda.get_device_properties(0).total_memory < 60e9:
        pytest.skip("needs a 60 GB+ device")
    n, m = N, M
    bpv = (n + 3) // 4
    d_bed = torch.empty(m * bpv, dtype=torch.uint8, device="cuda")
    for s in range(0, m, 32768):
        g.synth_bed_device(d_bed.data_ptr() + s * bpv, n, s, min(32768, m - s), SEED, 0.01)
    torch.cuda.synchronize()
    diag = np.cumsum(np.arange(1, n + 1, dtype=np.int64)) - 1

    def build(prec, lo, hi, async_finish=False):
        acc = g.Accumulator(n, m, method="SH", precision=prec)
        acc.push_device(d_bed.data_ptr() + lo * bpv, lo, hi - lo)
        (acc.finish_async if async_finish else acc.finish)()
        return acc

    # ---- whole matrix, bf16 operands, read out through the banded async finish
    whole = build(g.PREC_BF16, 0, m, async_finish=True)
    full, tr = whole.fetch(g.KIND_ADD, g.GROUP_WHOLE)
    assert full.size == g.tri(n) and np.isfinite(tr)
    d = full[diag].astype(np.float64)
    assert abs(d.sum() - tr) <= 1e-6 * tr
    # 1. diagonal ~ non-missing SNP count (1 % missing): mean within 0.5 %, every sample within 3 %
    assert abs(d.mean() / (0.99 * m) - 1.0) < 5e-3
    assert np.all(np.abs(d / (0.99 * m) - 1.0) < 3e-2)
    # off-diagonal entries of unrelated samples are O(sqrt(m)), not O(m)
    sample = full[g.tri(40_000):g.tri(40_000) + 40_000].astype(np.float64)   # row 40 000, columns < 40 000
    assert np.abs(sample).max() < 8 * np.sqrt(m)
    whole.finish()

    # 2a. blocking finish returns the same bits for a few row ranges crossing band boundaries
    again = build(g.PREC_BF16, 0, m)
    for r0, cnt in ((0, 300), (20_400, 200), (49_900, 100)):
        part, _ = again.fetch(g.KIND_ADD, g.GROUP_WHOLE, r0, cnt)
        assert np.array_equal(part, full[g.tri(r0):g.tri(r0 + cnt)])

    # 3. centring on a spread of rows: row i of the symmetric matrix = packed row i (columns <= i) + column i of the rows below
    for i in (0, 12_345, 25_000, 49_999):
        left = full[g.tri(i):g.tri(i + 1)].astype(np.float64)
        below = again.fetch_block(g.KIND_ADD, g.GROUP_WHOLE, i + 1, n - i - 1, i, 1).astype(np.float64).ravel() if i + 1 < n else np.zeros(0)
        total, mass = left.sum() + below.sum(), np.abs(left).sum() + np.abs(below).sum()
        assert abs(total) <= 2e-2 * mass, (i, total, mass)
    again.close()

    # 2b. linearity: two SNP shards summed by the peer-reduce kernel == the whole, entrywise on G/(trace/n)
    half = (m // 2 + 63) // 64 * 64
    a, b = build(g.PREC_BF16, 0, half), build(g.PREC_BF16, half, m)
    pb, nb, _ = b.accum_buffer(g.KIND_ADD, g.GROUP_WHOLE)
    a.reduce_peer(g.KIND_ADD, g.GROUP_WHOLE, [pb], 0, nb)
    b.close()
    _, tr_sum = a.fetch(g.KIND_ADD, g.GROUP_WHOLE, 0, n, out=np.empty(g.tri(n), dtype=np.float32))
    # raw sums carry the FP32-accumulate truncation bias of their slab length (-5.4e-9 x slab SNPs, uniform: DESIGN section 5), so
    # each matrix is normalised by its own trace, as every consumer of the .bin does
    assert abs(tr_sum / tr - 1.0) < 2e-5
    for r0, cnt in ((0, 2048), (30_000, 256), (49_744, 256)):
        part, _ = a.fetch(g.KIND_ADD, g.GROUP_WHOLE, r0, cnt)
        ref = full[g.tri(r0):g.tri(r0 + cnt)]
        assert np.max(np.abs(_norm(part, tr_sum, n) - _norm(ref, tr, n))) <= 5e-6
    a.close()

    # 4. fp16 vs bf16 operands on the normalised matrix
    h16 = build(g.PREC_FP16, 0, m)
    _, tr16 = h16.fetch(g.KIND_ADD, g.GROUP_WHOLE, 0, 1)  # warm the path; the trace of all rows comes from the loop below
    worst, tr16_all = 0.0, 0.0
    parts = []
    for r0 in range(0, n, 10_000):
        part, t = h16.fetch(g.KIND_ADD, g.GROUP_WHOLE, r0, min(10_000, n - r0))
        parts.append((r0, part))
        tr16_all += t
    for r0, part in parts:
        cnt = min(10_000, n - r0)
        worst = max(worst, float(np.max(np.abs(_norm(part, tr16_all, n) - _norm(full[g.tri(r0):g.tri(r0 + cnt)], tr, n)))))
    assert worst <= 2e-4, worst   # each is <= 1e-4 from FP64
    h16.close()
    whole.close()
