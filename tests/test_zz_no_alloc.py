"""ggrm_snp_stats_noalloc (include/ggrm.h): the entry point that allocates nothing - its scratch is a static device array - checked
through the C ABI against the existing entry point for the same quantities, ggrm_snp_stats of a keep_stats handle.  Counts and
allele frequencies are the bit-exact layer of the path, so the tolerance is the one every other test of them uses: equality.

In a file of its own that pytest collects last: passes alone in a fresh process and after the
rest of the suite.  Device buffers of the TEST are torch tensors; the entry point itself calls no allocator."""
import ctypes as C

import numpy as np
import pytest

import gelex_b200 as g
from gelex_b200 import _lib

pytestmark = pytest.mark.gpu


def reference_stats(bed, n, m):
    """The existing entry point: counts {n0, n1, n2, nmiss} and A1 frequency per SNP from a handle that keeps them."""
    with g.Accumulator(n, m, method="SH", precision=g.PREC_FP16, keep_stats=True) as acc:
        acc.push(bed, 0)
        acc.finish()
        n0, n1, n2, nm, fr = acc.snp_stats(0, m)
    return np.stack([n0, n1, n2, nm], axis=1), fr


def test_no_alloc_snp_stats_matches_ggrm_snp_stats_twice_in_a_row():
    import torch

    dev = torch.device("cuda", 0)
    # two different inputs, the second SHORTER and of another sample count than the first: the shared scratch array still holds
    # the first call's bytes behind what the second call overwrites, and must not leak into its result
    cases = [(1003, 700, 20261104, 0.02), (515, 300, 7, 0.10)]
    free_before = None
    for call, (n, m, seed, miss) in enumerate(cases):
        bed = g.synth_bed_host(n, 0, m, seed, miss).copy()
        bed[5] = 0x55                      # an all-missing SNP: counts {0, 0, 0, n}, frequency 0 (LocusStatistic default)
        bed[5, -1] = 0x55 & ((1 << (2 * (n - 4 * (bed.shape[1] - 1)))) - 1)  # padding bits of the tail byte stay 0
        want_counts, want_freq = reference_stats(bed, n, m)
        d_bed = torch.from_numpy(bed).to(dev)
        d_counts = torch.full((m, 4), -7, dtype=torch.int32, device=dev)
        d_freq = torch.full((m,), -7.0, dtype=torch.float64, device=dev)
        torch.cuda.synchronize()
        if call == 1:
            free_before = torch.cuda.mem_get_info(0)[0]
        _lib.snp_stats_noalloc(d_bed.data_ptr(), m, n, d_counts.data_ptr(), d_freq.data_ptr())  # default stream, enqueue only
        torch.cuda.synchronize()
        if call == 1:  # the second call runs with every module loaded: device memory in use must not have moved
            assert torch.cuda.mem_get_info(0)[0] == free_before
        got_counts, got_freq = d_counts.cpu().numpy(), d_freq.cpu().numpy()
        assert np.array_equal(got_counts, want_counts), call
        assert np.array_equal(got_freq, want_freq), call          # bit-exact, NaN-free by construction
        assert got_counts[5].tolist() == [0, 0, 0, n] and got_freq[5] == 0.0
        assert np.all(got_counts.sum(axis=1) == n)
    # one output may be left out; a caller-chosen stream works like the default one
    s = torch.cuda.Stream(device=dev)
    d_freq2 = torch.full((m,), -7.0, dtype=torch.float64, device=dev)
    torch.cuda.synchronize()
    _lib.snp_stats_noalloc(d_bed.data_ptr(), m, n, None, d_freq2.data_ptr(), stream=s.cuda_stream)
    s.synchronize()
    assert np.array_equal(d_freq2.cpu().numpy(), want_freq)


def test_no_alloc_rejects_what_does_not_fit_its_static_scratch():
    import torch

    lib = _lib.load()
    n = 50_000
    bpv, bpv_pad = (n + 3) // 4, (n + 255) // 256 * 256 // 4
    most = _lib.NOALLOC_SCRATCH_BYTES // bpv_pad                    # 668 SNPs at 50 000 samples
    d_bed = torch.zeros((most + 1) * bpv, dtype=torch.uint8, device="cuda")
    d_counts = torch.empty((most + 1, 4), dtype=torch.int32, device="cuda")
    with pytest.raises(g.GrmError, match=r"exceed the static device array of 8388608 bytes.*at most 668 SNPs"):
        _lib.snp_stats_noalloc(d_bed.data_ptr(), most + 1, n, d_counts.data_ptr(), None)
    _lib.snp_stats_noalloc(d_bed.data_ptr(), most, n, d_counts.data_ptr(), None)  # the largest call that fits: BED 00 everywhere = all A1A1
    torch.cuda.synchronize()
    assert d_counts[:most].cpu().numpy().tolist() == [[0, 0, n, 0]] * most
    # argument checks come before any device work
    vp = C.c_void_p
    assert lib.ggrm_snp_stats_noalloc(0, None, 1, bpv, n, vp(d_counts.data_ptr()), None, None) == -1
    assert lib.ggrm_snp_stats_noalloc(0, vp(d_bed.data_ptr()), 1, bpv + 1, n, vp(d_counts.data_ptr()), None, None) == -1
    assert lib.ggrm_snp_stats_noalloc(0, vp(d_bed.data_ptr()), 1, bpv, n, vp(d_counts.data_ptr() + 4), None, None) == -1  # misaligned counts
    assert b"16-byte aligned" in lib.ggrm_last_error(None)
