"""Multi-GPU partition logic on CPU: integer rules, and a world_size-2 gloo run that shards SNPs across two processes,
accumulates each shard with the CPU oracle and sum-reduces — the plumbing bench.py uses with NCCL on real GPUs."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from gelex_b200 import sharding


def test_snp_shards_cover_and_align():
    for m, w in ((500000, 8), (500000, 3), (131, 4), (63, 2), (64, 2), (1, 8), (600000, 7)):
        sh = sharding.snp_shards(m, w)
        assert len(sh) == w and sh[0][0] == 0 and sh[-1][1] == m
        for (b0, e0), (b1, e1) in zip(sh, sh[1:]):
            assert e0 == b1 and b0 <= e0
        assert all(b % 64 == 0 or b == m for b, _ in sh)


def test_chromosome_lpt_is_balanced_and_complete():
    sizes = [41000, 40000, 33000, 31000, 30000, 28000, 26000, 24000, 20000, 22000, 22000, 22000, 16000, 15000, 14000, 15000, 14000, 13000, 10000, 11000, 6500, 6500]
    groups, s = [], 0
    for z in sizes:
        groups.append((s, s + z))
        s += z
    for w in (2, 4, 8):
        sh = sharding.chromosome_shards(groups, w)
        assert sorted(g for r in sh for g in r) == list(range(22))
        loads = [sum(sizes[g] for g in r) for r in sh]
        assert max(loads) <= 1.12 * sum(sizes) / w
        assert all(r == sorted(r) for r in sh)


def test_supertile_ownership_partitions_the_triangle():
    for n, w in ((200000, 8), (50000, 4), (2300, 3)):
        supers, costs, owner = sharding.supertile_owners(n, w)
        t = (n + 255) // 256
        assert sum(costs) == t * (t + 1) // 2
        loads = [sum(c for c, o in zip(costs, owner) if o == r) for r in range(w)]
        assert max(loads) - min(loads) <= 64


def _worker(rank, world, port, tmpdir):
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__))))
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import gelex_b200 as g
    import oracle_lib as o

    n, m = 96, 400
    bed = g.synth_bed_host(n, 0, m, 5, 0.02)
    b, e = sharding.snp_shards(m, world)[rank]
    part, _, _ = o.grm_compute(bed, n, 1, 0, ranges=[(b, e)] if e > b else [(0, 1)])
    if e <= b:
        part = np.zeros_like(part)
    t = torch.from_numpy(np.ascontiguousarray(part))
    dist.reduce(t, dst=0, op=dist.ReduceOp.SUM)
    if rank == 0:
        whole, _, _ = o.grm_compute(bed, n, 1, 0)
        np.save(os.path.join(tmpdir, "err.npy"), np.array([np.max(np.abs(t.numpy() - whole))]))
    dist.barrier()
    dist.destroy_process_group()


def test_snp_shard_reduce_world2_gloo(tmp_path):
    port = 29500 + (os.getpid() % 2000)
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    assert np.load(tmp_path / "err.npy")[0] < 1e-9


def test_row_slices_cover_align_and_balance():
    for n in (1000, 4500, 50000, 200000):
        for w in (1, 2, 3, 4, 8):
            sl = sharding.row_slices(n, w)
            assert len(sl) == w and sl[0][0] == 0 and sl[-1][1] == n
            assert all(sl[r][1] == sl[r + 1][0] for r in range(w - 1))
            assert all((b % 2048 == 0 or b == n) and (e % 2048 == 0 or e == n) for b, e in sl)
    areas = [(e * (e + 1) - b * (b + 1)) / 2 for b, e in sharding.row_slices(50000, 8)]
    assert max(areas) <= 1.25 * sum(areas) / 8   # coarse (25 super-rows over 8 ranks) but bounded
    areas = [(e * (e + 1) - b * (b + 1)) / 2 for b, e in sharding.row_slices(200000, 8)]
    assert max(areas) <= 1.10 * sum(areas) / 8


def test_cpp_engine_uses_the_same_row_slice_rule():
    from gelex_b200 import engine

    for n in (1, 2047, 2048, 2049, 4500, 50000, 123457, 200000):
        for w in (1, 2, 3, 4, 7, 8):
            assert engine.row_slices(n, w) == sharding.row_slices(n, w), (n, w)


def test_cpp_engine_deals_strips_like_the_python_rule():
    # the band-pipelined SNP-shard run of the C++ engine (host/grm_engine.cpp::strip_owners) and bench.py must agree on who owns what
    from gelex_b200 import engine

    for n in (1, 2048, 2049, 4500, 50000, 123457):
        strips_total = (n + 2047) // 2048
        for nb in (1, 2, 3, 8):
            per = max(1, strips_total // nb)
            ends = sorted({min(n, (i + 1) * per * 2048) for i in range(nb - 1)} | {n})
            bands = [(0 if i == 0 else ends[i - 1], e) for i, e in enumerate(ends)]
            for w in (1, 2, 3, 4, 8):
                assert engine.strip_owners(n, w, bands) == sharding.strip_owners(n, w, bands), (n, nb, w)


def test_strip_owners_cover_balance_and_spread_the_last_band():
    for n, nb in ((4500, 3), (50000, 8), (200000, 8)):
        strips_total = (n + 2047) // 2048
        # bands as the library cuts them: whole strips, about equal tile counts (here: equal strip-area shares are close enough)
        area = [min(n, (s + 1) * 2048) * (min(n, (s + 1) * 2048) + 1) // 2 - (s * 2048) * (s * 2048 + 1) // 2 for s in range(strips_total)]
        tot, acc, ends = sum(area), 0, []
        for s_, a in enumerate(area):
            acc += a
            if len(ends) + 1 < nb and acc >= (len(ends) + 1) * tot / nb or s_ + 1 == strips_total:
                ends.append(min(n, (s_ + 1) * 2048))
        bands = [(0 if i == 0 else ends[i - 1], e) for i, e in enumerate(ends)]
        for w in (1, 2, 3, 4, 8):
            own = sharding.strip_owners(n, w, bands)
            flat = sorted((lo, hi) for r in own for _, lo, hi in r)
            assert flat == [(s * 2048, min(n, (s + 1) * 2048)) for s in range(strips_total)]       # every strip exactly once
            assert all(bands[b][0] <= lo and hi <= bands[b][1] for r in own for b, lo, hi in r)    # tagged with its band
            assert all(r == sorted(r, key=lambda t: t[1]) for r in own)                            # row (= completion) order
            loads = [sum(hi * (hi + 1) // 2 - lo * (lo + 1) // 2 for _, lo, hi in r) for r in own]
            if strips_total >= 3 * w:
                assert max(loads) <= 1.15 * sum(loads) / w, (n, w, loads)
            last = [r_i for r_i, r in enumerate(own) for b, _, _ in r if b == len(bands) - 1]
            # the last band's strips (the exposed tail) are spread over the ranks, not piled on one (the small ragged strip may double up)
            assert max(last.count(r_i) for r_i in set(last)) <= -(-len(last) // w) + 1, (n, w, last)
            if w > 1 and len(last) > 1:
                assert len(set(last)) >= min(len(last) - 1, w), (n, w, last)


def _bands_for(n, nb):
    strips_total = (n + 2047) // 2048
    area = [min(n, (s + 1) * 2048) * (min(n, (s + 1) * 2048) + 1) // 2 - (s * 2048) * (s * 2048 + 1) // 2 for s in range(strips_total)]
    tot, acc, ends = sum(area), 0, []
    for s_, a in enumerate(area):
        acc += a
        if len(ends) + 1 < nb and acc >= (len(ends) + 1) * tot / nb or s_ + 1 == strips_total:
            ends.append(min(n, (s_ + 1) * 2048))
    return [(0 if i == 0 else ends[i - 1], e) for i, e in enumerate(ends)]


def test_piece_owners_cover_every_supertile_once_and_spread_every_band():
    # the finer dealing of the band-pipelined reduce-scatter: column pieces of strips, every band over all ranks
    for n, nb in ((4500, 3), (50000, 8), (200000, 8), (2049, 1)):
        ns = (n + 2047) // 2048
        bands = _bands_for(n, nb)
        for w in (1, 2, 3, 4, 8):
            own = sharding.piece_owners(n, w, bands)
            seen = {}
            for r, lst in enumerate(own):
                assert lst == sorted(lst)
                for b, si, c0, c1 in lst:
                    assert 0 <= c0 < c1 <= si + 1 and bands[b][0] <= si * 2048 and min(n, (si + 1) * 2048) <= bands[b][1]
                    for c in range(c0, c1):
                        assert (si, c) not in seen
                        seen[(si, c)] = r
            assert len(seen) == ns * (ns + 1) // 2
            if n == 50000 and w == 8:
                # the exposed tail: bytes of the LAST band per rank (strip_owners put 393 MB on one rank)
                def area(si, c):
                    lo, hi = si * 2048, min(n, (si + 1) * 2048)
                    return (hi - lo) * 2048 if c < si else hi * (hi + 1) // 2 - lo * (lo + 1) // 2 - (hi - lo) * lo
                last = [0] * w
                for r, lst in enumerate(own):
                    for b, si, c0, c1 in lst:
                        if b == len(bands) - 1:
                            last[r] += sum(area(si, c) for c in range(c0, c1)) * 4
                assert max(last) < 120e6 and min(last) > 40e6, last


def test_cpp_engine_deals_pieces_like_the_python_rule():
    from gelex_b200 import engine

    for n in (1, 2048, 2049, 4500, 50000, 123457):
        for nb in (1, 3, 8):
            bands = _bands_for(n, nb)
            for w in (1, 2, 3, 8):
                assert engine.piece_owners(n, w, bands) == sharding.piece_owners(n, w, bands), (n, nb, w)


def test_cli_gpu_count_rule_on_the_named_configs():
    # --gpus is an upper bound in the CLI: the GPU count that minimises context creation (0.3 s each, serialised) + device time +
    # the fixed cost of a multi-GPU read-out; write-bound LOCO runs take a second GPU for its PCIe link and writer pool
    from gelex_b200 import _lib, engine

    assert engine.gpus_worth_using(2_000, 50_000) == 1                                   # C1
    assert engine.gpus_worth_using(20_000, 300_000, kinds=2) == 1                        # C3: 0.18 s of device time
    assert engine.gpus_worth_using(50_000, 500_000) == 1                                 # C2: measured 2.0 s on one GPU, 3.0 on two, 5.4 on eight
    assert engine.gpus_worth_using(50_000, 600_000, n_outputs=22) == 2                   # C4: measured 17.7 / 14.9 / 15.1 s on 1 / 2 / 8
    assert engine.gpus_worth_using(200_000, 600_000) == 8                                # C5: 18 s of device time on one GPU
    assert engine.gpus_worth_using(200_000, 600_000, offered=4) == 4
    assert engine.gpus_worth_using(50_000, 500_000, offered=1) == 1
    assert engine.gpus_worth_using(20_000, 300_000, precision=_lib.PREC_FP64) == 4       # the DMMA path is 40x slower (4 s on one GPU)
    assert engine.gpus_worth_using(400_000, 1_000, offered=8) >= 3                       # 320 GB of FP32 triangle: never fewer GPUs than hold it
