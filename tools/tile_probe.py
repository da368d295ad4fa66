"""C5 (200 000 samples x 600 000 SNPs) in 2D-tile mode on N GPUs of one box, BED resident in every GPU's HBM:
each rank owns the super-tiles ggrm deals it (tile_rank / tile_world), streams ALL SNPs, no collective.  One JSON line.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 tools/tile_probe.py [n m]
"""
import json
import os
import sys
import time

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import gelex_b200 as g  # noqa: E402


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 200_000
    m = int(sys.argv[2]) if len(sys.argv) > 2 else 600_000
    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    bpv = (n + 3) // 4
    bed = torch.empty(m * bpv, dtype=torch.uint8, device="cuda")
    for s in range(0, m, 32768):
        g.synth_bed_device(bed.data_ptr() + s * bpv, n, s, min(32768, m - s), 20261109, 0.01, device=local)
    torch.cuda.synchronize()
    acc = g.Accumulator(n, m, method="SH", precision=g.PREC_BF16, device=local, tile_rank=rank, tile_world=world)
    owned = len(acc.owned_supertiles())
    best = None
    prev = acc.timers()
    for p in range(2):
        if p:
            acc.reset()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        t0 = time.time()
        acc.push_device(bed.data_ptr(), 0, m)
        acc.finish()
        wall = time.time() - t0
        t = acc.timers()
        d = {k: t[k] - prev[k] for k in t}
        prev = t
        tt = torch.tensor([d["pass_ms"], wall * 1e3, d["syrk_ms"]], device="cuda", dtype=torch.float64)
        if world > 1:
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        rec = {"pass_ms_max": float(tt[0]), "wall_ms_max": float(tt[1]), "syrk_ms_max": float(tt[2])}
        if best is None or rec["pass_ms_max"] < best["pass_ms_max"]:
            best = rec
    # spot check on the rank that owns super-tile (0, 0): StandardizeHWE diagonal ~ non-missing SNP count
    chk = None
    if (0, 0) in acc.owned_supertiles():
        blk = acc.fetch_block(g.KIND_ADD, g.GROUP_WHOLE, 0, 512, 0, 512)
        chk = float(blk.diagonal().astype("float64").mean() / m)
    owned_all = [None] * world
    chk_all = [None] * world
    if world > 1:
        dist.all_gather_object(owned_all, owned)
        dist.all_gather_object(chk_all, chk)
    else:
        owned_all, chk_all = [owned], [chk]
    if rank == 0:
        fl = n * (n + 1) * m
        print(json.dumps({"config": "C5 2D tiles, HBM-resident BED on every GPU", "n": n, "m": m, "n_gpus": world, **best,
                          "effective_tflops": fl / best["pass_ms_max"] / 1e9, "per_gpu_tflops": fl / best["pass_ms_max"] / 1e9 / world,
                          "supertiles_per_rank": owned_all, "diag_mean_over_m": [c for c in chk_all if c is not None]}), flush=True)
    acc.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
