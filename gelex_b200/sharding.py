"""Multi-GPU partition rules of the GRM path, shared by bench.py, the tests and mirrored by the C++ host
(gelex_b200/csrc/host/grm_engine.cpp: run_multi_gpu).  Pure integer logic, no device code.

  SNP-shard   contiguous SNP ranges cut on 64-SNP (one UMMA K block) boundaries; every rank holds a partial GRM over all
              tiles; one sum-reduce of the FP32 accumulators finishes the matrix.
  LOCO-shard  whole chromosome runs dealt to ranks by LPT on their SNP counts: each partial is produced by exactly one
              rank and needs no reduce.
  row-slice   after an SNP-shard pass: GRM rows are cut on 2048-row (8-tile super-row) boundaries into `world` ranges of
              about equal triangle area; rank r sum-reduces ITS rows from every peer's partial (reduce-scatter over NVLink
              peer memory) and fetches / writes only those rows, so read-out and file writing scale with the GPU count.
  tile-shard  (n >= ~100k) 8x8-tile super-tiles of the lower triangle dealt by LPT; every rank streams all SNPs but owns
              disjoint output tiles; no collective on the output (ggrm_config_t.tile_rank / tile_world).
"""
from __future__ import annotations


def snp_shards(m: int, world: int, align: int = 64):
    """[(begin, end)] per rank: contiguous, disjoint, covering [0, m), interior cuts multiples of `align`."""
    cuts = [min(m, ((m * r // world) + align - 1) // align * align) for r in range(world)] + [m]
    return [(cuts[r], max(cuts[r], cuts[r + 1])) for r in range(world)]


def lpt_assign(costs, world: int):
    """Longest-processing-time-first: returns owner[i] for each cost; ties resolved like the C++ host (stable sort, first least-loaded rank)."""
    order = sorted(range(len(costs)), key=lambda i: -costs[i])
    load = [0] * world
    owner = [0] * len(costs)
    for i in order:
        r = min(range(world), key=lambda k: load[k])
        owner[i] = r
        load[r] += costs[i]
    return owner


def chromosome_shards(groups, world: int):
    """groups: [(begin, end)] chromosome runs -> per-rank ascending lists of group indices."""
    owner = lpt_assign([e - b for b, e in groups], world)
    return [[g for g in range(len(groups)) if owner[g] == r] for r in range(world)]


def supertile_owners(n: int, world: int, tile: int = 256, st: int = 8):
    """Owner rank of each (si, sj<=si) super-tile, in row-major super-tile order — the rule ggrm_create applies."""
    t = (n + tile - 1) // tile
    nst = (t + st - 1) // st
    supers, costs = [], []
    for si in range(nst):
        for sj in range(si + 1):
            c = sum(1 for i in range(si * st, min(t, (si + 1) * st)) for j in range(sj * st, min(t, (sj + 1) * st)) if j <= i)
            supers.append((si, sj))
            costs.append(c)
    owner = lpt_assign(costs, world) if world > 1 else [0] * len(costs)
    return supers, costs, owner


def row_slices(n: int, world: int, tile: int = 256, st: int = 8):
    """[(row_begin, row_end)] per rank: contiguous, covering [0, n), cuts on multiples of tile*st rows, balanced on the
    number of lower-triangle tiles (rank r gets the super-rows whose cumulative tile count first reaches (r+1)/world)."""
    t = (n + tile - 1) // tile
    nsr = (t + st - 1) // st
    cum = [min(t, (s + 1) * st) * (min(t, (s + 1) * st) + 1) // 2 for s in range(nsr)]  # tiles in super-rows 0..s
    total = cum[-1]
    cuts, s = [0], 0
    for r in range(1, world):
        target = total * r / world
        while s < nsr and cum[s] < target:
            s += 1
        # cut after super-row s or before it, whichever lands closer to the target
        lo = cum[s - 1] if s > 0 else 0
        c = s + 1 if s < nsr and (cum[s] - target) <= (target - lo) else s
        cuts.append(max(cuts[-1], min(nsr, c)))
    cuts.append(nsr)
    return [(min(n, cuts[r] * tile * st), min(n, cuts[r + 1] * tile * st)) for r in range(world)]


def strip_owners(n: int, world: int, bands, rows_per_strip: int = 2048):
    """Band-pipelined reduce-scatter: who reduces and reads out which 2048-row strip.
    `bands` = [(row_begin, row_end)] in completion order (ggrm_band_layout).  Strips are dealt by LPT on their triangle area
    (largest first, to the least loaded rank), so every rank moves about the same bytes and the strips of the LAST band -
    the only ones whose reduce + read-out cannot hide behind the accumulation - land on different ranks.
    Returns per rank a list of (band_index, row_begin, row_end) in row order."""
    strips = [(lo, min(n, lo + rows_per_strip)) for lo in range(0, n, rows_per_strip)]
    area = [hi * (hi + 1) // 2 - lo * (lo + 1) // 2 for lo, hi in strips]
    owner = lpt_assign(area, world) if world > 1 else [0] * len(strips)
    out = [[] for _ in range(world)]
    for (lo, hi), r in zip(strips, owner):
        b = next(i for i, (_, be) in enumerate(bands) if hi <= be)
        out[r].append((b, lo, hi))
    return out
