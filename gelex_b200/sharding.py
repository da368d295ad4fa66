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


def piece_owners(n: int, world: int, bands, rows_per_strip: int = 2048, target_elems: int = 16 << 20):
    """Band-pipelined reduce-scatter at the grain that keeps every band's tail short: each 2048-row strip (one super-row of
    2048 x 2048 super-tiles) is cut along its COLUMNS into pieces of `target_elems // (rows * 2048)` whole super-tile columns
    (about `target_elems` entries = 64 MB of FP32 each), and the pieces of every band are dealt over ALL ranks (largest first, to the rank with the least
    work in this band, ties to the least loaded overall).  With whole strips (strip_owners) the last band of a 50 000-sample
    matrix is two strips - 393 MB and 168 MB on two ranks - and their reduce + read-out is the exposed tail of the pass;
    with pieces it is ~70 MB on each of eight ranks.
    `bands` = [(row_begin, row_end)] in completion order (ggrm_band_layout).
    Returns per rank a list of (band, super_row, super_col_begin, super_col_end) in (band, super_row, column) order; the
    piece covers rows [super_row * 2048, min(n, (super_row + 1) * 2048)) x columns [super_col_begin * 2048, min(n, super_col_end * 2048)),
    lower triangle only."""
    def tri(x):
        return x * (x + 1) // 2

    pieces = []  # (band, si, c0, c1, area)
    for si, lo in enumerate(range(0, n, rows_per_strip)):
        hi = min(n, lo + rows_per_strip)
        b = next(i for i, (_, be) in enumerate(bands) if hi <= be)
        rows = hi - lo
        ncol = si + 1                                         # super-tile columns of this super-row
        col_area = [rows * rows_per_strip] * si + [tri(hi) - tri(lo) - rows * lo]  # left of the diagonal: full blocks; diagonal block: its triangle part
        w = max(1, target_elems // (rows * rows_per_strip))   # whole super-tile columns per piece (4 at 2048 rows and 64 MB)
        for c0 in range(0, ncol, w):
            c1 = min(ncol, c0 + w)
            pieces.append((b, si, c0, c1, sum(col_area[c0:c1])))
    total_load = [0] * world
    out = [[] for _ in range(world)]
    for b in range(len(bands)):
        band_load = [0] * world
        mine = sorted((p for p in pieces if p[0] == b), key=lambda p: -p[4])  # stable: equal areas keep (si, c0) order
        for p in mine:
            r = min(range(world), key=lambda q: (band_load[q], total_load[q], q))
            band_load[r] += p[4]
            total_load[r] += p[4]
            out[r].append(p[:4])
    for r in range(world):
        out[r].sort()
    return out
