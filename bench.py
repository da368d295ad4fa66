#!/usr/bin/env python
"""bench.py — headline benchmark of the `gelex grm` hot path (BASELINE.json).

  python bench.py --gpus N --steps K --warmup W                 # this repo: B200 engine through the C ABI, config C2
  python bench.py --config c4 --gpus N ...                      # C4: 22 LOCO chromosome partials + whole, 50 000 x 600 000
  python bench.py --impl reference --gpus N --steps K ...       # the reference's own CPU path (oracle/_ref) on the host cores

Workloads (synthetic packed PLINK BED, p~U(0.05,0.5), HWE, 1 % missing, counter-based generator):
  c2 (default)  configs[1]: Yang additive GRM (StandardizeHWE + Add), 50 000 samples x 500 000 SNPs.
                N > 1: SNP shards ("strong" scaling), partial GRMs reduce-scattered over NVLink peer memory, band-pipelined.
  c4            configs[3]: 22 per-chromosome partial GRMs + the whole-genome GRM (= their sum), 50 000 x 600 000, one pass.
                N > 1: whole chromosomes dealt to ranks by LPT (no reduce for the partials), one reduce-scatter for the whole.
A "step" = one full build (decode -> counts / frequencies -> standardise -> SYRK over every SNP).
  value : effective TFLOP/s = n(n+1)m / t with the packed BED already resident in HBM (6-7.5 GB > L2: no flush needed)
  e2e   : the same through ggrm_push from pinned HOST memory (H2D inside the timed region) .. ggrm_fetch of every output matrix's
          float32 lower triangle (the .bin payload) into pinned host memory
  wall  : `gelex-b200 grm` on real files, process start -> files closed (SURVEY 8d t_wall), with its phase split
  cpu_baseline / --impl reference : oracle/_ref/gelex_ref = the UNMODIFIED reference (GrmEngine::compute, OpenBLAS dsyrk) on the host
          cores, on a FIXED bounded sample of the workload, plus one like-for-like pair on configs[0] (C1), wall seconds both arms.
One JSON line on stdout (rank 0)."""
import argparse
import ctypes
import json
import os
import shutil
import subprocess
import sys
import tempfile
import threading
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))

MISSING = 0.01
METHOD = 1  # StandardizeHWE ("Yang" additive)
REF_BIN = ROOT / "oracle" / "_ref" / "gelex_ref"
CLI_BIN = ROOT / "gelex_b200" / "gelex-b200"
CHR_MB = [248, 242, 198, 190, 182, 171, 159, 145, 138, 134, 135, 133, 114, 107, 102, 90, 83, 80, 59, 64, 47, 51]  # GRCh38 autosomes

CONFIGS = {
    "c2": dict(tag="configs[1]", n=int(os.environ.get("GGRM_BENCH_N", 50000)), m=int(os.environ.get("GGRM_BENCH_M", 500000)), loco=False, seed=20261104 + 2,
               metric="GRM build effective TFLOP/s (n(n+1)m FLOPs / wall-s), Yang additive GRM, 50k samples x 500k SNPs",
               what="Yang additive GRM (StandardizeHWE+Add)"),
    "c4": dict(tag="configs[3]", n=int(os.environ.get("GGRM_BENCH_N", 50000)), m=int(os.environ.get("GGRM_BENCH_M", 600000)), loco=True, seed=20261104 + 4,
               metric="GRM build effective TFLOP/s (n(n+1)m FLOPs / wall-s), 22 LOCO chromosome partial GRMs + whole-genome Yang additive GRM, 50k samples x 600k SNPs",
               what="22 per-chromosome partial GRMs + whole-genome GRM (StandardizeHWE+Add), one pass"),
}
C1 = dict(n=2000, m=50000, seed=20261104 + 1)       # configs[0]: the like-for-like pair
MID = dict(n=20000, m=30000, seed=20261104 + 6)     # a second like-for-like pair in the regime the engine is for: C3's sample count, a tenth
                                                    # of its SNPs, ~17 s on the reference's CPU path with 16 cores (not a BASELINE.json config)
CPU_SAMPLE = (8000, 30000)                          # fixed bounded sample of the workload for the CPU arm (about 2-3 s per step)


def flops(n, m):
    return float(n) * (n + 1.0) * m


def chromosome_runs(m):
    """22 contiguous runs, lengths proportional to the human autosomes."""
    cuts, acc = [0], 0
    for mb in CHR_MB:
        acc += mb
        cuts.append(round(m * acc / sum(CHR_MB)))
    cuts[-1] = m
    return [(cuts[i], cuts[i + 1]) for i in range(22)]


# ---------------------------------------------------------------------------------------------------------------- clocks
class ClockSampler:
    FIELDS = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
              "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index=0):
        self.rows = []
        self.proc = None
        self.gpu = gpu_index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.gpu}", f"--query-gpu={self.FIELDS}", "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), [x.strip() for x in line.split(",")]))

    def stop(self, t0=None, t1=None):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        rows = [r for t, r in self.rows if (t0 is None or t >= t0) and (t1 is None or t <= t1)] or [r for _, r in self.rows]
        sm, mx, pw, reasons = [], [], [], set()
        for r in rows:
            try:
                sm.append(float(r[0])); mx.append(float(r[1])); pw.append(float(r[2]))
            except (ValueError, IndexError):
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(pw) if pw else None, "samples": len(sm), "reasons": sorted(reasons)}


def bind_to_gpu_numa_node(local_rank):
    """Multi-rank runs: keep this rank's threads (and, by first touch, its pinned buffers) on the NUMA node its GPU hangs off.
    Best effort; returns a note for the JSON line."""
    try:
        bus = subprocess.run(["nvidia-smi", f"--id={local_rank}", "--query-gpu=pci.bus_id", "--format=csv,noheader"], capture_output=True,
                             text=True, timeout=20).stdout.strip().lower()
        if bus.startswith("00000000:"):
            bus = bus[4:]  # sysfs uses a 4-digit PCI domain
        node = int((Path("/sys/bus/pci/devices") / bus / "numa_node").read_text())
        if node < 0:
            return "numa: single node"
        cpus = set()
        for part in (Path(f"/sys/devices/system/node/node{node}") / "cpulist").read_text().strip().split(","):
            a, _, b = part.partition("-")
            cpus.update(range(int(a), int(b or a) + 1))
        cpus &= os.sched_getaffinity(0)
        if not cpus:
            return f"numa: node {node} has no allowed cpus"
        os.sched_setaffinity(0, cpus)
        return f"numa: rank bound to node {node} ({len(cpus)} cpus)"
    except Exception as e:  # noqa: BLE001
        return f"numa: not bound ({type(e).__name__})"


# ---------------------------------------------------------------------------------------------------------------- host facts
def logical_cpus():
    return len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)


def physical_cores():
    """Distinct cores among the CPUs this process may run on (SMT siblings counted once), from the kernel's topology files."""
    cpus = sorted(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else list(range(os.cpu_count() or 1))
    seen = set()
    for c in cpus:
        try:
            sib = (Path(f"/sys/devices/system/cpu/cpu{c}/topology") / "thread_siblings_list").read_text().strip()
        except OSError:
            sib = str(c)
        seen.add(sib)
    return max(1, len(seen))


def cpu_model():
    try:
        for line in Path("/proc/cpuinfo").read_text().splitlines():
            if line.startswith("model name"):
                return line.split(":", 1)[1].strip()
    except OSError:
        pass
    return "unknown"


def fs_of(path):
    try:
        out = subprocess.run(["stat", "-f", "-c", "%T", str(path)], capture_output=True, text=True, timeout=10).stdout.strip()
        return out or "unknown"
    except Exception:  # noqa: BLE001
        return "unknown"


# ---------------------------------------------------------------------------------------------------------------- PLINK triples
def write_triple(prefix, n, m, payload_rows, chrom_ranges):
    """payload_rows: iterable of uint8 arrays (row blocks of the packed BED).  FAM in sorted order (the reference's dense path)."""
    prefix = Path(prefix)
    prefix.parent.mkdir(parents=True, exist_ok=True)
    with open(str(prefix) + ".bed", "wb") as f:
        f.write(bytes([0x6C, 0x1B, 0x01]))
        import numpy as np

        for block in payload_rows:
            f.write(np.ascontiguousarray(block, dtype=np.uint8))
    with open(str(prefix) + ".fam", "w") as f:
        f.write("".join(f"F\ts{i:07d}\t0\t0\t1\t-9\n" for i in range(n)))
    with open(str(prefix) + ".bim", "w") as f:
        for ci, (b, e) in enumerate(chrom_ranges):
            f.write("".join(f"{ci + 1}\tsnp{j + 1}\t0\t{1000 + j}\tA\tG\n" for j in range(b, e)))


def synth_rows_numpy(n, m, seed):
    """CPU arm: the generator restated in numpy (tests/synth_np.py), so this arm never maps the product library."""
    import synth_np

    step = 4096
    for s in range(0, m, step):
        yield synth_np.synth_bed(n, s, min(step, m - s), seed, MISSING, threads=min(16, logical_cpus()))


def synth_rows_device(n, m, seed, device=0):
    """GPU arm: the same bytes from the device generator (csrc/synth.h), 1 GiB at a time."""
    import torch

    import gelex_b200 as g

    bpv = (n + 3) // 4
    step = max(1, (1 << 30) // bpv)
    d = torch.empty(min(step, m) * bpv, dtype=torch.uint8, device=f"cuda:{device}")
    for s in range(0, m, step):
        c = min(step, m - s)
        g.synth_bed_device(d.data_ptr(), n, s, c, seed, MISSING, device=device)
        yield d[: c * bpv].cpu().numpy()
    del d
    torch.cuda.empty_cache()


# ---------------------------------------------------------------------------------------------------------------- CPU reference arm
def run_ref(args_list):
    r = subprocess.run([str(REF_BIN)] + [str(a) for a in args_list], capture_output=True, text=True, check=True)
    return json.loads(r.stdout.strip().splitlines()[-1])


def run_port_once(prefix, n, threads):
    """Fallback when oracle/_ref is absent: the plain-C oracle port (kind "port")."""
    import numpy as np

    import oracle_lib as o

    payload = np.fromfile(str(prefix) + ".bed", dtype=np.uint8)[3:]
    os.environ["OMP_NUM_THREADS"] = str(threads)
    t0 = time.time()
    o.grm_compute(payload, n, METHOD, 0, chunk=2000)
    return {"wall_s": time.time() - t0, "blas": "none (plain C oracle port)"}


def cpu_leg(steps, warmup, cfg, payload_rows_fn):
    """Times the reference's own CPU implementation (oracle/_ref, else the oracle port) on the fixed bounded sample, with the
    stage split, one run at gelex's default -t hw/2 (apps/cli/grm/grm_args.cpp:54-59) and the C1 like-for-like run."""
    have_ref = REF_BIN.exists()
    kind = "reference" if have_ref else "port"
    phys, logical = physical_cores(), logical_cpus()
    n_s, m_s = CPU_SAMPLE if have_ref else (1500, 3000)
    tmp = Path(tempfile.mkdtemp(prefix="ggrm_cpu_"))
    out = {}
    try:
        write_triple(tmp / "sample", n_s, m_s, payload_rows_fn(n_s, m_s, cfg["seed"]), [(round(m_s * i / 22), round(m_s * (i + 1) / 22)) for i in range(22)])
        walls, info = [], {}
        for i in range(warmup + steps):
            info = run_ref(["grm", tmp / "sample", tmp / "out", METHOD, "A", 0, 10000, phys]) if have_ref else run_port_once(tmp / "sample", n_s, phys)
            if i >= warmup:
                walls.append(info["wall_s"])
        wall = sum(walls) / len(walls)
        out = {"value": flops(n_s, m_s) / wall / 1e12, "unit": "TFLOP/s", "cores": phys, "kind": kind,
               "sample": f"{n_s} samples x {m_s} SNPs of the same synthetic workload (fixed shape, same generator and seed), GrmEngine::compute incl. "
                         f".bin/.id write, chunk 10000, {phys} threads = physical cores",
               "wall_s": wall, "steps_timed": len(walls), "physical_cores": phys, "logical_cpus": logical, "cpu_model": cpu_model(),
               "blas": info.get("blas", ""), "same_config": False}
        if have_ref:
            st = run_ref(["stages", tmp / "sample", METHOD, "A", 10000, phys])
            out["stages_s"] = {k: st[k] for k in ("decode_s", "standardise_s", "dsyrk_s", "wall_s")}
            out["stages_s"]["write_and_setup_s"] = max(0.0, wall - st["wall_s"])
            half = max(1, logical // 2)  # gelex's own default: -t hardware_concurrency / 2
            hw = run_ref(["grm", tmp / "sample", tmp / "out_half", METHOD, "A", 0, 10000, half])
            out["gelex_default_threads"] = {"threads": half, "wall_s": hw["wall_s"], "value": flops(n_s, m_s) / hw["wall_s"] / 1e12}
            # like for like: configs[0] (C1) is small enough for both arms; wall seconds of the whole command
            write_triple(tmp / "c1", C1["n"], C1["m"], payload_rows_fn(C1["n"], C1["m"], C1["seed"]), chromosome_runs(C1["m"]))
            t0 = time.time()
            c1 = run_ref(["grm", tmp / "c1", tmp / "c1_ref", METHOD, "A", 0, 10000, phys])
            out["like_for_like"] = {"config": f"configs[0] C1: Yang additive GRM, {C1['n']} samples x {C1['m']} SNPs, real files in and out", "same_config": True,
                                    "reference_cli_wall_s": time.time() - t0, "reference_compute_wall_s": c1["wall_s"], "reference_threads": phys,
                                    "fs": fs_of(tmp)}
            out["_c1_prefix"] = str(tmp / "c1")
            write_triple(tmp / "mid", MID["n"], MID["m"], payload_rows_fn(MID["n"], MID["m"], MID["seed"]), chromosome_runs(MID["m"]))
            t0 = time.time()
            mid = run_ref(["grm", tmp / "mid", tmp / "mid_ref", METHOD, "A", 0, 10000, phys])
            out["like_for_like_mid"] = {"config": f"Yang additive GRM, {MID['n']} samples x {MID['m']} SNPs (C3's sample count, a tenth of its SNPs), real files in and out",
                                        "same_config": True, "reference_cli_wall_s": time.time() - t0, "reference_compute_wall_s": mid["wall_s"],
                                        "reference_threads": phys, "fs": fs_of(tmp)}
            out["_mid_prefix"] = str(tmp / "mid")
            for junk in tmp.glob("*_ref.*"):  # the reference's 1.6 GB output is not needed beside ours
                junk.unlink()
        return out, tmp
    except Exception:
        shutil.rmtree(tmp, ignore_errors=True)
        raise


def parse_cli_report(stdout):
    """The CLI's closing lines: '  Device      : accumulate A ms (SYRK S ms), read-out + file write W s, engine wall E s' and
    '  Host phases : name a-b s | ...'."""
    import re

    rep = {}
    for ln in stdout.splitlines():
        if "Device      :" in ln:
            mm = re.search(r"accumulate ([\d.]+) ms \(SYRK ([\d.]+) ms\), read-out \+ file write ([\d.]+) s, engine wall ([\d.]+) s", ln)
            if mm:
                rep.update(accumulate_s=float(mm.group(1)) / 1e3, syrk_s=float(mm.group(2)) / 1e3, write_s=float(mm.group(3)), engine_wall_s=float(mm.group(4)))
        elif "Engine      :" in ln:
            mm = re.search(r"Engine      : (\d+) GPU\(s\)(?: of the (\d+) offered)?", ln)
            if mm:
                rep.update(gpus_used=int(mm.group(1)), gpus_offered=int(mm.group(2) or mm.group(1)))
        elif "Host phases :" in ln:
            ph = {}
            for part in ln.split(":", 1)[1].split("|"):
                mm = re.match(r"\s*(.+?) ([\d.]+)-([\d.]+) s\s*$", part)
                if mm:
                    ph[mm.group(1)] = [float(mm.group(2)), float(mm.group(3))]
            rep["phases_s"] = ph
    return rep


def run_cli(prefix, out_prefix, gpus, precision, loco=False):
    cmd = [str(CLI_BIN), "grm", "-b", str(prefix), "-o", str(out_prefix), "--gm", "SH", "--add", "--precision", precision, "--gpus", str(gpus)]
    if loco:
        cmd += ["--loco", "--with-whole"]
    t0 = time.time()
    r = subprocess.run(cmd, capture_output=True, text=True)
    wall = time.time() - t0
    rep = parse_cli_report(r.stdout)
    rep.update(cli_wall_s=wall, rc=r.returncode)
    if r.returncode != 0:
        rep["stderr"] = (r.stderr or r.stdout)[-600:]
    return rep


def cli_wall_leg(cfg, gpus, precision, device=0):
    """t_wall of SURVEY 8d: `gelex-b200 grm` on real files (tmpfs when it has room), process start -> files closed."""
    n, m = cfg["n"], cfg["m"]
    bed_bytes = m * ((n + 3) // 4)
    out_bytes = n * (n + 1) // 2 * 4 * (23 if cfg["loco"] else 1)
    base = None
    for cand in ("/dev/shm", tempfile.gettempdir()):
        try:
            st = os.statvfs(cand)
            if st.f_bavail * st.f_frsize > (bed_bytes + out_bytes) * 1.05 + (1 << 30):
                base = cand
                break
        except OSError:
            continue
    if base is None:
        return {"unavailable": f"no file system with {(bed_bytes + out_bytes) / 1e9:.0f} GB free for real files"}
    d = Path(tempfile.mkdtemp(prefix="ggrm_wall_", dir=base))
    try:
        t0 = time.time()
        write_triple(d / "in", n, m, synth_rows_device(n, m, cfg["seed"], device), chromosome_runs(m))
        gen_s = time.time() - t0
        rep = run_cli(d / "in", d / "out", gpus, precision, loco=cfg["loco"])
        sizes = [p.stat().st_size for p in d.glob("out.*.bin")]
        ph = rep.get("phases_s", {})
        rep.update({"fs": fs_of(d), "dir": base, "input_gen_s": gen_s, "bed_bytes": bed_bytes, "out_files": len(sizes), "out_bytes": int(sum(sizes)),
                    "all_sizes_ok": bool(sizes) and all(s == n * (n + 1) // 2 * 4 for s in sizes),
                    "startup_s": ph.get("push", [None])[0],
                    "effective_tflops_wall": flops(n, m) / rep["cli_wall_s"] / 1e12,
                    "command": f"gelex-b200 grm -b in -o out --gm SH --add --precision {precision} --gpus {gpus}" + (" --loco --with-whole" if cfg["loco"] else "")})
        return rep
    finally:
        shutil.rmtree(d, ignore_errors=True)


def c1_device_leg(g, torch, precision, device, sustained):
    """configs[0] through the C ABI with the BED resident in HBM: device pass (first push enqueued -> last accumulation done, CUDA
    events inside the library) of six passes over one handle; median and best.  The few-tile regime: split-K + the fused K1-K3 launch."""
    n, m = C1["n"], C1["m"]
    bpv = (n + 3) // 4
    bed = torch.empty(m * bpv, dtype=torch.uint8, device=f"cuda:{device}")
    g.synth_bed_device(bed.data_ptr(), n, 0, m, C1["seed"], MISSING, device=device)
    torch.cuda.synchronize(device)
    times = []
    with g.Accumulator(n, m, method="SH", precision=g.PREC_BF16 if precision == "bf16" else g.PREC_FP16, device=device) as acc:
        prev = acc.timers()
        for p in range(7):
            if p:
                acc.reset()
            acc.push_device(bed.data_ptr(), 0, m)
            acc.finish()
            t = acc.timers()
            if p:  # the first pass pays module loading
                times.append({k: t[k] - prev[k] for k in ("pass_ms", "prep_ms", "expand_ms", "syrk_ms", "kernel_launches")})
            prev = t
    times.sort(key=lambda d: d["pass_ms"])
    med, best = times[len(times) // 2], times[0]
    return {"device_pass_ms_median": round(med["pass_ms"], 4), "device_pass_ms_best": round(best["pass_ms"], 4), "prep_ms": round(med["prep_ms"] + med["expand_ms"], 4),
            "syrk_ms": round(med["syrk_ms"], 4), "kernel_launches_per_pass": int(med["kernel_launches"]), "precision": precision,
            "tflops_median": round(flops(n, m) / med["pass_ms"] / 1e9, 1), "frac_of_sustained": round(flops(n, m) / med["pass_ms"] / 1e9 / sustained, 3)}


# ---------------------------------------------------------------------------------------------------------------- GPU arm
class DevPtr:
    """Expose a raw device pointer to torch (zero-copy) through __cuda_array_interface__."""

    def __init__(self, ptr, n_elems, typestr):
        self.__cuda_array_interface__ = {"shape": (int(n_elems),), "typestr": typestr, "data": (int(ptr), False), "version": 3}


def peaks():
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        d = json.loads(p.read_text())
        return float(d.get("bf16_tflops_sustained", d.get("bf16_tflops", 1400.0))), float(d.get("bf16_tflops", 1590.0)), "measured"
    return 1400.0, 1590.0, "fallback"


DTYPE_NOTE = {"fp16": "f16 operands, f32 accumulate (tcgen05 kind::f16)", "bf16": "bf16 operands, f32 accumulate (tcgen05 kind::f16)", "fp64": "f64 (DMMA)"}


def gpu_arm(args):
    import numpy as np
    import torch
    import torch.distributed as dist

    import gelex_b200 as g
    from gelex_b200.sharding import chromosome_shards, row_slices, snp_shards

    cfg = CONFIGS[args.config]
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    if world != args.gpus:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}: launch with torch.distributed.run --nproc-per-node {args.gpus}")
    numa_note = bind_to_gpu_numa_node(local) if world > 1 else "numa: single rank, not bound"
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    n, m = cfg["n"], cfg["m"]
    bpv = (n + 3) // 4
    prec = {"fp16": g.PREC_FP16, "bf16": g.PREC_BF16, "fp64": g.PREC_FP64}[args.precision]
    loco = cfg["loco"]
    out_elems = g.tri(n)

    # ---- what this rank accumulates: SNP ranges [(begin, end)] in ascending order, and (LOCO) the groups they form
    if loco:
        chrom = chromosome_runs(m)
        my_groups = chromosome_shards(chrom, world)[rank]
        ranges = [chrom[gi] for gi in my_groups]
        per_rank_snps = [sum(chrom[gi][1] - chrom[gi][0] for gi in grp) for grp in chromosome_shards(chrom, world)]
    else:
        s0, s1 = snp_shards(m, world)[rank]
        ranges = [(s0, s1)]
        per_rank_snps = [e - b for b, e in snp_shards(m, world)]
    ms = sum(e - b for b, e in ranges)
    offs = np.cumsum([0] + [e - b for b, e in ranges])           # row offset of each range inside this rank's payload buffers
    d_bed = torch.empty(max(1, ms) * bpv, dtype=torch.uint8, device="cuda")
    for (b, e), o0 in zip(ranges, offs):
        for s in range(b, e, 32768):
            c = min(32768, e - s)
            g.synth_bed_device(d_bed.data_ptr() + int(o0 + s - b) * bpv, n, s, c, cfg["seed"], MISSING, device=local)
    h_bed = torch.empty(max(1, ms) * bpv, dtype=torch.uint8, pin_memory=True)
    h_bed.copy_(d_bed)
    torch.cuda.synchronize()
    acc = g.Accumulator(n, m, method=METHOD, precision=prec, device=local, groups=ranges if loco else None, keep_whole=True)
    if world > 1 and not loco:
        acc.set_push_end(ranges[0][1])  # this rank stops at the end of its SNP shard: equal slabs inside the shard
    h_full = torch.empty(out_elems, dtype=torch.float32, pin_memory=True) if (world > 1 and args.reduce == "nccl" and rank == 0 and not loco) else None

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def push_all(dev):
        for (b, e), o0 in zip(ranges, offs):
            if dev:
                acc.push_device(d_bed.data_ptr() + int(o0) * bpv, b, e - b)
            else:
                acc.push_ptr(h_bed.data_ptr() + int(o0) * bpv, b, e - b)

    # ---- N > 1: reduce-scatter by row slices over NVLink peer memory: rank r sums ITS rows from every peer's accumulator with one
    # kernel of plain P2P loads (ggrm_reduce_peer on CUDA-IPC mapped pointers) and, in the e2e leg, fetches only those rows, so the
    # D2H of the 5 GB result is spread over all ranks' PCIe links.  c2: the blocking form of the first warm-up pass (and of
    # --reduce peer); c4: how the whole-genome sum of the per-rank chromosome partials is formed.
    r0, r1 = row_slices(n, world)[rank]
    my_out_elems = g.tri(r1) - g.tri(r0)
    h_out = torch.empty(max(1, out_elems if loco else my_out_elems), dtype=torch.float32, pin_memory=True)
    peers = []
    state = {"ready": False, "slice": (0, 0)}

    def setup_peers():
        # accumulator addresses are stable across ggrm_reset; export once after the first finished pass
        mine = acc.ipc_export(g.KIND_ADD, g.GROUP_WHOLE)
        handles = [None] * world
        dist.all_gather_object(handles, mine)
        for r in range(world):
            if r != rank:
                peers.append(g.ipc_open(handles[r], device=local))
        state["slice"] = acc.row_slice(r0, r1)
        state["ready"] = True

    def reduce_partials():
        if world == 1:
            return
        if args.reduce == "nccl" and not loco:
            dptr, n_el, eb = acc.accum_buffer(g.KIND_ADD, g.GROUP_WHOLE)
            t = torch.as_tensor(DevPtr(dptr, n_el, "<f4" if eb == 4 else "<f8"), device="cuda")
            dist.reduce(t, dst=0, op=dist.ReduceOp.SUM)
            torch.cuda.synchronize()
            return
        if not state["ready"]:
            setup_peers()
        dist.barrier()                      # every rank's partial is complete (ggrm_finish returned everywhere)
        eb_, ee_ = state["slice"]
        t_r = time.time()
        if ee_ > eb_:
            acc.reduce_peer(g.KIND_ADD, g.GROUP_WHOLE, peers, eb_, ee_)   # synchronous: returns when this rank's rows are final
        state["reduce_s"] = state.get("reduce_s", 0.0) + (time.time() - t_r)
        state["reduce_n"] = state.get("reduce_n", 0) + 1
        dist.barrier()                      # peers have finished reading this rank's buffer before the next pass zeroes it

    # ---- c2, N > 1, default (--reduce pipe): the same reduce-scatter, pipelined band by band behind the accumulation.  The last
    # slabs of every shard are accumulated row band by row band (ggrm_finish_async); each rank's CUDA host callbacks raise a
    # flag per finished band in POSIX shared memory; the 2048-row strips of the matrix are dealt to the ranks by LPT on their area
    # (sharding.strip_owners), and the owner of a strip - as soon as all N flags of the strip's band are up - sums the peers'
    # rows into its own over NVLink (ggrm_reduce_rows_async, on the read-out stream) and, in the e2e leg, reads them out,
    # while every GPU is still on the tensor cores with the later bands.  No collective inside the pass.
    pipe = {"ready": False, "pass_id": 0}

    def all_ok(flag):
        t_ok = torch.tensor([1 if flag else 0], device="cuda", dtype=torch.int32)
        dist.all_reduce(t_ok, op=dist.ReduceOp.MIN)
        return int(t_ok.item()) == 1

    def setup_pipe():
        """Collectives are unconditional; everything that can fail locally is guarded and agreed on afterwards."""
        from multiprocessing import shared_memory

        name = [f"ggrm_bench_flags_{os.environ.get('MASTER_PORT', '0')}_{os.getpid()}"] if rank == 0 else [None]
        dist.broadcast_object_list(name, src=0)
        ok = True
        if rank == 0:
            try:
                pipe["shm"] = shared_memory.SharedMemory(name=name[0], create=True, size=world * 64 * 4)
                np.ndarray((world, 64), dtype=np.int32, buffer=pipe["shm"].buf)[:] = 0
            except Exception as e:  # noqa: BLE001  (e.g. no POSIX shared memory in this container)
                sys.stderr.write(f"[bench] band pipeline unavailable ({type(e).__name__}: {e}); using --reduce peer\n")
                ok = False
        if not all_ok(ok):
            return False
        if rank != 0:
            try:
                pipe["shm"] = shared_memory.SharedMemory(name=name[0])
                try:  # only the creator unlinks; keep this process's resource tracker from trying again at exit (Python < 3.13)
                    from multiprocessing import resource_tracker

                    resource_tracker.unregister(pipe["shm"]._name, "shared_memory")
                except Exception:  # noqa: BLE001
                    pass
            except Exception as e:  # noqa: BLE001
                sys.stderr.write(f"[bench] rank {rank}: cannot attach band flags ({type(e).__name__}: {e}); using --reduce peer\n")
                ok = False
        if ok:
            pipe["flags"] = np.ndarray((world, 64), dtype=np.int32, buffer=pipe["shm"].buf)
            from gelex_b200.sharding import piece_owners

            bands = acc.band_layout()
            # strips cut along their columns into ~64 MB pieces of whole 2048 x 2048 super-tile columns; every band's pieces are dealt
            # over all ranks, so the last band's reduce + read-out (the exposed tail) is ~70 MB per rank instead of 393 MB on one
            owners = piece_owners(n, world, bands)
            pipe["n_bands"] = len(bands)

            def geom(si, c0, c1):
                row0, col0 = si * 2048, c0 * 2048
                return row0, min(n, row0 + 2048) - row0, col0, min(n, c1 * 2048) - col0

            pipe["mine"] = [(b, si, c0, c1) + geom(si, c0, c1) for b, si, c0, c1 in owners[rank]]
            elems = sum(rows * cols for *_, rows, _, cols in pipe["mine"])
            pipe["h_out"] = torch.empty(max(1, elems), dtype=torch.float32, pin_memory=True)
            pipe["d2h_bytes_all_ranks"] = int(sum(geom(si, c0, c1)[1] * geom(si, c0, c1)[3] for lst in owners for _, si, c0, c1 in lst) * 4)
            T = (n + 255) // 256

            def stored_elems(si, c0, c1):  # accumulator elements of the piece (whole 256 x 256 tiles)
                ri = min(T, si * 8 + 8) - si * 8
                return sum(ri * 8 if c < si else ri * (ri + 1) // 2 for c in range(c0, c1)) * 65536

            pipe["nvlink_bytes"] = int(sum(stored_elems(si, c0, c1) for _, si, c0, c1, *_ in pipe["mine"]) * 4 * (world - 1))
        if not all_ok(ok):
            return False
        pipe["ready"] = True
        return True

    def pipelined_pass(dev, fetch):
        pipe["pass_id"] += 1
        pid = pipe["pass_id"]
        acc.reset()
        acc.set_band_flags(pipe["flags"][rank].ctypes.data, pid)
        push_all(dev)
        acc.finish_async()
        off = 0
        for b, si, c0, c1, row0, rows, col0, cols in pipe["mine"]:
            t_w = time.time()
            col = pipe["flags"][:, b]
            while not (col == pid).all():
                if time.time() - t_w > 120:
                    raise SystemExit(f"rank {rank}: band {b} flags {col.tolist()} never reached {pid}")
                time.sleep(2e-5)
            acc.reduce_block_async(g.KIND_ADD, g.GROUP_WHOLE, peers, si, c0, c1)
            if fetch:  # the piece as a dense rows x cols block (entries above the diagonal read 0); a writer stores its row segments at their .bin offsets
                acc._check(acc.lib.ggrm_fetch_block(acc.h, g.KIND_ADD, g.GROUP_WHOLE, row0, rows, col0, cols,
                                                    ctypes.c_void_p(pipe["h_out"].data_ptr() + 4 * off), cols))
                off += rows * cols
        acc.finish()      # waits for the accumulation and the read-out stream (the reduces), reports device faults
        dist.barrier()    # peers have finished reading this rank's rows before the next pass overwrites them

    use_pipe = world > 1 and args.reduce == "pipe" and not loco

    def fetch_rows(group, lo, hi, dst_off_elems=0):
        acc.lib.ggrm_fetch(acc.h, g.KIND_ADD, group, lo, hi - lo, ctypes.c_void_p(h_out.data_ptr() + 4 * dst_off_elems), None)

    def step_resident():
        if use_pipe and pipe["ready"]:
            pipelined_pass(True, fetch=False)
            return
        acc.reset()
        push_all(True)
        acc.finish()
        reduce_partials()
        if use_pipe and not pipe.get("failed"):  # first warm-up pass ran the blocking form (it maps the peers' accumulators); arm the pipeline now
            if not setup_pipe():
                pipe["failed"] = True
                args.reduce = "peer"

    def step_e2e():
        if use_pipe and pipe["ready"]:
            pipelined_pass(False, fetch=True)
            return
        acc.reset()
        if loco:
            # a chromosome's partial is read out (5 GB over PCIe) while the next chromosome is on the tensor cores: pushes from pinned
            # memory only enqueue, ggrm_fetch blocks this thread for the finished group alone
            for li, ((b, e), o0) in enumerate(zip(ranges, offs)):
                acc.push_ptr(h_bed.data_ptr() + int(o0) * bpv, b, e - b)
                if li > 0:
                    fetch_rows(li - 1, 0, n)
            acc.finish_async()
            if ranges:
                fetch_rows(len(ranges) - 1, 0, n)
            acc.finish()
            reduce_partials()               # whole = sum over ranks of their chromosomes' partial sums (N = 1: already folded)
            if r1 > r0:
                fetch_rows(g.GROUP_WHOLE, r0, r1)
            return
        push_all(False)
        if world == 1:
            # ggrm_finish_async: the last slab runs in row bands and ggrm_fetch (which blocks until the requested rows are on
            # the host) starts reading the first bands out while the later ones are still accumulating
            acc.finish_async()
        else:
            acc.finish()
        reduce_partials()
        if world > 1 and args.reduce == "nccl":
            if rank == 0:
                acc.lib.ggrm_fetch(acc.h, g.KIND_ADD, g.GROUP_WHOLE, 0, n, ctypes.c_void_p(h_full.data_ptr()), None)
        elif r1 > r0:
            fetch_rows(g.GROUP_WHOLE, r0, r1)

    wall_checks = []

    def timed(fn, steps, warmup):
        for _ in range(warmup):
            fn()
        barrier()
        before = acc.timers()
        w0 = time.time()
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ev0.record()
        for _ in range(steps):
            fn()
        ev1.record()
        barrier()
        w1 = time.time()
        after = acc.timers()
        # device clock: ev0/ev1 sit on torch's (idle) stream; every step ends host-synchronously (ggrm_finish / fetch / NCCL
        # + synchronize), so the two records bracket exactly the device work of the K steps.  Wall time is the cross-check.
        dt = ev0.elapsed_time(ev1) * 1e-3
        wall_checks.append((dt, w1 - w0))
        if world > 1:
            t = torch.tensor([dt], device="cuda", dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            dt = float(t.item())
        return dt, {k: after[k] - before[k] for k in after}, (w0, w1)

    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    dt, tm, win = timed(step_resident, args.steps, args.warmup)
    clocks = sampler.stop(*win) if rank == 0 else None
    dt_e, tm_e, _ = timed(step_e2e, args.steps, max(1, args.warmup // 2))
    fl = flops(n, m)
    # per-rank SYRK time of the resident leg (LPT balance / per-GPU roofline of the multi-GPU runs)
    syrk_all = [tm["syrk_ms"] / args.steps]
    if world > 1:
        t = torch.zeros(world, device="cuda", dtype=torch.float64)
        t[rank] = tm["syrk_ms"] / args.steps
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        syrk_all = [float(x) for x in t.tolist()]
    line = None
    if rank == 0:
        sustained, burst, how = peaks()
        syrk_ms = tm["syrk_ms"] / args.steps                      # this rank's SYRK kernel time per step (CUDA events)
        launches = max(1.0, tm["syrk_launches"] / args.steps)
        fl_rank = flops(n, ms)
        achieved = fl_rank / (syrk_ms * 1e-3) / 1e12 if syrk_ms > 0 else 0.0
        traffic, traffic_launch = None, None
        tp = ROOT / "profiles" / "roofline_traffic.json"
        if tp.exists():
            tj = json.loads(tp.read_text())
            traffic, traffic_launch = tj.get("syrk_dram_bytes_per_launch"), tj.get("launch")
        if world == 1:
            partition = "single GPU" + (": 22 partial accumulators + whole resident (23 x 5.06 GB), whole folded in the SYRK epilogue of each chromosome's last slab" if loco else "")
        elif loco:
            partition = (f"chromosome shards over {world} GPUs (whole chromosome runs dealt by LPT on SNP counts; SNPs per rank {per_rank_snps}, "
                         f"max/mean {max(per_rank_snps) / (sum(per_rank_snps) / world):.3f}): no reduce for the 22 partials; whole = reduce-scatter of the per-rank "
                         "sums by 2048-row slices over NVLink peer memory (CUDA IPC, P2P loads)")
        else:
            partition = {
                "nccl": f"SNP-sharded over {world} GPUs + NCCL sum-reduce of FP32 partial GRMs to rank 0",
                "peer": f"SNP-sharded over {world} GPUs + reduce-scatter of FP32 partial GRMs by 2048-row slices over NVLink peer memory (CUDA IPC, P2P loads) "
                        "after a blocking finish; each rank fetches its own rows",
                "pipe": f"SNP-sharded over {world} GPUs + reduce-scatter of FP32 partial GRMs over NVLink peer memory (CUDA IPC, P2P loads), pipelined "
                        f"band by band behind the accumulation: {pipe.get('n_bands')} row bands, 2048-row strips cut along their columns into ~64 MB pieces (whole super-tile "
                        "columns), every band's pieces dealt over all ranks, per-band completion flags raised by CUDA host callbacks in POSIX shared memory, no collective "
                        "inside the pass; each rank fetches its own pieces (ggrm_fetch_block)"}[args.reduce]
        d2h = int(out_elems * 4) * (23 if loco else 1)
        if use_pipe and pipe.get("ready"):
            d2h = pipe["d2h_bytes_all_ranks"]  # dense blocks: the diagonal super-tiles carry their zero upper halves too
        line = {
            "metric": cfg["metric"], "value": fl / (dt / args.steps) / 1e12, "unit": "TFLOP/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3, "wall_s_per_grm": dt / args.steps,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": DTYPE_NOTE[args.precision], "data": "synthetic",
            "config": {"workload": f"{cfg['tag']}: {cfg['what']}, {n} samples x {m} SNPs, packed 2-bit PLINK BED "
                                   f"({m * bpv / 1e9:.2f} GB), 1 % missing, seed {cfg['seed']}",
                       "partition": partition,
                       "l2": "inputs (packed BED 6-7.5 GB, 16-bit operand slabs 0.8 GB, accumulator 5 GB) are larger than the 126 MB L2; no flush needed",
                       "timing": "CUDA events around the K steps, barrier + cuda synchronize on both sides, max over ranks; kernel phases by CUDA events on the library's streams",
                       "chunk_snps": int(os.environ.get("GGRM_CHUNK_SNPS", 16384)), "host": numa_note},
            "device_pass_ms_per_step": tm["pass_ms"] / args.steps,
            "wall_clock_ms_per_step": wall_checks[0][1] / args.steps * 1e3,
            "phases_ms_per_step": {k: tm[k] / args.steps for k in ("h2d_ms", "prep_ms", "expand_ms", "syrk_ms")},
            "syrk_ms_per_rank": syrk_all, "snps_per_rank": per_rank_snps,
            "gpu_launches": int(tm["kernel_launches"]),
            "reduce_ms_per_call_rank0": (state.get("reduce_s", 0.0) / max(1, state.get("reduce_n", 0)) * 1e3) if world > 1 and not use_pipe else None,
            "reduce_bytes_over_nvlink_rank0": (pipe.get("nvlink_bytes") if use_pipe else
                                               int((state["slice"][1] - state["slice"][0]) * 4 * (world - 1)) if state["ready"] else None) if world > 1 else None,
            "clocks": clocks,
            "e2e": {"value": fl / (dt_e / args.steps) / 1e12, "unit": "TFLOP/s", "wall_s_per_grm": dt_e / args.steps,
                    "h2d_bytes_per_step": int(m * bpv), "d2h_bytes_per_step": d2h,
                    "path": ("ggrm_reset -> per chromosome: ggrm_push(pinned host BED) then ggrm_fetch of the previous chromosome's partial (float32 lower triangle to "
                             "pinned host memory, overlapping the next chromosome's accumulation) -> ggrm_finish -> [N>1: reduce-scatter of the whole over NVLink] -> "
                             "ggrm_fetch(whole; N>1: each rank its row slice); bytes summed over ranks") if loco else
                            "ggrm_reset -> ggrm_push(pinned host BED) -> ggrm_finish_async -> [N>1, per owned piece: wait for all ranks' band flags -> ggrm_reduce_block_async -> ggrm_fetch_block] / [N=1: ggrm_fetch(float32 lower-triangle rows)] to pinned host memory; bytes summed over ranks",
                    "phases_ms_per_step": {k: tm_e[k] / args.steps for k in ("h2d_ms", "prep_ms", "expand_ms", "syrk_ms", "fetch_ms")}},
            "roofline": {"bound": "tensor", "achieved": achieved, "peak": sustained, "unit": "TFLOP/s", "frac": achieved / sustained,
                         "traffic": traffic, "kernel": "syrk_tc_kernel<2> (tcgen05 cta_group::2, UMMA 256x256x16)",
                         "traffic_launch": traffic_launch,
                         "flops_per_launch": fl_rank / launches, "avg_launch_ms": syrk_ms / launches, "launches_per_step": launches,
                         "whole_job_frac_of_n_x_burst": fl / (dt / args.steps) / 1e12 / (burst * world),
                         "launch_note": "achieved = this rank's credited flops of a step / its summed SYRK launch time (CUDA events on the library's compute "
                                        "stream); a step is ceil(m_group/16384) equal slabs per group, the last four accumulated band-major (up to 8 row bands x 4 slabs); traffic is the ncu "
                                        "--set full DRAM bytes of ONE steady-state 16384-SNP slab launch (traffic_launch)",
                         "peak_source": f"MEASURED_PEAKS.json bf16_tflops_sustained ({how}); burst peak {burst} -> frac {achieved / burst:.3f}"},
        }
    for p_ in peers:
        g.ipc_close(p_, device=local)
    acc.close()
    del d_bed, h_bed, h_out
    torch.cuda.empty_cache()
    if pipe.get("shm") is not None:
        pipe["flags"] = None
        pipe["shm"].close()
        if world > 1:
            dist.barrier()
        if rank == 0:
            pipe["shm"].unlink()
    if world > 1:
        dist.barrier()  # every rank has released its accumulators: the CLI below gets the GPUs to itself
    if rank == 0 and (not args.no_wall) and (not loco or args.wall) and CLI_BIN.exists():
        try:
            line["wall"] = cli_wall_leg(cfg, world, args.precision, device=local)
        except Exception as e:  # noqa: BLE001
            line["wall"] = {"unavailable": f"{type(e).__name__}: {e}"}
    if rank == 0 and world == 1 and not args.no_cpu:
        cb, tmp = cpu_leg(1, 0, cfg, lambda n_, m_, seed: [g.synth_bed_host(n_, 0, m_, seed, MISSING)])
        try:
            c1p = cb.pop("_c1_prefix", None)
            if c1p is not None and CLI_BIN.exists():
                ours = run_cli(c1p, str(tmp / "c1_ours"), 1, args.precision)
                cb["like_for_like"].update({"ours_cli_wall_s": ours["cli_wall_s"], "ours_engine_wall_s": ours.get("engine_wall_s"),
                                            "ours_phases_s": ours.get("phases_s"), "ours_rc": ours["rc"]})
                try:
                    cb["like_for_like"]["ours_device"] = {pr: c1_device_leg(g, torch, pr, local, peaks()[0]) for pr in ("fp16", "bf16")}
                except Exception as e:  # noqa: BLE001
                    cb["like_for_like"]["ours_device"] = {"unavailable": f"{type(e).__name__}: {e}"}
                line["like_for_like"] = cb.pop("like_for_like")
            midp = cb.pop("_mid_prefix", None)
            if midp is not None and CLI_BIN.exists():
                ours = run_cli(midp, str(tmp / "mid_ours"), 1, args.precision)
                cb["like_for_like_mid"].update({"ours_cli_wall_s": ours["cli_wall_s"], "ours_engine_wall_s": ours.get("engine_wall_s"),
                                                "ours_phases_s": ours.get("phases_s"), "ours_rc": ours["rc"]})
                line["like_for_like_mid"] = cb.pop("like_for_like_mid")
        finally:
            shutil.rmtree(tmp, ignore_errors=True)
        line["cpu_baseline"] = cb
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if rank == 0:
        print(json.dumps(line), flush=True)


def reference_arm(args):
    rank = int(os.environ.get("RANK", 0))
    if rank != 0:
        return
    cfg = CONFIGS[args.config]
    cb, tmp = cpu_leg(args.steps, args.warmup, cfg, synth_rows_numpy)
    shutil.rmtree(tmp, ignore_errors=True)
    cb.pop("_c1_prefix", None)
    cb.pop("_mid_prefix", None)
    n_s, m_s = CPU_SAMPLE if REF_BIN.exists() else (1500, 3000)
    line = {"impl": "reference", "metric": cfg["metric"], "value": cb["value"], "unit": "TFLOP/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": cb["wall_s"] * 1e3, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic",
            "config": {"workload": f"{cfg['tag']}: {cfg['what']}, {cfg['n']} samples x {cfg['m']} SNPs; each step = the fixed bounded sample "
                                   f"{n_s} x {m_s} of it (the CPU path needs hours and 40 GB for the full shape)"},
            "like_for_like": cb.pop("like_for_like", None),
            "like_for_like_mid": cb.pop("like_for_like_mid", None),
            "cpu_baseline": cb,
            "e2e": {"value": cb["value"], "unit": "TFLOP/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


def main():
    # keep stdout to the one JSON line: NCCL prints its "NCCL version ..." banner to stdout at NCCL_DEBUG=VERSION and above
    # (WARN included), so drop the quiet levels and send whatever remains (INFO/TRACE asked for by the caller) to stderr
    if os.environ.get("NCCL_DEBUG", "").upper() in ("", "VERSION", "WARN"):
        os.environ.pop("NCCL_DEBUG", None)
    os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default=os.environ.get("GGRM_BENCH_CONFIG", "c2"), choices=sorted(CONFIGS),
                    help="c2 = configs[1] (the headline shape), c4 = configs[3] (22 LOCO partials + whole)")
    ap.add_argument("--precision", default=os.environ.get("GGRM_BENCH_PRECISION"), choices=["fp16", "bf16", "fp64"],
                    help="operand type of the tcgen05 path (same pipe and throughput for both 16-bit types).  Default: bf16 for c2 - inside the "
                         "1e-4 bound at 500k SNPs; fp16 (the engine default) for c4, whose chromosome partials are 9.8k-52k SNPs each: plain bf16 "
                         "misses the bound on those and the library would split its operands (3x the tensor time).  tests/test_named_configs_gpu.py")
    ap.add_argument("--no-cpu", action="store_true", help="skip the CPU baseline leg (N = 1)")
    ap.add_argument("--no-wall", action="store_true", help="skip the CLI wall-clock leg")
    ap.add_argument("--wall", action="store_true", help="c4: run the CLI wall-clock leg too (writes 23 x 5 GB)")
    ap.add_argument("--reduce", default="pipe", choices=["pipe", "peer", "nccl"],
                    help="c2, N>1: partial-GRM reduction (pipe = own P2P kernel pipelined band by band behind the accumulation, "
                         "peer = the same kernel after a blocking finish, nccl = dist.reduce to rank 0)")
    args = ap.parse_args()
    if args.precision is None:
        args.precision = "fp16" if args.config == "c4" else "bf16"
    if args.impl == "reference":
        reference_arm(args)
    else:
        gpu_arm(args)


if __name__ == "__main__":
    main()
