"""SASS mnemonic counts per kernel of libggrm_b200.so (what proves tcgen05 / TMEM / TMA code: B200_PROFILING.md, "What proves a
Blackwell-native kernel").  Runs without a GPU.

    python tools/sass_table.py > profiles/r01_sass_mnemonics.md
"""
import re
import subprocess
from pathlib import Path

LIB = Path(__file__).resolve().parents[1] / "gelex_b200" / "libggrm_b200.so"
PAT = {"UTCHMMA (tcgen05.mma)": r"\bUTCHMMA", "UTCBAR (tcgen05.commit)": r"\bUTCBAR", "LDTM (tcgen05.ld)": r"\bLDTM",
       "UTMALDG (TMA load)": r"\bUTMALDG", "SYNCS (mbarrier)": r"\bSYNCS", "DMMA (FP64 mma.sync)": r"\bDMMA", "HMMA (legacy)": r"\bHMMA",
       "PRMT": r"\bPRMT", "POPC": r"\bPOPC", "LDGSTS (cp.async)": r"\bLDGSTS", "SHFL": r"\bSHFL", "LDG": r"\bLDG", "STG": r"\bSTG"}


def main():
    txt = subprocess.run(["cuobjdump", "-sass", str(LIB)], capture_output=True, text=True, check=True).stdout
    rows = []
    for f in re.split(r"\n\s*Function : ", txt)[1:]:
        name = f.split("\n", 1)[0].strip()
        dem = subprocess.run(["c++filt", name], capture_output=True, text=True).stdout.strip()
        dem = re.sub(r"\(.*", "", dem).replace("void ", "").replace("ggrm::", "")
        rows.append((dem, len(re.findall(r"/\*[0-9a-f]{4}\*/", f)), {k: len(re.findall(v, f)) for k, v in PAT.items()}))
    print("# SASS mnemonics per kernel (`cuobjdump -sass gelex_b200/libggrm_b200.so`, sm_100a; `tools/sass_table.py`)\n")
    print("Static instruction counts (loops are not unrolled into them). `syrk_tc_kernel<2>` is the kernel `bench.py` runs: its four `UTCHMMA` are")
    print("the K = 16 steps of one 64-element K block (`tcgen05.mma.cta_group::2.kind::f16`, 256 x 256 x 16), `UTMALDG` the two TMA box")
    print("loads per stage, `LDTM` the `tcgen05.ld.32x32b.x32` of the epilogue; no `HMMA` anywhere (no legacy tensor path).\n")
    res = {}
    usage = subprocess.run(["cuobjdump", "-res-usage", str(LIB)], capture_output=True, text=True, check=True).stdout
    for mname, body in re.findall(r"Function ([^:\s]+):\s*\n\s*(REG:[^\n]*)", usage):
        d = subprocess.run(["c++filt", mname], capture_output=True, text=True).stdout.strip()
        d = re.sub(r"\(.*", "", d).replace("void ", "").replace("ggrm::", "")
        res[d] = (re.search(r"REG:(\d+)", body).group(1), re.search(r"STACK:(\d+)", body).group(1), re.search(r"SHARED:(\d+)", body).group(1))
    keys = list(PAT)
    print("Registers / stack bytes / static shared bytes per thread block are `cuobjdump -res-usage` (the SYRK ring is dynamic shared memory, 194 KB,")
    print("and does not show here); no kernel spills.\n")
    print("| kernel | regs | stack | static smem | SASS lines | " + " | ".join(keys) + " |")
    print("|---|---:|---:|---:|---:|" + "---:|" * len(keys))
    for dem, n, c in sorted(rows):
        r = res.get(dem, ("", "", ""))
        print(f"| `{dem}` | {r[0]} | {r[1]} | {r[2]} | {n} | " + " | ".join(str(c[k]) if c[k] else "" for k in keys) + " |")


if __name__ == "__main__":
    main()
