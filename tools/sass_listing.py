#!/usr/bin/env python
"""Write per-kernel SASS listings of libstencil_b200.so to profiles/sass/ and a summary of the Blackwell-specific
mnemonics each kernel contains (UTMALDG = TMA tensor load, SYNCS = mbarrier, LDG/STG widths, MEMBAR, SHFL...)."""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "stencil2_b200", "lib", "libstencil_b200.so")
OUT = os.path.join(ROOT, "profiles", "sass")
KEEP = ["jacobi7_tma_kernelILb1ELi2", "jacobi7_tma_kernelILb0ELi2", "astaroth_tma_kernelIdLb0", "astaroth_tma_kernelIdLb1", "astaroth_tma_kernelIfLb0",
        "astaroth_cross_kernelId", "box_copy_kernelILb1", "box_copy_kernelILb0", "box_copy_batch_kernel",
        "wait_flags_kernel", "signal_flags_kernel", "reduce_sum_f32_kernel", "astaroth_kernelIdLb1", "astaroth_kernelIfLb0", "jacobi7_kernelILb0",
        "fingerprint_kernel"]
INTEREST = ["UTMALDG", "UTMASTG", "UBLKCP", "SYNCS", "MEMBAR", "FENCE", "STG.E.ENL2.256", "LDG.E.ENL2.256", "LDG.E.128", "STG.E.128", "LDG.E.64", "STG.E.64", "LDG.E ", "STG.E ", "LDS.128", "LDS ",
            "SHFL", "ATOMG", "RED", "FFMA", "DFMA", "FADD", "DADD", "LD.E.STRONG.SYS", "ST.E.STRONG.SYS", "ERRBAR", "CCTL", "NANOSLEEP", "LDL", "STL"]


def main():
    os.makedirs(OUT, exist_ok=True)
    sass = subprocess.run(["cuobjdump", "-sass", LIB], stdout=subprocess.PIPE, text=True, check=True).stdout
    blocks = re.split(r"\n\s*Function : ", sass)
    summary = []
    for b in blocks[1:]:
        name = b.split("\n", 1)[0].strip()
        short = next((k for k in KEEP if k in name), None)
        if not short:
            continue
        body = b
        try:
            dem = subprocess.run(["cu++filt", name], stdout=subprocess.PIPE, text=True).stdout.strip()
        except Exception:
            dem = name
        lines = [l for l in body.splitlines() if re.search(r"/\*[0-9a-f]{4,6}\*/", l)]
        ops = collections.Counter()
        for l in lines:
            m = re.search(r"\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", l)
            if m:
                ops[m.group(1)] += 1
        path = os.path.join(OUT, short.replace("ILb", "_").replace("ELi", "_") + ".sass")
        with open(path, "w") as f:
            f.write("// " + dem + "\n// extracted with cuobjdump -sass from libstencil_b200.so (sm_100a)\n")
            f.write("\n".join(re.sub(r"\s*/\* 0x[0-9a-f]{16} \*/", "", l) for l in body.splitlines() if l.strip() and "0x" in l and "/*" in l))
            f.write("\n")
        counts = {k.strip(): sum(v for op, v in ops.items() if op.startswith(k.strip())) for k in INTEREST}
        summary.append((short, dem, len(lines), {k: v for k, v in counts.items() if v}))
    with open(os.path.join(OUT, "SUMMARY.md"), "w") as f:
        f.write("# SASS summary (cuobjdump -sass, sm_100a)\n\nInstruction counts of interest per kernel; full listings beside this file.\n"
                "`UTMALDG` = `cp.async.bulk.tensor` (TMA), `SYNCS` = mbarrier ops, `LD/ST.E.STRONG.SYS` = acquire/release flag accesses.\n\n")
        for short, dem, n, c in summary:
            f.write(f"## {short}\n`{dem}`\n\n{n} instructions; " + ", ".join(f"{k}: {v}" for k, v in c.items()) + "\n\n")
    print(open(os.path.join(OUT, "SUMMARY.md")).read())


if __name__ == "__main__":
    sys.exit(main())
