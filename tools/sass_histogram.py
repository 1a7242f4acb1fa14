#!/usr/bin/env python
"""Per-kernel SASS mnemonic counts of scorp_b200/libgsr_b200.so (cuobjdump -sass) plus registers / spills / shared
memory from the build log (-Xptxas -v): the evidence that the kernels are sm_100a code using cp.async staging (LDGSTS),
packed FP32x2 math (FFMA2 / FMUL2 / FADD2), warp shuffles and match (SHFL / MATCH), REDs and multimem (REDG / the
.MULTIMEM forms), and no tensor-core or TMA instructions (north_star: no stage is a contraction).

    python tools/sass_histogram.py > profiles/r02_sass_histogram.md
"""
import collections
import re
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
LIB = ROOT / "scorp_b200" / "libgsr_b200.so"
COLS = ["LDGSTS", "FFMA2", "FMUL2", "FADD2", "FFMA", "MUFU", "SHFL", "MATCH", "VOTE", "RED", "ATOM", "MULTIMEM", "LDS", "STS",
        "LDG", "STG", "BAR", "UTMA", "MMA"]


def demangle(names):
    out = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True).stdout.splitlines()
    return dict(zip(names, out))


def main():
    sass = subprocess.run(["cuobjdump", "-sass", str(LIB)], capture_output=True, text=True, check=True).stdout
    arch = sorted(set(re.findall(r"arch = (sm_\w+)", sass)))
    kernels, cur = collections.OrderedDict(), None
    for line in sass.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            cur = m.group(1); kernels[cur] = collections.Counter(); continue
        m = re.match(r"\s*/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d+\s+)?([A-Za-z0-9_.]+)", line)
        if m and cur:
            op = m.group(1)
            c = kernels[cur]
            c["total"] += 1
            base = op.split(".")[0]
            if base in ("REDG", "RED", "REDUX"):
                c["RED"] += 1
                if "F32x4" in op or "F32x2" in op:      # red.global.add.v4/.v2.f32 (backward blends, direct visits)
                    c["REDvec"] += 1
            elif base in ("ATOMG", "ATOMS", "ATOM"):
                c["ATOM"] += 1
            elif base in ("HMMA", "IMMA", "DMMA", "QMMA", "OMMA") or base.startswith("UTC") or "MMA" in base:
                c["MMA"] += 1
            elif base.startswith("UTMA") or base.startswith("UBLKCP"):
                c["UTMA"] += 1
            elif base in COLS:
                c[base] += 1
            if base == "LDG" and ".256" in op:
                c["LDG256"] += 1
            if base == "LDGMC" or "MULTIMEM" in op:     # multimem.ld_reduce -> LDGMC.E.ADD.F32 (multimem.st is a plain STG
                c["MULTIMEM"] += 1                      # .STRONG.SYS to the multicast address)
    log = (ROOT / "scorp_b200" / "build.log").read_text() if (ROOT / "scorp_b200" / "build.log").exists() else ""
    res = {}
    for m in re.finditer(r"Function properties for (\S+)\n\s+(\d+) bytes stack frame, (\d+) bytes spill stores, (\d+) bytes spill loads\n"
                         r"ptxas info\s+: Used (\d+) registers(?:, used \d+ barriers)?(?:, (\d+) bytes cumulative stack size)?(?:, (\d+) bytes smem)?", log):
        res[m.group(1)] = (int(m.group(5)), int(m.group(3)), int(m.group(7) or 0))
    names = demangle(list(kernels))
    print(f"# SASS histogram of scorp_b200/libgsr_b200.so ({', '.join(arch)}; cuobjdump -sass; registers / spill bytes / static smem from -Xptxas -v)\n")
    print("| kernel | regs | spill B | smem B | instr | " + " | ".join(COLS) + " |")
    print("|---|---|---|---|---|" + "---|" * len(COLS))
    for k, c in kernels.items():
        short = re.sub(r"\(.*", "", names.get(k, k)).replace("gsr::", "").replace("void ", "")
        r = res.get(k, ("", "", ""))
        print(f"| `{short}` | {r[0]} | {r[1]} | {r[2]} | {c['total']} | " + " | ".join(str(c[x]) if c[x] else "" for x in COLS) + " |")
    tot = collections.Counter()
    for c in kernels.values():
        tot.update(c)
    print(f"\nAll kernels: {len(kernels)} functions, {tot['total']} instructions; LDGSTS {tot['LDGSTS']}, FFMA2+FMUL2+FADD2 "
          f"{tot['FFMA2'] + tot['FMUL2'] + tot['FADD2']}, MATCH {tot['MATCH']}, multimem.ld_reduce (LDGMC) {tot['MULTIMEM']}, "
          f"vector REDG (F32x4 / F32x2) {tot['REDvec']}, 32-byte LDG (LDG.E.ENL2.256) {tot['LDG256']}, "
          f"TMA {tot['UTMA']}, tensor-core MMA {tot['MMA']}.")


if __name__ == "__main__":
    sys.exit(main())
