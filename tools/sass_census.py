"""Instruction census of the shipped library: `cuobjdump -sass libb200ops.so`, per kernel, counting the SASS
mnemonics that prove which hardware path a kernel uses (B200_PROFILING.md): UTCHMMA/UTCQMMA/UTCOMMA (tcgen05.mma
f16 / f8f6f4 / block-scaled), UTCCP (tcgen05.cp), LDTM/STTM (tcgen05.ld/st), UTMALDG/UTMASTG (TMA), UBLKCP (bulk
copy), HMMA (mma.sync), SYNCS (mbarrier), LDGMC (multimem.ld_reduce: in-switch reduction), LDG/STG/LDS/STS totals.
usage: python tools/sass_census.py [lib.so] > profiles/r2_sass_census.md   (no GPU needed)"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "vllm-fork_b200", "b200ops", "libb200ops.so")
WANT = ["UTCHMMA", "UTCQMMA", "UTCOMMA", "UTCCP", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UBLKCP", "UTMAPF", "HMMA",
        "SYNCS", "LDGMC", "LDG", "STG", "LDS", "STS", "ATOM", "RED", "MUFU", "total"]


def main():
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
    names = subprocess.run(["c++filt"], input="\n".join(re.findall(r"Function : (\S+)", sass)), capture_output=True,
                           text=True).stdout.splitlines()
    counts, order, cur, it = {}, [], None, iter(names)
    for line in sass.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = next(it)
            cur = re.sub(r"\(.*\)$", "", cur)           # drop the parameter list
            cur = re.sub(r"^void ", "", cur)
            counts[cur] = collections.Counter()
            order.append(cur)
            continue
        m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_]*)", line)
        if m and cur:
            op = m.group(1)
            counts[cur]["total"] += 1
            for w in WANT:
                if op.startswith(w):
                    counts[cur][w] += 1
                    break
    print(f"# SASS instruction census of `{os.path.relpath(LIB, ROOT)}` (cuobjdump -sass, sm_100a)\n")
    tot = collections.Counter()
    for k in order:
        tot.update(counts[k])
    print("Library totals: " + ", ".join(f"{w} {tot[w]}" for w in WANT if tot[w]) + f"; {len(order)} kernels\n")
    cols = [w for w in WANT if tot[w]]
    print("| kernel | " + " | ".join(cols) + " |\n|---|" + "---|" * len(cols))
    for k in sorted(order):
        c = counts[k]
        print(f"| `{k[:110]}` | " + " | ".join(str(c[w]) if c[w] else "" for w in cols) + " |")


if __name__ == "__main__":
    main()
