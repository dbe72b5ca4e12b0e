#!/usr/bin/env python
"""Instruction counts per kernel from the built library's SASS (no GPU needed):

    python benchmarks/sass_summary.py > profiles/sass_summary.txt

Columns are the mnemonics that prove which hardware path a kernel uses (B200_PROFILING.md): UTCHMMA = tcgen05.mma,
UTMALDG / UTMASTG = TMA load / store, LDTM = tcgen05.ld, UTCBAR = tcgen05.commit, SYNCS = mbarrier, LDGMC / STGMC =
multimem.ld_reduce / multimem.st, REDG = red.global (incl. .v4 and multimem.red), HMMA / LDSM = warp-level mma.sync /
ldmatrix (the one-pass tail kernel's small-N GEMMs), LDGSTS = cp.async, UBLKCP = bulk copy."""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "distributed_deep_learning_on_personal_computers_b200", "lib", "libb200seg.so")
COLS = ["UTCHMMA", "UTMALDG", "UTMASTG", "LDTM", "UTCBAR", "SYNCS", "LDGMC", "STGMC", "REDG", "HMMA", "LDSM", "LDGSTS", "UBLKCP"]


def strip_params(sig: str) -> str:
    """'void ns::k<(int)64, true>(const T *, int)' -> 'ns::k<64,true>' (drop the return type and the parameter list)."""
    depth, end = 0, len(sig)
    for i in range(len(sig) - 1, -1, -1):               # the parameter list is the last balanced (...) group
        if sig[i] == ")":
            depth += 1
        elif sig[i] == "(":
            depth -= 1
            if depth == 0:
                end = i
                break
    name = sig[:end].replace("(int)", "").replace("(bool)", "").replace(", ", ",")
    return name[5:] if name.startswith("void ") else name


def main():
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    names = {}
    for m in re.finditer(r"Function : (\S+)", sass):
        names[m.group(1)] = None
    dem = subprocess.run(["cu++filt"] + list(names), capture_output=True, text=True).stdout.splitlines()
    for k, d in zip(list(names), dem):
        names[k] = strip_params(d)
    counts, order, cur = collections.defaultdict(collections.Counter), [], None
    for line in sass.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = names[m.group(1)]
            if cur not in order:
                order.append(cur)
            continue
        m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\w+\s+)?([A-Z0-9_]+)", line)
        if m and cur:
            op = m.group(1)
            counts[cur]["total"] += 1
            if op in COLS:
                counts[cur][op] += 1
    head = subprocess.run(["git", "-C", ROOT, "rev-parse", "--short", "HEAD"], capture_output=True, text=True).stdout.strip()
    print(f"# instruction counts per kernel, cuobjdump -sass lib/libb200seg.so (sm_100a), commit {head}")
    print("# " + __doc__.split("Columns are", 1)[1].strip().replace("\n", "\n# "))
    print("kernel " + " ".join(COLS) + " total")
    for k in order:
        c = counts[k]
        print(k + " " + " ".join(str(c[x]) for x in COLS) + f" {c['total']}")


if __name__ == "__main__":
    sys.exit(main())
