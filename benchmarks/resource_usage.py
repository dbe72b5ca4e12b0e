#!/usr/bin/env python
"""Registers, static shared memory, stack (spills) and arch of every kernel in the built library (no GPU needed):

    python benchmarks/resource_usage.py > profiles/resource_usage.txt
"""
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "distributed_deep_learning_on_personal_computers_b200", "lib", "libb200seg.so")


def main():
    out = subprocess.run(["cuobjdump", "-res-usage", LIB], capture_output=True, text=True, check=True).stdout
    rows, arch, src = [], "?", "?"
    lines = out.splitlines()
    for i, l in enumerate(lines):
        if l.startswith("arch ="):
            arch = l.split("=")[1].strip()
        if l.startswith("identifier ="):
            src = os.path.relpath(l.split("=")[1].strip(), ROOT)
        m = re.match(r"\s*Function (\S+):", l)
        if m and i + 1 < len(lines):
            res = dict(kv.split(":") for kv in lines[i + 1].split() if ":" in kv)
            rows.append((m.group(1), arch, src, res))
    names = subprocess.run(["cu++filt"] + [r[0] for r in rows], capture_output=True, text=True).stdout.splitlines()
    head = subprocess.run(["git", "-C", ROOT, "rev-parse", "--short", "HEAD"], capture_output=True, text=True).stdout.strip()
    print(f"# cuobjdump -res-usage lib/libb200seg.so, commit {head}: every kernel is sm_100a code; REG = registers per thread, "
          "SHARED = static shared memory (dynamic is set at launch), STACK > 0 means spills or local arrays")
    print(f"{'REG':>4s} {'SHARED':>7s} {'STACK':>6s} {'arch':8s} {'source':28s} kernel")
    for (mangled, arch, src, res), name in zip(rows, names):
        name = re.sub(r"\((?:[^()]|\([^()]*\))*\)\s*$", "", name).replace("(int)", "").replace("(bool)", "")
        print(f"{res.get('REG', '?'):>4s} {res.get('SHARED', '?'):>7s} {res.get('STACK', '?'):>6s} {arch:8s} {src:28s} {name}")


if __name__ == "__main__":
    main()
