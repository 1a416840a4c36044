#!/usr/bin/env python
"""Instruction mix of a kernel's hot loop from `cuobjdump -sass` (no GPU needed).

    python tools/sass_loops.py <mangled-name-regex> [units-per-iteration] [label]

Takes every backward branch of the kernel as a loop [target, branch], picks the one with the most instructions that
contains no other backward branch (the innermost, unrolled body) and prints its mnemonic histogram -- in total and
per unit of work when `units-per-iteration` is given (e.g. 16 normals per iteration of the 2x unrolled GBM loop)."""
import collections
import re
import subprocess
import sys

LIB = "montecarlo_b200/libmc_b200.so"


def kernels(pattern):
    out = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
    cur, body = None, {}
    for line in out.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = m.group(1)
            body[cur] = []
            continue
        m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);", line)
        if m and cur:
            body[cur].append((int(m.group(1), 16), m.group(2).strip()))
    return {k: v for k, v in body.items() if re.search(pattern, k)}


def hot_loop(ins):
    loops = []
    for addr, text in ins:
        m = re.search(r"\bBRA(?:\.\w+)*\s+(?:\w+,\s*)?(0x[0-9a-f]+)", text)
        if m and int(m.group(1), 16) <= addr:
            loops.append((int(m.group(1), 16), addr))
    inner = [l for l in loops if not any(o != l and l[0] <= o[0] and o[1] <= l[1] for o in loops)]
    best = max(inner, key=lambda l: sum(1 for a, _ in ins if l[0] <= a <= l[1]))
    return best, [(a, t) for a, t in ins if best[0] <= a <= best[1]]


def mnemonic(text):
    t = re.sub(r"^@!?U?P\d+\s+", "", text)
    op = t.split()[0]
    parts = op.split(".")
    keep = parts[0]
    if keep in ("IMAD", "MUFU", "LDS", "STS", "LDG", "STG", "ATOMS", "RED", "ATOMG", "FFMA2", "FMUL2", "HFMA2", "LOP3", "SHF"):
        keep = ".".join(parts[:2]) if len(parts) > 1 and parts[1] in ("WIDE", "HI", "LG2", "EX2", "SQRT", "RSQ", "SIN", "COS", "RCP", "64", "128", "U16", "MOV") else parts[0]
    return keep


def main():
    pat = sys.argv[1]
    units = float(sys.argv[2]) if len(sys.argv) > 2 else None
    label = sys.argv[3] if len(sys.argv) > 3 else ""
    for name, ins in kernels(pat).items():
        (lo, hi), body = hot_loop(ins)
        hist = collections.Counter(mnemonic(t) for _, t in body)
        print(f"== {label or name}\n   {name}\n   hot loop 0x{lo:x}..0x{hi:x}: {len(body)} instructions" +
              (f" = {len(body) / units:.2f} per unit ({units:g} units per iteration)" if units else ""))
        for op, c in hist.most_common():
            print(f"   {op:<12}{c:5d}" + (f"   {c / units:6.2f} / unit" if units else ""))


if __name__ == "__main__":
    main()
