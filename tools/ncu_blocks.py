"""Instruction share per basic block from `ncu -i X.ncu-rep --page source --csv --print-source sass > src.csv`."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[1]; data = rows[2:]
ia = hdr.index("Instructions Executed"); isrc = hdr.index("Source"); iss = hdr.index("# Samples")
tot = sum(int(r[ia]) for r in data); stot = sum(int(r[iss]) for r in data)
print("total warp instructions", tot)
seg = 0; ss = 0; start = 0
for k, r in enumerate(data):
    seg += int(r[ia]); ss += int(r[iss])
    s = r[isrc].strip()
    if any(t in s for t in ("BRA", "EXIT", "BAR", "WARPSYNC")):
        if seg > tot * 0.005:
            print(f"{start:5d}-{k:5d} inst {seg / tot * 100:6.2f}%  samples {ss / stot * 100:6.2f}%  exec/inst {seg / (k - start + 1):.3g}  last: {s[:50]}")
        seg = 0; ss = 0; start = k + 1
