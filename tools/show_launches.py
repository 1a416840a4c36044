import csv, sys
for wl in sys.argv[1:]:
    rows = [r for r in csv.reader(open(f"gpurun_out/launches_{wl}.csv")) if len(r) > 10]
    hdr = rows[0]; ki = hdr.index("Kernel Name"); vi = hdr.index("Metric Value"); ui = hdr.index("Metric Unit")
    rows = rows[1:]
    names = [r[ki] for r in rows]
    sims = [i for i, n in enumerate(names) if "sim32_kernel" in n or "portfolio_kernel" in n or "sim_kernel" in n]
    start = sims[-1]
    end = len(rows)
    tot = 0.0
    print(wl)
    for r in rows[start:end]:
        v = float(r[vi].replace(",", "")); tot += v
        print(f"   {r[ki][:70]:70s} {v/1e3:10.1f} us")
    print(f"   {'sum':70s} {tot/1e3:10.1f} us")
