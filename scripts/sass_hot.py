"""Per-instruction executed counts / stall samples from `ncu --page source --csv` output.
usage: python scripts/sass_hot.py <csv> [top]   (csv made with: ncu -i rep --page source --csv --kernel-name regex:K --launch-count 1)"""
import csv, sys, collections
rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[1]; data = [r for r in rows[2:] if len(r) == len(hdr) and r[0] != "Address"]
ix = {h: i for i, h in enumerate(hdr)}
tot = sum(int(r[ix["Instructions Executed"]] or 0) for r in data)
samp = sum(int(r[ix["# Samples"]] or 0) for r in data)
by = collections.Counter(); bys = collections.Counter()
for r in data:
    op = r[ix["Source"]].split()
    op = [t for t in op if not t.startswith("@")][0].split(".")[0] if op else "?"
    by[op] += int(r[ix["Instructions Executed"]] or 0); bys[op] += int(r[ix["# Samples"]] or 0)
print("total warp inst", tot, "samples", samp)
for op, c in by.most_common(int(sys.argv[2]) if len(sys.argv) > 2 else 25):
    print(f"{op:10s} {c:12d} {100*c/tot:5.1f}%  samples {100*bys[op]/max(samp,1):5.1f}%")
if len(sys.argv) > 3:
    for n, r in enumerate(data):
        c = int(r[ix["Instructions Executed"]] or 0)
        print(f"{n:4d} {c:10d} {int(r[ix['# Samples']] or 0):6d} {float(r[ix['Avg. Predicated-On Threads Executed']] or 0):5.1f} {r[ix['Source']].strip()}")
