"""Per-source-line executed warp instructions and stall samples of one kernel from an ncu report (development helper).
usage: python scripts/src_hot.py <report.ncu-rep> <kernel regex> [top]"""
import collections, csv, subprocess, sys
rep, kern = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", f"regex:{kern}", "--launch-count", "1",
                      "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
cur = hdr = None
agg, samp, src = collections.Counter(), collections.Counter(), {}
for r in rows:
    if not r:
        continue
    if r[0] == "File Path":
        cur = r[1].split("/")[-1]
    elif r[0] == "Line No":
        hdr = r
        ix = {h: k for k, h in enumerate(hdr)}
    elif hdr and len(r) == len(hdr):
        try:
            key = (cur, int(r[0]))
        except ValueError:
            continue
        agg[key] += int(r[ix["Instructions Executed"]] or 0)
        samp[key] += int(r[ix["# Samples"]] or 0)
        src[key] = r[1].strip()[:120]
tot, ts = sum(agg.values()), max(sum(samp.values()), 1)
print("warp instructions", tot, "samples", ts)
for k, v in agg.most_common(top):
    print(f"{k[0]:18s} {k[1]:4d} {100 * v / tot:5.1f}% s{100 * samp[k] / ts:5.1f}%  {src[k]}")
