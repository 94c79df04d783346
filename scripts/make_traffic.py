"""profiles/traffic.json from an `ncu --set full` report: DRAM bytes (read + write) per launch of each kernel stage,
averaged over the captured launches.  python scripts/make_traffic.py <rep> [tag]"""
import csv, json, os, subprocess, sys
rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr, units, data = rows[0], rows[1], rows[2:]
ix = {h: i for i, h in enumerate(hdr)}
scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
stage_of = {"preprocess_forward_kernel": "preprocess", "bucket_scan_kernel": "bin_emit", "bucket_emit_kernel": "bin_emit",
            "tile_sort_gather_kernel": "sort", "big_tile_sort_gather_kernel": "sort", "block_lists_kernel": "ranges_gather",
            "render_forward_kernel": "render_forward", "render_backward_kernel": "render_backward",
            "render_backward2_kernel": "render_backward",
            "preprocess_backward_kernel": "preprocess_backward"}
per_kernel = {}
for r in data:
    name = r[ix["Kernel Name"]].split("(")[0].split("<")[0].replace("void ", "").replace("tsplat::", "")
    b = sum(float(r[ix[m]]) * scale[units[ix[m]]] for m in ("dram__bytes_read.sum", "dram__bytes_write.sum"))
    per_kernel.setdefault(name, []).append(b)
avg = {k: sum(v) / len(v) for k, v in per_kernel.items()}
# issue-slot utilisation and warp instructions per launch of the two render kernels (they are bound by instruction
# issue, not by a memory level: DESIGN.md 9)
issue = {}
for r in data:
    name = r[ix["Kernel Name"]].split("(")[0].split("<")[0].replace("void ", "").replace("tsplat::", "")
    if name in stage_of and "smsp__issue_active.avg.pct_of_peak_sustained_active" in ix:
        issue.setdefault(stage_of[name], []).append((float(r[ix["smsp__issue_active.avg.pct_of_peak_sustained_active"]]),
                                                     float(r[ix["smsp__inst_executed.sum"]]),
                                                     float(r[ix["sm__throughput.avg.pct_of_peak_sustained_elapsed"]])
                                                     if "sm__throughput.avg.pct_of_peak_sustained_elapsed" in ix else 0.0))
out = {}
for k, v in avg.items():
    if k in stage_of:
        out[stage_of[k]] = out.get(stage_of[k], 0.0) + v
out = {k: round(v) for k, v in out.items()}
out["_issue"] = {k: {"issue_slot_pct": round(sum(x[0] for x in v) / len(v), 1),
                     "warp_instructions": round(sum(x[1] for x in v) / len(v)),
                     "sm_throughput_pct": round(sum(x[2] for x in v) / len(v), 1)}
                 for k, v in issue.items() if k in ("render_forward", "render_backward")}
out["_source"] = f"{os.path.basename(rep)}: dram__bytes_read.sum + dram__bytes_write.sum per launch, ncu --set full --clock-control none; per kernel: " + \
                 ", ".join(f"{k} {v / 1e6:.1f} MB" for k, v in sorted(avg.items()))
json.dump(out, open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "profiles", "traffic.json"), "w"), indent=1)
print(out)
