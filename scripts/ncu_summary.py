"""Summarise an .ncu-rep (read here, no GPU): python scripts/ncu_summary.py <rep> [out.md]"""
import csv
import subprocess
import sys

rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr, units, data = rows[0], rows[1], rows[2:]
idx = {h: i for i, h in enumerate(hdr)}
want = [("Kernel Name", "kernel"), ("gpu__time_duration.sum", "time"), ("launch__grid_size", "grid"),
        ("launch__registers_per_thread", "regs"), ("launch__waves_per_multiprocessor", "waves"),
        ("dram__bytes_read.sum", "dram_rd"), ("dram__bytes_write.sum", "dram_wr"),
        ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram%"),
        ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm%"),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "occ%"),
        ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue%"),
        ("smsp__inst_executed.sum", "warp_inst"),
        ("smsp__thread_inst_executed_per_inst_executed.ratio", "thr/inst"),
        ("sm__cycles_elapsed.max", "cyc_max"), ("smsp__cycles_active.avg", "cyc_act"),
        ("l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "smem_wavefronts"),
        ("lts__t_sectors_op_red.sum", "l2_red_sectors"), ("lts__t_sectors_op_atom.sum", "l2_atom_sectors")]
lines = []
for r in data:
    parts = []
    for k, short in want:
        if k in idx:
            v = r[idx[k]]
            if short == "kernel":
                v = v.split("(")[0].replace("void tsplat::", "").replace("tsplat::", "")[:40]
            u = units[idx[k]] if short in ("time", "dram_rd", "dram_wr") else ""
            parts.append(f"{short}={v}{u}")
    lines.append(" | ".join(parts))
stall = [h for h in hdr if h.startswith("smsp__pcsamp_warps_issue_stalled_") and not h.endswith("_not_issued")]
for n, r in enumerate(data):
    vals = [(float(r[idx[h]] or 0), h.split("issue_stalled_")[1]) for h in stall]
    tot = sum(v for v, _ in vals) or 1.0
    top = sorted(vals, reverse=True)[:7]
    lines[n] += "\n    stall samples: " + ", ".join(f"{name} {100 * v / tot:.0f}%" for v, name in top)
out = "\n".join(lines)
print(out)
if len(sys.argv) > 2:
    open(sys.argv[2], "w").write(out + "\n")
