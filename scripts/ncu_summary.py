#!/usr/bin/env python
"""Summarise an .ncu-rep (raw + source pages) into a small markdown file for profiles/.

    python scripts/ncu_summary.py gpurun_out/prof.ncu-rep profiles/name.md "title"
"""
import collections
import csv
import io
import subprocess
import sys

rep, out, title = sys.argv[1], sys.argv[2], sys.argv[3]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, data = rows[0], rows[1], rows[2:]
keys = ["Kernel Name", "gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
        "launch__occupancy_limit_registers", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__warps_active.avg.per_cycle_active", "smsp__warps_eligible.avg.per_cycle_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_elapsed",
        "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "smsp__cycles_active.avg", "sm__cycles_elapsed.avg",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__t_bytes.sum", "lts__t_sector_hit_rate.pct", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__t_sector_hit_rate.pct", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
        "smsp__thread_inst_executed_per_inst_executed.ratio"]
lines = [f"# {title}", "", f"Source: `{rep}` (`ncu --set full --clock-control none --cache-control none --import-source on`), "
         "read with `ncu -i ... --page raw|source --csv`. One row per captured launch.", "", "## Raw metrics", "",
         "| metric | " + " | ".join(f"launch {i}" for i in range(len(data))) + " | unit |", "|---|" + "---|" * (len(data) + 1)]
for k in keys:
    if k in hdr:
        i = hdr.index(k)
        vals = [d[i] for d in data]
        if k == "Kernel Name":
            vals = [v.split("(")[0][-60:] for v in vals]
        lines.append(f"| `{k}` | " + " | ".join(vals) + f" | {units[i]} |")

src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-kernel-base", "function"],
                     capture_output=True, text=True).stdout
srows = list(csv.reader(io.StringIO(src)))
sec = None
for r in srows:
    if r and r[0] == "Address":
        shdr = r
        sec = []
        continue
    if sec is not None and r and r[0] != "Kernel Name":
        sec.append(r)
    elif sec:
        break
if sec:
    ix = {h: i for i, h in enumerate(shdr)}
    stalls = [h for h in shdr if h.startswith("stall_") and "Not Issued" not in h]
    ops, tot = collections.Counter(), collections.Counter()
    for r in sec:
        toks = r[ix["Source"]].split()
        op = toks[1] if toks[0].startswith("@") else toks[0]
        ops[op.split(".")[0]] += int(r[ix["Instructions Executed"]])
        for st in stalls:
            tot[st] += int(r[ix[st]] or 0)
    total = sum(ops.values())
    lines += ["", "## Executed warp-instruction mix (source page, first launch)", "", "| opcode | warp instructions | share |", "|---|---|---|"]
    for op, n in ops.most_common(14):
        lines.append(f"| `{op}` | {n} | {100 * n / total:.1f} % |")
    lines.append(f"| total | {total} | |")
    s_all = sum(tot.values())
    lines += ["", "## Warp stall samples (all samples)", "", "| reason | samples | share |", "|---|---|---|"]
    for st, n in tot.most_common(9):
        lines.append(f"| `{st}` | {n} | {100 * n / s_all:.1f} % |")
open(out, "w").write("\n".join(lines) + "\n")
print("wrote", out)
