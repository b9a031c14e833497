#!/usr/bin/env python
"""One table row per captured launch of an .ncu-rep (auxiliary kernels): scripts/ncu_table.py rep out.md "title" """
import csv
import io
import subprocess
import sys

rep, out, title = sys.argv[1], sys.argv[2], sys.argv[3]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, data = rows[0], rows[1], rows[2:]
keys = ["gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread", "dram__bytes_read.sum",
        "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__t_sector_hit_rate.pct"]
ix = {k: hdr.index(k) for k in keys if k in hdr}
name_i = hdr.index("Kernel Name")
lines = [f"# {title}", "", f"Source: `{rep}` (`ncu --set full --clock-control none --import-source on`), one row per captured launch.", "",
         "| kernel | " + " | ".join(f"`{k}` [{units[i]}]" for k, i in ix.items()) + " |", "|---|" + "---|" * len(ix)]
for d in data:
    nm = d[name_i].split("(")[0].replace("void ", "").replace("<unnamed>::", "")
    lines.append(f"| `{nm}` | " + " | ".join(d[i] for i in ix.values()) + " |")
open(out, "w").write("\n".join(lines) + "\n")
print("wrote", out)
