#!/usr/bin/env python
"""Per-kernel summary of an .ncu-rep (`ncu -i rep --page raw --csv`): duration, DRAM bytes and
throughput, tensor-pipe activity, L2 hit rate, registers, grid -- the lines copied into profiles/."""
import csv, io, subprocess, sys

WANT = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_tensor_op_hmma.avg.pct_of_peak_sustained_active",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct",
        "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "sm__cycles_elapsed.avg.per_second"]
raw = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
ki = hdr.index("Kernel Name")
seen = {}
for r in rows[2:]:
    name = r[ki].split("(")[0]
    key = (name, r[hdr.index("launch__grid_size")] if "launch__grid_size" in hdr else "")
    if key in seen:
        continue
    seen[key] = 1
    print(f"== {name}")
    for w in WANT:
        if w in hdr:
            j = hdr.index(w)
            print(f"   {w} [{units[j]}] {r[j]}")
