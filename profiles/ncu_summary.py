#!/usr/bin/env python
"""Summarise an .ncu-rep (read here, no GPU needed): one block of key metrics per captured launch.
usage: python profiles/ncu_summary.py gpurun_out/prof.ncu-rep > profiles/<name>.txt"""
import csv
import io
import re
import subprocess
import sys

raw = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
pat = re.compile(
    r"^(Kernel Name|gpu__time_duration\.sum|dram__bytes_(read|write)\.sum|gpu__dram_throughput\.avg\.pct_of_peak_sustained_elapsed"
    r"|lts__t_sector_hit_rate\.pct|l1tex__t_sector_hit_rate\.pct|sm__warps_active\.avg\.pct_of_peak_sustained_active"
    r"|launch__registers_per_thread|launch__grid_size|launch__block_size|launch__occupancy_limit_(registers|shared_mem|warps)"
    r"|sm__throughput\.avg\.pct_of_peak_sustained_elapsed|lts__throughput\.avg\.pct_of_peak_sustained_elapsed"
    r"|l1tex__throughput\.avg\.pct_of_peak_sustained_elapsed|smsp__issue_active\.avg\.pct_of_peak_sustained_active"
    r"|smsp__average_warps_issue_stalled_(long_scoreboard|short_scoreboard|barrier|wait|lg_throttle|mio_throttle|math_pipe_throttle|not_selected)_per_issue_active\.ratio"
    r"|smsp__inst_executed\.sum)$")
for r in rows[2:]:
    print("-" * 100)
    for i, h in enumerate(hdr):
        if pat.match(h):
            print(f"{h:92s} {r[i][:70]} {units[i]}")
