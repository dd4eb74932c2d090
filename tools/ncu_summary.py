#!/usr/bin/env python
"""Print the per-kernel metrics we track from an `ncu --page raw --csv` dump."""
import csv, sys
rows=list(csv.reader(open(sys.argv[1])))
hdr=rows[0]; data=rows[2:]
want=['Kernel Name','gpu__time_duration.sum','launch__registers_per_thread','launch__grid_size','sm__warps_active.avg.pct_of_peak_sustained_active','smsp__thread_inst_executed_per_inst_executed.ratio','sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active','sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active','smsp__issue_active.avg.pct','l1tex__t_sector_hit_rate.pct','lts__t_sector_hit_rate.pct','dram__bytes_read.sum','dram__bytes_write.sum','smsp__inst_executed.sum','sm__throughput.avg.pct_of_peak_sustained_elapsed']
want += [h for h in hdr if 'warp_issue_stalled' in h and h.endswith('per_warp_active.pct')]
idx={h:i for i,h in enumerate(hdr)}
for w in want:
    if w in idx:
        print(w.replace('smsp__warp_issue_stalled_','stall:').replace('_per_warp_active.pct','')[:60].ljust(62), [r[idx[w]][:22] for r in data])
