#!/bin/bash
# usage: tools/ncu_summary.sh X.ncu-rep [regions.txt] — headline metrics + per-line summary of one capture (run HERE)
rep=$1
ncu -i $rep --page raw --csv 2>/dev/null > /tmp/ncu_raw.csv
python - <<'P'
import csv
rows=list(csv.reader(open('/tmp/ncu_raw.csv')))
hdr=rows[0]; units=rows[1]; vals=rows[2]
want=['gpu__time_duration.sum','launch__grid_size','launch__block_size','launch__registers_per_thread','sm__cycles_elapsed.avg.per_second','smsp__inst_executed.sum','smsp__issue_active.avg.pct_of_peak_sustained_active','smsp__thread_inst_executed_per_inst_executed.ratio','sm__warps_active.avg.pct_of_peak_sustained_active','smsp__warps_eligible.avg.per_cycle_active','l1tex__t_sector_hit_rate.pct','lts__t_sector_hit_rate.pct','dram__bytes_read.sum','dram__bytes_write.sum','sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active','lts__t_sectors_op_red.sum','lts__t_sectors_op_atom.sum','l1tex__t_sectors_pipe_lsu_mem_global_op_red.sum','smsp__inst_executed_op_global_red.sum']
for w in want:
    if w in hdr:
        i=hdr.index(w); print(f"{w},{units[i]},{vals[i]}")
st=[(float(vals[i].replace(',','')) if vals[i] else 0,h) for i,h in enumerate(hdr) if h.startswith('smsp__pcsamp_warps_issue_stalled') and not h.endswith('_not_issued')]
tot=sum(x for x,_ in st) or 1
for x,h in sorted(st,reverse=True)[:8]: print(f"  {100*x/tot:4.1f}% {h}")
P
ncu -i $rep --page source --csv --print-source cuda,sass 2>/dev/null > /tmp/ncu_src.csv
python tools/ncu_lines.py /tmp/ncu_src.csv ${3:-40} $2
