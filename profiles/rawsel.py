"""Print selected metrics of an `ncu --page raw --csv` dump, one column per kernel.  usage: rawsel.py raw.csv"""
import csv, re, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[0]
pat = re.compile(r'^(sm__inst_executed_pipe_\w+\.avg\.pct_of_peak_sustained_active|sm__pipe_\w+_cycles_active\.avg\.pct_of_peak_sustained_active|smsp__average_warps_issue_stalled_\w+_per_issue_active\.ratio|gpu__time_duration.sum|dram__bytes_read.sum|dram__bytes_write.sum|sm__inst_executed.avg.per_cycle_active|smsp__thread_inst_executed_per_inst_executed.ratio|sm__warps_active.avg.pct_of_peak_sustained_active|smsp__inst_executed.sum|l1tex__data_pipe_lsu_wavefronts_mem_shared.sum|l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum|l1tex__data_pipe_lsu_wavefronts.sum|l1tex__throughput.avg.pct_of_peak_sustained_elapsed|sm__throughput.avg.pct_of_peak_sustained_elapsed|lts__t_sectors_op_red.sum|lts__t_sectors_op_atom.sum|launch__registers_per_thread|launch__grid_size|launch__block_size|smsp__inst_executed_op_\w+.sum|sm__cycles_active.avg|lts__t_sector_hit_rate.pct|l1tex__t_sector_hit_rate.pct)$')
ks = [k for k in hdr if pat.match(k)]
print(f"{'metric':92s}", *[f"{r[hdr.index('Kernel Name')][5:18]:>14s}" for r in rows[2:]])
for k in ks:
    vals = [r[hdr.index(k)] for r in rows[2:]]
    if all(v in ('', '0') for v in vals): continue
    print(f"{k:92s}", *[f"{v[:14]:>14s}" for v in vals])
