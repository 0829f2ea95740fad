#!/usr/bin/env python3
"""Summarise one kernel of an .ncu-rep (read here, no GPU): python tools/ncu_summary.py rep [out.txt] [title]"""
import csv, io, subprocess, sys
KEEP = ['Kernel Name','Block Size','Grid Size','gpu__time_duration.sum','dram__bytes_read.sum','dram__bytes_write.sum',
 'launch__registers_per_thread','launch__occupancy_limit_registers','launch__waves_per_multiprocessor',
 'sm__warps_active.avg.pct_of_peak_sustained_active','smsp__inst_executed.sum','smsp__issue_active.avg.pct_of_peak_sustained_active',
 'smsp__thread_inst_executed_per_inst_executed.ratio','l1tex__t_sector_hit_rate.pct','lts__t_sector_hit_rate.pct',
 'l1tex__throughput.avg.pct_of_peak_sustained_elapsed','lts__throughput.avg.pct_of_peak_sustained_elapsed',
 'sm__throughput.avg.pct_of_peak_sustained_elapsed','smsp__average_warp_latency_per_inst_issued.ratio','smsp__warps_eligible.avg.per_cycle_active',
 'smsp__average_warps_issue_stalled_wait_per_issue_active.ratio','smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio',
 'smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio','smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio',
 'smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio','smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio',
 'smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio','smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio',
 'l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum','l1tex__t_sectors_pipe_lsu_mem_global_op_ld_lookup_miss.sum','l1tex__t_sectors_pipe_lsu_mem_global_op_st.sum','l1tex__t_sectors_pipe_lsu_mem_local_op_ld.sum','l1tex__t_sectors_pipe_lsu_mem_local_op_st.sum','lts__t_sectors.sum','lts__t_sectors_srcunit_tex_op_read.sum','lts__t_sectors_srcunit_tex_op_read_lookup_miss.sum','l1tex__t_bytes.sum','lts__t_bytes.sum','l1tex__t_bytes.sum.per_second','lts__t_bytes.sum.per_second','sm__inst_executed_pipe_fma.sum','sm__inst_executed_pipe_alu.sum','sm__inst_executed_pipe_xu.sum','sm__inst_executed_pipe_lsu.sum',
 'smsp__inst_executed_pipe_fma.sum','smsp__inst_executed_pipe_alu.sum','smsp__inst_executed_pipe_xu.sum','smsp__inst_executed_pipe_lsu.sum','smsp__inst_executed_pipe_fmaheavy.sum','smsp__inst_executed_pipe_fp32.sum']
rep = sys.argv[1]
raw = subprocess.run(['ncu','-i',rep,'--page','raw','--csv'],capture_output=True,text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
lines = []
if len(sys.argv) > 3: lines.append('# ' + sys.argv[3])
for vals in rows[2:]:
    for h,u,v in zip(hdr,units,vals):
        if h in KEEP: lines.append('%s [%s] = %s' % (h,u,v))
    lines.append('')
txt = '\n'.join(lines)
print(txt)
if len(sys.argv) > 2: open(sys.argv[2],'w').write(txt + '\n')
