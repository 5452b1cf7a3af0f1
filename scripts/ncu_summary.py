#!/usr/bin/env python
"""Summarise an .ncu-rep (read with `ncu -i … --page raw --csv`) into the handful of metrics DESIGN.md quotes."""
import csv, subprocess, sys
WANT = ['gpu__time_duration.sum','launch__registers_per_thread','launch__grid_size','launch__occupancy_limit_registers','launch__occupancy_limit_shared_mem',
 'sm__warps_active.avg.pct_of_peak_sustained_active','smsp__thread_inst_executed_per_inst_executed.ratio','smsp__inst_executed.sum',
 'smsp__issue_active.avg.pct_of_peak_sustained_active','sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active',
 'sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active','sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active',
 'sm__throughput.avg.pct_of_peak_sustained_elapsed','l1tex__throughput.avg.pct_of_peak_sustained_elapsed','lts__throughput.avg.pct_of_peak_sustained_elapsed',
 'l1tex__t_sector_hit_rate.pct','lts__t_sector_hit_rate.pct','l1tex__t_bytes.sum','lts__t_bytes.sum','dram__bytes_read.sum','dram__bytes_write.sum',
 'smsp__average_warp_latency_per_inst_issued.ratio','smsp__sass_inst_executed_op_local_ld.sum','smsp__sass_inst_executed_op_local_st.sum',
 'smsp__sass_inst_executed_op_shared_ld.sum','smsp__sass_inst_executed_op_shared_st.sum','l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum']
def main(path):
    out = subprocess.run(['ncu','-i',path,'--page','raw','--csv'],capture_output=True,text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units = rows[0], rows[1]
    ki = hdr.index('Kernel Name')
    print('kernels:', [r[ki][:60] for r in rows[2:]])
    for w in WANT:
        if w in hdr:
            i = hdr.index(w); print(f'{w} [{units[i]}]', [r[i] for r in rows[2:]])
    for i,h in enumerate(hdr):
        if h not in WANT and (h.startswith('l1tex__data_pipe_lsu_wavefronts') or h.startswith('l1tex__t_sectors_pipe_lsu_mem_global_op') or h.startswith('l1tex__t_requests_pipe_lsu_mem_global_op')):
            print(f'{h} [{units[i]}]', [r[i] for r in rows[2:]])
    for i,h in enumerate(hdr):  # warps stalled per issue cycle, by reason (sums to the warp latency per instruction issued)
        if 'issue_stalled' in h and h.endswith('_per_issue_active.ratio') and 'not_issued' not in h:
            vals=[float(r[i] or 0) for r in rows[2:]]
            if max(vals) > 0.15: print('stall/issue:', h.replace('smsp__average_warps_issue_stalled_','').replace('smsp__average_warp_latency_issue_stalled_','').replace('_per_issue_active.ratio',''), [round(v,2) for v in vals])
    for i,h in enumerate(hdr):
        if 'warp_issue_stalled' in h and h.endswith('_per_warp_active.pct'):
            vals=[float(r[i]) for r in rows[2:]]
            if max(vals) > 5: print(h.replace('smsp__warp_issue_stalled_','stall:').replace('_per_warp_active.pct',''), [round(v,1) for v in vals])
if __name__ == '__main__':
    main(sys.argv[1])
