import csv,sys,subprocess
rep=sys.argv[1]
out=subprocess.run(['ncu','-i',rep,'--page','raw','--csv'],capture_output=True,text=True).stdout
rows=list(csv.reader(out.splitlines()))
hdr=rows[0]
want=['Kernel Name','gpu__time_duration.sum','launch__registers_per_thread','launch__occupancy_limit','sm__warps_active.avg.pct_of_peak_sustained_active','smsp__inst_executed.sum','smsp__issue_active.avg.pct_of_peak_sustained_active','sm__inst_executed_pipe_lsu.sum','l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed','l1tex__data_pipe_lsu_wavefronts_mem_shared.sum','l1tex__t_output_wavefronts_pipe_lsu_mem_global_op_ld.sum','l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum','l1tex__t_sector_hit_rate.pct','lts__t_sector_hit_rate.pct','dram__bytes_read.sum','dram__bytes_write.sum','gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed','l1tex__lsu_writeback_active','l1tex__throughput.avg.pct','lts__throughput.avg.pct','sm__throughput.avg.pct','smsp__thread_inst_executed_per_inst_executed.ratio','l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum','sm__pipe_fma_cycles_active.avg.pct','sm__inst_executed_pipe_alu','sm__inst_executed_pipe_fma','sm__inst_executed_pipe_xu','smsp__warp_issue_stalled']
for r in rows[2:]:
    for i,h in enumerate(hdr):
        if any(h.startswith(w) or w in h for w in want):
            if 'issue_stalled' in h and 'per_warp_active' not in h: continue
            if r[i] in ('','0'): continue
            print(h,'=',r[i],rows[1][i])
    print('-----')
