import csv,sys,subprocess
rep=sys.argv[1]; filt=sys.argv[2] if len(sys.argv)>2 else ''
out=subprocess.run(['ncu','-i',rep,'--page','raw','--csv'],capture_output=True,text=True).stdout
r=list(csv.reader(out.splitlines()))
h=r[0]
want=['gpu__time_duration.sum','dram__bytes_read.sum','dram__bytes_write.sum','gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed','sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active','l1tex__m_xbar2l1tex_read_bytes.sum','lts__t_sector_hit_rate.pct','sm__cycles_active.avg','smsp__inst_executed.sum','sm__throughput.avg.pct_of_peak_sustained_elapsed','launch__registers_per_thread','sm__warps_active.avg.pct_of_peak_sustained_active','launch__grid_size','smsp__warp_issue_stalled_long_scoreboard_per_warp_active.pct','smsp__issue_active.avg.pct_of_peak_sustained_active','l1tex__t_sector_hit_rate.pct','smsp__inst_executed_op_local_ld.sum','smsp__inst_executed_op_local_st.sum','lts__throughput.avg.pct_of_peak_sustained_elapsed','l1tex__throughput.avg.pct_of_peak_sustained_elapsed']
ki=h.index('Kernel Name')
for row in r[2:]:
    if filt and filt not in row[ki]: continue
    print('==',row[ki][:70])
    for w in want:
        if w in h: print('   %-75s %s %s'%(w,row[h.index(w)],r[1][h.index(w)]))
