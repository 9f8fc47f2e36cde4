import csv,sys,subprocess,io
out=subprocess.run(["ncu","-i",sys.argv[1],"--page","raw","--csv"],capture_output=True,text=True).stdout
rows=list(csv.reader(io.StringIO(out))); h=rows[0]
keys=['Kernel Name','gpu__time_duration.sum','dram__bytes_read.sum','dram__bytes_write.sum','dram__throughput.avg.pct_of_peak_sustained_elapsed','lts__t_sector_hit_rate.pct','l1tex__t_sector_hit_rate.pct','sm__warps_active.avg.pct_of_peak_sustained_active','smsp__issue_active.avg.pct_of_peak_sustained_active','l1tex__throughput.avg.pct_of_peak_sustained_elapsed','lts__throughput.avg.pct_of_peak_sustained_elapsed','smsp__inst_executed.sum','launch__registers_per_thread','launch__occupancy_limit_registers','launch__occupancy_limit_shared_mem','lts__t_sectors_srcunit_tex_op_read.sum','lts__t_sectors_srcunit_tex_op_write.sum','l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum']
seen=set()
for r in rows[2:]:
    name=r[h.index('Kernel Name')]
    if name in seen: continue
    seen.add(name)
    print('----')
    for k in keys:
        if k in h: print(f"{k:60s} {r[h.index(k)][:80]}")
    st=[(float(r[i]),k.replace('smsp__average_warps_issue_stalled_','').replace('_per_issue_active.ratio','')) for i,k in enumerate(h) if 'issue_stalled' in k and 'per_issue_active' in k and r[i]]
    print('stalls/issue:', ', '.join(f"{k} {v:.2f}" for v,k in sorted(st,reverse=True)[:8]))
