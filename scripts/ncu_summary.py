"""Summarise an ncu report (--set full) per kernel: duration, registers, occupancy limits, pipe utilisation, DRAM bytes,
shared-memory conflicts and the top warp-stall reasons.  Usage: python scripts/ncu_summary.py gpurun_out/prof.ncu-rep"""
import csv,sys,subprocess,collections
rep=sys.argv[1]
# a .csv argument is the `ncu -i <rep> --page raw --csv` export itself (a report over 64 MiB does not travel back from the GPU box)
raw = open(rep).read() if rep.endswith('.csv') else subprocess.run(['ncu', '-i', rep, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
r=list(csv.reader(raw.splitlines()))
hdr=r[0]
want=['gpu__time_duration.sum','launch__registers_per_thread','launch__grid_size','launch__occupancy_limit_shared_mem','launch__occupancy_limit_registers','sm__warps_active.avg.pct_of_peak_sustained_active','smsp__issue_active.avg.pct_of_peak_sustained_active','sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active','sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active','sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active','sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_active','sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active','smsp__inst_executed.sum','dram__bytes_read.sum','dram__bytes_write.sum','lts__t_sector_hit_rate.pct','l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum','l1tex__data_pipe_lsu_wavefronts_mem_shared.sum','gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed','smsp__inst_executed_op_local_ld.sum' ]
stall=[h for h in hdr if h.startswith('smsp__average_warps_issue_stalled') and h.endswith('per_issue_active.ratio')]
for row in r[2:]:
    print(row[hdr.index('Kernel Name')][:120])
    for w in want:
        if w in hdr: print('  ',w.ljust(66), row[hdr.index(w)], r[1][hdr.index(w)])
    st=sorted([(float(row[hdr.index(h)]),h.replace('smsp__average_warps_issue_stalled_','').replace('_per_issue_active.ratio','')) for h in stall],reverse=True)[:8]
    print('   stalls:', ', '.join('%s %.2f'%(n,v) for v,n in st))
