import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
h = rows[0]
keys = ['gpu__time_duration.sum', 'smsp__issue_active.avg.pct_of_peak_sustained_active', 'sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active',
        'sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_elapsed', 'smsp__inst_executed.sum', 'launch__registers_per_thread',
        'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum', 'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum']
for r in rows[2:]:
    print(r[h.index('Kernel Name')][:70])
    for k in keys:
        if k in h:
            print('  ', k, r[h.index(k)])
    st = [(c.replace('smsp__average_warps_issue_stalled_', '').replace('_per_issue_active.ratio', ''), float(r[i])) for i, c in enumerate(h)
          if c.startswith('smsp__average_warps_issue_stalled_') and c.endswith('_per_issue_active.ratio')]
    st.sort(key=lambda x: -x[1])
    print('   stalls:', ', '.join(f'{a} {b:.2f}' for a, b in st[:9]))
