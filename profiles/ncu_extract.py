"""Print the handful of ncu metrics used in profiles/*.md from a `--page raw --csv` export:
   ncu -i X.ncu-rep --page raw --csv > x.csv ; python profiles/ncu_extract.py x.csv"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[0]
keys = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'sm__inst_executed.avg.per_cycle_active',
        'lts__t_sector_hit_rate.pct', 'smsp__inst_executed.sum', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum', 'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum',
        'launch__registers_per_thread', 'dram__throughput.avg.pct_of_peak_sustained_elapsed',
        'lts__t_bytes.sum', 'sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active']
name = hdr.index('Kernel Name')
for r in rows[2:]:
    print(r[name][:60])
    for k in keys:
        if k in hdr:
            print(f"   {k:70s} {r[hdr.index(k)]} {rows[1][hdr.index(k)]}")
    stalls = [(float(r[i]), h.replace('smsp__average_warps_issue_stalled_', '').replace('_per_issue_active.ratio', ''))
              for i, h in enumerate(hdr) if 'smsp__average_warps_issue_stalled' in h and h.endswith('_per_issue_active.ratio')]
    print("   stalls:", ", ".join(f"{n}={v:.2f}" for v, n in sorted(stalls, reverse=True)[:7]))
