"""Prints the metrics we read from an ncu report (raw page csv): python tools/ncu_summary.py rep.ncu-rep [kernel-regex]"""
import csv, re, subprocess, sys
raw = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines())); hdr, units = rows[0], rows[1]
pat = re.compile(sys.argv[2]) if len(sys.argv) > 2 else None
WANT = ['gpu__time_duration.sum', 'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'smsp__issue_active.avg.pct_of_peak_sustained_active',
        'l1tex__throughput.avg.pct_of_peak_sustained_active', 'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum', 'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum',
        'dram__bytes_read.sum', 'dram__bytes_write.sum', 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'lts__throughput.avg.pct_of_peak_sustained_elapsed',
        'lts__t_sectors.sum', 'sm__warps_active.avg.pct_of_peak_sustained_active', 'smsp__inst_executed.sum', 'smsp__thread_inst_executed_per_inst_executed.ratio',
        'launch__registers_per_thread', 'launch__occupancy_limit_shared_mem', 'launch__occupancy_limit_registers', 'launch__grid_size']
seen = set()
for r in rows[2:]:
    name = r[hdr.index('Kernel Name')]
    short = name.split('(')[0]
    if (pat and not pat.search(name)) or short in seen: continue
    seen.add(short)
    print('----', short[:100])
    for w in WANT:
        if w in hdr: i = hdr.index(w); print(f"  {w} [{units[i]}] = {r[i][:40]}")
    st = [(float(r[i]), h) for i, h in enumerate(hdr) if 'issue_stalled' in h and h.endswith('per_issue_active.ratio') and 'average_warps_' in h and r[i]]
    for v, h in sorted(st, reverse=True)[:7]: print(f"  stall {h.split('issue_stalled_')[1].split('_per_issue')[0]} = {v:.2f}")
