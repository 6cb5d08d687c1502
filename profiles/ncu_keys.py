"""Print the metrics we track from an ncu report: python profiles/ncu_keys.py gpurun_out/x.ncu-rep [launch index]"""
import csv
import subprocess
import sys

KEYS = ['Kernel Name', 'Grid Size', 'Block Size', 'gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'lts__t_bytes.sum', 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed',
        'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum',
        'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'launch__registers_per_thread', 'launch__occupancy_limit_registers', 'launch__occupancy_limit_shared_mem',
        'launch__waves_per_multiprocessor', 'smsp__inst_executed.sum',
        'sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active',
        'smsp__issue_active.avg.pct_of_peak_sustained_active', 'sm__cycles_active.avg', 'sm__cycles_elapsed.max',
        'l1tex__t_sector_hit_rate.pct', 'lts__t_sector_hit_rate.pct', 'gpc__cycles_elapsed.avg.per_second',
        'l1tex__data_pipe_lsu_wavefronts.sum', 'sm__inst_executed_pipe_lsu.sum']
STALLS = 'smsp__average_warps_issue_stalled_'
rep = sys.argv[1]
rows = list(csv.reader(subprocess.run(['ncu', '-i', rep, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout.splitlines()))
hdr, units = rows[0], rows[1]
sel = rows[2:] if len(sys.argv) < 3 else [rows[2 + int(sys.argv[2])]]
for r in sel:
    for k in KEYS:
        if k in hdr:
            i = hdr.index(k)
            print(f'{k:70s} {r[i]:>24s} {units[i]}')
    st = [(float(r[i]), h[len(STALLS):-len('_per_issue_active.ratio')]) for i, h in enumerate(hdr)
          if h.startswith(STALLS) and h.endswith('_per_issue_active.ratio') and r[i]]
    print('stalls (warps per issue-active):', ', '.join(f'{n}={v:.2f}' for v, n in sorted(st, reverse=True)[:8]))
    print('---')
