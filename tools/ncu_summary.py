"""Prints a compact per-launch summary (duration, DRAM bytes, throughput, occupancy) of an .ncu-rep.
usage: python tools/ncu_summary.py report.ncu-rep > profiles/<name>.txt"""
import csv
import io
import subprocess
import sys

rep = sys.argv[1]
txt = subprocess.run(['ncu', '-i', rep, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(txt)))
H, U = rows[0], rows[1]
want = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'lts__t_bytes.sum',
        'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'sm__throughput.avg.pct_of_peak_sustained_elapsed',
        'smsp__issue_active.avg.pct_of_peak_sustained_active', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'smsp__inst_executed.sum', 'launch__registers_per_thread', 'launch__grid_size', 'launch__block_size',
        'launch__occupancy_limit_shared_mem', 'launch__occupancy_limit_registers',
        'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum']
ki = H.index('Kernel Name')
print(f'# {rep}')
for r in rows[2:]:
    print(f'\n{r[ki][:110]}')
    for w in want:
        if w in H:
            i = H.index(w)
            print(f'    {w:64s} {r[i]:>16s} {U[i]}')
