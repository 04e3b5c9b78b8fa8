"""Summarise an ncu report: python scratch/ncu_summary.py rep.ncu-rep [units_per_launch]  (development probe)"""
import csv, subprocess, sys, collections, io
rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, vals = rows[0], rows[1], rows[2]
keys = ['gpu__time_duration.sum', 'smsp__inst_executed.sum', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'smsp__issue_active.avg.pct_of_peak_sustained_active', 'launch__registers_per_thread', 'launch__occupancy_limit',
        'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum', 'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum',
        'dram__bytes_read.sum', 'dram__bytes_write.sum', 'launch__shared_mem_per_block_dynamic', 'launch__grid_size', 'launch__block_size',
        'sm__inst_executed_pipe_fma', 'sm__inst_executed_pipe_alu', 'sm__inst_executed_pipe_lsu', 'sm__inst_executed_pipe_xu', 'sm__pipe_fma_cycles_active', 'sm__throughput.avg.pct']
for i, h in enumerate(hdr):
    if any(k in h for k in keys) and 'per_second' not in h or ('issue_stalled' in h and 'ratio' in h and 'not_issued' not in h and float(vals[i] or 0) > 0.05):
        print(f"{h:90s} {units[i]:14s} {vals[i]}")
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
hdr = rows[1]; data = rows[2:]
isrc = hdr.index('Source'); iex = hdr.index('Instructions Executed'); isamp = hdr.index('# Samples')
tot = sum(int(r[iex]) for r in data)
op = collections.Counter()
for r in data:
    t = r[isrc].split()
    o = t[1] if t[0].startswith('@') else t[0]
    op[o.split('.')[0]] += int(r[iex])
print("total warp instructions", tot)
print("  ".join(f"{k}:{v/tot*100:.1f}%" for k, v in op.most_common(18)))
if len(sys.argv) > 2:
    print("warp instructions per unit:", tot / float(sys.argv[2]))
