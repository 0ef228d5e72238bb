"""Summarise an .ncu-rep (read on the CPU box): key raw metrics + top stall sites from the source page."""
import csv, subprocess, sys, io
rep = sys.argv[1]
top_n = int(sys.argv[2]) if len(sys.argv) > 2 else 14
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, vals = rows[0], rows[1], rows[2]
d = {h: (v, u) for h, u, v in zip(hdr, units, vals)}
keys = ['gpu__time_duration.sum', 'sm__cycles_elapsed.avg', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'sm__throughput.avg.pct_of_peak_sustained_elapsed',
        'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed', 'sm__inst_executed.avg.per_cycle_elapsed',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'launch__registers_per_thread', 'smsp__inst_executed.sum',
        'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum', 'smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio', 'smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio', 'smsp__average_warps_issue_stalled_wait_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio', 'smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio']
for k in keys:
    if k in d:
        print(f"{k:95s} {d[k][0]:>18s} {d[k][1]}")
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
h, data = rows[1], rows[2:]
ci = {n: i for i, n in enumerate(h)}
num = lambda x: float(x) if x.replace('.', '', 1).isdigit() else 0.0
tot = sum(num(r[ci['# Samples']]) for r in data) or 1
print(f"\n# top warp-stall sample sites (share of {int(tot)} samples | executions | dominant stalls | SASS)")
for r in sorted(data, key=lambda r: -num(r[ci['# Samples']]))[:top_n]:
    st = {k[6:]: num(r[ci[k]]) for k in h if k.startswith('stall_') and '(' not in k}
    dom = ",".join(f"{k}={int(v)}" for k, v in sorted(st.items(), key=lambda kv: -kv[1])[:2] if v > 0)
    print(f"{num(r[ci['# Samples']]) / tot * 100:5.1f}%  exec={r[ci['Instructions Executed']]:>9s}  {dom:34s} {r[ci['Source']][:90]}")
