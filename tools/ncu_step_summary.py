"""Per-kernel table from an .ncu-rep that holds one whole training step (tools/ncu_step.py).
usage: ncu_step_summary.py <rep> [hbm_gbs]   (reads MEASURED_PEAKS.json for the roofline denominator)"""
import csv, io, json, os, subprocess, sys
rep = sys.argv[1]
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
peaks = json.load(open(os.path.join(root, "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(root, "MEASURED_PEAKS.json")) else {}
hbm = float(sys.argv[2]) if len(sys.argv) > 2 else float(peaks.get("hbm_gbs", 6650.0))
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr = rows[0]
ci = {h: i for i, h in enumerate(hdr)}
def val(r, k):
    try:
        return float(r[ci[k]].replace(",", ""))
    except (KeyError, ValueError):
        return float("nan")
print(f"# per-kernel ncu summary of one training step ({os.path.basename(rep)}); HBM denominator {hbm:.0f} GB/s "
      f"({'measured copy, MEASURED_PEAKS.json' if peaks else 'fallback'}); times are ncu's serialised, cold-cache replays")
print(f"{'kernel':58s} {'us':>8s} {'rd MB':>8s} {'wr MB':>8s} {'GB/s':>7s} {'%HBM':>6s} {'dram%':>6s} {'tensor%':>8s} {'sm%':>6s} {'IPC':>5s} {'regs':>5s} {'smemKB':>7s} grid x block")
tot_us = tot_b = 0.0
for r in rows[2:]:
    name = r[ci["Kernel Name"]]
    us = val(r, "gpu__time_duration.sum") / 1e3
    rd, wr = val(r, "dram__bytes_read.sum"), val(r, "dram__bytes_write.sum")
    # ncu scales units per column; the units row tells which
    u_rd, u_wr, u_t = rows[1][ci["dram__bytes_read.sum"]], rows[1][ci["dram__bytes_write.sum"]], rows[1][ci["gpu__time_duration.sum"]]
    scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    rd *= scale.get(u_rd, 1); wr *= scale.get(u_wr, 1)
    us = val(r, "gpu__time_duration.sum") * {"ns": 1e-3, "us": 1, "ms": 1e3, "s": 1e6}.get(u_t, 1e-3)
    gbs = (rd + wr) / us / 1e3 if us > 0 else 0
    tot_us += us; tot_b += rd + wr
    smem = val(r, "launch__shared_mem_per_block_dynamic") * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6}.get(rows[1][ci.get("launch__shared_mem_per_block_dynamic", 0)], 1) / 1024 if "launch__shared_mem_per_block_dynamic" in ci else float("nan")
    print(f"{name[:58]:58s} {us:8.1f} {rd / 1e6:8.1f} {wr / 1e6:8.1f} {gbs:7.0f} {100 * gbs / hbm:6.1f} "
          f"{val(r, 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed'):6.1f} "
          f"{val(r, 'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed'):8.1f} "
          f"{val(r, 'sm__throughput.avg.pct_of_peak_sustained_elapsed'):6.1f} "
          f"{val(r, 'sm__inst_executed.avg.per_cycle_elapsed'):5.2f} {val(r, 'launch__registers_per_thread'):5.0f} {smem:7.1f} "
          f"{r[ci['Grid Size']] if 'Grid Size' in ci else ''} x {r[ci['Block Size']] if 'Block Size' in ci else ''}")
print(f"{'TOTAL':58s} {tot_us:8.1f} {'':8s} {'':8s} {tot_b / tot_us / 1e3:7.0f} {100 * tot_b / tot_us / 1e3 / hbm:6.1f}   "
      f"({tot_b / 1e9:.2f} GB of HBM traffic -> {tot_b / hbm / 1e6:.2f} ms at the HBM roofline)")
