"""Per-launch summary of an ncu report: duration, DRAM bytes, IPC, occupancy, registers, lane utilisation.
    python tools/ncu_summary.py gpurun_out/prof.ncu-rep"""
import csv, subprocess, sys, io
txt = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(txt)))
h = rows[0]; idx = {k: i for i, k in enumerate(h)}
cols = [("Kernel Name", "kernel", 34), ("gpu__time_duration.sum", "us", 8), ("dram__bytes_read.sum", "rdMB", 8), ("dram__bytes_write.sum", "wrMB", 8),
        ("smsp__inst_executed.sum", "winst", 10), ("sm__inst_executed.avg.per_cycle_elapsed", "ipc", 6),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "occ%", 6), ("launch__registers_per_thread", "regs", 5),
        ("smsp__thread_inst_executed_per_inst_executed.ratio", "lanes", 6), ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram%", 6),
        ("lts__t_sector_hit_rate.pct", "l2hit", 6), ("l1tex__t_sector_hit_rate.pct", "l1hit", 6)]
print(" ".join(f"{n:>{w}}" for _, n, w in cols))
for r in rows[2:]:
    out = []
    for k, n, w in cols:
        v = r[idx[k]] if k in idx else "-"
        if k == "Kernel Name": v = v.replace("void ", "")[:w]
        else:
            try: v = f"{float(v.replace(',', '')):.2f}" if "." in v else v
            except ValueError: pass
        out.append(f"{v:>{w}}")
    print(" ".join(out))
