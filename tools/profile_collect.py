"""Summarises gpurun_out/<tag>_{launches.csv,full.ncu-rep} into profiles/<tag>_* (launch list, per-kernel summary with the
hottest source lines, selected raw metrics, DRAM traffic per launch for bench.py's roofline.traffic).
    python tools/profile_collect.py r1_v10"""
import csv, io, json, os, shutil, subprocess, sys

tag = sys.argv[1]
rep = f"gpurun_out/{tag}_full.ncu-rep"
shutil.copy(f"gpurun_out/{tag}_launches.csv", f"profiles/{tag}_launches.csv")
with open(f"profiles/{tag}_ncu_summary.txt", "w") as f:
    f.write(subprocess.run([sys.executable, "tools/ncu_summary.py", rep], capture_output=True, text=True).stdout)
    for k in ("k_tile_forward", "k_setup_raster", "k_tile_shadow", "k_vertex_both", "k_tile_binned"):
        out = subprocess.run([sys.executable, "tools/ncu_lines.py", rep, k, "18"], capture_output=True, text=True).stdout
        f.write("\n".join(line[:160] for line in out.splitlines()) + "\n")
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
ix = {k: i for i, k in enumerate(rows[0])}
sel = [s for s in ('Kernel Name', 'gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'smsp__inst_executed.sum',
                   'sm__inst_executed.avg.per_cycle_elapsed', 'sm__warps_active.avg.pct_of_peak_sustained_active', 'launch__registers_per_thread',
                   'lts__t_sector_hit_rate.pct', 'l1tex__t_sector_hit_rate.pct', 'sm__throughput.avg.pct_of_peak_sustained_elapsed',
                   'lts__throughput.avg.pct_of_peak_sustained_elapsed', 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed') if s in ix]
names = {'k_vertex_both': 'k_vertex_both', 'k_setup_raster<1': 'k_setup_raster<shadow>', 'k_setup_raster<0': 'k_setup_raster<forward>',
         'k_tile_binned<1>': 'k_tile_binned<shadow>', 'k_tile_binned<0>': 'k_tile_binned<forward>',
         'k_tile_shadow': 'k_tile_shadow', 'k_tile_forward': 'k_tile_forward'}
mul = {'Mbyte': 1e6, 'Kbyte': 1e3, 'byte': 1, 'Gbyte': 1e9}
t = {}
with open(f"profiles/{tag}_ncu_raw_selected.csv", "w", newline="") as f:
    w = csv.writer(f)
    w.writerow(sel)
    w.writerow([rows[1][ix[s]] for s in sel])
    for r in rows[2:]:
        w.writerow([r[ix[s]] for s in sel])
        key = [v for k, v in names.items() if k in r[ix['Kernel Name']]][0]
        rd = float(r[ix['dram__bytes_read.sum']].replace(',', '')) * mul[rows[1][ix['dram__bytes_read.sum']]]
        wr = float(r[ix['dram__bytes_write.sum']].replace(',', '')) * mul[rows[1][ix['dram__bytes_write.sum']]]
        t[key] = int(rd + wr)
json.dump({'source': f'ncu --set full, profiles/{tag}_ncu_raw_selected.csv (C2, one launch each, cold L2, tools/prof_frame.py)',
           'dram_bytes_per_launch': t}, open(f"profiles/{tag}_traffic.json", "w"), indent=1)
print(t)
if os.path.exists(f"gpurun_out/{tag}_bench.json"):
    shutil.copy(f"gpurun_out/{tag}_bench.json", f"profiles/{tag}_bench.json")
