"""Aggregate an ncu report's per-instruction samples by CUDA source line.
    python tools/ncu_lines.py gpurun_out/prof.ncu-rep k_tile_forward [top]"""
import csv, subprocess, sys, io, os
rep, kern = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 30
extra = sys.argv[4:]   # e.g. --launch-skip 1 --launch-count 1 (applied after the name filter)
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass", "--kernel-name", "regex:" + kern] + extra,
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(txt)))
fname = ""; hdr = None; agg = {}; seen_launch = 0
tot_s = tot_i = 0
for r in rows:
    if not r: continue
    if r[0] == "File Path": fname = os.path.basename(r[1]); continue
    if r[0] == "Function Name": continue
    if r[0] == "Line No": hdr = r; si = hdr.index("# Samples"); ii = hdr.index("Instructions Executed"); ti = hdr.index("Thread Instructions Executed"); continue
    if hdr is None or len(r) <= ti: continue
    if r[2] != "-": continue   # only the per-source-line summary rows (Address == '-')
    try: s = int(r[si]); n = int(r[ii]); t = int(r[ti])
    except ValueError: continue
    k = (fname, r[0])
    a = agg.setdefault(k, [0, 0, 0, r[1].strip()[:100]])
    a[0] += s; a[1] += n; a[2] += t
    tot_s += s; tot_i += n
print(f"kernel {kern}: samples {tot_s}, warp-instructions {tot_i}")
for (f, ln), a in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    print(f"{100*a[0]/max(tot_s,1):5.1f}% smp {100*a[1]/max(tot_i,1):5.1f}% inst  lanes {a[2]/max(a[1],1):4.1f}  {f}:{ln}  {a[3]}")
