"""Sort-first split simulated on ONE GPU: renders the frame of every rank r of N in turn (its row bands of the forward
pass, its row slab of the shadow pass, per-draw culling on) and prints the device time per rank; the slowest rank bounds
the N-GPU step (the all-gather of the shadow slabs comes on top).  Saves 8-GPU box time while tuning the split.
    python tools/tiles_sim.py C4 8 [bands_per_rank ...]"""
import sys
sys.path.insert(0, '.')
import numpy as np, torch
from renderengine_b200 import capi, parallel, scene as sc
cfg = sys.argv[1] if len(sys.argv) > 1 else "C4"
N = int(sys.argv[2]) if len(sys.argv) > 2 else 8
bprs = [int(a) for a in sys.argv[3:]] or [1, 2, 4]
s = sc.make_scene(cfg)
S, H = s["shadow_size"], s["height"]
st = torch.cuda.Stream()
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
rows = parallel.slab_rows(S, N)
for bpr in bprs:
    band = parallel.contiguous_band(H, N * bpr) if bpr > 0 else 64
    per_rank = []
    for r in range(N):
        ctx = capi.Context(s["width"], s["height"], stream=st.cuda_stream, outputs=capi.REB_OUT_COLOR_SRGB8 | capi.REB_OUT_DEPTH_F32 | capi.REB_OUT_TRIID_U32,
                           rank=r, n_ranks=N, band_h=band)
        ctx.upload_scene(s)
        if rows:
            ctx.set_shadow_partition(r, N, rows)
        with torch.cuda.stream(st):
            def frame():
                ctx.render_shadow(s["draws"]); ctx.render_forward(s["draws"])
            for _ in range(3): flush.add_(1); frame()
            ctx.sync()
            ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(8)]
            for a, b in ev:
                flush.add_(1); a.record(st); frame(); b.record(st)
            ctx.sync()
            ts = sorted(a.elapsed_time(b) for a, b in ev)
            ctx.set_profiling(True)
            acc = {}
            for _ in range(4):
                flush.add_(1); frame(); ctx.sync()
                for k in (0, 1):
                    for n, v in ctx.kernel_times_ms(k).items():
                        acc[("sh." if k == 0 else "fw.") + n] = acc.get(("sh." if k == 0 else "fw.") + n, 0) + v / 4
        per_rank.append(1e3 * ts[len(ts) // 2])
        print(f"  bands/rank {bpr} band_h {band} rank {r}: {per_rank[-1]:.0f} us   " + " ".join(f"{k[:6]}={1e3*v:.0f}" for k, v in acc.items() if v > 4e-3))
        ctx.close(); del ctx
    print(f"bands/rank {bpr} band_h {band}: max {max(per_rank):.0f} us  mean {np.mean(per_rank):.0f} us")
