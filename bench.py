#!/usr/bin/env python
"""bench.py — headline benchmark of the raster hot path (BASELINE.json metric: Mtri/s and frames/s,
Phong + 3x3-PCF shadow, 1080p; achieved HBM GB/s as a fraction of the measured B200 peak).

A "step" is one full frame of config C2 (1,003,522 triangles, 1920x1080, 2048^2 shadow
map): shadow depth pass + forward pass, every step.  N > 1 (torchrun): view-parallel batch — every rank
renders a full frame (its own shadow pass + its own camera view) per step, no data-path collective (weak
scaling; `--shadow-share broadcast` renders the shadow map on rank 0 only and NCCL-broadcasts it over NVLink
instead, which pays off when many views share one map).  `--mode tiles` runs the sort-first row-band split of
ONE frame (strong scaling): shadow map rendered once on rank 0 and NCCL-broadcast, as BASELINE.json config 4 asks.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--config C2] [--mode views|tiles]
                    [--shadow-share replicate|broadcast]
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

from renderengine_b200 import scene as sc  # noqa: E402

METRIC, UNIT = "Mtri/s", "Mtri/s"


def profiled_traffic(kind, slot):
    """dram__bytes_read+write per launch of the named kernel from the newest committed ncu capture (profiles/*_traffic.json)."""
    import glob
    import re
    name = {"tile_stage": "k_tile_forward" if kind == 1 else "k_tile_shadow",
            "setup_raster": "k_setup_raster<%s>" % ("forward" if kind == 1 else "shadow"),
            "vertex": "k_vertex<%s>" % ("forward" if kind == 1 else "shadow")}.get(slot)

    def natural(path):   # r1_v10 after r1_v9
        return [int(t) if t.isdigit() else t for t in re.split(r"(\d+)", os.path.basename(path))]
    for path in sorted(glob.glob(os.path.join(ROOT, "profiles", "*_traffic.json")), key=natural, reverse=True):
        try:
            with open(path) as f:
                t = json.load(f)["dram_bytes_per_launch"]
            if slot == "vertex" and kind == 0 and "k_vertex_both" in t:   # fused two-pass vertex kernel (timed in the shadow slot)
                return int(t["k_vertex_both"]), os.path.basename(path)
            if name in t:
                return int(t[name]), os.path.basename(path)
        except Exception:
            pass
    return None, None


def measured_peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region (B200_PROFILING.md)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100",
                                          "-i", str(self.index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=lambda: self.lines.extend(self.proc.stdout), daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        self.t.join(timeout=2)
        sm, mx, reasons = [], None, set()
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx = float(f[2])
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons), "samples": len(sm)}


def cpu_reference_frames(scene, n_frames, warm=1, keep=None):
    """Times the CPU oracle (the port of the reference path) on all host cores: full frames of the same workload.
    `keep` (a dict) receives the last frame's buffers so the caller can compare them with the device's."""
    from oracle import orc
    o = orc.Oracle(scene)
    cores = os.cpu_count() or 1
    times = []
    for i in range(warm + n_frames):
        t0 = time.perf_counter()
        sm = o.render_shadow() if scene["shadow_size"] else None
        out = o.render_forward(sm)
        if i >= warm:
            times.append(time.perf_counter() - t0)
    if keep is not None:
        keep.update(out); keep["shadow"] = sm
    return times, cores


def build_scene(args):
    scene = sc.make_scene(args.config)
    if args.shading == "pbr":
        scene = sc.to_pbr(scene)
    if getattr(args, "translucent_every", 0) > 0:
        scene = sc.to_translucent(scene, every=args.translucent_every)
    return scene


def run_reference(args, rank, world):
    if rank != 0:
        return
    scene = build_scene(args)
    budget_s = 150.0
    probe, cores = cpu_reference_frames(scene, 1, warm=1)
    steps = max(1, min(args.steps, int(budget_s / max(probe[0], 1e-3))))
    warm = min(args.warmup, 2)
    times, cores = cpu_reference_frames(scene, steps, warm=warm)
    ms = 1e3 * float(np.mean(times))
    val = scene["n_triangles"] / (ms * 1e3)
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": world, "steps": steps, "warmup": warm,
        "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "frames_per_s": 1e3 / ms,
        "config": {"workload": f"{args.config}: {scene['n_triangles']} tris, {scene['width']}x{scene['height']}, shadow {scene['shadow_size']}^2 3x3 PCF, full frame (shadow + forward) per step" + (", PBR fragment path (3 lights)" if args.shading == "pbr" else ", Phong")},
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": cores, "kind": "port",
                         "sample": f"{steps} full frames of the same workload on the CPU oracle (reference Vulkan/lavapipe path unavailable on this box)"},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))


class _DevArray:
    def __init__(self, ptr, shape, typestr):
        self.__cuda_array_interface__ = {"data": (ptr, False), "shape": shape, "typestr": typestr, "version": 2}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default="C2")
    ap.add_argument("--mode", default="views", choices=["views", "tiles"])
    ap.add_argument("--shadow-share", default=None, choices=["replicate", "broadcast", "slabs"],
                    help="N > 1: every rank renders the map (replicate), rank 0 renders + NCCL broadcast, or every rank renders a 1/N row slab + NCCL all-gather")
    ap.add_argument("--bands-per-rank", type=int, default=2,
                    help="--mode tiles: row bands per rank, interleaved (1 = one contiguous band: least geometry per rank after per-draw "
                         "culling but unbalanced; 0 = 64-row bands: balanced, nothing to cull)")
    ap.add_argument("--shading", default="phong", choices=["phong", "pbr"],
                    help="phong = BASELINE.json's path (default); pbr = the reference's live surface_pbr.frag path (SURVEY 8f row 1)")
    ap.add_argument("--samples", type=int, default=1, choices=[1, 4],
                    help="samples per pixel of the base pass (4 = Appendix B.15 multisampling; the headline metric is quoted at 1)")
    ap.add_argument("--translucent-every", type=int, default=0,
                    help="make every K-th material AlphaMode::eBlend (alpha 0.45): its draws go through the translucent pass (SURVEY 8f row 3); 0 = all opaque (headline)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup

    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return

    import torch
    import torch.distributed as dist
    from renderengine_b200 import capi, parallel

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the product has no CPU path (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    scene = build_scene(args)
    T, W, H, S = scene["n_triangles"], scene["width"], scene["height"], scene["shadow_size"]
    tiles = args.mode == "tiles" and world > 1
    # measured: re-rendering the shadow map per rank beats render-on-0 + NCCL broadcast in both modes; in tiles mode the
    # row-slab split (each rank culls the draws outside its slab, one all-gather) beats both
    share = args.shadow_share or ("slabs" if tiles and parallel.slab_rows(S, world) else "replicate")
    if share == "slabs" and not (world > 1 and parallel.slab_rows(S, world)):
        share = "replicate"
    bcast = world > 1 and share == "broadcast"
    slabs = world > 1 and share == "slabs"
    band_h = parallel.contiguous_band(H, world * args.bands_per_rank) if (tiles and args.bands_per_rank > 0) else 64
    st = torch.cuda.Stream()
    outputs = capi.REB_OUT_COLOR_SRGB8 | capi.REB_OUT_DEPTH_F32 | capi.REB_OUT_TRIID_U32
    ctx = capi.Context(W, H, device=local_rank, stream=st.cuda_stream, outputs=outputs,
                       rank=rank if tiles else 0, n_ranks=world if tiles else 1, band_h=band_h)
    ctx.upload_scene(scene)
    if slabs:
        ctx.set_shadow_partition(rank, world, parallel.slab_rows(S, world))
    if args.samples != 1:
        ctx.set_sample_count(args.samples)
    if not tiles and world > 1:
        ctx.set_view(sc.view_for_azimuth(scene, 2 * np.pi * rank / world))   # one camera view per rank
    draws = scene["draws"]
    tl_draws = scene.get("translucent_draws")                    # drawn after the opaque list, blended (reb_render_forward2)
    shadow_draws = draws if tl_draws is None else np.concatenate([draws, tl_draws])   # RenderScene.cpp:191-208: both lists cast shadows
    shadow_t = None
    if S and (bcast or slabs):
        ptr, _ = ctx.output_ptr(capi.REB_SHADOW_F32)
        shadow_t = torch.as_tensor(_DevArray(ptr, (S, S), "<f4"), device=f"cuda:{local_rank}")

    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")   # > 126 MB L2

    def step():
        if S:
            if slabs:
                ctx.render_shadow(shadow_draws)
                ctx.flush()                                      # the recorded depth pass must be on the stream before the exchange
                parallel.allgather_shadow_slabs(shadow_t)        # NCCL over NVLink; ordered on the current stream
                ctx.shadow_mark_valid()
            elif not bcast:
                ctx.render_shadow(shadow_draws)
            else:
                if rank == 0:
                    ctx.render_shadow(shadow_draws)
                    ctx.flush()                                  # the recorded depth pass must be on the stream before the broadcast
                parallel.broadcast_shadow(shadow_t, src=0)   # NCCL over NVLink; ordered on the current stream
                if rank != 0:
                    ctx.shadow_mark_valid()
        ctx.render_forward(draws, tl_draws)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    with torch.cuda.stream(st):
        for _ in range(args.warmup):
            flush.add_(1); step()
        ctx.sync()
        barrier()
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
        sampler = ClockSampler(local_rank)
        if rank == 0:
            sampler.start()
        l0 = ctx.launch_count()
        for a, b in ev:
            flush.add_(1)                                 # L2 flush between timed iterations (outside the event pair)
            a.record(st); step(); b.record(st)
        ctx.sync()
        barrier()
        clocks = sampler.stop() if rank == 0 else None
        launches = ctx.launch_count() - l0
        per_frame = np.array([a.elapsed_time(b) for a, b in ev])
        total_ms = float(per_frame.sum())
        tt = torch.tensor([total_ms], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        ms = float(tt.item()) / args.steps
        frames_per_step = 1 if tiles else world
        value = frames_per_step * T / (ms * 1e3)

        # ---- forward-only frames with the cached shadow map: what the reference does after its first frame
        # (RenderScene.cpp:171-175 renders the shadow map once); SURVEY 8d asks for both figures ------------------------
        fwd_only = None
        if world == 1:
            evf = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(min(args.steps, 50))]
            for a, b in evf:
                flush.add_(1)
                a.record(st); ctx.render_forward(draws, tl_draws); b.record(st)
            ctx.sync()
            fwd_only = np.array([a.elapsed_time(b) for a, b in evf])

        # ---- per-kernel timing + roofline of the dominant kernel (separate, instrumented frames) ----------
        ctx.set_profiling(True)
        acc = [dict(), dict()]
        n_prof = 20
        for _ in range(n_prof):
            flush.add_(1)
            if S and (not bcast or rank == 0):
                ctx.render_shadow(shadow_draws)
            ctx.render_forward(draws, tl_draws)
            ctx.sync()
            for k in (0, 1):
                if k == 0 and not (S and (not bcast or rank == 0)):
                    continue
                for name, v in ctx.kernel_times_ms(k).items():
                    acc[k][name] = acc[k].get(name, 0.0) + v / n_prof
        ctx.set_profiling(False)
        counters = ctx.counters()

        # ---- end to end through the C ABI with host buffers (wall clock, includes D2H of the frame) -------
        # every frame: host view / light blocks in, the finished sRGB8 frame out to pinned host memory.  The target is a
        # two-image swapchain, so frame i's device->host copy (copy stream) overlaps frame i+1's rendering; a frame's
        # pixels are complete on the host after the sync that follows, all inside the timed region.
        pinned = [torch.empty((H, W, 4), dtype=torch.uint8).pin_memory() for _ in range(2)]
        view_host = np.array(scene["view"] if world == 1 or tiles else sc.view_for_azimuth(scene, 2 * np.pi * rank / world))
        n_e2e = max(10, min(args.steps, 100))
        for i in range(4):
            step(); ctx.readback_async_into(capi.REB_COLOR_SRGB8, pinned[i & 1].data_ptr())
        ctx.sync()
        barrier()
        t0 = time.perf_counter()
        for i in range(n_e2e):
            ctx.set_view(view_host)
            if S:
                ctx.shadow_set_transform(scene["shadow_view"], scene["shadow_proj"])
            step()
            ctx.readback_async_into(capi.REB_COLOR_SRGB8, pinned[i & 1].data_ptr())
        ctx.sync()
        barrier()
        e2e_ms = (time.perf_counter() - t0) * 1e3 / n_e2e
        te = torch.tensor([e2e_ms], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(te, op=dist.ReduceOp.MAX)
        e2e_ms = float(te.item())
    e2e_val = frames_per_step * T / (e2e_ms * 1e3)

    if rank == 0:
        peak, peak_src = measured_peak()
        b_shadow, b_forward = sc.algorithmic_bytes(scene)
        V = scene["n_vertices"]
        # algorithmic bytes owned by each kernel; their sum is B_shadow + B_forward of SURVEY.md §8d
        alg = {
            (0, "clear"): 0, (0, "vertex"): 12 * V, (0, "setup_raster"): 12 * T, (0, "bin_scatter"): 0, (0, "tile_stage"): 4 * S * S,
            (1, "vertex"): 32 * V, (1, "setup_raster"): 12 * T, (1, "bin_scatter"): 0,
            (1, "tile_stage"): (min(36 * W * H, 4 * S * S) if S else 0) + 12 * W * H,
        }
        if S and not bcast and acc[1].get("vertex", 1.0) < 0.2 * acc[0].get("vertex", 0.0):   # fused two-pass vertex kernel: timed in the shadow slot
            alg[(0, "vertex")] = 12 * V + 32 * V; alg[(1, "vertex")] = 0
        kt = {f"{'shadow' if k == 0 else 'forward'}.{n}": v for k in (0, 1) for n, v in acc[k].items()}
        dom = max(((k, n) for k in (0, 1) for n in acc[k]), key=lambda kn: acc[kn[0]][kn[1]])
        dom_ms = acc[dom[0]][dom[1]]
        dom_bytes = alg.get(dom, 0)
        achieved = dom_bytes / (dom_ms * 1e-3) / 1e9 if dom_ms > 0 else 0.0
        traffic, traffic_src = profiled_traffic(dom[0], dom[1])
        frame_bytes = (b_shadow + b_forward) * frames_per_step
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms, "higher_is_better": True, "scaling": "strong" if tiles else "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic", "frames_per_s": frames_per_step * 1e3 / ms,
            "config": {"workload": f"{args.config}: {T} tris / {V} verts, {W}x{H}, shadow {S}^2 3x3 PCF, full frame (shadow + forward) per step" + (", PBR fragment path (3 lights)" if args.shading == "pbr" else ", Phong") + (", 4x MSAA base pass" if args.samples == 4 else "") + (f", {len(tl_draws)} of {len(shadow_draws)} draws translucent" if tl_draws is not None else ""),
                       "parallelism": ("sort-first row bands x%d (band_h %d, per-draw culling), shadow map: %s" % (world, band_h, share)) if tiles else (("view-parallel x%d, " % world + ("shadow map rendered on rank 0 + NCCL broadcast" if bcast else "every rank renders its own shadow map (no collective)")) if world > 1 else "1 GPU"),
                       "l2": "256 MiB flush write between timed steps", "outputs": "sRGB8 + depth f32 + triID u32"},
            "hbm_frame": {"algorithmic_bytes": frame_bytes, "achieved_gbs": frame_bytes / (ms * 1e-3) / 1e9,
                          "frac_of_peak": frame_bytes / (ms * 1e-3) / 1e9 / peak / world, "peak_gbs": peak, "peak_source": peak_src},
            "roofline": {"bound": "hbm", "kernel": f"{'shadow' if dom[0] == 0 else 'forward'}.{dom[1]}", "achieved": achieved, "peak": peak,
                         "unit": "GB/s", "frac": achieved / peak, "traffic": traffic, "traffic_source": traffic_src, "algorithmic_bytes_per_launch": dom_bytes,
                         "kernel_ms": dom_ms, "peak_source": peak_src + " (MEASURED_PEAKS.json hbm_gbs, burst copy)"},
            "kernel_ms": kt,
            "e2e": {"value": e2e_val, "unit": UNIT, "ms_per_step": e2e_ms, "h2d_bytes_per_step": 160 + 128 + 2 * 16 * len(shadow_draws),
                    "d2h_bytes_per_step": W * H * 4, "note": "wall clock through reb_set_view/reb_render_*/reb_readback_async + final reb_sync with host buffers (pinned, 2-image swapchain: the copy of frame i overlaps frame i+1); geometry is static and resident as in the reference frame loop"},
            "gpu_launches": int(launches), "clocks": clocks,
            "frame_us": {"median": 1e3 * float(np.median(per_frame)), "p95": 1e3 * float(np.percentile(per_frame, 95)), "rank": 0},
            "counters": {"needs_clip": int(counters[0]), "binned_triangles": int(counters[1]), "tile_list_entries": int(counters[2]), "regrows": int(counters[3])},
        }
        if fwd_only is not None:
            fms = float(np.median(fwd_only))
            line["forward_only"] = {"ms_per_frame": fms, "value": T / (fms * 1e3), "unit": UNIT, "algorithmic_bytes": b_forward,
                                    "achieved_gbs": b_forward / (fms * 1e-3) / 1e9, "frac_of_peak": b_forward / (fms * 1e-3) / 1e9 / peak,
                                    "note": "forward pass alone against the cached shadow map (the reference renders the shadow map once, RenderScene.cpp:171-175)"}
        if world == 1 and not args.no_cpu_baseline and args.samples == 1 and tl_draws is None:
            keep = {}
            times, cores = cpu_reference_frames(scene, 5, warm=1, keep=keep)
            cms = 1e3 * float(np.mean(times))
            # the same frame on the device, compared bit for bit where the contract is bit-exact ("must hash-equal")
            ctx.set_view(np.array(scene["view"]))
            step(); ctx.sync()
            same = bool(np.array_equal(ctx.readback(capi.REB_DEPTH_F32).view(np.uint32), keep["depth"].view(np.uint32))
                        and np.array_equal(ctx.readback(capi.REB_TRIID_U32), keep["tri_id"])
                        and (not S or np.array_equal(ctx.readback(capi.REB_SHADOW_F32).view(np.uint32), np.asarray(keep["shadow"], np.float32).view(np.uint32))))
            d8 = np.abs(ctx.readback(capi.REB_COLOR_SRGB8).astype(np.int16) - keep["color_srgb8"].astype(np.int16))
            line["cpu_baseline"] = {"value": T / (cms * 1e3), "unit": UNIT, "cores": cores, "kind": "port", "ms_per_frame": cms,
                                    "sample": "5 full frames (shadow + forward) of the same workload on the CPU oracle, all host cores",
                                    "gpu_matches_bit_exact": {"depth_triid_shadow": same, "srgb8_max_lsb": int(d8.max())}}
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
