#!/usr/bin/env python
"""bench.py — headline benchmark of the raster hot path (BASELINE.json metric: Mtri/s and frames/s, Phong + 3x3-PCF shadow,
1080p / 4K; achieved HBM GB/s as a fraction of the measured B200 peak).  One JSON line on stdout (rank 0).

N = 1 (default): a "step" is one full frame of config C2 (1,003,522 triangles, 1920x1080, 2048^2 shadow map):
    shadow depth pass + forward pass, every step.  The line also carries sub-records measured in the same invocation:
    `c1`, `c3` (BASELINE configs 1 and 3: 720p / 4K), `lowpoly` (the C2 grid at 16 k triangles with the reference's default 4096^2 map:
    the binned tile path), `view_batch` (config 5's per-GPU share: 64 views of the 2 M-triangle scene against ONE shadow map) and the
    host-copy ceiling of the end-to-end figure.
N > 1 (torchrun, one rank per GPU): the two splits BASELINE.json names, in one invocation:
    headline  = the VIEW-PARALLEL batch as config 5 defines it: the 2 M-triangle scene, `--views-per-rank` (64) camera views per rank
                and step, the shadow map rendered ONCE per step on rank 0 and NCCL-broadcast over NVLink; the broadcast runs on its
                own stream and only the first view's tile stage waits for it (reb_shadow_wait_event), so that view's vertex stage
                and triangle setup overlap the transfer.  Weak scaling (views per rank fixed).  `view_batch` repeats the figure next
                to the same rank-local workload WITHOUT any exchange, measured in the same run (the single-GPU rate).
    sort_first = config 4: ONE 7680x4320 frame of the 10.5 M-triangle scene split into row bands (per-draw / per-block culling, the
                shadow pass split between the ranks + one NCCL all-reduce MIN / all-gather, overlapped the same way) — ms / frame at N next to the
                unsplit frame on one GPU in the same run.
    comm      = bytes and CUDA-event time of the collectives; parity_n = same-run bit comparisons of the exchanged / composed buffers.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--config C2] [--mode auto|replicas|tiles] ...
"""
import argparse
import hashlib
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


# ---------------------------------------------------------------------------------------------------------------------------
# small helpers
# ---------------------------------------------------------------------------------------------------------------------------
def profiled_traffic(kind, slot):
    """dram__bytes_read+write per launch of the named kernel from the newest committed ncu capture (profiles/*_traffic.json)."""
    import glob
    import re
    name = {"tile_stage": "k_tile_forward" if kind == 1 else "k_tile_shadow",
            "setup_raster": "k_setup_raster<%s>" % ("forward" if kind == 1 else "shadow"),
            "vertex": "k_vertex<%s>" % ("forward" if kind == 1 else "shadow")}.get(slot)

    def natural(path):   # r1_v10 after r1_v9, r2_* after r1_*
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


class _DevArray:
    def __init__(self, ptr, shape, typestr):
        self.__cuda_array_interface__ = {"data": (ptr, False), "shape": shape, "typestr": typestr, "version": 2}


def cpu_reference_frames(scene, n_frames, warm=1, keep=None, shadow_every_frame=True):
    """Times the CPU oracle (the port of the reference path) on all host cores: full frames of the same workload.
    `keep` (a dict) receives the last frame's buffers so the caller can compare them with the device's."""
    from oracle import orc
    o = orc.Oracle(scene)
    cores = os.cpu_count() or 1
    times = []
    sm = None
    for i in range(warm + n_frames):
        t0 = time.perf_counter()
        if scene["shadow_size"] and (shadow_every_frame or sm is None):
            sm = o.render_shadow()
        out = o.render_forward(sm)
        if i >= warm:
            times.append(time.perf_counter() - t0)
    if keep is not None:
        keep.update(out); keep["shadow"] = sm
    return times, cores


def build_scene(args, config=None):
    scene = sc.make_scene(config or args.config)
    if args.shading == "pbr":
        scene = sc.to_pbr(scene)
    if getattr(args, "translucent_every", 0) > 0:
        scene = sc.to_translucent(scene, every=args.translucent_every)
    return scene


def workload_string(args, world, scene, mode):
    """ONE description of the timed workload, shared by both arms (`--impl ours|reference`) so their lines compare."""
    T, V, W, H, S = scene["n_triangles"], scene["n_vertices"], scene["width"], scene["height"], scene["shadow_size"]
    tail = (", PBR fragment path (3 lights)" if args.shading == "pbr" else ", Phong") + (", 4x MSAA base pass" if args.samples == 4 else "")
    if scene.get("translucent_draws") is not None:
        tail += f", {len(scene['translucent_draws'])} of {len(scene['draws']) + len(scene['translucent_draws'])} draws translucent"
    if mode == "views":
        return (f"C5 view batch: {T} tris / {V} verts, {W}x{H} per view, {args.views_per_rank} views per GPU and step, shadow {S}^2 3x3 PCF rendered once per step" + tail)
    if mode == "tiles":
        return f"C4 sort-first: {T} tris / {V} verts, ONE {W}x{H} frame split into row bands, shadow {S}^2 3x3 PCF rendered once (shared between the ranks)" + tail
    return f"{scene['config']}: {T} tris / {V} verts, {W}x{H}, shadow {S}^2 3x3 PCF, full frame (shadow + forward) per step" + tail


def pin_to_gpu_numa(index):
    """Run this process (and so its pinned allocations, first-touched by the caller) on the CPUs local to GPU `index`."""
    info = {"numa_node": None, "cpus": None}
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(index)
        bus = pynvml.nvmlDeviceGetPciInfo(h).busId
        bus = bus.decode() if isinstance(bus, bytes) else bus
        dev = "/sys/bus/pci/devices/" + bus.lower()[-12:]
        with open(dev + "/numa_node") as f:
            info["numa_node"] = int(f.read())
        with open(dev + "/local_cpulist") as f:
            cl = f.read().strip()
        cpus = set()
        for part in cl.split(","):
            a, _, b = part.partition("-")
            cpus.update(range(int(a), int(b or a) + 1))
        cpus &= os.sched_getaffinity(0)
        if cpus:
            os.sched_setaffinity(0, cpus)
            info["cpus"] = cl
    except Exception as e:   # no NVML / sysfs: leave the affinity alone
        info["error"] = type(e).__name__
    return info


def resolve_mode(args, world):
    if args.mode != "auto":
        return args.mode
    return "views" if world > 1 else "frame"


# ---------------------------------------------------------------------------------------------------------------------------
# the reference arm: the reference's CPU path (its port, oracle/) on the box's host cores, rank 0 only
# ---------------------------------------------------------------------------------------------------------------------------
def run_reference(args, rank, world):
    if rank != 0:
        return
    mode = resolve_mode(args, world)
    budget_s = 150.0
    if mode == "views":
        scene = build_scene(args, "C5")
        from oracle import orc
        o = orc.Oracle(scene)
        cores = os.cpu_count() or 1
        t0 = time.perf_counter(); sm = o.render_shadow(); t_shadow = time.perf_counter() - t0
        t0 = time.perf_counter(); o.render_forward(sm); t_view = time.perf_counter() - t0
        n = max(1, min(args.steps * args.views_per_rank, int((budget_s - t_shadow) / max(t_view, 1e-3))))
        t0 = time.perf_counter()
        sm = o.render_shadow()
        for i in range(n):
            o.set_view(sc.view_for_azimuth(scene, 2 * np.pi * i / max(n, 1)))
            o.render_forward(sm)
        total = time.perf_counter() - t0
        steps, warm = 1, 1
        ms = 1e3 * total
        val = n * scene["n_triangles"] / (ms * 1e3)
        sample = f"one shadow pass + {n} camera views of the C5 scene on the CPU oracle (reference Vulkan/lavapipe path unavailable on this box)"
    else:
        scene = build_scene(args, "C4" if mode == "tiles" else None)
        probe, cores = cpu_reference_frames(scene, 1, warm=1)
        steps = max(1, min(args.steps, int(budget_s / max(probe[0], 1e-3))))
        warm = min(args.warmup, 2)
        times, cores = cpu_reference_frames(scene, steps, warm=warm)
        ms = 1e3 * float(np.mean(times))
        val = scene["n_triangles"] / (ms * 1e3)
        sample = f"{steps} full frames of the same workload on the CPU oracle (reference Vulkan/lavapipe path unavailable on this box)"
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": world, "steps": steps, "warmup": warm,
        "ms_per_step": ms, "higher_is_better": True, "scaling": "strong" if mode == "tiles" else "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": workload_string(args, world, scene, mode)},
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))


# ---------------------------------------------------------------------------------------------------------------------------
# device-side building blocks
# ---------------------------------------------------------------------------------------------------------------------------
class Rig:
    """One context + scene on this rank's GPU, on its own CUDA stream."""

    def __init__(self, scene, local_rank, rank=0, n_ranks=1, band_h=64, samples=1):
        import torch
        from renderengine_b200 import capi
        self.torch, self.capi, self.scene = torch, capi, scene
        self.st = torch.cuda.Stream()
        outputs = capi.REB_OUT_COLOR_SRGB8 | capi.REB_OUT_DEPTH_F32 | capi.REB_OUT_TRIID_U32
        self.ctx = capi.Context(scene["width"], scene["height"], device=local_rank, stream=self.st.cuda_stream, outputs=outputs,
                                rank=rank, n_ranks=n_ranks, band_h=band_h)
        self.ctx.upload_scene(scene)
        if samples != 1:
            self.ctx.set_sample_count(samples)
        self.draws = scene["draws"]
        self.tl = scene.get("translucent_draws")
        self.shadow_draws = self.draws if self.tl is None else np.concatenate([self.draws, self.tl])   # RenderScene.cpp:191-208: both lists cast shadows
        self.S = scene["shadow_size"]
        self.local_rank = local_rank

    def frame(self):
        if self.S:
            self.ctx.render_shadow(self.shadow_draws)
        self.ctx.render_forward(self.draws, self.tl)

    def forward(self):
        self.ctx.render_forward(self.draws, self.tl)

    def shadow_tensor(self):
        ptr, _ = self.ctx.output_ptr(self.capi.REB_SHADOW_F32)
        return self.torch.as_tensor(_DevArray(ptr, (self.S, self.S), "<f4"), device=f"cuda:{self.local_rank}")

    def timed(self, fn, steps, warmup, flush):
        """Device time of `steps` calls of fn (CUDA events on the context's stream, L2 flushed before each, outside the event pair)."""
        torch = self.torch
        with torch.cuda.stream(self.st):
            for _ in range(warmup):
                flush.add_(1); fn()
            self.ctx.sync()
            ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
            for a, b in ev:
                flush.add_(1)
                a.record(self.st); fn(); b.record(self.st)
            self.ctx.sync()
        return np.array([a.elapsed_time(b) for a, b in ev])

    def kernel_profile(self, flush, n_prof=20, shadow=True):
        """Per-kernel device time (instrumented frames: events between the kernels, so launches do not overlap here)."""
        torch = self.torch
        acc = [dict(), dict()]
        with torch.cuda.stream(self.st):
            self.ctx.set_profiling(True)
            for _ in range(n_prof):
                flush.add_(1)
                if self.S and shadow:
                    self.ctx.render_shadow(self.shadow_draws)
                self.ctx.render_forward(self.draws, self.tl)
                self.ctx.sync()
                for k in (0, 1):
                    if k == 0 and not (self.S and shadow):
                        continue
                    for name, v in self.ctx.kernel_times_ms(k).items():
                        acc[k][name] = acc[k].get(name, 0.0) + v / n_prof
            self.ctx.set_profiling(False)
        return acc

    def hashes(self, rows=None):
        """sha256 of the bit-exact outputs (depth + triangle IDs), optionally only of the given rows."""
        capi = self.capi
        d = self.ctx.readback(capi.REB_DEPTH_F32); t = self.ctx.readback(capi.REB_TRIID_U32)
        if rows is not None:
            d, t = d[rows], t[rows]
        h = hashlib.sha256(); h.update(np.ascontiguousarray(d).tobytes()); h.update(np.ascontiguousarray(t).tobytes())
        return h.hexdigest()

    def close(self):
        self.ctx.close()


def alg_table(scene, fused):
    """Algorithmic bytes owned by each kernel; their sum is B_shadow + B_forward of SURVEY.md §8d."""
    V, T, W, H, S = scene["n_vertices"], scene["n_triangles"], scene["width"], scene["height"], scene["shadow_size"]
    alg = {
        (0, "clear"): 0, (0, "vertex"): 12 * V, (0, "setup_raster"): 12 * T, (0, "bin_scatter"): 0, (0, "tile_stage"): 4 * S * S,
        (1, "clear"): 0, (1, "vertex"): 32 * V, (1, "setup_raster"): 12 * T, (1, "bin_scatter"): 0,
        (1, "tile_stage"): (min(36 * W * H, 4 * S * S) if S else 0) + 12 * W * H,
    }
    if fused:   # fused two-pass vertex kernel: timed in the shadow slot
        alg[(0, "vertex")] = 12 * V + 32 * V; alg[(1, "vertex")] = 0
    return alg


def roofline_of(scene, acc, peak, peak_src):
    fused = bool(scene["shadow_size"]) and bool(acc[0]) and acc[1].get("vertex", 1.0) < 0.2 * acc[0].get("vertex", 0.0)
    alg = alg_table(scene, fused)
    kt = {f"{'shadow' if k == 0 else 'forward'}.{n}": v for k in (0, 1) for n, v in acc[k].items()}
    dom = max(((k, n) for k in (0, 1) for n in acc[k]), key=lambda kn: acc[kn[0]][kn[1]])
    dom_ms = acc[dom[0]][dom[1]]
    dom_bytes = alg.get(dom, 0)
    achieved = dom_bytes / (dom_ms * 1e-3) / 1e9 if dom_ms > 0 else 0.0
    traffic, traffic_src = profiled_traffic(dom[0], dom[1]) if scene.get("config") == "C2" else (None, None)
    roof = {"bound": "hbm", "kernel": f"{'shadow' if dom[0] == 0 else 'forward'}.{dom[1]}", "achieved": achieved, "peak": peak,
            "unit": "GB/s", "frac": achieved / peak, "traffic": traffic, "traffic_source": traffic_src, "algorithmic_bytes_per_launch": dom_bytes,
            "kernel_ms": dom_ms, "peak_source": peak_src + " (MEASURED_PEAKS.json hbm_gbs, burst copy)"}
    return roof, kt


def bit_exact_vs_cpu(rig, keep):
    """The frame the CPU oracle was timed on, re-rendered on the device and compared bit for bit ("must hash-equal")."""
    capi, ctx = rig.capi, rig.ctx
    with rig.torch.cuda.stream(rig.st):
        ctx.set_view(np.array(rig.scene["view"]))
        rig.frame(); ctx.sync()
        same = bool(np.array_equal(ctx.readback(capi.REB_DEPTH_F32).view(np.uint32), keep["depth"].view(np.uint32))
                    and np.array_equal(ctx.readback(capi.REB_TRIID_U32), keep["tri_id"])
                    and (not rig.S or np.array_equal(ctx.readback(capi.REB_SHADOW_F32).view(np.uint32), np.asarray(keep["shadow"], np.float32).view(np.uint32))))
        d8 = np.abs(ctx.readback(capi.REB_COLOR_SRGB8).astype(np.int16) - keep["color_srgb8"].astype(np.int16))
    return {"depth_triid_shadow": same, "srgb8_max_lsb": int(d8.max())}


def sub_record(args, config, local_rank, flush, peak, steps, cpu_frames, scene=None, note=None):
    """One more configuration measured in the same invocation: device time, Mtri/s, fraction of the HBM roofline on the frame's
    algorithmic bytes, dominant kernel, and the CPU port beside it."""
    scene = scene or build_scene(args, config)
    rig = Rig(scene, local_rank, samples=args.samples)
    per = rig.timed(rig.frame, steps, 5, flush)
    acc = rig.kernel_profile(flush, n_prof=10)
    counters = rig.ctx.counters()
    roof, kt = roofline_of(scene, acc, peak, "measured")
    b_shadow, b_forward = sc.algorithmic_bytes(scene)
    ms = float(np.median(per))
    T = scene["n_triangles"]
    rec = {"workload": workload_string(args, 1, scene, "frame"), "ms_per_frame": ms, "value": T / (ms * 1e3), "unit": UNIT, "frames_per_s": 1e3 / ms,
           "frames_timed": steps, "algorithmic_bytes": b_shadow + b_forward, "achieved_gbs": (b_shadow + b_forward) / (ms * 1e-3) / 1e9,
           "frac_of_peak": (b_shadow + b_forward) / (ms * 1e-3) / 1e9 / peak,
           "dominant_kernel": {"name": roof["kernel"], "ms": roof["kernel_ms"], "frac": roof["frac"]}, "kernel_ms": kt,
           "counters": {"needs_clip": int(counters[0]), "binned_triangles": int(counters[1]), "tile_list_entries": int(counters[2]), "regrows": int(counters[3])}}
    if note:
        rec["note"] = note
    if cpu_frames and not args.no_cpu_baseline and args.samples == 1 and rig.tl is None:
        keep = {}
        times, cores = cpu_reference_frames(scene, cpu_frames, warm=0, keep=keep)
        cms = 1e3 * float(np.mean(times))
        rec["cpu_baseline"] = {"value": T / (cms * 1e3), "unit": UNIT, "cores": cores, "kind": "port", "ms_per_frame": cms,
                               "sample": f"{cpu_frames} full frame(s) of the same workload on the CPU oracle, all host cores",
                               "gpu_matches_bit_exact": bit_exact_vs_cpu(rig, keep)}
    rig.close()
    return rec


def host_copy_ceiling(torch, dist, world, nbytes, n=200):
    """Copy-only probe: every rank copies `nbytes` device -> pinned host `n` times on a side stream, nothing rendered: what the
    host side (PCIe root + host memory of the box) can take when all ranks deliver frames at once."""
    dev = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
    host = [torch.empty(nbytes, dtype=torch.uint8).pin_memory() for _ in range(2)]
    for h in host:
        h.zero_()
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        for i in range(8):
            host[i & 1].copy_(dev, non_blocking=True)
        s.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for i in range(n):
            host[i & 1].copy_(dev, non_blocking=True)
        s.synchronize()
        dt = time.perf_counter() - t0
    t = torch.tensor([dt], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dt = float(t.item())
    return {"per_rank_gbs": nbytes * n / dt / 1e9, "aggregate_gbs": world * nbytes * n / dt / 1e9, "copies": n, "bytes_per_copy": nbytes}


# ---------------------------------------------------------------------------------------------------------------------------
# N = 1
# ---------------------------------------------------------------------------------------------------------------------------
def view_batch_local(rig, views, flush, steps, warmup):
    """This rank's share of a view batch with no exchange at all: shadow map once per step, then every view against it."""
    def step():
        if rig.S:
            rig.ctx.render_shadow(rig.shadow_draws)
        for v in views:
            rig.ctx.set_view(v)
            rig.ctx.render_forward(rig.draws, rig.tl)
    return rig.timed(step, steps, warmup, flush)


def run_single(args, torch, dist, local_rank):
    from renderengine_b200 import capi
    numa = pin_to_gpu_numa(local_rank)
    scene = build_scene(args)
    T, W, H, S = scene["n_triangles"], scene["width"], scene["height"], scene["shadow_size"]
    rig = Rig(scene, local_rank, samples=args.samples)
    ctx, st = rig.ctx, rig.st
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")   # > 126 MB L2
    peak, peak_src = measured_peak()

    sampler = ClockSampler(local_rank)
    with torch.cuda.stream(st):
        for _ in range(args.warmup):
            flush.add_(1); rig.frame()
        ctx.sync()
        torch.cuda.synchronize()
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
        sampler.start()
        l0 = ctx.launch_count()
        for a, b in ev:
            flush.add_(1)                                 # L2 flush between timed iterations (outside the event pair)
            a.record(st); rig.frame(); b.record(st)
        ctx.sync()
        torch.cuda.synchronize()
        clocks = sampler.stop()
        launches = ctx.launch_count() - l0
        per_frame = np.array([a.elapsed_time(b) for a, b in ev])
    ms = float(per_frame.sum()) / args.steps
    value = T / (ms * 1e3)

    # ---- forward-only frames with the cached shadow map: what the reference does after its first frame (RenderScene.cpp:171-175) ----
    fwd_only = rig.timed(rig.forward, min(args.steps, 50), 2, flush)
    acc = rig.kernel_profile(flush)
    counters = ctx.counters()

    # ---- end to end through the C ABI with host buffers (wall clock, includes D2H of the frame) -------
    # every frame: host view / light blocks in, the finished sRGB8 frame out to pinned host memory.  The target is a three-image swapchain,
    # so frame i's device->host copy (copy stream) overlaps frame i+1's rendering; all inside the timed region.
    pinned = [torch.empty((H, W, 4), dtype=torch.uint8).pin_memory() for _ in range(3)]
    for p in pinned:
        p.zero_()                                          # first touch on the GPU-local NUMA node
    view_host = np.array(scene["view"])
    n_e2e = max(10, min(args.steps, 100))
    with torch.cuda.stream(st):
        for i in range(4):
            rig.frame(); ctx.readback_async_into(capi.REB_COLOR_SRGB8, pinned[i % 3].data_ptr())
        ctx.sync()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for i in range(n_e2e):
            ctx.set_view(view_host)
            if S:
                ctx.shadow_set_transform(scene["shadow_view"], scene["shadow_proj"])
            rig.frame()
            ctx.readback_async_into(capi.REB_COLOR_SRGB8, pinned[i % 3].data_ptr())
        ctx.sync()
        torch.cuda.synchronize()
        e2e_ms = (time.perf_counter() - t0) * 1e3 / n_e2e
    ceiling = host_copy_ceiling(torch, dist, 1, W * H * 4)
    e2e_val = T / (e2e_ms * 1e3)
    d2h_gbs = W * H * 4 / (e2e_ms * 1e-3) / 1e9

    b_shadow, b_forward = sc.algorithmic_bytes(scene)
    roof, kt = roofline_of(scene, acc, peak, peak_src)
    frame_bytes = b_shadow + b_forward
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": 1, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic", "frames_per_s": 1e3 / ms,
        "config": {"workload": workload_string(args, 1, scene, "frame"), "parallelism": "1 GPU",
                   "l2": "256 MiB flush write between timed steps", "outputs": "sRGB8 + depth f32 + triID u32"},
        "hbm_frame": {"algorithmic_bytes": frame_bytes, "achieved_gbs": frame_bytes / (ms * 1e-3) / 1e9,
                      "frac_of_peak": frame_bytes / (ms * 1e-3) / 1e9 / peak, "peak_gbs": peak, "peak_source": peak_src},
        "roofline": roof, "kernel_ms": kt,
        "e2e": {"value": e2e_val, "unit": UNIT, "ms_per_step": e2e_ms, "h2d_bytes_per_step": 160 + 128 + 2 * 16 * len(rig.shadow_draws),
                "d2h_bytes_per_step": W * H * 4, "d2h_gbs_achieved": d2h_gbs,
                "host_ceiling_gbs": ceiling["aggregate_gbs"], "host_ceiling": ceiling, "e2e_over_ceiling": d2h_gbs / ceiling["aggregate_gbs"], "numa": numa,
                "bound": "host copy" if d2h_gbs > 0.8 * ceiling["aggregate_gbs"] else "render",
                "note": "wall clock through reb_set_view/reb_render_*/reb_readback_async + final reb_sync with host buffers (pinned, 3-image swapchain: the copy of frame i overlaps the rendering of frames i+1 and i+2); geometry is static and resident as in the reference frame loop; host_ceiling_gbs = copy-only probe of the same 8.3-MB frames, nothing rendered"},
        "gpu_launches": int(launches), "clocks": clocks,
        "frame_us": {"median": 1e3 * float(np.median(per_frame)), "p95": 1e3 * float(np.percentile(per_frame, 95)), "rank": 0},
        "counters": {"needs_clip": int(counters[0]), "binned_triangles": int(counters[1]), "tile_list_entries": int(counters[2]), "regrows": int(counters[3])},
    }
    fms = float(np.median(fwd_only))
    line["forward_only"] = {"ms_per_frame": fms, "value": T / (fms * 1e3), "unit": UNIT, "algorithmic_bytes": b_forward,
                            "achieved_gbs": b_forward / (fms * 1e-3) / 1e9, "frac_of_peak": b_forward / (fms * 1e-3) / 1e9 / peak,
                            "note": "forward pass alone against the cached shadow map (the reference renders the shadow map once, RenderScene.cpp:171-175)"}
    if not args.no_cpu_baseline and args.samples == 1 and rig.tl is None:
        keep = {}
        times, cores = cpu_reference_frames(scene, 5, warm=1, keep=keep)
        cms = 1e3 * float(np.mean(times))
        line["cpu_baseline"] = {"value": T / (cms * 1e3), "unit": UNIT, "cores": cores, "kind": "port", "ms_per_frame": cms,
                                "sample": "5 full frames (shadow + forward) of the same workload on the CPU oracle, all host cores",
                                "gpu_matches_bit_exact": bit_exact_vs_cpu(rig, keep)}
    rig.close()
    del pinned

    if not args.no_extras and args.config == "C2" and args.shading == "phong" and args.samples == 1 and args.translucent_every == 0:
        # ---- the other single-GPU configurations of BASELINE.json, same invocation ("1080p/4K") -------------------------------
        line["c1"] = sub_record(args, "C1", local_rank, flush, peak, 50, 3)
        line["c3"] = sub_record(args, "C3", local_rank, flush, peak, 30, 1)
        # C3 with the reference's default 4096^2 map as well (RenderScene.cpp:26); device time only
        c3big = sc.make_scene("C3", shadow_size=4096)
        rb = Rig(c3big, local_rank)
        perb = rb.timed(rb.frame, 20, 3, flush)
        mb = float(np.median(perb))
        bsh, bfw = sc.algorithmic_bytes(c3big)
        line["c3"]["shadow_4096"] = {"workload": workload_string(args, 1, c3big, "frame"), "ms_per_frame": mb, "value": c3big["n_triangles"] / (mb * 1e3), "unit": UNIT,
                                     "algorithmic_bytes": bsh + bfw, "frac_of_peak": (bsh + bfw) / (mb * 1e-3) / 1e9 / peak}
        rb.close()
        # ---- the binned tile path on the record: the C2 grid at subdivision level 2 (16 k triangles of ~16 / 52 px) with the
        # reference's default 4096^2 shadow map (RenderScene.cpp:26) ------------------------------------------------------------
        low = sc.make_scene("C2", level=2, shadow_size=4096)
        low["config"] = "C2 grid at L=2 (16 k triangles) + 4096^2 map"
        line["lowpoly"] = sub_record(args, None, local_rank, flush, peak, 50, 2, scene=low,
                                     note="larger triangles: the shadow pass's 52-px triangles take the binned per-tile lists")
        # ---- config 5's per-GPU share: 64 views of the 2 M-triangle scene against one shadow map ---------------------------------
        s5 = build_scene(args, "C5")
        r5 = Rig(s5, local_rank)
        nv = args.views_per_rank
        views = [sc.view_for_azimuth(s5, 2 * np.pi * i / 512) for i in range(nv)]
        per = view_batch_local(r5, views, flush, max(3, min(args.steps, 10)), 2)
        bms = float(np.median(per))
        line["view_batch"] = {"workload": workload_string(args, 1, s5, "views"), "ms_per_step": bms, "views_per_step": nv,
                              "value": nv * s5["n_triangles"] / (bms * 1e3), "unit": UNIT, "ms_per_view": bms / nv,
                              "note": "single GPU, no exchange: the rate the N > 1 headline (same workload per rank + one NCCL broadcast per step) is compared with"}
        r5.close()
    return line


# ---------------------------------------------------------------------------------------------------------------------------
# N > 1
# ---------------------------------------------------------------------------------------------------------------------------
def all_true(torch, dist, flag):
    t = torch.tensor([1 if flag else 0], dtype=torch.int32, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MIN)
    return bool(t.item())


def max_over_ranks(torch, dist, x):
    t = torch.tensor([x], dtype=torch.float64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def run_views(args, torch, dist, rank, world, local_rank, flush, peak, peak_src):
    """Config 5: `views_per_rank` views per rank and step against ONE shadow map per step, rendered on rank 0 and broadcast."""
    from renderengine_b200 import capi
    scene = build_scene(args, "C5")
    T, W, H, S = scene["n_triangles"], scene["width"], scene["height"], scene["shadow_size"]
    nv = args.views_per_rank
    rig = Rig(scene, local_rank, samples=args.samples)
    ctx, st = rig.ctx, rig.st
    views = [sc.view_for_azimuth(scene, 2 * np.pi * (rank * nv + i) / (world * nv)) for i in range(nv)]   # the batch's views, dealt to the ranks in blocks
    shadow_t = rig.shadow_tensor()
    comm = torch.cuda.Stream()
    ready = torch.cuda.Event()
    comm_ev = []

    def step(record_comm=False):
        # rank 0 renders the map (ShadowSystem::Render once, RenderScene.cpp:171-175) ...
        if rank == 0:
            ctx.render_shadow(rig.shadow_draws)
            ctx.flush()                                   # the recorded depth pass must be on the stream before the exchange
        # ... the exchange runs on its own stream behind it (and behind the last reader of the previous step on every rank) ...
        comm.wait_stream(st)
        with torch.cuda.stream(comm):
            if record_comm:
                a = torch.cuda.Event(enable_timing=True); b = torch.cuda.Event(enable_timing=True)
                a.record(comm)
            dist.broadcast(shadow_t, src=0)               # NCCL over NVLink: 4 S^2 bytes to every other rank
            if record_comm:
                b.record(comm); comm_ev.append((a, b))
            ready.record(comm)
        # ... and only the first view's tile stage waits for it: K0 / K1 of that view overlap the transfer
        ctx.shadow_wait_event(ready.cuda_event)
        for v in views:
            ctx.set_view(v)
            ctx.render_forward(rig.draws, rig.tl)

    def barrier():
        torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()

    with torch.cuda.stream(st):
        for _ in range(max(args.warmup, 3)):
            flush.add_(1); step()
        ctx.sync(); barrier()
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
        sampler = ClockSampler(local_rank)
        if rank == 0:
            sampler.start()
        l0 = ctx.launch_count()
        for a, b in ev:
            flush.add_(1)
            a.record(st); step(record_comm=True); b.record(st)
        ctx.sync(); barrier()
        clocks = sampler.stop() if rank == 0 else None
        launches = ctx.launch_count() - l0
        per_step = np.array([a.elapsed_time(b) for a, b in ev])
    ms = max_over_ranks(torch, dist, float(per_step.sum())) / args.steps
    value = world * nv * T / (ms * 1e3)
    comm_ms = max_over_ranks(torch, dist, float(np.mean([a.elapsed_time(b) for a, b in comm_ev])))

    # ---- the same per-rank workload with NO exchange (every rank renders its own map): the single-GPU rate, same run ----------
    local = view_batch_local(rig, views, flush, max(3, min(args.steps, 10)), 2)
    local_ms = float(np.median(local))
    local_ms_max = max_over_ranks(torch, dist, local_ms)

    # ---- same-run parity at N: the broadcast map and frames rendered against it == the rank-local render ---------------------
    with torch.cuda.stream(st):
        ctx.render_shadow(rig.shadow_draws); ctx.sync()
        own_map = hashlib.sha256(ctx.readback(capi.REB_SHADOW_F32).tobytes()).hexdigest()
        own = []
        for v in (views[0], views[-1]):
            ctx.set_view(v); ctx.render_forward(rig.draws, rig.tl); ctx.sync()
            own.append(rig.hashes())
        step(); ctx.sync()
        got_map = hashlib.sha256(ctx.readback(capi.REB_SHADOW_F32).tobytes()).hexdigest()
        last_ok = rig.hashes() == own[1]
        ctx.set_view(views[0]); ctx.render_forward(rig.draws, rig.tl); ctx.sync()
        first_ok = rig.hashes() == own[0]
    parity = {"broadcast_map_equals_local_render": all_true(torch, dist, got_map == own_map),
              "views_against_broadcast_map_equal_local": all_true(torch, dist, first_ok and last_ok)}

    acc = rig.kernel_profile(flush, n_prof=10, shadow=(rank == 0))

    # ---- end to end: every view's finished frame to pinned host memory, host view blocks in -----------------------------------------
    numa = pin_to_gpu_numa(local_rank)
    pinned = [torch.empty((H, W, 4), dtype=torch.uint8).pin_memory() for _ in range(3)]
    for p in pinned:
        p.zero_()
    n_e2e = max(2, min(args.steps, 10))

    def e2e_step():
        if rank == 0:
            ctx.shadow_set_transform(scene["shadow_view"], scene["shadow_proj"])
            ctx.render_shadow(rig.shadow_draws); ctx.flush()
        comm.wait_stream(st)
        with torch.cuda.stream(comm):
            dist.broadcast(shadow_t, src=0)
            ready.record(comm)
        ctx.shadow_wait_event(ready.cuda_event)
        for i, v in enumerate(views):
            ctx.set_view(v)
            ctx.render_forward(rig.draws, rig.tl)
            ctx.readback_async_into(capi.REB_COLOR_SRGB8, pinned[i % 3].data_ptr())
    with torch.cuda.stream(st):
        e2e_step(); ctx.sync(); barrier()
        t0 = time.perf_counter()
        for _ in range(n_e2e):
            e2e_step()
        ctx.sync(); barrier()
        e2e_ms = max_over_ranks(torch, dist, (time.perf_counter() - t0) * 1e3 / n_e2e)
    ceiling = host_copy_ceiling(torch, dist, world, W * H * 4)
    d2h_gbs = world * nv * W * H * 4 / (e2e_ms * 1e-3) / 1e9

    b_shadow, b_forward = sc.algorithmic_bytes(scene)
    step_bytes = b_shadow + world * nv * b_forward                     # one shadow pass + the views
    roof, kt = roofline_of(scene, acc, peak, peak_src)
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
        "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "frames_per_s": world * nv * 1e3 / ms,
        "config": {"workload": workload_string(args, world, scene, "views"),
                   "parallelism": f"view-parallel x{world}: {nv} views per rank and step; shadow map rendered once per step on rank 0, NCCL broadcast on its own stream, first view's tile stage waits for it",
                   "l2": "256 MiB flush write between timed steps", "outputs": "sRGB8 + depth f32 + triID u32 per view"},
        "hbm_frame": {"algorithmic_bytes": step_bytes, "achieved_gbs": step_bytes / (ms * 1e-3) / 1e9,
                      "frac_of_peak": step_bytes / (ms * 1e-3) / 1e9 / peak / world, "peak_gbs": peak, "peak_source": peak_src},
        "roofline": roof, "kernel_ms": kt,
        "view_batch": {"value": value, "unit": UNIT, "ms_per_step": ms, "views_per_step": world * nv,
                       "single_gpu_same_run": {"ms_per_step": local_ms, "ms_per_step_slowest_rank": local_ms_max, "value": nv * T / (local_ms * 1e3),
                                               "note": "rank 0 rendering its own shadow map + its views, no exchange (every rank does the same at the same time)"},
                       "speedup_vs_single_gpu_same_run": value / (nv * T / (local_ms * 1e3))},
        "comm": {"collective": "ncclBroadcast (torch.distributed, NVLink)", "bytes_per_step": 4 * S * S * (world - 1), "payload_bytes": 4 * S * S,
                 "event_ms_per_step": comm_ms, "payload_gbs": 4 * S * S / (comm_ms * 1e-3) / 1e9 if comm_ms > 0 else None,
                 "overlap": "own stream; only the first view's tile stage waits (reb_shadow_wait_event)",
                 "note": "event time on the exchange stream, max over ranks; includes waiting for rank 0's shadow pass on the receivers"},
        "parity_n": parity,
        "e2e": {"value": world * nv * T / (e2e_ms * 1e3), "unit": UNIT, "ms_per_step": e2e_ms, "h2d_bytes_per_step": nv * 160 + 128 + 16 * len(rig.shadow_draws) + nv * 16 * len(rig.draws),
                "d2h_bytes_per_step": nv * W * H * 4, "d2h_gbs_achieved": d2h_gbs, "host_ceiling_gbs": ceiling["aggregate_gbs"], "host_ceiling": ceiling,
                "e2e_over_ceiling": d2h_gbs / ceiling["aggregate_gbs"], "numa": numa,
                "note": "per rank and step: host view block in + finished sRGB8 frame out to pinned host memory for EVERY view (3-image swapchain, copy stream); bytes are per rank; host_ceiling_gbs = copy-only probe, all ranks at once, nothing rendered"},
        "gpu_launches": int(launches), "clocks": clocks,
        "frame_us": {"median": 1e3 * float(np.median(per_step)) / nv, "p95": 1e3 * float(np.percentile(per_step, 95)) / nv, "rank": 0, "per": "view"},
    }
    rig.close()
    return line


def run_sort_first(args, torch, dist, rank, world, local_rank, flush, peak):
    """Config 4: ONE 8K frame of the 10.5 M-triangle scene over N GPUs (row bands), next to the unsplit frame on one GPU, same run."""
    from renderengine_b200 import capi, parallel
    scene = build_scene(args, "C4")
    T, W, H, S = scene["n_triangles"], scene["width"], scene["height"], scene["shadow_size"]
    steps = max(3, min(args.steps, args.sort_first_frames))
    rig = Rig(scene, local_rank)
    ctx, st = rig.ctx, rig.st
    # ---- (a) the whole frame on this GPU alone (every rank does the same, no exchange) ----------------------------------------
    single = rig.timed(rig.frame, steps, 2, flush)
    single_ms = float(np.median(single))
    with torch.cuda.stream(st):
        rig.frame(); ctx.sync()
    band_h = parallel.contiguous_band(H, world * args.bands_per_rank) if args.bands_per_rank > 0 else 64
    own = parallel.owned_rows(H, rank, world, band_h)
    want = rig.hashes(rows=own)
    want_map = hashlib.sha256(ctx.readback(capi.REB_SHADOW_F32).tobytes()).hexdigest()
    # ---- (b) the split frame ----------------------------------------------------------------------------------------------------
    slab = parallel.slab_rows(S, world) if args.shadow_split == "rows" else None
    by_tris = args.shadow_split == "triangles"
    ctx.set_partition(rank, world, band_h)
    if slab:
        ctx.set_shadow_partition(rank, world, slab)
    elif by_tris:
        ctx.set_shadow_split(rank, world)          # sort-last: an equal share of the triangle stream into a whole map on every rank
    shadow_t = rig.shadow_tensor()
    comm = torch.cuda.Stream()
    ready = torch.cuda.Event()
    comm_ev = []

    def step(record_comm=False):
        if slab or by_tris or rank == 0:
            ctx.render_shadow(rig.shadow_draws)            # this rank's row slab / triangle share of the map (or the whole map on rank 0)
            ctx.flush()
        comm.wait_stream(st)
        with torch.cuda.stream(comm):
            if record_comm:
                a = torch.cuda.Event(enable_timing=True); b = torch.cuda.Event(enable_timing=True); a.record(comm)
            if slab:
                parallel.allgather_shadow_slabs(shadow_t)  # NCCL all-gather in place: 4 S^2 / N bytes sent per rank
            elif by_tris:
                parallel.allreduce_shadow_min(shadow_t)    # NCCL all-reduce (MIN) in place: the minimum of the ranks' whole maps
            else:
                parallel.broadcast_shadow(shadow_t, src=0)
            if record_comm:
                b.record(comm); comm_ev.append((a, b))
            ready.record(comm)
        ctx.shadow_wait_event(ready.cuda_event)            # K0 / K1 of the forward pass overlap the exchange
        ctx.render_forward(rig.draws, rig.tl)

    with torch.cuda.stream(st):
        for _ in range(3):
            flush.add_(1); step()
        ctx.sync(); torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
        for a, b in ev:
            flush.add_(1)
            a.record(st); step(record_comm=True); b.record(st)
        ctx.sync(); torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
        per = np.array([a.elapsed_time(b) for a, b in ev])
        got = rig.hashes(rows=own)
        got_map = hashlib.sha256(ctx.readback(capi.REB_SHADOW_F32).tobytes()).hexdigest()
    ms = max_over_ranks(torch, dist, float(per.sum())) / steps
    t = torch.tensor([float(per.mean())], dtype=torch.float64, device="cuda"); dist.all_reduce(t); mean_ms = float(t.item()) / world
    comm_ms = max_over_ranks(torch, dist, float(np.mean([a.elapsed_time(b) for a, b in comm_ev])))
    single_max = max_over_ranks(torch, dist, single_ms)
    counters = ctx.counters()
    rec = {"workload": workload_string(args, world, scene, "tiles"), "ms_per_frame": ms, "ms_per_frame_mean_over_ranks": mean_ms,
           "value": T / (ms * 1e3), "unit": UNIT, "frames_timed": steps,
           "single_gpu_same_run": {"ms_per_frame": single_ms, "ms_per_frame_slowest_rank": single_max, "value": T / (single_ms * 1e3)},
           "speedup_vs_single_gpu_same_run": single_ms / ms,
           "parallelism": f"sort-first row bands x{world} (band_h {band_h}, {args.bands_per_rank} bands per rank, per-draw + per-block culling); shadow map: " +
                          (f"{slab}-row slabs + in-place NCCL all-gather" if slab else
                           "split by triangles (equal shares of the stream into whole maps) + in-place NCCL all-reduce MIN" if by_tris else "rank 0 + NCCL broadcast") +
                          ", on its own stream; the forward tile stage waits for it",
           "comm": {"collective": "ncclAllGather" if slab else "ncclAllReduce(min)" if by_tris else "ncclBroadcast",
                    "bytes_sent_per_rank": 4 * S * S // world if slab else 4 * S * S, "payload_bytes": 4 * S * S,
                    "event_ms_per_frame": comm_ms},
           "parity_n": {"own_bands_equal_unsplit_frame": all_true(torch, dist, got == want),
                        "gathered_map_equals_local_render": all_true(torch, dist, got_map == want_map)},
           "counters": {"needs_clip": int(counters[0]), "binned_triangles": int(counters[1]), "regrows": int(counters[3])}}
    rig.close()
    return rec


def run_replicas(args, torch, dist, rank, world, local_rank, flush, peak, peak_src):
    """`--mode replicas`: N independent replicas of the single-GPU headline frame (one C2 view per rank, own shadow map, no collective)."""
    scene = build_scene(args)
    T = scene["n_triangles"]
    rig = Rig(scene, local_rank, samples=args.samples)
    rig.ctx.set_view(sc.view_for_azimuth(scene, 2 * np.pi * rank / world))
    per = rig.timed(rig.frame, args.steps, max(args.warmup, 3), flush)
    ms = max_over_ranks(torch, dist, float(per.sum())) / args.steps
    rig.close()
    return {"metric": METRIC, "value": world * T / (ms * 1e3), "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": workload_string(args, world, scene, "frame"), "parallelism": f"{world} independent replicas (replicas only: no exchange)"}}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=None)
    ap.add_argument("--warmup", type=int, default=None)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default="C2")
    ap.add_argument("--mode", default="auto", choices=["auto", "frame", "views", "tiles", "replicas"],
                    help="auto: N = 1 -> the C2 frame; N > 1 -> the C5 view batch as headline + the C4 sort-first record")
    ap.add_argument("--views-per-rank", type=int, default=64, help="C5: 512 views over 8 GPUs")
    ap.add_argument("--shadow-split", choices=["rows", "triangles", "broadcast"], default="rows",
                    help="sort-first: how the shadow pass is shared: triangles = every rank rasterises an equal share of the triangle stream into a whole map, "
                         "all-reduce MIN (reb_set_shadow_split); rows = row slabs + all-gather (reb_set_shadow_partition); broadcast = rank 0 renders it")
    ap.add_argument("--bands-per-rank", type=int, default=3,
                    help="sort-first: row bands per rank, interleaved (1 = one contiguous band: least geometry per rank after culling but unbalanced; 0 = 64-row bands)")
    ap.add_argument("--sort-first-frames", type=int, default=10)
    ap.add_argument("--shading", default="phong", choices=["phong", "pbr"],
                    help="phong = BASELINE.json's path (default); pbr = the reference's live surface_pbr.frag path (SURVEY 8f row 1)")
    ap.add_argument("--samples", type=int, default=1, choices=[1, 4],
                    help="samples per pixel of the base pass (4 = Appendix B.15 multisampling; the headline metric is quoted at 1)")
    ap.add_argument("--translucent-every", type=int, default=0,
                    help="make every K-th material AlphaMode::eBlend (alpha 0.45): its draws go through the translucent pass (SURVEY 8f row 3); 0 = all opaque (headline)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="N = 1: only the headline record (no c1 / c3 / lowpoly / view_batch); N > 1: no sort_first record")
    args = ap.parse_args()

    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    mode = resolve_mode(args, world)
    if args.steps is None:
        args.steps = 200 if mode in ("frame", "replicas") else 20      # a views step is 64 frames, a tiles step one 8K frame
    if args.warmup is None:
        args.warmup = 20 if mode in ("frame", "replicas") else 3
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    if args.impl == "reference":
        run_reference(args, rank, world)
        return

    import torch
    import torch.distributed as dist

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the product has no CPU path (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    if world == 1:
        line = run_single(args, torch, dist, local_rank)
    else:
        flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
        peak, peak_src = measured_peak()
        if mode == "replicas":
            line = run_replicas(args, torch, dist, rank, world, local_rank, flush, peak, peak_src)
        elif mode == "tiles":
            rec = run_sort_first(args, torch, dist, rank, world, local_rank, flush, peak)
            line = {"metric": METRIC, "value": rec["value"], "unit": UNIT, "n_gpus": world, "steps": rec["frames_timed"], "warmup": 3, "ms_per_step": rec["ms_per_frame"],
                    "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                    "config": {"workload": rec["workload"], "parallelism": rec["parallelism"]}, "sort_first": rec}
        else:
            line = run_views(args, torch, dist, rank, world, local_rank, flush, peak, peak_src)
            if not args.no_extras:
                line["sort_first"] = run_sort_first(args, torch, dist, rank, world, local_rank, flush, peak)
    if rank == 0:
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
