#!/usr/bin/env python
"""bench.py - rendered Mpixel/s of batched camera views (BASELINE.json metric), ref/cuda vs ref_soft.

    python bench.py [--gpus N] [--steps K] [--warmup W]            # our arm (libref_cuda.so on the GPU)
    python bench.py --impl reference [--gpus N] ...                # the reference's CPU path (unmodified ref_soft)

A step = one batch of `--views` camera views per GPU through the whole frame path.  Default workload, at every N:
BASELINE.json configs[4] (c5: 4096 views per GPU at 320x240 on the synthetic Map S with water and 32 sorted
translucent sprites - the configuration of the metric's 1/2/4/8-GPU sweep); --config c2 / c3 / c4 select configs[1..3].
N > 1: one rank per GPU (torchrun), every rank renders its own slice of the camera list (weak scaling, no data-path
collective; NCCL only gathers the timing and the counters).

  value  = whole-job Mpixel/s with the batch's inputs resident in HBM (CUDA events around the device stages,
           L2 flushed between steps, max over ranks)
  e2e    = the same through the plugin's public C-ABI calls RefCuda_RenderViewsAsync + RefCuda_Wait with HOST buffers:
           entity state and view descriptors go host->device and the colour frames come back device->host (pinned,
           two alternating sets) every step; e2e.blocking_call_* = the blocking RefCuda_RenderViews
  roofline = the span drawing kernel (k_draw): 6 algorithmic bytes per pixel / its CUDA-event time vs the
           measured HBM copy bandwidth (MEASURED_PEAKS.json)
  cpu_baseline = the oracle (unmodified ref_soft, oracle/_ref) timed on one host core on a bounded sample
"""
from __future__ import annotations

import argparse
import os as _os

_os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")  # before torch creates the CUDA context: see rc_dev_init (rc_device.cu)
import importlib
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "rendered Mpixel/s (batched views)"
UNIT = "Mpixel/s"
BYTES_PER_PIXEL = 6  # 2 B colour write + 2 B z write + 2 B cache texel read


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="cuda", choices=["cuda", "reference"])
    ap.add_argument("--config", default=None, choices=["c2", "c3", "c4", "c5"],
                    help="BASELINE.json config: c2 = 256 views/GPU 320x240 Map S, c3 = 1920x1080 views of Map L with "
                         "light styles changing every step (surface cache rebuilt per step), c4 = c2's batch with 64 skinned 32-bone studio models, c5 = 4096 views/GPU 320x240 Map S with water and sprites (the default, at every N)")
    ap.add_argument("--views", type=int, default=None, help="views per GPU per step")
    ap.add_argument("--width", type=int, default=None)
    ap.add_argument("--height", type=int, default=None)
    ap.add_argument("--map", default=None, choices=["tiny", "small", "medium", "large"])
    ap.add_argument("--cpu-seconds", type=float, default=10.0, help="budget of the cpu_baseline sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    a = ap.parse_args()
    if a.config is None:
        # BASELINE.json quotes its metric "at 1/2/4/8 B200": that sweep is configs[4] (c5, 4096 views per GPU, water, sprites).
        # Per-GPU work must be the same at every N for the driver's weak-scaling figure, so c5 is the default at N = 1 too
        # (round 1 ran 256 views per GPU: 6 ms of GPU work per run); configs[1] is --config c2.
        a.config = "c5"
    preset = {"c2": (256, 320, 240, "small"), "c3": (8, 1920, 1080, "large"), "c4": (256, 320, 240, "small"), "c5": (4096, 320, 240, "small")}[a.config]
    a.views = a.views or preset[0]
    a.width = a.width or preset[1]
    a.height = a.height or preset[2]
    a.map = a.map or preset[3]
    a.map_kw = {"water_prob": (1, 5)} if a.config == "c5" else {}
    a.rebuild = a.config == "c3"
    a.models = 64 if a.config == "c4" else 0
    a.sprites = 32 if a.config == "c5" else 0
    a.cvars = {"sw_maxsurfs": 40000, "sw_maxedges": 160000} if a.config == "c3" else {}
    return a


C3_PATTERNS = {1: "mmnmmommommnonmmonqnmmo", 2: "abcdefghijklmnopqrstuvwxyzyxwvutsrqponmlkjihgfedcba", 3: "mamamamamama",
               4: "mmamammmmammamamaaamammma", 5: "jklmnopqrstuvwxyzyxwvutsrqponmlkj", 6: "nmonqnmomnmomomno",
               7: "mmmaaaabcdefgmmmmaaaammmaamm", 8: "mmmaaammmaaammmabcdefaaaammmmabcdefmmmaaaa", 9: "aaaaaaaazzzzzzzz",
               10: "mmamammmmammamamaaamammma", 11: "abcdefghijklmnopqrrqponmlkjihgfedcba"}


def advance_styles(r, step):
    """config c3: every step is a new timestamp - the animated light styles take new values, so every surface-cache
    block that lists one of them is rebuilt (D_CacheSurface's staleness test, ref/soft/r_surf.c:1090)"""
    for k, p in C3_PATTERNS.items():
        r.set_lightstyle(k, p, time=step / 60.0 * 7 + 0.05)
    r.set_time(1.0 + step / 60.0)
    r.run_lightstyles()


def populate(args, scene, r, m):
    """the entities of the config, the same on every rank and in every ref_soft process (the scene is replicated)"""
    if args.models:
        r.set_time(1.0)
        scene.add_studio_models(r, m.cameras, args.models)
    if args.sprites:
        scene.add_sprites(r, m.cameras, args.sprites)


def advance(args, r, step):
    """what changes between two steps: c3 = the light styles, c4 = the client time the models animate on (new bones per step)"""
    if args.rebuild:
        advance_styles(r, step)
    elif args.models:
        r.set_time(1.0 + step / 60.0)


def workload(args, world):
    what = {"c2": "static lightmaps (BASELINE.json configs[1])",
            "c3": "light styles 1..11 change every step: every surface-cache block that lists one is rebuilt per step (BASELINE.json configs[2])",
            "c4": "static lightmaps + 64 skinned MDL v10 studio models (32 bones, 1536 vertices, 2976 triangles each), animated: new bones every step (BASELINE.json configs[3])",
            "c5": "turbulent water on ~10 % of the floor + 32 translucent sprites sorted per view (BASELINE.json configs[4], per GPU)"}[args.config]
    mapname = {"small": "S (~1.7k faces; with pools ~2.2k)", "large": "L (~33k faces)", "medium": "M (~12.5k faces)", "tiny": "tiny"}[args.map]
    return {
        "workload": f"{args.config}: {args.views} batched camera views/GPU at {args.width}x{args.height}, synthetic BSP30 Map {mapname}, {what}",
        "config_id": args.config,
        "views_per_gpu": args.views, "width": args.width, "height": args.height, "map": args.map,
        "sharding": f"views x{world}, scene replicated, no data-path collective",
        "l2": "flushed with a 256 MB write between timed steps",
    }


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def measured_traffic(config_id, key="k_draw_dram_bytes_per_launch"):
    """dram bytes per launch of a kernel from the committed ncu capture of this workload (profiles/traffic.json), if one exists."""
    p = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(p):
        try:
            return json.load(open(p)).get(config_id, {}).get(key)
        except Exception:
            return None
    return None


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region (B200_PROFILING.md)."""

    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index=0):
        self.rows, self.proc = [], None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "20", "-i", str(index)],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], None, set()
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                mx = float(r[1])
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[3:7]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
            except Exception:
                pass
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons), "samples": len(sm)}


# ------------------------------------------------------------------------------------------------
# the reference's CPU implementation (oracle/_ref/libref_soft.so under the headless host)

def _soft_worker(conn, args_d, lo, hi, ncams):
    """one ref_soft instance (all-globals, single-threaded) per process, one process per host core"""
    try:
        try:
            os.sched_setaffinity(0, {args_d["cpu"]})
        except Exception:
            pass
        pkg = importlib.import_module("xash3d-fwgs_b200")
        scene = importlib.import_module("xash3d-fwgs_b200.scene")
        m = scene.make_map(args_d["map"], ncameras=ncams, **args_d["map_kw"])
        r = scene.Renderer(pkg.REF_SOFT_PATH, args_d["width"], args_d["height"], m, cvars=args_d["cvars"])
        cams = m.cameras[lo:hi]
        ns = argparse.Namespace(**args_d)
        populate(ns, scene, r, m)
        out = (np.empty((len(cams), args_d["height"], args_d["width"]), np.uint16), np.empty((len(cams), args_d["height"], args_d["width"]), np.int16))
        conn.send("ready")
        while True:
            cmd = conn.recv()
            if cmd == "quit":
                break
            advance(ns, r, cmd[1])
            t0 = time.perf_counter()
            r.render_views(cams, out=out)
            conn.send((time.perf_counter() - t0, r.last_seconds))
    except Exception as e:  # pragma: no cover
        conn.send(("error", repr(e)))


class SoftPool:
    def __init__(self, args, ncores):
        import multiprocessing as mp

        ctx = mp.get_context("spawn")
        self.n = min(ncores, args.views)
        cpus = sorted(os.sched_getaffinity(0))
        bounds = np.linspace(0, args.views, self.n + 1).astype(int)
        self.procs = []
        for i in range(self.n):
            a, b = ctx.Pipe()
            d = {"map": args.map, "width": args.width, "height": args.height, "cpu": cpus[i % len(cpus)], "map_kw": args.map_kw,
                 "cvars": args.cvars, "rebuild": args.rebuild, "models": args.models, "sprites": args.sprites}
            p = ctx.Process(target=_soft_worker, args=(b, d, int(bounds[i]), int(bounds[i + 1]), args.views), daemon=True)
            p.start()
            self.procs.append((p, a))
        for _, a in self.procs:
            msg = a.recv()
            if msg != "ready":
                raise RuntimeError(f"ref_soft worker failed: {msg}")

    def step(self, k=0):
        """all workers render their slice of the batch once; -> (wall seconds, max seconds inside R_RenderFrame)"""
        t0 = time.perf_counter()
        for _, a in self.procs:
            a.send(("go", k))
        res = [a.recv() for _, a in self.procs]
        wall = time.perf_counter() - t0
        if any(r[0] == "error" for r in res):
            raise RuntimeError(str(res))
        return wall, max(r[1] for r in res)

    def close(self):
        for p, a in self.procs:
            try:
                a.send("quit")
            except Exception:
                pass
        for p, _ in self.procs:
            p.join(timeout=5)


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    ncores = len(os.sched_getaffinity(0))
    pool = SoftPool(args, ncores)
    for k in range(args.warmup):
        pool.step(k)
    walls, inner = [], []
    for k in range(args.steps):
        w, i = pool.step(args.warmup + k)
        walls.append(w)
        inner.append(i)
    pool.close()
    px = args.views * args.width * args.height
    # time of a step = the slowest process's time inside GL_RenderFrame (R_RenderFrame only)
    ms = 1e3 * float(np.mean(inner))
    val = px / (ms * 1e3)
    sample = f"{args.steps} passes of the {args.views}-view batch split over {pool.n} processes (one ref_soft per core)"
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "int32/f32 (16-bit pixel_t)",
        "data": "synthetic", "config": workload(args, 1),
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": pool.n, "kind": "reference", "sample": sample,
                         "wall_ms_per_step_incl_ipc": 1e3 * float(np.mean(walls))},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def cpu_baseline_one_core(args):
    """the oracle on ONE host core, a bounded sample of the same workload (reported baseline, not the target)"""
    pkg = importlib.import_module("xash3d-fwgs_b200")
    scene = importlib.import_module("xash3d-fwgs_b200.scene")
    m = scene.make_map(args.map, ncameras=args.views, **args.map_kw)
    r = scene.Renderer(pkg.REF_SOFT_PATH, args.width, args.height, m, cvars=args.cvars)
    populate(args, scene, r, m)
    cams = m.cameras[: min(args.views, 256)]
    out = (np.empty((len(cams), args.height, args.width), np.uint16), np.empty((len(cams), args.height, args.width), np.int16))
    r.render_views(cams, out=out)  # warm: fills the surface cache
    secs, passes, t0 = 0.0, 0, time.perf_counter()
    while time.perf_counter() - t0 < args.cpu_seconds and passes < 1000:
        advance(args, r, passes + 1)
        r.render_views(cams, out=out)
        secs += r.last_seconds
        passes += 1
    px = passes * len(cams) * args.width * args.height
    return {"value": px / secs / 1e6, "unit": UNIT, "cores": 1, "kind": "reference",
            "sample": f"{passes} passes of {len(cams)} views of the batch ({secs:.1f} s inside GL_RenderFrame), "
                      + ("light styles advanced before every pass" if args.rebuild else "warm surface cache") + ", gcc -O2"}


# ------------------------------------------------------------------------------------------------
# our arm: libref_cuda.so

def run_cuda(args):
    import torch
    import torch.distributed as dist

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device - ref/cuda has no CPU path (use --impl reference for ref_soft)")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def allmax(x):
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    pkg = importlib.import_module("xash3d-fwgs_b200")
    scene = importlib.import_module("xash3d-fwgs_b200.scene")
    shard = importlib.import_module("xash3d-fwgs_b200.shard")
    m = scene.make_map(args.map, ncameras=args.views * world, **args.map_kw)
    lo, hi = shard.view_range(args.views * world, rank, world)  # this rank's slice of the global camera list
    cams = m.cameras[lo:hi]
    r = scene.Renderer(pkg.REF_CUDA_PATH, args.width, args.height, m, cvars={"cu_device": local})
    populate(args, scene, r, m)
    advance(args, r, 0)
    n, w, h = len(cams), args.width, args.height
    px = n * w * h

    # pinned host landing zone for the e2e leg (torch only provides the pinned allocation)
    pin = torch.empty((n, h, w), dtype=torch.int16).pin_memory()
    pin_z = torch.empty((n, h, w), dtype=torch.int16).pin_memory()
    vp = r.viewpasses(cams)
    r.render_prepared(vp, pin.data_ptr(), None)  # uploads the batch, builds the surface cache blocks it needs
    ctr0 = r.counters()
    overflow = {k: v for k, v in ctr0.items() if k.startswith("overflow") and v}
    if overflow:
        raise SystemExit(f"bench.py: capacity overflow {overflow} - numbers would be invalid")

    # -- value: device stages, inputs resident, CUDA events inside the plugin, L2 flushed between steps
    r.bench_resident(max(args.warmup, 3), flush_l2=True, rebuild_cache=args.rebuild)
    sampler = ClockSampler(local) if rank == 0 else None
    barrier()
    l0 = r.launch_count()
    ms, stages = r.bench_resident(args.steps, flush_l2=True, rebuild_cache=args.rebuild)
    launches = r.launch_count() - l0
    barrier()
    ms_step = allmax(ms)
    ms_scan = allmax(stages[4])
    allmax_entities = allmax(stages[5])

    # -- second HBM-bound stage: every surface-cache block the batch samples rebuilt (config C3's
    #    behaviour under animated light styles), same batch, outside the headline timing
    _, st_rb = r.bench_resident(max(3, args.steps // 2), rebuild_cache=True, flush_l2=True, serial_cache=True)  # cache kernels timed alone
    ms_cache = allmax(st_rb[2])
    cache_texels = r.counters()["cache_texels"]

    # -- display tail: the batch's frames, resident in HBM, converted to XRGB8888 by k_blit
    #    (2 B read + 4 B written per pixel), outside the headline timing; kernel time from CUDA events in the plugin
    import ctypes as _C
    blit_ms = []
    _ms = _C.c_float()
    for _ in range(max(args.warmup, 3) + max(3, args.steps // 4)):
        if r._blit(0, n, None, w, 4, 0x00FF0000, 0x0000FF00, 0x000000FF, _C.byref(_ms)) != 0:
            raise SystemExit("bench.py: RefCuda_BlitViews failed")
        blit_ms.append(_ms.value)
    ms_blit = allmax(float(np.median(blit_ms[max(args.warmup, 3):])))

    # -- e2e: the public C-ABI call with host buffers (H2D view descriptors, D2H colour frames), wall clock
    def e2e_leg(zptr, pipelined):
        """the public batch call with host buffers.  pipelined: RefCuda_RenderViewsAsync with two alternating sets of pinned
        buffers and RefCuda_Wait before a set is handed out again (every step's frames are complete in host memory one step
        later); else the blocking RefCuda_RenderViews."""
        outs = [(pin.data_ptr(), zptr), (pin_b.data_ptr(), zptr_b if zptr else None)]

        def step(k):
            advance(args, r, k)
            if pipelined:
                r.render_prepared_async(vp, *outs[k & 1])
            else:
                r.render_prepared(vp, pin.data_ptr(), zptr)

        for k in range(max(args.warmup, 3)):
            step(k + 1)
        r.wait()
        barrier()
        t0 = time.perf_counter()
        for k in range(args.steps):
            step(k + 100)
        r.wait()
        torch.cuda.synchronize()
        secs = allmax((time.perf_counter() - t0) / args.steps)
        barrier()
        return secs

    pin_b = torch.empty((n, h, w), dtype=torch.int16).pin_memory()
    zptr_b = torch.empty((n, h, w), dtype=torch.int16).pin_memory()
    zptr_b_keep = zptr_b
    zptr_b = zptr_b.data_ptr()
    e2e_sync_s = e2e_leg(None, False)
    e2e_s = e2e_leg(None, True)
    h2d = r.batch_h2d_bytes()
    e2e_z_s = e2e_leg(pin_z.data_ptr(), True)
    clocks = sampler.stop() if sampler else None

    # counters of the last batch, summed over ranks (the only collective: < 1 KiB, off the data path)
    names = list(ctr0.keys())
    ctr = torch.tensor([float(r.counters()[k]) for k in names], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(ctr, op=dist.ReduceOp.SUM)
    counters = {k: int(v) for k, v in zip(names, ctr.tolist())}

    if rank == 0:
        peak, peak_src = peaks()
        total_px = px * world
        achieved = BYTES_PER_PIXEL * px / (ms_scan * 1e-3) / 1e9
        line = {
            "metric": METRIC, "value": total_px / (ms_step * 1e3), "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "int32/f32 (16-bit pixel_t)", "data": "synthetic", "config": workload(args, world),
            "frames_per_s": n * world / (ms_step * 1e-3),
            "stage_ms": {"vis_walk": stages[0], "faces_edges": stages[1], "surface_cache": stages[2], "spans": stages[3], "draw": stages[4], "entities": stages[5],
                         "note": "the surface-cache kernels run on a side stream beside span generation: their time overlaps spans and is not part of the sum"},
            "e2e": {"value": total_px / e2e_s / 1e6, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": px * 2 + 512,
                    "ms_per_step": e2e_s * 1e3,
                    "blocking_call_ms_per_step": e2e_sync_s * 1e3, "blocking_call_value": total_px / e2e_sync_s / 1e6,
                    "note": "RefCuda_RenderViewsAsync + RefCuda_Wait, two alternating sets of pinned host buffers (batch k's copies run under batch k+1's stages; "
                            "blocking_call_* = the same batch through the blocking RefCuda_RenderViews): per-view constants (+ light styles when they change) host->device, colour frames "
                            "(what an observation consumer takes) device->host into pinned memory, 512 B of counters"},
            "e2e_colour_and_z": {"value": total_px / e2e_z_s / 1e6, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": px * 4 + 512,
                                 "ms_per_step": e2e_z_s * 1e3, "note": "the same call with the z buffers copied out as well (what the reference arm leaves in host memory)"},
            "gpu_launches": int(launches),
            "roofline": {"bound": "hbm", "kernel": "k_draw", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": measured_traffic(args.config) if args.views == {"c2": 256, "c3": 8, "c4": 256, "c5": 4096}[args.config] else None, "algorithmic_bytes_per_launch": BYTES_PER_PIXEL * px, "peak_source": peak_src,
                         "kernel_ms": ms_scan},
            "roofline_surface_cache": {"bound": "hbm", "kernel": "k_cache_build (+ k_cache_list)", "achieved": 4 * cache_texels / (ms_cache * 1e-3) / 1e9,
                                       "peak": peak, "unit": "GB/s", "frac": 4 * cache_texels / (ms_cache * 1e-3) / 1e9 / peak,
                                       "algorithmic_bytes_per_launch": 4 * cache_texels, "kernel_ms": ms_cache, "traffic": measured_traffic(args.config, "k_cache_build_dram_bytes_per_launch"),
                                       "note": "forced rebuild of every (face, mip) block this rank's batch samples: 4 B per cache texel (2 B texel read + 2 B written), luxel bytes not counted"},
            "value_note": ("value is timed with EVERY surface-cache block the batch samples invalidated before each step (a superset of the blocks whose light styles changed)" if args.rebuild else None),
            "roofline_display_tail": {"bound": "hbm", "kernel": "k_blit<4> (pixel_t -> XRGB8888)", "achieved": 6 * px / (ms_blit * 1e-3) / 1e9, "peak": peak,
                                      "unit": "GB/s", "frac": 6 * px / (ms_blit * 1e-3) / 1e9 / peak, "algorithmic_bytes_per_launch": 6 * px,
                                      "kernel_ms": ms_blit, "traffic": None, "note": "2 B pixel_t read + 4 B display pixel written; 256 KB screen map in L2; not part of value / e2e"},
            "clocks": clocks, "counters": counters, "device": r.device_name(),
        }
        if (args.models or args.sprites) and stages[5] > 0:
            # per tested fragment 2 B of z read; per written fragment z + colour written and a skin texel read (6 B)
            ent_bytes = 2 * counters["tri_pixels"] + 6 * max(counters["tri_written"], 0)
            ent_ms = allmax_entities
            line["roofline_entities"] = {
                "bound": "hbm", "kernel": "k_studio_points + k_tris_fast + k_tris_fast_big + k_tri_resolve (opaque run) / k_item_order + k_tri_polys + k_tris_ordered (sorted run)",
                "achieved": ent_bytes / world / (ent_ms * 1e-3) / 1e9, "peak": peak, "unit": "GB/s", "frac": ent_bytes / world / (ent_ms * 1e-3) / 1e9 / peak,
                "algorithmic_bytes_per_launch": ent_bytes // world, "kernel_ms": ent_ms, "traffic": None,
                "triangles_set_up_per_s": counters["tris"] / world / (ent_ms * 1e-3), "fragments_tested_per_s": counters["tri_pixels"] / world / (ent_ms * 1e-3),
                "submitted_triangle_view_pairs_per_s": (args.models * 2976 * n) / (ent_ms * 1e-3) if args.models else None,
                "note": "the triangle stage is set-up- and atomic-bound, not HBM-bound: 2 B per z-tested fragment + 6 B per fragment that ends up visible "
                        "(k_tri_resolve shades each pixel once; the reference writes every fragment that passes at its turn)"}
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_baseline_one_core(args)
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    a = parse_args()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_cuda(a)
