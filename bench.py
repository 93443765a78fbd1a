#!/usr/bin/env python3
"""bench.py — Mpaths/s of the wavefront path tracer on the C2 workload of BASELINE.json:
Book-3 Cornell box with glass sphere and aluminium box, 600x600, 1024 spp, depth 20.

  python bench.py --gpus N --steps K --warmup W            (N>1: launched under torchrun)
  python bench.py --impl reference ...                     (the CPU oracle on the host cores)

A "step" is one whole frame.  At N GPUs rank g renders the samples k = g (mod N) of every
pixel into its own fp32 accumulator, one NCCL reduce per frame combines them on rank 0
(SURVEY.md §8e), rank 0 tonemaps.  Total work is fixed → "scaling": "strong".
One JSON line on stdout (rank 0).
"""
import argparse
import importlib
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

NX, NY, SPP, DEPTH, SCENE = 600, 600, 1024, 20, 0
WORKLOAD = "C2: Book-3 Cornell box (glass sphere + aluminium box), 600x600, 1024 spp, depth 20"
METRIC = "Mpaths/s (camera samples completed per second, Cornell 600x600 1024spp)"


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


class ClockSampler:
    """SM clocks + throttle reasons DURING the timed region (B200_PROFILING.md's clocks line).  Read through NVML
    inside this process (the same counters nvidia-smi prints; initialised BEFORE the timed region) by rank 0 for every
    GPU of the run: one `nvidia-smi` process per rank starting inside a 50 ms timed region cost the 8-GPU run 5 %.
    Falls back to the nvidia-smi query of the recipe if NVML cannot be loaded."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
    BITS = (("sw_power_cap", 0x4), ("hw_slowdown", 0x8), ("sw_thermal_slowdown", 0x20), ("hw_thermal_slowdown", 0x40))

    def __init__(self, gpu_indices, enabled=True):
        self.rows = []          # (sm_mhz, sm_max_mhz, power_w, [reasons])
        self.gpus = list(gpu_indices)
        self.enabled = enabled
        self.stop = threading.Event()
        self.th = threading.Thread(target=self._run, daemon=True)
        self.nvml = None
        self.handles = []
        self.source = "none"
        if not enabled:
            return
        try:
            import pynvml
            pynvml.nvmlInit()
            self.handles = [pynvml.nvmlDeviceGetHandleByIndex(i) for i in self.gpus]
            self.nvml = pynvml
            self.source = "nvml"
        except Exception:
            self.source = "nvidia-smi"

    def _sample(self):
        if self.nvml is not None:
            n = self.nvml
            for h in self.handles:
                try:
                    sm = float(n.nvmlDeviceGetClockInfo(h, n.NVML_CLOCK_SM))
                    mx = float(n.nvmlDeviceGetMaxClockInfo(h, n.NVML_CLOCK_SM))
                    pw = n.nvmlDeviceGetPowerUsage(h) / 1000.0
                    get = getattr(n, "nvmlDeviceGetCurrentClocksEventReasons", None) or n.nvmlDeviceGetCurrentClocksThrottleReasons
                    mask = int(get(h))
                    self.rows.append((sm, mx, pw, [name for name, bit in self.BITS if mask & bit]))
                except Exception:
                    pass
            return
        try:
            out = subprocess.run(["nvidia-smi", "--query-gpu=" + self.Q, "--format=csv,noheader,nounits", "-i",
                                  ",".join(str(i) for i in self.gpus)], capture_output=True, text=True, timeout=5).stdout.strip()
            for line in out.splitlines():
                r = [c.strip() for c in line.split(",")]
                reasons = [name for name, v in zip(["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"], r[5:9])
                           if v.lower().startswith("active")]
                self.rows.append((float(r[1]), float(r[2]), float(r[3]), reasons))
        except Exception:
            pass

    def _run(self):
        while not self.stop.is_set():
            self._sample()
            self.stop.wait(0.02 if self.nvml is not None else 0.2)

    def __enter__(self):
        if self.enabled:
            self.th.start()
        return self

    def __exit__(self, *a):
        self.stop.set()
        if self.enabled:
            self.th.join(timeout=6)

    def summary(self):
        if not self.rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["unsampled"], "source": self.source}
        sm = sorted(r[0] for r in self.rows)
        reasons = set()
        for r in self.rows:
            reasons.update(r[3])
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": self.rows[0][1], "power_w_max": max(r[2] for r in self.rows),
                "samples": len(self.rows), "reasons": sorted(reasons), "source": self.source, "gpus": self.gpus}


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def oracle_scene(rtw):
    from oracle import oracle_py as orc
    s = rtw.Scene(SCENE, NX, NY)
    return s, orc.OracleScene(s.desc, s.desc.contents.n_prims)


def cpu_sample(rtw, budget_s):
    """Time the oracle (all host threads) on a bounded sample of the workload: whole-image
    passes of `spp` samples, sized from a 1 spp calibration pass to take about budget_s."""
    s, o = oracle_scene(rtw)
    cores = host_cores()
    t0 = time.perf_counter()
    _, c = o.render(NX, NY, 1, DEPTH, n_threads=cores, mode=0)
    t1 = time.perf_counter() - t0
    spp = int(max(1, min(64, budget_s / max(t1, 1e-3))))
    t0 = time.perf_counter()
    _, c = o.render(NX, NY, spp, DEPTH, n_threads=cores, mode=0, sample_first=1)
    dt = time.perf_counter() - t0
    s.close()
    return {"value": c["paths"] / dt / 1e6, "unit": "Mpaths/s", "cores": cores, "kind": "port",
            "sample": "%dx%d, %d spp (samples 1..%d of the 1024), depth %d, brute-force closest hit over 13 primitives, %.1f s"
                      % (NX, NY, spp, spp, DEPTH, dt),
            "mrays_per_s": (c["rays_closest"] + c["rays_shadow"]) / dt / 1e6}, dt, c


def run_reference(args):
    """--impl reference: the reference's OptiX build cannot be compiled here (no OptiX SDK,
    SURVEY.md §8c), so this arm times the CPU oracle — the restatement of the reference's
    own algorithm — on all host threads, same config, bounded sample per step."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    rtw = importlib.import_module("raytracing-weekend_b200")
    s, o = oracle_scene(rtw)
    cores = host_cores()
    spp = 32          # 11.5 M paths per step: about a second on 16 host threads
    for _ in range(args.warmup):
        o.render(NX, NY, 1, DEPTH, n_threads=cores, mode=0)
    paths = 0
    t0 = time.perf_counter()
    for i in range(args.steps):
        _, c = o.render(NX, NY, spp, DEPTH, n_threads=cores, mode=0, sample_first=spp * i)
        paths += c["paths"]
    dt = time.perf_counter() - t0
    v = paths / dt / 1e6
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": "Mpaths/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": WORKLOAD, "note": "CPU oracle (port of the reference algorithm); each step = a bounded "
                       "sample of %d spp of the 1024" % spp},
            "cpu_baseline": {"value": v, "unit": "Mpaths/s", "cores": cores, "kind": "port",
                             "sample": "%d steps x (%dx%d, %d spp)" % (args.steps, NX, NY, spp)},
            "e2e": {"value": v, "unit": "Mpaths/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)
    return 0


def main():
    global NX, NY, SCENE, WORKLOAD
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours")
    ap.add_argument("--spp", type=int, default=SPP, help="override for quick checks (not the benchmark config)")
    ap.add_argument("--pool", type=int, default=0, help="wavefront pool size override")
    ap.add_argument("--scene", type=int, default=SCENE, help="override for profiling other configs (not the benchmark config)")
    ap.add_argument("--nx", type=int, default=NX)
    ap.add_argument("--ny", type=int, default=NY)
    ap.add_argument("--leaf-size", type=int, default=0)
    ap.add_argument("--opt", action="append", default=[], help="name=value passed to rtw_set_option (tuning runs)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-relmse", action="store_true", help="skip the tier-2 relMSE measurement (N=1 only; ~10 s of GPU time)")
    args = ap.parse_args()
    if (args.scene, args.nx, args.ny) != (SCENE, NX, NY):
        SCENE, NX, NY = args.scene, args.nx, args.ny
        WORKLOAD = "OVERRIDE (not the benchmark config): scene %d, %dx%d, 1024 spp, depth 20" % (SCENE, NX, NY)
    if args.impl == "reference":
        return run_reference(args)

    # stdout carries exactly ONE JSON line: everything else that libraries write to fd 1 (NCCL's version
    # banner, extension build chatter) is sent to stderr; the line itself goes to the saved descriptor
    sys.stdout.flush()
    json_out = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    import numpy as np
    import torch
    import torch.distributed as dist
    rtw = importlib.import_module("raytracing-weekend_b200")
    sharding = importlib.import_module("raytracing-weekend_b200.sharding")

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            raise SystemExit("--gpus %d needs torchrun with %d ranks" % (args.gpus, args.gpus))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        # NCCL prints its version banner on stdout; stdout carries exactly one JSON line
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
        dist.init_process_group("nccl", device_id=dev)
    spp = args.spp
    # rank g renders samples k = g (mod world): counts differ by at most one
    my_first, my_stride, my_count = sharding.shard_samples(spp, rank, world)

    d = rtw.Director(local)
    d.init(NX, NY, spp)
    if args.leaf_size:
        d.set_option("leaf_size", args.leaf_size)
    for kv in args.opt:             # build-time options (walker, leaf_size, top_nodes) must precede createScene
        k, v = kv.split("=")
        d.set_option(k, int(v))
    d.createScene(SCENE)
    if args.pool:
        d.set_option("pool_paths", args.pool)
    d.set_option("profile", 1)
    for kv in args.opt:
        k, v = kv.split("=")
        d.set_option(k, int(v))
    scene_desc = d.scene.desc
    dsc = scene_desc.contents
    import ctypes as C
    h2d = (dsc.n_prims * C.sizeof(rtw.Prim) + dsc.n_transforms * C.sizeof(rtw.Transform) +
           dsc.n_materials * C.sizeof(rtw.Material) + dsc.n_textures * C.sizeof(rtw.Texture) + dsc.n_lights * C.sizeof(rtw.Light))

    accum = torch.zeros(NY, NX, 4, device=dev)
    out = torch.empty(NY, NX, 4, device=dev)
    host_img = torch.empty(NY, NX, 4).pin_memory()
    launches = {"n": 0}
    trace = {}      # host-side wall time of the phases of a frame on this rank (rank 0's goes into the line)
    agg = {"trav_ms": 0.0, "shadow_ms": 0.0, "shade_ms": 0.0, "logic_ms": 0.0, "render_ms": 0.0, "rays": 0, "shrays": 0, "iters": 0,
           "waves": 1, "pool": 0}

    def sync_all():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def frame(count_stats):
        """One frame with the scene resident: zero, render my samples, reduce, tonemap on root."""
        ts = [time.perf_counter()]

        def render_shard(f, sd, n):
            ts.append(time.perf_counter())          # after zero + synchronize
            d.render_accum(accum.data_ptr(), f, sd, n, DEPTH)
            ts.append(time.perf_counter())          # after this rank's samples
        sharding.render_frame_sharded(render_shard, accum, spp, rank, world)
        st = d.stats()
        n_l = st["kernel_launches"] + 1
        if world > 1:
            torch.cuda.synchronize()
        ts.append(time.perf_counter())              # after the reduce
        if rank == 0:
            d.tonemap(accum.data_ptr(), out.data_ptr(), spp)
            n_l += 1
        ts.append(time.perf_counter())
        if count_stats:
            for k, name in enumerate(("zero_ms", "render_call_ms", "reduce_ms", "tonemap_ms")):
                trace[name] = trace.get(name, 0.0) + (ts[k + 1] - ts[k]) * 1e3
        if count_stats:
            launches["n"] += n_l
            agg["trav_ms"] += st["traverse_ms"]; agg["shadow_ms"] += st["shadow_ms"]; agg["shade_ms"] += st["shade_ms"]
            agg["logic_ms"] += st["logic_ms"]; agg["render_ms"] += st["render_ms"]
            agg["rays"] += st["rays_closest"]; agg["shrays"] += st["rays_shadow"]; agg["iters"] += st["iterations"]
            agg["waves"] = st["waves"]; agg["pool"] = st["pool_paths"]

    def frame_e2e():
        """Same frame through the public API from HOST buffers: scene upload + LBVH build +
        render + reduce + tonemap + image back to pinned host memory."""
        d.upload(scene_desc)
        frame(False)
        if rank == 0:
            host_img.copy_(out, non_blocking=True)
            torch.cuda.synchronize()

    for _ in range(args.warmup):
        frame(False)
    sync_all()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with ClockSampler(list(range(world)) if world > 1 else [local], enabled=(rank == 0)) as clk:
        ev0.record()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            frame(True)
        ev1.record()
        sync_all()
        wall = time.perf_counter() - t0
    dev_ms = ev0.elapsed_time(ev1)
    t = torch.tensor([dev_ms], device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dev_ms = float(t.item())

    # end to end (host buffers in, host image out)
    frame_e2e()
    sync_all()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        frame_e2e()
    e1.record()
    sync_all()
    t = torch.tensor([e0.elapsed_time(e1)], device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_ms = float(t.item())

    # totals over ranks for the ray statistics
    tot = torch.tensor([agg["rays"], agg["shrays"], launches["n"]], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(tot)
    if rank == 0:
        paths = float(NX * NY * spp) * args.steps
        value = paths / (dev_ms * 1e-3) / 1e6
        e2e = paths / (e2e_ms * 1e-3) / 1e6
        peak, peak_src = measured_peaks()
        pool, waves = int(agg["pool"]), int(agg["waves"])   # what the library chose (api/rtw_render.cu) or --pool / --opt waves=
        line = {"metric": METRIC, "value": value, "unit": "Mpaths/s", "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": dev_ms / args.steps, "higher_is_better": True, "scaling": "strong",
                "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": {"workload": WORKLOAD if spp == SPP else WORKLOAD.replace("1024 spp", "%d spp (override)" % spp),
                           "parallelism": "sample-sharded x%d, one NCCL reduce per frame; on each GPU the samples are dealt to "
                                          "%d concurrent pools (waves) on separate streams" % (world, waves),
                           "l2_policy": "inputs larger than L2: the wavefront pools (%d x %d path slots per GPU x 160 B state + queues = %d MB) "
                                        "stream through HBM every iteration, no flush between steps" % (waves, pool, waves * pool * 160 >> 20),
                           "wall_ms_per_step": wall / args.steps * 1e3,
                           "frame_phases_ms_rank0": {k: v / args.steps for k, v in trace.items()}},
                "mrays_per_s": (tot[0].item() + tot[1].item()) / (dev_ms * 1e-3) / 1e6,
                "e2e": {"value": e2e, "unit": "Mpaths/s", "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": NX * NY * 16},
                "gpu_launches": int(tot[2].item()), "clocks": clk.summary()}
        # ---- roofline of the dominant kernel: k_traverse (closest-hit traversal) ----
        # HBM-visible ALGORITHMIC bytes per ray: the 32-byte ray record read,
        # hit 8 + queue entry 4 written = 44 B.  The BVH nodes / primitive records of SURVEY.md §8d
        # (64 B x nodes + 32 B x primitives, +48 B per transformed run) are served by L1 / shared memory
        # for this scene; they are reported beside it from the ORACLE's visit counters on its mirror
        # of the hierarchy, not from the GPU.
        s2, o2 = oracle_scene(rtw)
        _, c = o2.render(NX, NY, 1, DEPTH, n_threads=host_cores(), mode=1)
        s2.close()
        rc = max(c["rays_closest"], 1)
        onchip_per_ray = (64.0 * c["nodes_closest"] + 32.0 * (c["prims_closest"] - c["prims_xf_closest"]) +
                          80.0 * c["prims_xf_closest"]) / rc
        hbm_per_ray = 44.0
        iters = max(agg["iters"], 1)
        trav_s = agg["trav_ms"] * 1e-3
        tp = os.path.join(ROOT, "profiles", "traffic.json")      # dram bytes per launch from the committed ncu capture
        traffic_tab = json.load(open(tp)).get("dram_bytes_per_launch", {}) if os.path.exists(tp) else {}
        # per-stage picture of one wavefront iteration (HBM-visible algorithmic bytes per unit)
        n_paths = float(NX * NY * spp) * args.steps / world
        stages = {}
        fin = n_paths / max(agg["rays"], 1)          # share of shaded vertices that end their path
        # k_shade_all per vertex: queue entry 4 + ray 32 + path 32 + hit 8 read; a continuing path writes ray 32 +
        # path 32, an ending one its radiance record (16 R + 16 W) and a regen entry 4; + 48 per NEE ray emitted
        shade_bpu = 76.0 + 64.0 * (1.0 - fin) + 36.0 * fin + 48.0 * agg["shrays"] / max(agg["rays"], 1)
        for name, ms, units, bpu in (("k_traverse", agg["trav_ms"], agg["rays"], hbm_per_ray),
                                     ("k_shade*", agg["shade_ms"], agg["rays"], shade_bpu),
                                     ("k_shadow", agg["shadow_ms"], agg["shrays"], 72.0),
                                     ("k_regen", agg["logic_ms"], n_paths, 116.0)):
            gbs = units * bpu / max(ms * 1e-3, 1e-9) / 1e9
            stages[name] = {"ms_per_iteration": ms / iters, "units_per_iteration": units / iters, "bytes_per_unit": bpu,
                            "gbs": gbs, "frac": gbs / peak, "kernel_ms_over_render_ms": ms / max(agg["render_ms"], 1e-9)}
        ksum = sum(v["ms_per_iteration"] for v in stages.values())
        for name in stages:
            stages[name]["traffic"] = traffic_tab.get(name)
            stages[name]["share_of_kernel_time"] = stages[name]["ms_per_iteration"] / max(ksum, 1e-12)
        # the dominant kernel is the stage with the largest share of the kernel time
        dom = max(stages, key=lambda k: stages[k]["ms_per_iteration"])
        ds = stages[dom]
        # the whole step against the same roof: algorithmic HBM bytes of every stage / device time of the frames
        step_bytes = sum(v["units_per_iteration"] * v["bytes_per_unit"] for v in stages.values()) * iters
        step_gbs = step_bytes / max(agg["render_ms"] * 1e-3, 1e-9) / 1e9
        line["roofline"] = {"bound": "hbm", "kernel": dom, "achieved": ds["gbs"], "peak": peak, "unit": "GB/s",
                            "frac": ds["frac"], "traffic": ds["traffic"], "peak_source": peak_src,
                            "bytes_per_unit": ds["bytes_per_unit"], "units_per_launch": ds["units_per_iteration"],
                            "avg_launch_ms": ds["ms_per_iteration"],
                            "step": {"achieved": step_gbs, "frac": step_gbs / peak, "bytes_per_frame": step_bytes / max(args.steps, 1),
                                     "note": "all stages' algorithmic HBM bytes / device time of the timed frames"},
                            "traversal_on_chip_bytes_per_ray_survey_8d": onchip_per_ray + 96.0,
                            "traversal_on_chip_gbs": agg["rays"] * onchip_per_ray / max(trav_s, 1e-9) / 1e9,
                            "note": "achieved = algorithmic HBM bytes per launch (DESIGN.md 4.2) / average launch time measured "
                                    "live with CUDA events on wave 0's stream inside the timed frames; with %d waves a launch shares the "
                                    "device with the other waves' launches, so per-launch figures understate what the GPU moves: "
                                    "`step` is the whole frame against the same roof, `alone` the same kernels timed with one wave. "
                                    "traffic = dram bytes of one launch from the committed ncu capture. k_shade* gathers 64-byte slot "
                                    "records through queue indices; the traversal kernels are instruction-issue bound: a "
                                    "13-primitive scene is one leaf range passed by value, so their HBM share is only the state "
                                    "stream" % waves,
                            "stages": stages}
        if waves > 1:   # the same kernels timed without a concurrent wave (one untimed extra frame)
            d.set_option("waves", 1)
            for _ in range(2):      # the first frame re-allocates the pool
                d.render_accum(accum.data_ptr(), 0, 1, max(spp // world, 1), DEPTH)
            s1 = d.stats()
            d.set_option("waves", 0)
            it1 = max(s1["iterations"], 1)
            key = {"k_traverse": "traverse_ms", "k_shade*": "shade_ms", "k_shadow": "shadow_ms", "k_regen": "logic_ms"}[dom]
            units1 = {"k_traverse": s1["rays_closest"], "k_shade*": s1["rays_closest"], "k_shadow": s1["rays_shadow"],
                      "k_regen": s1["paths"]}[dom]
            g1 = units1 * ds["bytes_per_unit"] / max(s1[key] * 1e-3, 1e-9) / 1e9
            line["roofline"]["alone"] = {"waves": 1, "pool_paths": int(s1["pool_paths"]), "avg_launch_ms": s1[key] / it1,
                                         "units_per_launch": units1 / it1, "achieved": g1, "frac": g1 / peak,
                                         "frame_ms": s1["render_ms"]}
        if world == 1 and not args.no_cpu_baseline:
            cb, _, _ = cpu_sample(rtw, 15.0)
            line["cpu_baseline"] = cb
        json_out.write(json.dumps(line) + "\n")
        json_out.flush()
    d.destroy()
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
