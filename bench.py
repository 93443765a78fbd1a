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

# BASELINE.json's configurations as first-class bench lines (--config); C2 is the one `metric` is quoted on.
CONFIGS = {
    "C2": dict(scene=0, nx=600, ny=600, spp=1024,
               workload="C2: Book-3 Cornell box (glass sphere + aluminium box), 600x600, 1024 spp, depth 20",
               metric="Mpaths/s (camera samples completed per second, Cornell 600x600 1024spp)"),
    "C3": dict(scene=1, nx=1920, ny=1080, spp=500,
               workload="C3: Book-1 final random-spheres scene (528 primitives), 1920x1080, 500 spp, depth 20",
               metric="Mpaths/s (camera samples completed per second, Book-1 final 1920x1080 500spp)"),
    "C4": dict(scene=4, nx=800, ny=800, spp=10000,
               workload="C4: Book-2 final scene (3410 primitives: box grid, fog, Perlin, image texture), 800x800, 10000 spp, depth 20",
               metric="Mpaths/s (camera samples completed per second, Book-2 final 800x800 10000spp)"),
    "C5": dict(scene=5, nx=3840, ny=2160, spp=4096,
               workload="C5: stress scene (2^20 instanced spheres / quads, 1024 transforms), 3840x2160, 4096 spp, depth 20",
               metric="Mpaths/s (camera samples completed per second, stress scene 3840x2160 4096spp)"),
}
DEPTH = 20
NX, NY, SPP, SCENE = CONFIGS["C2"]["nx"], CONFIGS["C2"]["ny"], CONFIGS["C2"]["spp"], CONFIGS["C2"]["scene"]
WORKLOAD, METRIC = CONFIGS["C2"]["workload"], CONFIGS["C2"]["metric"]


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


STRESS_PRIMS = 0


def oracle_scene(rtw):
    from oracle import oracle_py as orc
    s = rtw.Scene(SCENE, NX, NY, n_stress_prims=STRESS_PRIMS)
    return s, orc.OracleScene(s.desc, s.desc.contents.n_prims)


def reference_available():
    """oracle/_ref/libref.so = the reference's own programs compiled for the host (oracle/Makefile `ref`).  Scenes 2 and 4
    read assets/earthmap.jpg from /root/reference, which the GPU box does not have."""
    try:
        from oracle import ref_py
    except Exception:
        return None
    if not ref_py.available(False):
        return None
    if SCENE in (2, 4) and not ref_py.assets_available():
        return None
    if SCENE not in (0, 3):
        # the stand-in's traversal is a loop over ALL instances (no hierarchy): representative for the 13 / 8 primitive
        # scenes only; larger scenes are timed with the CPU oracle on its host-mirror LBVH (kind "port")
        return None
    return ref_py


def reference_sample(budget_s, launches=None):
    """Time THE REFERENCE ITSELF (kind "reference") on all host threads: Director::init + createScene once, then
    renderFrame() repeatedly.  One launch = the reference's one and only frame, 1 spp of the workload's image
    (raygen/raygen.cu:129 seeds sample 0; its -ns flag is stored and never used), so `launches` launches
    trace launches x Nx x Ny paths of the same statistics as the workload's samples."""
    ref_py = reference_available()
    cores = host_cores()
    s = ref_py.Session(SCENE, NX, NY, dm=False, threads=cores)
    t0 = time.perf_counter()
    s.launch(1)
    t1 = time.perf_counter() - t0
    n = launches if launches else int(max(1, min(256, budget_s / max(t1, 1e-3))))
    t0 = time.perf_counter()
    rays = s.launch(n)
    dt = time.perf_counter() - t0
    paths = float(NX) * NY * n
    return {"value": paths / dt / 1e6, "unit": "Mpaths/s", "cores": cores, "kind": "reference",
            "sample": "%d launches of the reference's own programs (oracle/_ref/libref.so: its main.cpp / Director.cpp / device .cu "
                      "files compiled for the host on the OptiX stand-in of oracle/ref_shim, glibc libm), each the %dx%d frame at 1 spp, "
                      "depth %d, instance loop over the scene's primitives, %.1f s" % (n, NX, NY, DEPTH, dt),
            "mrays_per_s": rays / dt / 1e6}, dt, n


def cpu_sample(rtw, budget_s):
    """Time the oracle (all host threads) on a bounded sample of the workload: whole-image
    passes of `spp` samples, sized from a 1 spp calibration pass to take about budget_s."""
    s, o = oracle_scene(rtw)
    cores = host_cores()
    mode = 0 if s.desc.contents.n_prims <= 64 else 1
    t0 = time.perf_counter()
    _, c = o.render(NX, NY, 1, DEPTH, n_threads=cores, mode=mode)
    t1 = time.perf_counter() - t0
    spp = int(max(1, min(64, budget_s / max(t1, 1e-3))))
    t0 = time.perf_counter()
    _, c = o.render(NX, NY, spp, DEPTH, n_threads=cores, mode=mode, sample_first=1)
    dt = time.perf_counter() - t0
    s.close()
    return {"value": c["paths"] / dt / 1e6, "unit": "Mpaths/s", "cores": cores, "kind": "port",
            "sample": "%dx%d, %d spp (samples 1..%d of the %d), depth %d, %s, %.1f s"
                      % (NX, NY, spp, spp, SPP, DEPTH, "brute-force closest hit" if mode == 0 else "host-mirror LBVH", dt),
            "mrays_per_s": (c["rays_closest"] + c["rays_shadow"]) / dt / 1e6}, dt, c


def run_reference(args):
    """--impl reference: the reference's own CPU-runnable implementation of the path on the box's host cores.  The OptiX
    build needs the OptiX SDK and an RTX driver (unavailable here), but the reference's SOURCES compile for the host against
    the stand-in of oracle/ref_shim: oracle/_ref/libref.so is its own main.cpp / Director.cpp / scene headers / 26 device .cu
    files (kind "reference").  Where that library or its assets are missing the CPU oracle (kind "port") is timed instead.
    The scene is built by the reference's own host code: librtw.so is not loaded by this arm."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    cores = host_cores()
    if reference_available() is not None:
        launches = 32          # 32 x 360 000 paths per step on C2: about a second on 16 host threads
        ref_py = reference_available()
        sess = ref_py.Session(SCENE, NX, NY, dm=False, threads=cores)
        for _ in range(args.warmup):
            sess.launch(1)
        t0 = time.perf_counter()
        rays = 0
        for _ in range(args.steps):
            rays += sess.launch(launches)
        dt = time.perf_counter() - t0
        paths = float(NX) * NY * launches * args.steps
        kind = "reference"
        sample = ("%d steps x %d launches of the reference's own renderFrame() (oracle/_ref/libref.so), each the %dx%d frame at 1 spp "
                  "(the reference traces one sample per launch whatever -ns says)" % (args.steps, launches, NX, NY))
        note = "each step = a bounded sample: %d launches x 1 spp of the %d spp workload (%d-spp-equivalent path count)" % (launches, SPP, launches)
    else:
        rtw = importlib.import_module("raytracing-weekend_b200")
        s, o = oracle_scene(rtw)
        spp = 32
        mode = 0 if s.desc.contents.n_prims <= 64 else 1
        for _ in range(args.warmup):
            o.render(NX, NY, 1, DEPTH, n_threads=cores, mode=mode)
        paths = 0
        rays = 0
        t0 = time.perf_counter()
        for i in range(args.steps):
            _, c = o.render(NX, NY, spp, DEPTH, n_threads=cores, mode=mode, sample_first=spp * i)
            paths += c["paths"]
            rays += c["rays_closest"] + c["rays_shadow"]
        dt = time.perf_counter() - t0
        kind = "port"
        sample = "%d steps x (%dx%d, %d spp) of the CPU oracle" % (args.steps, NX, NY, spp)
        note = "each step = a bounded sample: %d spp of the %d spp workload (CPU oracle: oracle/_ref/libref.so or its assets are missing here)" % (spp, SPP)
    v = paths / dt / 1e6
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": "Mpaths/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": WORKLOAD, "note": note},
            "mrays_per_s": rays / dt / 1e6,
            "cpu_baseline": {"value": v, "unit": "Mpaths/s", "cores": cores, "kind": kind, "sample": sample},
            "e2e": {"value": v, "unit": "Mpaths/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)
    return 0


def main():
    global NX, NY, SCENE, SPP, WORKLOAD, METRIC, STRESS_PRIMS
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", default="C2", choices=sorted(CONFIGS), help="BASELINE.json configuration (C2 = the one `metric` is quoted on)")
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours")
    ap.add_argument("--spp", type=int, default=0, help="override for quick checks (not the benchmark config)")
    ap.add_argument("--pool", type=int, default=0, help="wavefront pool size override")
    ap.add_argument("--scene", type=int, default=-1, help="override for profiling (not a benchmark config)")
    ap.add_argument("--nx", type=int, default=0)
    ap.add_argument("--ny", type=int, default=0)
    ap.add_argument("--stress-prims", type=int, default=0, help="C5 / scene 5: primitive count (0 = 2^20)")
    ap.add_argument("--leaf-size", type=int, default=0)
    ap.add_argument("--opt", action="append", default=[], help="name=value passed to rtw_set_option (tuning runs)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-relmse", action="store_true", help="skip the tier-2 relMSE measurement (N=1 only; ~10 s of GPU time)")
    args = ap.parse_args()
    cfgd = CONFIGS[args.config]
    SCENE, NX, NY, SPP, WORKLOAD, METRIC = cfgd["scene"], cfgd["nx"], cfgd["ny"], cfgd["spp"], cfgd["workload"], cfgd["metric"]
    if args.scene >= 0 or args.nx or args.ny:
        SCENE = args.scene if args.scene >= 0 else SCENE
        NX, NY = args.nx or NX, args.ny or NY
        WORKLOAD = "OVERRIDE (not a benchmark config): scene %d, %dx%d, %d spp, depth 20" % (SCENE, NX, NY, args.spp or SPP)
    if not args.spp:
        args.spp = SPP
    STRESS_PRIMS = args.stress_prims
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
    d.createScene(SCENE, n_stress_prims=args.stress_prims)
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
           "waves": 1, "pool": 0, "unocc": 0, "build_ms": 0.0}

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
            agg["unocc"] += st.get("nee_unoccluded", 0); agg["build_ms"] = st["build_ms"]

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
                           "l2_policy": "inputs larger than L2: the wavefront pools (%d x %d path slots per GPU x 156 B state + queues = %d MB) "
                                        "stream through HBM every iteration, no flush between steps" % (waves, pool, waves * pool * 156 >> 20),
                           "wall_ms_per_step": wall / args.steps * 1e3,
                           "frame_phases_ms_rank0": {k: v / args.steps for k, v in trace.items()}},
                "mrays_per_s": (tot[0].item() + tot[1].item()) / (dev_ms * 1e-3) / 1e6,
                "e2e": {"value": e2e, "unit": "Mpaths/s", "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": NX * NY * 16},
                "gpu_launches": int(tot[2].item()), "clocks": clk.summary()}
        # ---- roofline: algorithmic bytes per unit x units / live kernel time ----
        # HBM-visible ALGORITHMIC bytes: closest-hit ray = the 32-byte ray record read + hit 8 + queue
        # entry 4 written = 44 B; shaded vertex = queue entry 4 + slot record 64 (the hit rides in it) read, then 64 written by a path
        # that continues or radiance record (16 R + 16 W) + regen entry 4 by one that ends, + 48 per NEE ray emitted;
        # shadow ray = 48 read + 32 per contribution added; finished path = 116.  Streaming wavefront (LBVH scenes
        # without NEE): a vertex = one 80-byte entry read + one written.  The BVH nodes / primitive records of §8d
        # (64 B x nodes + 32 B x primitives, 80 B with a transform) come from L1 / L2: they are counted by the ORACLE on
        # its mirror of the hierarchy (a bounded 1 spp pass at reduced size), not by the GPU, and set against the measured
        # L2 read peak (profiles/l2_peak.json) in `l2`.
        streaming = agg["shade_ms"] > 0.0 and agg["trav_ms"] < 0.02 * agg["shade_ms"]   # (the streaming variant records two events back to back there)
        n_paths = float(NX * NY * spp) * args.steps / world
        rays_c, rays_s = max(agg["rays"], 1), agg["shrays"]
        iters = max(agg["iters"], 1)
        fin = n_paths / rays_c          # share of shaded vertices that end their path
        unocc = (agg["unocc"] / rays_s) if (rays_s and agg["unocc"]) else 0.5
        stages = {}
        if streaming:
            table = (("k_stream_fill", agg["logic_ms"], n_paths, 80.0),
                     ("k_stream_shade", agg["shade_ms"], rays_c, 160.0),
                     ("k_stream_advance", agg["shadow_ms"], iters, 128.0))
        else:
            shade_bpu = 68.0 + 64.0 * (1.0 - fin) + 36.0 * fin + 48.0 * rays_s / rays_c
            table = (("k_traverse", agg["trav_ms"], rays_c, 44.0),
                     ("k_shade*", agg["shade_ms"], rays_c, shade_bpu),
                     ("k_shadow", agg["shadow_ms"], rays_s, 48.0 + 32.0 * unocc),
                     ("k_regen", agg["logic_ms"], n_paths, 116.0))
        tp = os.path.join(ROOT, "profiles", "traffic.json")      # dram bytes per launch from the committed ncu capture
        traffic_doc = json.load(open(tp)) if os.path.exists(tp) else {}
        traffic_tab = traffic_doc.get("dram_bytes_per_launch", {}) if SCENE == 0 else {}
        for name, ms, units, bpu in table:
            gbs = units * bpu / max(ms * 1e-3, 1e-9) / 1e9
            stages[name] = {"ms_per_iteration": ms / iters, "units_per_iteration": units / iters, "bytes_per_unit": bpu,
                            "gbs": gbs, "frac": gbs / peak, "kernel_ms_over_render_ms": ms / max(agg["render_ms"], 1e-9)}
        ksum = sum(v["ms_per_iteration"] for v in stages.values())
        for name in stages:
            t = traffic_tab.get(name)
            cap_pool = traffic_doc.get("pool_slots")
            # the capture was taken with pools of `pool_slots`; a steady-state launch covers a pool, so the bytes of a
            # launch scale with the pool this run used (C2 at 1024 spp: 2.75 M slots per wave, capture: 2^21)
            stages[name]["traffic"] = (int(t * pool / cap_pool) if (t and cap_pool and pool != cap_pool) else t)
            stages[name]["share_of_kernel_time"] = stages[name]["ms_per_iteration"] / max(ksum, 1e-12)
        # The kernel the HBM roof is reported for: the stage that moves the most algorithmic HBM bytes per iteration (on
        # C2 k_shade*: 147 B x 2.0 M vertices against 44 B x 2.0 M rays of k_traverse).  It is also the longest launch
        # under ncu (41 % of the serialised kernel time, profiles/r02/r02g_launch_list_summary.txt); live, with two waves,
        # its CUDA-event time and the (issue-bound) traversal launch's are within a few per cent of each other, and picking
        # "the longest" would flip between them from run to run.
        dom = max(stages, key=lambda k: (stages[k]["units_per_iteration"] * stages[k]["bytes_per_unit"], stages[k]["ms_per_iteration"]))
        ds = stages[dom]
        step_bytes = sum(v["units_per_iteration"] * v["bytes_per_unit"] for v in stages.values()) * iters
        step_gbs = step_bytes / max(agg["render_ms"] * 1e-3, 1e-9) / 1e9
        line["roofline"] = {"bound": "hbm", "kernel": dom, "achieved": ds["gbs"], "peak": peak, "unit": "GB/s",
                            "frac": ds["frac"], "traffic": ds["traffic"],
                            "traffic_source": (traffic_doc.get("source") or "") + ("; scaled by %d / %d slots per pool" % (pool, traffic_doc.get("pool_slots"))
                                                                                  if traffic_doc.get("pool_slots") and pool != traffic_doc.get("pool_slots") else ""),
                            "peak_source": peak_src, "bytes_per_unit": ds["bytes_per_unit"], "units_per_launch": ds["units_per_iteration"],
                            "avg_launch_ms": ds["ms_per_iteration"],
                            "step": {"achieved": step_gbs, "frac": step_gbs / peak, "bytes_per_frame": step_bytes / max(args.steps, 1),
                                     "note": "all stages' algorithmic HBM bytes / device time of the timed frames"},
                            "note": "kernel = the stage that moves the most algorithmic HBM bytes per iteration.  "
                                    "achieved = algorithmic HBM bytes per launch (DESIGN.md 4.2) / average launch time measured live with CUDA "
                                    "events on wave 0's stream inside the timed frames; with %d waves a launch shares the device with the other "
                                    "waves' launches, so per-launch figures understate what the GPU moves: `step` is the whole frame against the "
                                    "same roof, `alone` the same kernel timed with one wave.  traffic = dram bytes of one launch from the committed "
                                    "ncu capture (traffic_source), C2 only" % waves,
                            "stages": stages}
        # ---- the other roof of this path: instruction issue (C2) ----
        # warp instructions of one steady-state iteration of one wave, from the committed ncu capture
        # (smsp__inst_executed.sum per launch), scaled by the iterations the timed frames ran, against
        # SMs x 4 schedulers x the SM clock sampled during the run.
        wi = traffic_doc.get("warp_instructions_per_launch") if SCENE == 0 else None
        clk_now = line["clocks"].get("sm_mhz") or 1965.0
        if wi and not streaming:
            per_iter = float(sum(wi.values()))
            sms = torch.cuda.get_device_properties(dev).multi_processor_count
            peak_issue = sms * 4 * clk_now * 1e6
            achieved_issue = per_iter * iters / max(agg["render_ms"] * 1e-3, 1e-9)
            line["roofline"]["issue"] = {"warp_instructions_per_wave_iteration": per_iter, "wave_iterations": iters,
                                         "achieved": achieved_issue / 1e9, "peak": peak_issue / 1e9, "unit": "G warp-inst/s",
                                         "frac": achieved_issue / peak_issue,
                                         "note": "instructions from the committed capture (traffic_source), iterations and time from this run; "
                                                 "with `step.frac` it says how far the frame is from BOTH of its roofs"}
        # ---- the on-chip (L1 / L2) stream of the traversal, from the ORACLE's visit counters ----
        try:
            s2, o2 = oracle_scene(rtw)
            big = s2.desc.contents.n_prims > 100000
            sx, sy = (max(NX // 8, 32), max(NY // 8, 32)) if big else (max(NX // 2, 32), max(NY // 2, 32))
            _, c = o2.render(sx, sy, 1, DEPTH, n_threads=host_cores(), mode=1)
            s2.close()
            rc = max(c["rays_closest"], 1)
            rs = max(c["rays_shadow"], 1)
            per_c = (64.0 * c["nodes_closest"] + 32.0 * (c["prims_closest"] - c["prims_xf_closest"]) + 80.0 * c["prims_xf_closest"]) / rc
            per_s = (64.0 * c["nodes_shadow"] + 32.0 * (c["prims_shadow"] - c["prims_xf_shadow"]) + 80.0 * c["prims_xf_shadow"]) / rs if c["rays_shadow"] else 0.0
            l2p = os.path.join(ROOT, "profiles", "l2_peak.json")
            l2_peak = float(json.load(open(l2p))["l2_read_gbs"]) if os.path.exists(l2p) else None
            trav_ms = (agg["shade_ms"] if streaming else agg["trav_ms"] + agg["shadow_ms"])
            on_chip = (agg["rays"] * per_c + agg["shrays"] * per_s) / max(trav_ms * 1e-3, 1e-9) / 1e9
            line["roofline"]["l2"] = {"bytes_per_closest_ray_survey_8d": per_c + 96.0, "bytes_per_shadow_ray_survey_8d": per_s + 96.0,
                                      "nodes_per_closest_ray": c["nodes_closest"] / rc, "prims_per_closest_ray": c["prims_closest"] / rc,
                                      "achieved": on_chip, "peak": l2_peak, "unit": "GB/s", "frac": (on_chip / l2_peak) if l2_peak else None,
                                      "sample": "oracle, host-mirror LBVH (two-child nodes, leaf ranges as the product's), %dx%d x 1 spp" % (sx, sy),
                                      "note": "node + primitive bytes of SURVEY 8d per ray (oracle counters) x rays the GPU traced / live time of the "
                                              "traversal kernels; for a scene that is one leaf range (C2) the records come from the constant bank, "
                                              "not from L2"}
        except Exception as ex:      # the mirror of a very large scene may not fit the time budget
            line["roofline"]["l2"] = {"unavailable": str(ex)}
        if waves > 1:   # the same kernels timed without a concurrent wave (one untimed extra frame)
            d.set_option("waves", 1)
            for _ in range(2):      # the first frame re-allocates the pool
                d.render_accum(accum.data_ptr(), 0, 1, max(min(spp // world, max(1, (1 << 29) // (NX * NY))), 1), DEPTH)
            s1 = d.stats()
            d.set_option("waves", 0)
            it1 = max(s1["iterations"], 1)
            key = {"k_traverse": "traverse_ms", "k_shade*": "shade_ms", "k_shadow": "shadow_ms", "k_regen": "logic_ms",
                   "k_stream_fill": "logic_ms", "k_stream_shade": "shade_ms", "k_stream_advance": "shadow_ms"}[dom]
            units1 = {"k_traverse": s1["rays_closest"], "k_shade*": s1["rays_closest"], "k_shadow": s1["rays_shadow"],
                      "k_regen": s1["paths"], "k_stream_fill": s1["paths"], "k_stream_shade": s1["rays_closest"],
                      "k_stream_advance": it1}[dom]
            g1 = units1 * ds["bytes_per_unit"] / max(s1[key] * 1e-3, 1e-9) / 1e9
            line["roofline"]["alone"] = {"waves": 1, "pool_paths": int(s1["pool_paths"]), "avg_launch_ms": s1[key] / it1,
                                         "units_per_launch": units1 / it1, "achieved": g1, "frac": g1 / peak,
                                         "frame_ms": s1["render_ms"]}
        line["build"] = {"lbvh_build_ms": agg["build_ms"], "primitives": int(dsc.n_prims),
                         "note": "device time of rtw_build_accel (inside every e2e step, outside `value`)"}
        if world == 1 and not args.no_relmse:
            # tier 2 of BASELINE.json at THIS configuration's image (raytracing-weekend_b200/tier2.py): the frame at the
            # workload's sample count (and a ladder around it) against a converged frame of DISJOINT samples
            tier2 = importlib.import_module("raytracing-weekend_b200.tier2")
            budget_paths = 2.6e10 if SCENE == 0 else 6.0e9       # ~9 s on C2; bounded on the slower scenes
            conv = int(max(1024, min(65536, budget_paths / (NX * NY))))
            ladder = tuple(sorted({max(conv // 256, 1), max(conv // 64, 1), max(conv // 16, 1), max(conv // 4, 1)} | ({spp} if spp <= conv // 4 else set())))
            m = tier2.measure(d, torch, NX, NY, ladder=ladder, converged_spp=conv, max_depth=DEPTH)
            m["at_workload_spp"] = m["relmse"].get(str(spp))
            m["target"] = "BASELINE.json: relMSE <= 1e-3 against the converged image"
            line["relmse_vs_converged"] = m
        if world == 1 and not args.no_cpu_baseline:
            if reference_available() is not None:
                cb, _, _ = reference_sample(12.0)
                pb, _, _ = cpu_sample(rtw, 6.0)
                cb["port_value"] = pb["value"]
                cb["port_sample"] = pb["sample"]
            else:
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
