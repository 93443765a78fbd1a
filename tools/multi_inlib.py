"""Runs on the GPU box: the in-library multi-GPU path (rtw_create_multi; one process, one context and one
host thread per GPU) at BASELINE.json's C2 configuration, both shard modes, against one GPU through the
same entry points.  Wall clock around the C-ABI call (the call returns a HOST frame, so this is e2e).

    python tools/multi_inlib.py <n_gpus> [scene Nx Ny spp]
prints one JSON line per (devices, shard mode)."""
import importlib
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
rtw = importlib.import_module("raytracing-weekend_b200")


def run(devs, shard, sid, Nx, Ny, spp, reps=3):
    m = rtw.MultiDirector(devs)
    try:
        m.init(Nx, Ny, spp)
        m.set_option("shard", shard)
        m.createScene(sid)
        m.renderFrame(spp)                      # warm-up: pools, NCCL communicator
        ts = []
        for _ in range(reps):
            t0 = time.perf_counter()
            m.renderFrame(spp)
            ts.append(time.perf_counter() - t0)
        st = m.stats()
        t = min(ts)
        return {"devices": devs, "shard": st["shard"], "nccl": m.uses_nccl(), "scene": sid, "Nx": Nx, "Ny": Ny, "spp": spp,
                "s_per_frame": round(t, 4), "mpaths_per_s": round(st["paths"] / t / 1e6, 1),
                "exchange_ms": round(st["exchange_ms"], 3), "per_device_paths": [p["paths"] for p in st["per_device"]]}
    finally:
        m.destroy()


if __name__ == "__main__":
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 2
    sid, Nx, Ny, spp = (int(x) for x in sys.argv[2:6]) if len(sys.argv) >= 6 else (0, 1200, 800, 1024)
    have = rtw.device_count()
    print(json.dumps({"device_count": have}))
    sets = [[0]] + ([list(range(n))] if have >= n else [[0] * n])
    for devs in sets:
        for shard in ((1,) if len(devs) == 1 else (1, 2)):
            print(json.dumps(run(devs, shard, sid, Nx, Ny, spp)), flush=True)
