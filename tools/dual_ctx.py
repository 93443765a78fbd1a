"""GPU box experiment: two contexts (two streams, two pools) on ONE GPU, each rendering half of the samples of a
C2 frame into the same accumulator from its own host thread — do an issue-bound kernel of one context and a
DRAM-bound kernel of the other overlap?  Prints Mpaths/s for one context and for two, at several grid shares."""
import importlib, json, os, sys, threading, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
rtw = importlib.import_module("raytracing-weekend_b200")
Nx = Ny = 600
SPP = int(os.environ.get("SPP", "512"))

def run(n_ctx, pct, pool):
    ds = []
    for g in range(n_ctx):
        d = rtw.Director(0)
        d.init(Nx, Ny, SPP)
        d.set_option("grid_pct", pct)
        d.set_option("pool_paths", pool)
        d.createScene(0)
        ds.append(d)
    acc = torch.zeros(Ny, Nx, 4, device="cuda:0")
    torch.cuda.synchronize()
    best = 1e9
    for rep in range(4):
        acc.zero_()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        ths = [threading.Thread(target=ds[g].render_accum, args=(acc.data_ptr(), g, n_ctx, SPP // n_ctx)) for g in range(n_ctx)]
        for t in ths: t.start()
        for t in ths: t.join()
        torch.cuda.synchronize()
        best = min(best, time.perf_counter() - t0)
    ok = bool((acc[..., 3] == SPP).all().item())
    for d in ds: d.destroy()
    print(json.dumps({"contexts": n_ctx, "grid_pct": pct, "pool": pool, "ms": round(best * 1e3, 2), "Mpaths/s": round(Nx * Ny * SPP / best / 1e6, 1), "count_ok": ok}), flush=True)

run(1, 100, 1 << 22)
for pct in (100, 67, 50):
    for pool in (1 << 21, 1 << 22):
        run(2, pct, pool)
run(3, 34, 1 << 21)
