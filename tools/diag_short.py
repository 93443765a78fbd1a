"""GPU box: short frames (the per-GPU share of an 8-GPU run): C2 at 128 spp, waves x pool, frame ms and iteration count."""
import importlib, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
rtw = importlib.import_module("raytracing-weekend_b200")
d = rtw.Director(0)
d.init(600, 600, 128)
d.createScene(0)
for spp in (128, 1024):
    for waves, pool in ((1, 1 << 22), (2, 1 << 21), (2, 1 << 20), (3, 1 << 20), (4, 1 << 20), (4, 1 << 19), (2, 1 << 22)):
        d.set_option("waves", waves)
        d.set_option("pool_paths", pool)
        best = 1e9
        for rep in range(4):
            d.renderFrame(spp)
            st = d.stats()
            best = min(best, st["render_ms"])
        print(json.dumps({"spp": spp, "waves": waves, "pool": pool, "render_ms": round(best, 3), "Mpaths/s": round(600 * 600 * spp / best / 1e3, 1),
                          "iters": st["iterations"], "launches": st["kernel_launches"]}), flush=True)
d.destroy()
