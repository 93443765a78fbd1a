"""GPU box: where the fixed per-frame time goes (fill + drain tail) — C2 at several spp, stage sums vs frame time."""
import importlib, json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
rtw = importlib.import_module("raytracing-weekend_b200")
d = rtw.Director(0)
d.init(600, 600, 128)
d.createScene(0)
d.set_option("profile", 1)
for pool in (1 << 22, 1 << 23, 1 << 24):
    d.set_option("pool_paths", pool)
    for spp in (32, 128, 512, 1024):
        for rep in range(3):
            t0 = time.perf_counter()
            d.renderFrame(spp)
            wall = (time.perf_counter() - t0) * 1e3
        st = d.stats()
        stages = st["logic_ms"] + st["traverse_ms"] + st["shade_ms"] + st["shadow_ms"]
        print(json.dumps({"pool": pool, "spp": spp, "iters": st["iterations"], "render_ms": round(st["render_ms"], 3), "wall_ms": round(wall, 3),
                          "stage_sum_ms": round(stages, 3), "regen": round(st["logic_ms"], 3), "trav": round(st["traverse_ms"], 3),
                          "shade": round(st["shade_ms"], 3), "shadow": round(st["shadow_ms"], 3),
                          "rays_per_iter": int(st["rays_closest"] / st["iterations"])}))
d.destroy()
