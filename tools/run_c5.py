"""GPU box: the synthetic stress scene C5: LBVH build time and render throughput vs primitive count."""
import importlib, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
rtw = importlib.import_module("raytracing-weekend_b200")
d = rtw.Director(0)
Nx, Ny, spp = 3840, 2160, int(sys.argv[1]) if len(sys.argv) > 1 else 4
for n in (1 << 14, 1 << 17, 1 << 20):
    d.init(Nx, Ny, spp)
    d.createScene(5, n_stress_prims=n)
    d.upload(d.scene.desc)            # second build: warm (allocation-free) build time
    b = d.stats()["build_ms"]
    d.renderFrame(spp)
    d.renderFrame(spp)
    st = d.stats()
    info = d.bvh_info()
    print(json.dumps({"scene": 5, "n_prims": n, "Nx": Nx, "Ny": Ny, "spp": spp, "build_ms": round(b, 3),
                      "build_Mprims_per_s": round(n / b / 1e3, 1), "render_ms": round(st["render_ms"], 2),
                      "Mpaths_per_s": round(st["paths"] / st["render_ms"] / 1e3, 1),
                      "Mrays_per_s": round((st["rays_closest"] + st["rays_shadow"]) / st["render_ms"] / 1e3, 1),
                      "n_top": info.n_top, "iterations": st["iterations"]}), flush=True)
d.destroy()
