"""GPU box: LBVH build time of the stress scene (2^20 primitives by default), warm.  Under ncu
(--metrics gpu__time_duration.sum) the launch list gives the per-kernel split of the build."""
import importlib, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
rtw = importlib.import_module("raytracing-weekend_b200")
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 20
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 5
d = rtw.Director(0)
d.init(64, 64, 1)
d.createScene(5, n_stress_prims=n)
ms = []
for _ in range(reps):
    d.upload(d.scene.desc)
    ms.append(d.stats()["build_ms"])
print(json.dumps({"n_prims": n, "build_ms": [round(m, 4) for m in ms], "best_ms": round(min(ms), 4),
                  "Mprims_per_s": round(n / min(ms) / 1e3, 1)}), flush=True)
d.destroy()
