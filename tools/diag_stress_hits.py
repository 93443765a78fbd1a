"""GPU box: stress scene, GPU closest hits against the oracle's mirror and, where they differ, against brute force."""
import importlib, json, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
rtw = importlib.import_module("raytracing-weekend_b200")
from oracle import oracle_py as orc
import rayset
n_prims = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 15
d = rtw.Director(0)
Nx, Ny = 128, 72
d.init(Nx, Ny, 1)
for leaf in (4, 2, 1):
    d.set_option("leaf_size", leaf)
    d.createScene(5, n_stress_prims=n_prims)
    o = orc.OracleScene(d.scene.desc, n_prims)
    cam = d.camera_rays(0)
    box = o.prim_aabbs()
    rays = rayset.make_rays(cam, box[:, :3].min(axis=0), box[:, 3:].max(axis=0), 1 << 17)
    hg = d.trace_closest(rays)
    hm, _ = o.trace_closest(rays, 1)
    bad = np.nonzero((hg["prim"] != hm["prim"]) | ((hg["prim"] >= 0) & (hg["t"].view(np.uint32) != hm["t"].view(np.uint32))))[0]
    out = {"leaf_size": leaf, "rays": len(rays), "gpu_vs_mirror_mismatch": int(len(bad))}
    if len(bad):
        sel = bad[:64]
        hb, _ = o.trace_closest(rays[sel], 0)     # brute force on those rays
        out["of_first_64: gpu == brute force"] = int(np.sum((hg["prim"][sel] == hb["prim"]) & ((hb["prim"] < 0) | (hg["t"][sel].view(np.uint32) == hb["t"].view(np.uint32)))))
        out["of_first_64: mirror == brute force"] = int(np.sum((hm["prim"][sel] == hb["prim"]) & ((hb["prim"] < 0) | (hm["t"][sel].view(np.uint32) == hb["t"].view(np.uint32)))))
        k = sel[0]
        p = d.scene.desc.contents.prims[int(hb["prim"][0])] if hb["prim"][0] >= 0 else None
        out["example"] = {"ray_o": rays["origin"][k].tolist(), "ray_d": rays["dir"][k].tolist(), "gpu": [float(hg["t"][k]), int(hg["prim"][k])],
                          "mirror": [float(hm["t"][k]), int(hm["prim"][k])], "brute": [float(hb["t"][0]), int(hb["prim"][0])],
                          "prim_type": (p.type if p else None), "prim_xf": (p.transform if p else None), "box": box[int(hb["prim"][0])].tolist() if p else None}
    print(json.dumps(out), flush=True)
d.destroy()
