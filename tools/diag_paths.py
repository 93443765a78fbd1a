"""GPU diagnostic: per-path (1 spp) differences between the CUDA renderer and the oracle."""
import importlib, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
rtw = importlib.import_module("raytracing-weekend_b200")
from oracle import oracle_py as orc
from rtw_types import SCENE_CASES
d = rtw.Director(0)
for sid, Nx, Ny in SCENE_CASES:
    d.init(Nx, Ny, 1); d.createScene(sid)
    o = orc.OracleScene(d.scene.desc, d.scene.desc.contents.n_prims)
    ref, c = o.render(Nx, Ny, 1, mode=1)
    d.renderFrame(1)
    lin = d.linear()
    a, b = ref[..., :3], lin[..., :3]
    diff = np.abs(a - b).max(axis=2)
    scale = np.maximum(np.abs(a).max(axis=2), 1.0)
    rel = diff / scale
    st = d.stats()
    print("scene %d %dx%d: bit-identical %.4f  rel<=1e-6 %.4f  <=2e-4 %.4f  <=1e-2 %.4f  max %.3g | rays gpu %d/%d oracle %d/%d" % (
        sid, Nx, Ny, (diff == 0).mean(), (rel <= 1e-6).mean(), (rel <= 2e-4).mean(), (rel <= 1e-2).mean(), rel.max(),
        st["rays_closest"], st["rays_shadow"], c["rays_closest"], c["rays_shadow"]), flush=True)
    print("   mean gpu %.6f oracle %.6f" % (b.mean(), a.mean()))
d.destroy()
