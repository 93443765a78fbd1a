"""GPU box: what would sorting the rays of a wavefront iteration buy the LBVH walkers?  A ray set shaped like a
steady-state iteration (25 % camera rays, 75 % bounce rays leaving the first hit points in random directions) is
traced through rtw_trace_closest (the same persistent walkers as k_traverse) in random order and sorted by
(origin cell, direction bin).  Run under ncu (gpu__time_duration, lanes per instruction of k_trace) — the call itself
copies 48 B per ray from the host, which swamps the kernel in wall time.

    python tools/diag_sort_rays.py <scene> [n_rays] [stress_prims]"""
import importlib, json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
rtw = importlib.import_module("raytracing-weekend_b200")
sid = int(sys.argv[1]) if len(sys.argv) > 1 else 4
n = int(sys.argv[2]) if len(sys.argv) > 2 else 1 << 20
nprims = int(sys.argv[3]) if len(sys.argv) > 3 else 0
d = rtw.Director(0)
Nx = Ny = 1024
d.init(Nx, Ny, 1)
d.createScene(sid, n_stress_prims=nprims)
rng = np.random.RandomState(7)
cam = d.camera_rays(0)
cam = cam[rng.permutation(len(cam))[:n]]
hits = d.trace_closest(cam)
ok = hits["prim"] >= 0
src = cam[ok]
t = hits["t"][ok].astype(np.float64)
nb = min(len(src), (3 * n) // 4)
sel = rng.permutation(len(src))[:nb]
bounce = src[sel].copy()
p = src["origin"][sel].astype(np.float64) + (t[sel] * 0.999)[:, None] * src["dir"][sel].astype(np.float64)
v = rng.normal(size=(nb, 3))
v /= np.linalg.norm(v, axis=1)[:, None]
bounce["origin"] = p.astype(np.float32)
bounce["dir"] = v.astype(np.float32)
bounce["tmin"] = 1e-6
bounce["tmax"] = 1e27
rays = np.concatenate([cam[: n - nb], bounce])
rays = rays[rng.permutation(len(rays))]


def key_of(r, obits, dbits, dir_first=False):
    o = r["origin"].astype(np.float64)
    lo, hi = o.min(axis=0), o.max(axis=0)
    c = np.clip(((o - lo) / np.maximum(hi - lo, 1e-30) * (1 << obits)).astype(np.int64), 0, (1 << obits) - 1)
    dd = r["dir"].astype(np.float64)
    dd = dd / np.abs(dd).max(axis=1)[:, None]
    q = np.clip(((dd + 1.0) * (1 << (dbits - 1))).astype(np.int64), 0, (1 << dbits) - 1)
    mort = np.zeros(len(r), np.int64)
    for b in range(obits):
        for a in range(3):
            mort |= ((c[:, a] >> b) & 1) << (3 * b + a)
    dk = (q[:, 0] << (2 * dbits)) | (q[:, 1] << dbits) | q[:, 2]
    return (dk << (3 * obits)) | mort if dir_first else (mort << (3 * dbits)) | dk


out = {"scene": sid, "n_rays": len(rays), "camera_share": round((n - nb) / len(rays), 3)}
ref = None
for name, order in (("random", np.arange(len(rays))),
                    ("sorted_o5_d3", np.argsort(key_of(rays, 5, 3), kind="stable")),
                    ("sorted_o6_d2", np.argsort(key_of(rays, 6, 2), kind="stable")),
                    ("sorted_o4_d3", np.argsort(key_of(rays, 4, 3), kind="stable")),
                    ("sorted_dir_first", np.argsort(key_of(rays, 5, 3, True), kind="stable"))):
    rs = np.ascontiguousarray(rays[order])
    d.trace_closest(rs[:1024])
    t0 = time.perf_counter()
    h = d.trace_closest(rs)
    out[name + "_wall_ms"] = round((time.perf_counter() - t0) * 1e3, 2)
    back = np.empty_like(h)
    back[order] = h
    if ref is None:
        ref = back
    assert np.array_equal(ref, back)          # the order of the rays does not change a hit
print(json.dumps(out), flush=True)
d.destroy()
