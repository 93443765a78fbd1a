"""Writes tests/golden/oracle_kat.json: known answers of the CPU oracle (oracle/oracle.cpp) on fixed inputs —
closest-hit (t, primitive id) and any-hit answers on the tier-1 ray set, 1 spp linear images, ray counters —
as SHA-256 digests plus a few raw values.  The reference ships no vectors of its own and its device
code does not compile here (§8c), so these pin the ORACLE against drift (and, in `-m gpu`, the CUDA path against
the very same numbers); what pins the oracle to the reference is tests/golden/rng_kat.json (the reference's own
lib/random.cuh compiled where it lies) and the scene replays in tests/test_scenes.py.

    python tests/golden/make_golden.py        # rewrites oracle_kat.json (commit the result)
"""
import hashlib
import importlib
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

N_RAYS = 8192
CASES = [(0, 32, 32), (1, 32, 18), (2, 32, 16), (3, 32, 32), (4, 32, 32)]   # scene, Nx, Ny of the 1 spp image


def digest(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def known_answers(rtw, orc, rayset, sid, Nx, Ny):
    """Everything below is a pure function of (scene id, Nx, Ny); tests recompute it on both sides."""
    s = rtw.Scene(sid, Nx, Ny)
    n = s.desc.contents.n_prims
    o = orc.OracleScene(s.desc, n)
    box = o.prim_aabbs()
    rays = rayset.make_rays(o.camera_rays(Nx, Ny, 0), box[:, :3].min(axis=0), box[:, 3:].max(axis=0), N_RAYS)
    hits, _ = o.trace_closest(rays, 0)
    hits["t"][hits["prim"] < 0] = 0
    shadow = rayset.shadow_rays_from(rays, hits, [278.0, 554.0, 279.5] if sid in (0, 3) else [0.0, 300.0, 0.0], 4096)
    occ, _ = o.trace_any(shadow, 0)
    img, c = o.render(Nx, Ny, 1, mode=0)
    out = {"scene": sid, "Nx": Nx, "Ny": Ny, "n_prims": n, "n_rays": int(len(rays)),
           "rays_sha256": digest(rays), "hits_sha256": digest(hits), "n_hit": int((hits["prim"] >= 0).sum()),
           "n_shadow": int(len(shadow)), "occluded_sha256": digest(occ), "n_occluded": int(occ.sum()),
           "image_sha256": digest(img), "image_pixels_hex": {str(i): img.reshape(-1, 4)[i].tobytes().hex() for i in (Nx * Ny * np.arange(1, 5) // 5 + Nx // 2)},
           "rays_closest": c["rays_closest"], "rays_shadow": c["rays_shadow"]}
    return out, (s, rays, shadow)


def main():
    rtw = importlib.import_module("raytracing-weekend_b200")
    from oracle import oracle_py as orc
    import rayset
    kat = [known_answers(rtw, orc, rayset, *case)[0] for case in CASES]
    with open(os.path.join(HERE, "oracle_kat.json"), "w") as f:
        json.dump({"generator": "tests/golden/make_golden.py", "n_rays": N_RAYS, "cases": kat}, f, indent=1)
    print("wrote %d cases" % len(kat))


if __name__ == "__main__":
    main()
