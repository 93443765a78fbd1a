"""Tier-1 ray set: sample-0 camera rays followed by rays with origins
uniform in the scene box and directions from [-1,1]^3 (normalised), all drawn from the
SDK `lcg` recurrence seeded 0xC0FFEE.  The generator is committed, not the data."""
import numpy as np

from rtw_types import RAY_DTYPE

LCG_A = np.uint32(1664525)
LCG_C = np.uint32(1013904223)


def lcg_states(seed, n):
    """states s_1..s_n of prev = A*prev + C (mod 2^32), vectorised."""
    a = np.full(n, LCG_A, dtype=np.uint32)
    with np.errstate(over="ignore"):
        apow = np.cumprod(a, dtype=np.uint32)                       # A^1..A^n
        geo = np.cumsum(np.concatenate([[np.uint32(1)], apow[:-1]]), dtype=np.uint32)  # 1 + A + .. + A^(k-1)
        return apow * np.uint32(seed) + LCG_C * geo


def rnd_from_states(states):
    return (states & np.uint32(0x00FFFFFF)).astype(np.float32) / np.float32(0x01000000)


def make_rays(camera_rays, scene_min, scene_max, n_total, seed=0xC0FFEE):
    n_cam = min(len(camera_rays), n_total)
    rays = np.zeros(n_total, dtype=RAY_DTYPE)
    rays[:n_cam] = camera_rays[:n_cam]
    m = n_total - n_cam
    if m > 0:
        u = rnd_from_states(lcg_states(seed, m * 8)).reshape(m, 8)
        lo = np.asarray(scene_min, np.float32)
        hi = np.asarray(scene_max, np.float32)
        # keep the random origins inside a bounded box even for scenes with giant spheres
        lo = np.maximum(lo, np.float32(-2000.0))
        hi = np.minimum(hi, np.float32(2000.0))
        o = lo + u[:, 0:3] * (hi - lo)
        d = np.float32(2.0) * u[:, 3:6] - np.float32(1.0)
        ln = np.sqrt((d * d).sum(axis=1, keepdims=True))
        ln[ln == 0] = 1
        d = (d / ln).astype(np.float32)
        r = rays[n_cam:]
        r["origin"] = o.astype(np.float32)
        r["dir"] = d
        r["tmin"] = 1e-6
        r["tmax"] = 1e27
        r["time"] = u[:, 6]
        r["gather_time"] = u[:, 7]
        r["seed"] = lcg_states(seed ^ 0x5EED, m)
        rays[n_cam:] = r
    return rays


def shadow_rays_from(rays, hits, light_point, n_max=None):
    """Any-hit test set: from every closest hit towards a fixed point (like NEE rays)."""
    ok = hits["prim"] >= 0
    r = rays[ok]
    h = hits[ok]
    if n_max is not None:
        r, h = r[:n_max], h[:n_max]
    p = r["origin"] + h["t"][:, None] * r["dir"]
    d = np.asarray(light_point, np.float32)[None, :] - p
    dist = np.sqrt((d * d).sum(axis=1)).astype(np.float32)
    good = dist > 1e-3
    p, d, dist, r = p[good], d[good], dist[good], r[good]
    out = np.zeros(len(p), dtype=RAY_DTYPE)
    out["origin"] = p.astype(np.float32)
    out["dir"] = (d / dist[:, None]).astype(np.float32)
    out["tmin"] = np.float32(500) * np.float32(1.0e-7)
    out["tmax"] = dist - np.float32(500) * np.float32(1.0e-7)
    out["time"] = 0
    out["gather_time"] = r["gather_time"]
    out["seed"] = r["seed"] + np.uint32(12345)
    return out
