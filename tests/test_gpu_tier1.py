"""Tier 1 (bit-exact) parity of the CUDA path against the oracle, through the C-ABI:
RNG streams, the builder's radix sort, LBVH topology, camera rays, closest-hit primitive
ids + distances and any-hit flags on the fixed ray set."""
import json
import os

import numpy as np
import pytest

import rayset
from rtw_types import SCENE_CASES

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden", "rng_kat.json")


def test_device_rng_streams(director, orc):
    kat = json.load(open(GOLD))
    for row in kat["tea"]:
        t, lcg, rnd, xs, rf = director.rng_stream(row["s0"], row["s1"], 512)
        assert t == row["tea64"]
        o = orc.rng_stream(row["s0"], row["s1"], 512)
        assert np.array_equal(lcg, o[1]) and np.array_equal(xs, o[3])
        assert np.array_equal(rnd.view(np.uint32), o[2].view(np.uint32))
        assert np.array_equal(rf.view(np.uint32), o[4].view(np.uint32))


@pytest.mark.parametrize("n", [1, 2, 31, 32, 33, 2047, 2048, 2049, 100003, 1 << 20])
def test_radix_sort_stable(director, n):
    rng = np.random.default_rng(n)
    # few distinct keys → many ties → stability matters; plus full-range keys
    keys = rng.integers(0, 1 << 63, n, dtype=np.uint64)
    if n > 64:
        keys[: n // 2] = rng.integers(0, 50, n // 2, dtype=np.uint64) << np.uint64(40)
    vals = np.arange(n, dtype=np.uint32)
    k, v = director.sort_pairs(keys, vals)
    order = np.argsort(keys, kind="stable")
    assert np.array_equal(k, keys[order])
    assert np.array_equal(v, vals[order])


def test_radix_sort_empty(director):
    k, v = director.sort_pairs(np.zeros(0, np.uint64), np.zeros(0, np.uint32))
    assert len(k) == 0 and len(v) == 0


@pytest.mark.parametrize("sid,Nx,Ny", SCENE_CASES)
def test_lbvh_topology_and_hits(director, rtw, orc, sid, Nx, Ny):
    director.init(Nx, Ny, 1)
    director.createScene(sid)
    d = director.scene.desc.contents
    n = d.n_prims
    o = orc.OracleScene(director.scene.desc, n)
    mo = o.bvh_download()
    mg = director.bvh_download()
    for a, b, name in zip(mo, mg, ["morton", "leaf_prim", "left", "right", "bounds"]):
        assert np.array_equal(a.view(np.uint32) if a.dtype == np.float32 else a,
                              b.view(np.uint32) if b.dtype == np.float32 else b), name
    cam_o = o.camera_rays(Nx, Ny, 0)
    cam_g = director.camera_rays(0)
    assert cam_o.tobytes() == cam_g.tobytes()
    box = o.prim_aabbs()
    info = director.bvh_info()
    assert np.array_equal(np.array(info.scene_min, np.float32), box[:, :3].min(axis=0))
    assert np.array_equal(np.array(info.scene_max, np.float32), box[:, 3:].max(axis=0))
    rays = rayset.make_rays(cam_o, box[:, :3].min(axis=0), box[:, 3:].max(axis=0), 1 << 16)
    h0, _ = o.trace_closest(rays, 1 if n > 100 else 0)
    hg = director.trace_closest(rays)
    assert np.array_equal(h0["prim"], hg["prim"])
    assert np.array_equal(h0["t"].view(np.uint32), hg["t"].view(np.uint32))
    sh = rayset.shadow_rays_from(rays, h0, [278.0, 554.0, 279.5], 1 << 15)
    a0, _ = o.trace_any(sh, 1 if n > 100 else 0)
    assert np.array_equal(a0, director.trace_any(sh))


def test_top_staging_sizes_give_identical_hits(director, rtw, orc):
    Nx, Ny = 96, 96
    director.init(Nx, Ny, 1)
    ref = None
    director.set_option("wide", 0)      # the shared-memory staging belongs to the two-child walk
    # staged-node count x leaf-range size (1 = plain LBVH: every internal node is reachable)
    for top, leaf in ((0, 1), (7, 1), (512, 1), (3584, 1), (0, 4), (7, 8), (3584, 4), (1024, 32), (1024, 0)):
        director.set_option("top_nodes", top)
        director.set_option("leaf_size", leaf)
        director.createScene(4)
        info = director.bvh_info()
        n_top = info.n_top
        assert info.n_internal == 3407       # 3410 primitives, two of them in the oversize list (device/lbvh.h)
        assert n_top == min(top, 3407) if leaf == 1 else 0 <= n_top <= min(top, 3407)
        hits = director.trace_closest(director.camera_rays(0))
        if ref is None:
            ref = hits
        assert np.array_equal(ref, hits)
    director.set_option("top_nodes", 1024)
    # the 4-wide walk (the default) finds the same hits, for every leaf-range size
    director.set_option("wide", 1)
    for leaf in (1, 4, 32, 0):
        director.set_option("leaf_size", leaf)
        director.createScene(4)
        assert director.bvh_info().n_top == 0
        assert np.array_equal(ref, director.trace_closest(director.camera_rays(0)))
    director.set_option("leaf_size", 0)


def test_full_size_ray_set_cornell(director, rtw, orc):
    # 2^20 rays on the C2 scene: brute-force oracle ids, bit-exact
    Nx = Ny = 600
    director.init(Nx, Ny, 1)
    director.createScene(0)
    o = orc.OracleScene(director.scene.desc, 13)
    cam = director.camera_rays(0)
    assert cam.tobytes() == o.camera_rays(Nx, Ny, 0).tobytes()
    box = o.prim_aabbs()
    rays = rayset.make_rays(cam, box[:, :3].min(axis=0), box[:, 3:].max(axis=0), 1 << 20)
    h0, _ = o.trace_closest(rays, 0)
    hg = director.trace_closest(rays)
    assert np.array_equal(h0, hg)


def test_stress_scene_build_properties(director, rtw):
    # C5-style scene at 2^17 primitives: size-independent LBVH properties
    director.init(320, 200, 1)
    director.createScene(5, n_stress_prims=1 << 17)
    n = 1 << 17
    morton, leaf_prim, left, right, bounds = director.bvh_download()
    assert np.all(morton[:-1] <= morton[1:])
    assert np.array_equal(np.sort(leaf_prim), np.arange(n, dtype=np.int32))
    refs = np.concatenate([left, right])
    assert np.array_equal(np.sort(refs[refs >= 0]), np.arange(1, n - 1, dtype=np.int32))
    assert np.array_equal(np.sort(~refs[refs < 0]), np.arange(n, dtype=np.int32))
    # parent boxes contain child boxes
    for side, ref in ((0, left), (1, right)):
        idx = np.nonzero(ref >= 0)[0]
        child = bounds[ref[idx]]
        cmin = np.minimum(child[:, 0:3], child[:, 6:9])
        cmax = np.maximum(child[:, 3:6], child[:, 9:12])
        pb = bounds[idx][:, 6 * side: 6 * side + 6]
        assert np.array_equal(pb[:, :3], cmin) and np.array_equal(pb[:, 3:], cmax)
    hits = director.trace_closest(director.camera_rays(0))
    assert (hits["prim"] >= 0).mean() > 0.05


def test_api_error_paths(rtw):
    d = rtw.Director(0)
    d.init(32, 32, 1)
    with pytest.raises(rtw.RtwError):
        d.renderFrame(1)                      # render before createScene
    with pytest.raises(rtw.RtwError):
        d.set_option("no_such_option", 1)
    with pytest.raises(rtw.RtwError):
        d.set_option("pool_paths", 3)
    d.destroy()


def test_stress_scene_hits_match_the_oracle(director, orc):
    """C5 generator at 2^15 primitives: 1024 shared instance transforms, mixed spheres /
    quads; closest hits of the sample-0 camera rays + random rays, bit-exact against the oracle walking
    its own mirror of the hierarchy (brute force over 32 768 primitives would take minutes)."""
    Nx, Ny = 128, 72
    director.init(Nx, Ny, 1)
    director.createScene(5, n_stress_prims=1 << 15)
    n = director.scene.desc.contents.n_prims
    assert n == 1 << 15
    o = orc.OracleScene(director.scene.desc, n)
    cam = director.camera_rays(0)
    assert cam.tobytes() == o.camera_rays(Nx, Ny, 0).tobytes()
    box = o.prim_aabbs()
    rays = rayset.make_rays(cam, box[:, :3].min(axis=0), box[:, 3:].max(axis=0), 1 << 15)
    h0, _ = o.trace_closest(rays, 1)
    hg = director.trace_closest(rays)
    assert np.array_equal(h0["prim"], hg["prim"])
    assert np.array_equal(h0["t"].view(np.uint32), hg["t"].view(np.uint32))
    assert (hg["prim"] >= 0).sum() > 300          # the scene is sparse: ~2 % of these rays hit something
    # a brute-force spot check on a subset keeps the oracle's own BVH walk honest
    sub = rays[:: 64]
    hb, _ = o.trace_closest(sub, 0)
    assert np.array_equal(hb, h0[:: 64])
    ref, c = o.render(Nx, Ny, 1, mode=1)
    director.renderFrame(1)
    assert np.array_equal(ref.view(np.uint32), director.linear().view(np.uint32))
