"""The product's per-thread kernel bodies (csrc/device/*.h, host-compiled by tests/emu)
against the oracle: LBVH topology, camera rays, closest/any hit ids, textures, and the
whole wavefront state machine at 1 spp (bit-exact here: same libm on both sides)."""
import numpy as np
import pytest

import rayset
from rtw_types import SCENE_CASES


@pytest.mark.parametrize("sid,Nx,Ny", SCENE_CASES)
def test_lbvh_topology_matches_host_mirror(rtw, orc, emu, sid, Nx, Ny):
    s = rtw.Scene(sid, Nx, Ny)
    n = s.desc.contents.n_prims
    o = orc.OracleScene(s.desc, n)
    e = emu.EmuScene(s.desc, n, max_top=256)
    mo = o.bvh_download()
    me = e.bvh_download()
    for a, b, name in zip(mo, me, ["morton", "leaf_prim", "left", "right", "bounds"]):
        assert np.array_equal(a.view(np.uint32) if a.dtype == np.float32 else a,
                              b.view(np.uint32) if b.dtype == np.float32 else b), name
    s.close()


@pytest.mark.parametrize("sid,Nx,Ny", SCENE_CASES)
def test_rays_and_hits_bit_exact(rtw, orc, emu, sid, Nx, Ny):
    s = rtw.Scene(sid, Nx, Ny)
    n = s.desc.contents.n_prims
    o = orc.OracleScene(s.desc, n)
    # no staging / partial staging / everything staged  x  plain LBVH leaves / leaf ranges of
    # <= 4 / <= 32 primitives / the default (a <= 16-primitive scene is ONE range)
    for max_top, leaf_size in ((0, 1), (5, 4), (4096, 32), (64, 0)):
        e = emu.EmuScene(s.desc, n, max_top=max_top, leaf_size=leaf_size)
        o.set_leaf_size(leaf_size)
        cam_o = o.camera_rays(Nx, Ny, 3)
        cam_e = e.camera_rays(Nx, Ny, 3)
        assert cam_o.tobytes() == cam_e.tobytes()
        box = o.prim_aabbs()
        rays = rayset.make_rays(cam_o, box[:, :3].min(axis=0), box[:, 3:].max(axis=0), len(cam_o) + 6000)
        h0, _ = o.trace_closest(rays, 0)
        he, nv, pt = e.trace_closest(rays)
        assert np.array_equal(h0["prim"], he["prim"])
        assert np.array_equal(h0["t"].view(np.uint32), he["t"].view(np.uint32))
        sh = rayset.shadow_rays_from(rays, h0, [278.0, 554.0, 279.5], 6000)
        a0, _ = o.trace_any(sh, 0)
        assert np.array_equal(a0, e.trace_any(sh))
        # the oracle walking ITS mirror of the hierarchy (leaf ranges included) finds the same hits
        h1, _ = o.trace_closest(rays, 1)
        assert np.array_equal(h0, h1)
        e.close()
    s.close()


@pytest.mark.parametrize("sid,Nx,Ny", SCENE_CASES)
def test_wavefront_equals_scalar_path_loop_at_1spp(rtw, orc, emu, sid, Nx, Ny):
    s = rtw.Scene(sid, Nx, Ny)
    n = s.desc.contents.n_prims
    o = orc.OracleScene(s.desc, n)
    e = emu.EmuScene(s.desc, n, max_top=64)
    ref, c = o.render(Nx, Ny, 1, mode=0)
    for P in (1024, 100000):          # pool smaller / larger than the image
        img, st = e.render(Nx, Ny, 1, P=P)
        assert np.array_equal(ref.view(np.uint32), img.view(np.uint32))
        assert st["rays_closest"] == c["rays_closest"] and st["rays_shadow"] == c["rays_shadow"]
    # sample k of a pixel is the same walk whatever the sharding: samples {1,3} == first 1, stride 2
    a, _ = e.render(Nx, Ny, 2, P=4096, sample_first=1, sample_stride=2)
    b1, _ = o.render(Nx, Ny, 1, mode=1, sample_first=1)
    b3, _ = o.render(Nx, Ny, 1, mode=1, sample_first=3)
    assert np.allclose(a[..., :3], b1[..., :3] + b3[..., :3], rtol=1e-6, atol=1e-7)
    assert np.all(a[..., 3] == 2)
    s.close()


def test_depth_limit_and_empty_render(rtw, orc, emu):
    s = rtw.Scene(0, 48, 48)
    o = orc.OracleScene(s.desc, 13)
    e = emu.EmuScene(s.desc, 13)
    for depth in (0, 1, 3):
        ref, _ = o.render(48, 48, 1, max_depth=depth, mode=0)
        img, _ = e.render(48, 48, 1, max_depth=depth, P=512)
        assert np.array_equal(ref.view(np.uint32), img.view(np.uint32))
    img, st = e.render(48, 48, 0)
    assert not img.any() and st["rays_closest"] == 0
    s.close()


def test_textures_match(rtw, orc, emu):
    s = rtw.Scene(4, 64, 64)
    d = s.desc.contents
    o = orc.OracleScene(s.desc, d.n_prims)
    e = emu.EmuScene(s.desc, d.n_prims)
    rng = np.random.default_rng(1)
    for t in range(d.n_textures):
        for _ in range(200):
            u, v = rng.uniform(-0.2, 1.2, 2).astype(np.float32)
            p = rng.uniform(-500, 500, 3).astype(np.float32)
            a = o.texture_eval(t, float(u), float(v), p)
            b = e.texture_eval(t, float(u), float(v), p)
            assert np.array_equal(a.view(np.uint32), b.view(np.uint32)), (t, d.textures[t].type)
    s.close()


def test_checker_texture_dispatches_children(rtw, orc, emu):
    # Q18: no reference scene uses the checker; build one on top of Cornell's textures
    import ctypes as C
    s = rtw.Scene(0, 32, 32)
    d = rtw.SceneDesc.from_buffer_copy(s.desc.contents)
    texs = (rtw.Texture * (d.n_textures + 1))()
    for i in range(d.n_textures):
        texs[i] = d.textures[i]
    chk = rtw.Texture()
    chk.type = rtw.TEX_CHECKER
    chk.odd, chk.even, chk.perlin, chk.image = 0, 1, -1, -1
    texs[d.n_textures] = chk
    d.textures = C.cast(texs, C.POINTER(rtw.Texture))
    d.n_textures += 1
    o = orc.OracleScene(C.pointer(d), d.n_prims)
    e = emu.EmuScene(C.pointer(d), d.n_prims)
    seen = set()
    for p in ([0.1, 0.2, 0.3], [0.4, 3.0, 0.2], [1.0, 1.0, 1.0], [0.05, 9.9, 0.05]):
        a = o.texture_eval(d.n_textures - 1, 0.5, 0.5, p)
        b = e.texture_eval(d.n_textures - 1, 0.5, 0.5, p)
        assert np.array_equal(a, b)
        sines = np.sin(np.float32(10 * p[0])) * np.sin(np.float32(10 - p[1])) * np.sin(np.float32(10 * p[2]))
        expect = list(d.textures[0].color) if sines < 0 else list(d.textures[1].color)
        assert np.allclose(a, expect)
        seen.add(sines < 0)
    assert seen == {True, False}
    s.close()
