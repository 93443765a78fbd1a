"""Edge cases of the hot path through the C-ABI, each against the oracle on the same inputs: ragged and
one-pixel images, depth limits 0-3, empty and ragged ray sets, degenerate rays (zero direction, NaN, inverted
interval, huge magnitudes), a one-primitive scene, sample counts that do not divide by the number of waves."""
import ctypes as C

import numpy as np
import pytest

from rtw_types import RAY_DTYPE

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("sid,Nx,Ny", [(0, 1, 1), (0, 37, 3), (3, 5, 61), (1, 33, 7), (4, 17, 9)])
def test_ragged_image_sizes_bit_exact(director, orc, sid, Nx, Ny):
    director.init(Nx, Ny, 1)
    director.createScene(sid)
    o = orc.OracleScene(director.scene.desc, director.scene.desc.contents.n_prims)
    ref, c = o.render(Nx, Ny, 1, mode=1)
    director.renderFrame(1)
    assert np.array_equal(ref.view(np.uint32), director.linear().view(np.uint32))
    st = director.stats()
    assert st["paths"] == Nx * Ny and st["rays_closest"] == c["rays_closest"] and st["rays_shadow"] == c["rays_shadow"]


@pytest.mark.parametrize("sid", [0, 1, 4])
@pytest.mark.parametrize("depth", [0, 1, 2, 3])
def test_depth_limits_bit_exact(director, orc, sid, depth):
    """`while (depth < maxRayDepth)` (raygen/raygen.cu:36): limit 0 traces nothing (black, counted), 1 stops after
    the first hit, ..."""
    Nx, Ny = 40, 24
    director.init(Nx, Ny, 1)
    director.createScene(sid)
    o = orc.OracleScene(director.scene.desc, director.scene.desc.contents.n_prims)
    ref, c = o.render(Nx, Ny, 1, max_depth=depth, mode=1)
    out = director.renderFrame(1, max_depth=depth)
    lin = director.linear()
    assert np.array_equal(ref.view(np.uint32), lin.view(np.uint32))
    st = director.stats()
    assert st["rays_closest"] == c["rays_closest"] and st["rays_shadow"] == c["rays_shadow"]
    if depth == 0:
        assert not lin[..., :3].any() and np.all(lin[..., 3] == 1) and st["rays_closest"] == 0 and np.all(out[..., 3] == 1)


@pytest.mark.parametrize("sid", [0, 4])
def test_ragged_and_empty_ray_sets(director, orc, sid):
    director.init(64, 64, 1)
    director.createScene(sid)
    o = orc.OracleScene(director.scene.desc, director.scene.desc.contents.n_prims)
    cam = director.camera_rays(0)
    for n in (0, 1, 31, 32, 33, 255, 256, 257, 1000):
        rays = cam[:n]
        hits = director.trace_closest(rays)
        occ = director.trace_any(rays)
        assert len(hits) == n and len(occ) == n
        if n:
            rh, _ = o.trace_closest(rays, 0)
            ro, _ = o.trace_any(rays, 0)
            hits["t"][hits["prim"] < 0] = 0
            rh["t"][rh["prim"] < 0] = 0
            assert np.array_equal(hits, rh) and np.array_equal(occ, ro)


@pytest.mark.parametrize("sid", [0, 1, 3, 4])
def test_degenerate_rays_agree_with_the_oracle(director, orc, sid):
    """Zero and denormal directions, NaN / inf components, an inverted or empty [tmin, tmax], axis-parallel rays in
    the plane of a rect, huge origins: whatever the reference's arithmetic says, both sides must say the same."""
    director.init(32, 32, 1)
    director.createScene(sid)
    o = orc.OracleScene(director.scene.desc, director.scene.desc.contents.n_prims)
    cam = director.camera_rays(0)
    rays = np.zeros(0, dtype=RAY_DTYPE)
    base = cam[[5, 300, 700, 1000]]

    def variant(**kw):
        r = base.copy()
        for k, v in kw.items():
            r[k] = v
        return r
    parts = [variant(dir=(0.0, 0.0, 0.0)), variant(dir=(0.0, 1.0, 0.0)), variant(dir=(1.0, 0.0, 0.0)), variant(dir=(0.0, 0.0, -1.0)),
             variant(dir=(1e-42, 1.0, 0.0)), variant(dir=(np.nan, 1.0, 0.0)), variant(dir=(np.inf, 1.0, 0.0)),
             variant(origin=(np.nan, 0.0, 0.0)), variant(origin=(1e30, 1e30, 1e30)), variant(origin=(278.0, 554.0, 279.5)),
             variant(origin=(278.0, 0.0, 279.5), dir=(1.0, 0.0, 0.0)),       # in the plane of the floor
             variant(tmin=10.0, tmax=1.0), variant(tmin=0.0, tmax=0.0), variant(tmin=-5.0, tmax=np.inf),
             variant(dir=(1e20, 1e20, 1e20)), variant(dir=(-3e-39, -3e-39, 1.0))]
    rays = np.concatenate(parts)
    hits = director.trace_closest(rays)
    occ = director.trace_any(rays)
    rh, _ = o.trace_closest(rays, 0)
    ro, _ = o.trace_any(rays, 0)
    assert np.array_equal(hits["prim"], rh["prim"])
    m = rh["prim"] >= 0
    assert np.array_equal(hits["t"][m].view(np.uint32), rh["t"][m].view(np.uint32))
    assert np.array_equal(occ, ro)


def test_one_primitive_scene(director, rtw, orc):
    """One leaf: the LBVH has no internal node worth the name; hits, misses and a 1 spp frame still agree."""
    s = rtw.Scene(0, 32, 32)
    d1 = rtw.SceneDesc.from_buffer_copy(s.desc.contents)
    d1.n_prims = 1
    p1 = C.pointer(d1)
    director.init(32, 32, 1)
    director.scene = s
    director.upload(p1)
    o = orc.OracleScene(p1, 1)
    cam = director.camera_rays(0)
    hits = director.trace_closest(cam)
    rh, _ = o.trace_closest(cam, 0)
    hits["t"][hits["prim"] < 0] = 0
    rh["t"][rh["prim"] < 0] = 0
    assert np.array_equal(hits, rh) and (rh["prim"] == 0).any() and (rh["prim"] == -1).any()
    ref, c = o.render(32, 32, 1, mode=0)
    director.renderFrame(1)
    assert np.array_equal(ref.view(np.uint32), director.linear().view(np.uint32))
    assert director.stats()["rays_closest"] == c["rays_closest"]


def test_sample_counts_that_do_not_divide_by_the_waves(director, orc):
    torch = pytest.importorskip("torch")
    Nx = Ny = 32
    director.init(Nx, Ny, 7)
    director.createScene(0)
    o = orc.OracleScene(director.scene.desc, 13)
    acc = torch.zeros(Ny, Nx, 4, device="cuda:0")
    torch.cuda.synchronize()
    try:
        director.set_option("waves", 4)
        director.set_option("pool_paths", 1024)
        director.render_accum(acc.data_ptr(), 0, 1, 0)               # nothing to do: returns, adds nothing
        assert not acc.any().item()
        director.render_accum(acc.data_ptr(), 3, 2, 1)               # one sample (k = 3): fewer samples than waves
        ref, _ = o.render(Nx, Ny, 1, sample_first=3, sample_stride=2, mode=0)
        assert np.array_equal(acc.cpu().numpy().view(np.uint32), ref.view(np.uint32))
        acc.zero_()
        torch.cuda.synchronize()
        director.render_accum(acc.data_ptr(), 1, 3, 7)               # k = 1, 4, ..., 19 dealt to 4 waves
        ref, c = o.render(Nx, Ny, 7, sample_first=1, sample_stride=3, mode=0)
        a = acc.cpu().numpy()
        assert np.all(a[..., 3] == 7) and np.allclose(a, ref, rtol=1e-5, atol=1e-6)
        st = director.stats()
        assert st["waves"] == 4 and st["rays_closest"] == c["rays_closest"] and st["rays_shadow"] == c["rays_shadow"]
    finally:
        director.set_option("waves", 0)
        director.set_option("pool_paths", 0)


def test_upload_refuses_what_the_scene_file_loader_refuses(director, rtw):
    """rtw_upload_scene runs the one strict validator (advisor r01): negative indices, a textured material
    without a texture, nested checkers and a shutter interval outside a moving sphere's time range are
    refused with RTW_ERR_INVALID instead of being followed by a kernel."""
    import ctypes as C
    import scene_kit
    L = rtw.lib()

    def upload(k):
        return L.rtw_upload_scene(director._c, k.desc)

    k = scene_kit.build(rtw, "instances")
    assert upload(k) == rtw.RTW_OK
    for mutate in ("transform", "texture", "checker", "shutter", "odd"):
        k = scene_kit.build(rtw, "branches" if mutate in ("checker", "odd") else "instances")
        d = k.desc.contents
        if mutate == "transform":
            d.prims[0].transform = -2
        elif mutate == "texture":
            d.materials[0].texture = -1            # lambertian without a texture
        elif mutate == "checker":
            chk = [i for i in range(d.n_textures) if d.textures[i].type == rtw.TEX_CHECKER][0]
            d.textures[chk].even = chk             # nested checker
        elif mutate == "odd":
            chk = [i for i in range(d.n_textures) if d.textures[i].type == rtw.TEX_CHECKER][0]
            d.textures[chk].odd = -3
        else:
            d.camera.time1 = 2.0                   # moving sphere is defined on [0, 1]
        assert upload(k) == rtw.RTW_ERR_INVALID, mutate
        assert L.rtw_last_error(director._c)
    # an image of more than 2^31 pixels is refused before anything is sized
    out = np.zeros(4, np.float32)
    assert L.rtw_render(director._c, 70000, 70000, 1, 1, out.ctypes.data) == rtw.RTW_ERR_INVALID
    director.init(32, 32, 1)
    director.createScene(0)


def test_coincident_primitives_do_not_overflow_the_traversal_stack(director, rtw, orc):
    """4096 primitives with the SAME centroid (identical Morton keys: the Karras hierarchy is split on index bits
    only, the deepest tree this builder can make).  The stack-overflow counter stays 0 (a dropped push would fail
    the call with RTW_ERR_STATE) and the hits are the brute-force oracle's."""
    import ctypes as C
    import rayset
    import scene_kit
    k = scene_kit.KitScene(rtw)
    m = k.mat(rtw.MAT_LAMBERTIAN, k.constant(0.5, 0.5, 0.5))
    for i in range(4096):
        k.prim(rtw.PRIM_SPHERE, [1.0, 2.0, 3.0, 0.5 + 0.0001 * (i % 7)], m)
    k.sky = 1
    k.cam = scene_kit.camera(rtw, (1, 2, 8), (1, 2, 3), 30, 1.5)
    k.finish()
    director.init(48, 32, 1)
    director.set_option("leaf_size", 1)
    try:
        director.scene = k
        director.upload(k.desc)
        o = orc.OracleScene(k.desc, k.n_prims)
        cam = o.camera_rays(48, 32, 0)
        hits = director.trace_closest(cam)
        ref, _ = o.trace_closest(cam, 0)
        assert np.array_equal(hits["prim"], ref["prim"]) and (ref["prim"] >= 0).any()
        director.renderFrame(1)
        assert director.stats()["stack_overflows"] == 0
    finally:
        director.set_option("leaf_size", 0)


def test_subnormal_radiance_is_flushed_by_the_accumulator_atomics(director, rtw, orc):
    """The one place where the device does NOT keep the oracle's bits: finished paths are added to the
    accumulator with red/atom.global.add.v4.f32 (k_regen.sass: ATOMG.E.ADD.F32x4.FTZ.RN), and fp32 atomic adds on global
    memory flush subnormal inputs and results to zero — an architectural property of the instruction.  Cornell box
    with the light's radiance scaled to 1e-39: every path radiance is subnormal; the oracle keeps the sums, the device
    returns exactly 0 for them.  Normal-range sums are untouched (every other test)."""
    s = rtw.Scene(0, 48, 48)
    d1 = rtw.SceneDesc.from_buffer_copy(s.desc.contents)
    texs = (rtw.Texture * d1.n_textures)(*[d1.textures[i] for i in range(d1.n_textures)])
    lights = (rtw.Light * d1.n_lights)(*[d1.lights[i] for i in range(d1.n_lights)])
    scaled = 0
    for t in texs:
        if t.type == 1 and max(t.color) > 1.0:          # RTW_TEX_CONSTANT: the emitter's colour (15, 15, 15)
            for k in range(3):
                t.color[k] = 1e-39
            scaled += 1
    for l in lights:
        for k in range(3):
            l.emission[k] = 1e-39
    assert scaled == 1 and d1.n_lights == 1
    d1.textures = C.cast(texs, C.POINTER(rtw.Texture))
    d1.lights = C.cast(lights, C.POINTER(rtw.Light))
    p1 = C.pointer(d1)
    director.init(48, 48, 1)
    director.scene = s
    director.upload(p1)
    o = orc.OracleScene(p1, d1.n_prims)
    ref, c = o.render(48, 48, 1, mode=0)
    director.renderFrame(1)
    got = director.linear()
    assert director.stats()["rays_closest"] == c["rays_closest"] and director.stats()["rays_shadow"] == c["rays_shadow"]   # the same walks
    rgb_ref, rgb_got = ref[..., :3], got[..., :3]
    tiny = np.abs(rgb_ref) < np.float32(1.1754944e-38)
    assert tiny.all() and (rgb_ref != 0).mean() > 0.3          # the oracle's frame is lit, and all of it is subnormal
    assert np.all(rgb_got == 0.0)                               # ... the device flushed every one of those sums
    assert np.array_equal(ref[..., 3], got[..., 3])             # sample counts are ordinary numbers
