"""GPU-side scene flattening and texture baking (csrc/kernels/bake.cu behind rtw_bake_*): instance
transform inversion, Perlin table generation, image flip / RGBA expansion as device kernels, BIT-EQUAL to what the host
scene builder (rtw_scene_create: the restatement of Director.cpp:498-519, texture/ioTexture.h:181-216,228-266 — itself
pinned to the reference-built scene by tests/test_ref_pin.py) computes."""
import ctypes as C
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_transform_inversion_on_the_device(director, rtw):
    L = rtw.lib()
    for sid in (0, 3, 4, 5):
        s = rtw.Scene(sid, 64, 64, n_stress_prims=4096)
        d = s.desc.contents
        assert d.n_transforms >= 1
        o2w = np.zeros((d.n_transforms, 12), np.float32)
        for i in range(d.n_transforms):
            o2w[i] = list(d.transforms[i].o2w)
        out = (rtw.Transform * d.n_transforms)()
        assert L.rtw_bake_transforms(director._c, o2w.ctypes.data, d.n_transforms, out) == rtw.RTW_OK
        assert bytes(out) == C.string_at(d.transforms, d.n_transforms * C.sizeof(rtw.Transform)), sid
    # the hand-built scenes' transforms (scene_kit forms the same inverse in Python float64)
    import scene_kit
    k = scene_kit.build(rtw, "instances")
    d = k.desc.contents
    o2w = np.array([list(d.transforms[i].o2w) for i in range(d.n_transforms)], np.float32)
    out = (rtw.Transform * d.n_transforms)()
    assert L.rtw_bake_transforms(director._c, o2w.ctypes.data, d.n_transforms, out) == rtw.RTW_OK
    assert bytes(out) == C.string_at(d.transforms, d.n_transforms * C.sizeof(rtw.Transform))


def test_perlin_tables_on_the_device(director, rtw):
    L = rtw.lib()
    # scene 4 builds one table (the marble sphere's), scene 2 one: both are table 0 of mt19937(0)
    for sid in (2, 4):
        s = rtw.Scene(sid, 64, 64)
        d = s.desc.contents
        assert d.n_perlin == 1
        out = (rtw.Perlin * 1)()
        assert L.rtw_bake_perlin(director._c, 0, 0, 1, out) == rtw.RTW_OK
        assert bytes(out) == C.string_at(d.perlin, C.sizeof(rtw.Perlin)), sid
    # later tables continue the same stream: tables [0, 3) at once == table 0, 1, 2 one by one
    three = (rtw.Perlin * 3)()
    assert L.rtw_bake_perlin(director._c, 0, 0, 3, three) == rtw.RTW_OK
    for t in range(3):
        one = (rtw.Perlin * 1)()
        assert L.rtw_bake_perlin(director._c, 0, t, 1, one) == rtw.RTW_OK
        assert bytes(one) == bytes(three[t])
        perm = np.array(three[t].perm_y)
        assert sorted(perm.tolist()) == list(range(256))
        v = np.array(three[t].ranvec, np.float32)
        assert np.allclose((v * v).sum(axis=1), 1.0, atol=1e-5)
    assert bytes(three[0]) != bytes(three[1])


def test_image_flip_and_expand_on_the_device(director, rtw):
    L = rtw.lib()
    with open(os.path.join(ROOT, "assets", "earthmap.ppm"), "rb") as f:
        assert f.readline().strip() == b"P6"
        w, h = map(int, f.readline().split())
        assert f.readline().strip() == b"255"
        rgb = np.frombuffer(f.read(), np.uint8).copy()
    out = np.zeros(w * h * 4, np.uint8)
    assert L.rtw_bake_image(director._c, rgb.ctypes.data, w, h, 3, out.ctypes.data) == rtw.RTW_OK
    s = rtw.Scene(4, 64, 64)
    im = s.desc.contents.images[0]
    assert (im.width, im.height) == (w, h)
    assert out.tobytes() == C.string_at(im.rgba, w * h * 4)
    # 4-channel input keeps its alpha
    rng = np.random.RandomState(1)
    px = rng.randint(0, 256, (5, 7, 4)).astype(np.uint8)
    out = np.zeros(5 * 7 * 4, np.uint8)
    assert L.rtw_bake_image(director._c, px.ctypes.data, 7, 5, 4, out.ctypes.data) == rtw.RTW_OK
    assert np.array_equal(out.reshape(5, 7, 4), px[::-1])
    assert L.rtw_bake_image(director._c, px.ctypes.data, 7, 5, 2, out.ctypes.data) == rtw.RTW_ERR_INVALID


def test_a_scene_baked_on_the_device_renders_the_reference_frame(director, rtw, orc):
    """End to end: scene 4's description with its transforms, Perlin table and texels RE-MADE by the device kernels renders
    the frame the reference's own programs render (tests/golden/ref_frames.npz)."""
    import json
    L = rtw.lib()
    with open(os.path.join(ROOT, "tests", "golden", "ref_pin.json")) as f:
        pin = json.load(f)
    frames = np.load(os.path.join(ROOT, "tests", "golden", "ref_frames.npz"))
    Nx, Ny = pin["nx"], pin["ny"]
    s = rtw.Scene(4, Nx, Ny)
    d = rtw.SceneDesc.from_buffer_copy(s.desc.contents)
    o2w = np.array([list(d.transforms[i].o2w) for i in range(d.n_transforms)], np.float32)
    xf = (rtw.Transform * d.n_transforms)()
    assert L.rtw_bake_transforms(director._c, o2w.ctypes.data, d.n_transforms, xf) == rtw.RTW_OK
    pl = (rtw.Perlin * 1)()
    assert L.rtw_bake_perlin(director._c, 0, 0, 1, pl) == rtw.RTW_OK
    with open(os.path.join(ROOT, "assets", "earthmap.ppm"), "rb") as f:
        f.readline(); w, h = map(int, f.readline().split()); f.readline()
        rgb = np.frombuffer(f.read(), np.uint8).copy()
    texels = np.zeros(w * h * 4, np.uint8)
    assert L.rtw_bake_image(director._c, rgb.ctypes.data, w, h, 3, texels.ctypes.data) == rtw.RTW_OK
    img = (rtw.Image * 1)()
    img[0].width, img[0].height = w, h
    img[0].rgba = texels.ctypes.data_as(C.POINTER(C.c_uint8))
    d.transforms = C.cast(xf, C.POINTER(rtw.Transform))
    d.perlin = C.cast(pl, C.POINTER(rtw.Perlin))
    d.images = C.cast(img, C.POINTER(rtw.Image))
    director.init(Nx, Ny, 1)
    director.scene = s
    director.upload(C.pointer(d))
    assert np.array_equal(director.renderFrame(1)[..., :3].view(np.uint32), frames["scene4"])
