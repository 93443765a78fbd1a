"""Host scene flattening (scene/ioScene.h restated) and the C-ABI surface."""
import ctypes as C
import os
import re

import numpy as np
import pytest


def _randf_stream(seed):
    s = seed
    while True:
        s ^= (s << 13) & 0xffffffff
        s ^= s >> 17
        s ^= (s << 5) & 0xffffffff
        r = np.float32(s) / np.float32(4294967296.0)
        yield np.float32(0x3F7FFFFF) if r == 1.0 else r


def test_primitive_counts(rtw):
    # counts measured by replaying the reference's randf loops
    expect = {0: 13, 1: 528, 2: 520, 3: 8, 4: 3410}
    for sid, n in expect.items():
        s = rtw.Scene(sid, 640, 480)
        assert s.desc.contents.n_prims == n
        s.close()


def test_scene1_breakdown_matches_independent_replay(rtw):
    # independent python replay of scene/ioScene.h:200-252 (draw order right-to-left, Q33)
    g = _randf_stream(0x314759)
    f = np.float32
    kinds = {"moving": 0, "metal": 0, "glass": 0, "pair": 0}
    for a in range(-11, 11):
        for b in range(-11, 11):
            choose = next(g)
            x = f(a) + f(0.8) * next(g)
            z = f(b) + f(0.9) * next(g)
            zsq = z * z
            dist = np.sqrt((x - f(4.0)) * (x - f(4.0)) + zsq)
            if dist > f(0.9) or (zsq > f(0.7) and (x * x - f(16.0)) > f(-2.0)):
                if choose < f(0.70):
                    kinds["moving"] += 1
                    next(g); next(g); next(g)
                elif choose < f(0.85):
                    kinds["metal"] += 1
                    next(g); next(g); next(g); next(g)
                elif choose < f(0.93):
                    kinds["glass"] += 1
                else:
                    kinds["pair"] += 1
    assert kinds == {"moving": 333, "metal": 71, "glass": 36, "pair": 42}
    s = rtw.Scene(1, 640, 480)
    prims = s.prims()
    mats = [s.desc.contents.materials[p.material].type for p in prims]
    assert sum(p.type == rtw.PRIM_MOVING_SPHERE for p in prims) == 333
    assert sum(m == rtw.MAT_METAL for m in mats) == 71 + 1
    assert sum(m == rtw.MAT_DIELECTRIC for m in mats) == 36 + 2 * 42 + 1
    assert s.desc.contents.sky_light == 1 and s.desc.contents.n_lights == 0
    s.close()


def test_scene4_layout(rtw):
    s = rtw.Scene(4, 800, 800)
    d = s.desc.contents
    p = s.prims()
    assert p[0].type == rtw.PRIM_RECT_Y and p[0].flip == 1 and d.materials[p[0].material].type == rtw.MAT_DIFFUSE_LIGHT
    assert p[4].v[3] == 5000.0 and d.materials[p[4].material].type == rtw.MAT_DIELECTRIC          # Q34
    assert p[7].type == rtw.PRIM_MOVING_SPHERE
    assert all(q.type in (rtw.PRIM_RECT_X, rtw.PRIM_RECT_Y, rtw.PRIM_RECT_Z) for q in p[8:2408])
    assert p[2408].type == rtw.PRIM_VOLUME_SPHERE and abs(p[2408].v[4] - 0.2) < 1e-7
    assert p[2409].type == rtw.PRIM_VOLUME_SPHERE and p[2409].v[3] == 500.0
    assert all(q.type == rtw.PRIM_SPHERE and q.transform == 0 and q.v[3] == 10.0 for q in p[2410:3410])
    # first ground box height: 100 * (randf(0x6314759) + 0.01)
    g = _randf_stream(0x6314759)
    y1 = np.float32(100) * (next(g) + np.float32(0.01))
    assert p[8 + 3].v[4] == y1      # 4th rect of the box is the y = p1.y face (ioGeometryGroup.h:33-34)
    assert d.n_lights == 1 and d.has_light_rect == 1 and d.light_rect.axis == 1 and d.light_rect.k == 554.0
    assert d.n_perlin == 1 and d.n_images == 1 and d.images[0].width == 1024 and d.images[0].height == 512
    s.close()


def test_cornell_camera_and_light(rtw):
    s = rtw.Scene(0, 600, 600)
    d = s.desc.contents
    cam = d.camera
    assert list(cam.origin) == [278.0, 278.0, -800.0]
    assert np.allclose(list(cam.w), [0, 0, -1]) and np.allclose(list(cam.u), [-1, 0, 0]) and np.allclose(list(cam.v), [0, 1, 0])
    hh = np.tan(np.float32(40.0 * np.pi / 180.0) / 2)
    assert np.allclose(list(cam.horizontal), [-2 * hh * 10, 0, 0], rtol=1e-6)
    assert cam.lens_radius == 0.0 and cam.time0 == 0.0 and cam.time1 == 1.0      # Q3, Director.cpp:521-522
    L = d.lights[0]
    assert list(L.normal) == [0.0, -1.0, 0.0] and L.area == 13650.0 and list(L.emission) == [15.0] * 3   # Q28
    assert d.light_rect.k == np.float32(554.9) and L.position[1] == 554.0
    assert d.sky_light == 0
    s.close()


def test_transforms_are_inverse_pairs(rtw):
    for sid in (0, 3, 4):
        s = rtw.Scene(sid, 320, 240)
        d = s.desc.contents
        for i in range(d.n_transforms):
            a = np.vstack([np.array(d.transforms[i].o2w, np.float64).reshape(3, 4), [0, 0, 0, 1]])
            b = np.vstack([np.array(d.transforms[i].w2o, np.float64).reshape(3, 4), [0, 0, 0, 1]])
            assert np.allclose(a @ b, np.eye(4), atol=1e-4)
        s.close()


def test_perlin_tables_are_permutations(rtw):
    s = rtw.Scene(2, 320, 240)
    d = s.desc.contents
    t = d.perlin[0]
    for perm in (t.perm_x, t.perm_y, t.perm_z):
        assert sorted(perm) == list(range(256))
    v = np.array([list(r) for r in t.ranvec])
    assert np.allclose((v * v).sum(axis=1), 1, atol=1e-5)
    s.close()


def test_unknown_scene_is_an_error(rtw):
    with pytest.raises(rtw.RtwError):
        rtw.Scene(9, 320, 240)


def test_abi_exports_every_declared_symbol(rtw):
    hdr = open(os.path.join(rtw.REPO_ROOT, "include", "rtw.h")).read()
    declared = set(re.findall(r"\b(rtw_[a-z_0-9]+)\s*\(", hdr))
    assert declared == set(rtw.ABI_SYMBOLS), declared ^ set(rtw.ABI_SYMBOLS)
    L = rtw.lib()
    for name in declared:
        assert hasattr(L, name), name
    assert L.rtw_abi_version() == 2


def test_abi_struct_sizes(rtw):
    assert C.sizeof(rtw.Prim) == 64 and C.sizeof(rtw.Transform) == 96 and C.sizeof(rtw.Material) == 16
    assert C.sizeof(rtw.Texture) == 48 and C.sizeof(rtw.Perlin) == 256 * 12 + 3 * 1024
    assert C.sizeof(rtw.Ray) == 48 and C.sizeof(rtw.Hit) == 8


def test_no_cpu_fallback_without_device(rtw):
    if rtw.device_count() > 0:
        pytest.skip("a CUDA device is present")
    with pytest.raises(rtw.RtwError):
        rtw.Director(0)


def test_ppm_writer(rtw, tmp_path):
    # Director::printPPM, Director.cpp:1010-1031: P3, top row first, int(255.99*clamp)
    img = np.zeros((2, 2, 4), np.float32)
    img[0, 0] = [0, 0.5, 1.0, 1]      # bottom-left
    img[1, 1] = [2.0, -1.0, 0.25, 1]  # top-right
    p = str(tmp_path / "o.ppm")
    rc = rtw.lib().rtw_write_ppm(p.encode(), 2, 2, img.ctypes.data_as(C.c_void_p))
    assert rc == 0
    lines = open(p).read().split("\n")
    assert lines[:3] == ["P3", "2 2", "255"]
    assert lines[3] == "0 0 0" and lines[4] == "255 0 63" and lines[5] == "0 127 255" and lines[6] == "0 0 0"
