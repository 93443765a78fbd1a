"""Scene wire format (include/rtw.h rtw_scene_save / rtw_scene_load): a flattened
scene written to a file and read back is the same scene, byte for byte, and renders the same image;
damaged files are refused with a status code, never followed."""
import ctypes as C
import os
import struct
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _bytes(ptr, n, typ):
    return C.string_at(ptr, n * C.sizeof(typ)) if n else b""


def _blob(rtw, scene):
    """Every byte a kernel can see of a flattened scene."""
    d = scene.desc.contents
    parts = [_bytes(d.prims, d.n_prims, rtw.Prim), _bytes(d.transforms, d.n_transforms, rtw.Transform),
             _bytes(d.materials, d.n_materials, rtw.Material), _bytes(d.textures, d.n_textures, rtw.Texture),
             _bytes(d.perlin, d.n_perlin, rtw.Perlin), _bytes(d.lights, d.n_lights, rtw.Light),
             bytes(d.light_rect), bytes(d.camera), struct.pack("<ii", d.has_light_rect, d.sky_light)]
    for i in range(d.n_images):
        im = d.images[i]
        parts.append(struct.pack("<ii", im.width, im.height))
        parts.append(C.string_at(im.rgba, im.width * im.height * 4))
    return b"".join(parts)


@pytest.mark.parametrize("sid", [0, 1, 2, 3, 4, 5])
def test_round_trip_is_byte_exact(rtw, tmp_path, sid):
    s = rtw.Scene(sid, 320, 200, n_stress_prims=4096)
    path = tmp_path / ("scene%d.rtws" % sid)
    s.save(path)
    t = rtw.Scene.load(path)
    assert t.description == s.description
    assert _blob(rtw, t) == _blob(rtw, s)
    # saving the loaded scene gives the same file
    path2 = tmp_path / "again.rtws"
    t.save(path2)
    assert path.read_bytes() == path2.read_bytes()
    s.close()
    t.close()


def test_loaded_scene_renders_the_same_image_in_the_oracle(rtw, orc, tmp_path):
    s = rtw.Scene(0, 24, 24)
    s.save(tmp_path / "c.rtws")
    t = rtw.Scene.load(tmp_path / "c.rtws")
    a, ca = orc.OracleScene(s.desc, 13).render(24, 24, 2, mode=0)
    b, cb = orc.OracleScene(t.desc, 13).render(24, 24, 2, mode=0)
    assert a.tobytes() == b.tobytes() and ca == cb


def test_damaged_files_are_refused(rtw, tmp_path):
    s = rtw.Scene(3, 64, 64)
    path = tmp_path / "s.rtws"
    s.save(path)
    good = path.read_bytes()
    L = rtw.lib()

    def load(data):
        p = tmp_path / "x.rtws"
        p.write_bytes(data)
        h = C.c_void_p()
        rc = L.rtw_scene_load(str(p).encode(), C.byref(h))
        assert (rc == rtw.RTW_OK) == bool(h)
        if h:
            L.rtw_scene_free(h)
        return rc, L.rtw_scene_io_error().decode()

    assert load(good)[0] == rtw.RTW_OK
    assert load(b"")[0] != rtw.RTW_OK
    assert load(b"P3\n1 1\n255\n0 0 0\n")[0] != rtw.RTW_OK            # not a scene file
    assert load(good[:len(good) // 2])[0] != rtw.RTW_OK               # truncated
    assert load(good[:-8])[0] != rtw.RTW_OK                           # checksum missing
    flipped = bytearray(good)
    flipped[len(good) // 2] ^= 0x40
    rc, why = load(bytes(flipped))
    assert rc != rtw.RTW_OK and ("checksum" in why or "range" in why)
    # a well-formed file (checksum recomputed) whose first primitive points at a material that does not exist
    hdr = 12 + struct.calcsize("<10i") + C.sizeof(rtw.LightRect) + C.sizeof(rtw.Camera) + 4
    text = struct.unpack_from("<I", good, hdr - 4)[0]
    bad = bytearray(good[:-8])
    struct.pack_into("<i", bad, hdr + text + 4, 10 ** 6)              # rtw_prim.material of primitive 0
    h = 1469598103934665603
    for b in bad:
        h = ((h ^ b) * 1099511628211) & 0xffffffffffffffff
    rc, why = load(bytes(bad) + struct.pack("<Q", h))
    assert rc != rtw.RTW_OK and "material" in why
    # nothing to read, nothing to write
    hnd = C.c_void_p()
    assert L.rtw_scene_load(str(tmp_path / "missing.rtws").encode(), C.byref(hnd)) == rtw.RTW_ERR_IO
    assert L.rtw_scene_save(C.cast(s.desc, C.c_void_p), b"", str(tmp_path / "no_such_dir" / "s.rtws").encode()) == rtw.RTW_ERR_IO
    s.close()


@pytest.mark.gpu
def test_scene_file_renders_bit_exact_on_the_gpu(director, rtw, tmp_path):
    director.init(48, 48, 1)
    director.createScene(3)
    ref = director.renderFrame(1).copy()
    director.scene.save(tmp_path / "s3.rtws")
    director.loadScene(tmp_path / "s3.rtws")
    out = director.renderFrame(1)
    assert np.array_equal(ref.view(np.uint32), out.view(np.uint32))


@pytest.mark.gpu
def test_cli_reads_and_writes_scene_files(tmp_path):
    cli = os.path.join(ROOT, "raytracing-weekend_b200", "rtw_cli")
    a, b, f = tmp_path / "a.ppm", tmp_path / "b.ppm", tmp_path / "s.rtws"
    subprocess.check_call([cli, "-s", "0", "-dx", "64", "-dy", "64", "-ns", "2", "-o", str(a), "-w", str(f)])
    subprocess.check_call([cli, "-f", str(f), "-dx", "64", "-dy", "64", "-ns", "2", "-o", str(b)])
    assert a.read_bytes() == b.read_bytes()


def test_save_refuses_a_shutter_outside_a_moving_spheres_time_range(rtw, tmp_path):
    import scene_kit
    k = scene_kit.build(rtw, "instances")
    L = rtw.lib()
    assert L.rtw_scene_save(C.cast(k.desc, C.c_void_p), b"ok", str(tmp_path / "a.rtws").encode()) == rtw.RTW_OK
    k.desc.contents.camera.time0 = -0.5
    assert L.rtw_scene_save(C.cast(k.desc, C.c_void_p), b"bad", str(tmp_path / "b.rtws").encode()) == rtw.RTW_ERR_INVALID
    assert b"shutter" in L.rtw_scene_io_error()
    # the hand-built scenes survive the wire format byte for byte
    t = rtw.Scene.load(tmp_path / "a.rtws")
    assert t.desc.contents.n_prims == k.n_prims
    t.save(tmp_path / "c.rtws")
    assert (tmp_path / "a.rtws").read_bytes() == (tmp_path / "c.rtws").read_bytes()
