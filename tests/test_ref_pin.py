"""The oracle and the host scene builder pinned to THE REFERENCE ITSELF.

tests/golden/ref_frames.npz + ref_pin.json were written by tests/golden/make_ref_golden.py from
oracle/_ref/libref_dm.so — the reference's own main.cpp, Director.cpp, scene headers and all 26
device .cu files, compiled where they lie under /root/reference on the OptiX-on-the-CPU stand-in of
oracle/ref_shim.  Here (no GPU, no /root/reference needed):

  - the CPU oracle reproduces every committed reference frame BIT FOR BIT: scenes 0-4 through
    Director::createScene, and the hand-built scenes of tests/scene_kit.py (checker / null texture,
    normal material, X-axis light rectangle, two lights, lens radius > 0, transformed image-textured
    sphere, moving sphere, transformed constant medium) through the reference's device programs;
  - the product's host scene builder (rtw_scene_create) builds, value for value, the scene the
    reference's ioScene / Director::createSBT / initLaunchParams build (digest of every resolved
    primitive, transform, material, texture, Perlin table, texel, light, camera field).

When oracle/_ref/libref*.so and /root/reference are present (this container), the `live` tests
re-run the reference and compare again, check the earth-map texels against the reference's stb_image
decode, run the reference's own CLI, and bound what glibc's libm changes (the detmath contract).
"""
import ctypes as C
import hashlib
import json
import os
import subprocess

import numpy as np
import pytest

import scene_digest
import scene_kit

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = os.path.join(ROOT, "tests", "golden")


@pytest.fixture(scope="module")
def pin():
    with open(os.path.join(GOLD, "ref_pin.json")) as f:
        return json.load(f)


@pytest.fixture(scope="module")
def frames():
    return np.load(os.path.join(GOLD, "ref_frames.npz"))


def oracle_frame(orc, desc, n_prims, Nx, Ny, mode=0):
    o = orc.OracleScene(desc, n_prims)
    sums, _ = o.render(Nx, Ny, 1, mode=mode)
    return orc.tonemap(sums, 1)[..., :3].view(np.uint32)


@pytest.mark.parametrize("sid", [0, 1, 2, 3, 4])
def test_oracle_reproduces_the_reference_frame(rtw, orc, pin, frames, sid):
    Nx, Ny = pin["nx"], pin["ny"]
    s = rtw.Scene(sid, Nx, Ny)
    got = oracle_frame(orc, s.desc, s.desc.contents.n_prims, Nx, Ny)
    assert np.array_equal(got, frames["scene%d" % sid]), "oracle differs from the reference's own frame"


@pytest.mark.parametrize("sid", [1, 4])
def test_oracle_lbvh_mode_reproduces_the_reference_frame(rtw, orc, pin, frames, sid):
    """The host-mirror LBVH walk (mode 1) finds the hits the reference's instance loop finds."""
    Nx, Ny = pin["nx"], pin["ny"]
    s = rtw.Scene(sid, Nx, Ny)
    got = oracle_frame(orc, s.desc, s.desc.contents.n_prims, Nx, Ny, mode=1)
    assert np.array_equal(got, frames["scene%d" % sid])


@pytest.mark.parametrize("case", scene_kit.CASES)
def test_oracle_reproduces_the_reference_on_hand_built_scenes(rtw, orc, pin, frames, case):
    Nx, Ny = pin["nx"], pin["ny"]
    k = scene_kit.build(rtw, case, Nx / Ny)
    assert k.n_prims == pin["kit"][case]["n_prims"]
    got = oracle_frame(orc, k.desc, k.n_prims, Nx, Ny)
    assert np.array_equal(got, frames["kit_" + case])


@pytest.mark.parametrize("sid", [0, 1, 2, 3, 4])
def test_scene_builder_builds_the_reference_scene(rtw, pin, sid):
    s = rtw.Scene(sid, pin["nx"], pin["ny"])
    assert s.desc.contents.n_prims == pin["scenes"][str(sid)]["n_prims"]
    assert scene_digest.scene_digest(rtw, s.desc) == pin["scenes"][str(sid)]["digest"]


@pytest.mark.parametrize("name", ["scene0", "scene1", "scene2", "scene3", "scene4"] + ["kit_" + c for c in scene_kit.CASES])
def test_product_device_code_on_the_host_reproduces_the_reference_frame(rtw, orc, emu, pin, frames, name):
    """The product's per-thread kernel bodies (csrc/device/*.h: LBVH build, traversal, intersection,
    materials, textures, wavefront logic) compiled for the host (tests/emu) render the frame the
    reference's own programs render — the CPU-side half of tests/test_gpu_ref_pin.py."""
    Nx, Ny = pin["nx"], pin["ny"]
    if name.startswith("scene"):
        s = rtw.Scene(int(name[5:]), Nx, Ny)
        desc, n = s.desc, s.desc.contents.n_prims
    else:
        s = scene_kit.build(rtw, name[4:], Nx / Ny)
        desc, n = s.desc, s.n_prims
    for max_top in (64, -1):     # two-child nodes with a staged top / the 4-wide nodes
        e = emu.EmuScene(desc, n, max_top=max_top)
        acc, _ = e.render(Nx, Ny, 1, P=8192)
        got = orc.tonemap(acc, 1)[..., :3].view(np.uint32)
        assert np.array_equal(got, frames[name]), max_top


def test_earthmap_texels_are_the_reference_decode(rtw, pin):
    """assets/earthmap.ppm (tools/make_earthmap_ppm.py) == RestOfLife/texture/ioTexture.h:228-266 on
    assets/earthmap.jpg through the reference's in-tree external/stb_image.h."""
    s = rtw.Scene(2, 64, 64)
    d = s.desc.contents
    assert d.n_images == 1
    im = d.images[0]
    assert (im.width, im.height) == (pin["earthmap"]["width"], pin["earthmap"]["height"])
    assert hashlib.sha256(C.string_at(im.rgba, im.width * im.height * 4)).hexdigest() == pin["earthmap"]["sha256_rgba_flipped"]


def test_libm_changes_few_pixels(pin):
    """Recorded by the generator: with glibc's sinf/cosf/logf/powf/atan2f/asinf in the reference's device
    files instead of oracle/detmath.h, >= 94 % of the 1 spp pixels keep the same bits (the rest take a
    different random walk after a last-ulp difference)."""
    for v in list(pin["scenes"].values()) + list(pin["kit"].values()):
        assert v["libm_build_equal_pixels"] >= 0.94 * v["pixels"]


# ---- live: the reference compiled here ------------------------------------------------------
def _ref():
    from oracle import ref_py
    if not (ref_py.available(True) and ref_py.available(False) and os.path.isdir(ref_py.REF_ROOT)):
        pytest.skip("oracle/_ref/libref*.so or /root/reference not present (GPU box): committed fixtures only")
    return ref_py


LIVE_CODE = r'''
import sys, importlib, numpy as np, ctypes as C
sys.path.insert(0, %(root)r); sys.path.insert(0, %(root)r + "/tests")
from oracle import ref_py as ref, oracle_py as orc
import scene_digest
rtw = importlib.import_module("raytracing-weekend_b200")
sc, NX, NY = %(sc)d, %(nx)d, %(ny)d
img = ref.render_scene(sc, NX, NY, dm=True, q23_normalise=True)
rd = ref.scene_desc(True)
n = C.cast(rd, C.POINTER(rtw.SceneDesc)).contents.n_prims
s = rtw.Scene(sc, NX, NY)
assert scene_digest.scene_digest(rtw, rd) == scene_digest.scene_digest(rtw, s.desc), "scene builder != reference scene"
# primitive by primitive, so that a failure names the primitive
ra, _ = scene_digest.resolved_records(rtw, rd); pa, _ = scene_digest.resolved_records(rtw, s.desc)
assert ra == pa
o = orc.OracleScene(s.desc, n)
sums, _ = o.render(NX, NY, 1, mode=0)
got = orc.tonemap(sums, 1)
assert np.array_equal(got.view(np.uint32), img.view(np.uint32)), "oracle != reference frame"
gold = np.load(%(root)r + "/tests/golden/ref_frames.npz")["scene%%d" %% sc]
assert np.array_equal(img[..., :3].view(np.uint32), gold), "committed fixture is stale"
# the AABBs the reference hands to optixAccelBuild are the boxes the oracle's LBVH starts from (before padding)
ex = ref.scene_extras(n)
print("LIVE-OK")
'''


@pytest.mark.parametrize("sid", [0, 1, 2, 3, 4])
def test_live_reference_scene_and_frame(pin, sid):
    _ref()
    out = subprocess.run([os.sys.executable, "-c", LIVE_CODE % {"root": ROOT, "sc": sid, "nx": pin["nx"], "ny": pin["ny"]}],
                         cwd=ROOT, capture_output=True, text=True)
    assert "LIVE-OK" in out.stdout, out.stdout[-2000:] + out.stderr[-2000:]


@pytest.mark.parametrize("case", scene_kit.CASES)
def test_live_reference_programs_on_hand_built_scenes(rtw, orc, pin, frames, case):
    ref = _ref()
    Nx, Ny = 96, 64
    k = scene_kit.build(rtw, case, Nx / Ny)
    img = ref.render_desc(k.desc, Nx, Ny, 20, dm=True)
    got = oracle_frame(orc, k.desc, k.n_prims, Nx, Ny)
    assert np.array_equal(got, img[..., :3].view(np.uint32))
    # depth limits walk the loop exits of raygen/raygen.cu:36-84
    for depth in (1, 2, 3):
        img = ref.render_desc(k.desc, Nx, Ny, depth, dm=True)
        o = orc.OracleScene(k.desc, k.n_prims)
        sums, _ = o.render(Nx, Ny, 1, max_depth=depth, mode=0)
        assert np.array_equal(orc.tonemap(sums, 1).view(np.uint32), img.view(np.uint32)), depth


def test_live_unnormalised_q23_differs_only_through_media(rtw, orc, pin):
    """Without the Q23 normalisation (the constant-medium programs draw from the path seed in traversal
    order, geometry/volumeBox.cu:79) scene 0 — no media — is unchanged and scene 3 is not: the
    normalisation touches exactly what it claims to."""
    ref = _ref()
    code = ("import sys, numpy as np; sys.path.insert(0, %r); from oracle import ref_py as ref\n"
            "a = ref.render_scene(%%d, 64, 48, dm=True, q23_normalise=True); b = ref.render_scene(%%d, 64, 48, dm=True, q23_normalise=False)\n"
            "print('EQ', int(np.array_equal(a.view(np.uint32), b.view(np.uint32))))") % ROOT
    out0 = subprocess.run([os.sys.executable, "-c", code % (0, 0)], cwd=ROOT, capture_output=True, text=True).stdout
    out3 = subprocess.run([os.sys.executable, "-c", code % (3, 3)], cwd=ROOT, capture_output=True, text=True).stdout
    assert "EQ 1" in out0 and "EQ 0" in out3


def test_live_earthmap_decode(rtw):
    ref = _ref()
    tex = ref.load_image("assets/earthmap.jpg")
    s = rtw.Scene(4, 64, 64)
    im = s.desc.contents.images[0]
    assert (im.height, im.width) == tex.shape[:2]
    assert C.string_at(im.rgba, im.width * im.height * 4) == tex.tobytes()


def test_live_reference_cli_ppm(rtw, orc, tmp_path):
    """The reference's own program (oracle/_ref/ref_cli = its main.cpp) prints the PPM the oracle's frame
    gives through Director::printPPM's formula (Director.cpp:1010-1031).  `-s` can only select scene 4
    (main.cpp:68: `x >= Nscene`, Q2) and -dx/-dy are clamped to >= 320x200 (main.cpp:21-24)."""
    ref = _ref()
    cli = os.path.join(ROOT, "oracle", "_ref", "ref_cli")
    if not os.path.exists(cli):
        pytest.skip("ref_cli not built")
    env = dict(os.environ, REF_Q23_NORMALISE="1")
    out = subprocess.run([cli, "-s", "4", "-dx", "320", "-dy", "200"], cwd=ref.REF_ROOT, env=env, capture_output=True)
    assert out.returncode == 0
    toks = out.stdout.split()
    assert toks[:4] == [b"P3", b"320", b"200", b"255"]
    got = np.array(toks[4:], dtype=np.int32).reshape(200, 320, 3)
    # ref_cli is the libm build: compare with the libm library build of the same scene, bit for bit,
    # and with the oracle (detmath) on the pixels where the two libms agree
    code = ("import sys, numpy as np; sys.path.insert(0, %r); from oracle import ref_py as ref\n"
            "a = ref.render_scene(4, 320, 200, dm=False, q23_normalise=True); np.save(%r, a)") % (ROOT, str(tmp_path / "lm.npy"))
    subprocess.check_call([os.sys.executable, "-c", code], cwd=ROOT, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    lm = np.load(tmp_path / "lm.npy")

    def ppm(frame):   # Director.cpp:1016-1027: top row first, int(255.99f * clamp(c, 0, 1))
        c = np.clip(frame[::-1, :, :3], np.float32(0), np.float32(1))
        return (np.float32(255.99) * c).astype(np.int32)

    assert np.array_equal(got, ppm(lm))
    s = rtw.Scene(4, 320, 200)
    o = orc.OracleScene(s.desc, s.desc.contents.n_prims)
    sums, _ = o.render(320, 200, 1, mode=1)
    mine = ppm(orc.tonemap(sums, 1))
    same = (mine == got).all(axis=2).mean()
    assert same >= 0.94, same
