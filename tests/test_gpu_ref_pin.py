"""-m gpu: the CUDA path (through the C-ABI) reproduces, bit for bit, the frames THE REFERENCE'S OWN
PROGRAMS render (tests/golden/ref_frames.npz, written by tests/golden/make_ref_golden.py from the
reference compiled for the host, oracle/_ref): scenes 0-4 and the hand-built scenes that reach the
branches no built-in scene does (checker / null texture, normal material, X-axis light rectangle, two
lights, lens radius > 0, transformed image-textured sphere, moving sphere, transformed medium)."""
import json
import os

import numpy as np
import pytest

import scene_kit

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = os.path.join(ROOT, "tests", "golden")


@pytest.fixture(scope="module")
def pin():
    with open(os.path.join(GOLD, "ref_pin.json")) as f:
        return json.load(f)


@pytest.fixture(scope="module")
def frames():
    return np.load(os.path.join(GOLD, "ref_frames.npz"))


def _frame(director, Nx, Ny):
    """rtw_render's host image: sqrt-gamma float4, row 0 = bottom (raygen/raygen.cu:144-158)."""
    return director.renderFrame(1)[..., :3].view(np.uint32)


@pytest.mark.parametrize("sid", [0, 1, 2, 3, 4])
def test_cuda_reproduces_the_reference_frame(director, pin, frames, sid):
    Nx, Ny = pin["nx"], pin["ny"]
    director.init(Nx, Ny, 1)
    director.createScene(sid)
    assert np.array_equal(_frame(director, Nx, Ny), frames["scene%d" % sid])


@pytest.mark.parametrize("case", scene_kit.CASES)
def test_cuda_reproduces_the_reference_on_hand_built_scenes(director, rtw, orc, pin, frames, case):
    Nx, Ny = pin["nx"], pin["ny"]
    k = scene_kit.build(rtw, case, Nx / Ny)
    director.init(Nx, Ny, 1)
    director.scene = k
    director.upload(k.desc)
    assert np.array_equal(_frame(director, Nx, Ny), frames["kit_" + case])
    # and the traversal variants / several samples still agree with the oracle on these scenes
    o = orc.OracleScene(k.desc, k.n_prims)
    ref, c = o.render(Nx, Ny, 4, mode=0)
    director.renderFrame(4)
    lin = director.linear()[..., :3]
    r = ref[..., :3] / 4
    assert float(np.mean((lin - r) ** 2 / (r ** 2 + 1e-2))) <= 1e-10
    st = director.stats()
    assert st["rays_closest"] == c["rays_closest"] and st["rays_shadow"] == c["rays_shadow"]


@pytest.mark.parametrize("case", scene_kit.CASES)
def test_cuda_hits_on_hand_built_scenes(director, rtw, orc, case):
    """closest-hit (t, id) and any-hit answers on camera + random rays, against the oracle's brute force."""
    import rayset
    k = scene_kit.build(rtw, case, 1.5)
    director.init(96, 64, 1)
    director.scene = k
    director.upload(k.desc)
    o = orc.OracleScene(k.desc, k.n_prims)
    cam = o.camera_rays(96, 64, 0)
    assert cam.tobytes() == director.camera_rays(0).tobytes()     # lens radius > 0 included ("branches", "sky")
    box = o.prim_aabbs()
    rays = rayset.make_rays(cam, box[:, :3].min(axis=0), box[:, 3:].max(axis=0), 1 << 15)
    hits = director.trace_closest(rays)
    ref, _ = o.trace_closest(rays, 0)
    assert np.array_equal(hits["prim"], ref["prim"])
    ok = ref["prim"] >= 0
    assert np.array_equal(hits["t"][ok].view(np.uint32), ref["t"][ok].view(np.uint32))
    sh = rayset.shadow_rays_from(rays, ref, [5.0, 9.0, 5.0], 1 << 14)
    assert np.array_equal(director.trace_any(sh), o.trace_any(sh, 0)[0])
