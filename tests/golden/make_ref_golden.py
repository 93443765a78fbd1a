#!/usr/bin/env python3
"""Writes the fixtures that pin the oracle, the host scene builder and the CUDA path to THE
REFERENCE ITSELF:

  tests/golden/ref_frames.npz   the frame buffer the reference's own programs write (uint32 bit
                                patterns of the sqrt-gamma float4 frame, rgb only) for
                                  scene0..scene4  Director::init -> createScene -> renderFrame
                                                  (RestOfLife/main.cpp:148-157), 120x80, 1 spp
                                  kit_<case>      the reference's DEVICE programs on the hand-built
                                                  scenes of tests/scene_kit.py (branches no
                                                  built-in scene reaches), 120x80, 1 spp
  tests/golden/ref_pin.json     digest of each scene AS THE REFERENCE'S HOST CODE BUILT IT
                                (tests/scene_digest.py), primitive counts, sha256 of the earth-map
                                texels as the reference's stb_image decodes them, and — for the
                                record — how many pixels of the libm build (glibc sinf/cosf/... in
                                place of oracle/detmath.h) are bit-equal to the detmath build

The producer is oracle/_ref/libref_dm.so: the reference's main.cpp, Director.cpp, scene/geometry/
material/texture headers and all 26 device .cu files compiled where they lie under
/root/reference by `make -C oracle ref` (stand-in headers: oracle/ref_shim).  Needs
/root/reference; run from the repo root:  python tests/golden/make_ref_golden.py
"""
import hashlib
import importlib
import json
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

NX, NY = 120, 80


def fresh(code):
    """The reference keeps process-wide state (texture/ioTexture.h:21-26: a static mt19937 feeds the
    Perlin tables), so every scene is built in a fresh process, like the reference's own main()."""
    out = subprocess.check_output([sys.executable, "-c", code], cwd=ROOT)
    return out


SCENE_CODE = r'''
import sys, importlib, json, hashlib, numpy as np
sys.path.insert(0, %(root)r); sys.path.insert(0, %(root)r + "/tests")
from oracle import ref_py as ref
import scene_digest
rtw = importlib.import_module("raytracing-weekend_b200")
sc, NX, NY = %(sc)d, %(nx)d, %(ny)d
img = ref.render_scene(sc, NX, NY, dm=True, q23_normalise=True)
d = ref.scene_desc(True)
import ctypes as C
n = C.cast(d, C.POINTER(rtw.SceneDesc)).contents.n_prims
ex = ref.scene_extras(n)
assert (ex["instance_ids"] == np.arange(n)).all()
dig = scene_digest.scene_digest(rtw, d)
img_lm = ref.render_scene(sc, NX, NY, dm=False, q23_normalise=True)
eq = int((img.view(np.uint32) == img_lm.view(np.uint32)).all(axis=2).sum())
np.save(%(out)r, img)
print(json.dumps({"digest": dig, "n_prims": int(n), "libm_build_equal_pixels": eq, "pixels": NX * NY}))
'''


def main():
    rtw = importlib.import_module("raytracing-weekend_b200")
    from oracle import ref_py as ref
    import scene_kit
    subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle"), "ref"])
    frames, pin = {}, {"nx": NX, "ny": NY, "scenes": {}, "kit": {}}
    tmp = os.path.join(ROOT, "tests", "golden", "_tmp_frame.npy")
    for sc in range(5):
        info = json.loads(fresh(SCENE_CODE % {"root": ROOT, "sc": sc, "nx": NX, "ny": NY, "out": tmp}).decode().strip().splitlines()[-1])
        frames["scene%d" % sc] = np.load(tmp)[..., :3].view(np.uint32).copy()
        pin["scenes"][str(sc)] = info
        print("scene", sc, info)
    os.remove(tmp)
    for name in scene_kit.CASES:
        k = scene_kit.build(rtw, name, NX / NY)
        img = ref.render_desc(k.desc, NX, NY, 20, dm=True)
        img_lm = ref.render_desc(k.desc, NX, NY, 20, dm=False)
        frames["kit_" + name] = img[..., :3].view(np.uint32).copy()
        pin["kit"][name] = {"n_prims": k.n_prims,
                            "libm_build_equal_pixels": int((img.view(np.uint32) == img_lm.view(np.uint32)).all(axis=2).sum()),
                            "pixels": NX * NY}
        print("kit", name, pin["kit"][name])
    tex = ref.load_image("assets/earthmap.jpg")
    pin["earthmap"] = {"width": int(tex.shape[1]), "height": int(tex.shape[0]), "sha256_rgba_flipped": hashlib.sha256(tex.tobytes()).hexdigest()}
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "ref_frames.npz"), **frames)
    with open(os.path.join(ROOT, "tests", "golden", "ref_pin.json"), "w") as f:
        json.dump(pin, f, indent=1, sort_keys=True)
        f.write("\n")
    print("wrote tests/golden/ref_frames.npz, tests/golden/ref_pin.json")


if __name__ == "__main__":
    main()
