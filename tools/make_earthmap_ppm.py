#!/usr/bin/env python3
"""Decode the reference's earth-map JPEG into a binary PPM (P6) the C++ scene builder can
read without a JPEG decoder.

The texels come from the REFERENCE'S OWN decode path: oracle/_ref/libref.so contains the
reference's ioImageTexture (RestOfLife/texture/ioTexture.h:228-266) compiled where it lies
together with its in-tree external/stb_image.h (`make -C oracle ref`); ref_load_image returns
the RGBA8 array that code uploads (already flipped vertically, row 0 = bottom).  The PPM
stores the rows top-down again — the scene builder applies the flip itself — so the file is
byte for byte what stb_image decodes.  tests/test_ref_pin.py asserts the equality whenever
/root/reference is present.

Usage: python tools/make_earthmap_ppm.py [assets/earthmap.ppm]
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import ref_py  # noqa: E402

dst = sys.argv[1] if len(sys.argv) > 1 else "assets/earthmap.ppm"
rgba = ref_py.load_image("assets/earthmap.jpg")
h, w = rgba.shape[:2]
rgb = rgba[::-1, :, :3]
with open(dst, "wb") as f:
    f.write(b"P6\n%d %d\n255\n" % (w, h))
    f.write(rgb.tobytes())
print("wrote %s: %dx%d" % (dst, w, h))
