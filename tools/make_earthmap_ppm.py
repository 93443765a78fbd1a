#!/usr/bin/env python3
"""Decode the reference's earth-map JPEG into a binary PPM (P6) the C++ scene
builder can read without a JPEG decoder.

The reference decodes assets/earthmap.jpg with stb_image at run time
(RestOfLife/texture/ioTexture.h:228-313).  /root/reference does not exist on the
GPU box, and reference SOURCES (stb_image.h) are not copied into this repo, so
the texels are baked once here with PIL (libjpeg-turbo; may differ from stb's
IDCT by +-1 in some texels: documented in DESIGN.md, "image texture").

Usage: python tools/make_earthmap_ppm.py [/root/reference/RestOfLife/assets/earthmap.jpg] [assets/earthmap.ppm]
"""
import sys
from PIL import Image

src = sys.argv[1] if len(sys.argv) > 1 else "/root/reference/RestOfLife/assets/earthmap.jpg"
dst = sys.argv[2] if len(sys.argv) > 2 else "assets/earthmap.ppm"
im = Image.open(src).convert("RGB")
w, h = im.size
with open(dst, "wb") as f:
    f.write(b"P6\n%d %d\n255\n" % (w, h))
    f.write(im.tobytes())
print("wrote %s: %dx%d" % (dst, w, h))
