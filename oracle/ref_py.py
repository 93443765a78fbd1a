"""ORACLE — TEST INFRASTRUCTURE ONLY.

ctypes binding of oracle/_ref/libref.so and oracle/_ref/libref_dm.so: the REFERENCE ITSELF
(main.cpp, Director.cpp, scene/geometry/material/texture headers and all 26 device .cu files
of /root/reference/RestOfLife, compiled where they lie by `make -C oracle ref` against the
OptiX-on-the-CPU stand-in of oracle/ref_shim).  `dm=True` selects the build whose device files
take sin/cos/log/x^5/atan2/asin from oracle/detmath.h instead of glibc's libm.

Used by tests/golden/make_ref_golden.py (to write the committed fixtures), by the CPU tests
that can see /root/reference, and by bench.py's reference arm.  Never by the product.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
REF_ROOT = "/root/reference/RestOfLife"
_libs = {}


def lib_path(dm):
    return os.path.join(_HERE, "_ref", "libref_dm.so" if dm else "libref.so")


def available(dm=True):
    return os.path.exists(lib_path(dm))


def assets_available():
    return os.path.exists(os.path.join(REF_ROOT, "assets", "earthmap.jpg"))


def lib(dm=True):
    if dm in _libs:
        return _libs[dm]
    p = lib_path(dm)
    if not os.path.exists(p):
        raise RuntimeError("%s not built: run `make -C oracle ref` where /root/reference exists" % p)
    L = C.CDLL(p)
    i32, vp = C.c_int, C.c_void_p
    L.ref_set_threads.restype = i32
    L.ref_set_threads.argtypes = [i32]
    L.ref_set_q23_normalise.restype = i32
    L.ref_set_q23_normalise.argtypes = [i32]
    L.ref_traversal_count.restype = C.c_ulonglong
    L.ref_last_error.restype = C.c_char_p
    L.ref_render_scene.restype = i32
    L.ref_render_scene.argtypes = [i32, i32, i32, i32, C.c_char_p, i32, vp]
    L.ref_scene_desc.restype = vp
    L.ref_scene_extras.restype = i32
    L.ref_scene_extras.argtypes = [vp, vp, vp, vp, vp]
    L.ref_render_desc.restype = i32
    L.ref_render_desc.argtypes = [vp, i32, i32, i32, vp]
    L.ref_load_image.restype = i32
    L.ref_load_image.argtypes = [C.c_char_p, C.c_char_p, vp, vp, vp, i32]
    L.ref_run_main.restype = i32
    L.ref_open.restype = i32
    L.ref_open.argtypes = [i32, i32, i32, C.c_char_p, i32]
    L.ref_launch.restype = i32
    L.ref_launch.argtypes = [i32]
    L.ref_frame.restype = i32
    L.ref_frame.argtypes = [vp]
    _libs[dm] = L
    return L


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p)


def render_scene(scene, Nx, Ny, dm=True, q23_normalise=True, threads=0, launches=1, asset_root=REF_ROOT):
    """Director::init -> createScene(scene) -> renderFrame().  Returns the launch's frame buffer
    [Ny, Nx, 4] (sqrt-gamma, alpha 1, row 0 = bottom; raygen/raygen.cu:144-158)."""
    L = lib(dm)
    L.ref_set_threads(threads)
    L.ref_set_q23_normalise(1 if q23_normalise else 0)
    out = np.zeros((Ny, Nx, 4), np.float32)
    root = asset_root if (asset_root and os.path.isdir(asset_root)) else None
    rc = L.ref_render_scene(scene, Nx, Ny, launches, root.encode() if root else None, 1, _ptr(out))
    if rc != 0:
        raise RuntimeError("ref_render_scene(%d) failed: %s" % (scene, L.ref_last_error().decode()))
    return out


def scene_desc(dm=True):
    """The scene of the last render_scene as a pointer to an rtw_scene_desc (library-owned)."""
    p = lib(dm).ref_scene_desc()
    if not p:
        raise RuntimeError("no scene: call render_scene first")
    return p


def scene_extras(n_prims, dm=True):
    aabbs = np.zeros((n_prims, 6), np.float32)
    ids = np.zeros(n_prims, np.int32)
    sbt = np.zeros(n_prims, np.int32)
    g = np.zeros(2, np.int32)
    n = lib(dm).ref_scene_extras(_ptr(aabbs), _ptr(ids), _ptr(sbt), _ptr(g[0:1]), _ptr(g[1:2]))
    assert n == n_prims, (n, n_prims)
    return {"aabbs": aabbs, "instance_ids": ids, "sbt_offsets": sbt, "pdf_gen": int(g[0]), "p1_gen": int(g[1])}


def render_desc(desc_ptr, Nx, Ny, max_depth=20, dm=True, q23_normalise=True, threads=0):
    """The reference's device programs launched on an arbitrary rtw_scene_desc (1 spp, sample 0)."""
    L = lib(dm)
    L.ref_set_threads(threads)
    L.ref_set_q23_normalise(1 if q23_normalise else 0)
    out = np.zeros((Ny, Nx, 4), np.float32)
    addr = C.cast(desc_ptr, C.c_void_p)
    rc = L.ref_render_desc(addr, Nx, Ny, max_depth, _ptr(out))
    if rc != 0:
        raise RuntimeError("ref_render_desc failed: %s" % L.ref_last_error().decode())
    return out


def load_image(file="assets/earthmap.jpg", asset_root=REF_ROOT, dm=True):
    """texture/ioTexture.h:228-266 through the reference's ioImageTexture (external/stb_image.h)."""
    L = lib(dm)
    w, h = C.c_int(0), C.c_int(0)
    rc = L.ref_load_image(asset_root.encode(), file.encode(), C.byref(w), C.byref(h), None, 0)
    if rc != 0:
        raise RuntimeError("ref_load_image failed")
    out = np.zeros((h.value, w.value, 4), np.uint8)
    rc = L.ref_load_image(asset_root.encode(), file.encode(), C.byref(w), C.byref(h), _ptr(out), out.size)
    if rc != 0:
        raise RuntimeError("ref_load_image failed")
    return out


class Session:
    """Director::init + createScene once, renderFrame() on demand (bench.py's reference arm).  Every launch is
    the reference's one and only frame: 1 spp with the seeds of sample 0 (raygen/raygen.cu:129)."""

    def __init__(self, scene, Nx, Ny, dm=False, threads=0, q23_normalise=False, asset_root=REF_ROOT):
        self.L = lib(dm)
        self.Nx, self.Ny = Nx, Ny
        self.L.ref_set_threads(threads)
        self.L.ref_set_q23_normalise(1 if q23_normalise else 0)
        root = asset_root if (asset_root and os.path.isdir(asset_root)) else None
        if self.L.ref_open(scene, Nx, Ny, root.encode() if root else None, 1) != 0:
            raise RuntimeError("ref_open(%d) failed: %s" % (scene, self.L.ref_last_error().decode()))

    def launch(self, n=1):
        t0 = self.L.ref_traversal_count()
        if self.L.ref_launch(n) != 0:
            raise RuntimeError("ref_launch failed: %s" % self.L.ref_last_error().decode())
        return int(self.L.ref_traversal_count() - t0)       # optixTraverse calls (closest-hit + shadow rays)

    def frame(self):
        out = np.zeros((self.Ny, self.Nx, 4), np.float32)
        self.L.ref_frame(_ptr(out))
        return out
