"""TEST HARNESS ONLY: ctypes binding of tests/emu/libemu.so — the product's per-thread
kernel bodies (csrc/device/*.h) compiled for the host and driven by plain loops.
Built on demand with g++; never part of librtw.so."""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SRC = os.path.join(_HERE, "emu", "emu.cpp")
_LIB = os.path.join(_HERE, "emu", "libemu.so")
_DEV = os.path.join(os.path.dirname(_HERE), "raytracing-weekend_b200", "csrc", "device")

RAY_DTYPE = np.dtype([("origin", "<f4", 3), ("tmin", "<f4"), ("dir", "<f4", 3), ("tmax", "<f4"), ("time", "<f4"),
                      ("gather_time", "<f4"), ("seed", "<u4"), ("pad", "<u4")])
HIT_DTYPE = np.dtype([("t", "<f4"), ("prim", "<i4")])

_lib = None


def _stale():
    if not os.path.exists(_LIB):
        return True
    t = os.path.getmtime(_LIB)
    deps = [_SRC] + [os.path.join(_DEV, f) for f in os.listdir(_DEV)]
    return any(os.path.getmtime(d) > t for d in deps)


def lib():
    global _lib
    if _lib is not None:
        return _lib
    if _stale():
        cuda_inc = os.path.join(os.environ.get("CUDA_HOME", "/usr/local/cuda"), "include")
        subprocess.check_call(["g++", "-O2", "-std=c++17", "-fPIC", "-ffp-contract=off", "-mfma", "-Wno-unknown-pragmas",
                               "-I" + cuda_inc, "-shared", "-o", _LIB, _SRC])
    L = C.CDLL(_LIB)
    vp, i32 = C.c_void_p, C.c_int
    L.emu_create.restype = vp
    L.emu_create.argtypes = [vp, i32, i32]
    L.emu_destroy.argtypes = [vp]
    L.emu_bvh_download.argtypes = [vp] * 7
    L.emu_n_internal.restype = i32
    L.emu_n_internal.argtypes = [vp]
    L.emu_trace.argtypes = [vp, vp, i32, vp, vp, vp, vp]
    L.emu_render.restype = i32
    L.emu_render.argtypes = [vp, i32, i32, i32, i32, i32, i32, i32, vp, vp]
    L.emu_camera_rays.argtypes = [vp, i32, i32, i32, vp]
    L.emu_texture_eval.argtypes = [vp, i32, C.c_float, C.c_float, vp, vp]
    L.emu_math_eval.argtypes = [i32, vp, vp, i32, vp]
    L.emu_rng_stream.argtypes = [C.c_uint32, C.c_uint32, i32, vp, vp, vp, vp, vp]
    _lib = L
    return L


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p)


class EmuScene:
    def __init__(self, desc_ptr, n_prims, max_top=1024, leaf_size=0):
        self._h = lib().emu_create(C.cast(desc_ptr, C.c_void_p), max_top, leaf_size)
        self.n = n_prims

    def close(self):
        if self._h:
            lib().emu_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def bvh_download(self):
        n, ni = self.n, int(lib().emu_n_internal(self._h))
        morton = np.zeros(n, np.uint64)
        leaf_prim = np.zeros(n, np.int32)
        left = np.zeros(ni, np.int32)
        right = np.zeros(ni, np.int32)
        bounds = np.zeros((ni, 12), np.float32)
        box = np.zeros(6, np.float32)
        lib().emu_bvh_download(self._h, _ptr(morton), _ptr(leaf_prim), _ptr(left), _ptr(right), _ptr(bounds), _ptr(box))
        return morton, leaf_prim, left, right, bounds, box

    def trace_closest(self, rays):
        rays = np.ascontiguousarray(rays, dtype=RAY_DTYPE)
        hits = np.empty(len(rays), dtype=HIT_DTYPE)
        nv = np.zeros(1, np.uint64)
        pt = np.zeros(1, np.uint64)
        lib().emu_trace(self._h, _ptr(rays), len(rays), _ptr(hits), None, _ptr(nv), _ptr(pt))
        return hits, int(nv[0]), int(pt[0])

    def trace_any(self, rays):
        rays = np.ascontiguousarray(rays, dtype=RAY_DTYPE)
        occ = np.empty(len(rays), dtype=np.int32)
        lib().emu_trace(self._h, _ptr(rays), len(rays), None, _ptr(occ), None, None)
        return occ

    def render(self, Nx, Ny, spp, max_depth=20, P=4096, sample_first=0, sample_stride=1):
        acc = np.zeros((Ny, Nx, 4), np.float32)
        st = np.zeros(3, np.uint64)
        lib().emu_render(self._h, Nx, Ny, sample_first, sample_stride, spp, max_depth, P, _ptr(acc), _ptr(st))
        return acc, {"rays_closest": int(st[0]), "rays_shadow": int(st[1]), "iterations": int(st[2])}

    def camera_rays(self, Nx, Ny, sample=0):
        rays = np.empty(Nx * Ny, dtype=RAY_DTYPE)
        lib().emu_camera_rays(self._h, Nx, Ny, sample, _ptr(rays))
        return rays

    def texture_eval(self, tex, u, v, p):
        p = np.asarray(p, np.float32)
        out = np.zeros(3, np.float32)
        lib().emu_texture_eval(self._h, tex, u, v, _ptr(p), _ptr(out))
        return out


def rng_stream(seed0, seed1, n):
    t = np.zeros(1, np.uint32)
    lcg = np.zeros(n, np.uint32)
    rnd = np.zeros(n, np.float32)
    xs = np.zeros(n, np.uint32)
    rf = np.zeros(n, np.float32)
    lib().emu_rng_stream(seed0, seed1, n, _ptr(t), _ptr(lcg), _ptr(rnd), _ptr(xs), _ptr(rf))
    return int(t[0]), lcg, rnd, xs, rf


def math_eval(fn, x, y=None):
    """csrc/device/detmath.h compiled for the host (fn numbered like rtw_math_fn)."""
    x = np.ascontiguousarray(x, np.float32)
    out = np.empty_like(x)
    yy = np.ascontiguousarray(y, np.float32) if y is not None else None
    lib().emu_math_eval(fn, _ptr(x), _ptr(yy) if yy is not None else None, len(x), _ptr(out))
    return out
