"""ORACLE — TEST INFRASTRUCTURE ONLY.

ctypes binding of oracle/liboracle.so (the CPU restatement of the reference's
path-tracing launch).  Imported only by tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline / --impl reference legs — never by the product package.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "liboracle.so")


class Counters(C.Structure):
    _fields_ = [("paths", C.c_uint64), ("rays_closest", C.c_uint64), ("rays_shadow", C.c_uint64),
                ("nodes_visited", C.c_uint64), ("prims_tested", C.c_uint64), ("vertices_shaded", C.c_uint64),
                ("shadow_enqueued", C.c_uint64),
                ("nodes_closest", C.c_uint64), ("prims_closest", C.c_uint64), ("prims_xf_closest", C.c_uint64),
                ("nodes_shadow", C.c_uint64), ("prims_shadow", C.c_uint64), ("prims_xf_shadow", C.c_uint64)]

    def as_dict(self):
        return {k: int(getattr(self, k)) for k, _ in self._fields_}


_lib = None


def lib():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError("oracle/liboracle.so not built: run `make -C oracle`")
    L = C.CDLL(LIB_PATH)
    vp, i32, u32 = C.c_void_p, C.c_int, C.c_uint32
    L.orc_tea.restype = u32
    L.orc_tea.argtypes = [u32, u32, u32]
    L.orc_rng_stream.restype = None
    L.orc_rng_stream.argtypes = [u32, u32, i32, vp, vp, vp, vp, vp]
    L.orc_math_eval.restype = None
    L.orc_math_eval.argtypes = [i32, vp, vp, i32, vp]
    L.orc_scene_load.restype = i32
    L.orc_scene_load.argtypes = [vp, C.POINTER(vp)]
    L.orc_scene_free.restype = None
    L.orc_scene_free.argtypes = [vp]
    L.orc_camera_rays.restype = None
    L.orc_camera_rays.argtypes = [vp, i32, i32, i32, vp]
    L.orc_trace_closest.restype = None
    L.orc_trace_closest.argtypes = [vp, vp, i32, vp, i32, C.POINTER(Counters)]
    L.orc_trace_any.restype = None
    L.orc_trace_any.argtypes = [vp, vp, i32, vp, i32, C.POINTER(Counters)]
    L.orc_render.restype = i32
    L.orc_render.argtypes = [vp, i32, i32, i32, i32, i32, i32, i32, i32, vp, C.POINTER(Counters)]
    L.orc_render_rows.restype = i32
    L.orc_render_rows.argtypes = [vp, i32, i32, i32, i32, i32, i32, i32, i32, i32, i32, vp, C.POINTER(Counters)]
    L.orc_tonemap.restype = None
    L.orc_tonemap.argtypes = [i32, i32, vp, vp]
    L.orc_sample_radiance.restype = None
    L.orc_sample_radiance.argtypes = [vp, i32, i32, i32, i32, i32, i32, i32, vp]
    L.orc_bvh_build.restype = i32
    L.orc_bvh_build.argtypes = [vp]
    L.orc_bvh_set_leaf_size.restype = None
    L.orc_bvh_set_leaf_size.argtypes = [vp, i32]
    L.orc_bvh_info.restype = i32
    L.orc_bvh_info.argtypes = [vp, vp]
    L.orc_bvh_download.restype = i32
    L.orc_bvh_download.argtypes = [vp, vp, vp, vp, vp, vp]
    L.orc_prim_aabbs.restype = None
    L.orc_prim_aabbs.argtypes = [vp, vp]
    L.orc_denoise.restype = None
    L.orc_denoise.argtypes = [i32, i32, i32, C.c_float, vp, vp]
    L.orc_texture_eval.restype = None
    L.orc_texture_eval.argtypes = [vp, i32, C.c_float, C.c_float, vp, vp]
    _lib = L
    return L


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p)


def math_eval(fn, x, y=None):
    """oracle/detmath.h: fn numbered like rtw_math_fn (0 sin, 1 cos, 2 log, 3 pow5, 4 atan2, 5 asin)."""
    x = np.ascontiguousarray(x, np.float32)
    out = np.empty_like(x)
    yy = np.ascontiguousarray(y, np.float32) if y is not None else None
    lib().orc_math_eval(fn, _ptr(x), _ptr(yy) if yy is not None else None, len(x), _ptr(out))
    return out


RAY_DTYPE = np.dtype([("origin", "<f4", 3), ("tmin", "<f4"), ("dir", "<f4", 3), ("tmax", "<f4"), ("time", "<f4"),
                      ("gather_time", "<f4"), ("seed", "<u4"), ("pad", "<u4")])
HIT_DTYPE = np.dtype([("t", "<f4"), ("prim", "<i4")])


def tea(rounds, s0, s1):
    return int(lib().orc_tea(rounds, s0 & 0xffffffff, s1 & 0xffffffff))


def rng_stream(seed0, seed1, n):
    t = np.zeros(1, np.uint32)
    lcg = np.zeros(n, np.uint32)
    rnd = np.zeros(n, np.float32)
    xs = np.zeros(n, np.uint32)
    rf = np.zeros(n, np.float32)
    lib().orc_rng_stream(seed0, seed1, n, _ptr(t), _ptr(lcg), _ptr(rnd), _ptr(xs), _ptr(rf))
    return int(t[0]), lcg, rnd, xs, rf


class OracleScene:
    """Deep copy of an rtw_scene_desc (pass the ctypes pointer the product's scene builder returns)."""

    def __init__(self, desc_ptr, n_prims):
        self._h = C.c_void_p()
        addr = C.cast(desc_ptr, C.c_void_p)
        if lib().orc_scene_load(addr, C.byref(self._h)) != 0:
            raise RuntimeError("orc_scene_load failed")
        self.n_prims = n_prims
        self._bvh = False

    def close(self):
        if self._h:
            lib().orc_scene_free(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def build_bvh(self):
        if not self._bvh:
            if lib().orc_bvh_build(self._h) != 0:
                raise RuntimeError("orc_bvh_build failed")
            lib().orc_bvh_set_leaf_size(self._h, 0)     # the product's default leaf-range size
            self._bvh = True

    def set_leaf_size(self, leaf_size):
        """Leaf-range size used by the BVH-mode traversal counters (0 = product default)."""
        self.build_bvh()
        lib().orc_bvh_set_leaf_size(self._h, leaf_size)

    def camera_rays(self, Nx, Ny, sample=0):
        rays = np.empty(Nx * Ny, dtype=RAY_DTYPE)
        lib().orc_camera_rays(self._h, Nx, Ny, sample, _ptr(rays))
        return rays

    def trace_closest(self, rays, mode=0):
        if mode:
            self.build_bvh()
        rays = np.ascontiguousarray(rays, dtype=RAY_DTYPE)
        hits = np.empty(len(rays), dtype=HIT_DTYPE)
        c = Counters()
        lib().orc_trace_closest(self._h, _ptr(rays), len(rays), _ptr(hits), mode, C.byref(c))
        return hits, c.as_dict()

    def trace_any(self, rays, mode=0):
        if mode:
            self.build_bvh()
        rays = np.ascontiguousarray(rays, dtype=RAY_DTYPE)
        occ = np.empty(len(rays), dtype=np.int32)
        c = Counters()
        lib().orc_trace_any(self._h, _ptr(rays), len(rays), _ptr(occ), mode, C.byref(c))
        return occ, c.as_dict()

    def render(self, Nx, Ny, spp, max_depth=20, n_threads=0, mode=0, sample_first=0, sample_stride=1, rows=None):
        """Returns (linear sums [Ny,Nx,4] with w = sample count, counters)."""
        if mode:
            self.build_bvh()
        out = np.zeros((Ny, Nx, 4), dtype=np.float32)
        c = Counters()
        y0, y1 = rows if rows is not None else (0, Ny)
        rc = lib().orc_render_rows(self._h, Nx, Ny, y0, y1, sample_first, sample_stride, spp, max_depth, n_threads, mode,
                                   _ptr(out), C.byref(c))
        if rc != 0:
            raise RuntimeError("orc_render failed")
        return out, c.as_dict()

    def sample_radiance(self, Nx, Ny, x, y, k, max_depth=20, mode=0):
        out = np.zeros(3, np.float32)
        lib().orc_sample_radiance(self._h, Nx, Ny, x, y, k, max_depth, mode, _ptr(out))
        return out

    def bvh_download(self):
        self.build_bvh()
        n = self.n_prims
        class _Info(C.Structure):   # rtw_bvh_info (include/rtw.h)
            _fields_ = [("n_leaves", C.c_int32), ("n_internal", C.c_int32), ("root", C.c_int32), ("n_top", C.c_int32),
                        ("scene_min", C.c_float * 3), ("scene_max", C.c_float * 3)]
        info = _Info()
        assert lib().orc_bvh_info(self._h, C.byref(info)) == 0
        ni = info.n_internal     # n - 1, less the primitives of the oversize list (csrc/device/lbvh.h)
        morton = np.zeros(n, np.uint64)
        leaf_prim = np.zeros(n, np.int32)
        left = np.zeros(ni, np.int32)
        right = np.zeros(ni, np.int32)
        bounds = np.zeros((ni, 12), np.float32)
        lib().orc_bvh_download(self._h, _ptr(morton), _ptr(leaf_prim), _ptr(left), _ptr(right), _ptr(bounds))
        return morton, leaf_prim, left, right, bounds

    def prim_aabbs(self):
        out = np.zeros((self.n_prims, 6), np.float32)
        lib().orc_prim_aabbs(self._h, _ptr(out))
        return out

    def texture_eval(self, tex, u, v, p):
        p = np.asarray(p, np.float32)
        out = np.zeros(3, np.float32)
        lib().orc_texture_eval(self._h, tex, u, v, _ptr(p), _ptr(out))
        return out


def tonemap(sums, spp):
    out = np.empty_like(sums)
    lib().orc_tonemap(sums.shape[0] * sums.shape[1], spp, _ptr(np.ascontiguousarray(sums)), _ptr(out))
    return out


def denoise(frame, iterations=3, sigma=0.25):
    """The product's denoise stand-in restated on the CPU (oracle.cpp orc_denoise)."""
    frame = np.ascontiguousarray(frame, np.float32)
    out = np.empty_like(frame)
    lib().orc_denoise(frame.shape[1], frame.shape[0], iterations, sigma, _ptr(frame), _ptr(out))
    return out
