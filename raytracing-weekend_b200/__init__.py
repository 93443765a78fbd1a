"""ctypes binding of librtw.so — the C-ABI in include/rtw.h.

This is the Python face of the drop-in boundary.  It mirrors the
reference's ``Director`` call order (reference RestOfLife/Director.h:118-126,
main.cpp:145-164): ``init`` → ``createScene`` → ``renderFrame`` → ``printPPM`` →
``destroy``.  There is NO CPU fallback: if librtw.so is missing or no CUDA device
is usable the calls raise.

The directory name contains a hyphen, so import it with
``importlib.import_module("raytracing-weekend_b200")``.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
REPO_ROOT = os.path.dirname(_HERE)
LIB_PATH = os.environ.get("RTW_LIB") or os.path.join(_HERE, "librtw.so")   # RTW_LIB: tuning builds only
ASSET_DIR = os.path.join(REPO_ROOT, "assets")

RTW_OK, RTW_ERR_INVALID, RTW_ERR_CUDA, RTW_ERR_NO_DEVICE, RTW_ERR_STATE, RTW_ERR_IO = range(6)

PRIM_SPHERE, PRIM_MOVING_SPHERE, PRIM_RECT_X, PRIM_RECT_Y, PRIM_RECT_Z, PRIM_VOLUME_BOX, PRIM_VOLUME_SPHERE = range(7)
MAT_LAMBERTIAN, MAT_DIFFUSE_LIGHT, MAT_METAL, MAT_DIELECTRIC, MAT_ISOTROPIC, MAT_NORMAL = range(6)
TEX_CHECKER, TEX_CONSTANT, TEX_IMAGE, TEX_NOISE, TEX_NULL = range(5)


# ---- PODs of include/rtw.h -------------------------------------------------
class Prim(C.Structure):
    _fields_ = [("type", C.c_int32), ("material", C.c_int32), ("transform", C.c_int32), ("flip", C.c_int32),
                ("v", C.c_float * 12)]


class Transform(C.Structure):
    _fields_ = [("o2w", C.c_float * 12), ("w2o", C.c_float * 12)]


class Material(C.Structure):
    _fields_ = [("type", C.c_int32), ("texture", C.c_int32), ("param", C.c_float), ("pad", C.c_int32)]


class Texture(C.Structure):
    _fields_ = [("type", C.c_int32), ("color", C.c_float * 3), ("odd", C.c_int32), ("even", C.c_int32),
                ("scale", C.c_float), ("perlin", C.c_int32), ("image", C.c_int32), ("pad", C.c_int32 * 3)]


class Perlin(C.Structure):
    _fields_ = [("ranvec", (C.c_float * 3) * 256), ("perm_x", C.c_int32 * 256), ("perm_y", C.c_int32 * 256),
                ("perm_z", C.c_int32 * 256)]


class Image(C.Structure):
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("rgba", C.POINTER(C.c_uint8))]


class Light(C.Structure):
    _fields_ = [("position", C.c_float * 3), ("vecU", C.c_float * 3), ("vecV", C.c_float * 3),
                ("normal", C.c_float * 3), ("area", C.c_float), ("emission", C.c_float * 3)]


class LightRect(C.Structure):
    _fields_ = [("axis", C.c_int32), ("a0", C.c_float), ("a1", C.c_float), ("b0", C.c_float), ("b1", C.c_float),
                ("k", C.c_float)]


class Camera(C.Structure):
    _fields_ = [("origin", C.c_float * 3), ("u", C.c_float * 3), ("v", C.c_float * 3), ("w", C.c_float * 3),
                ("lower_left", C.c_float * 3), ("horizontal", C.c_float * 3), ("vertical", C.c_float * 3),
                ("lens_radius", C.c_float), ("time0", C.c_float), ("time1", C.c_float)]


class SceneDesc(C.Structure):
    _fields_ = [("n_prims", C.c_int32), ("prims", C.POINTER(Prim)),
                ("n_transforms", C.c_int32), ("transforms", C.POINTER(Transform)),
                ("n_materials", C.c_int32), ("materials", C.POINTER(Material)),
                ("n_textures", C.c_int32), ("textures", C.POINTER(Texture)),
                ("n_perlin", C.c_int32), ("perlin", C.POINTER(Perlin)),
                ("n_images", C.c_int32), ("images", C.POINTER(Image)),
                ("n_lights", C.c_int32), ("lights", C.POINTER(Light)),
                ("has_light_rect", C.c_int32), ("light_rect", LightRect),
                ("camera", Camera),
                ("sky_light", C.c_int32), ("reserved", C.c_int32)]


class Ray(C.Structure):
    _fields_ = [("origin", C.c_float * 3), ("tmin", C.c_float), ("dir", C.c_float * 3), ("tmax", C.c_float),
                ("time", C.c_float), ("gather_time", C.c_float), ("seed", C.c_uint32), ("pad", C.c_uint32)]


class Hit(C.Structure):
    _fields_ = [("t", C.c_float), ("prim", C.c_int32)]


class Stats(C.Structure):
    _fields_ = [("paths", C.c_uint64), ("rays_closest", C.c_uint64), ("rays_shadow", C.c_uint64),
                ("kernel_launches", C.c_uint64), ("iterations", C.c_uint64), ("render_ms", C.c_double),
                ("build_ms", C.c_double), ("traverse_ms", C.c_double), ("shadow_ms", C.c_double),
                ("shade_ms", C.c_double), ("logic_ms", C.c_double), ("waves", C.c_uint64), ("pool_paths", C.c_uint64),
                ("nee_unoccluded", C.c_uint64), ("stack_overflows", C.c_uint64)]


class BvhInfo(C.Structure):
    _fields_ = [("n_leaves", C.c_int32), ("n_internal", C.c_int32), ("root", C.c_int32), ("n_top", C.c_int32),
                ("scene_min", C.c_float * 3), ("scene_max", C.c_float * 3)]


RAY_DTYPE = np.dtype([("origin", "<f4", 3), ("tmin", "<f4"), ("dir", "<f4", 3), ("tmax", "<f4"), ("time", "<f4"),
                      ("gather_time", "<f4"), ("seed", "<u4"), ("pad", "<u4")])
HIT_DTYPE = np.dtype([("t", "<f4"), ("prim", "<i4")])
assert RAY_DTYPE.itemsize == C.sizeof(Ray) == 48 and HIT_DTYPE.itemsize == C.sizeof(Hit) == 8

# every symbol include/rtw.h declares (tests check that the library exports them all)
ABI_SYMBOLS = [
    "rtw_abi_version", "rtw_device_count", "rtw_create", "rtw_destroy", "rtw_last_error", "rtw_upload_scene",
    "rtw_build_accel", "rtw_render", "rtw_render_accum", "rtw_tonemap", "rtw_get_linear", "rtw_set_option",
    "rtw_get_stats", "rtw_trace_closest", "rtw_trace_any", "rtw_rng_stream", "rtw_camera_rays", "rtw_bvh_info_get",
    "rtw_bvh_download", "rtw_sort_pairs", "rtw_scene_create", "rtw_scene_get_desc", "rtw_scene_description",
    "rtw_scene_free", "rtw_write_ppm", "rtw_math_eval", "rtw_scene_save", "rtw_scene_load", "rtw_scene_io_error",
    "rtw_render_accum_rows", "rtw_create_multi", "rtw_multi_destroy", "rtw_multi_last_error", "rtw_multi_device_count",
    "rtw_multi_uses_nccl", "rtw_multi_set_option", "rtw_multi_upload_scene", "rtw_multi_build_accel", "rtw_multi_render",
    "rtw_multi_get_linear", "rtw_multi_get_stats", "rtw_denoise", "rtw_denoise_static",
    "rtw_bake_transforms", "rtw_bake_perlin", "rtw_bake_image",
]

DENOISE_STATIC_MAX_PIXELS = 1 << 19      # RTW_DENOISE_STATIC_MAX_PIXELS (include/rtw.h)

_lib = None


class RtwError(RuntimeError):
    pass


def lib():
    """Load librtw.so (built in-tree by `make -C raytracing-weekend_b200`); raises if missing."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RtwError("librtw.so not built: run `python -c 'import __graft_entry__ as g; g.build()'` "
                       "(there is no CPU fallback)")
    L = C.CDLL(LIB_PATH)
    vp, i32, u32, i64 = C.c_void_p, C.c_int, C.c_uint32, C.c_int64
    fp = C.POINTER(C.c_float)
    sig = {
        "rtw_abi_version": (i32, []),
        "rtw_device_count": (i32, [C.POINTER(i32)]),
        "rtw_create": (i32, [i32, C.POINTER(vp)]),
        "rtw_destroy": (i32, [vp]),
        "rtw_last_error": (C.c_char_p, [vp]),
        "rtw_upload_scene": (i32, [vp, C.POINTER(SceneDesc)]),
        "rtw_build_accel": (i32, [vp]),
        "rtw_render": (i32, [vp, i32, i32, i32, i32, vp]),
        "rtw_render_accum": (i32, [vp, i32, i32, i32, i32, i32, i32, vp]),
        "rtw_tonemap": (i32, [vp, i32, i32, i32, vp, vp]),
        "rtw_get_linear": (i32, [vp, vp, i32]),
        "rtw_set_option": (i32, [vp, C.c_char_p, i64]),
        "rtw_get_stats": (i32, [vp, C.POINTER(Stats)]),
        "rtw_trace_closest": (i32, [vp, vp, i32, vp]),
        "rtw_trace_any": (i32, [vp, vp, i32, vp]),
        "rtw_rng_stream": (i32, [vp, u32, u32, i32, vp, vp, vp, vp, vp]),
        "rtw_camera_rays": (i32, [vp, i32, i32, i32, vp]),
        "rtw_math_eval": (i32, [vp, i32, vp, vp, i32, vp]),
        "rtw_bvh_info_get": (i32, [vp, C.POINTER(BvhInfo)]),
        "rtw_bvh_download": (i32, [vp, vp, vp, vp, vp, vp]),
        "rtw_sort_pairs": (i32, [vp, vp, vp, i32]),
        "rtw_scene_create": (i32, [i32, i32, i32, C.c_char_p, i32, C.POINTER(vp)]),
        "rtw_scene_get_desc": (C.POINTER(SceneDesc), [vp]),
        "rtw_scene_description": (C.c_char_p, [vp]),
        "rtw_scene_free": (None, [vp]),
        "rtw_scene_save": (i32, [vp, C.c_char_p, C.c_char_p]),
        "rtw_scene_load": (i32, [C.c_char_p, C.POINTER(vp)]),
        "rtw_scene_io_error": (C.c_char_p, []),
        "rtw_write_ppm": (i32, [C.c_char_p, i32, i32, vp]),
        "rtw_bake_transforms": (i32, [vp, vp, i32, vp]),
        "rtw_bake_perlin": (i32, [vp, u32, i32, i32, vp]),
        "rtw_bake_image": (i32, [vp, vp, i32, i32, i32, vp]),
        "rtw_denoise": (i32, [vp, i32, i32, i32, C.c_float, vp, vp]),
        "rtw_denoise_static": (i32, [vp, i32, i32, i32, C.c_float, vp, vp, vp]),
        "rtw_render_accum_rows": (i32, [vp, i32, i32, i32, i32, i32, i32, i32, i32, vp]),
        "rtw_create_multi": (i32, [C.POINTER(i32), i32, C.POINTER(vp)]),
        "rtw_multi_destroy": (i32, [vp]),
        "rtw_multi_last_error": (C.c_char_p, [vp]),
        "rtw_multi_device_count": (i32, [vp]),
        "rtw_multi_uses_nccl": (i32, [vp]),
        "rtw_multi_set_option": (i32, [vp, C.c_char_p, i64]),
        "rtw_multi_upload_scene": (i32, [vp, C.POINTER(SceneDesc)]),
        "rtw_multi_build_accel": (i32, [vp]),
        "rtw_multi_render": (i32, [vp, i32, i32, i32, i32, vp]),
        "rtw_multi_get_linear": (i32, [vp, vp, i32, i32]),
        "rtw_multi_get_stats": (i32, [vp, C.POINTER(Stats), C.POINTER(Stats), C.POINTER(C.c_double), C.POINTER(i32)]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(L, name)
        fn.restype = res
        fn.argtypes = args
    _ = fp
    _lib = L
    return L


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p)


class Scene:
    """Host-side flattened scene (replaces ioScene::init, scene/ioScene.h:52-101)."""

    def __init__(self, scene_id, Nx, Ny, asset_dir=None, n_stress_prims=0):
        self._h = C.c_void_p()
        rc = lib().rtw_scene_create(scene_id, Nx, Ny, (asset_dir or ASSET_DIR).encode(), n_stress_prims,
                                    C.byref(self._h))
        if rc != RTW_OK:
            raise RtwError("rtw_scene_create(%d) failed: %d" % (scene_id, rc))
        self.scene_id, self.Nx, self.Ny = scene_id, Nx, Ny

    @classmethod
    def load(cls, path, Nx=0, Ny=0):
        """Scene wire format (rtw_scene_load): a flattened scene read back from a file."""
        self = cls.__new__(cls)
        self._h = C.c_void_p()
        rc = lib().rtw_scene_load(str(path).encode(), C.byref(self._h))
        if rc != RTW_OK:
            raise RtwError("rtw_scene_load(%s) failed: %d (%s)" % (path, rc, lib().rtw_scene_io_error().decode()))
        self.scene_id, self.Nx, self.Ny = -1, Nx, Ny
        return self

    def save(self, path):
        rc = lib().rtw_scene_save(C.cast(self.desc, C.c_void_p), self.description.encode(), str(path).encode())
        if rc != RTW_OK:
            raise RtwError("rtw_scene_save(%s) failed: %d (%s)" % (path, rc, lib().rtw_scene_io_error().decode()))

    @property
    def desc(self):
        return lib().rtw_scene_get_desc(self._h)

    @property
    def description(self):
        return lib().rtw_scene_description(self._h).decode()

    def prims(self):
        d = self.desc.contents
        return [d.prims[i] for i in range(d.n_prims)]

    def close(self):
        if self._h:
            lib().rtw_scene_free(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class Director:
    """Mirror of the reference's Director (Director.h:118-126) over the C-ABI."""

    def __init__(self, device=0):
        self._c = C.c_void_p()
        rc = lib().rtw_create(device, C.byref(self._c))
        if rc != RTW_OK:
            raise RtwError("rtw_create failed (%d): %s" % (rc, lib().rtw_last_error(None).decode()))
        self.Nx = self.Ny = self.Ns = 0
        self.max_depth = 20  # Director.cpp:42
        self.scene = None
        self.frame = None

    def _check(self, rc, what):
        if rc != RTW_OK:
            raise RtwError("%s failed (%d): %s" % (what, rc, lib().rtw_last_error(self._c).decode()))

    # Director::init(w, h, ns)
    def init(self, Nx, Ny, Ns):
        self.Nx, self.Ny, self.Ns = Nx, Ny, Ns

    # Director::createScene(n)
    def createScene(self, scene_number, asset_dir=None, n_stress_prims=0):
        self.scene = Scene(scene_number, self.Nx, self.Ny, asset_dir, n_stress_prims)
        self.upload(self.scene.desc)

    def loadScene(self, path):
        """createScene from a scene file (rtw_scene_load) instead of one of the built-in scene functions."""
        self.scene = Scene.load(path, self.Nx, self.Ny)
        self.upload(self.scene.desc)

    def upload(self, desc_ptr):
        self._check(lib().rtw_upload_scene(self._c, desc_ptr), "rtw_upload_scene")
        self._check(lib().rtw_build_accel(self._c), "rtw_build_accel")

    # Director::renderFrame()
    def renderFrame(self, spp=None, max_depth=None):
        spp = self.Ns if spp is None else spp
        out = np.empty((self.Ny, self.Nx, 4), dtype=np.float32)
        depth = self.max_depth if max_depth is None else max_depth
        self._check(lib().rtw_render(self._c, self.Nx, self.Ny, spp, depth, _ptr(out)),
                    "rtw_render")
        self.frame = out
        return out

    def linear(self):
        out = np.empty((self.Ny, self.Nx, 4), dtype=np.float32)
        self._check(lib().rtw_get_linear(self._c, _ptr(out), out.size), "rtw_get_linear")
        return out

    def render_accum(self, accum_dev_ptr, sample_first, sample_stride, sample_count, max_depth=None):
        depth = self.max_depth if max_depth is None else max_depth
        self._check(lib().rtw_render_accum(self._c, self.Nx, self.Ny, sample_first, sample_stride, sample_count,
                                           depth, C.c_void_p(accum_dev_ptr)), "rtw_render_accum")

    def tonemap(self, accum_dev_ptr, out_dev_ptr, spp_total):
        self._check(lib().rtw_tonemap(self._c, self.Nx, self.Ny, spp_total, C.c_void_p(accum_dev_ptr),
                                      C.c_void_p(out_dev_ptr)), "rtw_tonemap")

    # Director::printPPM()
    def denoise(self, frame, iterations=3, sigma=0.25):
        """The denoise STAND-IN (rtw_denoise): an a-trous filter, not the reference's OptiX AI denoiser."""
        frame = np.ascontiguousarray(frame, np.float32)
        out = np.empty_like(frame)
        self._check(lib().rtw_denoise(self._c, frame.shape[1], frame.shape[0], iterations, sigma, _ptr(frame), _ptr(out)), "rtw_denoise")
        return out

    def denoise_static(self, in_dev_ptr, out_dev_ptr, Nx, Ny, iterations=3, sigma=0.25, cuda_stream=None):
        """rtw_denoise_static: the same filter on caller-owned DEVICE buffers, no allocation, no synchronisation
        (scratch = a static __device__ array in librtw.so, at most DENOISE_STATIC_MAX_PIXELS pixels)."""
        self._check(lib().rtw_denoise_static(self._c, Nx, Ny, iterations, sigma, C.c_void_p(in_dev_ptr), C.c_void_p(out_dev_ptr),
                                             C.c_void_p(cuda_stream or 0)), "rtw_denoise_static")

    def printPPM(self, path="-"):
        self._check(lib().rtw_write_ppm(path.encode(), self.Nx, self.Ny, _ptr(self.frame)), "rtw_write_ppm")

    def set_option(self, name, value):
        self._check(lib().rtw_set_option(self._c, name.encode(), int(value)), "rtw_set_option(%s)" % name)

    def stats(self):
        s = Stats()
        self._check(lib().rtw_get_stats(self._c, C.byref(s)), "rtw_get_stats")
        return {k: getattr(s, k) for k, _ in Stats._fields_}

    # ---- test hooks ----
    def trace_closest(self, rays):
        rays = np.ascontiguousarray(rays, dtype=RAY_DTYPE)
        hits = np.empty(len(rays), dtype=HIT_DTYPE)
        self._check(lib().rtw_trace_closest(self._c, _ptr(rays), len(rays), _ptr(hits)), "rtw_trace_closest")
        return hits

    def trace_any(self, rays):
        rays = np.ascontiguousarray(rays, dtype=RAY_DTYPE)
        occ = np.empty(len(rays), dtype=np.int32)
        self._check(lib().rtw_trace_any(self._c, _ptr(rays), len(rays), _ptr(occ)), "rtw_trace_any")
        return occ

    def rng_stream(self, seed0, seed1, n):
        tea = np.zeros(1, np.uint32)
        lcg = np.zeros(n, np.uint32)
        rnd = np.zeros(n, np.float32)
        xs = np.zeros(n, np.uint32)
        rf = np.zeros(n, np.float32)
        self._check(lib().rtw_rng_stream(self._c, seed0, seed1, n, _ptr(tea), _ptr(lcg), _ptr(rnd), _ptr(xs), _ptr(rf)),
                    "rtw_rng_stream")
        return int(tea[0]), lcg, rnd, xs, rf

    def math_eval(self, fn, x, y=None):
        """Device evaluation of the arithmetic-contract functions (rtw_math_fn numbering)."""
        x = np.ascontiguousarray(x, np.float32)
        out = np.empty_like(x)
        yp = _ptr(np.ascontiguousarray(y, np.float32)) if y is not None else None
        self._check(lib().rtw_math_eval(self._c, fn, _ptr(x), yp, len(x), _ptr(out)), "rtw_math_eval")
        return out

    def camera_rays(self, sample=0):
        rays = np.empty(self.Nx * self.Ny, dtype=RAY_DTYPE)
        self._check(lib().rtw_camera_rays(self._c, self.Nx, self.Ny, sample, _ptr(rays)), "rtw_camera_rays")
        return rays

    def bvh_info(self):
        b = BvhInfo()
        self._check(lib().rtw_bvh_info_get(self._c, C.byref(b)), "rtw_bvh_info_get")
        return b

    def bvh_download(self):
        info = self.bvh_info()
        n, ni = info.n_leaves, info.n_internal
        morton = np.zeros(n, np.uint64)
        leaf_prim = np.zeros(n, np.int32)
        left = np.zeros(ni, np.int32)
        right = np.zeros(ni, np.int32)
        bounds = np.zeros((ni, 12), np.float32)
        self._check(lib().rtw_bvh_download(self._c, _ptr(morton), _ptr(leaf_prim), _ptr(left), _ptr(right), _ptr(bounds)),
                    "rtw_bvh_download")
        return morton, leaf_prim, left, right, bounds

    def sort_pairs(self, keys, values):
        keys = np.ascontiguousarray(keys, dtype=np.uint64).copy()
        values = np.ascontiguousarray(values, dtype=np.uint32).copy()
        self._check(lib().rtw_sort_pairs(self._c, _ptr(keys), _ptr(values), len(keys)), "rtw_sort_pairs")
        return keys, values

    # Director::destroy()
    def destroy(self):
        if self._c:
            lib().rtw_destroy(self._c)
            self._c = C.c_void_p()
        if self.scene is not None:
            self.scene.close()
            self.scene = None

    def __del__(self):
        try:
            self.destroy()
        except Exception:
            pass


class MultiDirector:
    """Several GPUs of one box behind the C-ABI (rtw_create_multi): the same call order as Director."""

    def __init__(self, devices):
        self._m = C.c_void_p()
        arr = (C.c_int * len(devices))(*devices)
        rc = lib().rtw_create_multi(arr, len(devices), C.byref(self._m))
        if rc != RTW_OK:
            raise RtwError("rtw_create_multi(%s) failed: %d (no CPU fallback)" % (list(devices), rc))
        self.devices = list(devices)
        self.Nx = self.Ny = self.Ns = 0
        self.max_depth = 20
        self.scene = None

    def _check(self, rc, what):
        if rc != RTW_OK:
            raise RtwError("%s failed (%d): %s" % (what, rc, lib().rtw_multi_last_error(self._m).decode()))

    def init(self, Nx, Ny, Ns):
        self.Nx, self.Ny, self.Ns = Nx, Ny, Ns

    def set_option(self, name, value):
        self._check(lib().rtw_multi_set_option(self._m, name.encode(), int(value)), "rtw_multi_set_option(%s)" % name)

    def upload(self, desc_ptr):
        self._check(lib().rtw_multi_upload_scene(self._m, desc_ptr), "rtw_multi_upload_scene")
        self._check(lib().rtw_multi_build_accel(self._m), "rtw_multi_build_accel")

    def createScene(self, scene_number, asset_dir=None, n_stress_prims=0):
        self.scene = Scene(scene_number, self.Nx, self.Ny, asset_dir, n_stress_prims)
        self.upload(self.scene.desc)

    def uses_nccl(self):
        return bool(lib().rtw_multi_uses_nccl(self._m))

    def renderFrame(self, spp=None, max_depth=None):
        spp = self.Ns if spp is None else spp
        out = np.empty((self.Ny, self.Nx, 4), dtype=np.float32)
        depth = self.max_depth if max_depth is None else max_depth
        self._check(lib().rtw_multi_render(self._m, self.Nx, self.Ny, spp, depth, _ptr(out)), "rtw_multi_render")
        self._spp = spp
        return out

    def linear(self):
        out = np.empty((self.Ny, self.Nx, 4), dtype=np.float32)
        self._check(lib().rtw_multi_get_linear(self._m, _ptr(out), out.size, self._spp), "rtw_multi_get_linear")
        return out

    def stats(self):
        tot = Stats()
        per = (Stats * len(self.devices))()
        ex = C.c_double(0)
        sh = C.c_int(0)
        self._check(lib().rtw_multi_get_stats(self._m, C.byref(tot), per, C.byref(ex), C.byref(sh)), "rtw_multi_get_stats")
        d = {k: getattr(tot, k) for k, _ in Stats._fields_}
        d["exchange_ms"] = ex.value
        d["shard"] = {1: "samples", 2: "tiles"}.get(sh.value, "?")
        d["per_device"] = [{k: getattr(p, k) for k, _ in Stats._fields_} for p in per]
        return d

    def destroy(self):
        if self._m:
            lib().rtw_multi_destroy(self._m)
            self._m = C.c_void_p()

    def __del__(self):
        try:
            self.destroy()
        except Exception:
            pass


def device_count():
    n = C.c_int(0)
    rc = lib().rtw_device_count(C.byref(n))
    return n.value if rc == RTW_OK else 0
