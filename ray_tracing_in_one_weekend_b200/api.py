"""ctypes binding of librtgpu.so (include/rtgpu.h) — the only way Python reaches the backend.

There is deliberately no fallback: if the CUDA library is missing or no CUDA device is usable,
calls raise `RtError`.  The oracle under oracle/ is test infrastructure and is never imported here.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(HERE)
LIB_PATH = os.environ.get("RTGPU_LIB") or os.path.join(HERE, "librtgpu.so")  # RTGPU_LIB: development override
TEST_LIB_PATH = os.path.join(HERE, "librtgpu_test.so")  # product objects + rt_test_* hooks (tests only)
RT_ERR_ARG, RT_ERR_CUDA, RT_ERR_IO, RT_ERR_UNSUPPORTED, RT_ERR_NCCL = -1, -2, -3, -4, -5
COMM_UNIQUE_ID_BYTES = 128
STATIC_MAX_RAYS = 65536
ASSET_DIR = os.path.join(REPO, "assets")

RT_ABI_VERSION = 2

# RtPrimKind
PRIM_SPHERE, PRIM_XYRECT, PRIM_XZRECT, PRIM_YZRECT, PRIM_MOVING_SPHERE, PRIM_MEDIUM = range(6)
PRIM_FLAG_BOUNDARY_ONLY = 1
XF_TRANSLATE, XF_ROTATE_Y = 0, 1
(MAT_EMISSION, MAT_DIFFUSE, MAT_LAMBERT, MAT_METAL, MAT_DIELECTRIC, MAT_ISOTROPIC, MAT_OREN_NAYAR, MAT_BURLEY_DIFFUSE,
 MAT_ROUGH_PLASTIC, MAT_DISNEY_DIFFUSE, MAT_DISNEY_METAL, MAT_DISNEY_SHEEN, MAT_DISNEY_CLEARCOAT) = range(13)
TEX_CONSTANT, TEX_CHECKER, TEX_PERLIN, TEX_IMAGE = range(4)
SKY_GRADIENT, SKY_BLACK, SKY_ENVMAP, SKY_CONSTANT = range(4)
MODE_AUTO, MODE_MEGAKERNEL, MODE_CTA_WAVEFRONT = 0, 1, 3  # (2 was the global-memory wavefront renderer of ABI 1)
RENDER_COUNTERS = 1
RENDER_RUSSIAN_ROULETTE = 2
RENDER_COSINE_SAMPLING = 4
RENDER_LOBE_SAMPLING = 8
SCENE_GPU_BVH = 1


class RtError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"rtgpu error {code}: {msg}")
        self.code = code


class RtPrim(C.Structure):
    _fields_ = [("kind", C.c_uint32), ("mat", C.c_uint32), ("xform", C.c_uint32), ("flags", C.c_uint32),
                ("object", C.c_uint32), ("aux0", C.c_uint32), ("aux1", C.c_uint32), ("reserved", C.c_uint32),
                ("g", C.c_float * 12)]


class RtXformOp(C.Structure):
    _fields_ = [("kind", C.c_uint32), ("v", C.c_float * 3)]


class RtXformChain(C.Structure):
    _fields_ = [("first_op", C.c_uint32), ("n_ops", C.c_uint32)]


class RtMaterial(C.Structure):
    _fields_ = [("kind", C.c_uint32), ("tex0", C.c_uint32), ("tex1", C.c_uint32), ("reserved", C.c_uint32),
                ("p", C.c_float * 4)]


class RtTexture(C.Structure):
    _fields_ = [("kind", C.c_uint32), ("a", C.c_uint32), ("b", C.c_uint32), ("reserved", C.c_uint32),
                ("p", C.c_float * 4)]


class RtPerlin(C.Structure):
    _fields_ = [("rand_vec", (C.c_float * 4) * 256), ("perm_x", C.c_uint8 * 256), ("perm_y", C.c_uint8 * 256),
                ("perm_z", C.c_uint8 * 256)]


class RtImage(C.Structure):
    _fields_ = [("width", C.c_uint32), ("height", C.c_uint32), ("rgb8", C.POINTER(C.c_uint8))]


class RtSky(C.Structure):
    _fields_ = [("kind", C.c_uint32), ("image", C.c_uint32), ("color", C.c_float * 3)]


class RtCamera(C.Structure):
    _fields_ = [("origin", C.c_float * 3), ("horizontal", C.c_float * 3), ("vertical", C.c_float * 3),
                ("lower_left_corner", C.c_float * 3), ("u", C.c_float * 3), ("v", C.c_float * 3),
                ("lens_radius", C.c_float), ("time0", C.c_float), ("time1", C.c_float)]


class RtFlatScene(C.Structure):
    _fields_ = [("abi_version", C.c_uint32), ("n_prims", C.c_uint32), ("n_xform_ops", C.c_uint32),
                ("n_xform_chains", C.c_uint32), ("n_materials", C.c_uint32), ("n_textures", C.c_uint32),
                ("n_perlin", C.c_uint32), ("n_images", C.c_uint32),
                ("prims", C.POINTER(RtPrim)), ("xform_ops", C.POINTER(RtXformOp)),
                ("xform_chains", C.POINTER(RtXformChain)), ("materials", C.POINTER(RtMaterial)),
                ("textures", C.POINTER(RtTexture)), ("perlin", C.POINTER(RtPerlin)), ("images", C.POINTER(RtImage)),
                ("sky", RtSky), ("camera", RtCamera)]


class RtBvhInfo(C.Structure):
    _fields_ = [("n_nodes", C.c_uint32), ("n_world_prims", C.c_uint32), ("max_depth", C.c_uint32),
                ("max_leaf_prims", C.c_uint32), ("sah_cost", C.c_float), ("node_bytes", C.c_uint32),
                ("builder", C.c_uint32), ("build_ms", C.c_float), ("upload_bytes", C.c_uint64)]


class RtHit(C.Structure):
    _fields_ = [("prim", C.c_int32), ("t", C.c_float), ("p", C.c_float * 3), ("n", C.c_float * 3),
                ("uv", C.c_float * 2), ("front_face", C.c_uint32)]


HIT_DTYPE = np.dtype([("prim", "<i4"), ("t", "<f4"), ("p", "<f4", 3), ("n", "<f4", 3), ("uv", "<f4", 2),
                      ("front_face", "<u4")])
assert HIT_DTYPE.itemsize == C.sizeof(RtHit) == 44


class RtRenderParams(C.Structure):
    _fields_ = [("width", C.c_uint32), ("height", C.c_uint32), ("spp_total", C.c_uint32),
                ("sample_begin", C.c_uint32), ("sample_end", C.c_uint32), ("max_depth", C.c_uint32),
                ("mode", C.c_uint32), ("flags", C.c_uint32), ("seed", C.c_uint64)]


class RtStats(C.Structure):
    _fields_ = [("paths", C.c_uint64), ("rays", C.c_uint64), ("node_visits", C.c_uint64), ("box_tests", C.c_uint64),
                ("prim_tests", C.c_uint64), ("shade_calls", C.c_uint64), ("ms_kernel", C.c_double),
                ("ms_total", C.c_double), ("launches", C.c_uint32), ("mode_used", C.c_uint32),
                ("class_counts", C.c_uint64 * 8)]

    def as_dict(self) -> dict:
        return {n: (list(getattr(self, n)) if n == "class_counts" else getattr(self, n)) for n, _ in self._fields_}


# every symbol include/rtgpu.h declares
EXPORTED_SYMBOLS = [
    "rt_scene_create", "rt_scene_create_ex", "rt_scene_destroy", "rt_scene_bvh_info", "rt_closest_hit", "rt_closest_hit_device",
    "rt_camera_get_rays", "rt_render_device", "rt_render", "rt_resolve_rgb8", "rt_write_png", "rt_measure_fp32_peak",
    "rt_flat_shade_queues", "rt_last_error", "rt_host_scene_build", "rt_host_scene_flat", "rt_host_scene_destroy",
    "rt_flat_validate", "rt_release_cached_memory", "rt_partition_samples", "rt_comm_unique_id", "rt_comm_create",
    "rt_comm_create_all", "rt_comm_info", "rt_comm_destroy", "rt_render_reduce", "rt_render_reduce_device",
    "rt_scene_kernel_times", "rt_closest_hit_static",
]

_lib: Optional[C.CDLL] = None
_test_lib: Optional[C.CDLL] = None


def _declare(lib: C.CDLL) -> None:
    vp, u32, sz = C.c_void_p, C.c_uint32, C.c_size_t
    lib.rt_last_error.restype = C.c_char_p
    lib.rt_scene_create.argtypes = [C.POINTER(RtFlatScene), C.c_int, C.POINTER(vp)]
    lib.rt_scene_create_ex.argtypes = [C.POINTER(RtFlatScene), C.c_int, C.c_uint32, C.POINTER(vp)]
    lib.rt_scene_destroy.argtypes = [vp]
    lib.rt_scene_destroy.restype = None
    lib.rt_scene_bvh_info.argtypes = [vp, C.POINTER(RtBvhInfo)]
    lib.rt_scene_kernel_times.argtypes = [vp, C.POINTER(C.c_double), C.c_int, C.POINTER(C.c_int)]
    lib.rt_flat_shade_queues.argtypes = [vp, C.POINTER(C.c_uint8), C.POINTER(C.c_uint32), C.POINTER(C.c_uint32)]
    lib.rt_flat_validate.argtypes = [C.POINTER(RtFlatScene)]
    lib.rt_release_cached_memory.argtypes = []
    lib.rt_release_cached_memory.restype = None
    lib.rt_closest_hit.argtypes = [vp, vp, vp, vp, sz, C.c_float, C.c_float, vp]
    lib.rt_closest_hit_static.argtypes = [vp, vp, vp, vp, sz, C.c_float, C.c_float, vp, C.POINTER(C.c_uint64)]
    lib.rt_closest_hit_device.argtypes = [vp, vp, vp, vp, sz, C.c_float, C.c_float, vp, vp]
    lib.rt_camera_get_rays.argtypes = [C.POINTER(RtCamera), vp, sz, vp, vp]
    lib.rt_render_device.argtypes = [vp, C.POINTER(RtRenderParams), vp, vp, C.POINTER(RtStats)]
    lib.rt_render.argtypes = [vp, C.POINTER(RtRenderParams), vp, C.POINTER(RtStats)]
    lib.rt_resolve_rgb8.argtypes = [vp, vp, u32, u32, vp]
    lib.rt_write_png.argtypes = [C.c_char_p, vp, u32, u32]
    lib.rt_measure_fp32_peak.argtypes = [C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_double)]
    lib.rt_host_scene_build.argtypes = [C.c_char_p, C.c_float, C.c_char_p, C.POINTER(vp)]
    lib.rt_host_scene_flat.argtypes = [vp]
    lib.rt_host_scene_flat.restype = C.POINTER(RtFlatScene)
    lib.rt_host_scene_destroy.argtypes = [vp]
    lib.rt_host_scene_destroy.restype = None
    lib.rt_partition_samples.argtypes = [u32, u32, C.c_int, C.c_int, C.POINTER(u32), C.POINTER(u32)]
    lib.rt_comm_unique_id.argtypes = [vp]
    lib.rt_comm_create.argtypes = [vp, C.c_int, C.c_int, C.c_int, C.POINTER(vp)]
    lib.rt_comm_create_all.argtypes = [C.POINTER(C.c_int), C.c_int, C.POINTER(vp)]
    lib.rt_comm_info.argtypes = [vp, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int)]
    lib.rt_comm_destroy.argtypes = [vp]
    lib.rt_comm_destroy.restype = None
    lib.rt_render_reduce.argtypes = [C.POINTER(vp), vp, C.POINTER(RtRenderParams), vp, C.POINTER(RtStats)]
    lib.rt_render_reduce_device.argtypes = [C.POINTER(vp), vp, C.POINTER(RtRenderParams), C.POINTER(vp), C.POINTER(vp)]


def load_library() -> C.CDLL:
    """dlopen librtgpu.so (built in-tree by ray_tracing_in_one_weekend_b200.build). Fails loudly."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RtError(-2, f"{LIB_PATH} is missing — run `python -m ray_tracing_in_one_weekend_b200.build` "
                          "(there is no CPU fallback)")
    lib = C.CDLL(LIB_PATH)
    _declare(lib)
    _lib = lib
    return lib


def load_test_library() -> C.CDLL:
    """dlopen librtgpu_test.so: the SAME objects as librtgpu.so plus the rt_test_* parity hooks (csrc/test_hooks.cu).
    A separate library instance: scene handles made through it must be used through it (`Scene(..., lib=...)`)."""
    global _test_lib
    if _test_lib is not None:
        return _test_lib
    if not os.path.exists(TEST_LIB_PATH):
        raise RtError(-2, f"{TEST_LIB_PATH} is missing — run `python -m ray_tracing_in_one_weekend_b200.build`")
    lib = C.CDLL(TEST_LIB_PATH)
    _declare(lib)
    vp, u32 = C.c_void_p, C.c_uint32
    lib.rt_test_shade_cases.argtypes = [vp, vp, vp, vp, vp, u32, u32, vp]
    lib.rt_test_texture_cases.argtypes = [vp, u32, vp, vp, u32, vp]
    lib.rt_test_medium_cases.argtypes = [vp, u32, vp, vp, vp, u32, vp]
    lib.rt_test_closest_hit_xi.argtypes = [vp, vp, vp, vp, u32, C.c_float, C.c_float, vp]
    lib.rt_test_bvh_validate.argtypes = [C.POINTER(RtFlatScene), u32, C.POINTER(RtBvhInfo)]
    lib.rt_test_div3_sweep.argtypes = [vp, u32, u32, u32, C.POINTER(C.c_uint64)]
    _test_lib = lib
    return lib


def _check(rc: int, lib: Optional[C.CDLL] = None) -> None:
    if rc != 0:
        raise RtError(rc, (lib or load_library()).rt_last_error().decode("utf-8", "replace"))


def _ptr(a: Optional[np.ndarray]):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


class HostScene:
    """A scene built by the host API (csrc/scene_api.hpp) and flattened to an RtFlatScene.

    `name` is one of the reference's scene functions (src/demo_scene.rs: test_sphere, sphere_scene,
    simple_light_scene, cornell_box, final_scene) or a BASELINE.json config scene (textured,
    cornell_box_solid, pbr_ibl, cover_defocus, cover_100k_motion, cover_4k_motion, furnace_<material>).
    """

    def __init__(self, name: str, aspect_ratio: float, asset_dir: str = ASSET_DIR):
        self._lib = load_library()
        h = C.c_void_p()
        _check(self._lib.rt_host_scene_build(name.encode(), C.c_float(aspect_ratio), asset_dir.encode(), C.byref(h)))
        self._h = h
        self.name = name
        self.aspect_ratio = aspect_ratio

    @property
    def flat(self) -> "C.POINTER(RtFlatScene)":
        return self._lib.rt_host_scene_flat(self._h)

    def shade_queues(self):
        """(queue of every material, class of every queue, drain order): the host-side planning step of the
        CTA-wavefront renderer's material-sorted shade pass (rt_flat_shade_queues; no device needed)."""
        flat = self.flat
        n = flat.contents.n_materials
        mq = (C.c_uint8 * max(n, 1))()
        qc, qo = C.c_uint32(), C.c_uint32()
        _check(self._lib.rt_flat_shade_queues(flat, mq, C.byref(qc), C.byref(qo)))
        return (list(mq)[:n], [(qc.value >> (4 * q)) & 15 for q in range(7)], [(qo.value >> (4 * q)) & 15 for q in range(7)])

    def close(self) -> None:
        if getattr(self, "_h", None):
            self._lib.rt_host_scene_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def flat_to_numpy(flat_ptr) -> dict:
    """Copy the POD arrays of an RtFlatScene into numpy structured arrays (for comparisons)."""
    f = flat_ptr.contents

    def arr(ptr, n, ctype):
        if n == 0:
            return np.zeros(0, dtype=np.uint8)
        buf = C.string_at(ptr, n * C.sizeof(ctype))
        return np.frombuffer(buf, dtype=np.uint8).copy()

    out = {
        "prims": arr(f.prims, f.n_prims, RtPrim), "xform_ops": arr(f.xform_ops, f.n_xform_ops, RtXformOp),
        "xform_chains": arr(f.xform_chains, f.n_xform_chains, RtXformChain),
        "materials": arr(f.materials, f.n_materials, RtMaterial), "textures": arr(f.textures, f.n_textures, RtTexture),
        "perlin": arr(f.perlin, f.n_perlin, RtPerlin),
        "sky": np.frombuffer(bytes(f.sky), dtype=np.uint8).copy(),
        "camera": np.frombuffer(bytes(f.camera), dtype=np.uint8).copy(),
    }
    imgs = []
    for i in range(f.n_images):
        im = f.images[i]
        imgs.append(np.frombuffer(C.string_at(im.rgb8, im.width * im.height * 3), dtype=np.uint8)
                    .reshape(im.height, im.width, 3).copy())
    out["images"] = imgs
    return out


def camera_get_rays(cam: RtCamera, uv: np.ndarray):
    """Host-side Camera::get_ray (camera.rs:40-46) for an (n,2) float32 array of (u,v)."""
    lib = load_library()
    uv = np.ascontiguousarray(uv, dtype=np.float32)
    n = uv.shape[0]
    o = np.empty((n, 3), np.float32)
    d = np.empty((n, 3), np.float32)
    _check(lib.rt_camera_get_rays(C.byref(cam), _ptr(uv), n, _ptr(o), _ptr(d)))
    return o, d


def pixel_center_uv(width: int, height: int) -> np.ndarray:
    """u=(i+0.5)/nx, v=(j+0.5)/ny for every pixel, j=0 bottom, row-major (j*width+i), float32 math."""
    i = (np.arange(width, dtype=np.float32) + np.float32(0.5)) / np.float32(width)
    j = (np.arange(height, dtype=np.float32) + np.float32(0.5)) / np.float32(height)
    uv = np.empty((height, width, 2), np.float32)
    uv[..., 0] = i[None, :]
    uv[..., 1] = j[:, None]
    return uv.reshape(-1, 2)


class Scene:
    """Device scene handle (rt_scene_create). Requires a CUDA device."""

    def __init__(self, flat_ptr, device: int = 0, flags: Optional[int] = None, lib: Optional[C.CDLL] = None):
        self._lib = lib or load_library()
        h = C.c_void_p()
        if flags is None:
            _check(self._lib.rt_scene_create(flat_ptr, device, C.byref(h)), self._lib)
        else:
            _check(self._lib.rt_scene_create_ex(flat_ptr, device, flags, C.byref(h)), self._lib)
        self._h = h
        self.device = device
        self.camera = RtCamera.from_buffer_copy(bytes(flat_ptr.contents.camera))

    @classmethod
    def from_name(cls, name: str, aspect_ratio: float, device: int = 0, flags: Optional[int] = None, lib=None) -> "Scene":
        hs = HostScene(name, aspect_ratio)
        try:
            return cls(hs.flat, device, flags, lib)
        finally:
            hs.close()

    def close(self) -> None:
        if getattr(self, "_h", None):
            self._lib.rt_scene_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def bvh_info(self) -> RtBvhInfo:
        info = RtBvhInfo()
        _check(self._lib.rt_scene_bvh_info(self._h, C.byref(info)), self._lib)
        return info

    def kernel_times(self, n: int = 32):
        """CUDA-event durations (ms) of the last n <= 32 render kernels on this scene, oldest first (waits for them)."""
        buf = (C.c_double * max(n, 1))()
        got = C.c_int()
        _check(self._lib.rt_scene_kernel_times(self._h, buf, n, C.byref(got)), self._lib)
        return [buf[i] for i in range(got.value)]

    def closest_hit(self, o: np.ndarray, d: np.ndarray, t_min: float = 1e-3, t_max: float = 3.4028234663852886e38,
                    times: Optional[np.ndarray] = None) -> np.ndarray:
        """world.hit over a batch of rays in HOST memory (main.rs:44). Returns a HIT_DTYPE array."""
        o = np.ascontiguousarray(o, dtype=np.float32)
        d = np.ascontiguousarray(d, dtype=np.float32)
        n = o.shape[0]
        if times is not None:
            times = np.ascontiguousarray(times, dtype=np.float32)
        out = np.zeros(n, dtype=HIT_DTYPE)
        _check(self._lib.rt_closest_hit(self._h, _ptr(o), _ptr(d), _ptr(times), n, C.c_float(t_min), C.c_float(t_max),
                                        _ptr(out)), self._lib)
        return out

    def closest_hit_static(self, o: np.ndarray, d: np.ndarray, t_min: float = 1e-3, t_max: float = 3.4028234663852886e38,
                           times: Optional[np.ndarray] = None):
        """rt_closest_hit_static: the same query with its device staging in the library's static array (no
        allocation; at most STATIC_MAX_RAYS rays).  Returns (HIT_DTYPE array, number of rays that hit)."""
        o = np.ascontiguousarray(o, dtype=np.float32)
        d = np.ascontiguousarray(d, dtype=np.float32)
        n = o.shape[0]
        if times is not None:
            times = np.ascontiguousarray(times, dtype=np.float32)
        out = np.zeros(n, dtype=HIT_DTYPE)
        n_hits = C.c_uint64()
        _check(self._lib.rt_closest_hit_static(self._h, _ptr(o), _ptr(d), _ptr(times), n, C.c_float(t_min), C.c_float(t_max),
                                               _ptr(out), C.byref(n_hits)), self._lib)
        return out, n_hits.value

    def make_params(self, width: int, height: int, spp: int, *, sample_begin: int = 0,
                    sample_end: Optional[int] = None, spp_total: Optional[int] = None, max_depth: int = 50,
                    seed: int = 95, mode: int = MODE_AUTO, flags: int = 0) -> RtRenderParams:
        se = spp if sample_end is None else sample_end
        return RtRenderParams(width, height, spp if spp_total is None else spp_total, sample_begin, se, max_depth,
                              mode, flags, seed)

    def render(self, width: int, height: int, spp: int, **kw):
        """Host-buffer render: returns (linear mean image (H,W,3) float32 with row 0 = bottom, RtStats)."""
        p = self.make_params(width, height, spp, **kw)
        out = np.empty((height, width, 3), np.float32)
        st = RtStats()
        _check(self._lib.rt_render(self._h, C.byref(p), _ptr(out), C.byref(st)), self._lib)
        return out, st

    def render_device(self, params: RtRenderParams, accum_ptr: int, stream_ptr: int = 0, want_stats: bool = True):
        """Adds radiance sums of the sample range to a DEVICE buffer (e.g. a torch tensor's data_ptr())."""
        st = RtStats() if want_stats else None
        _check(self._lib.rt_render_device(self._h, C.byref(params), C.c_void_p(accum_ptr),
                                          C.c_void_p(stream_ptr) if stream_ptr else None,
                                          C.byref(st) if st is not None else None), self._lib)
        return st

    @staticmethod
    def write_png(path: str, rgb8: np.ndarray) -> None:
        """image::save_buffer (main.rs:128): (H,W,3) uint8, row 0 = top -> PNG file."""
        rgb8 = np.ascontiguousarray(rgb8, dtype=np.uint8)
        h, w = rgb8.shape[:2]
        _check(load_library().rt_write_png(os.fsencode(path), _ptr(rgb8), w, h))

    def resolve_rgb8(self, mean: np.ndarray) -> np.ndarray:
        """powf(0.5), *255.99 as u8, vertical flip (main.rs:98-105,127). Returns (H,W,3) uint8, row 0 = top."""
        mean = np.ascontiguousarray(mean, dtype=np.float32)
        h, w = mean.shape[:2]
        out = np.empty((h, w, 3), np.uint8)
        _check(self._lib.rt_resolve_rgb8(self._h, _ptr(mean), w, h, _ptr(out)), self._lib)
        return out


def measure_fp32_peak(device: int = 0):
    lib = load_library()
    a, b = C.c_double(), C.c_double()
    _check(lib.rt_measure_fp32_peak(device, C.byref(a), C.byref(b)))
    return a.value, b.value


def release_cached_memory() -> None:
    load_library().rt_release_cached_memory()


def partition_samples(sample_begin: int, sample_end: int, n_ranks: int, rank: int):
    """rt_partition_samples: global sample indices [begin, end) of `rank` (host arithmetic, no device needed)."""
    b, e = C.c_uint32(), C.c_uint32()
    _check(load_library().rt_partition_samples(sample_begin, sample_end, n_ranks, rank, C.byref(b), C.byref(e)))
    return b.value, e.value


def flat_validate(flat_ptr) -> None:
    """rt_flat_validate: raises RtError(RT_ERR_ARG | RT_ERR_UNSUPPORTED) exactly as rt_scene_create would."""
    _check(load_library().rt_flat_validate(flat_ptr))


class Comm:
    """RtComm: the NCCL communicator(s) of the spp-partitioned multi-GPU frame (rt_comm_* / rt_render_reduce*)."""

    def __init__(self, handle, lib):
        self._h, self._lib = handle, lib

    @staticmethod
    def unique_id() -> bytes:
        buf = (C.c_uint8 * COMM_UNIQUE_ID_BYTES)()
        _check(load_library().rt_comm_unique_id(buf))
        return bytes(buf)

    @classmethod
    def from_id(cls, uid: bytes, n_ranks: int, rank: int, device: int) -> "Comm":
        """One process per GPU (collective: every rank calls it with rank 0's id)."""
        lib = load_library()
        if len(uid) != COMM_UNIQUE_ID_BYTES:
            raise ValueError("unique id must be 128 bytes")
        buf = (C.c_uint8 * COMM_UNIQUE_ID_BYTES).from_buffer_copy(uid)
        h = C.c_void_p()
        _check(lib.rt_comm_create(buf, n_ranks, rank, device, C.byref(h)))
        return cls(h, lib)

    @classmethod
    def all_local(cls, devices) -> "Comm":
        """One process that drives several devices (ncclCommInitAll)."""
        lib = load_library()
        arr = (C.c_int * len(devices))(*devices)
        h = C.c_void_p()
        _check(lib.rt_comm_create_all(arr, len(devices), C.byref(h)))
        return cls(h, lib)

    def info(self):
        a, b, c = C.c_int(), C.c_int(), C.c_int()
        _check(self._lib.rt_comm_info(self._h, C.byref(a), C.byref(b), C.byref(c)))
        return {"n_ranks": a.value, "first_rank": b.value, "n_local": c.value}

    def render_reduce(self, scenes, params: RtRenderParams, out: Optional[np.ndarray] = None, want_stats: bool = True):
        """rt_render_reduce: `params` describes the WHOLE frame; returns RtStats of this process's ranks.  `out` is
        the (H,W,3) float32 host image (required on the process that owns rank 0)."""
        arr = (C.c_void_p * len(scenes))(*[s._h for s in scenes])
        st = RtStats() if want_stats else None
        _check(self._lib.rt_render_reduce(arr, self._h, C.byref(params), _ptr(out), C.byref(st) if st is not None else None))
        return st

    def render_reduce_device(self, scenes, params: RtRenderParams, accum_ptrs, stream_ptrs=None) -> None:
        """rt_render_reduce_device: asynchronous on the given streams; rank 0's buffer holds the mean afterwards."""
        n = len(scenes)
        arr = (C.c_void_p * n)(*[s._h for s in scenes])
        acc = (C.c_void_p * n)(*accum_ptrs)
        sts = (C.c_void_p * n)(*stream_ptrs) if stream_ptrs is not None else None
        _check(self._lib.rt_render_reduce_device(arr, self._h, C.byref(params), acc, sts))

    def close(self) -> None:
        if getattr(self, "_h", None):
            self._lib.rt_comm_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
