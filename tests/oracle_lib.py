"""ctypes wrapper around oracle/liboracle.so — TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import
this module.  The oracle is the CPU restatement of the reference (see oracle/vec.hpp header);
it is the checker, never the thing measured or shipped.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(REPO, "oracle")
ORACLE_LIB = os.path.join(ORACLE_DIR, "liboracle.so")
ASSET_DIR = os.path.join(REPO, "assets")

from ray_tracing_in_one_weekend_b200.api import HIT_DTYPE, RtCamera, RtFlatScene  # noqa: E402  (POD layouts only)


class OrcStats(C.Structure):
    _fields_ = [("paths", C.c_uint64), ("rays", C.c_uint64), ("box_tests", C.c_uint64), ("prim_tests", C.c_uint64),
                ("shade_calls", C.c_uint64), ("unit_dir_violations", C.c_uint64), ("seconds", C.c_double),
                ("threads", C.c_uint32)]

    def as_dict(self):
        return {n: getattr(self, n) for n, _ in self._fields_}


def build_oracle(force: bool = False) -> str:
    srcs = [os.path.join(ORACLE_DIR, f) for f in ("oracle_capi.cpp", "scene.hpp", "rt.hpp", "vec.hpp")]
    srcs.append(os.path.join(REPO, "include", "rtgpu.h"))
    stale = force or not os.path.exists(ORACLE_LIB) or any(
        os.path.getmtime(s) > os.path.getmtime(ORACLE_LIB) for s in srcs)
    if stale:
        subprocess.run(["make", "-C", ORACLE_DIR, "-B", "liboracle.so"], check=True, capture_output=True)
    return ORACLE_LIB


_lib = None


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        L = C.CDLL(build_oracle())
        vp = C.c_void_p
        L.orc_last_error.restype = C.c_char_p
        L.orc_scene_builtin.restype = vp
        L.orc_scene_builtin.argtypes = [C.c_char_p, C.c_float, C.c_char_p]
        L.orc_scene_from_flat.restype = vp
        L.orc_scene_from_flat.argtypes = [C.POINTER(RtFlatScene)]
        L.orc_scene_flat.restype = C.POINTER(RtFlatScene)
        L.orc_scene_flat.argtypes = [vp]
        L.orc_scene_free.argtypes = [vp]
        L.orc_scene_free.restype = None
        L.orc_closest_hit.argtypes = [vp, vp, vp, vp, C.c_size_t, C.c_float, C.c_float, C.c_int, vp]
        L.orc_render.argtypes = [vp, C.c_uint32, C.c_uint32, C.c_uint32, C.c_int, C.c_int, C.c_uint64, C.c_uint32,
                                 C.c_uint32, C.c_uint32, vp, C.POINTER(OrcStats)]
        L.orc_resolve_rgb8.argtypes = [vp, C.c_uint32, C.c_uint32, vp]
        L.orc_resolve_rgb8.restype = None
        L.orc_scatter.argtypes = [vp, C.c_uint32, vp, vp, vp, vp, C.c_size_t, vp, C.POINTER(C.c_uint32)]
        L.orc_texture_value.argtypes = [vp, C.c_uint32, vp, vp, vp]
        L.orc_camera_new.argtypes = [vp, vp, vp, C.c_float, C.c_float, C.POINTER(RtCamera)]
        L.orc_camera_new.restype = None
        L.orc_camera_get_rays.argtypes = [C.POINTER(RtCamera), vp, C.c_size_t, vp, vp]
        L.orc_camera_get_rays.restype = None
        for name in ("orc_reflectance", "orc_fresnel_dielectric_2", "orc_gtr1", "orc_gtr2", "orc_smith_geo_ggx"):
            getattr(L, name).restype = C.c_float
            getattr(L, name).argtypes = [C.c_float, C.c_float]
        L.orc_schlick_fresnel.restype = C.c_float
        L.orc_schlick_fresnel.argtypes = [C.c_float]
        L.orc_refract.argtypes = [vp, vp, C.c_float, vp]
        L.orc_aabb_hit.argtypes = [vp, vp, vp, vp, C.c_float, C.c_float]
        L.orc_sky.argtypes = [C.c_uint32, vp, vp]
        L.orc_scene_sky.argtypes = [vp, vp, vp]
        L.orc_perlin_table.argtypes = [C.c_uint64, C.c_uint32, vp, C.c_uint32, vp, vp, vp]
        L.orc_perlin_table.restype = None
        L.orc_closest_hit_xi.argtypes = [vp, vp, vp, vp, C.c_size_t, C.c_float, C.c_float, vp]
        L.orc_medium_cases.argtypes = [vp, C.c_uint32, vp, vp, vp, C.c_size_t, vp]
        _lib = L
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def f3(x, y, z):
    return np.array([x, y, z], dtype=np.float32)


class OracleScene:
    def __init__(self, handle):
        if not handle:
            raise RuntimeError("oracle: " + lib().orc_last_error().decode())
        self.h = C.c_void_p(handle)

    @classmethod
    def builtin(cls, name: str, aspect: float, asset_dir: str = ASSET_DIR) -> "OracleScene":
        """The oracle's own restatement of src/demo_scene.rs: nested graph with the reference's BvhNode."""
        return cls(lib().orc_scene_builtin(name.encode(), C.c_float(aspect), asset_dir.encode()))

    @classmethod
    def from_flat(cls, flat_ptr) -> "OracleScene":
        """List-mode world (HitableList::hit over every primitive) rebuilt from an RtFlatScene."""
        return cls(lib().orc_scene_from_flat(flat_ptr))

    @property
    def flat(self):
        return lib().orc_scene_flat(self.h)

    def close(self):
        if self.h:
            lib().orc_scene_free(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def closest_hit(self, o, d, t_min=1e-3, t_max=3.4028234663852886e38, times=None, skip_media=True):
        o = np.ascontiguousarray(o, np.float32)
        d = np.ascontiguousarray(d, np.float32)
        n = o.shape[0]
        if times is not None:
            times = np.ascontiguousarray(times, np.float32)
        out = np.zeros(n, dtype=HIT_DTYPE)
        rc = lib().orc_closest_hit(self.h, _p(o), _p(d), _p(times), n, C.c_float(t_min), C.c_float(t_max),
                                   1 if skip_media else 0, _p(out))
        assert rc == 0
        return out

    def render(self, width, height, spp, max_depth=50, threads=0, seed_base=95, column_begin=0, column_end=None,
               column_stride=1):
        out = np.zeros((height, width, 3), np.float32)
        st = OrcStats()
        ce = width if column_end is None else column_end
        rc = lib().orc_render(self.h, width, height, spp, max_depth, threads, seed_base, column_begin, ce, column_stride,
                              _p(out), C.byref(st))
        assert rc == 0
        return out, st

    def scatter(self, mat_id, ray_o, ray_d, rec13, xi):
        ray_o = np.ascontiguousarray(ray_o, np.float32)
        ray_d = np.ascontiguousarray(ray_d, np.float32)
        rec13 = np.ascontiguousarray(rec13, np.float32)
        xi = np.ascontiguousarray(xi, np.float32)
        out = np.zeros(12, np.float32)
        used = C.c_uint32()
        rc = lib().orc_scatter(self.h, mat_id, _p(ray_o), _p(ray_d), _p(rec13), _p(xi), xi.size, _p(out), C.byref(used))
        assert rc >= 0
        return rc, out, used.value

    def sky(self, d):
        d = np.ascontiguousarray(d, np.float32)
        out = np.zeros(3, np.float32)
        assert lib().orc_scene_sky(self.h, _p(d), _p(out)) == 0
        return out

    def closest_hit_xi(self, o, d, xi, t_min=1e-3, t_max=3.4028234663852886e38):
        """world.hit with the media included, every random draw of ray i = xi[i]."""
        o = np.ascontiguousarray(o, np.float32)
        d = np.ascontiguousarray(d, np.float32)
        xi = np.ascontiguousarray(xi, np.float32)
        out = np.zeros(o.shape[0], dtype=HIT_DTYPE)
        assert lib().orc_closest_hit_xi(self.h, _p(o), _p(d), _p(xi), o.shape[0], C.c_float(t_min), C.c_float(t_max), _p(out)) == 0
        return out

    def medium_cases(self, flat_prim, rays, tmm, xi):
        rays = np.ascontiguousarray(rays, np.float32)
        tmm = np.ascontiguousarray(tmm, np.float32)
        xi = np.ascontiguousarray(xi, np.float32)
        out = np.zeros((rays.shape[0], 11), np.float32)
        rc = lib().orc_medium_cases(self.h, flat_prim, _p(rays), _p(tmm), _p(xi), rays.shape[0], _p(out))
        assert rc == 0, rc
        return out

    def texture_value(self, tex_id, uv, p):
        uv = np.ascontiguousarray(uv, np.float32)
        p = np.ascontiguousarray(p, np.float32)
        out = np.zeros(3, np.float32)
        rc = lib().orc_texture_value(self.h, tex_id, _p(uv), _p(p), _p(out))
        assert rc == 0
        return out


def resolve_rgb8(mean: np.ndarray) -> np.ndarray:
    mean = np.ascontiguousarray(mean, np.float32)
    h, w = mean.shape[:2]
    out = np.zeros((h, w, 3), np.uint8)
    lib().orc_resolve_rgb8(_p(mean), w, h, _p(out))
    return out


def camera_get_rays(cam: RtCamera, uv: np.ndarray):
    uv = np.ascontiguousarray(uv, np.float32)
    n = uv.shape[0]
    o = np.zeros((n, 3), np.float32)
    d = np.zeros((n, 3), np.float32)
    lib().orc_camera_get_rays(C.byref(cam), _p(uv), n, _p(o), _p(d))
    return o, d
