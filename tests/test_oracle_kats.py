"""Known-answer tests K1-K12 pinning the CPU oracle — no GPU needed.

The reference ships no tests or golden vectors; these values were derived by hand from the cited
reference lines (float32 arithmetic).  Parity with a real Rust build stays UNPINNED at the glam / rand /
libm boundaries (oracle/vec.hpp header); K12 is xoshiro256++'s published reference vector.
"""
import ctypes as C

import numpy as np
import pytest


_KEEP = []


def f3(*v):
    """float32[3]; kept alive for the whole test session so `.ctypes.data` of a temporary stays valid."""
    a = np.array(v, dtype=np.float32)
    _KEEP.append(a)
    return a


def test_k1_k2_sphere_hits(oracle):
    sc = oracle.OracleScene.builtin("test_sphere", 2.0)
    o = np.zeros((2, 3), np.float32)
    d = np.array([[0, 0, -1], [0, -1, 0]], np.float32)
    h = sc.closest_hit(o, d)
    assert h["prim"].tolist() == [1, 0]                       # K1: "Test" sphere is list index 1
    assert h["t"][0] == np.float32(0.5)
    np.testing.assert_array_equal(h["p"][0], [0, 0, -0.5])
    np.testing.assert_array_equal(h["n"][0], [0, 0, 1])
    assert h["front_face"][0] == 1
    np.testing.assert_allclose(h["uv"][0], [0.25, 0.5], atol=1e-7)
    assert h["t"][1].view(np.uint32) == 0x3F014780            # K2: 0.50499725f


def test_k3_get_uv_axes(oracle):
    L = oracle.lib()
    cases = [((1, 0, 0), (0.5, 0.5)), ((0, 0, 1), (0.25, 0.5)), ((0, 0, -1), (0.75, 0.5)), ((0, 1, 0), (0.5, 1.0)),
             ((0, -1, 0), (0.5, 0.0))]
    for n, want in cases:
        out = np.zeros(2, np.float32)
        L.orc_get_uv(f3(*n).ctypes.data, out.ctypes.data)
        np.testing.assert_allclose(out, want, atol=1e-7)
    # seam is sign-of-zero sensitive: n=(-1,+0,0): -n.z = -0.0 -> atan2(-0,-1) = -pi -> u = 0; n.z=-0.0 -> u = 1
    out = np.zeros(2, np.float32)
    L.orc_get_uv(f3(-1, 0, 0.0).ctypes.data, out.ctypes.data)
    assert abs(out[0]) < 1e-7
    L.orc_get_uv(f3(-1, 0, -0.0).ctypes.data, out.ctypes.data)
    assert abs(out[0] - 1.0) < 1e-7


def test_k4_camera(oracle, rtlib):
    from ray_tracing_in_one_weekend_b200.api import RtCamera
    cam = RtCamera()
    lookfrom, lookat, vup = f3(0, 0, 0), f3(0, 0, -1), f3(0, 1, 0)  # keep the arrays alive across the call
    oracle.lib().orc_camera_new(lookfrom.ctypes.data, lookat.ctypes.data, vup.ctypes.data,
                                C.c_float(90.0), C.c_float(2.0), C.byref(cam))
    np.testing.assert_allclose(list(cam.horizontal), [4, 0, 0], atol=1e-6)
    np.testing.assert_allclose(list(cam.vertical), [0, 2, 0], atol=1e-6)
    np.testing.assert_allclose(list(cam.lower_left_corner), [-2, -1, -1], atol=1e-6)
    uv = np.array([[0.5, 0.5], [0, 0]], np.float32)
    o, d = oracle.camera_get_rays(cam, uv)
    np.testing.assert_allclose(d[0], [0, 0, -1], atol=1e-7)
    np.testing.assert_allclose(d[1], [-0.81649655, -0.40824828, -0.40824828], atol=1e-7)
    # the product's host-side Camera::get_ray is bit-identical
    o2, d2 = rtlib.camera_get_rays(cam, uv)
    assert np.array_equal(d.view(np.uint32), d2.view(np.uint32))


def test_k5_offset_hit_point(oracle):
    L = oracle.lib()
    out = np.zeros(3, np.float32)
    L.orc_offset_hit_point(f3(1, 2, 3).ctypes.data, f3(0, 0, 1).ctypes.data, out.ctypes.data)
    assert out.tolist() == [1.0, 2.0, 3.00006103515625]      # bits(3.0) + 256
    L.orc_offset_hit_point(f3(0.01, 0, 0).ctypes.data, f3(1, 0, 0).ctypes.data, out.ctypes.data)
    assert out[0] == np.float32(0.01) + np.float32(1.0 / 65536.0)
    # negative coordinate: integer offset is sign-corrected
    L.orc_offset_hit_point(f3(-5, 0, 0).ctypes.data, f3(-1, 0, 0).ctypes.data, out.ctypes.data)
    assert out[0] < -5 and out[0].view(np.uint32) == np.float32(-5).view(np.uint32) + 256


def test_k6_optics(oracle):
    L = oracle.lib()
    s = np.float32(1.0 / np.sqrt(2.0))
    out = np.zeros(3, np.float32)
    L.orc_reflect(f3(s, -s, 0).ctypes.data, f3(0, 1, 0).ctypes.data, out.ctypes.data)
    np.testing.assert_allclose(out, [s, s, 0], atol=1e-7)
    L.orc_refract(f3(0, -1, 0).ctypes.data, f3(0, 1, 0).ctypes.data, C.c_float(1 / 1.5), out.ctypes.data)
    np.testing.assert_allclose(out, [0, -1, 0], atol=1e-7)
    assert abs(L.orc_reflectance(1.0, 1 / 1.5) - 0.04) < 1e-6
    assert abs(L.orc_fresnel_dielectric_2(1.0, 1.5) - 0.04) < 1e-6
    a = 0.3
    assert abs(L.orc_gtr2(1.0, a) - 1 / (np.pi * a * a)) < 1e-4
    assert abs(L.orc_gtr1(0.7, 1.0) - 1 / np.pi) < 1e-7
    assert abs(L.orc_gtr1(0.7, 2.5) - 1 / np.pi) < 1e-7
    assert abs(L.orc_schlick_fresnel(0.5) - 0.5 ** 5) < 1e-8


def test_k7_sky(oracle):
    L = oracle.lib()
    out = np.zeros(3, np.float32)
    L.orc_sky(0, f3(0, 1, 0).ctypes.data, out.ctypes.data)
    np.testing.assert_allclose(out, [0.5, 0.7, 1.0], atol=1e-7)
    L.orc_sky(0, f3(0, -1, 0).ctypes.data, out.ctypes.data)
    np.testing.assert_allclose(out, [1, 1, 1], atol=1e-7)
    L.orc_sky(0, f3(1, 0, 0).ctypes.data, out.ctypes.data)
    np.testing.assert_allclose(out, [0.75, 0.85, 1.0], atol=1e-7)
    L.orc_sky(1, f3(0, 1, 0).ctypes.data, out.ctypes.data)
    assert out.tolist() == [0, 0, 0]


def test_k8_checker(oracle):
    L = oracle.lib()
    assert L.orc_checker_is_odd(f3(0.1, 0.1, 0.1).ctypes.data) == 0   # even
    assert L.orc_checker_is_odd(f3(-0.1, 0.1, 0.1).ctypes.data) == 1  # odd


def test_k9_aabb_edge_cases(oracle):
    L = oracle.lib()
    mn, mx = f3(-1, -1, -1), f3(1, 1, 1)
    hit = lambda o, d, a=1e-3, b=3.4e38: L.orc_aabb_hit(mn.ctypes.data, mx.ctypes.data, f3(*o).ctypes.data,
                                                          f3(*d).ctypes.data, C.c_float(a), C.c_float(b))
    assert hit((0, 0, -5), (0, 0, 1)) == 1          # d.x = d.y = 0 -> inv_d = inf, still hits
    assert hit((2, 0, -5), (0, 0, 1)) == 0
    assert hit((-1, 0, -5), (0, 0, 1)) == 1         # min == o on an axis with d = 0: 0*inf = NaN lane is ignored
    assert hit((0, 0, -5), (0, 0, -1)) == 0         # box behind
    # zero-thickness box never hits (t_max <= t_min): why rect boxes are padded +-1e-4 (hitable.rs:274-277)
    mn2, mx2 = f3(-1, -1, 0), f3(1, 1, 0)
    assert L.orc_aabb_hit(mn2.ctypes.data, mx2.ctypes.data, f3(0, 0, -5).ctypes.data, f3(0, 0, 1).ctypes.data,
                          C.c_float(1e-3), C.c_float(3.4e38)) == 0


def test_k10_furnace_oracle(oracle):
    """Statistical: one convex sphere under a constant white sky -> albedo for Diffuse/Lambert/Metal, 1 for glass."""
    for mat, want in [("diffuse", (0.6, 0.5, 0.4)), ("lambert", (0.6, 0.5, 0.4)), ("metal", (0.6, 0.5, 0.4)),
                      ("dielectric", (1, 1, 1))]:
        sc = oracle.OracleScene.builtin("furnace_" + mat, 1.0)
        img, _ = sc.render(5, 5, 2048)
        np.testing.assert_allclose(img[2, 2], want, rtol=0.03)


def test_k11_tie_rule(oracle, rtlib):
    # covered against the GPU in test_gpu_closest_hit.py; here: list semantics of the oracle alone
    from ray_tracing_in_one_weekend_b200 import api
    prims = (api.RtPrim * 2)()
    for i in range(2):
        prims[i].kind = api.PRIM_SPHERE
        prims[i].g[0:4] = [0.0, 0.0, -1.0, 0.5]
    mats = (api.RtMaterial * 1)()
    mats[0].kind = api.MAT_DIELECTRIC
    mats[0].p[0] = 1.5
    chains = (api.RtXformChain * 1)()
    fs = api.RtFlatScene()
    fs.abi_version, fs.n_prims, fs.prims = 1, 2, prims
    fs.n_xform_chains, fs.xform_chains, fs.n_materials, fs.materials = 1, chains, 1, mats
    sc = oracle.OracleScene.from_flat(C.pointer(fs))
    h = sc.closest_hit(np.zeros((1, 3), np.float32), np.array([[0, 0, -1]], np.float32))
    assert h["prim"][0] == 1


def test_k12_xoshiro256pp_and_seeding(oracle):
    L = oracle.lib()
    s = (C.c_uint64 * 4)(1, 2, 3, 4)
    out = (C.c_uint64 * 10)()
    L.orc_xoshiro_next(s, 10, out)
    assert list(out) == [41943041, 58720359, 3588806011781223, 3591011842654386, 9228616714210784205,
                         9973669472204895162, 14011001112246962877, 12406186145184390807, 15849039046786891736,
                         10450023813501588000]
    # SplitMix64 seeding (seed_from_u64): first output of splitmix64(0) is the published 0xe220a8397b1dcdaf
    st = (C.c_uint64 * 4)()
    L.orc_seed_from_u64(C.c_uint64(0), st)
    assert st[0] == 0xE220A8397B1DCDAF and st[1] == 0x6E789E6AA1B965F4
    # gen::<f32>() in [0,1) with 24-bit resolution
    f = np.zeros(4096, np.float32)
    L.orc_gen_f32(C.c_uint64(95), 4096, f.ctypes.data)
    assert f.min() >= 0 and f.max() < 1 and np.all(f * 16777216 == np.floor(f * 16777216))
    assert 0.45 < f.mean() < 0.55
    # gen_range(0..3) and shuffle are unbiased permutations
    r = (C.c_uint64 * 3000)()
    L.orc_gen_range(C.c_uint64(7), C.c_uint64(3), 3000, r)
    cnt = np.bincount(np.array(list(r)), minlength=3)
    assert cnt.sum() == 3000 and cnt.min() > 900
    p = np.zeros(256, np.uint32)
    L.orc_shuffle(C.c_uint64(1995), 256, p.ctypes.data)
    assert sorted(p.tolist()) == list(range(256)) and p.tolist() != list(range(256))


def test_reference_bvh_node_counts(oracle):
    """BvhNode::new shape facts from SURVEY.md 3.3 (N=533 -> visitation covers every object once)."""
    sc = oracle.OracleScene.builtin("sphere_scene", 1.5)
    assert sc.flat.contents.n_prims == 533
    sc = oracle.OracleScene.builtin("cornell_box", 1.0)
    f = sc.flat.contents
    assert f.n_prims == 6 + 2 * (6 + 1)   # 6 walls + 2 x (6 boundary rects + the medium)
    kinds = [f.prims[i].kind for i in range(f.n_prims)]
    assert kinds.count(5) == 2
    sc = oracle.OracleScene.builtin("final_scene", 1.0)
    assert sc.flat.contents.n_prims == 1 + 1000 + 5 + 400 * 6 + 1 + 1 + 1
