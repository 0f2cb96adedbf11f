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


# ------------------------------------------------------------------------------------------------------------------
# K13-K15: known-answer vectors for the rand 0.8.5 pieces SURVEY App. B restates from memory — UniformInt's zone
# rejection (`gen_range(0..3)`, hitable.rs:183), `SliceRandom::shuffle` (texture.rs:72,76,80) — and for the Perlin
# table construction order (texture.rs:62-82).  The literals below were derived with the INDEPENDENT pure-Python
# restatement of that specification kept in this file (`_PyXoshiro`), from the K12-pinned xoshiro256++ stream
# seeded with SplitMix64(1995); they pin the oracle (C++), the product's host RNG (C++) and the Python restatement
# to one another and to the written specification.  (They cannot pin the specification to the crate itself: no Rust
# toolchain / crate source here — "parity unpinned" stands for that last step.)
class _PyXoshiro:
    M = (1 << 64) - 1

    def __init__(self, seed):
        self.s = []
        state = seed
        for _ in range(4):  # SplitMix64 (App. B seed_from_u64)
            state = (state + 0x9E3779B97F4A7C15) & self.M
            z = state
            z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & self.M
            z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & self.M
            self.s.append(z ^ (z >> 31))

    @staticmethod
    def _rotl(x, k):
        return ((x << k) | (x >> (64 - k))) & _PyXoshiro.M

    def next_u64(self):
        s = self.s
        r = (self._rotl((s[0] + s[3]) & self.M, 23) + s[0]) & self.M
        t = (s[1] << 17) & self.M
        s[2] ^= s[0]
        s[3] ^= s[1]
        s[1] ^= s[2]
        s[0] ^= s[3]
        s[2] ^= t
        s[3] = self._rotl(s[3], 45)
        return r

    def next_u32(self):
        return self.next_u64() >> 32

    def gen_f32(self):
        return np.float32(self.next_u32() >> 8) * np.float32(2.0 ** -24)

    def below_u64(self, rng):  # UniformInt<usize>::sample_single
        zone = ((rng << (64 - rng.bit_length())) & self.M) - 1
        while True:
            m = self.next_u64() * rng
            if (m & self.M) <= zone:
                return m >> 64

    def below_u32(self, rng):  # gen_index -> UniformInt<u32>::sample_single
        zone = ((rng << (32 - rng.bit_length())) & 0xFFFFFFFF) - 1
        while True:
            m = self.next_u32() * rng
            if (m & 0xFFFFFFFF) <= zone:
                return m >> 32


K13_GEN_RANGE_3 = [2, 0, 1, 1, 0, 1, 0, 1, 1, 2, 2, 1, 1, 2, 2, 1]          # seed 1995, gen_range(0..3) x 16
K14_SHUFFLE_10 = [9, 8, 5, 4, 6, 2, 1, 7, 3, 0]                              # seed 1995, (0..10).shuffle()
K15_RAND_VEC = [[0x3F740566, 0x3F16BF0A, 0xBED31424], [0x3EA90238, 0xBF64C7C0, 0xBE130BC8], [0x3DA4FD90, 0xBF539CAA, 0xBE9A34C0]]
K15_PERM_X = [201, 114, 245, 225, 76, 60, 50, 128]
K15_PERM_Y = [20, 101, 92, 233, 255, 183, 238, 139]
K15_PERM_Z = [188, 220, 71, 59, 219, 92, 202, 124]


def test_k13_gen_range_known_answer(oracle):
    L = oracle.lib()
    out = (C.c_uint64 * 16)()
    L.orc_gen_range(C.c_uint64(1995), C.c_uint64(3), 16, out)
    assert list(out) == K13_GEN_RANGE_3
    r = _PyXoshiro(1995)
    assert [r.below_u64(3) for _ in range(16)] == K13_GEN_RANGE_3


def test_k14_shuffle_known_answer(oracle):
    L = oracle.lib()
    out = (C.c_uint32 * 10)()
    L.orc_shuffle(C.c_uint64(1995), 10, out)
    assert list(out) == K14_SHUFFLE_10
    r, p = _PyXoshiro(1995), list(range(10))
    for i in range(9, 0, -1):
        j = r.below_u32(i + 1)
        p[i], p[j] = p[j], p[i]
    assert p == K14_SHUFFLE_10


def test_k15_perlin_table_known_answer(oracle, rtlib):
    """texture.rs:62-82: 256 x (x, y, z) draws in [-1,1), then three shuffles that CONTINUE on the same index vector.
    Oracle (orc_perlin_table), product host API (Perlin table 0 of sphere_scene = the first consumer of the main
    thread's RNG, seed 1995, demo_scene.rs:41) and the Python restatement agree with the literals."""
    L = oracle.lib()
    vec = np.zeros((3, 3), np.float32)
    px, py, pz = ((C.c_uint32 * 8)() for _ in range(3))
    L.orc_perlin_table(C.c_uint64(1995), 3, vec.ctypes.data, 8, px, py, pz)
    assert vec.view(np.uint32).tolist() == K15_RAND_VEC
    assert (list(px), list(py), list(pz)) == (K15_PERM_X, K15_PERM_Y, K15_PERM_Z)
    hs = rtlib.HostScene("sphere_scene", 1.5)
    tab = hs.flat.contents.perlin[0]
    got = np.array([[tab.rand_vec[i][k] for k in range(3)] for i in range(3)], np.float32)
    assert got.view(np.uint32).tolist() == K15_RAND_VEC
    assert list(tab.perm_x)[:8] == K15_PERM_X and list(tab.perm_y)[:8] == K15_PERM_Y and list(tab.perm_z)[:8] == K15_PERM_Z
    # the independent restatement
    r = _PyXoshiro(1995)
    v = [[r.gen_f32() * np.float32(2.0) + np.float32(-1.0) for _ in range(3)] for _ in range(256)]
    assert np.array(v[:3], np.float32).view(np.uint32).tolist() == K15_RAND_VEC
    p, perms = list(range(256)), []
    for _ in range(3):
        for i in range(255, 0, -1):
            j = r.below_u32(i + 1)
            p[i], p[j] = p[j], p[i]
        perms.append(list(p))
    assert [perms[0][:8], perms[1][:8], perms[2][:8]] == [K15_PERM_X, K15_PERM_Y, K15_PERM_Z]
    assert sorted(perms[2]) == list(range(256))


# ------------------------------------------------------------------------------------------------------------------
# K16: ConstantMedium::hit (hitable.rs:536-579) by hand.  Medium of density 0.5 (neg_inv_density = -2) bounded by the
# unit sphere at the origin; unit-length ray directions, so distances along the ray are t.
#   ray from (-3,0,0) along +x: boundary roots t = 2 and t = 4 (oc = (-3,0,0): half_b = -3, c = 8, disc = 1), the second
#   query starts at 2 + 1e-4 and finds 4.  xi = 0.5: hit_dist = -2 ln 0.5 = 1.3862944 <= (4 - 2) -> t = 3.3862944,
#   p = (0.3862944, 0, 0), norm = (1,0,0), front_face = true, ONE draw.  xi = 0.25: hit_dist = 2.7725887 > 2 -> miss,
#   one draw.  A ray that misses the sphere: no hit, NO draw.  Origin inside (0,0,0): roots -1 and 1; t1 = -1 is clamped
#   to t_min = 0.001, t2 = 1; xi = 0.9: hit_dist = -2 ln 0.9 = 0.21072103 <= 0.999 -> t = 0.21172103.
#   [t_min, t_max] = [0.001, 3]: t2 clamped to 3, xi = 0.5: 1.3862944 > (3 - 2) -> miss.  [2.5, inf): t1 clamped to 2.5,
#   xi = 0.5: 1.3862944 <= 1.5 -> t = 3.8862944.  t_max = 1.5 (< t1): t2 clamped to 1.5 <= t1 -> miss before any draw.
def _medium_flat(api):
    prims = (api.RtPrim * 2)()
    prims[0].kind, prims[0].mat, prims[0].flags = api.PRIM_SPHERE, 0, api.PRIM_FLAG_BOUNDARY_ONLY
    prims[0].g[0:4] = [0.0, 0.0, 0.0, 1.0]
    prims[1].kind, prims[1].mat, prims[1].aux0, prims[1].aux1 = api.PRIM_MEDIUM, 1, 0, 1
    prims[1].g[0] = -2.0
    tex = (api.RtTexture * 1)()
    tex[0].kind = api.TEX_CONSTANT
    tex[0].p[0:3] = [0.5, 0.5, 0.5]
    mats = (api.RtMaterial * 2)()
    mats[0].kind, mats[0].tex0 = api.MAT_DIFFUSE, 0
    mats[1].kind, mats[1].tex0 = api.MAT_ISOTROPIC, 0
    chains = (api.RtXformChain * 1)()
    fs = api.RtFlatScene()
    fs.abi_version = api.RT_ABI_VERSION
    fs.n_prims, fs.prims = 2, prims
    fs.n_xform_chains, fs.xform_chains = 1, chains
    fs.n_materials, fs.materials = 2, mats
    fs.n_textures, fs.textures = 1, tex
    return fs, (prims, tex, mats, chains)


def test_k16_constant_medium_known_answers(oracle, rtlib):
    api = rtlib.api
    fs, keep = _medium_flat(api)
    rtlib.flat_validate(C.pointer(fs))  # the hand-built description is well formed
    sc = oracle.OracleScene.from_flat(C.pointer(fs))
    INF = np.float32(np.inf)
    f32 = np.float32

    def t_of(t1, xi):  # the reference's binary32 steps: hit_dist = neg_inv_density * ln(xi); t = t1 + hit_dist / |d|
        return f32(t1) + f32(-2.0) * np.log(f32(xi)) / f32(1.0)

    cases = [  # origin, direction, t_min, t_max, xi -> hit, draws, t
        ((-3, 0, 0), (1, 0, 0), 1e-3, INF, 0.5, 1, 1, t_of(2.0, 0.5)),
        ((-3, 0, 0), (1, 0, 0), 1e-3, INF, 0.25, 0, 1, None),
        ((-3, 2, 0), (1, 0, 0), 1e-3, INF, 0.5, 0, 0, None),
        ((0, 0, 0), (1, 0, 0), 1e-3, INF, 0.9, 1, 1, t_of(1e-3, 0.9)),
        ((-3, 0, 0), (1, 0, 0), 1e-3, 3.0, 0.5, 0, 1, None),
        ((-3, 0, 0), (1, 0, 0), 2.5, INF, 0.5, 1, 1, t_of(2.5, 0.5)),
        ((-3, 0, 0), (1, 0, 0), 1e-3, 1.5, 0.5, 0, 0, None),
    ]
    rays = np.array([list(o) + list(d) for o, d, *_ in cases], np.float32)
    tmm = np.array([[c[2], c[3]] for c in cases], np.float32)
    xi = np.array([c[4] for c in cases], np.float32)
    got = sc.medium_cases(1, rays, tmm, xi)
    for k, (o, d, _, _, _, hit, draws, t) in enumerate(cases):
        assert int(got[k, 0]) == hit and int(got[k, 2]) == draws, (k, got[k])
        if hit:
            np.testing.assert_allclose(got[k, 1], t, rtol=2.5e-7)  # (numpy's float32 log vs glibc logf: <= 1 ulp)
            np.testing.assert_allclose(got[k, 3:6], np.array(o, np.float64) + t * np.array(d, np.float64), rtol=0, atol=5e-7)
            assert got[k, 6:10].tolist() == [1.0, 0.0, 0.0, 1.0]  # norm (1,0,0), front_face = true (hitable.rs:570-571)
    np.testing.assert_allclose(got[0, 1], 3.3862944, rtol=2e-7)
    np.testing.assert_allclose(got[3, 1], 0.21172103, rtol=5e-7)
