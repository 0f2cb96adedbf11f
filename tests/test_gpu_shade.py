"""Shading parity: Material::emitted/scatter (material.rs, pbr.rs) and Texture::value (texture.rs) on the
GPU vs the oracle, fed IDENTICAL hit records and IDENTICAL injected random numbers.

All 13 material kinds x 4 texture kinds.  +,-,*,/,sqrt are bit-exact by construction; sin/cos/acos/
atan2/log differ between glibc and CUDA by a few ulp, so the tolerance is 2e-5 relative + 2e-6
absolute on attenuation / direction and exact on the scatter() boolean and draw count.
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _random_cases(rng, n):
    """Unit normals, incoming unit directions against the normal, tangents, uv, points."""
    nrm = rng.normal(size=(n, 3)).astype(np.float32)
    nrm /= np.linalg.norm(nrm, axis=1, keepdims=True).astype(np.float32)
    d = rng.normal(size=(n, 3)).astype(np.float32)
    d /= np.linalg.norm(d, axis=1, keepdims=True).astype(np.float32)
    flip = (d * nrm).sum(1) > 0
    d[flip] *= -1  # rec.norm always faces against the incoming ray
    tang = np.cross(np.array([0, 1, 0], np.float32), nrm).astype(np.float32)
    tang /= np.maximum(np.linalg.norm(tang, axis=1, keepdims=True), 1e-20).astype(np.float32)
    p = (rng.random((n, 3), dtype=np.float32) * 20 - 10).astype(np.float32)
    uv = rng.random((n, 2), dtype=np.float32)
    rec = np.zeros((n, 13), np.float32)
    rec[:, 0:3] = p
    rec[:, 3:6] = nrm
    rec[:, 6:9] = tang
    rec[:, 9:11] = uv
    rec[:, 11] = 1.0
    rec[:, 12] = (rng.random(n) < 0.5).astype(np.float32)
    rays = np.zeros((n, 6), np.float32)
    rays[:, 0:3] = p - d
    rays[:, 3:6] = d
    return rays, rec


@pytest.mark.parametrize("scene_name", ["sphere_scene", "textured", "pbr_ibl", "final_scene", "cornell_box", "test_sphere",
                                        "furnace_disney_metal", "furnace_oren_nayar", "furnace_disney_sheen"])
def test_scatter_matches_oracle(rtlib, oracle, scene_name):
    rt = rtlib
    lib = rt.load_test_library()
    hs = rt.HostScene(scene_name, 1.5)
    flat = hs.flat.contents
    scene = rt.Scene(hs.flat, 0, lib=lib)
    ref = oracle.OracleScene.from_flat(hs.flat)
    rng = np.random.default_rng(1234)
    # every material kind present (at most 40 materials per kind)
    kinds = {}
    for i in range(flat.n_materials):
        kinds.setdefault(flat.materials[i].kind, []).append(i)
    mat_ids = np.array([m for k in sorted(kinds) for m in kinds[k][:40]], np.uint32)
    reps = 8
    mat_ids = np.repeat(mat_ids, reps)
    n = mat_ids.size
    rays, rec = _random_cases(rng, n)
    XI = 64
    xi = rng.random((n, XI), dtype=np.float32)
    out = np.zeros((n, 14), np.float32)
    rc = lib.rt_test_shade_cases(scene._h, mat_ids.ctypes.data, rays.ctypes.data, rec.ctypes.data, xi.ctypes.data, XI, n,
                                 out.ctypes.data)
    assert rc == 0, lib.rt_last_error()
    n_checked = 0
    for i in range(n):
        ok, want, used = ref.scatter(int(mat_ids[i]), rays[i, 0:3], rays[i, 3:6], rec[i], xi[i])
        if used > XI:
            continue  # rejection loop ran past the injected sequence (both sides then see 0.5s) — skip
        assert int(out[i, 12]) == ok, (scene_name, i, flat.materials[int(mat_ids[i])].kind)
        assert int(out[i, 13]) == used
        np.testing.assert_allclose(out[i, 0:12], want, rtol=2e-5, atol=2e-6,
                                   err_msg=f"{scene_name} case {i} material kind {flat.materials[int(mat_ids[i])].kind}")
        n_checked += 1
    assert n_checked > n * 0.9
    scene.close()


@pytest.mark.parametrize("scene_name", ["sphere_scene", "textured", "pbr_ibl"])
def test_textures_and_sky_match_oracle(rtlib, oracle, scene_name):
    rt = rtlib
    lib = rt.load_test_library()
    hs = rt.HostScene(scene_name, 1.5)
    flat = hs.flat.contents
    scene = rt.Scene(hs.flat, 0, lib=lib)
    ref = oracle.OracleScene.from_flat(hs.flat)
    rng = np.random.default_rng(99)
    n = 2000
    uv = rng.random((n, 2), dtype=np.float32)
    uv[:8] = [[0, 0], [1, 1], [0, 1], [1, 0], [-0.5, 2.0], [0.5, 0.5], [0.999999, 0.000001], [np.nan, 0.3]]
    p = (rng.random((n, 3), dtype=np.float32) * 30 - 15).astype(np.float32)
    # lattice coordinates beyond 2^31 / 2^63, infinities and NaN: `floor() as isize` saturates, `& 255` keeps 8 bits
    p[8:16] = [[3e9, 0.5, -7e9], [1e20, -1e20, 2.5], [np.inf, 0.25, 0.5], [np.nan, 1.0, 2.0], [-2147483648.0, 2147483648.0, 0.25],
               [2147483520.0, -2147483520.0, 1.5], [-np.inf, 3.0, -3.0], [9.3e18, -9.3e18, 0.75]]
    seen = set()
    for t in range(flat.n_textures):
        kind = flat.textures[t].kind
        if kind in seen and kind == 0:
            continue
        seen.add(kind)
        out = np.zeros((n, 3), np.float32)
        rc = lib.rt_test_texture_cases(scene._h, t, uv.ctypes.data, p.ctypes.data, n, out.ctypes.data)
        assert rc == 0
        want = np.stack([ref.texture_value(t, uv[i], p[i]) for i in range(n)])
        if kind == 3:  # ImageTex: integer texel selection, u8/255 -> exact
            assert np.array_equal(out.view(np.uint32), want.view(np.uint32)), f"image texture {t}"
        elif kind == 2:  # Perlin marble: sin() of a sum of 7 octaves; abs tolerance
            np.testing.assert_allclose(out, want, rtol=0, atol=5e-5)
        else:
            # checker flips where sin(10x)sin(10y)sin(10z) ~ 0: allow a handful of sign-boundary points
            neq = np.count_nonzero(np.abs(out - want).max(1) > 1e-6)
            assert neq <= 2, f"texture {t} kind {kind}: {neq} mismatches"
    # sky (tex id 0xffffffff, p = direction)
    dirs = rng.normal(size=(n, 3)).astype(np.float32)
    dirs /= np.linalg.norm(dirs, axis=1, keepdims=True).astype(np.float32)
    out = np.zeros((n, 3), np.float32)
    rc = lib.rt_test_texture_cases(scene._h, 0xFFFFFFFF, uv.ctypes.data, dirs.ctypes.data, n, out.ctypes.data)
    assert rc == 0
    want = np.stack([ref.sky(dirs[i]) for i in range(n)])
    if flat.sky.kind == 2:
        # tex_sky_color (demo_scene.rs:22-26): uv through acos/atan2 (libm vs libdevice ulps) selects a TEXEL, so a
        # direction within an ulp of a texel edge may pick the neighbour: exact on all but a handful of directions
        same = np.all(out.view(np.uint32) == want.view(np.uint32), axis=1)
        assert (~same).sum() <= 4, f"env-map sky: {(~same).sum()} of {n} directions differ"
    else:
        np.testing.assert_allclose(out, want, rtol=1e-6, atol=1e-7)
    scene.close()


def test_div3_is_the_ieee_quotient(rtlib):
    """`V3f / float` (rt_device.cuh: the three quotients of a normalize through ONE refined reciprocal) equals the IEEE
    division bit for bit: every one of the 2^23 divisor mantissas at several exponents and both signs, against
    numerators that cover the exponent range of the fast path and beyond it (zero, denormal, huge, inf, NaN take the
    plain divisions), with adversarial mantissas (all ones, all zeros, one bit)."""
    rt = rtlib
    lib = rt.load_test_library()
    rng = np.random.default_rng(7)
    mant = np.array([0x000000, 0x7FFFFF, 0x7FFFFE, 0x000001, 0x400000, 0x3FFFFF, 0x555555, 0x2AAAAA], np.uint32)
    vals = []
    for e in (60, 67, 100, 126, 127, 128, 150, 187, 188):   # 2^-67 .. 2^61: both sides of the fast path's range
        vals.append(((np.uint32(e) << np.uint32(23)) | mant).view(np.float32))
    vals.append(rng.standard_normal(64).astype(np.float32))
    vals.append((rng.random(32, dtype=np.float32) * 2 - 1) * np.float32(1e-3))
    vals.append(np.array([0.0, -0.0, 1e-40, -1e-42, np.inf, -np.inf, np.nan, 3.0e38, 1.0, -1.0, 0.5, 1.5], np.float32))
    a = np.concatenate(vals).astype(np.float32)
    a[1::2] *= np.float32(-1)   # both signs
    a = np.ascontiguousarray(a)
    for exp_s, neg in ((127, 0), (126, 1), (100, 0), (150, 1), (66, 0), (188, 0), (30, 0), (230, 1)):
        m = C.c_uint64(123)
        rc = lib.rt_test_div3_sweep(a.ctypes.data, a.size, exp_s, neg, C.byref(m))
        assert rc == 0, lib.rt_last_error()
        assert m.value == 0, f"divisor exponent {exp_s - 127}, negate {neg}: {m.value} quotient components differ from IEEE division"
