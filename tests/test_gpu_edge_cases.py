"""Edge cases and full-size properties of the CUDA path (through the C ABI).

The reference ships no tests; the cases below are the ones its code paths make possible: empty and
ragged ray batches, 1x1 and odd-sized frames, empty sample ranges, depth limits 0 and 1 (main.rs:40-42), and — at
BASELINE.json's full 1920x1080 size, where the oracle is too slow — size-independent properties: path/ray
bookkeeping, additivity over sample ranges (the multi-GPU contract), agreement of the three renderers, and
consistency of the full-size frame with a small frame of the same scene."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

LUM = np.array([0.2126, 0.7152, 0.0722], np.float32)
MODES = [1, 3]


def test_empty_and_ragged_ray_batches(rtlib, oracle):
    rt = rtlib
    hs = rt.HostScene("sphere_scene", 1.5)
    scene = rt.Scene(hs.flat, 0)
    ref = oracle.OracleScene.from_flat(hs.flat)
    cam = rt.RtCamera.from_buffer_copy(bytes(hs.flat.contents.camera))
    rng = np.random.default_rng(3)
    for n in (0, 1, 31, 32, 33, 127, 1000, 4099):
        uv = rng.random((n, 2), dtype=np.float32)
        o, d = rt.camera_get_rays(cam, uv) if n else (np.zeros((0, 3), np.float32), np.zeros((0, 3), np.float32))
        got = scene.closest_hit(o, d)
        assert got.shape == (n,)
        if n:
            want = ref.closest_hit(o, d)
            assert np.array_equal(got["prim"], want["prim"])
            assert np.array_equal(got["t"].view(np.uint32), want["t"].view(np.uint32))
    scene.close()


def test_degenerate_rays(rtlib, oracle):
    """Axis-parallel directions (1/d = inf in the slab test), rays starting inside / on a sphere, a ray that only
    grazes.  (A zero direction is outside the contract: Sphere::hit then yields t = 0/0 = NaN for every sphere,
    hitable.rs:85, and the winner depends on visiting order; main.rs:39 never traces such a ray.)"""
    rt = rtlib
    hs = rt.HostScene("test_sphere", 2.0)
    scene = rt.Scene(hs.flat, 0)
    ref = oracle.OracleScene.from_flat(hs.flat)
    o = np.array([[0, 0, 0], [0, 0, 0], [0, 0, 0], [0, 0, -1], [0, 0, -0.5], [0, 5, -1], [0.5, 0, 0], [3, 0, -1]], np.float32)
    d = np.array([[0, 0, -1], [0, -1, 0], [1, 0, 0], [0, 1, 0], [0, 0, -1], [0, -1, 0], [0, 0, -1], [-1, 0, 0]], np.float32)
    got, want = scene.closest_hit(o, d), ref.closest_hit(o, d)
    assert np.array_equal(got["prim"], want["prim"])
    hit = want["prim"] >= 0
    assert np.array_equal(got["t"].view(np.uint32)[hit], want["t"].view(np.uint32)[hit])
    scene.close()


@pytest.mark.parametrize("mode", MODES)
@pytest.mark.parametrize("w,h,spp", [(1, 1, 1), (1, 1, 257), (7, 5, 3), (33, 1, 2), (1, 40, 2), (97, 61, 1)])
def test_small_and_odd_frames(rtlib, w, h, spp, mode):
    rt = rtlib
    scene = rt.Scene.from_name("sphere_scene", max(w / h, 0.05), 0)
    img, st = scene.render(w, h, spp, mode=mode)
    assert st.paths == w * h * spp and st.rays >= st.paths
    assert img.shape == (h, w, 3) and np.isfinite(img).all() and (img >= 0).all()
    ref, sr = scene.render(w, h, spp, mode=1)
    assert st.rays == sr.rays                       # same per-path RNG streams in every renderer
    np.testing.assert_allclose(img, ref, rtol=1e-4, atol=1e-5)
    scene.close()


@pytest.mark.parametrize("mode", MODES)
def test_empty_sample_range_and_depth_limits(rtlib, oracle, mode):
    rt = rtlib
    hs = rt.HostScene("sphere_scene", 1.5)
    scene = rt.Scene(hs.flat, 0)
    img, st = scene.render(48, 32, 8, sample_begin=5, sample_end=5, spp_total=8, mode=mode)
    assert st.paths == 0 and st.rays == 0 and not img.any()
    # main.rs:40-42: depth > MAX_DEPTH returns black -> a path traces at most max_depth + 1 rays
    ref = oracle.OracleScene.from_flat(hs.flat)
    for depth in (0, 1, 3):
        img, st = scene.render(48, 32, 64, max_depth=depth, mode=mode)
        assert st.paths <= st.rays <= st.paths * (depth + 1)
        want, so = ref.render(48, 32, 64, max_depth=depth)
        assert abs(st.rays / st.paths - so.rays / so.paths) < 0.02
        lg, lc = float((img * LUM).sum(-1).mean()), float((want * LUM).sum(-1).mean())
        assert abs(lg - lc) <= 0.02 * lc + 1e-4, (depth, lg, lc)
    scene.close()


def test_full_size_properties(rtlib):
    """BASELINE.json size (1920x1080): bookkeeping, additivity over sample ranges, renderer agreement,
    consistency with a small frame."""
    rt = rtlib
    w, h, spp = 1920, 1080, 4
    scene = rt.Scene.from_name("sphere_scene", w / h, 0)
    full, st = scene.render(w, h, spp, mode=3)
    assert st.paths == w * h * spp
    assert 2.3 < st.rays / st.paths < 2.8            # cover scene: 2.53 rays per path
    assert np.isfinite(full).all()
    a, sa = scene.render(w, h, spp, sample_begin=0, sample_end=1, spp_total=spp, mode=3)
    b, sb = scene.render(w, h, spp, sample_begin=1, sample_end=spp, spp_total=spp, mode=3)
    assert sa.rays + sb.rays == st.rays
    np.testing.assert_allclose(a + b, full, rtol=2e-5, atol=1e-6)
    for mode in (1,):
        other, so = scene.render(w, h, spp, mode=mode)
        assert so.rays == st.rays and so.shade_calls == st.shade_calls
        np.testing.assert_allclose(other, full, rtol=1e-4, atol=1e-5)
    small, _ = scene.render(480, 270, 64, mode=3)
    lf, ls = float((full * LUM).sum(-1).mean()), float((small * LUM).sum(-1).mean())
    assert abs(lf - ls) / ls < 0.01
    # 4x4 box-filtered full frame vs the quarter-size frame: same picture
    down = full.reshape(270, 4, 480, 4, 3).mean((1, 3))
    assert np.sqrt(np.mean((down - small) ** 2)) < 0.08
    scene.close()


def test_resolve_saturation_and_nan(rtlib):
    """main.rs:98-105: powf(0.5), (c*255.99) as u8 — Rust's saturating cast: >255 -> 255, <0 -> 0, NaN -> 0; flip."""
    rt = rtlib
    scene = rt.Scene.from_name("test_sphere", 2.0, 0)
    mean = np.zeros((2, 3, 3), np.float32)
    mean[0, 0] = (0.25, 1.0, 4.0)
    mean[0, 1] = (np.nan, -1.0, np.inf)
    mean[1, 2] = (0.0, 1e-9, 0.999)
    out = scene.resolve_rgb8(mean)
    assert out.shape == (2, 3, 3)
    assert tuple(out[1, 0]) == (127, 255, 255)      # row 0 (bottom) lands in the last output row
    assert tuple(out[1, 1]) == (0, 0, 255)
    assert tuple(out[0, 2]) == (0, int(np.float32(np.sqrt(np.float32(1e-9))) * np.float32(255.99)), int(np.float32(np.sqrt(np.float32(0.999))) * np.float32(255.99)))
    scene.close()
