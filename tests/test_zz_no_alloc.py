"""rt_closest_hit_static — the no-allocation variant of the batch closest hit
(no counterpart in the reference).  Collected LAST on purpose.

Called through the C ABI and compared with the existing entry point rt_closest_hit on the same rays: the two run the
same device functions, so every field of every record is BIT-equal (the tolerance of tests/test_gpu_closest_hit.py for
t / p / n, applied to uv as well).  Two calls in a row on one scene — the second smaller than the first, so stale
records and the stale hit counter of the first call would show — then a second scene on the same device (the array
is shared by every call on a GPU).  Oversized batches are refused with a clear error.  Passes alone in a fresh
process and after the other tests."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _rays(rt, hs, w, h):
    cam = rt.RtCamera.from_buffer_copy(bytes(hs.flat.contents.camera))
    return rt.camera_get_rays(cam, rt.pixel_center_uv(w, h))


def _same(a, b):
    return a.tobytes() == b.tobytes()


def test_static_scratch_variant_equals_rt_closest_hit(rtlib):
    rt = rtlib
    hs = rt.HostScene("sphere_scene", 1.5)
    scene = rt.Scene(hs.flat, 0)
    o, d = _rays(rt, hs, 240, 160)                      # 38 400 rays <= RT_STATIC_MAX_RAYS
    want = scene.closest_hit(o, d)
    got, n_hits = scene.closest_hit_static(o, d)        # call 1
    assert _same(got, want)
    assert n_hits == int((want["prim"] >= 0).sum()) > 0
    # call 2, straight after: fewer rays, other interval -> nothing of call 1 may survive (records, hit counter)
    o2, d2 = o[5000:12345], d[5000:12345]
    want2 = scene.closest_hit(o2, d2, t_min=0.5, t_max=20.0)
    got2, n_hits2 = scene.closest_hit_static(o2, d2, t_min=0.5, t_max=20.0)
    assert _same(got2, want2)
    assert n_hits2 == int((want2["prim"] >= 0).sum())
    # call 3: the first batch again gives the first answer again
    got3, n_hits3 = scene.closest_hit_static(o, d)
    assert _same(got3, want) and n_hits3 == n_hits
    # per-ray times (moving spheres) and another scene on the same device share the same array
    hs2 = rt.HostScene("cover_4k_motion", 1.5)
    scene2 = rt.Scene(hs2.flat, 0)
    om, dm = _rays(rt, hs2, 160, 100)
    tm = np.random.default_rng(7).random(om.shape[0], dtype=np.float32)
    want_m = scene2.closest_hit(om, dm, times=tm)
    got_m, n_m = scene2.closest_hit_static(om, dm, times=tm)
    assert _same(got_m, want_m) and n_m == int((want_m["prim"] >= 0).sum())
    hs3 = rt.HostScene("cornell_box_solid", 1.0)
    scene3 = rt.Scene(hs3.flat, 0)
    oc, dc = _rays(rt, hs3, 128, 128)
    want_c = scene3.closest_hit(oc, dc)
    got_c, n_c = scene3.closest_hit_static(oc, dc)
    assert _same(got_c, want_c) and n_c == int((want_c["prim"] >= 0).sum())
    for s in (scene, scene2, scene3):
        s.close()


def test_static_scratch_variant_limits(rtlib):
    rt = rtlib
    lib = rt.load_library()
    hs = rt.HostScene("test_sphere", 2.0)
    scene = rt.Scene(hs.flat, 0)
    n = rt.api.STATIC_MAX_RAYS
    o = np.zeros((n + 1, 3), np.float32)
    d = np.tile(np.array([[0, 0, -1]], np.float32), (n + 1, 1))
    with pytest.raises(rt.RtError) as e:
        scene.closest_hit_static(o, d)
    assert e.value.code == rt.RT_ERR_ARG and "RT_STATIC_MAX_RAYS" in str(e.value) and str(n + 1) in str(e.value)
    got, n_hits = scene.closest_hit_static(o[:n], d[:n])            # exactly the maximum is accepted
    assert n_hits == n and (got["prim"] == 1).all() and (got["t"] == np.float32(0.5)).all()
    got0, n0 = scene.closest_hit_static(o[:0], d[:0])               # empty batch
    assert got0.shape == (0,) and n0 == 0
    nh = C.c_uint64(99)
    assert lib.rt_closest_hit_static(None, None, None, None, 1, C.c_float(0), C.c_float(1), None, C.byref(nh)) == rt.RT_ERR_ARG
    scene.close()
