"""GPU LBVH builder (RT_SCENE_GPU_BVH, csrc/gpu_bvh.cu).

Any tree with conservative boxes must give the SAME closest hits (the reference's tree only fixes the primitive
order, which lives in the flat indices): the scenes built on the device are therefore checked bit for bit against
the committed golden vectors, against the host-SAH-built scene on a big random ray batch, and by rendering."""
import glob
import os
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "golden"))
import make_golden  # noqa: E402

FILES = sorted(glob.glob(os.path.join(HERE, "golden", "*.npz")))


@pytest.mark.parametrize("path", FILES, ids=lambda p: os.path.basename(p)[:-4])
def test_gpu_built_tree_matches_golden(rtlib, path):
    rt = rtlib
    g = np.load(path)
    name, w, h = os.path.basename(path)[:-4], int(g["w"]), int(g["h"])
    hs = rt.HostScene(name, w / h)
    o, d, times = make_golden.batch(rt, hs, name, w, h)
    scene = rt.Scene(hs.flat, 0, flags=rt.SCENE_GPU_BVH)
    info = scene.bvh_info()
    assert info.builder == 1 and info.max_leaf_prims == 1 and info.n_nodes >= 1 and info.node_bytes == 128
    got = scene.closest_hit(o, d, times=times)
    assert np.array_equal(got["prim"], g["prim"])
    hit = g["prim"] >= 0
    for f in ("t", "p", "n"):
        assert np.array_equal(got[f].view(np.uint32)[hit], g[f][hit]), f"{name}: {f} not bit-equal"
    scene.close()
    hs.close()


@pytest.mark.parametrize("name", ["sphere_scene", "final_scene", "cover_100k_motion"])
def test_gpu_built_tree_equals_host_built_tree_on_random_rays(rtlib, name):
    rt = rtlib
    hs = rt.HostScene(name, 1.5)
    a = rt.Scene(hs.flat, 0, flags=0)
    b = rt.Scene(hs.flat, 0, flags=rt.SCENE_GPU_BVH)
    ia, ib = a.bvh_info(), b.bvh_info()
    assert ia.builder == 0 and ib.builder == 1 and ia.n_world_prims == ib.n_world_prims
    rng = np.random.default_rng(11)
    n = 200_000
    cam = rt.RtCamera.from_buffer_copy(bytes(hs.flat.contents.camera))
    o1, d1 = rt.camera_get_rays(cam, rng.random((n // 2, 2), dtype=np.float32))
    # incoherent rays from points around the primary hits
    h1 = a.closest_hit(o1, d1)
    o2 = np.where((h1["prim"] >= 0)[:, None], h1["p"] + h1["n"] * np.float32(1e-2), o1).astype(np.float32)
    d2 = rng.normal(size=(n // 2, 3)).astype(np.float32)
    d2 /= np.linalg.norm(d2, axis=1, keepdims=True).astype(np.float32)
    o, d = np.concatenate([o1, o2]), np.concatenate([d1, d2.astype(np.float32)])
    times = rng.random(n, dtype=np.float32) if "motion" in name else None
    ha, hb = a.closest_hit(o, d, times=times), b.closest_hit(o, d, times=times)
    assert np.array_equal(ha["prim"], hb["prim"])
    hit = ha["prim"] >= 0
    assert hit.sum() > n // 4
    assert np.array_equal(ha["t"].view(np.uint32)[hit], hb["t"].view(np.uint32)[hit])
    print(f"{name}: {ia.n_world_prims} prims; host SAH {ia.n_nodes} nodes depth {ia.max_depth} in {ia.build_ms:.1f} ms; "
          f"GPU LBVH {ib.n_nodes} nodes depth {ib.max_depth} in {ib.build_ms:.1f} ms")
    a.close()
    b.close()
    hs.close()


def test_gpu_built_tree_renders_the_same_image(rtlib):
    """Same per-path RNG streams and the same closest hits -> the same rays and the same image, whatever the tree."""
    rt = rtlib
    hs = rt.HostScene("sphere_scene", 1.5)
    a = rt.Scene(hs.flat, 0, flags=0)
    b = rt.Scene(hs.flat, 0, flags=rt.SCENE_GPU_BVH)
    ia, sa = a.render(96, 64, 32)
    ib, sb = b.render(96, 64, 32)
    assert sa.rays == sb.rays and sa.shade_calls == sb.shade_calls
    np.testing.assert_allclose(ia, ib, rtol=1e-4, atol=1e-5)
    a.close()
    b.close()
    hs.close()


def test_single_primitive_and_flag_errors(rtlib):
    import ctypes as C
    rt = rtlib
    scene = rt.Scene.from_name("furnace_diffuse", 1.0, 0, flags=rt.SCENE_GPU_BVH)
    img, st = scene.render(9, 9, 64)
    assert st.paths == 81 * 64 and np.isfinite(img).all()
    scene.close()
    hs = rt.HostScene("test_sphere", 2.0)
    out = C.c_void_p()
    assert rt.load_library().rt_scene_create_ex(hs.flat, 0, 0x80, C.byref(out)) == -1
    hs.close()
