"""Host-side logic, no GPU: the C-ABI library loads and exports every symbol of include/rtgpu.h, the host
scene API re-expresses src/demo_scene.rs bit-identically to the oracle's restatement, the BVH builder's
structural invariants hold, argument errors come back as negative codes (never aborts)."""
import ctypes as C
import os
import re

import numpy as np
import pytest

SCENES = ["test_sphere", "sphere_scene", "textured", "cover_defocus", "cover_4k_motion", "simple_light_scene",
          "cornell_box", "cornell_box_solid", "final_scene", "pbr_ibl", "furnace_diffuse", "furnace_disney_metal",
          "smoke_instances"]


def _dynamic_symbols(path):
    """Defined symbols of the shared object's dynamic symbol table (`nm -D --defined-only`)."""
    import subprocess
    out = subprocess.check_output(["nm", "-D", "--defined-only", path], text=True)
    return {ln.split()[-1] for ln in out.splitlines() if ln.strip()}


def test_library_exports_exactly_the_header(rtlib):
    """librtgpu.so's dynamic symbol table == the functions include/rtgpu.h declares: nothing missing, no rt_test_*
    hook, no C++ internal (csrc/exports.map).  The ctypes mirror's list is pinned to the same set."""
    lib = rtlib.load_library()
    header = open(os.path.join(os.path.dirname(rtlib.api.HERE), "include", "rtgpu.h")).read()
    declared = set(re.findall(r"\b(rt_[a-z0-9_]+)\s*\(", header))
    exported = _dynamic_symbols(rtlib.LIB_PATH)
    assert exported == declared, exported ^ declared
    assert declared == set(rtlib.EXPORTED_SYMBOLS), declared ^ set(rtlib.EXPORTED_SYMBOLS)
    for sym in declared:
        assert getattr(lib, sym) is not None
    # the test library = the same ABI plus the hooks, nothing else
    extra = _dynamic_symbols(rtlib.api.TEST_LIB_PATH) - declared
    assert extra == {"rt_test_shade_cases", "rt_test_texture_cases", "rt_test_medium_cases", "rt_test_closest_hit_xi",
                     "rt_test_bvh_validate", "rt_test_div3_sweep"}, extra


def test_library_reads_no_environment(rtlib):
    """The shipped library never looks at the environment: the one getenv in csrc/ sits inside dev_knob() behind
    #ifdef RT_DEV_KNOBS (tools/build_variant.py builds), and the product build does not define it."""
    csrc = os.path.join(rtlib.api.HERE, "csrc")
    hits = []
    for f in sorted(os.listdir(csrc)):
        if f.endswith((".cu", ".cuh", ".cpp", ".hpp", ".h")):
            for n, line in enumerate(open(os.path.join(csrc, f), errors="replace"), 1):
                if "getenv" in line and not line.lstrip().startswith("//"):
                    hits.append((f, n, line.strip()))
    assert len(hits) == 1 and hits[0][0] == "rtgpu.cu", hits
    src = open(os.path.join(csrc, "rtgpu.cu")).read()
    i = src.index("getenv(")
    assert "#ifdef RT_DEV_KNOBS" in src[max(0, i - 200):i]
    from ray_tracing_in_one_weekend_b200 import build
    assert not any("RT_DEV_KNOBS" in f for f in build.NVCC_FLAGS + build.HOST_FLAGS)


def test_struct_layouts_match_header(rtlib):
    api = rtlib.api
    assert C.sizeof(api.RtPrim) == 80 and C.sizeof(api.RtMaterial) == 32 and C.sizeof(api.RtTexture) == 32
    assert C.sizeof(api.RtHit) == 44 and C.sizeof(api.RtPerlin) == 4096 + 768
    assert C.sizeof(api.RtXformOp) == 16 and C.sizeof(api.RtXformChain) == 8


@pytest.mark.parametrize("name", SCENES)
def test_host_scene_equals_oracle_restatement(rtlib, oracle, name):
    """Product flattener output == oracle flattener output, byte for byte (prims in the reference's
    visitation order incl. BvhNode::new's RNG-driven sorts, transforms, materials, textures, Perlin tables,
    decoded images, sky, camera)."""
    hs = rtlib.HostScene(name, 1.5)
    a = rtlib.flat_to_numpy(hs.flat)
    ref = oracle.OracleScene.builtin(name, 1.5)
    b = rtlib.flat_to_numpy(ref.flat)
    for k in a:
        if k == "images":
            assert len(a[k]) == len(b[k]) and all(np.array_equal(x, y) for x, y in zip(a[k], b[k]))
        else:
            assert np.array_equal(a[k], b[k]), k


def test_cover_100k_motion_scene_shape(rtlib):
    hs = rtlib.HostScene("cover_100k_motion", 16 / 9)
    f = hs.flat.contents
    assert f.n_prims == 4 + 316 * 316
    kinds = np.array([f.prims[i].kind for i in range(0, f.n_prims, 97)])
    assert (kinds == 4).mean() > 0.7      # ~80 % moving Diffuse spheres
    assert f.camera.time0 == 0.0 and f.camera.time1 == 1.0


@pytest.mark.parametrize("name", ["test_sphere", "sphere_scene", "cornell_box", "final_scene", "cover_4k_motion", "pbr_ibl",
                                  "smoke_instances", "cornell_box_solid"])
@pytest.mark.parametrize("leaf_max", [1, 4, 8])
def test_bvh_invariants(rtlib, name, leaf_max):
    lib = rtlib.load_test_library()
    hs = rtlib.HostScene(name, 1.5)
    info = rtlib.api.RtBvhInfo()
    rc = lib.rt_test_bvh_validate(hs.flat, leaf_max, C.byref(info))
    assert rc == 0, lib.rt_last_error()
    assert info.node_bytes == 128 and info.max_leaf_prims <= leaf_max and info.n_nodes >= 1
    assert info.max_depth * 3 + 4 <= 64


def test_errors_are_codes_not_aborts(rtlib):
    lib = rtlib.load_library()
    h = C.c_void_p()
    assert lib.rt_host_scene_build(b"no_such_scene", C.c_float(1.0), b"assets", C.byref(h)) == -1
    assert b"unknown scene" in lib.rt_last_error()
    assert lib.rt_host_scene_build(b"sphere_scene", C.c_float(1.0), b"/nonexistent", C.byref(h)) == -3
    # malformed flat scenes
    fs = rtlib.RtFlatScene()
    out = C.c_void_p()
    assert lib.rt_scene_create(C.byref(fs), 0, C.byref(out)) == -1          # abi_version 0
    assert lib.rt_scene_create(None, 0, C.byref(out)) == -1
    hs = rtlib.HostScene("test_sphere", 2.0)
    bad = rtlib.RtFlatScene.from_buffer_copy(bytes(hs.flat.contents))
    prims = (rtlib.api.RtPrim * 2)()
    C.memmove(prims, hs.flat.contents.prims, C.sizeof(prims))
    prims[1].mat = 99
    bad.prims = prims
    assert lib.rt_scene_create(C.byref(bad), 0, C.byref(out)) == -1
    assert b"material" in lib.rt_last_error()
    assert lib.rt_render(None, None, None, None) == -1


def test_no_cpu_fallback_without_gpu(rtlib):
    """On a box without a CUDA device the product path must fail loudly (RT_ERR_CUDA), not fall back."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    hs = rtlib.HostScene("test_sphere", 2.0)
    with pytest.raises(rtlib.RtError) as e:
        rtlib.Scene(hs.flat, 0)
    assert e.value.code == -2


def test_product_does_not_route_through_the_oracle():
    """No file of the product package mentions the oracle library or imports the oracle wrapper."""
    root = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "ray_tracing_in_one_weekend_b200")
    for dp, _, files in os.walk(root):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".hpp", ".h")):
                txt = open(os.path.join(dp, f), errors="replace").read()
                assert "liboracle" not in txt and "oracle_lib" not in txt and "orc_" not in txt, f


def test_png_writer_round_trip(rtlib, tmp_path):
    """rt_write_png replaces image::save_buffer (main.rs:128): decode the file by hand (signature, IHDR, one
    zlib stream of filter-0 rows, CRCs) and compare the pixels."""
    import struct
    import zlib
    a = (np.random.default_rng(1).random((37, 53, 3)) * 255).astype(np.uint8)
    path = str(tmp_path / "out.png")
    rtlib.Scene.write_png(path, a)
    raw = open(path, "rb").read()
    assert raw[:8] == b"\x89PNG\r\n\x1a\n"
    pos, chunks = 8, []
    while pos < len(raw):
        n, typ = struct.unpack(">I4s", raw[pos:pos + 8])
        data = raw[pos + 8:pos + 8 + n]
        assert struct.unpack(">I", raw[pos + 8 + n:pos + 12 + n])[0] == zlib.crc32(typ + data)
        chunks.append((typ, data))
        pos += 12 + n
    assert [c[0] for c in chunks] == [b"IHDR", b"IDAT", b"IEND"]
    assert struct.unpack(">IIBBBBB", chunks[0][1]) == (53, 37, 8, 2, 0, 0, 0)
    rows = np.frombuffer(zlib.decompress(chunks[1][1]), np.uint8).reshape(37, 1 + 53 * 3)
    assert not rows[:, 0].any()
    assert np.array_equal(rows[:, 1:].reshape(37, 53, 3), a)
    lib = rtlib.load_library()
    assert lib.rt_write_png(b"/nonexistent_dir/x.png", a.ctypes.data_as(C.c_void_p), 53, 37) == -3
    assert lib.rt_write_png(None, None, 0, 0) == -3


def _material_class(m, texs):
    """rt_device.cuh CLS_*: 1 Diffuse(Constant) 2 Metal 3 Dielectric 4 Diffuse(Perlin) 5 other material.rs 6 pbr.rs."""
    if m.kind >= 6: return 6
    if m.kind == 3: return 2
    if m.kind == 4: return 3
    if m.kind == 1 and texs[m.tex0].kind == 0: return 1
    if m.kind == 1 and texs[m.tex0].kind == 2: return 4
    return 5


@pytest.mark.parametrize("name", SCENES)
def test_shade_queue_assignment(rtlib, name):
    """rt_flat_shade_queues (the planning step of the material-sorted shade pass): queue 0 is kept for the misses,
    a queue runs exactly one shading class, materials of the same (class, kind, texture kind) share a queue, free
    queues are used to separate the kinds of a mixed class, and the drain order is a permutation with the misses
    after every queue in use."""
    hs = rtlib.HostScene(name, 1.5)
    f = hs.flat.contents
    mats = [f.materials[i] for i in range(f.n_materials)]
    texs = [f.textures[i] for i in range(f.n_textures)]
    used = sorted({f.prims[i].mat for i in range(f.n_prims)})
    mq, q_cls, q_order = hs.shade_queues()
    assert sorted(q_order) == list(range(7))
    assert q_cls[0] == 0
    sig_queue = {}
    for i in used:
        m = mats[i]
        cls = _material_class(m, texs)
        assert 1 <= mq[i] <= 6 and q_cls[mq[i]] == cls, (i, mq[i], q_cls, cls)
        has_tex = m.kind not in (3, 4, 12)
        sig = (cls, m.kind, texs[m.tex0].kind if has_tex else -1)
        assert sig_queue.setdefault(sig, mq[i]) == mq[i]
    n_sig, n_queues = len(sig_queue), len(set(sig_queue.values()))
    assert n_queues == min(n_sig, 6), (sig_queue, q_cls)  # every free queue is used while kinds still share one
    in_use = set(sig_queue.values())
    pos = {q: k for k, q in enumerate(q_order)}
    assert all(pos[q] < pos[0] for q in in_use)           # the misses are drained after every hit queue in use
    rank = {4: 0, 6: 1, 5: 2, 3: 3, 1: 4, 2: 5}            # Perlin, pbr.rs, classic, dielectric, diffuse, metal
    ranks = [rank[q_cls[q]] for q in sorted(in_use)]
    assert ranks == sorted(ranks) and sorted(in_use) == list(range(1, len(in_use) + 1))  # numbered by cost, no gaps
    hs.close()


def test_shade_queues_split_the_pbr_class(rtlib):
    """pbr_ibl holds seven pbr.rs material kinds (one class): they are spread over the six hit queues, the ground's
    RoughPlastic (by far the largest surface) in a queue of its own."""
    hs = rtlib.HostScene("pbr_ibl", 1.5)
    f = hs.flat.contents
    mq, q_cls, _ = hs.shade_queues()
    assert q_cls[1:7] == [6] * 6
    ground = max(range(f.n_prims), key=lambda i: f.prims[i].g[3])
    gq = mq[f.prims[ground].mat]
    kinds_in_gq = {f.materials[i].kind for i in range(f.n_materials) if mq[i] == gq}
    assert kinds_in_gq == {8}
    hs.close()


def _build_c_example(tmp_path):
    import subprocess
    repo = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    libdir = os.path.join(repo, "ray_tracing_in_one_weekend_b200")
    exe = str(tmp_path / "render_scene")
    subprocess.check_call(["gcc", "-std=c99", "-pedantic", "-Wall", "-Wextra", "-Werror", "-I" + os.path.join(repo, "include"),
                           os.path.join(repo, "examples", "render_scene.c"), "-L" + libdir, "-lrtgpu",
                           "-Wl,-rpath," + libdir, "-o", exe])
    return repo, exe


def test_header_is_plain_c99_and_the_c_example_links(rtlib, tmp_path):
    """The boundary is a C ABI: include/rtgpu.h compiles as pedantic C99 and examples/render_scene.c (a C caller
    with no C++/CUDA/torch on its side) links against librtgpu.so.  Without a GPU it must fail with the library's
    error code, not abort and not fall back to the CPU."""
    import subprocess
    repo, exe = _build_c_example(tmp_path)
    if os.path.exists("/dev/nvidiactl"):
        pytest.skip("the no-device half of this test needs a box without a GPU")
    r = subprocess.run([exe, "test_sphere", "32", "16", "1", os.path.join(repo, "assets"), str(tmp_path / "o.png")],
                       capture_output=True, text=True)
    assert r.returncode == 1 and "rt_scene_create" in r.stderr and "failed: -2" in r.stderr, (r.returncode, r.stderr)
    assert not (tmp_path / "o.png").exists()


def _synthetic_flat(api, mats, prim_mats, radii):
    """A minimal RtFlatScene: one ConstantTex, the given materials (kind per entry), one sphere per prim_mats entry."""
    tex = (api.RtTexture * 1)()
    tex[0].kind = 0
    m = (api.RtMaterial * len(mats))()
    for i, k in enumerate(mats):
        m[i].kind, m[i].tex0 = k, 0
    p = (api.RtPrim * len(prim_mats))()
    for i, (mi, r) in enumerate(zip(prim_mats, radii)):
        p[i].kind, p[i].mat = 0, mi  # RT_PRIM_SPHERE
        p[i].g[0], p[i].g[1], p[i].g[2], p[i].g[3] = float(i), 0.0, 0.0, r
    f = api.RtFlatScene()
    f.abi_version = api.RT_ABI_VERSION
    f.n_prims, f.n_materials, f.n_textures = len(prim_mats), len(mats), 1
    f.prims, f.materials, f.textures = p, m, tex
    return f, (tex, m, p)


def test_shade_queues_synthetic_edge_cases(rtlib):
    """More kinds than queues (the lightest pbr.rs kinds share the class queue, the heaviest by surface get their
    own), an unreferenced material (lands in its class's queue), a bad material index (error code, no crash)."""
    api = rtlib.api
    lib = rtlib.load_library()
    # materials: Diffuse, Metal, then seven pbr.rs kinds (6..12) and one unreferenced Dielectric
    mats = [1, 3, 6, 7, 8, 9, 10, 11, 12, 4]
    prim_mats = [0, 1, 2, 3, 4, 5, 6, 7, 8]
    radii = [1.0, 1.0, 9.0, 8.0, 7.0, 0.3, 0.2, 0.1, 0.05]
    f, keep = _synthetic_flat(api, mats, prim_mats, radii)
    mq = (C.c_uint8 * len(mats))()
    qc, qo = C.c_uint32(), C.c_uint32()
    assert lib.rt_flat_shade_queues(C.byref(f), mq, C.byref(qc), C.byref(qo)) == 0
    mq = list(mq)
    q_cls = [(qc.value >> (4 * q)) & 15 for q in range(7)]
    # classes in use: Diffuse(Constant)=1, Metal=2, pbr=6 -> 3 queues, 3 free ones go to the three largest pbr kinds
    assert sorted(q_cls[1:]) == [1, 2, 6, 6, 6, 6]
    own = [mq[2], mq[3], mq[4]]                     # radii 9, 8, 7
    assert len(set(own)) == 3 and all(q_cls[q] == 6 for q in own)
    shared = {mq[i] for i in (5, 6, 7, 8)}          # the four small ones share the class queue
    assert len(shared) == 1 and shared.isdisjoint(own) and q_cls[shared.pop()] == 6
    assert q_cls[mq[0]] == 1 and q_cls[mq[1]] == 2
    assert mq[9] == 0 or q_cls[mq[9]] == 3          # unreferenced Dielectric: no queue of its class exists -> 0
    # pbr queues come before Diffuse and Metal in the drain order (numbered by cost)
    assert max(q for q in range(1, 7) if q_cls[q] == 6) < min(mq[0], mq[1])
    # bad material index
    f2, keep2 = _synthetic_flat(api, [1], [5], [1.0])
    assert lib.rt_flat_shade_queues(C.byref(f2), mq if False else (C.c_uint8 * 1)(), C.byref(qc), C.byref(qo)) < 0
    assert b"material" in lib.rt_last_error()


# ------------------------------------------------------------------------------------------- ABI 2 additions
@pytest.mark.parametrize("name,needle", [("unsupported_nested_medium", b"boundary of a ConstantMedium"),
                                         ("unsupported_disney_metal_rect", b"stale HitRecord.tang"),
                                         ("unsupported_image_isotropic_medium", b"stale HitRecord.uv")])
def test_unreproducible_scenes_are_refused(rtlib, name, needle):
    """Quirk A4 and stochastic boundaries: the flattener accepts what the reference's API accepts,
    rt_flat_validate / rt_scene_create answer RT_ERR_UNSUPPORTED with the reason instead of rendering a picture the
    reference would not produce."""
    lib = rtlib.load_library()
    hs = rtlib.HostScene(name, 1.5)
    assert lib.rt_flat_validate(hs.flat) == rtlib.RT_ERR_UNSUPPORTED
    assert needle in lib.rt_last_error()
    out = C.c_void_p()
    assert lib.rt_scene_create(hs.flat, 0, C.byref(out)) == rtlib.RT_ERR_UNSUPPORTED  # checked before any device call
    assert not out.value


def test_media_under_wrappers_flatten_with_prefixed_chains(rtlib):
    """smoke_instances: Translate(ConstantMedium(Sphere)), Translate(RotateY(ConstantMedium(GBox))),
    RotateY(ConstantMedium(Translate(GBox))) — the medium record carries the wrappers AROUND it, every boundary
    primitive's chain starts with exactly those ops, and the description validates."""
    hs = rtlib.HostScene("smoke_instances", 1.0)
    f = hs.flat.contents
    rtlib.flat_validate(hs.flat)
    media = [f.prims[i] for i in range(f.n_prims) if f.prims[i].kind == 5]
    assert len(media) == 3 and all(m.xform != 0 for m in media)
    depth = []
    for m in media:
        mc = f.xform_chains[m.xform]
        for k in range(m.aux1):
            b = f.prims[m.aux0 + k]
            assert b.flags & 1
            bc = f.xform_chains[b.xform]
            assert bc.n_ops >= mc.n_ops
            for j in range(mc.n_ops):
                x, y = f.xform_ops[bc.first_op + j], f.xform_ops[mc.first_op + j]
                assert (x.kind, list(x.v)) == (y.kind, list(y.v))
        depth.append((mc.n_ops, f.xform_chains[f.prims[m.aux0].xform].n_ops))
    assert sorted(depth) == [(1, 1), (1, 2), (2, 2)]
    # a boundary chain that does not start with the medium's chain is a malformed description
    bad = rtlib.RtFlatScene.from_buffer_copy(bytes(f))
    prims = (rtlib.api.RtPrim * f.n_prims)()
    C.memmove(prims, f.prims, C.sizeof(prims))
    mi = [i for i in range(f.n_prims) if prims[i].kind == 5][0]
    prims[prims[mi].aux0].xform = 0
    bad.prims = prims
    lib = rtlib.load_library()
    assert lib.rt_flat_validate(C.byref(bad)) == rtlib.RT_ERR_ARG and b"chain" in lib.rt_last_error()


def test_partition_samples_matches_the_reference_arithmetic(rtlib):
    lib = rtlib.load_library()
    b, e = C.c_uint32(), C.c_uint32()
    for begin, end, n in [(0, 256, 8), (0, 37, 2), (5, 5, 3), (3, 10, 4), (0, 1, 8), (100, 4196, 8)]:
        spans = []
        for r in range(n):
            assert lib.rt_partition_samples(begin, end, n, r, C.byref(b), C.byref(e)) == 0
            spans.append((b.value, e.value))
        assert spans[0][0] == begin and spans[-1][1] == end
        assert all(spans[k][1] == spans[k + 1][0] for k in range(n - 1))
        sizes = [y - x for x, y in spans]
        assert max(sizes) - min(sizes) <= 1 and sizes == sorted(sizes, reverse=True)
    assert lib.rt_partition_samples(0, 8, 0, 0, C.byref(b), C.byref(e)) == -1
    assert lib.rt_partition_samples(0, 8, 2, 2, C.byref(b), C.byref(e)) == -1
    assert lib.rt_partition_samples(9, 8, 2, 0, C.byref(b), C.byref(e)) == -1


def test_comm_errors_are_codes(rtlib):
    """The multi-GPU entry points fail with codes, never abort: bad arguments -> RT_ERR_ARG; without a CUDA device
    -> RT_ERR_CUDA (no CPU fallback for the reduce either)."""
    import torch
    lib = rtlib.load_library()
    out = C.c_void_p()
    assert lib.rt_comm_create(None, 2, 0, 0, C.byref(out)) == -1
    uid = (C.c_uint8 * 128)()
    assert lib.rt_comm_create(uid, 2, 2, 0, C.byref(out)) == -1
    assert lib.rt_comm_create(uid, 0, 0, 0, C.byref(out)) == -1
    assert lib.rt_comm_create_all(None, 1, C.byref(out)) == -1
    devs = (C.c_int * 2)(0, 0)
    assert lib.rt_comm_create_all(devs, 0, C.byref(out)) == -1
    assert lib.rt_render_reduce(None, None, None, None, None) == -1
    assert lib.rt_render_reduce_device(None, None, None, None, None) == -1
    assert lib.rt_comm_info(None, None, None, None) == -1
    lib.rt_comm_destroy(None)
    lib.rt_release_cached_memory()
    if not torch.cuda.is_available():
        one = (C.c_int * 1)(0)
        assert lib.rt_comm_create_all(one, 1, C.byref(out)) == rtlib.RT_ERR_CUDA
        assert lib.rt_comm_create(uid, 1, 0, 0, C.byref(out)) == rtlib.RT_ERR_CUDA
        assert not out.value
