#!/usr/bin/env python3
"""Generate the golden closest-hit fixtures under tests/golden/ (run from the repo root, CPU only).

    python tests/golden/make_golden.py

The reference is a Rust crate and neither rustc nor cargo exists in this image, so the vectors cannot be
produced by the reference itself ("parity unpinned").  They are produced by the CPU oracle
(oracle/, the C++ restatement of src/hitable.rs / src/camera.rs) in LIST mode — the semantic ground truth
of HitableList::hit (src/hitable.rs:117-132) — on a fixed primary-ray batch per scene: all pixel centres
u=(i+0.5)/w, v=(j+0.5)/h through Camera::get_ray (src/camera.rs:40-46).  The fixtures freeze today's
oracle output so that (a) any later change of the oracle or of the scene builders shows up on the CPU
(tests/test_golden.py::test_oracle_reproduces_golden) and (b) the CUDA path is compared with committed
vectors that do not depend on the oracle being rebuilt on the GPU box
(tests/test_gpu_golden.py).

Each <scene>.npz holds: w, h, times (float32[n], motion scenes only), prim (int32[n], -1 = miss),
t / p / n as uint32 BIT PATTERNS of the float32 results (parity is bit-exact), front_face (uint8),
flat_sha256 (hash of the flattened scene's POD arrays: pins scene construction, i.e. the xoshiro256++ /
SplitMix64 / Perlin-table / BvhNode-order restatement).
"""
import hashlib
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, REPO)
sys.path.insert(0, os.path.join(REPO, "tests"))

# scene, width, height  (small batches: a few thousand rays each; the whole directory stays < 1 MB)
CASES = [
    ("test_sphere", 64, 32),
    ("sphere_scene", 96, 64),
    ("textured", 72, 48),
    ("simple_light_scene", 72, 48),
    ("cornell_box", 64, 64),
    ("cornell_box_solid", 64, 64),
    ("final_scene", 64, 64),
    ("pbr_ibl", 96, 54),
    ("cover_4k_motion", 72, 48),
]


def flat_digest(rt, flat) -> str:
    """sha256 over the POD arrays of an RtFlatScene (prims, transform ops/chains, materials, textures, Perlin
    tables, camera, sky) — not over pointers or image payloads."""
    import ctypes as C
    f = flat.contents
    h = hashlib.sha256()

    def add(ptr, n):
        if n and ptr:
            h.update(C.string_at(ptr, n * C.sizeof(ptr._type_)))

    add(f.prims, f.n_prims)
    add(f.xform_ops, f.n_xform_ops)
    add(f.xform_chains, f.n_xform_chains)
    add(f.materials, f.n_materials)
    add(f.textures, f.n_textures)
    add(f.perlin, f.n_perlin)
    h.update(bytes(f.camera))
    h.update(bytes(f.sky))
    return h.hexdigest()


def batch(rt, hs, name, w, h):
    uv = rt.pixel_center_uv(w, h)
    cam = rt.RtCamera.from_buffer_copy(bytes(hs.flat.contents.camera))
    o, d = rt.camera_get_rays(cam, uv)
    times = np.random.default_rng(7).random(o.shape[0], dtype=np.float32) if "motion" in name else None
    return o, d, times


def main():
    import oracle_lib
    import ray_tracing_in_one_weekend_b200 as rt
    rt.load_library()  # host-side scene builders + camera only; no device call
    for name, w, h in CASES:
        hs = rt.HostScene(name, w / h)
        o, d, times = batch(rt, hs, name, w, h)
        want = oracle_lib.OracleScene.from_flat(hs.flat).closest_hit(o, d, times=times)
        out = dict(w=np.int32(w), h=np.int32(h), prim=want["prim"].astype(np.int32),
                   t=want["t"].view(np.uint32), p=want["p"].view(np.uint32), n=want["n"].view(np.uint32),
                   front_face=want["front_face"].astype(np.uint8), flat_sha256=np.array(flat_digest(rt, hs.flat)))
        if times is not None:
            out["times"] = times
        path = os.path.join(HERE, name + ".npz")
        np.savez_compressed(path, **out)
        print(f"{name}: {o.shape[0]} rays, {int((want['prim'] >= 0).sum())} hits -> {os.path.relpath(path, REPO)} "
              f"({os.path.getsize(path)} B)")
        hs.close()


if __name__ == "__main__":
    main()
