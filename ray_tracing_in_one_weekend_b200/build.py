"""Build recipe for librtgpu.so (in-tree, sm_100a only).

    python -m ray_tracing_in_one_weekend_b200.build        # or build_all() from __graft_entry__.build()

Host translation units are compiled with -ffp-contract=off and the CUDA unit with -fmad=false so
that every float operation is rounded individually, in the reference's operation order (see
csrc/rt_device.cuh "NUMERICS CONTRACT").  nvcc cross-compiles without a GPU.
"""
from __future__ import annotations

import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "librtgpu.so")
TEST_LIB = os.path.join(HERE, "librtgpu_test.so")  # the product objects + csrc/test_hooks.cu (rt_test_* entry points)

NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
CXX = os.environ.get("CXX", "g++")

HOST_SRCS = ["scene_api.cpp", "demo_scene.cpp", "bvh_build.cpp", "png_writer.cpp"]
CUDA_SRCS = ["rtgpu.cu", "rt_comm.cu", "rt_static.cu", "gpu_bvh.cu"]
TEST_SRCS = ["test_hooks.cu"]
CTA_FEATURE_SETS = [0, 4, 8, 9, 11, 15]  # cta_inst.cu is compiled once per set (kernels_cta.cuh)
CTA_NO_PBR_SETS = [0, 4]  # ... and once more without the pbr.rs shading class for the sphere-only sets
CTA_FLAT_SETS = [0, 8, 9, 11]  # ... and the tiny-scene builds (warp-uniform primitive loop instead of the BVH traversal)
HEADERS = ["scene_api.hpp", "bvh_build.hpp", "device_types.h", "rt_device.cuh", "kernels.cuh", "kernels_cta.cuh", "gpu_bvh.hpp", "rt_internal.hpp",
           os.path.join("..", "..", "include", "rtgpu.h")]

HOST_FLAGS = ["-std=c++17", "-O2", "-ffp-contract=off", "-fno-fast-math", "-fPIC", "-Wall"]
NVCC_FLAGS = [
    "-std=c++17", "-O3", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo",
    "-fmad=false", "-prec-div=true", "-prec-sqrt=true", "-ftz=false",
    "-Xcompiler", "-fPIC,-ffp-contract=off,-fno-fast-math", "-Xptxas", "-v",
]


def _stale(target: str, deps: list[str]) -> bool:
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps if os.path.exists(d))


def _run(cmd: list[str]) -> str:
    r = subprocess.run(cmd, cwd=CSRC, capture_output=True, text=True)
    if r.returncode != 0:
        sys.stderr.write(r.stdout + r.stderr)
        raise RuntimeError("build failed: " + " ".join(cmd))
    return r.stdout + r.stderr


def build_all(force: bool = False, verbose: bool = False, extra: tuple = (), out_dir: str | None = None) -> str:
    """Builds librtgpu.so (+ librtgpu_test.so).  `extra` nvcc flags and `out_dir` are for development variants
    (tools/build_variant.py): objects and libraries then go to `out_dir` instead of the package directory."""
    hdrs = [os.path.join(CSRC, h) for h in HEADERS]
    obj_dir = out_dir or CSRC
    lib_dir = out_dir or HERE
    os.makedirs(obj_dir, exist_ok=True)
    lib, test_lib = os.path.join(lib_dir, "librtgpu.so"), os.path.join(lib_dir, "librtgpu_test.so")
    nvcc_flags = NVCC_FLAGS + list(extra)
    objs = []
    log = []
    for src in HOST_SRCS:
        obj = os.path.join(obj_dir, src.replace(".cpp", ".o"))
        if force or _stale(obj, [os.path.join(CSRC, src)] + hdrs):
            log.append(_run([CXX] + HOST_FLAGS + ["-c", src, "-o", obj]))
        objs.append(obj)
    jobs = []
    test_objs = []
    for src in CUDA_SRCS + TEST_SRCS:
        obj = os.path.join(obj_dir, src.replace(".cu", ".o"))
        if force or _stale(obj, [os.path.join(CSRC, src)] + hdrs):
            jobs.append([NVCC] + nvcc_flags + ["-c", src, "-o", obj])
        (test_objs if src in TEST_SRCS else objs).append(obj)
    for f in CTA_FEATURE_SETS:
        obj = os.path.join(obj_dir, f"cta_f{f}.o")
        if force or _stale(obj, [os.path.join(CSRC, "cta_inst.cu")] + hdrs):
            jobs.append([NVCC] + nvcc_flags + [f"-DRT_CTA_F={f}", "-c", "cta_inst.cu", "-o", obj])
        objs.append(obj)
    for f in CTA_NO_PBR_SETS:
        obj = os.path.join(obj_dir, f"cta_f{f}_m3f.o")
        if force or _stale(obj, [os.path.join(CSRC, "cta_inst.cu")] + hdrs):
            jobs.append([NVCC] + nvcc_flags + [f"-DRT_CTA_F={f}", "-DRT_CTA_CLS_MASK=0x3f", "-DRT_CTA_SUFFIX=_m3f", "-c", "cta_inst.cu", "-o", obj])
        objs.append(obj)
    for f in CTA_FLAT_SETS:
        obj = os.path.join(obj_dir, f"cta_f{f}_flat.o")
        if force or _stale(obj, [os.path.join(CSRC, "cta_inst.cu")] + hdrs):
            jobs.append([NVCC] + nvcc_flags + [f"-DRT_CTA_F={f}", "-DRT_CTA_FLAT=1", "-DRT_CTA_SUFFIX=_flat", "-c", "cta_inst.cu", "-o", obj])
        objs.append(obj)
    if jobs:  # the CUDA units are independent: compile them side by side
        from concurrent.futures import ThreadPoolExecutor
        with ThreadPoolExecutor(max_workers=min(len(jobs), os.cpu_count() or 4)) as ex:
            log.extend(ex.map(_run, jobs))
    # exports.map: the dynamic symbol table holds the rt_* entry points and nothing else (no C++ internals)
    link = ["-gencode", "arch=compute_100a,code=sm_100a", "-lz", "-ldl", "-Xlinker", "--version-script=" + os.path.join(CSRC, "exports.map")]
    if force or _stale(lib, objs + [os.path.join(CSRC, "exports.map")]):
        log.append(_run([NVCC, "-shared", "-o", lib] + objs + link))
    if force or _stale(test_lib, objs + test_objs + [os.path.join(CSRC, "exports.map")]):
        log.append(_run([NVCC, "-shared", "-o", test_lib] + objs + test_objs + link))
    out = "".join(log)
    if verbose and out:
        print(out)
    return lib


if __name__ == "__main__":
    build_all(force="--force" in sys.argv, verbose=True)
    print(LIB)
