#!/usr/bin/env python3
"""Development tool: build a variant of librtgpu.so with extra nvcc flags into variants/<name>/librtgpu.so.

    python tools/build_variant.py t224b4 -DRT_CTA_THREADS=224 -DRT_CTA_MIN_BLOCKS=4
    RTGPU_LIB=variants/t224b4/librtgpu.so python tools/mode_sweep.py ...
"""
import os, subprocess, sys
sys.path.insert(0, os.getcwd())
from ray_tracing_in_one_weekend_b200 import build as b

name, extra = sys.argv[1], sys.argv[2:]
b.build_all()
out = os.path.join("variants", name)
os.makedirs(out, exist_ok=True)
obj = os.path.join(os.path.abspath(out), "rtgpu.o")
r = subprocess.run([b.NVCC] + b.NVCC_FLAGS + extra + ["-c", "rtgpu.cu", "-o", obj], cwd=b.CSRC, capture_output=True, text=True)
if r.returncode:
    sys.stderr.write(r.stderr)
    raise SystemExit(1)
for l in (r.stdout + r.stderr).splitlines():
    if "k_render_cta" in l and "Compiling" in l:
        want = True
        print(l.split("'")[1][:40], end=" ")
    elif "registers" in l and locals().get("want"):
        print(l.strip()); want = False
    elif "spill" in l and locals().get("want"):
        print(l.strip(), end=" | ")
objs = [os.path.join(b.CSRC, s.replace(".cpp", ".o")) for s in b.HOST_SRCS] + [obj, os.path.join(b.CSRC, "gpu_bvh.o")]
for f in b.CTA_FEATURE_SETS:  # the CTA renderer lives in its own translation units
    o = os.path.join(os.path.abspath(out), f"cta_f{f}.o")
    subprocess.check_call([b.NVCC] + [x for x in b.NVCC_FLAGS if x not in ("-Xptxas", "-v")] + extra + [f"-DRT_CTA_F={f}", "-c", "cta_inst.cu", "-o", o], cwd=b.CSRC)
    objs.append(o)
for f in b.CTA_NO_PBR_SETS:
    o = os.path.join(os.path.abspath(out), f"cta_f{f}_m3f.o")
    subprocess.check_call([b.NVCC] + [x for x in b.NVCC_FLAGS if x not in ("-Xptxas", "-v")] + extra + [f"-DRT_CTA_F={f}", "-DRT_CTA_CLS_MASK=0x3f", "-DRT_CTA_SUFFIX=_m3f", "-c", "cta_inst.cu", "-o", o], cwd=b.CSRC)
    objs.append(o)
subprocess.check_call([b.NVCC, "-shared", "-o", os.path.join(out, "librtgpu.so")] + objs + ["-gencode", "arch=compute_100a,code=sm_100a", "-lz"])
os.remove(obj)
print(os.path.join(out, "librtgpu.so"))
