#!/usr/bin/env python3
"""Development tool: build a variant of librtgpu.so with extra nvcc flags into variants/<name>/librtgpu.so.
Variants are built with -DRT_DEV_KNOBS: they (and only they) read the RT_* development knobs from the environment,
once per rt_scene_create (csrc/rtgpu.cu dev_knob).

    python tools/build_variant.py t224b4 -DRT_CTA_THREADS=224 -DRT_CTA_MIN_BLOCKS=4
    RTGPU_LIB=variants/t224b4/librtgpu.so python tools/mode_sweep.py ...
"""
import os, re, sys
sys.path.insert(0, os.getcwd())
from ray_tracing_in_one_weekend_b200 import build as b

name, extra = sys.argv[1], sys.argv[2:]
out = os.path.abspath(os.path.join("variants", name))
log_path = os.path.join(out, "build.log")
os.makedirs(out, exist_ok=True)
import io, contextlib
buf = io.StringIO()
with contextlib.redirect_stdout(buf):
    lib = b.build_all(force=True, verbose=True, extra=tuple(["-DRT_DEV_KNOBS"] + extra), out_dir=out)
log = buf.getvalue()
open(log_path, "w").write(log)
# registers / spills of the CTA kernels (ptxas -v)
want = None
for l in log.splitlines():
    if "Compiling entry function" in l and "k_render_cta" in l:
        m = re.search(r"k_render_ctaILj(\d+)ELb(\d)ELj(\d+)E", l)
        want = f"k_render_cta<F={m.group(1)},COUNT={m.group(2)},MASK={m.group(3)}>" if m else l.split("'")[1][:48]
    elif want and "bytes stack frame" in l:
        spill = l.strip()
    elif want and "Used" in l and "registers" in l:
        print(want, "|", locals().get("spill", ""), "|", l.strip())
        want = None
for f in os.listdir(out):
    if f.endswith(".o"):
        os.remove(os.path.join(out, f))
print(lib)
