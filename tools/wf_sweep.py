#!/usr/bin/env python3
"""Tuning sweep of the renderer's environment knobs on one GPU (development tool, not a bench).

    python tools/wf_sweep.py "RT_WF_POOL=262144,524288 RT_WF_REFILL=16,24" [--mode 2] [--spp 64] [--scene sphere_scene]

Each combination runs in its own process (the knobs are read when the pool / graph is built) and prints
kernel ms, Mpaths/s and Mrays/s of the best of 3 frames.
"""
import itertools
import os
import subprocess
import sys

CHILD = r"""
import sys, os
sys.path.insert(0, os.getcwd())
import ray_tracing_in_one_weekend_b200 as rt
scene_name, w, h, spp, mode = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4]), int(sys.argv[5])
hs = rt.HostScene(scene_name, w / h)
sc = rt.Scene(hs.flat, 0)
best = None
for i in range(4):
    img, st = sc.render(w, h, spp, mode=mode)
    if i and (best is None or st.ms_kernel < best.ms_kernel):
        best = st
print(f"{best.ms_kernel:9.3f} ms  {best.paths / best.ms_kernel / 1e3:9.1f} Mpaths/s  {best.rays / best.ms_kernel / 1e3:9.1f} Mrays/s  launches {best.launches}")
"""


def main():
    spec = sys.argv[1] if len(sys.argv) > 1 else ""
    args = sys.argv[2:]

    def opt(name, dflt):
        return args[args.index(name) + 1] if name in args else dflt

    mode, spp, scene = opt("--mode", "2"), opt("--spp", "64"), opt("--scene", "sphere_scene")
    w, h = opt("--width", "1920"), opt("--height", "1080")
    keys, vals = [], []
    for tok in spec.split():
        k, v = tok.split("=")
        keys.append(k)
        vals.append(v.split(","))
    for combo in itertools.product(*vals):
        env = dict(os.environ)
        env.update(dict(zip(keys, combo)))
        r = subprocess.run([sys.executable, "-c", CHILD, scene, w, h, spp, mode], env=env, capture_output=True, text=True)
        tag = " ".join(f"{k}={v}" for k, v in zip(keys, combo))
        print(f"{tag:60s} {r.stdout.strip() or r.stderr.strip()[-300:]}", flush=True)


if __name__ == "__main__":
    main()
