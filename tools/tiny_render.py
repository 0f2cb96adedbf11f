"""Smallest useful invocation of every renderer (for compute-sanitizer): python tools/tiny_render.py [scene] [mode]"""
import os, sys
sys.path.insert(0, os.getcwd())
import numpy as np
import ray_tracing_in_one_weekend_b200 as rt
name = sys.argv[1] if len(sys.argv) > 1 else "sphere_scene"
mode = int(sys.argv[2]) if len(sys.argv) > 2 else 3
sc = rt.Scene.from_name(name, 1.5, 0)
img, st = sc.render(96, 64, 4, mode=mode)
print(name, "mode", mode, "paths", st.paths, "rays", st.rays, "mean", float(np.nanmean(img)))
sc.close()
