"""Time the three renderers on one scene: python tools/mode_sweep.py [scene] [w] [h] [spp]"""
import os, sys
sys.path.insert(0, os.getcwd())
import ray_tracing_in_one_weekend_b200 as rt

name = sys.argv[1] if len(sys.argv) > 1 else "sphere_scene"
w = int(sys.argv[2]) if len(sys.argv) > 2 else 1920
h = int(sys.argv[3]) if len(sys.argv) > 3 else 1080
spp = int(sys.argv[4]) if len(sys.argv) > 4 else 64
modes = [int(m) for m in (sys.argv[5].split(",") if len(sys.argv) > 5 else "1,3".split(","))]
hs = rt.HostScene(name, w / h)
sc = rt.Scene(hs.flat, 0)
for mode in modes:
    best = None
    for i in range(3):
        img, st = sc.render(w, h, spp, mode=mode)
        best = st.ms_kernel if best is None else min(best, st.ms_kernel)
    print(f"{name} mode {mode} {dict((k, v) for k, v in os.environ.items() if k.startswith('RT_'))}: {best:9.3f} ms "
          f"{st.paths / best / 1e3:9.1f} Mpaths/s {st.rays / best / 1e3:9.1f} Mrays/s launches {st.launches}", flush=True)
