"""Host SAH tree vs device LBVH tree on one scene: python tools/builder_ab.py [scene] [w] [h] [spp]"""
import os, sys, time
sys.path.insert(0, os.getcwd())
import ray_tracing_in_one_weekend_b200 as rt

name = sys.argv[1] if len(sys.argv) > 1 else "cover_100k_motion"
w = int(sys.argv[2]) if len(sys.argv) > 2 else 1920
h = int(sys.argv[3]) if len(sys.argv) > 3 else 1080
spp = int(sys.argv[4]) if len(sys.argv) > 4 else 32
hs = rt.HostScene(name, w / h)
for label, flags in (("device LBVH", rt.api.SCENE_GPU_BVH), ("host SAH", 0)):
    t0 = time.perf_counter()
    sc = rt.Scene(hs.flat, 0, flags=flags)
    t_create = time.perf_counter() - t0
    info = sc.bvh_info()
    best = None
    for i in range(3):
        img, st = sc.render(w, h, spp)
        best = st.ms_kernel if best is None else min(best, st.ms_kernel)
    print(f"{name} {label:12s}: create {t_create * 1e3:8.1f} ms, {info.n_nodes} nodes depth {info.max_depth}, "
          f"{best:9.3f} ms {st.paths / best / 1e3:9.1f} Mpaths/s {st.rays / best / 1e3:9.1f} Mrays/s", flush=True)
    del sc
