"""A/B of the per-path generator's output function against the oracle: RMSE(gpu, oracle A) / RMSE(oracle A, oracle B)
per channel (clipped at the oracle's 99.5th percentile, as tests/test_gpu_render.py does) and the mean-luminance
difference, for several GPU seeds.  RTGPU_LIB selects the library variant.
    RTGPU_LIB=variants/pcg/librtgpu.so python tools/rng_ab.py pbr_ibl 96 54 512 8"""
import os, sys
sys.path.insert(0, os.getcwd()); sys.path.insert(0, os.path.join(os.getcwd(), "tests"))
import numpy as np
import oracle_lib
import ray_tracing_in_one_weekend_b200 as rt
name, w, h, spp, n_seeds = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4]), int(sys.argv[5])
LUM = np.array([0.2126, 0.7152, 0.0722], np.float32)
lum = lambda a: float(np.nanmean((a * LUM).sum(-1)))
hs = rt.HostScene(name, w / h)
ref = oracle_lib.OracleScene.from_flat(hs.flat)
a, _ = ref.render(w, h, spp, seed_base=95)
b, _ = ref.render(w, h, spp, seed_base=100095)
cap = float(np.nanpercentile(a, 99.5))
ca, cb = np.clip(a, 0, cap), np.clip(b, 0, cap)
noise = np.sqrt(np.nanmean((ca - cb) ** 2, axis=(0, 1)))
sc = rt.Scene(hs.flat, 0)
rows = []
for seed in [95] + list(range(1, n_seeds)):
    img, _ = sc.render(w, h, spp, seed=seed)
    r = np.sqrt(np.nanmean((np.clip(img, 0, cap) - ca) ** 2, axis=(0, 1))) / noise
    rows.append((seed, r, abs(lum(img) - lum(a)) / lum(a)))
    print(f"{os.environ.get('RTGPU_LIB', 'default'):>28} {name} seed {seed:3d}: rmse/noise {np.round(r, 3)}  max {r.max():.3f}  dlum {rows[-1][2]:.3%}", flush=True)
m = np.array([r for _, r, _ in rows])
print(f"  -> mean ratio {np.round(m.mean(0), 3)}, max over seeds {m.max():.3f}, seeds above 1.25: {(m.max(1) > 1.25).sum()} of {len(rows)}")
