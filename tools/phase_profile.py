"""Phase timing of k_render_cta from a -DRT_CTA_PROFILE=1 build (tools/build_variant.py prof -DRT_CTA_PROFILE=1):
RTGPU_LIB=variants/prof/librtgpu.so python tools/phase_profile.py [scene] [spp]"""
import os, sys
sys.path.insert(0, os.getcwd())
import ray_tracing_in_one_weekend_b200 as rt
name = sys.argv[1] if len(sys.argv) > 1 else "sphere_scene"
spp = int(sys.argv[2]) if len(sys.argv) > 2 else 32
w, h = 1920, 1080
sc = rt.Scene.from_name(name, w / h, 0)
for _ in range(2):
    img, st = sc.render(w, h, spp, mode=3)
c = list(st.class_counts)
blocks, warps = 296, 12
regen, trav, shade, iters, wt, ws = c[0], c[1], c[2], c[3], c[4], c[5]
tot = regen + trav + shade
print(f"{name}: kernel {st.ms_kernel:.2f} ms, {iters / blocks:.0f} iterations/block, {tot / iters:.0f} cycles/iteration")
print(f"  phase share of block time: D+regen {regen / tot:.1%}, traverse {trav / tot:.1%}, shade {shade / tot:.1%}")
print(f"  warp utilisation inside traverse {wt / (warps * trav):.1%}, inside shade {ws / (warps * shade):.1%} (own work / phase incl. barrier wait)")
