import sys, os
sys.path.insert(0, os.getcwd())
import ray_tracing_in_one_weekend_b200 as rt
w,h,spp,mode = 1920,1080,int(sys.argv[1]),2
hs = rt.HostScene("sphere_scene", w / h)
sc = rt.Scene(hs.flat, 0)
for i in range(2):
    img, st = sc.render(w, h, spp, mode=mode)
print(st.ms_kernel, st.launches)
