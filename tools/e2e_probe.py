"""Where does the end-to-end step spend its time?  (development tool)"""
import os, sys, time
sys.path.insert(0, os.getcwd())
import torch
import ray_tracing_in_one_weekend_b200 as rt
W, H, SPP = 1920, 1080, int(sys.argv[1]) if len(sys.argv) > 1 else 64
hs = rt.HostScene("sphere_scene", W / H)
dev = torch.device("cuda", 0)
accum = torch.zeros(H * W * 3, dtype=torch.float32, device=dev)
host = torch.empty(H * W * 3, dtype=torch.float32).pin_memory()
for it in range(4):
    torch.cuda.synchronize()
    t0 = time.perf_counter(); sc = rt.Scene(hs.flat, 0); t1 = time.perf_counter()
    p = sc.make_params(W, H, SPP)
    accum.zero_()
    sc.render_device(p, accum.data_ptr(), 0, want_stats=False); t2 = time.perf_counter()
    torch.cuda.synchronize(); t3 = time.perf_counter()
    host.copy_(accum, non_blocking=True); torch.cuda.synchronize(); t4 = time.perf_counter()
    sc.close(); t5 = time.perf_counter()
    print(f"iter {it}: create {1e3*(t1-t0):.2f} ms, launch {1e3*(t2-t1):.2f}, kernel wait {1e3*(t3-t2):.2f}, d2h {1e3*(t4-t3):.2f}, close {1e3*(t5-t4):.2f}")
