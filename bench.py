#!/usr/bin/env python3
"""bench.py — throughput of the path-tracing hot path on N B200s (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

Workload (config.workload): BASELINE.json configs[1]-class headline — the reference's random-spheres
"cover" scene (`sphere_scene`, src/demo_scene.rs:37-86: 533 spheres, Perlin ground, emissive earth
sphere, glass, metal) at 1920x1080, depth limit 50, `--spp` samples per pixel PER GPU (default 256).
One "step" = one frame: every rank renders its own global sample range of every pixel (scene
replicated, spp partitioned — weak scaling: per-GPU work is fixed as N grows), then ONE NCCL reduce
of the float32 accumulation buffers to rank 0, which divides by the total spp.

metric  = Mpaths/s (paths = depth-0 ray_color calls = W*H*spp, src/main.rs:96); Mrays/s is reported
beside it (rays = world.hit calls, src/main.rs:44).
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

REPO = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, REPO)
sys.path.insert(0, os.path.join(REPO, "tests"))

SCENE = "sphere_scene"
WIDTH, HEIGHT = 1920, 1080
MAX_DEPTH = 50


KERNEL_NAMES = {1: "k_render_mega", 2: "k_wf_traverse + k_wf_shade (CUDA graph)", 3: "k_render_cta"}


SCENE_NOTES = {
    "sphere_scene": "reference cover scene, 533 spheres, demo_scene.rs:37-86",
    "textured": "cover-scene generator with checker ground, Perlin and earthmap spheres (BASELINE configs[1])",
    "cornell_box": "demo_scene.rs:112-148, smoke boxes (BASELINE configs[2])",
    "cornell_box_solid": "Cornell box with solid rotated boxes (BASELINE configs[2] variant)",
    "pbr_ibl": "pbr.rs materials under the newport_loft environment map (BASELINE configs[3])",
    "cover_100k_motion": "cover-scene generator at 99 860 spheres with moving spheres (BASELINE configs[4])",
    "final_scene": "demo_scene.rs:150-221",
    "simple_light_scene": "demo_scene.rs:88-110",
    "test_sphere": "demo_scene.rs:229-244, the scene main() renders",
}


def scene_note(name):
    return SCENE_NOTES.get(name, "extension scene")


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--spp", type=int, default=256, help="samples per pixel per GPU and step")
    ap.add_argument("--scene", default=SCENE)
    ap.add_argument("--width", type=int, default=WIDTH)
    ap.add_argument("--height", type=int, default=HEIGHT)
    ap.add_argument("--mode", type=int, default=0, help="RtRenderMode: 0 auto (= 3), 1 megakernel, 2 wavefront, 3 CTA-local wavefront")
    ap.add_argument("--cpu-seconds", type=float, default=15.0, help="target CPU time of the cpu_baseline sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    return ap.parse_args()


# ---------------------------------------------------------------------------------------------- clocks
class ClockSampler:
    """Samples nvidia-smi SM clocks + throttle reasons DURING the timed region."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.gpu = gpu_index
        self.rows = []
        self.proc = None
        self.thread = None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100", "-i", str(self.gpu)],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except Exception:
            self.proc = None
            return
        def pump():
            for line in self.proc.stdout:
                self.rows.append(line.strip())
        self.thread = threading.Thread(target=pump, daemon=True)
        self.thread.start()

    def stop(self) -> dict:
        if self.proc is not None:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except Exception:
                self.proc.kill()
        sm, smax, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            f = [x.strip() for x in r.split(",")]
            if len(f) < 8:
                continue
            try:
                sm.append(float(f[1]))
                smax.append(float(f[2]))
            except ValueError:
                continue
            for k, nm in enumerate(names):
                if f[4 + k].lower().startswith("active"):
                    reasons.add(nm)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(smax) if smax else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ------------------------------------------------------------------------------------ CPU reference arm
class OracleTimer:
    """Times the oracle (the C++ restatement of the reference's CPU renderer: all host threads, one job
    per image column like src/main.rs:77-80, the reference's own BvhNode) on a bounded, strided sample
    of the frame's columns."""

    def __init__(self, scene_name, width, height, spp, threads=0):
        import oracle_lib
        self.w, self.h, self.spp, self.threads = width, height, spp, threads
        self.ref = oracle_lib.OracleScene.builtin(scene_name, width / height)  # nested graph + reference BVH
        self.ncpu = os.cpu_count() or 1
        stride = max(1, width // max(self.ncpu, 8))
        _, st = self.ref.render(width, height, spp, MAX_DEPTH, threads, 95, 0, width, stride)  # probe
        self.rate = st.paths / st.seconds

    def sample(self, target_seconds):
        n_cols = int(max(self.ncpu, min(self.w, target_seconds * self.rate / (self.h * self.spp))))
        stride = max(1, self.w // n_cols)
        _, st = self.ref.render(self.w, self.h, self.spp, MAX_DEPTH, self.threads, 95, 0, self.w, stride)
        return {
            "mpaths_s": st.paths / st.seconds / 1e6, "mrays_s": st.rays / st.seconds / 1e6, "seconds": st.seconds,
            "paths": int(st.paths), "rays": int(st.rays), "threads": int(st.threads),
            "box_tests_per_ray": st.box_tests / max(1, st.rays), "prim_tests_per_ray": st.prim_tests / max(1, st.rays),
            "sample": f"every {stride}th column of the {self.w}x{self.h} frame ({(self.w + stride - 1) // stride} columns x "
                      f"{self.h} px x {self.spp} spp = {int(st.paths)} paths), depth 50, reference BvhNode, "
                      f"{int(st.threads)} threads",
        }


def run_reference(args, rank):
    """--impl reference: the reference's own CPU implementation of the path (oracle port — no Rust
    toolchain exists here, so there is no oracle/_ref), all host threads, same config/metric.
    Each step is a bounded sample of the frame (CPU throughput is spp-independent: 16 spp, as configs[0])."""
    if rank != 0:
        return
    t0 = time.perf_counter()
    timer = OracleTimer(args.scene, args.width, args.height, 16)
    per_step = max(1.0, min(30.0, 4.0 * args.cpu_seconds / max(1, args.steps)))
    for _ in range(args.warmup):
        timer.sample(1.0)
    vals = [timer.sample(per_step) for _ in range(args.steps)]
    total_paths = sum(v["paths"] for v in vals)
    total_s = sum(v["seconds"] for v in vals)
    value = total_paths / total_s / 1e6
    last = vals[-1]
    line = {
        "impl": "reference", "metric": "Mpaths/s", "value": value, "unit": "Mpaths/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * total_s / max(1, len(vals)),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "mrays_s": sum(v["rays"] for v in vals) / total_s / 1e6,
        "config": {"workload": f"{args.scene} ({scene_note(args.scene)}) {args.width}x{args.height}, depth {MAX_DEPTH}; "
                               "each step = a bounded strided column sample at 16 spp",
                   "scene": args.scene, "width": args.width, "height": args.height, "max_depth": MAX_DEPTH},
        "cpu_baseline": {"value": value, "unit": "Mpaths/s", "cores": last["threads"], "kind": "port",
                         "sample": last["sample"]},
        "e2e": {"value": value, "unit": "Mpaths/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
        "wall_s": time.perf_counter() - t0,
    }
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------- our arm
def run_ours(args, rank, world, local_rank):
    import numpy as np
    import torch
    import torch.distributed as dist
    import ray_tracing_in_one_weekend_b200 as rt

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the backend has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    W, H, SPP = args.width, args.height, args.spp
    spp_total = SPP * world
    hs = rt.HostScene(args.scene, W / H)
    scene = rt.Scene(hs.flat, local_rank)
    info = scene.bvh_info()
    accum = torch.zeros(H * W * 3, dtype=torch.float32, device=dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2
    host_out = torch.empty(H * W * 3, dtype=torch.float32).pin_memory()
    # one explicit (non-default) stream carries torch's ops, our kernels and the timing events alike
    stream = torch.cuda.Stream(dev)
    torch.cuda.synchronize(dev)
    torch.cuda.set_stream(stream)
    assert stream.cuda_stream != 0
    params = scene.make_params(W, H, SPP, sample_begin=rank * SPP, sample_end=(rank + 1) * SPP, spp_total=spp_total,
                               max_depth=MAX_DEPTH, seed=95, mode=args.mode)

    def frame(want_stats=False):
        accum.zero_()
        st = scene.render_device(params, accum.data_ptr(), stream.cuda_stream, want_stats=want_stats)
        if world > 1:
            dist.reduce(accum, dst=0, op=dist.ReduceOp.SUM)
        if rank == 0:
            accum.div_(float(spp_total))
        return st

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    # ---- counters pass (untimed): per-ray work figures for the roofline ---------------------------
    pc = scene.make_params(W, H, 8, sample_begin=rank * SPP, sample_end=rank * SPP + 8, spp_total=8, max_depth=MAX_DEPTH,
                           seed=95, mode=args.mode, flags=rt.RENDER_COUNTERS)
    accum.zero_()
    cst = scene.render_device(pc, accum.data_ptr(), stream.cuda_stream, want_stats=True)

    # ---- warm-up ------------------------------------------------------------------------------------
    for _ in range(max(3, args.warmup)):
        frame()
        flush.fill_(1)
    barrier()

    # ---- timed region: EXACTLY K steps, CUDA events on the launching stream, max over ranks ---------
    sampler = ClockSampler(local_rank)
    sampler.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    kev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    barrier()
    ev0.record(stream)
    for k in range(args.steps):
        accum.zero_()
        kev[k][0].record(stream)
        scene.render_device(params, accum.data_ptr(), stream.cuda_stream, want_stats=False)
        kev[k][1].record(stream)
        if world > 1:
            dist.reduce(accum, dst=0, op=dist.ReduceOp.SUM)
        if rank == 0:
            accum.div_(float(spp_total))
        flush.fill_(1)  # L2 flush between timed iterations
    ev1.record(stream)
    barrier()
    clocks = sampler.stop()
    ms_total = ev0.elapsed_time(ev1)
    ms_kernel = sum(a.elapsed_time(b) for a, b in kev) / args.steps
    t = torch.tensor([ms_total, ms_kernel], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_total, ms_kernel = float(t[0]), float(t[1])

    # stats of one frame (rays per path etc.), untimed
    st = frame(want_stats=True)
    rays = torch.tensor([float(st.rays)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(rays, op=dist.ReduceOp.SUM)
    rays_per_step = float(rays[0])
    paths_per_step = float(W) * H * spp_total

    # ---- e2e: through the public host-buffer API, scene upload + render + reduce + D2H per step ------
    flat = hs.flat.contents
    h2d = (flat.n_prims * 32 + info.n_nodes * info.node_bytes + flat.n_materials * 32 + flat.n_textures * 32 +
           flat.n_perlin * (4096 + 768) + sum(flat.images[i].width * flat.images[i].height * 4 for i in range(flat.n_images)))
    def e2e_step(spp_params):
        sc2 = rt.Scene(hs.flat, local_rank)  # flattened scene (host) -> device: H2D of every table + BVH build
        accum.zero_()
        sc2.render_device(spp_params, accum.data_ptr(), stream.cuda_stream, want_stats=False)
        if world > 1:
            dist.reduce(accum, dst=0, op=dist.ReduceOp.SUM)
        if rank == 0:
            accum.div_(float(spp_total))
            host_out.copy_(accum, non_blocking=True)  # D2H read of the frame
        torch.cuda.synchronize(dev)
        sc2.close()

    # one untimed warm-up of the create/destroy path (the first scene teardown of a process costs ~0.7 s of
    # one-time driver work), at 1 spp
    e2e_step(scene.make_params(W, H, 1, sample_begin=rank, sample_end=rank + 1, spp_total=spp_total,
                               max_depth=MAX_DEPTH, seed=95, mode=args.mode))
    barrier()
    e2e_steps = max(1, min(args.steps, 3))
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        e2e_step(params)
    barrier()
    e2e_s = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(e2e_s, op=dist.ReduceOp.MAX)
    e2e_value = paths_per_step * e2e_steps / float(e2e_s[0]) / 1e6

    if rank == 0:
        value = paths_per_step * args.steps / (ms_total * 1e-3) / 1e6
        mrays = rays_per_step * args.steps / (ms_total * 1e-3) / 1e6
        # ---- roofline of the dominant kernel (k_render_mega): FP32 lane/issue bound ----
        peak_ffma, eff_mhz = rt.measure_fp32_peak(local_rank)
        per_ray = {
            "node_visits": cst.node_visits / max(1, cst.rays), "box_tests": cst.box_tests / max(1, cst.rays),
            "prim_tests": cst.prim_tests / max(1, cst.rays), "p_hit": cst.shade_calls / max(1, cst.rays),
        }
        # algorithmic FP32 ops per ray, constants of SURVEY.md 8(d): C_box 27 per slab test, C_sph 24 per sphere
        # test (+10 when a root is evaluated ~ per hit), C_fill 46 per hit, C_shade per shading class (counted by
        # the RT_RENDER_COUNTERS build: rays per class): sky 11, Diffuse 80, Metal 60, Dielectric 70,
        # Diffuse+Perlin 80 + 7*(8*(3+11)+30) = 1074, other material.rs combinations >= 5 (Emission; lower bound),
        # pbr.rs 170.
        C_CLS = [11.0, 80.0, 60.0, 70.0, 1074.0, 5.0, 170.0]
        cls = [int(x) for x in cst.class_counts][:7]
        if sum(cls) == cst.rays and cst.rays:
            shade_ops = sum((c * (C_CLS[i] + (56.0 if i else 0.0))) for i, c in enumerate(cls)) / cst.rays
            per_ray["class_share"] = [c / cst.rays for c in cls]
        else:  # renderer without class counters: Diffuse-dominated estimate, Perlin not counted (lower bound)
            shade_ops = (10.0 + 46.0 + 80.0) * per_ray["p_hit"]
        f_ray = 27.0 * per_ray["box_tests"] + 24.0 * per_ray["prim_tests"] + shade_ops
        rays_per_launch = rays_per_step / world
        achieved = rays_per_launch * f_ray / (ms_kernel * 1e-3) / 1e12
        peak = peak_ffma / 1e12
        try:
            peaks = json.load(open(os.path.join(REPO, "MEASURED_PEAKS.json")))
            hbm_peak = peaks.get("hbm_gbs")
        except Exception:
            hbm_peak = None
        hbm_bytes = W * H * 12.0 + h2d  # accumulation write + scene read once
        # DRAM traffic of one launch of this workload from the committed ncu --set full capture (null when the
        # workload differs from the captured one)
        traffic = None
        try:
            tj = json.load(open(os.path.join(REPO, "profiles", "r01_cta_traffic.json")))
            m = tj["match"]
            if (m["scene"], m["width"], m["height"], m["spp_per_gpu"], m["mode"]) == (args.scene, W, H, SPP, int(st.mode_used)):
                traffic = tj["dram_bytes_read"] + tj["dram_bytes_write"]
        except Exception:
            traffic = None
        line = {
            "metric": "Mpaths/s", "value": value, "unit": "Mpaths/s", "n_gpus": world, "steps": args.steps,
            "warmup": max(3, args.warmup), "ms_per_step": ms_total / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "mrays_s": mrays, "rays_per_path": rays_per_step / paths_per_step,
            "config": {"workload": f"{args.scene} ({scene_note(args.scene)}) {W}x{H}, {SPP} spp per GPU "
                                   f"({spp_total} spp frame), depth {MAX_DEPTH}, spp-partitioned + one NCCL reduce",
                       "scene": args.scene, "width": W, "height": H, "spp_per_gpu": SPP, "spp_total": spp_total,
                       "max_depth": MAX_DEPTH, "mode": int(st.mode_used), "parallelism": f"spp-split x{world}",
                       "l2": "256 MiB buffer written between timed steps (L2 flush); scene working set is L1/L2-resident by design",
                       "bvh": {"nodes": info.n_nodes, "node_bytes": info.node_bytes, "depth": info.max_depth,
                               "max_leaf": info.max_leaf_prims, "sah": info.sah_cost}},
            "clocks": clocks,
            "e2e": {"value": e2e_value, "unit": "Mpaths/s", "h2d_bytes_per_step": int(h2d),
                    "d2h_bytes_per_step": int(W * H * 12), "steps": e2e_steps,
                    "what": "rt_scene_create (H2D of all scene tables + host BVH build) + render + reduce + divide + D2H of the float frame, per step"},
            "gpu_launches": int(args.steps * st.launches),
            "roofline": {"bound": "fp32-issue", "achieved": achieved, "peak": peak, "unit": "TFLOP/s", "frac": achieved / peak,
                         "traffic": traffic,
                         "note": KERNEL_NAMES.get(int(st.mode_used), "?") + "; achieved = algorithmic FP32 ops per ray (SURVEY 8d constants, counted on this BVH) "
                                 "x rays per launch / CUDA-event kernel time; peak = MEASURED lane-FFMA issue rate of this GPU "
                                 "(rt_measure_fp32_peak, 1 op per lane-slot), of measured",
                         "ops_per_ray": f_ray, "per_ray": per_ray, "ms_kernel": ms_kernel,
                         "peak_effective_sm_mhz": eff_mhz,
                         "hbm": {"algorithmic_bytes_per_launch": hbm_bytes,
                                 "achieved_gbs": hbm_bytes / (ms_kernel * 1e-3) / 1e9, "peak_gbs": hbm_peak,
                                 "frac": (hbm_bytes / (ms_kernel * 1e-3) / 1e9 / hbm_peak) if hbm_peak else None}},
        }
        if world == 1 and not args.no_cpu_baseline:
            cb = OracleTimer(args.scene, W, H, 16).sample(args.cpu_seconds)
            line["cpu_baseline"] = {"value": cb["mpaths_s"], "unit": "Mpaths/s", "cores": cb["threads"], "kind": "port",
                                    "sample": cb["sample"], "mrays_s": cb["mrays_s"],
                                    "box_tests_per_ray": cb["box_tests_per_ray"], "prim_tests_per_ray": cb["prim_tests_per_ray"]}
        print(json.dumps(line), flush=True)
    scene.close()
    hs.close()
    if world > 1:
        dist.destroy_process_group()


def main():
    args = parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank)
        return
    run_ours(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
