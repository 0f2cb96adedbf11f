#!/usr/bin/env python3
"""bench.py — throughput of the path-tracing hot path on N B200s (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

Headline workload (config.workload): the reference's random-spheres "cover" scene (`sphere_scene`,
src/demo_scene.rs:37-86: 533 spheres, Perlin ground, emissive earth sphere, glass, metal) at 1920x1080, depth
limit 50, `--spp` samples per pixel PER GPU (default 256).  One "step" = one frame through the C ABI's
rt_render_reduce_device: every rank renders its own global sample range of every pixel (scene replicated, spp
partitioned — WEAK scaling: per-GPU work is fixed as N grows), ONE ncclReduce of the float32 accumulation buffers
to rank 0 inside librtgpu.so, rank 0 divides by the total spp.  torch.distributed only launches: it ferries the
128-byte NCCL id, provides the barriers and the max-over-ranks of the timings; no torch collective touches the frame.

Beside the headline the same JSON line carries
  e2e      the same metric through the HOST-buffer call (rt_scene_create from the flattened host scene = H2D of every
           table, rt_render_reduce into a pinned host image = D2H), >= 10 individually timed steps, median
  strong   one 256-spp frame split N ways (per-GPU work shrinks with N)
  configs  the five BASELINE.json configs, each split over the N GPUs
metric  = Mpaths/s (paths = depth-0 ray_color calls = W*H*spp, src/main.rs:96); Mrays/s is reported
beside it (rays = world.hit calls, src/main.rs:44).
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
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

KERNEL_NAMES = {1: "k_render_mega", 3: "k_render_cta"}

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

# The five BASELINE.json configs: (name, scene, width, height, frame spp).  Every one is split over the N GPUs
# (a STRONG split of the frame's samples).  cfg5 is defined at 4096 spp split over 8 GPUs: with fewer than 8 ranks
# the run renders the 512-spp share per GPU of that split (N x 512 spp frame) and says so.
CONFIGS = [
    ("cfg1_cover_1200x800_16spp", "sphere_scene", 1200, 800, 16),
    ("cfg2_textured_1080p_256spp", "textured", 1920, 1080, 256),
    ("cfg3_cornell_1024_1024spp", "cornell_box", 1024, 1024, 1024),
    ("cfg4_pbr_ibl_1080p_1024spp", "pbr_ibl", 1920, 1080, 1024),
    ("cfg5_cover100k_4k_4096spp", "cover_100k_motion", 3840, 2160, 4096),
]

# algorithmic FP32 ops per ray, constants of SURVEY.md 8(d): C_box 27 per slab test, C_sph 24 per sphere test
# (+10 when a root is evaluated ~ per hit), C_fill 46 per hit, C_shade per shading class (counted by the
# RT_RENDER_COUNTERS build: rays per class): sky 11, Diffuse 80, Metal 60, Dielectric 70, Diffuse+Perlin
# 80 + 7*(8*(3+11)+30) = 1074, other material.rs combinations >= 5 (Emission; lower bound), pbr.rs 170.
C_CLS = [11.0, 80.0, 60.0, 70.0, 1074.0, 5.0, 170.0]


def scene_note(name):
    return SCENE_NOTES.get(name, "extension scene")


def workload_config(scene, w, h, spp, world, l2_note):
    """The `config` block — identical for both arms (`--impl ours` / `--impl reference`)."""
    return {"workload": f"{scene} ({scene_note(scene)}) {w}x{h}, {spp} spp per GPU ({spp * world} spp frame), depth {MAX_DEPTH}, "
                        "samples of every pixel partitioned over the GPUs + one sum-reduce per frame",
            "scene": scene, "width": w, "height": h, "spp_per_gpu": spp, "spp_total": spp * world, "max_depth": MAX_DEPTH,
            "parallelism": f"spp-split x{world}", "l2": l2_note}


L2_NOTE = ("256 MiB buffer written between timed steps (L2 flush) on the GPU arm; the scene working set is L1/L2-resident by design; "
           "the CPU arm has no device cache to flush")


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
    ap.add_argument("--mode", type=int, default=0, help="RtRenderMode: 0 auto (= 3), 1 megakernel, 3 CTA-local wavefront")
    ap.add_argument("--cpu-seconds", type=float, default=15.0, help="target CPU time of the cpu_baseline sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-configs", action="store_true", help="skip the five-config block")
    ap.add_argument("--no-strong", action="store_true", help="skip the strong-scaling block")
    ap.add_argument("--e2e-steps", type=int, default=10)
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
        "config": workload_config(args.scene, args.width, args.height, args.spp, max(1, args.gpus), L2_NOTE),
        "details": {"timed": "each step = a bounded strided column sample of that frame at 16 spp on the host cores (CPU throughput in "
                             "Mpaths/s does not depend on spp); rank 0 only"},
        "cpu_baseline": {"value": value, "unit": "Mpaths/s", "cores": last["threads"], "kind": "port",
                         "sample": last["sample"]},
        "e2e": {"value": value, "unit": "Mpaths/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
        "wall_s": time.perf_counter() - t0,
    }
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------- our arm
def ops_per_ray(cst):
    """Algorithmic FP32 ops per ray from a RT_RENDER_COUNTERS pass (SURVEY.md 8d constants)."""
    rays = max(1, cst.rays)
    per_ray = {"node_visits": cst.node_visits / rays, "box_tests": cst.box_tests / rays,
               "prim_tests": cst.prim_tests / rays, "p_hit": cst.shade_calls / rays}
    cls = [int(x) for x in cst.class_counts][:7]
    if sum(cls) == cst.rays and cst.rays:
        shade_ops = sum((c * (C_CLS[i] + (56.0 if i else 0.0))) for i, c in enumerate(cls)) / rays
        per_ray["class_share"] = [c / rays for c in cls]
    else:  # renderer without class counters: Diffuse-dominated estimate, Perlin not counted (lower bound)
        shade_ops = (10.0 + 46.0 + 80.0) * per_ray["p_hit"]
    return 27.0 * per_ray["box_tests"] + 24.0 * per_ray["prim_tests"] + shade_ops, per_ray


class Rig:
    """One rank's device state: stream, accumulation buffer, L2-flush buffer, communicator."""

    def __init__(self, args, rank, world, local_rank):
        import torch
        import torch.distributed as dist
        import ray_tracing_in_one_weekend_b200 as rt
        from ray_tracing_in_one_weekend_b200 import distributed as rtd
        self.torch, self.dist, self.rt = torch, dist, rt
        self.rank, self.world, self.local_rank = rank, world, local_rank
        self.dev = torch.device("cuda", local_rank)
        if world > 1:
            dist.init_process_group("nccl", device_id=self.dev)
            self.comm = rtd.make_comm(local_rank)          # ncclCommInitRank inside librtgpu.so
        else:
            self.comm = rt.Comm.all_local([local_rank])     # one rank: no collective is issued
        self.flush = torch.empty(256 << 20, dtype=torch.uint8, device=self.dev)  # > 126 MB L2
        # one explicit (non-default) stream carries torch's ops, our kernels, NCCL and the timing events alike
        self.stream = torch.cuda.Stream(self.dev)
        torch.cuda.synchronize(self.dev)
        torch.cuda.set_stream(self.stream)
        assert self.stream.cuda_stream != 0
        self.accum = None

    def accum_for(self, w, h):
        n = w * h * 3
        if self.accum is None or self.accum.numel() < n:
            self.accum = self.torch.zeros(n, dtype=self.torch.float32, device=self.dev)
        return self.accum

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize(self.dev)

    def max_over_ranks(self, vals):
        t = self.torch.tensor(list(vals), dtype=self.torch.float64, device=self.dev)
        if self.world > 1:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return [float(x) for x in t]

    def sum_over_ranks(self, vals):
        t = self.torch.tensor(list(vals), dtype=self.torch.float64, device=self.dev)
        if self.world > 1:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.SUM)
        return [float(x) for x in t]

    def timed_frames(self, scene, frame, w, h, steps, warmup):
        """`warmup` untimed + `steps` timed frames through rt_render_reduce_device, an L2 flush after every frame.
        Returns (ms per step over the whole region, mean render-kernel ms) — both max over ranks, CUDA events."""
        torch = self.torch
        accum = self.accum_for(w, h)
        for _ in range(warmup):
            self.comm.render_reduce_device([scene], frame, [accum.data_ptr()], [self.stream.cuda_stream])
            self.flush.fill_(1)
        self.barrier()
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ev0.record(self.stream)
        for _ in range(steps):
            self.comm.render_reduce_device([scene], frame, [accum.data_ptr()], [self.stream.cuda_stream])
            self.flush.fill_(1)  # L2 flush between timed iterations
        ev1.record(self.stream)
        self.barrier()
        ms_total = ev0.elapsed_time(ev1)
        kt = scene.kernel_times(steps)  # the library's own event pairs around each render kernel, on this stream
        ms_kernel = sum(kt) / max(1, len(kt))
        ms_total, ms_kernel = self.max_over_ranks([ms_total, ms_kernel])
        return ms_total / steps, ms_kernel


def frame_stats(rig, scene, frame):
    """Rays of one frame summed over the ranks (untimed stats pass through rt_render_device on this rank's share)."""
    rt = rig.rt
    b, e = rt.partition_samples(frame.sample_begin, frame.sample_end, rig.world, rig.rank)
    p = rt.RtRenderParams(frame.width, frame.height, frame.spp_total, b, e, frame.max_depth, frame.mode, frame.flags, frame.seed)
    accum = rig.accum_for(frame.width, frame.height)
    st = scene.render_device(p, accum.data_ptr(), rig.stream.cuda_stream, want_stats=True)
    rays, = rig.sum_over_ranks([float(st.rays)])
    return rays, st


def counters_pass(rig, scene, w, h, mode):
    rt = rig.rt
    pc = scene.make_params(w, h, 8, sample_begin=rig.rank * 8, sample_end=rig.rank * 8 + 8, spp_total=8, max_depth=MAX_DEPTH,
                           seed=95, mode=mode, flags=rt.RENDER_COUNTERS)
    accum = rig.accum_for(w, h)
    return scene.render_device(pc, accum.data_ptr(), rig.stream.cuda_stream, want_stats=True)


def run_ours(args, rank, world, local_rank):
    import numpy as np
    import torch
    import ray_tracing_in_one_weekend_b200 as rt

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the backend has no CPU fallback")
    torch.cuda.set_device(local_rank)
    rig = Rig(args, rank, world, local_rank)
    dev = rig.dev

    W, H, SPP = args.width, args.height, args.spp
    spp_total = SPP * world
    hs = rt.HostScene(args.scene, W / H)
    scene = rt.Scene(hs.flat, local_rank)
    info = scene.bvh_info()
    # the WHOLE frame: rt_render_reduce* gives rank r the sample range rt_partition_samples(0, spp_total, world, r)
    frame = scene.make_params(W, H, spp_total, sample_begin=0, sample_end=spp_total, spp_total=spp_total,
                              max_depth=MAX_DEPTH, seed=95, mode=args.mode)

    # ---- counters pass (untimed): per-ray work figures for the roofline ---------------------------
    cst = counters_pass(rig, scene, W, H, args.mode)

    # ---- timed region: W warm-up frames, then EXACTLY K frames, CUDA events on the launching stream, max over ranks
    warm = max(3, args.warmup)
    sampler = ClockSampler(local_rank)
    sampler.start()
    ms_step, ms_kernel = rig.timed_frames(scene, frame, W, H, args.steps, warm)
    clocks = sampler.stop()
    rays_per_step, st = frame_stats(rig, scene, frame)
    paths_per_step = float(W) * H * spp_total

    # ---- strong scaling: ONE 256-spp frame split over the N ranks ------------------------------------
    strong = None
    if not args.no_strong:
        sframe = scene.make_params(W, H, 256, sample_begin=0, sample_end=256, spp_total=256, max_depth=MAX_DEPTH, seed=95, mode=args.mode)
        s_ms, s_kernel = rig.timed_frames(scene, sframe, W, H, max(3, args.steps), 2)
        s_rays, _ = frame_stats(rig, scene, sframe)
        strong = {"spp_frame": 256, "spp_per_gpu": 256 / world, "value": W * H * 256.0 / (s_ms * 1e-3) / 1e6,
                  "mrays_s": s_rays / (s_ms * 1e-3) / 1e6, "ms_per_step": s_ms, "ms_kernel": s_kernel,
                  "ms_reduce_and_divide": max(0.0, s_ms - s_kernel),
                  "what": "one 1920x1080 frame of 256 spp split over the ranks with rt_partition_samples, ncclReduce, divide; "
                          "ms_kernel = slowest rank's render kernel, the rest of ms_per_step is memset + reduce + divide + L2 flush"}

    # ---- e2e: the HOST-buffer call, scene upload + render + reduce + D2H per step, every step timed on its own ----
    host_out = torch.empty(H * W * 3, dtype=torch.float32).pin_memory()
    host_np = host_out.numpy().reshape(H, W, 3)

    def e2e_step():
        sc2 = rt.Scene(hs.flat, local_rank)  # flattened host scene -> device: ONE H2D of every table (+ host BVH build)
        rig.comm.render_reduce([sc2], frame, host_np if rank == 0 else None, want_stats=False)  # ... -> pinned host image
        sc2.close()

    for _ in range(2):  # untimed: first-use costs of the create/destroy path (buffer cache, pinned staging)
        e2e_step()
    e2e_ms = []
    for _ in range(max(1, args.e2e_steps)):
        rig.barrier()
        t0 = time.perf_counter()
        e2e_step()
        t1 = time.perf_counter()
        e2e_ms.append(rig.max_over_ranks([(t1 - t0) * 1e3])[0])
    e2e_med = statistics.median(e2e_ms)
    e2e_value = paths_per_step / (e2e_med * 1e-3) / 1e6

    # ---- the five BASELINE.json configs, each split over the N ranks -----------------------------------
    configs = []
    if not args.no_configs:
        for name, sc_name, cw, ch, cspp in CONFIGS:
            note = None
            if name.startswith("cfg5") and world < 8:
                cspp = 512 * world
                note = f"{world} x 512 spp: the per-GPU share of the config's 8-way split of 4096 spp"
            chs = rt.HostScene(sc_name, cw / ch)
            csc = rt.Scene(chs.flat, local_rank)
            cfr = csc.make_params(cw, ch, cspp, sample_begin=0, sample_end=cspp, spp_total=cspp, max_depth=MAX_DEPTH, seed=95, mode=args.mode)
            ccnt = counters_pass(rig, csc, cw, ch, args.mode)
            wfr = csc.make_params(cw, ch, 8 * world, sample_begin=0, sample_end=8 * world, spp_total=8 * world, max_depth=MAX_DEPTH, seed=95, mode=args.mode)
            rig.timed_frames(csc, wfr, cw, ch, 1, 0)  # warm-up at 8 spp per rank
            c_ms, c_kernel = rig.timed_frames(csc, cfr, cw, ch, 2, 0)
            c_rays, _ = frame_stats(rig, csc, cfr)
            f_ray, _ = ops_per_ray(ccnt)
            configs.append({"name": name, "scene": sc_name, "width": cw, "height": ch, "spp_frame": cspp, "note": note,
                            "value": cw * ch * float(cspp) / (c_ms * 1e-3) / 1e6, "mrays_s": c_rays / (c_ms * 1e-3) / 1e6,
                            "ms_per_step": c_ms, "ms_kernel": c_kernel, "rays_per_path": c_rays / (cw * ch * float(cspp)),
                            "ops_per_ray": f_ray, "_rays_per_launch": c_rays / world})
            csc.close()
            chs.close()

    if rank == 0:
        value = paths_per_step / (ms_step * 1e-3) / 1e6
        mrays = rays_per_step / (ms_step * 1e-3) / 1e6
        # ---- roofline of the dominant kernel: FP32 lane/issue bound ------------------------
        peak_ffma, eff_mhz = rt.measure_fp32_peak(local_rank)
        peak = peak_ffma / 1e12
        f_ray, per_ray = ops_per_ray(cst)
        rays_per_launch = rays_per_step / world
        achieved = rays_per_launch * f_ray / (ms_kernel * 1e-3) / 1e12
        for c in configs:
            c["roofline_frac"] = c.pop("_rays_per_launch") * c["ops_per_ray"] / (c["ms_kernel"] * 1e-3) / 1e12 / peak
        try:
            peaks = json.load(open(os.path.join(REPO, "MEASURED_PEAKS.json")))
            hbm_peak = peaks.get("hbm_gbs")
        except Exception:
            hbm_peak = None
        h2d = int(info.upload_bytes)
        hbm_bytes = W * H * 12.0 + h2d  # accumulation write + scene read once
        # DRAM traffic / issue / L2 figures of one launch of this workload from the committed ncu --set full capture
        # (null when the workload differs from the captured one)
        traffic = issue_pct = l2_pct = lanes = None
        try:
            tj = json.load(open(os.path.join(REPO, "profiles", "cta_ncu_headline.json")))
            m = tj["match"]
            if (m["scene"], m["width"], m["height"], m["spp_per_gpu"], m["mode"]) == (args.scene, W, H, SPP, int(st.mode_used)):
                traffic = tj["dram_bytes_read"] + tj["dram_bytes_write"]
                issue_pct, l2_pct, lanes = tj.get("issue_active_pct"), tj.get("l2_throughput_pct"), tj.get("threads_per_inst")
        except Exception:
            pass
        line = {
            "metric": "Mpaths/s", "value": value, "unit": "Mpaths/s", "n_gpus": world, "steps": args.steps,
            "warmup": warm, "ms_per_step": ms_step, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "mrays_s": mrays, "rays_per_path": rays_per_step / paths_per_step,
            "config": workload_config(args.scene, W, H, SPP, world, L2_NOTE),
            "details": {"mode": int(st.mode_used), "frame_call": "rt_render_reduce_device (spp partition + ncclReduce inside librtgpu.so)",
                        "bvh": {"nodes": info.n_nodes, "node_bytes": info.node_bytes, "depth": info.max_depth,
                                "max_leaf": info.max_leaf_prims, "sah": info.sah_cost}},
            "clocks": clocks,
            "e2e": {"value": e2e_value, "unit": "Mpaths/s", "h2d_bytes_per_step": h2d,
                    "d2h_bytes_per_step": int(W * H * 12), "steps": len(e2e_ms), "ms_median": e2e_med,
                    "ms_min": min(e2e_ms), "ms_max": max(e2e_ms), "ms_all": [round(x, 3) for x in e2e_ms],
                    "what": "per step: rt_scene_create from the flattened HOST scene (one H2D of every table from pinned staging "
                            "+ host BVH build) + rt_render_reduce (render, ncclReduce, divide, D2H into a pinned host image) + "
                            "rt_scene_destroy; every step timed with perf_counter, max over ranks, MEDIAN reported"},
            "gpu_launches": int(args.steps * 2),  # per frame: k_render_cta + k_divide (the NCCL kernel of N > 1 is not ours)
            "roofline": {"bound": "fp32-issue", "achieved": achieved, "peak": peak, "unit": "T lane-op/s", "frac": achieved / peak,
                         "traffic": traffic, "issue_active_pct": issue_pct, "l2_pct": l2_pct, "threads_per_inst": lanes,
                         "note": KERNEL_NAMES.get(int(st.mode_used), "?") + "; achieved = algorithmic FP32 ops per ray (SURVEY 8d constants, counted on this BVH) "
                                 "x rays per launch / CUDA-event kernel time (rt_scene_kernel_times: the library's event pair around every "
                                 "render kernel of the timed region); peak = MEASURED lane-FFMA issue rate of this GPU "
                                 "(rt_measure_fp32_peak, 1 op per lane-slot per cycle), of measured; traffic / issue_active_pct / l2_pct from the "
                                 "committed ncu --set full capture of this workload (profiles/cta_ncu_headline.json)",
                         "ops_per_ray": f_ray, "per_ray": per_ray, "ms_kernel": ms_kernel,
                         "peak_effective_sm_mhz": eff_mhz,
                         "hbm": {"algorithmic_bytes_per_launch": hbm_bytes,
                                 "achieved_gbs": hbm_bytes / (ms_kernel * 1e-3) / 1e9, "peak_gbs": hbm_peak,
                                 "frac": (hbm_bytes / (ms_kernel * 1e-3) / 1e9 / hbm_peak) if hbm_peak else None}},
        }
        if strong is not None:
            line["strong"] = strong
        if configs:
            line["configs"] = configs
        if world == 1 and not args.no_cpu_baseline:
            cb = OracleTimer(args.scene, W, H, 16).sample(args.cpu_seconds)
            line["cpu_baseline"] = {"value": cb["mpaths_s"], "unit": "Mpaths/s", "cores": cb["threads"], "kind": "port",
                                    "sample": cb["sample"], "mrays_s": cb["mrays_s"],
                                    "box_tests_per_ray": cb["box_tests_per_ray"], "prim_tests_per_ray": cb["prim_tests_per_ray"]}
        print(json.dumps(line), flush=True)
    scene.close()
    hs.close()
    rig.comm.close()
    if world > 1:
        rig.dist.barrier()
        rig.dist.destroy_process_group()


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
