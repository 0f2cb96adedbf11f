#!/usr/bin/env python3
"""profiles/r02_bench_{1,2,4,8}gpu.json (bench.py lines) -> profiles/r02_scaling_1_2_4_8.md"""
import json
runs = {n: json.loads(open(f"profiles/r02_bench_{n}gpu.json").read()) for n in (1, 2, 4, 8)}
v1 = runs[1]["value"]
out = ["# Round 2: throughput of every named scene at 1 / 2 / 4 / 8 B200 (builder runs, `bench.py`, 5 timed steps)\n",
       "Raw lines: `profiles/r02_bench_{1,2,4,8}gpu.json`.  Every frame goes through `rt_render_reduce_device` (spp partitioned with",
       "`rt_partition_samples`, one `ncclReduce` inside librtgpu.so, divide on rank 0); CUDA events, max over ranks, L2 flushed between steps.\n",
       "## Cover scene (`sphere_scene`, 1920x1080, depth 50)\n",
       "| GPUs | weak: 256 spp per GPU, Mpaths/s | efficiency | e2e (host-buffer call), Mpaths/s | strong: ONE 256-spp frame split N ways, Mpaths/s | efficiency | ms/frame | slowest rank's kernel ms | memset + reduce + divide ms |",
       "|---|---|---|---|---|---|---|---|---|"]
for n, d in runs.items():
    s = d["strong"]
    out.append(f"| {n} | {d['value']:.0f} | {d['value'] / (n * v1):.4f} | {d['e2e']['value']:.0f} | {s['value']:.0f} | {s['value'] / (n * runs[1]['strong']['value']):.4f} | "
               f"{s['ms_per_step']:.3f} | {s['ms_kernel']:.3f} | {s['ms_reduce_and_divide']:.3f} |")
k1 = runs[1]["strong"]["ms_kernel"]
s8 = runs[8]["strong"]
out += ["",
        f"Strong-scaling limiter at N = 8 (measured, not modelled): the frame takes {s8['ms_per_step']:.2f} ms; an ideal split of the one-GPU kernel is "
        f"{k1 / 8:.2f} ms; the slowest rank's kernel takes {s8['ms_kernel']:.2f} ms (+{(s8['ms_kernel'] / (k1 / 8) - 1) * 100:.1f} %: the tail of the persistent kernel — "
        f"the last pools of a 32-spp share drain with idle blocks — and rank-to-rank spread); memset + ncclReduce of 24.9 MB + divide add "
        f"{s8['ms_reduce_and_divide']:.2f} ms ({s8['ms_reduce_and_divide'] / s8['ms_per_step'] * 100:.1f} %).  The reduce is not what limits the curve.\n",
        "## The five BASELINE.json configs (strong split of each frame over the N GPUs), Mpaths/s (Mrays/s)\n",
        "| config | 1 GPU | 2 GPUs | 4 GPUs | 8 GPUs | 8-GPU speed-up |",
        "|---|---|---|---|---|---|"]
for i, c1 in enumerate(runs[1]["configs"]):
    row = [c1["name"] + f" ({c1['scene']} {c1['width']}x{c1['height']})"]
    for n in (1, 2, 4, 8):
        c = runs[n]["configs"][i]
        row.append(f"{c['value']:.0f} ({c['mrays_s']:.0f})" + (f" @{c['spp_frame']} spp" if c["name"].startswith("cfg5") else ""))
    row.append(f"{runs[8]['configs'][i]['value'] / c1['value']:.2f}x")
    out.append("| " + " | ".join(row) + " |")
out += ["",
        "cfg5 is DEFINED as 4096 spp split over 8 GPUs; with fewer ranks the run renders the 512-spp-per-GPU share (N x 512 spp frame), so its row is a weak",
        "series by construction.  cfg1 is a 15-Mpath frame (4.3 ms on one GPU, under 0.9 ms of kernel per GPU at N = 8): its curve is bounded by the fixed per-frame",
        "cost (launch, memset, reduce, divide: 0.2 ms), not by the renderer.\n"]
open("profiles/r02_scaling_1_2_4_8.md", "w").write("\n".join(out))
print("\n".join(out[6:11]))
