#!/usr/bin/env python3
"""Summarise an Nsight Compute report (.ncu-rep) into a small markdown file for profiles/.

    python tools/ncu_summary.py gpurun_out/prof.ncu-rep profiles/r01_xxx.md [--launch-list gpurun_out/launches.csv]

Reads the report here (no GPU needed) with `ncu -i ... --page raw/source --csv`, keeps the metrics the
design decisions hinge on (issue-slot utilisation, warp execution efficiency, L1/L2 hit rates, DRAM bytes,
occupancy, stall reasons) and the hottest source lines with their SIMT efficiency.
"""
import csv
import io
import subprocess
import sys

RAW_KEYS = [
    "gpu__time_duration.sum", "launch__registers_per_thread", "launch__occupancy_limit_registers",
    "launch__grid_size", "launch__block_size", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__warps_eligible.avg.per_cycle_active",
    "smsp__inst_executed.sum", "smsp__thread_inst_executed_per_inst_executed.ratio",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "l1tex__t_sector_hit_rate.pct",
    "lts__t_sector_hit_rate.pct", "l1tex__throughput.avg.pct_of_peak_sustained_active",
    "lts__throughput.avg.pct_of_peak_sustained_elapsed", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sass__inst_executed_local_loads",
    "sass__inst_executed_local_stores", "sm__cycles_elapsed.avg.per_second",
]


def ncu_csv(rep, page, extra=()):
    out = subprocess.run(["ncu", "-i", rep, "--page", page, "--csv", *extra], capture_output=True, text=True).stdout
    return list(csv.reader(io.StringIO(out)))


def main():
    rep, dst = sys.argv[1], sys.argv[2]
    launch_list = sys.argv[sys.argv.index("--launch-list") + 1] if "--launch-list" in sys.argv else None
    md = [f"# ncu summary of `{rep}`", ""]
    rows = ncu_csv(rep, "raw")
    hdr, units, data = rows[0], rows[1], rows[2:]
    name_col = hdr.index("Kernel Name")
    md.append("## Raw metrics (one column per profiled launch; `--set full --clock-control none`)")
    md.append("")
    md.append("| metric | unit | " + " | ".join(f"launch {i}" for i in range(len(data))) + " |")
    md.append("|---|---|" + "---|" * len(data))
    md.append("| kernel | | " + " | ".join(r[name_col][:60] for r in data) + " |")
    for i, h in enumerate(hdr):
        if h in RAW_KEYS:
            md.append(f"| {h} | {units[i]} | " + " | ".join(r[i] for r in data) + " |")
    stall_cols = [i for i, h in enumerate(hdr) if h.startswith("smsp__average_warp") and h.endswith("_per_issue_active.ratio")
                  or h.startswith("smsp__average_warps_issue_stalled") and h.endswith("per_issue_active.ratio")]
    if stall_cols:
        md.append("")
        md.append("## Warp stall reasons (cycles per issued instruction)")
        md.append("")
        pairs = sorted(((float(data[0][i] or 0), hdr[i]) for i in stall_cols), reverse=True)[:10]
        for v, h in pairs:
            md.append(f"* {h}: {v:.3f}")
    # source page: hottest lines
    src = ncu_csv(rep, "source", ("--print-source", "sass,cuda"))
    hdr_idx = [i for i, r in enumerate(src) if r and r[0] == "Line No"]
    secs, seen = [], set()
    for hi in hdr_idx:
        f = src[hi - 2][1]
        if f in seen:
            break
        seen.add(f)
        secs.append(hi)
    ends = secs[1:] + [hdr_idx[len(secs)] if len(hdr_idx) > len(secs) else len(src)]
    lines, tot, tot_thr = [], 0, 0
    for hi, e in zip(secs, ends):
        f = src[hi - 2][1].split("/")[-1]
        col = {n: i for i, n in enumerate(src[hi])}
        for r in src[hi + 1:e - 2]:
            if len(r) < 12 or r[0] == "":
                continue
            try:
                inst = int(r[col["Instructions Executed"]])
                thr = int(r[col["Thread Instructions Executed"]])
                smp = int(r[col["# Samples"]])
            except ValueError:
                continue
            lines.append((inst, thr, smp, f, r[0], r[1].strip()[:80]))
            tot += inst
            tot_thr += thr
    if tot:
        lines.sort(reverse=True)
        md += ["", "## Hottest source lines of the first profiled launch (needs -lineinfo)", "",
               f"total warp instructions {tot}, average active threads per instruction {tot_thr / tot:.2f} / 32", "",
               "| % warp inst | threads/inst | stall samples | where | source |", "|---|---|---|---|---|"]
        for inst, thr, smp, f, ln, s in lines[:40]:
            md.append(f"| {inst / tot:.2%} | {thr / max(1, inst):.1f} | {smp} | {f}:{ln} | `{s.replace('|', '/')}` |")
    if launch_list:
        md += ["", "## Launch list (`--metrics gpu__time_duration.sum`, cold-cache, serialised: compare SHARES)", ""]
        tot_t, by = 0.0, {}
        for r in csv.reader(open(launch_list)):
            if len(r) < 15 or r[0] == "ID":
                continue
            try:
                t = float(r[14])
            except ValueError:
                continue
            k = r[4].split("(")[0][:70]
            by.setdefault(k, [0, 0.0])
            by[k][0] += 1
            by[k][1] += t
            tot_t += t
        md += ["| kernel | launches | total ms | share |", "|---|---|---|---|"]
        for k, (n, t) in sorted(by.items(), key=lambda kv: -kv[1][1]):
            md.append(f"| `{k}` | {n} | {t / 1e6:.3f} | {t / tot_t:.1%} |")
    open(dst, "w").write("\n".join(md) + "\n")
    print("wrote", dst)


if __name__ == "__main__":
    main()
