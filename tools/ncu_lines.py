#!/usr/bin/env python3
"""Top source lines of one profiled launch by warp-stall samples, with the dominant stall reasons.

    python tools/ncu_lines.py gpurun_out/prof.ncu-rep [launch_index=0] [top=25]

Reads the report here (no GPU needed) through `ncu -i ... --page source --csv --print-source cuda`.
"""
import csv
import io
import subprocess
import sys


def main():
    rep = sys.argv[1]
    launch = int(sys.argv[2]) if len(sys.argv) > 2 else 0
    top = int(sys.argv[3]) if len(sys.argv) > 3 else 25
    # launch index -> ncu result id filter (0-based range on import)
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass,cuda"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    lines = []
    fname, func = "?", "?"
    col = None
    cur_launch, seen_files, want_func = -1, set(), "?"
    for r in rows:
        if not r:
            continue
        if r[0] in ("File Name", "File Path"):
            fname = r[1].split("/")[-1]
            continue
        if r[0] == "Function Name":
            # a launch's sections list each file once; a repeated (function, file) pair starts the next launch
            if r[1] != func or fname in seen_files:
                cur_launch += 1
                seen_files = set()
            func = r[1]
            seen_files.add(fname)
            if cur_launch == launch:
                want_func = func
            continue
        if cur_launch != launch:
            continue
        if r[0] == "Line No":
            col = {n: i for i, n in enumerate(r)}
            continue
        if col is None or len(r) < len(col) or not r[0].isdigit():
            continue  # SASS rows under a source line have an empty line number
        try:
            smp = int(r[col["# Samples"]])
            inst = int(r[col["Instructions Executed"]])
            thr = int(r[col["Thread Instructions Executed"]])
        except (ValueError, KeyError):
            continue
        stalls = []
        for n, i in col.items():
            if n.startswith("stall_") and "Not Issued" not in n:
                try:
                    v = int(r[i])
                except ValueError:
                    v = 0
                if v:
                    stalls.append((v, n[6:]))
        stalls.sort(reverse=True)
        lines.append((smp, inst, thr, fname, r[0], r[1].strip()[:70], stalls[:3]))
    tot_s = sum(l[0] for l in lines) or 1
    tot_i = sum(l[1] for l in lines) or 1
    print(f"{want_func[:100]}: {tot_s} samples, {tot_i} warp inst, {sum(l[2] for l in lines) / tot_i:.1f} thr/inst")
    lines.sort(key=lambda l: l[1 if "--by-inst" in sys.argv else 0], reverse=True)
    for smp, inst, thr, f, ln, src, st in lines[:top]:
        ststr = " ".join(f"{n}:{v}" for v, n in st)
        print(f"{smp / tot_s:6.1%} smp {inst / tot_i:6.1%} inst {thr / max(1, inst):5.1f}thr {f}:{ln} | {src} | {ststr}")


if __name__ == "__main__":
    main()
