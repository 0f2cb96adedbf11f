#!/usr/bin/env python3
"""Split one profiled kernel into phases at its BAR.SYNC instructions (SASS layout order) and print, per
phase, the share of warp instructions, thread instructions, active threads per instruction and stall samples.

    python tools/ncu_phases.py gpurun_out/prof.ncu-rep
"""
import csv, io, subprocess, sys

out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "source", "--csv", "--print-source", "sass"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr = rows[1]
c = {n: i for i, n in enumerate(hdr)}
ph = []
cur = {"start": 0, "inst": 0, "thr": 0, "smp": 0, "n": 0, "ops": {}}
tot_i = tot_s = 0
for k, r in enumerate(rows[2:]):
    if len(r) < len(hdr):
        continue
    src = r[c["Source"]].strip()
    inst, thr, smp = int(r[c["Instructions Executed"]]), int(r[c["Thread Instructions Executed"]]), int(r[c["# Samples"]])
    cur["inst"] += inst; cur["thr"] += thr; cur["smp"] += smp; cur["n"] += 1
    op = src.split()[0] if not src.startswith("@") else src.split()[1]
    op = op.split(".")[0]
    cur["ops"][op] = cur["ops"].get(op, 0) + inst
    tot_i += inst; tot_s += smp
    if "BAR.SYNC" in src or src.startswith("EXIT"):
        cur["end"] = k
        ph.append(cur)
        cur = {"start": k + 1, "inst": 0, "thr": 0, "smp": 0, "n": 0, "ops": {}}
ph.append(cur)
print(f"total warp inst {tot_i}, samples {tot_s}")
for p in ph:
    if p["inst"] == 0:
        continue
    top = sorted(p["ops"].items(), key=lambda x: -x[1])[:8]
    print(f"sass[{p['start']:5d}..{p.get('end', -1):5d}] n={p['n']:5d}  inst {100 * p['inst'] / tot_i:5.1f}%  thr/inst {p['thr'] / p['inst']:5.1f}  "
          f"samples {100 * p['smp'] / max(1, tot_s):5.1f}%  " + " ".join(f"{o}:{100 * v / p['inst']:.0f}%" for o, v in top))
