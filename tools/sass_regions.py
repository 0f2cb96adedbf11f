#!/usr/bin/env python3
"""Development tool: static SASS of one kernel grouped by source region.  Reads `nvdisasm -gi -c` output (inline
chains per instruction) and lists the instructions whose inline chain passes through file:line.

    nvdisasm -gi -c cta_inst.sm_100a.cubin > dis.txt
    python tools/sass_regions.py dis.txt 'k_render_ctaILj0ELb0E' rt_device.cuh:907        # one Perlin octave
    python tools/sass_regions.py dis.txt 'k_render_ctaILj0ELb0E' --by-frame kernels_cta.cuh   # counts per line of that file
"""
import re, sys, collections
path, kern = sys.argv[1], sys.argv[2]
want = sys.argv[3:]
on = False
chain = []
new_chain = []
ins = []  # (addr, text, chain)
for l in open(path):
    if l.startswith('.text.') or l.startswith('\t.section\t.text.') or '.section' in l and '.text.' in l:
        on = kern in l
        continue
    if not on:
        continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m:
        new_chain.append((m.group(1).split('/')[-1], int(m.group(2))))
        continue
    m = re.match(r'\s+/\*([0-9a-f]+)\*/\s+(.*?);', l)
    if m:
        if new_chain:
            chain = new_chain
            new_chain = []
        ins.append((m.group(1), m.group(2).strip(), tuple(chain)))
print(f"{len(ins)} instructions in kernel")
if want and want[0] == '--by-frame':
    f = want[1]
    c = collections.Counter()
    for a, t, ch in ins:
        for fl, ln in reversed(ch):
            if fl == f:
                c[ln] += 1
                break
    for ln, n in sorted(c.items()):
        print(f"{f}:{ln}  {n}")
    sys.exit(0)
for w in want:
    f, ln = w.split(':')
    ln = int(ln)
    sel = [(a, t, ch) for a, t, ch in ins if (f, ln) in ch]
    ops = collections.Counter()
    for a, t, ch in sel:
        t2 = re.sub(r'^@!?U?P\d\s+', '', t)
        ops[t2.split()[0].split('.')[0]] += 1
    print(f"== {w}: {len(sel)} instructions: " + ' '.join(f"{k}:{v}" for k, v in ops.most_common()))
    if '-v' in sys.argv:
        for a, t, ch in sel:
            print(f"  /*{a}*/ {t:60s}  {ch[0][0]}:{ch[0][1]}")
