"""CPU: the development tools that DESIGN.md quotes numbers from still build and still say what DESIGN.md says.

tools/bvh_sim.cpp = traversal statistics of the host SAH builder's tree + the SIMT scheduling model of the traverse
phase ("last pass"): node visits per ray of the cover scene pin the builder's tree quality without
a GPU (the GPU's RT_RENDER_COUNTERS figure for the same tree is 5.20 node visits and 19.8 box tests per ray)."""
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "ray_tracing_in_one_weekend_b200")


def test_bvh_sim_builds_and_reports_tree_quality(tmp_path, rtlib):
    exe = str(tmp_path / "bvh_sim")
    cmd = ["g++", "-O2", "-std=c++17", "-I.", "tools/bvh_sim.cpp", "-o", exe, "-L" + PKG, "-lrtgpu", "-Wl,-rpath," + PKG]
    r = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-2000:]
    r = subprocess.run([exe, "sphere_scene", "240", "135", "1"], cwd=ROOT, capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stderr[-2000:]
    out = r.stdout
    m = re.search(r"(\d+) world prims in the tree, (\d+) giants, (\d+) nodes, depth (\d+)", out)
    assert m, out[:500]
    prims, giants, nodes, depth = map(int, m.groups())
    assert prims == 532 and giants == 1            # the r = 1000 ground sphere is tested before the tree
    assert nodes <= 300 and depth <= 7
    m = re.search(r"^all\s+rays\s+\d+\s+visits/ray\s+([\d.]+)\s+boxes/ray\s+([\d.]+)", out, re.M)
    assert m, out[:800]
    visits, boxes = float(m.group(1)), float(m.group(2))
    assert 4.6 < visits < 5.6 and 17.5 < boxes < 21.5, (visits, boxes)
    # the scheduling model: topping a warp's idle lanes up with camera rays costs fewer issue slots than the two lists
    sep = re.search(r"^A2 refill, then A1 coherent\s+[\d.]+\s+[\d.]+\s+[\d.]+\s+[\d.]+%\s+[\d.]+%\s+([\d.]+)", out, re.M)
    uni = re.search(r"^one list, refill throughout\s+[\d.]+\s+[\d.]+\s+[\d.]+\s+[\d.]+%\s+[\d.]+%\s+([\d.]+)", out, re.M)
    assert sep and uni, out[-1500:]
    assert float(uni.group(1)) < float(sep.group(1))


def _res_usage(obj, kernel_substr):
    out = subprocess.run(["cuobjdump", "-res-usage", os.path.join(PKG, "csrc", obj)], capture_output=True, text=True).stdout
    lines = out.splitlines()
    for i, l in enumerate(lines):
        if "Function" in l and kernel_substr in l:
            return {k: int(v) for k, v in re.findall(r"(REG|STACK|SHARED):(\d+)", lines[i + 1])}
    raise AssertionError(f"{kernel_substr} not found in {obj}")


def _sass(obj, kernel_substr):
    out = subprocess.run(["cuobjdump", "-sass", os.path.join(PKG, "csrc", obj)], capture_output=True, text=True).stdout
    keep, on = [], False
    for l in out.splitlines():
        if "Function :" in l:
            on = kernel_substr in l
        elif on and re.match(r"\s+/\*[0-9a-f]{4,5}\*/", l):
            keep.append(l)
    assert keep, f"{kernel_substr} not found in {obj}"
    return keep


def test_render_kernels_keep_their_register_budgets(rtlib):
    """The block shapes in kernels_cta.cuh are chosen for these register counts (two blocks per SM): 448 threads x 72
    registers for the sphere-only builds without the pbr.rs class, 384 x 80 for the others and the flat builds,
    320 x 96 for the media sets.  A build that needs more would silently drop to one block per SM."""
    for obj, kern, regs, stack in (("cta_f0_m3f.o", "ILj0ELb0ELj63ELb0E", 72, 640), ("cta_f4_m3f.o", "ILj4ELb0ELj63ELb0E", 72, 640),
                                   ("cta_f0.o", "ILj0ELb0ELj127ELb0E", 80, 640), ("cta_f0_flat.o", "ILj0ELb0ELj127ELb1E", 80, 128),
                                   ("cta_f11_flat.o", "ILj11ELb0ELj127ELb1E", 80, 128), ("cta_f11.o", "ILj11ELb0ELj127ELb0E", 96, 640),
                                   ("cta_f15.o", "ILj15ELb0ELj127ELb0E", 96, 640)):
        u = _res_usage(obj, kern)
        assert u["REG"] <= regs and u["STACK"] <= stack, (obj, u)


def test_headline_kernel_codegen(rtlib):
    """SASS properties DESIGN.md section 5 relies on, for the headline instantiation (spheres only, no pbr.rs class):
    three-input min/max in the slab tests, 128-bit node loads through the read-only path, and one-lane counter takes
    that stay ONE shared-memory atomic (ptxas expands a provably warp-uniform address into a 17-instruction
    vote / popcount / shuffle sequence: kernels_cta.cuh warp_take)."""
    sass = _sass("cta_f0_m3f.o", "ILj0ELb0ELj63ELb0E")
    n = lambda pat: sum(1 for l in sass if re.search(pat, l))
    assert 7000 < len(sass) < 9000, len(sass)
    assert n(r"FMNMX3") >= 16 and n(r"LDG\.E\.128\.CONSTANT") >= 14
    assert n(r"ATOMS") >= 9
    # aggregation sequences left: the per-lane appends to the todo / dead lists (wanted there), none for the takes
    assert n(r"VOTEU\.ANY") <= 12, n(r"VOTEU\.ANY")
    assert n(r"\bHMMA|\bWGMMA|UTCMMA") == 0   # nothing on this path is a dense contraction
