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
