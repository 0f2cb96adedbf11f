"""GPU: rt_closest_hit (C ABI, host buffers) against the committed golden vectors of tests/golden/ — hit
primitive id identical on every ray, t / p / n BIT-equal (stated tolerance 1e-5 relative), front_face equal.
Nothing here touches the oracle: the fixtures travel with the repo."""
import glob
import os
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "golden"))
import make_golden  # noqa: E402

FILES = sorted(glob.glob(os.path.join(HERE, "golden", "*.npz")))


@pytest.mark.parametrize("path", FILES, ids=lambda p: os.path.basename(p)[:-4])
def test_cuda_closest_hit_matches_golden(rtlib, path):
    rt = rtlib
    g = np.load(path)
    name, w, h = os.path.basename(path)[:-4], int(g["w"]), int(g["h"])
    hs = rt.HostScene(name, w / h)
    assert make_golden.flat_digest(rt, hs.flat) == str(g["flat_sha256"]), "flattened scene changed"
    o, d, times = make_golden.batch(rt, hs, name, w, h)
    scene = rt.Scene(hs.flat, 0)
    got = scene.closest_hit(o, d, times=times)
    assert np.array_equal(got["prim"], g["prim"])
    hit = g["prim"] >= 0
    for f in ("t", "p", "n"):
        assert np.array_equal(got[f].view(np.uint32)[hit], g[f][hit]), f"{name}: {f} not bit-equal"
        np.testing.assert_allclose(got[f][hit], g[f].view(np.float32)[hit], rtol=1e-5, atol=0)
    assert np.array_equal(got["front_face"].astype(np.uint8)[hit], g["front_face"][hit])
    scene.close()
    hs.close()
