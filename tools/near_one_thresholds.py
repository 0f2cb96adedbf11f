#!/usr/bin/env python3
"""Derives the squared-length interval used by vec3a_near_one (csrc/rt_device.cuh): every float32 x around 1 is
tested with the reference's predicate |sqrt(x) - 1| < 1e-6 (math.rs:13-15) in float32 arithmetic; the passing set
must be contiguous, and its end points are printed as bit patterns."""
import numpy as np

lo = np.float32(1 - 1e-5).view(np.uint32)
hi = np.float32(1 + 1e-5).view(np.uint32)
bits = np.arange(int(lo), int(hi) + 1, dtype=np.uint32)
x = bits.view(np.float32)
ok = np.abs(np.sqrt(x) - np.float32(1.0)) < np.float32(1e-6)
idx = np.nonzero(ok)[0]
assert np.all(np.diff(idx) == 1), "passing set is not an interval"
assert not ok[0] and not ok[-1]
print(hex(bits[idx[0]]), hex(bits[idx[-1]]))
