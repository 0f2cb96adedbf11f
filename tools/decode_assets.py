#!/usr/bin/env python3
"""Decode the reference's two JPEG assets ONCE into raw RGB8 texel arrays (binary PPM, P6).

Why: the reference loads `res/earthmap.jpg` (demo_scene.rs:42,160) and `res/newport_loft.jpg`
(main.rs:63) through image 0.24.3 + jpeg-decoder 0.2.6 and converts to f32 = u8/255
(texture.rs:175-181).  JPEG decoders differ by +-1..2 LSB (SURVEY.md section 7 "hard parts"), so
the texel arrays are decoded once here (Pillow / libjpeg-turbo) and the SAME bytes are fed to the
CPU oracle and to the CUDA backend.  The GPU box has no /root/reference, so the decoded arrays
are committed under assets/.

Run in the build container only:  python tools/decode_assets.py
"""
import os
import sys

import numpy as np
from PIL import Image

REF = "/root/reference/res"
OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "assets")


def main() -> int:
    os.makedirs(OUT, exist_ok=True)
    for name in ("earthmap", "newport_loft"):
        im = Image.open(os.path.join(REF, name + ".jpg")).convert("RGB")
        a = np.asarray(im, dtype=np.uint8)
        with open(os.path.join(OUT, name + ".ppm"), "wb") as fh:
            fh.write(b"P6\n%d %d\n255\n" % (a.shape[1], a.shape[0]))
            fh.write(a.tobytes())
        print(name, a.shape, "mean", float(a.mean()))
    return 0


if __name__ == "__main__":
    sys.exit(main())
