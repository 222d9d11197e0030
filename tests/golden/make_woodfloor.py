"""Decodes the reference's WoodFloor.jpg (the texture BASELINE.json config 3 names) into a committed fixture.

The GPU box has no /root/reference, so the decoded 8-bit pixels travel as mitsuba3_experiments_b200/data/woodfloor_u8.npz
(benchmark input data, read by scenes.woodfloor_texture); the host layer
applies Mitsuba's sRGB -> linear conversion (bitmap.cpp:286-294) when it builds the fp32 parameter tensor.
Run in the build container:  python tests/golden/make_woodfloor.py
"""
import os
import numpy as np
from PIL import Image

src = "/root/reference/WoodFloor.jpg"
img = np.asarray(Image.open(src).convert("RGB"), dtype=np.uint8)
out = os.path.join(os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))), "mitsuba3_experiments_b200", "data", "woodfloor_u8.npz")
np.savez_compressed(out, rgb=img)
print(img.shape, img.dtype, os.path.getsize(out), "bytes")
