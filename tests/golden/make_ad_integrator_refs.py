"""Extracts the reference's stored PRB goldens (mitsuba3/resources/data/tests/integrators/*_ref.exr) that are in the hot
path's scope (triangle-only configs of test_ad_integrators.py) into tests/golden/ref_ad_integrators.npz.

The container has no OpenEXR; the files of the 32x32 configs are PIZ-flagged EXRs whose single 32-scanline chunk the
writer stored uncompressed (the OpenEXR rule: a chunk is written raw when compression does not shrink it), so a minimal
reader suffices: header attributes, one offset-table entry, one chunk = scanlines x channels (alphabetical) x float32.

    python tests/golden/make_ad_integrator_refs.py      # needs /root/reference (this container only)
"""
import os
import struct

import numpy as np

REF = "/root/reference/mitsuba3/resources/data/tests/integrators"
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "ref_ad_integrators.npz")
CONFIGS = ["crop_window"]          # rectangle + diffuse + constant emitter (test_ad_integrators.py:250-277)


def read_exr_raw(path):
    b = open(path, "rb").read()
    magic, version = struct.unpack("<II", b[:8])
    assert magic == 20000630 and (version & 0xFF) == 2 and not (version & 0x200), "single-part scanline EXR expected"
    p, hdr = 8, {}
    while b[p] != 0:
        e = b.index(0, p); name = b[p:e].decode(); p = e + 1
        e = b.index(0, p); p = e + 1
        (sz,) = struct.unpack("<I", b[p:p + 4]); p += 4
        hdr[name] = b[p:p + sz]; p += sz
    p += 1
    x0, y0, x1, y1 = struct.unpack("<4i", hdr["dataWindow"])
    w, h = x1 - x0 + 1, y1 - y0 + 1
    chans, q, ch = [], 0, hdr["channels"]
    while ch[q] != 0:
        e = ch.index(0, q); name = ch[q:e].decode()
        (ptype,) = struct.unpack("<i", ch[e + 1:e + 5]); assert ptype == 2, "float32 channels expected"
        chans.append(name); q = e + 1 + 16
    lines_per_chunk = {0: 1, 1: 1, 2: 1, 3: 16, 4: 32, 5: 16}[hdr["compression"][0]]
    n_chunks = (h + lines_per_chunk - 1) // lines_per_chunk
    offsets = struct.unpack(f"<{n_chunks}Q", b[p:p + 8 * n_chunks])
    img = np.zeros((h, w, len(chans)), np.float32)
    for off in offsets:
        y, size = struct.unpack("<iI", b[off:off + 8])
        rows = min(lines_per_chunk, y1 + 1 - y)
        assert size == rows * w * len(chans) * 4, f"{path}: chunk is compressed ({size} bytes) - not readable without OpenEXR"
        a = np.frombuffer(b, np.float32, rows * w * len(chans), off + 8).reshape(rows, len(chans), w)
        img[y - y0:y - y0 + rows] = a.transpose(0, 2, 1)
    return img[..., [chans.index(c) for c in "RGB"]]


if __name__ == "__main__":
    out = {}
    for name in CONFIGS:
        for kind in ("primal", "fwd"):
            out[f"{name}_{kind}"] = read_exr_raw(os.path.join(REF, f"test_{name}_image_{kind}_ref.exr"))
            print(name, kind, out[f"{name}_{kind}"].shape, float(out[f"{name}_{kind}"].mean()))
    np.savez_compressed(OUT, **out)
    print("wrote", OUT)
