"""Minimal OpenEXR reader for the reference's stored test images (test infrastructure; this container has no OpenEXR).

Single-part scanline files, float32 / float16 channels, compression NONE, ZIPS/ZIP or PIZ.  The PIZ path restates the
published OpenEXR algorithm (ImfPizCompressor / ImfHuf / ImfWav): a 16-bit value bitmap + LUT, a canonical Huffman code
with a run-length symbol, and a 2-D Haar-like wavelet over each channel's half-words.
"""
import struct
import zlib

import numpy as np


# ----------------------------------------------------------------------------------------------- Huffman (ImfHuf.cpp)
class _Bits:
    def __init__(self, data, pos=0):
        self.d, self.p, self.c, self.lc = data, pos, 0, 0

    def get(self, n):
        while self.lc < n:
            self.c = (self.c << 8) | self.d[self.p]; self.p += 1; self.lc += 8
        self.lc -= n
        return (self.c >> self.lc) & ((1 << n) - 1)


def _huf_uncompress(data, n_raw):
    im, iM, _table_len, n_bits, _ = struct.unpack("<5I", data[:20])
    enc_size = (1 << 16) + 1
    lens = np.zeros(enc_size, np.int64)
    br = _Bits(data, 20)
    i = im
    while i <= iM:                                   # hufUnpackEncTable: 6-bit code lengths with zero-run escapes
        l = br.get(6)
        if l == 63:
            i += br.get(8) + 6
        elif l >= 59:
            i += l - 59 + 2
        else:
            lens[i] = l; i += 1
    # hufCanonicalCodeTable: codes are assigned from the longest length down, increasing with the symbol index
    n = np.bincount(lens, minlength=59).astype(np.int64)
    start, c = {}, 0
    for l in range(58, 0, -1):
        nc = (c + int(n[l])) >> 1
        start[l] = c; c = nc
    table = {}
    for sym in np.nonzero(lens)[0]:
        l = int(lens[sym]); table[(l, start[l])] = int(sym); start[l] += 1
    pos = br.p                                       # the bit stream starts at the next whole byte after the table
    bits = np.unpackbits(np.frombuffer(data, np.uint8, (n_bits + 7) // 8, pos))[:n_bits]
    out = np.zeros(n_raw, np.uint16); o = 0; k = 0; code = 0; l = 0; nb = int(n_bits)
    bl = bits.tolist()
    while k < nb:
        code = (code << 1) | bl[k]; k += 1; l += 1
        sym = table.get((l, code))
        if sym is None:
            continue
        if sym == iM:                                # run-length symbol: repeat the previous value `count` (8 bits) times
            cnt = 0
            for _ in range(8):
                cnt = (cnt << 1) | bl[k]; k += 1
            out[o:o + cnt] = out[o - 1]; o += cnt
        else:
            out[o] = sym; o += 1
        code = 0; l = 0
    assert o == n_raw, (o, n_raw)
    return out


# ----------------------------------------------------------------------------------------------- wavelet (ImfWav.cpp)
def _wdec14(l, h):
    ls = l.astype(np.int16).astype(np.int32); hs = h.astype(np.int16).astype(np.int32)
    ai = ls + (hs & 1) + (hs >> 1)
    a = ai.astype(np.int16); b = (ai - hs).astype(np.int16)
    return a.astype(np.uint16), b.astype(np.uint16)


def _wdec16(l, h):
    m = l.astype(np.int32); d = h.astype(np.int32)
    bb = (m - (d >> 1)) & 0xFFFF
    aa = (d + bb - (1 << 15)) & 0xFFFF
    return aa.astype(np.uint16), bb.astype(np.uint16)


def _wav2_decode(a, mx):
    """In-place inverse transform of a [ny, nx] uint16 array."""
    ny, nx = a.shape
    dec = _wdec14 if mx < (1 << 14) else _wdec16
    n = min(nx, ny); p = 1
    while p <= n:
        p <<= 1
    p >>= 1; p2 = p; p >>= 1
    while p >= 1:
        ys = np.arange(0, ny - p2 + 1, p2); xs = np.arange(0, nx - p2 + 1, p2)
        if len(ys) and len(xs):
            Y, X = np.meshgrid(ys, xs, indexing="ij")
            i00, i10 = dec(a[Y, X], a[Y + p, X]); i01, i11 = dec(a[Y, X + p], a[Y + p, X + p])
            a[Y, X], a[Y, X + p] = dec(i00, i01); a[Y + p, X], a[Y + p, X + p] = dec(i10, i11)
        if nx & p and len(ys):                       # odd column left over at this level
            x = (xs[-1] + p2) if len(xs) else 0
            a[ys, x], a[ys + p, x] = dec(a[ys, x], a[ys + p, x])
        if ny & p and len(xs):                       # odd row
            y = (ys[-1] + p2) if len(ys) else 0
            a[y, xs], a[y, xs + p] = dec(a[y, xs], a[y, xs + p])
        p2 = p; p >>= 1


def _piz_chunk(data, nx, ny, sizes):
    """sizes: half-words per pixel of each channel (2 for float32, 1 for half).  Returns the scanline-interleaved uint16 buffer."""
    mn, mxnz = struct.unpack("<HH", data[:4]); p = 4
    bitmap = np.zeros(8192, np.uint8)
    if mn <= mxnz:
        bitmap[mn:mxnz + 1] = np.frombuffer(data, np.uint8, mxnz - mn + 1, p); p += mxnz - mn + 1
    present = np.unpackbits(bitmap, bitorder="little").astype(bool); present[0] = True
    lut = np.zeros(65536, np.uint16); vals = np.nonzero(present)[0]; lut[:len(vals)] = vals
    max_value = len(vals) - 1
    (length,) = struct.unpack("<i", data[p:p + 4]); p += 4
    total = nx * ny * sum(sizes)
    raw = _huf_uncompress(data[p:p + length], total)
    out_rows = [[None] * len(sizes) for _ in range(ny)]
    off = 0
    for c, size in enumerate(sizes):
        block = raw[off:off + nx * ny * size].reshape(ny, nx, size).copy(); off += nx * ny * size
        for j in range(size):
            plane = np.ascontiguousarray(block[:, :, j]); _wav2_decode(plane, max_value); block[:, :, j] = plane
        block = lut[block]
        for y in range(ny):
            out_rows[y][c] = block[y].reshape(-1)
    return out_rows


# ----------------------------------------------------------------------------------------------- file
def read_exr(path):
    """-> (image [H, W, C] float32, channel names)."""
    b = open(path, "rb").read()
    magic, version = struct.unpack("<II", b[:8])
    assert magic == 20000630 and (version & 0xFF) == 2 and not (version & 0x1A00), "single-part scanline EXR expected"
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
        (ptype,) = struct.unpack("<i", ch[e + 1:e + 5]); assert ptype in (1, 2), "half / float channels expected"
        chans.append((name, ptype)); q = e + 1 + 16
    comp = hdr["compression"][0]
    lines = {0: 1, 2: 1, 3: 16, 4: 32}[comp]
    n_chunks = (h + lines - 1) // lines
    offsets = struct.unpack(f"<{n_chunks}Q", b[p:p + 8 * n_chunks])
    sizes = [2 if t == 2 else 1 for _, t in chans]
    img = np.zeros((h, w, len(chans)), np.float32)
    for off in offsets:
        y, size = struct.unpack("<iI", b[off:off + 8])
        rows = min(lines, y1 + 1 - y)
        raw_bytes = rows * w * sum(sizes) * 2
        data = b[off + 8:off + 8 + size]
        if size == raw_bytes or comp == 0:           # stored raw (OpenEXR writes a chunk uncompressed when that is smaller)
            u16 = np.frombuffer(data, np.uint16).reshape(rows, -1)
            per_row = [[u16[r, sum(sizes[:c]) * w:sum(sizes[:c + 1]) * w] for c in range(len(chans))] for r in range(rows)]
        elif comp in (2, 3):
            d = np.frombuffer(zlib.decompress(data), np.uint8).astype(np.int32)
            d = np.cumsum(np.concatenate([d[:1], d[1:] - 128])) & 0xFF         # predictor
            half = (len(d) + 1) // 2; out = np.zeros(len(d), np.uint8); out[0::2] = d[:half]; out[1::2] = d[half:]
            u16 = out.view(np.uint16).reshape(rows, -1)
            per_row = [[u16[r, sum(sizes[:c]) * w:sum(sizes[:c + 1]) * w] for c in range(len(chans))] for r in range(rows)]
        else:
            per_row = _piz_chunk(data, w, rows, sizes)
        for r in range(rows):
            for c, (_, t) in enumerate(chans):
                v = np.ascontiguousarray(per_row[r][c])
                img[y - y0 + r, :, c] = v.view(np.float32) if t == 2 else v.view(np.float16).astype(np.float32)
    return img, [n for n, _ in chans]


def read_rgb(path):
    img, names = read_exr(path)
    return img[..., [names.index(c) for c in "RGB"]]
