"""Generate tests/golden/ref_pin.npz from the REFERENCE's own headers (oracle/_ref/libref_pin.so).

Run in the build container only (needs /root/reference): `make -C oracle ref && python tests/golden/make_golden.py`.
The fixture travels with the repo; tests never read /root/reference.
Inputs are drawn from numpy's PCG64 with fixed seeds; outputs are what the reference code computes in scalar
float mode (Float = float), i.e. the arithmetic llvm_ad_rgb executes per lane.
"""
import ctypes as C
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
r = C.CDLL(os.path.join(ROOT, "oracle", "_ref", "libref_pin.so"))
vp = C.c_void_p
r.ref_tea32.argtypes = [C.c_uint32, C.c_uint32, C.c_int, vp]
r.ref_tea_float32.argtypes = [C.c_uint32, C.c_uint32, C.c_int]; r.ref_tea_float32.restype = C.c_float
r.ref_pcg32_floats.argtypes = [C.c_uint64, C.c_uint64, C.c_int, vp]
r.ref_pcg32_default_floats.argtypes = [C.c_int, vp]
r.ref_warp.argtypes = [C.c_int, vp, vp]
r.ref_fresnel.argtypes = [C.c_float, C.c_float, vp]
r.ref_fresnel_conductor.argtypes = [C.c_float] * 3; r.ref_fresnel_conductor.restype = C.c_float
r.ref_microfacet.argtypes = [C.c_int, C.c_float, C.c_float, C.c_int, C.c_int, vp, vp]
r.ref_h2d_create.argtypes = [vp, C.c_uint32, C.c_uint32]; r.ref_h2d_create.restype = vp
r.ref_h2d_sample.argtypes = [vp, C.c_float, C.c_float, vp]
r.ref_h2d_eval.argtypes = [vp, C.c_float, C.c_float]; r.ref_h2d_eval.restype = C.c_float


def p(a):
    return a.ctypes.data_as(vp)


def main():
    rng = np.random.default_rng(1234)
    out = {}
    # --- TEA / PCG32 (integer: bit exact)
    v0 = rng.integers(0, 2 ** 32, 256, dtype=np.uint64).astype(np.uint32); v1 = rng.integers(0, 2 ** 32, 256, dtype=np.uint64).astype(np.uint32)
    tea = np.zeros((256, 2), np.uint32)
    for i in range(256):
        r.ref_tea32(int(v0[i]), int(v1[i]), 4, p(tea[i]))
    out.update(tea_in0=v0, tea_in1=v1, tea_out=tea)
    out["tea_float"] = np.array([r.ref_tea_float32(1, i, 4) for i in range(1, 6)] + [r.ref_tea_float32(i, 1, 4) for i in (2, 3, 4)], np.float32)
    st = rng.integers(0, 2 ** 63, 16, dtype=np.uint64); sq = rng.integers(0, 2 ** 63, 16, dtype=np.uint64)
    pf = np.zeros((16, 32), np.float32)
    for i in range(16):
        r.ref_pcg32_floats(int(st[i]), int(sq[i]), 32, p(pf[i]))
    d = np.zeros(8, np.float32); r.ref_pcg32_default_floats(8, p(d))
    out.update(pcg_state=st, pcg_seq=sq, pcg_floats=pf, pcg_default=d)
    # --- warps
    n = 512
    u = rng.random((n, 2), dtype=np.float32)
    for which, name, ni, no in [(0, "disk", 2, 2), (1, "coshemi", 2, 3), (2, "tri", 2, 2)]:
        o = np.zeros((n, no), np.float32)
        for i in range(n):
            r.ref_warp(which, p(u[i]), p(o[i]))
        out["warp_" + name] = o
    out["warp_u"] = u
    itl_in = np.concatenate([rng.random((n, 2), dtype=np.float32) * 4, rng.random((n, 1), dtype=np.float32)], axis=1).astype(np.float32)
    itl = np.zeros(n, np.float32)
    for i in range(n):
        o = np.zeros(1, np.float32); r.ref_warp(3, p(itl_in[i]), p(o)); itl[i] = o[0]
    out.update(itl_in=itl_in, itl_out=itl)
    ei = (rng.random(n, dtype=np.float32) * 1.98 - 0.99).astype(np.float32); eo = np.zeros(n, np.float32)
    for i in range(n):
        o = np.zeros(1, np.float32); r.ref_warp(4, p(ei[i:i + 1]), p(o)); eo[i] = o[0]
    out.update(erfinv_in=ei, erfinv_out=eo)
    nv = rng.normal(size=(n, 3)).astype(np.float32); nv /= np.linalg.norm(nv, axis=1, keepdims=True)
    cs = np.zeros((n, 6), np.float32)
    for i in range(n):
        r.ref_warp(5, p(nv[i]), p(cs[i]))
    out.update(cs_in=nv, cs_out=cs)
    # --- fresnel
    ci = (rng.random(n, dtype=np.float32) * 2 - 1).astype(np.float32); eta = (1.0 + rng.random(n, dtype=np.float32)).astype(np.float32)
    fr = np.zeros((n, 4), np.float32)
    for i in range(n):
        r.ref_fresnel(float(ci[i]), float(eta[i]), p(fr[i]))
    out.update(fresnel_cos=ci, fresnel_eta=eta, fresnel_out=fr)
    # the dielectric spot check of mitsuba3/src/bsdfs/tests/test_dielectric.py:138-158 (wi at 80 deg, eta 1.5)
    spot = np.zeros(4, np.float32); r.ref_fresnel(float(np.cos(np.radians(80.0))), 1.5, p(spot)); out["fresnel_spot"] = spot
    ce, ck = (rng.random(n, dtype=np.float32) * 3).astype(np.float32), (rng.random(n, dtype=np.float32) * 9).astype(np.float32)
    cc = rng.random(n, dtype=np.float32)
    out.update(fc_cos=cc, fc_eta=ce, fc_k=ck, fc_out=np.array([r.ref_fresnel_conductor(float(cc[i]), float(ce[i]), float(ck[i])) for i in range(n)], np.float32))
    # --- microfacet
    wi = rng.normal(size=(n, 3)).astype(np.float32); wi[:, 2] = np.abs(wi[:, 2]) + 0.05; wi /= np.linalg.norm(wi, axis=1, keepdims=True)
    su = rng.random((n, 2), dtype=np.float32)
    out.update(mf_wi=wi, mf_u=su)
    for t, tn in ((0, "beckmann"), (1, "ggx")):
        for (au, av) in ((0.1, 0.1), (0.2, 0.2), (0.3, 0.15)):
            for vis in (1, 0):
                key = f"mf_{tn}_{au}_{av}_{vis}"
                smp = np.zeros((n, 4), np.float32); ev = np.zeros(n, np.float32); g1 = np.zeros(n, np.float32); pdf = np.zeros(n, np.float32)
                for i in range(n):
                    inp = np.concatenate([wi[i], su[i]]).astype(np.float32)
                    r.ref_microfacet(t, au, av, vis, 0, p(inp), p(smp[i]))
                    m = smp[i, :3].copy(); o = np.zeros(1, np.float32)
                    r.ref_microfacet(t, au, av, vis, 1, p(m), p(o)); ev[i] = o[0]
                    inp2 = np.concatenate([wi[i], m]).astype(np.float32)
                    r.ref_microfacet(t, au, av, vis, 2, p(inp2), p(o)); g1[i] = o[0]
                    r.ref_microfacet(t, au, av, vis, 3, p(inp2), p(o)); pdf[i] = o[0]
                out.update({key + "_sample": smp, key + "_eval": ev, key + "_g1": g1, key + "_pdf": pdf})
    # --- Hierarchical2D on a 33x17 luminance map with a hot spot
    h, w = 17, 33
    lum = (rng.random((h, w), dtype=np.float32) + 0.05).astype(np.float32); lum[5, 7] = 50.0
    lum[:, -1] = lum[:, 0]
    hd = r.ref_h2d_create(p(np.ascontiguousarray(lum)), w, h)
    su = rng.random((n, 2), dtype=np.float32); hs = np.zeros((n, 3), np.float32)
    for i in range(n):
        r.ref_h2d_sample(hd, float(su[i, 0]), float(su[i, 1]), p(hs[i]))
    he = np.array([r.ref_h2d_eval(hd, float(su[i, 0]), float(su[i, 1])) for i in range(n)], np.float32)
    out.update(h2d_lum=lum, h2d_u=su, h2d_sample=hs, h2d_eval=he)
    np.savez_compressed(os.path.join(HERE, "ref_pin.npz"), **out)
    print("wrote", os.path.join(HERE, "ref_pin.npz"), len(out), "arrays")


if __name__ == "__main__":
    sys.exit(main())
