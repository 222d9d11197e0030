"""Golden vectors for the rough-plastic precomputation: the REFERENCE's own eval_transmittance / eval_reflectance
(mitsuba/render/microfacet.h:463-570) run exactly as RoughPlastic::parameters_changed does (roughplastic.cpp:232-262),
through oracle/_ref/libref_pin.so (reference headers compiled in scalar mode, `make -C oracle ref`).
Writes tests/golden/ref_rough_tables.npz.  Needs /root/reference (this container only)."""
import ctypes as C
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
CASES = [(1, 0.15, 1.5), (0, 0.15, 1.5), (1, 0.4, 1.33), (0, 0.05, 1.49 / 1.000277)]      # (0 beckmann | 1 ggx, alpha, eta)


def main():
    r = C.CDLL(os.path.join(HERE, "..", "..", "oracle", "_ref", "libref_pin.so"))
    r.ref_rough_tables.argtypes = [C.c_int, C.c_float, C.c_float, C.c_void_p, C.c_void_p]
    out = {"cases": np.array(CASES, np.float32)}
    for i, (t, a, e) in enumerate(CASES):
        tr = np.zeros(64, np.float32); ir = np.zeros(1, np.float32)
        r.ref_rough_tables(t, a, e, tr.ctypes.data, ir.ctypes.data)
        out[f"trans_{i}"] = tr; out[f"int_refl_{i}"] = ir
        print(t, a, e, tr[[0, 1, 16, 32, 63]], ir)
    np.savez_compressed(os.path.join(HERE, "ref_rough_tables.npz"), **out)


if __name__ == "__main__":
    sys.exit(main())
