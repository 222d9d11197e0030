"""Extracts the reference's stored PRB goldens (mitsuba3/resources/data/tests/integrators/*_ref.exr) of the basic
test_ad_integrators.py configs that are in the hot path's scope into tests/golden/ref_ad_integrators.npz.

The container has no OpenEXR: tests/golden/exr_min.py reads the files (raw-stored and PIZ chunks).

    python tests/golden/make_ad_integrator_refs.py      # needs /root/reference (this container only)
"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import exr_min  # noqa: E402

REF = "/root/reference/mitsuba3/resources/data/tests/integrators"
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "ref_ad_integrators.npz")
# BASIC_CONFIGS_LIST (test_ad_integrators.py:694-705) minus the two that move geometry (translate_textured_plane,
# rotate_shading_normals_plane: shape derivatives are outside PRB's reflectance-parameter scope)
CONFIGS = ["crop_window", "diffuse_albedo", "diffuse_albedo_g_i", "area_light_radiance", "directly_visible_area_light_radiance"]

if __name__ == "__main__":
    out = {}
    for name in CONFIGS:
        for kind in ("primal", "fwd"):
            img = exr_min.read_rgb(os.path.join(REF, f"test_{name}_image_{kind}_ref.exr"))
            assert np.isfinite(img).all()
            out[f"{name}_{kind}"] = img
            print(name, kind, img.shape, float(img.mean()))
    np.savez_compressed(OUT, **out)
    print("wrote", OUT, os.path.getsize(OUT), "bytes")
