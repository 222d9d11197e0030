"""Fixtures for the reference's render Z-tests of the rough-plastic / rough-dielectric spheres
(mitsuba3/resources/data/tests/scenes/bsdf_spheres/test_{roughplastic_linear,roughplastic_nonlinear,roughdielectric}_path.xml
over fragments/single_sphere.xml, and scenes/infinite_mirrors/scene_path.xml; procedure: src/render/tests/test_renders.py:158-226).

Written to tests/golden/ref_bsdf_spheres.npz:
* `<scene>_mean`, `<scene>_var`: the stored reference image and its per-sample variance (refs/*_ref_rgb{,_var}.exr);
* `envmap`: common/textures/museum.exr (half precision, as stored);
* `light_radiance`, `light_reflectance`: the scene's `<spectrum value="400:.., 500:..">` entries converted the way the rgb
  variants do it (xml.cpp:1229-1250 -> spectrum_list_to_srgb, spectrum.cpp:85-134): 1000-step Riemann sum of the piecewise
  linear spectrum against the CIE 1931 observer (x D65 for reflectances), XYZ -> sRGB, clamped.  The CIE tables are read
  from the reference's spectrum.cpp / spectrum.h at generation time and are not stored here.  The observer is Y-normalised
  (MI_CIE_Y_NORMALIZATION, spectrum.h:132): the stored reference images were rendered that way -- with the unnormalised
  tables the area light would be 106.75 x brighter and the images 8 x brighter than the stored ones (checked with the oracle).

    python tests/golden/make_bsdf_spheres_refs.py      # needs /root/reference (this container only)
"""
import os
import re
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import exr_min  # noqa: E402

MI = "/root/reference/mitsuba3"
SCENES = ["roughplastic_linear", "roughplastic_nonlinear", "roughdielectric"]
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "ref_bsdf_spheres.npz")
CIE_MIN, CIE_MAX, CIE_SAMPLES = 360.0, 830.0, 95


def _table(text, name, n):
    body = text[text.index(name):]; body = body[body.index("{") + 1:body.index("};")]
    vals = [float(x) for x in re.findall(r"(?:Float\()?(-?\d+\.\d+(?:[eE][-+]?\d+)?)f?\)?", body)]
    assert len(vals) == n, (name, len(vals))
    return np.asarray(vals, np.float32)


def _lerp_table(tbl, w):
    t = (np.float32(w) - np.float32(CIE_MIN)) * np.float32((CIE_SAMPLES - 1) / (CIE_MAX - CIE_MIN))
    i0 = int(np.clip(int(t), 0, CIE_SAMPLES - 2)); w1 = np.float32(t) - np.float32(i0)
    return (np.float32(1) - w1) * tbl[..., i0] + w1 * tbl[..., i0 + 1]


def spectrum_list_to_srgb(wavelengths, values, bounded, d65, xyz_tbl, d65_tbl):
    xyz = np.zeros(3, np.float32); steps = 1000
    for i in range(steps):
        w = np.float32(CIE_MIN) + np.float32(i / np.float32(steps - 1)) * np.float32(CIE_MAX - CIE_MIN)
        if w < wavelengths[0] or w > wavelengths[-1]:
            continue
        k = int(np.clip(np.searchsorted(wavelengths, w, side="right") - 1, 0, len(wavelengths) - 2))
        w0, w1, v0, v1 = wavelengths[k], wavelengths[k + 1], values[k], values[k + 1]
        v = np.float32(((w - w1) * v0 + (w0 - w) * v1) / (w0 - w1))
        xyz += _lerp_table(xyz_tbl, w) * v * (_lerp_table(d65_tbl, w) * np.float32(1.0 / 98.99741751876255) if d65 else np.float32(1))
    xyz *= np.float32((CIE_MAX - CIE_MIN) / steps) * np.float32(1.0 / 106.7502593994140625)
    m = np.array([[3.240479, -1.537150, -0.498535], [-0.969256, 1.875991, 0.041556], [0.055648, -0.204043, 1.057311]], np.float32)
    rgb = m @ xyz
    return np.clip(rgb, 0.0, 1.0) if bounded else np.maximum(rgb, 0.0)


if __name__ == "__main__":
    xyz_tbl = _table(open(f"{MI}/src/core/spectrum.cpp").read(), "cie1931_tbl[", 3 * CIE_SAMPLES).reshape(3, CIE_SAMPLES)
    d65_tbl = _table(open(f"{MI}/include/mitsuba/core/spectrum.h").read(), "d65_table[", CIE_SAMPLES)
    wl = np.array([400, 500, 600, 700], np.float32)
    out = {
        "light_radiance": spectrum_list_to_srgb(wl, np.array([0, 8, 15.6, 18.4], np.float32), False, False, xyz_tbl, d65_tbl),
        "light_reflectance": spectrum_list_to_srgb(wl, np.array([0.78] * 4, np.float32), True, True, xyz_tbl, d65_tbl),
    }
    print({k: v.tolist() for k, v in out.items()})
    env = exr_min.read_rgb(f"{MI}/resources/data/common/textures/museum.exr")
    assert np.array_equal(env.astype(np.float16).astype(np.float32), env), "museum.exr is expected to hold half-precision values"
    out["envmap"] = env.astype(np.float16)
    refs = f"{MI}/resources/data/tests/scenes/bsdf_spheres/refs"
    for s in SCENES:
        out[f"{s}_mean"] = exr_min.read_rgb(f"{refs}/test_{s}_ref_rgb.exr")
        out[f"{s}_var"] = exr_min.read_rgb(f"{refs}/test_{s}_ref_rgb_var.exr")
        print(s, out[f"{s}_mean"].shape, float(out[f"{s}_mean"].mean()), float(out[f"{s}_var"].mean()))
    # scenes/infinite_mirrors/scene_path.xml: three touching conductor spheres under the same envmap, max_depth 8, 32x32
    mirrors = f"{MI}/resources/data/tests/scenes/infinite_mirrors/refs"
    out["infinite_mirrors_mean"] = exr_min.read_rgb(f"{mirrors}/scene_ref_rgb.exr")
    out["infinite_mirrors_var"] = exr_min.read_rgb(f"{mirrors}/scene_ref_rgb_var.exr")
    np.savez_compressed(OUT, **out)
    print("wrote", OUT, os.path.getsize(OUT), "bytes")
