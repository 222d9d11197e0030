"""Small scene dictionaries shared by the CPU and GPU tests (all generated from integer seeds)."""
import numpy as np

from mitsuba3_experiments_b200.scene import T
from mitsuba3_experiments_b200.scenes import cornell_box


def value_noise_texture(h=64, w=96, seed=2, dtype=np.float32):
    rng = np.random.RandomState(seed)
    t = rng.uniform(0.15, 0.95, size=(h, w, 3)).astype(np.float32)
    return t.astype(dtype)


def sky_envmap(h=32, w=64):
    """Closed-form lat-long map: sky gradient plus a bright lobe (kept small for the tests)."""
    v, u = np.meshgrid((np.arange(h) + 0.5) / h, (np.arange(w) + 0.5) / w, indexing="ij")
    sky = 0.3 + 0.7 * (1.0 - v)
    sun = 40.0 * np.exp(-((u - 0.3) ** 2 + (v - 0.25) ** 2) / 0.004)
    img = np.stack([sky * 0.6 + sun, sky * 0.8 + sun * 0.9, sky + sun * 0.7], axis=-1)
    return img.astype(np.float32)


def random_mesh(n_tris, seed, extent=1.0, size=0.15):
    rng = np.random.RandomState(seed)
    c = rng.uniform(-extent, extent, size=(n_tris, 1, 3))
    v = c + rng.uniform(-size, size, size=(n_tris, 3, 3))
    pos = v.reshape(-1, 3).astype(np.float32)
    faces = np.arange(3 * n_tris, dtype=np.uint32).reshape(-1, 3)
    return pos, faces


def materials_scene(width=48, height=40, spp=8, envmap=True, filter_type="gaussian", mesh_light=True):
    """Cornell box variant exercising every BSDF / emitter / texture path of the scope table."""
    d = cornell_box(width, height, spp)
    d["sensor"]["film"]["rfilter"] = {"type": filter_type}
    d["metal"] = {"type": "twosided", "nested": {"type": "roughconductor", "distribution": "ggx", "alpha": 0.2,
                                                  "eta": [1.657, 0.880, 0.521], "k": [9.224, 6.270, 4.837]}}
    d["beck"] = {"type": "roughconductor", "distribution": "beckmann", "alpha_u": 0.1, "alpha_v": 0.25,
                 "eta": [0.2, 0.92, 1.1], "k": [3.9, 2.45, 2.14]}
    d["glass"] = {"type": "dielectric", "int_ior": 1.5, "ext_ior": 1.0}
    d["tex"] = {"type": "twosided", "nested": {"type": "diffuse", "reflectance": {"type": "bitmap", "data": value_noise_texture(), "raw": True}}}
    d["tex16"] = {"type": "diffuse", "reflectance": {"type": "bitmap", "data": value_noise_texture(32, 32, 5, np.float16),
                                                     "wrap_mode": "mirror", "to_uv": [[2.5, 0, 0.1], [0, 2.5, -0.2], [0, 0, 1]]}}
    d["small-box"]["bsdf"] = {"type": "ref", "id": "glass"}
    # lift the glass box off the floor: mi.cornell_box() rests it exactly ON the floor (coplanar triangles), and a ray
    # inside the glass that reaches that plane resolves the z-fight by the last ulp of t -- chaotic in any implementation
    d["small-box"]["to_world"] = T().translate([0.335, -0.62, 0.38]).rotate([0, 1, 0], -17).scale(0.3)
    d["large-box"]["bsdf"] = {"type": "ref", "id": "metal"}
    d["back"]["bsdf"] = {"type": "ref", "id": "tex"}
    d["floor"]["bsdf"] = {"type": "ref", "id": "tex16"}
    d["red-wall"]["bsdf"] = {"type": "ref", "id": "beck"}
    # smooth conductor (delta reflection), copper-like eta/k, tinted specular_reflectance, twosided
    d["mirror"] = {"type": "twosided", "nested": {"type": "conductor", "eta": [0.27, 0.68, 1.32], "k": [3.6, 2.62, 2.29],
                                                   "specular_reflectance": [0.9, 0.95, 1.0]}}
    d["green-wall"]["bsdf"] = {"type": "ref", "id": "mirror"}
    # the bedroom's curtain (scene.xml:199-206): opacity mask over a twosided diffuse, hanging in front of the back wall
    d["curtain-bsdf"] = {"type": "mask", "opacity": 0.531049, "nested": {"type": "twosided", "nested": {"type": "diffuse", "reflectance": [0.8, 0.7, 0.6]}}}
    d["curtain"] = {"type": "rectangle", "to_world": T().translate([-0.3, 0.1, -0.55]).rotate([0, 1, 0], 20).scale(0.45),
                    "bsdf": {"type": "ref", "id": "curtain-bsdf"}}
    # the bedroom's rough glass (scene.xml: roughdielectric, GGX): a frosted block floating above the boxes
    d["frosted"] = {"type": "roughdielectric", "distribution": "ggx", "alpha": 0.15, "int_ior": 1.5, "ext_ior": 1.0,
                    "specular_transmittance": [0.9, 0.95, 1.0]}
    d["frosted-box"] = {"type": "cube", "to_world": T().translate([0.25, 0.45, 0.1]).rotate([0, 1, 0], 30).scale([0.2, 0.12, 0.2]),
                        "bsdf": {"type": "ref", "id": "frosted"}}
    if mesh_light:
        pos = np.array([[-0.6, 0.2, -0.6], [-0.3, 0.2, -0.6], [-0.45, 0.45, -0.5], [-0.45, 0.3, -0.3]], np.float32)
        faces = np.array([[0, 1, 2], [0, 3, 1], [1, 3, 2], [0, 2, 3]], np.uint32)
        d["tetra-light"] = {"type": "mesh", "positions": pos, "faces": faces, "bsdf": {"type": "ref", "id": "white"},
                            "emitter": {"type": "area", "radiance": {"type": "rgb", "value": [6.0, 8.0, 5.0]}}}
    if envmap:
        del d["ceiling"]
        d["env"] = {"type": "envmap", "data": sky_envmap(), "scale": 0.8,
                    "to_world": T().rotate([0, 1, 0], 35.0).rotate([1, 0, 0], 10.0)}
    return d


def triangles_scene(n_tris=20000, width=64, height=48, spp=4, seed=3):
    """Random-triangle stress scene (C4 in miniature): one rectangle light, constant environment."""
    pos, faces = random_mesh(n_tris, seed)
    return {
        "type": "scene",
        "sensor": {"type": "perspective", "fov": 40, "near_clip": 0.01, "far_clip": 100.0,
                   "to_world": T().look_at([0, 0, 3.5], [0, 0, 0], [0, 1, 0]),
                   "sampler": {"type": "independent", "sample_count": spp},
                   "film": {"type": "hdrfilm", "width": width, "height": height, "rfilter": {"type": "tent"}}},
        "soup": {"type": "mesh", "positions": pos, "faces": faces,
                 "bsdf": {"type": "twosided", "nested": {"type": "diffuse", "reflectance": {"type": "rgb", "value": [0.7, 0.6, 0.5]}}}},
        "light": {"type": "rectangle", "to_world": T().translate([0, 1.4, 0]).rotate([1, 0, 0], 90).scale(0.5),
                  "emitter": {"type": "area", "radiance": {"type": "rgb", "value": [50, 50, 50]}}},
        "sky": {"type": "constant", "radiance": {"type": "rgb", "value": [0.4, 0.5, 0.6]}},
    }


def mirror_scene(width=16, height=16, spp=4, tilt_deg=30.0, eta=(0.2, 0.92, 1.1), k=(3.9, 2.45, 2.14), refl=(0.9, 0.8, 0.7), radiance=(1.0, 2.0, 3.0), opacity=None):
    """Known-answer scene for the smooth conductor: a huge mirror under a constant environment.  Every camera ray hits the
    mirror once and escapes, so radiance(pixel) = specular_reflectance * F_conductor(cos_i; eta, k) * radiance, with
    cos_i = dot(-d, n) known in closed form per pixel (conductor.cpp:249-307, fresnel.h:92-117)."""
    return {
        "type": "scene",
        "sensor": {"type": "perspective", "fov": 20, "near_clip": 0.01, "far_clip": 100.0,
                   "to_world": T().look_at([0, 0, 4], [0, 0, 0], [0, 1, 0]),
                   "sampler": {"type": "independent", "sample_count": spp},
                   "film": {"type": "hdrfilm", "width": width, "height": height, "rfilter": {"type": "box"}}},
        "m": ({"type": "conductor", "eta": list(eta), "k": list(k), "specular_reflectance": list(refl)} if opacity is None else
              {"type": "mask", "opacity": opacity, "nested": {"type": "conductor", "eta": list(eta), "k": list(k), "specular_reflectance": list(refl)}}),
        "plane": {"type": "rectangle", "to_world": T().rotate([1, 0, 0], tilt_deg).scale(50.0), "bsdf": {"type": "ref", "id": "m"}},
        "env": {"type": "constant", "radiance": {"type": "rgb", "value": list(radiance)}},
    }


def crop_window_scene(filter_type="gaussian"):
    """CropWindowConfig of the reference's PRB acceptance tests (test_ad_integrators.py:250-277 over ConfigBase :39-66):
    a diffuse rectangle under a constant emitter, 64x64 film cropped to 32x32 at (32, 20), gaussian(0.5), sample_border."""
    rf = {"type": "gaussian", "stddev": 0.5} if filter_type == "gaussian" else {"type": filter_type}
    return {"type": "scene",
            "plane": {"type": "rectangle", "bsdf": {"type": "diffuse"}},
            "light": {"type": "constant"},
            "sensor": {"type": "perspective", "to_world": T().look_at(origin=[0, 0, 4], target=[0, 0, 0], up=[0, 1, 0]),
                       "film": {"type": "hdrfilm", "rfilter": rf, "width": 64, "height": 64, "sample_border": True,
                                "crop_width": 32, "crop_height": 32, "crop_offset_x": 32, "crop_offset_y": 20,
                                "pixel_format": "rgb", "component_format": "float32"}}}


def ad_config_scene(name, res=128):
    """The basic configs of the reference's PRB acceptance tests (test_ad_integrators.py:116-277) that stay inside the hot
    path's scope -> (scene dict, max_depth, differentiated parameter key | None).  Spheres are tessellated (scene.py)."""
    film = {"type": "hdrfilm", "rfilter": {"type": "gaussian", "stddev": 0.5}, "width": res, "height": res, "sample_border": True,
            "pixel_format": "rgb", "component_format": "float32"}
    d = {"type": "scene",
         "sensor": {"type": "perspective", "to_world": T().look_at(origin=[0, 0, 4], target=[0, 0, 0], up=[0, 1, 0]), "film": film}}
    side = T().translate([1.25, 0.0, 1.0]) @ T().rotate([0, 1, 0], -90)
    if name == "crop_window":
        return crop_window_scene(), 2, "plane.bsdf.reflectance.value"
    if name == "diffuse_albedo":
        d.update(plane={"type": "rectangle", "bsdf": {"type": "diffuse"}}, sphere={"type": "sphere", "to_world": T().scale(0.25)},
                 light={"type": "constant"})
        return d, 2, "plane.bsdf.reflectance.value"
    if name == "diffuse_albedo_g_i":
        d.update(plane={"type": "rectangle"}, sphere={"type": "sphere", "to_world": T().scale(0.25)},
                 green={"type": "rectangle", "bsdf": {"type": "diffuse", "reflectance": {"type": "rgb", "value": [0.1, 1.0, 0.1]}}, "to_world": side},
                 light={"type": "constant", "radiance": 3.0})
        return d, 3, "green.bsdf.reflectance.value"
    if name == "area_light_radiance":
        d.update(plane={"type": "rectangle", "bsdf": {"type": "diffuse", "reflectance": {"type": "rgb", "value": [1.0, 1.0, 1.0]}}},
                 sphere={"type": "sphere", "to_world": T().scale(0.25)},
                 light={"type": "rectangle", "to_world": side, "emitter": {"type": "area", "radiance": {"type": "rgb", "value": [3.0, 3.0, 3.0]}}})
        return d, 2, None
    if name == "directly_visible_area_light_radiance":
        d.update(light={"type": "rectangle", "emitter": {"type": "area", "radiance": {"type": "rgb", "value": [1.0, 1.0, 1.0]}}})
        return d, 2, None
    raise KeyError(name)


AD_CONFIGS = ["crop_window", "diffuse_albedo", "diffuse_albedo_g_i", "area_light_radiance", "directly_visible_area_light_radiance"]


def bsdf_sphere_scene(name, fx, light_scale=1.0, spp=64):
    """bsdf_spheres/test_<name>_path.xml over fragments/single_sphere.xml of the reference's render tests, rebuilt from the
    fixture `fx` (tests/golden/ref_bsdf_spheres.npz): unit spheres (meshes/sphere.obj -> the tessellated sphere, both within
    1e-4 of the analytic one), the rotated 4x4 area light, museum.exr and the point light; 48x48 box-filtered film."""
    def rp(alpha, y, z, nonlinear):
        return {"type": "sphere", "tessellation": 127, "to_world": T().translate([4, y, z]),
                "bsdf": {"type": "roughplastic", "distribution": "ggx", "alpha": alpha, "int_ior": "bk7", "ext_ior": "air",
                         "nonlinear": nonlinear, "diffuse_reflectance": {"type": "rgb", "value": [0.02, 0.2, 0.1]}}}
    if name == "infinite_mirrors":       # scenes/infinite_mirrors/scene_path.xml: 8 bounces between three conductor spheres
        def mirror(y, z):
            return {"type": "sphere", "tessellation": 127, "to_world": T().translate([0, y, z]), "bsdf": {"type": "conductor", "eta": 0.789, "k": 6.15}}
        return {"type": "scene", "integrator": {"type": "path", "max_depth": 8}, "s0": mirror(-1.05, -0.9), "s1": mirror(1.05, -0.9), "s2": mirror(0.0, 0.9),
                "env": {"type": "envmap", "data": np.asarray(fx["envmap"], np.float32)},
                "sensor": {"type": "perspective", "near_clip": 1, "far_clip": 1000,
                           "to_world": T().look_at(origin=[-5, 0, 0], target=[0, 0, 0], up=[0, 0, 1]),
                           "film": {"type": "hdrfilm", "rfilter": {"type": "box"}, "width": 32, "height": 32},
                           "sampler": {"type": "independent", "sample_count": spp}}}
    d = {"type": "scene", "integrator": {"type": "path", "max_depth": 2}}
    if name.startswith("roughplastic"):
        nl = name.endswith("nonlinear")
        d.update(s0=rp(0.01, 1.2, 1.2, nl), s1=rp(0.1, -1.2, 1.2, nl), s2=rp(0.4, 1.2, -1.2, nl), s3=rp(1.0, -1.2, -1.2, nl))
    elif name == "roughdielectric":
        d["s0"] = {"type": "sphere", "tessellation": 127,
                   "bsdf": {"type": "roughdielectric", "distribution": "beckmann", "alpha": 0.02, "int_ior": "bk7", "ext_ior": "air"}}
    else:
        raise KeyError(name)
    d["sensor"] = {"type": "perspective", "near_clip": 1, "far_clip": 1000,
                   "to_world": T().look_at(origin=[-4, 0, 0], target=[0, 0, 0], up=[0, 0, 1]),
                   "film": {"type": "hdrfilm", "rfilter": {"type": "box"}, "width": 48, "height": 48},
                   "sampler": {"type": "independent", "sample_count": spp}}
    # <rotate y=1 angle=120/> <scale 2/> <translate -4 0 6/>: applied in document order
    to_world = T().translate([-4, 0, 6]) @ T().scale([2, 2, 2]) @ T().rotate([0, 1, 0], 120)
    d["light"] = {"type": "mesh", "positions": [[-1, -1, 0], [1, -1, 0], [-1, 1, 0], [1, 1, 0]], "faces": [[1, 2, 0], [1, 3, 2]],
                  "to_world": to_world,
                  "bsdf": {"type": "diffuse", "reflectance": {"type": "rgb", "value": np.asarray(fx["light_reflectance"]).tolist()}},
                  "emitter": {"type": "area", "radiance": {"type": "rgb", "value": (np.asarray(fx["light_radiance"]) * light_scale).tolist()}}}
    d["env"] = {"type": "envmap", "data": np.asarray(fx["envmap"], np.float32)}
    d["point"] = {"type": "point", "position": [-4, 0, -6], "intensity": {"type": "rgb", "value": [300, 300, 300]}}
    return d


def render_z_test(img_rgb, spp, ref_rgb, ref_var_rgb, significance_level=0.01):
    """The acceptance test of the reference's render tests (src/render/tests/test_renders.py:158-226): per-value Z-test of the
    rendered mean against the stored reference mean / variance in XYZ, Sidak-corrected, passes when >= 99.75 % of the pixels
    accept the null hypothesis.  Returns (passed, fraction of accepted values)."""
    from math import erf, sqrt
    m = np.array([[0.412453, 0.357580, 0.180423], [0.212671, 0.715160, 0.072169], [0.019334, 0.119193, 0.950227]], np.float32)
    img, ref, var = (np.asarray(a, np.float32) @ m.T for a in (img_rgb, ref_rgb, ref_var_rgb))
    var = np.maximum(var, 1e-4)
    z = np.abs(img - ref) * np.sqrt(spp / var)
    cdf = np.vectorize(lambda x: (1.0 - erf(-x / sqrt(2.0))) * 0.5)(z.astype(np.float64))
    p_value = 2.0 * (1.0 - cdf)
    pixel_count = img.shape[0] * img.shape[1]
    alpha = 1.0 - (1.0 - significance_level) ** (1.0 / pixel_count)
    success = p_value > alpha
    return (np.count_nonzero(success) / 3) >= 0.9975 * pixel_count, np.count_nonzero(success) / success.size
