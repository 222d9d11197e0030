"""Scene dictionaries for the benchmark configurations.

``cornell_box()`` carries the same values as ``mi.cornell_box()``
(mitsuba3/src/python/python/util.py:568-703): it is scene *data*, the schema is Mitsuba's.
The bedroom stand-in and the random-triangle stress scene are generated procedurally from integer seeds
because the reference's bedroom meshes/textures are git-LFS pointers.
"""
from __future__ import annotations

import os

import numpy as np

from .scene import T


def cornell_box(width: int = 256, height: int = 256, spp: int = 64) -> dict:
    white = {"type": "ref", "id": "white"}
    return {
        "type": "scene",
        "integrator": {"type": "path", "max_depth": 8},
        "sensor": {
            "type": "perspective", "fov_axis": "smaller", "near_clip": 0.001, "far_clip": 100.0,
            "focus_distance": 1000, "fov": 39.3077,
            "to_world": T().look_at(origin=[0, 0, 3.90], target=[0, 0, 0], up=[0, 1, 0]),
            "sampler": {"type": "independent", "sample_count": spp},
            "film": {"type": "hdrfilm", "width": width, "height": height, "rfilter": {"type": "gaussian"},
                     "pixel_format": "rgb", "component_format": "float32"},
        },
        "white": {"type": "diffuse", "reflectance": {"type": "rgb", "value": [0.885809, 0.698859, 0.666422]}},
        "green": {"type": "diffuse", "reflectance": {"type": "rgb", "value": [0.105421, 0.37798, 0.076425]}},
        "red": {"type": "diffuse", "reflectance": {"type": "rgb", "value": [0.570068, 0.0430135, 0.0443706]}},
        "light": {"type": "rectangle",
                  "to_world": T().translate([0.0, 0.99, 0.01]).rotate([1, 0, 0], 90).scale([0.23, 0.19, 0.19]),
                  "bsdf": white,
                  "emitter": {"type": "area", "radiance": {"type": "rgb", "value": [18.387, 13.9873, 6.75357]}}},
        "floor": {"type": "rectangle", "to_world": T().translate([0.0, -1.0, 0.0]).rotate([1, 0, 0], -90), "bsdf": white},
        "ceiling": {"type": "rectangle", "to_world": T().translate([0.0, 1.0, 0.0]).rotate([1, 0, 0], 90), "bsdf": white},
        "back": {"type": "rectangle", "to_world": T().translate([0.0, 0.0, -1.0]), "bsdf": white},
        "green-wall": {"type": "rectangle", "to_world": T().translate([1.0, 0.0, 0.0]).rotate([0, 1, 0], -90),
                       "bsdf": {"type": "ref", "id": "green"}},
        "red-wall": {"type": "rectangle", "to_world": T().translate([-1.0, 0.0, 0.0]).rotate([0, 1, 0], 90),
                     "bsdf": {"type": "ref", "id": "red"}},
        "small-box": {"type": "cube", "to_world": T().translate([0.335, -0.7, 0.38]).rotate([0, 1, 0], -17).scale(0.3),
                      "bsdf": white},
        "large-box": {"type": "cube",
                      "to_world": T().translate([-0.33, -0.4, -0.28]).rotate([0, 1, 0], 18.25).scale([0.3, 0.61, 0.3]),
                      "bsdf": white},
    }


# ---------------------------------------------------------------------------------------------
# Procedural meshes (all closed form / integer seeded): used by the bedroom stand-in and the stress scene
# ---------------------------------------------------------------------------------------------
def _vertex_normals(pos, faces):
    p = pos.astype(np.float64)
    fn = np.cross(p[faces[:, 1]] - p[faces[:, 0]], p[faces[:, 2]] - p[faces[:, 0]])
    n = np.zeros_like(p)
    for k in range(3):
        np.add.at(n, faces[:, k], fn)
    n /= np.maximum(np.linalg.norm(n, axis=1, keepdims=True), 1e-20)
    return n.astype(np.float32)


def icosphere(level: int):
    t = (1.0 + 5.0 ** 0.5) / 2.0
    v = np.array([[-1, t, 0], [1, t, 0], [-1, -t, 0], [1, -t, 0], [0, -1, t], [0, 1, t], [0, -1, -t], [0, 1, -t],
                  [t, 0, -1], [t, 0, 1], [-t, 0, -1], [-t, 0, 1]], dtype=np.float64)
    v /= np.linalg.norm(v, axis=1, keepdims=True)
    f = np.array([[0, 11, 5], [0, 5, 1], [0, 1, 7], [0, 7, 10], [0, 10, 11], [1, 5, 9], [5, 11, 4], [11, 10, 2], [10, 7, 6],
                  [7, 1, 8], [3, 9, 4], [3, 4, 2], [3, 2, 6], [3, 6, 8], [3, 8, 9], [4, 9, 5], [2, 4, 11], [6, 2, 10],
                  [8, 6, 7], [9, 8, 1]], dtype=np.int64)
    for _ in range(level):
        e = np.concatenate([f[:, [0, 1]], f[:, [1, 2]], f[:, [2, 0]]], axis=0)
        key = np.minimum(e[:, 0], e[:, 1]) * (len(v) + 1) + np.maximum(e[:, 0], e[:, 1])
        uk, inv = np.unique(key, return_inverse=True)
        a, b = uk // (len(v) + 1), uk % (len(v) + 1)
        mid = v[a] + v[b]; mid /= np.linalg.norm(mid, axis=1, keepdims=True)
        m = inv.reshape(3, -1) + len(v)
        v = np.concatenate([v, mid], axis=0)
        f = np.concatenate([np.stack([f[:, 0], m[0], m[2]], 1), np.stack([f[:, 1], m[1], m[0]], 1),
                            np.stack([f[:, 2], m[2], m[1]], 1), np.stack([m[0], m[1], m[2]], 1)], axis=0)
    return v, f.astype(np.uint32)


def displaced_icosphere(level, radius, amp, phase):
    v, f = icosphere(level)
    d = 1.0 + amp * (np.sin(5.0 * v[:, 0] + phase) * np.sin(4.0 * v[:, 1] + 2.0 * phase) + 0.5 * np.sin(9.0 * v[:, 2] + 3.0 * phase))
    pos = (v * d[:, None] * radius).astype(np.float32)
    uv = np.stack([np.arctan2(v[:, 0], v[:, 2]) / (2 * np.pi) + 0.5, np.arccos(np.clip(v[:, 1], -1, 1)) / np.pi], axis=1).astype(np.float32)
    return pos, _vertex_normals(pos, f), uv, f


def torus(nu, nv, R, r):
    u, w = np.meshgrid(np.arange(nu) / nu * 2 * np.pi, np.arange(nv) / nv * 2 * np.pi, indexing="ij")
    pos = np.stack([(R + r * np.cos(w)) * np.cos(u), r * np.sin(w), (R + r * np.cos(w)) * np.sin(u)], axis=-1).reshape(-1, 3).astype(np.float32)
    nrm = np.stack([np.cos(w) * np.cos(u), np.sin(w), np.cos(w) * np.sin(u)], axis=-1).reshape(-1, 3).astype(np.float32)
    uv = np.stack([u / (2 * np.pi) * 4.0, w / (2 * np.pi)], axis=-1).reshape(-1, 2).astype(np.float32)
    i, j = np.meshgrid(np.arange(nu), np.arange(nv), indexing="ij")
    a = (i * nv + j).ravel(); b = (((i + 1) % nu) * nv + j).ravel(); c = (((i + 1) % nu) * nv + (j + 1) % nv).ravel(); d = (i * nv + (j + 1) % nv).ravel()
    faces = np.concatenate([np.stack([a, c, b], 1), np.stack([a, d, c], 1)], axis=0).astype(np.uint32)
    return pos, nrm, uv, faces


def grid_box(n, half):
    """Axis-aligned box [-half, half] with n x n quads per side (outward faces, per-face uv)."""
    hx, hy, hz = half
    s, t = np.meshgrid(np.linspace(-1, 1, n + 1), np.linspace(-1, 1, n + 1), indexing="ij")
    s, t = s.ravel(), t.ravel()
    i, j = np.meshgrid(np.arange(n), np.arange(n), indexing="ij")
    a = (i * (n + 1) + j).ravel(); b = a + (n + 1); c = b + 1; d = a + 1
    quad = np.concatenate([np.stack([a, b, c], 1), np.stack([a, c, d], 1)], axis=0)
    P, N, U, F = [], [], [], []
    one = np.ones_like(s)
    sides = [(np.stack([one, s, t], 1), [1, 0, 0], False), (np.stack([-one, s, t], 1), [-1, 0, 0], True),
             (np.stack([t, one, s], 1), [0, 1, 0], False), (np.stack([t, -one, s], 1), [0, -1, 0], True),
             (np.stack([s, t, one], 1), [0, 0, 1], False), (np.stack([s, t, -one], 1), [0, 0, -1], True)]
    for k, (p, nrm, flip) in enumerate(sides):
        P.append(p * np.array([hx, hy, hz])); N.append(np.tile(np.array(nrm, np.float64), (len(s), 1)))
        U.append(np.stack([(s + 1) / 2, (t + 1) / 2], 1))
        q = quad[:, ::-1] if flip else quad
        F.append(q + k * len(s))
    return (np.concatenate(P).astype(np.float32), np.concatenate(N).astype(np.float32), np.concatenate(U).astype(np.float32),
            np.concatenate(F).astype(np.uint32))


def value_noise_rgb(res, seed, base, octaves=4):
    rng = np.random.RandomState(seed)
    out = np.zeros((res, res), np.float64)
    for o in range(octaves):
        n = 8 << o
        g = rng.uniform(0, 1, size=(n + 1, n + 1))
        x = np.linspace(0, n, res, endpoint=False); i = x.astype(int); fx = x - i
        gx = g[:, i] * (1 - fx) + g[:, i + 1] * fx
        out += (gx[i, :] * (1 - fx)[:, None] + gx[i + 1, :] * fx[:, None]) / (2 ** o)
    out /= out.max()
    return (np.asarray(base, np.float64)[None, None, :] * (0.35 + 0.65 * out[:, :, None])).astype(np.float16)


def sky_envmap(width=1024, height=512, sun_radiance=5000.0, sun_elevation_deg=45.0, sun_azimuth_u=0.25, sun_radius_deg=5.0):
    """Lat-long RGB f32: sky gradient + one sun disc (closed form)."""
    v, u = np.meshgrid((np.arange(height) + 0.5) / height, (np.arange(width) + 0.5) / width, indexing="ij")
    theta, phi = v * np.pi, u * 2 * np.pi
    d = np.stack([np.sin(theta) * np.sin(phi), np.cos(theta), -np.sin(theta) * np.cos(phi)], axis=-1)
    up = np.clip(d[..., 1], 0, 1)
    sky = np.stack([0.35 + 0.35 * (1 - up), 0.5 + 0.3 * (1 - up), 0.9 * np.ones_like(up)], axis=-1) * (0.25 + 0.75 * (d[..., 1:2] > 0))
    el, az = np.radians(sun_elevation_deg), sun_azimuth_u * 2 * np.pi
    sd = np.array([np.cos(el) * np.sin(az), np.sin(el), -np.cos(el) * np.cos(az)])
    sun = (d @ sd) > np.cos(np.radians(sun_radius_deg))
    img = sky + sun[..., None] * sun_radiance * np.array([1.0, 0.95, 0.85])
    return img.astype(np.float32)


# BSDF palette of data/bedroom/scene.xml restricted to the north-star set
_BEDROOM_ALBEDOS = [[0.483044, 0.384664, 0.301561], [0.558543, 0.558543, 0.558543], [0.64, 0.64, 0.64], [0.485435, 0.456263, 0.428075],
                    [0.111567, 0.037068, 0.017016], [0.350827, 0.242986, 0.17883], [0.098964, 0.098964, 0.098964],
                    [0.893289, 0.893289, 0.893289], [0.799999, 0.799999, 0.799999], [0.567027, 0.567027, 0.567027]]
_AL_ETA, _AL_K = [1.65746, 0.880369, 0.521229], [9.22387, 6.26952, 4.837]


def bedroom_standin(width=1280, height=720, spp=64, n_tris=1_800_000, seed=1, envmap=True, tex_res=1024, palette="north_star"):
    """Bedroom *stand-in* (the reference's meshes/textures are git-LFS pointers): camera, film, filter and the two area
    lights are verbatim from data/bedroom/scene.xml:10-25,706-731; room shell + procedural furniture of ~n_tris
    triangles.  ``palette="north_star"`` (BASELINE config C2 as SURVEY.md 8d restates it): 65 % diffuse, 10 % textured diffuse,
    15 % rough conductor, 10 % dielectric.  ``palette="scene_xml"``: the BSDF definitions of data/bedroom/scene.xml:26-217
    themselves -- nonlinear rough plastic with a bitmap diffuse_reflectance on the floor and the wooden furniture (WoodFloorBSDF,
    WoodFurnitureBSDF), smooth conductor mirrors, rough conductors (ggx 0.2 / 0.1, beckmann 0.1), a rough dielectric lamp glass,
    smooth glass, a 0.53-opacity curtain mask and the diffuse albedos, by object in roughly the file's proportions."""
    rng = np.random.RandomState(seed)
    cam = np.array([[-0.653592, -0.0128556, -0.756738, 3.45558], [2.84986e-09, 0.999856, -0.0169858, 1.21244],
                    [0.756847, -0.0111018, -0.653498, 3.29897], [0, 0, 0, 1]])
    d = {"type": "scene", "integrator": {"type": "path", "max_depth": 8},
         "sensor": {"type": "perspective", "fov": 65.0, "fov_axis": "x", "near_clip": 0.01, "far_clip": 1000.0, "to_world": cam,
                    "sampler": {"type": "independent", "sample_count": spp},
                    "film": {"type": "hdrfilm", "width": width, "height": height, "rfilter": {"type": "tent"}, "pixel_format": "rgb"}}}
    black = {"type": "twosided", "nested": {"type": "diffuse", "reflectance": {"type": "rgb", "value": [0, 0, 0]}}}
    for k, x in enumerate((-1.47544, 1.44379)):
        m = np.array([[0.53241, 0, 0, x], [0, 0.907794, 1.78252e-08, 1.18968], [0, 1.78252e-08, 0.5, -1.26735], [0, 0, 0, 1]])
        d[f"light{k}"] = {"type": "rectangle", "to_world": m, "bsdf": black,
                          "emitter": {"type": "area", "radiance": {"type": "rgb", "value": [16.4648] * 3}}}
    # palette
    textures = [value_noise_rgb(tex_res, 2 + i, _BEDROOM_ALBEDOS[i]) for i in range(3)]
    def two(b): return {"type": "twosided", "nested": b}
    for i, a in enumerate(_BEDROOM_ALBEDOS):
        d[f"diffuse{i}"] = two({"type": "diffuse", "reflectance": {"type": "rgb", "value": a}})
    for i, t in enumerate(textures):
        d[f"textured{i}"] = two({"type": "diffuse", "reflectance": {"type": "bitmap", "data": t, "raw": True}})
    d["metal_ggx"] = two({"type": "roughconductor", "distribution": "ggx", "alpha": 0.2, "eta": _AL_ETA, "k": _AL_K})
    d["metal_beckmann"] = two({"type": "roughconductor", "distribution": "beckmann", "alpha": 0.1, "eta": _AL_ETA, "k": _AL_K})
    d["glass"] = {"type": "dielectric", "int_ior": 1.5, "ext_ior": 1.0}
    xml = palette == "scene_xml"
    if palette not in ("north_star", "scene_xml"):
        raise ValueError("palette: 'north_star' or 'scene_xml'")
    if xml:     # data/bedroom/scene.xml:108-217
        d["wood_floor"] = two({"type": "roughplastic", "alpha": 0.15, "distribution": "ggx", "int_ior": 1.5, "ext_ior": 1.0, "nonlinear": True,
                               "diffuse_reflectance": {"type": "bitmap", "data": textures[0], "raw": True}})
        d["wood_furniture"] = two({"type": "roughplastic", "alpha": 0.15, "distribution": "beckmann", "int_ior": 1.5, "ext_ior": 1.0, "nonlinear": True,
                                   "diffuse_reflectance": {"type": "bitmap", "data": textures[1], "raw": True}})
        d["mirror"] = two({"type": "conductor", "material": "none"})
        d["lamp_metal"] = two({"type": "roughconductor", "distribution": "ggx", "alpha": 0.1, "eta": _AL_ETA, "k": _AL_K})
        d["curtain_rod"] = two({"type": "roughconductor", "distribution": "ggx", "alpha": 0.1, "eta": _AL_ETA, "k": _AL_K,
                                "specular_reflectance": {"type": "rgb", "value": [0.5, 0.5, 0.5]}})
        d["lamp_glass"] = {"type": "roughdielectric", "alpha": 0.1, "distribution": "beckmann", "int_ior": 1.5, "ext_ior": 1.0}
        d["curtains"] = {"type": "mask", "opacity": 0.531049, "nested": two({"type": "diffuse", "reflectance": {"type": "rgb", "value": [1, 1, 1]}})}
    def ref(n): return {"type": "ref", "id": n}
    # room shell x in [-2.6, 4], y in [0, 2.6], z in [-1.4, 4.2]; +x wall has a 1.2 x 1.2 window hole (y 0.9..2.1, z 0.8..2.0)
    X0, X1, Y0, Y1, Z0, Z1 = -2.6, 4.0, 0.0, 2.6, -1.4, 4.2
    def quad(p0, e1, e2, name, bsdf):
        p0, e1, e2 = (np.asarray(a, np.float64) for a in (p0, e1, e2))
        pos = np.stack([p0, p0 + e1, p0 + e1 + e2, p0 + e2]).astype(np.float32)
        uv = np.array([[0, 0], [1, 0], [1, 1], [0, 1]], np.float32) * 4.0
        d[name] = {"type": "mesh", "positions": pos, "texcoords": uv, "faces": np.array([[0, 1, 2], [0, 2, 3]], np.uint32), "bsdf": ref(bsdf)}
    quad([X0, Y0, Z0], [X1 - X0, 0, 0], [0, 0, Z1 - Z0], "floor", "wood_floor" if xml else "textured0")
    quad([X0, Y1, Z0], [0, 0, Z1 - Z0], [X1 - X0, 0, 0], "ceiling", "diffuse7")
    quad([X0, Y0, Z0], [0, Y1 - Y0, 0], [X1 - X0, 0, 0], "wall_back", "diffuse2")
    quad([X0, Y0, Z1], [X1 - X0, 0, 0], [0, Y1 - Y0, 0], "wall_front", "diffuse2")
    quad([X0, Y0, Z0], [0, 0, Z1 - Z0], [0, Y1 - Y0, 0], "wall_left", "diffuse1")
    if envmap:   # +x wall around the window hole
        quad([X1, Y0, Z0], [0, 0.9, 0], [0, 0, Z1 - Z0], "wall_right_a", "diffuse1")
        quad([X1, 2.1, Z0], [0, Y1 - 2.1, 0], [0, 0, Z1 - Z0], "wall_right_b", "diffuse1")
        quad([X1, 0.9, Z0], [0, 1.2, 0], [0, 0, 0.8 - Z0], "wall_right_c", "diffuse1")
        quad([X1, 0.9, 2.0], [0, 1.2, 0], [0, 0, Z1 - 2.0], "wall_right_d", "diffuse1")
    else:
        quad([X1, Y0, Z0], [0, Y1 - Y0, 0], [0, 0, Z1 - Z0], "wall_right", "diffuse1")
    # furniture: displaced icospheres (81 920 tris at level 6), tori (2*nu*nv) and grid boxes (12*n^2), scaled to n_tris
    scale = n_tris / 1_800_000.0
    ico_level = 6 if scale >= 0.5 else (5 if scale >= 0.1 else 3)
    n_ico, n_tor, n_box = 10, 8, 10
    tor_nu = max(8, int(160 * scale ** 0.5)); tor_nv = max(8, int(256 * scale ** 0.5)); box_n = max(2, int(52 * scale ** 0.5))
    if ico_level < 6:
        n_ico = max(1, int(round(n_ico * scale * 81920 / (20 * 4 ** ico_level))))
        n_ico = min(n_ico, 24)
    kinds = ["ico"] * n_ico + ["torus"] * n_tor + ["box"] * n_box
    n_obj = len(kinds)
    # material assignment by object: 65 % diffuse, 10 % textured, 15 % rough conductor, 10 % glass
    mats = []
    for i in range(n_obj):
        u = (i * 0.6180339887 + 0.31) % 1.0
        if xml:     # 34 BSDFs of the file: 22 diffuse, 4 rough conductor, 2 rough plastic, 2 conductor, 2 dielectric, 1 rough dielectric, 1 mask
            mats.append(f"diffuse{i % len(_BEDROOM_ALBEDOS)}" if u < 0.56 else "wood_furniture" if u < 0.68 else
                        ("metal_ggx", "lamp_metal", "metal_beckmann", "curtain_rod")[i % 4] if u < 0.80 else "mirror" if u < 0.86 else
                        "glass" if u < 0.92 else "lamp_glass" if u < 0.96 else "curtains")
            continue
        mats.append(f"diffuse{i % len(_BEDROOM_ALBEDOS)}" if u < 0.65 else (f"textured{i % 3}" if u < 0.75 else
                    (("metal_ggx" if i % 2 else "metal_beckmann") if u < 0.90 else "glass")))
    # placement on a jittered grid over the floor, away from the camera corner
    gx, gz = np.meshgrid(np.linspace(-2.0, 2.4, 7), np.linspace(-0.8, 2.6, 4), indexing="ij")
    slots = np.stack([gx.ravel(), gz.ravel()], 1)
    order = np.resize(rng.permutation(len(slots)), n_obj)     # (more objects than slots only for tiny n_tris: slots are reused)
    for i, kind in enumerate(kinds):
        cx, cz = slots[order[i]] + rng.uniform(-0.12, 0.12, 2)
        rot = rng.uniform(0, 360)
        if kind == "ico":
            r = rng.uniform(0.22, 0.32)
            pos, nrm, uv, faces = displaced_icosphere(ico_level, r, 0.08, float(rng.uniform(0, 6)))
            y = r * 1.15 + rng.uniform(0.0, 0.9)
        elif kind == "torus":
            R, r = rng.uniform(0.2, 0.28), rng.uniform(0.06, 0.09)
            pos, nrm, uv, faces = torus(tor_nu, tor_nv, R, r)
            y = r + 0.02 + rng.uniform(0.0, 1.2)
        else:
            h = rng.uniform(0.15, 0.3, 3)
            pos, nrm, uv, faces = grid_box(box_n, h)
            y = h[1] + 0.01
        to_world = T().translate([cx, y, cz]).rotate([0, 1, 0], rot)
        if kind == "torus":
            to_world = to_world.rotate([1, 0, 0], rng.uniform(0, 60))
        d[f"obj{i:02d}_{kind}"] = {"type": "mesh", "positions": pos, "normals": nrm, "texcoords": uv, "faces": faces,
                                  "to_world": to_world, "bsdf": ref(mats[i])}
    if envmap:
        d["env"] = {"type": "envmap", "data": sky_envmap(), "scale": 1.0,
                    "to_world": T().rotate([0, 1, 0], 90.0)}
    return d


def random_triangles(n_tris=10_000_000, width=1920, height=1080, spp=128, seed=3):
    """C4 stress scene: n_tris random triangles in [-1, 1]^3, edge length log-uniform in [0.002, 0.02], 8 albedos by
    triangle id & 7, one 0.5 x 0.5 rectangle light at y = 0.99 (radiance 50), camera (0, 0, 3.5) fov 40, gaussian filter."""
    rng = np.random.RandomState(seed)
    d = {"type": "scene", "integrator": {"type": "path", "max_depth": 8},
         "sensor": {"type": "perspective", "fov": 40.0, "near_clip": 0.01, "far_clip": 100.0,
                    "to_world": T().look_at([0, 0, 3.5], [0, 0, 0], [0, 1, 0]),
                    "sampler": {"type": "independent", "sample_count": spp},
                    "film": {"type": "hdrfilm", "width": width, "height": height, "rfilter": {"type": "gaussian"}}},
         "light": {"type": "rectangle", "to_world": T().translate([0, 0.99, 0]).rotate([1, 0, 0], 90).scale(0.25),
                   "emitter": {"type": "area", "radiance": {"type": "rgb", "value": [50, 50, 50]}}}}
    albedos = rng.uniform(0.2, 0.9, size=(8, 3))
    per = n_tris // 8
    for k in range(8):
        n = per if k < 7 else n_tris - 7 * per
        c = rng.uniform(-1, 1, size=(n, 1, 3)).astype(np.float32)
        edge = np.exp(rng.uniform(np.log(0.002), np.log(0.02), size=(n, 1, 1))).astype(np.float32)
        v = c + (rng.normal(size=(n, 3, 3)).astype(np.float32) * edge * 0.5)
        d[f"soup{k}"] = {"type": "mesh", "positions": v.reshape(-1, 3), "faces": np.arange(3 * n, dtype=np.uint32).reshape(-1, 3),
                         "bsdf": {"type": "twosided", "nested": {"type": "diffuse", "reflectance": {"type": "rgb", "value": albedos[k].tolist()}}}}
    return d


def woodfloor_texture() -> np.ndarray:
    """The reference's ``WoodFloor.jpg`` (960 x 870) as the fp32 linear-RGB tensor Mitsuba would hold for an optimised
    ``bitmap`` parameter (sRGB -> linear, bitmap.cpp:286-294; fp32 because it is a differentiable tensor, :208-218).
    The pixels come from the package's data file ``data/woodfloor_u8.npz`` (decoded from the reference checkout by
    ``tests/golden/make_woodfloor.py``); without the fixture a procedural stand-in of the same shape is returned."""
    from .scene import srgb_to_linear
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "data", "woodfloor_u8.npz")
    if os.path.exists(path):
        rgb = np.load(path)["rgb"]
        return np.ascontiguousarray(srgb_to_linear(rgb.astype(np.float32) / np.float32(255.0)))
    return np.ascontiguousarray(value_noise_rgb(960, 7, [0.6, 0.45, 0.3]).astype(np.float32)[:870])


def cornell_prb_scene(width=512, height=512, spp=32) -> dict:
    """BASELINE config 3: the Cornell box whose ``white`` BSDF carries the WoodFloor texture as its reflectance."""
    d = cornell_box(width, height, spp)
    d["white"] = {"type": "diffuse", "reflectance": {"type": "bitmap", "data": woodfloor_texture(), "raw": True}}
    return d

