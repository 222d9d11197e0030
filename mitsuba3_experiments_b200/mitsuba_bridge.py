"""Front door 1: plug the CUDA integrators into a live Mitsuba 3 (only reachable when ``import mitsuba`` works).

``register_with_mitsuba(mi)`` subclasses ``mi.SamplingIntegrator`` and registers the result through ``mi.register_integrator``
exactly as path.py:305 / path-mis.py:158 do (``register_typed_plugin``, mitsuba3/src/render/python/scene_v.cpp:26-43,166-171).
The overridden trampoline slots are ``render`` / ``render_forward`` / ``render_backward`` / ``sample`` / ``aov_names`` /
``to_string`` (integrator_v.cpp:59-124,286-371); film and parameter buffers cross via DLPack.

``scene_to_dict`` flattens a live ``mi.Scene`` into the dictionary schema ``scene.SceneDesc`` understands.  What a live scene
exposes to Python (Mitsuba 3.6): geometry through ``Mesh.vertex_positions_buffer() / vertex_normals_buffer() /
vertex_texcoords_buffer() / faces_buffer()`` (shape_v.cpp:372-375; rectangles and cubes are meshes, scene_v.cpp:119-131),
everything else through ``mi.traverse(object)`` with the key names each plugin registers in its ``traverse()``
(diffuse.cpp:97, twosided.cpp:106-107, mask.cpp:119-120, roughconductor.cpp:214-224, conductor.cpp:242-244,
dielectric.cpp:238-242, roughdielectric.cpp:223-234, roughplastic.cpp:209-214, bitmap.cpp:397-398, area.cpp:80,
envmap.cpp:240-242, constant.cpp:71, point.cpp:89-90, perspective.cpp:157-160, film.cpp:57-59; pattern of
reproject.py:9-21), ``Object.class_name()`` for the plugin type, and -- for the three switches no parameter exposes
(microfacet ``distribution``, ``sample_visible``, rough plastic ``nonlinear``) -- the plugin's ``to_string()``
(roughconductor.cpp / roughdielectric.cpp / roughplastic.cpp ``to_string``).  A bitmap's ``wrap_mode`` / ``filter_type`` are not
observable from Python at all (bitmap.cpp:396-399 exposes ``data`` and ``to_uv`` only): the plugin defaults (repeat, bilinear) are
assumed.  Anything outside the hot-path scope raises instead of being dropped.

A ``DeviceScene`` is cached on the ``mi.Scene``; before every render the current reflectance parameters are re-read from the
scene and pushed to the device (``SceneParameters.update()`` mutates the scene objects in an optimisation loop), so renders
and gradients are never taken at stale values.  Geometry, emitters and the sensor are read once.

Mitsuba is not installed on the build or GPU box: this module is exercised through the recording fake in
``tests/fake_mitsuba.py`` (``tests/test_host.py`` on CPU, ``tests/test_gpu_parity.py::test_mitsuba_front_door`` on the GPU).
"""
from __future__ import annotations

import re

import numpy as np


def _np(x, dtype=np.float32):
    if hasattr(x, "numpy"):
        x = x.numpy()
    return np.array(x, dtype=dtype)


def _class_name(obj) -> str:
    if hasattr(obj, "class_name"):
        return str(obj.class_name())
    if hasattr(obj, "class_"):                       # Mitsuba < 3.6
        return str(obj.class_().name())
    return type(obj).__name__


def _strip(params, prefix: str) -> dict:
    n = len(prefix)
    return {k[n:]: params[k] for k in params.keys() if k.startswith(prefix)}


def _spectrum(p: dict, name: str, default=None):
    """A texture slot as the scene dictionary wants it: rgb constant or raw bitmap tensor."""
    if f"{name}.data" in p:
        d = {"type": "bitmap", "data": _np(p[f"{name}.data"]), "raw": True}
        if f"{name}.to_uv" in p:
            m = p[f"{name}.to_uv"]
            d["to_uv"] = _np(getattr(m, "matrix", m), np.float64).reshape(3, 3)
        return d
    if f"{name}.value" in p:
        v = _np(p[f"{name}.value"]).reshape(-1)
        return {"type": "rgb", "value": (v[:3] if v.size >= 3 else np.repeat(v[:1], 3)).tolist()}
    if name in p:
        v = _np(p[name]).reshape(-1)
        return {"type": "rgb", "value": (v[:3] if v.size >= 3 else np.repeat(v[:1], 3)).tolist()}
    return default


def _scalar(p: dict, name: str, default=None):
    for k in (name, f"{name}.value"):
        if k in p:
            return float(_np(p[k], np.float64).reshape(-1)[0])
    return default


def _switch(text: str, name: str, default: str) -> str:
    m = re.search(rf"\b{name}\s*=\s*([A-Za-z0-9_]+)", text or "")
    return m.group(1).lower() if m else default


def bsdf_params_to_dict(p: dict, text: str = "", name: str = "") -> dict:
    """One BSDF from its (prefix-stripped) ``mi.traverse`` keys; the plugin type follows from the key set."""
    if any(k.startswith("nested_bsdf.") for k in p):                      # mask.cpp:119-120
        op = _spectrum(p, "opacity", {"type": "rgb", "value": [0.5] * 3})
        if op.get("type") == "bitmap":
            raise RuntimeError(f"b2path: BSDF {name!r}: textured mask opacity is outside the hot-path scope")
        return {"type": "mask", "opacity": float(np.mean(op["value"])), "nested": bsdf_params_to_dict(_strip(p, "nested_bsdf."), text, name)}
    if any(k.startswith("brdf_0.") for k in p):                          # twosided.cpp:106-107 (one nested BRDF for both sides)
        a, b = _strip(p, "brdf_0."), _strip(p, "brdf_1.")
        if b and set(a) != set(b):
            raise RuntimeError(f"b2path: BSDF {name!r}: twosided with two different BRDFs is outside the hot-path scope")
        return {"type": "twosided", "nested": bsdf_params_to_dict(a, text, name)}
    keys = {k.split(".")[0] for k in p}
    dist = _switch(text, "distribution", "beckmann")
    vis = _switch(text, "sample_visible", "1") not in ("0", "false")
    def alpha(d):
        if "alpha_u" in keys:
            d["alpha_u"], d["alpha_v"] = _scalar(p, "alpha_u"), _scalar(p, "alpha_v")
        else:
            d["alpha"] = _scalar(p, "alpha", 0.1)
        d["distribution"], d["sample_visible"] = dist, vis
        return d
    if "reflectance" in keys:
        return {"type": "diffuse", "reflectance": _spectrum(p, "reflectance")}
    if "diffuse_reflectance" in keys:                                     # roughplastic.cpp:209-214
        d = alpha({"type": "roughplastic", "diffuse_reflectance": _spectrum(p, "diffuse_reflectance"),
                   "int_ior": _scalar(p, "eta", 1.49), "ext_ior": 1.0, "nonlinear": _switch(text, "nonlinear", "0") not in ("0", "false")})
        if "specular_reflectance" in keys:
            d["specular_reflectance"] = _spectrum(p, "specular_reflectance")
        return d
    if "k" in keys:
        d = {"eta": _spectrum(p, "eta")["value"], "k": _spectrum(p, "k")["value"]}
        if "specular_reflectance" in keys:
            d["specular_reflectance"] = _spectrum(p, "specular_reflectance")
        if "alpha" in keys or "alpha_u" in keys:
            return alpha(dict(d, type="roughconductor"))
        return dict(d, type="conductor")
    if "eta" in keys:
        d = {"int_ior": _scalar(p, "eta"), "ext_ior": 1.0}
        if "specular_transmittance" in keys:
            d["specular_transmittance"] = _spectrum(p, "specular_transmittance")
        if "specular_reflectance" in keys:
            d["specular_reflectance"] = _spectrum(p, "specular_reflectance")
        if "alpha" in keys or "alpha_u" in keys:
            return alpha(dict(d, type="roughdielectric"))
        return dict(d, type="dielectric")
    raise RuntimeError(f"b2path: BSDF {name!r} with parameters {sorted(keys)} is outside the hot-path scope")


def _rfilter(rf) -> dict:
    cn = _class_name(rf).lower()
    r = float(rf.radius())
    if "box" in cn:
        return {"type": "box"}
    if "tent" in cn:
        return {"type": "tent", "radius": r}
    if "gaussian" in cn:
        return {"type": "gaussian", "stddev": r / 4.0}       # gaussian.cpp:48-55: radius = 4 sigma
    raise RuntimeError(f"b2path: reconstruction filter {cn!r} is outside the hot-path scope")


def scene_to_dict(scene, mi=None) -> dict:
    if mi is None:
        import mitsuba as mi  # type: ignore
    out: dict = {"type": "scene"}
    sensors = scene.sensors()
    if len(sensors) != 1:
        raise RuntimeError("b2path: exactly one sensor is supported")
    sensor = sensors[0]
    if "perspective" not in _class_name(sensor).lower():
        raise RuntimeError(f"b2path: sensor {_class_name(sensor)!r} is outside the hot-path scope (perspective only)")
    sp = mi.traverse(sensor)
    film = sensor.film()
    size, crop, off = film.size(), film.crop_size(), film.crop_offset()
    tw = sp["to_world"]
    out["sensor"] = {
        "type": "perspective", "fov": float(_np(sp["x_fov"], np.float64).reshape(-1)[0]), "fov_axis": "x",
        "near_clip": float(sensor.near_clip()), "far_clip": float(sensor.far_clip()),
        "to_world": _np(getattr(tw, "matrix", tw), np.float64).reshape(4, 4),
        "sampler": {"type": "independent", "sample_count": int(sensor.sampler().sample_count())},
        "film": {"type": "hdrfilm", "width": int(size[0]), "height": int(size[1]), "crop_width": int(crop[0]),
                 "crop_height": int(crop[1]), "crop_offset_x": int(off[0]), "crop_offset_y": int(off[1]),
                 "sample_border": bool(film.sample_border()), "rfilter": _rfilter(film.rfilter())},
    }
    # Emitter order matters: Scene::sample_emitter picks emitter floor(u * n) (scene.cpp:252-275), so our flat table must list the
    # emitters as scene.emitters() does.  SceneDesc appends an area light when it ingests its shape, hence the dictionary is
    # emitted in emitter order: each area light's shape at its emitter's position, the non-emitting shapes afterwards.
    shapes = list(scene.shapes())
    shape_ids = [(s.id() or f"shape{i}") for i, s in enumerate(shapes)]
    seen_bsdf = {}

    def add_shape(i):
        shape, sid = shapes[i], shape_ids[i]
        if _class_name(shape) == "Rectangle":
            # rectangle.cpp:85,110: a Mesh subclass with shape_type Rectangle (is_mesh() is False) whose emitter sampling is the
            # closed form of rectangle.cpp:182-200, not the mesh CDF: exported as a rectangle so that the samples match
            tw = mi.traverse(shape)["to_world"]
            entry = {"type": "rectangle", "to_world": _np(getattr(tw, "matrix", tw), np.float64).reshape(4, 4)}
        elif hasattr(shape, "vertex_positions_buffer") and (shape.is_mesh() or _class_name(shape) in ("Cube",)):
            entry = {"type": "mesh",
                     "positions": _np(shape.vertex_positions_buffer()).reshape(-1, 3),
                     "faces": _np(shape.faces_buffer(), np.uint32).reshape(-1, 3)}
            if shape.has_vertex_normals():
                entry["normals"] = _np(shape.vertex_normals_buffer()).reshape(-1, 3)
            if shape.has_vertex_texcoords():
                entry["texcoords"] = _np(shape.vertex_texcoords_buffer()).reshape(-1, 2)
        else:
            raise RuntimeError(f"b2path: shape {sid!r} ({_class_name(shape)}) is not a triangle mesh (hot-path scope is triangles only)")
        bsdf = shape.bsdf()
        bid = bsdf.id() or f"{sid}.bsdf"
        if bid not in seen_bsdf:                         # shared BSDFs become one top-level entry + refs (util.py:646-700 pattern)
            out[bid] = seen_bsdf[bid] = bsdf_params_to_dict(dict(mi.traverse(bsdf).items()), str(bsdf), bid)
        entry["bsdf"] = {"type": "ref", "id": bid}
        if shape.is_emitter():
            em = shape.emitter()
            if _class_name(em) != "AreaLight":
                raise RuntimeError(f"b2path: emitter {_class_name(em)!r} on shape {sid!r} is outside the hot-path scope")
            entry["emitter"] = {"type": "area", "radiance": _spectrum(dict(mi.traverse(em).items()), "radiance")}
        out[sid] = entry

    done = set()
    for i, em in enumerate(scene.emitters()):
        cn = _class_name(em)
        if cn == "AreaLight":
            owner = [j for j, s in enumerate(shapes) if s.is_emitter() and s.emitter() is em]
            if len(owner) != 1:
                raise RuntimeError("b2path: an area light without (or with more than) one shape")
            add_shape(owner[0]); done.add(owner[0])
            continue
        ep = dict(mi.traverse(em).items())
        eid = em.id() or f"emitter{i}"
        if cn == "EnvironmentMapEmitter":                # envmap.cpp:240-242: data is [H, W + 1, C] (the first column repeated)
            data = _np(ep["data"])
            tw = ep["to_world"]
            out[eid] = {"type": "envmap", "data": np.ascontiguousarray(data[:, :-1, :3]), "scale": _scalar(ep, "scale", 1.0),
                        "to_world": _np(getattr(tw, "matrix", tw), np.float64).reshape(4, 4)}
        elif cn == "ConstantBackgroundEmitter":
            out[eid] = {"type": "constant", "radiance": _spectrum(ep, "radiance")}
        elif cn == "PointLight":
            out[eid] = {"type": "point", "position": _np(ep["position"], np.float64).reshape(3).tolist(), "intensity": _spectrum(ep, "intensity")}
        else:
            raise RuntimeError(f"b2path: emitter {cn!r} is outside the hot-path scope")
    for j in range(len(shapes)):
        if j not in done:
            add_shape(j)
    return out


def _bsdf_value_candidates(key: str):
    """Our parameter key '<bsdf>.reflectance.value' -> the keys mi.traverse(scene) may use for it (twosided / mask nesting)."""
    bid, rest = key.split(".", 1)
    slots = [rest] + [rest.replace("reflectance", s, 1) for s in ("specular_reflectance", "diffuse_reflectance")]
    for nest in ("", "brdf_0.", "nested_bsdf.", "nested_bsdf.brdf_0."):
        for s in slots:
            yield f"{bid}.{nest}{s}"


def map_param_keys(ours, theirs) -> dict:
    """{our key: mitsuba key} for every parameter of ours that the live scene also exposes."""
    out = {}
    for k in ours:
        for c in _bsdf_value_candidates(k):
            if c in theirs:
                out[k] = c
                break
    return out


def device_scene_for(scene, mi=None):
    """The cached ``DeviceScene`` of a live ``mi.Scene``, with its reflectance parameters refreshed from the scene."""
    from .integrator import traverse
    from .runtime import DeviceScene
    if mi is None:
        import mitsuba as mi  # type: ignore
    ds = getattr(scene, "_b2_device_scene", None)
    if ds is None:
        ds = DeviceScene(scene_to_dict(scene, mi))
        ds._b2_params = traverse(ds)
        try:
            scene._b2_device_scene = ds
        except Exception:
            pass
        return ds
    live = mi.traverse(scene)
    ours = ds._b2_params
    for k, mk in map_param_keys(ours.keys(), live).items():
        i, what = ours._index[k]
        v = live[mk]
        ds.set_param(i, what, _np(v).reshape(tuple(ours[k].shape)) if not hasattr(v, "__dlpack__") else v)
    return ds


def register_with_mitsuba(mi):
    """Create ``mi.SamplingIntegrator`` subclasses around the CUDA integrators and register them."""
    from . import integrator as I

    def dr_module():
        d = getattr(mi, "_dr", None)
        if d is not None:
            return d
        import drjit  # type: ignore
        return drjit

    def wrap(base_cls, name):
        class Wrapped(mi.SamplingIntegrator):
            def __init__(self, props):
                super().__init__(props)
                keys = ("max_depth", "rr_depth", "samples_per_pass", "kernel", "env_warp")
                self._impl = base_cls({k: props.get(k) for k in keys if props.has_property(k)})

            def _ds(self, scene):
                return device_scene_for(scene, mi)

            def render(self, scene, sensor=0, seed=0, spp=0, develop=True, evaluate=True):
                img = self._impl.render(self._ds(scene), 0, seed, spp, develop, evaluate)
                return mi.TensorXf(img)          # DLPack: zero-copy torch -> Dr.Jit

            def _params(self, ds, params):
                p = I.traverse(ds)
                kmap = map_param_keys(p.keys(), params)
                p.keep(list(kmap))
                return p, kmap

            def render_forward(self, scene, params, sensor=0, seed=0, spp=0):
                if not hasattr(self._impl, "render_forward"):
                    raise RuntimeError(f"{name}: render_forward needs the PRB integrator (b2prb)")
                ds = self._ds(scene)
                p, kmap = self._params(ds, params)
                dr = dr_module()
                for k, mk in kmap.items():               # the tangent of every parameter is its forward gradient (dr.set_grad)
                    t = dr.grad(params[mk])
                    p.set_tangent(k, I.as_device_tensor(t, ds.device) if hasattr(t, "__dlpack__") else _np(t))
                return mi.TensorXf(self._impl.render_forward(ds, p, 0, seed, spp))

            def render_backward(self, scene, params, grad_in, sensor=0, seed=0, spp=0):
                if not hasattr(self._impl, "render_backward"):
                    raise RuntimeError(f"{name}: render_backward needs the PRB integrator (b2prb)")
                ds = self._ds(scene)
                p, kmap = self._params(ds, params)
                self._impl.render_backward(ds, p, grad_in, 0, seed, spp)
                dr = dr_module()
                for k, mk in kmap.items():               # side effect of the slot: accumulate into dr.grad(params[key])
                    dr.accum_grad(params[mk], type(params[mk])(p.grad[k]))

            def sample(self, scene, sampler, ray, medium=None, active=True):
                return self._impl.sample(scene, sampler, ray, medium, active)

            def aov_names(self):
                return self._impl.aov_names()

            def to_string(self):
                return self._impl.to_string()

        Wrapped.__name__ = f"B2{base_cls.__name__}"
        mi.register_integrator(name, lambda props: Wrapped(props))
        return Wrapped

    return {n: wrap(c, n) for n, c in (("b2path", I.Path), ("b2path_mis", I.PathIntegrator), ("b2prb", I.PRBIntegrator))}
