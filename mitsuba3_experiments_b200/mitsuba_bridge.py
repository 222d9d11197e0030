"""Front door 1: plug the CUDA integrators into a live Mitsuba 3 (only reachable when ``import mitsuba`` works).

``register_with_mitsuba`` subclasses ``mi.SamplingIntegrator`` and registers the result through
``mi.register_integrator`` exactly as path.py:305 / path-mis.py:158 do; ``render`` is the overridden slot
(mitsuba3/src/render/python/integrator_v.cpp:72-79), film and parameter buffers cross via DLPack.
``scene_to_dict`` flattens a live ``mi.Scene`` into the dictionary schema ``scene.SceneDesc`` understands, reading
geometry through ``Mesh.*_buffer()`` (shape_v.cpp:372-375) and everything else through ``mi.traverse``
(pattern of reproject.py:9-21).  Mitsuba is not installed on the build or GPU box, so this module is exercised by
tests only through a recording fake ``mi`` (tests/test_host.py).
"""
from __future__ import annotations

import numpy as np


def _np(x):
    return np.array(x)


def scene_to_dict(scene) -> dict:
    import mitsuba as mi  # type: ignore
    params = mi.traverse(scene)
    out: dict = {"type": "scene"}
    sensor = scene.sensors()[0]
    film = sensor.film()
    size, crop, off = film.size(), film.crop_size(), film.crop_offset()
    rf = film.rfilter()
    rname = type(rf).__name__.lower()
    rfilter = {"type": "box"} if "box" in rname else ({"type": "tent", "radius": float(rf.radius())} if "tent" in rname
                                                       else {"type": "gaussian", "stddev": float(rf.radius()) / 4.0})
    out["sensor"] = {
        "type": "perspective", "fov": float(_np(params["sensor.x_fov"]).reshape(-1)[0]), "fov_axis": "x",
        "near_clip": float(params["sensor.near_clip"]), "far_clip": float(params["sensor.far_clip"]),
        "to_world": _np(params["sensor.to_world"].matrix).reshape(4, 4),
        "sampler": {"type": "independent", "sample_count": int(sensor.sampler().sample_count())},
        "film": {"type": "hdrfilm", "width": int(size[0]), "height": int(size[1]), "crop_width": int(crop[0]),
                 "crop_height": int(crop[1]), "crop_offset_x": int(off[0]), "crop_offset_y": int(off[1]), "rfilter": rfilter},
    }
    for i, shape in enumerate(scene.shapes()):
        sid = shape.id() or f"shape{i}"
        if not shape.is_mesh():
            raise RuntimeError(f"b2path: shape {sid!r} is not a triangle mesh (hot-path scope is triangles only)")
        entry = {"type": "mesh",
                 "positions": _np(shape.vertex_positions_buffer()).reshape(-1, 3),
                 "faces": _np(shape.faces_buffer()).reshape(-1, 3)}
        if shape.has_vertex_normals():
            entry["normals"] = _np(shape.vertex_normals_buffer()).reshape(-1, 3)
        if shape.has_vertex_texcoords():
            entry["texcoords"] = _np(shape.vertex_texcoords_buffer()).reshape(-1, 2)
        entry["bsdf"] = bsdf_to_dict(shape.bsdf(), params)
        if shape.is_emitter():
            entry["emitter"] = {"type": "area", "radiance": _np(params[f"{shape.emitter().id() or sid + '.emitter'}.radiance.value"]).reshape(3)}
        out[sid] = entry
    return out


def bsdf_to_dict(bsdf, params) -> dict:
    bid, name = bsdf.id(), type(bsdf).__name__.lower()

    def refl(key):
        if f"{bid}.{key}.data" in params:
            return {"type": "bitmap", "data": _np(params[f"{bid}.{key}.data"]).astype(np.float32), "raw": True}
        return {"type": "rgb", "value": _np(params[f"{bid}.{key}.value"]).reshape(-1)[:3]}

    if "twosided" in name:
        raise RuntimeError("b2path: pass twosided BSDFs through the dictionary front door")
    if "diffuse" in name:
        return {"type": "diffuse", "id": bid, "reflectance": refl("reflectance")}
    if "roughconductor" in name:
        return {"type": "roughconductor", "id": bid, "alpha": float(params[f"{bid}.alpha.value"]),
                "eta": _np(params[f"{bid}.eta.value"]).reshape(-1)[:3], "k": _np(params[f"{bid}.k.value"]).reshape(-1)[:3],
                "specular_reflectance": refl("specular_reflectance")}
    if "dielectric" in name:
        return {"type": "dielectric", "id": bid, "int_ior": float(params[f"{bid}.eta"]), "ext_ior": 1.0}
    raise RuntimeError(f"b2path: BSDF {name!r} is outside the hot-path scope")


def register_with_mitsuba(mi):
    """Create ``mi.SamplingIntegrator`` subclasses around the CUDA integrators and register them."""
    from . import integrator as I

    def wrap(base_cls, name):
        class Wrapped(mi.SamplingIntegrator):
            def __init__(self, props):
                super().__init__(props)
                self._impl = base_cls({k: props.get(k) for k in ("max_depth", "rr_depth", "samples_per_pass") if props.has_property(k)})

            def render(self, scene, sensor=0, seed=0, spp=0, develop=True, evaluate=True):
                img = self._impl.render(scene, 0, seed, spp, develop, evaluate)
                return mi.TensorXf(img)          # DLPack: zero-copy torch -> Dr.Jit

            def render_backward(self, scene, params, grad_in, sensor=0, seed=0, spp=0):
                ds = I.B2Integrator._scene(scene)
                p = I.traverse(ds)
                p.keep([k for k in p.keys() if k in params])
                self._impl.render_backward(ds, p, grad_in, 0, seed, spp)
                import drjit as dr  # type: ignore
                for k, g in p.grad.items():
                    dr.accum_grad(params[k], type(params[k])(g))

            def sample(self, scene, sampler, ray, medium=None, active=True):
                return self._impl.sample(scene, sampler, ray, medium, active)

            def to_string(self):
                return self._impl.to_string()

        Wrapped.__name__ = f"B2{base_cls.__name__}"
        mi.register_integrator(name, lambda props: Wrapped(props))
        return Wrapped

    return {n: wrap(c, n) for n, c in (("b2path", I.Path), ("b2path_mis", I.PathIntegrator), ("b2prb", I.PRBIntegrator))}
