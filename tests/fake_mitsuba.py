"""A recording stand-in for the parts of ``mitsuba`` the front door touches (test infrastructure; Mitsuba itself cannot be
installed on these boxes).  It presents a scene *dictionary* the way a live ``mi.Scene`` presents itself to Python in
Mitsuba 3.6 -- and nothing more:

* ``register_integrator(name, ctor)`` / ``load_dict({'type': name, ...})`` through a plugin table   scene_v.cpp:26-43,166-171
* ``SamplingIntegrator`` base whose ctor takes ``Properties`` (``has_property`` / ``get``)            integrator_v.cpp:59-124
* ``Scene.shapes() / emitters() / sensors()``; ``Mesh.*_buffer()``, ``is_mesh``, ``bsdf``, ``emitter``   shape_v.cpp:372-375
* ``traverse(obj)`` with each plugin's own key names (diffuse.cpp:97, twosided.cpp:106-107, mask.cpp:119-120, ...)
* ``Object.class_name()``, ``id()``, ``to_string`` carrying distribution / sample_visible / nonlinear as the plugins print them
* ``TensorXf`` (DLPack pass-through) and a ``_dr`` module recording ``accum_grad`` / ``grad``

Rectangles, cubes and spheres are handed out as meshes with the tessellation of scene.SceneDesc (rectangle.cpp:102-110,
cube.cpp:105-135): that is what ``scene.shapes()`` returns for them (scene_v.cpp:119-131).
"""
import numpy as np

from mitsuba3_experiments_b200.scene import SceneDesc, _as_matrix, _rgb, parse_fov


class Transform4f:
    def __init__(self, m):
        self.matrix = np.asarray(m, np.float64).reshape(4, 4)


class Properties(dict):
    def has_property(self, k):
        return k in self

    def get(self, k, default=None):
        return dict.get(self, k, default)


class Object:
    _cn = "Object"

    def __init__(self, oid=""):
        self._id = oid

    def id(self):
        return self._id

    def class_name(self):
        return self._cn


class SamplingIntegrator(Object):
    _cn = "SamplingIntegrator"

    def __init__(self, props):
        super().__init__("")
        self.props = props


class TensorXf:
    def __init__(self, x):
        self.array = x

    def torch(self):
        return self.array

    def __dlpack__(self, *a, **k):
        return self.array.__dlpack__(*a, **k)

    def __dlpack_device__(self):
        return self.array.__dlpack_device__()


class Color3f(np.ndarray):
    def __new__(cls, v):
        a = np.asarray(v.detach().cpu().numpy() if hasattr(v, "detach") else v, dtype=np.float32).reshape(-1)[:3]
        return a.view(cls)


class Tensor(np.ndarray):
    def __new__(cls, v):
        a = np.asarray(v.detach().cpu().numpy() if hasattr(v, "detach") else v, dtype=np.float32)
        return a.view(cls)


class _Dr:
    """Records what the integrator does to gradients (dr.accum_grad / dr.grad)."""

    def __init__(self):
        self.grads, self.tangents = {}, {}

    def accum_grad(self, param, value):
        k = id(param)
        self.grads[k] = self.grads.get(k, 0) + np.asarray(value)

    def grad(self, param):
        return self.tangents.get(id(param), np.zeros_like(np.asarray(param)))

    def set_grad(self, param, value):
        self.tangents[id(param)] = np.asarray(value, np.float32)

    def grad_of(self, param):
        return self.grads.get(id(param))


_dr = _Dr()
_plugins = {}
calls = []


def register_integrator(name, ctor):
    calls.append(("register_integrator", name))
    _plugins[name] = ctor


def variant():
    return "cuda_ad_rgb"


# ---------------------------------------------------------------------------------------------- parameter trees
def _texture_params(spec, default):
    if isinstance(spec, dict) and spec.get("type") == "bitmap":
        out = {"data": Tensor(np.asarray(spec["data"], np.float32))}
        out["to_uv"] = np.asarray(spec.get("to_uv", np.eye(3)), np.float64)
        return out
    return {"value": Color3f(_rgb(spec, default))}


def _put(tree, name, sub):
    for k, v in sub.items():
        tree[f"{name}.{k}"] = v


class BSDF(Object):
    def __init__(self, spec, oid):
        super().__init__(oid)
        self.spec = spec
        self.tree, self.text = self._build(spec)
        self._cn = {"diffuse": "SmoothDiffuse", "twosided": "TwoSidedBRDF", "roughconductor": "RoughConductor", "conductor": "SmoothConductor",
                    "dielectric": "SmoothDielectric", "roughdielectric": "RoughDielectric", "roughplastic": "RoughPlastic", "mask": "MaskBSDF"}[spec["type"]]

    @staticmethod
    def _nested(spec):
        return [v for k, v in spec.items() if isinstance(v, dict) and k != "opacity" and "type" in v and v["type"] not in ("rgb", "bitmap", "srgb", "spectrum")][0]

    def _build(self, spec):
        t, tree = spec["type"], {}
        mf = f"  distribution = {spec.get('distribution', 'beckmann')},\n  sample_visible = {int(bool(spec.get('sample_visible', True)))},\n"
        if t == "diffuse":
            _put(tree, "reflectance", _texture_params(spec.get("reflectance"), 0.5))
            return tree, "SmoothDiffuse[\n  reflectance = ...\n]"
        if t == "twosided":
            sub, text = self._build(self._nested(spec))
            _put(tree, "brdf_0", sub); _put(tree, "brdf_1", sub)
            return tree, f"TwoSided[\n  brdf[0] = {text},\n  brdf[1] = {text}\n]"
        if t == "mask":
            sub, text = self._build(self._nested(spec))
            tree["opacity.value"] = np.float32(np.mean(_rgb(spec.get("opacity", 0.5))))
            _put(tree, "nested_bsdf", sub)
            return tree, f"Mask[\n  opacity = ...,\n  nested_bsdf = {text}\n]"
        if t in ("roughconductor", "conductor"):
            if t == "roughconductor":
                if "alpha_u" in spec:
                    tree["alpha_u.value"], tree["alpha_v.value"] = np.float32(spec["alpha_u"]), np.float32(spec["alpha_v"])
                else:
                    tree["alpha.value"] = np.float32(spec.get("alpha", 0.1))
            tree["eta.value"] = Color3f(_rgb(spec.get("eta"), 0.0)); tree["k.value"] = Color3f(_rgb(spec.get("k"), 1.0))
            if "specular_reflectance" in spec:
                _put(tree, "specular_reflectance", _texture_params(spec["specular_reflectance"], 1.0))
            return tree, ("RoughConductor[\n" + mf + "]") if t == "roughconductor" else "SmoothConductor[\n]"
        if t in ("dielectric", "roughdielectric"):
            iors = {"bk7": 1.5046, "air": 1.000277, "vacuum": 1.0, "water": 1.3330, "diamond": 2.419}
            f = lambda v: float(iors[v]) if isinstance(v, str) else float(v)
            tree["eta"] = np.float32(f(spec.get("int_ior", "bk7")) / f(spec.get("ext_ior", "air")))
            if t == "roughdielectric":
                tree["alpha.value"] = np.float32(spec.get("alpha", 0.1))
            for slot in ("specular_reflectance", "specular_transmittance"):
                if slot in spec:
                    _put(tree, slot, _texture_params(spec[slot], 1.0))
            return tree, ("RoughDielectric[\n" + mf + "]") if t == "roughdielectric" else "SmoothDielectric[\n]"
        if t == "roughplastic":
            iors = {"polypropylene": 1.49, "air": 1.000277, "bk7": 1.5046}
            f = lambda v: float(iors[v]) if isinstance(v, str) else float(v)
            _put(tree, "diffuse_reflectance", _texture_params(spec.get("diffuse_reflectance"), 0.5))
            tree["alpha"] = np.float32(spec.get("alpha", 0.1))
            tree["eta"] = np.float32(f(spec.get("int_ior", "polypropylene")) / f(spec.get("ext_ior", "air")))
            if "specular_reflectance" in spec:
                _put(tree, "specular_reflectance", _texture_params(spec["specular_reflectance"], 1.0))
            return tree, "RoughPlastic[\n" + mf + f"  nonlinear = {int(bool(spec.get('nonlinear', False)))}\n]"
        raise ValueError(t)

    def __str__(self):
        return self.text


class Emitter(Object):
    def __init__(self, spec, oid):
        super().__init__(oid)
        t, self.tree = spec["type"], {}
        self._cn = {"area": "AreaLight", "envmap": "EnvironmentMapEmitter", "constant": "ConstantBackgroundEmitter", "point": "PointLight"}[t]
        if t == "area":
            _put(self.tree, "radiance", _texture_params(spec.get("radiance"), 1.0))
        elif t == "constant":
            _put(self.tree, "radiance", _texture_params(spec.get("radiance"), 1.0))
        elif t == "point":
            pos = np.asarray(spec["position"], np.float64) if "position" in spec else _as_matrix(spec.get("to_world"))[:3, 3]
            self.tree["position"] = np.asarray(pos, np.float32)
            _put(self.tree, "intensity", _texture_params(spec.get("intensity"), 1.0))
        elif t == "envmap":
            data = np.asarray(spec["data"], np.float32)
            self.tree["data"] = Tensor(np.concatenate([data, data[:, :1]], axis=1))       # envmap.cpp:147-149: [H, W + 1, C]
            self.tree["scale"] = np.float32(spec.get("scale", 1.0))
            self.tree["to_world"] = Transform4f(_as_matrix(spec.get("to_world")))


class Mesh(Object):
    _cn = "Mesh"

    def __init__(self, oid, pos, nrm, uv, faces, bsdf, emitter):
        super().__init__(oid)
        self._pos, self._nrm, self._uv, self._faces, self._bsdf, self._emitter = pos, nrm, uv, faces, bsdf, emitter

    def is_mesh(self):
        return self._cn != "Rectangle"

    def vertex_positions_buffer(self):
        return self._pos.reshape(-1)

    def faces_buffer(self):
        return self._faces.reshape(-1)

    def has_vertex_normals(self):
        return self._nrm is not None

    def vertex_normals_buffer(self):
        return self._nrm.reshape(-1)

    def has_vertex_texcoords(self):
        return self._uv is not None

    def vertex_texcoords_buffer(self):
        return self._uv.reshape(-1)

    def bsdf(self):
        return self._bsdf

    def is_emitter(self):
        return self._emitter is not None

    def emitter(self):
        return self._emitter


class _RFilter(Object):
    def __init__(self, spec):
        super().__init__("")
        t = spec.get("type", "gaussian")
        self._cn = {"box": "BoxFilter", "tent": "TentFilter", "gaussian": "GaussianFilter"}[t]
        self._r = 0.5 if t == "box" else (float(spec.get("radius", 1.0)) if t == "tent" else 4.0 * float(spec.get("stddev", 0.5)))

    def radius(self):
        return self._r


class _Film(Object):
    _cn = "HDRFilm"

    def __init__(self, spec):
        super().__init__("")
        self.s = spec

    def size(self):
        return (self.s["width"], self.s["height"])

    def crop_size(self):
        return (self.s.get("crop_width", self.s["width"]), self.s.get("crop_height", self.s["height"]))

    def crop_offset(self):
        return (self.s.get("crop_offset_x", 0), self.s.get("crop_offset_y", 0))

    def sample_border(self):
        return bool(self.s.get("sample_border", False))

    def rfilter(self):
        return _RFilter(self.s.get("rfilter", {"type": "gaussian"}))


class _Sampler:
    def __init__(self, n):
        self.n = n

    def sample_count(self):
        return self.n


class Sensor(Object):
    _cn = "PerspectiveCamera"

    def __init__(self, desc: SceneDesc, spec):
        super().__init__("sensor")
        self.spec, self.desc = spec, desc
        film = spec.get("film", {})
        self.tree = {"x_fov": np.float32(parse_fov(spec, int(film.get("width", 768)) / int(film.get("height", 576)))), "to_world": Transform4f(_as_matrix(spec.get("to_world")))}

    def film(self):
        return _Film(self.spec["film"])

    def sampler(self):
        return _Sampler(int(self.spec.get("sampler", {}).get("sample_count", 4)))

    def near_clip(self):
        return float(self.spec.get("near_clip", 1e-2))

    def far_clip(self):
        return float(self.spec.get("far_clip", 1e4))


class Scene(Object):
    """Built from a scene dictionary; hands out what a live mi.Scene would."""
    _cn = "Scene"

    def __init__(self, d):
        super().__init__("")
        self.dict = d
        desc = SceneDesc(d)
        self._sensor = Sensor(desc, d["sensor"])
        named = {k: BSDF(v, k) for k, v in d.items() if isinstance(v, dict) and v.get("type") in
                 ("diffuse", "twosided", "roughconductor", "conductor", "dielectric", "roughdielectric", "roughplastic", "mask")}
        self._shapes, self._emitters = [], []
        from mitsuba3_experiments_b200 import capi
        by_name = dict(zip(desc.mesh_names, desc.meshes))
        for name, spec in d.items():                     # declaration order, as Scene's constructor walks its children (scene.cpp:22-111)
            if not isinstance(spec, dict):
                continue
            if spec.get("type") in ("envmap", "constant", "point"):
                self._emitters.append(Emitter(spec, name))
                continue
            if name not in by_name:
                continue
            m = by_name[name]
            v0, nv, f0, nf = m.first_vertex, m.n_vertices, m.first_face, m.n_faces
            pos = desc.positions[v0:v0 + nv]; faces = desc.faces[f0:f0 + nf] - np.uint32(v0)
            nrm = desc.normals[v0:v0 + nv] if (m.flags & capi.MESH_HAS_NORMALS) else None
            uv = desc.texcoords[v0:v0 + nv] if (m.flags & capi.MESH_HAS_UV) else None
            b = spec.get("bsdf", {"type": "diffuse"})
            bsdf = named[b["id"]] if b.get("type") == "ref" else BSDF(b, "")
            em = Emitter(spec["emitter"], "") if "emitter" in spec else None
            if em is not None:
                self._emitters.append(em)
            shape = Mesh(name, pos, nrm, uv, faces, bsdf, em)
            if spec["type"] == "rectangle":      # rectangle.cpp:85,110: a Mesh subclass whose shape_type is Rectangle (is_mesh() is False)
                shape._cn = "Rectangle"; shape.tree = {"to_world": Transform4f(_as_matrix(spec.get("to_world")))}
                shape._flip = bool(spec.get("flip_normals", False))
            self._shapes.append(shape)

    def sensors(self):
        return [self._sensor]

    def shapes(self):
        return self._shapes

    def emitters(self):
        return self._emitters


class SceneParameters(dict):
    def update(self, values=None):       # noqa: A003 - mirrors mi.SceneParameters.update (the objects ARE the scene's)
        if values:
            for k, v in values.items():
                self[k][...] = np.asarray(v, np.float32).reshape(self[k].shape)


def traverse(obj):
    calls.append(("traverse", obj.class_name()))
    if isinstance(obj, Scene):
        out = SceneParameters()
        for s in obj.shapes():
            b = s.bsdf()
            _put(out, b.id() or f"{s.id()}.bsdf", b.tree)
            if s.is_emitter():
                _put(out, f"{s.id()}.emitter", s.emitter().tree)
        _put(out, "sensor", obj.sensors()[0].tree)
        return out
    return SceneParameters(obj.tree)


def load_dict(d):
    if d.get("type") == "scene":
        return Scene(d)
    props = Properties({k: v for k, v in d.items() if k != "type"})
    calls.append(("instantiate", d["type"]))
    return _plugins[d["type"]](props)
