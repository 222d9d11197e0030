"""Drop-in integrators: the host-side mirror of the reference's plugin interface for the hot path.

Reference interface mirrored (file:line relative to /root/reference):

* ``Path.__init__(props)`` / ``Path.sample`` / registration under a name    path.py:21-25,194-305
* ``PathIntegrator`` (``max_depth=8``, ``rr_depth=2`` defaults)               path-mis.py:18-22,158
* ``PRBIntegrator`` / ``RBIntegrator.render_backward`` (``max_depth=6``,       mitsuba3/src/python/python/ad/integrators/
  ``rr_depth=5`` defaults)                                                   common.py:28-40,623-781, prb.py:63-265
* ``Integrator.render(scene, sensor, seed, spp, develop, evaluate)``         mitsuba3/src/render/python/integrator_v.cpp:286-308
* ``Integrator.render_backward(scene, params, grad_in, sensor, seed, spp)``  integrator_v.cpp:348-371
* ``mi.render`` custom op (``seed_grad = sample_tea_32(seed, 1)[0]``)         mitsuba3/src/python/python/util.py:343-526
* ``mi.traverse`` -> ``SceneParameters`` (``<bsdf>.reflectance.value|data``)  util.py:271-337, bitmap.cpp:396-399

Two front doors onto one back door (the C-ABI): when ``mitsuba`` is importable the classes below are registered
through ``mi.register_integrator`` and accept a live ``mi.Scene``; otherwise (no Mitsuba on this box) they accept
the scene *dictionary* itself (same schema as ``mi.cornell_box()``).  There is no CPU fallback: without the CUDA
extension every call raises.
"""
from __future__ import annotations

import numpy as np

from . import capi
from .runtime import DeviceScene, as_device_tensor
from .scene import SceneDesc


def sample_tea_32(v0: int, v1: int, rounds: int = 4):
    """mitsuba3/include/mitsuba/core/random.h:77-90 (host copy used only to derive ``seed_grad``)."""
    s, m = 0, 0xffffffff
    for _ in range(rounds):
        s = (s + 0x9e3779b9) & m
        v0 = (v0 + ((((v1 << 4) & m) + 0xa341316c) ^ ((v1 + s) & m) ^ ((v1 >> 5) + 0xc8013ea4))) & m
        v1 = (v1 + ((((v0 << 4) & m) + 0xad90777d) ^ ((v0 + s) & m) ^ ((v0 >> 5) + 0x7e95761e))) & m
    return v0, v1


class SceneParameters(dict):
    """``mi.traverse(scene)`` stand-in: differentiable reflectance parameters keyed like Mitsuba's
    (``'<bsdf id>.reflectance.value'`` for constants, ``'<bsdf id>.reflectance.data'`` for bitmaps)."""

    def __init__(self, scene: "DeviceScene"):
        super().__init__()
        import torch
        self.scene = scene
        self.grad: dict[str, "torch.Tensor"] = {}
        self.tangent: dict[str, "torch.Tensor"] = {}      # forward mode: d(param), see set_tangent / render_forward
        self._index: dict[str, tuple[int, str]] = {}
        d = scene.desc
        # the differentiable reflectance slot of every BSDF type, under the name its plugin registers (diffuse.cpp:97,
        # roughconductor.cpp:214, roughplastic.cpp:209, roughdielectric.cpp:232); delta BSDFs have no eval() and no PRB gradient
        slots = {capi.BSDF_DIFFUSE: "reflectance", capi.BSDF_ROUGHCONDUCTOR: "reflectance", capi.BSDF_ROUGHPLASTIC: "diffuse_reflectance",
                 capi.BSDF_ROUGHDIELECTRIC: "specular_reflectance"}
        for i, (name, b) in enumerate(zip(d.bsdf_names, d.bsdfs)):
            if b.type not in slots:
                continue
            slot = slots[b.type]
            if b.reflectance_tex >= 0:
                arr = d.texture_arrays[b.reflectance_tex]
                if arr.dtype != np.float32:
                    continue                          # fp16 storage is not a differentiable parameter (bitmap.cpp:208-218)
                key, val = f"{name}.{slot}.data", torch.from_numpy(arr.copy())
                self._index[key] = (i, "data")
            else:
                key, val = f"{name}.{slot}.value", torch.tensor(list(b.reflectance), dtype=torch.float32)
                self._index[key] = (i, "value")
            super().__setitem__(key, val.to(scene.device))

    def keep(self, keys):
        keys = [keys] if isinstance(keys, str) else list(keys)
        for k in list(self.keys()):
            if k not in keys:
                super().__delitem__(k)

    def update(self, values: dict | None = None):  # noqa: A003 - mirrors SceneParameters.update
        """Push (changed) parameters to the device scene."""
        if values:
            for k, v in values.items():
                self[k] = v
        for k, v in self.items():
            i, what = self._index[k]
            self.scene.set_param(i, what, v.detach())

    def __setitem__(self, key, value):
        if key not in self._index:
            raise KeyError(f"unknown scene parameter {key!r}")
        super().__setitem__(key, as_device_tensor(value, self.scene.device))

    def targets(self):
        """(bsdf index, kind, grad tensor) triples for ``b2_render_backward``; allocates zeroed grads on first use."""
        import torch
        out = []
        for k, v in self.items():
            i, what = self._index[k]
            if k not in self.grad:
                self.grad[k] = torch.zeros_like(v.detach(), dtype=torch.float32).contiguous()
            out.append((i, 1 if what == "data" else 0, self.grad[k]))
        return out

    def set_tangent(self, key, value):
        """Forward mode: the direction d(param) along which ``render_forward`` differentiates (``dr.set_grad`` + ``dr.forward``)."""
        import torch
        if key not in self._index or key not in self:
            raise KeyError(f"unknown scene parameter {key!r}")
        t = as_device_tensor(value, self.scene.device, torch.float32)
        self.tangent[key] = torch.broadcast_to(t, self[key].shape).contiguous()

    def tangents(self):
        out = []
        for k, t in self.tangent.items():
            i, what = self._index[k]
            out.append((i, 1 if what == "data" else 0, t))
        return out

    def zero_grad(self):
        for g in self.grad.values():
            g.zero_()


def traverse(scene: "DeviceScene") -> SceneParameters:
    return SceneParameters(scene)


def load_dict(scene_dict: dict, device: int = 0, base_path: str = ".") -> DeviceScene:
    """``mi.load_dict`` stand-in for the plugin types on the hot path."""
    return DeviceScene(scene_dict, device=device, base_path=base_path)


def load_file(path: str, device: int = 0, **params) -> DeviceScene:
    """``mi.load_file(path, **params)`` stand-in: the XML subset of xml_loader.py; keyword arguments override ``<default>``
    parameters (``spp=16, resx=640`` ...), relative file names resolve against the directory of ``path``."""
    import os
    from .xml_loader import load_file as _xml
    return DeviceScene(_xml(path, **params), device=device, base_path=os.path.dirname(os.path.abspath(path)))


class _Props(dict):
    """Minimal ``mi.Properties``: ``props.get(name, default)``."""


class B2Integrator:
    """Common driver (SamplingIntegrator::render, integrator.cpp:122-367): derives passes, launches, develops."""

    variant = capi.VARIANT_PATH
    default_max_depth, default_rr_depth = 8, 4

    def __init__(self, props=None):
        props = props or {}
        get = props.get
        self.max_depth = int(get("max_depth", self.default_max_depth))
        self.rr_depth = int(get("rr_depth", self.default_rr_depth))
        self.samples_per_pass = int(get("samples_per_pass", 0) or 0)
        self.kernel = {"wavefront": capi.KERNEL_WAVEFRONT, "megakernel": capi.KERNEL_MEGAKERNEL}[get("kernel", "wavefront")]
        self.env_warp = {"hierarchical": capi.ENV_WARP_HIERARCHICAL, "alias": capi.ENV_WARP_ALIAS}[get("env_warp", "hierarchical")]
        if self.max_depth < 0:
            if self.max_depth != -1:
                raise RuntimeError('"max_depth" must be set to -1 (infinite) or a value >= 0')
            self.max_depth = 1 << 20     # -1 = unbounded (integrator.cpp:505-515): Russian roulette ends every path
        if self.rr_depth <= 0:
            raise RuntimeError('"rr_depth" must be set to a value greater than zero!')

    # -- the reference records sample() symbolically once and JIT-compiles it; here the loop is a CUDA kernel
    def sample(self, scene, sampler, ray, medium=None, active=True):
        raise RuntimeError("b2path integrators run the whole render loop in CUDA kernels: call render() "
                           "(SamplingIntegrator.render is the overridden slot, integrator_v.cpp:72-79)")

    def aov_names(self):
        return []

    def to_string(self):
        return f"{type(self).__name__}[max_depth = {self.max_depth}, rr_depth = {self.rr_depth}]"

    @staticmethod
    def _scene(scene) -> DeviceScene:
        if isinstance(scene, DeviceScene):
            return scene
        if isinstance(scene, (dict, SceneDesc)):
            return DeviceScene(scene)
        from .mitsuba_bridge import device_scene_for    # live mi.Scene (only reachable when Mitsuba is installed): cached per scene,
        return device_scene_for(scene)                  # reflectance parameters re-read on every call

    def _args(self, scene: DeviceScene, seed, spp, **kw):
        return dict(seed=seed, spp=spp or scene.desc.spp, max_depth=self.max_depth, rr_depth=self.rr_depth,
                    variant=self.variant, spp_per_pass=self.samples_per_pass, env_warp=self.env_warp, kernel=self.kernel, **kw)

    def render(self, scene, sensor=0, seed=0, spp=0, develop=True, evaluate=True, film=None, lane_begin=0, lane_end=0,
               counters=False, sample_begin=0, sample_count=0):
        """-> [H, W, 3] float32 torch tensor on the device (``develop=True``) or the RGBW film (``develop=False``).

        ``lane_begin/lane_end`` (image bands) or ``sample_begin/sample_count`` (a slice of every pixel's samples) restrict
        the render to a shard of the wavefront (multi-GPU, see ``dist.py``)."""
        if sensor not in (0, None):
            raise RuntimeError("only sensor 0 is supported")
        ds = self._scene(scene)
        film, cnt = ds.render_film(film=film, counters=counters, **self._args(ds, seed, spp, lane_begin=lane_begin, lane_end=lane_end,
                                                                              sample_begin=sample_begin, sample_count=sample_count))
        self.last_counters = cnt
        out = ds.develop(film) if develop else film
        if evaluate:
            import torch
            torch.cuda.current_stream(ds.device).synchronize()
        return out


class Path(B2Integrator):
    """``mypath`` of path.py (defaults path.py:23-24: max_depth 8, rr_depth 4)."""
    variant = capi.VARIANT_PATH
    default_max_depth, default_rr_depth = 8, 4


class PathIntegrator(B2Integrator):
    """``path_test`` of path-mis.py == built-in ``path`` (defaults path-mis.py:21-22: max_depth 8, rr_depth 2)."""
    variant = capi.VARIANT_PATH_MIS
    default_max_depth, default_rr_depth = 8, 2


class PRBIntegrator(B2Integrator):
    """``prb`` (defaults common.py:28-40: max_depth 6, rr_depth 5); adds ``render_backward``."""
    variant = capi.VARIANT_PRB
    default_max_depth, default_rr_depth = 6, 5

    def render_backward(self, scene, params: SceneParameters, grad_in, sensor=0, seed=0, spp=0, lane_begin=0, lane_end=0,
                        counters=False, reuse_weights=False, sample_begin=0, sample_count=0):
        """Accumulates d(loss)/d(param) into ``params.grad[key]`` (side effect, like ``dr.grad(params[key])``).
        A shard (lane range or sample slice) adds its share; the weight film is always that of the whole sample set."""
        ds = self._scene(scene)
        a = self._args(ds, seed, spp, lane_begin=lane_begin, lane_end=lane_end, reuse_weights=reuse_weights,
                       sample_begin=sample_begin, sample_count=sample_count)
        a.pop("variant")
        self.last_counters = ds.render_backward(grad_in, params.targets(), counters=counters, **a)
        return None

    def render_forward(self, scene, params: SceneParameters, sensor=0, seed=0, spp=0, counters=False):
        """``RBIntegrator.render_forward`` (common.py:495-621): the gradient image for the parameter tangents set with
        ``params.set_tangent(key, d_value)`` -- one primal pass, one replay that gathers d L per lane, one splat."""
        ds = self._scene(scene)
        a = self._args(ds, seed, spp)
        a.pop("variant")
        film, self.last_counters = ds.render_forward(params.tangents(), counters=counters, **a)
        return ds.develop(film)


_REGISTRY = {"mypath": Path, "path_test": PathIntegrator, "path": PathIntegrator, "prb": PRBIntegrator,
             "b2path": Path, "b2path_mis": PathIntegrator, "b2prb": PRBIntegrator}


def register_integrator(name: str, constructor):
    """``mi.register_integrator`` (scene_v.cpp:166-171): name -> ``constructor(props)``."""
    _REGISTRY[name] = constructor


def make_integrator(spec: dict | str | None, scene: DeviceScene | None = None) -> B2Integrator:
    """Instantiate from ``{'type': name, ...props}`` (what ``mi.load_dict`` does for the ``integrator`` entry)."""
    if spec is None and scene is not None:
        spec = scene.desc.integrator_props or {"type": "path"}
    if isinstance(spec, str):
        spec = {"type": spec}
    spec = dict(spec)
    ctor = _REGISTRY.get(spec.pop("type"))
    if ctor is None:
        raise RuntimeError("unknown integrator plugin")
    return ctor(_Props(spec))


def render(scene, params: SceneParameters | None = None, sensor=0, integrator=None, seed=0, seed_grad=0, spp=0, spp_grad=0):
    """``mi.render`` (util.py:395-526): differentiable w.r.t. the tensors in ``params`` through torch.autograd.

    The primal image uses ``seed``; the adjoint pass uses ``seed_grad`` (default ``sample_tea_32(seed, 1)[0]``,
    util.py:504-509) so that the gradient sample set is independent of the one the loss was computed on."""
    import torch
    ds = B2Integrator._scene(scene)
    integ = integrator or make_integrator(None, ds)
    if seed_grad == 0:
        seed_grad = sample_tea_32(seed, 1)[0]
    spp_grad = spp_grad or spp
    if params is None or not any(v.requires_grad for v in params.values()):
        return integ.render(ds, sensor, seed, spp)
    keys = [k for k, v in params.items() if v.requires_grad]

    class _RenderOp(torch.autograd.Function):
        @staticmethod
        def forward(ctx, *tensors):
            params.update()
            return integ.render(ds, sensor, seed, spp)

        @staticmethod
        def backward(ctx, grad_out):
            params.zero_grad()
            sub = SceneParameters.__new__(SceneParameters)
            dict.__init__(sub, {k: params[k] for k in keys})
            sub.scene, sub.grad, sub._index = params.scene, params.grad, params._index
            integ.render_backward(ds, sub, grad_out.contiguous(), sensor, seed_grad, spp_grad)
            return tuple(params.grad[k].clone() for k in keys)

    return _RenderOp.apply(*[params[k] for k in keys])


# ---- front door 1: register with a live Mitsuba when there is one (never the case on this box; kept import-guarded)
try:  # pragma: no cover
    import mitsuba as _mi  # type: ignore
    if _mi.variant() is not None:
        from .mitsuba_bridge import register_with_mitsuba
        register_with_mitsuba(_mi)
except Exception:  # Mitsuba absent: the standalone classes above are the interface
    _mi = None
