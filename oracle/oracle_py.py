"""ORACLE loader (test infrastructure, NOT product code).

ctypes access to ``oracle/liboracle.so``: the CPU restatement of the reference path.  Only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s cpu_baseline / ``--impl reference`` legs may import this.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(_HERE, "liboracle.so")
_lib = None


def build(force: bool = False) -> str:
    if force or not os.path.exists(LIB):
        subprocess.run(["make", "-C", _HERE, "liboracle.so"], check=True, capture_output=True)
    return LIB


def lib():
    global _lib
    if _lib is None:
        build()
        from mitsuba3_experiments_b200 import capi
        l = C.CDLL(LIB)
        vp = C.c_void_p
        l.orc_scene_create.argtypes = [C.POINTER(capi.SceneDesc), C.POINTER(vp)]
        l.orc_scene_destroy.argtypes = [vp]; l.orc_scene_destroy.restype = None
        l.orc_render.argtypes = [vp, C.POINTER(capi.RenderArgs), vp, C.POINTER(capi.Counters)]
        l.orc_develop.argtypes = [vp, vp, vp]
        l.orc_render_forward.argtypes = [vp, C.POINTER(capi.RenderArgs), C.POINTER(capi.GradTarget), C.c_int, vp, C.POINTER(capi.Counters)]
        l.orc_render_backward.argtypes = [vp, C.POINTER(capi.RenderArgs), vp, C.POINTER(capi.GradTarget), C.c_int, C.POINTER(capi.Counters)]
        l.orc_trace.argtypes = [vp, vp, C.c_uint64, C.c_int, vp]
        l.orc_last_error.restype = C.c_char_p
        l.orc_rfilter_eval.argtypes = [C.POINTER(capi.Sensor), C.c_float]; l.orc_rfilter_eval.restype = C.c_float
        l.orc_fresnel_conductor.argtypes = [C.c_float] * 3; l.orc_fresnel_conductor.restype = C.c_float
        l.orc_fresnel.argtypes = [C.c_float, C.c_float, vp]; l.orc_fresnel.restype = None
        l.orc_env_warp_eval.argtypes = [C.POINTER(capi.Envmap), C.c_float, C.c_float]; l.orc_env_warp_eval.restype = C.c_float
        l.orc_env_sample.argtypes = [C.POINTER(capi.Envmap), C.c_float, C.c_float, vp]; l.orc_env_sample.restype = None
        l.orc_env_alias_sample.argtypes = [C.POINTER(capi.Envmap), C.c_float, C.c_float, vp]; l.orc_env_alias_sample.restype = None
        l.orc_env_eval_dir.argtypes = [C.POINTER(capi.Envmap), vp, vp]; l.orc_env_eval_dir.restype = None
        l.orc_camera_ray.argtypes = [C.POINTER(capi.Sensor), C.c_float, C.c_float, vp]; l.orc_camera_ray.restype = None
        l.orc_warp.argtypes = [C.c_int, vp, vp]; l.orc_warp.restype = None
        l.orc_tea32.argtypes = [C.c_uint32, C.c_uint32, C.c_int, vp]; l.orc_tea32.restype = None
        l.orc_pcg32_floats.argtypes = [C.c_uint64, C.c_uint64, C.c_int, vp]; l.orc_pcg32_floats.restype = None
        l.orc_pcg32_uints.argtypes = [C.c_uint64, C.c_uint64, C.c_int, vp]; l.orc_pcg32_uints.restype = None
        l.orc_sampler_floats.argtypes = [C.c_uint32, C.c_uint32, C.c_int, vp]; l.orc_sampler_floats.restype = None
        l.orc_microfacet.argtypes = [C.c_int, C.c_float, C.c_float, C.c_int, C.c_int, vp, vp]; l.orc_microfacet.restype = None
        l.orc_rough_tables.argtypes = [C.c_int, C.c_float, C.c_float, vp, vp]; l.orc_rough_tables.restype = None
        l.orc_bsdf_sample.argtypes = [C.POINTER(capi.Bsdf), vp, C.c_float, vp, vp]; l.orc_bsdf_sample.restype = None
        l.orc_bsdf_eval_pdf.argtypes = [C.POINTER(capi.Bsdf), vp, vp, vp]; l.orc_bsdf_eval_pdf.restype = None
        l.orc_texture_eval.argtypes = [C.POINTER(capi.Texture), C.c_float, C.c_float, vp, vp, vp]; l.orc_texture_eval.restype = None
        l.orc_surface_interaction.argtypes = [vp, vp, vp]; l.orc_surface_interaction.restype = None
        l.orc_splat.argtypes = [C.POINTER(capi.Sensor), vp, vp, C.c_uint64, vp]; l.orc_splat.restype = None
        l.orc_debug_lane.argtypes = [vp, C.POINTER(capi.RenderArgs), C.c_uint32, vp, C.c_int, vp]; l.orc_debug_lane.restype = C.c_int
        l.orc_num_threads.restype = C.c_int
        l.orc_set_num_threads.argtypes = [C.c_int]; l.orc_set_num_threads.restype = None
        _lib = l
    return _lib


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p)


class OracleScene:
    """Owns an oracle handle for a ``SceneDesc`` (which must stay alive)."""

    def __init__(self, desc):
        self.desc = desc
        self.h = C.c_void_p()
        if lib().orc_scene_create(C.byref(desc.c), C.byref(self.h)) != 0:
            raise RuntimeError(lib().orc_last_error().decode())

    def __del__(self):
        if getattr(self, "h", None):
            lib().orc_scene_destroy(self.h); self.h = None

    def make_args(self, seed=0, spp=None, max_depth=8, rr_depth=4, variant=0, spp_per_pass=0, lane_begin=0, lane_end=0, kernel=0, all_shadow_rays=False, env_warp=0):
        from mitsuba3_experiments_b200 import capi
        a = capi.RenderArgs()
        a.seed, a.spp, a.spp_per_pass = seed, spp or self.desc.spp, spp_per_pass
        a.max_depth, a.rr_depth, a.variant = max_depth, rr_depth, variant
        a.lane_begin, a.lane_end = lane_begin, lane_end
        a.env_warp = env_warp
        return a

    def render_film(self, **kw):
        from mitsuba3_experiments_b200 import capi
        a = self.make_args(**kw)
        h, w = self.desc.film_shape
        film = np.zeros((h, w, 4), dtype=np.float32)
        cnt = capi.Counters()
        if lib().orc_render(self.h, C.byref(a), _ptr(film), C.byref(cnt)) != 0:
            raise RuntimeError(lib().orc_last_error().decode())
        return film, cnt.as_dict()

    def develop(self, film):
        h, w = self.desc.film_shape
        img = np.zeros((h, w, 3), dtype=np.float32)
        lib().orc_develop(self.h, _ptr(np.ascontiguousarray(film)), _ptr(img))
        return img

    def render(self, **kw):
        film, cnt = self.render_film(**kw)
        return self.develop(film), cnt

    def render_backward(self, grad_in, targets, **kw):
        """targets: list of (bsdf_index, kind, grad ndarray float32).  Gradients are accumulated in place."""
        from mitsuba3_experiments_b200 import capi
        a = self.make_args(variant=2, **kw)
        arr = (capi.GradTarget * len(targets))()
        for i, (b, k, g) in enumerate(targets):
            arr[i].bsdf, arr[i].kind, arr[i].grad = b, k, g.ctypes.data
        cnt = capi.Counters()
        gi = np.ascontiguousarray(grad_in, dtype=np.float32)
        if lib().orc_render_backward(self.h, C.byref(a), _ptr(gi), arr, len(targets), C.byref(cnt)) != 0:
            raise RuntimeError(lib().orc_last_error().decode())
        return cnt.as_dict()

    def render_forward(self, tangents, **kw):
        """tangents: list of (bsdf_index, kind, tangent ndarray float32) -> (gradient image [H, W, 3], counters)."""
        from mitsuba3_experiments_b200 import capi
        a = self.make_args(variant=2, **kw)
        arr = (capi.GradTarget * max(1, len(tangents)))(); keep = []
        for i, (b, k, g) in enumerate(tangents):
            g = np.ascontiguousarray(g, dtype=np.float32); keep.append(g)
            arr[i].bsdf, arr[i].kind, arr[i].grad = b, k, g.ctypes.data
        h, w = self.desc.film_shape
        film = np.zeros((h, w, 4), dtype=np.float32); cnt = capi.Counters()
        if lib().orc_render_forward(self.h, C.byref(a), arr, len(tangents), _ptr(film), C.byref(cnt)) != 0:
            raise RuntimeError(lib().orc_last_error().decode())
        return self.develop(film), cnt.as_dict()

    def debug_lane(self, lane, max_records=64, **kw):
        """-> (records [n][11] float32: ray(7), t, u, v, face-bits ; L rgb) for one lane (diagnostics)."""
        a = self.make_args(**kw)
        out = np.zeros((max_records, 11), np.float32); rgb = np.zeros(3, np.float32)
        n = lib().orc_debug_lane(self.h, C.byref(a), int(lane), _ptr(out), max_records, _ptr(rgb))
        return out[:n], rgb

    def trace(self, rays, any_hit=False):
        from mitsuba3_experiments_b200 import capi
        rays = np.ascontiguousarray(rays, dtype=np.float32).reshape(-1, 8)
        hits = np.zeros(len(rays), dtype=[("t", "f4"), ("u", "f4"), ("v", "f4"), ("prim", "u4")])
        lib().orc_trace(self.h, _ptr(rays), len(rays), int(any_hit), _ptr(hits))
        return hits

    def surface_interaction(self, o, d, maxt=3.4e38):
        ray = np.array(list(o) + list(d) + [maxt], dtype=np.float32)
        out = np.zeros(25, dtype=np.float32)
        lib().orc_surface_interaction(self.h, _ptr(ray), _ptr(out))
        keys = [("valid", 0, 1), ("t", 1, 2), ("p", 2, 5), ("n", 5, 8), ("uv", 8, 10), ("dp_du", 10, 13), ("dp_dv", 13, 16),
                ("sh_n", 16, 19), ("wi", 19, 22), ("prim_uv", 22, 24), ("face", 24, 25)]
        return {k: out[a:b] if b - a > 1 else float(out[a]) for k, a, b in keys}
