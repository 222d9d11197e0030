"""Device-side scene handle: thin Python wrapper over the C-ABI (``capi``) using PyTorch only for device
memory, streams and DLPack exchange.

Mirrors what the reference's integrators get from a live ``mi.Scene``: an acceleration structure that is built once
(``Scene`` constructor, mitsuba3/src/render/scene.cpp:22-111) and parameters that can be updated in place
(``SceneParameters.update``, mitsuba3/src/python/python/util.py:271-337).  All numerics live in ``csrc``.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import capi
from .scene import SceneDesc


def _torch():
    import torch
    return torch


def as_device_tensor(x, device, dtype=None):
    """Accept torch tensors, numpy arrays and anything exporting ``__dlpack__`` (e.g. Dr.Jit ``TensorXf``)."""
    torch = _torch()
    if isinstance(x, torch.Tensor):
        t = x
    elif hasattr(x, "__dlpack__"):
        t = torch.from_dlpack(x)
    else:
        t = torch.as_tensor(np.asarray(x))
    if dtype is not None and t.dtype != dtype:
        t = t.to(dtype)
    return t.to(device).contiguous()


class DeviceScene:
    """Owns a ``b2_scene`` on one CUDA device."""

    def __init__(self, scene, device: int = 0, base_path: str = ".", builder: int = -1):
        torch = _torch()
        self.lib = capi.load_library()
        if not torch.cuda.is_available():
            raise capi.B2Error("b2path needs a CUDA device (sm_100a); there is no CPU fallback")
        self.desc = scene if isinstance(scene, SceneDesc) else SceneDesc(scene, base_path)
        self.device = torch.device("cuda", device)
        self.ctx = C.c_void_p()
        capi.check(self.lib.b2_init(device, C.byref(self.ctx)))
        self.handle = C.c_void_p()
        capi.check(self.lib.b2_scene_create(self.ctx, C.byref(self.desc.c), C.byref(self.handle)))
        self.build_stats = capi.BuildStats()
        capi.check(self.lib.b2_scene_build_accel(self.handle, builder, self._stream(), C.byref(self.build_stats)))

    def close(self):
        if getattr(self, "handle", None):
            self.lib.b2_scene_destroy(self.handle); self.handle = None
        if getattr(self, "ctx", None):
            self.lib.b2_shutdown(self.ctx); self.ctx = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _stream(self) -> int:
        return int(_torch().cuda.current_stream(self.device).cuda_stream)

    @property
    def film_shape(self):
        return self.desc.film_shape

    @property
    def lane_grid(self):
        return self.desc.lane_grid

    def rebuild_accel(self, builder=-1):
        capi.check(self.lib.b2_scene_build_accel(self.handle, builder, self._stream(), C.byref(self.build_stats)))
        return self.build_stats.as_dict()

    def make_args(self, seed=0, spp=None, max_depth=8, rr_depth=4, variant=capi.VARIANT_PATH, spp_per_pass=0,
                  lane_begin=0, lane_end=0, chunk_lanes=0, env_warp=capi.ENV_WARP_HIERARCHICAL, reuse_weights=False,
                  kernel=capi.KERNEL_WAVEFRONT, all_shadow_rays=False, sample_begin=0, sample_count=0):
        a = capi.RenderArgs()
        a.seed = int(seed) & 0xffffffff
        a.spp = int(spp or self.desc.spp); a.spp_per_pass = int(spp_per_pass)
        a.max_depth, a.rr_depth, a.variant = int(max_depth), int(rr_depth), int(variant)
        a.kernel = int(kernel); a.env_warp = int(env_warp)
        a.lane_begin, a.lane_end, a.chunk_lanes = int(lane_begin), int(lane_end), int(chunk_lanes)
        a.reserved = (1 if reuse_weights else 0) | (2 if all_shadow_rays else 0)
        a.sample_begin, a.sample_count = int(sample_begin), int(sample_count)
        return a

    def wavefront_size(self, spp, spp_per_pass=0):
        """Lanes per pass (integrator.cpp:247-265)."""
        h, w = self.lane_grid
        spp_pp = min(spp_per_pass, spp) if spp_per_pass else spp
        n = h * w * spp_pp
        if n > 0xffffffff:
            spp_pp //= (n + 0xffffffff - 1) // 0xffffffff
            n = h * w * spp_pp
        return n

    def new_film(self):
        h, w = self.film_shape
        return _torch().zeros((h, w, 4), dtype=_torch().float32, device=self.device)

    def render_film(self, film=None, counters=False, **kw):
        """Accumulates RGBW into ``film`` ([H, W, 4] float32 on this device) and returns (film, counters | None)."""
        a = self.make_args(**kw)
        if film is None:
            film = self.new_film()
        cnt = capi.Counters() if counters else None
        capi.check(self.lib.b2_render(self.handle, C.byref(a), film.data_ptr(), self._stream(), C.byref(cnt) if counters else None))
        return film, (cnt.as_dict() if counters else None)

    def develop(self, film):
        h, w = self.film_shape
        img = _torch().empty((h, w, 3), dtype=_torch().float32, device=self.device)
        capi.check(self.lib.b2_develop(self.handle, film.data_ptr(), img.data_ptr(), self._stream()))
        return img

    def render(self, **kw):
        film, cnt = self.render_film(**kw)
        return self.develop(film), cnt

    def render_host(self, film_host: np.ndarray, image_host: np.ndarray | None = None, **kw):
        """Host-buffer entry point (``b2_render_host``): upload, render, develop and download inside the call."""
        a = self.make_args(**kw)
        capi.check(self.lib.b2_render_host(self.handle, C.byref(a), film_host.ctypes.data,
                                           image_host.ctypes.data if image_host is not None else None, None))
        return film_host, image_host

    def render_backward(self, grad_in, targets, counters=False, **kw):
        """targets: list of (bsdf_index, kind, grad tensor on this device); gradients are accumulated in place."""
        kw.pop("variant", None)
        a = self.make_args(variant=capi.VARIANT_PRB, **kw)
        gi = as_device_tensor(grad_in, self.device, _torch().float32)
        arr = (capi.GradTarget * max(1, len(targets)))()
        for i, (b, k, g) in enumerate(targets):
            assert g.is_cuda and g.dtype == _torch().float32 and g.is_contiguous()
            arr[i].bsdf, arr[i].kind, arr[i].grad = int(b), int(k), g.data_ptr()
        cnt = capi.Counters() if counters else None
        capi.check(self.lib.b2_render_backward(self.handle, C.byref(a), gi.data_ptr(), arr, len(targets), self._stream(),
                                               C.byref(cnt) if counters else None))
        return cnt.as_dict() if counters else None

    def render_forward(self, tangents, film=None, counters=False, **kw):
        """tangents: list of (bsdf_index, kind, tangent tensor on this device, read only).  Splats the lanes' d L into ``film``
        (RGBW, accumulated) and returns (film, counters | None); ``develop(film)`` is the gradient image."""
        kw.pop("variant", None)
        a = self.make_args(variant=capi.VARIANT_PRB, **kw)
        if film is None:
            film = self.new_film()
        arr = (capi.GradTarget * max(1, len(tangents)))()
        for i, (b, k, g) in enumerate(tangents):
            assert g.is_cuda and g.dtype == _torch().float32 and g.is_contiguous()
            arr[i].bsdf, arr[i].kind, arr[i].grad = int(b), int(k), g.data_ptr()
        cnt = capi.Counters() if counters else None
        capi.check(self.lib.b2_render_forward(self.handle, C.byref(a), arr, len(tangents), film.data_ptr(), self._stream(),
                                              C.byref(cnt) if counters else None))
        return film, (cnt.as_dict() if counters else None)

    def trace(self, rays, any_hit=False, kernel=capi.KERNEL_WAVEFRONT):
        """Unit-test hook: closest hits (or arrival tests) of caller-supplied rays, through the wavefront trace kernel (default)
        or the megakernel's per-thread traversal."""
        torch = _torch()
        r = as_device_tensor(rays, self.device, torch.float32).reshape(-1, 8)
        hits = torch.empty((r.shape[0], 4), dtype=torch.float32, device=self.device)
        flags = int(bool(any_hit)) | (2 if kernel == capi.KERNEL_MEGAKERNEL else 0)
        capi.check(self.lib.b2_trace(self.handle, r.data_ptr(), r.shape[0], flags, hits.data_ptr(), self._stream()))
        torch.cuda.synchronize(self.device)
        out = hits.cpu().numpy()
        return out.view(np.dtype([("t", "f4"), ("u", "f4"), ("v", "f4"), ("prim", "u4")])).reshape(-1)

    def trace_noalloc(self, rays, hits, any_hit=False):
        """``b2_trace_noalloc``: the wavefront trace kernel over at most ``capi.TRACE_NOALLOC_MAX_RAYS`` caller-supplied rays with
        the kernel's lane state in the library's static device array -- no device allocation inside the call.  ``rays`` [n, 8]
        and ``hits`` [n, 4] are float32 CUDA tensors the caller owns; the launch is asynchronous on the current stream.  Calls
        share the array: do not let two overlap on the GPU.  (No counterpart in the reference.)"""
        torch = _torch()
        assert rays.is_cuda and hits.is_cuda and rays.dtype == torch.float32 and hits.dtype == torch.float32
        assert rays.is_contiguous() and hits.is_contiguous() and rays.shape[-1] == 8 and hits.numel() == rays.numel() // 2
        capi.check(self.lib.b2_trace_noalloc(self.handle, rays.data_ptr(), rays.numel() // 8, int(bool(any_hit)), hits.data_ptr(), self._stream()))
        return hits

    def debug_sampler(self, lane_begin, n_lanes, n_draws, seed=0, spp=None):
        """Unit-test hook: (draws [n_lanes, n_draws], sample positions [n_lanes, 2]) exactly as the render kernels derive them."""
        torch = _torch()
        a = self.make_args(seed=seed, spp=spp)
        out = torch.empty((n_lanes, n_draws), dtype=torch.float32, device=self.device)
        pos = torch.empty((n_lanes, 2), dtype=torch.float32, device=self.device)
        capi.check(self.lib.b2_debug_sampler(self.handle, C.byref(a), int(lane_begin), int(n_lanes), int(n_draws), out.data_ptr(), pos.data_ptr(), self._stream()))
        torch.cuda.synchronize(self.device)
        return out.cpu().numpy(), pos.cpu().numpy()

    def debug_primary_hits(self, lane_begin, n_lanes, kernel=1, **kw):
        """Unit-test hook: closest hits of the primary rays of lanes ``lane_begin .. lane_begin + n_lanes`` (pass 0), through the
        packet kernel renders use for bounce 0 (``kernel=1`` / ``2``: packets of 32 / 64 lanes) or the per-ray trace kernel (``kernel=0``)."""
        torch = _torch()
        a = self.make_args(**kw)
        hits = torch.empty((n_lanes, 4), dtype=torch.float32, device=self.device)
        capi.check(self.lib.b2_debug_primary_hits(self.handle, C.byref(a), int(lane_begin), int(n_lanes), int(kernel), hits.data_ptr(), self._stream()))
        torch.cuda.synchronize(self.device)
        return hits.cpu().numpy().view(np.dtype([("t", "f4"), ("u", "f4"), ("v", "f4"), ("prim", "u4")])).reshape(-1)

    def set_kernel_timing(self, enabled: bool):
        """Kernel history (``dr.kernel_history`` analogue): bracket every wavefront launch with CUDA events."""
        capi.check(self.lib.b2_set_kernel_timing(self.handle, int(bool(enabled))))

    def kernel_times(self):
        """-> {kind: (milliseconds, launches)} accumulated since the previous call (synchronises the device)."""
        ms = (C.c_double * 4)(); n = (C.c_uint64 * 4)()
        capi.check(self.lib.b2_kernel_times(self.handle, ms, n))
        return {k: (float(ms[i]), int(n[i])) for i, k in enumerate(("trace", "shade", "splat", "other"))}

    def set_param(self, bsdf_index: int, what: str, value):
        """what = 'value' (3 floats) | 'data' ([H, W, 3] fp32 texture)."""
        torch = _torch()
        t = as_device_tensor(value, self.device, torch.float32)
        key = f"{int(bsdf_index)}.reflectance.{what}".encode()
        capi.check(self.lib.b2_scene_set_param(self.handle, key, t.data_ptr(), t.numel() * 4, self._stream()))
        b = self.desc.bsdfs[int(bsdf_index)]
        if b.type == capi.BSDF_ROUGHPLASTIC:
            # RoughPlastic::parameters_changed (roughplastic.cpp:216-238): the specular sampling weight follows the texture means
            # (bitmap: mean luminance, bitmap.cpp:829-853; rgb: mean of the channels, srgb.cpp:117-122)
            d_mean = (t.reshape(-1, 3) @ torch.tensor([0.212671, 0.715160, 0.072169], device=self.device)).mean() if what == "data" else t.mean()
            s_mean = float(np.mean(np.asarray(b.transmittance[:], np.float32)))
            w = (s_mean / (d_mean + s_mean)).to(torch.float32).reshape(1).contiguous()
            capi.check(self.lib.b2_scene_set_param(self.handle, f"{int(bsdf_index)}.spec_sampling_weight".encode(), w.data_ptr(), 4, self._stream()))
