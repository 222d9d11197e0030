"""ctypes mirror of ``include/b2path.h`` and loader of the CUDA product library ``libb2path.so``.

This is the *only* module that talks to the C-ABI.  It is the Python-side stub a maintainer
of the reference would add; ``integrator.py`` sits above it.  There is
deliberately NO CPU fallback: if the CUDA extension is missing the loader raises.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
import sys

B2_ABI_VERSION = 2
B2_MAX_WARP_LEVELS = 24

# enums (keep in sync with b2path.h)
BSDF_DIFFUSE, BSDF_ROUGHCONDUCTOR, BSDF_DIELECTRIC, BSDF_CONDUCTOR, BSDF_ROUGHDIELECTRIC, BSDF_ROUGHPLASTIC = 0, 1, 2, 3, 4, 5
BSDF_FLAG_TWOSIDED = 1
BSDF_FLAG_MASK = 2
MF_BECKMANN, MF_GGX = 0, 1
TEX_F32, TEX_F16 = 0, 1
WRAP_REPEAT, WRAP_MIRROR, WRAP_CLAMP = 0, 1, 2
FILTER_BILINEAR, FILTER_NEAREST = 0, 1
EMITTER_AREA, EMITTER_ENVMAP, EMITTER_CONSTANT, EMITTER_POINT = 0, 1, 2, 3
SHAPE_RECTANGLE, SHAPE_MESH = 0, 1
RFILTER_BOX, RFILTER_TENT, RFILTER_GAUSSIAN = 0, 1, 2
MESH_HAS_NORMALS, MESH_HAS_UV, MESH_FLIP_NORMALS = 1, 2, 4
VARIANT_PATH, VARIANT_PATH_MIS, VARIANT_PRB = 0, 1, 2
KERNEL_WAVEFRONT, KERNEL_MEGAKERNEL = 0, 1
TRACE_NOALLOC_MAX_RAYS = 32768        # B2_TRACE_NOALLOC_MAX_RAYS (include/b2path.h)
ENV_WARP_HIERARCHICAL, ENV_WARP_ALIAS = 0, 1

f32, i32, u32, u64 = C.c_float, C.c_int32, C.c_uint32, C.c_uint64


class Texture(C.Structure):
    _fields_ = [("data", C.c_void_p), ("width", i32), ("height", i32), ("format", i32),
                ("wrap", i32), ("filter", i32), ("to_uv", f32 * 9)]


class Bsdf(C.Structure):
    _fields_ = [("type", i32), ("flags", i32), ("reflectance_tex", i32), ("distribution", i32),
                ("sample_visible", i32), ("reflectance", f32 * 3), ("transmittance", f32 * 3),
                ("eta", f32 * 3), ("k", f32 * 3), ("alpha_u", f32), ("alpha_v", f32), ("ior", f32), ("opacity", f32), ("nonlinear", i32), ("spec_sampling_weight", f32)]


class Mesh(C.Structure):
    _fields_ = [("first_vertex", u32), ("n_vertices", u32), ("first_face", u32), ("n_faces", u32),
                ("bsdf", i32), ("emitter", i32), ("flags", u32)]


class Emitter(C.Structure):
    _fields_ = [("type", i32), ("mesh", i32), ("shape_kind", i32), ("radiance", f32 * 3),
                ("to_world", f32 * 16), ("normal", f32 * 3), ("inv_area", f32), ("cdf_offset", u32)]


class Envmap(C.Structure):
    _fields_ = [("data", C.c_void_p), ("width", i32), ("height", i32), ("scale", f32),
                ("to_world", f32 * 16), ("from_world", f32 * 16), ("bsphere_center", f32 * 3),
                ("bsphere_radius", f32), ("n_levels", i32), ("warp_data", C.c_void_p),
                ("level_offset", u32 * B2_MAX_WARP_LEVELS), ("level_width", u32 * B2_MAX_WARP_LEVELS),
                ("level_size", u32 * B2_MAX_WARP_LEVELS), ("alias_prob", C.c_void_p), ("alias_idx", C.c_void_p)]


class Sensor(C.Structure):
    _fields_ = [("sample_to_camera", f32 * 16), ("to_world", f32 * 16), ("near_clip", f32), ("far_clip", f32),
                ("film_width", i32), ("film_height", i32), ("crop_offset_x", i32), ("crop_offset_y", i32),
                ("crop_width", i32), ("crop_height", i32), ("rfilter", i32), ("rfilter_radius", f32),
                ("gaussian_coeff", f32 * 10), ("sample_border", i32), ("reserved", i32)]


class SceneDesc(C.Structure):
    _fields_ = [("abi_version", u32), ("n_vertices", u32), ("n_faces", u32), ("n_meshes", u32), ("n_bsdfs", u32),
                ("n_textures", u32), ("n_emitters", u32),
                ("positions", C.c_void_p), ("normals", C.c_void_p), ("texcoords", C.c_void_p), ("faces", C.c_void_p),
                ("meshes", C.POINTER(Mesh)), ("bsdfs", C.POINTER(Bsdf)), ("textures", C.POINTER(Texture)),
                ("emitters", C.POINTER(Emitter)), ("area_cdf", C.c_void_p), ("n_area_cdf", u32),
                ("envmap", C.POINTER(Envmap)), ("sensor", Sensor)]


class RenderArgs(C.Structure):
    _fields_ = [("seed", u32), ("spp", u32), ("spp_per_pass", u32), ("max_depth", i32), ("rr_depth", i32),
                ("variant", i32), ("kernel", i32), ("env_warp", i32), ("lane_begin", u64), ("lane_end", u64),
                ("chunk_lanes", u32), ("reserved", u32), ("sample_begin", u32), ("sample_count", u32)]


class Counters(C.Structure):
    _fields_ = [(n, u64) for n in ("samples", "closest_rays", "shadow_rays", "vertices", "tex_evals", "nee_area",
                                   "nee_env", "env_escapes", "splat_taps", "grad_atomics", "node_fetches",
                                   "tri_fetches", "launches")]

    def as_dict(self):
        return {n: int(getattr(self, n)) for n, _ in self._fields_}


class BuildStats(C.Structure):
    _fields_ = [("ms_morton_sort", f32), ("ms_lbvh", f32), ("ms_collapse", f32), ("ms_total", f32),
                ("n_nodes8", u32), ("n_tris", u32), ("max_depth", u32), ("builder", u32), ("node_bytes", u64), ("tri_bytes", u64),
                ("sah_cost", f32), ("reserved", f32)]

    def as_dict(self):
        return {n: getattr(self, n) for n, _ in self._fields_}


class Hit(C.Structure):
    _fields_ = [("t", f32), ("u", f32), ("v", f32), ("prim", u32)]


class GradTarget(C.Structure):
    _fields_ = [("bsdf", i32), ("kind", i32), ("grad", C.c_void_p)]


# every symbol include/b2path.h declares for the product library
EXPORTS = ["b2_init", "b2_shutdown", "b2_scene_create", "b2_scene_destroy", "b2_scene_build_accel",
           "b2_scene_set_param", "b2_render", "b2_develop", "b2_render_backward", "b2_render_forward", "b2_trace", "b2_trace_noalloc", "b2_debug_sampler", "b2_debug_primary_hits",
           "b2_render_host", "b2_set_kernel_timing", "b2_kernel_times", "b2_last_error", "b2_abi_version"]

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("B2PATH_LIB") or os.path.join(_HERE, "csrc", "libb2path.so")   # B2PATH_LIB: tuning builds only
_lib = None


class B2Error(RuntimeError):
    """Raised for any non-zero status of the C-ABI (mirrors Mitsuba's RuntimeError on Throw)."""


def build_library(verbose: bool = False) -> str:
    """Compile csrc/*.cu for sm_100a into csrc/libb2path.so (nvcc cross-compiles without a GPU)."""
    cmd = ["make", "-C", os.path.join(_HERE, "csrc")]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if verbose or res.returncode != 0:
        sys.stderr.write(res.stdout + res.stderr)
    if res.returncode != 0:
        raise B2Error("building libb2path.so failed")
    return LIB_PATH


def load_library() -> C.CDLL:
    """Load libb2path.so (no CUDA calls are made by loading).  Fails loudly if it is missing."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise B2Error(f"CUDA extension {LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                      "(there is no CPU fallback)")
    lib = C.CDLL(LIB_PATH)
    vp, up = C.c_void_p, C.c_size_t
    lib.b2_init.argtypes = [C.c_int, C.POINTER(vp)]
    lib.b2_shutdown.argtypes = [vp]
    lib.b2_shutdown.restype = None
    lib.b2_scene_create.argtypes = [vp, C.POINTER(SceneDesc), C.POINTER(vp)]
    lib.b2_scene_destroy.argtypes = [vp]
    lib.b2_scene_destroy.restype = None
    lib.b2_scene_build_accel.argtypes = [vp, C.c_int, up, C.POINTER(BuildStats)]
    lib.b2_scene_set_param.argtypes = [vp, C.c_char_p, vp, C.c_size_t, up]
    lib.b2_render.argtypes = [vp, C.POINTER(RenderArgs), vp, up, C.POINTER(Counters)]
    lib.b2_develop.argtypes = [vp, vp, vp, up]
    lib.b2_render_backward.argtypes = [vp, C.POINTER(RenderArgs), vp, C.POINTER(GradTarget), C.c_int, up,
                                       C.POINTER(Counters)]
    lib.b2_render_forward.argtypes = [vp, C.POINTER(RenderArgs), C.POINTER(GradTarget), C.c_int, vp, up, C.POINTER(Counters)]
    lib.b2_trace.argtypes = [vp, vp, u64, C.c_int, vp, up]
    lib.b2_trace_noalloc.argtypes = [vp, vp, u64, C.c_int, vp, up]
    lib.b2_debug_sampler.argtypes = [vp, C.POINTER(RenderArgs), u64, C.c_uint32, C.c_uint32, vp, vp, up]
    lib.b2_debug_primary_hits.argtypes = [vp, C.POINTER(RenderArgs), u64, C.c_uint32, C.c_int, vp, up]
    lib.b2_render_host.argtypes = [vp, C.POINTER(RenderArgs), vp, vp, C.POINTER(Counters)]
    lib.b2_set_kernel_timing.argtypes = [vp, C.c_int]
    lib.b2_kernel_times.argtypes = [vp, C.POINTER(C.c_double), C.POINTER(u64)]
    lib.b2_last_error.restype = C.c_char_p
    lib.b2_abi_version.restype = C.c_int
    for name in ("b2_init", "b2_scene_create", "b2_scene_build_accel", "b2_scene_set_param", "b2_render",
                 "b2_develop", "b2_render_backward", "b2_render_forward", "b2_trace", "b2_trace_noalloc", "b2_debug_sampler", "b2_debug_primary_hits", "b2_render_host", "b2_set_kernel_timing", "b2_kernel_times"):
        getattr(lib, name).restype = C.c_int
    if lib.b2_abi_version() != B2_ABI_VERSION:
        raise B2Error("libb2path.so ABI version mismatch")
    _lib = lib
    return lib


def check(status: int):
    if status != 0:
        msg = load_library().b2_last_error()
        raise B2Error((msg or b"unknown error").decode() + f" (status {status})")
