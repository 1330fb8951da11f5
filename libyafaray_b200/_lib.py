"""ctypes binding of the C-ABI in include/y4gpu.h (liby4gpu.so, built in-tree for sm_100a).

There is no CPU fallback: if the shared library is missing or no CUDA device is usable,
loading / context creation raises.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "liby4gpu.so")

OK, ERR_ARG, ERR_CUDA, ERR_UNSUPPORTED = 0, -1, -2, -3
TRACE_CLOSEST, TRACE_SHADOW, TRACE_TSHADOW = 0, 1, 2
TRI_VIS_CLOSEST, TRI_VIS_SHADOW, TRI_VIS_TSHADOW, TRI_TRANSPARENT, TRI_MAT_SHIFT = 1, 2, 4, 8, 8

RAY_DTYPE = np.dtype([("o", "f4", 3), ("tmin", "f4"), ("d", "f4", 3), ("tmax", "f4")])
HIT_DTYPE = np.dtype([("t", "f4"), ("bu", "f4"), ("bv", "f4"), ("prim", "i4")])
TSHIT_DTYPE = np.dtype([("t", "f4"), ("bu", "f4"), ("bv", "f4"), ("prim", "i4"), ("filter", "f4", 3), ("pad", "f4")])


class BuildParams(C.Structure):
    _fields_ = [("leaf_max", C.c_int32), ("reserved", C.c_int32 * 7)]


class Stats(C.Structure):
    _fields_ = [
        ("n_tris", C.c_uint64), ("n_nodes", C.c_uint64),
        ("node_bytes", C.c_uint64), ("tri_bytes", C.c_uint64),
        ("build_ms", C.c_double), ("last_trace_ms", C.c_double),
        ("last_trace_rays", C.c_uint64),
        ("nodes_visited", C.c_uint64), ("tris_tested", C.c_uint64),
        ("kernel_launches", C.c_uint64), ("bvh_levels", C.c_uint64),
        ("last_trace_sorted", C.c_uint64), ("last_probe_distinct", C.c_double),
    ]  # fmt: skip


class Mesh(C.Structure):          # y4gpu_mesh
    _fields_ = [("v_xyz", C.c_void_p), ("flags", C.c_void_p), ("n_tri", C.c_uint64)]


class Instance(C.Structure):      # y4gpu_instance
    _fields_ = [("mesh", C.c_uint32), ("identity", C.c_uint32), ("obj_to_world", C.c_float * 12)]


class PtMaterial(C.Structure):
    _fields_ = [("type", C.c_int32), ("color", C.c_float * 3), ("diffuse_reflect", C.c_float), ("emit", C.c_float),
                ("oren_nayar", C.c_int32), ("sigma", C.c_float), ("double_sided", C.c_int32), ("receive_shadows", C.c_int32),
                ("flat_material", C.c_int32), ("layer_first", C.c_int32), ("layer_count", C.c_int32), ("specular_reflect", C.c_float),
                ("mirror_color", C.c_float * 3), ("fresnel_effect", C.c_int32), ("ior", C.c_float), ("transparency", C.c_float), ("transmit_filter", C.c_float),
                ("translucency", C.c_float), ("reserved", C.c_int32 * 3),
                ("absorb", C.c_int32), ("absorption", C.c_float * 3), ("absorption_dist", C.c_double)]  # fmt: skip


class PtTexture(C.Structure):
    _fields_ = [("texels", C.c_void_p), ("width", C.c_int32), ("height", C.c_int32), ("clip", C.c_int32), ("interpolate", C.c_int32),
                ("xrepeat", C.c_int32), ("yrepeat", C.c_int32), ("mirror_x", C.c_int32), ("mirror_y", C.c_int32), ("rot90", C.c_int32),
                ("cropmin_x", C.c_float), ("cropmin_y", C.c_float), ("cropmax_x", C.c_float), ("cropmax_y", C.c_float),
                ("even_tiles", C.c_int32), ("odd_tiles", C.c_int32), ("checker_dist", C.c_float),
                ("adj_intensity", C.c_float), ("adj_contrast", C.c_float), ("adj_mult_factor_red", C.c_float),
                ("adj_mult_factor_green", C.c_float), ("adj_mult_factor_blue", C.c_float), ("adj_clamp", C.c_int32),
                ("reserved", C.c_int32 * 4)]  # fmt: skip


class PtLayer(C.Structure):
    _fields_ = [("texture", C.c_int32), ("texco", C.c_int32), ("mapping", C.c_int32), ("proj_x", C.c_int32), ("proj_y", C.c_int32),
                ("proj_z", C.c_int32), ("scale", C.c_float * 3), ("offset", C.c_float * 3), ("flags", C.c_int32), ("blend_mode", C.c_int32),
                ("do_color", C.c_int32), ("colfac", C.c_float), ("def_col", C.c_float * 3), ("upper_color", C.c_float * 4),
                ("reserved", C.c_int32 * 4)]  # fmt: skip


class PtLight(C.Structure):
    _fields_ = [("type", C.c_int32), ("from_", C.c_float * 3), ("point1", C.c_float * 3), ("point2", C.c_float * 3),
                ("color", C.c_float * 3), ("power", C.c_float), ("samples", C.c_int32), ("cast_shadows", C.c_int32),
                ("radius", C.c_float), ("angle", C.c_float), ("infinite", C.c_int32), ("blend", C.c_float),
                ("double_sided", C.c_int32), ("object_n_tris", C.c_uint32), ("object_tris", C.c_void_p)]  # fmt: skip


class PtCamera(C.Structure):
    _fields_ = [("from_", C.c_float * 3), ("to", C.c_float * 3), ("up", C.c_float * 3), ("resx", C.c_int32), ("resy", C.c_int32),
                ("aspect_ratio", C.c_float), ("focal", C.c_float), ("aperture", C.c_float), ("dof_distance", C.c_float),
                ("near_clip", C.c_float), ("far_clip", C.c_float), ("bokeh_type", C.c_int32), ("bokeh_bias", C.c_int32),
                ("bokeh_rotation", C.c_float), ("type", C.c_int32),
                ("angle_degrees", C.c_double), ("max_angle_degrees", C.c_double), ("circular", C.c_int32), ("mirrored", C.c_int32),
                ("projection", C.c_int32)]  # fmt: skip


class PtParams(C.Structure):
    _fields_ = [("bounces", C.c_int32), ("path_samples", C.c_int32), ("russian_roulette_min_bounces", C.c_int32), ("no_recursive", C.c_int32),
                ("transp_background", C.c_int32), ("background", C.c_float * 3),
                ("width", C.c_int32), ("height", C.c_int32), ("xstart", C.c_int32), ("ystart", C.c_int32),
                ("aa_passes", C.c_int32), ("aa_samples", C.c_int32), ("aa_inc_samples", C.c_int32),
                ("aa_sample_multiplier_factor", C.c_float), ("aa_light_sample_multiplier_factor", C.c_float),
                ("aa_pixelwidth", C.c_float), ("aa_clamp_samples", C.c_float), ("filter_type", C.c_int32),
                ("shadow_bias_auto", C.c_int32), ("shadow_bias", C.c_float), ("min_raydist_auto", C.c_int32), ("min_raydist", C.c_float),
                ("base_sampling_offset", C.c_uint32), ("rr_seed", C.c_uint32), ("wave_size", C.c_uint32), ("profile", C.c_int32),
                ("transp_shadows", C.c_int32), ("shadow_depth", C.c_int32), ("raydepth", C.c_int32), ("trace_caustics", C.c_int32),
                ("transp_refracted_background", C.c_int32), ("background_gradient", C.c_int32), ("bg_gradient", (C.c_float * 3) * 4),
                ("ao_samples", C.c_int32), ("ao_distance", C.c_float), ("ao_color", C.c_float * 3),
                ("bg_texture", C.c_int32), ("bg_texture_angular", C.c_int32), ("bg_texture_rotation", C.c_float), ("bg_texture_power", C.c_float),
                ("bg_ibl_clamp", C.c_float), ("bg_sun_from", C.c_float * 3), ("bg_sun_turbidity", C.c_float), ("bg_sun_var", C.c_float * 5),
                ("bg_sun_power", C.c_float), ("bg_dark_altitude", C.c_float), ("bg_dark_bright", C.c_float), ("bg_dark_exposure", C.c_float),
                ("bg_dark_night", C.c_int32), ("bg_dark_color_space", C.c_int32)]  # fmt: skip


class PtStats(C.Structure):
    _fields_ = [("samples", C.c_uint64), ("rays_closest", C.c_uint64), ("rays_shadow", C.c_uint64), ("kernel_launches", C.c_uint64),
                ("pass_ms", C.c_double), ("class_ms", C.c_double * 8), ("class_launches", C.c_uint64 * 8),
                ("nodes_closest", C.c_uint64), ("tris_closest", C.c_uint64), ("nodes_shadow", C.c_uint64), ("tris_shadow", C.c_uint64),
                ("live_closest", C.c_uint64), ("live_shadow", C.c_uint64), ("camera_hits", C.c_uint64), ("path_entries", C.c_uint64),
                ("reserved", C.c_uint64 * 2)]  # fmt: skip


class PtAaNoise(C.Structure):
    _fields_ = [("threshold", C.c_float), ("detect_color_noise", C.c_int32), ("dark_detection_type", C.c_int32), ("dark_threshold_factor", C.c_float),
                ("variance_edge_size", C.c_int32), ("variance_pixels", C.c_int32), ("reserved", C.c_int32 * 6)]  # fmt: skip


MAT_SHINYDIFFUSE, MAT_LIGHT, MAT_MIRROR, MAT_GLASS = 0, 1, 2, 3
LIGHT_POINT, LIGHT_AREA, LIGHT_DIRECTIONAL, LIGHT_SUN, LIGHT_SPHERE, LIGHT_SPOT, LIGHT_OBJECT, LIGHT_BACKGROUND, LIGHT_SPOT_SOFT, LIGHT_PORTAL = 0, 1, 2, 3, 4, 5, 6, 7, 8, 9
FILTERS = {"box": 0, "gauss": 1, "mitchell": 2, "lanczos": 3}
CAMERA_TYPES = {"perspective": 0, "architect": 1, "orthographic": 2, "equirectangular": 3, "angular": 4}
DARKSKY_COLOR_SPACES = {"CIE (E)": 0, "CIE (D50)": 1, "sRGB (D65)": 2, "sRGB (D50)": 3}
ANGULAR_PROJECTIONS = {"equidistant": 0, "orthographic": 1, "stereographic": 2, "equisolid_angle": 3, "rectilinear": 4}
BOKEH_TYPES = {"disk1": 0, "disk2": 1, "triangle": 3, "square": 4, "pentagon": 5, "hexagon": 6, "ring": 7}
BOKEH_BIAS = {"uniform": 0, "center": 1, "edge": 2}
CLIP_MODES = {"extend": 0, "clip": 1, "clipcube": 2, "repeat": 3, "checker": 4}
INTERPOLATIONS = {"none": 0, "bilinear": 1}
TEXCO = {"global": 0, "uv": 1, "orco": 2}
PROJECTIONS = {"plain": 0, "cube": 1}
BLEND_MODES = {"mix": 0, "add": 1, "multiply": 2, "subtract": 3, "screen": 4, "divide": 5, "difference": 6, "darken": 7, "lighten": 8}
LAYER_RGB_TO_INT, LAYER_STENCIL, LAYER_NEGATIVE = 1, 2, 4


NOALLOC_MAX_RAYS = 65536  # Y4GPU_NOALLOC_MAX_RAYS


class Y4Error(RuntimeError):
    pass


_lib = None

# every symbol include/y4gpu.h declares (checked by tests/test_cabi.py)
EXPORTED = [
    "y4gpu_create", "y4gpu_destroy", "y4gpu_last_error", "y4gpu_device_count", "y4gpu_struct_size",
    "y4gpu_set_stream", "y4gpu_host_alloc", "y4gpu_host_free",
    "y4gpu_upload_triangles", "y4gpu_upload_instanced", "y4gpu_build", "y4gpu_get_bound",
    "y4gpu_trace", "y4gpu_trace_counted", "y4gpu_trace_tshadow", "y4gpu_trace_tshadow_inner", "y4gpu_trace_noalloc",
    "y4gpu_set_ray_sort", "y4gpu_get_stats", "y4gpu_synchronize",
    "y4gpu_pt_set_textures", "y4gpu_pt_set_scene", "y4gpu_pt_set_film_buffers", "y4gpu_pt_clear_film", "y4gpu_pt_render_pass",
    "y4gpu_pt_read_film", "y4gpu_pt_write_film", "y4gpu_pt_get_stats", "y4gpu_pt_next_pass", "y4gpu_libm_eval",
]  # fmt: skip


# index = y4gpu_struct_size's argument (rays / hits travel as plain float arrays: no mirror)
MIRRORED_STRUCTS = [BuildParams, Stats, Mesh, Instance, PtMaterial, PtTexture, PtLayer, PtLight, PtCamera, PtParams, PtStats, PtAaNoise, None, None, None]


def lib() -> C.CDLL:
    """Load liby4gpu.so (raises if it was not built: run `python -c 'import __graft_entry__ as g; g.build()'`)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise Y4Error(f"{LIB_PATH} is missing: the CUDA extension was not built and there is no CPU fallback")
    L = C.CDLL(LIB_PATH)
    vp, u64, i32, u32 = C.c_void_p, C.c_uint64, C.c_int, C.c_uint32
    L.y4gpu_create.argtypes = [i32, C.POINTER(vp)]
    L.y4gpu_create.restype = i32
    L.y4gpu_destroy.argtypes = [vp]
    L.y4gpu_destroy.restype = None
    L.y4gpu_last_error.argtypes = [vp]
    L.y4gpu_last_error.restype = C.c_char_p
    L.y4gpu_device_count.argtypes = []
    L.y4gpu_device_count.restype = i32
    L.y4gpu_struct_size.argtypes = [i32]
    L.y4gpu_struct_size.restype = u64
    # the ctypes mirrors above are written by hand: a library built from another header revision must not be driven with them
    for which, cls in enumerate(MIRRORED_STRUCTS):
        if cls is not None and C.sizeof(cls) != L.y4gpu_struct_size(which):
            raise Y4Error(f"{LIB_PATH}: sizeof({cls.__name__}) is {L.y4gpu_struct_size(which)} in the library but {C.sizeof(cls)} in libyafaray_b200/_lib.py "
                          "(rebuild: python -c 'import __graft_entry__ as g; g.build()')")
    L.y4gpu_set_stream.argtypes = [vp, vp]
    L.y4gpu_set_stream.restype = i32
    L.y4gpu_host_alloc.argtypes = [u64]
    L.y4gpu_host_alloc.restype = vp
    L.y4gpu_host_free.argtypes = [vp]
    L.y4gpu_host_free.restype = None
    L.y4gpu_upload_triangles.argtypes = [vp, vp, vp, u64]
    L.y4gpu_upload_triangles.restype = i32
    L.y4gpu_upload_instanced.argtypes = [vp, C.POINTER(Mesh), u32, C.POINTER(Instance), u32]
    L.y4gpu_upload_instanced.restype = i32
    L.y4gpu_build.argtypes = [vp, C.POINTER(BuildParams)]
    L.y4gpu_build.restype = i32
    L.y4gpu_get_bound.argtypes = [vp, C.POINTER(C.c_float * 6)]
    L.y4gpu_get_bound.restype = i32
    for name in ("y4gpu_trace", "y4gpu_trace_counted", "y4gpu_trace_noalloc"):
        f = getattr(L, name)
        f.argtypes = [vp, i32, vp, u64, vp]
        f.restype = i32
    L.y4gpu_trace_tshadow.argtypes = [vp, vp, u64, i32, vp, u32, vp]
    L.y4gpu_trace_tshadow.restype = i32
    L.y4gpu_trace_tshadow_inner.argtypes = [vp, vp, u64, i32, vp, u32, vp]
    L.y4gpu_trace_tshadow_inner.restype = i32
    L.y4gpu_set_ray_sort.argtypes = [vp, i32]
    L.y4gpu_set_ray_sort.restype = i32
    L.y4gpu_get_stats.argtypes = [vp, C.POINTER(Stats)]
    L.y4gpu_get_stats.restype = i32
    L.y4gpu_synchronize.argtypes = [vp]
    L.y4gpu_synchronize.restype = i32
    L.y4gpu_pt_set_scene.argtypes = [vp, C.POINTER(PtCamera), C.POINTER(PtMaterial), u32, C.POINTER(PtLight), u32, C.POINTER(PtParams), vp]
    L.y4gpu_pt_set_scene.restype = i32
    L.y4gpu_pt_set_textures.argtypes = [vp, C.POINTER(PtTexture), u32, C.POINTER(PtLayer), u32, vp, vp]
    L.y4gpu_pt_set_textures.restype = i32
    L.y4gpu_pt_set_film_buffers.argtypes = [vp, vp, vp]
    L.y4gpu_pt_set_film_buffers.restype = i32
    L.y4gpu_pt_clear_film.argtypes = [vp]
    L.y4gpu_pt_clear_film.restype = i32
    L.y4gpu_pt_render_pass.argtypes = [vp, i32, i32, i32, vp, u64, i32, i32]
    L.y4gpu_pt_render_pass.restype = i32
    L.y4gpu_pt_next_pass.argtypes = [vp, vp, vp, vp, C.POINTER(vp), C.POINTER(C.c_uint64)]
    L.y4gpu_pt_next_pass.restype = i32
    L.y4gpu_libm_eval.argtypes = [vp, i32, vp, vp, vp, u64]
    L.y4gpu_libm_eval.restype = i32
    L.y4gpu_pt_read_film.argtypes = [vp, vp, vp]
    L.y4gpu_pt_read_film.restype = i32
    L.y4gpu_pt_write_film.argtypes = [vp, vp, vp]
    L.y4gpu_pt_write_film.restype = i32
    L.y4gpu_pt_get_stats.argtypes = [vp, C.POINTER(PtStats)]
    L.y4gpu_pt_get_stats.restype = i32
    _lib = L
    return L


def check(ctx, rc: int, what: str) -> None:
    if rc != OK:
        msg = lib().y4gpu_last_error(ctx)
        raise Y4Error(f"{what} failed (rc={rc}): {msg.decode() if msg else ''}")


def pinned_empty(n: int, dtype) -> np.ndarray:
    """numpy array backed by page-locked host memory (freed when the array is collected)."""
    dtype = np.dtype(dtype)
    nbytes = max(1, n * dtype.itemsize)
    L = lib()
    p = L.y4gpu_host_alloc(nbytes)
    if not p:
        raise Y4Error("y4gpu_host_alloc failed")
    buf = (C.c_char * nbytes).from_address(p)
    arr = np.frombuffer(buf, dtype=dtype, count=n)

    import weakref

    weakref.finalize(buf, L.y4gpu_host_free, p)  # views keep `buf` alive through .base
    return arr

