"""ctypes binding of libsrb200.so (include/srb200.h).

The library is the product; this module only declares its C ABI.  There is no
Python/NumPy fallback: if the shared object is missing, or there is no CUDA
device when a context is created, the error is raised to the caller.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
# SRB200_LIB selects an alternative build of the same library (kernel A/B experiments); never a fallback.
LIB_PATH = os.environ.get("SRB200_LIB") or os.path.join(_HERE, "libsrb200.so")

SRB_OK = 0
SRB_ERR_ARG, SRB_ERR_CUDA, SRB_ERR_HANDLE, SRB_ERR_STATE, SRB_ERR_UNSUPPORTED = -1, -2, -3, -4, -5
SRB_CLEAR_COLOR, SRB_CLEAR_DEPTH = 1, 2
SRB_DRAW_FRUSTUM_CULL = 1
SRB_PIPELINE_AUTO, SRB_PIPELINE_ORDERED, SRB_PIPELINE_VISIBILITY = 0, 1, 2
SRB_RAYCAST_STATIC_MAX_RAYS = 4096          # include/srb200.h


class SrbUniforms(C.Structure):
    _fields_ = [
        ("light_dir", C.c_float * 4),
        ("light_color", C.c_float * 4),
        ("fog_color", C.c_float * 4),
        ("fog_start", C.c_float),
        ("fog_end", C.c_float),
        ("near_clip", C.c_float),
        ("numerics_fma", C.c_int32),
        ("tex_filter", C.c_int32),
        ("n_extra_lights", C.c_int32),
        ("debug_mode", C.c_int32),
        ("_pad", C.c_int32),
        ("extra_dir", (C.c_float * 4) * 4),
        ("extra_color", (C.c_float * 4) * 4),
        ("extra_atten", (C.c_float * 4) * 4),
    ]


class SrbStats(C.Structure):
    _fields_ = [
        ("draws_submitted", C.c_int64),
        ("draws_visible", C.c_int64),
        ("triangles_submitted", C.c_int64),
        ("triangles_setup", C.c_int64),
        ("tile_pairs", C.c_int64),
        ("clipped_triangles", C.c_int64),
        ("pipeline_used", C.c_int32),
        ("kernel_launches", C.c_int32),
        ("gpu_ms", C.c_float),
        ("triangles_no_coverage", C.c_int32),
        ("split_parts", C.c_int32),
        ("_pad", C.c_int32),
    ]


# numpy mirror of SrbDrawCmd (248 bytes) for batched submission.
DRAW_CMD_DTYPE = np.dtype(
    [
        ("mesh", "<i4"),
        ("texture", "<i4"),
        ("model", "<f4", (16,)),
        ("view", "<f4", (16,)),
        ("proj", "<f4", (16,)),
        ("sphere", "<f4", (4,)),
        ("cull", "<i4"),
        ("depth_test", "<i4"),
        ("blend", "<i4"),
        ("flags", "<u4"),
        ("submission_index", "<i4"),
        ("_pad", "<i4", (3,)),
    ],
    align=False,
)
assert DRAW_CMD_DTYPE.itemsize == 248, DRAW_CMD_DTYPE.itemsize

# Every symbol include/srb200.h declares (checked by tests/test_abi.py).
EXPORTS = [
    "srb_create", "srb_destroy", "srb_resize", "srb_get_size", "srb_last_error", "srb_set_band",
    "srb_set_pipeline", "srb_set_bin_split", "srb_set_pipeline_depth", "srb_set_graphs", "srb_export_color_handle", "srb_set_color_target", "srb_gather_rgba8_export", "srb_gather_rgba8_target",
    "srb_read_gathered_rgba8", "srb_device_gathered_rgba8_ptr", "srb_present_gathered_rgba8_async", "srb_get_tile_row_pairs",
    "srb_texture_create", "srb_texture_destroy", "srb_texture_sample",
    "srb_mesh_create", "srb_mesh_destroy", "srb_set_uniforms", "srb_begin_frame", "srb_draw",
    "srb_submit_draws", "srb_end_frame", "srb_replay_frame", "srb_sync", "srb_read_color",
    "srb_read_depth", "srb_read_color_rgba8", "srb_present_rgb", "srb_present_rgb_async", "srb_present_rgb_rows_async",
    "srb_present_rgba8", "srb_present_rgba8_async", "srb_present_rgba8_rows_async", "srb_present_wait", "srb_present_fence", "srb_barrier_export", "srb_barrier_connect", "srb_frame_barrier",
    "srb_read_color_rows", "srb_get_pixel", "srb_set_pixel",
    "srb_get_depth", "srb_set_depth", "srb_device_color_ptr", "srb_device_depth_ptr", "srb_stream",
    "srb_frustum_test", "srb_raycast", "srb_raycast_static", "srb_get_stats", "srb_version", "srb_set_profiling", "srb_get_stage_ms",
]

_lib = None


class SrbRayHit(C.Structure):
    _fields_ = [("hit", C.c_int32), ("distance", C.c_float), ("point", C.c_float * 3), ("normal", C.c_float * 3),
                ("triangle", C.c_int32), ("_pad", C.c_int32 * 3)]


RAY_HIT_DTYPE = np.dtype([("hit", "<i4"), ("distance", "<f4"), ("point", "<f4", (3,)), ("normal", "<f4", (3,)),
                          ("triangle", "<i4"), ("_pad", "<i4", (3,))])


class SrbError(RuntimeError):
    def __init__(self, code: int, message: str):
        super().__init__(f"srb200 error {code}: {message}")
        self.code = code


def lib() -> C.CDLL:
    """Load libsrb200.so; raise (never fall back) if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(softwarerenderer_b200 has no CPU fallback)"
        )
    L = C.CDLL(LIB_PATH)
    vp, i32, f32p = C.c_void_p, C.c_int, C.POINTER(C.c_float)
    L.srb_create.argtypes = [i32, i32, i32, C.POINTER(vp)]
    L.srb_destroy.argtypes = [vp]
    L.srb_resize.argtypes = [vp, i32, i32]
    L.srb_get_size.argtypes = [vp, C.POINTER(i32), C.POINTER(i32)]
    L.srb_last_error.argtypes = [vp]
    L.srb_last_error.restype = C.c_char_p
    L.srb_set_band.argtypes = [vp, i32, i32]
    L.srb_set_pipeline.argtypes = [vp, i32]
    L.srb_set_bin_split.argtypes = [vp, i32, i32, i32]
    L.srb_set_pipeline_depth.argtypes = [vp, i32]
    L.srb_set_graphs.argtypes = [vp, i32]
    L.srb_export_color_handle.argtypes = [vp, vp]
    L.srb_set_color_target.argtypes = [vp, vp]
    L.srb_get_tile_row_pairs.argtypes = [vp, vp, i32]
    L.srb_texture_create.argtypes = [vp, vp, i32, i32, C.POINTER(C.c_int32)]
    L.srb_texture_destroy.argtypes = [vp, C.c_int32]
    L.srb_texture_sample.argtypes = [vp, C.c_int32, vp, i32, i32, vp]
    L.srb_mesh_create.argtypes = [vp, vp, i32, vp, i32, C.POINTER(C.c_int32)]
    L.srb_mesh_destroy.argtypes = [vp, C.c_int32]
    L.srb_set_uniforms.argtypes = [vp, C.POINTER(SrbUniforms)]
    L.srb_begin_frame.argtypes = [vp, f32p, C.c_uint32]
    L.srb_draw.argtypes = [vp, C.c_int32, vp, vp, vp, C.c_int32, i32, i32, i32, i32]
    L.srb_submit_draws.argtypes = [vp, vp, i32]
    L.srb_end_frame.argtypes = [vp]
    L.srb_replay_frame.argtypes = [vp]
    L.srb_sync.argtypes = [vp]
    L.srb_read_color.argtypes = [vp, vp]
    L.srb_read_depth.argtypes = [vp, vp]
    L.srb_read_color_rgba8.argtypes = [vp, vp]
    L.srb_present_rgb.argtypes = [vp, vp]
    L.srb_present_rgb_async.argtypes = [vp, vp]
    L.srb_present_rgb_rows_async.argtypes = [vp, vp, C.c_int, C.c_int]
    L.srb_present_rgba8.argtypes = [vp, vp]
    L.srb_present_rgba8_async.argtypes = [vp, vp]
    L.srb_present_rgba8_rows_async.argtypes = [vp, vp, C.c_int, C.c_int]
    L.srb_present_wait.argtypes = [vp]
    L.srb_raycast.argtypes = [vp, C.c_int32, vp, vp, C.c_int, C.c_int, vp]
    L.srb_raycast_static.argtypes = [vp, C.c_int32, vp, vp, C.c_int, C.c_int, vp]
    L.srb_present_fence.argtypes = [vp]
    L.srb_barrier_export.argtypes = [vp, vp]
    L.srb_barrier_connect.argtypes = [vp, i32, i32, vp]
    L.srb_frame_barrier.argtypes = [vp, i32]
    L.srb_read_color_rows.argtypes = [vp, vp, i32, i32]
    L.srb_get_pixel.argtypes = [vp, i32, i32, f32p]
    L.srb_set_pixel.argtypes = [vp, i32, i32, f32p]
    L.srb_get_depth.argtypes = [vp, i32, i32, f32p]
    L.srb_set_depth.argtypes = [vp, i32, i32, C.c_float]
    L.srb_gather_rgba8_export.argtypes = [vp, vp]
    L.srb_gather_rgba8_target.argtypes = [vp, vp]
    L.srb_read_gathered_rgba8.argtypes = [vp, vp]
    L.srb_present_gathered_rgba8_async.argtypes = [vp, vp]
    for name in ("srb_device_color_ptr", "srb_device_depth_ptr", "srb_stream", "srb_device_gathered_rgba8_ptr"):
        getattr(L, name).argtypes = [vp]
        getattr(L, name).restype = vp
    L.srb_frustum_test.argtypes = [vp, vp, vp, vp, vp, i32, vp]
    L.srb_get_stats.argtypes = [vp, C.POINTER(SrbStats)]
    L.srb_version.argtypes = []
    L.srb_set_profiling.argtypes = [vp, i32]
    L.srb_get_stage_ms.argtypes = [vp, f32p, i32]
    _lib = L
    return L


def check(ctx, rc: int) -> None:
    if rc != SRB_OK:
        msg = lib().srb_last_error(ctx)
        raise SrbError(rc, msg.decode("utf-8", "replace") if msg else "")


def ptr(a: np.ndarray) -> C.c_void_p:
    return C.c_void_p(a.ctypes.data)
