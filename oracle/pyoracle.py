"""ctypes wrapper of liboracle.so -- TEST INFRASTRUCTURE (see oracle.h).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
reference legs import this module.  PARITY UNPINNED by upstream (no reference
tests or goldens exist); pinned by tests/test_oracle_kat.py.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "liboracle.so")


class OrcUniforms(C.Structure):
    _fields_ = [
        ("light_dir", C.c_float * 4), ("light_color", C.c_float * 4), ("fog_color", C.c_float * 4),
        ("fog_start", C.c_float), ("fog_end", C.c_float), ("near_clip", C.c_float),
        ("numerics_fma", C.c_int32), ("tex_filter", C.c_int32), ("n_extra_lights", C.c_int32),
        ("debug_mode", C.c_int32), ("_pad", C.c_int32), ("extra_dir", (C.c_float * 4) * 4), ("extra_color", (C.c_float * 4) * 4),
        ("extra_atten", (C.c_float * 4) * 4),
    ]


class OrcDraw(C.Structure):
    _fields_ = [
        ("vertices", C.c_void_p), ("indices", C.c_void_p), ("tex_rgba8", C.c_void_p),
        ("n_vertices", C.c_int32), ("n_indices", C.c_int32), ("tex_w", C.c_int32), ("tex_h", C.c_int32),
        ("model", C.c_float * 16), ("view", C.c_float * 16), ("proj", C.c_float * 16),
        ("cull", C.c_int32), ("depth_test", C.c_int32), ("blend", C.c_int32), ("_pad", C.c_int32),
    ]


class OrcStats(C.Structure):
    _fields_ = [("triangles_in", C.c_int64), ("triangles_drawn", C.c_int64),
                ("fragments_tested", C.c_int64), ("fragments_shaded", C.c_int64)]


_lib = None


def build(force: bool = False) -> str:
    if force or not os.path.exists(LIB_PATH) or os.path.getmtime(LIB_PATH) < os.path.getmtime(os.path.join(_HERE, "oracle.cpp")):
        subprocess.check_call(["make", "-C", _HERE, "-s"])
    return LIB_PATH


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            build()
        L = C.CDLL(LIB_PATH)
        vp = C.c_void_p
        L.orc_create.argtypes = [C.c_int, C.c_int]
        L.orc_create.restype = vp
        L.orc_destroy.argtypes = [vp]
        L.orc_color.argtypes = [vp]
        L.orc_color.restype = C.POINTER(C.c_float)
        L.orc_depth.argtypes = [vp]
        L.orc_depth.restype = C.POINTER(C.c_float)
        L.orc_clear_color.argtypes = [vp, C.POINTER(C.c_float)]
        L.orc_clear_depth.argtypes = [vp]
        L.orc_clear_frame_mt.argtypes = [vp, C.POINTER(C.c_float), C.c_int]
        L.orc_render_mesh.argtypes = [vp, C.POINTER(OrcDraw), C.POINTER(OrcUniforms)]
        L.orc_render_frame_mt.argtypes = [vp, C.POINTER(OrcDraw), C.c_int, C.POINTER(OrcUniforms), C.c_int, C.c_int, C.c_int]
        L.orc_sphere_in_frustum.argtypes = [vp, vp, vp, vp, C.c_int]
        L.orc_texture_sample.argtypes = [vp, C.c_int, C.c_int, C.c_float, C.c_float, C.c_int, vp]
        L.orc_vertex_shader.argtypes = [vp, vp, vp, vp, C.c_int, vp, vp]
        L.orc_clip_near.argtypes = [vp, C.c_float, C.c_int, vp]
        L.orc_get_stats.argtypes = [vp, C.POINTER(OrcStats)]
        _lib = L
    return _lib


def _f32(a, n):
    return np.ascontiguousarray(np.asarray(a, dtype=np.float32).reshape(n))


def make_uniforms(scene) -> OrcUniforms:
    u = OrcUniforms()
    for i in range(3):
        u.light_dir[i] = float(scene.light_dir[i])
    for i in range(4):
        u.light_color[i] = float(scene.light_color[i])
        u.fog_color[i] = float(scene.fog_color[i])
    u.fog_start, u.fog_end, u.near_clip = scene.fog_start, scene.fog_end, scene.near_clip
    u.numerics_fma, u.tex_filter = scene.numerics_fma, scene.tex_filter
    u.debug_mode = int(getattr(scene, "debug_mode", 0))
    from softwarerenderer_b200.shaders import light_rows       # plain data marshalling shared with the product's binding
    rows = light_rows(scene.extra_lights)
    u.n_extra_lights = len(rows)
    for k, (d, col, att) in enumerate(rows):
        for i in range(4):
            u.extra_dir[k][i] = float(d[i])
        for i in range(3):
            u.extra_color[k][i] = float(col[i])
            u.extra_atten[k][i] = float(att[i])
    return u


def make_draws(scene, visible=None):
    """Returns (OrcDraw array, keepalive list).  `visible` optionally filters draws (frustum result)."""
    items = [d for i, d in enumerate(scene.draws) if visible is None or visible[i]]
    arr = (OrcDraw * max(1, len(items)))()
    keep = []
    for k, d in enumerate(items):
        v = np.ascontiguousarray(d.mesh.Vertices)
        ix = np.ascontiguousarray(d.mesh.Indices, dtype=np.uint16)
        keep += [v, ix]
        o = arr[k]
        o.vertices, o.indices = v.ctypes.data, ix.ctypes.data
        o.n_vertices, o.n_indices = v.shape[0], ix.size
        if d.mesh.Texture is not None:
            px = d.mesh.Texture.pixels
            keep.append(px)
            o.tex_rgba8, o.tex_w, o.tex_h = px.ctypes.data, px.shape[1], px.shape[0]
        else:
            o.tex_rgba8 = None
        for name, m in (("model", d.model), ("view", d.view), ("proj", d.proj)):
            f = _f32(m, 16)
            for i in range(16):
                getattr(o, name)[i] = float(f[i])
        o.cull, o.depth_test, o.blend = int(d.cull), int(d.depth), int(d.blend)
    return arr, len(items), keep


def sphere_in_frustum(bounds, model, view, proj, fma=1) -> bool:
    s = _f32((*bounds.Center, bounds.Radius), 4)
    m, v, p = _f32(model, 16), _f32(view, 16), _f32(proj, 16)
    return bool(lib().orc_sphere_in_frustum(s.ctypes.data, m.ctypes.data, v.ctypes.data, p.ctypes.data, fma))


def frustum_flags(scene):
    """The caller-side cull of Renderer.RenderDust2 etc. (Renderer.cs:446) for draws that ask for it."""
    return [(not d.frustum_cull) or sphere_in_frustum(d.mesh.SphereBounds, d.model, d.view, d.proj, scene.numerics_fma)
            for d in scene.draws]


def render_scene(scene, threads: int = 0, tile_rows=None, clear=(True, True), window=None, return_window=False):
    """Render `scene` with the oracle.  threads=0: literal single-thread path
    (orc_render_mesh per draw); threads>=1: tile-row-parallel variant.
    Returns (color[H,W,4], depth[H,W], stats)."""
    L = lib()
    w = window if window is not None else L.orc_create(scene.width, scene.height)
    if not w:
        raise ValueError("Width and height must be positive non-zero values.")
    if clear[1]:
        L.orc_clear_depth(w)
    if clear[0]:
        L.orc_clear_color(w, (C.c_float * 4)(*[float(x) for x in scene.clear_color]))
    u = make_uniforms(scene)
    arr, n, keep = make_draws(scene, frustum_flags(scene))
    if threads <= 0:
        for k in range(n):
            L.orc_render_mesh(w, C.byref(arr[k]), C.byref(u))
    else:
        r0, r1 = tile_rows if tile_rows is not None else (0, 1 << 30)
        L.orc_render_frame_mt(w, arr, n, C.byref(u), threads, r0, r1)
    H, W = scene.height, scene.width
    color = np.ctypeslib.as_array(L.orc_color(w), shape=(H, W, 4)).copy()
    depth = np.ctypeslib.as_array(L.orc_depth(w), shape=(H, W)).copy()
    st = OrcStats()
    L.orc_get_stats(w, C.byref(st))
    del keep
    if return_window:
        return color, depth, st, w
    if window is None:
        L.orc_destroy(w)
    return color, depth, st


def raycast(origin, direction, vertices, indices, model, face_mask=1, fma=1):
    """Physics.Raycast on the CPU oracle -> (hit, distance, point[3], normal[3], triangle)."""
    L = lib()
    o, d, m = _f32(origin, 3), _f32(direction, 3), _f32(model, 16)
    v = np.ascontiguousarray(vertices).view(np.float32).reshape(-1, 12)
    ix = np.ascontiguousarray(indices, dtype=np.uint16)
    dist, tri = C.c_float(), C.c_int()
    pt, nr = np.zeros(3, np.float32), np.zeros(3, np.float32)
    L.orc_raycast.restype = C.c_int
    hit = L.orc_raycast(C.c_void_p(o.ctypes.data), C.c_void_p(d.ctypes.data), C.c_void_p(v.ctypes.data), v.shape[0],
                        C.c_void_p(ix.ctypes.data), ix.size, C.c_void_p(m.ctypes.data), int(face_mask), int(fma),
                        C.byref(dist), C.c_void_p(pt.ctypes.data), C.c_void_p(nr.ctypes.data), C.byref(tri))
    return bool(hit), dist.value, pt, nr, tri.value


def texture_sample(pixels: np.ndarray, u: float, v: float, filter: int = 0) -> np.ndarray:
    px = np.ascontiguousarray(pixels, dtype=np.uint8)
    out = np.zeros(4, dtype=np.float32)
    lib().orc_texture_sample(px.ctypes.data, px.shape[1], px.shape[0], float(u), float(v), filter, out.ctypes.data)
    return out


def vertex_shader(vtx12, model, view, proj, fma=1):
    v = _f32(vtx12, 12)
    m, vv, p = _f32(model, 16), _f32(view, 16), _f32(proj, 16)
    clip, wn = np.zeros(4, np.float32), np.zeros(3, np.float32)
    lib().orc_vertex_shader(v.ctypes.data, m.ctypes.data, vv.ctypes.data, p.ctypes.data, fma, clip.ctypes.data, wn.ctypes.data)
    return clip, wn


def clip_near(clip3x4, near=0.1, fma=1):
    c = _f32(clip3x4, 12)
    out = np.zeros(24, np.float32)
    n = lib().orc_clip_near(c.ctypes.data, float(near), fma, out.ctypes.data)
    return out.reshape(2, 3, 4)[:n]
