"""MainWindow -- headless owner of the device framebuffer (MainWindow.cs:17-437).

The reference's window/GL present is replaced by a device buffer; the public
buffer surface (RenderWidth/Height, ColorBuffer, DepthBuffer, clears, pixel
accessors, RenderFrame) is kept.  Draws recorded by Rasterizer.RenderMesh are
deferred: RenderFrame() submits them and launches the CUDA pipeline.
"""
from __future__ import annotations

import ctypes as C
import hashlib
import weakref
from collections import OrderedDict
from typing import Optional

import numpy as np

from . import _native as N
from .numerics import euler_to_direction


class Uniforms:
    """Captured state of the stock shaders (Renderer.cs:39-44) + Rasterizer.NearClip."""

    def __init__(self):
        self.LightDirection = euler_to_direction((-45.0, -45.0, 0.0))
        self.LightColor = np.ones(4, dtype=np.float32)
        self.FogColor = np.array([1.0, 0.62, 0.5, 1.0], dtype=np.float32)
        self.FogStart = 1.0
        self.FogEnd = 25.0
        self.NumericsFma = 1
        self.TexFilter = 0            # 0 nearest (reference), 1 bilinear (extension)
        self.ExtraLights = []         # extension: Light objects (directional or point) or (direction3, color3) pairs

    def to_struct(self, near_clip: float, debug_mode: int = 0) -> N.SrbUniforms:
        u = N.SrbUniforms()
        for i in range(3):
            u.light_dir[i] = float(self.LightDirection[i])
        for i in range(4):
            u.light_color[i] = float(self.LightColor[i])
            u.fog_color[i] = float(self.FogColor[i])
        u.fog_start, u.fog_end, u.near_clip = float(self.FogStart), float(self.FogEnd), float(near_clip)
        u.numerics_fma, u.tex_filter = int(self.NumericsFma), int(self.TexFilter)
        u.debug_mode = int(debug_mode)
        from .shaders import light_rows
        rows = light_rows(self.ExtraLights)
        u.n_extra_lights = len(rows)
        for k, (d, col, att) in enumerate(rows):
            for i in range(4):
                u.extra_dir[k][i] = float(d[i])
            for i in range(3):
                u.extra_color[k][i] = float(col[i])
                u.extra_atten[k][i] = float(att[i])
        return u


class MeshHandle:
    """A mesh resident on one window's device (MainWindow.CreateMesh)."""
    __slots__ = ("window_uid", "handle")

    def __init__(self, window_uid: int, handle: int):
        self.window_uid, self.handle = window_uid, handle


class MainWindow:
    _next_uid = 0

    def __init__(self, width: int, height: int, device: int = 0, render_scale: float = 1.0):
        if width <= 0 or height <= 0:
            raise ValueError("Width and height must be positive non-zero values.")   # Rasterizer.cs:71-74
        self.RenderScale = render_scale
        w, h = max(1, int(width * render_scale)), max(1, int(height * render_scale))
        ctx = C.c_void_p()
        rc = N.lib().srb_create(device, w, h, C.byref(ctx))
        if rc != N.SRB_OK:
            N.check(None, rc)
        self._ctx = ctx
        MainWindow._next_uid += 1
        self._uid = MainWindow._next_uid     # key for per-context resource handles
        self.RenderWidth, self.RenderHeight = w, h
        self.Uniforms = Uniforms()
        self._clear_flags = 0
        self._clear_color = np.zeros(4, dtype=np.float32)
        self._draws = []               # recorded SrbDrawCmd rows
        self._mesh_by_content = OrderedDict()    # blake2b of the arrays -> device mesh handle, LRU order
        self._mesh_by_identity = {}              # (id(vertices), id(indices)) -> (weakrefs, layout, content key)
        self.MeshCacheCapacity = 8192            # resident meshes kept before the least recently used is destroyed

    # -- lifetime -----------------------------------------------------------
    def Close(self) -> None:
        if self._ctx:
            N.lib().srb_destroy(self._ctx)
            self._ctx = None

    def __del__(self):
        try:
            self.Close()
        except Exception:
            pass

    def Resize(self, width: int, height: int) -> None:
        """MainWindow.HandleResize (MainWindow.cs:298-361)."""
        w, h = max(1, int(width * self.RenderScale)), max(1, int(height * self.RenderScale))
        N.check(self._ctx, N.lib().srb_resize(self._ctx, w, h))
        self.RenderWidth, self.RenderHeight = w, h

    def SetBand(self, tile_row0: int, tile_row1: int) -> None:
        N.check(self._ctx, N.lib().srb_set_band(self._ctx, tile_row0, tile_row1))

    # -- multi-GPU band gather fused into the tile kernel -------
    def ExportColorHandle(self) -> bytes:
        """CUDA IPC handle of this window's colour buffer (rank 0 hands it to the band ranks)."""
        buf = (C.c_uint8 * 64)()
        N.check(self._ctx, N.lib().srb_export_color_handle(self._ctx, buf))
        return bytes(buf)

    def SetColorTarget(self, handle: Optional[bytes]) -> None:
        """Store final colour into the buffer behind `handle` (a peer GPU's) instead of the local one."""
        if handle is None:
            N.check(self._ctx, N.lib().srb_set_color_target(self._ctx, None))
        else:
            buf = (C.c_uint8 * 64).from_buffer_copy(handle)
            N.check(self._ctx, N.lib().srb_set_color_target(self._ctx, buf))

    # -- the same gather at display precision: 4 B per pixel over NVLink --------------
    def ExportGatherRgba8(self) -> bytes:
        """Rank 0: create the RGBA8 gather frame; this window's tile kernel fills its own band of it."""
        buf = (C.c_uint8 * 64)()
        N.check(self._ctx, N.lib().srb_gather_rgba8_export(self._ctx, buf))
        return bytes(buf)

    def SetGatherRgba8Target(self, handle: Optional[bytes]) -> None:
        """Band rank: store the band's 8-bit colour into the frame behind `handle` (float4 colour stays local)."""
        if handle is None:
            N.check(self._ctx, N.lib().srb_gather_rgba8_target(self._ctx, None))
        else:
            buf = (C.c_uint8 * 64).from_buffer_copy(handle)
            N.check(self._ctx, N.lib().srb_gather_rgba8_target(self._ctx, buf))

    def GatheredRgba8(self) -> np.ndarray:
        out = np.empty((self.RenderHeight, self.RenderWidth, 4), dtype=np.uint8)
        N.check(self._ctx, N.lib().srb_read_gathered_rgba8(self._ctx, N.ptr(out)))
        return out

    # -- frame barrier in peer memory (one window per GPU) ------------------------------------
    def BarrierExport(self) -> bytes:
        buf = (C.c_uint8 * 64)()
        N.check(self._ctx, N.lib().srb_barrier_export(self._ctx, buf))
        return bytes(buf)

    def BarrierConnect(self, rank: int, world: int, handles) -> None:
        """handles: every rank's BarrierExport() in rank order (None disconnects)."""
        if handles is None:
            N.check(self._ctx, N.lib().srb_barrier_connect(self._ctx, 0, 0, None))
            return
        blob = (C.c_uint8 * (64 * world)).from_buffer_copy(b"".join(handles))
        N.check(self._ctx, N.lib().srb_barrier_connect(self._ctx, rank, world, blob))

    def FrameBarrier(self, slack: int = 0) -> None:
        N.check(self._ctx, N.lib().srb_frame_barrier(self._ctx, slack))

    def TileRowPairs(self) -> np.ndarray:
        """(triangle, tile) pairs binned into each 16-px tile row by the last frame."""
        n = (self.RenderHeight + 15) // 16
        out = np.zeros(n, dtype=np.uint32)
        N.check(self._ctx, N.lib().srb_get_tile_row_pairs(self._ctx, N.ptr(out), n))
        return out

    def SetBinSplit(self, min_entries_full_frame: int, min_entries_band: int, part_entries: int = 128) -> None:
        """Tuning only (never changes the image): bins of at least this many entries are rasterized by several CTAs."""
        N.check(self._ctx, N.lib().srb_set_bin_split(self._ctx, min_entries_full_frame, min_entries_band, part_entries))

    def SetPipelineDepth(self, buffer_sets: int) -> None:
        """Tuning only: frames in flight on the device (2..4, 0 = chosen per frame)."""
        N.check(self._ctx, N.lib().srb_set_pipeline_depth(self._ctx, buffer_sets))

    def SetGraphs(self, enabled: bool) -> None:
        """Tuning only: CUDA-graph replay of the geometry chain of repeating frames (on by default)."""
        N.check(self._ctx, N.lib().srb_set_graphs(self._ctx, 1 if enabled else 0))

    def SetPipeline(self, pipeline: int) -> None:
        N.check(self._ctx, N.lib().srb_set_pipeline(self._ctx, pipeline))

    # -- clears (MainWindow.cs:400-407, 429-436) ------------------------------
    def ClearColorBuffer(self, clear_color) -> None:
        if self._draws:
            self.RenderFrame()
        self._clear_color = np.asarray(clear_color, dtype=np.float32).reshape(4).copy()
        self._clear_flags |= N.SRB_CLEAR_COLOR

    def ClearDepthBuffer(self) -> None:
        if self._draws:
            self.RenderFrame()
        self._clear_flags |= N.SRB_CLEAR_DEPTH

    # -- mesh upload cache (the reference re-passes ToArray() copies every call, Renderer.cs:452-453) --
    # Keyed by array identity: an array object seen before costs a dictionary lookup and two
    # weak-reference checks per RenderMesh call.  An array seen for the first time is hashed once (blake2b, so that
    # equal copies share one device mesh and a collision cannot silently swap geometry); the least recently used
    # device meshes are destroyed beyond MeshCacheCapacity.  Arrays are assumed immutable while cached, as the
    # reference's List<VertexInput> are after load; InvalidateMesh() drops the entry of one that was edited in place.
    @staticmethod
    def _layout(a: np.ndarray):
        return (a.__array_interface__["data"][0], a.nbytes, a.dtype.str, a.strides)

    def mesh_handle(self, vertices, indices) -> int:
        if isinstance(vertices, MeshHandle):                     # explicit resident handle: no hashing at all
            if vertices.window_uid != self._uid:
                raise ValueError("mesh handle belongs to another window")
            return vertices.handle
        ident = (id(vertices), id(indices))
        hit = self._mesh_by_identity.get(ident)
        if hit is not None:
            wv, wi, lay, ckey = hit
            if wv() is vertices and wi() is indices and lay == (self._layout(vertices), self._layout(indices)):
                h = self._mesh_by_content.get(ckey)
                if h is not None:
                    self._mesh_by_content.move_to_end(ckey)
                    return h
            del self._mesh_by_identity[ident]
        v = np.ascontiguousarray(vertices)
        i = np.ascontiguousarray(indices, dtype=np.uint16)
        dig = hashlib.blake2b(digest_size=20)
        dig.update(v.view(np.uint8).reshape(-1))
        dig.update(i.view(np.uint8).reshape(-1))
        ckey = (v.nbytes, i.nbytes, dig.digest())
        h = self._mesh_by_content.get(ckey)
        if h is None:
            flat = v.view(np.float32).reshape(-1)
            if flat.size % 12:
                raise ValueError("vertices must be VertexInput[] (12 floats each)")
            out = C.c_int32(0)
            N.check(self._ctx, N.lib().srb_mesh_create(self._ctx, N.ptr(flat), flat.size // 12, N.ptr(i), i.size,
                                                       C.byref(out)))
            h = out.value
            self._mesh_by_content[ckey] = h
            self._trim_mesh_cache()
        else:
            self._mesh_by_content.move_to_end(ckey)
        if isinstance(vertices, np.ndarray) and isinstance(indices, np.ndarray):
            try:
                if len(self._mesh_by_identity) > 4 * max(1, int(self.MeshCacheCapacity)):
                    self._mesh_by_identity = {k: e for k, e in self._mesh_by_identity.items()
                                              if e[0]() is not None and e[1]() is not None}
                self._mesh_by_identity[ident] = (weakref.ref(vertices), weakref.ref(indices),
                                                 (self._layout(vertices), self._layout(indices)), ckey)
            except TypeError:
                pass
        return h

    def _trim_mesh_cache(self) -> None:
        """Destroy the least recently used device meshes beyond MeshCacheCapacity.  Only between frames: recorded
        draws may name the victim (RenderFrame calls this again once they are submitted)."""
        if self._draws:
            return
        while len(self._mesh_by_content) > max(1, int(self.MeshCacheCapacity)):
            _, victim = self._mesh_by_content.popitem(last=False)
            N.check(self._ctx, N.lib().srb_mesh_destroy(self._ctx, victim))

    def CreateMesh(self, vertices, indices) -> "MeshHandle":
        """Upload once, keep the handle: RenderMesh / Physics.Raycast accept it in place of the vertex array
        (indices=None), and a resident mesh then costs no lookup at all."""
        return MeshHandle(self._uid, self.mesh_handle(vertices, indices))

    def InvalidateMesh(self, vertices, indices) -> None:
        """Forget the identity entry of arrays that were modified in place."""
        self._mesh_by_identity.pop((id(vertices), id(indices)), None)

    def record_draw(self, row: np.ndarray) -> None:
        self._draws.append(row)

    # -- frame ----------------------------------------------------------------
    def RenderFrame(self, near_clip: Optional[float] = None) -> None:
        """MainWindow.RenderFrame (MainWindow.cs:209-212): submit the recorded draws, run the pipeline."""
        from .rasterizer import Rasterizer

        L = N.lib()
        # Rasterizer.RenderDebugMode is a static the reference reads while rasterizing (Rasterizer.cs:419);
        # here it is captured when the frame is submitted.
        u = self.Uniforms.to_struct(Rasterizer.NearClip if near_clip is None else near_clip,
                                    int(Rasterizer.RenderDebugMode))
        N.check(self._ctx, L.srb_set_uniforms(self._ctx, C.byref(u)))
        cc = (C.c_float * 4)(*[float(x) for x in self._clear_color])
        N.check(self._ctx, L.srb_begin_frame(self._ctx, cc, self._clear_flags))
        if self._draws:
            cmds = np.concatenate(self._draws)
            N.check(self._ctx, L.srb_submit_draws(self._ctx, N.ptr(cmds), cmds.shape[0]))
        N.check(self._ctx, L.srb_end_frame(self._ctx))
        self._draws = []
        self._clear_flags = 0
        self._trim_mesh_cache()

    def ApplyUniforms(self, near_clip: Optional[float] = None) -> None:
        """Push the captured shader state (and the numerics_fma switch) to the device without submitting a frame:
        srb_raycast and srb_frustum_test read it."""
        from .rasterizer import Rasterizer
        u = self.Uniforms.to_struct(Rasterizer.NearClip if near_clip is None else near_clip, int(Rasterizer.RenderDebugMode))
        N.check(self._ctx, N.lib().srb_set_uniforms(self._ctx, C.byref(u)))

    def ReplayFrame(self) -> None:
        N.check(self._ctx, N.lib().srb_replay_frame(self._ctx))

    def Sync(self) -> None:
        N.check(self._ctx, N.lib().srb_sync(self._ctx))

    # -- buffers (MainWindow.cs:30-31) -----------------------------------------
    @property
    def ColorBuffer(self) -> np.ndarray:
        out = np.empty((self.RenderHeight, self.RenderWidth, 4), dtype=np.float32)
        N.check(self._ctx, N.lib().srb_read_color(self._ctx, N.ptr(out)))
        return out

    @property
    def DepthBuffer(self) -> np.ndarray:
        out = np.empty((self.RenderHeight, self.RenderWidth), dtype=np.float32)
        N.check(self._ctx, N.lib().srb_read_depth(self._ctx, N.ptr(out)))
        return out

    def ColorBufferRgba8(self) -> np.ndarray:
        out = np.empty((self.RenderHeight, self.RenderWidth, 4), dtype=np.uint8)
        N.check(self._ctx, N.lib().srb_read_color_rgba8(self._ctx, N.ptr(out)))
        return out

    def PresentRgb(self, out: np.ndarray = None, wait: bool = True) -> np.ndarray:
        """The array OnRender uploads (MainWindow.cs:232-246): ColorBuffer flattened to packed RGB floats.
        wait=False queues the copy and returns (pass a pinned `out`; call PresentWait before reading it)."""
        if out is None:
            out = np.empty((self.RenderHeight, self.RenderWidth, 3), dtype=np.float32)
        fn = N.lib().srb_present_rgb if wait else N.lib().srb_present_rgb_async
        N.check(self._ctx, fn(self._ctx, N.ptr(out)))
        return out

    def PresentRgbRows(self, out: np.ndarray, y0: int, y1: int) -> None:
        """Queue the copy of rows [y0,y1) into the full-frame array `out` (band ranks, shared host frame)."""
        N.check(self._ctx, N.lib().srb_present_rgb_rows_async(self._ctx, N.ptr(out), int(y0), int(y1)))

    def PresentRgba8(self, out: np.ndarray = None, wait: bool = True, rows=None) -> np.ndarray:
        """The presented frame at display precision (what the reference's blit leaves in the window's 8-bit
        framebuffer, MainWindow.cs:251-253): uint8 [H, W, 4].  rows=(y0, y1) copies only those rows, at their
        place in `out`."""
        if out is None:
            out = np.empty((self.RenderHeight, self.RenderWidth, 4), dtype=np.uint8)
        L = N.lib()
        if rows is not None:
            N.check(self._ctx, L.srb_present_rgba8_rows_async(self._ctx, N.ptr(out), int(rows[0]), int(rows[1])))
            if wait:
                self.PresentWait()
        else:
            fn = L.srb_present_rgba8 if wait else L.srb_present_rgba8_async
            N.check(self._ctx, fn(self._ctx, N.ptr(out)))
        return out

    def PresentWait(self) -> None:
        N.check(self._ctx, N.lib().srb_present_wait(self._ctx))

    # -- accessors (MainWindow.cs:378-427) ----------------------------------------
    def GetPixel(self, x: int, y: int) -> np.ndarray:
        out = (C.c_float * 4)()
        N.check(self._ctx, N.lib().srb_get_pixel(self._ctx, x, y, out))
        return np.array(list(out), dtype=np.float32)

    def SetPixel(self, x: int, y: int, color) -> None:
        c = (C.c_float * 4)(*[float(v) for v in color])
        N.check(self._ctx, N.lib().srb_set_pixel(self._ctx, x, y, c))

    def GetDepth(self, x: int, y: int) -> float:
        out = C.c_float()
        N.check(self._ctx, N.lib().srb_get_depth(self._ctx, x, y, C.byref(out)))
        return out.value

    def SetDepth(self, x: int, y: int, depth: float) -> None:
        N.check(self._ctx, N.lib().srb_set_depth(self._ctx, x, y, float(depth)))

    def Stats(self) -> N.SrbStats:
        self.Sync()
        s = N.SrbStats()
        N.check(self._ctx, N.lib().srb_get_stats(self._ctx, C.byref(s)))
        return s
