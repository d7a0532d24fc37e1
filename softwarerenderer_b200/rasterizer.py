"""Rasterizer / FrustumCuller -- host-side mirror of the reference's static classes.

Rasterizer.RenderMesh keeps the reference signature (Rasterizer.cs:163-174); the
body is the CUDA pipeline behind libsrb200.so.  Only the stock shader pair
(Renderer.cs:830-860) is executable: any other delegate raises
NotImplementedError -- there is no CPU fallback by design.
"""
from __future__ import annotations

import enum
from dataclasses import dataclass

import numpy as np

from . import _native as N
from .shaders import Shaders


class BlendMode(enum.IntEnum):      # Rasterizer.cs:25-31
    None_ = 0
    Alpha = 1
    Additive = 2
    Multiply = 3


class DepthTest(enum.IntEnum):      # Rasterizer.cs:33-43
    Disabled = 0
    Less = 1
    LessEqual = 2
    Greater = 3
    GreaterEqual = 4
    Equal = 5
    NotEqual = 6
    Always = 7


class CullMode(enum.IntEnum):       # Rasterizer.cs:45-50
    None_ = 0
    Back = 1
    Front = 2


class DebugMode(enum.IntEnum):      # Rasterizer.cs:14-18
    None_ = 0
    Wireframe = 1


@dataclass(frozen=True)
class BoundingSphere:               # FrustumCuller.cs:7-17
    Center: tuple
    Radius: float


def _mat(m) -> np.ndarray:
    a = np.ascontiguousarray(np.asarray(m, dtype=np.float32).reshape(16))
    return a


class Rasterizer:
    BlendMode = BlendMode
    DepthTest = DepthTest
    CullMode = CullMode
    DebugMode = DebugMode

    NearClip = 0.1                  # Rasterizer.cs:20
    FarClip = 1000.0                # Rasterizer.cs:21
    RenderDebugMode = DebugMode.None_

    @staticmethod
    def RenderMesh(window, vertices, indices, model, view, projection, vertexShader, fragmentShader,
                   cullMode=CullMode.Back, depthTest=DepthTest.LessEqual, blendMode=BlendMode.Alpha,
                   submissionIndex: int = -1, sphereBounds: BoundingSphere = None) -> None:
        """Rasterizer.RenderMesh (Rasterizer.cs:163-230).  Records one draw on `window`.

        `sphereBounds`, when given, moves the caller's FrustumCuller.IsSphereInFrustum test
        (Renderer.cs:446) onto the device for this draw."""
        if window.RenderWidth <= 0 or window.RenderHeight <= 0:      # :176 silent return
            return
        vs_ok = vertexShader is Shaders.StockVertexShader or isinstance(vertexShader, Shaders.StockVertexShader)
        if not vs_ok or not isinstance(fragmentShader, Shaders.StockFragmentShader):
            raise NotImplementedError(
                "only the stock shader pair (Renderer.cs:830-860) runs on the GPU; arbitrary delegates are "
                "unsupported and there is no CPU fallback")
        row = np.zeros(1, dtype=N.DRAW_CMD_DTYPE)
        row["mesh"] = window.mesh_handle(vertices, indices)
        tex = fragmentShader.texture
        row["texture"] = 0 if tex is None else tex.handle(window)
        row["model"], row["view"], row["proj"] = _mat(model), _mat(view), _mat(projection)
        row["cull"], row["depth_test"], row["blend"] = int(cullMode), int(depthTest), int(blendMode)
        row["submission_index"] = submissionIndex
        if sphereBounds is not None:
            row["sphere"] = (*sphereBounds.Center, sphereBounds.Radius)
            row["flags"] = N.SRB_DRAW_FRUSTUM_CULL
        window.record_draw(row)


class FrustumCuller:
    @staticmethod
    def IsSphereInFrustum(window, bounds: BoundingSphere, modelMatrix, viewMatrix, projectionMatrix) -> bool:
        """FrustumCuller.IsSphereInFrustum (FrustumCuller.cs:201-218), evaluated on the device."""
        return bool(FrustumCuller.TestSpheres(window, [bounds], [modelMatrix], viewMatrix, projectionMatrix)[0])

    @staticmethod
    def TestSpheres(window, bounds, models, viewMatrix, projectionMatrix) -> np.ndarray:
        n = len(bounds)
        sph = np.array([(*b.Center, b.Radius) for b in bounds], dtype=np.float32).reshape(n, 4)
        mod = np.ascontiguousarray(np.asarray(models, dtype=np.float32).reshape(n, 16))
        out = np.zeros(n, dtype=np.int32)
        # the converted matrices are bound to locals: N.ptr() keeps no reference, and a temporary would be freed
        # (and its block reused by the next conversion) before the C call reads it
        v, p = _mat(viewMatrix), _mat(projectionMatrix)
        window.ApplyUniforms()                      # numerics_fma selects the form of Vector4.Transform
        N.check(window._ctx, N.lib().srb_frustum_test(window._ctx, N.ptr(sph), N.ptr(mod), N.ptr(v), N.ptr(p), n, N.ptr(out)))
        return out

    @staticmethod
    def CalculateBoundingSphere(vertices) -> BoundingSphere:
        """Load-time helper (FrustumCuller.cs:59-151 is an order-dependent parallel Ritter sphere;
        out of the hot path).  A deterministic Ritter pass is used for scene construction."""
        p = np.asarray(vertices["Position"], dtype=np.float64)
        if p.shape[0] == 0:
            return BoundingSphere((0.0, 0.0, 0.0), 0.0)
        a = p[np.argmax(np.linalg.norm(p - p[0], axis=1))]
        b = p[np.argmax(np.linalg.norm(p - a, axis=1))]
        c = (a + b) / 2
        r = np.linalg.norm(b - a) / 2
        for q in p:
            d = np.linalg.norm(q - c)
            if d > r:
                nr = (r + d) / 2
                c = c + (q - c) * ((nr - r) / d)
                r = nr
        return BoundingSphere(tuple(float(x) for x in c), float(r) * 1.0001)
