"""Physics -- host-side mirror of the reference's static class (Physics.cs:16-179).

Physics.Raycast keeps the reference's argument meaning (ray, VertexInput[], ushort[], model, face mask) and
returns its four results; the body is the CUDA batch kernel behind srb_raycast.  RaycastBatch is the same
call for many rays at once (the mesh is transformed and tested once per batch instead of once per ray).
"""
from __future__ import annotations

import enum

import numpy as np

from . import _native as N


class RaycastFaceMask(enum.IntFlag):       # Physics.cs:8-14
    None_ = 0
    IgnoreBackfaces = 1
    IgnoreFrontfaces = 2


class Physics:
    RaycastFaceMask = RaycastFaceMask

    @staticmethod
    def RaycastBatch(window, rayOrigins, rayDirections, vertices, indices, model,
                     faceMask=RaycastFaceMask.IgnoreBackfaces) -> np.ndarray:
        """n rays -> structured array (hit, distance, point, normal, triangle), one row per ray."""
        o = np.asarray(rayOrigins, dtype=np.float32).reshape(-1, 3)
        d = np.asarray(rayDirections, dtype=np.float32).reshape(-1, 3)
        if o.shape != d.shape:
            raise ValueError("rayOrigins and rayDirections must have the same length")
        rays = np.ascontiguousarray(np.concatenate([o, d], axis=1))
        m = np.ascontiguousarray(np.asarray(model, dtype=np.float32).reshape(16))
        out = np.zeros(rays.shape[0], dtype=N.RAY_HIT_DTYPE)
        window.ApplyUniforms()                  # numerics_fma selects the form of Vector3.Transform (Physics.cs:56-58)
        N.check(window._ctx, N.lib().srb_raycast(window._ctx, window.mesh_handle(vertices, indices), N.ptr(m), N.ptr(rays),
                                                 rays.shape[0], int(faceMask), N.ptr(out)))
        return out

    @staticmethod
    def Raycast(window, rayOrigin, rayDirection, vertices, indices, model,
                faceMask=RaycastFaceMask.IgnoreBackfaces):
        """Physics.Raycast (Physics.cs:19-52) -> (hit, hitDistance, hitPoint, hitNormal)."""
        h = Physics.RaycastBatch(window, [rayOrigin], [rayDirection], vertices, indices, model, faceMask)[0]
        return bool(h["hit"]), float(h["distance"]), h["point"].copy(), h["normal"].copy()
