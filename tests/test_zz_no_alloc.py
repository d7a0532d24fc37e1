"""-m gpu, collected last: srb_raycast_static -- srb_raycast for callers that cannot allocate (its scratch is a fixed
device array inside the library; no counterpart in the reference).
Through the C ABI, against the existing entry point, on two calls in a row: same hit flags, triangles, distances,
points and normals bit for bit (both run the same kernels, so the raycast tests' tolerances hold with equality)."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _both(win, handle, M, rays, mask):
    from softwarerenderer_b200 import _native as N
    L = N.lib()
    m = np.ascontiguousarray(np.asarray(M, dtype=np.float32).reshape(16))
    rays = np.ascontiguousarray(rays, dtype=np.float32)
    n = rays.shape[0]
    a = np.zeros(n, dtype=N.RAY_HIT_DTYPE)
    b = np.zeros(n, dtype=N.RAY_HIT_DTYPE)
    N.check(win._ctx, L.srb_raycast(win._ctx, handle, N.ptr(m), N.ptr(rays), n, mask, N.ptr(a)))
    N.check(win._ctx, L.srb_raycast_static(win._ctx, handle, N.ptr(m), N.ptr(rays), n, mask, N.ptr(b)))
    return a, b


def test_raycast_static_matches_raycast_on_two_calls_in_a_row():
    import softwarerenderer_b200 as srb
    from softwarerenderer_b200 import _native as N
    from softwarerenderer_b200 import numerics as nm
    from softwarerenderer_b200 import scene as S
    mesh = S.load_model("Gun").Meshes[0]
    c, r = np.array(mesh.SphereBounds.Center, dtype=np.float64), mesh.SphereBounds.Radius
    M = (nm.create_scale(1.5) @ nm.create_rotation_y(0.7)).astype(np.float32)
    M[3, :3] = (3.0, -1.0, 2.0)
    cw = (np.r_[c, 1.0] @ M.astype(np.float64))[:3]
    rng = np.random.default_rng(77)
    win = srb.MainWindow(64, 64)
    try:
        win.ApplyUniforms()
        h = win.mesh_handle(mesh.Vertices, mesh.Indices)
        # first call: a large batch, most rays aimed at the mesh
        n1 = 3000
        o1 = (cw + rng.normal(size=(n1, 3)) * 6 * r).astype(np.float32)
        d1 = (cw + rng.uniform(-0.7, 0.7, (n1, 3)) * r * 1.5 - o1).astype(np.float32)
        a, b = _both(win, h, M, np.concatenate([o1, d1], axis=1), 0)
        assert a["hit"].sum() > 100
        assert a.tobytes() == b.tobytes()
        # second call right after it: fewer rays, half of them pointing away (the scratch still holds the first
        # call's keys and hits beyond and below n2 -- none of it may show)
        n2 = 517
        o2 = (cw + rng.normal(size=(n2, 3)) * 6 * r).astype(np.float32)
        d2 = (cw - o2).astype(np.float32)
        d2[::2] *= -1.0
        a2, b2 = _both(win, h, M, np.concatenate([o2, d2], axis=1), 1)
        assert 0 < a2["hit"].sum() < n2
        assert a2.tobytes() == b2.tobytes()
        # the limit is enforced with a clear error, and nothing is written
        L = N.lib()
        too_many = N.SRB_RAYCAST_STATIC_MAX_RAYS + 1
        rays = np.zeros((too_many, 6), dtype=np.float32)
        out = np.zeros(too_many, dtype=N.RAY_HIT_DTYPE)
        m = np.ascontiguousarray(M.reshape(16))
        rc = L.srb_raycast_static(win._ctx, h, N.ptr(m), N.ptr(rays), too_many, 0, N.ptr(out))
        assert rc == N.SRB_ERR_ARG
        assert b"at most 4096 rays" in L.srb_last_error(win._ctx)
        # the full-size batch is accepted
        full = N.SRB_RAYCAST_STATIC_MAX_RAYS
        o3 = (cw + rng.normal(size=(full, 3)) * 6 * r).astype(np.float32)
        d3 = (cw - o3).astype(np.float32)
        a3, b3 = _both(win, h, M, np.concatenate([o3, d3], axis=1), 0)
        assert a3.tobytes() == b3.tobytes() and a3["hit"].sum() > full // 2
    finally:
        win.Close()
