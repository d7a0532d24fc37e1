"""Physics.Raycast (Physics.cs:19-179): known-answer tests that pin the oracle (CPU), and the
CUDA batch kernel against the oracle (-m gpu)."""
import math

import numpy as np
import pytest

F = np.float32
I = np.eye(4, dtype=np.float32)


def _tri_mesh(tris, normal=(0, 0, 1)):
    from softwarerenderer_b200.shaders import Shaders
    pos = np.asarray(tris, dtype=np.float32).reshape(-1, 3)
    v = Shaders.make_vertices(pos, normal=np.tile(np.asarray(normal, np.float32), (pos.shape[0], 1)))
    return v, np.arange(pos.shape[0], dtype=np.uint16)


FRONT = [(-1, -1, -5), (1, -1, -5), (0, 1, -5)]          # counter-clockwise seen from the origin: det > 0
BACK = [(-1, -1, -5), (0, 1, -5), (1, -1, -5)]


def test_front_face_hit_distance_point_normal(oracle):
    v, ix = _tri_mesh([FRONT])
    hit, dist, pt, nr, tri = oracle.raycast((0, 0, 0), (0, 0, -1), v, ix, I)
    assert hit and tri == 0 and dist == 5.0
    assert np.array_equal(pt, [0, 0, -5]) and np.allclose(nr, [0, 0, 1])
    # the direction is normalised first (Physics.cs:70): the distance is in world units whatever its length
    hit, dist, pt, _, _ = oracle.raycast((0, 0, 0), (0, 0, -10), v, ix, I)
    assert hit and dist == 5.0
    # pointing away: distance < 0 is no hit (:171-172); defaults float.MaxValue / zero vectors (:66-68)
    hit, dist, pt, nr, tri = oracle.raycast((0, 0, 0), (0, 0, 1), v, ix, I)
    assert not hit and dist == np.finfo(np.float32).max and not pt.any() and not nr.any() and tri == -1


def test_face_mask(oracle):
    vf, ix = _tri_mesh([FRONT])
    vb, _ = _tri_mesh([BACK])
    ray = ((0, 0, 0), (0, 0, -1))
    assert oracle.raycast(*ray, vf, ix, I, face_mask=1)[0] and not oracle.raycast(*ray, vb, ix, I, face_mask=1)[0]   # IgnoreBackfaces (:156)
    assert not oracle.raycast(*ray, vf, ix, I, face_mask=2)[0] and oracle.raycast(*ray, vb, ix, I, face_mask=2)[0]   # IgnoreFrontfaces (:157)
    assert oracle.raycast(*ray, vf, ix, I, face_mask=0)[0] and oracle.raycast(*ray, vb, ix, I, face_mask=0)[0]
    assert not oracle.raycast(*ray, vf, ix, I, face_mask=3)[0]
    # a ray in the triangle's plane: |det| < 1e-8 (:158)
    assert not oracle.raycast((-5, 0, -5), (1, 0, 0), vf, ix, I, face_mask=0)[0]


def test_closest_hit_and_tie_goes_to_the_lowest_index(oracle):
    far = [(x, y, -9) for x, y, _ in FRONT]
    v, ix = _tri_mesh([far, FRONT, FRONT])
    hit, dist, _, _, tri = oracle.raycast((0, 0, 0), (0, 0, -1), v, ix, I)
    assert hit and dist == 5.0 and tri == 1           # strict `<` (:104): the first of two equal distances stays


def test_edges_are_inclusive_u_v_bounds(oracle):
    v, ix = _tri_mesh([FRONT])
    assert oracle.raycast((0, -1, 0), (0, 0, -1), v, ix, I)[0]          # on the edge v0-v1: v == 0 passes (v < 0 fails only)
    assert oracle.raycast((-1, -1, 0), (0, 0, -1), v, ix, I)[0]         # at v0: u == 0, v == 0
    assert not oracle.raycast((0, -1.001, 0), (0, 0, -1), v, ix, I)[0]


def test_model_matrix_moves_the_mesh_and_normal_uses_inverse_transpose(oracle):
    from softwarerenderer_b200 import numerics as nm
    n = np.array([1, 1, 0], dtype=np.float32) / F(math.sqrt(2))
    v, ix = _tri_mesh([FRONT], normal=n)
    M = np.diag([2, 1, 1, 1]).astype(np.float32)
    M[3, :3] = (10, 0, 0)                                                # row-vector convention: translation in row 4
    hit, dist, pt, nr, _ = oracle.raycast((10, 0, 0), (0, 0, -1), v, ix, M)
    assert hit and dist == 5.0 and np.array_equal(pt, [10, 0, -5])
    want = np.array([0.5, 1, 0]) / np.linalg.norm([0.5, 1, 0])          # n * transpose(inverse(M)), renormalised (:47-48)
    assert np.allclose(nr, want, atol=1e-6)
    assert not oracle.raycast((0, 0, 0), (0, 0, -1), v, ix, M)[0]       # the mesh is no longer in front of the origin
    S0 = np.diag([1, 0, 1, 1]).astype(np.float32)                        # singular: Matrix4x4.Invert fails (:30-36)
    hit, dist, _, _, _ = oracle.raycast((0, 0, 0), (0, 0, -1), v, ix, S0)
    assert not hit and dist == np.finfo(np.float32).max
    assert nm.identity().shape == (4, 4)


def _brute_force_f64(o, d, v, ix, M, mask):
    pos = np.c_[v["Position"].astype(np.float64), np.ones(len(v))] @ M.astype(np.float64)
    p = pos[:, :3]
    d = d / np.linalg.norm(d)
    t = ix.reshape(-1, 3)
    v0, e1, e2 = p[t[:, 0]], p[t[:, 1]] - p[t[:, 0]], p[t[:, 2]] - p[t[:, 0]]
    pv = np.cross(d, e2)
    det = (e1 * pv).sum(1)
    ok = np.abs(det) >= 1e-8
    if mask & 1:
        ok &= det >= 1e-8
    if mask & 2:
        ok &= det <= -1e-8
    with np.errstate(divide="ignore", invalid="ignore"):
        inv = 1.0 / det
        tv = o - v0
        u = (tv * pv).sum(1) * inv
        q = np.cross(tv, e1)
        w = (q * d).sum(1) * inv
        dist = (e2 * q).sum(1) * inv
    ok &= (u >= 0) & (u <= 1) & (w >= 0) & (u + w <= 1) & (dist >= 0)
    if not ok.any():
        return False, np.inf, -1
    k = int(np.argmin(np.where(ok, dist, np.inf)))
    return True, float(dist[k]), k


def test_against_a_float64_brute_force_on_the_gun(oracle):
    from softwarerenderer_b200 import scene as S
    gun = S.load_model("Gun")
    rng = np.random.default_rng(3)
    checked = hits = 0
    for m in gun.Meshes:
        c, r = np.array(m.SphereBounds.Center), m.SphereBounds.Radius
        for _ in range(40):
            o = c + rng.normal(size=3) * 3 * r
            tgt = c + rng.uniform(-0.6, 0.6, 3) * r
            d = tgt - o
            for mask in (0, 1):
                hit, dist, pt, nr, tri = oracle.raycast(o, d, m.Vertices, m.Indices, I, face_mask=mask)
                bh, bd, bt = _brute_force_f64(o.astype(np.float32).astype(np.float64), d.astype(np.float32).astype(np.float64),
                                              m.Vertices, m.Indices, I, mask)
                checked += 1
                if hit != bh:                        # grazing an edge: float32 and float64 may disagree; must be rare
                    continue
                if hit:
                    hits += 1
                    assert abs(dist - bd) <= 1e-4 * max(1.0, bd)
                    assert abs(np.linalg.norm(nr) - 1.0) < 1e-5
                    assert np.allclose(pt, o.astype(np.float32) + (d / np.linalg.norm(d)).astype(np.float32) * dist, atol=1e-4)
    assert hits > 50 and checked == 400


# ---------------------------------------------------------------------------
# CUDA batch kernel vs the oracle
# ---------------------------------------------------------------------------
def _check_batch(oracle, win, mesh, M, origins, dirs, mask, fma=1):
    import softwarerenderer_b200 as srb
    win.Uniforms.NumericsFma = fma                      # both forms of Vector3.Transform
    got = srb.Physics.RaycastBatch(win, origins, dirs, mesh.Vertices, mesh.Indices, M, srb.RaycastFaceMask(mask))
    nhit = 0
    for k in range(len(origins)):
        hit, dist, pt, nr, tri = oracle.raycast(origins[k], dirs[k], mesh.Vertices, mesh.Indices, M, face_mask=mask, fma=fma)
        g = got[k]
        assert bool(g["hit"]) == hit, (k, g, hit, dist, tri)
        assert g["triangle"] == tri
        assert g["distance"] == F(dist)                                 # same single-rounded operations: bit-exact
        assert np.array_equal(g["point"], pt)
        assert np.allclose(g["normal"], nr, atol=1e-6)                  # stated tolerance: normal within 1e-6 per component
        nhit += hit
    return nhit


@pytest.mark.gpu
@pytest.mark.parametrize("fma", [0, 1])
@pytest.mark.parametrize("mask", [0, 1, 2])
def test_cuda_raycast_matches_the_oracle_on_the_bundled_models(oracle, mask, fma):
    import softwarerenderer_b200 as srb
    from softwarerenderer_b200 import numerics as nm
    from softwarerenderer_b200 import scene as S
    rng = np.random.default_rng(10 + mask)
    win = srb.MainWindow(64, 64)
    try:
        total = 0
        for name in ("Gun", "gordon_freeman"):
            for m in S.load_model(name).Meshes[:4]:
                c, r = np.array(m.SphereBounds.Center, dtype=np.float64), m.SphereBounds.Radius
                M = (nm.create_scale(1.5) @ nm.create_rotation_y(0.7)).astype(np.float32)
                M[3, :3] = (3.0, -1.0, 2.0)
                cw = (np.r_[c, 1.0] @ M.astype(np.float64))[:3]
                n = 150                                                      # > 2 ray chunks of 64
                origins = (cw + rng.normal(size=(n, 3)) * 4 * r * 1.5).astype(np.float32)
                dirs = (cw + rng.uniform(-0.7, 0.7, (n, 3)) * r * 1.5 - origins).astype(np.float32)
                total += _check_batch(oracle, win, m, M, origins, dirs, mask, fma)
        assert total > 100
    finally:
        win.Close()


@pytest.mark.gpu
def test_cuda_raycast_single_call_defaults_and_singular_model():
    import softwarerenderer_b200 as srb
    v, ix = _tri_mesh([FRONT, [(x, y, -9) for x, y, _ in FRONT], FRONT])
    win = srb.MainWindow(64, 64)
    try:
        hit, dist, pt, nr = srb.Physics.Raycast(win, (0, 0, 0), (0, 0, -3), v, ix, I)
        assert hit and dist == 5.0 and np.array_equal(pt, [0, 0, -5]) and np.allclose(nr, [0, 0, 1])
        assert srb.Physics.RaycastBatch(win, [(0, 0, 0)], [(0, 0, -1)], v, ix, I)[0]["triangle"] == 0    # tie -> lowest index
        hit, dist, pt, nr = srb.Physics.Raycast(win, (0, 0, 0), (0, 0, 1), v, ix, I)
        assert not hit and dist == np.finfo(np.float32).max and not pt.any() and not nr.any()
        S0 = np.diag([1, 0, 1, 1]).astype(np.float32)
        hit, dist, _, _ = srb.Physics.Raycast(win, (0, 0, 0), (0, 0, -1), v, ix, S0)
        assert not hit and dist == np.finfo(np.float32).max
        with pytest.raises(srb.SrbError):
            srb.Physics.RaycastBatch(win, [(0, 0, 0)], [(0, 0, -1)], v, ix, I, 7)
    finally:
        win.Close()


@pytest.mark.gpu
@pytest.mark.parametrize("fma", [0, 1])
def test_cuda_raycast_large_batch_on_a_dense_grid(oracle, fma):
    """4,096 rays against a 64,800-triangle grid mesh of the stress scene; a sample is checked ray by ray."""
    import softwarerenderer_b200 as srb
    from softwarerenderer_b200 import scene as S
    sc = S.scene_stress(256, 144, n_meshes=1, grid=181)
    m = sc.draws[0].mesh
    M = np.asarray(sc.draws[0].model, dtype=np.float32).reshape(4, 4)
    p = np.c_[m.Vertices["Position"].astype(np.float64), np.ones(len(m.Vertices))] @ M.astype(np.float64)
    lo, hi = p[:, :3].min(0), p[:, :3].max(0)
    rng = np.random.default_rng(5)
    n = 4096
    tgt = rng.uniform(lo, hi, (n, 3))
    origins = (tgt + np.array([0, 0, 5.0]) + rng.normal(size=(n, 3)) * 0.3).astype(np.float32)
    dirs = (tgt - origins).astype(np.float32)
    win = srb.MainWindow(64, 64)
    try:
        win.Uniforms.NumericsFma = fma
        got = srb.Physics.RaycastBatch(win, origins, dirs, m.Vertices, m.Indices, M, srb.RaycastFaceMask.None_)
        assert got["hit"].mean() > 0.5
        for k in range(0, n, 64):
            hit, dist, pt, nr, tri = oracle.raycast(origins[k], dirs[k], m.Vertices, m.Indices, M, face_mask=0, fma=fma)
            assert bool(got[k]["hit"]) == hit and got[k]["triangle"] == tri and got[k]["distance"] == F(dist)
    finally:
        win.Close()
