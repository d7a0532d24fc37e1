"""-m gpu: edge cases of the path against the oracle -- empty frames, targets smaller than a tile, degenerate,
off-screen, NaN/Inf and w == 0 geometry, a draw with no triangles, everything frustum-culled, resize."""
import numpy as np
import pytest

from tests.test_gpu_parity import _render_gpu
from tests.util import CLEAR_DEPTH, compare, screen_tri_scene

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("size", [(1, 1), (7, 3), (16, 16), (17, 33), (250, 1)])
@pytest.mark.parametrize("pipeline", [1, 2])
def test_frame_without_draws_is_the_clear(size, pipeline):
    from softwarerenderer_b200.scene import Scene
    W, H = size
    sc = Scene(W, H, [], clear_color=(0.25, 0.5, 0.75, 1.0))
    c, d, st = _render_gpu(sc, pipeline)
    assert c.shape == (H, W, 4) and (d == CLEAR_DEPTH).all()
    assert np.array_equal(c, np.broadcast_to(np.array([0.25, 0.5, 0.75, 1.0], np.float32), c.shape))
    assert st.triangles_submitted == 0


@pytest.mark.parametrize("size", [(1, 1), (5, 9), (15, 15), (31, 17)])
@pytest.mark.parametrize("pipeline", [1, 2])
def test_targets_smaller_than_or_straddling_a_tile(oracle, size, pipeline):
    W, H = size
    tris = [[(-4, -4), (3 * W + 8, -4), (-4, 3 * H + 8)], [(0.25, 0.25), (W - 0.5, 0.5), (W * 0.5, H - 0.25)]]
    sc = screen_tri_scene_any(W, H, tris)
    oc, od, _ = oracle.render_scene(sc, threads=2)
    gc, gd, _ = _render_gpu(sc, pipeline)
    compare(gc, gd, oc, od, min_frac=0.0 if W * H < 100 else 0.999)
    assert np.array_equal(gd, od)


def screen_tri_scene_any(W, H, tris, **kw):
    """screen_tri_scene for any W, H: vertices are given in pixels and mapped through the same float32 viewport
    arithmetic both sides use, so exact invertibility does not matter -- oracle and GPU get identical inputs."""
    return screen_tri_scene(W, H, tris, cull=0, colors=[(1.0, 0.5, 0.25, 1.0), (0.2, 0.9, 0.4, 1.0)][: len(tris)],
                            ndc_z=[0.25, 0.5][: len(tris)], **kw)


@pytest.mark.parametrize("pipeline", [1, 2])
def test_degenerate_offscreen_and_non_finite_triangles(oracle, pipeline):
    from softwarerenderer_b200 import numerics as nm
    from softwarerenderer_b200.scene import DrawItem, Mesh, Scene
    from softwarerenderer_b200.shaders import Material, Shaders
    W, H = 64, 48
    nan, inf = float("nan"), float("inf")
    pos = [
        (-0.5, -0.5, 0.5), (0.5, -0.5, 0.5), (0.0, 0.5, 0.5),          # 0: ordinary
        (0.1, 0.1, 0.5), (0.1, 0.1, 0.5), (0.1, 0.1, 0.5),             # 1: a point (area 0, :396)
        (-0.9, -0.9, 0.5), (0.0, 0.0, 0.5), (0.9, 0.9, 0.5),           # 2: collinear (area 0)
        (5.0, 5.0, 0.5), (6.0, 5.0, 0.5), (5.0, 6.0, 0.5),             # 3: entirely off-screen (empty bbox)
        (nan, 0.0, 0.5), (0.5, 0.5, 0.5), (0.0, 0.5, 0.5),             # 4: NaN position (:378-380)
        (inf, 0.0, 0.5), (0.5, 0.5, 0.5), (0.0, 0.5, 0.5),             # 5: Inf position
        (-1e15, -1e15, 0.5), (1e15, -1e15, 0.5), (0.0, 1e15, 0.5),     # 6: huge but finite edge products: covers the whole target
        (-1e30, -1e30, 0.4), (1e30, -1e30, 0.4), (0.0, 1e30, 0.4),     # 6b: edge products overflow -> NaN area, covers nothing
        (-0.2, -0.2, 0.2), (0.2, -0.2, 0.2), (0.0, 0.2, 0.2),          # 7: small, nearer
    ]
    col = np.repeat(np.array([(1, 0, 0, 1), (0, 1, 0, 1), (0, 0, 1, 1), (1, 1, 0, 1), (0, 1, 1, 1), (1, 0, 1, 1),
                              (0.5, 0.5, 0.5, 1), (0.3, 0.3, 0.3, 1), (1, 1, 1, 1)], dtype=np.float32), 3, axis=0)
    v = Shaders.make_vertices(pos, color=col)
    I = nm.identity()
    mesh = Mesh(v, np.arange(len(pos), dtype=np.uint16), Material())
    empty = Mesh(Shaders.make_vertices([(0, 0, 0)]), np.zeros(0, dtype=np.uint16), Material())     # a draw with no triangles
    sc = Scene(W, H, [DrawItem(empty, I, I, I, cull=0), DrawItem(mesh, I, I, I, cull=0), DrawItem(empty, I, I, I, cull=0)],
               clear_color=(0, 0, 0, 0), fog_start=10.0, fog_end=20.0)
    oc, od, ost = oracle.render_scene(sc, threads=2)
    gc, gd, st = _render_gpu(sc, pipeline)
    compare(gc, gd, oc, od)
    assert (od != CLEAR_DEPTH).all()                                   # the huge triangle covers everything
    assert st.triangles_submitted == 9
    assert abs(float(od[0, 0]) + 0.75) < 1e-6 and abs(float(od[H - 1, W - 1]) + 0.75) < 1e-6   # (0.5 + 1) / 2, negated: triangle 6, not 6b


@pytest.mark.parametrize("fma", [0, 1])
def test_w_zero_and_behind_the_camera(oracle, fma):
    """Corners with w <= 0 go through the near clip (:208-224); all three behind: dropped; w == 0 exactly (:393).
    Both forms of Shaders.Lerp / Vector4.Transform (numerics_fma)."""
    from softwarerenderer_b200 import numerics as nm
    from softwarerenderer_b200.scene import DrawItem, Mesh, Scene, default_projection
    from softwarerenderer_b200.shaders import Material, Shaders
    W, H = 160, 96
    pos = [(-1, -1, -3), (1, -1, -3), (0, 1, 2),          # one corner behind the camera
           (-1, -1, 2), (1, -1, 3), (0, 1, 4),            # all behind
           (-2, -0.5, 0.0), (2, -0.5, 0.0), (0, 0.5, -4), # two corners at z == 0: w == 0 under the projection
           (-1, 0.2, -2), (1, 0.2, -2), (0, 1.0, -2)]     # ordinary
    v = Shaders.make_vertices(pos)
    mesh = Mesh(v, np.arange(12, dtype=np.uint16), Material())
    I = nm.identity()
    sc = Scene(W, H, [DrawItem(mesh, I, I, default_projection(W, H), cull=0)])
    sc.numerics_fma = fma
    oc, od, _ = oracle.render_scene(sc, threads=2)
    for pipeline in (1, 2):
        gc, gd, st = _render_gpu(sc, pipeline)
        compare(gc, gd, oc, od)
        assert st.clipped_triangles >= 1


def test_everything_frustum_culled_leaves_the_clear(oracle):
    from softwarerenderer_b200 import scene as S
    sc = S.scene_city(320, 180, n_instances=6)
    for d in sc.draws:                                    # move every instance far behind the camera
        m = np.array(d.model, dtype=np.float32).reshape(4, 4).copy()
        m[3, 2] += 5000.0
        d.model = m
    assert not any(oracle.frustum_flags(sc))
    gc, gd, st = _render_gpu(sc, 0)
    assert st.draws_visible == 0 and (gd == CLEAR_DEPTH).all()


def test_resize_between_frames(oracle):
    """MainWindow.HandleResize (MainWindow.cs:298-361): new buffers, same meshes and textures."""
    import softwarerenderer_b200 as srb
    from softwarerenderer_b200 import scene as S
    win = srb.MainWindow(320, 180)
    try:
        for W, H in ((320, 180), (640, 361), (100, 40), (320, 180)):
            win.Resize(W, H)
            assert (win.RenderWidth, win.RenderHeight) == (W, H)
            sc = S.scene_model("Gun", W, H)
            S.submit_scene(win, sc)
            oc, od, _ = oracle.render_scene(sc, threads=4)
            compare(win.ColorBuffer, win.DepthBuffer, oc, od)
    finally:
        win.Close()


@pytest.mark.parametrize("pipeline", [1, 2])
@pytest.mark.parametrize("normals", [[(0, 0, 1), (0, 0, 1), (0, 0, -1)], [(0, 0, 1), (0, 0, -1), (0, 0, 1)],
                                     [(1, 0, 0), (-1, 0, 0), (0, 1e-4, 0)]])
def test_cancelling_vertex_normals(oracle, pipeline, normals):
    """InterpolateData's Vector3 branch (Rasterizer.cs:680-688): the interpolated WorldNormal is renormalised only
    above a squared length of 1e-6; where the corner normals cancel the fragment keeps the short vector (lighting
    floor, no NaN).  Same scenes as the known-answer tests in test_oracle_kat.py."""
    from tests.test_oracle_kat import _one_tri_scene
    from softwarerenderer_b200 import numerics as nm
    W = H = 32
    t = [(0, 0), (16, 0), (0, 16)]
    pos = [(2.0 * x / W - 1.0, 1.0 - 2.0 * y / H, 0.5) for x, y in t]
    sc = _one_tri_scene(W, H, pos, nm.identity(), normal=normals, color=[(0.8, 0.6, 0.4, 1.0)] * 3)
    sc.light_dir = np.array((0, 0, -1), np.float32)
    oc, od, _ = oracle.render_scene(sc, threads=2)
    gc, gd, _ = _render_gpu(sc, pipeline)
    assert np.isfinite(gc).all()
    assert np.array_equal(gd, od)
    compare(gc, gd, oc, od)


@pytest.mark.parametrize("pipeline", [1, 2])
def test_triangle_behind_the_eye_draws_nothing(oracle, pipeline):
    """All three clip w <= 0 (Rasterizer.cs:208-215)."""
    from tests.test_oracle_kat import _one_tri_scene
    from softwarerenderer_b200.scene import default_projection
    W, H = 64, 48
    sc = _one_tri_scene(W, H, [(-1, -1, 2.0), (1, -1, 2.0), (0, 1, 2.0)], default_projection(W, H))
    gc, gd, st = _render_gpu(sc, pipeline)
    oc, od, _ = oracle.render_scene(sc, threads=2)
    assert (gd == CLEAR_DEPTH).all() and np.array_equal(gd, od) and np.array_equal(gc, oc)


def test_frame_that_keeps_the_framebuffer_grows_its_buffers(oracle):
    """More near-clipped triangles than the clip side buffer's first capacity (4096), in a frame that does NOT
    clear: every raster kernel returns at once when a capacity overflowed, so the framebuffer is untouched and the
    frame is simply run again with larger buffers."""
    import softwarerenderer_b200 as srb
    from softwarerenderer_b200 import numerics as nm
    from softwarerenderer_b200.scene import DrawItem, Mesh, Scene, default_projection, submit_scene
    from softwarerenderer_b200.shaders import Material, Shaders
    W, H = 320, 200
    rng = np.random.RandomState(3)
    n = 6000
    pos = []
    for k in range(n):                                    # one corner behind the camera, two in front: always clipped
        cx, cy = rng.uniform(-3, 3), rng.uniform(-2, 2)
        pos += [(cx - 0.2, cy - 0.2, -3.0), (cx + 0.2, cy - 0.2, -3.0), (cx, cy + 0.2, 2.0)]
    v = Shaders.make_vertices(pos, color=np.tile(rng.uniform(0.2, 1.0, (n, 1, 4)), (1, 3, 1)).reshape(-1, 4))
    v["Color"][:, 3] = 0.5                                # translucent: ordered pipeline, blending with what is there
    mesh = Mesh(v, np.arange(3 * n, dtype=np.uint16), Material())
    I = nm.identity()
    from softwarerenderer_b200 import scene as S
    first = S.scene_model("Gun", W, H)
    second = Scene(W, H, [DrawItem(mesh, I, I, default_projection(W, H), cull=0)])
    win = srb.MainWindow(W, H)
    try:
        submit_scene(win, first)
        # second frame without clears: draws over the first one's colour and depth
        from softwarerenderer_b200.rasterizer import Rasterizer
        d = second.draws[0]
        Rasterizer.RenderMesh(win, d.mesh.Vertices, d.mesh.Indices, d.model, d.view, d.proj, Shaders.StockVertexShader,
                              Shaders.StockFragmentShader(None), d.cull, d.depth, d.blend)
        win.RenderFrame()
        st = win.Stats()
        assert st.clipped_triangles == n and st.pipeline_used == 1
        gc, gd = win.ColorBuffer, win.DepthBuffer
    finally:
        win.Close()
    _, _, _, w = oracle.render_scene(first, threads=4, return_window=True)
    oc, od, ost = oracle.render_scene(second, threads=4, clear=(False, False), window=w)
    compare(gc, gd, oc, od)


def test_bands_with_a_projection_that_keeps_negative_w(oracle):
    """Band and block culling assume clip w > 0 for what they drop (for w < 0 the NDC position flips sides).  With
    a projection whose clip z is constant 1, the reference's clip test z >= NearClip * w keeps every vertex with
    w <= 10 -- also those behind the eye, w < 0 -- so triangles that straddle w = 0 are drawn with flipped corners.
    Bands must still reproduce the full frame, and the full frame the oracle."""
    from softwarerenderer_b200 import numerics as nm
    from softwarerenderer_b200.scene import DrawItem, Mesh, Scene, default_projection
    from softwarerenderer_b200.shaders import Material, Shaders
    W, H = 640, 368
    rng = np.random.RandomState(5)
    pos = []
    for k in range(600):
        c = np.array([rng.uniform(-6, 6), rng.uniform(-4, 4), rng.uniform(-7, 3)])
        pos += [tuple(c + rng.uniform(-2.5, 2.5, 3)) for _ in range(3)]
    col = np.tile(np.c_[rng.uniform(0.2, 1, (600, 3)), np.ones(600)][:, None, :], (1, 3, 1)).reshape(-1, 4)
    v = Shaders.make_vertices(pos, color=col)
    p = np.asarray(default_projection(W, H), dtype=np.float32).reshape(4, 4).copy()
    p[:, 2] = 0.0
    p[3, 2] = 1.0
    I = nm.identity()
    draws = [DrawItem(Mesh(v[i:i + 450], np.arange(450, dtype=np.uint16), Material()), I, I, p, cull=0)
             for i in range(0, 1800, 450)]
    sc = Scene(W, H, draws)
    wv = -np.array(pos)[:, 2].reshape(-1, 3)
    assert ((wv < 0).any(1) & (wv > 0).any(1)).sum() > 100          # triangles that straddle w = 0
    oc, od, _ = oracle.render_scene(sc, threads=8)
    fc, fd, _ = _render_gpu(sc, 0)
    compare(fc, fd, oc, od)
    tiles_y = (H + 15) // 16
    cuts = [0, 5, 11, 17, tiles_y]
    for r0, r1 in zip(cuts[:-1], cuts[1:]):
        bc, bd, _ = _render_gpu(sc, 0, band=(r0, r1))
        y0, y1 = r0 * 16, min(r1 * 16, H)
        assert np.array_equal(bd[y0:y1], fd[y0:y1]) and np.array_equal(bc[y0:y1], fc[y0:y1])


@pytest.mark.parametrize("seed", [1, 2, 3])
def test_band_block_culling_by_boxes_on_random_slabs(oracle, seed):
    """Band contexts drop 128-triangle blocks whose object-space bounding box lies outside the band (block_outside).
    Flat, long and rotated blocks -- the shapes a box bounds tightly and a sphere does not -- under rotated, scaled
    and translated model matrices, some of them through the near plane and behind the eye: every band must reproduce
    its rows of the full frame bit for bit, and the full frame the oracle."""
    from softwarerenderer_b200 import numerics as nm
    from softwarerenderer_b200.scene import DrawItem, Mesh, Scene, default_projection
    from softwarerenderer_b200.shaders import Material, Shaders
    W, H = 480, 272
    rng = np.random.RandomState(100 + seed)
    proj = default_projection(W, H)
    view = nm.create_look_at((0.0, 1.5, 6.0), (0.0, 0.5, 0.0), (0.0, 1.0, 0.0))
    draws = []
    for _ in range(7):
        pos, col = [], []
        for _blk in range(rng.randint(2, 5)):             # every run of 128 triangles is a slab of its own
            ext = np.array([rng.uniform(0.5, 6.0), rng.uniform(0.02, 0.4), rng.uniform(0.3, 4.0)])[rng.permutation(3)]
            c0 = rng.uniform(-5, 5, 3) * (1.0, 0.5, 1.0)
            for _t in range(128):
                a = c0 + rng.uniform(-0.5, 0.5, 3) * ext
                pos += [a, a + rng.uniform(-0.25, 0.25, 3) * ext, a + rng.uniform(-0.25, 0.25, 3) * ext]
                col += [(*rng.uniform(0.2, 1.0, 3), 1.0)] * 3
        v = Shaders.make_vertices(pos, color=col)
        model = nm.matmul(nm.create_scale(rng.uniform(0.4, 1.6)),
                          nm.create_from_yaw_pitch_roll(*rng.uniform(-3.1, 3.1, 3)),
                          nm.create_translation(rng.uniform(-3, 3), rng.uniform(-1, 2), rng.uniform(-8, 7)))
        draws.append(DrawItem(Mesh(v, np.arange(len(pos), dtype=np.uint16), Material()), model, view, proj, cull=0))
    sc = Scene(W, H, draws)
    oc, od, _ = oracle.render_scene(sc, threads=8)
    fc, fd, _ = _render_gpu(sc, 0)
    compare(fc, fd, oc, od)
    assert (fd != fd.min()).mean() > 0.05                  # something is on screen
    tiles_y = (H + 15) // 16
    cuts = [0] + sorted(rng.choice(np.arange(1, tiles_y), 4, replace=False).tolist()) + [tiles_y]
    for r0, r1 in zip(cuts[:-1], cuts[1:]):
        bc, bd, _ = _render_gpu(sc, 0, band=(r0, r1))
        y0, y1 = r0 * 16, min(r1 * 16, H)
        assert np.array_equal(bd[y0:y1], fd[y0:y1]) and np.array_equal(bc[y0:y1], fc[y0:y1])
