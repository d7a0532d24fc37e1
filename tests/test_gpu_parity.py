"""-m gpu: CUDA path (through the C ABI) vs the oracle on identical scenes."""
import numpy as np
import pytest

from tests.util import compare

pytestmark = pytest.mark.gpu


def _render_gpu(scene, pipeline=0, band=None):
    import softwarerenderer_b200 as srb
    from softwarerenderer_b200.scene import submit_scene
    win = srb.MainWindow(scene.width, scene.height)
    try:
        win.SetPipeline(pipeline)
        if band is not None:
            win.SetBand(*band)
        submit_scene(win, scene)
        c, d = win.ColorBuffer, win.DepthBuffer
        st = win.Stats()
        return c, d, st
    finally:
        win.Close()


# numerics_fma in {0, 1}: whether Vector4.Transform / Vector*.Lerp fuse on a given .NET runtime is unknown
# (Renderer.cs:832-834, Shaders.cs:52-55), so both forms of the vertex stage, the near clipper and
# the fragment shader's Lerp are held to the oracle under the same switch.
@pytest.mark.parametrize("fma", [0, 1])
@pytest.mark.parametrize("pipeline", [1, 2])
@pytest.mark.parametrize("cam", [0, 2, 5])
def test_gun_1080p(oracle, pipeline, cam, fma):
    """BASELINE config 2: bundled Gun at 1920x1080 (partial last tile row), pixel-diff gate."""
    from softwarerenderer_b200 import scene as S
    if fma == 0 and cam == 2:
        pytest.skip("fma=0 is covered by cameras 0 and 5")
    sc = S.scene_model("Gun", 1920, 1080, view=S.gun_camera(cam))
    sc.numerics_fma = fma
    oc, od, _ = oracle.render_scene(sc, threads=8)
    gc, gd, st = _render_gpu(sc, pipeline)
    stats = compare(gc, gd, oc, od)
    assert st.pipeline_used == pipeline
    print(stats)


@pytest.mark.parametrize("fma", [0, 1])
@pytest.mark.parametrize("pipeline", [1, 2])
def test_city_small(oracle, pipeline, fma):
    """Config 3 shape at reduced size: multi-model scene, device FrustumCuller, near-plane clipping."""
    from softwarerenderer_b200 import scene as S
    sc = S.scene_city(1280, 720, n_instances=24)
    sc.numerics_fma = fma
    oc, od, ost = oracle.render_scene(sc, threads=8)
    gc, gd, st = _render_gpu(sc, pipeline)
    stats = compare(gc, gd, oc, od)
    assert st.draws_visible == sum(oracle.frustum_flags(sc))
    assert st.triangles_setup + st.triangles_no_coverage == ost.triangles_drawn
    print(stats, st.draws_visible, st.triangles_setup, st.clipped_triangles, st.tile_pairs)


@pytest.mark.parametrize("pipeline", [1, 2])
def test_stress_small(oracle, pipeline):
    """Config 5 shape at reduced size: many tiny triangles, repeat-wrapped texture."""
    from softwarerenderer_b200 import scene as S
    sc = S.scene_stress(1024, 576, n_meshes=12, grid=91)
    oc, od, _ = oracle.render_scene(sc, threads=8)
    gc, gd, st = _render_gpu(sc, pipeline)
    print(compare(gc, gd, oc, od))


@pytest.mark.parametrize("pipeline", [1, 2])
def test_stress_extensions(oracle, pipeline):
    """Bilinear + extra lights, directional and point (extensions; oracle = our restatement only)."""
    from softwarerenderer_b200 import scene as S
    sc = S.scene_stress(640, 360, n_meshes=6, grid=61, extensions=True)
    oc, od, _ = oracle.render_scene(sc, threads=8)
    plain = S.scene_stress(640, 360, n_meshes=6, grid=61)
    pc, _, _ = oracle.render_scene(plain, threads=8)
    assert np.abs(oc - pc).max() > 0.05                    # the lights do something
    gc, gd, st = _render_gpu(sc, pipeline)
    stats = compare(gc, gd, oc, od)
    assert stats["float_max_err"] < 1e-5, stats


def test_point_lights_on_near_clipped_and_wireframe_geometry(oracle):
    """The world-position varying of the point-light extension goes through Shaders.Lerp in the near clipper and
    through the DrawLine interpolation (corner 2 = corner 0) like every other varying."""
    from softwarerenderer_b200 import scene as S
    from softwarerenderer_b200.shaders import Light
    for fma in (0, 1):
        sc = S.scene_city(640, 360, n_instances=24)
        sc.numerics_fma = fma
        sc.fog_end = 600.0                                  # the default fog (1 -> 25) hides everything the light reaches
        plain, _, _ = oracle.render_scene(sc, threads=8)
        sc.extra_lights = [Light(Type="Point", Position=(182.0, 6.0, 100.0), Color=(2.0, 1.5, 1.0), AttenuationLinear=0.02,
                                 AttenuationQuadratic=0.001),
                           Light(Direction=(0.3, -0.8, 0.2), Color=(0.2, 0.2, 0.4))]
        oc, od, _ = oracle.render_scene(sc, threads=8)
        assert np.abs(oc - plain).max() > 0.2
        for pipeline in (1, 2):
            gc, gd, st = _render_gpu(sc, pipeline)
            stats = compare(gc, gd, oc, od)
            assert stats["float_max_err"] < 1e-5, stats
        # triangles through the near clipper (one corner behind the eye), lit by a point light close to them
        from softwarerenderer_b200 import numerics as nm
        from softwarerenderer_b200.scene import DrawItem, Mesh, Scene, default_projection
        from softwarerenderer_b200.shaders import Material, Shaders
        pos = [(-1, -1, -3), (1, -1, -3), (0, 1, 2), (-2, -0.5, 0.0), (2, -0.5, 0.0), (0, 0.5, -4),
               (-1, 0.2, -2), (1, 0.2, -2), (0, 1.0, -2), (-3, -2, -6), (3, -2, -6), (0, 0.3, 1.0)]
        mesh = Mesh(Shaders.make_vertices(pos), np.arange(12, dtype=np.uint16), Material())
        I = nm.identity()
        cs = Scene(320, 192, [DrawItem(mesh, I, I, default_projection(320, 192), cull=0)], fog_end=600.0)
        cs.numerics_fma = fma
        cs.extra_lights = [Light(Type="Point", Position=(0.3, 0.5, -1.0), Color=(1.5, 1.0, 0.5), AttenuationQuadratic=0.1)]
        oc, od, _ = oracle.render_scene(cs, threads=2)
        for pipeline in (1, 2):
            gc, gd, st = _render_gpu(cs, pipeline)
            stats = compare(gc, gd, oc, od)
            assert st.clipped_triangles >= 2 and stats["float_max_err"] < 1e-5, (st.clipped_triangles, stats)
    w = S.scene_model("Gun", 640, 360)
    w.debug_mode = 1
    w.extra_lights = [Light(Type="Point", Position=(0.5, 0.5, 4.0), Color=(3.0, 3.0, 3.0), AttenuationQuadratic=0.2)]
    oc, od, _ = oracle.render_scene(w, threads=4)
    gc, gd, _ = _render_gpu(w, 0)
    assert compare(gc, gd, oc, od)["float_max_err"] < 1e-5


def test_stress8k_ext_full_size(oracle):
    """BASELINE config 5 AS STATED, full size on one GPU: 10,044,000 small triangles at 7680x4320 with the bilinear
    filter and four lights (two of them point lights).  Extensions: the oracle is our restatement only."""
    import os
    from softwarerenderer_b200 import scene as S
    sc = S.scene_stress(7680, 4320, extensions=True)
    oc, od, ost = oracle.render_scene(sc, threads=os.cpu_count() or 8)
    gc, gd, st = _render_gpu(sc, 0)
    stats = compare(gc, gd, oc, od)
    assert st.triangles_setup + st.triangles_no_coverage == ost.triangles_drawn
    print(stats)


def test_bands_equal_full_frame(oracle):
    """Tile-row bands (multi-GPU partition) reproduce the full frame exactly."""
    from softwarerenderer_b200 import scene as S
    sc = S.scene_city(1280, 720, n_instances=12)
    fc, fd, _ = _render_gpu(sc, 0)
    tiles_y = (720 + 15) // 16
    cuts = [0, 11, 23, 34, tiles_y]
    for r0, r1 in zip(cuts[:-1], cuts[1:]):
        bc, bd, _ = _render_gpu(sc, 0, band=(r0, r1))
        y0, y1 = r0 * 16, min(r1 * 16, 720)
        assert np.array_equal(bc[y0:y1], fc[y0:y1])
        assert np.array_equal(bd[y0:y1], fd[y0:y1])


def _random_triangle_scene(seed, W=512, H=384, n=400):
    """Triangles of every shape straight in screen space: slivers, huge ones reaching far outside
    the target, needle-thin diagonals, tiny ones -- the cases the conservative tile rejection and
    the binning have to get exactly right."""
    from tests.util import screen_tri_scene
    rng = np.random.RandomState(seed)
    tris, cols, zs = [], [], []
    for k in range(n):
        kind = k % 5
        c = rng.uniform(-0.2, 1.2, 2) * (W, H)
        if kind == 0:                                   # big
            t = c + rng.uniform(-1.0, 1.0, (3, 2)) * (W, H)
        elif kind == 1:                                 # thin diagonal needle
            d = rng.uniform(-1, 1, 2) * (W, H)
            nrm = np.array([-d[1], d[0]]) / (np.linalg.norm(d) + 1e-9)
            t = np.stack([c, c + d, c + 0.5 * d + nrm * rng.uniform(0.05, 3.0)])
        elif kind == 2:                                 # medium
            t = c + rng.uniform(-40, 40, (3, 2))
        elif kind == 3:                                 # tiny
            t = c + rng.uniform(-2, 2, (3, 2))
        else:                                           # huge, vertices far off-screen
            t = c + rng.uniform(-30, 30, (3, 2)) * (W, H)
        tris.append(np.round(t * 16) / 16)
        cols.append((*rng.uniform(0.2, 1.0, 3), 1.0))
        zs.append(rng.uniform(0.05, 0.95))
    return screen_tri_scene(W, H, tris, colors=cols, ndc_z=zs, cull=0)


@pytest.mark.parametrize("pipeline", [1, 2])
@pytest.mark.parametrize("seed", [1, 2, 3])
def test_random_triangles_all_shapes(oracle, pipeline, seed):
    sc = _random_triangle_scene(seed)
    oc, od, _ = oracle.render_scene(sc, threads=8)
    gc, gd, st = _render_gpu(sc, pipeline)
    print(compare(gc, gd, oc, od), st.tile_pairs)


def test_city4k_full_size(oracle):
    """BASELINE config 3 at its full size: 1,007,140 triangles, 1,332 draws, 3840x2160, device FrustumCuller."""
    import os
    from softwarerenderer_b200 import scene as S
    sc = S.scene_city(3840, 2160)
    oc, od, ost = oracle.render_scene(sc, threads=os.cpu_count() or 8)
    gc, gd, st = _render_gpu(sc, 0)
    stats = compare(gc, gd, oc, od)
    assert st.triangles_setup + st.triangles_no_coverage == ost.triangles_drawn and st.draws_visible == sum(oracle.frustum_flags(sc))
    print(stats)


def test_stress8k_full_size(oracle):
    """BASELINE config 5 at its full size on ONE GPU: 10,044,000 small triangles, 7680x4320, reference semantics."""
    import os
    from softwarerenderer_b200 import scene as S
    sc = S.scene_stress(7680, 4320)
    oc, od, ost = oracle.render_scene(sc, threads=os.cpu_count() or 8)
    gc, gd, st = _render_gpu(sc, 0)
    stats = compare(gc, gd, oc, od)
    assert st.triangles_setup + st.triangles_no_coverage == ost.triangles_drawn
    print(stats)


def test_present_rgb_matches_colorbuffer():
    """srb_present_rgb[_async] delivers ColorBuffer[..., :3] (MainWindow.cs:232-246), also when two copies are
    in flight and the next frame is already rendering."""
    import softwarerenderer_b200 as srb
    from softwarerenderer_b200 import scene as S
    from softwarerenderer_b200.scene import submit_scene
    frames = [S.scene_model("Gun", 642, 361, view=S.gun_camera(k)) for k in (0, 3)]     # odd size: scalar tail path
    win = srb.MainWindow(642, 361)
    try:
        want = []
        for sc in frames:
            submit_scene(win, sc)
            want.append(win.ColorBuffer[..., :3].copy())
            assert np.array_equal(win.PresentRgb(), want[-1])
        outs = [np.zeros((361, 642, 3), np.float32) for _ in range(2)]
        for k in range(6):                              # async, slots reused three times
            submit_scene(win, frames[k & 1])
            win.PresentRgb(outs[k & 1], wait=False)
        win.PresentWait()
        assert np.array_equal(outs[0], want[0]) and np.array_equal(outs[1], want[1])
    finally:
        win.Close()


def test_present_rgba8_is_the_quantised_colorbuffer(oracle):
    """srb_present_rgba8[_rows][_async]: clamp(c,0,1)*255 rounded to nearest, per channel --
    equal to the host-side quantisation of ColorBuffer, and within the parity gate of the oracle's image."""
    import softwarerenderer_b200 as srb
    from softwarerenderer_b200 import scene as S
    from softwarerenderer_b200.scene import submit_scene
    sc = S.scene_model("Gun", 642, 361)                  # odd size: scalar tail path
    oc, _, _ = oracle.render_scene(sc, threads=4)
    win = srb.MainWindow(642, 361)
    try:
        submit_scene(win, sc)
        cb = win.ColorBuffer
        want = np.rint(np.clip(cb, 0.0, 1.0) * np.float32(255.0)).astype(np.uint8)
        got = win.PresentRgba8()
        assert got.dtype == np.uint8 and np.array_equal(got, want)
        assert np.array_equal(win.ColorBufferRgba8(), want)
        o8 = np.rint(np.clip(oc, 0.0, 1.0) * np.float32(255.0)).astype(np.int32)
        assert (np.abs(got.astype(np.int32) - o8).max(axis=-1) <= 1).mean() >= 0.999
        rows = np.zeros_like(want)
        win.PresentRgba8(rows, wait=False, rows=(0, 100))
        win.PresentRgba8(rows, wait=True, rows=(100, 361))
        assert np.array_equal(rows, want)
        outs = [np.zeros_like(want) for _ in range(2)]
        for k in range(4):
            win.ReplayFrame()
            win.PresentRgba8(outs[k & 1], wait=False)
        win.PresentWait()
        assert np.array_equal(outs[0], want) and np.array_equal(outs[1], want)
    finally:
        win.Close()


def test_pipelined_frames_do_not_disturb_each_other():
    """Frames are pipelined (geometry of frame N+1 beside the tile kernel of frame N, two buffer sets): a burst
    of frames without any synchronisation in between -- replays of one scene, then two different scenes
    alternating -- must leave exactly the image of the last frame."""
    import softwarerenderer_b200 as srb
    from softwarerenderer_b200 import scene as S
    from softwarerenderer_b200.scene import submit_scene
    a = S.scene_city(1280, 720, n_instances=12)
    b = S.scene_model("Gun", 1280, 720, view=S.gun_camera(4))
    win = srb.MainWindow(1280, 720)
    try:
        submit_scene(win, a)
        want_a = (win.ColorBuffer.copy(), win.DepthBuffer.copy())
        submit_scene(win, b)
        want_b = (win.ColorBuffer.copy(), win.DepthBuffer.copy())
        submit_scene(win, a)
        for _ in range(25):
            win.ReplayFrame()
        assert np.array_equal(win.ColorBuffer, want_a[0]) and np.array_equal(win.DepthBuffer, want_a[1])
        for k in range(9):                                  # a b a b ... a b a, no read-back in between
            submit_scene(win, a if k % 2 == 0 else b)
        assert np.array_equal(win.ColorBuffer, want_a[0]) and np.array_equal(win.DepthBuffer, want_a[1])
        submit_scene(win, b)
        for _ in range(7):
            win.ReplayFrame()
        assert np.array_equal(win.ColorBuffer, want_b[0]) and np.array_equal(win.DepthBuffer, want_b[1])
    finally:
        win.Close()


@pytest.mark.parametrize("depth", [0, 2, 3, 4])
def test_pipeline_depth_leaves_the_image_alone(oracle, depth):
    """2..4 frames in flight (srb_set_pipeline_depth: one buffer set and one geometry stream per frame in flight):
    bursts of replays and of alternating scenes without host synchronisation leave exactly the last frame."""
    import softwarerenderer_b200 as srb
    from softwarerenderer_b200 import scene as S
    from softwarerenderer_b200.scene import submit_scene
    a = S.scene_city(1280, 720, n_instances=12)
    b = S.scene_model("Gun", 1280, 720, view=S.gun_camera(4))
    oa, ob = oracle.render_scene(a, threads=8), oracle.render_scene(b, threads=8)
    win = srb.MainWindow(1280, 720)
    try:
        win.SetPipelineDepth(depth)
        submit_scene(win, a)
        for _ in range(11):
            win.ReplayFrame()
        compare(win.ColorBuffer, win.DepthBuffer, oa[0], oa[1])
        for k in range(9):                                      # a b a b ... a, no read-back in between
            submit_scene(win, a if k % 2 == 0 else b)
        compare(win.ColorBuffer, win.DepthBuffer, oa[0], oa[1])
        submit_scene(win, b)
        for _ in range(7):
            win.ReplayFrame()
        compare(win.ColorBuffer, win.DepthBuffer, ob[0], ob[1])
        win.SetBand(10, 30)                                     # a band context, deeper pipeline, split bins
        win.SetBinSplit(64, 64, 32)
        submit_scene(win, a)
        for _ in range(9):
            win.ReplayFrame()
        assert np.array_equal(win.DepthBuffer[160:480], oa[1][160:480])
    finally:
        win.Close()


def test_graph_replay_of_the_geometry_chain(oracle):
    """Frames that repeat the shape of the frame before them on the same buffer set replay their chain of geometry
    launches as a CUDA graph (srb_set_graphs, on by default): replays, re-submissions of the same scene, a change of
    uniforms, of the scene and of the band in between, graphs off -- always the oracle's image."""
    import softwarerenderer_b200 as srb
    from softwarerenderer_b200 import scene as S
    from softwarerenderer_b200.scene import submit_scene
    a = S.scene_city(1280, 720, n_instances=12)
    b = S.scene_model("Gun", 1280, 720, view=S.gun_camera(2))
    a2 = S.scene_city(1280, 720, n_instances=12)
    a2.fog_end = 90.0                                         # same draws, other uniforms
    want = {id(s): oracle.render_scene(s, threads=8) for s in (a, b, a2)}
    win = srb.MainWindow(1280, 720)
    try:
        for graphs in (True, False, True):
            win.SetGraphs(graphs)
            for sc in (a, a, a, b, b, a2, a2, a2, a):
                submit_scene(win, sc)                         # per set: first sight, capture, replay
                for _ in range(9):
                    win.ReplayFrame()
                compare(win.ColorBuffer, win.DepthBuffer, want[id(sc)][0], want[id(sc)][1])
        launches = win.Stats().kernel_launches
        win.SetGraphs(False)
        win.ReplayFrame()
        assert win.Stats().kernel_launches == launches        # the statistic counts kernels, replayed or launched
        win.SetGraphs(True)
        win.SetBand(10, 30)
        submit_scene(win, a)
        for _ in range(12):
            win.ReplayFrame()
        assert np.array_equal(win.DepthBuffer[160:480], want[id(a)][1][160:480])
    finally:
        win.Close()


def test_mesh_cache_identity_content_and_invalidate(oracle):
    """The mirror's mesh cache (window.mesh_handle): equal copies share one resident mesh, an array edited in place
    keeps rendering the old upload until InvalidateMesh (documented), after which it matches the oracle again, and
    the least recently used meshes are destroyed beyond MeshCacheCapacity without changing any result."""
    import softwarerenderer_b200 as srb
    from softwarerenderer_b200 import scene as S
    from softwarerenderer_b200.scene import submit_scene
    sc = S.scene_stress(320, 192, n_meshes=3, grid=21)
    win = srb.MainWindow(sc.width, sc.height)
    try:
        submit_scene(win, sc)
        c0 = win.ColorBuffer.copy()
        n0 = len(win._mesh_by_content)
        # equal copies of every array: new identities, same content -> no new device mesh
        for d in sc.draws:
            d.mesh.Vertices = d.mesh.Vertices.copy()
            d.mesh.Indices = d.mesh.Indices.copy()
        submit_scene(win, sc)
        assert len(win._mesh_by_content) == n0
        assert np.array_equal(win.ColorBuffer, c0)
        # edit in place: stale until invalidated
        m = sc.draws[0].mesh
        m.Vertices["Color"][:, :3] *= np.float32(0.25)       # darken the vertex colours of one mesh
        submit_scene(win, sc)
        assert np.array_equal(win.ColorBuffer, c0)
        win.InvalidateMesh(m.Vertices, m.Indices)
        submit_scene(win, sc)
        oc, od, _ = oracle.render_scene(sc, threads=4)
        compare(win.ColorBuffer, win.DepthBuffer, oc, od)
        assert not np.array_equal(win.ColorBuffer, c0)
        # eviction: beyond the capacity the least recently used meshes are destroyed between frames, and a frame
        # that needs them again uploads them again
        assert len(win._mesh_by_content) == n0 + 1               # the edited mesh and its stale predecessor
        win.MeshCacheCapacity = 1
        submit_scene(win, sc)
        assert len(win._mesh_by_content) == 1
        submit_scene(win, sc)
        assert len(win._mesh_by_content) == 1
        compare(win.ColorBuffer, win.DepthBuffer, oc, od)
    finally:
        win.Close()
