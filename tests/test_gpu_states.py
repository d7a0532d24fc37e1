"""-m gpu: the rest of the Rasterizer surface against the oracle:
every DepthTest x BlendMode x CullMode (Rasterizer.cs:25-50, 58-65, 415-417, 542-559) on
translucent, overlapping geometry, frames that mix states per draw, frames that keep the
previous framebuffer, and DebugMode.Wireframe (Rasterizer.cs:419-425, 232-340)."""
import numpy as np
import pytest

from tests.test_gpu_parity import _random_triangle_scene, _render_gpu
from tests.util import compare, screen_tri_scene

pytestmark = pytest.mark.gpu


def _translucent_scene(seed, depth, blend, cull, W=256, H=192, n=160, separate=False):
    """Overlapping triangles with alpha in [-0.2, 1]: zero and negative alphas exercise the alpha <= 0
    discard (:511) and the BlendMode.None row break (:520-523); repeated depths exercise Equal/NotEqual."""
    rng = np.random.RandomState(seed)
    tris, cols, zs = [], [], []
    for k in range(n):
        c = rng.uniform(0.0, 1.0, 2) * (W, H)
        size = (6, 40, 160)[k % 3]
        t = c + rng.uniform(-size, size, (3, 2))
        tris.append(np.round(t * 8) / 8)
        a = rng.choice([1.0, 1.0, 0.5, 0.25, 0.0, -0.2, rng.uniform(0.05, 0.95)])
        per_vertex = rng.rand() < 0.3
        if per_vertex:
            cols.append(np.concatenate([rng.uniform(0.1, 1.0, (3, 3)), rng.uniform(-0.1, 1.0, (3, 1))], axis=1))
        else:
            cols.append((*rng.uniform(0.1, 1.0, 3), a))
        zs.append(float(rng.choice([0.125, 0.25, 0.5, 0.75, rng.uniform(0.05, 0.95)])))
    return screen_tri_scene(W, H, tris, colors=cols, ndc_z=zs, cull=cull, depth=depth, blend=blend,
                            separate_draws=separate)


@pytest.mark.parametrize("blend", [0, 1, 2, 3])
@pytest.mark.parametrize("depth", [0, 1, 2, 3, 4, 5, 6, 7])
def test_every_depth_function_and_blend_mode(oracle, depth, blend):
    for cull in (0, 1, 2):
        # a LessEqual/Alpha layer first, so that Equal/NotEqual/Multiply/... meet a non-trivial framebuffer
        sc = _translucent_scene(100 * depth + 10 * blend + cull, 2, 1, cull, n=80)
        sc.draws += _translucent_scene(100 * depth + 10 * blend + cull + 5000, depth, blend, cull, n=80).draws
        oc, od, _ = oracle.render_scene(sc, threads=4)
        gc, gd, st = _render_gpu(sc, 0)
        stats = compare(gc, gd, oc, od)
        assert st.pipeline_used == 1, "translucent frames must take the submission-ordered path"
        assert stats["float_max_err"] < 1e-5, stats


def test_states_mixed_per_draw(oracle):
    """One draw per triangle, every draw with its own (cull, depth, blend)."""
    sc = _translucent_scene(7, 2, 1, 0, n=120, separate=True)
    rng = np.random.RandomState(11)
    for d in sc.draws:
        d.cull, d.depth, d.blend = int(rng.randint(0, 3)), int(rng.randint(0, 8)), int(rng.randint(0, 4))
    oc, od, _ = oracle.render_scene(sc, threads=4)
    gc, gd, st = _render_gpu(sc, 0)
    stats = compare(gc, gd, oc, od)
    assert st.pipeline_used == 1 and stats["float_max_err"] < 1e-5, stats


def test_frame_without_clear_continues_from_the_framebuffer(oracle):
    """Renderer draws in several passes between clears; a frame that does not clear must blend and
    depth-test against what is already there (MainWindow.cs:378-436 accessors)."""
    import softwarerenderer_b200 as srb
    from softwarerenderer_b200.rasterizer import Rasterizer
    from softwarerenderer_b200.scene import submit_scene
    from softwarerenderer_b200.shaders import Shaders
    a = _translucent_scene(21, 2, 1, 0, n=60)
    b = _translucent_scene(22, 2, 1, 0, n=60)
    oc, od, _, w = oracle.render_scene(a, threads=2, return_window=True)
    oc, od, _ = oracle.render_scene(b, threads=2, clear=(False, False), window=w)
    win = srb.MainWindow(a.width, a.height)
    try:
        submit_scene(win, a)
        for i, d in enumerate(b.draws):                 # second pass: no ClearColorBuffer / ClearDepthBuffer
            Rasterizer.RenderMesh(win, d.mesh.Vertices, d.mesh.Indices, d.model, d.view, d.proj,
                                  Shaders.StockVertexShader, Shaders.StockFragmentShader(d.mesh.Texture),
                                  d.cull, d.depth, d.blend, submissionIndex=i)
        win.RenderFrame(near_clip=b.near_clip)
        gc, gd = win.ColorBuffer, win.DepthBuffer
    finally:
        win.Close()
    stats = compare(gc, gd, oc, od)
    assert stats["float_max_err"] < 1e-5, stats


# ---------------------------------------------------------------------------
# DebugMode.Wireframe
# ---------------------------------------------------------------------------
@pytest.mark.parametrize("seed", [1, 2])
def test_wireframe_random_triangles(oracle, seed):
    sc = _random_triangle_scene(seed, n=200)
    sc.debug_mode = 1
    oc, od, _ = oracle.render_scene(sc, threads=8)
    gc, gd, st = _render_gpu(sc, 0)
    stats = compare(gc, gd, oc, od)
    assert st.pipeline_used == 1
    assert 0.01 < stats["coverage"] < 0.9
    print(stats)


@pytest.mark.parametrize("depth,blend", [(2, 1), (7, 0), (1, 2), (6, 3), (0, 1)])
def test_wireframe_states(oracle, depth, blend):
    sc = _translucent_scene(5 + depth, depth, blend, 0, n=80)
    sc.debug_mode = 1
    oc, od, _ = oracle.render_scene(sc, threads=4)
    gc, gd, st = _render_gpu(sc, 0)
    stats = compare(gc, gd, oc, od)
    assert stats["float_max_err"] < 1e-5, stats


def test_wireframe_models(oracle):
    """The bundled Gun at 1080p and the small city (near-clipped triangles included) as wireframes."""
    from softwarerenderer_b200 import scene as S
    for sc in (S.scene_model("Gun", 1920, 1080), S.scene_city(1280, 720, n_instances=12)):
        sc.debug_mode = 1
        oc, od, _ = oracle.render_scene(sc, threads=8)
        gc, gd, st = _render_gpu(sc, 0)
        print(compare(gc, gd, oc, od), st.tile_pairs)


def test_wireframe_is_refused_by_the_forced_visibility_pipeline():
    import softwarerenderer_b200 as srb
    from softwarerenderer_b200 import scene as S
    sc = S.scene_model("Gun", 320, 180)
    sc.debug_mode = 1
    win = srb.MainWindow(sc.width, sc.height)
    try:
        win.SetPipeline(2)
        with pytest.raises(Exception):
            S.submit_scene(win, sc)
    finally:
        win.Close()


def test_racing_host_threads_with_submission_indices_give_the_ordered_frame(oracle):
    """Renderer calls RenderMesh from Parallel.ForEach threads (Renderer.cs:444, 479): srb_draw is thread-safe and
    the explicit submission index, not the arrival order, decides the order of blending."""
    import ctypes as C
    import threading

    import softwarerenderer_b200 as srb
    from softwarerenderer_b200 import _native as N
    from softwarerenderer_b200.scene import submit_scene
    sc = _translucent_scene(31, 2, 1, 0, n=96, separate=True)
    oc, od, _ = oracle.render_scene(sc, threads=4)
    win = srb.MainWindow(sc.width, sc.height)
    try:
        submit_scene(win, sc)                        # uploads meshes, sets uniforms; reference image of the ordered path
        want_c, want_d = win.ColorBuffer, win.DepthBuffer
        compare(want_c, want_d, oc, od)
        L, ctx = N.lib(), win._ctx
        handles = [win.mesh_handle(d.mesh.Vertices, d.mesh.Indices) for d in sc.draws]
        mats = [[np.ascontiguousarray(np.asarray(m, np.float32).reshape(16)) for m in (d.model, d.view, d.proj)] for d in sc.draws]
        cc = (C.c_float * 4)(*[float(x) for x in sc.clear_color])
        for trial in range(3):
            N.check(ctx, L.srb_begin_frame(ctx, cc, N.SRB_CLEAR_COLOR | N.SRB_CLEAR_DEPTH))
            order = np.random.RandomState(trial).permutation(len(sc.draws))
            errs = []

            def worker(chunk):
                for i in chunk:
                    d = sc.draws[i]
                    rc = L.srb_draw(ctx, handles[i], N.ptr(mats[i][0]), N.ptr(mats[i][1]), N.ptr(mats[i][2]), 0,
                                    int(d.cull), int(d.depth), int(d.blend), int(i))
                    if rc != 0:
                        errs.append(rc)

            ts = [threading.Thread(target=worker, args=(order[k::8],)) for k in range(8)]
            [t.start() for t in ts]
            [t.join() for t in ts]
            assert not errs
            N.check(ctx, L.srb_end_frame(ctx))
            assert np.array_equal(win.ColorBuffer, want_c) and np.array_equal(win.DepthBuffer, want_d)
    finally:
        win.Close()
