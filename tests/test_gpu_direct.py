"""-m gpu: the direct path of the visibility pipeline -- unclipped triangles of at most 4x4 pixels are rasterized
by the setup kernel straight into the per-pixel winner keys (no setup record, no bin entry) and rebuilt from the
mesh when they win.  Same gate as everywhere: coverage and depth winner bit-exact against the oracle
(Rasterizer.cs:432-535), RGBA8 within +-1."""
import numpy as np
import pytest

from tests.test_gpu_parity import _render_gpu
from tests.util import compare, screen_tri_scene

pytestmark = pytest.mark.gpu


def _tiny_scene(seed, depth, W=256, H=160, n=900, straddle=False):
    """Mostly 1-4 pixel triangles (many on tile corners when `straddle`), a few large ones behind and in front of
    them, and exact duplicates at equal depth so that the tie rule (later / earlier submission wins) decides."""
    rng = np.random.RandomState(seed)
    tris, cols, zs = [], [], []
    for k in range(n):
        if straddle:
            c = np.array([16.0 * rng.randint(1, W // 16), 16.0 * rng.randint(1, H // 16)]) + rng.uniform(-1.5, 1.5, 2)
        else:
            c = rng.uniform(0.0, 1.0, 2) * (W, H)
        kind = k % 10
        if kind == 0:
            t = c + rng.uniform(-60, 60, (3, 2))                    # large: goes through records and bins
        elif kind == 1:
            t = c + rng.uniform(-6, 6, (3, 2))                      # medium
        else:
            t = c + rng.uniform(-1.9, 1.9, (3, 2))                  # tiny: the direct path
        tris.append(np.round(t * 16) / 16)
        cols.append((*rng.uniform(0.2, 1.0, 3), 1.0))
        zs.append(float(rng.choice([0.25, 0.5, rng.uniform(0.05, 0.95)])))
        if kind in (2, 3):                                           # exact duplicate, other colour: a tie in depth
            tris.append(tris[-1].copy())
            cols.append((*rng.uniform(0.2, 1.0, 3), 1.0))
            zs.append(zs[-1])
    return screen_tri_scene(W, H, tris, colors=cols, ndc_z=zs, cull=0, depth=depth, blend=1)


@pytest.mark.parametrize("straddle", [False, True])
@pytest.mark.parametrize("depth", [1, 2, 3, 4])
def test_direct_path_every_visibility_mode(oracle, depth, straddle):
    sc = _tiny_scene(40 + depth, depth, straddle=straddle)
    oc, od, ost = oracle.render_scene(sc, threads=4)
    gc, gd, st = _render_gpu(sc, 2)
    stats = compare(gc, gd, oc, od)
    assert st.pipeline_used == 2
    if depth in (1, 2):          # Greater / GreaterEqual never beat the cleared depth (float.MinValue): nothing to draw
        assert st.triangles_direct > 300, st.triangles_direct
    assert st.triangles_setup + st.triangles_no_coverage == ost.triangles_drawn
    assert stats["float_max_err"] < 1e-5, stats
    # the ordered pipeline on the same frame: no direct path there, same image
    gc1, gd1, st1 = _render_gpu(sc, 1)
    assert st1.triangles_direct == 0
    assert np.array_equal(gd1, gd)


def test_direct_path_on_odd_sized_targets(oracle):
    """Partial last tile row / column and targets smaller than a tile."""
    for (W, H) in ((250, 131), (33, 17), (7, 5)):
        rng = np.random.RandomState(W)
        tris = [np.round((rng.uniform(0, 1, 2) * (W, H) + rng.uniform(-1.9, 1.9, (3, 2))) * 16) / 16 for _ in range(400)]
        cols = [(*rng.uniform(0.2, 1.0, 3), 1.0) for _ in tris]
        zs = [float(rng.uniform(0.05, 0.95)) for _ in tris]
        sc = screen_tri_scene(W, H, tris, colors=cols, ndc_z=zs, cull=0)
        oc, od, _ = oracle.render_scene(sc, threads=2)
        gc, gd, st = _render_gpu(sc, 2)
        compare(gc, gd, oc, od)
        assert st.triangles_direct > 0


def test_direct_keys_do_not_leak_between_frames_or_pipelines(oracle):
    """The global key buffer is handed back empty by every tile that used it: tiny-triangle frames, a frame of the
    ordered pipeline, an empty frame and replays in any order leave exactly the last frame's image."""
    import softwarerenderer_b200 as srb
    from softwarerenderer_b200.scene import submit_scene
    a, b = _tiny_scene(1, 2), _tiny_scene(2, 2, straddle=True)
    wire = _tiny_scene(3, 2, n=120)
    wire.debug_mode = 1                                   # ordered pipeline
    from softwarerenderer_b200.scene import Scene
    empty = Scene(256, 160, [], clear_color=(0.0, 0.0, 0.0, 0.0), fog_start=10.0, fog_end=20.0)
    want = {}
    for name, sc in (("a", a), ("b", b), ("wire", wire), ("empty", empty)):
        oc, od, _ = oracle.render_scene(sc, threads=4)
        want[name] = (oc, od)
    win = srb.MainWindow(256, 160)
    try:
        for name, sc in (("a", a), ("b", b), ("a", a), ("wire", wire), ("b", b), ("empty", empty), ("a", a), ("b", b)):
            submit_scene(win, sc)
            if name in ("a", "b"):
                for _ in range(3):
                    win.ReplayFrame()
            compare(win.ColorBuffer, win.DepthBuffer, *want[name])
    finally:
        win.Close()


def test_direct_path_in_bands(oracle):
    """Tile-row bands (multi-GPU partition): a tiny triangle straddling a band cut is rasterized by both bands,
    each inside its own rows only."""
    sc = _tiny_scene(9, 2, straddle=True)
    oc, od, _ = oracle.render_scene(sc, threads=4)
    tiles_y = (sc.height + 15) // 16
    for cuts in ([0, 3, 4, 7, tiles_y], [0, 1, tiles_y]):
        for r0, r1 in zip(cuts[:-1], cuts[1:]):
            bc, bd, _ = _render_gpu(sc, 2, band=(r0, r1))
            y0, y1 = r0 * 16, min(r1 * 16, sc.height)
            assert np.array_equal(bd[y0:y1], od[y0:y1])
            compare(bc[y0:y1], bd[y0:y1], oc[y0:y1], od[y0:y1])


def test_stress_scene_goes_direct(oracle):
    """Config 5 shape: nearly every triangle is 1-3 px; the bins stay (almost) empty."""
    from softwarerenderer_b200 import scene as S
    sc = S.scene_stress(1024, 576, n_meshes=12, grid=91)
    oc, od, ost = oracle.render_scene(sc, threads=8)
    gc, gd, st = _render_gpu(sc, 2)
    compare(gc, gd, oc, od)
    assert st.triangles_direct > 0.8 * st.triangles_setup
    assert st.triangles_setup + st.triangles_no_coverage == ost.triangles_drawn
