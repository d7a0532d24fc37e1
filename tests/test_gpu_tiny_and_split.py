"""-m gpu: crowds of tiny triangles and split bins.  Frames of mostly 1-4 pixel triangles (many on tile corners,
exact duplicates at equal depth) under every depth function the visibility pipeline takes, and the same frames with
bin splitting forced on: a bin of many entries is rasterized by several CTAs whose per-pixel winner keys are merged
with 64-bit atomic max, the last CTA resolving the tile.  Same gate as everywhere: coverage and depth winner
bit-exact against the oracle (Rasterizer.cs:432-535), RGBA8 within +-1."""
import numpy as np
import pytest

from tests.test_gpu_parity import _render_gpu
from tests.util import compare, screen_tri_scene

pytestmark = pytest.mark.gpu


def _render_split(scene, split, band=None, pipeline=2):
    """_render_gpu with bin splitting forced on: bins of >= split[0] entries, parts of split[1] entries."""
    import softwarerenderer_b200 as srb
    from softwarerenderer_b200.scene import submit_scene
    win = srb.MainWindow(scene.width, scene.height)
    try:
        win.SetPipeline(pipeline)
        win.SetBinSplit(split[0], split[0], split[1])
        if band is not None:
            win.SetBand(*band)
        submit_scene(win, scene)
        c, d = win.ColorBuffer, win.DepthBuffer
        return c, d, win.Stats()
    finally:
        win.Close()


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
def test_tiny_triangles_every_visibility_mode(oracle, depth, straddle):
    """Less / LessEqual / Greater / GreaterEqual (the last two never beat the cleared depth, float.MinValue:
    an empty frame), ties in depth decided by submission order, with and without split bins."""
    sc = _tiny_scene(40 + depth, depth, straddle=straddle)
    oc, od, ost = oracle.render_scene(sc, threads=4)
    gc, gd, st = _render_gpu(sc, 2)
    stats = compare(gc, gd, oc, od)
    assert st.pipeline_used == 2 and st.split_parts == 0
    assert st.triangles_setup + st.triangles_no_coverage == ost.triangles_drawn
    assert stats["float_max_err"] < 1e-5, stats
    s_c, s_d, _ = _render_split(sc, (64, 128))
    assert np.array_equal(s_d, gd) and np.array_equal(s_c, gc)
    # the ordered pipeline on the same frame: same winners
    gc1, gd1, st1 = _render_gpu(sc, 1)
    assert np.array_equal(gd1, gd)


@pytest.mark.parametrize("part", [32, 128, 256])
@pytest.mark.parametrize("depth", [1, 2])
def test_split_bins_merge_to_the_same_winners(oracle, depth, part):
    """A crowd: 6,000 tiny triangles (a third of them exact duplicates in depth) over 12 tiles plus large ones across
    them -- bins of 300-700 entries, split into parts of 128 / 256 entries."""
    rng = np.random.RandomState(7 * depth + part)
    W, H = 192, 96
    tris, cols, zs = [], [], []
    for k in range(6000):
        c = np.array([rng.uniform(40, 104), rng.uniform(30, 78)])
        t = c + (rng.uniform(-70, 70, (3, 2)) if k % 200 == 0 else rng.uniform(-2.4, 2.4, (3, 2)))
        tris.append(np.round(t * 16) / 16)
        cols.append((*rng.uniform(0.2, 1.0, 3), 1.0))
        zs.append(float(rng.choice([0.25, 0.5, rng.uniform(0.05, 0.95)])))
    sc = screen_tri_scene(W, H, tris, colors=cols, ndc_z=zs, cull=0, depth=depth, blend=1)
    oc, od, ost = oracle.render_scene(sc, threads=4)
    gc, gd, st = _render_split(sc, (128, part))
    stats = compare(gc, gd, oc, od)
    assert st.split_parts >= 8, st.split_parts
    assert stats["float_max_err"] < 1e-5, stats
    # every rank of a band partition splits its own bins
    for r0, r1 in ((0, 3), (3, 6)):
        bc, bd, bst = _render_split(sc, (128, part), band=(r0, r1))
        assert np.array_equal(bd[r0 * 16:r1 * 16], od[r0 * 16:r1 * 16]) and np.array_equal(bc[r0 * 16:r1 * 16], gc[r0 * 16:r1 * 16])
    # replays reuse the key slices: they must come back empty every time
    import softwarerenderer_b200 as srb
    from softwarerenderer_b200.scene import submit_scene
    win = srb.MainWindow(W, H)
    try:
        win.SetBinSplit(128, 128, part)
        submit_scene(win, sc)
        for _ in range(6):
            win.ReplayFrame()
        assert np.array_equal(win.DepthBuffer, od) and np.array_equal(win.ColorBuffer, gc)
    finally:
        win.Close()


def test_tiny_triangles_on_odd_sized_targets(oracle):
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


def test_key_slices_do_not_leak_between_frames_or_pipelines(oracle):
    """The key slices of split bins are handed back empty by the CTA that resolves the tile: tiny-triangle frames, a
    frame of the ordered pipeline, an empty frame and replays in any order leave exactly the last frame's image."""
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
        win.SetBinSplit(64, 64, 128)
        for name, sc in (("a", a), ("b", b), ("a", a), ("wire", wire), ("b", b), ("empty", empty), ("a", a), ("b", b)):
            submit_scene(win, sc)
            if name in ("a", "b"):
                for _ in range(3):
                    win.ReplayFrame()
            compare(win.ColorBuffer, win.DepthBuffer, *want[name])
    finally:
        win.Close()


def test_tiny_triangles_in_bands(oracle):
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


def test_extra_ctas_follow_the_frame(oracle):
    """The tile kernel is launched with as many extra CTAs as the last frame's split bins wanted.  A context's first
    split frame has 256 of them whatever it needs (the scan kernel then makes the parts larger), later frames get what
    the one before asked for, and a frame that needs fewer again shrinks the grid: same image every time."""
    import softwarerenderer_b200 as srb
    from softwarerenderer_b200.scene import submit_scene
    rng = np.random.RandomState(11)
    W, H = 160, 96

    def crowd(n):
        tris, cols, zs = [], [], []
        for k in range(n):
            c = np.array([rng.uniform(8, W - 8), rng.uniform(8, H - 8)])
            t = c + (rng.uniform(-60, 60, (3, 2)) if k % 400 == 0 else rng.uniform(-2.2, 2.2, (3, 2)))
            tris.append(np.round(t * 16) / 16)
            cols.append((*rng.uniform(0.2, 1.0, 3), 1.0))
            zs.append(float(rng.choice([0.25, 0.5, rng.uniform(0.05, 0.95)])))
        return screen_tri_scene(W, H, tris, colors=cols, ndc_z=zs, cull=0, depth=2, blend=1)

    heavy, light = crowd(20000), crowd(2400)             # ~400 / ~50 entries per tile over 60 tiles (one mesh: ushort indices)
    oh = oracle.render_scene(heavy, threads=8)
    ol = oracle.render_scene(light, threads=4)
    win = srb.MainWindow(W, H)
    try:
        win.SetPipeline(2)
        win.SetBinSplit(64, 64, 32)                       # heavy: ~60 x 12 parts wanted, far beyond the first 256
        parts = []
        for sc, (oc, od, _) in ((heavy, oh), (heavy, oh), (heavy, oh), (light, ol), (light, ol), (heavy, oh)):
            submit_scene(win, sc)
            compare(win.ColorBuffer, win.DepthBuffer, oc, od)
            parts.append(win.Stats().split_parts)
        assert parts[0] <= 256 < parts[1] == parts[2], parts
        assert parts[4] <= parts[3] < parts[1], parts
        assert parts[5] <= 256, parts
    finally:
        win.Close()


def test_band_context_follows_sparse_and_dense_frames(oracle):
    """One band context (with the RGBA8 gather frame, i.e. the tile kernel's OUT8 instance) through frames that leave
    most of its tiles empty and frames that fill all of them, back and forth: the grids sized from the frame before
    never change the image -- float colour and the 8-bit gather frame against the oracle's band every time."""
    import softwarerenderer_b200 as srb
    from softwarerenderer_b200.scene import submit_scene
    rng = np.random.RandomState(5)
    W, H = 512, 256                                       # 32 x 16 tiles (power-of-two target: exact screen coordinates)
    band = (2, 14)

    def crowd(n, x1, y1):
        tris, cols, zs = [], [], []
        for k in range(n):
            c = np.array([rng.uniform(4, x1), rng.uniform(4, y1)])
            t = c + (rng.uniform(-40, 40, (3, 2)) if k % 100 == 0 else rng.uniform(-3.0, 3.0, (3, 2)))
            tris.append(np.round(t * 16) / 16)
            cols.append((*rng.uniform(0.2, 1.0, 3), 1.0))
            zs.append(float(rng.uniform(0.05, 0.95)))
        return screen_tri_scene(W, H, tris, colors=cols, ndc_z=zs, cull=0, depth=2, blend=1)

    sparse, dense = crowd(600, 120, 110), crowd(6000, W - 4, H - 4)
    refs = {id(sparse): oracle.render_scene(sparse, threads=4), id(dense): oracle.render_scene(dense, threads=4)}
    rows = slice(band[0] * 16, band[1] * 16)
    win = srb.MainWindow(W, H)
    try:
        win.SetPipeline(2)
        win.SetBand(*band)
        win.ExportGatherRgba8()                            # the tile kernel's OUT8 instance, as on a band rank
        for sc in (sparse, sparse, sparse, dense, dense, sparse, sparse):
            submit_scene(win, sc)
            oc, od, _ = refs[id(sc)]
            gc, gd = win.ColorBuffer, win.DepthBuffer
            compare(gc[rows], gd[rows], oc[rows], od[rows])
            want8 = np.rint(np.clip(gc[rows], 0.0, 1.0) * np.float32(255.0)).astype(np.uint8)
            assert np.array_equal(win.GatheredRgba8()[rows], want8)
    finally:
        win.Close()
