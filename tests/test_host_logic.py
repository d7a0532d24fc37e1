"""CPU: host-side mirror logic (API surface, scene builders, band partition, byte accounting)."""
import numpy as np
import pytest


def test_band_rows_partition_is_exact_and_balanced():
    from softwarerenderer_b200.bands import band_pixels, band_rows
    for tiles_y in (45, 68, 135, 270, 1, 7):
        for n in (1, 2, 3, 4, 8):
            rows = [band_rows(tiles_y, n, r) for r in range(n)]
            assert rows[0][0] == 0 and rows[-1][1] == tiles_y
            assert all(a[1] == b[0] for a, b in zip(rows, rows[1:]))
            sizes = [b - a for a, b in rows]
            assert max(sizes) - min(sizes) <= 1
    # 4K on 8 GPUs: 135 tile rows -> 7 bands of 17 and 1 of 16
    assert sorted(b - a for a, b in (band_rows(135, 8, r) for r in range(8))) == [16] + [17] * 7
    # 1080p: the last tile row is partial (8 px)
    assert band_pixels(1080, 2, 1) == (34 * 16, 1080)
    with pytest.raises(ValueError):
        band_rows(10, 2, 2)


def test_vertex_input_layout_is_the_reference_struct():
    from softwarerenderer_b200 import Shaders
    dt = Shaders.VertexInput                       # Shaders.cs:10-24: pos3, uv2, normal3, color4
    assert dt.itemsize == 48
    assert [dt.fields[k][1] for k in ("Position", "UV", "Normal", "Color")] == [0, 12, 20, 32]
    v = Shaders.make_vertices([(1, 2, 3)])
    assert v["Color"].tolist() == [[1, 1, 1, 1]]   # ModelLoader.cs:192-194 default colour


def test_enums_match_the_reference_values():
    from softwarerenderer_b200 import BlendMode, CullMode, DepthTest
    assert [int(x) for x in BlendMode] == [0, 1, 2, 3]          # None, Alpha, Additive, Multiply (Rasterizer.cs:25-31)
    assert [int(x) for x in CullMode] == [0, 1, 2]              # None, Back, Front (:45-50)
    assert int(DepthTest.LessEqual) == 2 and int(DepthTest.Always) == 7


def test_non_stock_shaders_fail_loudly():
    """Arbitrary delegates cannot run on the GPU and there is no CPU fallback."""
    from softwarerenderer_b200 import Rasterizer, Shaders

    class FakeWindow:
        RenderWidth, RenderHeight = 64, 64
    v = Shaders.make_vertices([(0, 0, 0), (1, 0, 0), (0, 1, 0)])
    i = np.array([0, 1, 2], np.uint16)
    I = np.eye(4, dtype=np.float32)
    with pytest.raises(NotImplementedError):
        Rasterizer.RenderMesh(FakeWindow(), v, i, I, I, I, lambda x: x, Shaders.StockFragmentShader(None))
    with pytest.raises(NotImplementedError):
        Rasterizer.RenderMesh(FakeWindow(), v, i, I, I, I, Shaders.StockVertexShader, lambda x: x)

    class Zero:
        RenderWidth, RenderHeight = 0, 64
    # RenderWidth/Height <= 0: silent return (Rasterizer.cs:176)
    assert Rasterizer.RenderMesh(Zero(), v, i, I, I, I, lambda x: x, lambda x: x) is None


def test_numerics_conventions():
    from softwarerenderer_b200 import numerics as nm
    # row vectors: translation lives in the fourth ROW
    T = nm.create_translation(1, 2, 3)
    assert (np.array([0, 0, 0, 1], np.float32) @ T).tolist() == [1, 2, 3, 1]
    P = nm.create_perspective_fov(np.pi / 2, 16 / 9, 0.1, 1000.0)
    clip = np.array([0, 0, -0.1, 1], np.float32) @ P
    assert abs(clip[2] / clip[3]) < 1e-6 and abs(clip[3] - 0.1) < 1e-7        # near plane -> NDC z 0, w = -z_view
    clip = np.array([0, 0, -1000, 1], np.float32) @ P
    assert abs(clip[2] / clip[3] - 1.0) < 1e-5                               # far plane -> NDC z 1
    L = nm.euler_to_direction((-45.0, -45.0, 0.0))                           # Renderer.cs:40, 967-972
    assert np.allclose(L, (0.5, -0.70710677, -0.5), atol=1e-6)
    V = nm.Camera((0.5, -0.4, 5.0)).GetViewMatrix()
    assert np.allclose(np.array([0.5, -0.4, 5.0, 1.0], np.float32) @ V, (0, 0, 0, 1), atol=1e-6)


def test_bundled_scene_shapes():
    from softwarerenderer_b200 import scene as S
    gun = S.load_model("Gun")
    assert len(gun.Meshes) == 5 and gun.triangle_count == 3910
    assert sum(m.Vertices.shape[0] for m in gun.Meshes) == 3120
    assert gun.Meshes[0].Texture.Width == 2048
    assert S.load_model("dust2").triangle_count == 9061 and S.load_model("gordon_freeman").triangle_count == 639
    assert all(m.Texture is None for m in S.load_model("dust2").Meshes)      # the JPEGs are missing upstream
    city = S.scene_city(3840, 2160)
    assert city.triangle_count == 1_007_140 and len(city.draws) == 1332
    st = S.scene_stress(7680, 4320, n_meshes=2)
    assert st.draws[0].mesh.Indices.size // 3 == 64_800 and st.draws[0].mesh.Vertices.shape[0] == 32_761


def test_algorithmic_bytes_formula():
    import bench
    from softwarerenderer_b200 import scene as S
    sc = S.scene_model("Gun", 1920, 1080)
    b = bench.algorithmic_bytes(sc, [True] * len(sc.draws), p_cov=10**9)
    assert b["framebuffer"] == 20 * 1920 * 1080
    assert b["geometry"] == 48 * 3120 + 2 * 3910 * 3 + 192 * 5
    assert b["texels"] == 4 * 2048 * 2048            # one unique texture bound by the five draws
    assert bench.algorithmic_bytes(sc, [True] * 5, p_cov=1000)["texels"] == 4000      # min(texture bytes, 4*P_cov)
    assert bench.algorithmic_bytes(sc, [False] * 5, p_cov=0)["geometry"] == 0


def test_balanced_cuts():
    from softwarerenderer_b200.bands import balanced_cuts
    assert balanced_cuts([1] * 8, 4) == [0, 2, 4, 6, 8]
    # all the work in the lower half: the upper bands shrink to keep every band's weight about equal
    w = [0.1] * 60 + [10.0] * 8 + [0.1] * 67
    for n in (2, 4, 8):
        c = balanced_cuts(w, n)
        assert c[0] == 0 and c[-1] == len(w) and all(b > a for a, b in zip(c, c[1:]))
        loads = [sum(w[a:b]) for a, b in zip(c, c[1:])]
        assert max(loads) <= sum(w) / n + 10.0            # within one heavy row of the ideal
    assert balanced_cuts([0, 0, 0], 2) == [0, 2, 3]       # no information -> plain row split
    assert balanced_cuts([5.0, 1.0], 2) == [0, 1, 2]
    assert len(balanced_cuts([1.0] * 135, 8)) == 9


def test_rebalance_from_periods_converges_on_a_hidden_cost():
    """bench.py --refine-cuts: the model's row weights miss part of the true cost (here: a constant per row nine times
    the model's, and a fixed cost per band); re-balancing on the bands' measured periods must bring the worst band
    down, keep the cuts valid, be a pure function of its inputs (every rank computes the same cuts) and reach a fixed
    point."""
    import numpy as np
    import pytest
    from softwarerenderer_b200.bands import balanced_cuts, rebalance_from_periods
    rng = np.random.RandomState(3)
    rows = 135
    pairs = np.r_[rng.uniform(0, 40, 66), [4000, 9000, 8000, 5000], rng.uniform(200, 900, 20), rng.uniform(0, 30, 45)]
    model = pairs + 0.85 * 240                        # what bench.py starts from
    true_row = 0.34e-3 * pairs + 2.5e-3 * 240         # the tile kernel's real cost per row (us), see DESIGN.md 6

    def periods(c):
        return [20.0 + float(true_row[a:b].sum()) for a, b in zip(c, c[1:])]

    for n in (2, 4, 8):
        cuts = balanced_cuts(model, n)
        first = max(periods(cuts))
        best = first
        for _ in range(6):
            new = rebalance_from_periods(cuts, periods(cuts), model)
            assert new == rebalance_from_periods(list(cuts), list(periods(cuts)), model.copy())
            assert new[0] == 0 and new[-1] == rows and all(b > a for a, b in zip(new, new[1:]))
            if new == cuts:
                break
            cuts = new
            best = min(best, max(periods(cuts)))
        assert best <= first + 1e-9
        if n == 8:
            assert best < 0.9 * first, (first, best)   # the sky band sheds rows: the hidden per-tile cost is found
    with pytest.raises(ValueError):
        rebalance_from_periods([0, 3, 3, 5], [1.0, 1.0, 1.0], [1.0] * 5)
    with pytest.raises(ValueError):
        rebalance_from_periods([0, 5], [1.0, 2.0], [1.0] * 5)


def test_relieve_worst_band_local_search():
    """The local search that follows the model-based rounds in bench.py: one tile row at a time off the slowest band,
    faster neighbour first, never from a single-row band; with a cost the proportional model cannot see (a fixed part
    per band) it still walks to cuts no single move improves."""
    import numpy as np
    from softwarerenderer_b200.bands import relieve_worst_band
    assert relieve_worst_band([0, 70, 135], [116.0, 93.0]) == [[0, 69, 135]]
    assert relieve_worst_band([0, 10, 20, 30], [1.0, 3.0, 2.0]) == [[0, 11, 20, 30], [0, 10, 19, 30]]
    assert relieve_worst_band([0, 10, 20, 30], [2.0, 3.0, 1.0]) == [[0, 10, 19, 30], [0, 11, 20, 30]]
    assert relieve_worst_band([0, 10, 11, 30], [1.0, 3.0, 2.0]) == []          # a tile row cannot be divided
    row = np.r_[np.full(60, 0.6), [30.0, 36.0, 32.0, 20.0], np.full(20, 2.0), np.full(51, 0.5)]

    def periods(c):
        return [45.0 + float(row[a:b].sum()) for a, b in zip(c, c[1:])]

    cuts = [0, 70, 135]
    worst = max(periods(cuts))
    for _ in range(20):
        moved = False
        for cand in relieve_worst_band(cuts, periods(cuts)):
            if max(periods(cand)) < worst:
                cuts, worst, moved = cand, max(periods(cand)), True
                break
        if not moved:
            break
    assert all(max(periods(c)) >= worst for c in ([0, cuts[1] - 1, 135], [0, cuts[1] + 1, 135]))
    assert cuts[0] == 0 and cuts[-1] == 135 and 0 < cuts[1] < 135
