"""Known-answer tests that pin the CPU oracle (CPU only, no GPU).

The reference ships no tests, golden images or fixtures, and its
C# cannot run in this image, so upstream parity is UNPINNED.  These cases are
hand-derived from the reference source lines they cite (paths under
/root/reference); each expected value is computed in the test independently of
the oracle (closed form / float64), never by calling the oracle twice.
"""
import math

import numpy as np
import pytest

from tests.util import CLEAR_DEPTH, screen_tri_scene

F = np.float32


def _render(oracle, sc, threads=0):
    return oracle.render_scene(sc, threads=threads)


# --------------------------------------------------------------------------
# A.5: winding sign and the negated stored depth (Rasterizer.cs:365-398, 411-417, 498-502)
# --------------------------------------------------------------------------
@pytest.mark.parametrize("z,expected", [(-2.0, -0.9750475), (-5.0, -0.99004906)])
def test_ccw_triangle_is_front_facing_and_depth_is_negated(oracle, z, expected):
    from softwarerenderer_b200 import numerics as nm
    from softwarerenderer_b200.scene import DrawItem, Mesh, Scene, default_projection
    from softwarerenderer_b200.shaders import Material, Shaders
    W, H = 1280, 720
    # CCW seen from the camera at the origin looking down -Z
    v = Shaders.make_vertices([(-1, -1, z), (1, -1, z), (0, 1, z)])
    mesh = Mesh(v, np.array([0, 1, 2], np.uint16), Material())
    I = nm.identity()
    sc = Scene(W, H, [DrawItem(mesh, I, I, default_projection(W, H), cull=1)])
    c, d, st = _render(oracle, sc)
    assert st.triangles_drawn == 1            # front-facing under CullMode.Back <=> area < 0
    covered = d != CLEAR_DEPTH
    assert covered.sum() > 1000
    # every covered pixel of a constant-z triangle stores -(ndc.z+1)/2
    far, near = 1000.0, 0.1
    ndc_z = (z * far / (near - far) + near * far / (near - far)) / (-z)
    want = -(ndc_z + 1.0) * 0.5
    assert abs(want - expected) < 1e-6
    assert np.allclose(d[covered], want, atol=2e-7)
    # the same triangle wound clockwise is culled
    mesh_cw = Mesh(v, np.array([0, 2, 1], np.uint16), Material())
    _, d2, st2 = _render(oracle, Scene(W, H, [DrawItem(mesh_cw, I, I, default_projection(W, H), cull=1)]))
    assert st2.triangles_drawn == 0 and not (d2 != CLEAR_DEPTH).any()
    # ... and drawn under CullMode.Front (Rasterizer.cs:415-417)
    _, d3, st3 = _render(oracle, Scene(W, H, [DrawItem(mesh_cw, I, I, default_projection(W, H), cull=2)]))
    # (edge pixels may differ by rounding: the accumulation chains run over differently ordered edges)
    assert st3.triangles_drawn == 1 and abs(int((d3 != CLEAR_DEPTH).sum()) - int(covered.sum())) < 0.01 * covered.sum()


# --------------------------------------------------------------------------
# A.6: inclusive edges, no top-left rule, sample at integer (x, y)
# --------------------------------------------------------------------------
def test_shared_edge_is_rasterised_by_both_triangles(oracle):
    # square (0,0)-(8,8) split along the diagonal x == y, 16x16 target
    A = [(0, 0), (8, 0), (8, 8)]
    B = [(0, 0), (8, 8), (0, 8)]
    sc = screen_tri_scene(16, 16, [A, B], cull=0)
    c, d, st = _render(oracle, sc)
    covered = d != CLEAR_DEPTH
    ys, xs = np.nonzero(covered)
    assert covered.sum() == 81 and xs.max() == 8 and ys.max() == 8      # 9x9 integer lattice points
    assert st.fragments_tested == 45 + 45                                # the 9 diagonal pixels twice (:493-494)


def test_sample_point_is_the_integer_pixel_coordinate(oracle):
    # contains the lattice point (3,3) but not the "pixel centre" (3.5,3.5)
    t_in = [(2.75, 2.75), (3.5, 2.75), (2.75, 3.5)]
    # contains (3.5,3.5)-ish area but no lattice point
    t_out = [(3.25, 3.25), (3.75, 3.25), (3.25, 3.75)]
    _, d, _ = _render(oracle, screen_tri_scene(16, 16, [t_in], cull=0))
    assert np.argwhere(d != CLEAR_DEPTH).tolist() == [[3, 3]]
    _, d, _ = _render(oracle, screen_tri_scene(16, 16, [t_out], cull=0))
    assert not (d != CLEAR_DEPTH).any()


def test_bbox_is_floor_ceil_clamped_to_the_target(oracle):
    # triangle far larger than the 32x16 target: every pixel is covered, nothing outside is touched
    big = [(-64, -64), (256, -64), (-64, 256)]
    _, d, st = _render(oracle, screen_tri_scene(32, 16, [big], cull=0))
    assert (d != CLEAR_DEPTH).all()
    assert st.fragments_tested == 32 * 16          # bbox clamped to [0,W-1]x[0,H-1] (:437-442)
    # fully off-screen: rejected by the empty clamped bbox
    off = [(40, 2), (60, 2), (40, 12)]
    _, d, st = _render(oracle, screen_tri_scene(32, 16, [off], cull=0))
    assert not (d != CLEAR_DEPTH).any() and st.fragments_tested == 0


def test_both_windings_cover_when_cull_none(oracle):
    t = [(1, 1), (9, 1), (1, 9)]
    t_rev = [t[0], t[2], t[1]]
    _, d1, _ = _render(oracle, screen_tri_scene(16, 16, [t], cull=0))
    _, d2, _ = _render(oracle, screen_tri_scene(16, 16, [t_rev], cull=0))
    assert np.array_equal(d1 != CLEAR_DEPTH, d2 != CLEAR_DEPTH)
    assert (d1 != CLEAR_DEPTH).sum() == 45       # lattice points with x>=1, y>=1, x+y<=10


# --------------------------------------------------------------------------
# A.7/A.8: depth compare order, ties, alpha discard, BlendMode.None row break
# --------------------------------------------------------------------------
def test_depth_tie_later_submission_wins(oracle):
    t = [(0, 0), (12, 0), (0, 12)]
    red, green = (1, 0, 0, 1), (0, 1, 0, 1)
    for order, winner in (([red, green], 1), ([green, red], 0)):
        sc = screen_tri_scene(16, 16, [t, t], colors=order, cull=0, separate_draws=True)
        c, d, _ = _render(oracle, sc)
        cov = d != CLEAR_DEPTH
        assert (np.argmax(c[cov][:, :3], axis=1) == winner).all()      # LessEqual is `new >= old` (:545)


@pytest.mark.parametrize("depth_test,front_wins", [(1, True), (2, True), (3, False), (4, False), (7, False), (0, False)])
def test_depth_functions_are_flipped_for_negated_depth(oracle, depth_test, front_wins):
    t = [(0, 0), (12, 0), (0, 12)]
    near_red, far_green = (1, 0, 0, 1), (0, 1, 0, 1)
    # submission order: near first, far second; NDC z 0.2 (near) vs 0.8 (far)
    sc = screen_tri_scene(16, 16, [t, t], colors=[near_red, far_green], ndc_z=[0.2, 0.8], cull=0, depth=depth_test,
                          separate_draws=True)
    # Greater/GreaterEqual/(Always, Disabled) need a cleared buffer that lets the first fragment in:
    # clear is float.MinValue, so `new < old` never passes for the first draw under Greater (:547-548)
    c, d, _ = _render(oracle, sc)
    cov = (c[..., 3] != 0)
    if depth_test in (3, 4):
        assert not cov.any()                       # nothing can be "greater" than the cleared depth
        return
    assert cov.sum() == 91
    is_red = c[cov][:, 0] > c[cov][:, 1]
    assert is_red.all() == front_wins and (~is_red).all() == (not front_wins)
    if depth_test == 0:
        assert (d == CLEAR_DEPTH).all()            # Disabled never writes depth (:517-518)


def test_equal_and_notequal_use_absolute_epsilon(oracle):
    t = [(0, 0), (12, 0), (0, 12)]
    # Equal: second draw at the same depth passes, at a different depth fails (:554-555)
    for z2, passes in ((0.5, True), (0.6, False)):
        sc = screen_tri_scene(16, 16, [t, t], colors=[(1, 0, 0, 1), (0, 1, 0, 1)], ndc_z=[0.5, z2], cull=0,
                              separate_draws=True)
        sc.draws[1].depth = 5
        c, _, _ = _render(oracle, sc)
        cov = c[..., 3] != 0
        assert (c[cov][:, 1] > 0.5).all() == passes
    sc = screen_tri_scene(16, 16, [t, t], colors=[(1, 0, 0, 1), (0, 1, 0, 1)], ndc_z=[0.5, 0.5], cull=0,
                          separate_draws=True)
    sc.draws[1].depth = 6                           # NotEqual at equal depth: rejected
    c, _, _ = _render(oracle, sc)
    assert (c[c[..., 3] != 0][:, 0] > 0.5).all()


def _alpha_hole_texture():
    from softwarerenderer_b200.shaders import Texture
    px = np.full((1, 32, 4), 255, np.uint8)
    px[0, 5, 3] = 0                                  # texel column 5 is fully transparent
    return Texture(px)


def _strip_scene(blend):
    # one triangle over a 64x16 target; u = (x+0.5)/32 so that texel column == pixel x mod 32 (Texture.cs:45-49)
    tri = [(0, 0), (64, 0), (0, 16)]
    uv = [(0.5 / 32, 0.5), (64.5 / 32, 0.5), (0.5 / 32, 0.5)]
    return screen_tri_scene(64, 16, [tri], uvs=[uv], cull=0, blend=blend, texture=_alpha_hole_texture())


def test_alpha_zero_fragment_writes_neither_colour_nor_depth(oracle):
    c, d, _ = _render(oracle, _strip_scene(blend=1))
    row = d[1] != CLEAR_DEPTH                                   # row 1 of the triangle spans x = 0..60
    assert not row[5] and not row[37] and c[1, 5, 3] == 0       # alpha <= 0 -> no write (:511)
    assert row[:5].all() and row[6:37].all() and row[38:61].all()   # Alpha blending does not break the row
    assert not row[61:].any()


def test_blend_none_breaks_the_rest_of_the_tile_row(oracle):
    c, d, _ = _render(oracle, _strip_scene(blend=0))
    row = d[1] != CLEAR_DEPTH
    assert row[:5].all() and not row[5:16].any()     # `break` leaves the x loop of THIS tile (:520-523)
    assert row[16:32].all()                          # the next tile starts a fresh loop
    assert row[32:37].all() and not row[37:48].any()
    assert row[48:61].all() and not row[61:].any()


def test_blend_modes(oracle):
    t = [(0, 0), (12, 0), (0, 12)]
    base, top = (0.5, 0.25, 1.0, 1.0), (0.5, 0.5, 0.5, 0.5)
    out = {}
    for mode in (0, 1, 2, 3):
        sc = screen_tri_scene(16, 16, [t, t], colors=[base, top], ndc_z=[0.5, 0.4], cull=0, separate_draws=True)
        sc.draws[1].blend = mode
        sc.light_dir = np.array([0, 0, -1], np.float32)   # N=(0,0,1) -> diffuse 1 -> lit colour == base colour
        c, _, _ = _render(oracle, sc)
        out[mode] = c[2, 2]
    s, d = np.array(top, F), np.array(base, F)
    assert np.allclose(out[0], s, atol=1e-6)                                   # None    (:60)
    assert np.allclose(out[1], s * s[3] + d * (1 - s[3]), atol=1e-6)           # Alpha   (:61) on all four channels
    assert np.allclose(out[2], np.minimum(s + d, 1.0), atol=1e-6)              # Additive(:62)
    assert np.allclose(out[3], s * d, atol=1e-6)                               # Multiply(:63)


# --------------------------------------------------------------------------
# A.4 near clip (Rasterizer.cs:95-160)
# --------------------------------------------------------------------------
def test_near_clip_t_solve_and_fan(oracle):
    n = 0.1
    inside = (0.0, 0.0, 1.0, 1.0)            # z >= n*w
    out1 = (1.0, 0.0, -0.5, 1.0)             # z <  n*w
    out2 = (0.0, 1.0, -0.2, 0.5)
    # one vertex inside -> one triangle
    tris = oracle.clip_near([inside, out1, out2], near=n)
    assert tris.shape == (1, 3, 4)

    def tsolve(a, b):
        z0, w0, z1, w1 = a[2], a[3], b[2], b[3]
        return (z0 - n * w0) / (n * (w1 - w0) - (z1 - z0))
    t01, t20 = tsolve(inside, out1), tsolve(out2, inside)
    p01 = np.array(inside) + (np.array(out1) - np.array(inside)) * t01
    p20 = np.array(out2) + (np.array(inside) - np.array(out2)) * t20
    assert np.allclose(tris[0, 0], inside, atol=1e-6)
    assert np.allclose(tris[0, 1], p01, atol=1e-6)
    assert np.allclose(tris[0, 2], p20, atol=1e-6)
    for p in (tris[0, 1], tris[0, 2]):
        assert abs(p[2] - n * p[3]) < 1e-6           # new vertices lie on the plane z = n*w
    # two vertices inside -> quad -> fan of two triangles sharing vertex 0 (:152-158)
    in2 = (0.5, 0.5, 2.0, 1.5)
    tris = oracle.clip_near([inside, in2, out1], near=n)
    assert tris.shape == (2, 3, 4)
    assert np.array_equal(tris[0, 0], tris[1, 0]) and np.array_equal(tris[0, 2], tris[1, 1])
    # all inside -> unchanged; all outside -> nothing
    assert oracle.clip_near([inside, in2, (0, 0, 5, 2)], near=n).shape[0] == 1
    assert oracle.clip_near([out1, out2, (0, 0, -1, 1)], near=n).shape[0] == 0


def test_near_clip_degenerate_denominator_uses_half(oracle):
    n = 0.1
    a = (0.0, 0.0, 0.1000004, 1.0)       # inside by a hair
    b = (2.0, 0.0, 0.0999996, 1.0)       # outside by a hair: |denom| = 8e-7 < 1e-6 -> t = 0.5 (:133-136)
    c = (0.0, 2.0, 0.1000004, 1.0)
    tris = oracle.clip_near([a, b, c], near=n)
    assert tris.shape[0] == 2
    assert abs(tris[0, 1, 0] - 1.0) < 1e-6           # midpoint of a-b in x


# --------------------------------------------------------------------------
# A.11 Texture.Sample (Texture.cs:42-63)
# --------------------------------------------------------------------------
def test_texture_sample_wrap_and_scale(oracle):
    px = np.zeros((4, 8, 4), np.uint8)
    px[..., 0] = np.arange(8)[None, :] * 10        # R encodes x
    px[..., 1] = np.arange(4)[:, None] * 20        # G encodes y
    px[..., 2], px[..., 3] = 255, 128

    def s(u, v):
        o = oracle.texture_sample(px, u, v)
        return int(round(o[0] * 255 / 10)), int(round(o[1] * 255 / 20))
    assert s(0.0, 0.0) == (0, 0)
    assert s(0.124, 0.24) == (0, 0) and s(0.126, 0.26) == (1, 1)       # nearest = truncation, no +0.5
    assert s(0.999, 0.999) == (7, 3)
    assert s(1.0, 1.0) == (0, 0) and s(1.25, 2.5) == (2, 2)            # u -= (int)u
    assert s(-0.25, -0.25) == (6, 3)                                   # negative: +1 after truncation
    assert s(-0.0, -0.0) == (0, 0)
    assert s(-1e-9, 0.0) == (0, 0)       # u+1 rounds to exactly 1.0f -> (int)(1*W) % W == 0 (:49-53)
    assert s(-3.75, 5.5) == (2, 2)
    o = oracle.texture_sample(px, 0.0, 0.0)
    assert o[2] == F(255) * (F(1) / F(255)) and o[3] == F(128) * (F(1) / F(255))     # x * (1/255f), not x/255f


def test_rgba8_255_maps_to_exactly_one(oracle):
    px = np.full((1, 1, 4), 255, np.uint8)
    assert (oracle.texture_sample(px, 0.3, 0.3) == 1.0).all()       # alpha is exactly 1.0


# --------------------------------------------------------------------------
# A.2 vertex stage, a17 frustum
# --------------------------------------------------------------------------
def test_vertex_shader_is_three_chained_row_vector_products(oracle):
    from softwarerenderer_b200 import numerics as nm
    rng = np.random.RandomState(7)
    M = nm.matmul(nm.create_scale(1.5, 0.5, 2.0), nm.create_rotation_y(0.7), nm.create_translation(3, -2, 1))
    V = nm.create_look_at((1, 2, 8), (0, 0, 0), (0, 1, 0))
    P = nm.create_perspective_fov(math.pi / 2, 16 / 9, 0.1, 1000.0)
    for _ in range(20):
        vtx = np.zeros(12, F)
        vtx[:3] = rng.uniform(-3, 3, 3)
        vtx[5:8] = rng.normal(size=3)
        for fma in (0, 1):
            clip, wn = oracle.vertex_shader(vtx, M, V, P, fma)
            p = np.r_[vtx[:3].astype(np.float64), 1.0]
            want = p @ M.astype(np.float64) @ V.astype(np.float64) @ P.astype(np.float64)
            assert np.allclose(clip, want, rtol=2e-5, atol=2e-5)
            n = vtx[5:8].astype(np.float64) @ M[:3, :3].astype(np.float64)       # TransformNormal: no inverse-transpose
            assert np.allclose(wn, n / np.linalg.norm(n), atol=2e-6)
    # exact small case: translation only, all sums exact in fp32
    clip, _ = oracle.vertex_shader(np.array([1, 2, 3, 0, 0, 0, 0, 1, 0, 1, 1, 1], F), nm.create_translation(4, 5, 6),
                                   nm.identity(), nm.identity(), 1)
    assert clip.tolist() == [5.0, 7.0, 9.0, 1.0]


def test_sphere_in_frustum(oracle):
    from softwarerenderer_b200 import numerics as nm
    from softwarerenderer_b200.rasterizer import BoundingSphere
    I = nm.identity()
    V = nm.create_look_at((0, 0, 0), (0, 0, -1), (0, 1, 0))
    P = nm.create_perspective_fov(math.pi / 2, 1.0, 0.1, 100.0)
    assert oracle.sphere_in_frustum(BoundingSphere((0, 0, -10), 1.0), I, V, P)
    assert not oracle.sphere_in_frustum(BoundingSphere((0, 0, 10), 1.0), I, V, P)        # behind
    assert not oracle.sphere_in_frustum(BoundingSphere((0, 0, -200), 1.0), I, V, P)      # beyond far
    assert oracle.sphere_in_frustum(BoundingSphere((0, 0, -100.5), 1.0), I, V, P)        # straddles far
    # 90 degree FOV: the side planes are x = +-z; the sphere at x=12,z=-10 has distance sqrt(2) to the right plane
    assert not oracle.sphere_in_frustum(BoundingSphere((12, 0, -10), 1.40), I, V, P)
    assert oracle.sphere_in_frustum(BoundingSphere((12, 0, -10), 1.43), I, V, P)
    # radius scales with the largest row norm of the model's 3x3 (FrustumCuller.cs:204-208)
    assert oracle.sphere_in_frustum(BoundingSphere((6, 0, -5), 0.5), nm.create_scale(2.0, 1.0, 1.0) @ I, V, P) == \
        oracle.sphere_in_frustum(BoundingSphere((12, 0, -5), 1.0), I, V, P)


# --------------------------------------------------------------------------
# a13 fragment shader (Renderer.cs:848-860)
# --------------------------------------------------------------------------
def test_fragment_shader_lighting_floor_fog_and_ambient(oracle):
    t = [(0, 0), (12, 0), (0, 12)]
    col = (0.8, 0.6, 0.4, 1.0)
    # N = (0,0,1); light pointing along -N direction gives diffuse 1; opposite gives the 0.25 floor
    for L, diff in (((0, 0, -1), 1.0), ((0, 0, 1), 0.25), ((1, 0, 0), 0.25)):
        sc = screen_tri_scene(16, 16, [t], colors=[col], cull=0)
        sc.light_dir = np.array(L, np.float32)
        c, _, _ = _render(oracle, sc)
        want = np.array(col[:3]) * (0.1 + 0.9 * diff)
        assert np.allclose(c[2, 2, :3], want, atol=1e-6) and c[2, 2, 3] == 1.0       # z=0.5 < FogStart: factor 1
        sc.fog_start, sc.fog_end = -2.0, -1.0                # (end - z)/(end - start) = -1.5 -> clamp 0 -> FogColor
        c, _, _ = _render(oracle, sc)
        assert np.allclose(c[2, 2, :3], (1.0, 0.62, 0.5), atol=1e-6)
    # smoothstep at the midpoint: f = 0.5 -> 0.5*0.5*(3-1) = 0.5
    sc = screen_tri_scene(16, 16, [t], colors=[col], cull=0)
    sc.light_dir = np.array((0, 0, -1), np.float32)
    sc.fog_start, sc.fog_end = 0.0, 1.0
    c, _, _ = _render(oracle, sc)
    assert np.allclose(c[2, 2, :3], 0.5 * np.array((1.0, 0.62, 0.5)) + 0.5 * np.array(col[:3]), atol=1e-6)


def test_extension_point_light_attenuation_and_direction(oracle):
    """EXTENSION (BASELINE config 5; not in the reference's shader): a point light adds
    Color * max(0, N.L) / (c + l*d + q*d*d) * base with L from the fragment's world position to Light.Position
    (Light.cs:9, 13-15).  Identity matrices: world position = the NDC position of the pixel."""
    from softwarerenderer_b200.shaders import Light
    t = [(0, 0), (15, 0), (0, 15)]
    col = (0.8, 0.6, 0.4, 1.0)
    sc = screen_tri_scene(16, 16, [t], colors=[col], cull=0)
    sc.light_dir = np.array((0, 0, 1), np.float32)            # stock light from behind: the 0.25 floor, k = 0.325
    lc = (0.9, 0.6, 0.3)
    sc.extra_lights = [Light(Type="Point", Position=(-0.75, 0.75, 2.5), Color=lc, AttenuationConstant=1.0,
                             AttenuationLinear=0.5, AttenuationQuadratic=0.25)]
    c, _, _ = _render(oracle, sc)
    # pixel (2,2): world (-0.75, 0.75, 0.5) straight below the light: d = 2, N.L = 1, attenuation 1/(1+1+1)
    want = np.array(col[:3]) * (0.325 + np.array(lc) / 3.0)
    assert np.allclose(c[2, 2, :3], want, atol=1e-6)
    # pixel (6,3): off axis, in float64
    wp = np.array([2 * 6 / 16 - 1, 1 - 2 * 3 / 16, 0.5])
    L = np.array([-0.75, 0.75, 2.5]) - wp
    d = np.linalg.norm(L)
    want = np.array(col[:3]) * (0.325 + np.array(lc) * (L[2] / d) / (1.0 + 0.5 * d + 0.25 * d * d))
    assert np.allclose(c[3, 6, :3], want, atol=2e-6)
    # a light behind the surface adds nothing; a directional extra light ignores position and attenuation
    sc.extra_lights = [Light(Type="Point", Position=(-0.75, 0.75, -2.5), Color=lc)]
    c, _, _ = _render(oracle, sc)
    assert np.allclose(c[2, 2, :3], np.array(col[:3]) * 0.325, atol=1e-6)
    sc.extra_lights = [Light(Direction=(0, 0, -1), Color=lc, AttenuationLinear=5.0)]
    c, _, _ = _render(oracle, sc)
    assert np.allclose(c[2, 2, :3], np.array(col[:3]) * (0.325 + np.array(lc)), atol=1e-6)


def test_perspective_correct_interpolation(oracle):
    """Interpolate (Rasterizer.cs:566-640): attribute = sum(A_i * r_i) / sum(r_i), r_i = bary_i / w_i."""
    from softwarerenderer_b200 import numerics as nm
    from softwarerenderer_b200.scene import DrawItem, Mesh, Scene
    from softwarerenderer_b200.shaders import Material, Shaders
    W = H = 64
    # clip coordinates written directly through an identity pipeline, with per-vertex w via a custom projection:
    # proj maps (x,y,z,1) -> (x, y, z, -z)  => w = -z
    P = np.zeros((4, 4), F)
    P[0, 0] = P[1, 1] = P[2, 2] = 1.0
    P[2, 3] = -1.0
    pos = [(-1.0, -1.0, -1.0), (4.0, -4.0, -4.0), (-1.0, 1.0, -1.0)]      # w = 1, 4, 1 ; ndc x = -1, 1, -1
    cols = [(1, 0, 0, 1), (0, 1, 0, 1), (1, 0, 0, 1)]
    v = Shaders.make_vertices(pos, color=cols)
    sc = Scene(W, H, [DrawItem(Mesh(v, np.array([0, 1, 2], np.uint16), Material()), nm.identity(), nm.identity(), P, cull=0)],
               clear_color=(0, 0, 0, 0), fog_start=-200.0, fog_end=-100.0)
    sc.fog_start, sc.fog_end = 100.0, 200.0
    sc.light_dir = np.array((0, 0, -1), np.float32)
    c, d, _ = _render(oracle, sc)
    y, x = 32, 16                      # on the row through vertex 1 (sy = 32): screen-linear s = x/64 from v0/v2 edge to v1
    assert d[y, x] != CLEAR_DEPTH
    s = x / 64.0                       # screen-space weight of vertex 1
    g_persp = (s / 4.0) / ((1 - s) / 1.0 + s / 4.0)
    assert abs(c[y, x, 1] - g_persp) < 1e-5 and abs(c[y, x, 1] - s) > 0.1      # perspective-correct, not screen-linear


# --------------------------------------------------------------------------
# tile semantics and variants
# --------------------------------------------------------------------------
def test_accumulation_restarts_at_tile_boundaries(oracle):
    """Edge values are re-seeded by the direct formula at each tile's clipped start pixel
    (Rasterizer.cs:481-483) and accumulated inside the tile (:527-534): a sliver whose edge
    function is tiny relative to its rounding error exposes the difference to one global walk."""
    from tests.ref_walk import coverage_by_tile_walk
    rng = np.random.RandomState(11)
    for _ in range(6):
        tri = rng.uniform(0, 63.75, size=(3, 2))
        tri = np.round(tri * 4) / 4
        sc = screen_tri_scene(64, 64, [tri], cull=0)
        _, d, _ = _render(oracle, sc)
        assert np.array_equal(d != CLEAR_DEPTH, coverage_by_tile_walk(tri, 64, 64))


def test_tile_parallel_variant_is_bit_identical_to_the_literal_one(oracle):
    from softwarerenderer_b200 import scene as S
    for sc in (S.scene_model("Gun", 320, 180), S.scene_city(320, 180, n_instances=6),
               S.scene_stress(256, 144, n_meshes=4, grid=41)):
        c0, d0, s0 = oracle.render_scene(sc, threads=0)
        c1, d1, s1 = oracle.render_scene(sc, threads=3)
        assert np.array_equal(d0, d1) and np.array_equal(c0, c1)
        assert s0.triangles_drawn == s1.triangles_drawn


def test_bands_of_the_parallel_variant_tile_the_frame(oracle):
    from softwarerenderer_b200 import scene as S
    sc = S.scene_model("Gun", 320, 200)            # 13 tile rows, last one partial (8 px)
    cf, df, _ = oracle.render_scene(sc, threads=2)
    for r0, r1 in ((0, 4), (4, 9), (9, 13)):
        cb, db, _ = oracle.render_scene(sc, threads=2, tile_rows=(r0, r1))
        y0, y1 = r0 * 16, min(r1 * 16, 200)
        assert np.array_equal(cb[y0:y1], cf[y0:y1]) and np.array_equal(db[y0:y1], df[y0:y1])
        assert (db[:y0] == CLEAR_DEPTH).all() and (db[y1:] == CLEAR_DEPTH).all()


def test_numerics_switch_changes_only_low_bits(oracle):
    from softwarerenderer_b200 import scene as S
    sc = S.scene_model("Gun", 320, 180)
    c1, d1, _ = oracle.render_scene(sc, threads=2)
    sc.numerics_fma = 0
    c0, d0, _ = oracle.render_scene(sc, threads=2)
    both = (d0 != CLEAR_DEPTH) & (d1 != CLEAR_DEPTH)
    assert ((d0 != CLEAR_DEPTH) != (d1 != CLEAR_DEPTH)).mean() < 1e-3
    assert np.abs(d0[both] - d1[both]).max() < 1e-5


# --------------------------------------------------------------------------
# N4: DebugMode.Wireframe = three DrawLine calls per triangle (Rasterizer.cs:419-425, 232-340)
# --------------------------------------------------------------------------
def _ndc_z_to_depth(z):
    return F((F(z) + F(1.0)) * F(0.5))          # depths[i] = (ndc.z + 1) * 0.5 (:388)


def test_wireframe_draws_only_the_three_edges(oracle):
    tri = [(5.0, 5.0), (58.0, 9.0), (30.0, 56.0)]
    sc = screen_tri_scene(64, 64, [tri], cull=0)
    sc.debug_mode = 1
    c, d, st = _render(oracle, sc)
    covered = d != CLEAR_DEPTH
    fill = screen_tri_scene(64, 64, [tri], cull=0)
    _, dfill, _ = _render(oracle, fill)
    filled = dfill != CLEAR_DEPTH
    assert 100 < covered.sum() < 0.25 * filled.sum()          # an outline, not the interior
    assert not covered[30, 30] and filled[30, 30]
    # a line fragment lies within half a pixel of its segment, measured from the pixel CENTRE (:296-313)
    ys, xs = np.nonzero(covered)
    P = np.array(tri, dtype=np.float64)
    def seg_dist(px, py, a, b):
        ab, ap = b - a, np.array([px, py]) - a
        t = min(1.0, max(0.0, ap.dot(ab) / ab.dot(ab)))
        return np.linalg.norm(ap - t * ab)
    for x, y in zip(xs, ys):
        dmin = min(seg_dist(x + 0.5, y + 0.5, P[i], P[(i + 1) % 3]) for i in range(3))
        assert dmin <= 0.5 + 1e-4
    # the line depth is 1/lerp(depths[0], depths[1], t): positive, unlike the negated fill depth (:315 vs :502)
    assert (d[covered] > 0).all() and (dfill[filled] < 0).all()
    assert np.allclose(d[covered], 1.0 / _ndc_z_to_depth(0.5), rtol=1e-6)


@pytest.mark.parametrize("seed", [1, 2, 3, 4])
def test_wireframe_matches_an_independent_restatement(oracle, seed):
    from tests.ref_walk import wireframe_depth
    rng = np.random.default_rng(seed)
    for _ in range(6):
        tri = np.round(rng.uniform(-6, 70, size=(3, 2)) * 4) / 4          # partly off-target: bbox clamps and truncation
        z = np.round(rng.uniform(-0.5, 0.9, size=3) * 64) / 64
        sc = screen_tri_scene(64, 64, [np.concatenate([tri, z[:, None]], axis=1)], cull=0)
        sc.debug_mode = 1
        _, d, st = _render(oracle, sc)
        want, hits = wireframe_depth(tri, [_ndc_z_to_depth(v) for v in z], 64, 64)
        if st.triangles_drawn == 0:                                           # zero area: rejected before the lines (:396)
            assert not (d != CLEAR_DEPTH).any()
            continue
        assert np.array_equal(d, want)
        assert st.fragments_tested == hits.sum()


def test_wireframe_every_edge_uses_the_first_two_depths_and_outputs(oracle):
    # DrawLine is handed the triangle's depths[]/outputs[] and always reads elements 0 and 1 (:315, :321):
    # on the edge (s2,s0) = (v0,v2) the depth at the v0 end is v2's, not v0's.  Kept as the reference has it.
    tri = [(8.5, 40.5, -0.5), (40.5, 40.5, 0.25), (24.5, 8.5, 0.75)]         # v0, v1, v2 with distinct depths
    red, green, blue = (1, 0, 0, 1), (0, 1, 0, 1), (0, 0, 1, 1)
    sc = screen_tri_scene(64, 64, [tri], colors=[[red, green, blue]], cull=0, depth=7)   # Always: last line wins
    sc.debug_mode = 1
    c, d, _ = _render(oracle, sc)
    d_v2, d_v1 = _ndc_z_to_depth(0.75), _ndc_z_to_depth(0.25)
    # pixel (8,40) has its centre exactly on v0; it is touched by line 1 (s1,s2)=(v1,v0) at t=1 and then by
    # line 2 (s2,s0)=(v0,v2) at t=0, which wins under Always: depth 1/depths[0] = 1/d(v2), colour of v2.
    assert d[40, 8] == F(1.0) / d_v2
    # before line 2, the same pixel held 1/depths[1] (line 1 at t=1): check with LessEqual, where 1/d(v1) > 1/d(v2) stays
    sc2 = screen_tri_scene(64, 64, [tri], colors=[[red, green, blue]], cull=0, depth=2)
    sc2.debug_mode = 1
    c2, d2, _ = _render(oracle, sc2)
    assert d2[40, 8] == F(1.0) / d_v1
    k = 0.1 + 0.9 * 0.5                           # normal (0,0,1), L = (0.5,-0.7071,-0.5): dot(N,-L) = 0.5 (Renderer.cs:851-857)
    assert np.allclose(c[40, 8, :3], np.array(blue[:3]) * k, atol=1e-6)
    assert np.allclose(c2[40, 8, :3], np.array(green[:3]) * k, atol=1e-6)


def test_wireframe_writes_when_alpha_is_nonzero_and_has_no_row_break(oracle):
    # fill: write iff alpha > 0 (:511); line: iff alpha != 0 (:325) -- a negative alpha is written by DrawLine only
    tri = [(4.5, 4.5), (28.5, 4.5), (16.5, 26.5)]
    sc = screen_tri_scene(32, 32, [tri], colors=[(1, 1, 1, -0.5)], cull=0, blend=0)
    sc.debug_mode = 1
    _, d, _ = _render(oracle, sc)
    assert (d != CLEAR_DEPTH).sum() > 40
    sc0 = screen_tri_scene(32, 32, [tri], colors=[(1, 1, 1, 0.0)], cull=0, blend=0)
    sc0.debug_mode = 1
    _, d0, st0 = _render(oracle, sc0)
    assert not (d0 != CLEAR_DEPTH).any() and st0.fragments_shaded == st0.fragments_tested > 40   # no early-out: all shaded
    fill = screen_tri_scene(32, 32, [tri], colors=[(1, 1, 1, -0.5)], cull=0, blend=0)
    _, df, _ = _render(oracle, fill)
    assert not (df != CLEAR_DEPTH).any()


def test_wireframe_respects_cull_mode(oracle):
    tri = [(5.0, 5.0), (58.0, 9.0), (30.0, 56.0)]
    drawn = []
    for cull in (1, 2):
        sc = screen_tri_scene(64, 64, [tri], cull=cull)
        sc.debug_mode = 1
        _, d, _ = _render(oracle, sc)
        drawn.append(bool((d != CLEAR_DEPTH).any()))
    assert sorted(drawn) == [False, True]          # the cull test precedes the wireframe branch (:414-425)


def test_wireframe_tile_parallel_variant_is_bit_identical(oracle):
    from softwarerenderer_b200 import scene as S
    sc = S.scene_model("Gun", 320, 180)
    sc.debug_mode = 1
    c0, d0, _ = oracle.render_scene(sc, threads=0)
    c1, d1, _ = oracle.render_scene(sc, threads=3)
    assert np.array_equal(d0, d1) and np.array_equal(c0, c1)
    assert 0.005 < (d0 != CLEAR_DEPTH).mean() < 0.5


# --------------------------------------------------------------------------
# RenderMesh / DrawTriangle rejections (Rasterizer.cs:208-215, 378-380) and the interpolated WorldNormal
# (InterpolateData, Rasterizer.cs:680-688)
# --------------------------------------------------------------------------
def _one_tri_scene(W, H, pos, proj, normal=None, color=None, cull=0):
    from softwarerenderer_b200 import numerics as nm
    from softwarerenderer_b200.scene import DrawItem, Mesh, Scene
    from softwarerenderer_b200.shaders import Material, Shaders
    I = nm.identity()
    v = Shaders.make_vertices(pos, normal=normal, color=color)
    mesh = Mesh(v, np.array([0, 1, 2], np.uint16), Material())
    return Scene(W, H, [DrawItem(mesh, I, I, proj, cull=cull)], clear_color=(0.0, 0.0, 0.0, 0.0), fog_start=1e6, fog_end=2e6)


def test_triangle_entirely_behind_the_eye_is_rejected(oracle):
    """All three clip w <= 0 -> the triangle is skipped before clipping (Rasterizer.cs:208-215).  With the reference's
    perspective matrix w = -z_view, so a triangle at z = +2 (behind a camera looking down -Z) has w = -2 at every corner."""
    from softwarerenderer_b200.scene import default_projection
    W, H = 64, 48
    sc = _one_tri_scene(W, H, [(-1, -1, 2.0), (1, -1, 2.0), (0, 1, 2.0)], default_projection(W, H))
    c, d, st = _render(oracle, sc)
    assert st.triangles_drawn == 0
    assert not (d != CLEAR_DEPTH).any() and not c.any()
    # the same triangle mirrored to z = -2 is in front of the eye and is drawn
    sc = _one_tri_scene(W, H, [(-1, -1, -2.0), (1, -1, -2.0), (0, 1, -2.0)], default_projection(W, H))
    _, d, st = _render(oracle, sc)
    assert st.triangles_drawn == 1 and (d != CLEAR_DEPTH).any()


@pytest.mark.parametrize("bad", [float("nan"), float("inf"), float("-inf")])
def test_non_finite_ndc_rejects_the_triangle(oracle, bad):
    """DrawTriangle returns when any NDC component is NaN or infinite (Rasterizer.cs:378-380)."""
    from softwarerenderer_b200 import numerics as nm
    W, H = 32, 32
    sc = _one_tri_scene(W, H, [(-0.5, -0.5, 0.5), (0.5, -0.5, 0.5), (bad, 0.5, 0.5)], nm.identity())
    c, d, st = _render(oracle, sc)
    assert st.triangles_drawn == 0
    assert not (d != CLEAR_DEPTH).any() and not c.any()


def test_interpolated_world_normal_is_renormalised(oracle):
    """InterpolateData normalises a Vector3 entry after the weighted sum when its squared length exceeds 1e-6
    (Rasterizer.cs:680-688).  Affine triangle (w = 1): weights are the barycentrics.  Pixel (4,4) of the triangle
    (0,0),(16,0),(0,16) has weights 0.5 / 0.25 / 0.25; normals (0,0,1),(0,0,1),(0,0,-1) sum to (0,0,0.5), which
    becomes (0,0,1): diffuse 1.0, not 0.5."""
    W = H = 32
    col = (0.8, 0.6, 0.4, 1.0)

    def scene(normals):
        t = [(0, 0), (16, 0), (0, 16)]
        pos = [(2.0 * x / W - 1.0, 1.0 - 2.0 * y / H, 0.5) for x, y in t]
        from softwarerenderer_b200 import numerics as nm
        sc = _one_tri_scene(W, H, pos, nm.identity(), normal=normals, color=[col] * 3)
        sc.light_dir = np.array((0, 0, -1), np.float32)
        return sc

    c, d, _ = _render(oracle, scene([(0, 0, 1), (0, 0, 1), (0, 0, -1)]))
    assert d[4, 4] != CLEAR_DEPTH
    assert np.allclose(c[4, 4, :3], np.array(col[:3]) * (0.1 + 0.9 * 1.0), atol=1e-6)
    # ... and where the sum points away from the light (weights 0.125/0.125/0.75 at (2,12)) the floor applies
    assert np.allclose(c[12, 2, :3], np.array(col[:3]) * (0.1 + 0.9 * 0.25), atol=1e-6)


def test_vanishing_world_normal_is_left_unnormalised(oracle):
    """Below the 1e-6 squared-length threshold the vector is NOT normalised (Rasterizer.cs:680-688): at pixel (8,4)
    the weights are 0.25 / 0.5 / 0.25 and normals (0,0,1),(0,0,-1),(0,0,1) cancel; the fragment gets the lighting
    floor max(0.25, ~0) and no NaN."""
    W = H = 32
    col = (0.8, 0.6, 0.4, 1.0)
    t = [(0, 0), (16, 0), (0, 16)]
    pos = [(2.0 * x / W - 1.0, 1.0 - 2.0 * y / H, 0.5) for x, y in t]
    from softwarerenderer_b200 import numerics as nm
    sc = _one_tri_scene(W, H, pos, nm.identity(), normal=[(0, 0, 1), (0, 0, -1), (0, 0, 1)], color=[col] * 3)
    sc.light_dir = np.array((0, 0, -1), np.float32)
    c, d, _ = _render(oracle, sc)
    assert d[4, 8] != CLEAR_DEPTH and np.isfinite(c).all()
    assert np.allclose(c[4, 8, :3], np.array(col[:3]) * (0.1 + 0.9 * 0.25), atol=1e-6)


def test_partial_last_tile_row_and_column_are_rasterised(oracle):
    """Tile loops clamp to the bounding box, not to whole tiles (Rasterizer.cs:449-476): a target whose size is not a
    multiple of TileSize = 16 (1080 rows leave 8) is covered to its last row and column."""
    W, H = 40, 24                                     # 2.5 x 1.5 tiles
    sc = screen_tri_scene(W, H, [[(-50, -50), (200, -50), (-50, 200)]], cull=0)
    _, d, _ = _render(oracle, sc)
    assert (d != CLEAR_DEPTH).all()
