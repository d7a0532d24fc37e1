"""CPU-only: the ingest step in front of the path: glTF -> VertexInput[] / ushort[] per
primitive, as ModelLoader.LoadSingleModel hands them to Rasterizer.RenderMesh (ModelLoader.cs:137-217), and
Texture.LoadTexture (Texture.cs:70-94).  The model is written by the test; nothing is read from the reference."""
import json
import os

import numpy as np
import pytest


def _write_gltf(tmp_path, with_texture=True):
    pos = np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0], [1, 1, 0]], dtype=np.float32)
    nrm = np.array([[0, 0, 1]] * 4, dtype=np.float32)
    uv = np.array([[0, 0], [1, 0], [0, 0.25], [1, 1]], dtype=np.float32)
    idx = np.array([0, 1, 2, 2, 1, 3], dtype=np.uint16)
    blob = pos.tobytes() + nrm.tobytes() + uv.tobytes() + idx.tobytes()
    (tmp_path / "scene.bin").write_bytes(blob)
    views, accs, off = [], [], 0
    for arr, typ, ct in ((pos, "VEC3", 5126), (nrm, "VEC3", 5126), (uv, "VEC2", 5126), (idx, "SCALAR", 5123)):
        views.append({"buffer": 0, "byteOffset": off, "byteLength": arr.nbytes})
        accs.append({"bufferView": len(views) - 1, "componentType": ct, "count": arr.shape[0], "type": typ})
        off += arr.nbytes
    g = {
        "asset": {"version": "2.0"}, "scene": 0, "scenes": [{"nodes": [0]}],
        # parent translates by (10,0,0); child scales by 2 and rotates 90 deg about Z (quaternion x,y,z,w)
        "nodes": [{"translation": [10, 0, 0], "children": [1]},
                  {"mesh": 0, "scale": [2, 2, 2], "rotation": [0, 0, float(np.sin(np.pi / 4)), float(np.cos(np.pi / 4))]}],
        "meshes": [{"primitives": [{"attributes": {"POSITION": 0, "NORMAL": 1, "TEXCOORD_0": 2}, "indices": 3,
                                    "material": 0}]}],
        "materials": [{"pbrMetallicRoughness": {"baseColorFactor": [0.5, 0.25, 1, 1], "baseColorTexture": {"index": 0}}}],
        "textures": [{"source": 0}], "images": [{"uri": "tex.png"}],
        "buffers": [{"uri": "scene.bin", "byteLength": len(blob)}], "bufferViews": views, "accessors": accs,
    }
    (tmp_path / "scene.gltf").write_text(json.dumps(g))
    if with_texture:
        from PIL import Image
        px = np.zeros((4, 8, 4), dtype=np.uint8)
        px[..., 0] = np.arange(8)[None, :] * 30
        px[..., 1] = np.arange(4)[:, None] * 60
        px[..., 3] = 255
        Image.fromarray(px, "RGBA").save(tmp_path / "tex.png")
        return px
    return None


def test_gltf_ingest_bakes_node_transforms_and_flips_v(tmp_path):
    from softwarerenderer_b200.scene import Model
    px = _write_gltf(tmp_path)
    model = Model().LoadModel(str(tmp_path / "scene.gltf"))
    assert len(model.Meshes) == 1 and model.triangle_count == 2
    m = model.Meshes[0]
    assert m.Indices.dtype == np.uint16 and m.Indices.tolist() == [0, 1, 2, 2, 1, 3]
    assert m.Vertices.dtype.itemsize == 48                                        # Shaders.VertexInput (Shaders.cs:10-24)
    # row-vector convention: p' = p * (S * R * T_child... ) * T_parent -> scale 2, rotate +90 deg about Z, then +10 in x
    want = np.array([[10, 0, 0], [10, 2, 0], [8, 0, 0], [8, 2, 0]], dtype=np.float32)
    assert np.allclose(m.Vertices["Position"], want, atol=1e-5)
    assert np.allclose(m.Vertices["Normal"], [[0, 0, 1]] * 4, atol=1e-6)          # renormalised after the scaled 3x3 (:198-200)
    assert np.allclose(m.Vertices["UV"], [[0, 1], [1, 1], [0, 0.75], [1, 0]])     # PostProcessSteps.FlipUVs (:148)
    assert np.all(m.Vertices["Color"] == 1.0)                                     # no vertex colours -> (1,1,1,1) (:192-194)
    assert m.Material.BaseColor == (0.5, 0.25, 1, 1)
    assert m.Texture is not None and m.Texture.Width == 8 and m.Texture.Height == 4
    assert np.array_equal(m.Texture.pixels, px)                                   # row 0 = top of the image
    c, r = np.array(m.SphereBounds.Center), m.SphereBounds.Radius
    assert (np.linalg.norm(want - c, axis=1) <= r + 1e-5).all()


def test_missing_texture_file_gives_an_untextured_mesh(tmp_path):
    from softwarerenderer_b200.scene import load_gltf
    from softwarerenderer_b200.shaders import Texture
    _write_gltf(tmp_path, with_texture=False)
    meshes = load_gltf(str(tmp_path / "scene.gltf"))
    assert meshes[0].Texture is None                                              # Texture.cs:89-93: null on failure
    assert "Diffuse" in meshes[0].Material.TexturePaths
    assert Texture.LoadTexture(os.path.join(str(tmp_path), "nope.png")) is None


def test_load_texture_downscales_to_max_resolution(tmp_path):
    from PIL import Image
    from softwarerenderer_b200.shaders import Texture
    Image.fromarray(np.full((64, 128, 4), 200, dtype=np.uint8), "RGBA").save(tmp_path / "big.png")
    t = Texture.LoadTexture(str(tmp_path / "big.png"), max_resolution=32)
    assert (t.Width, t.Height) == (32, 16)                                        # Texture.cs:76-84 keeps the aspect ratio
    assert t.pixels.shape == (16, 32, 4) and t.pixels.dtype == np.uint8


def test_packed_assets_are_what_the_loader_produces():
    """assets/*.npz were baked by tools/pack_assets.py from the bundled models with this loader."""
    from softwarerenderer_b200 import scene as S
    gun = S.load_model("Gun")
    assert len(gun.Meshes) == 5 and gun.triangle_count == 3910
    assert all(m.Vertices.dtype.itemsize == 48 and m.Indices.dtype == np.uint16 for m in gun.Meshes)
    tex = [m.Texture for m in gun.Meshes if m.Texture is not None]
    assert tex and tex[0].Width == 2048 and tex[0].Height == 2048


FIXTURE = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "gltf", "gordon_freeman", "scene.gltf")


def test_bundled_gltf_fixture_loads_to_the_packed_arrays():
    """tests/golden/gltf/gordon_freeman is the reference's own OutputAssets model (CC-BY-NC, license.txt beside it),
    kept as a fixture so that Model.LoadModel runs on real exporter output where /root/reference does not exist.
    The loader's arrays must be the packed ones every other test and the bench use."""
    from softwarerenderer_b200 import scene as S
    from softwarerenderer_b200.scene import Model, load_pack
    loaded = Model().LoadModel(FIXTURE).Meshes
    packed = load_pack(os.path.join(S._repo_assets(), "gordon_freeman.npz"))
    assert len(loaded) == len(packed) and len(loaded) >= 2
    for a, b in zip(loaded, packed):
        assert np.array_equal(np.ascontiguousarray(a.Vertices).view(np.float32), np.ascontiguousarray(b.Vertices).view(np.float32))
        assert np.array_equal(a.Indices, b.Indices)
        assert (a.Texture is None) == (b.Texture is None)
        if a.Texture is not None:
            assert np.array_equal(a.Texture.pixels, b.Texture.pixels)


@pytest.mark.gpu
@pytest.mark.parametrize("pipeline", [1, 2])
def test_gltf_to_device_to_oracle(oracle, pipeline):
    """Ingest on the device side of the boundary: Model().LoadModel(glTF) -> Rasterizer.RenderMesh
    with the loader's own arrays -> CUDA frame, against the oracle fed the same arrays (ModelLoader.cs:145-217,
    Renderer.cs:444-459)."""
    import softwarerenderer_b200 as srb
    from softwarerenderer_b200 import numerics as nm
    from softwarerenderer_b200 import scene as S
    from softwarerenderer_b200.scene import DrawItem, Model, Scene, default_projection, submit_scene
    from tests.util import compare
    model = Model().LoadModel(FIXTURE)
    W, H = 960, 540
    view = nm.create_look_at((0.5, 1.0, 1.1), (0.0, 0.9, 0.0), (0, 1, 0))
    M = nm.create_rotation_y(0.4)
    sc = Scene(W, H, [DrawItem(m, M, view, default_projection(W, H)) for m in model.Meshes], fog_end=60.0)
    oc, od, ost = oracle.render_scene(sc, threads=8)
    assert (od != od.min()).mean() > 0.02
    win = srb.MainWindow(W, H)
    try:
        win.SetPipeline(pipeline)
        submit_scene(win, sc)
        stats = compare(win.ColorBuffer, win.DepthBuffer, oc, od)
        # the same arrays again: served from the identity cache (no second upload), same image
        n_meshes = len(win._mesh_by_content)
        submit_scene(win, sc)
        assert len(win._mesh_by_content) == n_meshes == len(model.Meshes)
        compare(win.ColorBuffer, win.DepthBuffer, oc, od)
        # explicit resident handles in place of the arrays
        handles = [win.CreateMesh(m.Vertices, m.Indices) for m in model.Meshes]
        assert [h.handle for h in handles] == [win.mesh_handle(m.Vertices, m.Indices) for m in model.Meshes]
    finally:
        win.Close()
    print(stats)
