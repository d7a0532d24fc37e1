"""Golden fixtures (tests/golden/, frozen from the oracle by make_golden.py).

CPU: the oracle (both variants) must reproduce them bit for bit.
GPU (-m gpu): the CUDA path, through the C ABI, must match them under the north-star gate
without the oracle in the loop."""
import importlib.util
import json
import os

import numpy as np
import pytest

from tests.util import compare

HERE = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
_spec = importlib.util.spec_from_file_location("make_golden", os.path.join(HERE, "make_golden.py"))
G = importlib.util.module_from_spec(_spec)
_spec.loader.exec_module(G)
HASHES = json.load(open(os.path.join(HERE, "hashes.json")))
IMAGE_CASES = sorted(n for n in HASHES if os.path.exists(os.path.join(HERE, n + ".npz")))
HASH_CASES = sorted(n for n in HASHES if n not in IMAGE_CASES)


@pytest.fixture(scope="module")
def scenes():
    d = dict(G.cases())
    d.update(G.hash_cases())
    return d


@pytest.mark.parametrize("name", IMAGE_CASES)
@pytest.mark.parametrize("threads", [0, 2])
def test_oracle_reproduces_golden_images(oracle, scenes, name, threads):
    z = np.load(os.path.join(HERE, name + ".npz"))
    c, d, _ = oracle.render_scene(scenes[name], threads=threads)
    assert np.array_equal(d, z["depth"])
    assert np.array_equal(c, z["color"])


@pytest.mark.parametrize("name", HASH_CASES)
def test_oracle_reproduces_config1_hashes(oracle, scenes, name):
    """BASELINE config 1: bundled Gun at 1280x720, the nine fixed cameras (CPU reference path)."""
    c, d, _ = oracle.render_scene(scenes[name], threads=4)
    assert int((d != np.float32(-3.4028235e38)).sum()) == HASHES[name]["covered"]
    assert G.sha(d) == HASHES[name]["depth"]
    assert G.sha(G.to_u8(c)) == HASHES[name]["rgba8"]


RAYS = json.load(open(os.path.join(HERE, "raycast.json")))


def _ray_inputs():
    from softwarerenderer_b200 import scene as S
    gun = S.load_model("Gun")
    return {mi: (gun.Meshes[mi], M, o, d) for mi, M, o, d in G.raycast_cases()}


def test_oracle_reproduces_golden_raycasts(oracle):
    """Physics.Raycast fixtures: hit flag, triangle, distance and point bit for bit; normal within 1e-6."""
    inp = _ray_inputs()
    assert sum(r["hit"] for r in RAYS) > 30
    for r in RAYS:
        m, M, o, d = inp[r["mesh"]]
        hit, dist, pt, nr, tri = oracle.raycast(o[r["ray"]], d[r["ray"]], m.Vertices, m.Indices, M, face_mask=r["mask"])
        assert hit == r["hit"] and tri == r["triangle"]
        assert int(np.float32(dist).view(np.uint32)) == r["distance_bits"]
        assert [int(x) for x in pt.view(np.uint32)] == r["point_bits"]
        assert np.allclose(nr, r["normal"], atol=1e-6)


@pytest.mark.gpu
def test_cuda_matches_golden_raycasts():
    import softwarerenderer_b200 as srb
    inp = _ray_inputs()
    win = srb.MainWindow(64, 64)
    try:
        for mi, (m, M, o, d) in inp.items():
            for mask in (0, 1):
                got = srb.Physics.RaycastBatch(win, o, d, m.Vertices, m.Indices, M, srb.RaycastFaceMask(mask))
                for r in (x for x in RAYS if x["mesh"] == mi and x["mask"] == mask):
                    g = got[r["ray"]]
                    assert bool(g["hit"]) == r["hit"] and int(g["triangle"]) == r["triangle"]
                    assert int(g["distance"].view(np.uint32)) == r["distance_bits"]
                    assert [int(x) for x in g["point"].view(np.uint32)] == r["point_bits"]
                    assert np.allclose(g["normal"], r["normal"], atol=1e-6)
    finally:
        win.Close()


@pytest.mark.gpu
@pytest.mark.parametrize("pipeline", [1, 2])
@pytest.mark.parametrize("name", IMAGE_CASES)
def test_cuda_matches_golden_images(scenes, name, pipeline):
    import softwarerenderer_b200 as srb
    from softwarerenderer_b200.scene import submit_scene
    z = np.load(os.path.join(HERE, name + ".npz"))
    sc = scenes[name]
    win = srb.MainWindow(sc.width, sc.height)
    try:
        win.SetPipeline(pipeline)
        submit_scene(win, sc)
        gc, gd = win.ColorBuffer, win.DepthBuffer
    finally:
        win.Close()
    stats = compare(gc, gd, z["color"], z["depth"])          # coverage + depth bit-exact, RGBA8 +-1 on >= 99.9 %
    assert stats["rgba8_max_err"] <= 1, stats


@pytest.mark.gpu
@pytest.mark.parametrize("name", HASH_CASES)
def test_cuda_depth_matches_config1_hashes(scenes, name):
    import softwarerenderer_b200 as srb
    from softwarerenderer_b200.scene import submit_scene
    sc = scenes[name]
    win = srb.MainWindow(sc.width, sc.height)
    try:
        submit_scene(win, sc)
        gd = win.DepthBuffer
    finally:
        win.Close()
    assert G.sha(gd) == HASHES[name]["depth"]                # coverage mask and depth winner: bit-exact
