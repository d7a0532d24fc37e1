"""Scene data for the hot path: glTF ingest (ModelLoader.cs:137-326 semantics
without Assimp), Mesh/Model containers, and the synthetic BASELINE.json scenes.

Everything here is load-time: it produces the VertexInput[] / ushort[] / RGBA8
arrays and matrices that are *inputs* to the rasterization path, fed identically
to the CUDA library and to the oracle.
"""
from __future__ import annotations

import json
import math
import os
from dataclasses import dataclass, field
from typing import List, Optional

import numpy as np

from . import numerics as nm
from .rasterizer import BoundingSphere, FrustumCuller
from .shaders import VERTEX_INPUT_DTYPE, Material, Texture

_COMP = {5120: np.int8, 5121: np.uint8, 5122: np.int16, 5123: np.uint16, 5125: np.uint32, 5126: np.float32}
_NCOMP = {"SCALAR": 1, "VEC2": 2, "VEC3": 3, "VEC4": 4, "MAT4": 16}


@dataclass
class Mesh:                       # ModelLoader.cs:42-58
    Vertices: np.ndarray
    Indices: np.ndarray
    Material: Optional[Material] = None
    SphereBounds: Optional[BoundingSphere] = None
    Texture: Optional[Texture] = None      # resolved Diffuse texture (Renderer.LoadTexture, :821-828)


@dataclass
class Model:                      # ModelLoader.cs:60-67
    Meshes: List[Mesh] = field(default_factory=list)

    def LoadModel(self, path: str) -> "Model":
        self.Meshes = load_gltf(path)
        return self

    @property
    def triangle_count(self) -> int:
        return sum(m.Indices.size // 3 for m in self.Meshes)


def _accessor(g, buffers, idx) -> np.ndarray:
    acc = g["accessors"][idx]
    bv = g["bufferViews"][acc["bufferView"]]
    dt = np.dtype(_COMP[acc["componentType"]])
    nc = _NCOMP[acc["type"]]
    off = bv.get("byteOffset", 0) + acc.get("byteOffset", 0)
    stride = bv.get("byteStride", 0) or dt.itemsize * nc
    raw = buffers[bv["buffer"]]
    return np.ndarray(shape=(acc["count"], nc), dtype=dt, buffer=raw, offset=off,
                      strides=(stride, dt.itemsize)).copy()


def _node_matrix(node) -> np.ndarray:
    """Local transform as a System.Numerics row-vector matrix (float64)."""
    if "matrix" in node:
        return np.array(node["matrix"], dtype=np.float64).reshape(4, 4)   # column-major == row-vector layout
    m = np.eye(4)
    if "scale" in node:
        m = m @ np.diag([*node["scale"], 1.0])
    if "rotation" in node:
        m = m @ nm.create_from_quaternion(node["rotation"]).astype(np.float64)
    if "translation" in node:
        t = np.eye(4)
        t[3, :3] = node["translation"]
        m = m @ t
    return m


def load_gltf(path: str, load_textures: bool = True) -> List[Mesh]:
    """Per-primitive Mesh list in node-traversal order, node transforms baked into
    positions/normals (ModelLoader.cs:159-217), V flipped (PostProcessSteps.FlipUVs, :148)."""
    with open(path, "r") as f:
        g = json.load(f)
    root = os.path.dirname(path)
    buffers = []
    for b in g["buffers"]:
        with open(os.path.join(root, b["uri"]), "rb") as f:
            buffers.append(f.read())
    tex_cache = {}

    def material_of(prim):
        mi = prim.get("material")
        if mi is None:
            return Material(), None
        m = g["materials"][mi]
        pbr = m.get("pbrMetallicRoughness", {})
        paths, tex = {}, None
        bct = pbr.get("baseColorTexture")
        if bct is not None:
            img = g["images"][g["textures"][bct["index"]]["source"]]
            p = os.path.join(root, img["uri"])
            paths["Diffuse"] = p
            if load_textures:
                if p not in tex_cache:
                    tex_cache[p] = Texture.LoadTexture(p)       # None when the file is missing (Texture.cs:89-93)
                tex = tex_cache[p]
        mat = Material(BaseColor=tuple(pbr.get("baseColorFactor", (1, 1, 1, 1))),
                       Metallic=pbr.get("metallicFactor", 0.0), Roughness=pbr.get("roughnessFactor", 0.5),
                       Emissive=tuple(m.get("emissiveFactor", (0, 0, 0))), TexturePaths=paths)
        return mat, tex

    meshes: List[Mesh] = []

    def visit(ni, parent):
        node = g["nodes"][ni]
        glob = _node_matrix(node) @ parent                 # nodeTransform * parentTransform (:162)
        if "mesh" in node:
            for prim in g["meshes"][node["mesh"]]["primitives"]:
                if prim.get("mode", 4) != 4:
                    continue
                at = prim["attributes"]
                pos = _accessor(g, buffers, at["POSITION"]).astype(np.float64)
                n = pos.shape[0]
                nrm = _accessor(g, buffers, at["NORMAL"]).astype(np.float64) if "NORMAL" in at else np.zeros((n, 3))
                uv = _accessor(g, buffers, at["TEXCOORD_0"]).astype(np.float32) if "TEXCOORD_0" in at else np.zeros((n, 2), np.float32)
                col = np.ones((n, 4), dtype=np.float32)
                if "COLOR_0" in at:
                    c = _accessor(g, buffers, at["COLOR_0"]).astype(np.float32)
                    col[:, : c.shape[1]] = c
                idx = _accessor(g, buffers, prim["indices"]).reshape(-1) if "indices" in prim else np.arange(n)
                if n > 65536:
                    raise ValueError("mesh exceeds the ushort index range (ModelLoader.cs:206)")
                wpos = np.c_[pos, np.ones(n)] @ glob
                wn = nrm @ glob[:3, :3]                            # rotationOnly keeps scale (:164-168)
                ln = np.linalg.norm(wn, axis=1, keepdims=True)
                wn = np.divide(wn, ln, out=np.zeros_like(wn), where=ln > 0)
                v = np.zeros(n, dtype=VERTEX_INPUT_DTYPE)
                v["Position"] = wpos[:, :3].astype(np.float32)
                v["UV"] = np.c_[uv[:, 0], 1.0 - uv[:, 1]].astype(np.float32)     # FlipUVs
                v["Normal"] = wn.astype(np.float32)
                v["Color"] = col
                mat, tex = material_of(prim)
                ntri = idx.size // 3
                mesh = Mesh(v, idx[: ntri * 3].astype(np.uint16), mat, None, tex)
                mesh.SphereBounds = FrustumCuller.CalculateBoundingSphere(v)
                meshes.append(mesh)
        for ch in node.get("children", []):
            visit(ch, glob)

    scene = g["scenes"][g.get("scene", 0)]
    for ni in scene["nodes"]:
        visit(ni, np.eye(4))
    return meshes


# ===========================================================================
# Scene description shared by the CUDA path and the oracle harness
# ===========================================================================
ASSET_ROOT_ENV = "SRB_ASSET_ROOT"
CLEAR_COLOR = (0.9137, 0.7098, 0.6588, 1.0)          # Renderer.cs:45, 413


@dataclass
class DrawItem:
    """One Rasterizer.RenderMesh call."""
    mesh: Mesh
    model: np.ndarray
    view: np.ndarray
    proj: np.ndarray
    cull: int = 1          # CullMode.Back   (Renderer.cs:459)
    depth: int = 2         # DepthTest.LessEqual (Rasterizer.cs:173)
    blend: int = 1         # BlendMode.Alpha (Rasterizer.cs:174)
    frustum_cull: bool = False


@dataclass
class Scene:
    width: int
    height: int
    draws: List[DrawItem]
    clear_color: tuple = CLEAR_COLOR
    near_clip: float = 0.1
    light_dir: Optional[np.ndarray] = None
    light_color: tuple = (1.0, 1.0, 1.0, 1.0)
    fog_color: tuple = (1.0, 0.62, 0.5, 1.0)
    fog_start: float = 1.0
    fog_end: float = 25.0
    numerics_fma: int = 1
    tex_filter: int = 0
    extra_lights: list = field(default_factory=list)
    name: str = ""
    debug_mode: int = 0          # Rasterizer.RenderDebugMode (0 None, 1 Wireframe)

    def __post_init__(self):
        if self.light_dir is None:
            self.light_dir = nm.euler_to_direction((-45.0, -45.0, 0.0))

    @property
    def triangle_count(self) -> int:
        return sum(d.mesh.Indices.size // 3 for d in self.draws)


def default_projection(width: int, height: int, fov_deg: float = 90.0, near: float = 0.1, far: float = 1000.0):
    """Renderer.RenderScene (Renderer.cs:406-410)."""
    return nm.create_perspective_fov(fov_deg * math.pi / 180.0, width / height, near, far)


def submit_scene(window, scene: Scene) -> None:
    """Renderer.RenderScene call pattern (Renderer.cs:404-419) against the mirror API."""
    from .rasterizer import Rasterizer
    from .shaders import Shaders

    u = window.Uniforms
    u.LightDirection = np.asarray(scene.light_dir, dtype=np.float32)
    u.LightColor = np.asarray(scene.light_color, dtype=np.float32)
    u.FogColor = np.asarray(scene.fog_color, dtype=np.float32)
    u.FogStart, u.FogEnd = scene.fog_start, scene.fog_end
    u.NumericsFma, u.TexFilter, u.ExtraLights = scene.numerics_fma, scene.tex_filter, list(scene.extra_lights)
    window.ClearDepthBuffer()
    window.ClearColorBuffer(scene.clear_color)
    saved_mode = Rasterizer.RenderDebugMode
    Rasterizer.RenderDebugMode = Rasterizer.DebugMode(scene.debug_mode)
    try:
        _submit_draws(window, scene)
        window.RenderFrame(near_clip=scene.near_clip)
    finally:
        Rasterizer.RenderDebugMode = saved_mode


def _submit_draws(window, scene: Scene) -> None:
    from .rasterizer import Rasterizer
    from .shaders import Shaders

    for i, d in enumerate(scene.draws):
        Rasterizer.RenderMesh(window, d.mesh.Vertices, d.mesh.Indices, d.model, d.view, d.proj,
                              Shaders.StockVertexShader, Shaders.StockFragmentShader(d.mesh.Texture),
                              d.cull, d.depth, d.blend, submissionIndex=i,
                              sphereBounds=d.mesh.SphereBounds if d.frustum_cull else None)


def _repo_assets() -> str:
    return os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "assets")


def load_pack(path: str) -> List[Mesh]:
    """Read a flat scene pack written by tools/pack_assets.py."""
    z = np.load(path)
    textures = {}
    meshes = []
    for i in range(int(z["n_meshes"])):
        v = np.ascontiguousarray(z[f"v{i}"]).view(VERTEX_INPUT_DTYPE).reshape(-1)
        s = z[f"s{i}"]
        ti = int(z[f"t{i}"])
        if ti >= 0 and ti not in textures:
            textures[ti] = Texture(z[f"tex{ti}"])
        meshes.append(Mesh(v, np.ascontiguousarray(z[f"i{i}"]), Material(),
                           BoundingSphere(tuple(float(x) for x in s[:3]), float(s[3])), textures.get(ti)))
    return meshes


_model_cache = {}


def load_model(name: str) -> Model:
    """Bundled model by name: the baked pack under assets/ (travels to the GPU box), else the
    reference's glTF via $SRB_ASSET_ROOT or /root/reference (build container only)."""
    if name in _model_cache:
        return _model_cache[name]
    pack = os.path.join(_repo_assets(), f"{name}.npz")
    if os.path.exists(pack):
        m = Model(load_pack(pack))
    else:
        root = os.environ.get(ASSET_ROOT_ENV, "/root/reference/OutputAssets/Assets")
        m = Model().LoadModel(os.path.join(root, name, "scene.gltf"))
    _model_cache[name] = m
    return m


def gun_camera(k: int):
    """Config 1/2 cameras: k=0 base pose, k=1..8 orbit poses."""
    if k == 0:
        return nm.Camera((0.5, -0.4, 5.0)).GetViewMatrix()
    yaw = (k - 1) * math.pi / 4
    centre = np.array([0.5, -0.4, 2.0])
    pos = centre + 5.0 * np.array([math.sin(yaw), 0.0, math.cos(yaw)])
    return nm.create_look_at(pos, centre, (0, 1, 0))


def scene_model(name: str, width: int, height: int, view=None, model=None) -> Scene:
    """Configs 1 and 2: one bundled model, one directional light, textured material."""
    m = load_model(name)
    view = gun_camera(0) if view is None else view
    model = nm.identity() if model is None else model
    proj = default_projection(width, height)
    return Scene(width, height, [DrawItem(mesh, model, view, proj) for mesh in m.Meshes], name=f"{name}-{width}x{height}")


def scene_city(width: int, height: int, n_instances: int = 74, seed: int = 0x5EED0003) -> Scene:
    """Config 3: n instances of {dust2, Gun, gordon_freeman} on a 10x8 grid, FrustumCuller active."""
    rng = np.random.RandomState(seed)
    parts = [(load_model("dust2"), nm.create_scale(0.5)),
             (load_model("Gun"), nm.create_translation(0.0, 6.0, 0.0)),
             (load_model("gordon_freeman"), nm.create_translation(2.0, 0.0, 2.0))]
    cam = nm.Camera((40.0 * 4.5, 3.0, 40.0 * 3.5))
    view = cam.GetViewMatrix()
    proj = default_projection(width, height)
    draws = []
    for i in range(n_instances):
        gx, gz = i % 10, i // 10
        s, th = rng.uniform(0.5, 1.5), rng.uniform(0.0, 2 * math.pi)
        inst = nm.matmul(nm.create_scale(s), nm.create_rotation_y(th), nm.create_translation(40.0 * gx, 0.0, 40.0 * gz))
        for mdl, pre in parts:
            mm = nm.matmul(pre, inst)
            for mesh in mdl.Meshes:
                draws.append(DrawItem(mesh, mm, view, proj, frustum_cull=True))
    return Scene(width, height, draws, name=f"city{n_instances}-{width}x{height}")


def noise_texture(size: int = 1024, seed: int = 0x5EED0005, alpha: bool = False) -> Texture:
    rng = np.random.RandomState(seed)
    px = rng.randint(0, 256, size=(size, size, 4)).astype(np.uint8)
    if not alpha:
        px[..., 3] = 255
    return Texture(px)


def scene_stress(width: int, height: int, n_meshes: int = 155, grid: int = 181, seed: int = 0x5EED0005,
                 cell_px: float = 2.0, extensions: bool = False, tex_size: int = 1024) -> Scene:
    """Config 5: n_meshes jittered grid patches of (grid-1)^2*2 small triangles stacked in depth."""
    rng = np.random.RandomState(seed)
    tex = noise_texture(tex_size, seed)
    proj = default_projection(width, height)
    view = nm.identity()
    gi, gj = np.meshgrid(np.arange(grid), np.arange(grid), indexing="xy")
    q = (gj[:-1, :-1] * grid + gi[:-1, :-1]).reshape(-1)
    # two CCW (seen from +z) triangles per cell
    idx = np.stack([q, q + 1, q + grid + 1, q, q + grid + 1, q + grid], axis=1).reshape(-1).astype(np.uint16)
    draws = []
    for k in range(n_meshes):
        depth = 5.0 + 55.0 * (k / max(1, n_meshes - 1))
        cell = cell_px * 2.0 * depth / height                    # world size of cell_px pixels at this depth
        half_w, half_h = depth * width / height, depth
        ext = cell * (grid - 1)
        cx = rng.uniform(-half_w, max(-half_w, half_w - ext))
        cy = rng.uniform(-half_h, max(-half_h, half_h - ext))
        jit = rng.uniform(-0.3, 0.3, size=(grid, grid, 2)) * cell
        v = np.zeros(grid * grid, dtype=VERTEX_INPUT_DTYPE)
        px = cx + gi * cell + jit[..., 0]
        py = cy + gj * cell + jit[..., 1]
        pz = -depth + rng.uniform(-0.2, 0.2, size=(grid, grid)) * cell
        v["Position"] = np.stack([px, py, pz], axis=-1).reshape(-1, 3).astype(np.float32)
        v["UV"] = (np.stack([gi, gj], axis=-1).reshape(-1, 2) * (8.0 / (grid - 1)) - 2.0).astype(np.float32)
        v["Normal"] = (0.0, 0.0, 1.0)
        col = rng.uniform(0.5, 1.0, size=(grid * grid, 4)).astype(np.float32)
        col[:, 3] = 1.0
        v["Color"] = col
        mesh = Mesh(v, idx, Material(), None, tex)
        draws.append(DrawItem(mesh, nm.identity(), view, proj))
    sc = Scene(width, height, draws, fog_end=80.0, name=f"stress{n_meshes}x{grid}-{width}x{height}")
    if extensions:
        sc.tex_filter = 1
        # 4 lights in all: the stock directional one + a directional and two point lights with
        # the reference's attenuation fields (Light.cs:13-15), placed among the stacked meshes
        from .shaders import Light
        sc.extra_lights = [Light(Direction=(0.0, -1.0, 0.0), Color=(0.3, 0.3, 0.5)),
                           Light(Type="Point", Position=(3.0, 2.0, -8.0), Color=(1.6, 0.8, 0.4),
                                 AttenuationConstant=1.0, AttenuationLinear=0.09, AttenuationQuadratic=0.032),
                           Light(Type="Point", Position=(-12.0, -6.0, -30.0), Color=(0.4, 1.2, 2.0),
                                 AttenuationConstant=1.0, AttenuationLinear=0.045, AttenuationQuadratic=0.0075)]
        sc.name += "-ext"
    return sc
