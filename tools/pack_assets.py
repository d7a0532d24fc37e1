"""Bake the reference's bundled glTF models into flat .npz scene packs.

Run in the build container (reads /root/reference, which does not exist on the
GPU box):  python tools/pack_assets.py
Each pack holds, per mesh, the VertexInput[] / ushort[] arrays exactly as
ModelLoader would hand them to Rasterizer.RenderMesh (node transforms baked,
V flipped), the bounding sphere, and the diffuse texture as RGBA8.
"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from softwarerenderer_b200.scene import load_gltf  # noqa: E402

SRC = "/root/reference/OutputAssets/Assets"
DST = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "assets")


def pack(name: str) -> None:
    meshes = load_gltf(os.path.join(SRC, name, "scene.gltf"))
    out, textures = {"n_meshes": np.int32(len(meshes))}, []
    for i, m in enumerate(meshes):
        out[f"v{i}"] = m.Vertices.view(np.float32).reshape(-1, 12)
        out[f"i{i}"] = m.Indices
        out[f"s{i}"] = np.array([*m.SphereBounds.Center, m.SphereBounds.Radius], dtype=np.float32)
        ti = -1
        if m.Texture is not None:
            for k, t in enumerate(textures):
                if t is m.Texture:
                    ti = k
            if ti < 0:
                textures.append(m.Texture)
                ti = len(textures) - 1
        out[f"t{i}"] = np.int32(ti)
    for k, t in enumerate(textures):
        out[f"tex{k}"] = t.pixels
    np.savez_compressed(os.path.join(DST, f"{name}.npz"), **out)
    print(name, len(meshes), "meshes,", len(textures), "textures ->", os.path.getsize(os.path.join(DST, f"{name}.npz")), "bytes")


if __name__ == "__main__":
    os.makedirs(DST, exist_ok=True)
    for n in ("Gun", "dust2", "gordon_freeman"):
        pack(n)
