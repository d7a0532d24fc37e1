"""Generate the golden fixtures under tests/golden/ from the CPU oracle.

    python tests/golden/make_golden.py

PARITY UNPINNED upstream: the reference has no golden data and cannot run here, so these
fixtures freeze the oracle's output (the C++ restatement of the C# path, literal
single-thread variant).  They pin (a) the oracle against regressions and (b) the CUDA
path against a committed artefact that does not depend on building the oracle.
"""
import hashlib
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

from oracle import pyoracle as O  # noqa: E402
from softwarerenderer_b200 import scene as S  # noqa: E402


def to_u8(c):
    return (np.clip(np.nan_to_num(c, nan=0.0), 0.0, 1.0) * 255.0 + 0.5).astype(np.uint8)


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def fit_camera(name):
    """Look at the model's bounding-box centre from 1.2 radii away along +Z."""
    from softwarerenderer_b200 import numerics as nm
    p = np.concatenate([m.Vertices["Position"] for m in S.load_model(name).Meshes])
    c, r = (p.min(0) + p.max(0)) / 2, float(np.linalg.norm(p.max(0) - p.min(0))) / 2
    return nm.create_look_at(c + np.array([0.3 * r, 0.2 * r, 1.2 * r]), c, (0, 1, 0))


def cases():
    """name -> Scene (small image fixtures, stored in full)."""
    return {
        "gun_320x180_cam0": S.scene_model("Gun", 320, 180),
        "gun_320x200_cam3": S.scene_model("Gun", 320, 200, view=S.gun_camera(3)),     # partial last tile row
        "gordon_256x144": S.scene_model("gordon_freeman", 256, 144, view=fit_camera("gordon_freeman")),
        "city24_480x270": S.scene_city(480, 270, n_instances=24),
        "stress4_256x144": S.scene_stress(256, 144, n_meshes=4, grid=41),
    }


def hash_cases():
    """name -> Scene (BASELINE config 1 at full size: hashes only)."""
    out = {f"gun_1280x720_cam{k}": S.scene_model("Gun", 1280, 720, view=S.gun_camera(k)) for k in range(9)}
    out["city24_1280x720"] = S.scene_city(1280, 720, n_instances=24)
    # DebugMode.Wireframe (three DrawLine calls per triangle, Rasterizer.cs:419-425, 232-340)
    for name, sc in (("gun_1280x720_cam0_wireframe", S.scene_model("Gun", 1280, 720)),
                     ("city12_640x360_wireframe", S.scene_city(640, 360, n_instances=12))):
        sc.debug_mode = 1
        out[name] = sc
    return out


def raycast_cases():
    """Physics.Raycast (Physics.cs:19-179): fixed rays against the first meshes of the Gun under a fixed model matrix."""
    from softwarerenderer_b200 import numerics as nm
    rng = np.random.default_rng(0x5EED0007)
    M = (nm.create_scale(1.25) @ nm.create_rotation_y(0.6)).astype(np.float32)
    M[3, :3] = (1.0, -0.5, 2.0)
    out = []
    for mi, m in enumerate(S.load_model("Gun").Meshes[:3]):
        c, r = np.array(m.SphereBounds.Center, dtype=np.float64), m.SphereBounds.Radius
        cw = (np.r_[c, 1.0] @ M.astype(np.float64))[:3]
        o = (cw + rng.normal(size=(24, 3)) * 4 * r).astype(np.float32)
        d = (cw + rng.uniform(-0.7, 0.7, (24, 3)) * r * 1.25 - o).astype(np.float32)
        out.append((mi, M, o, d))
    return out


def main():
    hashes = {}
    for name, sc in cases().items():
        c, d, st = O.render_scene(sc, threads=0)
        np.savez_compressed(os.path.join(HERE, name + ".npz"), color=c, depth=d)
        hashes[name] = {"depth": sha(d), "rgba8": sha(to_u8(c)), "covered": int((d != np.float32(-3.4028235e38)).sum())}
        print(name, hashes[name]["covered"], os.path.getsize(os.path.join(HERE, name + ".npz")))
    for name, sc in hash_cases().items():
        c, d, st = O.render_scene(sc, threads=0)
        hashes[name] = {"depth": sha(d), "rgba8": sha(to_u8(c)), "covered": int((d != np.float32(-3.4028235e38)).sum())}
        print(name, hashes[name]["covered"])
    with open(os.path.join(HERE, "hashes.json"), "w") as f:
        json.dump(hashes, f, indent=1, sort_keys=True)
    gun = S.load_model("Gun")
    rays = []
    for mi, M, o, d in raycast_cases():
        m = gun.Meshes[mi]
        for k in range(o.shape[0]):
            for mask in (0, 1):
                hit, dist, pt, nr, tri = O.raycast(o[k], d[k], m.Vertices, m.Indices, M, face_mask=mask)
                rays.append({"mesh": mi, "ray": k, "mask": mask, "hit": bool(hit), "triangle": int(tri),
                             "distance_bits": int(np.float32(dist).view(np.uint32)),
                             "point_bits": [int(x) for x in np.asarray(pt, np.float32).view(np.uint32)],
                             "normal": [float(x) for x in nr]})
    with open(os.path.join(HERE, "raycast.json"), "w") as f:
        json.dump(rays, f, indent=0)
    print("raycast", len(rays), sum(r["hit"] for r in rays), "hits")


if __name__ == "__main__":
    main()
