"""Shaders / Material / Light / Texture -- host-side mirror of the reference types.

Shaders.cs:8-99 (VertexInput layout, delegate types), Renderer.cs:830-860 (the
stock shader pair, the only one executable on the GPU), Material.cs:6-57,
Light.cs:7-33, Texture.cs:31-94.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Dict, Optional

import numpy as np

from . import _native as N

# Shaders.VertexInput (Shaders.cs:10-24): 12 floats = 48 bytes, blittable.
VERTEX_INPUT_DTYPE = np.dtype(
    [("Position", "<f4", (3,)), ("UV", "<f4", (2,)), ("Normal", "<f4", (3,)), ("Color", "<f4", (4,))]
)
assert VERTEX_INPUT_DTYPE.itemsize == 48


class Shaders:
    VertexInput = VERTEX_INPUT_DTYPE

    @staticmethod
    def make_vertices(position, uv=None, normal=None, color=None) -> np.ndarray:
        """Build a VertexInput[] array; colour defaults to (1,1,1,1) as ModelLoader.cs:192-194."""
        position = np.asarray(position, dtype=np.float32).reshape(-1, 3)
        n = position.shape[0]
        v = np.zeros(n, dtype=VERTEX_INPUT_DTYPE)
        v["Position"] = position
        v["UV"] = 0 if uv is None else np.asarray(uv, dtype=np.float32).reshape(n, 2)
        v["Normal"] = (0, 0, 1) if normal is None else np.asarray(normal, dtype=np.float32).reshape(n, 3)
        v["Color"] = 1 if color is None else np.asarray(color, dtype=np.float32).reshape(n, 4)
        return v

    class StockVertexShader:
        """Marker for Renderer.VertexShader (Renderer.cs:830-846)."""

    class StockFragmentShader:
        """Marker for `input => Renderer.FragmentShader(input, texture)` (Renderer.cs:458, 848-860).

        The C# closure captures the texture; here it is explicit."""

        def __init__(self, texture: Optional["Texture"] = None):
            self.texture = texture


@dataclass(frozen=True)
class Material:
    """Material.cs:6-57 -- only TexturePaths[Diffuse] is read by the renderer (Renderer.cs:821-828)."""
    BaseColor: tuple = (1.0, 1.0, 1.0, 1.0)
    Metallic: float = 0.0
    Roughness: float = 0.5
    Emissive: tuple = (0.0, 0.0, 0.0)
    TexturePaths: Dict[str, str] = field(default_factory=dict)


@dataclass(frozen=True)
class Light:
    """Light.cs:7-33 -- data holder; unused by the reference's shading, used by the lights extension."""
    Position: tuple = (0.0, 0.0, 0.0)
    Direction: tuple = (0.0, -1.0, 0.0)
    Color: tuple = (1.0, 1.0, 1.0)
    Type: str = "Directional"
    AttenuationConstant: float = 1.0
    AttenuationLinear: float = 0.0
    AttenuationQuadratic: float = 0.0
    InnerConeAngle: float = 0.0
    OuterConeAngle: float = 0.0


class Texture:
    """Texture.cs:31-94.  Pixels are RGBA8, row 0 = top of the image."""

    def __init__(self, rgba8: np.ndarray):
        a = np.ascontiguousarray(rgba8, dtype=np.uint8)
        if a.ndim != 3 or a.shape[2] != 4:
            raise ValueError("Texture expects an (H, W, 4) uint8 array")
        self._pixels = a
        self._handles = {}      # context id -> SrbTexture

    @property
    def Width(self) -> int:
        return self._pixels.shape[1]

    @property
    def Height(self) -> int:
        return self._pixels.shape[0]

    @property
    def pixels(self) -> np.ndarray:
        return self._pixels

    @staticmethod
    def LoadTexture(file_path: str, max_resolution: int = 2048) -> Optional["Texture"]:
        """Texture.cs:70-94: decode, downscale to <= maxResolution, None on failure."""
        try:
            from PIL import Image

            im = Image.open(file_path).convert("RGBA")
            if im.width > max_resolution or im.height > max_resolution:
                s = min(max_resolution / im.width, max_resolution / im.height)
                im = im.resize((int(im.width * s), int(im.height * s)))
            return Texture(np.asarray(im, dtype=np.uint8))
        except Exception:
            return None

    def handle(self, window) -> int:
        key = window._uid
        h = self._handles.get(key)
        if h is None:
            import ctypes as C

            out = C.c_int32(0)
            N.check(window._ctx, N.lib().srb_texture_create(window._ctx, N.ptr(self._pixels), self.Width,
                                                            self.Height, C.byref(out)))
            h = out.value
            self._handles[key] = h
        return h

    def Sample(self, window, uv, filter: int = 0) -> np.ndarray:
        """Texture.Sample (Texture.cs:42-63), evaluated by the device kernel for one or many uv."""
        uv = np.ascontiguousarray(np.asarray(uv, dtype=np.float32).reshape(-1, 2))
        out = np.empty((uv.shape[0], 4), dtype=np.float32)
        N.check(window._ctx, N.lib().srb_texture_sample(window._ctx, self.handle(window), N.ptr(uv),
                                                        uv.shape[0], filter, N.ptr(out)))
        return out

    def Dispose(self) -> None:
        self._handles.clear()


class Light:
    """Light (Light.cs:7-33): the fields the extension lights of BASELINE config 5 use.  The reference's stock
    fragment shader has ONE directional light; extra lights are an extension: directional
    (`Type="Directional"`, `Direction`) or point (`Type="Point"`, `Position`, attenuation 1/(c + l*d + q*d*d))."""

    def __init__(self, Position=(0.0, 0.0, 0.0), Direction=(0.0, -1.0, 0.0), Color=(1.0, 1.0, 1.0), Type="Directional",
                 AttenuationConstant=1.0, AttenuationLinear=0.0, AttenuationQuadratic=0.0):
        self.Position, self.Direction, self.Color, self.Type = tuple(Position), tuple(Direction), tuple(Color), Type
        self.AttenuationConstant, self.AttenuationLinear, self.AttenuationQuadratic = (
            float(AttenuationConstant), float(AttenuationLinear), float(AttenuationQuadratic))


def light_rows(extra_lights):
    """Extra lights -> rows (dir_or_pos4, color3, atten3) of SrbUniforms / OrcUniforms.  Entries are Light objects or
    (direction3, color3) pairs (directional)."""
    rows = []
    for l in list(extra_lights)[:4]:
        if isinstance(l, Light):
            if l.Type == "Point":
                rows.append(((*l.Position, 1.0), l.Color, (l.AttenuationConstant, l.AttenuationLinear, l.AttenuationQuadratic)))
            else:
                rows.append(((*l.Direction, 0.0), l.Color, (1.0, 0.0, 0.0)))
        else:
            d, col = l
            rows.append(((*d, 0.0), col, (1.0, 0.0, 0.0)))
    return rows
