"""softwarerenderer_b200 -- B200-native rasterization backend for OCSYT/SoftwareRenderer.

Host-side mirror of the reference's Renderer/Rasterizer/Shaders/Texture/Material
API over libsrb200.so (hand-written sm_100a CUDA).  No CPU fallback.
"""
from ._native import LIB_PATH, SrbError  # noqa: F401
from .numerics import Camera  # noqa: F401
from .rasterizer import (BlendMode, BoundingSphere, CullMode, DebugMode, DepthTest, FrustumCuller,  # noqa: F401
                         Rasterizer)
from .physics import Physics, RaycastFaceMask  # noqa: F401
from .shaders import Light, Material, Shaders, Texture  # noqa: F401
from .window import MainWindow, Uniforms  # noqa: F401

__all__ = ["MainWindow", "Uniforms", "Rasterizer", "FrustumCuller", "BoundingSphere", "BlendMode", "CullMode",
           "DepthTest", "DebugMode", "Shaders", "Texture", "Material", "Light", "Camera", "Physics", "RaycastFaceMask", "SrbError", "LIB_PATH"]
