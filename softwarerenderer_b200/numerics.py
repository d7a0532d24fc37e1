"""float32 helpers with System.Numerics conventions (row-major, row vectors: v' = v @ M).

These build the *inputs* of the hot path (model / view / projection matrices,
light direction) the way Renderer.cs / Camera.cs do; the same arrays are fed to
the CUDA path and to the oracle, so they need not be bit-identical to the BCL.
"""
from __future__ import annotations

import math

import numpy as np

F = np.float32


def identity() -> np.ndarray:
    return np.eye(4, dtype=F)


def create_scale(sx, sy=None, sz=None) -> np.ndarray:
    if sy is None:
        sy = sz = sx
    m = np.eye(4, dtype=F)
    m[0, 0], m[1, 1], m[2, 2] = F(sx), F(sy), F(sz)
    return m


def create_translation(x, y, z) -> np.ndarray:
    m = np.eye(4, dtype=F)
    m[3, 0], m[3, 1], m[3, 2] = F(x), F(y), F(z)
    return m


def create_rotation_y(rad) -> np.ndarray:
    c, s = F(math.cos(rad)), F(math.sin(rad))
    m = np.eye(4, dtype=F)
    m[0, 0], m[0, 2], m[2, 0], m[2, 2] = c, -s, s, c
    return m


def create_rotation_x(rad) -> np.ndarray:
    c, s = F(math.cos(rad)), F(math.sin(rad))
    m = np.eye(4, dtype=F)
    m[1, 1], m[1, 2], m[2, 1], m[2, 2] = c, s, -s, c
    return m


def quaternion_from_yaw_pitch_roll(yaw, pitch, roll) -> np.ndarray:
    """Quaternion.CreateFromYawPitchRoll -> (x, y, z, w)."""
    sr, cr = math.sin(roll * 0.5), math.cos(roll * 0.5)
    sp, cp = math.sin(pitch * 0.5), math.cos(pitch * 0.5)
    sy, cy = math.sin(yaw * 0.5), math.cos(yaw * 0.5)
    return np.array(
        [cy * sp * cr + sy * cp * sr, sy * cp * cr - cy * sp * sr, cy * cp * sr - sy * sp * cr,
         cy * cp * cr + sy * sp * sr], dtype=F)


def create_from_quaternion(q) -> np.ndarray:
    x, y, z, w = [float(v) for v in q]
    xx, yy, zz = x * x, y * y, z * z
    xy, wz, xz, wy, yz, wx = x * y, z * w, z * x, y * w, y * z, x * w
    m = np.eye(4, dtype=F)
    m[0, 0], m[0, 1], m[0, 2] = 1 - 2 * (yy + zz), 2 * (xy + wz), 2 * (xz - wy)
    m[1, 0], m[1, 1], m[1, 2] = 2 * (xy - wz), 1 - 2 * (zz + xx), 2 * (yz + wx)
    m[2, 0], m[2, 1], m[2, 2] = 2 * (xz + wy), 2 * (yz - wx), 1 - 2 * (yy + xx)
    return m


def create_from_yaw_pitch_roll(yaw, pitch, roll) -> np.ndarray:
    return create_from_quaternion(quaternion_from_yaw_pitch_roll(yaw, pitch, roll))


def transform_by_quaternion(v, q) -> np.ndarray:
    """Vector3.Transform(v, Quaternion)."""
    m = create_from_quaternion(q)
    return (np.asarray(v, dtype=F) @ m[:3, :3]).astype(F)


def _normalize(v):
    v = np.asarray(v, dtype=np.float64)
    return v / np.linalg.norm(v)


def create_look_at(position, target, up) -> np.ndarray:
    """Matrix4x4.CreateLookAt (right-handed), as used by Camera.GetViewMatrix (Camera.cs:12-17)."""
    position = np.asarray(position, dtype=np.float64)
    zaxis = _normalize(position - np.asarray(target, dtype=np.float64))
    xaxis = _normalize(np.cross(np.asarray(up, dtype=np.float64), zaxis))
    yaxis = np.cross(zaxis, xaxis)
    m = np.eye(4, dtype=np.float64)
    m[0, :3] = [xaxis[0], yaxis[0], zaxis[0]]
    m[1, :3] = [xaxis[1], yaxis[1], zaxis[1]]
    m[2, :3] = [xaxis[2], yaxis[2], zaxis[2]]
    m[3, :3] = [-np.dot(xaxis, position), -np.dot(yaxis, position), -np.dot(zaxis, position)]
    return m.astype(F)


def create_perspective_fov(fov, aspect, near, far) -> np.ndarray:
    """Matrix4x4.CreatePerspectiveFieldOfView (right-handed, NDC z in [0,1]); Renderer.cs:406-410."""
    y_scale = 1.0 / math.tan(fov * 0.5)
    x_scale = y_scale / aspect
    m = np.zeros((4, 4), dtype=F)
    m[0, 0] = F(x_scale)
    m[1, 1] = F(y_scale)
    rng = far / (near - far)
    m[2, 2] = F(rng)
    m[2, 3] = F(-1.0)
    m[3, 2] = F(near * rng)
    return m


def matmul(*ms) -> np.ndarray:
    """Matrix4x4 operator* chain in float32 (A * B applies A first under row vectors)."""
    out = np.asarray(ms[0], dtype=F)
    for m in ms[1:]:
        out = (out @ np.asarray(m, dtype=F)).astype(F)
    return out


def euler_to_direction(euler_degrees) -> np.ndarray:
    """Renderer.EulerToDirection (Renderer.cs:967-972)."""
    r = np.asarray(euler_degrees, dtype=np.float64) * math.pi / 180.0
    rot = create_from_yaw_pitch_roll(r[1], r[0], r[2])
    d = np.array([0.0, 0.0, -1.0], dtype=F) @ rot[:3, :3]
    return (d / np.linalg.norm(d)).astype(F)


class Camera:
    """Camera.cs:6-62 (caller-side; produces the view matrix input)."""

    def __init__(self, position=(0.0, 0.0, 0.0), rotation=(0.0, 0.0, 0.0, 1.0)):
        self.Position = np.asarray(position, dtype=F)
        self.Rotation = np.asarray(rotation, dtype=F)   # quaternion x,y,z,w

    def GetFront(self):
        return transform_by_quaternion([0, 0, -1], self.Rotation)

    def GetUp(self):
        return transform_by_quaternion([0, 1, 0], self.Rotation)

    def GetViewMatrix(self) -> np.ndarray:
        return create_look_at(self.Position, self.Position + self.GetFront(), self.GetUp())
