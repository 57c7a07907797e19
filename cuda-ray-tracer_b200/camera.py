"""Fly camera -> camera matrix, the caller side of update() (reference src/ray-tracer.cpp:25-32,44-58).

The reference keeps position / yaw / pitch (degrees) and hands `inverse(lookAt(position,
position - direction, up))` to update() each frame.  The matrix layout handed to rt_b200_update is
GLM's: 16 doubles, column-major (m[col*4 + row]).
"""
from __future__ import annotations

import math

import numpy as np

UP = np.array([0.0, 1.0, 0.0])
_DEG = 0.01745329251994329576923690768489  # glm::radians

# ray-tracer.cpp:25-32: initial pose
INITIAL_POSITION = (0.0, 0.0, 0.0)
INITIAL_YAW = 90.0
INITIAL_PITCH = 0.0


def direction(yaw_deg: float, pitch_deg: float) -> np.ndarray:
    """ray-tracer.cpp:44-48 (update_direction)."""
    y, p = yaw_deg * _DEG, pitch_deg * _DEG
    return np.array([math.cos(y) * math.cos(p), math.sin(p), math.sin(y) * math.cos(p)])


def _normalize(v):
    return v * (1.0 / math.sqrt(float(np.dot(v, v))))


def look_at_rh(eye, center, up) -> np.ndarray:
    """GLM's right-handed lookAt; returns a 4x4 array indexed [row, col]."""
    eye, center, up = (np.asarray(a, dtype=np.float64) for a in (eye, center, up))
    f = _normalize(center - eye)
    s = _normalize(np.cross(f, up))
    u = np.cross(s, f)
    m = np.eye(4)
    m[0, :3] = s
    m[1, :3] = u
    m[2, :3] = -f
    m[0, 3] = -np.dot(s, eye)
    m[1, 3] = -np.dot(u, eye)
    m[2, 3] = np.dot(f, eye)
    return m


def camera_matrix(position=INITIAL_POSITION, yaw_deg=INITIAL_YAW, pitch_deg=INITIAL_PITCH) -> np.ndarray:
    """ray-tracer.cpp:54-58.  Returns 16 doubles, column-major, ready for rt_b200_update.

    The inverse is numpy's LU inverse rather than GLM's cofactor formula; the last bits of the matrix
    may differ from the reference viewer's, which is immaterial here because the oracle and the GPU
    are always handed the very same 16 doubles."""
    pos = np.asarray(position, dtype=np.float64)
    world_to_camera = look_at_rh(pos, pos - direction(yaw_deg, pitch_deg), UP)
    cam = np.linalg.inv(world_to_camera)
    return np.ascontiguousarray(cam.T).reshape(16)  # [row, col] -> column-major


IDENTITY = np.eye(4).reshape(16)

# Fixed poses used by the parity tests and the benchmark: the reference's initial pose
# (identity to ~6e-17) plus two non-trivial ones.
TEST_POSES = {
    "identity": IDENTITY,
    "initial": camera_matrix(),
    "pose_a": camera_matrix((3.0, 2.0, -5.0), 80.0, -10.0),
    "pose_b": camera_matrix((-4.0, 1.5, -2.0), 105.0, 7.5),
}
