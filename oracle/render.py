"""Oracle: CPU restatement of the planar-scene RGB-D camera and of the rigid-body
integration of the camera twist.

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).  PARITY UNPINNED by the reference: the reference
gets its frames from Gazebo / OGRE (a graphics renderer - "unavailable on this box"); what is restated
here is the geometry the reference's world files define - a 0.8 x 0.6 m textured plane
(models/viso/meshes/resized.dae:113), a 640 x 480 pinhole camera with the intrinsics of config.yaml,
the depth encoding of RealSensePlugin.cpp:252-262 (uint16 mm, 0 outside [0.2, 10] m) - and the
first-order kinematic integration gazebo_vel_broadcaster.py:32-46 asks Gazebo for.  fp64 numpy, one
rounding per operation, in the operation order csrc/render.cu uses (the kernel is bit-identical).
"""
from __future__ import annotations

import numpy as np

PLANE_W, PLANE_H = 0.8, 0.6


def render_plane(tex: np.ndarray, pose: np.ndarray, width=640, height=480, fx=502.3016357421875,
                 fy=502.3016357421875, cx=320.0, cy=240.0, plane_w=PLANE_W, plane_h=PLANE_H):
    """tex [th, tw, 3] u8; pose [12] = camera-to-world R (row-major 9) + t (3).
    -> rgb [height, width, 3] u8, depth [height, width] u16 (mm)."""
    th, tw = tex.shape[:2]
    P = np.asarray(pose, np.float64)
    u = np.arange(width, dtype=np.float64)[None, :].repeat(height, 0)
    v = np.arange(height, dtype=np.float64)[:, None].repeat(width, 1)
    dx, dy = (u - cx) / fx, (v - cy) / fy
    wx = (P[0] * dx + P[1] * dy) + P[2]
    wy = (P[3] * dx + P[4] * dy) + P[5]
    wz = (P[6] * dx + P[7] * dy) + P[8]
    rgb = np.full((height, width, 3), 128, np.uint8)
    depth = np.zeros((height, width), np.uint16)
    fwd = wz > 1e-12
    with np.errstate(divide="ignore", invalid="ignore"):
        s = np.where(fwd, -P[11] / np.where(fwd, wz, 1.0), -1.0)
    hit = fwd & (s > 0.0)
    px, py = P[9] + s * wx, P[10] + s * wy
    in_range = hit & (s >= 0.2) & (s <= 10.0)
    depth[in_range] = np.floor(s[in_range] * 1000.0 + 0.5).astype(np.uint16)
    tx = ((px + 0.5 * plane_w) / plane_w) * tw - 0.5
    ty = ((py + 0.5 * plane_h) / plane_h) * th - 0.5
    on = hit & (tx >= -0.5) & (tx <= tw - 0.5) & (ty >= -0.5) & (ty <= th - 0.5)
    fx0, fy0 = np.floor(tx[on]), np.floor(ty[on])
    ax, ay = tx[on] - fx0, ty[on] - fy0
    x0 = np.clip(fx0.astype(np.int64), 0, tw - 1)
    x1 = np.clip(fx0.astype(np.int64) + 1, 0, tw - 1)
    y0 = np.clip(fy0.astype(np.int64), 0, th - 1)
    y1 = np.clip(fy0.astype(np.int64) + 1, 0, th - 1)
    t = tex.astype(np.float64)
    out = np.empty((len(fx0), 3), np.uint8)
    for k in range(3):
        c00, c01, c10, c11 = t[y0, x0, k], t[y0, x1, k], t[y1, x0, k], t[y1, x1, k]
        top = c00 + ax * (c01 - c00)
        bot = c10 + ax * (c11 - c10)
        val = top + ay * (bot - top)
        out[:, k] = np.clip(np.floor(val + 0.5), 0.0, 255.0).astype(np.uint8)
    rgb[on] = out
    return rgb, depth


# ------------------------------------------------------------------ rigid-body kinematics (host side of the loop)
def skew(w):
    return np.array([[0.0, -w[2], w[1]], [w[2], 0.0, -w[0]], [-w[1], w[0], 0.0]])


def exp_so3(w):
    """Rodrigues formula."""
    th = float(np.linalg.norm(w))
    K = skew(w)
    if th < 1e-12:
        return np.eye(3) + K
    return np.eye(3) + (np.sin(th) / th) * K + ((1.0 - np.cos(th)) / (th * th)) * (K @ K)


def integrate_twist(pose: np.ndarray, twist: np.ndarray, dt: float) -> np.ndarray:
    """Camera-to-world pose after applying the camera-frame twist (v, w) for dt seconds, first order in the
    translation (what gazebo_vel_broadcaster.py:32-46 delegates to Gazebo's SetModelState twist)."""
    P = np.asarray(pose, np.float64)
    R, t = P[:9].reshape(3, 3), P[9:]
    v, w = np.asarray(twist[:3], np.float64), np.asarray(twist[3:], np.float64)
    t2 = t + R @ (v * dt)
    R2 = R @ exp_so3(w * dt)
    return np.concatenate([R2.reshape(-1), t2])


def pose_error(pose: np.ndarray, goal: np.ndarray):
    """(translation error in metres, rotation error in degrees) between two camera-to-world poses - the two
    quantities Controller.calculate_end_error reports (vitvs_v2.py:351)."""
    Ra, ta = np.asarray(pose[:9]).reshape(3, 3), np.asarray(pose[9:])
    Rb, tb = np.asarray(goal[:9]).reshape(3, 3), np.asarray(goal[9:])
    c = (np.trace(Ra.T @ Rb) - 1.0) / 2.0
    return float(np.linalg.norm(ta - tb)), float(np.degrees(np.arccos(np.clip(c, -1.0, 1.0))))
