"""Seeded synthetic RGB-D frame pairs and descriptor sets (SURVEY §8d "Synthetic inputs").

Host-side input generation only (torch CPU); shared by bench.py, the tests and the oracle
so that both sides of every parity check see identical bytes.

  * goal frame    : band-limited noise texture - uniform noise at 60x80, bicubic-upsampled to
                    480x640x3, u8 (seed 1234+i)
  * current frame : goal warped by a random similarity-ish homography (+-15 deg in-plane rotation,
                    +-10 % scale, +-40 px shift, small perspective terms; seed 4321+i) + N(0, 2/255)
  * depth         : u16 millimetres, plane at 0.61 m (vitvs_v2.py:1391 desired height) tilted
                    +-10 deg, 2 % of the pixels zeroed (exercises the Z=100 branch, vitvs_v2.py:582-584)
  * resize        : PIL bicubic 640x480 -> SxS exactly as vitvs_v2.py:474-475 (input generation for the
                    pre-resized entry points; the product resizes raw frames on the GPU, csrc/resize.cu)
  * perturbed     : generate_perturbed_models.py:66-105 recipe on the current frame
                    (RandomErasing / ColorJitter / Gaussian 0.05, seed 489+i), BASELINE config 4
"""
from __future__ import annotations

import math

import numpy as np
import torch
import torch.nn.functional as F
from PIL import Image

FRAME_W, FRAME_H = 640, 480


def goal_frame(i: int) -> np.ndarray:
    g = torch.Generator().manual_seed(1234 + i)
    low = torch.rand(1, 3, 60, 80, generator=g)
    img = F.interpolate(low, size=(FRAME_H, FRAME_W), mode="bicubic", align_corners=False)
    return (img[0].clamp(0, 1) * 255).round().byte().permute(1, 2, 0).contiguous().numpy()


def _homography(i: int) -> torch.Tensor:
    g = torch.Generator().manual_seed(4321 + i)
    r = torch.rand(6, generator=g) * 2 - 1
    ang = math.radians(15.0) * r[0].item()
    sc = 1.0 + 0.10 * r[1].item()
    tx, ty = 40.0 * r[2].item(), 40.0 * r[3].item()
    cx, cy = FRAME_W / 2, FRAME_H / 2
    c, s = math.cos(ang) * sc, math.sin(ang) * sc
    H = torch.tensor([[c, -s, cx - c * cx + s * cy + tx],
                      [s, c, cy - s * cx - c * cy + ty],
                      [1e-5 * r[4].item(), 1e-5 * r[5].item(), 1.0]], dtype=torch.float64)
    return H


def current_frame(i: int, goal: np.ndarray | None = None) -> np.ndarray:
    goal = goal_frame(i) if goal is None else goal
    H = _homography(i)
    ys, xs = torch.meshgrid(torch.arange(FRAME_H, dtype=torch.float64),
                            torch.arange(FRAME_W, dtype=torch.float64), indexing="ij")
    pts = torch.stack([xs, ys, torch.ones_like(xs)], dim=-1) @ H.T     # current px -> goal px
    gx = pts[..., 0] / pts[..., 2]
    gy = pts[..., 1] / pts[..., 2]
    grid = torch.stack([(gx + 0.5) / FRAME_W * 2 - 1, (gy + 0.5) / FRAME_H * 2 - 1], dim=-1).float()[None]
    src = torch.from_numpy(goal).permute(2, 0, 1).float()[None] / 255.0
    warped = F.grid_sample(src, grid, mode="bilinear", padding_mode="reflection", align_corners=False)
    g = torch.Generator().manual_seed(99991 + i)
    warped = warped + torch.randn(warped.shape, generator=g) * (2.0 / 255.0)
    return (warped[0].clamp(0, 1) * 255).round().byte().permute(1, 2, 0).contiguous().numpy()


def depth_frame(i: int) -> np.ndarray:
    g = torch.Generator().manual_seed(777 + i)
    r = torch.rand(2, generator=g) * 2 - 1
    tx, ty = math.tan(math.radians(10.0) * r[0].item()), math.tan(math.radians(10.0) * r[1].item())
    f = 502.3016357421875
    ys, xs = torch.meshgrid(torch.arange(FRAME_H, dtype=torch.float64),
                            torch.arange(FRAME_W, dtype=torch.float64), indexing="ij")
    x = (xs - FRAME_W / 2) / f
    y = (ys - FRAME_H / 2) / f
    z = 0.61 / (1.0 - tx * x - ty * y)                                  # tilted plane through (0,0,0.61)
    mm = (z * 1000.0).round().clamp(0, 65535)
    holes = torch.rand(FRAME_H, FRAME_W, generator=g) < 0.02
    mm[holes] = 0
    return mm.to(torch.int32).numpy().astype(np.uint16)


def perturb(frame_u8: np.ndarray, i: int) -> np.ndarray:
    """generate_perturbed_models.py:68-77 recipe (tensor form), seeded per frame."""
    from torchvision import transforms
    torch.manual_seed(489 + i)
    x = torch.from_numpy(frame_u8).permute(2, 0, 1).float() / 255.0
    x = transforms.RandomErasing(p=0.5, scale=(0.02, 0.33), ratio=(0.3, 3.3))(x)
    x = transforms.ColorJitter(brightness=0.6, contrast=0.4)(x)
    x = x + torch.randn_like(x) * 0.05
    return (x.clamp(0, 1) * 255).round().byte().permute(1, 2, 0).contiguous().numpy()


def resize_square(frame_u8: np.ndarray, size: int) -> np.ndarray:
    """vitvs_v2.py:474-475: PIL Image.resize((S,S)) - Pillow default resample (bicubic)."""
    return np.asarray(Image.fromarray(frame_u8).resize((size, size)))


def frame_pair(i: int, size: int = 308, perturbed: bool = False):
    """-> goal_u8[S,S,3], cur_u8[S,S,3], depth_u16[480,640] for pair ``i``."""
    g = goal_frame(i)
    c = current_frame(i, g)
    if perturbed:
        c = perturb(c, i)
    return resize_square(g, size), resize_square(c, size), depth_frame(i)


def raw_frame_pair(i: int, perturbed: bool = False):
    """-> goal_u8[480,640,3], cur_u8[480,640,3], depth_u16[480,640]: the camera frames before vitvs_v2.py:474-475."""
    g = goal_frame(i)
    c = current_frame(i, g)
    if perturbed:
        c = perturb(c, i)
    return g, c, depth_frame(i)


def raw_frame_batch(n: int, perturbed: bool = False, distinct: int | None = None):
    """n pairs of un-resized camera frames as stacked arrays (``distinct`` as in frame_batch)."""
    distinct = n if distinct is None else min(distinct, n)
    pairs = [raw_frame_pair(i, perturbed) for i in range(distinct)]
    idx = [i % distinct for i in range(n)]
    return (np.stack([pairs[j][0] for j in idx]), np.stack([pairs[j][1] for j in idx]),
            np.stack([pairs[j][2] for j in idx]))


def frame_batch(n: int, size: int = 308, perturbed: bool = False, distinct: int | None = None):
    """n pairs as stacked arrays; only ``distinct`` different pairs are generated (then tiled)
    to keep host-side generation time bounded for large benchmark batches."""
    distinct = n if distinct is None else min(distinct, n)
    pairs = [frame_pair(i, size, perturbed) for i in range(distinct)]
    idx = [i % distinct for i in range(n)]
    goal = np.stack([pairs[j][0] for j in idx])
    cur = np.stack([pairs[j][1] for j in idx])
    depth = np.stack([pairs[j][2] for j in idx])
    return goal, cur, depth


def descriptor_sets(tokens: int, dim: int, seed: int = 7, noise: float = 0.5):
    """BASELINE config 5 / SURVEY §8d C5: A = i.i.d. Gaussian descriptors (not normalised);
    B = a random permutation of A + N(0, noise^2) so that planted matches exist.
    -> A[T,d], B[T,d] fp32, perm[T] (B[j] = A[perm[j]] + noise)."""
    g = torch.Generator().manual_seed(seed)
    a = torch.randn(tokens, dim, generator=g)
    perm = torch.randperm(tokens, generator=g)
    b = a[perm] + noise * torch.randn(tokens, dim, generator=g)
    return a, b, perm
