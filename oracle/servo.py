"""Oracle: CPU restatement of the descriptor -> matching -> IBVS stages.

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).  PARITY PINNED: every function
here is checked against the reference's own code executed from /root/reference
(tests/test_oracle_vs_reference.py) and against tests/golden/*.npz made by
oracle/make_golden.py.  All citations are /root/reference/catkin_ws/ibvs/src/.

torch fp32 on CPU for the tensor stages (what the reference computes in),
numpy fp64 for the control law (what the reference computes in).
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field

import numpy as np
import torch
import torch.nn.functional as F

IMAGENET_MEAN = (0.485, 0.456, 0.406)   # dinov2_extractor.py:49
IMAGENET_STD = (0.229, 0.224, 0.225)    # dinov2_extractor.py:50

DEFAULT_CFG = dict(                      # config/config.yaml:1-17,38 + hard-coded constants
    u_max=640, v_max=480, f_x=502.3016357421875, f_y=502.3016357421875,
    lambda_=0.03, num_pairs=24, dino_input_size=308, use_feature_binning=True,
    ema_alpha=0.8, max_velocity=1.0,
    layer=11, facet="token",             # vitvs_v2.py:484-485
    distance_threshold=1,                # vitvs_v2.py:72
    same_image_threshold=0.99,           # vitvs_v2.py:84
)


# --------------------------------------------------------------------------- a3
def preprocess_u8(img_u8: np.ndarray) -> torch.Tensor:
    """dinov2_extractor.py:177-191: ToTensor (x/255, HWC->CHW) + Normalize. [1,3,H,W] fp32."""
    x = torch.from_numpy(np.array(img_u8, copy=True)).permute(2, 0, 1).float().div(255.0)
    mean = torch.tensor(IMAGENET_MEAN).view(3, 1, 1)
    std = torch.tensor(IMAGENET_STD).view(3, 1, 1)
    return ((x - mean) / std)[None]


# ------------------------------------------------------------------------ a7/a8
@torch.no_grad()
def block_tokens(model, batch: torch.Tensor, layer: int = 11) -> torch.Tensor:
    """dinov2_extractor.py:193-201,226-229,245-263 with facet 'token': the OUTPUT of
    ``blocks[layer]`` (residual stream, before the final norm).  [B, 1+T, D].
    Stops after ``layer`` (identical values; the reference runs the deeper blocks and
    discards them)."""
    x = model.prepare_tokens(batch)
    for blk in model.blocks[: layer + 1]:
        x = blk(x)
    return x


def log_bin(tokens_no_cls: torch.Tensor, grid: tuple[int, int]) -> torch.Tensor:
    """dinov2_extractor.py:265-311 with hierarchy=1: the nine 3x3 neighbours (row-major
    (dy,dx) order, own position in the middle), replicate-clamped at the border,
    concatenated channel-minor.  [B,T,D] -> [B,T,9D].  (k=0 avg-pool is the identity.)"""
    B, T, D = tokens_no_cls.shape
    hp, wp = grid
    g = tokens_no_cls.reshape(B, hp, wp, D)
    ys = torch.arange(hp)
    xs = torch.arange(wp)
    parts = []
    for dy in (-1, 0, 1):
        for dx in (-1, 0, 1):
            yy = (ys + dy).clamp(0, hp - 1)
            xx = (xs + dx).clamp(0, wp - 1)
            parts.append(g[:, yy][:, :, xx])
    return torch.cat(parts, dim=-1).reshape(B, T, 9 * D)


def log_bin_hierarchy(tokens_no_cls: torch.Tensor, grid: tuple[int, int], hierarchy: int) -> torch.Tensor:
    """dinov2_extractor.py:265-311 for any ``hierarchy``: level k holds the 3x3 neighbourhood at
    distance 3^k of the 3^k x 3^k average-pooled map (AvgPool2d stride 1, padding 3^k // 2, count_include_pad=False),
    centre skipped for k > 0, out-of-grid neighbours replicate-clamped.  [B,T,D] -> [B,T,(1+8h)D], level-major then
    (dy,dx) row-major.  PINNED against the reference's own _log_bin (tests/test_oracle_vs_reference.py)."""
    B, T, D = tokens_no_cls.shape
    hp, wp = grid
    g = tokens_no_cls.reshape(B, hp, wp, D).permute(0, 3, 1, 2)                     # B D hp wp
    parts = []
    for k in range(hierarchy):
        win = 3 ** k
        pooled = torch.nn.AvgPool2d(win, stride=1, padding=win // 2, count_include_pad=False)(g)
        for dy in (-win, 0, win):
            for dx in (-win, 0, win):
                if dy == 0 and dx == 0 and k != 0:
                    continue
                yy = (torch.arange(hp) + dy).clamp(0, hp - 1)
                xx = (torch.arange(wp) + dx).clamp(0, wp - 1)
                parts.append(pooled[:, :, yy][:, :, :, xx])
    out = torch.cat(parts, dim=1)                                                    # B (bins*D) hp wp
    return out.flatten(start_dim=-2).permute(0, 2, 1)


@torch.no_grad()
def extract_descriptors(model, batch: torch.Tensor, layer: int = 11, bin: bool = True) -> torch.Tensor:
    """dinov2_extractor.py:313-337 for facet='token', include_cls=False. -> [B,1,T,d']."""
    x = block_tokens(model, batch, layer)[:, 1:, :]
    hp, wp = batch.shape[2] // 14, batch.shape[3] // 14
    if bin:
        x = log_bin(x, (hp, wp))
    return x.unsqueeze(1)


@torch.no_grad()
def extract_facet_descriptors(model, batch: torch.Tensor, layer: int = 11, facet: str = "key", bin: bool = False,
                              include_cls: bool = False) -> torch.Tensor:
    """dinov2_extractor.py:203-218 + :313-337 for facet in {'key','query','value'}: the hook on
    ``blocks[layer].attn`` recomputes ``qkv(input)`` of that block's attention input (= norm1 of the residual
    stream entering the block), takes q / k / v as [B, h, t, d], drops CLS unless asked to keep it, and flattens
    to channel order (d major, h minor) - ``permute(0, 2, 3, 1).flatten(-2)`` - before the optional 3x3 binning.
    -> [B, 1, t, d'] (PINNED: tests/test_oracle_vs_reference.py runs the reference's ViTExtractor against it)."""
    idx = {"query": 0, "key": 1, "value": 2}[facet]
    x = model.prepare_tokens(batch)
    for blk in model.blocks[:layer]:
        x = blk(x)
    blk = model.blocks[layer]
    inp = blk.norm1(x)
    B, N, C = inp.shape
    h = blk.attn.num_heads
    qkv = blk.attn.qkv(inp).reshape(B, N, 3, h, C // h).permute(2, 0, 3, 1, 4)[idx]     # [B, h, t, d]
    if not include_cls:
        qkv = qkv[:, :, 1:, :]
    else:
        assert not bin
    flat = qkv.permute(0, 2, 3, 1).flatten(start_dim=-2, end_dim=-1)                     # [B, t, d*h]
    if bin:
        hp, wp = batch.shape[2] // 14, batch.shape[3] // 14
        flat = log_bin(flat, (hp, wp))
    return flat.unsqueeze(1)


# --------------------------------------------------------------------------- a9
def cosine_similarity_matrix(desc1: torch.Tensor, desc2: torch.Tensor) -> torch.Tensor:
    """vitvs_v2.py:49-56 (chunk_cosine_sim): S[i,j] = cos(desc1[i], desc2[j]), eps 1e-8.
    Vectorised form of the reference's per-goal-token loop. [B,1,T,d]x2 -> [B,1,T,T]."""
    a = desc1 / desc1.norm(dim=-1, keepdim=True).clamp_min(1e-8)
    b = desc2 / desc2.norm(dim=-1, keepdim=True).clamp_min(1e-8)
    return a @ b.transpose(-1, -2)


def cosine_similarity_matrix_literal(desc1: torch.Tensor, desc2: torch.Tensor) -> torch.Tensor:
    """Same as above in the reference's own loop structure (one CosineSimilarity call per goal
    token, vitvs_v2.py:52-56) - this is what the CPU baseline times."""
    rows = []
    cos = torch.nn.CosineSimilarity(dim=3)
    for i in range(desc1.shape[2]):
        rows.append(cos(desc1[:, :, i:i + 1, :], desc2))
    return torch.stack(rows, dim=2)


# -------------------------------------------------------------------------- a10
@dataclass
class MatchResult:
    nn1: np.ndarray           # [T] int64  argmax_j S[i,j]   (vitvs_v2.py:80)
    nn2: np.ndarray           # [T] int64  argmax_i S[i,j]   (vitvs_v2.py:81)
    sim1: np.ndarray          # [T] f32    max_j S[i,j]
    mean_sim1: float
    same_image: bool          # vitvs_v2.py:84
    candidates: np.ndarray    # goal-token indices that pass `>= distance_threshold`, ascending
    mutual: np.ndarray = field(default=None)  # [T] bool  nn2[nn1[i]] == i


def cyclic_distance_f32(dr: np.ndarray, dc: np.ndarray) -> np.ndarray:
    """vitvs_v2.py:117-120: PairwiseDistance(p=2) on int64 (row,col) pairs = ||a - b + 1e-6||_2
    evaluated in fp32 (int64 promoted by the float eps).  torch's CPU 2-norm accumulates with
    fused multiply-adds: acc = fl(a*a); acc = fma(b, b, acc); sqrt - bit-exact over every offset of
    a 46x46 grid (tests/test_oracle_vs_reference.py)."""
    a = np.asarray(dr).astype(np.float32) + np.float32(1e-6)
    b = np.asarray(dc).astype(np.float32) + np.float32(1e-6)
    aa = (a * a).astype(np.float32).astype(np.float64)
    acc = (b.astype(np.float64) * b.astype(np.float64) + aa).astype(np.float32)   # exact product, one rounding
    return np.sqrt(acc).astype(np.float32)


def match(S: torch.Tensor, cfg=DEFAULT_CFG) -> MatchResult:
    """vitvs_v2.py:80-131 for B=1, up to (not including) the random selection.
    S: [T,T] fp32, rows = goal tokens, cols = current tokens."""
    T = S.shape[0]
    width = int(np.sqrt(T))                                   # :75 (square grid assumption)
    sim1, nn1 = torch.max(S, dim=-1)                          # ties -> lowest index (CPU)
    _, nn2 = torch.max(S, dim=-2)
    sim1, nn1, nn2 = sim1.numpy(), nn1.numpy(), nn2.numpy()
    mean_sim1 = float(torch.from_numpy(sim1).mean().item())
    same = mean_sim1 > cfg["same_image_threshold"]
    cyc = nn2[nn1]                                            # :106
    idx = np.arange(T)
    mutual = cyc == idx
    d = -cyclic_distance_f32(cyc // width - idx // width, cyc % width - idx % width)
    dn = d - d.min()                                          # :123
    dn = (dn / (dn.max() + np.float32(1e-8))).astype(np.float32)   # :124
    cand = np.nonzero(dn >= np.float32(cfg["distance_threshold"]))[0]   # :130-131 (as a set)
    return MatchResult(nn1, nn2, sim1, mean_sim1, same, cand, mutual)


def to_cartesian(idx: np.ndarray, width: int) -> np.ndarray:
    """vitvs_v2.py:58-70: raveled index -> (row, col)."""
    return np.stack([idx // width, idx % width], axis=-1)


def select_pairs(m: MatchResult, chosen_goal_tokens: np.ndarray):
    """vitvs_v2.py:143-151 given the goal tokens the random selection picked:
    points1 (goal, row/col), points2 (current), sim."""
    width = int(np.sqrt(len(m.nn1)))
    p1 = to_cartesian(chosen_goal_tokens, width)
    p2 = to_cartesian(m.nn1[chosen_goal_tokens], width)
    return p1, p2, m.sim1[chosen_goal_tokens]


def top_k_by_similarity(m: MatchResult, k: int) -> np.ndarray:
    """The deterministic selection policy of the B200 path: the k candidates with
    the highest sim1, ties -> lower token index; returned in that order."""
    c = m.candidates
    order = np.lexsort((c, -m.sim1[c].astype(np.float64)))
    return c[order[:k]]


# -------------------------------------------------------------------------- a11
def py_round_half_even(x: float) -> int:
    return int(round(x))      # Python round() == banker's rounding (vitvs_v2.py:548-551)


def calculate_uv(points1_rc, points2_rc, T: int, cfg=DEFAULT_CFG):
    """vitvs_v2.py:511-513 + :525-553.  (row,col) patch coords -> patch-centre pixels of the
    S x S image -> (u,v) of the camera frame.  Returns (s_uv_star, s_uv) int [num_pairs,2],
    zero rows beyond the supplied count; all-zero when fewer than 4 pairs (:539-541)."""
    S = cfg["dino_input_size"]
    K = cfg["num_pairs"]
    scale = S / int(np.sqrt(T))
    # vitvs_v2.py:511-513 `points1 * scale + scale / 2`: an int64 tensor times a Python float is evaluated by torch in
    # the default dtype, fp32 - the scalar is rounded to fp32, product and sum round to fp32 separately - and
    # `.tolist()` then widens those fp32 values to Python floats for calculate_uv.  For S / grid not exactly
    # representable (stride 7: 112 / 15, 308 / 43) the fp64 form would round some key-points the other way.
    def centres(p):
        p = np.asarray(p, dtype=np.int64).reshape(-1, 2).astype(np.float32)
        return (p * np.float32(scale) + np.float32(scale / 2)).astype(np.float64)
    g = np.flip(centres(points1_rc), 1)
    c = np.flip(centres(points2_rc), 1)
    s_uv = np.zeros([K, 2], dtype=int)
    s_uv_star = np.zeros([K, 2], dtype=int)
    n = len(g)
    if n != K and n < 4:
        return s_uv_star, s_uv
    sx, sy = cfg["u_max"] / S, cfg["v_max"] / S
    for i in range(min(n, K)):
        s_uv_star[i] = (py_round_half_even(g[i, 0] * sx), py_round_half_even(g[i, 1] * sy))
        s_uv[i] = (py_round_half_even(c[i, 0] * sx), py_round_half_even(c[i, 1] * sy))
    return s_uv_star, s_uv


# ---------------------------------------------------------------------- a12-a16
def get_depth(s_uv: np.ndarray, depth_u16: np.ndarray) -> np.ndarray:
    """vitvs_v2.py:566-586: Z = depth[v,u]/1000 m; 0 or out of bounds -> 100."""
    h, w = depth_u16.shape
    z = np.zeros((len(s_uv), 1))
    for i, (u, v) in enumerate(s_uv):
        u, v = int(u), int(v)
        if 0 <= u < w and 0 <= v < h and depth_u16[v, u] != 0:
            z[i] = depth_u16[v, u] / 1000.0
        else:
            z[i] = 100
    return z


def to_normalised(s_uv: np.ndarray, cfg=DEFAULT_CFG) -> np.ndarray:
    """vitvs_v2.py:634-648 with c = (u_max/2, v_max/2) (:282-283)."""
    cx, cy = cfg["u_max"] / 2, cfg["v_max"] / 2
    return np.stack([(s_uv[:, 0] - cx) / cfg["f_x"], (s_uv[:, 1] - cy) / cfg["f_y"]], axis=1)


def interaction_matrix(s_xy: np.ndarray, Z: np.ndarray) -> np.ndarray:
    """vitvs_v2.py:650-659."""
    L = np.zeros([2 * len(s_xy), 6])
    for i in range(len(s_xy)):
        x, y, z = s_xy[i, 0], s_xy[i, 1], Z[i, 0]
        L[2 * i] = [-1 / z, 0, x / z, x * y, -(1 + x * x), y]
        L[2 * i + 1] = [0, -1 / z, y / z, 1 + y * y, -x * y, -x]
    return L


def ibvs_twist(s_uv: np.ndarray, s_uv_star: np.ndarray, depth_u16: np.ndarray, cfg=DEFAULT_CFG) -> np.ndarray:
    """vitvs_v2.py:610-622: v = -lambda * pinv(L) @ e.  [6] fp64 (before EMA)."""
    s_xy, s_star = to_normalised(s_uv, cfg), to_normalised(s_uv_star, cfg)
    e = (s_xy - s_star).reshape(2 * len(s_xy), 1)
    L = interaction_matrix(s_xy, get_depth(s_uv, depth_u16))
    return (-cfg["lambda_"] * np.linalg.pinv(L) @ e).flatten()


class Ema:
    """vitvs_v2.py:325-343: first call copies, then a*v + (1-a)*prev per component."""

    def __init__(self, alpha: float):
        self.alpha = alpha
        self.state = None

    def update(self, v: np.ndarray) -> np.ndarray:
        self.state = v.copy() if self.state is None else self.alpha * v + (1 - self.alpha) * self.state
        return self.state.copy()


def to_gazebo_twist(v_c: np.ndarray, max_velocity: float) -> np.ndarray:
    """vitvs_v2.py:668-676: (v2,-v0,-v1, v5,-v3,-v4) clipped to +-max_velocity."""
    t = np.array([v_c[2], -v_c[0], -v_c[1], v_c[5], -v_c[3], -v_c[4]])
    return np.clip(t, -max_velocity, max_velocity)


# ------------------------------------------------------------------- end-to-end
@torch.no_grad()
def frame_pair_to_twist(model, goal_u8: np.ndarray, cur_u8: np.ndarray, depth_u16: np.ndarray,
                        cfg=DEFAULT_CFG, literal_similarity: bool = False):
    """One pass of Controller.ibvs() (vitvs_v2.py:588-623) on S x S uint8 inputs that are
    already resized (a2 is host-side), with the deterministic top-K selection.
    Returns dict of every intermediate the parity tests look at."""
    d1 = extract_descriptors(model, preprocess_u8(goal_u8), cfg["layer"], cfg["use_feature_binning"])
    d2 = extract_descriptors(model, preprocess_u8(cur_u8), cfg["layer"], cfg["use_feature_binning"])
    sim_fn = cosine_similarity_matrix_literal if literal_similarity else cosine_similarity_matrix
    S = sim_fn(d1, d2)[0, 0]
    m = match(S, cfg)
    T = S.shape[0]
    out = dict(desc1=d1, desc2=d2, S=S, match=m, twist=None, s_uv=None, s_uv_star=None, chosen=None)
    if m.same_image:
        # reference takes K random identical points, e = 0 => twist = 0 whatever the choice
        chosen = np.argsort(-m.sim1.astype(np.float64), kind="stable")[: cfg["num_pairs"]]
        p1 = to_cartesian(chosen, int(np.sqrt(T)))
        p2 = p1.copy()
    else:
        if len(m.candidates) == 0:
            return out
        chosen = top_k_by_similarity(m, cfg["num_pairs"])
        p1, p2, _ = select_pairs(m, chosen)
    s_uv_star, s_uv = calculate_uv(p1, p2, T, cfg)
    out.update(chosen=chosen, s_uv=s_uv, s_uv_star=s_uv_star)
    # < 4 matches: calculate_uv returned all-zero arrays of length num_pairs, which pass the
    # `len(s_uv_star) < 4` test at :602 => e = 0 => twist = 0 (a reference quirk, kept).
    out["twist"] = ibvs_twist(s_uv, s_uv_star, depth_u16, cfg)
    return out
