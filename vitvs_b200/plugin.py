"""Host-side mirror of the reference's feature-method plugin surface and control-law API for the
`--method dino` path, backed by the CUDA library.  Same names, argument meaning, return
conventions and error behaviour as the reference, so parity tests read like calls into the reference:

  * ``B200ViTExtractor``            <-> ``ViTExtractor``              (dinov2_extractor.py:14-337)
  * ``find_correspondences_batch``  <-> ``find_correspondences_batch`` (vitvs_v2.py:72-155)
  * ``DinoServoMixin``              <-> ``Controller.detect_features / calculate_uv / get_depth /
                                        transform_to_real_world / calculate_interaction_matrix / ibvs /
                                        update_ema``                   (vitvs_v2.py:329-343, 464-659)

Everything numerical runs in hand-written CUDA; the Python here only moves pointers and keeps the
reference's return shapes.  No CPU fallback: without the library / a B200 construction raises.
"""
from __future__ import annotations

import math
from typing import Optional

import numpy as np
import torch

from .config import ServoConfig
from .engine import Engine

_default_engine: Optional[Engine] = None


def set_default_engine(engine: Engine) -> None:
    """Engine the module-level ``find_correspondences_batch`` uses (the reference's function is a free
    function with no state; ours needs a handle that owns the workspace)."""
    global _default_engine
    _default_engine = engine


class B200ViTExtractor:
    """Drop-in for ``ViTExtractor(model_type, stride, model, device)`` (dinov2_extractor.py:25).
    ``model`` supplies the weights (an nn.Module or a state_dict with upstream DINOv2 names)."""

    def __init__(self, model_type: str = "dinov2_vits14", stride: int = 14, model=None, device="cuda",
                 input_size: int = 308, max_images: int = 2, cfg: ServoConfig | None = None):
        if model is None:
            raise ValueError("pass model= (nn.Module or state_dict): torch.hub is unreachable offline "
                             "(dinov2_extractor.py:65-67)")
        self.model_type = model_type
        self.device = device
        cfg = cfg or ServoConfig(model_type=model_type, stride=stride, dino_input_size=input_size)
        self.engine = Engine(cfg, max_images=max_images)
        self.engine.load_weights(model)
        self.p = 14                                            # dinov2_extractor.py:44-46
        self.stride = (stride, stride)                         # dinov2_extractor.py:47
        self.mean = (0.485, 0.456, 0.406)                      # dinov2_extractor.py:49
        self.std = (0.229, 0.224, 0.225)                       # dinov2_extractor.py:50
        self.load_size = None
        self.num_patches = None

    def preprocess_pil(self, pil_image) -> torch.Tensor:
        """dinov2_extractor.py:177-191 (ToTensor + Normalize) -> [1,3,H,W] fp32 on the host.  Kept for
        API compatibility; the fused path takes the uint8 image and normalises on the GPU instead."""
        x = torch.from_numpy(np.array(pil_image, copy=True)).permute(2, 0, 1).float().div(255.0)
        mean = torch.tensor(self.mean).view(3, 1, 1)
        std = torch.tensor(self.std).view(3, 1, 1)
        return ((x - mean) / std)[None]

    def extract_descriptors(self, batch: torch.Tensor, layer: int = 11, facet: str = "key", bin: bool = False,
                            include_cls: bool = False) -> torch.Tensor:
        """dinov2_extractor.py:313-337.  batch: [B,3,H,W] fp32 (or uint8 [B,H,W,3]) -> [B,1,T,d'] fp32 (cuda)."""
        assert facet in ["key", "query", "value", "token"], f"{facet} is not a supported facet for descriptors."
        if include_cls:
            assert not bin, "bin = True and include_cls = True are not supported together, set one of them False."
        batch = batch.to(self.engine.device)
        if batch.dtype == torch.uint8:
            H, W = batch.shape[1:3]
        else:
            H, W = batch.shape[2:4]
        self.load_size = (H, W)                                                        # dinov2_extractor.py:261
        self.num_patches = (1 + (H - self.p) // self.stride[0], 1 + (W - self.p) // self.stride[1])  # :262
        if facet != "token":                                                           # :203-218
            self.engine.forward_facet(batch, layer)
            return self.engine.facet_descriptors(facet, bin, include_cls)
        self.engine.forward_tokens(batch, layer)
        if include_cls:
            return self.engine.tokens_out().unsqueeze(1)                               # :328-334 with the CLS row kept
        return self.engine.descriptors(bin)


def _to_cartesian(coords: torch.Tensor, shape):
    """vitvs_v2.py:58-70."""
    coords = coords.long()
    width = shape[1]
    return torch.stack([coords // width, coords % width], dim=-1)


def find_correspondences_batch(descriptors1, descriptors2, num_pairs=18, distance_threshold=1, engine=None):
    """vitvs_v2.py:72-155 with the same signature and returns: ``(points1[K,2], points2[K,2], sim[1,K])``
    (row, col) int64 patch coordinates, or ``(None, None, None)`` when nothing passes the filter.
    Similarity, arg-max, same-image test and the cyclic-distance candidate rule run on the GPU; like the
    reference the K pairs are then a ``torch.randperm`` draw from the candidates (:139-141)."""
    eng = engine or _default_engine
    if eng is None:
        raise RuntimeError("no engine: pass engine= or call set_default_engine()")
    if distance_threshold != 1:
        raise NotImplementedError("distance_threshold is hard-wired to the reference's default of 1")
    B, _, t_m_1, d_h = descriptors1.size()
    if B != 1:
        raise NotImplementedError("the reference's function is effectively B=1 (SURVEY a10); use Engine.match")
    num_patches = (int(np.sqrt(t_m_1)), int(np.sqrt(t_m_1)))                       # :75
    r = eng.match(descriptors1[:, 0], descriptors2[:, 0], grid_w=num_patches[1], num_pairs=min(num_pairs, 256))
    dev = r["nn1"].device
    if int(r["flags"][0].item()) & 1:                                              # :84-101 same image
        num_points = min(num_pairs, t_m_1)
        indices = torch.randperm(t_m_1, device=dev)[:num_points]
        points1 = _to_cartesian(indices, num_patches)
        return points1, points1.clone(), torch.ones(num_points, device=dev)
    filtered_points = torch.nonzero(r["candidate"][0]).flatten()                   # :130-131 (as a set)
    num_available = filtered_points.numel()
    num_to_select = min(num_pairs, num_available)
    if num_to_select <= 0:
        return None, None, None                                                     # :155
    perm = torch.randperm(num_available, device=dev)                                # :139
    sel1 = filtered_points[perm[:num_to_select]].unsqueeze(0)
    nn_1 = r["nn1"].long()
    sel2 = torch.gather(nn_1, dim=-1, index=sel1)                                   # :144
    sim = torch.gather(r["sim1"], dim=-1, index=sel1)                               # :147
    return _to_cartesian(sel1[0], num_patches), _to_cartesian(sel2[0], num_patches), sim


class DinoServoMixin:
    """Methods with the reference Controller's names and contracts, for a class that provides
    ``self.cfg`` (ServoConfig), ``self.engine``, ``self.goal_image`` / ``self.latest_pil_image`` (PIL) and
    ``self.latest_image_depth`` (uint16 HxW array).  ``ibvs()`` sets ``self.v_c`` like the reference."""

    feature_failure_count = 0

    def initialize_ema(self):                                                        # vitvs_v2.py:325-327
        self.ema_velocities = [None] * 6

    def update_ema(self, index, new_value):                                          # vitvs_v2.py:329-343
        if self.ema_velocities[index] is None:
            self.ema_velocities[index] = new_value
        else:
            a = self.cfg.ema_alpha
            self.ema_velocities[index] = a * new_value + (1 - a) * self.ema_velocities[index]
        return self.ema_velocities[index]

    def _camera_frame_u8(self, pil_image) -> np.ndarray:
        """RGB bytes of a PIL frame in the camera geometry the engine was built for (u_max x v_max)."""
        if pil_image.size != (self.cfg.u_max, self.cfg.v_max):
            raise ValueError(f"frame size {pil_image.size} != configured camera ({self.cfg.u_max}, {self.cfg.v_max})")
        return np.array(pil_image.convert("RGB"), copy=True)

    def detect_features(self):
        """vitvs_v2.py:464-523: ``((s_uv_star, s_uv), sim)`` or ``(None, None)``; RuntimeError after 10
        consecutive failures (:500-508).  One fused GPU call: PIL-exact resize of both frames (:474-475),
        forward x2, descriptors, matching, keypoints."""
        if getattr(self, "latest_pil_image", None) is None:
            return None, None
        dev = self.engine.device
        goal = torch.from_numpy(self._camera_frame_u8(self.goal_image))[None].to(dev)
        cur = torch.from_numpy(self._camera_frame_u8(self.latest_pil_image))[None].to(dev)
        # selection: the reference draws num_pairs of the mutual nearest neighbours at random (vitvs_v2.py:139-141);
        # ``random_selection = True`` on the controller reproduces that with a host randperm handed to the fused call,
        # the default takes the num_pairs most similar ones (deterministic)
        order = torch.randperm(self.engine.tokens) if getattr(self, "random_selection", False) else None
        out = self.engine.frames_to_twist(goal, cur, self._depth_on_device(), want_match=True, sel_order=order)
        self._last_fused = out
        n = int(out["n_selected"][0].item())
        if n == 0:
            self.feature_failure_count += 1
            if self.feature_failure_count >= 10:
                raise RuntimeError("Persistent feature detection failure")
            return None, None
        self.feature_failure_count = 0
        s_uv_star = out["uv_star"][0].cpu().numpy().astype(int)
        s_uv = out["uv"][0].cpu().numpy().astype(int)
        return (s_uv_star, s_uv), out["sel_sim"][:1, :n]

    def calculate_uv(self, goal_features, current_features):
        """vitvs_v2.py:525-553 (host restatement for callers that bring their own points)."""
        K = self.cfg.num_pairs
        g = np.flip(np.asarray(goal_features, dtype=float), 1)
        c = np.flip(np.asarray(current_features, dtype=float), 1)
        s_uv = np.zeros([K, 2], dtype=int)
        s_uv_star = np.zeros([K, 2], dtype=int)
        n = len(goal_features)
        if n != K and n < 4:
            return s_uv_star, s_uv
        sx, sy = self.cfg.u_max / self.cfg.dino_input_size, self.cfg.v_max / self.cfg.dino_input_size
        for i in range(min(n, K)):
            s_uv_star[i] = (round(g[i, 0] * sx), round(g[i, 1] * sy))
            s_uv[i] = (round(c[i, 0] * sx), round(c[i, 1] * sy))
        return s_uv_star, s_uv

    def _depth_on_device(self) -> torch.Tensor:
        """The current depth frame on the GPU, uploaded once per frame object (the camera callback REPLACES
        ``latest_image_depth`` with a new array per message, vitvs_v2.py:460-462, so identity tells frames apart)."""
        src = self.latest_image_depth
        cached = getattr(self, "_depth_cache", None)
        if cached is None or cached[0] is not src:
            dev = torch.from_numpy(np.ascontiguousarray(src))[None].to(self.engine.device)
            self._depth_cache = cached = (src, dev)
        return cached[1]

    def get_depth(self, current_points):
        """vitvs_v2.py:566-586 -> Z [K,1] metres (0 / out of bounds -> 100), evaluated by the IBVS kernel."""
        if self.latest_image_depth is None:
            return None
        uv = torch.from_numpy(np.asarray(current_points, dtype=np.int32))[None].to(self.engine.device)
        return self.engine.ibvs_terms(uv, uv, self._depth_on_device())["Z"][0].cpu().numpy().reshape(-1, 1)

    def transform_to_real_world(self, s_uv, s_uv_star):
        """vitvs_v2.py:634-648 -> (s_xy, s_star_xy) normalised image coordinates (no depth frame involved)."""
        xy, xys = self.engine.transform_to_real_world(torch.from_numpy(np.asarray(s_uv, dtype=np.int32)),
                                                      torch.from_numpy(np.asarray(s_uv_star, dtype=np.int32)))
        return xy.cpu().numpy(), xys.cpu().numpy()

    def calculate_interaction_matrix(self, s_xy, Z):
        """vitvs_v2.py:650-659 -> L [2K,6] from normalised coordinates and depths, the reference's two arguments."""
        L = self.engine.interaction_matrix(torch.from_numpy(np.asarray(s_xy, dtype=np.float64)),
                                           torch.from_numpy(np.asarray(Z, dtype=np.float64)))
        return L.cpu().numpy()

    def ibvs(self):
        """vitvs_v2.py:588-623: sets ``self.v_c`` (EMA-smoothed twist) or returns early on failure."""
        if getattr(self, "latest_pil_image", None) is None:
            return
        result = self.detect_features()
        if result is None or result[0] is None:
            return
        (s_uv_star, s_uv), _sim = result
        if s_uv_star is None or s_uv is None or len(s_uv_star) < 4:
            return
        v_c = self._last_fused["twist"][0].cpu().numpy()          # -lambda * pinv(L) @ e from the fused call
        self.v_c = np.array([self.update_ema(i, v) for i, v in enumerate(v_c.flatten())])
        # velocity history read by is_visual_servoing_done (vitvs_v2.py:373-374, :626-628)
        if not hasattr(self, "velocity_vector_history"):
            self.velocity_vector_history = []
        self.velocity_vector_history.append(self.v_c)
        if len(self.velocity_vector_history) > getattr(self, "max_velocity_vector_history",
                                                        self.cfg.max_velocity_vector_history):
            self.velocity_vector_history.pop(0)


class DinoServo(DinoServoMixin):
    """Minimal concrete holder (no ROS): what Controller.__init__ sets up for the hot path."""

    def __init__(self, cfg: ServoConfig, model, goal_image, max_images: int = 2):
        self.cfg = cfg
        self.engine = Engine(cfg, max_images=max_images)
        self.engine.load_weights(model)
        self.goal_image = goal_image
        self.latest_pil_image = None
        self.latest_image_depth = None
        self.v_c = None
        self.velocity_vector_history = []
        self.max_velocity_vector_history = 200
        self.initialize_ema()
