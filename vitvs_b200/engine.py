"""Host-side owner of one libvitvs_b200 handle (one per GPU).  PyTorch is plumbing only: device
memory, streams and `tensor.data_ptr()`; every computation on the hot path happens inside the
hand-written CUDA library.  There is no fallback - construction raises when the library or an
sm_100 device is missing."""
from __future__ import annotations

import ctypes as C
import math
from typing import Mapping

import torch
import torch.nn.functional as F

from . import _lib
from ._lib import VvsError, check
from .config import ARCHS, ServoConfig

PATCH = 14


def _ptr(t: torch.Tensor | None):
    return None if t is None else C.c_void_p(t.data_ptr())


def _stream(device) -> C.c_void_p:
    return C.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def interpolate_pos_embed(pos_embed: torch.Tensor, grid: int) -> torch.Tensor:
    """pos_embed [1, 1+M*M, D] -> [1+grid*grid, D] fp32.  Same maths as the reference's statement of the
    DINO interpolation (dinov2_extractor.py:94-120; bicubic, +0.1 offset, antialias off).  Depends only on
    the weights and the geometry, so it runs once per (weights, image_size) on the host, not per frame."""
    pos = pos_embed.detach().float().cpu()
    n = pos.shape[1] - 1
    m = int(math.sqrt(n))
    if grid * grid == n:
        return pos[0].contiguous()
    dim = pos.shape[-1]
    s = float(grid + 0.1) / m
    patch = F.interpolate(pos[:, 1:].reshape(1, m, m, dim).permute(0, 3, 1, 2), scale_factor=(s, s),
                          mode="bicubic", antialias=False)
    assert patch.shape[-2:] == (grid, grid)
    patch = patch.permute(0, 2, 3, 1).reshape(-1, dim)
    return torch.cat((pos[0, :1], patch), dim=0).contiguous()


class Engine:
    def __init__(self, cfg: ServoConfig | None = None, max_images: int = 2, device: int | str | None = None,
                 max_match_tokens: int = 0, max_match_dim: int = 0):
        self.cfg = cfg or ServoConfig()
        self.lib = _lib.load()                      # raises VvsError when the .so is missing
        if not torch.cuda.is_available():
            raise VvsError("vitvs_b200 needs a CUDA device (B200, sm_100a); there is no CPU fallback")
        self.device = torch.device("cuda", torch.cuda.current_device() if device is None else
                                   (device if isinstance(device, int) else torch.device(device).index or 0))
        if self.cfg.model_type not in ARCHS:
            raise ValueError(f"unknown model_type {self.cfg.model_type}")
        if PATCH % self.cfg.stride != 0 or (self.cfg.dino_input_size - PATCH) % self.cfg.stride != 0:
            raise ValueError(f"stride {self.cfg.stride} should divide patch_size {PATCH} (dinov2_extractor.py:136-137) "
                             f"and dino_input_size - {PATCH}")
        if self.cfg.facet != "token":
            raise NotImplementedError("the servo path matches 'token' descriptors (vitvs_v2.py:485); key / query / "
                                      "value facets are served by forward_facet / facet_descriptors")
        dim, heads, depth, hidden, swiglu = ARCHS[self.cfg.model_type]
        self.dim, self.heads, self.depth, self.hidden, self.swiglu = dim, heads, depth, hidden, swiglu
        self.grid = 1 + (self.cfg.dino_input_size - PATCH) // self.cfg.stride      # dinov2_extractor.py:262
        self.tokens = self.grid * self.grid
        self.max_images = max_images
        c = _lib.Config(
            dim=dim, heads=heads, depth=depth, ffn_hidden=hidden, swiglu=int(swiglu),
            image_size=self.cfg.dino_input_size, max_images=max_images,
            frame_width=self.cfg.u_max, frame_height=self.cfg.v_max,
            f_x=self.cfg.f_x, f_y=self.cfg.f_y, c_x=self.cfg.c_x, c_y=self.cfg.c_y,
            lambda_=self.cfg.lambda_, ema_alpha=self.cfg.ema_alpha,
            num_pairs=self.cfg.num_pairs, layer=min(self.cfg.layer, depth - 1),
            use_feature_binning=int(self.cfg.use_feature_binning),
            same_image_threshold=self.cfg.same_image_threshold,
            max_match_tokens=max_match_tokens, max_match_dim=max_match_dim, patch_stride=self.cfg.stride,
            damping_mu=self.cfg.damping_mu)
        self._h = C.c_void_p()
        with torch.cuda.device(self.device):
            rc = self.lib.vvs_create(C.byref(c), C.byref(self._h))
        if rc != 0:
            msg = self.lib.vvs_last_error(None)
            raise VvsError(f"vvs_create failed ({rc}): {msg.decode() if msg else '?'}")
        self.num_sms = self.lib.vvs_num_sms(self._h)

    # ------------------------------------------------------------------ lifetime
    def close(self):
        if getattr(self, "_h", None) is not None and self._h.value:
            self.lib.vvs_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc, what):
        check(self._h, rc, what)

    # ------------------------------------------------------------------ weights
    def load_weights(self, state: Mapping[str, torch.Tensor] | torch.nn.Module):
        """Upload a DINOv2 state_dict (upstream parameter names).  Blocks deeper than cfg.layer are not
        needed: the forward stops after blocks[layer] (identical values)."""
        if isinstance(state, torch.nn.Module):
            state = state.state_dict()
        last = min(self.cfg.layer, self.depth - 1)
        s = _stream(self.device)

        def put(name, t):
            t = t.detach().to(self.device, torch.float32).contiguous()
            shape = (C.c_int64 * t.dim())(*t.shape)
            self._check(self.lib.vvs_set_weight(self._h, name.encode(), _ptr(t), _lib.F32, shape, t.dim(), s),
                        f"vvs_set_weight({name})")

        put("patch_embed.proj.weight", state["patch_embed.proj.weight"])
        put("patch_embed.proj.bias", state["patch_embed.proj.bias"])
        put("cls_token", state["cls_token"].reshape(-1))
        put("pos_embed_interp", interpolate_pos_embed(state["pos_embed"], self.grid))
        for name, t in state.items():
            if not name.startswith("blocks."):
                continue
            idx = int(name.split(".")[1])
            if idx <= last:
                put(name, t)
        torch.cuda.synchronize(self.device)
        self._check(self.lib.vvs_weights_ready(self._h), "vvs_weights_ready")

    # ------------------------------------------------------------------ stages
    def forward_tokens(self, images: torch.Tensor, layer: int | None = None) -> None:
        """images: uint8 [n,S,S,3] or float32 [n,3,S,S] on this device."""
        layer = min(self.cfg.layer, self.depth - 1) if layer is None else layer
        S = self.cfg.dino_input_size
        if images.dtype == torch.uint8:
            assert images.shape[1:] == (S, S, 3), images.shape
            kind = _lib.IMAGE_U8_HWC
        elif images.dtype == torch.float32:
            assert images.shape[1:] == (3, S, S), images.shape
            kind = _lib.IMAGE_F32_NCHW
        else:
            raise TypeError("images must be uint8 HWC or float32 NCHW")
        images = images.contiguous()
        self._check(self.lib.vvs_forward_tokens(self._h, _ptr(images), kind, images.shape[0], layer,
                                                _stream(self.device)), "vvs_forward_tokens")
        self._n_img = images.shape[0]

    def forward_facet(self, images: torch.Tensor, layer: int | None = None) -> None:
        """blocks[0..layer-1] + norm1 + qkv of blocks[layer]: what the key / query / value hooks see
        (dinov2_extractor.py:203-218).  images as in forward_tokens."""
        layer = min(self.cfg.layer, self.depth - 1) if layer is None else layer
        S = self.cfg.dino_input_size
        if images.dtype == torch.uint8:
            assert images.shape[1:] == (S, S, 3), images.shape
            kind = _lib.IMAGE_U8_HWC
        elif images.dtype == torch.float32:
            assert images.shape[1:] == (3, S, S), images.shape
            kind = _lib.IMAGE_F32_NCHW
        else:
            raise TypeError("images must be uint8 HWC or float32 NCHW")
        images = images.contiguous()
        self._check(self.lib.vvs_forward_facet(self._h, _ptr(images), kind, images.shape[0], layer,
                                               _stream(self.device)), "vvs_forward_facet")
        self._n_facet = images.shape[0]

    def facet_descriptors(self, facet: str, bin: bool = False, include_cls: bool = False) -> torch.Tensor:
        """[n, 1, T (+1), d'] fp32 descriptors of the last forward_facet pass; facet in {'query','key','value'}."""
        idx = {"query": 0, "key": 1, "value": 2}[facet]
        n = self._n_facet
        out = torch.empty(n, 1, self.tokens + int(include_cls), self.dim * (9 if bin else 1), device=self.device,
                          dtype=torch.float32)
        self._check(self.lib.vvs_facet_descriptors(self._h, n, idx, int(bin), int(include_cls), _ptr(out),
                                                   _stream(self.device)), "vvs_facet_descriptors")
        return out

    def tokens_out(self) -> torch.Tensor:
        out = torch.empty(self._n_img, self.tokens + 1, self.dim, device=self.device, dtype=torch.float32)
        self._check(self.lib.vvs_get_tokens(self._h, _ptr(out), self._n_img, _stream(self.device)), "vvs_get_tokens")
        return out

    def descriptors(self, bin: bool) -> torch.Tensor:
        d = self.dim * (9 if bin else 1)
        out = torch.empty(self._n_img, 1, self.tokens, d, device=self.device, dtype=torch.float32)
        self._check(self.lib.vvs_descriptors(self._h, self._n_img, int(bin), _ptr(out), _stream(self.device)),
                    "vvs_descriptors")
        return out

    def descriptors_hierarchy(self, hierarchy: int) -> torch.Tensor:
        """``_log_bin(x, hierarchy)`` (dinov2_extractor.py:265-311) of the last forward_tokens pass:
        [n, 1, T, (1 + 8 * hierarchy) * dim] fp32."""
        out = torch.empty(self._n_img, 1, self.tokens, self.dim * (1 + 8 * hierarchy), device=self.device,
                          dtype=torch.float32)
        self._check(self.lib.vvs_descriptors_hierarchy(self._h, self._n_img, int(hierarchy), _ptr(out),
                                                       _stream(self.device)), "vvs_descriptors_hierarchy")
        return out

    def _match_buffers(self, n_pairs: int, T: int, K: int, sel_order: torch.Tensor | None = None):
        dev = self.device
        bufs = dict(
            nn1=torch.empty(n_pairs, T, dtype=torch.int32, device=dev),
            nn2=torch.empty(n_pairs, T, dtype=torch.int32, device=dev),
            sim1=torch.empty(n_pairs, T, dtype=torch.float32, device=dev),
            candidate=torch.empty(n_pairs, T, dtype=torch.uint8, device=dev),
            mean_sim1=torch.empty(n_pairs, dtype=torch.float32, device=dev),
            n_candidates=torch.empty(n_pairs, dtype=torch.int32, device=dev),
            n_selected=torch.empty(n_pairs, dtype=torch.int32, device=dev),
            flags=torch.empty(n_pairs, dtype=torch.int32, device=dev),
            sel_tok=torch.empty(n_pairs, K, dtype=torch.int32, device=dev),
            sel_sim=torch.empty(n_pairs, K, dtype=torch.float32, device=dev),
            uv_star=torch.empty(n_pairs, K, 2, dtype=torch.int32, device=dev),
            uv=torch.empty(n_pairs, K, 2, dtype=torch.int32, device=dev),
        )
        mo = _lib.MatchOut(**{k: v.data_ptr() for k, v in bufs.items()})
        if sel_order is not None:
            # reference-faithful selection (vitvs_v2.py:139-141): a priority list of goal tokens per pair, e.g. randperm(T)
            order = sel_order.to(dev, torch.int32).contiguous()
            if order.dim() == 1:
                order = order[None].expand(n_pairs, -1).contiguous()
            assert order.shape == (n_pairs, T), order.shape
            bufs["_sel_order"] = order                # keeps the buffer alive until the caller drops the dict
            mo.sel_order = order.data_ptr()
        return bufs, mo

    def match(self, desc1: torch.Tensor, desc2: torch.Tensor, grid_w: int | None = None,
              num_pairs: int | None = None, sel_order: torch.Tensor | None = None) -> dict:
        """desc1/desc2: fp32 [n_pairs, T, d] (goal / current).  Returns the dict of vvs_match_out tensors.
        sel_order (optional, int [T] or [n_pairs, T]): priority list of goal tokens - the first num_pairs candidates
        in that order are selected instead of the num_pairs most similar ones."""
        assert desc1.shape == desc2.shape and desc1.dim() == 3
        n_pairs, T, d = desc1.shape
        K = self.cfg.num_pairs if num_pairs is None else num_pairs
        grid_w = int(math.sqrt(T)) if grid_w is None else grid_w          # vitvs_v2.py:75
        d1 = desc1.to(self.device, torch.float32).contiguous()
        d2 = desc2.to(self.device, torch.float32).contiguous()
        bufs, mo = self._match_buffers(n_pairs, T, K, sel_order)
        self._check(self.lib.vvs_match(self._h, _ptr(d1), _ptr(d2), n_pairs, T, d, grid_w, K, C.byref(mo),
                                       _stream(self.device)), "vvs_match")
        return bufs

    def similarity(self, desc1: torch.Tensor, desc2: torch.Tensor) -> torch.Tensor:
        n_pairs, T, d = desc1.shape
        d1 = desc1.to(self.device, torch.float32).contiguous()
        d2 = desc2.to(self.device, torch.float32).contiguous()
        out = torch.empty(n_pairs, T, T, dtype=torch.float32, device=self.device)
        self._check(self.lib.vvs_similarity(self._h, _ptr(d1), _ptr(d2), n_pairs, T, d, _ptr(out),
                                            _stream(self.device)), "vvs_similarity")
        return out

    def ibvs(self, uv: torch.Tensor, uv_star: torch.Tensor, depth: torch.Tensor,
             ema_state: torch.Tensor | None = None, ema_init: torch.Tensor | None = None):
        """uv/uv_star int32 [n,K,2]; depth uint16 [n,H,W] -> (twist [n,6] f64, twist_raw [n,6] f64)."""
        n, K, _ = uv.shape
        uv = uv.to(self.device, torch.int32).contiguous()
        uv_star = uv_star.to(self.device, torch.int32).contiguous()
        depth = depth.to(self.device).contiguous()
        assert depth.dtype == torch.uint16 and depth.shape == (n, self.cfg.v_max, self.cfg.u_max)
        twist = torch.empty(n, 6, dtype=torch.float64, device=self.device)
        raw = torch.empty(n, 6, dtype=torch.float64, device=self.device)
        self._check(self.lib.vvs_ibvs(self._h, _ptr(uv), _ptr(uv_star), _ptr(depth), n, K, _ptr(twist), _ptr(raw),
                                      _ptr(ema_state), _ptr(ema_init), _stream(self.device)), "vvs_ibvs")
        return twist, raw

    def ibvs_terms(self, uv: torch.Tensor, uv_star: torch.Tensor, depth: torch.Tensor) -> dict:
        """Per-point terms of the control law: Z [n,K], s_xy / s_star_xy [n,K,2], L [n,2K,6], twist [n,6]."""
        n, K, _ = uv.shape
        uv = uv.to(self.device, torch.int32).contiguous()
        uv_star = uv_star.to(self.device, torch.int32).contiguous()
        depth = depth.to(self.device).contiguous()
        f64 = dict(dtype=torch.float64, device=self.device)
        out = dict(twist=torch.empty(n, 6, **f64), Z=torch.empty(n, K, **f64), s_xy=torch.empty(n, K, 2, **f64),
                   s_star_xy=torch.empty(n, K, 2, **f64), L=torch.empty(n, 2 * K, 6, **f64))
        self._check(self.lib.vvs_ibvs_terms(self._h, _ptr(uv), _ptr(uv_star), _ptr(depth), n, K, _ptr(out["twist"]),
                                            _ptr(out["Z"]), _ptr(out["s_xy"]), _ptr(out["s_star_xy"]), _ptr(out["L"]),
                                            _stream(self.device)), "vvs_ibvs_terms")
        return out

    def frame_pairs_to_twist(self, goal: torch.Tensor, current: torch.Tensor, depth: torch.Tensor,
                             want_match: bool = False, sel_order: torch.Tensor | None = None) -> dict:
        """Device-resident batched Controller.ibvs(): goal/current uint8 [B,S,S,3], depth uint16 [B,H,W].
        sel_order: see match() - e.g. torch.randperm(T) for the reference's random choice of key-points."""
        B = goal.shape[0]
        K = self.cfg.num_pairs
        goal, current, depth = goal.contiguous(), current.contiguous(), depth.contiguous()
        out = dict(twist=torch.empty(B, 6, dtype=torch.float64, device=self.device),
                   n_selected=torch.empty(B, dtype=torch.int32, device=self.device),
                   mean_sim_sel=torch.empty(B, dtype=torch.float32, device=self.device))
        po = _lib.PairOut(twist=out["twist"].data_ptr(), n_selected=out["n_selected"].data_ptr(),
                          mean_sim_sel=out["mean_sim_sel"].data_ptr())
        if want_match or sel_order is not None:
            bufs, mo = self._match_buffers(B, self.tokens, K, sel_order)
            po.match = mo
            out.update(bufs)
        self._check(self.lib.vvs_frame_pairs_to_twist(self._h, _ptr(goal), _ptr(current), _ptr(depth), B,
                                                      C.byref(po), _stream(self.device)),
                    "vvs_frame_pairs_to_twist")
        self._n_img = 2 * B
        return out

    def transform_to_real_world(self, uv: torch.Tensor, uv_star: torch.Tensor):
        """Controller.transform_to_real_world (vitvs_v2.py:634-648): int [..., 2] pixel key-points -> fp64 normalised
        image coordinates (s_xy, s_star_xy), same leading shape.  No depth frame involved."""
        uv = uv.to(self.device, torch.int32).contiguous()
        uv_star = uv_star.to(self.device, torch.int32).contiguous()
        assert uv.shape == uv_star.shape and uv.shape[-1] == 2
        xy = torch.empty(uv.shape, dtype=torch.float64, device=self.device)
        xys = torch.empty(uv.shape, dtype=torch.float64, device=self.device)
        self._check(self.lib.vvs_transform_to_real_world(self._h, _ptr(uv), _ptr(uv_star), uv.numel() // 2, _ptr(xy),
                                                         _ptr(xys), _stream(self.device)), "vvs_transform_to_real_world")
        return xy, xys

    def interaction_matrix(self, s_xy: torch.Tensor, Z: torch.Tensor) -> torch.Tensor:
        """Controller.calculate_interaction_matrix(s_xy, Z) (vitvs_v2.py:650-659): s_xy fp64 [K, 2], Z fp64 [K] or
        [K, 1] -> L fp64 [2K, 6]."""
        xy = s_xy.to(self.device, torch.float64).contiguous()
        z = Z.to(self.device, torch.float64).reshape(-1).contiguous()
        assert xy.dim() == 2 and xy.shape[1] == 2 and z.numel() == xy.shape[0]
        L = torch.empty(2 * xy.shape[0], 6, dtype=torch.float64, device=self.device)
        self._check(self.lib.vvs_interaction_matrix(self._h, _ptr(xy), _ptr(z), xy.shape[0], _ptr(L),
                                                    _stream(self.device)), "vvs_interaction_matrix")
        return L

    @property
    def build_id(self) -> str:
        return self.lib.vvs_build_id().decode()

    def frame_pairs_to_twist_host(self, goal: torch.Tensor, current: torch.Tensor, depth: torch.Tensor,
                                  twist_out: torch.Tensor, n_selected_out: torch.Tensor | None = None,
                                  mean_sim_out: torch.Tensor | None = None, wait: bool = True) -> None:
        """Same through HOST buffers (pinned for async copies); synchronises the stream before returning.
        wait=False: returns once enqueued - results are valid after self.wait(); the next call's copies then run under
        this call's compute (two landing zones).  At most two calls in flight; their host buffers must stay untouched."""
        assert not goal.is_cuda and not current.is_cuda and not depth.is_cuda and not twist_out.is_cuda
        B = goal.shape[0]
        fn = self.lib.vvs_frame_pairs_to_twist_host if wait else self.lib.vvs_frame_pairs_to_twist_host_async
        self._check(fn(self._h, _ptr(goal), _ptr(current), _ptr(depth), B, _ptr(twist_out), _ptr(n_selected_out),
                       _ptr(mean_sim_out), _stream(self.device)), "vvs_frame_pairs_to_twist_host")
        self._n_img = 2 * B

    def wait(self, calls_back: int = 0) -> None:
        """Block until the latest (0) / the one-before-latest (1) wait=False host call has delivered its results."""
        self._check(self.lib.vvs_wait(self._h, calls_back), "vvs_wait")

    # ------------------------------------------------------------------ camera frames (resize on the GPU)
    def resize_frames(self, frames: torch.Tensor) -> torch.Tensor:
        """PIL ``Image.resize((S, S))`` (vitvs_v2.py:474-475) of uint8 [n, v_max, u_max, 3] frames on this
        device -> uint8 [n, S, S, 3], bit-identical to Pillow's antialiased bicubic."""
        S = self.cfg.dino_input_size
        assert frames.dtype == torch.uint8 and frames.shape[1:] == (self.cfg.v_max, self.cfg.u_max, 3), frames.shape
        frames = frames.to(self.device).contiguous()
        out = torch.empty(frames.shape[0], S, S, 3, dtype=torch.uint8, device=self.device)
        self._check(self.lib.vvs_resize_frames(self._h, _ptr(frames), frames.shape[0], _ptr(out), _stream(self.device)),
                    "vvs_resize_frames")
        return out

    def frames_to_twist(self, goal_frames: torch.Tensor, current_frames: torch.Tensor, depth: torch.Tensor,
                        want_match: bool = False, sel_order: torch.Tensor | None = None) -> dict:
        """frame_pairs_to_twist on raw camera frames: uint8 [B, v_max, u_max, 3] x2 + depth uint16 [B, v_max, u_max]."""
        B = goal_frames.shape[0]
        K = self.cfg.num_pairs
        goal_frames, current_frames, depth = goal_frames.contiguous(), current_frames.contiguous(), depth.contiguous()
        assert goal_frames.shape[1:] == (self.cfg.v_max, self.cfg.u_max, 3) and goal_frames.shape == current_frames.shape
        out = dict(twist=torch.empty(B, 6, dtype=torch.float64, device=self.device),
                   n_selected=torch.empty(B, dtype=torch.int32, device=self.device),
                   mean_sim_sel=torch.empty(B, dtype=torch.float32, device=self.device))
        po = _lib.PairOut(twist=out["twist"].data_ptr(), n_selected=out["n_selected"].data_ptr(),
                          mean_sim_sel=out["mean_sim_sel"].data_ptr())
        if want_match or sel_order is not None:
            bufs, mo = self._match_buffers(B, self.tokens, K, sel_order)
            po.match = mo
            out.update(bufs)
        self._check(self.lib.vvs_frames_to_twist(self._h, _ptr(goal_frames), _ptr(current_frames), _ptr(depth), B,
                                                 C.byref(po), _stream(self.device)), "vvs_frames_to_twist")
        self._n_img = 2 * B
        return out

    # ------------------------------------------------------------------ cached-goal mode (an extra, SURVEY 7 design note)
    def _is_frames(self, t: torch.Tensor) -> bool:
        S = self.cfg.dino_input_size
        if tuple(t.shape[1:]) == (self.cfg.v_max, self.cfg.u_max, 3):
            return True
        assert tuple(t.shape[1:]) == (S, S, 3), "expected [n, S, S, 3] images or [n, v_max, u_max, 3] camera frames"
        return False

    def set_goal(self, goal: torch.Tensor) -> None:
        """Run the goal views (uint8 [n, S, S, 3] images or [n, v_max, u_max, 3] camera frames, on this device) through
        the backbone once and keep their descriptors: the reference recomputes them on every iteration
        (vitvs_v2.py:482-487) although the goal never changes inside an episode."""
        goal = goal.contiguous()
        self._check(self.lib.vvs_set_goal(self._h, _ptr(goal), goal.shape[0], int(self._is_frames(goal)),
                                          _stream(self.device)), "vvs_set_goal")
        self._n_img = 0

    def current_to_twist(self, current: torch.Tensor, depth: torch.Tensor, want_match: bool = False,
                         sel_order: torch.Tensor | None = None) -> dict:
        """frame_pairs_to_twist / frames_to_twist against the goal views of set_goal (current view i with goal i):
        only the current views are forwarded; the results are bit-identical to the pair call."""
        B = current.shape[0]
        K = self.cfg.num_pairs
        current, depth = current.contiguous(), depth.contiguous()
        out = dict(twist=torch.empty(B, 6, dtype=torch.float64, device=self.device),
                   n_selected=torch.empty(B, dtype=torch.int32, device=self.device),
                   mean_sim_sel=torch.empty(B, dtype=torch.float32, device=self.device))
        po = _lib.PairOut(twist=out["twist"].data_ptr(), n_selected=out["n_selected"].data_ptr(),
                          mean_sim_sel=out["mean_sim_sel"].data_ptr())
        if want_match or sel_order is not None:
            bufs, mo = self._match_buffers(B, self.tokens, K, sel_order)
            po.match = mo
            out.update(bufs)
        self._check(self.lib.vvs_current_to_twist(self._h, _ptr(current), _ptr(depth), B, int(self._is_frames(current)),
                                                  C.byref(po), _stream(self.device)), "vvs_current_to_twist")
        self._n_img = B
        return out

    def frames_to_twist_host(self, goal_frames: torch.Tensor, current_frames: torch.Tensor, depth: torch.Tensor,
                             twist_out: torch.Tensor, n_selected_out: torch.Tensor | None = None,
                             mean_sim_out: torch.Tensor | None = None, wait: bool = True) -> None:
        """Same through HOST buffers (pinned for async copies); synchronises before returning unless wait=False."""
        assert not goal_frames.is_cuda and not current_frames.is_cuda and not depth.is_cuda and not twist_out.is_cuda
        B = goal_frames.shape[0]
        fn = self.lib.vvs_frames_to_twist_host if wait else self.lib.vvs_frames_to_twist_host_async
        self._check(fn(self._h, _ptr(goal_frames), _ptr(current_frames), _ptr(depth), B, _ptr(twist_out),
                       _ptr(n_selected_out), _ptr(mean_sim_out), _stream(self.device)), "vvs_frames_to_twist_host")
        self._n_img = 2 * B

    # ------------------------------------------------------------------ headless planar-scene camera (N3)
    def render_plane(self, texture: torch.Tensor, cam_to_world: torch.Tensor, plane_w: float = 0.8,
                     plane_h: float = 0.6):
        """texture uint8 [th, tw, 3] and poses fp64 [n, 12] (R row-major + t, camera-to-world) on this device ->
        (rgb uint8 [n, v_max, u_max, 3], depth uint16 [n, v_max, u_max]) as the reference's Gazebo camera would."""
        assert texture.dtype == torch.uint8 and texture.dim() == 3 and texture.shape[2] == 3 and texture.is_cuda
        poses = cam_to_world.to(self.device, torch.float64).contiguous()
        assert poses.dim() == 2 and poses.shape[1] == 12
        n = poses.shape[0]
        rgb = torch.empty(n, self.cfg.v_max, self.cfg.u_max, 3, dtype=torch.uint8, device=self.device)
        depth = torch.empty(n, self.cfg.v_max, self.cfg.u_max, dtype=torch.uint16, device=self.device)
        texture = texture.contiguous()
        self._check(self.lib.vvs_render_plane(self._h, _ptr(texture), texture.shape[0], texture.shape[1], plane_w,
                                              plane_h, _ptr(poses), n, _ptr(rgb), _ptr(depth), _stream(self.device)),
                    "vvs_render_plane")
        return rgb, depth

    def profile(self, on: bool) -> None:
        self._check(self.lib.vvs_profile_enable(self._h, int(on)), "vvs_profile_enable")

    def profile_read(self) -> dict:
        """{kernel class: (total_ms, launches)} since the last read (CUDA events around every launch)."""
        n = len(_lib.KERNEL_CLASSES)
        ms = (C.c_double * n)()
        cnt = (C.c_int32 * n)()
        self._check(self.lib.vvs_profile_read(self._h, ms, cnt, n), "vvs_profile_read")
        return {k: (ms[i], cnt[i]) for i, k in enumerate(_lib.KERNEL_CLASSES) if cnt[i]}

    def trace_step(self, n_pairs: int, replays: int = 3) -> list[tuple[str, int, int]]:
        """[(kernel class, begin_ns, end_ns)] of every kernel of one traced graph replay of the frame-pair path (the
        last of `replays` back-to-back ones), on the inputs of the previous frame_pairs_to_twist call."""
        cap = 1024
        cls = (C.c_int32 * cap)()
        b = (C.c_uint64 * cap)()
        e = (C.c_uint64 * cap)()
        n = C.c_int32(0)
        self._check(self.lib.vvs_trace_step(self._h, n_pairs, replays, cls, b, e, cap, C.byref(n), _stream(self.device)),
                    "vvs_trace_step")
        return [(_lib.KERNEL_CLASSES[cls[i]], int(b[i]), int(e[i])) for i in range(n.value)]

    @property
    def gram_stencil(self) -> bool:
        """True when the fused path matches binned descriptors through the token Gram matrix (never materialised)."""
        return bool(self.lib.vvs_uses_gram_stencil(self._h))

    def set_num_pairs(self, num_pairs: int) -> None:
        """`controller.num_pairs = K` (vitvs_v2.py:1158): K of the fused entry points from now on."""
        self._check(self.lib.vvs_set_num_pairs(self._h, int(num_pairs)), "vvs_set_num_pairs")
        self.cfg.num_pairs = int(num_pairs)

    def last_launch_count(self) -> int:
        return self.lib.vvs_last_launch_count(self._h)

    # ------------------------------------------------------------------ kernel-level (tests / per-kernel bench)
    def dbg_gemm(self, A, B, epilogue, bias=None, gamma=None, x_f32=None, out_bf16=None, ldo=0, q_cols=0,
                 q_scale=1.0, pos=None, T=0):
        M, K = A.shape
        N = B.shape[0]
        self._check(self.lib.vvs_dbg_gemm(self._h, _ptr(A), _ptr(B), M, N, K, epilogue, _ptr(bias), _ptr(gamma),
                                          _ptr(x_f32), _ptr(out_bf16), ldo, q_cols, q_scale, _ptr(pos), T,
                                          _stream(self.device)), "vvs_dbg_gemm")

    def dbg_gemm_ln(self, A, B, epilogue, bias=None, gamma=None, x_f32=None, out_bf16=None, ldo=0, q_cols=0,
                    q_scale=1.0, pos=None, T=0, ln_part_out=None, xb_out=None, ln_part_in=None, colsum=None, ln_eps=1e-6):
        """GEMM with a folded-LayerNorm epilogue (csrc/vvs_internal.h EPI_*_LN_*)."""
        M, K = A.shape
        N = B.shape[0]
        self._check(self.lib.vvs_dbg_gemm_ln(self._h, _ptr(A), _ptr(B), M, N, K, epilogue, _ptr(bias), _ptr(gamma),
                                             _ptr(x_f32), _ptr(out_bf16), ldo, q_cols, q_scale, _ptr(pos), T,
                                             _ptr(ln_part_out), _ptr(xb_out), _ptr(ln_part_in), _ptr(colsum), ln_eps,
                                             _stream(self.device)), "vvs_dbg_gemm_ln")

    def dbg_attention(self, qkv, n_img, n_tok, heads):
        out = torch.empty(n_img * n_tok, heads * 64, dtype=torch.bfloat16, device=self.device)
        self._check(self.lib.vvs_dbg_attention(self._h, _ptr(qkv), _ptr(out), n_img, n_tok, heads,
                                               _stream(self.device)), "vvs_dbg_attention")
        return out

    def dbg_layernorm(self, x, w, b):
        out = torch.empty(x.shape, dtype=torch.bfloat16, device=self.device)
        self._check(self.lib.vvs_dbg_layernorm(self._h, _ptr(x), _ptr(w), _ptr(b), _ptr(out), x.shape[0], x.shape[1],
                                               _stream(self.device)), "vvs_dbg_layernorm")
        return out
