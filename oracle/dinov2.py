"""Oracle: plain-PyTorch fp32 restatement of the DINOv2 ViT backbone (S/B/L/g, patch 14).

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).  PARITY UNPINNED at this
boundary: the reference obtains the model through
``torch.hub.load('facebookresearch/dinov2', model_type)``
(/root/reference/catkin_ws/ibvs/src/dinov2_extractor.py:65-67) - third party,
not vendored, hub default branch (no pinned version), unreachable offline.
What the reference *does* vendor, and what this file follows:

  * dino_patch/attention.py:37-80   Attention: fused qkv Linear -> SDPA(scale hd^-0.5) -> proj
  * dino_patch/attention.py:83-100  MemEffAttention: xFormers absent => Attention.forward
  * dino_patch/block.py:44-88       Block members: norm1, attn, ls1, norm2, mlp, ls2
  * dino_patch/block.py:90-95,112-114  inference branch: x += ls1(attn(norm1 x)); x += ls2(mlp(norm2 x))
  * catkin_ws/ibvs/src/dinov2_extractor.py:94-120  the repo's own statement of the DINO
    pos-embed interpolation (bicubic, +0.1 offset); at stride 14 the hub model's
    own method runs (:133 early return) which is the same maths.

The rest restates the published upstream architecture (PatchEmbed = Conv2d(3,D,14,14);
cls_token concat; pos_embed [1,1+37*37,D]; LayerNorm eps 1e-6; LayerScale init 1.0;
Mlp GELU(erf) ratio 4; ViT-g SwiGLUFFNFused hidden 4096; all biases on; final norm)
and is cross-checked against transformers' Dinov2Model in tests/test_oracle_backbone.py.

The attribute surface is what the reference's ViTExtractor touches
(dinov2_extractor.py:41-47,226-235,259): ``patch_embed.patch_size``,
``patch_embed.proj.stride``, ``blocks[i]`` (forward hooks; ``blocks[i].attn.qkv``,
``.attn.num_heads``), ``interpolate_pos_encoding``, ``pos_embed``.
"""
from __future__ import annotations

import math
from dataclasses import dataclass

import torch
from torch import nn
import torch.nn.functional as F


@dataclass(frozen=True)
class Arch:
    name: str
    dim: int
    heads: int
    depth: int
    swiglu: bool

    @property
    def ffn_hidden(self) -> int:
        if self.swiglu:  # SwiGLUFFNFused: (int(4D * 2/3) + 7) // 8 * 8
            return (int(self.dim * 4 * 2 / 3) + 7) // 8 * 8
        return 4 * self.dim


ARCHS = {
    "dinov2_vits14": Arch("dinov2_vits14", 384, 6, 12, False),
    "dinov2_vitb14": Arch("dinov2_vitb14", 768, 12, 12, False),
    "dinov2_vitl14": Arch("dinov2_vitl14", 1024, 16, 24, False),
    "dinov2_vitg14": Arch("dinov2_vitg14", 1536, 24, 40, True),
}

PATCH = 14
TRAIN_GRID = 37  # pos_embed trained at 518 = 37*14


class PatchEmbed(nn.Module):
    def __init__(self, dim: int):
        super().__init__()
        self.patch_size = (PATCH, PATCH)
        self.proj = nn.Conv2d(3, dim, kernel_size=PATCH, stride=PATCH)

    def forward(self, x):
        x = self.proj(x)                      # B D Hp Wp
        return x.flatten(2).transpose(1, 2)   # B T D


class Attention(nn.Module):
    """dino_patch/attention.py:37-80 (inference: no dropout)."""

    def __init__(self, dim: int, num_heads: int):
        super().__init__()
        self.num_heads = num_heads
        self.scale = (dim // num_heads) ** -0.5
        self.qkv = nn.Linear(dim, dim * 3, bias=True)
        self.attn_drop = nn.Identity()  # hook point for facet 'attn' (dinov2_extractor.py:234)
        self.proj = nn.Linear(dim, dim, bias=True)

    def forward(self, x):
        B, N, C = x.shape
        qkv = self.qkv(x).reshape(B, N, 3, self.num_heads, C // self.num_heads)
        q, k, v = (t.transpose(1, 2) for t in torch.unbind(qkv, 2))
        o = F.scaled_dot_product_attention(q, k, v)      # scale = hd^-0.5, non causal
        return self.proj(o.transpose(1, 2).reshape(B, N, C))


class LayerScale(nn.Module):
    def __init__(self, dim: int, init_values: float = 1.0):
        super().__init__()
        self.gamma = nn.Parameter(init_values * torch.ones(dim))

    def forward(self, x):
        return x * self.gamma


class Mlp(nn.Module):
    def __init__(self, dim: int, hidden: int):
        super().__init__()
        self.fc1 = nn.Linear(dim, hidden)
        self.fc2 = nn.Linear(hidden, dim)

    def forward(self, x):
        return self.fc2(F.gelu(self.fc1(x)))   # exact (erf) GELU


class SwiGLUFFNFused(nn.Module):
    def __init__(self, dim: int, hidden: int):
        super().__init__()
        self.w12 = nn.Linear(dim, 2 * hidden)
        self.w3 = nn.Linear(hidden, dim)

    def forward(self, x):
        x1, x2 = self.w12(x).chunk(2, dim=-1)
        return self.w3(F.silu(x1) * x2)


class Block(nn.Module):
    """dino_patch/block.py:44-95,112-114 (eval branch)."""

    def __init__(self, arch: Arch):
        super().__init__()
        self.norm1 = nn.LayerNorm(arch.dim, eps=1e-6)
        self.attn = Attention(arch.dim, arch.heads)
        self.ls1 = LayerScale(arch.dim)
        self.norm2 = nn.LayerNorm(arch.dim, eps=1e-6)
        self.mlp = (SwiGLUFFNFused if arch.swiglu else Mlp)(arch.dim, arch.ffn_hidden)
        self.ls2 = LayerScale(arch.dim)

    def forward(self, x):
        x = x + self.ls1(self.attn(self.norm1(x)))
        x = x + self.ls2(self.mlp(self.norm2(x)))
        return x


class DinoV2(nn.Module):
    def __init__(self, arch: Arch):
        super().__init__()
        self.arch = arch
        self.embed_dim = arch.dim
        self.patch_size = PATCH
        self.interpolate_offset = 0.1
        self.patch_embed = PatchEmbed(arch.dim)
        self.cls_token = nn.Parameter(torch.zeros(1, 1, arch.dim))
        self.pos_embed = nn.Parameter(torch.zeros(1, 1 + TRAIN_GRID * TRAIN_GRID, arch.dim))
        self.blocks = nn.ModuleList([Block(arch) for _ in range(arch.depth)])
        self.norm = nn.LayerNorm(arch.dim, eps=1e-6)

    def interpolate_pos_encoding(self, x, w, h):
        """Same maths as dinov2_extractor.py:94-120 at stride == patch."""
        npatch = x.shape[1] - 1
        N = self.pos_embed.shape[1] - 1
        if npatch == N and w == h:
            return self.pos_embed
        pos = self.pos_embed.float()
        cls_pos, patch_pos = pos[:, 0], pos[:, 1:]
        dim = x.shape[-1]
        w0, h0 = w // PATCH, h // PATCH
        M = int(math.sqrt(N))
        sx = float(w0 + self.interpolate_offset) / M
        sy = float(h0 + self.interpolate_offset) / M
        patch_pos = F.interpolate(
            patch_pos.reshape(1, M, M, dim).permute(0, 3, 1, 2),
            scale_factor=(sx, sy), mode="bicubic", antialias=False)
        assert (w0, h0) == tuple(patch_pos.shape[-2:])
        patch_pos = patch_pos.permute(0, 2, 3, 1).reshape(1, -1, dim)
        return torch.cat((cls_pos.unsqueeze(0), patch_pos), dim=1).to(x.dtype)

    def prepare_tokens(self, x):
        B, _, w, h = x.shape
        x = self.patch_embed(x)
        x = torch.cat((self.cls_token.expand(B, -1, -1), x), dim=1)
        return x + self.interpolate_pos_encoding(x, w, h)

    def forward(self, x):
        x = self.prepare_tokens(x)
        for blk in self.blocks:
            x = blk(x)
        return self.norm(x)[:, 0]   # head = Identity on the normed CLS token


def build(model_type: str = "dinov2_vits14", seed: int = 0, rich: bool = True, max_depth: int | None = None) -> DinoV2:
    """Random-init weights of the named architecture:
    Linear/Conv trunc-normal sigma 0.02, biases 0, LN (1,0), LayerScale 1.0,
    pos_embed trunc-normal 0.02, cls_token N(0, 1e-6).  fp32, CPU, eval mode.
    ``rich=True`` (default) additionally randomises every bias (N(0,0.02)), LN
    affine (1+N(0,0.1), N(0,0.05)) and LayerScale gamma (U(0.5,1.5)) so that parity
    tests exercise every epilogue term instead of multiplying by 1 / adding 0."""
    arch = ARCHS[model_type]
    if max_depth is not None and max_depth < arch.depth:
        # the servo path hooks blocks[11] (vitvs_v2.py:484): deeper blocks never influence its output, so tests of
        # the 24 / 40-block models may instantiate only the blocks that matter (same seed => same leading weights
        # is NOT guaranteed vs the full model; each variant is self-consistent between oracle and GPU path)
        arch = Arch(arch.name, arch.dim, arch.heads, max_depth, arch.swiglu)
    g = torch.Generator().manual_seed(seed)
    m = DinoV2(arch)
    with torch.no_grad():
        for mod in m.modules():
            if isinstance(mod, (nn.Linear, nn.Conv2d)):
                nn.init.trunc_normal_(mod.weight, std=0.02, generator=g)
                nn.init.zeros_(mod.bias)
        nn.init.trunc_normal_(m.pos_embed, std=0.02, generator=g)
        m.cls_token.normal_(0.0, 1e-6, generator=g)
        if rich:
            for mod in m.modules():
                if isinstance(mod, (nn.Linear, nn.Conv2d)):
                    mod.bias.normal_(0.0, 0.02, generator=g)
                elif isinstance(mod, nn.LayerNorm):
                    mod.weight.copy_(1.0 + 0.1 * torch.randn(mod.weight.shape, generator=g))
                    mod.bias.normal_(0.0, 0.05, generator=g)
                elif isinstance(mod, LayerScale):
                    mod.gamma.copy_(0.5 + torch.rand(mod.gamma.shape, generator=g))
    return m.eval()


def with_stride(model: "DinoV2", stride: int) -> "DinoV2":
    """ViTExtractor.patch_vit_resolution + _fix_pos_enc (dinov2_extractor.py:85-144) restated on a COPY of the model:
    the patch-embedding convolution runs at ``stride`` (a divisor of 14: overlapping patches) and the positional
    encoding is interpolated to the 1 + (S - 14) // stride grid with the same bicubic call (+0.1 offset).
    PINNED: tests/test_oracle_vs_reference.py lets the reference's own ViTExtractor patch a copy and compares."""
    import copy
    import types
    m = copy.deepcopy(model)
    if stride == PATCH:
        return m
    assert (PATCH // stride) * stride == PATCH, f"stride {stride} should divide patch_size {PATCH}"
    m.patch_embed.proj.stride = (stride, stride)

    def interpolate_pos_encoding(self, x, w, h):
        npatch = x.shape[1] - 1
        N = self.pos_embed.shape[1] - 1
        if npatch == N and w == h:
            return self.pos_embed
        class_pos_embed = self.pos_embed[:, 0]
        patch_pos_embed = self.pos_embed[:, 1:]
        dim = x.shape[-1]
        w0 = 1 + (w - PATCH) // stride
        h0 = 1 + (h - PATCH) // stride
        assert w0 * h0 == npatch
        w0, h0 = w0 + 0.1, h0 + 0.1
        M = int(math.sqrt(N))
        patch_pos_embed = F.interpolate(patch_pos_embed.reshape(1, M, M, dim).permute(0, 3, 1, 2),
                                        scale_factor=(w0 / math.sqrt(N), h0 / math.sqrt(N)), mode="bicubic",
                                        align_corners=False, recompute_scale_factor=False)
        assert int(w0) == patch_pos_embed.shape[-2] and int(h0) == patch_pos_embed.shape[-1]
        patch_pos_embed = patch_pos_embed.permute(0, 2, 3, 1).reshape(1, -1, dim)
        return torch.cat((class_pos_embed.unsqueeze(0), patch_pos_embed), dim=1)

    m.interpolate_pos_encoding = types.MethodType(interpolate_pos_encoding, m)
    return m


def flops_per_image(arch: Arch, tokens: int, layer: int) -> float:
    """2*588*D*T + (layer+1)*(24*N*D^2 + 4*N^2*D), N = T+1."""
    n = tokens + 1
    d = arch.dim
    return 2.0 * 588 * d * tokens + (layer + 1) * (24.0 * n * d * d + 4.0 * n * n * d)
