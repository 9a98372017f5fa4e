"""bf16-operand emulation of the oracle backbone.  TEST INFRASTRUCTURE ONLY (imported by tests/ alone).

Why.  The GPU path multiplies bf16 operands and accumulates in fp32; the oracle (oracle/dinov2.py) is fp32
throughout, like the reference on CPU.  When the two disagree on an arg-max index there are two possible causes: the
rounding of the operands (arithmetic, allowed by north_star's tolerances) or a defect in a kernel (not allowed).  This
module restates the GPU path's STATED arithmetic on the CPU with plain torch fp32 ops on bf16-rounded values, so that

  * GPU vs emulation isolates kernel defects: the two share the rounding of x, W, qkv, the attention output and the
    GELU / SwiGLU output, and differ in accumulation order, in the exp / erf / rsqrt implementations and in ONE rounding
    point - the flash-attention kernel rounds P = exp(s - running max) block by block and rescales, the emulation rounds
    exp(s - final max) - so they agree tighter than GPU vs fp32 (measured: 1 - cosine 0.42 - 0.66 of the fp32 figure);
  * emulation vs fp32 oracle shows what the rounding alone costs - e.g. the 1 - 3 % arg-max flips of random-init ViT-g.

Rounding points emulated (file:line of what each stands for is in DESIGN.md 3 and csrc/api.cu forward_tokens):
  patch embedding   bf16(normalised pixels) x bf16(W), fp32 accumulate, + bias + pos in fp32
  norm1 -> qkv      LayerNorm folded: rstd * (bf16(x) . bf16(W diag(gamma)) - mean * colsum(bf16(W diag(gamma)))) + (b + W beta),
                    statistics of the fp32 row with var = E[x^2] - mean^2; q scaled by hd^-0.5; qkv stored bf16
  attention         S = q k^T in fp32, P = bf16(exp(S - rowmax)), row sum over the ROUNDED P, O = P v / sum, stored bf16
  proj, fc2 / w3    bf16 input x bf16(W), fp32: x += gamma_ls * (acc + b) on the fp32 residual stream
  norm2 -> fc1/w12  folded like qkv; GELU(erf) or SiLU(x1) * x2 in fp32, stored bf16
  similarity        descriptors L2-normalised in fp32, rounded to bf16, fp32 product
It follows dino_patch/block.py:90-95,112-114 and attention.py:70-80 exactly as oracle/dinov2.py does; only the
rounding is added.  Parity status: the emulation is derived from the oracle, so it is pinned wherever the oracle is."""
from __future__ import annotations

import torch
import torch.nn.functional as F
from torch import nn

from . import servo


def r16(t: torch.Tensor) -> torch.Tensor:
    """round-to-nearest-even to bf16 and back: the value a bf16 operand / stored activation carries"""
    return t.to(torch.bfloat16).to(torch.float32)


def _linear(x16: torch.Tensor, lin: nn.Linear) -> torch.Tensor:
    return x16 @ r16(lin.weight).T + lin.bias


def _folded(x: torch.Tensor, norm: nn.LayerNorm, lin: nn.Linear) -> torch.Tensor:
    """Linear(LayerNorm(x)) the way the consumer GEMM computes it (csrc/gemm_tc.cu EPI_*_LN_*, api.cu refold_block)."""
    mean = x.mean(-1, keepdim=True)
    var = ((x * x).mean(-1, keepdim=True) - mean * mean).clamp_min(0.0)
    rstd = torch.rsqrt(var + norm.eps)
    wf = r16(lin.weight * norm.weight)                 # W diag(gamma), rounded once from the fp32 product
    colsum = wf.sum(1)                                 # over the ROUNDED values, so the mean term cancels exactly
    bias = lin.bias + lin.weight @ norm.bias
    return rstd * (r16(x) @ wf.T - mean * colsum) + bias


class EmulatedBlock(nn.Module):
    def __init__(self, blk):
        super().__init__()
        self.b = blk

    def forward(self, x):
        b = self.b
        B, N, C = x.shape
        H = b.attn.num_heads
        qkv = _folded(x, b.norm1, b.attn.qkv)
        qkv[..., :C] *= b.attn.scale                    # q pre-scaled in the QKV epilogue (a power of two for hd = 64)
        qkv = r16(qkv).reshape(B, N, 3, H, C // H)
        q, k, v = (t.transpose(1, 2) for t in torch.unbind(qkv, 2))
        s = q @ k.transpose(-1, -2)
        p = r16(torch.exp(s - s.amax(-1, keepdim=True)))
        o = r16((p @ v) / p.sum(-1, keepdim=True))
        o = o.transpose(1, 2).reshape(B, N, C)
        x = x + b.ls1.gamma * _linear(o, b.attn.proj)
        if hasattr(b.mlp, "w12"):
            x1, x2 = _folded(x, b.norm2, b.mlp.w12).chunk(2, dim=-1)
            h = r16(F.silu(x1) * x2)
            x = x + b.ls2.gamma * _linear(h, b.mlp.w3)
        else:
            h = r16(F.gelu(_folded(x, b.norm2, b.mlp.fc1)))
            x = x + b.ls2.gamma * _linear(h, b.mlp.fc2)
        return x


class EmulatedDinoV2(nn.Module):
    """Same surface as oracle.dinov2.DinoV2 for what oracle/servo.py calls: prepare_tokens + blocks."""

    def __init__(self, model):
        super().__init__()
        self.model = model
        self.blocks = nn.ModuleList([EmulatedBlock(b) for b in model.blocks])

    def prepare_tokens(self, x):
        m = self.model
        B, _, w, h = x.shape
        conv = m.patch_embed.proj
        t = F.conv2d(r16(x), r16(conv.weight), conv.bias, stride=conv.stride).flatten(2).transpose(1, 2)
        t = torch.cat((m.cls_token.expand(B, -1, -1), t), dim=1)
        return t + m.interpolate_pos_encoding(t, w, h)


def emulate(model) -> EmulatedDinoV2:
    return EmulatedDinoV2(model).eval()


def similarity_bf16(desc1: torch.Tensor, desc2: torch.Tensor) -> torch.Tensor:
    """vitvs_v2.py:49-56 on unit descriptors rounded to bf16 (what EPI_SIM multiplies): [.., T, d] x2 -> [T, T] fp32"""
    a = r16(F.normalize(desc1.reshape(-1, desc1.shape[-1]), dim=-1, eps=1e-8))
    b = r16(F.normalize(desc2.reshape(-1, desc2.shape[-1]), dim=-1, eps=1e-8))
    return a @ b.T


def similarity_matrices(model, goal_u8, cur_u8, layer: int = 11, bin: bool = True):
    """-> (tokens [2, 1+T, D], S [T, T]) of one frame pair through the emulated backbone and similarity"""
    with torch.no_grad():
        pre = torch.cat([servo.preprocess_u8(goal_u8), servo.preprocess_u8(cur_u8)])
        em = emulate(model)
        tok = servo.block_tokens(em, pre, layer)
        d = servo.extract_descriptors(em, pre, layer, bin)
        return tok, similarity_bf16(d[0], d[1])
