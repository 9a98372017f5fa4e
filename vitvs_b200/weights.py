"""Random-init DINOv2 state_dicts with the upstream parameter names (there is no network for
checkpoints: benchmarks run on random weights of the named architecture).  Host-side only."""
from __future__ import annotations

import torch

from .config import ARCHS


def random_state_dict(model_type: str, seed: int = 0, max_block: int | None = None) -> dict:
    """trunc-normal(0.02) matrices, N(0,0.02) biases, LN ~ (1 +- 0.1, 0.05), LayerScale U(0.5,1.5),
    pos_embed trunc-normal(0.02) on the 37x37 training grid, cls_token N(0,1e-6)."""
    dim, heads, depth, hidden, swiglu = ARCHS[model_type]
    g = torch.Generator().manual_seed(seed)
    last = depth - 1 if max_block is None else min(max_block, depth - 1)

    def tn(*shape):
        return torch.nn.init.trunc_normal_(torch.empty(*shape), std=0.02, generator=g)

    def nb(n):
        return torch.randn(n, generator=g) * 0.02

    sd = {
        "patch_embed.proj.weight": tn(dim, 3, 14, 14),
        "patch_embed.proj.bias": nb(dim),
        "cls_token": (torch.randn(1, 1, dim, generator=g) * 1e-6),
        "pos_embed": tn(1, 1 + 37 * 37, dim),
    }
    for i in range(last + 1):
        p = f"blocks.{i}."
        sd[p + "norm1.weight"] = 1.0 + 0.1 * torch.randn(dim, generator=g)
        sd[p + "norm1.bias"] = 0.05 * torch.randn(dim, generator=g)
        sd[p + "attn.qkv.weight"] = tn(3 * dim, dim)
        sd[p + "attn.qkv.bias"] = nb(3 * dim)
        sd[p + "attn.proj.weight"] = tn(dim, dim)
        sd[p + "attn.proj.bias"] = nb(dim)
        sd[p + "ls1.gamma"] = 0.5 + torch.rand(dim, generator=g)
        sd[p + "norm2.weight"] = 1.0 + 0.1 * torch.randn(dim, generator=g)
        sd[p + "norm2.bias"] = 0.05 * torch.randn(dim, generator=g)
        if swiglu:
            sd[p + "mlp.w12.weight"] = tn(2 * hidden, dim)
            sd[p + "mlp.w12.bias"] = nb(2 * hidden)
            sd[p + "mlp.w3.weight"] = tn(dim, hidden)
            sd[p + "mlp.w3.bias"] = nb(dim)
        else:
            sd[p + "mlp.fc1.weight"] = tn(hidden, dim)
            sd[p + "mlp.fc1.bias"] = nb(hidden)
            sd[p + "mlp.fc2.weight"] = tn(dim, hidden)
            sd[p + "mlp.fc2.bias"] = nb(dim)
        sd[p + "ls2.gamma"] = 0.5 + torch.rand(dim, generator=g)
    return sd
