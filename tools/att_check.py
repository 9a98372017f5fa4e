"""Attention parity sweep against torch SDPA (GPU box)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from vitvs_b200 import Engine, ServoConfig
eng = Engine(ServoConfig(model_type="dinov2_vitb14"), max_images=2)
cases = [tuple(int(x) for x in a.split("x")) for a in sys.argv[1:]] or [(1, 64, 1), (1, 320, 1), (1, 384, 1), (1, 385, 1), (1, 448, 1), (1, 485, 1), (1, 512, 1)]
for n_img, n_tok, heads in cases:
    D = heads * 64
    g = torch.Generator(device="cuda").manual_seed(n_tok)
    qkv = torch.randn(n_img * n_tok, 3 * D, device="cuda", generator=g).to(torch.bfloat16)
    qkv[:, :D] *= 0.125
    out = eng.dbg_attention(qkv, n_img, n_tok, heads)
    torch.cuda.synchronize()
    x = qkv.float().reshape(n_img, n_tok, 3, heads, 64)
    q, k, v = (t.transpose(1, 2) for t in x.unbind(2))
    ref = torch.nn.functional.scaled_dot_product_attention(q, k, v, scale=1.0).transpose(1, 2).reshape(n_img * n_tok, D)
    err = (out.float() - ref).abs()
    bad_rows = (err.amax(dim=1) > 0.05).nonzero().flatten()
    print(f"{n_img}x{n_tok}x{heads}: max err {err.max().item():.4f} rel {err.max().item() / ref.abs().max().item():.4f} "
          f"bad rows {bad_rows.numel()} first {bad_rows[:6].tolist()} last {bad_rows[-3:].tolist()}", flush=True)
