"""GEMM parity at sizes that take the two-pair cluster path (GPU box): bias/bf16, residual fp32, ragged M."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from vitvs_b200 import Engine, ServoConfig, _lib
eng = Engine(ServoConfig(model_type="dinov2_vitb14"), max_images=2)
def rnd(shape, seed, scale=1.0, dtype=torch.float32):
    g = torch.Generator(device="cuda").manual_seed(seed)
    return (torch.randn(*shape, device="cuda", generator=g) * scale).to(dtype)
ok = True
for M, N, K in [(4096, 2304, 768), (62080, 768, 768), (5000, 3072, 768), (8277, 768, 3072)]:
    A = rnd((M, K), 1, dtype=torch.bfloat16); B = rnd((N, K), 2, 0.05, dtype=torch.bfloat16); bias = rnd((N,), 3, 0.1)
    out = torch.zeros(M, N, dtype=torch.bfloat16, device="cuda")
    eng.dbg_gemm(A, B, _lib.EPI_BIAS_BF16, bias=bias, out_bf16=out, ldo=N, q_cols=0, q_scale=1.0)
    torch.cuda.synchronize()
    ref = A.float() @ B.float().T + bias
    rel = ((out.float() - ref).abs().max() / ref.abs().max()).item()
    gamma = rnd((N,), 4); x = rnd((M, N), 5); x0 = x.clone()
    eng.dbg_gemm(A, B, _lib.EPI_RESID_F32, bias=bias, gamma=gamma, x_f32=x, ldo=N)
    torch.cuda.synchronize()
    ref2 = x0 + gamma * ref
    rel2 = ((x - ref2).abs().max() / ref2.abs().max()).item()
    print(f"M={M} N={N} K={K}: bias-bf16 rel {rel:.2e}  resid-f32 rel {rel2:.2e}", flush=True)
    ok &= rel < 1e-2 and rel2 < 1e-4
print("OK" if ok else "MISMATCH")
