"""Run single kernels at the ViT-B / 64-pair shapes (for ncu and quick CUDA-event timing on the GPU box).
usage: python tools/prof_kernels.py [attention|gemm_ffn1|gemm_qkv|gemm_proj|gemm_ffn2|layernorm|all] [reps]"""
import sys
import os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from vitvs_b200 import Engine, ServoConfig, _lib

which = sys.argv[1] if len(sys.argv) > 1 else "all"
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 5
eng = Engine(ServoConfig(model_type="dinov2_vitb14"), max_images=2)
n_img, n_tok, heads, D, H = 128, int(os.environ.get('NTOK', '485')), 12, 768, 3072
M = n_img * n_tok
g = torch.Generator(device="cuda").manual_seed(0)


def timeit(name, fn, flops=None, bytes_=None):
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    extra = ""
    if flops:
        extra += f"  {flops / ms / 1e9:.1f} TFLOP/s"
    if bytes_:
        extra += f"  {bytes_ / ms / 1e6:.1f} GB/s"
    print(f"{name:12s} {ms * 1e3:9.1f} us{extra}", flush=True)


if which in ("attention", "all"):
    qkv = (torch.randn(M, 3 * D, device="cuda", generator=g) * 1.0).to(torch.bfloat16)
    qkv[:, :D] *= 0.125
    timeit("attention", lambda: eng.dbg_attention(qkv, n_img, n_tok, heads), flops=4.0 * n_tok * n_tok * D * n_img)
if which in ("gemm_ffn1", "all"):
    A = torch.randn(M, D, device="cuda", generator=g).to(torch.bfloat16)
    W = (torch.randn(H, D, device="cuda", generator=g) * 0.02).to(torch.bfloat16)
    bias = torch.randn(H, device="cuda", generator=g)
    out = torch.empty(M, H, dtype=torch.bfloat16, device="cuda")
    timeit("gemm_ffn1", lambda: eng.dbg_gemm(A, W, _lib.EPI_GELU_BF16, bias=bias, out_bf16=out, ldo=H), flops=2.0 * M * D * H)
if which in ("gemm_qkv", "all"):
    A = torch.randn(M, D, device="cuda", generator=g).to(torch.bfloat16)
    W = (torch.randn(3 * D, D, device="cuda", generator=g) * 0.02).to(torch.bfloat16)
    bias = torch.randn(3 * D, device="cuda", generator=g)
    out = torch.empty(M, 3 * D, dtype=torch.bfloat16, device="cuda")
    timeit("gemm_qkv", lambda: eng.dbg_gemm(A, W, _lib.EPI_BIAS_BF16, bias=bias, out_bf16=out, ldo=3 * D, q_cols=D, q_scale=0.125),
           flops=2.0 * M * D * 3 * D)
if which in ("gemm_proj", "all"):
    A = torch.randn(M, D, device="cuda", generator=g).to(torch.bfloat16)
    W = (torch.randn(D, D, device="cuda", generator=g) * 0.02).to(torch.bfloat16)
    bias, gamma = torch.randn(D, device="cuda", generator=g), torch.randn(D, device="cuda", generator=g)
    x = torch.randn(M, D, device="cuda", generator=g)
    timeit("gemm_proj", lambda: eng.dbg_gemm(A, W, _lib.EPI_RESID_F32, bias=bias, gamma=gamma, x_f32=x, ldo=D),
           flops=2.0 * M * D * D, bytes_=M * D * (2 + 8))
if which in ("gemm_ffn2", "all"):
    A = torch.randn(M, H, device="cuda", generator=g).to(torch.bfloat16)
    W = (torch.randn(D, H, device="cuda", generator=g) * 0.02).to(torch.bfloat16)
    bias, gamma = torch.randn(D, device="cuda", generator=g), torch.randn(D, device="cuda", generator=g)
    x = torch.randn(M, D, device="cuda", generator=g)
    timeit("gemm_ffn2", lambda: eng.dbg_gemm(A, W, _lib.EPI_RESID_F32, bias=bias, gamma=gamma, x_f32=x, ldo=D),
           flops=2.0 * M * D * H)
if which in ("layernorm", "all"):
    x = torch.randn(M, D, device="cuda", generator=g)
    w, b = torch.randn(D, device="cuda", generator=g), torch.randn(D, device="cuda", generator=g)
    timeit("layernorm", lambda: eng.dbg_layernorm(x, w, b), bytes_=M * D * 6)
