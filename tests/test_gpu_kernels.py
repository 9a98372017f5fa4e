"""Kernel-level parity on the B200 (through the C ABI's vvs_dbg_* entry points): each hand-written
kernel against a plain PyTorch fp32 reference of the same op on identical (bf16-rounded) operands."""
import math

import pytest
import torch

pytestmark = pytest.mark.gpu

from vitvs_b200 import _lib  # noqa: E402


@pytest.fixture(scope="module")
def eng():
    from vitvs_b200 import Engine, ServoConfig
    e = Engine(ServoConfig(model_type="dinov2_vits14"), max_images=4)
    yield e
    e.close()


def _rand(shape, seed, scale=1.0, dtype=torch.float32):
    g = torch.Generator().manual_seed(seed)
    return (torch.randn(*shape, generator=g) * scale).to(dtype).cuda()


def _report(name, got, ref):
    err = (got.float() - ref.float()).abs().max().item()
    rel = err / max(ref.float().abs().max().item(), 1e-30)
    print(f"[{name}] max abs err {err:.3e}  rel-to-max {rel:.3e}")
    return err, rel


# (M, N, K): tails in M, N tile 128 and 256 paths, K = 1..many k-blocks, multi-wave persistence
GEMM_SHAPES = [(128, 256, 64), (200, 384, 128), (485 * 2, 1152, 384), (970, 768, 768), (300, 256, 3072),
               (128 * 150 + 37, 512, 256), (64, 128, 640)]


@pytest.mark.parametrize("M,N,K", GEMM_SHAPES)
def test_gemm_store_f32(eng, M, N, K):
    A = _rand((M, K), 1, dtype=torch.bfloat16)
    B = _rand((N, K), 2, dtype=torch.bfloat16)
    out = torch.full((M, N), float("nan"), device="cuda")
    eng.dbg_gemm(A, B, _lib.EPI_STORE_F32, x_f32=out, ldo=N)
    torch.cuda.synchronize()
    ref = A.float() @ B.float().T
    err, rel = _report(f"gemm f32 {M}x{N}x{K}", out, ref)
    assert rel < 1e-5 * math.sqrt(K)


def test_gemm_store_f32_ragged_n(eng):
    M, N, K = 70, 484, 192            # similarity-like: N not a multiple of 32
    A = _rand((M, K), 3, dtype=torch.bfloat16)
    B = _rand((N, K), 4, dtype=torch.bfloat16)
    out = torch.full((M, N), float("nan"), device="cuda")
    eng.dbg_gemm(A, B, _lib.EPI_STORE_F32, x_f32=out, ldo=N)
    torch.cuda.synchronize()
    _, rel = _report("gemm f32 ragged", out, A.float() @ B.float().T)
    assert rel < 1e-4


@pytest.mark.parametrize("M,N,K", [(485 * 2, 1152, 384), (970, 2304, 768)])
def test_gemm_bias_bf16_qscale(eng, M, N, K):
    A = _rand((M, K), 5, dtype=torch.bfloat16)
    B = _rand((N, K), 6, 0.05, dtype=torch.bfloat16)
    bias = _rand((N,), 7, 0.1)
    out = torch.zeros(M, N, dtype=torch.bfloat16, device="cuda")
    D = N // 3
    eng.dbg_gemm(A, B, _lib.EPI_BIAS_BF16, bias=bias, out_bf16=out, ldo=N, q_cols=D, q_scale=0.125)
    torch.cuda.synchronize()
    ref = A.float() @ B.float().T + bias
    ref[:, :D] *= 0.125
    _, rel = _report("gemm bias bf16", out, ref)
    assert rel < 1e-2                     # bf16 output rounding (2^-9)


def test_gemm_bias_bf16_ragged_tails(eng):
    """M and N tails of the TMA-store epilogue: 37 rows (one partial 32-row slab), N = 96 (half-filled slab)."""
    M, N, K = 37, 96, 128
    A = _rand((M, K), 50, dtype=torch.bfloat16)
    B = _rand((N, K), 51, 0.05, dtype=torch.bfloat16)
    bias = _rand((N,), 52, 0.1)
    out = torch.full((M + 3, N), 9.0, dtype=torch.bfloat16, device="cuda")      # 3 guard rows
    eng.dbg_gemm(A, B, _lib.EPI_BIAS_BF16, bias=bias, out_bf16=out, ldo=N, q_cols=0, q_scale=1.0)
    torch.cuda.synchronize()
    _, rel = _report("gemm bias bf16 tails", out[:M], A.float() @ B.float().T + bias)
    assert rel < 1e-2
    assert torch.all(out[M:] == 9.0)              # rows past M untouched (TMA clips)


def test_gemm_residual_f32_exact_add(eng):
    """The residual epilogue reduce-adds d = gamma*(acc+bias) into x at the L2: with x = 0 the result must be
    exactly d, and adding into a non-zero x must equal fp32 x + d bit for bit."""
    M, N, K = 300, 768, 256
    A = _rand((M, K), 53, dtype=torch.bfloat16)
    B = _rand((N, K), 54, 0.05, dtype=torch.bfloat16)
    bias, gamma = _rand((N,), 55, 0.1), _rand((N,), 56)
    d = torch.zeros(M, N, device="cuda")
    eng.dbg_gemm(A, B, _lib.EPI_RESID_F32, bias=bias, gamma=gamma, x_f32=d, ldo=N)
    x0 = _rand((M, N), 57)
    x = x0.clone()
    eng.dbg_gemm(A, B, _lib.EPI_RESID_F32, bias=bias, gamma=gamma, x_f32=x, ldo=N)
    torch.cuda.synchronize()
    assert torch.equal(x, x0 + d)


def test_gemm_gelu_bf16(eng):
    M, N, K = 970, 1536, 384
    A = _rand((M, K), 8, dtype=torch.bfloat16)
    B = _rand((N, K), 9, 0.05, dtype=torch.bfloat16)
    bias = _rand((N,), 10, 0.1)
    out = torch.zeros(M, N, dtype=torch.bfloat16, device="cuda")
    eng.dbg_gemm(A, B, _lib.EPI_GELU_BF16, bias=bias, out_bf16=out, ldo=N)
    torch.cuda.synchronize()
    ref = torch.nn.functional.gelu(A.float() @ B.float().T + bias)
    _, rel = _report("gemm gelu bf16", out, ref)
    assert rel < 1e-2


def test_gemm_residual_f32(eng):
    M, N, K = 970, 384, 1536
    A = _rand((M, K), 11, dtype=torch.bfloat16)
    B = _rand((N, K), 12, 0.05, dtype=torch.bfloat16)
    bias, gamma = _rand((N,), 13, 0.1), _rand((N,), 14)
    x = _rand((M, N), 15)
    ref = x + gamma * (A.float() @ B.float().T + bias)
    eng.dbg_gemm(A, B, _lib.EPI_RESID_F32, bias=bias, gamma=gamma, x_f32=x, ldo=N)
    torch.cuda.synchronize()
    _, rel = _report("gemm residual f32", x, ref)
    assert rel < 1e-4


def test_gemm_patch_embed_rows(eng):
    """EPI_PATCH_F32: A carries one all-zero row per CLS slot, so GEMM row r is residual-stream row r and
    x[r] = acc + bias + pos'[r % (T + 1)] leaves through a TMA store (ragged last tile: M = 150 rows)."""
    n_img, T, D, K = 3, 49, 384, 640
    A = _rand((n_img, T + 1, K), 16, dtype=torch.bfloat16)
    A[:, 0] = 0
    A = A.reshape(n_img * (T + 1), K).contiguous()
    B = _rand((D, K), 17, 0.05, dtype=torch.bfloat16)
    bias, pos = _rand((D,), 18, 0.1), _rand((T + 1, D), 19)
    x = torch.full((n_img * (T + 1), D), 7.0, device="cuda")
    eng.dbg_gemm(A, B, _lib.EPI_PATCH_F32, bias=bias, x_f32=x, ldo=D, pos=pos, T=T)
    torch.cuda.synchronize()
    ref = (A.float() @ B.float().T + bias).reshape(n_img, T + 1, D) + pos
    got = x.reshape(n_img, T + 1, D)
    _, rel = _report("gemm patch rows", got, ref)
    assert rel < 1e-4
    assert torch.equal(got[:, 0], (bias + pos[0]).expand(n_img, D))        # zero accumulator on the CLS rows


def test_gemm_swiglu(eng):
    M, D, H = 300, 1536, 4096
    A = _rand((M, D), 20, dtype=torch.bfloat16)
    W12 = _rand((2 * H, D), 21, 0.03, dtype=torch.bfloat16)
    b12 = _rand((2 * H,), 22, 0.1)
    bn = 256                                      # tile the launcher picks for N = 8192
    half = bn // 2
    r = torch.arange(2 * H)
    n, wi = r // bn, r % bn
    src = torch.where(wi < half, n * half + wi, H + n * half + (wi - half)).cuda()
    W12i, b12i = W12[src].contiguous(), b12[src].contiguous()
    out = torch.zeros(M, H, dtype=torch.bfloat16, device="cuda")
    eng.dbg_gemm(A, W12i, _lib.EPI_SWIGLU_BF16, bias=b12i, out_bf16=out, ldo=H)
    torch.cuda.synchronize()
    x12 = A.float() @ W12.float().T + b12
    ref = torch.nn.functional.silu(x12[:, :H]) * x12[:, H:]
    _, rel = _report("gemm swiglu", out, ref)
    assert rel < 1e-2


# ---- LayerNorm folded into the GEMMs on either side of it (dino_patch/block.py:90-95: x + ls(attn(norm1(x))), ...) ----
def _group_stats(x):
    """(sum, sum of squares) of every 64-column group of the fp32 rows: what a producer epilogue emits."""
    g = x.float().reshape(x.shape[0], -1, 64)
    return torch.stack([g.sum(-1), (g * g).sum(-1)], dim=-1).contiguous()


# small tiles (1 CTA, 128-wide), CTA pairs (256-wide), mixed clusters (M >= 4096), K = 1 and many k-blocks, ragged M
@pytest.mark.parametrize("M,N,K", [(300, 768, 256), (970, 384, 1536), (3400, 768, 768), (4400 + 37, 768, 768)])
def test_gemm_residual_ln_producer(eng, M, N, K):
    """EPI_RESID_LN_F32: the new residual rows must equal the reduce-add epilogue's bit for bit, xb must be their
    bf16 rounding, and the per-group statistics must be those of the stored rows."""
    A = _rand((M, K), 60, dtype=torch.bfloat16)
    B = _rand((N, K), 61, 0.05, dtype=torch.bfloat16)
    bias, gamma = _rand((N,), 62, 0.1), _rand((N,), 63)
    x0 = _rand((M, N), 64) + 0.3
    x_ref = x0.clone()
    eng.dbg_gemm(A, B, _lib.EPI_RESID_F32, bias=bias, gamma=gamma, x_f32=x_ref, ldo=N)
    x = x0.clone()
    xb = torch.full((M + 2, N), 5.0, dtype=torch.bfloat16, device="cuda")       # 2 guard rows
    part = torch.full((M + 2, N // 64, 2), -1.0, device="cuda")
    eng.dbg_gemm_ln(A, B, _lib.EPI_RESID_LN_F32, bias=bias, gamma=gamma, x_f32=x, ldo=N, xb_out=xb, ln_part_out=part)
    torch.cuda.synchronize()
    assert torch.equal(x, x_ref)
    assert torch.equal(xb[:M], x.to(torch.bfloat16))
    assert torch.all(xb[M:] == 5.0) and torch.all(part[M:] == -1.0)             # rows past M untouched
    ref = _group_stats(x)
    assert torch.allclose(part[:M], ref, rtol=2e-5, atol=1e-4), (part[:M] - ref).abs().max().item()


def test_gemm_patch_ln_producer(eng):
    n_img, T, D, K = 3, 49, 384, 640
    A = _rand((n_img, T + 1, K), 16, dtype=torch.bfloat16)
    A[:, 0] = 0
    A = A.reshape(n_img * (T + 1), K).contiguous()
    B = _rand((D, K), 17, 0.05, dtype=torch.bfloat16)
    bias, pos = _rand((D,), 18, 0.1), _rand((T + 1, D), 19)
    M = n_img * (T + 1)
    x_ref = torch.zeros(M, D, device="cuda")
    eng.dbg_gemm(A, B, _lib.EPI_PATCH_F32, bias=bias, x_f32=x_ref, ldo=D, pos=pos, T=T)
    x = torch.zeros(M, D, device="cuda")
    xb = torch.zeros(M, D, dtype=torch.bfloat16, device="cuda")
    part = torch.zeros(M, D // 64, 2, device="cuda")
    pos_ext = torch.cat([pos, pos[:31]]).contiguous()      # the epilogue fetches 32-row boxes: periodic extension
    eng.dbg_gemm_ln(A, B, _lib.EPI_PATCH_LN_F32, bias=bias, x_f32=x, ldo=D, pos=pos_ext, T=T, xb_out=xb, ln_part_out=part)
    torch.cuda.synchronize()
    assert torch.equal(x, x_ref) and torch.equal(xb, x.to(torch.bfloat16))
    assert torch.allclose(part, _group_stats(x), rtol=2e-5, atol=1e-4)


def _ln_consumer_case(M, N, K, seed, mean_shift=0.5):
    """x rows with a non-zero mean, LayerNorm(gamma, beta), Linear(W, b): the fp32 reference of Linear(LN(x)) and the
    operands of the folded form (what api.cu's fold_layernorm_kernel and a producer epilogue hand to the consumer)."""
    x = _rand((M, K), seed) * 1.7 + mean_shift
    g, be = 1.0 + _rand((K,), seed + 1, 0.1), _rand((K,), seed + 2, 0.05)
    W, b = _rand((N, K), seed + 3, 0.03), _rand((N,), seed + 4, 0.1)
    ref = torch.nn.functional.layer_norm(x, (K,), g, be, 1e-6) @ W.T + b
    Wf = (W * g).to(torch.bfloat16)
    return dict(xb=x.to(torch.bfloat16), part=_group_stats(x), Wf=Wf, cs=Wf.float().sum(1).contiguous(),
                bf=(b + W @ be).contiguous(), ref=ref)


@pytest.mark.parametrize("M,N,K", [(970, 1152, 384), (300, 2304, 768), (4400 + 37, 2304, 768), (3400, 768, 1024)])
def test_gemm_bias_ln_consumer(eng, M, N, K):
    c = _ln_consumer_case(M, N, K, 70)
    out = torch.zeros(M, N, dtype=torch.bfloat16, device="cuda")
    D = N // 3
    eng.dbg_gemm_ln(c["xb"], c["Wf"], _lib.EPI_BIAS_LN_BF16, bias=c["bf"], out_bf16=out, ldo=N, q_cols=D, q_scale=0.125,
                    ln_part_in=c["part"], colsum=c["cs"])
    torch.cuda.synchronize()
    ref = c["ref"].clone()
    ref[:, :D] *= 0.125
    _, rel = _report(f"gemm bias+LN {M}x{N}x{K}", out, ref)
    assert rel < 1e-2


def test_gemm_gelu_ln_consumer(eng):
    M, N, K = 970, 1536, 384
    c = _ln_consumer_case(M, N, K, 80)
    out = torch.zeros(M, N, dtype=torch.bfloat16, device="cuda")
    eng.dbg_gemm_ln(c["xb"], c["Wf"], _lib.EPI_GELU_LN_BF16, bias=c["bf"], out_bf16=out, ldo=N, ln_part_in=c["part"],
                    colsum=c["cs"])
    torch.cuda.synchronize()
    _, rel = _report("gemm gelu+LN", out, torch.nn.functional.gelu(c["ref"]))
    assert rel < 1e-2


def test_gemm_swiglu_ln_consumer(eng):
    M, D, H = 300, 1536, 4096
    c = _ln_consumer_case(M, 2 * H, D, 90)
    bn, half = 256, 128
    r = torch.arange(2 * H)
    n, wi = r // bn, r % bn
    src = torch.where(wi < half, n * half + wi, H + n * half + (wi - half)).cuda()
    out = torch.zeros(M, H, dtype=torch.bfloat16, device="cuda")
    eng.dbg_gemm_ln(c["xb"], c["Wf"][src].contiguous(), _lib.EPI_SWIGLU_LN_BF16, bias=c["bf"][src].contiguous(),
                    out_bf16=out, ldo=H, ln_part_in=c["part"], colsum=c["cs"][src].contiguous())
    torch.cuda.synchronize()
    ref = torch.nn.functional.silu(c["ref"][:, :H]) * c["ref"][:, H:]
    _, rel = _report("gemm swiglu+LN", out, ref)
    assert rel < 1e-2


def _rms(a, ref):
    return ((a.double() - ref) ** 2).mean().sqrt().item()


@pytest.mark.parametrize("offset,outlier", [(0.3, 300.0), (10.0, 0.0), (30.0, 300.0)])
def test_ln_fold_rows_with_offset_and_outlier_channel(eng, offset, outlier):
    """Rows unlike the random-init ones of the model tests: a common offset of `offset` standard deviations and one
    channel `outlier` standard deviations away (the massive-activation channels of trained ViTs).  The folded form feeds
    bf16(x) to the tensor cores, the unfused form (LayerNorm kernel, then EPI_BIAS_BF16) feeds bf16(LN(x)).  bf16 rounding
    is relative, so a large channel costs nothing; an offset does, because x_k carries it and LN(x)_k does not: the
    operand error of the fold grows like sqrt(1 + (mean/std)^2) of the row.  The test pins exactly that law against the
    unfused path ON THE SAME INPUT - a fold that loses more than the law allows fails - and keeps the variance
    E[x^2] - mean^2 honest at 30 sigma."""
    M, N, K = 970, 1152, 768
    x = _rand((M, K), 170) + offset
    x[:, 17] += outlier
    g, be = 1.0 + _rand((K,), 171, 0.1), _rand((K,), 172, 0.05)
    W, b = _rand((N, K), 173, 0.03), _rand((N,), 174, 0.1)
    ref = torch.nn.functional.layer_norm(x.double(), (K,), g.double(), be.double(), 1e-6) @ W.double().T + b.double()
    # unfused: the LayerNorm kernel rounds LN(x) to bf16, the plain GEMM multiplies it
    out_u = torch.zeros(M, N, dtype=torch.bfloat16, device="cuda")
    eng.dbg_gemm(eng.dbg_layernorm(x, g, be), W.to(torch.bfloat16), _lib.EPI_BIAS_BF16, bias=b, out_bf16=out_u, ldo=N)
    # folded: raw bf16 rows, row statistics, W' = W diag(gamma), colsum(W'), b' = b + W beta
    Wf = (W * g).to(torch.bfloat16)
    out_f = torch.zeros(M, N, dtype=torch.bfloat16, device="cuda")
    eng.dbg_gemm_ln(x.to(torch.bfloat16), Wf, _lib.EPI_BIAS_LN_BF16, bias=(b + W @ be).contiguous(), out_bf16=out_f, ldo=N,
                    ln_part_in=_group_stats(x), colsum=Wf.float().sum(1).contiguous())
    torch.cuda.synchronize()
    err_u, err_f, scale = _rms(out_u, ref), _rms(out_f, ref), ref.pow(2).mean().sqrt().item()
    ratio = (x.mean(1).abs() / x.std(1)).max().item()
    print(f"[ln fold offset {offset} outlier {outlier}] row |mean|/std {ratio:.2f}: rms err unfused {err_u:.3e} "
          f"folded {err_f:.3e} (output rms {scale:.3e}) -> folded / unfused {err_f / err_u:.2f}, "
          f"law {math.sqrt(1 + ratio * ratio):.2f}")
    assert err_u < 1e-2 * scale
    assert err_f <= 1.5 * math.sqrt(1 + ratio * ratio) * err_u


@pytest.mark.parametrize("n_img,n_tok,heads", [(1, 128, 1), (2, 485, 6), (3, 50, 2), (1, 1370, 12), (2, 257, 3)])
def test_attention(eng, n_img, n_tok, heads):
    D = heads * 64
    qkv = _rand((n_img * n_tok, 3 * D), 30 + n_tok, 1.0, dtype=torch.bfloat16)
    qkv[:, :D] *= 0.125                           # q arrives pre-scaled by hd^-0.5 (QKV GEMM epilogue)
    out = eng.dbg_attention(qkv, n_img, n_tok, heads)
    torch.cuda.synchronize()
    x = qkv.float().reshape(n_img, n_tok, 3, heads, 64)
    q, k, v = (t.transpose(1, 2) for t in x.unbind(2))
    ref = torch.nn.functional.scaled_dot_product_attention(q, k, v, scale=1.0)
    ref = ref.transpose(1, 2).reshape(n_img * n_tok, D)
    err, rel = _report(f"attention {n_img}x{n_tok}x{heads}", out, ref)
    assert rel < 2e-2                             # bf16 P and bf16 output


@pytest.mark.parametrize("D", [384, 768, 1024, 1536])
def test_layernorm(eng, D):
    rows = 971
    x = _rand((rows, D), 40, 3.0) + 0.5
    w, b = _rand((D,), 41) + 1.0, _rand((D,), 42, 0.1)
    out = eng.dbg_layernorm(x, w, b)
    torch.cuda.synchronize()
    ref = torch.nn.functional.layer_norm(x, (D,), w, b, eps=1e-6)
    _, rel = _report(f"layernorm {D}", out, ref)
    assert rel < 6e-3


def test_gemm_two_pair_clusters_with_multicast():
    """VITVS_GEMM_CTAS=4 (experimental): clusters of two CTA pairs sharing the B tile through TMA multicast; the
    parity script covers bias/bf16 and residual/fp32 epilogues at sizes that take that path, ragged M included."""
    import os, subprocess, sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    env = dict(os.environ, VITVS_GEMM_CTAS="4")
    r = subprocess.run([sys.executable, os.path.join(root, "tools", "gemm_check.py")], capture_output=True, text=True,
                       timeout=300, env=env, cwd=root)
    assert r.returncode == 0 and r.stdout.strip().endswith("OK"), r.stdout[-1500:] + r.stderr[-1500:]
