"""Backbone restatement cross-checked by weight copy against transformers' Dinov2Model
(the only independent DINOv2 implementation in the image; the hub model is unobtainable)."""
import pytest
import torch

from oracle import dinov2


def _hf_from_oracle(m, arch):
    tr = pytest.importorskip("transformers")
    cfg = tr.Dinov2Config(hidden_size=arch.dim, num_hidden_layers=arch.depth, num_attention_heads=arch.heads,
                          mlp_ratio=4, image_size=518, patch_size=14, layerscale_value=1.0,
                          use_swiglu_ffn=arch.swiglu, qkv_bias=True, layer_norm_eps=1e-6,
                          hidden_act="gelu", attn_implementation="eager")
    hf = tr.Dinov2Model(cfg).eval()
    sd = {}
    sd["embeddings.cls_token"] = m.cls_token
    sd["embeddings.mask_token"] = torch.zeros(1, arch.dim)
    sd["embeddings.position_embeddings"] = m.pos_embed
    sd["embeddings.patch_embeddings.projection.weight"] = m.patch_embed.proj.weight
    sd["embeddings.patch_embeddings.projection.bias"] = m.patch_embed.proj.bias
    for i, b in enumerate(m.blocks):
        p = f"encoder.layer.{i}."
        D = arch.dim
        for j, n in enumerate(("query", "key", "value")):
            sd[p + f"attention.attention.{n}.weight"] = b.attn.qkv.weight[j * D:(j + 1) * D]
            sd[p + f"attention.attention.{n}.bias"] = b.attn.qkv.bias[j * D:(j + 1) * D]
        sd[p + "attention.output.dense.weight"] = b.attn.proj.weight
        sd[p + "attention.output.dense.bias"] = b.attn.proj.bias
        sd[p + "norm1.weight"], sd[p + "norm1.bias"] = b.norm1.weight, b.norm1.bias
        sd[p + "norm2.weight"], sd[p + "norm2.bias"] = b.norm2.weight, b.norm2.bias
        sd[p + "layer_scale1.lambda1"] = b.ls1.gamma
        sd[p + "layer_scale2.lambda1"] = b.ls2.gamma
        if arch.swiglu:
            sd[p + "mlp.weights_in.weight"], sd[p + "mlp.weights_in.bias"] = b.mlp.w12.weight, b.mlp.w12.bias
            sd[p + "mlp.weights_out.weight"], sd[p + "mlp.weights_out.bias"] = b.mlp.w3.weight, b.mlp.w3.bias
        else:
            sd[p + "mlp.fc1.weight"], sd[p + "mlp.fc1.bias"] = b.mlp.fc1.weight, b.mlp.fc1.bias
            sd[p + "mlp.fc2.weight"], sd[p + "mlp.fc2.bias"] = b.mlp.fc2.weight, b.mlp.fc2.bias
    sd["layernorm.weight"], sd["layernorm.bias"] = m.norm.weight, m.norm.bias
    missing, unexpected = hf.load_state_dict({k: v.detach().clone() for k, v in sd.items()}, strict=False)
    assert not unexpected, unexpected
    assert not missing, missing
    return hf


def test_vits_matches_hf(size=518):
    arch = dinov2.ARCHS["dinov2_vits14"]
    m = dinov2.build("dinov2_vits14", seed=0)
    hf = _hf_from_oracle(m, arch)
    x = torch.randn(1, 3, size, size, generator=torch.Generator().manual_seed(0))
    with torch.no_grad():
        out = hf(pixel_values=x, output_hidden_states=True, interpolate_pos_encoding=True)
        mine = m.prepare_tokens(x)
        # native 518 grid: no pos-embed interpolation => exact architecture check (the
        # interpolation itself is pinned by the reference's own _fix_pos_enc, see
        # test_oracle_vs_reference.py; HF resamples with size= instead of the +0.1 kludge).
        tol = 2e-5
        assert torch.allclose(out.hidden_states[0], mine, atol=tol)
        for i, blk in enumerate(m.blocks):
            mine = blk(mine)
        assert torch.allclose(out.hidden_states[-1], mine, atol=tol * 10)
        assert torch.allclose(out.last_hidden_state[:, 0], m(x), atol=tol * 10)


def test_swiglu_block_matches_hf():
    """ViT-g FFN (SwiGLUFFNFused, hidden 4096) on a 2-block giant-width model."""
    arch = dinov2.Arch("tiny_g", 1536, 24, 2, True)
    assert arch.ffn_hidden == 4096
    torch.manual_seed(0)
    m = dinov2.DinoV2(arch).eval()
    with torch.no_grad():
        for p in m.parameters():
            p.normal_(0, 0.02)
    hf = _hf_from_oracle(m, arch)
    x = torch.randn(1, 3, 518, 518, generator=torch.Generator().manual_seed(0))[:, :, :70, :70]
    x = torch.nn.functional.pad(x, (0, 518 - 70, 0, 518 - 70))
    with torch.no_grad():
        out = hf(pixel_values=x, output_hidden_states=True)
        mine = m.prepare_tokens(x)
        for blk in m.blocks:
            mine = blk(mine)
    assert torch.allclose(out.hidden_states[-1], mine, atol=2e-4)


def test_bf16_emulation_stays_next_to_the_fp32_oracle():
    """oracle/bf16_emulation.py: rounding the operands where the GPU path rounds them moves the tokens of 12 ViT-S blocks by
    what bf16 predicts (cosine 0.99998) and no more - an emulation that drifted further would blame kernels for nothing."""
    import numpy as np
    from oracle import bf16_emulation as be, dinov2, servo
    from vitvs_b200 import synth
    m = dinov2.build("dinov2_vits14", seed=0)
    g, c, d = synth.frame_pair(1)
    tok, S16 = be.similarity_matrices(m, g, c)
    with torch.no_grad():
        ref = servo.block_tokens(m, torch.cat([servo.preprocess_u8(g), servo.preprocess_u8(c)]), 11)
    S = servo.frame_pair_to_twist(m, g, c, d)["S"]
    cos = torch.nn.functional.cosine_similarity(tok, ref, dim=-1).min().item()
    assert 0.9999 <= cos < 0.999999                      # rounded, but only by bf16
    assert (S16 - S).abs().max() < 2e-3
    assert (S16.argmax(1) == S.argmax(1)).float().mean() >= 0.99
