"""Full-depth parity of the deep DINOv2 sizes (ViT-L layer = 23, ViT-g layer = 39): bf16 operands /
fp32 residual stream through 24 and 40 blocks against the fp32 oracle, one synthetic frame pair each.  The token
cosine is printed per depth so that a failure names the block at which 0.999 breaks."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from oracle import dinov2, servo  # noqa: E402
from vitvs_b200 import synth  # noqa: E402

from _parity import check_argmax, check_keypoints, margin_histogram  # noqa: E402

# North_star: >= 99 % index agreement, disagreements only at near-ties.  ViT-B / ViT-L meet it at every depth.  RANDOM-INIT
# ViT-g does not: its similarity landscape is nearly flat (every S entry within 0.39 .. 0.47), 3 - 6 % of
# the rows have a best-vs-second margin under 2e-4 - inside the noise bf16 operands leave on S (token cosine 0.99987 after
# 40 blocks) - so 1 - 3 % of the arg-max indices flip, every one of them at a margin < 2e-4 (the test still demands that).
MAX_BAD = {"dinov2_vitl14": 0.01, "dinov2_vitg14": 0.035}
# token cosine: north_star's limit is 0.999; regression gates next to the measurements (L-23 0.99997, g-39 0.99983)
MIN_COS = {"dinov2_vitl14": 0.9999, "dinov2_vitg14": 0.9995}


@pytest.mark.parametrize("model_type,layers", [("dinov2_vitl14", (11, 17, 23)), ("dinov2_vitg14", (11, 23, 31, 39))])
def test_full_depth_parity(model_type, layers):
    from vitvs_b200 import Engine, ServoConfig
    last = layers[-1]
    model = dinov2.build(model_type, seed=1)                      # every block of the architecture
    cfg = ServoConfig(model_type=model_type, layer=last)
    eng = Engine(cfg, max_images=2)
    eng.load_weights(model)
    goal, cur, depth = synth.frame_pair(7)
    imgs = torch.from_numpy(np.stack([goal, cur])).cuda()
    pre = torch.cat([servo.preprocess_u8(im) for im in (goal, cur)])
    report = []
    for layer in layers:
        eng.forward_tokens(imgs, layer=layer)
        tok = eng.tokens_out().cpu()
        ref = servo.block_tokens(model, pre, layer)
        cos = torch.nn.functional.cosine_similarity(tok, ref, dim=-1)
        report.append((layer, cos.min().item(), cos.mean().item()))
    print(f"[{model_type}] token cosine by depth: " + "  ".join(f"L{l}: min {a:.6f} mean {b:.6f}" for l, a, b in report))
    for layer, cmin, _ in report:
        assert cmin >= MIN_COS[model_type], f"{model_type}: descriptor cosine {cmin:.6f} < {MIN_COS[model_type]} at layer {layer}"
    # the whole path at the last layer
    out = eng.frame_pairs_to_twist(imgs[:1], imgs[1:], torch.from_numpy(depth[None].copy()).cuda(), want_match=True)
    torch.cuda.synchronize()
    ocfg = dict(servo.DEFAULT_CFG, layer=last)
    ref = servo.frame_pair_to_twist(model, goal, cur, depth, ocfg)
    print(f"[{model_type}] layer {last}: oracle best-vs-second margins below threshold: {margin_histogram(ref['S'])}")
    a1, a2, marg = check_argmax(out["nn1"][0].cpu().numpy(), out["nn2"][0].cpu().numpy(), ref["S"], MAX_BAD[model_type])
    print(f"[{model_type}] layer {last}: nn1 / nn2 agreement {a1:.4f} / {a2:.4f}, worst margin at a disagreement {marg:.1e}")
    check_keypoints(out, 0, 22)
    uv, uvs = out["uv"][0].cpu().numpy().astype(int), out["uv_star"][0].cpu().numpy().astype(int)
    tw_ref = servo.ibvs_twist(uv, uvs, depth)
    tw = out["twist"][0].cpu().numpy()
    assert np.abs(tw - tw_ref).max() <= 1e-9 * np.abs(tw_ref).max()
    if ref["chosen"] is not None and np.array_equal(out["sel_tok"][0].cpu().numpy(), ref["chosen"]):
        rel = np.abs(tw - ref["twist"]).max() / np.abs(ref["twist"]).max()
        print(f"[{model_type}] identical selection, twist rel err {rel:.2e}")
        assert rel <= 1e-3
    eng.close()
