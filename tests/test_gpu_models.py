"""Whole-path parity for the other DINOv2 sizes of BASELINE.json's configs (ViT-B/14, ViT-L/14, ViT-g/14 with
the SwiGLU FFN), one synthetic frame pair each, against the fp32 oracle on the host; plus the matching +
IBVS-only sweep of config 5 at a large token count."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from oracle import dinov2, servo  # noqa: E402
from vitvs_b200 import synth  # noqa: E402

from _parity import check_keypoints  # noqa: E402

TIE_EPS = 1e-3     # regression gate (tests/_parity.py); worst measured margin at a disagreement: 4e-4


@pytest.mark.parametrize("model_type", ["dinov2_vitb14", "dinov2_vitl14", "dinov2_vitg14"])
def test_model_family_parity(model_type):
    from vitvs_b200 import Engine, ServoConfig
    torch.set_num_threads(max(1, torch.get_num_threads()))
    model = dinov2.build(model_type, seed=1, max_depth=12)        # blocks[0..11] are all the path evaluates
    eng = Engine(ServoConfig(model_type=model_type), max_images=2)
    eng.load_weights(model)
    goal, cur, depth = synth.frame_pair(7)
    out = eng.frame_pairs_to_twist(torch.from_numpy(goal[None].copy()).cuda(), torch.from_numpy(cur[None].copy()).cuda(),
                                   torch.from_numpy(depth[None].copy()).cuda(), want_match=True)
    torch.cuda.synchronize()
    tok = eng.tokens_out().cpu()
    ref_tok = torch.cat([servo.block_tokens(model, servo.preprocess_u8(im), 11) for im in (goal, cur)])
    cos = torch.nn.functional.cosine_similarity(tok, ref_tok, dim=-1)
    print(f"[{model_type}] token cosine min {cos.min():.6f}")
    # north_star: descriptor cosine >= 0.999; regression gate: 0.9999 (measured 0.99998), ViT-g 0.9995 (0.99990)
    assert cos.min() >= (0.9995 if model_type == "dinov2_vitg14" else 0.9999)
    ref = servo.frame_pair_to_twist(model, goal, cur, depth)
    S, m = ref["S"], ref["match"]
    nn1, nn2 = out["nn1"][0].cpu().numpy(), out["nn2"][0].cpu().numpy()
    rows = np.arange(S.shape[0])
    bad1, bad2 = nn1 != m.nn1, nn2 != m.nn2
    marg1 = (S[rows, m.nn1] - S[rows, nn1]).numpy()
    marg2 = (S[m.nn2, rows] - S[nn2, rows]).numpy()
    print(f"[{model_type}] nn1 agree {1 - bad1.mean():.4f} nn2 agree {1 - bad2.mean():.4f} "
          f"worst margin {max(marg1.max(), marg2.max()):.2e}")
    # >= 99 % mutual-NN index agreement, per direction (north_star).  Random-init ViT-g is the stated exception (DESIGN.md 5,
    # tests/test_gpu_depth.py) at FULL depth only; at layer 11 it measures 0.6 - 0.8 % and is held to the same 1 %
    from _parity import margin_histogram
    print(f"[{model_type}] oracle best-vs-second margins below threshold: {margin_histogram(S)}")
    limit = 0.01
    assert bad1.mean() <= limit and bad2.mean() <= limit
    assert np.all(marg1[bad1] < TIE_EPS) and np.all(marg2[bad2] < TIE_EPS)   # ... disagreements only at near-ties
    uv, uvs = out["uv"][0].cpu().numpy().astype(int), out["uv_star"][0].cpu().numpy().astype(int)
    tw_ref = servo.ibvs_twist(uv, uvs, depth)
    tw = out["twist"][0].cpu().numpy()
    assert np.abs(tw - tw_ref).max() <= 1e-9 * np.abs(tw_ref).max()
    check_keypoints(out, 0, 22)
    if np.array_equal(out["sel_tok"][0].cpu().numpy(), ref["chosen"]):
        rel = np.abs(tw - ref["twist"]).max() / np.abs(ref["twist"]).max()
        print(f"[{model_type}] identical selection, twist rel err {rel:.2e}")
        assert rel <= 1e-3                                                   # twist relative error <= 1e-3
    eng.close()


@pytest.mark.parametrize("T,d", [(4096, 384), (2304, 768)])
def test_matching_sweep_large_grid(T, d):
    """BASELINE config 5: i.i.d. Gaussian descriptors, planted permutation + noise, unbinned."""
    from vitvs_b200 import Engine, ServoConfig
    eng = Engine(ServoConfig(model_type="dinov2_vits14"), max_images=4, max_match_tokens=2 * T, max_match_dim=d)
    a, b, perm = synth.descriptor_sets(T, d, seed=7, noise=0.5)
    r = eng.match(torch.stack([a, a]), torch.stack([b, b]), num_pairs=24)          # two identical episodes
    torch.cuda.synchronize()
    S = servo.cosine_similarity_matrix(a[None, None], b[None, None])[0, 0]
    m = servo.match(S)
    nn1, nn2 = r["nn1"][0].cpu().numpy(), r["nn2"][0].cpu().numpy()
    assert (nn1 == m.nn1).mean() >= 0.99 and (nn2 == m.nn2).mean() >= 0.99
    inv = np.empty(T, dtype=np.int64)
    inv[perm.numpy()] = np.arange(T)
    assert (nn1 == inv).mean() >= 0.99                                           # the planted matches are recovered
    assert torch.equal(r["nn1"][0], r["nn1"][1]) and torch.equal(r["uv"][0], r["uv"][1])
    grid = int(np.sqrt(T))
    uv = r["uv"][0].cpu().numpy()
    assert uv[:, 0].max() < 640 and uv[:, 1].max() < 480 and uv.min() >= 0
    check_keypoints(r, 0, grid)
    eng.close()


@pytest.mark.parametrize("T,d", [(16384, 768), (8192, 3456)])
def test_matching_maximum_grid(T, d):
    """SURVEY §8d C5 at its largest sizes (128 x 128 tokens unbinned ViT-B; 8192 tokens with ViT-S binned width):
    the similarity matrix (1 GiB at T = 16 384) is never written; arg-max parity against the CPU oracle computed in
    row blocks, and the planted permutation must come back."""
    from vitvs_b200 import Engine, ServoConfig
    eng = Engine(ServoConfig(model_type="dinov2_vits14"), max_images=4, max_match_tokens=T, max_match_dim=d)
    a, b, perm = synth.descriptor_sets(T, d, seed=7, noise=0.5)
    # 16 current tokens carry unrelated content: with every token mutual the reference's min/max normalisation
    # divides by zero and it returns (None, None, None) (vitvs_v2.py:123-124,155) - covered by the "all_mutual" fixture
    b[:16] = torch.randn(16, d, generator=torch.Generator().manual_seed(99))
    r = eng.match(a[None], b[None], num_pairs=24)
    torch.cuda.synchronize()
    nn1, nn2 = r["nn1"][0].cpu().numpy(), r["nn2"][0].cpu().numpy()
    inv = np.empty(T, dtype=np.int64)
    inv[perm.numpy()] = np.arange(T)
    planted = np.ones(T, dtype=bool)
    planted[perm.numpy()[:16]] = False
    assert (nn1 == inv)[planted].mean() >= 0.999 and (nn2 == perm.numpy())[16:].mean() >= 0.999
    assert T - 48 <= int(r["n_candidates"][0]) < T
    # oracle arg-max (fp32 cosine similarity, vitvs_v2.py:49-56,80-81) in row blocks of 2048
    an = a / a.norm(dim=1, keepdim=True).clamp_min(1e-8)
    bn = b / b.norm(dim=1, keepdim=True).clamp_min(1e-8)
    ref1 = torch.cat([(an[i:i + 2048] @ bn.T).argmax(dim=1) for i in range(0, T, 2048)]).numpy()
    ref2 = torch.cat([(bn[i:i + 2048] @ an.T).argmax(dim=1) for i in range(0, T, 2048)]).numpy()
    assert (nn1 == ref1).mean() >= 0.99 and (nn2 == ref2).mean() >= 0.99
    # every selected pair is mutual and its key-points lie inside the frame
    sel = r["sel_tok"][0].cpu().numpy()
    sel = sel[sel >= 0]
    assert len(sel) == 24 and (nn2[nn1[sel]] == sel).all()
    uv = r["uv"][0].cpu().numpy()
    assert uv[:, 0].max() < 640 and uv[:, 1].max() < 480 and uv.min() >= 0
    check_keypoints(r, 0, int(np.sqrt(T)))          # 128 x 128: scale 2.40625; T = 8192: width int(sqrt(T)) = 90 (:75)
    eng.close()


@pytest.mark.parametrize("model_type,pair", [("dinov2_vitb14", 1), ("dinov2_vitb14", 3), ("dinov2_vitg14", 2)])
def test_perturbed_frame_parity(model_type, pair):
    """BASELINE config 4's input distribution (generate_perturbed_models.py:68-77 through synth.perturb): the current frame
    carries an erased rectangle (pairs 1 and 2: a quarter / a sixth of the frame blanked before jitter and noise - large
    areas of near-identical patches, the worst case for arg-max ties) or only colour jitter + noise (pair 3).  Same gates
    as the clean frames: token cosine, >= 99 % arg-max agreement per direction with disagreements at near-ties only, and
    the reference's key-point / twist arithmetic on the GPU's own selection."""
    from vitvs_b200 import Engine, ServoConfig
    from _parity import check_argmax, margin_histogram
    model = dinov2.build(model_type, seed=1, max_depth=12)
    eng = Engine(ServoConfig(model_type=model_type), max_images=2)
    eng.load_weights(model)
    goal, cur, depth = synth.frame_pair(pair, perturbed=True)
    clean = synth.frame_pair(pair)[1]
    print(f"[{model_type} perturbed pair {pair}] pixels changed by > 64 levels: "
          f"{(np.abs(cur.astype(int) - clean.astype(int)).max(-1) > 64).mean():.3f}")
    out = eng.frame_pairs_to_twist(torch.from_numpy(goal[None].copy()).cuda(), torch.from_numpy(cur[None].copy()).cuda(),
                                   torch.from_numpy(depth[None].copy()).cuda(), want_match=True)
    torch.cuda.synchronize()
    tok = eng.tokens_out().cpu()
    ref_tok = torch.cat([servo.block_tokens(model, servo.preprocess_u8(im), 11) for im in (goal, cur)])
    cos = torch.nn.functional.cosine_similarity(tok, ref_tok, dim=-1)
    assert cos.min() >= (0.9995 if model_type == "dinov2_vitg14" else 0.9999), cos.min()
    ref = servo.frame_pair_to_twist(model, goal, cur, depth)
    a1, a2, marg = check_argmax(out["nn1"][0].cpu().numpy(), out["nn2"][0].cpu().numpy(), ref["S"], 0.01)
    print(f"[{model_type} perturbed pair {pair}] token cosine min {cos.min():.6f}  nn1 / nn2 agreement {a1:.4f} / {a2:.4f} "
          f"(worst margin at a disagreement {marg:.1e}); oracle margins {margin_histogram(ref['S'])}")
    check_keypoints(out, 0, 22)
    uv, uvs = out["uv"][0].cpu().numpy().astype(int), out["uv_star"][0].cpu().numpy().astype(int)
    n_sel = int(out["n_selected"][0])
    if n_sel >= 4:
        tw_ref = servo.ibvs_twist(uv, uvs, depth)
        tw = out["twist"][0].cpu().numpy()
        assert np.abs(tw - tw_ref).max() <= 1e-9 * np.abs(tw_ref).max()
    if ref["chosen"] is not None and np.array_equal(out["sel_tok"][0].cpu().numpy(), ref["chosen"]):
        rel = np.abs(out["twist"][0].cpu().numpy() - ref["twist"]).max() / np.abs(ref["twist"]).max()
        assert rel <= 1e-3
    eng.close()


def _flips(nn_a1, nn_a2, S):
    """arg-max disagreement of index lists (nn_a1 rows, nn_a2 columns) with similarity matrix S, and the worst margin S
    itself sees at a disagreement"""
    S = S.numpy() if hasattr(S, "numpy") else S
    rows = np.arange(S.shape[0])
    r1, r2 = S.argmax(1), S.argmax(0)
    b1, b2 = nn_a1 != r1, nn_a2 != r2
    m1, m2 = S[rows, r1] - S[rows, nn_a1], S[r2, rows] - S[nn_a2, rows]
    return float(b1.mean()), float(b2.mean()), float(max(m1[b1].max(initial=0), m2[b2].max(initial=0)))


@pytest.mark.parametrize("model_type,layer", [("dinov2_vitb14", 11), ("dinov2_vitl14", 23), ("dinov2_vitg14", 11),
                                              ("dinov2_vitg14", 39)])
def test_gpu_against_bf16_emulated_oracle(model_type, layer):
    """Attribution of the fp32-vs-GPU differences (oracle/bf16_emulation.py): the CPU emulation rounds where the GPU
    path rounds.  (1) GPU vs emulation must be much tighter than GPU vs fp32 - what is left is accumulation order and
    exp / erf implementations, so a kernel defect shows here even where the fp32 comparison is inside its tolerance.
    (2) emulation vs fp32 must flip about as many arg-max indices as the GPU does - the ViT-g exception to north_star's
    99 % (tests/test_gpu_depth.py) is then arithmetic, not a kernel."""
    from oracle import bf16_emulation as be
    from vitvs_b200 import Engine, ServoConfig
    model = dinov2.build(model_type, seed=1, max_depth=layer + 1)
    eng = Engine(ServoConfig(model_type=model_type, layer=layer), max_images=2)
    eng.load_weights(model)
    goal, cur, depth = synth.frame_pair(7)
    out = eng.frame_pairs_to_twist(torch.from_numpy(goal[None].copy()).cuda(), torch.from_numpy(cur[None].copy()).cuda(),
                                   torch.from_numpy(depth[None].copy()).cuda(), want_match=True)
    torch.cuda.synchronize()
    tok = eng.tokens_out().cpu()
    nn1, nn2 = out["nn1"][0].cpu().numpy(), out["nn2"][0].cpu().numpy()
    eng.close()
    tok_e, S_e = be.similarity_matrices(model, goal, cur, layer)
    ref = servo.frame_pair_to_twist(model, goal, cur, depth, dict(servo.DEFAULT_CFG, layer=layer))
    tok_f = torch.cat([servo.block_tokens(model, servo.preprocess_u8(im), layer) for im in (goal, cur)])
    cosine = torch.nn.functional.cosine_similarity
    cos_ge, cos_gf, cos_ef = (cosine(a, b, dim=-1).min().item() for a, b in ((tok, tok_e), (tok, tok_f), (tok_e, tok_f)))
    ge1, ge2, ge_m = _flips(nn1, nn2, S_e)                                        # GPU indices judged by the emulation's S
    gf1, gf2, gf_m = _flips(nn1, nn2, ref["S"])                                   # ... by the fp32 oracle's S
    ef1, ef2, ef_m = _flips(S_e.argmax(1).numpy(), S_e.argmax(0).numpy(), ref["S"])   # emulation's indices by fp32's S
    print(f"[{model_type} L{layer}] token cosine min: GPU~emul {cos_ge:.7f}  GPU~fp32 {cos_gf:.7f}  emul~fp32 {cos_ef:.7f}")
    print(f"[{model_type} L{layer}] arg-max flips nn1/nn2: GPU~emul {ge1:.4f}/{ge2:.4f} (worst margin {ge_m:.1e})  "
          f"GPU~fp32 {gf1:.4f}/{gf2:.4f} ({gf_m:.1e})  emul~fp32 {ef1:.4f}/{ef2:.4f} ({ef_m:.1e})")
    # (1) The GPU is closer to the emulation than to fp32 (measured 1 - cos: 0.42 - 0.66 of the fp32 figure).  Not 10x closer:
    # the flash-attention kernel rounds P = exp(s - RUNNING max) and rescales later, the emulation rounds exp(s - final max),
    # so the P rounding noise is independent between the two; x, W, qkv, attention-output and GELU roundings are shared.
    assert 1 - cos_ge <= 0.8 * (1 - cos_gf)
    assert ge1 + ge2 <= gf1 + gf2 + 3 / 484 and ge_m < TIE_EPS
    # (2) rounding alone (no kernel involved) flips as many indices against fp32 as the GPU does
    assert ef1 + ef2 >= 0.5 * (gf1 + gf2) - 2 / 484
    if (model_type, layer) == ("dinov2_vitg14", 39):
        # the stated exception to north_star's 99 %: the CPU emulation misses it too, by the same amount (2.3 % / 2.3 %
        # against the GPU's 2.9 % / 2.3 %), every flip at a margin the fp32 oracle itself calls a near-tie
        assert max(ef1, ef2) > 0.01 and ef_m < TIE_EPS
