"""Parity AT the benchmarked configuration (BASELINE.json configs[1]: ViT-B/14, 64 frame pairs per call, S = 308,
layer 11, binned, K = 24) - the shapes only the bench used to exercise: M = 62 080 rows (243 persistent tiles per CTA
pair), the evict-last pinning of the residual stream (active above 100 MB), reversed image order in attention, CUDA
graph replay and the host-buffer pipeline.

  * every one of the 64 pairs, through the device entry (eager, graph capture, graph replay), the match-reporting
    (eager) entry and the host-buffer entry, is BIT-IDENTICAL to the same pair run alone through a 1-pair handle;
  * four of them (first, two middle, last) are checked against the fp32 oracle with the north_star tolerances;
  * key-points (a11) are re-derived from the GPU's own selection with oracle/servo.calculate_uv for all 64.
"""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from oracle import dinov2, servo  # noqa: E402  (the checker)
from vitvs_b200 import synth  # noqa: E402

from _parity import check_argmax, check_keypoints  # noqa: E402

N_PAIRS = 64


@pytest.fixture(scope="module")
def vitb():
    return dinov2.build("dinov2_vitb14", seed=0)


@pytest.fixture(scope="module")
def batch():
    pairs = [synth.frame_pair(i) for i in range(N_PAIRS)]              # 64 DISTINCT pairs
    goal = torch.from_numpy(np.stack([p[0] for p in pairs]))
    cur = torch.from_numpy(np.stack([p[1] for p in pairs]))
    depth = torch.from_numpy(np.stack([p[2] for p in pairs]))
    return pairs, goal, cur, depth


def test_bench_configuration_equals_single_pair_calls_and_oracle(vitb, batch):
    from vitvs_b200 import Engine, ServoConfig
    pairs, goal, cur, depth = batch
    cfg = ServoConfig(model_type="dinov2_vitb14")
    big = Engine(cfg, max_images=2 * N_PAIRS)
    big.load_weights(vitb)
    g, c, d = goal.cuda(), cur.cuda(), depth.cuda()

    # device entry: 1st call eager, 2nd captures the graphs, 3rd replays them
    runs = [big.frame_pairs_to_twist(g, c, d) for _ in range(3)]
    torch.cuda.synchronize()
    launches_graph = big.last_launch_count()
    for r in runs[1:]:
        for k in ("twist", "n_selected", "mean_sim_sel"):
            assert torch.equal(r[k], runs[0][k]), f"graph replay differs from the eager pass in {k}"
    # the match-reporting entry (eager launches, per-token outputs)
    full = big.frame_pairs_to_twist(g, c, d, want_match=True)
    torch.cuda.synchronize()
    tok_big = big.tokens_out().cpu()                                    # [128, 485, 768]: goal images, then current
    for k in ("twist", "n_selected", "mean_sim_sel"):
        assert torch.equal(full[k], runs[0][k]), k
    # host-buffer entry (pinned memory, H2D + path + D2H inside the call)
    tw = torch.empty(N_PAIRS, 6, dtype=torch.float64).pin_memory()
    ns = torch.empty(N_PAIRS, dtype=torch.int32).pin_memory()
    ms = torch.empty(N_PAIRS, dtype=torch.float32).pin_memory()
    gh, ch, dh = goal.pin_memory(), cur.pin_memory(), depth.pin_memory()
    for _ in range(2):
        big.frame_pairs_to_twist_host(gh, ch, dh, tw, ns, ms)
    assert torch.equal(tw, runs[0]["twist"].cpu()) and torch.equal(ns, runs[0]["n_selected"].cpu())
    assert torch.equal(ms, runs[0]["mean_sim_sel"].cpu())

    # the same 64 pairs, one at a time, through a 1-pair handle (no pinning branch, 4-tile GEMMs, its own graph)
    one = Engine(cfg, max_images=2)
    one.load_weights(vitb)
    keys = ("twist", "n_selected", "mean_sim_sel", "nn1", "nn2", "sim1", "candidate", "sel_tok", "sel_sim", "uv", "uv_star",
            "flags", "n_candidates")
    worst = 0.0
    for i in range(N_PAIRS):
        s = one.frame_pairs_to_twist(g[i:i + 1], c[i:i + 1], d[i:i + 1], want_match=True)
        for k in keys:
            assert torch.equal(s[k][0], full[k][i]), f"pair {i}: {k} differs between the 64-pair call and the 1-pair call"
        tw1 = s["twist"][0].cpu().numpy()
        worst = max(worst, float(np.abs(tw1 - full["twist"][i].cpu().numpy()).max()))
        n = check_keypoints(full, i, 22)                                 # a11, unconditional
        assert n == 24
    print(f"[bench config] 64-pair call == 64 single-pair calls bit for bit (max twist diff {worst:.1e}); "
          f"{launches_graph} kernels per 64-pair graph replay")

    # four pairs against the fp32 oracle
    for i in (0, 21, 42, 63):
        gi, ci, di = pairs[i]
        ref_tok = torch.cat([servo.block_tokens(vitb, servo.preprocess_u8(im), 11) for im in (gi, ci)])
        mine = torch.stack([tok_big[i], tok_big[N_PAIRS + i]])
        cos = torch.nn.functional.cosine_similarity(mine, ref_tok, dim=-1).min().item()
        ref = servo.frame_pair_to_twist(vitb, gi, ci, di)
        a1, a2, marg = check_argmax(full["nn1"][i].cpu().numpy(), full["nn2"][i].cpu().numpy(), ref["S"])
        uv, uvs = full["uv"][i].cpu().numpy().astype(int), full["uv_star"][i].cpu().numpy().astype(int)
        tw_ref = servo.ibvs_twist(uv, uvs, di)
        twi = full["twist"][i].cpu().numpy()
        rel_same_kp = float(np.abs(twi - tw_ref).max() / np.abs(tw_ref).max())
        same_sel = np.array_equal(full["sel_tok"][i].cpu().numpy(), ref["chosen"])
        rel = float(np.abs(twi - ref["twist"]).max() / np.abs(ref["twist"]).max())
        print(f"[bench config] pair {i}: token cosine min {cos:.6f}  nn1/nn2 agree {a1:.4f}/{a2:.4f} (worst margin "
              f"{marg:.1e})  twist rel err on the GPU's key-points {rel_same_kp:.1e}; vs the oracle's own twist "
              f"{rel:.1e} ({'identical' if same_sel else 'different'} selection)")
        assert cos >= 0.9999          # north_star 0.999; regression gate (measured 0.99998)
        assert rel_same_kp <= 1e-9
        if same_sel:
            assert rel <= 1e-3
    big.close()
    one.close()


def test_repeated_replays_are_bit_identical(vitb, batch):
    """Race evidence in place of compute-sanitizer (closed on this pool, profiles/r2_sanitizer_unavailable.txt): hand-rolled
    mbarrier rings, tcgen05.commit multicast, TMA reduce-adds into the live fp32 residual stream and smem staging reused
    behind bulk_wait_read are all order-sensitive - a lost wait shows up as run-to-run differences.  Twelve replays of the
    full 64-pair step (mixed-cluster GEMMs, 4 CTAs / SM attention) must reproduce every residual-stream bit."""
    from vitvs_b200 import Engine, ServoConfig
    _, goal, cur, depth = batch
    eng = Engine(ServoConfig(model_type="dinov2_vitb14"), max_images=2 * N_PAIRS)
    eng.load_weights(vitb)
    g, c, d = goal.cuda(), cur.cuda(), depth.cuda()
    ref_tok = ref_tw = None
    for i in range(12):
        out = eng.frame_pairs_to_twist(g, c, d)
        tok = eng.tokens_out()
        if ref_tok is None:
            ref_tok, ref_tw = tok.clone(), out["twist"].clone()
        else:
            assert torch.equal(tok, ref_tok), f"replay {i}: residual stream differs"
            assert torch.equal(out["twist"], ref_tw), f"replay {i}: twists differ"
    assert torch.isfinite(ref_tok).all()
    eng.close()
