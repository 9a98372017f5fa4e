"""Gram-stencil matching (csrc/gram_stencil.cu): the fused path's similarity of 3x3-binned descriptors computed as a
9-point stencil over the token Gram matrix must agree with (a) the materialised-descriptor path of the same library
and (b) the oracle's cosine similarity of the reference's `_log_bin` descriptors, up to near-tie arg-max flips."""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from oracle import dinov2, servo  # noqa: E402  (the checker)
from vitvs_b200 import synth  # noqa: E402

TIE_EPS = 2e-3


@pytest.fixture(scope="module")
def vits():
    return dinov2.build("dinov2_vits14", seed=0)


def _engine(vits, gram: bool, **kw):
    from vitvs_b200 import Engine, ServoConfig
    old = os.environ.get("VITVS_GRAM")
    os.environ["VITVS_GRAM"] = "1" if gram else "0"
    try:
        e = Engine(ServoConfig(model_type="dinov2_vits14", **kw), max_images=8)
    finally:
        if old is None:
            del os.environ["VITVS_GRAM"]
        else:
            os.environ["VITVS_GRAM"] = old
    e.load_weights(vits)
    return e


def test_gram_stencil_vs_materialised_and_oracle(vits):
    eg, ed = _engine(vits, True), _engine(vits, False)
    assert eg.gram_stencil and not ed.gram_stencil
    n = 4
    pairs = [synth.frame_pair(10 + i) for i in range(n)]
    g = torch.from_numpy(np.stack([p[0] for p in pairs])).cuda()
    c = torch.from_numpy(np.stack([p[1] for p in pairs])).cuda()
    d = torch.from_numpy(np.stack([p[2] for p in pairs])).cuda()
    a = eg.frame_pairs_to_twist(g, c, d, want_match=True)
    b = ed.frame_pairs_to_twist(g, c, d, want_match=True)
    sim_a, sim_b = a["sim1"].cpu().numpy(), b["sim1"].cpu().numpy()
    assert np.abs(sim_a - sim_b).max() < 2e-3                  # bf16 operand rounding on either side
    for k in ("nn1", "nn2"):
        agree = (a[k] == b[k]).float().mean().item()
        print(f"[gram vs materialised] {k} agreement {agree:.4f}")
        assert agree >= 0.99
    # against the oracle (fp32 reference arithmetic): >= 99 % arg-max agreement over the batch, disagreements only at
    # near-ties - and the stencil path is not less accurate than the materialised one (same bf16 operand rounding)
    wrong = {"gram": 0, "materialised": 0}
    for i in range(n):
        ref = servo.frame_pair_to_twist(vits, pairs[i][0], pairs[i][1], pairs[i][2])
        m = ref["match"]
        S = ref["S"].numpy()
        for name, out, sims in (("gram", a, sim_a), ("materialised", b, sim_b)):
            nn1 = out["nn1"][i].cpu().numpy()
            diff = np.nonzero(nn1 != m.nn1)[0]
            wrong[name] += len(diff)
            for t in diff:
                assert S[t, m.nn1[t]] - S[t, nn1[t]] < TIE_EPS, (name, i, t)
            assert np.abs(sims[i] - np.asarray(m.sim1)).max() < 5e-3
    total = n * a["nn1"].shape[1]
    print(f"[vs oracle] arg-max disagreements of {total}: {wrong}")
    assert wrong["gram"] <= 0.01 * total and wrong["materialised"] <= 0.01 * total
    assert wrong["gram"] <= 2 * wrong["materialised"] + 3
    # twists: identical keypoints => identical control law; where a near-tie flipped a keypoint they may differ
    same_kp = (a["uv"] == b["uv"]).all(dim=(1, 2)) & (a["uv_star"] == b["uv_star"]).all(dim=(1, 2))
    assert torch.equal(a["twist"][same_kp], b["twist"][same_kp])
    eg.close()
    ed.close()


def test_gram_stencil_tiny_and_unbinned(vits):
    """1x1 ... 3x3 grids (every neighbour clamps) and the un-binned configuration (falls back to the direct product)."""
    for size in (14, 28, 42):
        eg = _engine(vits, True, dino_input_size=size)
        ed = _engine(vits, False, dino_input_size=size)
        assert eg.gram_stencil
        rng = np.random.default_rng(size)
        g = torch.from_numpy(rng.integers(0, 256, (2, size, size, 3), dtype=np.uint8)).cuda()
        c = torch.from_numpy(rng.integers(0, 256, (2, size, size, 3), dtype=np.uint8)).cuda()
        d = torch.from_numpy(np.stack([synth.depth_frame(0)] * 2)).cuda()
        a = eg.frame_pairs_to_twist(g, c, d, want_match=True)
        b = ed.frame_pairs_to_twist(g, c, d, want_match=True)
        assert (a["sim1"] - b["sim1"]).abs().max().item() < 2e-3
        eg.close()
        ed.close()
    e = _engine(vits, True, use_feature_binning=False)
    assert not e.gram_stencil
    e.close()
