"""Shared checks of the GPU path's outputs against the oracle (imported by the -m gpu tests only)."""
import numpy as np

from oracle import servo

# Two limits: north_star's acceptance numbers (cosine >= 0.999, near-tie) and, tighter, REGRESSION gates a factor ~2-5 off
# what the path measures (cosine 0.99998 S/B, 0.99997 L-23, 0.99983 g-39; worst margin at a disagreement 4e-4), so that a
# defect costing a factor 10 in accuracy (wrong ln_eps, a positional row off, a dropped k-block) cannot stay green.
TIE_EPS = 1e-3     # similarity margin below which an arg-max disagreement counts as a near-tie (bf16 operands)


def check_keypoints(out: dict, p: int, grid_w: int, cfg: dict | None = None) -> int:
    """a11, unconditionally: the key-points the GPU reports for pair ``p`` must be exactly what the reference's
    arithmetic (oracle/servo.calculate_uv, pinned against vitvs_v2.py:511-513 + :525-553) makes of the GPU's OWN
    selected goal tokens and their arg-max partners.  Returns the number of selected pairs."""
    cfg = dict(servo.DEFAULT_CFG) if cfg is None else cfg
    sel = out["sel_tok"][p].cpu().numpy()
    n = int(out["n_selected"][p])
    K = sel.shape[0]
    assert (sel[:n] >= 0).all() and (sel[n:] == -1).all()
    nn1 = out["nn1"][p].cpu().numpy()
    same = bool(int(out["flags"][p]) & 1)
    t1 = sel[:n].astype(np.int64)
    t2 = t1 if same else nn1[t1].astype(np.int64)
    T = nn1.shape[0]
    p1 = np.stack([t1 // grid_w, t1 % grid_w], -1)
    p2 = np.stack([t2 // grid_w, t2 % grid_w], -1)
    # oracle/servo.calculate_uv derives the scale from int(sqrt(T)) like the reference (:75, :511)
    assert int(np.sqrt(T)) == grid_w
    uv_star_ref, uv_ref = servo.calculate_uv(p1, p2, T, dict(cfg, num_pairs=K))
    assert np.array_equal(out["uv_star"][p].cpu().numpy(), uv_star_ref), "goal key-points differ from calculate_uv"
    assert np.array_equal(out["uv"][p].cpu().numpy(), uv_ref), "current key-points differ from calculate_uv"
    return n


def check_argmax(nn1, nn2, S_ref, max_bad_frac: float = 0.01):
    """>= 99 % index agreement per direction, disagreements only where the oracle's own margin is a near-tie."""
    S = S_ref.numpy() if hasattr(S_ref, "numpy") else np.asarray(S_ref)
    T = S.shape[0]
    rows = np.arange(T)
    r1, r2 = S.argmax(1), S.argmax(0)
    bad1, bad2 = nn1 != r1, nn2 != r2
    marg1 = S[rows, r1] - S[rows, nn1]
    marg2 = S[r2, rows] - S[nn2, rows]
    allowed = max(1, int(np.ceil(max_bad_frac * T)))
    assert np.all(marg1[bad1] < TIE_EPS) and np.all(marg2[bad2] < TIE_EPS)       # only at near-ties, whatever the count
    assert bad1.sum() <= allowed and bad2.sum() <= allowed, (int(bad1.sum()), int(bad2.sum()), allowed)
    return 1 - bad1.mean(), 1 - bad2.mean(), float(max(marg1[bad1].max(initial=0), marg2[bad2].max(initial=0)))


def margin_histogram(S_ref) -> str:
    """How flat the oracle's own similarity landscape is: share of rows / columns whose best-vs-second-best margin lies
    below a few thresholds (a disagreement can only happen where that margin is within the bf16 noise of S)."""
    S = S_ref.numpy() if hasattr(S_ref, "numpy") else np.asarray(S_ref)
    top_r = np.sort(S, axis=1)[:, -2:]
    top_c = np.sort(S, axis=0)[-2:, :]
    m = np.concatenate([top_r[:, 1] - top_r[:, 0], top_c[1] - top_c[0]])
    return "  ".join(f"<{t:.0e}: {100 * (m < t).mean():.1f}%" for t in (5e-5, 1e-4, 2e-4, 5e-4, 1e-3, 2e-3))
