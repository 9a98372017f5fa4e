"""Generate tests/golden/*.npz by EXECUTING THE REFERENCE'S OWN CODE from /root/reference.

TEST INFRASTRUCTURE ONLY.  Run in the build container (the only place /root/reference exists):

    python -m oracle.make_golden

The reference has no tests / golden vectors of its own, so these fixtures are
outputs of the reference itself (vitvs_v2.py / dinov2_extractor.py imported unmodified through
oracle/ref_loader.py) on seeded inputs.  They travel to the GPU box, where /root/reference
does not exist, and pin the oracle (tests/test_oracle_golden.py) and the CUDA path
(tests/test_gpu_*.py).  The backbone under the reference's ViTExtractor is the restated
DINOv2 (oracle/dinov2.py) because the hub model is unobtainable - see oracle/__init__.py.
"""
from __future__ import annotations

import os
import warnings

import numpy as np
import torch

from oracle import dinov2, ref_loader, servo
from vitvs_b200 import synth

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")


def _ref_full_match(desc1, desc2):
    """Reference find_correspondences_batch with num_pairs = T => the full candidate set
    (vitvs_v2.py:72-155); order is random (randperm) so it is stored sorted by goal token."""
    v = ref_loader.vitvs()
    T = desc1.shape[2]
    w = int(np.sqrt(T))
    S = v.chunk_cosine_sim(desc1, desc2)
    sim1, nn1 = torch.max(S, dim=-1)
    _, nn2 = torch.max(S, dim=-2)
    p1, p2, sim = v.find_correspondences_batch(desc1, desc2, num_pairs=T)
    if p1 is None:
        tok1 = tok2 = np.zeros(0, np.int64)
        simsel = np.zeros(0, np.float32)
        none = True
    else:
        tok1 = (p1[:, 0] * w + p1[:, 1]).numpy()
        tok2 = (p2[:, 0] * w + p2[:, 1]).numpy()
        simsel = sim.reshape(-1).numpy()
        o = np.argsort(tok1, kind="stable")
        tok1, tok2, simsel = tok1[o], tok2[o], simsel[o]
        none = False
    return dict(S=S[0, 0].numpy(), nn1=nn1[0, 0].numpy(), nn2=nn2[0, 0].numpy(), sim1=sim1[0, 0].numpy(),
                mean_sim1=np.float32(sim1.mean().item()), sel_tok1=tok1, sel_tok2=tok2, sel_sim=simsel,
                returned_none=np.bool_(none))


def golden_matching():
    cases = {}
    g = torch.Generator().manual_seed(11)
    # generic: planted permutation + noise, 8x8 grid
    a = torch.randn(64, 48, generator=g)
    perm = torch.randperm(64, generator=g)
    b = a[perm] + 2.0 * torch.randn(64, 48, generator=g)
    cases["planted"] = (a, b)
    # unrelated sets: few mutual NNs
    cases["unrelated"] = (torch.randn(100, 16, generator=g), torch.randn(100, 16, generator=g))
    # all tokens mutual (pure permutation, low similarity elsewhere) => reference returns None (SURVEY §4 v)
    a = torch.randn(49, 64, generator=g)
    cases["all_mutual"] = (a, a[torch.randperm(49, generator=g)] + 0.7 * torch.randn(49, 64, generator=g))
    # same image => random identical points, sim = 1 (vitvs_v2.py:84-101)
    a = torch.randn(36, 32, generator=g)
    cases["same_image"] = (a, a.clone())
    # exact ties: duplicated descriptors => argmax must take the lowest index
    a = torch.randn(64, 24, generator=g)
    b = torch.randn(64, 24, generator=g)
    b[10] = b[3]
    b[40] = b[3]
    a[7] = a[5]
    cases["ties"] = (a, b)
    out = {}
    for name, (d1, d2) in cases.items():
        r = _ref_full_match(d1[None, None].contiguous(), d2[None, None].contiguous())
        out[f"{name}/desc1"] = d1.numpy()
        out[f"{name}/desc2"] = d2.numpy()
        for k, val in r.items():
            out[f"{name}/{k}"] = val
    np.savez_compressed(os.path.join(OUT, "matching.npz"), **out)
    return out


def golden_logbin():
    g = torch.Generator().manual_seed(5)
    m = dinov2.build("dinov2_vits14", seed=0)
    ex = ref_loader.make_extractor(m, "dinov2_vits14")
    out = {}
    for name, (hp, wp, d) in {"5x6": (5, 6, 8), "3x3": (3, 3, 4), "1x4": (1, 4, 4), "22x22": (22, 22, 6)}.items():
        x = torch.randn(2, 1, hp * wp, d, generator=g)
        ex.num_patches = (hp, wp)
        y = ex._log_bin(x)                                   # dinov2_extractor.py:265-311
        out[f"{name}/x"] = x.numpy()
        out[f"{name}/y"] = y.numpy()
        out[f"{name}/grid"] = np.array([hp, wp])
    np.savez_compressed(os.path.join(OUT, "logbin.npz"), **out)
    return out


def golden_ibvs():
    cfg = dict(servo.DEFAULT_CFG)
    rng = np.random.default_rng(3)
    out = {}
    depth = synth.depth_frame(0)
    for name, n in {"k24": 24, "k48": 48, "k7_padded": 7, "k3_too_few": 3, "k24_oob_depth": 24}.items():
        K = 48 if name == "k48" else 24
        cfg["num_pairs"] = K
        c = ref_loader.make_controller(cfg, depth)
        p1 = rng.integers(0, 22, size=(n, 2))
        p2 = np.clip(p1 + rng.integers(-3, 4, size=(n, 2)), 0, 21)
        scale = cfg["dino_input_size"] / 22
        s_uv_star, s_uv = c.calculate_uv((p1 * scale + scale / 2).tolist(), (p2 * scale + scale / 2).tolist())
        if name == "k24_oob_depth":      # not reachable from calculate_uv; exercises get_depth's bounds branch
            s_uv = s_uv.copy()
            s_uv[0] = (700, 10)
            s_uv[1] = (5, -3)
        Z = c.get_depth(s_uv)
        s_xy, s_star = c.transform_to_real_world(s_uv, s_uv_star)
        L = c.calculate_interaction_matrix(s_xy, Z)
        e = (s_xy - s_star).reshape((len(s_xy) * 2, 1))
        v = -c.lambda_ * np.linalg.pinv(L.astype("float")) @ e            # vitvs_v2.py:622
        ema1 = np.array([c.update_ema(i, x) for i, x in enumerate(v.flatten())])
        ema2 = np.array([c.update_ema(i, x) for i, x in enumerate((0.5 * v).flatten())])
        for k, val in dict(p1=p1, p2=p2, s_uv=s_uv, s_uv_star=s_uv_star, Z=Z, L=L, e=e, v=v.flatten(),
                           ema1=ema1, ema2=ema2, num_pairs=np.array(K)).items():
            out[f"{name}/{k}"] = val
    out["depth"] = depth
    np.savez_compressed(os.path.join(OUT, "ibvs.npz"), **out)
    return out


@torch.no_grad()
def golden_e2e(model_type="dinov2_vits14", n_pairs=2):
    """Whole path through the reference's own functions on synthetic pairs (config C1)."""
    cfg = dict(servo.DEFAULT_CFG)
    v = ref_loader.vitvs()
    m = dinov2.build(model_type, seed=0)
    ex = ref_loader.make_extractor(m, model_type)
    out = {}
    for i in range(n_pairs):
        g, c, depth = synth.frame_pair(i, cfg["dino_input_size"])
        d1 = ex.extract_descriptors(ex.preprocess_pil(__import__("PIL.Image").Image.fromarray(g)), layer=11,
                                    facet="token", bin=True)
        d2 = ex.extract_descriptors(ex.preprocess_pil(__import__("PIL.Image").Image.fromarray(c)), layer=11,
                                    facet="token", bin=True)
        r = _ref_full_match(d1, d2)
        mo = servo.match(torch.from_numpy(r["S"]), cfg)
        chosen = servo.top_k_by_similarity(mo, cfg["num_pairs"])
        w = 22
        ctrl = ref_loader.make_controller(cfg, depth)
        scale = cfg["dino_input_size"] / w
        p1 = servo.to_cartesian(chosen, w)
        p2 = servo.to_cartesian(r["nn1"][chosen], w)
        s_uv_star, s_uv = ctrl.calculate_uv((p1 * scale + scale / 2).tolist(), (p2 * scale + scale / 2).tolist())
        Z = ctrl.get_depth(s_uv)
        s_xy, s_star = ctrl.transform_to_real_world(s_uv, s_uv_star)
        L = ctrl.calculate_interaction_matrix(s_xy, Z)
        e = (s_xy - s_star).reshape((len(s_xy) * 2, 1))
        tw = (-ctrl.lambda_ * np.linalg.pinv(L.astype("float")) @ e).flatten()
        top2 = torch.from_numpy(r["S"]).topk(2, dim=-1).values
        top2c = torch.from_numpy(r["S"]).topk(2, dim=-2).values
        pre = f"pair{i}/"
        out[pre + "desc1_rows"] = d1[0, 0, ::37, :].numpy().astype(np.float32)     # 14 sampled tokens, full 9D
        out[pre + "desc2_rows"] = d2[0, 0, ::37, :].numpy().astype(np.float32)
        out[pre + "nn1"] = r["nn1"]
        out[pre + "nn2"] = r["nn2"]
        out[pre + "sim1"] = r["sim1"]
        out[pre + "row_margin"] = (top2[:, 0] - top2[:, 1]).numpy()
        out[pre + "col_margin"] = (top2c[0] - top2c[1]).numpy()
        out[pre + "mean_sim1"] = r["mean_sim1"]
        out[pre + "sel_tok1"] = r["sel_tok1"]
        out[pre + "chosen"] = chosen
        out[pre + "s_uv"] = s_uv
        out[pre + "s_uv_star"] = s_uv_star
        out[pre + "twist"] = tw
    out["model_type"] = np.array(model_type)
    np.savez_compressed(os.path.join(OUT, f"e2e_{model_type}.npz"), **out)
    return out


def main():
    assert ref_loader.available(), "run in the build container: /root/reference is required"
    os.makedirs(OUT, exist_ok=True)
    torch.manual_seed(121)      # vitvs_v2.py:1397 (randperm inside find_correspondences_batch)
    np.random.seed(41)          # vitvs_v2.py:1398
    torch.set_num_threads(os.cpu_count() or 1)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        golden_matching()
        golden_logbin()
        golden_ibvs()
        golden_e2e("dinov2_vits14", 2)
    for f in sorted(os.listdir(OUT)):
        print(f, os.path.getsize(os.path.join(OUT, f)))


if __name__ == "__main__":
    main()
