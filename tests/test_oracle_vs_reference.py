"""Live cross-check of the oracle against the reference's own functions, executed from
/root/reference (build container only; skipped on the GPU box where it does not exist)."""
import copy
import sys
import warnings

import numpy as np
import pytest
import torch

from oracle import dinov2, ref_loader, servo

pytestmark = pytest.mark.skipif(not ref_loader.available(), reason="/root/reference not present")


@pytest.fixture(scope="module")
def vits():
    return dinov2.build("dinov2_vits14", seed=0)


def test_extractor_descriptors_bit_equal(vits):
    ex = ref_loader.make_extractor(vits, "dinov2_vits14")
    g = torch.Generator().manual_seed(0)
    batch = torch.randn(1, 3, 98, 98, generator=g)                  # 7x7 patches: fast
    with torch.no_grad(), warnings.catch_warnings():
        warnings.simplefilter("ignore")
        for bin_ in (False, True):
            ref = ex.extract_descriptors(batch, layer=11, facet="token", bin=bin_)
            mine = servo.extract_descriptors(vits, batch, 11, bin_)
            assert torch.equal(ref, mine)
        ref = ex.extract_descriptors(batch, layer=5, facet="token", bin=False)
        assert torch.equal(ref, servo.extract_descriptors(vits, batch, 5, False))


def test_attention_matches_reference_module(vits):
    """oracle Attention == dino_patch/attention.py MemEffAttention (xFormers absent => SDPA)."""
    A = ref_loader.attention()
    blk = vits.blocks[3].attn
    ref = A.MemEffAttention(384, num_heads=6, qkv_bias=True, proj_bias=True).eval()
    ref.qkv.load_state_dict(blk.qkv.state_dict())
    ref.proj.load_state_dict(blk.proj.state_dict())
    x = torch.randn(2, 50, 384, generator=torch.Generator().manual_seed(1))
    with torch.no_grad():
        assert torch.equal(ref(x), blk(x))


@pytest.mark.parametrize("cls_name", ["Block", "NestedTensorBlock"])
def test_block_matches_reference_module(vits, cls_name):
    """oracle Block == dino_patch/block.py Block / NestedTensorBlock (inference branch, :90-95,112-114) with the reference's
    own MemEffAttention inside: LayerNorm placement, residual order and LayerScale position are the reference file's.
    LayerScale / Mlp are stand-ins for the un-vendored upstream siblings (oracle/ref_loader.block)."""
    from functools import partial
    B = ref_loader.block()
    A = sys.modules["_ref_dino_patch.attention"]
    for idx in (0, 7):
        blk = vits.blocks[idx]
        ref = getattr(B, cls_name)(384, num_heads=6, mlp_ratio=4.0, qkv_bias=True, proj_bias=True, ffn_bias=True,
                                   init_values=1.0, norm_layer=partial(torch.nn.LayerNorm, eps=1e-6),
                                   attn_class=A.MemEffAttention).eval()
        sd = {k: v.clone() for k, v in blk.state_dict().items()}
        sd["ls1.gamma"] = torch.linspace(0.5, 1.5, 384)          # LayerScale must sit on the branch, not on x
        sd["ls2.gamma"] = torch.linspace(1.5, 0.5, 384)
        ref.load_state_dict(sd, strict=True)
        mine = copy.deepcopy(blk)
        mine.load_state_dict(sd, strict=True)
        x = torch.randn(2, 50, 384, generator=torch.Generator().manual_seed(idx))
        with torch.no_grad():
            assert torch.equal(ref(x), mine(x))


def test_preprocess_matches_reference(vits):
    from PIL import Image
    from vitvs_b200 import synth
    ex = ref_loader.make_extractor(vits, "dinov2_vits14")
    g, _, _ = synth.frame_pair(1, 308)
    assert torch.equal(ex.preprocess_pil(Image.fromarray(g)), servo.preprocess_u8(g))


def test_pos_embed_interpolation_matches_reference_patch(vits):
    """dinov2_extractor.py:94-120 (_fix_pos_enc) evaluated at stride 14 == oracle's method."""
    ex_mod = ref_loader.extractor()
    fn = ex_mod.ViTExtractor._fix_pos_enc(14, (14, 14))
    x = torch.zeros(1, 1 + 22 * 22, 384)
    with torch.no_grad():
        ref = fn(vits, x, 308, 308)
        mine = vits.interpolate_pos_encoding(x, 308, 308)
    assert torch.allclose(ref, mine, atol=1e-6)


def test_full_candidate_set_is_mutual_nn_set():
    v = ref_loader.vitvs()
    g = torch.Generator().manual_seed(2)
    for T, d in ((64, 32), (121, 16)):
        a = torch.randn(1, 1, T, d, generator=g)
        b = a[:, :, torch.randperm(T, generator=g)] + 2.5 * torch.randn(1, 1, T, d, generator=g)
        p1, p2, sim = v.find_correspondences_batch(a, b, num_pairs=T)
        w = int(np.sqrt(T))
        tok1 = np.sort((p1[:, 0] * w + p1[:, 1]).numpy())
        m = servo.match(servo.cosine_similarity_matrix_literal(a, b)[0, 0])
        assert np.array_equal(tok1, m.candidates)
        assert np.array_equal(m.candidates, np.nonzero(m.mutual)[0])


@pytest.mark.parametrize("S,grid", [(308, 22), (112, 15), (308, 43), (518, 37), (308, 128), (308, 64), (322, 45)])
def test_keypoint_recovery_equals_reference_lines(S, grid):
    """a11 on grids whose scale S / grid is not a power-of-two fraction: the reference's own three lines
    (vitvs_v2.py:511-513, evaluated literally: int64 tensor * Python float in torch => fp32) followed by the
    reference's own Controller.calculate_uv (:525-553, Python round = half-to-even), for EVERY patch of the grid."""
    cfg = dict(servo.DEFAULT_CFG, dino_input_size=S, num_pairs=grid * grid)
    ctrl = ref_loader.make_controller(cfg)
    ctrl.dino_input_size, ctrl.num_pairs = S, grid * grid
    idx = torch.arange(grid * grid)
    points1 = torch.stack([idx // grid, idx % grid], dim=-1)                     # _to_cartesian (:58-70)
    points2 = points1.flip(0).contiguous()
    scale = S / int(np.sqrt(grid * grid))                                        # :511
    points1_scaled = points1 * scale + scale / 2                                 # :512
    points2_scaled = points2 * scale + scale / 2                                 # :513
    assert points1_scaled.dtype == torch.float32
    ref_star, ref_uv = ctrl.calculate_uv(points1_scaled.tolist(), points2_scaled.tolist())
    mine_star, mine_uv = servo.calculate_uv(points1.numpy(), points2.numpy(), grid * grid, cfg)
    assert np.array_equal(ref_star, mine_star) and np.array_equal(ref_uv, mine_uv)


def test_cyclic_distance_bit_equal_to_pairwise_distance():
    """Exhaustive over every (drow, dcol) offset of a 46x46 grid (covers 22x22 and 34x45)."""
    r = torch.arange(-45, 46)
    dr, dc = torch.meshgrid(r, r, indexing="ij")
    a = torch.stack([dr.reshape(-1), dc.reshape(-1)], -1)
    ref = torch.nn.PairwiseDistance(p=2)(a, torch.zeros_like(a)).numpy()
    mine = servo.cyclic_distance_f32(dr.reshape(-1).numpy(), dc.reshape(-1).numpy())
    assert np.array_equal(ref, mine)


def test_experiment_sampling_equals_reference():
    """vitvs_b200.episode restates the start-pose sampling of the reference's main() (vitvs_v2.py:1385-1418): same
    numbers from the same seed, then mapped into the renderer's camera convention."""
    from vitvs_b200 import episode
    V = ref_loader.vitvs()
    n, circles, aug = 40, 4, 0.08
    np.random.seed(41)
    pos_r = V.sample_camera_positions(np.array([1.2, 1.2, 0.3]), n, np.array([0, 0, 0.61]))
    foc_r = V.sample_focal_points_original(n, np.array([0.0, 0.0, 0.01]), circles, aug)
    mats_r, _ = V.calculate_look_at_orientation(pos_r, foc_r)
    quat_r = V.apply_z_axis_rotation(mats_r, circles, n // circles)
    np.random.seed(41)
    pos = episode.sample_camera_positions(np.array([1.2, 1.2, 0.3]), n, np.array([0, 0, 0.61]))
    foc = episode.sample_focal_points(n, np.array([0.0, 0.0, 0.01]), circles, aug)
    mats = episode.look_at_matrices(pos, foc)
    rot = episode.apply_optical_axis_rotation(mats, circles, n // circles)
    assert np.array_equal(pos, pos_r) and np.array_equal(foc, foc_r) and np.array_equal(mats, mats_r)
    from scipy.spatial.transform import Rotation as R
    assert np.allclose(R.from_matrix(rot).as_quat(), quat_r, atol=1e-12)
    poses = episode.sample_start_poses(n, circles, aug, seed=41)
    assert poses.shape == (n, 12)
    for i in range(n):
        assert np.allclose(poses[i], episode.to_camera_pose(pos_r[i], R.from_quat(quat_r[i]).as_matrix()), atol=1e-12)
    # the reference's desired pose (position (0,0,0.61), quaternion [0, .7071068, 0, .7071068]: vitvs_v2.py:1391-1392)
    # is the renderer's goal view: identity rotation, 0.61 m in front of the poster
    goal = episode.to_camera_pose([0, 0, 0.61], R.from_quat([0, 0.7071068, 0, 0.7071068]).as_matrix())
    assert np.allclose(goal, episode.make_pose((0, 0, -0.61)), atol=1e-7)
    # every sampled camera looks at the poster: its optical axis hits z = 0 within the circles' radius
    for p in poses:
        Rm, t = p[:9].reshape(3, 3), p[9:]
        s = -t[2] / Rm[2, 2]
        hit = t + s * Rm[:, 2]
        assert s > 0 and np.hypot(hit[0], hit[1]) < circles * aug + 0.02
    # rotate_camera_x_axis (:1125-1148) == a rotation about our optical axis
    q0 = quat_r[3]
    for ang in (90, 180, 270):
        ref_pose = episode.to_camera_pose(pos_r[3], R.from_quat(V.rotate_camera_x_axis(q0, ang)).as_matrix())
        assert np.allclose(episode.rotate_about_optical_axis(poses[3], ang), ref_pose, atol=1e-12)


def test_facet_descriptors_bit_equal(vits):
    """key / query / value facets (dinov2_extractor.py:203-218), with and without binning / CLS."""
    ex = ref_loader.make_extractor(vits, "dinov2_vits14")
    batch = torch.randn(2, 3, 70, 98, generator=torch.Generator().manual_seed(2))     # 5 x 7 patches
    with torch.no_grad(), warnings.catch_warnings():
        warnings.simplefilter("ignore")
        for facet in ("key", "query", "value"):
            for bin_, cls_ in ((False, False), (True, False), (False, True)):
                ref = ex.extract_descriptors(batch, layer=9, facet=facet, bin=bin_, include_cls=cls_)
                mine = servo.extract_facet_descriptors(vits, batch, 9, facet, bin_, cls_)
                assert torch.equal(ref, mine), (facet, bin_, cls_)


def test_stride_patching_equals_reference(vits):
    """stride < patch size (dinov2_extractor.py:85-144).  The reference's ViTExtractor patches a copy of
    the oracle model itself (patch_vit_resolution + _fix_pos_enc); oracle.dinov2.with_stride must give the same tokens."""
    import copy
    batch = torch.randn(1, 3, 56, 56, generator=torch.Generator().manual_seed(4))
    for stride in (7, 2):
        ex = ref_loader.extractor().ViTExtractor(model_type="dinov2_vits14", stride=stride, model=copy.deepcopy(vits),
                                                 device="cpu")
        mine = dinov2.with_stride(vits, stride)
        with torch.no_grad(), warnings.catch_warnings():
            warnings.simplefilter("ignore")
            ref = ex.extract_descriptors(batch, layer=3, facet="token", bin=False)
            grid = 1 + (56 - 14) // stride
            assert ex.num_patches == (grid, grid) and ref.shape == (1, 1, grid * grid, 384)
            x = mine.prepare_tokens(batch)
            for blk in mine.blocks[:4]:
                x = blk(x)
            assert torch.equal(ref[:, 0], x[:, 1:])
    assert vits.patch_embed.proj.stride == (14, 14)          # the shared fixture was not patched


def test_log_bin_hierarchy_equals_reference(vits):
    """_log_bin with hierarchy > 1 (dinov2_extractor.py:265-311; only reachable by calling it directly)."""
    ex = ref_loader.make_extractor(vits, "dinov2_vits14")
    g = torch.Generator().manual_seed(7)
    for hp, wp in ((7, 7), (5, 9), (2, 3)):
        x = torch.randn(2, 1, hp * wp, 24, generator=g)            # B x h x t x d as the reference's hook delivers it
        ex.num_patches = (hp, wp)
        for hierarchy in (1, 2, 3):
            ref = ex._log_bin(x, hierarchy)[:, 0]
            mine = servo.log_bin_hierarchy(x[:, 0], (hp, wp), hierarchy)
            assert ref.shape == mine.shape == (2, hp * wp, (1 + 8 * hierarchy) * 24)
            assert torch.equal(ref, mine), (hp, wp, hierarchy)
        assert torch.equal(servo.log_bin_hierarchy(x[:, 0], (hp, wp), 1), servo.log_bin(x[:, 0], (hp, wp)))
