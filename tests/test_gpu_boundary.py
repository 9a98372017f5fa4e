"""The drop-in seam: the mix-in's control-law methods carry the reference's own signatures, so the body of
the reference's Controller.ibvs() (vitvs_v2.py:588-628) can be run line for line on top of them; selection can follow
the reference's random draw through a caller-supplied priority list; handles on two devices coexist in one process."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from oracle import dinov2, servo  # noqa: E402
from vitvs_b200 import synth  # noqa: E402

from _parity import check_keypoints  # noqa: E402


@pytest.fixture(scope="module")
def vits():
    return dinov2.build("dinov2_vits14", seed=0)


@pytest.fixture(scope="module")
def eng(vits):
    from vitvs_b200 import Engine, ServoConfig
    e = Engine(ServoConfig(model_type="dinov2_vits14"), max_images=8)
    e.load_weights(vits)
    yield e
    e.close()


class _StubReferenceController:
    """What Controller.__init__ / load_parameters give the reference's ibvs() besides the mix-in's methods."""
    lambda_ = 0.03

    def reference_ibvs_body(self):
        """vitvs_v2.py:597-623 in the reference's own order and with its own argument lists."""
        result = self.detect_features()
        assert result is not None and result[0] is not None
        (s_uv_star, s_uv), sim_selected_12 = result
        assert len(s_uv_star) >= 4
        s_xy, s_star_xy = self.transform_to_real_world(s_uv, s_uv_star)            # :610
        e = s_xy - s_star_xy                                                       # :613
        e = e.reshape((len(s_xy) * 2, 1))                                          # :614
        Z = self.get_depth(s_uv)                                                   # :616
        L = self.calculate_interaction_matrix(s_xy, Z)                             # :621
        v_c = -self.lambda_ * np.linalg.pinv(L.astype('float')) @ e                # :622
        return v_c.flatten(), (s_uv_star, s_uv, s_xy, s_star_xy, Z, L)


def test_reference_ibvs_body_on_the_mixin(eng):
    from PIL import Image
    from vitvs_b200 import ServoConfig, plugin

    class C(plugin.DinoServoMixin, _StubReferenceController):
        pass
    s = C()
    s.cfg, s.engine = ServoConfig(), eng
    s.initialize_ema()
    goal = synth.goal_frame(1)
    s.goal_image = Image.fromarray(goal)
    s.latest_pil_image = Image.fromarray(synth.current_frame(1, goal))
    s.latest_image_depth = None
    # transform_to_real_world needs no depth frame (vitvs_v2.py:634-648)
    xy, xys = s.transform_to_real_world(np.array([[320, 240], [0, 0]]), np.array([[100, 50], [639, 479]]))
    assert np.array_equal(xy, servo.to_normalised(np.array([[320, 240], [0, 0]])))
    assert np.array_equal(xys, servo.to_normalised(np.array([[100, 50], [639, 479]])))
    assert s.get_depth(np.array([[1, 1]])) is None                                # :568-570
    s.latest_image_depth = synth.depth_frame(1)
    v_ref_order, (s_uv_star, s_uv, s_xy, s_star, Z, L) = s.reference_ibvs_body()
    fused = s._last_fused["twist"][0].cpu().numpy()
    assert np.abs(v_ref_order - fused).max() <= 1e-9 * np.abs(fused).max()
    # each term against the oracle (pinned to the reference's methods): bit-exact
    assert np.array_equal(s_xy, servo.to_normalised(s_uv)) and np.array_equal(s_star, servo.to_normalised(s_uv_star))
    assert np.array_equal(Z, servo.get_depth(s_uv, s.latest_image_depth))
    assert np.array_equal(L, servo.interaction_matrix(s_xy, Z))
    assert L.shape == (48, 6) and Z.shape == (24, 1)
    # ibvs() keeps the velocity window the stopping rule reads (:626-628)
    s.max_velocity_vector_history = 3
    for _ in range(5):
        s.ibvs()
    assert len(s.velocity_vector_history) == 3 and np.array_equal(s.velocity_vector_history[-1], s.v_c)
    # the depth frame is uploaded once per frame object
    d0 = s._depth_on_device()
    assert s._depth_on_device() is d0
    s.latest_image_depth = s.latest_image_depth.copy()
    assert s._depth_on_device() is not d0


def test_selection_order_input(eng):
    """vvs_match_out.sel_order: the first num_pairs candidates of the caller's priority list, in list order
    (vitvs_v2.py:139-141); in the same-image branch exactly perm[:num_points] (:90-91)."""
    pairs = [synth.frame_pair(i) for i in range(3)]
    goal = torch.from_numpy(np.stack([p[0] for p in pairs])).cuda()
    cur = torch.from_numpy(np.stack([p[1] for p in pairs])).cuda()
    depth = torch.from_numpy(np.stack([p[2] for p in pairs])).cuda()
    cur[2] = goal[2]                                                    # third pair: the same image twice
    T = eng.tokens
    gen = torch.Generator().manual_seed(121)
    order = torch.stack([torch.randperm(T, generator=gen) for _ in range(3)]).to(torch.int32)
    base = eng.frame_pairs_to_twist(goal, cur, depth, want_match=True)
    out = eng.frame_pairs_to_twist(goal, cur, depth, want_match=True, sel_order=order)
    torch.cuda.synchronize()
    for k in ("nn1", "nn2", "sim1", "candidate", "n_candidates", "flags"):
        assert torch.equal(base[k], out[k]), k                           # selection does not touch the matching
    for p in range(3):
        cand = out["candidate"][p].cpu().numpy().astype(bool)
        same = bool(int(out["flags"][p]) & 1)
        assert same == (p == 2)
        o = order[p].numpy()
        expect = o[:24] if same else o[cand[o]][:24]
        n = int(out["n_selected"][p])
        assert n == len(expect) == 24
        assert np.array_equal(out["sel_tok"][p].cpu().numpy(), expect)
        sim1 = out["sim1"][p].cpu().numpy()
        assert np.array_equal(out["sel_sim"][p].cpu().numpy(), np.ones(24, np.float32) if same else sim1[expect])
        check_keypoints(out, p, 22)
        uv, uvs = out["uv"][p].cpu().numpy().astype(int), out["uv_star"][p].cpu().numpy().astype(int)
        tw_ref = servo.ibvs_twist(uv, uvs, pairs[p][2])
        tw = out["twist"][p].cpu().numpy()
        assert np.abs(tw - tw_ref).max() <= 1e-9 * max(np.abs(tw_ref).max(), 1e-300) + 1e-18
    assert not torch.equal(base["sel_tok"][0], out["sel_tok"][0])
    # a 1-D list is shared by every pair; Engine.match takes it too
    a, b, _ = synth.descriptor_sets(484, 384, seed=3, noise=8.0)
    perm = torch.randperm(484, generator=gen)
    r = eng.match(a[None], b[None], num_pairs=24, sel_order=perm)
    cand = r["candidate"][0].cpu().numpy().astype(bool)
    assert np.array_equal(r["sel_tok"][0].cpu().numpy(), perm.numpy()[cand[perm.numpy()]][:24])


def test_nan_descriptors_do_not_fault(eng):
    """a descriptor row of NaNs leaves its arg-max key unset; indices decode to 0, flags bit 1 is raised."""
    a, b, _ = synth.descriptor_sets(196, 128, seed=5)
    a[7] = float("nan")
    r = eng.match(a[None], b[None], num_pairs=24)
    torch.cuda.synchronize()
    nn1, nn2 = r["nn1"][0].cpu().numpy(), r["nn2"][0].cpu().numpy()
    assert nn1.min() >= 0 and nn1.max() < 196 and nn2.min() >= 0 and nn2.max() < 196
    assert int(r["flags"][0]) & 2
    assert nn1[7] == 0 and np.isnan(r["sim1"][0, 7].item())
    c, d, _ = synth.descriptor_sets(196, 128, seed=5)
    assert not int(eng.match(c[None], d[None], num_pairs=24)["flags"][0]) & 2


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs in one process")
def test_handles_on_two_devices_in_one_process(vits):
    """include/vitvs_b200.h: any number of handles per process and device.  Function attributes (> 48 KB dynamic shared
    memory) are per device; every entry switches to its handle's device whatever the caller's current device is."""
    from vitvs_b200 import Engine, ServoConfig
    g, c, d = synth.frame_pair(2)
    outs = []
    engines = [Engine(ServoConfig(model_type="dinov2_vits14"), max_images=2, device=i) for i in (0, 1)]
    for e in engines:
        e.load_weights(vits)
    torch.cuda.set_device(0)                                            # the caller's device stays 0 throughout
    for i, e in enumerate(engines):
        dev = torch.device("cuda", i)
        o = e.frame_pairs_to_twist(torch.from_numpy(g[None]).to(dev), torch.from_numpy(c[None]).to(dev),
                                   torch.from_numpy(d[None]).to(dev), want_match=True)
        torch.cuda.synchronize(dev)
        outs.append({k: v.cpu() for k, v in o.items() if not k.startswith("_")})
    assert torch.cuda.current_device() == 0
    for k in outs[0]:
        assert torch.equal(outs[0][k], outs[1][k]), k
    for e in engines:
        e.close()


def test_fold_sources_released_after_weights_ready(vits):
    """vvs_weights_ready hands the fp32 fold sources back to the device (include/vitvs_b200.h): a second load of the same
    state_dict gives the same twists bit for bit; changing only norm1.weight afterwards makes the block incomplete (loud,
    not stale) until its attn.qkv.weight is set again, and the result then equals that of a fresh handle."""
    import ctypes as C
    from vitvs_b200 import Engine, ServoConfig, VvsError, _lib
    g, c, d = synth.frame_batch(2)
    args = (torch.from_numpy(g).cuda(), torch.from_numpy(c).cuda(), torch.from_numpy(d).cuda())
    e = Engine(ServoConfig(model_type="dinov2_vits14"), max_images=4)
    e.load_weights(vits)
    base = e.frame_pairs_to_twist(*args)["twist"].clone()
    e.load_weights(vits)                                               # whole state_dict again, sources re-uploaded
    assert torch.equal(e.frame_pairs_to_twist(*args)["twist"], base)

    state = {k: v.clone() for k, v in vits.state_dict().items()}
    state["blocks.3.norm1.weight"] = state["blocks.3.norm1.weight"] * 1.25

    def put(name):
        t = state[name].detach().to("cuda", torch.float32).contiguous()
        shape = (C.c_int64 * t.dim())(*t.shape)
        rc = e.lib.vvs_set_weight(e._h, name.encode(), t.data_ptr(), _lib.F32, shape, t.dim(), None)
        torch.cuda.synchronize()
        assert rc == 0, e.lib.vvs_last_error(e._h)

    put("blocks.3.norm1.weight")
    assert e.lib.vvs_weights_ready(e._h) != 0 and b"blocks.3 incomplete" in e.lib.vvs_last_error(e._h)
    with pytest.raises(VvsError):
        e.frame_pairs_to_twist(*args)
    put("blocks.3.attn.qkv.weight")
    assert e.lib.vvs_weights_ready(e._h) == 0
    changed = e.frame_pairs_to_twist(*args)["twist"].clone()
    fresh = Engine(ServoConfig(model_type="dinov2_vits14"), max_images=4)
    fresh.load_weights(state)
    assert torch.equal(changed, fresh.frame_pairs_to_twist(*args)["twist"])
    assert not torch.equal(changed, base)
    fresh.close()
    e.close()
