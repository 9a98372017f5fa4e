"""SURVEY §8f N3 + N2: the headless planar-scene camera is bit-identical to its CPU restatement (oracle/render.py),
and closed-loop servo episodes driven by the hot path's own twists converge on the goal view - the end-to-end check
that every sign and frame convention of a9-a16 (keypoint order, interaction matrix, v = -lambda L+ e) is the
reference's: a wrong one makes the loop diverge."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from oracle import dinov2, render  # noqa: E402  (the checker)
from vitvs_b200 import episode  # noqa: E402


def _texture(seed=0, th=600, tw=800):
    """band-limited noise poster (same recipe as synth.goal_frame, texture-sized)"""
    g = torch.Generator().manual_seed(seed)
    low = torch.rand(1, 3, 75, 100, generator=g)
    img = torch.nn.functional.interpolate(low, size=(th, tw), mode="bicubic", align_corners=False)
    return (img[0].clamp(0, 1) * 255).round().byte().permute(1, 2, 0).contiguous().numpy()


@pytest.fixture(scope="module")
def vits():
    return dinov2.build("dinov2_vits14", seed=0)


@pytest.fixture(scope="module")
def eng(vits):
    from vitvs_b200 import Engine, ServoConfig
    e = Engine(ServoConfig(model_type="dinov2_vits14", lambda_=0.5), max_images=16)
    e.load_weights(vits)
    yield e
    e.close()


def test_render_bit_exact_vs_oracle(eng):
    tex = _texture()
    poses = np.stack([
        episode.make_pose((0.0, 0.0, -0.61)),                                   # the goal view
        episode.make_pose((0.05, -0.03, -0.70), (0.05, -0.08, 0.3)),            # tilted, rotated, partly off the poster
        episode.make_pose((0.3, 0.25, -0.25), (0.0, 0.0, -1.0)),                # close: bilinear magnification
        episode.make_pose((0.0, 0.0, -0.15)),                                    # nearer than 0.2 m: depth 0
        episode.make_pose((0.0, 0.0, -0.61), (np.pi / 2 + 0.2, 0.0, 0.0)),       # looking past the plane: rays that miss
    ])
    rgb, depth = eng.render_plane(torch.from_numpy(tex).cuda(), torch.from_numpy(poses))
    rgb, depth = rgb.cpu().numpy(), depth.cpu().numpy()
    for i, p in enumerate(poses):
        r, d = render.render_plane(tex, p)
        assert np.array_equal(depth[i], d), f"depth, pose {i}"
        assert np.array_equal(rgb[i], r), f"rgb, pose {i}: {np.count_nonzero(rgb[i] != r)} bytes differ"
    assert depth[0].min() == 610 and depth[0].max() == 610 and depth[3].max() == 0
    assert (rgb[4] == 128).any()


def test_kinematics_match_oracle():
    rng = np.random.default_rng(0)
    pose = episode.make_pose((0.02, -0.04, -0.6), (0.1, -0.2, 0.3))
    for _ in range(20):
        tw = rng.normal(size=6) * 0.05
        a, b = episode.integrate_twist(pose, tw, 0.7), render.integrate_twist(pose, tw, 0.7)
        assert np.allclose(a, b, rtol=0, atol=1e-15)
        pose = a
    R = pose[:9].reshape(3, 3)
    assert np.allclose(R @ R.T, np.eye(3), atol=1e-12)
    et, er = episode.end_error(pose, episode.make_pose((0, 0, -0.61)))
    ot, orr = render.pose_error(pose, episode.make_pose((0, 0, -0.61)))
    assert abs(et - 100 * ot) < 1e-9 and abs(er - orr) < 1e-9


def test_closed_loop_episodes_converge(eng, tmp_path):
    tex = _texture(1)
    goal = episode.make_pose((0.0, 0.0, -0.61))
    starts = [episode.make_pose((0.04, -0.03, -0.66), (0.0, 0.0, 0.12)),
              episode.make_pose((-0.05, 0.02, -0.58), (0.03, -0.02, -0.15)),
              episode.make_pose((0.02, 0.05, -0.70), (-0.04, 0.03, 0.08)),
              episode.make_pose((0.0, 0.0, -0.61))]                                 # starts at the goal: same-image branch
    batch = episode.EpisodeBatch(eng, tex, goal, starts, dt=0.3, min_iterations=60, max_iterations=100)
    res = batch.run()
    for r in res[:3]:
        print(f"[episode] {r['initial_position_error_cm']:.2f} cm / {r['initial_orientation_error_deg']:.2f} deg -> "
              f"{r['position_error_cm']:.3f} cm / {r['orientation_error_deg']:.3f} deg in {r['iterations']} iterations "
              f"(converged {r['converged']})")
        # random-init weights and 14-px patch-centre keypoints bound the final accuracy; a sign or frame error in
        # a9-a16 would make the error GROW (the 2x divergence abort of vitvs_v2.py:359 would fire)
        assert r["position_error_cm"] < 0.5 * r["initial_position_error_cm"]
        assert r["orientation_error_deg"] < 0.5 * r["initial_orientation_error_deg"]
    assert res[3]["position_error_cm"] < 0.5                                       # stays at the goal
    path = str(tmp_path / "results.npz")
    batch.save_npz(path)
    z = np.load(path, allow_pickle=True)
    for key in ("initial_positions", "final_positions", "final_quaternions", "convergence_flags", "position_errors",
                "orientation_errors", "all_position_histories", "all_iteration_histories", "lowest_position_errors",
                "all_applied_velocity_x", "all_applied_velocity_yaw", "total_execution_time"):
        assert key in z.files
    assert z["final_positions"].shape == (4, 3) and z["convergence_flags"].shape == (4,)


def test_best_pose_search_and_num_pairs(eng):
    """find_and_set_best_pose (vitvs_v2.py:1151-1189), batched: starts turned by a multiple of 90 degrees about the
    optical axis are turned back by the search; num_pairs is 48 during the search and restored after it."""
    tex = torch.from_numpy(_texture(2)).cuda()
    goal_pose = episode.make_pose((0.0, 0.0, -0.61))
    goal_frame, _ = eng.render_plane(tex, torch.from_numpy(goal_pose[None]))
    base = [episode.make_pose((0.02, -0.01, -0.63), (0.0, 0.0, 0.05)), episode.make_pose((-0.03, 0.02, -0.60), (0.01, 0.0, -0.04))]
    turned = np.stack([episode.rotate_about_optical_axis(base[0], 180), episode.rotate_about_optical_axis(base[1], 90),
                       base[0]])
    assert eng.cfg.num_pairs == 24
    best = episode.find_best_poses(eng, tex, goal_frame, turned)
    assert eng.cfg.num_pairs == 24
    for b, want in zip(best, (base[0], base[1], base[0])):
        _, err_deg = episode.end_error(b, want)
        assert err_deg < 1e-6, err_deg
    # K really changes: 48 keypoints come back while it is raised
    eng.set_num_pairs(48)
    frames, depth = eng.render_plane(tex, torch.from_numpy(np.stack(base)))
    out = eng.frames_to_twist(goal_frame.expand(2, -1, -1, -1).contiguous(), frames, depth, want_match=True)
    assert out["uv"].shape == (2, 48, 2) and int(out["n_selected"].max()) > 24
    eng.set_num_pairs(24)
    from vitvs_b200 import VvsError
    with pytest.raises(VvsError):
        eng.set_num_pairs(1000)


def test_cached_goal_matches_pair_call(eng):
    """vvs_set_goal + vvs_current_to_twist (an extra: the goal view's descriptors computed once) return, bit for bit, what
    the pair entries return for the same (goal, current) views - images and raw camera frames, per-token outputs,
    fewer current views than cached goals."""
    from vitvs_b200 import synth
    fr = [synth.frame_pair(i) for i in range(4)]
    g, c, d = (torch.from_numpy(np.stack([f[k] for f in fr])).cuda() for k in range(3))
    ref = eng.frame_pairs_to_twist(g, c, d, want_match=True)
    eng.set_goal(g)
    out = eng.current_to_twist(c, d, want_match=True)
    torch.cuda.synchronize()
    for k in ("twist", "n_selected", "mean_sim_sel", "nn1", "nn2", "sim1", "sel_tok", "uv", "uv_star"):
        assert torch.equal(out[k], ref[k]), k
    for rep in range(4):                            # eager, graph capture, graph replays
        o = eng.current_to_twist(c, d)
        assert torch.equal(o["twist"], ref["twist"]) and torch.equal(o["n_selected"], ref["n_selected"]), rep
        assert torch.equal(o["mean_sim_sel"], ref["mean_sim_sel"]), rep
    out2 = eng.current_to_twist(c[:2], d[:2])       # a prefix of the cached goals
    assert torch.equal(out2["twist"], ref["twist"][:2])
    graw, craw, depth = synth.raw_frame_batch(3, distinct=3)
    graw, craw, depth = (torch.from_numpy(x).cuda() for x in (graw, craw, depth))
    ref = eng.frames_to_twist(graw, craw, depth)
    eng.set_goal(graw)
    for rep in range(3):
        out = eng.current_to_twist(craw, depth)
        assert torch.equal(out["twist"], ref["twist"]) and torch.equal(out["n_selected"], ref["n_selected"]), rep
    with pytest.raises(Exception):
        eng.current_to_twist(torch.cat([craw, craw]), torch.cat([depth, depth]))   # more views than cached goals


def test_episode_batch_with_cached_goal(eng):
    """EpisodeBatch(cache_goal=True) walks the same trajectory as the default driver."""
    tex = _texture()
    goal = episode.make_pose((0.0, 0.0, -0.61))
    starts = [episode.make_pose((0.03, -0.02, -0.66), (0.0, 0.0, 0.06)), episode.make_pose((-0.04, 0.03, -0.58), (0.02, -0.03, -0.05))]
    runs = []
    for cache in (False, True):
        eb = episode.EpisodeBatch(eng, tex, goal, starts, min_iterations=5, max_iterations=6, cache_goal=cache)
        for _ in range(4):
            eb.step()
        runs.append(np.stack([e.pose for e in eb.episodes]))
    assert np.array_equal(runs[0], runs[1])
