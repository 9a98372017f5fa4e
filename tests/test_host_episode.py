"""CPU tests of the episode driver's host logic and of the oracle renderer's geometry: kinematics,
EMA, clipping, the stopping rule of vitvs_v2.py:345-421 and the results file, with a stand-in engine whose twist
is an ideal pose-error controller (no GPU, no kernels)."""
import numpy as np
import torch

from oracle import render
from vitvs_b200 import ServoConfig, episode


class _IdealEngine:
    """render_plane / frames_to_twist with the shapes of Engine; the twist drives the pose error to zero."""
    device = torch.device("cpu")
    max_images = 64

    tokens = 484

    def __init__(self, goal_pose, gain=0.2, fail_from=None, noise=0.0, fail_until=None, fail_nsel=0):
        self.cfg = ServoConfig()
        self.fail_until, self.fail_nsel = fail_until, fail_nsel
        self.goal = goal_pose
        self.gain, self.fail_from, self.calls = gain, fail_from, 0
        self._poses = None
        self.noise, self.rng = noise, np.random.default_rng(3)

    def render_plane(self, texture, poses, plane_w=0.8, plane_h=0.6):
        self._poses = poses.numpy()
        n = poses.shape[0]
        return torch.zeros(n, 4, 4, 3, dtype=torch.uint8), torch.zeros(n, 4, 4, dtype=torch.uint16)

    def frames_to_twist(self, goal, cur, depth, want_match=False, sel_order=None):
        self.calls += 1
        self.last_order = sel_order
        tw = []
        for p in self._poses:
            R, t = p[:9].reshape(3, 3), p[9:]
            Rg, tg = self.goal[:9].reshape(3, 3), self.goal[9:]
            v = R.T @ (tg - t) * self.gain
            dR = R.T @ Rg
            w = np.array([dR[2, 1] - dR[1, 2], dR[0, 2] - dR[2, 0], dR[1, 0] - dR[0, 1]]) / 2 * self.gain
            tw.append(np.concatenate([v, w]) + self.noise * self.rng.normal(size=6))
        n = len(tw)
        nsel = 24 if self.fail_from is None or self.calls < self.fail_from else 0
        if self.fail_until is not None and self.fail_from <= self.calls < self.fail_until:
            nsel = self.fail_nsel
        elif self.fail_until is not None:
            nsel = 24
        if 0 < nsel < 4:
            tw = [np.zeros(6) for _ in tw]            # the kernel's answer for 1 <= n < 4 pairs: all-zero key-points, zero twist
        return dict(twist=torch.from_numpy(np.array(tw)), n_selected=torch.full((n,), nsel, dtype=torch.int32),
                    mean_sim_sel=torch.ones(n))


def test_kinematics_and_errors_match_oracle():
    rng = np.random.default_rng(1)
    pose = episode.make_pose((0.02, -0.04, -0.6), (0.1, -0.2, 0.3))
    goal = episode.make_pose((0, 0, -0.61))
    for _ in range(10):
        tw = rng.normal(size=6) * 0.05
        a, b = episode.integrate_twist(pose, tw, 0.5), render.integrate_twist(pose, tw, 0.5)
        assert np.allclose(a, b, rtol=0, atol=1e-15)
        pose = a
    et, er = episode.end_error(pose, goal)
    ot, orr = render.pose_error(pose, goal)
    assert abs(et - 100 * ot) < 1e-9 and abs(er - orr) < 1e-9
    q = episode.quaternion_xyzw(pose)
    assert abs(np.linalg.norm(q) - 1) < 1e-12


def test_ideal_controller_converges_and_stops_by_rule(tmp_path):
    goal = episode.make_pose((0, 0, -0.61))
    starts = [episode.make_pose((0.05, -0.03, -0.7), (0.02, -0.03, 0.2)), episode.make_pose((-0.04, 0.04, -0.55), (0, 0, -0.3))]
    eng = _IdealEngine(goal)
    b = episode.EpisodeBatch(eng, np.zeros((4, 4, 3), np.uint8), goal, starts, dt=1.0, min_iterations=300, max_iterations=400)
    res = b.run()
    for r in res:
        # a noiseless controller's velocity only ever falls, so the velocity-trend stop (second window above the
        # first, :402-405) never fires: the episode ends at max_iterations, converged by the 90 % rule (:408-414)
        assert r["converged"] and r["iterations"] == 400
        assert r["position_error_cm"] < 0.01 and r["orientation_error_deg"] < 0.01
    # with a noise floor the velocity-trend rule stops the episode after the 300-iteration floor
    eng2 = _IdealEngine(goal, noise=1e-4)
    b2 = episode.EpisodeBatch(eng2, np.zeros((4, 4, 3), np.uint8), goal, starts, dt=1.0, min_iterations=300, max_iterations=2000)
    for r in b2.run():
        assert r["converged"] and 300 <= r["iterations"] < 2000
    e = b.episodes[0]
    assert e.iteration_count == len(e.applied) == len(e.average_velocities) == 400
    assert len(e.velocity_vector_history) == 200                    # trimmed at max_velocity_vector_history (:627-628)
    # EMA (vitvs_v2.py:329-343): first sample copied, then alpha * new + (1 - alpha) * old
    a = ServoConfig().ema_alpha
    assert a == 0.8
    # ROS twist re-mapping of publish_twist (vitvs_v2.py:668-676)
    v = e.velocity_vector_history[-1]                               # the last iteration's twist (below the clip limit)
    assert np.allclose(e.applied[-1], [v[2], -v[0], -v[1], v[5], -v[3], -v[4]])
    path = str(tmp_path / "r.npz")
    b.save_npz(path, 1.5)
    z = np.load(path, allow_pickle=True)
    assert z["convergence_flags"].all() and z["all_position_histories"].shape == (2,)
    assert float(z["total_execution_time"]) == 1.5


def test_divergence_and_feature_failure_abort():
    goal = episode.make_pose((0, 0, -0.61))
    start = [episode.make_pose((0.05, 0.0, -0.61))]
    eng = _IdealEngine(goal, gain=-0.05)                                    # wrong sign: error grows
    b = episode.EpisodeBatch(eng, np.zeros((4, 4, 3), np.uint8), goal, start, dt=1.0, min_iterations=5, max_iterations=500)
    r = b.run()[0]
    assert not r["converged"] and r["position_error_cm"] > 2 * r["initial_position_error_cm"]   # :359 abort
    eng = _IdealEngine(goal, fail_from=3)
    b = episode.EpisodeBatch(eng, np.zeros((4, 4, 3), np.uint8), goal, start, dt=1.0, min_iterations=5, max_iterations=500)
    b.run()
    # calls 1-2 succeed, calls 3-11 fail but the loop goes on with the last twist, the 10th failure (call 12) raises
    # inside ibvs() BEFORE that iteration is counted (vitvs_v2.py:745-758)
    assert b.episodes[0].aborted == "Persistent feature detection failure" and b.episodes[0].iteration_count == 11


def test_feature_failure_keeps_the_camera_moving():
    """vitvs_v2.py:740-793: a failed detection makes ibvs() return early, but run() still republishes the
    previous self.v_c, integrates the pose, appends the histories and evaluates the stopping rule; only the failure
    counter notices.  The velocity window of the stopping rule gets no new entry (it is appended inside ibvs())."""
    goal = episode.make_pose((0, 0, -0.61))
    start = [episode.make_pose((0.05, 0.0, -0.7))]
    eng = _IdealEngine(goal, fail_from=3, fail_until=6)                       # calls 3, 4, 5 fail, then recover
    b = episode.EpisodeBatch(eng, np.zeros((4, 4, 3), np.uint8), goal, start, dt=1.0, min_iterations=5, max_iterations=8)
    e = b.episodes[0]
    for _ in range(2):
        b.step()
    v_before, pose_before = e.v_c.copy(), e.pose.copy()
    b.step()                                                                  # call 3: no features
    assert e.failures == 1 and e.iteration_count == 3 and np.array_equal(e.v_c, v_before)
    assert np.allclose(e.pose, episode.integrate_twist(pose_before, v_before, 1.0))     # moved by the PREVIOUS twist
    assert len(e.velocity_vector_history) == 2 and len(e.applied) == 3 and len(e.position_history) == 3
    b.step(); b.step()
    assert e.failures == 3
    b.step()                                                                  # call 6: features again
    assert e.failures == 0 and len(e.velocity_vector_history) == 3 and not e.aborted
    # a failure before any twist exists: run() trips over np.abs(None) and returns its error tuple (:762, :819-821)
    eng = _IdealEngine(goal, fail_from=1)
    b = episode.EpisodeBatch(eng, np.zeros((4, 4, 3), np.uint8), goal, start, dt=1.0, min_iterations=5, max_iterations=8)
    b.step()
    assert b.episodes[0].done and b.episodes[0].aborted.startswith("Error in run loop") and b.episodes[0].iteration_count == 0


def test_fewer_than_four_pairs_is_a_zero_twist_through_the_ema():
    """1 <= n < 4 matches: calculate_uv returns num_pairs all-zero rows (vitvs_v2.py:539-541), len >= 4 passes :603,
    the error vector is zero, v = 0, and that zero goes through update_ema and is published."""
    goal = episode.make_pose((0, 0, -0.61))
    start = [episode.make_pose((0.05, 0.0, -0.7))]
    eng = _IdealEngine(goal, fail_from=3, fail_until=4, fail_nsel=2)
    b = episode.EpisodeBatch(eng, np.zeros((4, 4, 3), np.uint8), goal, start, dt=1.0, min_iterations=5, max_iterations=8)
    e = b.episodes[0]
    b.step(); b.step()
    v2 = e.v_c.copy()
    b.step()                                                                  # two pairs only
    a = ServoConfig().ema_alpha
    assert e.failures == 0 and np.allclose(e.v_c, (1 - a) * v2)               # alpha * 0 + (1 - alpha) * previous
    assert len(e.velocity_vector_history) == 3 and e.iteration_count == 3


def test_random_selection_hands_a_seeded_permutation_to_the_fused_call():
    goal = episode.make_pose((0, 0, -0.61))
    starts = [episode.make_pose((0.05, 0.0, -0.7)), episode.make_pose((0.0, 0.03, -0.6))]
    orders = []
    for _ in range(2):
        eng = _IdealEngine(goal)
        b = episode.EpisodeBatch(eng, np.zeros((4, 4, 3), np.uint8), goal, starts, random_selection=True)
        b.step()
        orders.append(eng.last_order.clone())
    assert orders[0].shape == (2, 484) and orders[0].dtype == torch.int32 and torch.equal(orders[0], orders[1])
    assert torch.equal(orders[0][0].sort().values, torch.arange(484, dtype=torch.int32))
    assert not torch.equal(orders[0][0], orders[0][1])


def test_cache_goal_sets_the_goal_once_and_forwards_only_current_frames():
    """EpisodeBatch(cache_goal=True): one Engine.set_goal call with one goal view per episode at construction, then
    Engine.current_to_twist per iteration (never frames_to_twist), the random-selection list still handed over, and the
    same trajectory as the default driver."""
    class _CachingEngine(_IdealEngine):
        def __init__(self, *a, **k):
            super().__init__(*a, **k)
            self.goal_sets, self.current_calls, self.pair_calls = [], 0, 0

        def set_goal(self, goal):
            self.goal_sets.append(tuple(goal.shape))

        def current_to_twist(self, cur, depth, want_match=False, sel_order=None):
            self.current_calls += 1
            return _IdealEngine.frames_to_twist(self, None, cur, depth, want_match, sel_order)

        def frames_to_twist(self, *a, **k):
            self.pair_calls += 1
            return super().frames_to_twist(*a, **k)

    goal = episode.make_pose((0, 0, -0.61))
    starts = [episode.make_pose((0.05, 0.0, -0.7)), episode.make_pose((0.0, 0.03, -0.6)), episode.make_pose((0.02, 0.01, -0.65))]
    poses = []
    for cache in (False, True):
        eng = _CachingEngine(goal)
        b = episode.EpisodeBatch(eng, np.zeros((4, 4, 3), np.uint8), goal, starts, cache_goal=cache, random_selection=True)
        for _ in range(5):
            b.step()
        poses.append(np.stack([e.pose for e in b.episodes]))
        if cache:
            assert eng.goal_sets == [(3, 4, 4, 3)] and eng.current_calls == 5 and eng.pair_calls == 0
            assert eng.last_order.shape == (3, 484)
        else:
            assert eng.goal_sets == [] and eng.current_calls == 0 and eng.pair_calls == 5
    assert np.array_equal(poses[0], poses[1])


def test_velocity_clip():
    goal = episode.make_pose((0, 0, -0.61))
    eng = _IdealEngine(goal, gain=50.0)
    eng.cfg.max_velocity = 0.1
    b = episode.EpisodeBatch(eng, np.zeros((4, 4, 3), np.uint8), goal, [episode.make_pose((0.3, 0, -0.61))], dt=0.1,
                             min_iterations=1, max_iterations=3, cfg=eng.cfg)
    b.step()
    e = b.episodes[0]
    assert np.abs(e.velocity_vector_history[0]).max() > 0.1          # the history keeps the UNCLIPPED EMA twist (:623-626)
    assert np.abs(np.array(e.applied[0])).max() <= 0.1 + 1e-15       # what is published (and integrated) is clipped


def test_oracle_render_geometry():
    tex = np.zeros((60, 80, 3), np.uint8)
    tex[:, :40] = 200                                                         # left half bright
    rgb, depth = render.render_plane(tex, episode.make_pose((0, 0, -0.61)))
    assert depth.min() == 610 and depth.max() == 610
    assert rgb[240, 100, 0] == 200 and rgb[240, 540, 0] == 0                  # left of the image = left of the poster
    # moving the camera to +x shifts the scene to the left in the image
    rgb2, _ = render.render_plane(tex, episode.make_pose((0.1, 0, -0.61)))
    assert (rgb2[240] == 200).sum() < (rgb[240] == 200).sum()
    # farther away: the poster no longer fills the frame, grey background appears; depth grows
    rgb3, depth3 = render.render_plane(tex, episode.make_pose((0, 0, -1.5)))
    assert (rgb3 == 128).any() and depth3.max() == 1500
