"""Closed-loop servo episodes, batched on one GPU (SURVEY §8f N2 + N3): the callers of the hot path.

Mirrors the reference's servo loop around ``Controller.ibvs()`` - EMA smoothing (vitvs_v2.py:329-343, :623),
velocity clipping (:668-676), the stopping rule of ``is_visual_servoing_done`` (:345-421) and the per-sample
records ``run()`` returns (:700-842) - but B episodes advance together: every iteration is ONE
``vvs_render_plane`` call (the headless stand-in for the Gazebo camera) and ONE ``vvs_frames_to_twist`` call for
all episodes still running.  Host code here is bookkeeping only (poses, histories, stopping rule); the frames
never leave the device.  ``save_npz`` writes the keys the reference's results file holds (:1493-1516) so that
``src/eval/*.py`` can read it.

Pose convention (csrc/render.cu): camera-to-world [R | t] as 12 doubles; the poster lies in z = 0; the camera at
(0, 0, -h) with R = I looks straight at it from h metres.  The twist is the camera-frame velocity
v_c = (vx, vy, vz, wx, wy, wz) the control law returns (vitvs_v2.py:622).
"""
from __future__ import annotations

from dataclasses import dataclass, field

import numpy as np
import torch

from .config import ServoConfig


# ----------------------------------------------------------------------------------------- rigid-body kinematics
def _skew(w):
    return np.array([[0.0, -w[2], w[1]], [w[2], 0.0, -w[0]], [-w[1], w[0], 0.0]])


def rotation_from_rotvec(w) -> np.ndarray:
    th = float(np.linalg.norm(w))
    K = _skew(w)
    if th < 1e-12:
        return np.eye(3) + K
    return np.eye(3) + (np.sin(th) / th) * K + ((1.0 - np.cos(th)) / (th * th)) * (K @ K)


def make_pose(position, rotvec=(0.0, 0.0, 0.0)) -> np.ndarray:
    """camera-to-world pose [12] from a camera position and a rotation vector (world frame)."""
    return np.concatenate([rotation_from_rotvec(np.asarray(rotvec, float)).reshape(-1), np.asarray(position, float)])


def integrate_twist(pose: np.ndarray, twist: np.ndarray, dt: float) -> np.ndarray:
    """Apply the camera-frame twist for dt seconds (what gazebo_vel_broadcaster.py:32-46 asks Gazebo to do)."""
    R, t = pose[:9].reshape(3, 3), pose[9:]
    return np.concatenate([(R @ rotation_from_rotvec(twist[3:] * dt)).reshape(-1), t + R @ (twist[:3] * dt)])


def end_error(pose: np.ndarray, goal: np.ndarray):
    """Controller.calculate_end_error (vitvs_v2.py:843-861): (position error in cm, orientation error in degrees)."""
    Ra, Rb = pose[:9].reshape(3, 3), goal[:9].reshape(3, 3)
    c = (np.trace(Ra.T @ Rb) - 1.0) / 2.0
    return float(np.linalg.norm(pose[9:] - goal[9:]) * 100.0), float(np.degrees(np.arccos(np.clip(c, -1.0, 1.0))))


def quaternion_xyzw(pose: np.ndarray) -> np.ndarray:
    R = pose[:9].reshape(3, 3)
    w = np.sqrt(max(0.0, 1.0 + R[0, 0] + R[1, 1] + R[2, 2])) / 2.0
    if w > 1e-8:
        return np.array([(R[2, 1] - R[1, 2]) / (4 * w), (R[0, 2] - R[2, 0]) / (4 * w), (R[1, 0] - R[0, 1]) / (4 * w), w])
    i = int(np.argmax(np.diag(R)))
    j, k = (i + 1) % 3, (i + 2) % 3
    s = np.sqrt(max(0.0, 1.0 + R[i, i] - R[j, j] - R[k, k])) * 2.0
    q = np.zeros(4)
    q[i] = s / 4.0
    q[j] = (R[j, i] + R[i, j]) / s
    q[k] = (R[k, i] + R[i, k]) / s
    q[3] = (R[k, j] - R[j, k]) / s
    return q


# ----------------------------------------------------------------------------------------- experiment sampling
# The reference samples in ITS world / body frames (Z up, camera body x = optical axis; vitvs_v2.py:1391-1418) - the
# functions below restate that sampling number for number (numpy legacy RNG, seed 41) and map the result into the
# optical camera-to-world convention of csrc/render.cu.
_REF_WORLD_TO_OURS = np.array([[0.0, -1.0, 0.0], [-1.0, 0.0, 0.0], [0.0, 0.0, -1.0]])   # x = -Y_ref, y = -X_ref, z = -Z_ref
_OPTICAL_IN_BODY = np.array([[0.0, 0.0, 1.0], [-1.0, 0.0, 0.0], [0.0, -1.0, 0.0]])      # columns: optical x, y, z in body axes


def sample_camera_positions(volume_dimensions, num_samples, desired_position):
    """vitvs_v2.py:959-978: uniform positions in a box centred on the desired position (uses np.random's global state)."""
    half = np.asarray(volume_dimensions, float) / 2
    return np.random.uniform(np.asarray(desired_position) - half, np.asarray(desired_position) + half, size=(num_samples, 3))


def sample_focal_points(num_samples, reference_point, num_circles, circle_radius_aug):
    """vitvs_v2.py:981-1010: looked-at points on concentric circles around the poster centre."""
    per = num_samples // num_circles
    pts = np.empty((num_samples, 3))
    for cn in range(num_circles):
        radius = circle_radius_aug * (cn + 1)
        theta = np.random.uniform(-np.pi, np.pi, size=per)
        pts[cn * per:(cn + 1) * per] = np.column_stack((np.cos(theta) * radius + reference_point[0],
                                                        np.sin(theta) * radius + reference_point[1],
                                                        np.repeat(reference_point[2], per)))
    return pts


def look_at_matrices(camera_positions, focal_points):
    """vitvs_v2.py:1253-1292: body-to-world rotations, body x = forward, world_up = (-1, 0, 0)."""
    out = np.empty((len(camera_positions), 3, 3))
    for i in range(len(camera_positions)):
        forward = focal_points[i] - camera_positions[i]
        forward = forward / np.linalg.norm(forward)
        right = -np.cross(forward, np.array([-1, 0, 0]))
        right = right / np.linalg.norm(right)
        up = np.cross(right, forward)
        out[i] = np.column_stack((forward, -right, up))
    return out


def apply_optical_axis_rotation(rotation_matrices, num_circles, samples_per_circle, rz_max=np.radians(120)):
    """vitvs_v2.py:1295-1342: per circle, rotations about the body x axis swept linearly over [-rz_max, rz_max]."""
    out = []
    for cn in range(num_circles):
        rz_values = np.linspace(-rz_max, rz_max, num=samples_per_circle)
        for i in range(samples_per_circle):
            idx = cn * samples_per_circle + i
            if idx >= len(rotation_matrices):
                break
            c, s_ = np.cos(rz_values[i]), np.sin(rz_values[i])
            out.append(np.dot(rotation_matrices[idx], np.array([[1, 0, 0], [0, c, -s_], [0, s_, c]])))
    return np.array(out)


def to_camera_pose(position_ref, body_to_world_ref) -> np.ndarray:
    """A reference (world Z-up, body-frame) camera pose -> our camera-to-world [R | t] (optical axes, poster in z = 0)."""
    R = _REF_WORLD_TO_OURS @ np.asarray(body_to_world_ref, float) @ _OPTICAL_IN_BODY
    return np.concatenate([R.reshape(-1), _REF_WORLD_TO_OURS @ np.asarray(position_ref, float)])


def sample_start_poses(num_samples=500, num_circles=4, circle_radius_aug=0.08, seed=41,
                       desired_position=(0.0, 0.0, 0.61), box_sample_size=(1.2, 1.2, 0.3),
                       reference_point=(0.0, 0.0, 0.01)) -> np.ndarray:
    """The start poses of the reference's experiment (main(), vitvs_v2.py:1385-1418; config.yaml num_samples /
    num_circles / circle_radius_aug; np.random.seed(41)) as [n, 12] camera-to-world poses in our convention."""
    per = num_samples // num_circles
    n = num_circles * per
    np.random.seed(seed)
    pos = sample_camera_positions(np.array(box_sample_size), n, np.array(desired_position))
    foc = sample_focal_points(n, np.array(reference_point), num_circles, circle_radius_aug)
    rot = apply_optical_axis_rotation(look_at_matrices(pos, foc), num_circles, per)
    return np.stack([to_camera_pose(pos[i], rot[i]) for i in range(n)])


def rotate_about_optical_axis(pose: np.ndarray, angle_degrees: float) -> np.ndarray:
    """rotate_camera_x_axis (vitvs_v2.py:1125-1148): the body x axis is the optical axis = our camera z."""
    a = np.radians(angle_degrees)
    Rz = np.array([[np.cos(a), -np.sin(a), 0.0], [np.sin(a), np.cos(a), 0.0], [0.0, 0.0, 1.0]])
    return np.concatenate([(pose[:9].reshape(3, 3) @ Rz).reshape(-1), pose[9:]])


def find_best_poses(engine, texture, goal_frame, poses, search_num_pairs: int = 48) -> np.ndarray:
    """find_and_set_best_pose (vitvs_v2.py:1151-1189), batched: every start pose is tried at 0 / 90 / 180 / 270 degrees
    about the optical axis with num_pairs raised to 48; the rotation whose selected pairs have the highest mean
    similarity wins (the first of equals, as the reference's strict `>`); a rotation with no match is skipped and a
    pose with none keeps its initial orientation.  One render + one frames_to_twist call per rotation."""
    poses = np.asarray(poses, float)
    n = len(poses)
    best_mean = np.full(n, -np.inf)
    best = poses.copy()
    K0 = engine.cfg.num_pairs
    engine.set_num_pairs(search_num_pairs)
    try:
        for angle in (0, 90, 180, 270):
            cand = poses if angle == 0 else np.stack([rotate_about_optical_axis(p, angle) for p in poses])
            for i0 in range(0, n, engine.max_images // 2):
                sl = slice(i0, min(n, i0 + engine.max_images // 2))
                frames, depth = engine.render_plane(texture, torch.from_numpy(cand[sl]))
                goal = goal_frame.expand(frames.shape[0], -1, -1, -1).contiguous()
                out = engine.frames_to_twist(goal, frames, depth)
                mean = out["mean_sim_sel"].double().cpu().numpy()
                ok = out["n_selected"].cpu().numpy() > 0
                better = ok & (mean > best_mean[sl])
                idx = np.nonzero(better)[0] + i0
                best_mean[idx] = mean[better]
                best[idx] = cand[idx]
    finally:
        engine.set_num_pairs(K0)
    return best


# ----------------------------------------------------------------------------------------- per-episode state
@dataclass
class Episode:
    pose: np.ndarray
    start_pose: np.ndarray
    ema: list = field(default_factory=lambda: [None] * 6)         # Controller.initialize_ema (vitvs_v2.py:325-327)
    v_c: np.ndarray | None = None                                  # Controller.v_c: last EMA twist (None until one exists, :223)
    iteration_count: int = 0
    failures: int = 0                                              # feature_failure_count (:500-508)
    done: bool = False
    converged: bool = False
    aborted: str = ""
    velocity_vector_history: list = field(default_factory=list)    # unclipped EMA twists, last 200 (:626-628): the stopping rule's window
    average_velocities: list = field(default_factory=list)         # mean |v_c| per iteration (:762-764)
    velocity_mean_100: list = field(default_factory=list)          # running means of the above (:767-775)
    velocity_mean_10: list = field(default_factory=list)
    position_history: list = field(default_factory=list)
    orientation_history: list = field(default_factory=list)
    applied: list = field(default_factory=list)                    # clipped, re-mapped twist published on /camera_vel
    lowest_position_error: float = float("inf")
    lowest_orientation_error: float = float("inf")
    initial_error_translation: float = 0.0
    initial_error_rotation: float = 0.0


class EpisodeBatch:
    """B closed-loop servo episodes against one goal view, advanced together on one Engine."""

    def __init__(self, engine, texture_u8, goal_pose: np.ndarray, start_poses, dt: float = 1.0,
                 min_iterations: int = 300, max_iterations: int = 2000, cfg: ServoConfig | None = None,
                 random_selection: bool = False, selection_seed: int = 121, cache_goal: bool = False):
        self.engine = engine
        # cache_goal: the goal view's descriptors are computed once (Engine.set_goal) and every iteration forwards only
        # the current frames - same twists bit for bit, about half the work.  The reference recomputes them each time
        # (vitvs_v2.py:482-487); False keeps its work per iteration.
        self.cache_goal = cache_goal
        self.cfg = cfg or engine.cfg
        # random_selection: the reference's choice of key-points - num_pairs mutual nearest neighbours drawn with
        # torch.randperm (vitvs_v2.py:139-141; main() seeds torch with 121, :1397) - handed to the fused call as a
        # per-episode priority list.  False: the num_pairs most similar ones (deterministic).
        self.random_selection = random_selection
        self._sel_gen = torch.Generator().manual_seed(selection_seed)
        self.dt, self.min_iterations, self.max_iterations = dt, min_iterations, max_iterations
        dev = engine.device
        self.texture = torch.as_tensor(np.ascontiguousarray(texture_u8)).to(dev)
        self.goal_pose = np.asarray(goal_pose, float)
        self.episodes = [Episode(pose=np.asarray(p, float).copy(), start_pose=np.asarray(p, float).copy())
                         for p in start_poses]
        if 2 * len(self.episodes) > engine.max_images:
            raise ValueError("engine.max_images must be >= 2 x the number of episodes")
        for e in self.episodes:
            e.initial_error_translation, e.initial_error_rotation = end_error(e.pose, self.goal_pose)
        # the goal view never changes: rendered once (the reference loads goalrgb.jpg once, vitvs_v2.py:255)
        self.goal_frame, _ = engine.render_plane(self.texture, torch.from_numpy(self.goal_pose[None]))
        if cache_goal:
            engine.set_goal(self.goal_frame.expand(len(self.episodes), -1, -1, -1).contiguous())

    # -- vitvs_v2.py:329-343
    def _update_ema(self, e: Episode, i: int, v: float) -> float:
        a = self.cfg.ema_alpha
        e.ema[i] = v if e.ema[i] is None else a * v + (1 - a) * e.ema[i]
        return e.ema[i]

    # -- vitvs_v2.py:345-421, with the iteration floor as a parameter (300 in the reference)
    def _is_done(self, e: Episode):
        if e.iteration_count < self.min_iterations:
            return False, False
        err_t, err_r = end_error(e.pose, self.goal_pose)
        if err_t > 2 * e.initial_error_translation:
            return True, False
        reduced = (err_t / max(e.initial_error_translation, 1e-12) < 0.1 and
                   err_r / max(e.initial_error_rotation, 1e-12) < 0.1)
        if len(e.velocity_vector_history) >= 200:
            recent = np.array(e.velocity_vector_history[-200:])
            first, second = recent[:100], recent[100:]
            ft = np.linalg.norm(first[:, :3] * 1000.0, axis=1).mean()
            fr = np.linalg.norm(np.degrees(first[:, 3:]), axis=1).mean()
            st = np.linalg.norm(second[:, :3] * 1000.0, axis=1).mean()
            sr = np.linalg.norm(np.degrees(second[:, 3:]), axis=1).mean()
            if ft < 1.0 and fr < 0.1 and st > ft and sr > fr:
                return True, reduced
        if e.iteration_count >= self.max_iterations:
            return True, reduced
        return False, False

    def step(self) -> int:
        """One iteration of every running episode; returns how many are still running."""
        live = [e for e in self.episodes if not e.done]
        if not live:
            return 0
        n = len(live)
        poses = torch.from_numpy(np.stack([e.pose for e in live]))
        frames, depth = self.engine.render_plane(self.texture, poses)
        order = None
        if self.random_selection:
            T = self.engine.tokens
            order = torch.stack([torch.randperm(T, generator=self._sel_gen) for _ in range(n)]).to(torch.int32)
        if self.cache_goal:      # every cached goal set is the same view: the first n of them serve any n live episodes
            out = self.engine.current_to_twist(frames, depth, sel_order=order)
        else:
            goal = self.goal_frame.expand(n, -1, -1, -1).contiguous()
            out = self.engine.frames_to_twist(goal, frames, depth, sel_order=order)
        twist = out["twist"].cpu().numpy()
        n_sel = out["n_selected"].cpu().numpy()
        mv = self.cfg.max_velocity
        hist_cap = getattr(self.cfg, "max_velocity_vector_history", 200)
        for k, e in enumerate(live):
            # --- Controller.ibvs() (vitvs_v2.py:588-628)
            if n_sel[k] == 0:                                     # detect_features failure (:500-508)
                e.failures += 1
                if e.failures >= 10:                               # RuntimeError -> run() aborts the sample (:746-756)
                    e.done, e.aborted = True, "Persistent feature detection failure"
                    continue
                # ibvs() returned early: self.v_c keeps the previous EMA twist and run() goes on with it (:758-795) -
                # the camera keeps moving, so the next frame differs and detection gets another chance
            else:
                e.failures = 0
                # 1 <= n < 4 pairs: calculate_uv hands back num_pairs all-zero rows (:539-541), len >= 4 passes :603,
                # e = 0 => v = 0, and that zero twist goes through the EMA like any other (the kernel returns it)
                e.v_c = np.array([self._update_ema(e, i, float(v)) for i, v in enumerate(twist[k])])
                e.velocity_vector_history.append(e.v_c.copy())     # the UNCLIPPED EMA twist (:623-626)
                if len(e.velocity_vector_history) > hist_cap:
                    e.velocity_vector_history.pop(0)
            if e.v_c is None:
                # no twist yet and none computed: run() trips over np.abs(None) and returns its error tuple (:762, :819-821)
                e.done, e.aborted = True, "Error in run loop: no velocity before the first successful detection"
                continue
            # --- the rest of one run() iteration (:758-795)
            e.iteration_count += 1
            e.average_velocities.append(float(np.mean(np.abs(e.v_c))))
            e.velocity_mean_100.append(float(np.mean(e.average_velocities[-100:])))
            e.velocity_mean_10.append(float(np.mean(e.average_velocities[-10:])))
            v_pub = np.clip(e.v_c, -mv, mv)                        # publish_twist clips every component (:668-676)
            e.applied.append([v_pub[2], -v_pub[0], -v_pub[1], v_pub[5], -v_pub[3], -v_pub[4]])
            e.pose = integrate_twist(e.pose, v_pub, self.dt)
            err_t, err_r = end_error(e.pose, self.goal_pose)
            e.position_history.append(e.pose[9:].copy())
            e.orientation_history.append(quaternion_xyzw(e.pose))
            e.lowest_position_error = min(e.lowest_position_error, err_t)
            e.lowest_orientation_error = min(e.lowest_orientation_error, err_r)
            done, conv = self._is_done(e)
            if done:
                e.done, e.converged = True, bool(conv)
        return sum(not e.done for e in self.episodes)

    def run(self) -> list[dict]:
        while self.step():
            pass
        return self.results()

    def results(self) -> list[dict]:
        out = []
        for e in self.episodes:
            err_t, err_r = end_error(e.pose, self.goal_pose)
            out.append(dict(converged=e.converged, aborted=e.aborted, iterations=e.iteration_count,
                            position_error_cm=err_t, orientation_error_deg=err_r,
                            initial_position_error_cm=e.initial_error_translation,
                            initial_orientation_error_deg=e.initial_error_rotation,
                            lowest_position_error_cm=e.lowest_position_error,
                            lowest_orientation_error_deg=e.lowest_orientation_error))
        return out

    def save_npz(self, path: str, total_execution_time: float = 0.0) -> None:
        """The keys of the reference's results file (vitvs_v2.py:1493-1516)."""
        E = self.episodes

        def obj(rows):                      # ragged per-episode histories: a 1-D object array, one entry per episode
            a = np.empty(len(rows), dtype=object)
            for i, r in enumerate(rows):
                a[i] = np.asarray(r)
            return a
        res = self.results()
        applied = [np.asarray(e.applied).reshape(-1, 6) for e in E]
        np.savez(path,
                 initial_positions=np.array([e.start_pose[9:] for e in E]),
                 initial_orientations=np.array([quaternion_xyzw(e.start_pose) for e in E]),
                 final_positions=np.array([e.pose[9:] for e in E]),
                 final_quaternions=np.array([quaternion_xyzw(e.pose) for e in E]),
                 convergence_flags=np.array([e.converged for e in E]),
                 position_errors=np.array([r["position_error_cm"] for r in res]),
                 orientation_errors=np.array([r["orientation_error_deg"] for r in res]),
                 best_poses=obj([e.start_pose for e in E]),
                 all_position_histories=obj([e.position_history for e in E]),
                 all_orientation_histories=obj([e.orientation_history for e in E]),
                 all_iteration_histories=np.array([e.iteration_count for e in E]),
                 lowest_position_errors=np.array([e.lowest_position_error for e in E]),
                 lowest_orientation_errors=np.array([e.lowest_orientation_error for e in E]),
                 all_average_velocities=obj([e.average_velocities for e in E]),
                 all_velocity_mean_100=obj([e.velocity_mean_100 for e in E]),
                 all_velocity_mean_10=obj([e.velocity_mean_10 for e in E]),
                 all_applied_velocity_x=obj([a[:, 0] for a in applied]),
                 all_applied_velocity_y=obj([a[:, 1] for a in applied]),
                 all_applied_velocity_z=obj([a[:, 2] for a in applied]),
                 all_applied_velocity_roll=obj([a[:, 3] for a in applied]),
                 all_applied_velocity_pitch=obj([a[:, 4] for a in applied]),
                 all_applied_velocity_yaw=obj([a[:, 5] for a in applied]),
                 total_execution_time=total_execution_time)
