"""TrackingEnv -- headless env shell around the fused sm_100a step.

The reference's env shell is `IGParkourEnv` (PARC/motion_tracker/envs/ig_parkour/ig_parkour_env.py):
it owns the simulator tensors, hands views of them to `DeepMimicEnv` and runs, after every physics
step, `_post_physics_step` (PARC/motion_tracker/envs/ig_env.py:368-377): heights -> time ->
reference motion -> observation -> reward -> termination, then `reset(done_ids)`.
The physics integrator (Isaac Gym) stays external, so this class keeps exactly that half: it owns
buffers with Isaac Gym's layout (`root_state [N,13]`, `dof_state [N,D,2]`, rigid-body and contact
tensors), exposes them for an integrator (or a host upload) to fill, and performs the whole
post-physics step as ONE kernel launch (`parc_b200_env_step`) plus one for the fail-rate EMA, and
the reference-state-init reset as one or two more. Config keys are the reference's yaml keys
(data/configs/tracker_config/dm_env_default.yaml).
"""
import copy

import numpy as np
import torch

from . import _lib
from .kin_char_model import KinCharModel
from .motion_lib import MotionLib
from .terrain import SubTerrain


def default_env_config():
    """The `env:` section of data/configs/tracker_config/dm_env_default.yaml (hot-path keys)."""
    return {
        "contact_bodies": [],
        "contact_detection_eps": 1.0e-05,
        "contact_weights": [5.0] * 15,
        "control_freq": 30,
        "demo_mode": False,
        "dm": {"min_motion_weight": 0.01, "terrains_per_motion": 1, "ignore_fail_rates": False},
        "enable_early_termination": True,
        "enable_tar_obs": True,
        "episode_length": 10.0,
        "global_obs": False,
        "global_root_height_obs": False,
        "joint_err_w": [1.0, 0.6, 0.6, 0.4, 0.0, 0.6, 0.4, 0.0, 1.0, 0.6, 0.4, 1.0, 0.6, 0.4],
        "key_bodies": ["right_hand", "left_hand", "right_foot", "left_foot"],
        "key_pos_w": 0.15,
        "max_obs_h": 3.0,
        "min_obs_h": -3.0,
        "never_done": False,
        "pose_termination": True,
        "pose_termination_dist": [0.7, 1.0, 0.7, 0.7, 0.7, 0.7, 0.7, 0.7, 1.0, 1.2, 10.0, 1.0, 1.2, 10.0],
        "pose_w": 0.5,
        "rand_reset": True,
        "rand_root_pos_offset_scale": 0.075,
        "root_pos_termination_dist": 0.6,
        "root_pos_w": 0.15,
        "root_rot_termination_angle": 1.309,
        "root_vel_w": 0.1,
        "tar_obs_steps": [1, 2, 3, 10, 20, 30],
        "termination_height": 0.15,
        "track_root": True,
        "track_root_h": True,
        "use_contact_info": True,
        "vel_w": 0.1,
    }


def _f32(x):
    return float(np.float32(x))


def build_config_struct(env_config, char_model: KinCharModel, num_ray_points: int, ray_xy_points=None) -> _lib.EnvConfigStruct:
    """yaml `env:` dict -> parc_env_config_t, evaluating every derived constant the way the
    reference's Python does (fp32 tensors times Python doubles, ig_parkour_env.py:43-149)."""
    ec = env_config
    J, D = char_model.get_num_joints(), char_model.get_dof_size()
    c = _lib.EnvConfigStruct()
    steps = list(ec.get("tar_obs_steps", [1]))
    assert len(steps) <= _lib.MAX_TAR_STEPS, "at most %d target steps" % _lib.MAX_TAR_STEPS
    timestep = 1.0 / ec["control_freq"]
    c.num_tar_steps = len(steps)
    # fetch_tar_obs_data: `timestep * tar_obs_steps` with an int32 tensor -> fp32 (mgdm_dm_util.py:232)
    tdt = (timestep * torch.tensor(steps, dtype=torch.int32)).to(torch.float32)
    for i, v in enumerate(tdt.tolist()):
        c.tar_dt[i] = v
    keys = [char_model.get_body_id(b) for b in ec.get("key_bodies", [])]
    assert len(keys) <= _lib.MAX_KEY_BODIES
    c.num_key_bodies = len(keys)
    for i, k in enumerate(keys):
        c.key_body_ids[i] = k
    c.num_ray_points = int(num_ray_points)
    if ray_xy_points is not None:                                  # lattice ray sets (get_xy_grid_points) take the shared-product path
        assert int(ray_xy_points.shape[0]) == c.num_ray_points
        c.ray_grid_nx, c.ray_grid_ny = _lib.ray_points_lattice(ray_xy_points)
    c.global_obs = int(bool(ec["global_obs"]))
    c.root_height_obs = int(bool(ec["global_root_height_obs"]))
    c.enable_tar_obs = int(bool(ec.get("enable_tar_obs", True)))
    c.use_contact_info = int(bool(ec["use_contact_info"]))
    c.track_root = int(bool(ec["track_root"]))
    c.track_root_h = int(bool(ec["track_root_h"]))
    c.enable_early_termination = int(bool(ec["enable_early_termination"]))
    c.pose_termination = int(bool(ec.get("pose_termination", False)))
    cb = [char_model.get_body_id(b) for b in ec.get("contact_bodies", [])]
    c.num_fall_contact_bodies = len(cb)
    mask = 0
    for b in cb:
        mask |= 1 << b
    c.fall_contact_body_mask = mask
    c.never_done = int(bool(ec.get("never_done", False)))
    c.timestep = _f32(timestep)                                    # ig_env.py:393: fp32 tensor * python float
    c.episode_length = _f32(ec["episode_length"])
    c.termination_height = _f32(ec["termination_height"])
    d = float(ec["root_pos_termination_dist"])
    c.root_pos_termination_dist = _f32(d * d)                      # compute_done compares against dist*dist (double -> fp32)
    c.root_rot_termination_angle = _f32(ec["root_rot_termination_angle"])
    ptd = list(ec["pose_termination_dist"])
    assert len(ptd) == J - 1, "pose_termination_dist needs one entry per non-root body"
    for i, v in enumerate(ptd):
        c.pose_termination_dist[i] = v
    jw = ec.get("joint_err_w")
    jw = [1.0] * (J - 1) if jw is None else list(jw)
    assert len(jw) == J - 1
    for i, v in enumerate(jw):
        c.joint_err_w[i] = v
    for j in range(1, J):                                          # _parse_joint_err_weights, :1134-1153
        jt = char_model.get_joint(j)
        for k in range(jt.get_dof_dim()):
            c.dof_err_w[jt.dof_idx + k] = jw[j - 1]
    cw = list(ec.get("contact_weights", [0.0] * J))
    for i, v in enumerate(cw[:J]):
        c.contact_weights[i] = v
    tw = ec["pose_w"] + ec["vel_w"] + ec["root_pos_w"] + ec["root_vel_w"] + ec["key_pos_w"]
    c.pose_w, c.vel_w = ec["pose_w"] / tw, ec["vel_w"] / tw
    c.root_pos_w, c.root_vel_w, c.key_pos_w = ec["root_pos_w"] / tw, ec["root_vel_w"] / tw, ec["key_pos_w"] / tw
    c.contact_eps = 1e-5                                           # _get_char_contact_state default eps (:655)
    c.min_obs_h, c.max_obs_h = ec["min_obs_h"], ec["max_obs_h"]
    ema_w = 0.01                                                   # DeepMimicEnv._ema_weight, dm_env.py:90
    c.ema_weight = ema_w
    c.ema_keep = _f32(1.0 - ema_w)
    c.min_motion_weight = ec.get("dm", {}).get("min_motion_weight", 0.01)
    c.rand_root_pos_offset_scale = ec.get("rand_root_pos_offset_scale", 0.0)
    return c


REWARD_TERM_NAMES = ("pose_r", "vel_r", "root_pos_r", "root_vel_r", "key_pos_r", "contact_penalty", "total_r")


def make_buffers_struct(num_envs, *, obs_buf, reward_buf, done_buf, time_buf, timestep_buf, env_offsets, motion_ids,
                        motion_terrain_ids, motion_time_offsets, dm_motion_offsets, terrains_per_motion=1, reward_terms=None,
                        actors_need_reset=None, ray_xy_points=None, ray_hfs=None, motion_id_fail_rates=None, motion_weights=None,
                        done_count=None, done_list=None, done_code=None, **views) -> _lib.EnvBuffersStruct:
    """parc_env_buffers_t from the tensors the env shell injects into DeepMimicEnv (get_sim_tensor_views /
    get_data_buffer_views / get_obs_buf_view, mgdm_dm_util.py:37-84). `views` are the simulator and ref_* tensors by the
    names of include/parc_b200_env.h (char_root_pos, ..., ref_contacts): column slices and stride-2 slices are passed as
    they are -- the kernels read and write through (ptr, row_stride, mid_stride, elem_stride)."""
    unknown = set(views) - set(_lib._VIEW_NAMES)
    assert not unknown, "unknown state views: %s" % sorted(unknown)
    b = _lib.EnvBuffersStruct()
    b.num_envs = int(num_envs)
    for name in _lib._VIEW_NAMES:
        setattr(b, name, _lib.view3(views.get(name)))
    p = _lib.ptr
    b.reward_buf, b.reward_terms, b.done_buf = p(reward_buf), p(reward_terms), p(done_buf)
    b.time_buf, b.timestep_buf = p(time_buf), p(timestep_buf)
    b.actors_need_reset = p(actors_need_reset)
    b.actors_per_env = int(actors_need_reset.shape[1]) if actors_need_reset is not None else 0
    b.env_offsets, b.ray_xy_points, b.ray_hfs = p(env_offsets), p(ray_xy_points), p(ray_hfs)
    b.obs_buf, b.obs_row_stride = p(obs_buf), int(obs_buf.stride(0))
    b.motion_ids, b.motion_terrain_ids = p(motion_ids), p(motion_terrain_ids)
    b.motion_time_offsets, b.motion_id_fail_rates = p(motion_time_offsets), p(motion_id_fail_rates)
    b.motion_weights, b.dm_motion_offsets = p(motion_weights), p(dm_motion_offsets)
    b.terrains_per_motion = int(terrains_per_motion)
    b.done_count, b.done_list, b.done_code = p(done_count), p(done_list), p(done_code)
    return b


class TrackingEnv:
    """Post-physics half of the tracking env on one GPU. `env_offset_rows` lets a rank of a
    sharded job own envs [first_env, first_env + num_envs) of a larger grid."""

    def __init__(self, env_config, num_envs, device, char_model: KinCharModel, motion_lib: MotionLib,
                 terrain: SubTerrain, dm_motion_offsets, ray_xy_points, env_offsets=None, seed=0,
                 keep_ray_hfs=False, interleaved_state=False):
        dev = torch.device(device)
        assert dev.type == "cuda", "TrackingEnv runs on a CUDA device only (no CPU fallback)"
        self._device = dev
        self._env_config = copy.deepcopy(env_config)
        self._num_envs = N = int(num_envs)
        self._kin_char_model = char_model
        self._mlib = self._motion_lib = motion_lib
        self._terrain = terrain
        J, D = char_model.get_num_joints(), char_model.get_dof_size()
        self._J, self._D = J, D
        f32, i32, i64 = torch.float32, torch.int32, torch.int64
        z = lambda *s, dt=f32: torch.zeros(*s, dtype=dt, device=dev)

        self._ray_xy_points = ray_xy_points.to(device=dev, dtype=f32).contiguous()
        P = int(self._ray_xy_points.shape[0])
        self._cfg = build_config_struct(env_config, char_model, P, self._ray_xy_points)
        self._dm_motion_offsets = dm_motion_offsets.to(device=dev, dtype=f32).contiguous()
        self._terrains_per_motion = int(self._dm_motion_offsets.shape[1])
        assert self._dm_motion_offsets.shape[0] >= motion_lib.num_motions()
        self._env_offsets = (z(N, 3) if env_offsets is None else env_offsets.to(device=dev, dtype=f32)).contiguous()
        assert self._env_offsets.shape == (N, 3)

        # simulator tensors in Isaac Gym's layout (ig_char_env.py:172-196): the char_* names are views.
        # interleaved_state: the four tensors a simulator hands over every step are column blocks of ONE [N, 13 + 2D + 6J] row
        # (root 13 | dof D x 2 | rigid-body pos 3J | contact forces 3J; 636 bytes for the humanoid), so a host-fed step uploads a
        # row range with one copy instead of four. The kernels read through strided views either way.
        self._interleaved = bool(interleaved_state)
        if self._interleaved:
            W = 13 + 2 * D + 6 * J
            self._sim_state = z(N, W)
            self._root_state = self._sim_state[:, 0:13]
            self._dof_state = self._sim_state[:, 13:13 + 2 * D].unflatten(1, (D, 2))
            self._char_rigid_body_pos = self._sim_state[:, 13 + 2 * D:13 + 2 * D + 3 * J].unflatten(1, (J, 3))
            self._char_contact_forces = self._sim_state[:, 13 + 2 * D + 3 * J:W].unflatten(1, (J, 3))
        else:
            self._sim_state = None
            self._root_state = z(N, 13)
            self._dof_state = z(N, D, 2)
            self._char_rigid_body_pos = z(N, J, 3)
            self._char_contact_forces = z(N, J, 3)
        self._root_state[:, 6] = 1.0
        self._char_root_pos = self._root_state[:, 0:3]
        self._char_root_rot = self._root_state[:, 3:7]
        self._char_root_vel = self._root_state[:, 7:10]
        self._char_root_ang_vel = self._root_state[:, 10:13]
        self._char_dof_pos = self._dof_state[:, :, 0]
        self._char_dof_vel = self._dof_state[:, :, 1]
        self._char_rigid_body_vel = z(N, J, 3)
        self._char_rigid_body_ang_vel = z(N, J, 3)
        # reference-motion state
        self._ref_root_pos, self._ref_root_rot = z(N, 3), z(N, 4)
        self._ref_root_vel, self._ref_root_ang_vel = z(N, 3), z(N, 3)
        self._ref_body_pos, self._ref_joint_rot = z(N, J, 3), z(N, J - 1, 4)
        self._ref_dof_pos, self._ref_dof_vel, self._ref_contacts = z(N, D), z(N, D), z(N, J)
        # data buffers
        self._obs_dim = _lib.lib().parc_b200_obs_dim(char_model._cstruct(), self._cfg)
        assert self._obs_dim > 0
        self._obs_buf = z(N, self._obs_dim)
        self._reward_buf, self._time_buf = z(N), z(N)
        self._reward_terms = z(N, 8)
        self._done_buf, self._timestep_buf = z(N, dt=i32), z(N, dt=i32)
        self._actors_need_reset = torch.zeros(N, 1, dtype=torch.bool, device=dev)
        self._ray_hfs = z(N, P) if keep_ray_hfs else None
        self._motion_ids, self._motion_terrain_ids = z(N, dt=i64), z(N, dt=i64)
        self._motion_time_offsets = z(N)
        M = motion_lib.num_motions()
        self._motion_id_fail_rates = torch.ones(M, dtype=f32, device=dev)
        self._motion_start_time_fraction = z(N)
        # per-step scratch
        self._done_count = z(1, dt=i32)
        self._done_list = z(N, dt=i32)
        self._done_code = torch.zeros(N, dtype=torch.uint8, device=dev)
        # RSI sampling CDF (dm_env.py:485-490): clamp(fail_rate, min_motion_weight) * motion_weight, inclusive prefix sum. Valid from
        # the start (fail rates are 1) -- parc_b200_fail_rate_ema refreshes it whenever the rates change
        self._cdf = torch.cumsum(torch.clamp(self._motion_id_fail_rates, min=float(self._cfg.min_motion_weight)) * motion_lib._motion_weights, 0).contiguous()
        self._cdf_weights_only = torch.cumsum(motion_lib._motion_weights, 0).contiguous()      # dm.ignore_fail_rates
        self._ignore_fail_rates = bool(env_config.get("dm", {}).get("ignore_fail_rates", False))
        self._ema_scratch = z(2 * M + N, dt=i32)
        self._seed = int(seed)
        self._step_count = 0
        self._rand_reset = bool(env_config.get("rand_reset", True))
        self._demo_mode = bool(env_config.get("demo_mode", False))
        self._root_pos_offset = None
        self._root_rot_offset = None
        self._fail_rate_sync = None       # sharding.FailRateSync: cross-rank combination of the fail rates (enable_fail_rate_sync)
        self._kernel_events = None        # a list: post_physics_step appends (start, end) CUDA events around the fused step launch
        self._build_buffers_struct()

    # ---------------------------------------------------------------- descriptors
    def _build_buffers_struct(self):
        names = list(_lib._VIEW_NAMES) + ["reward_buf", "reward_terms", "done_buf", "time_buf", "timestep_buf", "actors_need_reset",
                                          "env_offsets", "ray_xy_points", "ray_hfs", "obs_buf", "motion_ids", "motion_terrain_ids",
                                          "motion_time_offsets", "motion_id_fail_rates", "dm_motion_offsets", "done_count",
                                          "done_list", "done_code"]
        self._bufs = make_buffers_struct(self._num_envs, motion_weights=self._mlib._motion_weights,
                                         terrains_per_motion=self._terrains_per_motion,
                                         **{n: getattr(self, "_" + n) for n in names})
        self._terrain_cs = self._terrain.cstruct()

    def _stream(self):
        return _lib.stream_ptr(self._device)

    # ---------------------------------------------------------------- accessors
    def get_num_envs(self):
        return self._num_envs

    def get_obs_dim(self):
        return self._obs_dim

    def get_obs(self):
        return self._obs_buf

    def get_reward(self):
        return self._reward_buf

    def get_done(self):
        return self._done_buf

    def get_reward_terms(self):
        """dict of [N] views named like IGParkourEnv._info["rewards"] (ig_parkour_env.py:1015-1044)."""
        return {n: self._reward_terms[:, i] for i, n in enumerate(REWARD_TERM_NAMES)}

    def compute_tracking_error(self):
        """IGParkourEnv._update_info with report_tracking_error (ig_parkour_env.py:1066-1095): FK of the
        simulated and of the reference character, then compute_tracking_error -> [N,7] (what the
        reference stores in info["tracking_error"])."""
        from . import dm_util
        m = self._kin_char_model
        char_body_pos, char_body_rot = m.forward_kinematics(self._char_root_pos, self._char_root_rot, m.dof_to_rot(self._char_dof_pos))
        ref_body_pos, ref_body_rot = m.forward_kinematics(self._ref_root_pos, self._ref_root_rot, self._ref_joint_rot)
        return dm_util.compute_tracking_error(
            self._char_root_pos, self._char_root_rot, char_body_rot, char_body_pos, self._ref_root_pos, self._ref_root_rot, ref_body_rot,
            ref_body_pos, self._char_root_vel, self._char_root_ang_vel, self._char_dof_vel, self._ref_root_vel, self._ref_root_ang_vel,
            self._ref_dof_vel)

    def sim_state_tensors(self):
        """The tensors an external integrator (or a host upload) writes every step: one interleaved [N, W] row per env
        (interleaved_state=True) or Isaac Gym's four tensors."""
        if self._interleaved:
            return {"sim_state": self._sim_state}
        return {"root_state": self._root_state, "dof_state": self._dof_state,
                "rigid_body_pos": self._char_rigid_body_pos, "contact_forces": self._char_contact_forces}

    def sim_state_parts(self):
        """The same state by Isaac Gym's tensor names (views of the interleaved row when interleaved_state=True)."""
        return {"root_state": self._root_state, "dof_state": self._dof_state,
                "rigid_body_pos": self._char_rigid_body_pos, "contact_forces": self._char_contact_forces}

    def set_demo_mode(self, val=None):
        self._demo_mode = (not self._demo_mode) if val is None else bool(val)
        return self._demo_mode

    def set_motion_start_time_fraction(self, val):
        assert val.shape == (self._num_envs,)
        self._motion_start_time_fraction = val.to(device=self._device, dtype=torch.float32).contiguous()

    def set_rand_root_pos_offset_scale(self, val):
        self._env_config["rand_root_pos_offset_scale"] = val
        self._cfg.rand_root_pos_offset_scale = val

    def set_root_pos_offset(self, val=None):
        if val is not None:
            assert val.shape == (self._num_envs, 3)
            val = val.to(device=self._device, dtype=torch.float32).contiguous()
        self._root_pos_offset = val

    def set_root_rot_offset(self, val=None):
        if val is not None:
            assert val.shape == (self._num_envs, 4)
            val = val.to(device=self._device, dtype=torch.float32).contiguous()
        self._root_rot_offset = val

    # ---------------------------------------------------------------- step
    def post_physics_step(self, parts=_lib.STEP_ALL, update_fail_rates=True):
        """IGEnv._post_physics_step (ig_env.py:368-377) in one launch, then the fail-rate EMA +
        RSI sampling CDF (dm_env.py:648-660, 487-490) in one more."""
        km = self._kin_char_model._cstruct()
        L = _lib.lib()
        if parts & _lib.STEP_DONE:
            self._done_count.zero_()      # the step kernel appends finished envs to done_list
        ev = None
        if self._kernel_events is not None:
            ev = (torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
            ev[0].record()
        _lib.check(L.parc_b200_env_step(km, self._mlib.cstruct(), self._terrain_cs, self._cfg, self._bufs, parts, None, 0,
                                        self._stream()))
        if ev is not None:
            ev[1].record()
            self._kernel_events.append(ev)
        if (parts & _lib.STEP_DONE) and update_fail_rates:
            _lib.check(L.parc_b200_fail_rate_ema(self._cfg, self._bufs, self._mlib.num_motions(), _lib.ptr(self._cdf),
                                                 _lib.ptr(self._ema_scratch), self._stream()))
            self._sync_fail_rates()
        self._step_count += 1

    def enable_fail_rate_sync(self, interval=1, group=None):
        """Env-sharded runs (one process per GPU): combine the per-motion fail / pass counts of all ranks every `interval` steps
        with one all-reduce, so that every rank samples RSI motions from the same fail rates (sharding.FailRateSync)."""
        from . import sharding
        self._fail_rate_sync = sharding.FailRateSync(self._motion_id_fail_rates, interval, float(self._cfg.ema_weight), group)

    def _sync_fail_rates(self):
        if self._fail_rate_sync is None:
            return
        M = int(self._motion_id_fail_rates.shape[0])
        count = lambda acc: _lib.check(_lib.lib().parc_b200_outcome_counts(self._bufs, M, _lib.ptr(acc), self._stream()))
        if self._fail_rate_sync.after_step(self._motion_id_fail_rates, self._done_code, self._motion_ids, accumulate=count):
            w = torch.clamp(self._motion_id_fail_rates, min=float(self._cfg.min_motion_weight)) * self._mlib._motion_weights
            self._cdf.copy_(torch.cumsum(w, 0))           # the RSI sampling CDF follows the replaced rates

    def reset_done(self):
        """env.reset(done_ids) of the rollout loop (base_agent.py:366-370) without leaving the
        device: RSI draws from a Philox counter keyed by (seed, env id, step), state reset, heights
        and observation rows of the reset envs. Consumes the done list of the last step."""
        opt = _lib.ResetOptionsStruct()
        opt.root_pos_offset, opt.root_rot_offset = _lib.ptr(self._root_pos_offset), _lib.ptr(self._root_rot_offset)
        opt.motion_start_time_fraction = _lib.ptr(self._motion_start_time_fraction)
        opt.demo_mode = int(self._demo_mode)
        cdf = self._cdf_weights_only if self._ignore_fail_rates else self._cdf
        _lib.check(_lib.lib().parc_b200_env_reset_sampled(
            self._kin_char_model._cstruct(), self._mlib.cstruct(), self._terrain_cs, self._cfg, self._bufs,
            _lib.ptr(cdf), self._mlib.num_motions(), self._seed, self._step_count, int(self._rand_reset), opt,
            self._stream()))

    def sim_standin(self, noise):
        """Benchmark utility (not a reference method): stands in for the simulator between two steps --
        character state := current reference state + a fixed per-env perturbation (synth.standin_noise)."""
        assert noise.shape == (self._num_envs, 13 + 2 * self._D + 3 * self._J) and noise.dtype == torch.float32
        _lib.check(_lib.lib().parc_b200_sim_standin(self._kin_char_model._cstruct(), self._bufs, _lib.ptr(noise), self._stream()))

    def step(self):
        """One full control step after physics: post_physics_step + reset of the finished envs."""
        self.post_physics_step()
        self.reset_done()

    # ---------------------------------------------------------------- reset with the reference's RNG
    def draw_reset(self, env_ids):
        """The random draws of DeepMimicEnv._reset_ref_motion / add_noise_to_char_state
        (dm_env.py:473-502, mgdm_dm_util.py:102-106) from torch's generator, in the reference's
        call order: multinomial, randint, rand (sample_time), rand (xy noise)."""
        n = env_ids.shape[0]
        dev = self._device
        mlib = self._mlib
        if self._demo_mode:
            motion_ids = env_ids % mlib.num_motions()
        elif self._ignore_fail_rates:
            motion_ids = mlib.sample_motions(n)
        else:
            w = torch.clamp(self._motion_id_fail_rates, min=self._cfg.min_motion_weight) * mlib._motion_weights
            motion_ids = mlib.sample_motions(n, w)
        terrain_ids = torch.randint(high=self._terrains_per_motion, size=(n,), dtype=torch.int64, device=dev)
        if self._rand_reset:
            times = mlib.sample_time(motion_ids)
        else:
            times = mlib.get_motion_length(motion_ids) * self._motion_start_time_fraction[env_ids]
        noise = torch.rand(size=[n, 2], device=dev, dtype=torch.float32) * 2.0 - 1.0
        return motion_ids, terrain_ids, times, noise

    def reset(self, env_ids, draws=None, refresh_obs=True):
        """IGParkourEnv.reset (ig_parkour_env.py:809-829) -> DeepMimicEnv.reset (dm_env.py:567-592).
        `draws` = (motion_ids, terrain_ids, times, xy_noise) to bypass torch's RNG (parity tests)."""
        n = int(env_ids.shape[0])
        if n == 0:
            return
        env_ids = env_ids.to(device=self._device, dtype=torch.int64).contiguous()
        mids, tids, times, noise = self.draw_reset(env_ids) if draws is None else draws
        dev, i64, f32 = self._device, torch.int64, torch.float32
        mids = mids.to(device=dev, dtype=i64).contiguous()
        tids = tids.to(device=dev, dtype=i64).contiguous()
        times = times.to(device=dev, dtype=f32).contiguous()
        noise = noise.to(device=dev, dtype=f32).contiguous()
        assert mids.shape[0] == n and tids.shape[0] == n and times.shape[0] == n and noise.shape == (n, 2)
        p = _lib.ptr
        L = _lib.lib()
        km = self._kin_char_model._cstruct()
        _lib.check(L.parc_b200_env_reset(km, self._mlib.cstruct(), self._cfg, self._bufs, p(env_ids), n, p(mids), p(tids),
                                         p(times), p(noise), p(self._root_pos_offset), p(self._root_rot_offset),
                                         self._stream()))
        if refresh_obs:
            # _refresh_sim_tensors refreshes ALL heights, _update_observations only the reset rows
            if self._ray_hfs is not None:
                _lib.check(L.parc_b200_env_step(km, self._mlib.cstruct(), self._terrain_cs, self._cfg, self._bufs,
                                                _lib.STEP_HEIGHTS, None, 0, self._stream()))
            _lib.check(L.parc_b200_env_step(km, self._mlib.cstruct(), self._terrain_cs, self._cfg, self._bufs,
                                            _lib.STEP_HEIGHTS | _lib.STEP_OBS, p(env_ids), n, self._stream()))

    # ---------------------------------------------------------------- recording path (stage 4)
    def build_agent_states_dict(self, name_suffix="", record_obs=False, max_frames=None, record_ref=False, bypass_record_fail=False):
        """IGParkourEnv.build_agent_states_dict (ig_parkour_env.py:1153-1182): start recording every env. The reference
        keeps per-env Python lists of per-step tensors; here the tape is [N, max_frames, ...] in HBM (max_frames defaults
        to one episode + 2 frames) and write_agent_states is one launch."""
        N, dev = self._num_envs, self._device
        if max_frames is None:
            max_frames = int(round(self._cfg.episode_length / self._cfg.timestep)) + 2
        C = self._mlib.cstruct().num_contacts if record_ref else self._J
        z = lambda *s, dt=torch.float32: torch.zeros(s, dtype=dt, device=dev)
        self._tape = dict(frames=z(N, max_frames, 6 + self._D), contacts=z(N, max_frames, C),
                          obs=z(N, max_frames, self._obs_dim) if record_obs else None, cursor=z(N, dt=torch.int32),
                          writing=torch.ones(N, dtype=torch.uint8, device=dev), success=z(N, dt=torch.uint8),
                          saved_list=z(N, dt=torch.int32), saved_count=z(1, dt=torch.int32),
                          still_recording=torch.ones(1, dtype=torch.int32, device=dev), overflow=z(1, dt=torch.int32))
        t = _lib.RecordTapeStruct()
        for k in ("frames", "contacts", "obs", "cursor", "writing", "success", "saved_list", "saved_count", "still_recording", "overflow"):
            setattr(t, k, _lib.ptr(self._tape[k]))
        t.max_frames, t.obs_dim = int(max_frames), int(self._obs_dim if record_obs else 0)
        self._tape_cs = t
        self._record_obs, self._record_ref, self._bypass_record_fail = bool(record_obs), bool(record_ref), bool(bypass_record_fail)
        self._save_motion_name_suffix = name_suffix

    def write_agent_states(self):
        """IGParkourEnv.write_agent_states (:759-796) for all envs in one launch; call it after post_physics_step and
        before the reset, as the reference does (:696). Nothing is copied to the host."""
        _lib.check(_lib.lib().parc_b200_record_agent_states(
            self._kin_char_model._cstruct(), self._mlib.cstruct(), self._cfg, self._bufs, self._tape_cs, int(self._record_ref),
            int(self._bypass_record_fail), 1.0 / float(self._env_config["control_freq"]) if "control_freq" in self._env_config
            else float(self._cfg.timestep), self._stream()))

    def is_writing_agent_states(self):
        """:742-743 -- reads the device flag (synchronises)."""
        return bool(self._tape["still_recording"].item())

    def get_env_success_states(self):
        return self._tape["success"].bool()

    def envs_finished_this_step(self):
        """Env ids the reference would pass to save_agent_states_to_file in the last write_agent_states call, ascending."""
        n = int(self._tape["saved_count"].item())
        return torch.sort(self._tape["saved_list"][:n]).values

    def get_agent_motion(self, env_id):
        """The recorded dict of one env as save_agent_states_to_file builds it (:698-735) before the terrain slice:
        frames [T, 6+D] with global xy (:707-708), contacts [T, C], optionally obs [T, obs_dim]."""
        T = int(self._tape["cursor"][env_id].item())
        frames = self._tape["frames"][env_id, :T].clone()
        frames[:, 0:2] += self._env_offsets[env_id, 0:2]                      # _get_global_xy_pos, :1184-1188
        out = {"fps": int(round(1.0 / self._cfg.timestep)), "loop_mode": "CLAMP", "frames": frames,
               "contacts": self._tape["contacts"][env_id, :T].clone()}
        if self._record_obs:
            out["obs"] = self._tape["obs"][env_id, :T].clone()
        return out

    # ---------------------------------------------------------------- host <-> device (e2e path)
    def make_host_state(self):
        """Pinned host staging buffers with the layout of sim_state_tensors()."""
        return {k: torch.empty(v.shape, dtype=v.dtype, pin_memory=True) for k, v in self.sim_state_tensors().items()}

    def upload_sim_state(self, host_state):
        for k, v in self.sim_state_tensors().items():
            v.copy_(host_state[k], non_blocking=True)

    def post_physics_step_host(self, host_state, num_chunks=6, standin_noise=None, parts=_lib.STEP_ALL, update_fail_rates=True):
        """post_physics_step for simulator state that lives in pinned HOST memory (make_host_state):
        parc_b200_env_step_host uploads the envs in `num_chunks` row ranges on a copy stream and runs
        the fused step on each range as soon as its rows have arrived, so the PCIe transfer overlaps
        the computation. Results equal upload_sim_state + post_physics_step."""
        sim = self.sim_state_tensors()
        rows = (_lib.HostRowsStruct * len(sim))()
        for i, (k, v) in enumerate(sim.items()):
            h = host_state[k]
            assert h.shape == v.shape and h.dtype == v.dtype and h.is_contiguous() and v.is_contiguous() and not h.is_cuda
            assert h.is_pinned(), "host state must be pinned (make_host_state) for the upload to overlap"
            rows[i].dst_device, rows[i].src_host = v.data_ptr(), h.data_ptr()
            rows[i].row_bytes = v[0].numel() * v.element_size()
        L = _lib.lib()
        if parts & _lib.STEP_DONE:
            self._done_count.zero_()
        _lib.check(L.parc_b200_env_step_host(self._kin_char_model._cstruct(), self._mlib.cstruct(), self._terrain_cs, self._cfg,
                                             self._bufs, parts, rows, len(sim), int(num_chunks), _lib.ptr(standin_noise), self._stream()))
        if (parts & _lib.STEP_DONE) and update_fail_rates:
            _lib.check(L.parc_b200_fail_rate_ema(self._cfg, self._bufs, self._mlib.num_motions(), _lib.ptr(self._cdf),
                                                 _lib.ptr(self._ema_scratch), self._stream()))
            self._sync_fail_rates()
        self._step_count += 1

    def step_host(self, host_state, host_results, num_chunks=6, standin_noise=None):
        """One full control step from host buffers to host buffers: chunked upload + fused step, fail-rate
        EMA, device-side RSI reset, reward/done back to the host (asynchronous: synchronise the current
        stream before reading host_results)."""
        self.post_physics_step_host(host_state, num_chunks=num_chunks, standin_noise=standin_noise)
        self.download_results(host_results)     # the rollout reads reward/done of THIS step; the reset follows
        self.reset_done()

    def make_host_results(self):
        return {"reward": torch.empty(self._num_envs, dtype=torch.float32, pin_memory=True),
                "done": torch.empty(self._num_envs, dtype=torch.int32, pin_memory=True)}

    def download_results(self, host_results):
        host_results["reward"].copy_(self._reward_buf, non_blocking=True)
        host_results["done"].copy_(self._done_buf, non_blocking=True)
