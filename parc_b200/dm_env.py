"""DeepMimicEnv / RefCharEnv -- drop-in mirror of the reference's reference-motion env for the B200 kernels.

Reference: PARC/motion_tracker/envs/ig_parkour/dm_env.py (class DeepMimicEnv) on top of
PARC/motion_tracker/envs/ig_parkour/mgdm_dm_util.py (class RefCharEnv). The env shell (IGParkourEnv) owns the simulator
tensors and the data buffers and INJECTS views of them (get_sim_tensor_views / get_data_buffer_views / get_obs_buf_view,
mgdm_dm_util.py:37-84): column slices of Isaac Gym's [N*actors, 13] root state, stride-2 slices of its [N*dofs, 2] dof state.
This class keeps that contract -- it owns only what DeepMimicEnv owns (dm_env.py:80-92) -- and every method keeps the
reference's name, arguments and in-place semantics; each body is one launch of the C ABI (include/parc_b200_env.h) with the
matching PARC_STEP_* parts mask, reading and writing the injected views through their strides.

Differences, all additive: `config["env"]["dm"]["motion_file"]` may also be an already built parc_b200.MotionLib (tests /
benchmarks without files); reset() takes an optional `draws=` to bypass torch's RNG; build_terrain_square does not build the
PhysX triangle meshes (the physics integrator is outside this path) and returns (None, None).
"""
import copy

import numpy as np
import torch

from . import _lib, file_io, synth
from .kin_char_model import KinCharModel
from .motion_lib import LoopMode, MotionLib
from .terrain import SubTerrain
from .tracking_env import build_config_struct, make_buffers_struct

SIM_CHAR_IDX = 0          # ig_parkour/dm_env.py:17
DONE_NULL, DONE_FAIL, DONE_SUCC, DONE_TIME = 0, 1, 2, 3      # base_env.DoneFlags

_SIM_VIEW_NAMES = ("ref_root_pos", "ref_root_rot", "ref_root_vel", "ref_root_ang_vel", "ref_body_pos", "ref_joint_rot", "ref_dof_pos",
                   "ref_dof_vel", "ref_contacts", "char_root_pos", "char_root_rot", "char_root_vel", "char_root_ang_vel", "char_dof_pos",
                   "char_dof_vel", "char_contact_forces", "char_rigid_body_pos", "char_rigid_body_vel", "char_rigid_body_ang_vel")


class RefCharEnv:
    """mgdm_dm_util.py:13-218."""

    def __init__(self, config, num_envs, device, visualize, char_model: KinCharModel):
        env_config = config["env"]
        dev = torch.device(device)
        assert dev.type == "cuda", "the B200 env runs on a CUDA device only (no CPU fallback)"
        self._num_envs = int(num_envs)
        self._device = dev
        self._visualize = visualize
        self._kin_char_model = char_model
        self._env_config = copy.deepcopy({k: v for k, v in env_config.items() if k != "dm"})
        self._timestep = 1.0 / env_config["control_freq"]
        self.set_rand_root_pos_offset_scale(env_config["rand_root_pos_offset_scale"])
        self._root_pos_offset = None
        self._root_rot_offset = None
        self._root_vel_offset = None
        self._root_ang_vel_offset = None
        self._dof_pos_offset = None
        self._dof_vel_offset = None
        self._max_obs_h = env_config["max_obs_h"]
        self._min_obs_h = env_config["min_obs_h"]
        self._bufs = None          # parc_env_buffers_t of the injected views, rebuilt when they change
        self._cfg = None
        self._obs_buf = None

    # ---- injected views (mgdm_dm_util.py:37-84) ------------------------------------------------------------------
    def get_sim_tensor_views(self, ref_root_pos, ref_root_rot, ref_root_vel, ref_root_ang_vel, ref_body_pos, ref_joint_rot, ref_dof_pos,
                             ref_dof_vel, ref_contacts, char_root_pos, char_root_rot, char_root_vel, char_root_ang_vel, char_dof_pos,
                             char_dof_vel, char_contact_forces, char_rigid_body_pos, char_rigid_body_vel, char_rigid_body_ang_vel):
        loc = locals()
        for name in _SIM_VIEW_NAMES:
            setattr(self, "_" + name, loc[name])
        self._bufs = None

    def get_data_buffer_views(self, reward_buf, done_buf, time_buf, timestep_buf, actors_need_reset, env_offsets, key_body_ids, ray_xy_points,
                              ray_hfs):
        self._reward_buf = reward_buf
        self._done_buf = done_buf
        self._time_buf = time_buf
        self._timestep_buf = timestep_buf
        self._actors_need_reset = actors_need_reset
        self._env_offsets = env_offsets
        self._key_body_ids = key_body_ids
        self._ray_xy_points = ray_xy_points
        self._ray_hfs = ray_hfs
        self._bufs = None
        self._cfg = None

    def get_obs_buf_view(self, obs_buf):
        self._obs_buf = obs_buf
        self._bufs = None

    def _has_key_bodies(self):
        return len(self._key_body_ids) > 0

    # ---- setters (mgdm_dm_util.py:174-218) -------------------------------------------------------------------------
    def set_rand_root_pos_offset_scale(self, val):
        self._rand_root_pos_offset_scale = val
        if getattr(self, "_cfg", None) is not None:
            self._cfg.rand_root_pos_offset_scale = val

    def _check_offset(self, val, width):
        if isinstance(val, torch.Tensor):
            assert val.shape[0] == self._num_envs
            assert val.shape[1] == width
            val = val.to(device=self._device, dtype=torch.float32).contiguous()
        return val

    def set_root_pos_offset(self, val=None):
        self._root_pos_offset = self._check_offset(val, 3)

    def set_root_rot_offset(self, val=None):
        self._root_rot_offset = self._check_offset(val, 4)

    def set_root_vel_offset(self, val=None):
        self._root_vel_offset = self._check_offset(val, 3)

    def set_root_ang_vel_offset(self, val=None):
        self._root_ang_vel_offset = self._check_offset(val, 3)

    def set_dof_pos_offset(self, val=None):
        self._dof_pos_offset = self._check_offset(val, self._kin_char_model.get_dof_size())

    def set_dof_vel_offset(self, val=None):
        self._dof_vel_offset = self._check_offset(val, self._kin_char_model.get_dof_size())

    # ---- heights (mgdm_dm_util.py:128-145) ---------------------------------------------------------------------------
    def _refresh_ray_obs_hfs(self, char_root_pos_xyz, char_heading):
        """ray_hfs[...] = clamp(hf[cell(rotate(ray_xy, heading) + root_xy)] - root_z, min_obs_h, max_obs_h), in place."""
        N, P = self._num_envs, int(self._ray_xy_points.shape[0])
        assert char_root_pos_xyz.shape == (N, 3) and char_heading.shape == (N,)
        assert self._ray_hfs.is_contiguous() or self._ray_hfs.stride(1) == 1
        f32 = torch.float32
        _lib.check(_lib.lib().parc_b200_ray_heights(
            self._terrain.cstruct(), _lib.cptr(char_root_pos_xyz.contiguous(), f32), _lib.cptr(char_heading.contiguous(), f32),
            _lib.cptr(self._ray_xy_points.contiguous(), f32), P, N, float(self._min_obs_h), float(self._max_obs_h), _lib.ptr(self._ray_hfs),
            int(self._ray_hfs.stride(0)), None, _lib.stream_ptr(self._device)))


class DeepMimicEnv(RefCharEnv):
    """ig_parkour/dm_env.py:19-760 (the methods on the hot path; GUI / recording-history helpers are not mirrored)."""

    def __init__(self, config, num_envs, device, visualize, char_model):
        env_config = config["env"]
        dm_config = env_config["dm"]
        self._random_reset_pos = dm_config.get("random_reset_pos", False)
        assert not self._random_reset_pos, "dm.random_reset_pos is a single-env debugging aid of the reference; not built here"
        super().__init__(config, num_envs, device, visualize, char_model)
        dev = self._device
        self._min_motion_weight = dm_config.get("min_motion_weight", 0.01)
        self._demo_mode = env_config["demo_mode"]
        self._rand_reset = env_config["rand_reset"]
        self._ignore_fail_rates = dm_config.get("ignore_fail_rates", False)
        self._terrains_per_motion = dm_config["terrains_per_motion"]
        self._fail_rate_quantiles = torch.tensor(dm_config.get("fail_rate_quantiles", [0.25, 0.5, 0.75]), dtype=torch.float32, device=dev)
        self._one_motion_mode = False
        self._selected_motion_id = 0
        self._terrain_build_mode = dm_config.get("terrain_build_mode", "square")
        motion_file = dm_config["motion_file"]
        if isinstance(motion_file, MotionLib):
            self._mlib = motion_file
        else:
            self._mlib = MotionLib.from_file(motion_file=motion_file, char_model=char_model, device=dev, contact_info=True)
        # the reference's IGParkourEnv / DMMotionRecorder read `_motion_lib` (ig_parkour_env.py:362, dm_motion_recorder.py:100);
        # the attribute the reference actually sets is `_mlib` (dm_env.py:73): both names are offered
        self._motion_lib = self._mlib
        M = self._mlib.num_motions()
        N = self._num_envs
        self._motion_ids = torch.zeros(size=(N,), device=dev, dtype=torch.int64)
        self._motion_terrain_ids = torch.zeros(size=(N,), device=dev, dtype=torch.int64)
        self._motion_time_offsets = torch.zeros(size=(N,), device=dev, dtype=torch.float32)
        if "fail_rates_path" in dm_config:
            self._motion_id_fail_rates = torch.load(dm_config["fail_rates_path"]).to(dtype=torch.float32, device=dev)
        else:
            self._motion_id_fail_rates = torch.ones(size=[M], dtype=torch.float32, device=dev)
        self._ema_weight = 0.01
        self._motion_start_time_fraction = torch.zeros(size=[N], dtype=torch.float32, device=dev)
        # device-side scratch of the step / EMA kernels (not part of the reference's state)
        self._done_count = torch.zeros(1, dtype=torch.int32, device=dev)
        self._done_list = torch.zeros(N, dtype=torch.int32, device=dev)
        self._done_code = torch.zeros(N, dtype=torch.uint8, device=dev)
        self._cdf = torch.zeros(M, dtype=torch.float32, device=dev)
        self._ema_scratch = torch.zeros(2 * M + N, dtype=torch.int32, device=dev)
        self._terrain = None
        self._dm_motion_offsets = None

    # ---- terrain (dm_env.py:95-316, 447-463) ----------------------------------------------------------------------------
    def build_terrain(self, env_config, terrain_save_path, x_offset=0.0, y_offset=0.0):
        if self._terrain_build_mode == "square":
            return self.build_terrain_square(env_config, terrain_save_path, x_offset=x_offset, y_offset=y_offset)
        elif self._terrain_build_mode == "file":
            return self.load_motion_terrain_file(env_config, terrain_save_path)
        else:
            assert False, "unsupported terrain build mode"       # "wide" is an unused variant upstream (dm_env.py:318-445)

    def load_motion_terrain_file(self, env_config, terrain_save_path):
        """dm_env.py:105-155: the terrain already loaded with the (single) clip."""
        self._terrain = self._mlib._terrains[0]
        self._terrains_per_motion = 1
        self._dm_motion_offsets = torch.zeros(size=[self._num_envs, self._terrains_per_motion, 2], dtype=torch.float32, device=self._device)
        self._bufs = None
        return None, None

    def build_terrain_square(self, env_config, terrain_save_path, x_offset=0.0, y_offset=0.0):
        """dm_env.py:157-316: ceil(sqrt(M))^2 tiles of (max clip dims + 2 pad) cells; every clip's terrain is padded IN PLACE with
        its own minimum height (as upstream: MotionLib._terrains changes), dm_motion_offsets[m, 0] = tile origin - padded clip
        min_point. Pinned bit for bit to the reference by tests/test_golden_setup_inputs.py (through synth.build_terrain_square)."""
        hmap_config = env_config["dm"]["heightmap"]
        dx = float(hmap_config["horizontal_scale"])
        num_padding_cells = synth.num_padding_cells(hmap_config["padding"], dx)
        assert self._terrains_per_motion == 1
        terrains = self._mlib._terrains
        clips = [dict(hf=t.hf.detach().cpu(), min_point=t.min_point.detach().cpu(), dx=dx) for t in terrains]
        hf, min_point, _, offs = synth.build_terrain_square(clips, padding_cells=num_padding_cells, x_offset=x_offset, y_offset=y_offset)
        for t in terrains:
            t.pad(num_padding_cells, torch.min(t.hf).item())
        self._dm_motion_offsets = offs.to(self._device)
        self._terrain = SubTerrain("heightmap", int(hf.shape[0]), int(hf.shape[1]), dx, dx, float(min_point[0]), float(min_point[1]),
                                   device=self._device)
        self._terrain.hf = hf.to(self._device).contiguous()
        self._bufs = None
        if terrain_save_path is not None:
            # the reference's terrain file (dm_env.py:299-312) without the PhysX meshes
            file_io.save_ms_file(file_io.MSFileData(motion_data=None, terrain_data=self._terrain.to_ms_terrain_data(),
                                                    misc_data={"terrains_per_motion": self._terrains_per_motion,
                                                               "motion_offsets": self._dm_motion_offsets.cpu().numpy(),
                                                               "all_terrain_verts": [], "all_terrain_tris": []}), terrain_save_path)
        return None, None

    def load_terrain(self, terrain_save_path):
        """dm_env.py:447-463."""
        data = file_io.load_ms_file(terrain_save_path, device=self._device)
        self._terrain = SubTerrain.from_ms_terrain_data(terrain_data=data.terrain_data, device=self._device)
        misc = data.misc_data
        self._terrains_per_motion = misc["terrains_per_motion"]
        offs = misc["motion_offsets"]
        if isinstance(offs, np.ndarray):
            offs = torch.from_numpy(offs)
        self._dm_motion_offsets = offs.to(device=self._device)
        self._bufs = None
        return misc.get("all_terrain_verts"), misc.get("all_terrain_tris")

    # ---- descriptors ------------------------------------------------------------------------------------------------------
    def _config_struct(self):
        if self._cfg is None:
            self._cfg = build_config_struct(self._full_env_config(), self._kin_char_model, int(self._ray_xy_points.shape[0]), self._ray_xy_points)
            kb = [int(k) for k in self._key_body_ids.tolist()]
            assert len(kb) <= _lib.MAX_KEY_BODIES
            self._cfg.num_key_bodies = len(kb)
            for i, k in enumerate(kb):
                self._cfg.key_body_ids[i] = k
            self._cfg.rand_root_pos_offset_scale = self._rand_root_pos_offset_scale
            self._cfg.min_motion_weight = self._min_motion_weight
        return self._cfg

    def _full_env_config(self):
        """The env shell's yaml keys this class was not handed explicitly default to dm_env_default.yaml's values; the ones that
        matter for a call are overwritten from the call's arguments (update_done, compute_tar_obs)."""
        from .tracking_env import default_env_config
        ec = default_env_config()
        ec.update(self._env_config)
        ec["key_bodies"] = []                                   # key bodies come as ids through get_data_buffer_views
        ec["dm"] = {"min_motion_weight": self._min_motion_weight, "terrains_per_motion": self._terrains_per_motion}
        J = self._kin_char_model.get_num_joints()
        for k in ("pose_termination_dist", "joint_err_w"):
            if ec.get(k) is not None and len(ec[k]) != J - 1:
                ec[k] = [1.0] * (J - 1)
        ec["contact_weights"] = (list(ec.get("contact_weights", [])) + [0.0] * J)[:J]
        return ec

    def _buffers_struct(self):
        if self._bufs is None:
            assert self._terrain is not None and self._dm_motion_offsets is not None, "build_terrain() / load_terrain() first"
            dummy_obs = self._obs_buf if self._obs_buf is not None else torch.zeros(1, 1, dtype=torch.float32, device=self._device)
            self._dummy_obs = dummy_obs
            anr = self._actors_need_reset
            self._bufs = make_buffers_struct(
                self._num_envs, obs_buf=dummy_obs, reward_buf=self._reward_buf, done_buf=self._done_buf, time_buf=self._time_buf,
                timestep_buf=self._timestep_buf, env_offsets=self._env_offsets.contiguous(), motion_ids=self._motion_ids,
                motion_terrain_ids=self._motion_terrain_ids, motion_time_offsets=self._motion_time_offsets,
                dm_motion_offsets=self._dm_motion_offsets.contiguous(), terrains_per_motion=self._terrains_per_motion,
                actors_need_reset=anr.view(torch.uint8) if anr.dtype == torch.bool else anr, ray_xy_points=self._ray_xy_points.contiguous(),
                ray_hfs=self._ray_hfs, motion_id_fail_rates=self._motion_id_fail_rates, motion_weights=self._mlib._motion_weights,
                done_count=self._done_count, done_list=self._done_list, done_code=self._done_code,
                **{n: getattr(self, "_" + n) for n in _SIM_VIEW_NAMES})
            self._terrain_cs = self._terrain.cstruct()
        return self._bufs

    def _step(self, parts, env_ids=None):
        km = self._kin_char_model._cstruct()
        n = 0 if env_ids is None else int(env_ids.shape[0])
        bufs = self._buffers_struct()
        _lib.check(_lib.lib().parc_b200_env_step(km, self._mlib.cstruct(), self._terrain_cs, self._config_struct(), bufs, parts,
                                                 _lib.ptr(env_ids) if env_ids is not None else None, n, _lib.stream_ptr(self._device)))

    # ---- per-step methods ---------------------------------------------------------------------------------------------------
    def _refresh_obs_hfs(self, char_root_pos_xyz, char_heading):
        """dm_env.py:465-467."""
        self._refresh_ray_obs_hfs(char_root_pos_xyz, char_heading)

    def _get_motion_times(self, env_ids=None):
        """dm_env.py:547-552."""
        if env_ids is None:
            return self._time_buf + self._motion_time_offsets
        return self._time_buf[env_ids] + self._motion_time_offsets[env_ids]

    def _move_to_motion_terrain(self, pos, env_ids=None):
        """dm_env.py:554-565 (torch; for callers outside the step -- the kernels apply the same offset themselves)."""
        if env_ids is None:
            offset = self._dm_motion_offsets[self._motion_ids, self._motion_terrain_ids] - self._env_offsets[..., 0:2]
        else:
            offset = self._dm_motion_offsets[self._motion_ids[env_ids], self._motion_terrain_ids[env_ids]] - self._env_offsets[env_ids, 0:2]
        assert pos.shape[0] == offset.shape[0], str(pos.shape[0]) + " " + str(offset.shape[0])
        assert pos.shape[-1] == offset.shape[-1]
        for _ in range(len(pos.shape) - len(offset.shape)):
            offset = offset.unsqueeze(1)
        return pos + offset

    def _update_ref_motion(self):
        """dm_env.py:523-545: calc_motion_frame at time_buf + motion_time_offsets, terrain offset, the nine ref_* views in place
        (ref_body_pos by forward kinematics, ref_dof_pos by rot_to_dof) -- parc_b200_env_step(PARC_STEP_REF)."""
        self._step(_lib.STEP_REF)

    def compute_tar_obs(self, tar_obs_steps, env_ids=None):
        """dm_env.py:594-626 + fetch_tar_obs_data (mgdm_dm_util.py:221-244) -> (tar_root_pos [n,S,3], tar_root_rot [n,S,4],
        tar_joint_rot [n,S,J-1,4], tar_key_pos [n,S,K,3], tar_contacts [n,S,C]); freshly allocated, owned by the caller."""
        S = int(tar_obs_steps.shape[0])
        assert S > 0
        cfg = self._config_struct()
        # `timestep * tar_obs_steps` with an int tensor -> fp32 (mgdm_dm_util.py:232)
        tdt = (self._timestep * tar_obs_steps.to(device="cpu")).to(torch.float32)
        cfg.num_tar_steps = S
        for i, v in enumerate(tdt.tolist()):
            cfg.tar_dt[i] = v
        n = self._num_envs if env_ids is None else int(env_ids.shape[0])
        J, K = self._kin_char_model.get_num_joints(), len(self._key_body_ids)
        C = self._mlib.cstruct().num_contacts
        dev, f32 = self._device, torch.float32
        out = (torch.empty(n, S, 3, device=dev, dtype=f32), torch.empty(n, S, 4, device=dev, dtype=f32),
               torch.empty(n, S, J - 1, 4, device=dev, dtype=f32),
               torch.empty(n, S, K, 3, device=dev, dtype=f32) if K > 0 else torch.zeros([0], device=dev),
               torch.empty(n, S, C, device=dev, dtype=f32))
        ids = None if env_ids is None else env_ids.to(device=dev, dtype=torch.int64).contiguous()
        _lib.check(_lib.lib().parc_b200_compute_tar_obs_data(
            self._kin_char_model._cstruct(), self._mlib.cstruct(), cfg, self._buffers_struct(), _lib.ptr(ids), n, _lib.ptr(out[0]), _lib.ptr(out[1]),
            _lib.ptr(out[2]), _lib.ptr(out[3]) if K > 0 else None, _lib.ptr(out[4]), _lib.stream_ptr(dev)))
        return out

    def update_done(self, termination_height, episode_length, contact_body_ids, pose_termination, pose_termination_dist, global_obs,
                    enable_early_termination, track_root, root_pos_termination_dist, root_rot_termination_angle):
        """RefCharEnv.update_done (mgdm_dm_util.py:147-172, compute_done :334-402) + the DeepMimicEnv tail (dm_env.py:628-665):
        done flags in place, fail-rate EMA over the finished envs in env-id order, motion end -> FAIL."""
        cfg = self._config_struct()
        f = lambda x: float(x.item()) if isinstance(x, torch.Tensor) else float(x)
        cfg.termination_height = float(np.float32(f(termination_height)))
        cfg.episode_length = float(np.float32(f(episode_length)))
        cb = [int(b) for b in contact_body_ids.tolist()] if isinstance(contact_body_ids, torch.Tensor) else [int(b) for b in contact_body_ids]
        cfg.num_fall_contact_bodies = len(cb)
        mask = 0
        for b in cb:
            mask |= 1 << b
        cfg.fall_contact_body_mask = mask
        cfg.pose_termination = int(bool(pose_termination))
        ptd = pose_termination_dist.tolist() if isinstance(pose_termination_dist, torch.Tensor) else list(pose_termination_dist)
        for i, v in enumerate(ptd):
            cfg.pose_termination_dist[i] = v
        cfg.global_obs = int(bool(global_obs))
        cfg.enable_early_termination = int(bool(enable_early_termination))
        cfg.track_root = int(bool(track_root))
        d = float(root_pos_termination_dist)
        cfg.root_pos_termination_dist = float(np.float32(d * d))      # compute_done compares with dist * dist (double -> fp32)
        cfg.root_rot_termination_angle = float(np.float32(f(root_rot_termination_angle)))
        cfg.never_done = 0
        self._done_count.zero_()
        self._step(_lib.STEP_DONE)
        _lib.check(_lib.lib().parc_b200_fail_rate_ema(cfg, self._buffers_struct(), self._mlib.num_motions(), _lib.ptr(self._cdf),
                                                      _lib.ptr(self._ema_scratch), _lib.stream_ptr(self._device)))

    # ---- reset (dm_env.py:473-521, 567-592; mgdm_dm_util.py:89-126) -----------------------------------------------------------
    def _draw_reset(self, env_ids):
        """The random draws in the reference's call order: sample_motions (multinomial), randint, sample_time (rand), xy noise (rand)."""
        n = len(env_ids)
        mlib, dev = self._mlib, self._device
        if self._demo_mode:
            motion_ids = env_ids % mlib.num_motions()
        elif self._one_motion_mode:
            motion_ids = torch.ones_like(env_ids)
            motion_ids[...] = self._selected_motion_id
        elif self._ignore_fail_rates:
            motion_ids = mlib.sample_motions(n)
        else:
            weights_modifier = torch.clamp(self._motion_id_fail_rates, min=self._min_motion_weight)
            motion_ids = mlib.sample_motions(n, weights_modifier * mlib._motion_weights)
        motion_terrain_ids = torch.randint(high=self._terrains_per_motion, size=(n,), dtype=torch.int64, device=dev)
        if self._rand_reset:
            motion_times = mlib.sample_time(motion_ids)
        else:
            motion_times = mlib.get_motion_length(motion_ids) * self._motion_start_time_fraction[env_ids]
        noise = torch.rand(size=[n, 2], device=dev, dtype=torch.float32) * 2.0 - 1.0
        return motion_ids, motion_terrain_ids, motion_times, noise

    def reset(self, env_ids, draws=None):
        if len(env_ids) == 0:
            return
        dev, i64, f32 = self._device, torch.int64, torch.float32
        env_ids = env_ids.to(device=dev, dtype=i64).contiguous()
        n = int(env_ids.shape[0])
        mids, tids, times, noise = self._draw_reset(env_ids) if draws is None else draws
        mids, tids = mids.to(device=dev, dtype=i64).contiguous(), tids.to(device=dev, dtype=i64).contiguous()
        times, noise = times.to(device=dev, dtype=f32).contiguous(), noise.to(device=dev, dtype=f32).contiguous()
        p = _lib.ptr
        _lib.check(_lib.lib().parc_b200_env_reset(self._kin_char_model._cstruct(), self._mlib.cstruct(), self._config_struct(), self._buffers_struct(),
                                                  p(env_ids), n, p(mids), p(tids), p(times), p(noise), p(self._root_pos_offset), p(self._root_rot_offset),
                                                  _lib.stream_ptr(dev)))
        # the remaining offsets of apply_offsets_to_char_state (mgdm_dm_util.py:118-126): whole-tensor adds, as upstream writes them
        if self._root_vel_offset is not None:
            self._char_root_vel[env_ids] += self._root_vel_offset
        if self._root_ang_vel_offset is not None:
            self._char_root_ang_vel[env_ids] += self._root_ang_vel_offset
        if self._dof_pos_offset is not None:
            self._char_dof_pos[env_ids] += self._dof_pos_offset
        if self._dof_vel_offset is not None:
            self._char_dof_vel[env_ids] += self._dof_vel_offset

    # ---- small accessors (dm_env.py:737-760) ------------------------------------------------------------------------------------
    def set_demo_mode(self, val=None):
        if val is None:
            val = not self._demo_mode
        self._demo_mode = val
        return self._demo_mode

    def set_motion_start_time_fraction(self, val: torch.Tensor):
        self._motion_start_time_fraction = val.to(self._device)

    def get_env_motion_length(self, env_ids):
        return self._mlib.get_motion_length(self._motion_ids[env_ids])

    def get_env_motion_time(self, env_ids):
        return self._get_motion_times(env_ids)

    def get_env_motion_name(self, env_id):
        return self._mlib.get_motion_names()[self._motion_ids[env_id].item()]

    def get_extra_log_info(self):
        """dm_env.py:667-722 without the console table: fail rates by motion name and their quantiles."""
        names = self._mlib.get_motion_names()
        rates = self._motion_id_fail_rates
        top, _ = torch.sort(rates, descending=True)
        q = torch.quantile(top, self._fail_rate_quantiles)
        info = {"MOTION_FAIL_RATES": {names[i]: rates[i].item() * 100.0 for i in range(len(names))},
                "Misc": {"top fail rate": top[0].item() * 100.0}}
        for i in range(self._fail_rate_quantiles.shape[0]):
            info["Misc"]["Fail Rate at " + str(round(self._fail_rate_quantiles[i].item() * 100.0)) + "% Quantile"] = q[i].item() * 100.0
        return info


DeepMimicEnvB200 = DeepMimicEnv
