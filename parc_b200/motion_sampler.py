"""MDM dataset sampler on the GPU -- host-side mirror of the reference's
`MotionSampler` (PARC/motion_generator/motion_sampler.py:9-54) and
`MDMHeightfieldContactMotionSampler` (PARC/motion_generator/mdm_heightfield_contact_motion_sampler.py:24-302):
same constructor config keys, same `sample_motion_data` signature and return structure, with the body
(`_extract_motion_data`, `get_hfs_from_data`, canonicalisation, forward kinematics, `_sample_target_info`)
done by `parc_b200_mdm_sample_motion_data` (include/parc_b200_sampler.h) in two launches instead of a
Python loop over samples. SURVEY.md section 8(f) rank 1.

Not covered (raises NotImplementedError): the heightfield augmenter (`use_hf_augmentation`, :304-330 -- Python
`random` + per-sample loops around maxpool/box edits) and the FLOOR_HEIGHTS / JOINT_VEL frame components.
There is no CPU fallback.
"""
import enum

import torch

from . import _lib
from .kin_char_model import KinCharModel
from .motion_lib import LoopMode, MotionLib


class MDMFrameType(enum.Enum):           # diffusion_util.py:13-20
    ROOT_POS = 0
    ROOT_ROT = 1
    JOINT_POS = 2
    JOINT_VEL = 3
    JOINT_ROT = 4
    CONTACTS = 5
    FLOOR_HEIGHTS = 6


class RelativeZStyle(enum.Enum):         # diffusion_util.py:9-11
    RELATIVE_TO_ROOT = 0
    RELATIVE_TO_ROOT_FLOOR = 1


class HFAugmentationMode(enum.Enum):     # mdm_heightfield_contact_motion_sampler.py:19-22
    NOISE = 0
    MAXPOOL_AND_BOXES = 1
    NONE = 2


class TargetInfo:                        # diffusion_util.py:40-44
    def __init__(self, future_pos: torch.Tensor, future_rot: torch.Tensor):
        self.future_pos = future_pos
        self.future_rot = future_rot


def get_xy_grid_points(center, dx, dy, num_x_neg, num_x_pos, num_y_neg, num_y_pos):
    """geom_util.get_xy_grid_points (PARC/util/geom_util.py:212-223)."""
    device = center.device
    x_points = torch.linspace(center[0] - dx * num_x_neg, center[0] + dx * num_x_pos, num_x_neg + num_x_pos + 1, device=device)
    y_points = torch.linspace(center[1] - dy * num_y_neg, center[1] + dy * num_y_pos, num_y_neg + num_y_pos + 1, device=device)
    x, y = torch.meshgrid(x_points, y_points, indexing="ij")
    return torch.stack([x, y], dim=-1)


class MotionSampler:
    """motion_sampler.py:9-54"""

    def __init__(self, cfg):
        self._device = cfg["device"]
        self._seq_len = None
        self._fps = None
        self._kin_char_model = KinCharModel(self._device)
        self._kin_char_model.load_char_file(cfg["char_file"])
        self._motion_lib_file = cfg["motion_lib_file"]
        if isinstance(self._motion_lib_file, str):
            self._mlib = MotionLib.from_file(motion_file=self._motion_lib_file, char_model=self._kin_char_model, device=self._device,
                                             contact_info=True)
        elif isinstance(self._motion_lib_file, MotionLib):
            self._mlib = self._motion_lib_file
        else:
            assert False
        print("num_motions =", self._mlib.num_motions())

    def check_init(self):
        assert self._seq_len is not None
        assert self._fps is not None

    def get_seq_len(self):
        return self._seq_len

    def _sample_motion_start_times(self, motion_ids, seq_duration):
        """:39-54 -- torch draws in the reference's order (two rand_like)."""
        motion_full_lengths = self._mlib.get_motion_length(motion_ids)
        if self._random_start_times:
            loop = torch.rand_like(motion_full_lengths) * motion_full_lengths
            clamp = torch.rand_like(motion_full_lengths) * (motion_full_lengths - seq_duration)
            modes = self._mlib.get_motion_loop_mode(motion_ids)
            return torch.where(modes == LoopMode.WRAP.value, loop, clamp)
        return torch.zeros_like(motion_full_lengths)


class MDMHeightfieldContactMotionSampler(MotionSampler):
    def __init__(self, cfg):
        super().__init__(cfg)
        self._canonicalize_samples = cfg["features"]["canonicalize_samples"]
        self._frame_components = [MDMFrameType[c] for c in cfg["features"]["frame_components"]]
        for c in self._frame_components:
            if c in (MDMFrameType.FLOOR_HEIGHTS, MDMFrameType.JOINT_VEL):
                raise NotImplementedError("frame component %s is not produced by the CUDA sampler" % c.name)
        self._num_rb = self._kin_char_model.get_num_joints()
        self._fps = cfg["sequence_fps"]
        self._timestep = 1.0 / self._fps
        self._random_start_times = cfg["autoregressive"]
        self._sample_seq_time = cfg["sequence_duration"]
        # built on the CPU and moved: torch.arange / linspace differ in the last bit between CPU and CUDA, and frame
        # indices / grid cells are compared bit for bit with the reference's CPU run
        self._motion_times = torch.arange(start=0.0, end=self._sample_seq_time, step=self._timestep, dtype=torch.float32).to(self._device)
        self._seq_len = self._motion_times.shape[0]
        self._num_prev_states = cfg["num_prev_states"]
        assert self._num_prev_states >= 1
        self._canon_idx = self._num_prev_states - 1
        self._ref_frame_idx = self._num_prev_states - 1
        self._future_pos_noise_scale = cfg["future_pos_noise_scale"]
        self._relative_z_style = RelativeZStyle[cfg["relative_z_style"]]
        self._hf_augmentation_mode = HFAugmentationMode[cfg["hf_augmentation_mode"]]
        hmap = cfg["heightmap"]
        dx = hmap["horizontal_scale"]
        self._dx = dx
        g = hmap["local_grid"]
        self._num_x_neg, self._num_x_pos, self._num_y_neg, self._num_y_pos = g["num_x_neg"], g["num_x_pos"], g["num_y_neg"], g["num_y_pos"]
        self._grid_dim_x = self._num_x_neg + 1 + self._num_x_pos
        self._grid_dim_y = self._num_y_neg + 1 + self._num_y_pos
        self._num_hf_points = self._grid_dim_x * self._grid_dim_y
        zero = torch.zeros(size=(2,), dtype=torch.float32)
        self._generic_heightmap = get_xy_grid_points(zero, dx, dx, self._num_x_neg, self._num_x_pos, self._num_y_neg,
                                                     self._num_y_pos).contiguous().to(self._device)
        self._max_h = hmap["max_h"]
        self._min_h = -self._max_h
        self._use_hf_augmentation = cfg["use_hf_augmentation"]
        if self._use_hf_augmentation and self._hf_augmentation_mode != HFAugmentationMode.NONE:
            raise NotImplementedError("heightfield augmentation (mode %s) is outside the CUDA sampler path" % self._hf_augmentation_mode.name)
        self._future_window_min = cfg["future_window_min"]
        self._future_window_max = cfg["future_window_max"]
        self.check_init()
        self._cfg = _lib.SamplerConfigStruct(self._seq_len, self._ref_frame_idx, self._relative_z_style.value, self._grid_dim_x,
                                        self._grid_dim_y, self._num_x_neg, self._num_y_neg, self._min_h, self._max_h,
                                        self._future_pos_noise_scale)
        self._terrain_table = None

    # ---------------------------------------------------------------- per-motion terrain table on the device
    def _terrains_device(self):
        """[M] parc_terrain_t records (pointer, dims, origin, cell size) of MotionLib._terrains, as device bytes."""
        if self._terrain_table is None:
            ts = self._mlib._terrains
            assert all(t is not None for t in ts), "every motion needs its terrain for the local heightmaps"
            arr = (_lib.TerrainStruct * len(ts))(*[t.cstruct() for t in ts])
            host = torch.frombuffer(bytearray(bytes(arr)), dtype=torch.uint8)
            self._terrain_table = (host.to(self._device), [t.hf for t in ts])      # keep the heightfields alive
        return self._terrain_table[0]

    def _sample_motion_future_times(self, motion_ids, start_times, future_window_min, future_window_max):
        """:133-144"""
        motion_full_lengths = self._mlib.get_motion_length(motion_ids)
        remaining_time = torch.clamp_max(motion_full_lengths - start_times, future_window_max - future_window_min)
        return torch.rand_like(start_times) * remaining_time + start_times + future_window_min

    def sample_motion_data(self, motion_ids=None, num_samples=None, motion_start_times=None, ret_hf_obs=True, ret_target_info=True,
                           future_times=None, future_pos_noise=None):
        """:197-302. `future_times` [n] / `future_pos_noise` [n,3] (standard normal) replace the draws of
        _sample_target_info (:146-160) when given (parity tests); otherwise they are drawn from torch's generator
        on the device in the reference's order."""
        if motion_ids is None:
            assert num_samples is not None
            motion_ids = self._mlib.sample_motions(num_samples)
        else:
            num_samples = motion_ids.shape[0]
        if motion_start_times is None:
            motion_start_times = self._sample_motion_start_times(motion_ids, self._sample_seq_time)
        dev, f32 = self._device, torch.float32
        motion_ids = motion_ids.to(device=dev, dtype=torch.int64).contiguous()
        motion_start_times = motion_start_times.to(device=dev, dtype=f32).contiguous()
        n, T, J = num_samples, self._seq_len, self._kin_char_model.get_num_joints()
        Cn = self._mlib._frame_contacts.shape[-1]
        if ret_target_info:
            if future_times is None:
                future_times = self._sample_motion_future_times(motion_ids, motion_start_times, self._future_window_min,
                                                                self._future_window_max)
            if future_pos_noise is None:
                future_pos_noise = torch.randn(n, 3, device=dev, dtype=f32)
            future_times = future_times.to(device=dev, dtype=f32).contiguous()
            future_pos_noise = future_pos_noise.to(device=dev, dtype=f32).contiguous()
            assert future_times.shape == (n,) and future_pos_noise.shape == (n, 3)
        e = lambda *s: torch.empty(s, dtype=f32, device=dev)
        want = set(self._frame_components)
        root_pos, root_rot = e(n, T, 3), e(n, T, 4)
        joint_pos = e(n, T, J - 1, 3) if MDMFrameType.JOINT_POS in want else None
        joint_rot = e(n, T, J - 1, 4) if MDMFrameType.JOINT_ROT in want else None
        contacts = e(n, T, Cn) if MDMFrameType.CONTACTS in want else None
        hfs = e(n, self._grid_dim_x, self._grid_dim_y) if ret_hf_obs else None
        fpos, frot = (e(n, 3), e(n, 4)) if ret_target_info else (None, None)
        canon = e(n, 12)
        p = _lib.ptr
        _lib.check(_lib.lib().parc_b200_mdm_sample_motion_data(
            self._kin_char_model._cstruct(), self._mlib.cstruct(), p(self._terrains_device()) if ret_hf_obs else None, self._cfg,
            p(motion_ids), p(motion_start_times), n, p(self._motion_times), p(self._generic_heightmap),
            p(future_times) if ret_target_info else None, p(future_pos_noise) if ret_target_info else None,
            p(root_pos), p(root_rot), p(joint_pos), p(joint_rot), p(contacts), p(hfs), p(fpos), p(frot), p(canon),
            _lib.stream_ptr(dev)))
        self._last_canon = canon
        motion_ret = dict()
        for frame_type in self._frame_components:
            if frame_type == MDMFrameType.ROOT_POS:
                motion_ret[MDMFrameType.ROOT_POS] = root_pos
            if frame_type == MDMFrameType.ROOT_ROT:
                motion_ret[MDMFrameType.ROOT_ROT] = root_rot
            if frame_type == MDMFrameType.JOINT_POS:
                motion_ret[MDMFrameType.JOINT_POS] = joint_pos
            if frame_type == MDMFrameType.JOINT_ROT:
                motion_ret[MDMFrameType.JOINT_ROT] = joint_rot
            if frame_type == MDMFrameType.CONTACTS:
                motion_ret[MDMFrameType.CONTACTS] = contacts
        if ret_hf_obs or ret_target_info:
            ret = [motion_ret]
            if ret_hf_obs:
                ret.append(hfs)
            if ret_target_info:
                ret.append(TargetInfo(fpos, frot))
            return tuple(ret)
        return motion_ret

    def get_motion_sequences_for_id(self, motion_id: int):
        """:109-131 -- every admissible window of one motion."""
        num_frames = self._mlib._motion_num_frames[motion_id].item()
        starts = torch.arange(start=0, end=num_frames - self._seq_len, dtype=torch.float32, device=self._device) * (1.0 / self._fps)
        assert len(starts) > 0, "motion_start_times tensor is empty"
        ids = torch.full_like(starts, motion_id, dtype=torch.int64)
        return self.sample_motion_data(ids, motion_start_times=starts, ret_hf_obs=False, ret_target_info=False)
