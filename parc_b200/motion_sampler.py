"""MDM dataset sampler on the GPU -- host-side mirror of the reference's
`MotionSampler` (PARC/motion_generator/motion_sampler.py:9-54) and
`MDMHeightfieldContactMotionSampler` (PARC/motion_generator/mdm_heightfield_contact_motion_sampler.py:24-302):
same constructor config keys, same `sample_motion_data` signature and return structure, with the body
(`_extract_motion_data`, `get_hfs_from_data`, canonicalisation, forward kinematics, `_sample_target_info`)
done by `parc_b200_mdm_sample_motion_data` (include/parc_b200_sampler.h) in two launches instead of a
Python loop over samples. SURVEY.md section 8(f) rank 1.

The heightfield augmenter (`use_hf_augmentation`, :304-364, :390-400) is a third launch, `parc_b200_mdm_augment_hfs`: the
host draws what the reference draws (Python `random` for the pool / box schedule, torch's CPU generator for the boxes, in the
reference's order, so equal seeds give equal augmentations) and ships them as one record per window. FLOOR_HEIGHTS
(:267-278) is `parc_b200_mdm_floor_heights`. JOINT_VEL is accepted and, as upstream (whose loop :247-278 has no branch for
it), produces no entry. There is no CPU fallback.
"""
import enum
import random

import numpy as np
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
        if self._use_hf_augmentation:                                   # :90-97
            self._max_num_boxes = cfg["max_num_boxes"]
            self._box_min_len = cfg["box_min_len"]
            self._box_max_len = cfg["box_max_len"]
            self._hf_maxpool_chance = cfg["hf_maxpool_chance"]
            self._hf_max_maxpool_size = cfg["hf_max_maxpool_size"]
            self._hf_change_height_chance = cfg["hf_change_height_chance"]
        self._angle_noise_scale = cfg["angle_noise_scale"]
        self._pos_noise_scale = cfg["pos_noise_scale"]
        if "custom_num_contacts" in cfg:
            self._custom_num_contacts = cfg["custom_num_contacts"]
        self._aug_tables = None
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

    # ---------------------------------------------------------------- heightfield augmenter
    def _augmenter_tables(self):
        """What get_hfs_from_data_helper (:332-364) reads per motion, flattened for the kernel: the terrains' hf_maxmin
        concatenated, and MotionLib._hf_mask_inds as one CSR (row = frame) of packed cell indices."""
        if self._aug_tables is None:
            ts, masks = self._mlib._terrains, self._mlib._hf_mask_inds
            M = len(ts)
            assert len(masks) >= M and all(m is not None for m in masks[:M]), \
                "use_hf_augmentation needs hf_mask_inds for every motion (the reference slices MotionLib._hf_mask_inds[id], :349)"
            mm_off, o = [], 0
            for t in ts:
                mm_off.append(o)
                o += t.hf_maxmin.numel()
            maxmin = torch.cat([t.hf_maxmin.to(dtype=torch.float32).reshape(-1) for t in ts]).to(self._device).contiguous()
            frame_off, starts, cells = [0], [0], []
            for m in range(M):
                dim_y = int(ts[m].hf.shape[1])
                for fr in masks[m]:
                    fr = torch.as_tensor(fr).reshape(-1, 2).to(dtype=torch.int64).cpu()
                    cells.append((fr[:, 0] * dim_y + fr[:, 1]).to(torch.int32))
                    starts.append(starts[-1] + int(fr.shape[0]))
                frame_off.append(len(starts) - 1)
            dev = self._device
            self._aug_tables = dict(
                maxmin=maxmin, maxmin_offset=torch.tensor(mm_off, dtype=torch.int64, device=dev),
                frame_off=torch.tensor(frame_off, dtype=torch.int32, device=dev), starts=torch.tensor(starts, dtype=torch.int32, device=dev),
                cells=(torch.cat(cells) if cells and starts[-1] > 0 else torch.zeros(1, dtype=torch.int32)).to(dev),
                max_cells=max(int(t.hf.numel()) for t in ts))
        return self._aug_tables

    def _draw_box_aug_records(self, n):
        """The draws of _box_hf_augmentation (:304-325) and add_boxes_to_hf2 (terrain_util.py:861-914) for n windows, window
        after window: Python `random` for the schedule, torch's CPU generator for the boxes (upstream draws the boxes on
        the sampler's device; here always on the host, so a CPU run of the reference with the same seeds is reproduced).
        -> rec_i [n,8] int32, rec_f [n, 1 + 7 max_num_boxes] fp32 (layout: include/parc_b200_sampler.h)."""
        nbmax = max(int(self._max_num_boxes), 1)
        rec_i = np.full((n, _lib.AUG_REC_INTS), -1, dtype=np.int32)
        rec_f = np.zeros((n, 1 + _lib.AUG_BOX_FLOATS * nbmax), dtype=np.float32)
        shape = torch.tensor([float(self._grid_dim_x), float(self._grid_dim_y)], dtype=torch.float32)
        ones = torch.ones(self._grid_dim_x, self._grid_dim_y)
        span = self._max_h - self._min_h
        for i in range(n):
            change = random.random() < self._hf_change_height_chance
            rec_i[i, 0] = int(change)
            if change:
                rec_f[i, 0] = random.random() * span + self._min_h
            use = [random.random() < self._hf_maxpool_chance, random.random() < self._hf_maxpool_chance,
                   random.random() < self._hf_maxpool_chance]
            order = [0, 1, 2]                      # PARC_POOL_2D, _X, _Y = maxpool_hf, maxpool_hf_1d_x, maxpool_hf_1d_y (:314-316)
            random.shuffle(order)
            for k in range(3):
                rec_i[i, 2 + 2 * k] = 0
                if use[k]:
                    rec_i[i, 1 + 2 * k] = order[k]
                    rec_i[i, 2 + 2 * k] = random.randint(0, self._hf_max_maxpool_size)
            nb = random.randint(0, self._max_num_boxes)
            rec_i[i, 7] = nb
            for b in range(nb):
                center = torch.rand(size=(2,), dtype=torch.float32) * shape
                side = torch.rand(size=(2,), dtype=torch.float32) * (self._box_max_len - self._box_min_len) + self._box_min_len
                angle = (torch.rand(1) * (2.0 * torch.pi - 0.0) + 0.0) * ones          # upstream takes cos / sin over the whole grid
                h = torch.rand(1) * span + self._min_h
                o = 1 + _lib.AUG_BOX_FLOATS * b
                rec_f[i, o:o + 7] = [float(center[0]), float(center[1]), float(side[0]), float(side[1]),
                                     float(torch.cos(angle)[0, 0]), float(torch.sin(angle)[0, 0]), float(h[0])]
        return torch.from_numpy(rec_i), torch.from_numpy(rec_f)

    def _augment_hfs(self, motion_ids, motion_start_times, n, canon, hfs, hf_aug_records, hf_noise):
        mode = self._hf_augmentation_mode
        if mode == HFAugmentationMode.NONE:
            return
        tb = self._augmenter_tables()
        dev, p = self._device, _lib.ptr
        aug = _lib.HfAugStruct()
        aug.mode, aug.max_terrain_cells, aug.timestep = mode.value, tb["max_cells"], self._timestep
        aug.hf_maxmin, aug.maxmin_offset = p(tb["maxmin"]), p(tb["maxmin_offset"])
        aug.mask_frame_off, aug.mask_start, aug.mask_cells = p(tb["frame_off"]), p(tb["starts"]), p(tb["cells"])
        keep = []
        if mode == HFAugmentationMode.NOISE:
            if hf_noise is None:
                hf_noise = torch.rand(n, self._grid_dim_x, self._grid_dim_y, device=dev, dtype=torch.float32)     # rand_like(hf), :328
            hf_noise = hf_noise.to(device=dev, dtype=torch.float32).contiguous()
            assert hf_noise.shape == (n, self._grid_dim_x, self._grid_dim_y)
            aug.noise = p(hf_noise)
            keep.append(hf_noise)
        else:
            rec_i, rec_f = self._draw_box_aug_records(n) if hf_aug_records is None else hf_aug_records
            rec_i = rec_i.to(device=dev, dtype=torch.int32).contiguous()
            rec_f = rec_f.to(device=dev, dtype=torch.float32).contiguous()
            assert rec_i.shape == (n, _lib.AUG_REC_INTS) and rec_f.shape[0] == n and (rec_f.shape[1] - 1) % _lib.AUG_BOX_FLOATS == 0
            aug.max_boxes = (rec_f.shape[1] - 1) // _lib.AUG_BOX_FLOATS
            aug.rec_i, aug.rec_f = p(rec_i), p(rec_f)
            keep += [rec_i, rec_f]
        _lib.check(_lib.lib().parc_b200_mdm_augment_hfs(p(self._terrains_device()), self._cfg, aug, p(motion_ids), p(motion_start_times), n,
                                                       p(self._generic_heightmap), p(canon), p(hfs), _lib.stream_ptr(dev)))
        self._last_aug_inputs = keep                      # alive until the next call (the launch is asynchronous)

    def _sample_motion_future_times(self, motion_ids, start_times, future_window_min, future_window_max):
        """:133-144"""
        motion_full_lengths = self._mlib.get_motion_length(motion_ids)
        remaining_time = torch.clamp_max(motion_full_lengths - start_times, future_window_max - future_window_min)
        return torch.rand_like(start_times) * remaining_time + start_times + future_window_min

    def sample_motion_data(self, motion_ids=None, num_samples=None, motion_start_times=None, ret_hf_obs=True, ret_target_info=True,
                           future_times=None, future_pos_noise=None, hf_aug_records=None, hf_noise=None):
        """:197-302. `future_times` [n] / `future_pos_noise` [n,3] (standard normal) replace the draws of
        _sample_target_info (:146-160) when given (parity tests); otherwise they are drawn from torch's generator
        on the device in the reference's order. Likewise `hf_aug_records` = (rec_i, rec_f) / `hf_noise` [n,gx,gy] replace
        the augmenter's draws (_draw_box_aug_records / rand_like)."""
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
        if ret_hf_obs and self._use_hf_augmentation and self._hf_augmentation_mode == HFAugmentationMode.MAXPOOL_AND_BOXES \
                and hf_aug_records is None:
            hf_aug_records = self._draw_box_aug_records(n)          # get_hfs_from_data (:226) draws before _sample_target_info (:297)
        if ret_hf_obs and self._use_hf_augmentation and self._hf_augmentation_mode == HFAugmentationMode.NOISE and hf_noise is None:
            hf_noise = torch.rand(n, self._grid_dim_x, self._grid_dim_y, device=dev, dtype=f32)
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
        if ret_hf_obs and self._use_hf_augmentation:
            self._augment_hfs(motion_ids, motion_start_times, n, canon, hfs, hf_aug_records, hf_noise)
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
            if frame_type == MDMFrameType.FLOOR_HEIGHTS:            # :267-278, [n, n, T, 1] as upstream
                assert ret_hf_obs
                fh = e(n, n, T)
                _lib.check(_lib.lib().parc_b200_mdm_floor_heights(self._cfg, float(self._dx), p(hfs), p(root_pos), n, p(fh),
                                                                 _lib.stream_ptr(dev)))
                motion_ret[MDMFrameType.FLOOR_HEIGHTS] = fh.unsqueeze(-1)
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
