"""TEST INFRASTRUCTURE ONLY -- CPU restatement (torch, fp32, eager) of the deterministic part of
PARC's MDM dataset sampler, MDMHeightfieldContactMotionSampler.sample_motion_data
(PARC/motion_generator/mdm_heightfield_contact_motion_sampler.py:197-302) with its helpers
get_hfs_from_data (:366-401), get_hfs_from_data_helper (:332-364, the height lookup only: the
hf_maxmin bookkeeping feeds the augmenter alone) and _sample_target_info (:146-160).

SURVEY.md section 8(f) rank 1. Like oracle/parc_oracle.py this file is the checker for the CUDA path
(parc_b200/motion_sampler.py) and is imported by tests only. Random draws (motion ids, start times,
future times, future-position noise) are INPUTS here: the reference draws them from torch's
generator (motion_sampler.py:39-54, :133-153), which a CUDA generator cannot reproduce.

Parity pin: tests/golden/make_golden_sampler.py runs the unmodified reference sampler on the two
bundled clips that carry hf_mask_inds and commits inputs + outputs as tests/golden/ref_mdm_sampler.npz;
tests/test_oracle_golden_sampler.py checks this restatement against them.
"""
import torch

from . import parc_oracle as O

RELATIVE_TO_ROOT, RELATIVE_TO_ROOT_FLOOR = 0, 1      # RelativeZStyle, diffusion_util.py:9-11


def xy_grid_points(dx, dy, num_x_neg, num_x_pos, num_y_neg, num_y_pos):
    """geom_util.get_xy_grid_points (geom_util.py:212-223) around the origin -> [gx, gy, 2]"""
    x = torch.linspace(0.0 - dx * num_x_neg, 0.0 + dx * num_x_pos, num_x_neg + num_x_pos + 1)
    y = torch.linspace(0.0 - dy * num_y_neg, 0.0 + dy * num_y_pos, num_y_neg + num_y_pos + 1)
    gx, gy = torch.meshgrid(x, y, indexing="ij")
    return torch.stack([gx, gy], dim=-1)


class SamplerConfig:
    def __init__(self, sequence_duration=0.5, sequence_fps=30, num_prev_states=2, relative_z_style=RELATIVE_TO_ROOT,
                 dx=0.2, num_x_neg=10, num_x_pos=20, num_y_neg=15, num_y_pos=15, max_h=3.0, future_pos_noise_scale=0.05,
                 future_window_min=0.4, future_window_max=1.5):
        self.fps = sequence_fps
        self.timestep = 1.0 / sequence_fps                                                     # :40-41
        self.motion_times = torch.arange(start=0.0, end=sequence_duration, step=self.timestep, dtype=torch.float32)   # :50-51
        self.seq_len = self.motion_times.shape[0]
        self.ref_frame_idx = num_prev_states - 1                                              # :99
        self.relative_z_style = relative_z_style
        self.dx = dx
        self.num_x_neg, self.num_x_pos, self.num_y_neg, self.num_y_pos = num_x_neg, num_x_pos, num_y_neg, num_y_pos
        self.grid = xy_grid_points(dx, dx, num_x_neg, num_x_pos, num_y_neg, num_y_pos)         # :83
        self.max_h, self.min_h = max_h, -max_h                                                 # :86-87
        self.future_pos_noise_scale = future_pos_noise_scale
        self.future_window_min, self.future_window_max = future_window_min, future_window_max


def local_heightmaps(lib: O.Library, cfg: SamplerConfig, motion_ids, canon_pos, canon_rot):
    """get_hfs_from_data (:366-401) without augmentation -> hfs [n, gx, gy] (clamped, :228), center_h [n]"""
    n = motion_ids.shape[0]
    pts = cfg.grid.unsqueeze(0).expand(n, -1, -1, -1)
    heading = O.calc_heading(canon_rot).unsqueeze(-1).unsqueeze(-1).expand(-1, cfg.grid.shape[0], cfg.grid.shape[1])
    pts = O.rotate_2d_vec(pts, heading) + canon_pos[:, 0:2].unsqueeze(1).unsqueeze(1)
    hfs = torch.stack([lib.terrains[int(motion_ids[i])].height_at(pts[i]) for i in range(n)], dim=0)   # :336-339
    center_h = hfs[:, cfg.num_x_neg, cfg.num_y_neg].clone()
    if cfg.relative_z_style == RELATIVE_TO_ROOT_FLOOR:
        hfs = hfs - center_h.unsqueeze(-1).unsqueeze(-1)
    else:
        hfs = hfs - canon_pos[:, 2].unsqueeze(-1).unsqueeze(-1)
    return torch.clamp(hfs, min=cfg.min_h, max=cfg.max_h), center_h


def future_times(lib: O.Library, cfg: SamplerConfig, motion_ids, start_times, u):
    """_sample_motion_future_times (:133-144) with the uniform draw u in [0,1) supplied"""
    remaining = torch.clamp_max(lib.lengths[motion_ids] - start_times, cfg.future_window_max - cfg.future_window_min)
    return u * remaining + start_times + cfg.future_window_min


def sample_motion_data(model: O.CharModel, lib: O.Library, cfg: SamplerConfig, motion_ids, start_times, future_t=None,
                       future_noise=None, ret_hf_obs=True):
    """sample_motion_data (:197-302) for given ids / start times; target info when future_t [n] (and the
    standard-normal draw future_noise [n,3]) are given. Returns a dict."""
    n, T = motion_ids.shape[0], cfg.seq_len
    times = (start_times.unsqueeze(-1) + cfg.motion_times).flatten()                           # :168-170
    ids = motion_ids.unsqueeze(-1).expand(-1, T).flatten()
    root_pos, root_rot, root_vel, _, joint_rot, _, contacts = lib.calc_motion_frame(ids, times)
    root_pos = root_pos.reshape(n, T, 3).clone()
    root_rot = root_rot.reshape(n, T, 4)
    root_vel = root_vel.reshape(n, T, 3)
    joint_rot = joint_rot.reshape(n, T, -1, 4)
    contacts = contacts.reshape(n, T, -1)
    canon_pos = root_pos[:, cfg.ref_frame_idx].clone()                                        # :222-224
    canon_rot = root_rot[:, cfg.ref_frame_idx].clone()
    out = {}
    if ret_hf_obs:
        hfs, center_h = local_heightmaps(lib, cfg, motion_ids, canon_pos, canon_rot)
        out["hfs"], out["center_h"] = hfs, center_h
        if cfg.relative_z_style == RELATIVE_TO_ROOT_FLOOR:                                    # :230-232
            root_pos[..., 2] = root_pos[..., 2] - center_h.unsqueeze(-1)
    root_pos = root_pos - canon_pos.unsqueeze(1)                                              # :235-239
    hinv = O.calc_heading_quat_inv(canon_rot)
    he = hinv.unsqueeze(1).expand(-1, T, -1)
    root_rot = O.quat_multiply(he, root_rot)
    root_vel = O.quat_rotate(he, root_vel)
    root_pos = O.quat_rotate(he, root_pos)
    body_pos, _ = model.forward_kinematics(root_pos, root_rot, joint_rot)                     # :241
    out.update(root_pos=root_pos, root_rot=root_rot, root_vel=root_vel, joint_pos=body_pos[..., 1:, :], joint_rot=joint_rot,
               contacts=contacts, canon_pos=canon_pos, canon_heading_quat_inv=hinv)
    if future_t is not None:                                                                  # :146-160
        fp, fr, *_ = lib.calc_motion_frame(motion_ids, future_t)
        if future_noise is not None:
            fp = fp + cfg.future_pos_noise_scale * future_noise
        fp = O.quat_rotate(hinv, fp - canon_pos)
        out["future_pos"], out["future_rot"] = fp, O.quat_multiply(hinv, fr)
    return out
