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
                       future_noise=None, ret_hf_obs=True, aug=None, want_floor_heights=False):
    """sample_motion_data (:197-302) for given ids / start times; target info when future_t [n] (and the
    standard-normal draw future_noise [n,3]) are given. `aug` (use_hf_augmentation): dict(mode, hf_maxmins, mask_inds and
    rec_i / rec_f or noise), see local_heightmaps_augmented. Returns a dict."""
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
        if aug is None:
            hfs, center_h = local_heightmaps(lib, cfg, motion_ids, canon_pos, canon_rot)
        else:
            hfs, center_h = local_heightmaps_augmented(lib, aug["hf_maxmins"], aug["mask_inds"], cfg, motion_ids, start_times, canon_pos,
                                                       canon_rot, aug["mode"], aug.get("rec_i"), aug.get("rec_f"), aug.get("noise"))
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
               contacts=contacts, canon_pos=canon_pos, canon_rot=canon_rot, canon_heading_quat_inv=hinv)
    if want_floor_heights:
        out["floor_heights"] = floor_heights(cfg, out["hfs"], root_pos)                       # :267-278
    if future_t is not None:                                                                  # :146-160
        fp, fr, *_ = lib.calc_motion_frame(motion_ids, future_t)
        if future_noise is not None:
            fp = fp + cfg.future_pos_noise_scale * future_noise
        fp = O.quat_rotate(hinv, fp - canon_pos)
        out["future_pos"], out["future_rot"] = fp, O.quat_multiply(hinv, fr)
    return out


# ------------------------------------------------------------------------------------------------
# heightfield augmenter (:304-330, :332-364, :390-400) and FLOOR_HEIGHTS (:267-278)
# Pinned by the aug_* cases of tests/golden/ref_mdm_sampler.npz (tests/test_oracle_golden_sampler.py).
# ------------------------------------------------------------------------------------------------
POOL_2D, POOL_X, POOL_Y = 0, 1, 2          # terrain_util.maxpool_hf, maxpool_hf_1d_x, maxpool_hf_1d_y (terrain_util.py:1509-1536)
BOX_FLOATS = 7                             # centre x, y | side x, y | cos, sin of the angle | height


def draw_box_aug_records(n, grid_dims, max_h, max_num_boxes, box_min_len, box_max_len, maxpool_chance, max_maxpool_size,
                         change_height_chance):
    """The draws of _box_hf_augmentation (:304-330) + add_boxes_to_hf2 (terrain_util.py:861-914) for n windows in a row,
    consuming Python's `random` and torch's CPU generator exactly as the reference does.
    -> rec_i [n,8] int32 (change_height | pool kind, size x3 in applied order, kind -1 = not applied | num_boxes),
       rec_f [n, 1 + 7*max_num_boxes] fp32 (new_height | boxes)."""
    import random
    min_h = -max_h
    rec_i = torch.full((n, 8), -1, dtype=torch.int32)
    rec_f = torch.zeros(n, 1 + BOX_FLOATS * max(max_num_boxes, 1), dtype=torch.float32)
    shape = torch.tensor([float(grid_dims[0]), float(grid_dims[1])], dtype=torch.float32)
    ones = torch.ones(grid_dims[0], grid_dims[1])
    for i in range(n):
        change = random.random() < change_height_chance
        rec_i[i, 0] = int(change)
        if change:
            rec_f[i, 0] = random.random() * (max_h - min_h) + min_h
        use = [random.random() < maxpool_chance for _ in range(3)]
        fns = [POOL_2D, POOL_X, POOL_Y]
        random.shuffle(fns)
        for k in range(3):
            if use[k]:
                rec_i[i, 1 + 2 * k] = fns[k]
                rec_i[i, 2 + 2 * k] = random.randint(0, max_maxpool_size)
            else:
                rec_i[i, 2 + 2 * k] = 0
        nb = random.randint(0, max_num_boxes)
        rec_i[i, 7] = nb
        for b in range(nb):
            center = torch.rand(size=(2,), dtype=torch.float32) * shape
            side = torch.rand(size=(2,), dtype=torch.float32) * (box_max_len - box_min_len) + box_min_len
            angle = (torch.rand(1) * (2.0 * torch.pi - 0.0) + 0.0) * ones          # the reference takes cos / sin of the full grid
            h = torch.rand(1) * (max_h - min_h) + min_h
            o = 1 + BOX_FLOATS * b
            rec_f[i, o:o + 2], rec_f[i, o + 2:o + 4] = center, side
            rec_f[i, o + 4], rec_f[i, o + 5], rec_f[i, o + 6] = torch.cos(angle)[0, 0], torch.sin(angle)[0, 0], h[0]
    return rec_i, rec_f


def window_maxmins(lib_terrains, hf_maxmins, mask_inds, cfg: SamplerConfig, motion_ids, start_times, pts):
    """get_hfs_from_data_helper (:332-364): per window the (max, min) bounds of every grid point -- the terrain's own
    hf_maxmin where a body passes over the cell during the window's frames, (2 max_h, 2 min_h) elsewhere.
    mask_inds[m] = list over frames of [k,2] cell indices; pts [n,gx,gy,2] world points. -> [n,gx,gy,2]"""
    out = []
    idx0 = torch.round(start_times / cfg.timestep).to(torch.int64)                               # :213
    for i in range(motion_ids.shape[0]):
        m = int(motion_ids[i])
        ter = lib_terrains[m]
        f0, f1 = int(idx0[i]), int(idx0[i]) + cfg.seq_len - 1
        mask = torch.zeros(ter.hf.shape, dtype=torch.bool)
        for cells in mask_inds[m][slice(f0, f1 + 1)]:                                           # :348-350 (Python slicing clips)
            mask[cells[..., 0].long(), cells[..., 1].long()] = True
        mm = torch.zeros(*ter.hf.shape, 2)
        mm[..., 0], mm[..., 1] = cfg.max_h * 2.0, cfg.min_h * 2.0                               # :353-355
        mm[mask] = hf_maxmins[m][mask]                                                          # :357
        idx = ter.grid_index(pts[i])
        out.append(mm[idx[..., 0], idx[..., 1]])
    return torch.stack(out)


def augment_boxes(hf, mm, rec_i, rec_f):
    """_box_hf_augmentation (:304-330) on one window with its draws given. hf [gx,gy], mm [gx,gy,2] = (max, min)."""
    F = torch.nn.functional
    hf = hf.clone()
    if int(rec_i[0]):
        hf[...] = rec_f[0]
    for k in range(3):
        kind, size = int(rec_i[1 + 2 * k]), int(rec_i[2 + 2 * k])
        if kind < 0:
            continue
        w = 2 * size + 1
        x = hf[None, None]
        if kind == POOL_2D:
            x = F.max_pool2d(x, kernel_size=w, stride=1, padding=size)
        elif kind == POOL_X:
            x = F.max_pool2d(x, kernel_size=(w, 1), stride=1, padding=(size, 0))
        else:
            x = F.max_pool2d(x, kernel_size=(1, w), stride=1, padding=(0, size))
        hf = torch.clamp(x[0, 0], mm[..., 1], mm[..., 0])                                       # terrain_util.py:1513-1514
    gx, gy = hf.shape
    X, Y = torch.meshgrid(torch.arange(gx, dtype=torch.float32), torch.arange(gy, dtype=torch.float32), indexing="ij")
    for b in range(int(rec_i[7])):
        cx, cy, lx, ly, c, s, h = [rec_f[1 + BOX_FLOATS * b + q] for q in range(BOX_FLOATS)]
        dx_, dy_ = X - cx, Y - cy
        rx = (dx_ * c - dy_ * s) + cx                                                           # rotate_2d_vec + box_center (:890)
        ry = (dx_ * s + dy_ * c) + cy
        inside = (rx < cx + lx / 2) & (rx > cx - lx / 2) & (ry < cy + ly / 2) & (ry > cy - ly / 2)
        hf = torch.where(inside, h.expand_as(hf), hf)
    return torch.clamp(hf, mm[..., 1], mm[..., 0])                                              # terrain_util.py:911-912


def augment_noise(mm, noise, cfg: SamplerConfig):
    """_noise_hf_augmentation (:332-335): clamp(noise, min=mm[...,0], max=mm[...,1]) -- the bounds arrive swapped, and
    torch.clamp with min > max returns max, so the result is mm[...,1] wherever the bounds are a proper interval."""
    return torch.clamp(noise * (cfg.max_h - cfg.min_h) + cfg.min_h, mm[..., 0], mm[..., 1])


def grid_points_world(cfg: SamplerConfig, canon_pos, canon_rot):
    """:367-374: the generic grid rotated by the heading of the reference frame and moved to its root -> [n,gx,gy,2]"""
    n = canon_pos.shape[0]
    pts = cfg.grid.unsqueeze(0).expand(n, -1, -1, -1)
    heading = O.calc_heading(canon_rot).unsqueeze(-1).unsqueeze(-1).expand(-1, cfg.grid.shape[0], cfg.grid.shape[1])
    return O.rotate_2d_vec(pts, heading) + canon_pos[:, 0:2].unsqueeze(1).unsqueeze(1)


def local_heightmaps_augmented(lib: O.Library, hf_maxmins, mask_inds, cfg: SamplerConfig, motion_ids, start_times, canon_pos, canon_rot,
                               mode, rec_i=None, rec_f=None, noise=None):
    """get_hfs_from_data (:366-401) with the augmenter, then the clamp of sample_motion_data (:228). mode: "NOISE" |
    "MAXPOOL_AND_BOXES" | "NONE". -> hfs [n,gx,gy], center_h [n]"""
    n = motion_ids.shape[0]
    pts = grid_points_world(cfg, canon_pos, canon_rot)
    hfs = torch.stack([lib.terrains[int(motion_ids[i])].height_at(pts[i]) for i in range(n)], dim=0)
    mms = window_maxmins(lib.terrains, hf_maxmins, mask_inds, cfg, motion_ids, start_times, pts)
    center_h = hfs[:, cfg.num_x_neg, cfg.num_y_neg].clone()
    base = center_h if cfg.relative_z_style == RELATIVE_TO_ROOT_FLOOR else canon_pos[:, 2]
    hfs = hfs - base[:, None, None]
    mms = mms - base[:, None, None, None]
    out = []
    for i in range(n):
        if mode == "NOISE":
            out.append(augment_noise(mms[i], noise[i], cfg))
        elif mode == "MAXPOOL_AND_BOXES":
            out.append(augment_boxes(hfs[i], mms[i], rec_i[i], rec_f[i]))
        else:
            out.append(hfs[i])
    return torch.clamp(torch.stack(out), min=cfg.min_h, max=cfg.max_h), center_h


def floor_heights(cfg: SamplerConfig, hfs, root_pos):
    """FLOOR_HEIGHTS (:267-278): for EVERY window's heightmap, the heights under the canonical root positions of ALL
    windows of the batch (the reference indexes the whole batch inside its per-window loop) -> [n, n, T, 1]."""
    gmin = torch.tensor([-cfg.num_x_neg, -cfg.num_y_neg], dtype=torch.float32) * cfg.dx        # :72
    dims = torch.tensor([cfg.grid.shape[0], cfg.grid.shape[1]], dtype=torch.int64)
    inds = torch.round((root_pos[:, :, 0:2] - gmin) / cfg.dx).to(torch.int64)
    inds = torch.clamp(inds, torch.zeros_like(dims), dims - 1)
    return torch.stack([hf[inds[..., 0], inds[..., 1]] for hf in hfs]).unsqueeze(-1)
