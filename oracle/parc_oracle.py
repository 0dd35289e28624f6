"""TEST INFRASTRUCTURE ONLY -- CPU restatement (torch, fp32, eager) of PARC's tracking hot path.

This is the ORACLE the CUDA kernels are checked against and the `cpu_baseline` / `--impl reference`
arm of bench.py. Nothing under `parc_b200/` imports it; only `tests/`, `__graft_entry__.smoke()`
and bench.py's CPU legs do. The product path has no CPU fallback.

Parity pin: every function here is checked against the UNMODIFIED reference (imported from
/root/reference through oracle/ref_import.py) by tests/golden/make_golden.py, which also writes the
golden vectors under tests/golden/ that `pytest -m "not gpu"` re-checks on any box, plus the one
golden vector the reference itself ships (the recorded observation tape of
data/motion_terrains/dec2024_teaser_717_1_opt_dm.pkl).

The reference is a torch program, so the restatement uses torch CPU ops in the reference's
operation ORDER (quat_mul's 8-multiplication form, separately rounded phase arithmetic, ...): frame
indices and done flags then agree bit for bit, floats to ~1 ulp. What differs from the reference is
structure: no per-joint Python loops, no boolean-mask scatters, one pass per level of the tree.
Citations are file:line into /root/reference.
"""
import math

import numpy as np
import torch

ROOT, HINGE, SPHERICAL, FIXED = 0, 1, 2, 3      # JointType, PARC/anim/kin_char_model.py:15-19
CLAMP, WRAP = 0, 1                              # LoopMode, PARC/anim/motion_lib.py:14-16
DONE_NULL, DONE_FAIL, DONE_SUCC, DONE_TIME = 0, 1, 2, 3   # PARC/motion_tracker/envs/base_env.py:14-18


# ------------------------------------------------------------------------------------------------
# quaternion / exp-map arithmetic (PARC/util/torch_util.py); quaternions are (x, y, z, w)
# ------------------------------------------------------------------------------------------------
def normalize(x, eps=1e-9):
    """torch_util.py:10-13"""
    return x / x.norm(p=2, dim=-1).clamp(min=eps).unsqueeze(-1)


def quat_conj(q):
    """torch_util.py:30-32"""
    return torch.cat([-q[..., :3], q[..., 3:]], dim=-1)


def quat_pos(q):
    """torch_util.py:34-39"""
    return (1 - 2 * (q[..., 3:] < 0).float()) * q


def quat_mul(a, b):
    """torch_util.py:41-59 (the 8-multiplication factoring, same op order)."""
    a, b = torch.broadcast_tensors(a, b)
    x1, y1, z1, w1 = a.unbind(-1)
    x2, y2, z2, w2 = b.unbind(-1)
    ww = (z1 + x1) * (x2 + y2)
    yy = (w1 - y1) * (w2 + z2)
    zz = (w1 + y1) * (w2 - z2)
    xx = ww + yy + zz
    qq = 0.5 * (xx + (z1 - x1) * (x2 - y2))
    w = qq - ww + (z1 - y1) * (y2 - z2)
    x = qq - xx + (x1 + w1) * (x2 + w2)
    y = qq - yy + (w1 - x1) * (y2 + z2)
    z = qq - zz + (z1 + y1) * (w2 - x2)
    return torch.stack([x, y, z, w], dim=-1)


def quat_rotate(q, v):
    """torch_util.py:61-67"""
    lead = torch.broadcast_shapes(q.shape[:-1], v.shape[:-1])
    qv = q[..., :3].expand(*lead, 3)
    v = v.expand(*lead, 3)
    t = 2 * torch.cross(qv, v, dim=-1)
    return v + q[..., 3:] * t + torch.cross(qv, t, dim=-1)


def quat_to_axis_angle(q):
    """torch_util.py:69-91"""
    q = quat_pos(q)
    length = torch.norm(q[..., 0:3], dim=-1, p=2)
    angle = 2.0 * torch.atan2(length, q[..., 3])
    axis = q[..., 0:3] / length.unsqueeze(-1).clamp(min=1e-6)
    default_axis = torch.zeros_like(axis)
    default_axis[..., -1] = 1
    mask = length > 1e-5
    angle = torch.where(mask, angle, torch.zeros_like(angle))
    axis = torch.where(mask.unsqueeze(-1), axis, default_axis)
    return axis, angle


def quat_to_exp_map(q):
    """torch_util.py:371-376 (axis_angle_to_exp_map = angle * axis, :344-349)"""
    axis, angle = quat_to_axis_angle(q)
    return angle.unsqueeze(-1) * axis


def normalize_angle(x):
    """torch_util.py:5-8"""
    return torch.atan2(torch.sin(x), torch.cos(x))


def axis_angle_to_quat(axis, angle):
    """torch_util.py:336-342"""
    theta = (angle / 2).unsqueeze(-1)
    xyz = normalize(axis) * theta.sin()
    return normalize(torch.cat([xyz, theta.cos()], dim=-1))


def exp_map_to_quat(e):
    """torch_util.py:425-450"""
    angle = torch.norm(e, dim=-1)
    axis = e / angle.unsqueeze(-1)
    angle = normalize_angle(angle)
    default_axis = torch.zeros_like(e)
    default_axis[..., -1] = 1
    mask = torch.abs(angle) > 1e-5
    angle = torch.where(mask, angle, torch.zeros_like(angle))
    axis = torch.where(mask.unsqueeze(-1), axis, default_axis)
    return axis_angle_to_quat(axis, angle)


def quat_diff_angle(q0, q1):
    """torch_util.py:453-462"""
    return quat_to_axis_angle(quat_mul(q1, quat_conj(q0)))[1]


def quat_normalize(q):
    """torch_util.py:469-472"""
    return normalize(quat_pos(q))


def slerp(q0, q1, t):
    """torch_util.py:474-499; t has one dim less than q0 or broadcasts against it."""
    c = torch.sum(q0 * q1, dim=-1)
    q1 = torch.where((c < 0).unsqueeze(-1), -q1, q1)
    c = torch.abs(c).unsqueeze(-1)
    half = torch.acos(c)
    s = torch.sqrt(1.0 - c * c)
    if t.dim() == q0.dim() - 1:
        t = t.unsqueeze(-1)
    ra = torch.sin((1 - t) * half) / s
    rb = torch.sin(t * half) / s
    out = ra * q0 + rb * q1
    out = torch.where(torch.abs(s) < 0.001, 0.5 * q0 + 0.5 * q1, out)
    out = torch.where(torch.abs(c) >= 1, q0, out)
    return out


def calc_heading(q):
    """torch_util.py:501-510"""
    ref = torch.zeros_like(q[..., 0:3])
    ref[..., 0] = 1
    d = quat_rotate(q, ref)
    return torch.atan2(d[..., 1], d[..., 0])


def calc_heading_quat_inv(q):
    """torch_util.py:522-530"""
    axis = torch.zeros_like(q[..., 0:3])
    axis[..., 2] = 1
    return axis_angle_to_quat(axis, -calc_heading(q))


def quat_to_tan_norm(q):
    """torch_util.py:392-404"""
    tan = torch.zeros_like(q[..., 0:3])
    tan[..., 0] = 1
    nrm = torch.zeros_like(q[..., 0:3])
    nrm[..., -1] = 1
    return torch.cat([quat_rotate(q, tan), quat_rotate(q, nrm)], dim=-1)


def rotate_2d_vec(vec, angle):
    """torch_util.py:650-662"""
    x, y = vec[..., 0], vec[..., 1]
    c, s = torch.cos(angle), torch.sin(angle)
    return torch.stack([x * c - y * s, x * s + y * c], dim=-1)


def quat_multiply(q1, q2):
    """torch_util.py:607-625 (plain Hamilton product, used for the reset rotation offset)."""
    x1, y1, z1, w1 = q1.unbind(-1)
    x2, y2, z2, w2 = q2.unbind(-1)
    w = w1 * w2 - x1 * x2 - y1 * y2 - z1 * z2
    x = w1 * x2 + x1 * w2 + y1 * z2 - z1 * y2
    y = w1 * y2 - x1 * z2 + y1 * w2 + z1 * x2
    z = w1 * z2 + x1 * y2 - y1 * x2 + z1 * w2
    return torch.stack([x, y, z, w], dim=-1)


# ------------------------------------------------------------------------------------------------
# kinematic tree (PARC/anim/kin_char_model.py)
# ------------------------------------------------------------------------------------------------
class CharModel:
    """Built from the JSON skeleton spec (parc_b200.KinCharModel.to_spec): body_names,
    parent_indices, local_translation, local_rotation, joints[{name,type,axis,limits}]."""

    def __init__(self, spec):
        self.body_names = list(spec["body_names"])
        self.parent = [int(p) for p in spec["parent_indices"]]
        self.J = len(self.parent)
        self.local_trans = torch.tensor(spec["local_translation"], dtype=torch.float32)
        self.local_rot = torch.tensor(spec["local_rotation"], dtype=torch.float32)
        names = {"ROOT": ROOT, "HINGE": HINGE, "SPHERICAL": SPHERICAL, "FIXED": FIXED}
        self.jtype = [names[j["type"]] for j in spec["joints"]]
        self.axis = [None if j["axis"] is None else torch.tensor(j["axis"], dtype=torch.float32) for j in spec["joints"]]
        self.limits = [None if j["limits"] is None else torch.tensor(j["limits"], dtype=torch.float32) for j in spec["joints"]]
        self.dof_idx, d = [], 0
        for t in self.jtype:                           # _label_dof_indices, :936-943
            self.dof_idx.append(d)
            d += 3 if t == SPHERICAL else (1 if t == HINGE else 0)
        self.D = d
        depth = [0] * self.J
        for j in range(1, self.J):
            depth[j] = depth[self.parent[j]] + 1
        self.levels = [[j for j in range(self.J) if depth[j] == d] for d in range(1, max(depth) + 1)]
        self.sph = [j for j in range(self.J) if self.jtype[j] == SPHERICAL]
        self.hin = [j for j in range(self.J) if self.jtype[j] == HINGE]

    def body_id(self, name):
        return self.body_names.index(name)

    def dof_limits(self):
        """-> lower [D], upper [D] (kin_char_model.py:199-217)"""
        lo, hi = torch.zeros(self.D), torch.zeros(self.D)
        for j in range(1, self.J):
            if self.jtype[j] == HINGE:
                lo[self.dof_idx[j]], hi[self.dof_idx[j]] = self.limits[j][0], self.limits[j][1]
            elif self.jtype[j] == SPHERICAL:
                lo[self.dof_idx[j]:self.dof_idx[j] + 3] = self.limits[j][:, 0]
                hi[self.dof_idx[j]:self.dof_idx[j] + 3] = self.limits[j][:, 1]
        return lo, hi

    def forward_kinematics(self, root_pos, root_rot, joint_rot):
        """kin_char_model.py:617-649, one pass per tree level instead of one per joint."""
        lead = root_pos.shape[:-1]
        pos = torch.zeros(*lead, self.J, 3)
        rot = torch.zeros(*lead, self.J, 4)
        pos[..., 0, :] = root_pos
        rot[..., 0, :] = root_rot
        for lv in self.levels:
            par = [self.parent[j] for j in lv]
            ppos, prot = pos[..., par, :], rot[..., par, :]
            jq = joint_rot[..., [j - 1 for j in lv], :]
            pos[..., lv, :] = ppos + quat_rotate(prot, self.local_trans[lv].expand(*lead, len(lv), 3))
            rot[..., lv, :] = quat_mul(prot, quat_mul(self.local_rot[lv].expand(*lead, len(lv), 4), jq))
        return pos, rot

    def dof_to_rot(self, dof):
        """kin_char_model.py:586-599 / Joint.dof_to_rot :61-81"""
        lead = dof.shape[:-1]
        rot = torch.zeros(*lead, self.J - 1, 4)
        rot[..., 3] = 1.0
        if self.sph:
            e = torch.stack([dof[..., self.dof_idx[j]:self.dof_idx[j] + 3] for j in self.sph], dim=-2)
            rot[..., [j - 1 for j in self.sph], :] = exp_map_to_quat(e)
        if self.hin:
            ang = torch.stack([dof[..., self.dof_idx[j]] for j in self.hin], dim=-1)
            ax = torch.stack([self.axis[j] for j in self.hin], dim=0).expand(*lead, len(self.hin), 3)
            rot[..., [j - 1 for j in self.hin], :] = axis_angle_to_quat(ax, ang)
        return rot

    def rot_to_dof(self, rot):
        """kin_char_model.py:601-615 / Joint.rot_to_dof :83-104"""
        lead = rot.shape[:-2]
        dof = torch.zeros(*lead, self.D)
        if self.sph:
            e = quat_to_exp_map(rot[..., [j - 1 for j in self.sph], :])
            for k, j in enumerate(self.sph):
                dof[..., self.dof_idx[j]:self.dof_idx[j] + 3] = e[..., k, :]
        if self.hin:
            axis, angle = quat_to_axis_angle(rot[..., [j - 1 for j in self.hin], :])
            jax = torch.stack([self.axis[j] for j in self.hin], dim=0)
            sign = torch.where(torch.sum(jax * axis, dim=-1) < 0, -1.0, 1.0)
            for k, j in enumerate(self.hin):
                dof[..., self.dof_idx[j]] = (sign * angle)[..., k]
        return dof

    def compute_frame_dof_vel(self, joint_rot, dt):
        """kin_char_model.py:651-690: [T,J-1,4] -> [T,D], last frame repeats the one before."""
        T = joint_rot.shape[0]
        out = torch.zeros(T, self.D)
        if T < 2:
            return out
        drot = quat_normalize(quat_mul(quat_conj(joint_rot[:-1]), joint_rot[1:]))
        vel = quat_to_exp_map(drot) / dt                         # [T-1, J-1, 3]
        for j in self.sph:
            out[:-1, self.dof_idx[j]:self.dof_idx[j] + 3] = vel[:, j - 1]
        for j in self.hin:
            out[:-1, self.dof_idx[j]] = torch.sum(self.axis[j] * vel[:, j - 1], dim=-1)
        out[-1] = out[-2]
        return out


# ------------------------------------------------------------------------------------------------
# terrain (PARC/util/terrain_util.py:18-156)
# ------------------------------------------------------------------------------------------------
class Terrain:
    def __init__(self, hf, min_point, dxdy):
        self.hf = torch.as_tensor(hf, dtype=torch.float32).contiguous()
        self.min_point = torch.as_tensor(min_point, dtype=torch.float32)
        self.dxdy = torch.as_tensor(dxdy, dtype=torch.float32)
        self.dims = torch.tensor(self.hf.shape, dtype=torch.int64)

    def grid_index(self, p):
        """:143-152: clamp(round_half_even((p - min) / dxdy), 0, dims - 1)"""
        idx = torch.round((p - self.min_point) / self.dxdy).long()
        return torch.minimum(torch.clamp(idx, min=0), self.dims - 1)

    def height_at(self, p):
        """:154-156"""
        idx = self.grid_index(p)
        return self.hf[idx[..., 0], idx[..., 1]]


# ------------------------------------------------------------------------------------------------
# motion library (PARC/anim/motion_lib.py)
# ------------------------------------------------------------------------------------------------
class Library:
    """clips: list of dicts root_pos [nf,3], root_rot [nf,4], joint_rot [nf,J-1,4],
    contacts [nf,C] | None, fps, loop_mode (int), terrain (Terrain | None). from_file semantics
    (motion_lib.py:255-401): rotations are NOT renormalised, velocities are finite differences."""

    def __init__(self, model: CharModel, clips, weights=None):
        self.model = model
        rp, rr, rv, rav, jr, dv, con = [], [], [], [], [], [], []
        nfs, lens, loops, deltas, fpss = [], [], [], [], []
        self.terrains = []
        for c in clips:
            p = torch.as_tensor(c["root_pos"], dtype=torch.float32)
            r = torch.as_tensor(c["root_rot"], dtype=torch.float32)
            j = torch.as_tensor(c["joint_rot"], dtype=torch.float32)
            nf, fps = p.shape[0], c["fps"]
            delta = p[-1] - p[0]                                   # :315-316
            delta[-1] = 0.0
            vel = torch.zeros_like(p)                              # :318-320
            ang = torch.zeros_like(p)                              # :322-325
            if nf > 1:
                vel[:-1] = fps * (p[1:] - p[:-1])
                vel[-1] = vel[-2]
                ang[:-1] = fps * quat_to_exp_map(quat_mul(r[1:], quat_conj(r[:-1])))
                ang[-1] = ang[-2]
            rp.append(p); rr.append(r); jr.append(j); rv.append(vel); rav.append(ang)
            dv.append(model.compute_frame_dof_vel(j, 1.0 / fps))   # :327
            cc = c.get("contacts")
            con.append(torch.zeros(nf, model.J) if cc is None else torch.as_tensor(cc, dtype=torch.float32))
            nfs.append(nf); fpss.append(fps); loops.append(int(c["loop_mode"]))
            lens.append(1.0 / fps * (nf - 1))                      # :313 (python double, cast once)
            deltas.append(delta)
            self.terrains.append(c.get("terrain"))
        M = len(clips)
        w = torch.tensor([1.0] * M if weights is None else weights, dtype=torch.float32)
        self.motion_weights = w / w.sum()                          # :374-375
        self.num_frames = torch.tensor(nfs, dtype=torch.int64)
        self.lengths = torch.tensor(lens, dtype=torch.float32)
        self.loop_modes = torch.tensor(loops, dtype=torch.int32)
        self.fps = torch.tensor(fpss, dtype=torch.float32)
        self.root_pos_delta = torch.stack(deltas)
        self.start_idx = torch.cumsum(self.num_frames, 0) - self.num_frames    # :394-396
        self.frame_root_pos = torch.cat(rp); self.frame_root_rot = torch.cat(rr)
        self.frame_root_vel = torch.cat(rv); self.frame_root_ang_vel = torch.cat(rav)
        self.frame_joint_rot = torch.cat(jr); self.frame_dof_vel = torch.cat(dv)
        self.frame_contacts = torch.cat(con)

    def num_motions(self):
        return self.lengths.shape[0]

    def frame_blend(self, ids, times):
        """_calc_frame_blend :425-438 + calc_phase :519-530 (each fp32 op rounded on its own)."""
        nf = self.num_frames[ids]
        phase = times / self.lengths[ids]
        wrap = self.loop_modes[ids] == WRAP
        phase = torch.where(wrap, phase - torch.floor(phase), phase)
        phase = torch.clip(phase, 0.0, 1.0)
        scaled = phase * (nf - 1)
        i0 = scaled.long()
        i1 = torch.min(i0 + 1, nf - 1)
        blend = scaled - i0
        s = self.start_idx[ids]
        return i0 + s, i1 + s, blend

    def calc_motion_frame(self, ids, times):
        """:94-126 -> root_pos, root_rot, root_vel, root_ang_vel, joint_rot, dof_vel, contacts"""
        i0, i1, blend = self.frame_blend(ids, times)
        b = blend.unsqueeze(-1)
        root_pos = (1.0 - b) * self.frame_root_pos[i0] + b * self.frame_root_pos[i1]
        root_rot = slerp(self.frame_root_rot[i0], self.frame_root_rot[i1], blend)
        joint_rot = slerp(self.frame_joint_rot[i0], self.frame_joint_rot[i1], b)
        # _calc_loop_offset :440-457
        wrap = (self.loop_modes[ids] == WRAP).unsqueeze(-1)
        cycles = torch.floor(times / self.lengths[ids]).unsqueeze(-1)
        root_pos = root_pos + torch.where(wrap, cycles * self.root_pos_delta[ids], torch.zeros(1))
        contacts = (1.0 - b) * self.frame_contacts[i0] + b * self.frame_contacts[i1]
        return (root_pos, root_rot, self.frame_root_vel[i0], self.frame_root_ang_vel[i0], joint_rot,
                self.frame_dof_vel[i0], contacts)


# ------------------------------------------------------------------------------------------------
# observation / reward / termination (free jit functions of the reference)
# ------------------------------------------------------------------------------------------------
def compute_char_obs(root_pos, root_rot, root_vel, root_ang_vel, joint_rot, dof_vel, key_pos, global_obs,
                     root_height_obs):
    """PARC/motion_tracker/envs/ig_char_env.py:581-625"""
    hinv = calc_heading_quat_inv(root_rot)
    if global_obs:
        r_rot, r_vel, r_av = root_rot, root_vel, root_ang_vel
    else:
        r_rot = quat_mul(hinv, root_rot)
        r_vel = quat_rotate(hinv, root_vel)
        r_av = quat_rotate(hinv, root_ang_vel)
    n = root_pos.shape[0]
    parts = [root_pos[:, 2:3]] if root_height_obs else []
    parts += [quat_to_tan_norm(r_rot), r_vel, r_av, quat_to_tan_norm(joint_rot).reshape(n, -1), dof_vel]
    if key_pos.numel() > 0:
        kp = key_pos - root_pos.unsqueeze(-2)
        if not global_obs:
            kp = quat_rotate(hinv.unsqueeze(-2), kp)
        parts.append(kp.reshape(n, -1))
    return torch.cat(parts, dim=-1)


def compute_tar_obs(ref_root_pos, ref_root_rot, tar_root_pos, tar_root_rot, joint_rot, tar_key_pos, global_obs,
                    global_tar_root_h_obs):
    """PARC/motion_tracker/envs/ig_parkour/mgdm_dm_util.py:404-461 -> [n, S, 9 + 6(J-1) + 3K]"""
    n, S = tar_root_pos.shape[0], tar_root_pos.shape[1]
    root_pos_obs = tar_root_pos - ref_root_pos.unsqueeze(-2)
    has_key = tar_key_pos.numel() > 0
    if has_key:
        tar_key_pos = tar_key_pos - tar_root_pos.unsqueeze(-2)
    if not global_obs:
        hinv = calc_heading_quat_inv(ref_root_rot).unsqueeze(-2)            # [n,1,4]
        root_pos_obs = quat_rotate(hinv, root_pos_obs)
        tar_root_rot = quat_mul(hinv.expand(n, S, 4), tar_root_rot)
        if has_key:
            tar_key_pos = quat_rotate(hinv.unsqueeze(-2), tar_key_pos) + root_pos_obs.unsqueeze(2)
    if global_tar_root_h_obs:
        root_pos_obs = root_pos_obs.clone()
        root_pos_obs[..., 2] = tar_root_pos[..., 2]
    parts = [root_pos_obs, quat_to_tan_norm(tar_root_rot), quat_to_tan_norm(joint_rot).reshape(n, S, -1)]
    if has_key:
        parts.append(tar_key_pos.reshape(n, S, -1))
    return torch.cat(parts, dim=-1)


def _to_local(root_rot, root_vel, root_ang_vel, key_pos):
    """convert_to_local, mgdm_dm_util.py:246-267"""
    hinv = calc_heading_quat_inv(root_rot)
    kp = quat_rotate(hinv.unsqueeze(-2), key_pos) if key_pos.numel() > 0 else key_pos
    return quat_mul(hinv, root_rot), quat_rotate(hinv, root_vel), quat_rotate(hinv, root_ang_vel), kp


def compute_deepmimic_reward(root_pos, root_rot, root_vel, root_ang_vel, joint_rot, dof_vel, key_pos, tar_root_pos,
                             tar_root_rot, tar_root_vel, tar_root_ang_vel, tar_joint_rot, tar_dof_vel, tar_key_pos,
                             joint_rot_err_w, dof_err_w, track_root_h, track_root):
    """mgdm_dm_util.py:269-332 -> [n,5]"""
    pose_diff = quat_diff_angle(joint_rot, tar_joint_rot)
    pose_err = torch.sum(joint_rot_err_w * pose_diff * pose_diff, dim=-1)
    vel_diff = tar_dof_vel - dof_vel
    vel_err = torch.sum(dof_err_w * vel_diff * vel_diff, dim=-1)
    rpd = tar_root_pos - root_pos
    if not track_root:
        rpd[..., 0:2] = 0
    if not track_root_h:
        rpd[..., 2] = 0
    root_pos_err = torch.sum(rpd * rpd, dim=-1)
    has_key = key_pos.numel() > 0
    if has_key:
        key_pos = key_pos - root_pos.unsqueeze(-2)
        tar_key_pos = tar_key_pos - tar_root_pos.unsqueeze(-2)
    if not track_root:
        root_rot, root_vel, root_ang_vel, key_pos = _to_local(root_rot, root_vel, root_ang_vel, key_pos)
        tar_root_rot, tar_root_vel, tar_root_ang_vel, tar_key_pos = _to_local(tar_root_rot, tar_root_vel,
                                                                               tar_root_ang_vel, tar_key_pos)
    rre = quat_diff_angle(root_rot, tar_root_rot)
    rre = rre * rre
    rvd = tar_root_vel - root_vel
    root_vel_err = torch.sum(rvd * rvd, dim=-1)
    rad = tar_root_ang_vel - root_ang_vel
    root_ang_vel_err = torch.sum(rad * rad, dim=-1)
    if has_key:
        kd = tar_key_pos - key_pos
        key_err = torch.sum(torch.sum(kd * kd, dim=-1), dim=-1)
    else:
        key_err = torch.zeros_like(pose_err)
    return torch.stack([torch.exp(-0.25 * pose_err), torch.exp(-0.01 * vel_err),
                        torch.exp(-5.0 * (root_pos_err + 0.1 * rre)),
                        torch.exp(-1.0 * (root_vel_err + 0.1 * root_ang_vel_err)),
                        torch.exp(-10.0 * key_err)], dim=1)


def compute_contact_reward(tar_contacts, contact_forces, contact_weights):
    """mgdm_dm_util.py:497-518"""
    forces = torch.clamp_max(torch.norm(contact_forces, dim=2), 1.0)
    r = -(1.0 - tar_contacts) * forces
    r = r + tar_contacts * forces
    return contact_weights * r


def compute_done(time, ep_len, root_rot, body_pos, tar_root_rot, tar_body_pos, contact_force, contact_body_ids,
                 termination_heights, pose_termination, pose_termination_dist, enable_early_termination, track_root,
                 root_pos_termination_dist, root_rot_termination_angle):
    """mgdm_dm_util.py:334-402 -> int32 [n]"""
    n = time.shape[0]
    done = torch.full((n,), DONE_NULL, dtype=torch.int32)
    done[time >= ep_len] = DONE_TIME
    if enable_early_termination:
        failed = torch.zeros(n, dtype=torch.bool)
        if len(contact_body_ids) > 0:
            masked = contact_force.clone()
            masked[:, contact_body_ids, :] = 0
            fall_contact = torch.any(torch.any(torch.abs(masked) > 0.1, dim=-1), dim=-1)
            fall_height = body_pos[..., 2] < termination_heights
            fall_height[:, contact_body_ids] = False
            failed = failed | (fall_contact & torch.any(fall_height, dim=-1))
        if pose_termination:
            rp, trp = body_pos[..., 0:1, :], tar_body_pos[..., 0:1, :]
            d = (tar_body_pos[..., 1:, :] - trp) - (body_pos[..., 1:, :] - rp)
            pose_fail = torch.any(torch.sum(d * d, dim=-1) > pose_termination_dist * pose_termination_dist, dim=-1)
            if track_root:
                rd = rp - trp
                pose_fail = pose_fail | (torch.sum(rd * rd, dim=-1).squeeze(-1) >
                                         root_pos_termination_dist * root_pos_termination_dist)
                pose_fail = pose_fail | (torch.abs(quat_diff_angle(root_rot, tar_root_rot)) > root_rot_termination_angle)
            failed = failed | pose_fail
        failed = failed & (time > 1e-5)
        done[failed] = DONE_FAIL
    return done


def compute_tracking_error(root_pos, root_rot, body_rot, body_pos, tar_root_pos, tar_root_rot, tar_body_rot, tar_body_pos, root_vel,
                           root_ang_vel, dof_vel, tar_root_vel, tar_root_ang_vel, tar_dof_vel):
    """mgdm_dm_util.py:520-553 -> [n,7]"""
    pose_err = torch.mean(torch.abs(quat_diff_angle(body_rot, tar_body_rot)), dim=-1)
    root_pos_l2 = torch.linalg.vector_norm(tar_root_pos - root_pos, dim=-1)
    d = (tar_body_pos - tar_root_pos.unsqueeze(-2)) - (body_pos - root_pos.unsqueeze(-2))
    body_pos_err = torch.mean(torch.linalg.vector_norm(d, dim=-1), dim=-1)
    root_rot_err = torch.abs(quat_diff_angle(root_rot, tar_root_rot))
    dof_vel_err = torch.mean(torch.abs(tar_dof_vel - dof_vel), dim=-1)
    root_vel_err = torch.mean(torch.abs(tar_root_vel - root_vel), dim=-1)
    root_ang_vel_err = torch.mean(torch.abs(tar_root_ang_vel - root_ang_vel), dim=-1)
    return torch.stack([root_pos_l2, root_rot_err, body_pos_err, pose_err, dof_vel_err, root_vel_err, root_ang_vel_err], dim=-1)


# ------------------------------------------------------------------------------------------------
# device-RSI restatement: Philox4x32-10 + inverse-CDF (our own sampler; validated bit-exact)
# ------------------------------------------------------------------------------------------------
def philox4x32_10(ctr, key):
    """ctr: 4 python ints (uint32), key: 2 -> 4 uint32 (Salmon et al. 2011, 10 rounds)."""
    M0, M1, W0, W1, MASK = 0xD2511F53, 0xCD9E8D57, 0x9E3779B9, 0xBB67AE85, 0xFFFFFFFF
    c0, c1, c2, c3 = ctr
    k0, k1 = key
    for _ in range(10):
        p0, p1 = M0 * c0, M1 * c2
        c0, c1, c2, c3 = ((p1 >> 32) ^ c1 ^ k0) & MASK, p1 & MASK, ((p0 >> 32) ^ c3 ^ k1) & MASK, p0 & MASK
        k0, k1 = (k0 + W0) & MASK, (k1 + W1) & MASK
    return c0, c1, c2, c3


def u01(x):
    return np.float32(x >> 8) * np.float32(1.0 / 16777216.0)


def sample_cdf(cdf, u):
    """first m with cdf[m] > u * cdf[M-1] (fp32), cdf a float32 numpy array"""
    target = np.float32(u) * cdf[-1]
    lo, hi = 0, len(cdf) - 1
    while lo < hi:
        mid = (lo + hi) >> 1
        if cdf[mid] > target:
            hi = mid
        else:
            lo = mid + 1
    return lo


def rsi_draws(env_id, seed, step, cdf, terrains_per_motion, lengths, rand_reset=True):
    """The draws parc_b200_env_reset_sampled makes for one env -> (motion_id, terrain_id, time, nx, ny)."""
    key = (seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF)
    ctr = (env_id & 0xFFFFFFFF, (env_id >> 32) & 0xFFFFFFFF, step & 0xFFFFFFFF, (step >> 32) & 0xFFFFFFFF)
    r0 = philox4x32_10(ctr, key)
    mid = sample_cdf(cdf, u01(r0[0]))
    tid = r0[1] % terrains_per_motion
    t = np.float32(u01(r0[2]) * np.float32(lengths[mid])) if rand_reset else np.float32(0.0)
    r1 = philox4x32_10((ctr[0], ctr[1], ctr[2], ctr[3] ^ 0x80000000), key)
    nx = np.float32(u01(r1[0]) * np.float32(2.0) - np.float32(1.0))
    ny = np.float32(u01(r1[1]) * np.float32(2.0) - np.float32(1.0))
    return mid, tid, t, nx, ny


# ------------------------------------------------------------------------------------------------
# the tracking env: IGEnv._post_physics_step order (PARC/motion_tracker/envs/ig_env.py:368-377)
# with IGParkourEnv's orchestration (ig_parkour_env.py:842-1044) and DeepMimicEnv (dm_env.py)
# ------------------------------------------------------------------------------------------------
class EnvConfig:
    """The knobs of data/configs/tracker_config/dm_env_default.yaml that change arithmetic."""

    def __init__(self, model: CharModel, **kw):
        J = model.J
        self.control_freq = 30
        self.episode_length = 10.0
        self.tar_obs_steps = [1, 2, 3, 10, 20, 30]
        self.key_bodies = ["right_hand", "left_hand", "right_foot", "left_foot"]
        self.global_obs = False
        self.root_height_obs = False
        self.enable_tar_obs = True
        self.use_contact_info = True
        self.track_root = True
        self.track_root_h = True
        self.enable_early_termination = True
        self.pose_termination = True
        self.contact_bodies = []                      # names; fall test disabled when empty
        self.never_done = False
        self.termination_height = 0.15
        self.root_pos_termination_dist = 0.6
        self.root_rot_termination_angle = 1.309
        self.pose_termination_dist = [0.7, 1.0, 0.7, 0.7, 0.7, 0.7, 0.7, 0.7, 1.0, 1.2, 10.0, 1.0, 1.2, 10.0]
        self.joint_err_w = [1.0, 0.6, 0.6, 0.4, 0.0, 0.6, 0.4, 0.0, 1.0, 0.6, 0.4, 1.0, 0.6, 0.4]
        self.contact_weights = [5.0] * J
        self.pose_w, self.vel_w, self.root_pos_w, self.root_vel_w, self.key_pos_w = 0.5, 0.1, 0.15, 0.1, 0.15
        self.min_obs_h, self.max_obs_h = -3.0, 3.0
        self.min_motion_weight = 0.01
        self.rand_root_pos_offset_scale = 0.0
        self.terrains_per_motion = 1
        for k, v in kw.items():
            assert hasattr(self, k), "unknown env config key " + k
            setattr(self, k, v)


def make_dof_err_w(model, joint_err_w):
    """IGParkourEnv._parse_joint_err_weights, ig_parkour_env.py:1134-1153"""
    w = torch.zeros(model.D)
    for j in range(1, model.J):
        n = 3 if model.jtype[j] == SPHERICAL else (1 if model.jtype[j] == HINGE else 0)
        w[model.dof_idx[j]:model.dof_idx[j] + n] = joint_err_w[j - 1]
    return w


class TrackingEnv:
    """State layout = the tensors IGParkourEnv hands DeepMimicEnv (mgdm_dm_util.py:37-84)."""

    def __init__(self, model, lib, terrain, dm_motion_offsets, cfg: EnvConfig, num_envs, env_offsets, ray_xy_points):
        N, J, D = num_envs, model.J, model.D
        self.model, self.lib, self.terrain, self.cfg, self.N = model, lib, terrain, cfg, N
        self.dm_motion_offsets = torch.as_tensor(dm_motion_offsets, dtype=torch.float32)   # [M, Tpm, 2]
        self.env_offsets = torch.as_tensor(env_offsets, dtype=torch.float32)
        self.ray_xy_points = torch.as_tensor(ray_xy_points, dtype=torch.float32)
        self.key_body_ids = [model.body_id(b) for b in cfg.key_bodies]
        self.contact_body_ids = [model.body_id(b) for b in cfg.contact_bodies]
        self.timestep = 1.0 / cfg.control_freq
        self.tar_obs_steps = torch.tensor(cfg.tar_obs_steps, dtype=torch.int32)
        tw = cfg.pose_w + cfg.vel_w + cfg.root_pos_w + cfg.root_vel_w + cfg.key_pos_w
        self.w = [cfg.pose_w / tw, cfg.vel_w / tw, cfg.root_pos_w / tw, cfg.root_vel_w / tw, cfg.key_pos_w / tw]
        self.joint_err_w = torch.tensor(cfg.joint_err_w, dtype=torch.float32)
        self.dof_err_w = make_dof_err_w(model, self.joint_err_w)
        self.contact_weights = torch.tensor(cfg.contact_weights, dtype=torch.float32)
        self.pose_termination_dist = torch.tensor(cfg.pose_termination_dist, dtype=torch.float32)
        z = lambda *s: torch.zeros(*s, dtype=torch.float32)
        self.char_root_pos, self.char_root_rot, self.char_root_vel, self.char_root_ang_vel = z(N, 3), z(N, 4), z(N, 3), z(N, 3)
        self.char_root_rot[:, 3] = 1
        self.char_dof_pos, self.char_dof_vel = z(N, D), z(N, D)
        self.char_contact_forces, self.char_rigid_body_pos = z(N, J, 3), z(N, J, 3)
        self.char_rigid_body_vel, self.char_rigid_body_ang_vel = z(N, J, 3), z(N, J, 3)
        self.ref_root_pos, self.ref_root_rot, self.ref_root_vel, self.ref_root_ang_vel = z(N, 3), z(N, 4), z(N, 3), z(N, 3)
        self.ref_body_pos, self.ref_joint_rot = z(N, J, 3), z(N, J - 1, 4)
        self.ref_dof_pos, self.ref_dof_vel, self.ref_contacts = z(N, D), z(N, D), z(N, J)
        self.reward_buf, self.time_buf = z(N), z(N)
        self.reward_terms = z(N, 7)
        self.done_buf = torch.zeros(N, dtype=torch.int32)
        self.timestep_buf = torch.zeros(N, dtype=torch.int32)
        self.actors_need_reset = torch.zeros(N, 1, dtype=torch.bool)
        self.ray_hfs = z(N, self.ray_xy_points.shape[0])
        self.motion_ids = torch.zeros(N, dtype=torch.int64)
        self.motion_terrain_ids = torch.zeros(N, dtype=torch.int64)
        self.motion_time_offsets = z(N)
        self.fail_rates = torch.ones(lib.num_motions(), dtype=torch.float32)
        self.ema_weight = 0.01
        self.obs_buf = None

    # ---- pieces --------------------------------------------------------------------------------
    def _terrain_offset(self, ids=None):
        """_move_to_motion_terrain, dm_env.py:554-565"""
        if ids is None:
            return self.dm_motion_offsets[self.motion_ids, self.motion_terrain_ids] - self.env_offsets[:, 0:2]
        return self.dm_motion_offsets[self.motion_ids[ids], self.motion_terrain_ids[ids]] - self.env_offsets[ids, 0:2]

    def refresh_heights(self):
        """ig_parkour_env.py:518-533 + mgdm_dm_util.py:128-145"""
        xyz = self.char_root_pos + self.env_offsets
        heading = calc_heading(self.char_root_rot)
        pts = rotate_2d_vec(self.ray_xy_points.unsqueeze(0), heading.unsqueeze(-1)) + xyz[:, None, 0:2]
        h = self.terrain.height_at(pts) - xyz[:, 2:3]
        self.ray_hfs[:] = torch.clamp(h, min=self.cfg.min_obs_h, max=self.cfg.max_obs_h)

    def update_time(self):
        """ig_env.py:391-394"""
        self.timestep_buf += 1
        self.time_buf[:] = self.timestep * self.timestep_buf

    def update_ref_motion(self):
        """dm_env.py:523-545"""
        out = self.lib.calc_motion_frame(self.motion_ids, self.time_buf + self.motion_time_offsets)
        root_pos = out[0].clone()
        root_pos[:, 0:2] = root_pos[:, 0:2] + self._terrain_offset()
        self.ref_root_pos[:], self.ref_root_rot[:], self.ref_root_vel[:], self.ref_root_ang_vel[:] = root_pos, out[1], out[2], out[3]
        self.ref_joint_rot[:], self.ref_dof_vel[:], self.ref_contacts[:] = out[4], out[5], out[6]
        self.ref_body_pos[:] = self.model.forward_kinematics(self.ref_root_pos, self.ref_root_rot, self.ref_joint_rot)[0]
        self.ref_dof_pos[:] = self.model.rot_to_dof(out[4])

    def tar_obs_data(self, ids=None):
        """dm_env.py:594-626 + fetch_tar_obs_data mgdm_dm_util.py:221-244"""
        sel = slice(None) if ids is None else ids
        mids = self.motion_ids[sel]
        t = (self.time_buf[sel] + self.motion_time_offsets[sel]).unsqueeze(-1) + self.timestep * self.tar_obs_steps
        n, S = t.shape
        out = self.lib.calc_motion_frame(mids.unsqueeze(-1).expand(n, S).flatten(), t.flatten())
        root_pos = out[0].reshape(n, S, 3).clone()
        root_pos[..., 0:2] = root_pos[..., 0:2] + self._terrain_offset(ids).unsqueeze(1)
        root_rot = out[1].reshape(n, S, 4)
        joint_rot = out[4].reshape(n, S, self.model.J - 1, 4)
        contacts = out[6].reshape(n, S, -1)
        body_pos = self.model.forward_kinematics(root_pos, root_rot, joint_rot)[0]
        return root_pos, root_rot, joint_rot, body_pos[..., self.key_body_ids, :], contacts

    def compute_obs(self, ids=None):
        """IGParkourEnv._compute_obs, ig_parkour_env.py:842-965"""
        sel = slice(None) if ids is None else ids
        cfg = self.cfg
        root_pos, root_rot = self.char_root_pos[sel], self.char_root_rot[sel]
        joint_rot = self.model.dof_to_rot(self.char_dof_pos[sel])
        body_pos = self.model.forward_kinematics(root_pos, root_rot, joint_rot)[0]
        key_pos = body_pos[..., self.key_body_ids, :]
        parts = [compute_char_obs(root_pos, root_rot, self.char_root_vel[sel], self.char_root_ang_vel[sel], joint_rot,
                                  self.char_dof_vel[sel], key_pos, cfg.global_obs, cfg.root_height_obs)]
        n = root_pos.shape[0]
        if cfg.enable_tar_obs:
            trp, trr, tjr, tkp, tcon = self.tar_obs_data(ids)
            parts.append(compute_tar_obs(root_pos, root_rot, trp, trr, tjr, tkp, cfg.global_obs, False).reshape(n, -1))
        if cfg.use_contact_info:
            if cfg.enable_tar_obs:
                parts.append(tcon.reshape(n, -1))
            parts.append((torch.norm(self.char_contact_forces[sel], dim=-1) > 1e-5).float())   # :655-662
        parts.append(self.ray_hfs[sel])
        return torch.cat(parts, dim=-1)

    def update_reward(self):
        """IGParkourEnv._update_reward, ig_parkour_env.py:984-1044"""
        joint_rot = self.model.dof_to_rot(self.char_dof_pos)
        kb = self.key_body_ids
        r5 = compute_deepmimic_reward(self.char_root_pos, self.char_root_rot, self.char_root_vel, self.char_root_ang_vel,
                                      joint_rot, self.char_dof_vel, self.char_rigid_body_pos[:, kb], self.ref_root_pos,
                                      self.ref_root_rot, self.ref_root_vel, self.ref_root_ang_vel, self.ref_joint_rot,
                                      self.ref_dof_vel, self.ref_body_pos[:, kb], self.joint_err_w, self.dof_err_w,
                                      self.cfg.track_root_h, self.cfg.track_root)
        w = self.w
        total = w[0] * r5[:, 0] + w[1] * r5[:, 1] + w[2] * r5[:, 2] + w[3] * r5[:, 3] + w[4] * r5[:, 4]
        pen = torch.zeros_like(total)
        if self.cfg.use_contact_info:
            pen = torch.mean(compute_contact_reward(self.ref_contacts, self.char_contact_forces, self.contact_weights), dim=-1)
            total = total + pen
        self.reward_buf[:] = total
        self.reward_terms[:, 0:5] = r5
        self.reward_terms[:, 5] = pen
        self.reward_terms[:, 6] = total

    def update_done(self):
        """_update_done ig_parkour_env.py:967-982; RefCharEnv.update_done mgdm_dm_util.py:147-172;
        DeepMimicEnv.update_done dm_env.py:628-665 (fail-rate EMA in ascending env id)."""
        cfg = self.cfg
        gpos = self.char_rigid_body_pos[..., 0:2] + self.env_offsets[:, None, 0:2]
        term_h = self.terrain.height_at(gpos) + torch.tensor(cfg.termination_height, dtype=torch.float32)
        self.done_buf[:] = compute_done(self.time_buf, cfg.episode_length, self.char_root_rot, self.char_rigid_body_pos,
                                        self.ref_root_rot, self.ref_body_pos, self.char_contact_forces,
                                        self.contact_body_ids, term_h, cfg.pose_termination, self.pose_termination_dist,
                                        cfg.enable_early_termination, cfg.track_root, cfg.root_pos_termination_dist,
                                        cfg.root_rot_termination_angle)
        t = self.time_buf + self.motion_time_offsets
        motion_end = (t >= self.lib.lengths[self.motion_ids]) & (self.lib.loop_modes[self.motion_ids] != WRAP)
        finished = torch.nonzero((self.done_buf != DONE_NULL) | motion_end).flatten().tolist()
        keep = torch.tensor(1.0 - self.ema_weight, dtype=torch.float32)     # python double -> fp32 at the multiply
        w = torch.tensor(self.ema_weight, dtype=torch.float32)
        rates = self.fail_rates
        flags = self.done_buf.tolist()
        mids = self.motion_ids.tolist()
        for e in finished:
            m = mids[e]
            rates[m] = rates[m] * keep + w if flags[e] == DONE_FAIL else rates[m] * keep
        self.done_buf[motion_end] = DONE_FAIL
        if cfg.never_done:
            self.done_buf[:] = DONE_NULL

    # ---- whole step / reset ---------------------------------------------------------------------
    def post_physics_step(self):
        """IGEnv._post_physics_step (ig_env.py:368-377) minus the PhysX refresh."""
        self.refresh_heights()
        self.update_time()
        self.update_ref_motion()
        self.obs_buf = self.compute_obs()
        self.update_reward()
        self.update_done()

    def sampling_weights(self):
        """dm_env.py:487-490"""
        return torch.clamp(self.fail_rates, min=self.cfg.min_motion_weight) * self.lib.motion_weights

    def reset(self, env_ids, motion_ids, terrain_ids, motion_times, xy_noise, root_pos_offset=None, root_rot_offset=None):
        """DeepMimicEnv.reset (dm_env.py:567-592) given the random draws; IGParkourEnv.reset then
        refreshes ALL heights and recomputes the observation rows of env_ids (ig_parkour_env.py:809-829)."""
        if len(env_ids) == 0:
            return
        ids = env_ids
        self.motion_ids[ids], self.motion_terrain_ids[ids], self.motion_time_offsets[ids] = motion_ids, terrain_ids, motion_times
        out = self.lib.calc_motion_frame(motion_ids, motion_times)
        root_pos = out[0].clone()
        root_pos[:, 0:2] = root_pos[:, 0:2] + self._terrain_offset(ids)
        self.ref_contacts[ids], self.ref_root_pos[ids], self.ref_root_rot[ids] = out[6], root_pos, out[1]
        self.ref_root_vel[ids], self.ref_root_ang_vel[ids] = out[2], out[3]
        self.ref_joint_rot[ids], self.ref_dof_vel[ids] = out[4], out[5]
        self.ref_dof_pos[ids] = self.model.rot_to_dof(out[4])
        # _char_state_init_from_ref, mgdm_dm_util.py:89-100
        self.char_rigid_body_vel[ids] = 0.0
        self.char_rigid_body_ang_vel[ids] = 0.0
        self.char_root_pos[ids], self.char_root_rot[ids] = self.ref_root_pos[ids], self.ref_root_rot[ids]
        self.char_root_vel[ids], self.char_root_ang_vel[ids] = self.ref_root_vel[ids], self.ref_root_ang_vel[ids]
        self.char_dof_pos[ids], self.char_dof_vel[ids] = self.ref_dof_pos[ids], self.ref_dof_vel[ids]
        # add_noise_to_char_state, :102-106
        noise = self.cfg.rand_root_pos_offset_scale * torch.as_tensor(xy_noise, dtype=torch.float32)
        cp = self.char_root_pos[ids]
        cp[:, 0:2] += noise
        self.char_root_pos[ids] = cp
        if root_pos_offset is not None:                                    # :108-113
            self.char_root_pos[ids] += root_pos_offset[ids]
        if root_rot_offset is not None:
            self.char_root_rot[ids] = quat_multiply(root_rot_offset[ids], self.char_root_rot[ids])
        self.timestep_buf[ids] = 0
        self.time_buf[ids] = 0.0
        self.done_buf[ids] = DONE_NULL
        self.actors_need_reset[ids, 0] = True

    def refresh_obs_after_reset(self, env_ids):
        self.refresh_heights()
        if self.obs_buf is not None and len(env_ids) > 0:
            self.obs_buf[env_ids] = self.compute_obs(env_ids)
