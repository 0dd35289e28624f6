"""The kinematic motion optimizer's loss on the GPU (reference:
PARC/motion_synthesis/motion_opt/motion_optimization.py). `motion_terrain_contact_loss_localized`
keeps the reference's signature and return value `(loss, {LossType: float})`; the loss tensor carries
autograd history to `tgt_root_pos`, `tgt_root_rot`, `tgt_joint_dof`, so the reference's Adam loop
(:772-797) runs unchanged on top of it. `motion_loss_batched` evaluates many motions in one call
(BASELINE config 4: 1,024 motions x 300 frames).

Everything numeric happens in `parc_b200_motion_loss` (include/parc_b200_opt.h): value and gradient in
one pass; `backward` only scales the stored gradients."""
import enum
from dataclasses import dataclass
from typing import List

import numpy as np
import torch

from . import _lib
from .kin_char_model import GeomType, KinCharModel


class LossType(enum.Enum):      # motion_optimization.py:15-25
    ROOT_POS_LOSS = 0
    ROOT_ROT_LOSS = 1
    JOINT_ROT_LOSS = 2
    SMOOTHNESS_LOSS = 3
    PENETRATION_LOSS = 4
    CONTACT_LOSS = 5
    SLIDING_LOSS = 6
    BODY_CONSTRAINT_LOSS = 7
    JERK_LOSS = 8
    LOOPING_LOSS = 9


@dataclass
class BodyConstraint:           # motion_optimization.py:27-31
    start_frame_idx = 0
    end_frame_idx = 0
    constraint_point = None     # 3D torch vector


SOLE_POINTS = 18                # the box (foot) term reads the body's first 18 sample points (:600)


class PackedBodyConstraints:
    """The reference's per-body Python lists of BodyConstraint (one list of lists per motion), flattened into the device
    arrays parc_opt_constraints_t points at (include/parc_b200_opt.h). Pack once per optimisation, reuse every iteration."""

    def __init__(self, per_motion: list, char_model: KinCharModel, device):
        J = char_model.get_num_joints()
        rows, starts = [], [0]
        for body_constraints in per_motion:
            if body_constraints is not None:
                assert len(body_constraints) == J, "body_constraints holds one list per body"
                for b, lst in enumerate(body_constraints):
                    for c in lst:
                        pt = c.constraint_point.detach().reshape(3).cpu().numpy().astype(np.float32)
                        rows.append((b, int(c.start_frame_idx), int(c.end_frame_idx), pt))
            starts.append(len(rows))
        n = len(rows)
        B = len(per_motion)
        # one int32 buffer = one host-to-device copy: motion_start [B+1] | body [n] | start [n] | end [n] | point bits [n,3]
        buf = np.zeros(B + 1 + 6 * max(n, 1), dtype=np.int32)
        buf[:B + 1] = starts
        o = B + 1
        if n:
            buf[o:o + n] = [r[0] for r in rows]
            buf[o + n:o + 2 * n] = [r[1] for r in rows]
            buf[o + 2 * n:o + 3 * n] = [r[2] for r in rows]
            buf[o + 3 * n:o + 6 * n] = np.stack([r[3] for r in rows]).reshape(-1).view(np.int32)
        self.num, self.num_motions = n, B
        self.buffer = torch.from_numpy(buf).to(device)
        st = _lib.OptConstraintsStruct()
        st.num = n
        base = self.buffer.data_ptr()
        st.motion_start = base
        st.body = base + 4 * o
        st.start_frame = base + 4 * (o + n)
        st.end_frame = base + 4 * (o + 2 * n)
        st.point = base + 4 * (o + 3 * n)
        for b in range(J):
            geoms = char_model.get_geoms(b)
            kind = _lib.GEOM_NONE
            if geoms:
                g = geoms[0]
                if g._shape_type == GeomType.SPHERE:        # :590-592
                    kind = _lib.GEOM_SPHERE
                    st.geom_radius[b] = float(g._dims.reshape(-1)[0])
                    for k in range(3):
                        st.geom_offset[b][k] = float(g._offset[k])
                elif g._shape_type == GeomType.BOX:         # :598-599
                    kind = _lib.GEOM_BOX
                    st.geom_radius[b] = float(torch.norm(g._dims.float()) * 1.25)
            st.geom_kind[b] = kind
        st.sole_points = SOLE_POINTS
        self.struct = st


def compute_approx_body_constraints(root_pos, root_rot, joint_rot, contacts, char_model: KinCharModel, terrain):
    """motion_optimization.py:70-205. For each foot and hand: every run of consecutive frames whose contact label exceeds
    0.9 becomes one BodyConstraint whose point is the body's mean position over the run (feet: the centre of the box
    geom), pulled onto the terrain surface by 1,000 SGD steps (lr 0.01) on the squared heightfield distance.
    Returns one list of BodyConstraint per body."""
    def extract_consecutive_trues(flags):                     # :74-104
        idx = torch.nonzero(flags.flatten(), as_tuple=True)[0]
        if len(idx) == 0:
            return []
        breaks = [0] + (torch.nonzero(idx[1:] - idx[:-1] > 1, as_tuple=True)[0] + 1).tolist()
        groups = [idx[breaks[i]:breaks[i + 1]].clone() for i in range(len(breaks) - 1)]
        if breaks[-1] < idx.shape[0] - 1:                     # upstream drops a trailing run of a single frame
            groups.append(idx[breaks[-1]:].clone())
        return groups

    body_pos, body_rot = char_model.forward_kinematics(root_pos, root_rot, joint_rot)
    from .geom_util import _quat_rotate
    dev = body_pos.device
    ids = [char_model.get_body_id(n) for n in ("left_foot", "right_foot", "left_hand", "right_hand")]
    points, frame_groups = [], []
    for k, b in enumerate(ids):
        pos = body_pos[:, b]
        if k < 2:                                             # feet: centre of the box geom (:145-152)
            off = char_model.get_geoms(b)[0]._offset.to(dev)
            pos = pos + _quat_rotate(body_rot[:, b], off.unsqueeze(0).expand(pos.shape[0], -1))     # torch_util.quat_rotate, torch_util.py:62-67
        groups = extract_consecutive_trues(contacts[:, b] > 0.9)          # :106-108
        frame_groups.append(groups)
        points.extend(torch.mean(pos[g], dim=0) for g in groups)         # :115
    if points:
        from . import terrain as _terrain
        # every point descends independently (:123-143), so all of them share the 1,000 steps here
        pts = torch.stack(points).detach().clone().unsqueeze(0).contiguous().requires_grad_(True)
        base_z = terrain.hf.min().item() - 10.0
        opt = torch.optim.SGD([pts], lr=0.01)
        hf, mp = terrain.hf.unsqueeze(0), terrain.min_point.unsqueeze(0)
        for _ in range(1000):
            opt.zero_grad()
            sd = _terrain.points_hf_sdf(pts, hf, mp, terrain.dxdy, base_z=base_z, inverted=False)
            torch.sum(torch.square(sd)).backward()
            opt.step()
        pts = pts.detach()[0]
    body_constraints = [[] for _ in range(char_model.get_num_joints())]
    k = 0
    for b, groups in zip(ids, frame_groups):
        out = []
        for g in groups:                                      # :185-193
            c = BodyConstraint()
            c.start_frame_idx, c.end_frame_idx = g[0].item(), g[-1].item()
            c.constraint_point = pts[k].clone()
            k += 1
            out.append(c)
        body_constraints[b] = out
    return body_constraints


def build_opt_config(char_model: KinCharModel, body_points: List[torch.Tensor], num_contact_cols: int, w_root_pos, w_root_rot,
                     w_joint_rot, w_smoothness, w_penetration, w_contact, w_sliding, w_body_constraints, w_jerk, max_jerk,
                     base_z=-10.0, whole_terrain=False) -> _lib.OptConfigStruct:
    J = char_model.get_num_joints()
    assert len(body_points) == J, "body_points needs one tensor per body"
    cfg = _lib.OptConfigStruct()
    for k, v in enumerate((w_root_pos, w_root_rot, w_joint_rot, w_smoothness, w_penetration, w_contact, w_sliding, w_body_constraints,
                           w_jerk)):
        cfg.w[k] = float(v)
    cfg.max_jerk, cfg.base_z = float(max_jerk), float(base_z)
    start = 0
    for b in range(J):
        cfg.point_start[b] = start
        start += int(body_points[b].shape[0])
    for b in range(J, _lib.MAX_BODIES + 1):
        cfg.point_start[b] = start
    cfg.num_points = start
    for b in range(_lib.MAX_BODIES):
        cfg.contact_col[b] = -1
    for b in range(J):          # :575-579: only bodies listed as contact bodies carry a contact label
        name = char_model.get_body_name(b)
        if char_model.is_contact_body(name):
            cfg.contact_col[b] = int(char_model.get_contact_body_id(name))
    cfg.num_contact_cols = int(num_contact_cols)
    cfg.whole_terrain = int(bool(whole_terrain))
    return cfg


class _MotionLossFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, tgt_root_pos, tgt_root_rot, tgt_joint_dof, char_model, cfg, consts, terrain_args, points, packed=None):
        (src_root_pos, src_root_rot_quat, src_joint_rot, src_body_vels, src_body_rot_vels, contacts) = consts
        hf, hf_stride, X, Y, min_point, mp_stride, dx, dy = terrain_args
        B, T = tgt_root_pos.shape[0], tgt_root_pos.shape[1]
        dev = tgt_root_pos.device
        f32 = torch.float32
        c = lambda t: t.detach().to(device=dev, dtype=f32).contiguous()
        p = [c(tgt_root_pos), c(tgt_root_rot), c(tgt_joint_dof)]
        need_grad = any(ctx.needs_input_grad[:3])
        L = _lib.lib()
        km = char_model._cstruct()
        ws_bytes = int(L.parc_b200_motion_loss_workspace_bytes(km, B, T))
        ws = torch.empty(ws_bytes // 4 + 1, dtype=f32, device=dev)
        terms = torch.empty(B, _lib.PARC_NUM_LOSS_TERMS, dtype=f32, device=dev)
        loss = torch.empty(B, dtype=f32, device=dev)
        grads = [torch.empty_like(x) for x in p] if need_grad else [None, None, None]
        keep = [c(x) for x in (src_root_pos, src_root_rot_quat, src_joint_rot, src_body_vels, src_body_rot_vels, contacts)]
        if packed is not None:
            assert packed.num_motions == B and packed.buffer.device == dev, "constraints packed for another batch / device"
        _lib.check(L.parc_b200_motion_loss_constrained(km, cfg, packed.struct if packed is not None else None, B, T, _lib.ptr(p[0]), _lib.ptr(p[1]), _lib.ptr(p[2]), *[_lib.ptr(x) for x in keep],
                                                       _lib.ptr(hf), hf_stride, X, Y, _lib.ptr(min_point), mp_stride, dx, dy,
                                                       _lib.ptr(points), _lib.ptr(ws), ws_bytes, _lib.ptr(terms), _lib.ptr(loss),
                                                       _lib.ptr(grads[0]), _lib.ptr(grads[1]), _lib.ptr(grads[2]), _lib.stream_ptr(dev)))
        ctx.save_for_backward(*[g for g in grads if g is not None])
        ctx.has_grads = need_grad
        ctx.mark_non_differentiable(terms)
        return loss, terms

    @staticmethod
    def backward(ctx, g_loss, _g_terms):
        if not ctx.has_grads:
            return (None,) * 9
        g0, g1, g2 = ctx.saved_tensors
        s = g_loss.reshape(-1, 1, 1)
        return s * g0, s * g1, s * g2, None, None, None, None, None, None


def motion_loss_batched(tgt_root_pos, tgt_root_rot, tgt_joint_dof, src_root_pos, src_root_rot_quat, src_joint_rot, src_body_vels,
                        src_body_rot_vels, contacts, hf, min_point, dxdy, body_points: List[torch.Tensor], char_model: KinCharModel,
                        w_root_pos, w_root_rot, w_joint_rot, w_smoothness, w_penetration, w_contact, w_sliding, w_body_constraints,
                        w_jerk, max_jerk, body_constraints=None, localized=True):
    """B independent motions of T frames: all motion tensors carry a leading [B, T(-1), ...]; `hf` is
    [B, X, Y] with `min_point` [B, 2], or one shared [X, Y] / [2]. `body_constraints`: None, a PackedBodyConstraints, or a
    list with one entry per motion (None or the reference's list of BodyConstraint lists per body). `localized`: per-frame
    heightfield windows (motion_terrain_contact_loss_localized) or every cell of the terrain (motion_terrain_contact_loss).
    -> (loss [B], terms [B, 9]) with autograd to the three tgt_* tensors."""
    dev = tgt_root_pos.device
    assert tgt_root_pos.dim() == 3 and tgt_root_rot.shape == tgt_root_pos.shape and tgt_joint_dof.dim() == 3
    B, T = tgt_root_pos.shape[:2]
    J = char_model.get_num_joints()
    assert src_joint_rot.shape == (B, T, J - 1, 4) and src_body_vels.shape == (B, T - 1, J, 3) and src_body_rot_vels.shape == (B, T - 1, J)
    assert contacts.shape[:2] == (B, T)
    cfg = build_opt_config(char_model, body_points, contacts.shape[-1], w_root_pos, w_root_rot, w_joint_rot, w_smoothness, w_penetration,
                           w_contact, w_sliding, w_body_constraints, w_jerk, max_jerk, whole_terrain=not localized)
    hf = hf.to(device=dev, dtype=torch.float32).contiguous()
    min_point = min_point.to(device=dev, dtype=torch.float32).contiguous()
    if hf.dim() == 2:
        X, Y, hf_stride, mp_stride = hf.shape[0], hf.shape[1], 0, 0
        assert min_point.shape == (2,)
    else:
        assert hf.shape[0] == B and min_point.shape == (B, 2)
        X, Y, hf_stride, mp_stride = hf.shape[1], hf.shape[2], hf.shape[1] * hf.shape[2], 2
    points = torch.cat([p.to(device=dev, dtype=torch.float32) for p in body_points], dim=0).contiguous()
    if body_constraints is not None and not isinstance(body_constraints, PackedBodyConstraints):
        assert len(body_constraints) == B, "one body_constraints entry per motion"
        body_constraints = PackedBodyConstraints(body_constraints, char_model, dev)
    return _MotionLossFn.apply(tgt_root_pos, tgt_root_rot, tgt_joint_dof, char_model, cfg,
                               (src_root_pos, src_root_rot_quat, src_joint_rot, src_body_vels, src_body_rot_vels, contacts),
                               (hf, hf_stride, X, Y, min_point, mp_stride, float(dxdy[0]), float(dxdy[1])), points, body_constraints)


def motion_terrain_contact_loss_localized(tgt_root_pos, tgt_root_rot, tgt_joint_dof, src_root_pos, src_root_rot_quat, src_joint_rot,
                                          src_body_vels, src_body_rot_vels, contacts, terrain, body_points: List[torch.Tensor],
                                          char_model: KinCharModel, w_root_pos: float, w_root_rot: float, w_joint_rot: float,
                                          w_smoothness: float, w_penetration: float, w_contact: float, w_sliding: float,
                                          w_body_constraints: float, w_jerk: float, body_constraints: list, max_jerk: float,
                                          _localized=True):
    """motion_optimization.py:426-656, one motion of T frames. Returns (loss, {LossType: float}). `body_constraints` is the
    reference's list (one list of BodyConstraint per body), None, or a PackedBodyConstraints made once for the whole run."""
    if body_constraints is not None and not isinstance(body_constraints, PackedBodyConstraints):
        body_constraints = PackedBodyConstraints([body_constraints], char_model, tgt_root_pos.device)
    u = lambda t: t.unsqueeze(0)
    loss, terms = motion_loss_batched(u(tgt_root_pos), u(tgt_root_rot), u(tgt_joint_dof), u(src_root_pos), u(src_root_rot_quat),
                                      u(src_joint_rot), u(src_body_vels), u(src_body_rot_vels), u(contacts), terrain.hf,
                                      terrain.min_point, terrain.dxdy, body_points, char_model, w_root_pos, w_root_rot, w_joint_rot,
                                      w_smoothness, w_penetration, w_contact, w_sliding, w_body_constraints, w_jerk, max_jerk,
                                      body_constraints=body_constraints, localized=_localized)
    vals = terms[0].tolist()       # the reference also reads every term back with .item() (:626-643)
    losses = {t: vals[t.value] for t in list(LossType)[:9]}
    return loss[0], losses


def motion_terrain_contact_loss(tgt_root_pos, tgt_root_rot, tgt_joint_dof, src_root_pos, src_root_rot_quat, src_joint_rot,
                                src_body_vels, src_body_rot_vels, contacts, terrain, body_points: List[torch.Tensor],
                                char_model: KinCharModel, w_root_pos: float, w_root_rot: float, w_joint_rot: float,
                                w_smoothness: float, w_penetration: float, w_contact: float, w_sliding: float,
                                w_body_constraints: float, w_jerk: float, body_constraints: list, max_jerk: float):
    """motion_optimization.py:207-423: the same loss with every body point measured against EVERY cell of the terrain
    instead of a per-frame window (the two differ when the nearest column lies outside the window)."""
    return motion_terrain_contact_loss_localized(tgt_root_pos, tgt_root_rot, tgt_joint_dof, src_root_pos, src_root_rot_quat, src_joint_rot,
                                                 src_body_vels, src_body_rot_vels, contacts, terrain, body_points, char_model, w_root_pos,
                                                 w_root_rot, w_joint_rot, w_smoothness, w_penetration, w_contact, w_sliding,
                                                 w_body_constraints, w_jerk, body_constraints, max_jerk, _localized=False)


# ---- set-up conversions of the optimisation loop (run once per motion, elementwise torch on the caller's device) --------------
def _quat_axis_angle(q):
    """torch_util.quat_to_axis_angle (torch_util.py:69-91): w made non-negative, angle = 2 atan2(|xyz|, w), axis z below 1e-5."""
    q = torch.where(q[..., 3:] < 0, -q, q)
    ln = torch.linalg.vector_norm(q[..., :3], dim=-1)
    angle = 2.0 * torch.atan2(ln, q[..., 3])
    axis = q[..., :3] / ln.unsqueeze(-1).clamp(min=1e-6)
    small = ln <= 1e-5
    z = torch.zeros_like(axis)
    z[..., 2] = 1.0
    return torch.where(small.unsqueeze(-1), z, axis), torch.where(small, torch.zeros_like(angle), angle)


def _quat_to_exp_map(q):                     # torch_util.py:372-376
    axis, angle = _quat_axis_angle(q)
    return axis * angle.unsqueeze(-1)


def _quat_mul(a, b):                         # Hamilton product, xyzw
    ax, ay, az, aw = a.unbind(-1)
    bx, by, bz, bw = b.unbind(-1)
    return torch.stack([aw * bx + ax * bw + ay * bz - az * by, aw * by - ax * bz + ay * bw + az * bx,
                        aw * bz + ax * by - ay * bx + az * bw, aw * bw - ax * bx - ay * by - az * bz], dim=-1)


def _quat_diff_angle(q0, q1):                # torch_util.py:453-462: angle of q1 * conj(q0)
    conj = torch.cat([-q0[..., :3], q0[..., 3:]], dim=-1)
    return _quat_axis_angle(_quat_mul(q1, conj))[1]


def _exp_map_to_quat(e):
    """torch_util.exp_map_to_quat (:426-450): angle = |e| normalised to (-pi, pi], axis e / |e| (z below 1e-5), unit quaternion."""
    angle = torch.linalg.vector_norm(e, dim=-1)
    axis = e / angle.unsqueeze(-1)
    angle = torch.atan2(torch.sin(angle), torch.cos(angle))
    small = angle.abs() <= 1e-5
    z = torch.zeros_like(e)
    z[..., 2] = 1.0
    axis = torch.where(small.unsqueeze(-1), z, axis)
    angle = torch.where(small, torch.zeros_like(angle), angle)
    axis = axis / torch.linalg.vector_norm(axis, dim=-1, keepdim=True).clamp(min=1e-9)
    q = torch.cat([axis * torch.sin(angle / 2).unsqueeze(-1), torch.cos(angle / 2).unsqueeze(-1)], dim=-1)
    return q / torch.linalg.vector_norm(q, dim=-1, keepdim=True).clamp(min=1e-9)


def motion_contact_optimization(src_motion_frames, body_points: list, terrain, char_model: KinCharModel, num_iters: int, step_size: float,
                                w_root_pos: float, w_root_rot: float, w_joint_rot: float, w_smoothness: float, w_penetration: float,
                                w_contact: float, w_sliding: float, w_body_constraints: float, w_jerk: float, body_constraints: list,
                                max_jerk: float, exp_name: str = None, use_wandb: bool = False, log_file: str = None, log=None):
    """motion_optimization.py:721-823: Adam on (root position, root exp map, joint dofs) of one motion against the CUDA loss.
    Same arguments and return (a MotionFrames with the optimised root_pos / root_rot / joint_rot and the source contacts).
    Every iteration is one call of parc_b200_motion_loss_constrained (value + gradient) and torch's Adam step; the constraint
    lists are packed once. Logging: every 25 iterations `log(iteration, loss, {LossType: value})` if given, and a line per log
    step into `log_file`; `use_wandb` is accepted and ignored (wandb is outside this path)."""
    from .file_io import MotionFrames
    src_root_pos, src_root_rot_quat, src_joint_rot = src_motion_frames.root_pos, src_motion_frames.root_rot, src_motion_frames.joint_rot
    assert src_root_pos.dim() == 2 and src_root_rot_quat.dim() == 2 and src_joint_rot.dim() == 3, "one motion: [T,3], [T,4], [T,J-1,4]"
    contacts = src_motion_frames.contacts
    src_root_rot = _quat_to_exp_map(src_root_rot_quat)
    src_joint_dof = char_model.rot_to_dof(src_joint_rot)
    src_body_pos, src_body_rot = char_model.forward_kinematics(src_root_pos, src_root_rot_quat, src_joint_rot)
    src_body_vels = src_body_pos[1:] - src_body_pos[:-1]
    src_body_rot_vels = _quat_diff_angle(src_body_rot[1:], src_body_rot[:-1])
    params = [src_root_pos.clone().requires_grad_(True), src_root_rot.clone().requires_grad_(True), src_joint_dof.clone().requires_grad_(True)]
    optimizer = torch.optim.Adam(params, lr=step_size)
    packed = None
    if body_constraints is not None:
        packed = body_constraints if isinstance(body_constraints, PackedBodyConstraints) else \
            PackedBodyConstraints([body_constraints], char_model, src_root_pos.device)
    out = open(log_file, "w") if log_file is not None else None
    for it in range(num_iters):
        optimizer.zero_grad()
        loss, loss_dict = motion_terrain_contact_loss_localized(
            params[0], params[1], params[2], src_root_pos, src_root_rot_quat, src_joint_rot, src_body_vels, src_body_rot_vels, contacts,
            terrain, body_points, char_model, w_root_pos, w_root_rot, w_joint_rot, w_smoothness, w_penetration, w_contact, w_sliding,
            w_body_constraints, w_jerk, packed, max_jerk)
        loss.backward()
        optimizer.step()
        if it % 25 == 0 and (log is not None or out is not None):
            if log is not None:
                log(it, float(loss.detach()), loss_dict)
            if out is not None:
                out.write("%d %.6f %s\n" % (it, float(loss.detach()), " ".join("%s=%.6f" % (k.name, v) for k, v in loss_dict.items())))
    if out is not None:
        out.close()
    return MotionFrames(root_pos=params[0].detach(), root_rot=_exp_map_to_quat(params[1].detach()),
                        joint_rot=char_model.dof_to_rot(params[2].detach()), contacts=contacts.clone())
