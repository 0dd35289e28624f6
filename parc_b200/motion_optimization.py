"""The kinematic motion optimizer's loss on the GPU (reference:
PARC/motion_synthesis/motion_opt/motion_optimization.py). `motion_terrain_contact_loss_localized`
keeps the reference's signature and return value `(loss, {LossType: float})`; the loss tensor carries
autograd history to `tgt_root_pos`, `tgt_root_rot`, `tgt_joint_dof`, so the reference's Adam loop
(:772-797) runs unchanged on top of it. `motion_loss_batched` evaluates many motions in one call
(BASELINE config 4: 1,024 motions x 300 frames).

Everything numeric happens in `parc_b200_motion_loss` (include/parc_b200_opt.h): value and gradient in
one pass; `backward` only scales the stored gradients."""
import enum
from typing import List

import torch

from . import _lib
from .kin_char_model import KinCharModel


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


def build_opt_config(char_model: KinCharModel, body_points: List[torch.Tensor], num_contact_cols: int, w_root_pos, w_root_rot,
                     w_joint_rot, w_smoothness, w_penetration, w_contact, w_sliding, w_body_constraints, w_jerk, max_jerk,
                     base_z=-10.0) -> _lib.OptConfigStruct:
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
    return cfg


class _MotionLossFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, tgt_root_pos, tgt_root_rot, tgt_joint_dof, char_model, cfg, consts, terrain_args, points):
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
        _lib.check(L.parc_b200_motion_loss(km, cfg, B, T, _lib.ptr(p[0]), _lib.ptr(p[1]), _lib.ptr(p[2]), *[_lib.ptr(x) for x in keep],
                                           _lib.ptr(hf), hf_stride, X, Y, _lib.ptr(min_point), mp_stride, dx, dy, _lib.ptr(points),
                                           _lib.ptr(ws), ws_bytes, _lib.ptr(terms), _lib.ptr(loss), _lib.ptr(grads[0]), _lib.ptr(grads[1]),
                                           _lib.ptr(grads[2]), _lib.stream_ptr(dev)))
        ctx.save_for_backward(*[g for g in grads if g is not None])
        ctx.has_grads = need_grad
        ctx.mark_non_differentiable(terms)
        return loss, terms

    @staticmethod
    def backward(ctx, g_loss, _g_terms):
        if not ctx.has_grads:
            return (None,) * 8
        g0, g1, g2 = ctx.saved_tensors
        s = g_loss.reshape(-1, 1, 1)
        return s * g0, s * g1, s * g2, None, None, None, None, None


def motion_loss_batched(tgt_root_pos, tgt_root_rot, tgt_joint_dof, src_root_pos, src_root_rot_quat, src_joint_rot, src_body_vels,
                        src_body_rot_vels, contacts, hf, min_point, dxdy, body_points: List[torch.Tensor], char_model: KinCharModel,
                        w_root_pos, w_root_rot, w_joint_rot, w_smoothness, w_penetration, w_contact, w_sliding, w_body_constraints,
                        w_jerk, max_jerk):
    """B independent motions of T frames: all motion tensors carry a leading [B, T(-1), ...]; `hf` is
    [B, X, Y] with `min_point` [B, 2], or one shared [X, Y] / [2]. -> (loss [B], terms [B, 9]) with
    autograd to the three tgt_* tensors."""
    dev = tgt_root_pos.device
    assert tgt_root_pos.dim() == 3 and tgt_root_rot.shape == tgt_root_pos.shape and tgt_joint_dof.dim() == 3
    B, T = tgt_root_pos.shape[:2]
    J = char_model.get_num_joints()
    assert src_joint_rot.shape == (B, T, J - 1, 4) and src_body_vels.shape == (B, T - 1, J, 3) and src_body_rot_vels.shape == (B, T - 1, J)
    assert contacts.shape[:2] == (B, T)
    cfg = build_opt_config(char_model, body_points, contacts.shape[-1], w_root_pos, w_root_rot, w_joint_rot, w_smoothness, w_penetration,
                           w_contact, w_sliding, w_body_constraints, w_jerk, max_jerk)
    hf = hf.to(device=dev, dtype=torch.float32).contiguous()
    min_point = min_point.to(device=dev, dtype=torch.float32).contiguous()
    if hf.dim() == 2:
        X, Y, hf_stride, mp_stride = hf.shape[0], hf.shape[1], 0, 0
        assert min_point.shape == (2,)
    else:
        assert hf.shape[0] == B and min_point.shape == (B, 2)
        X, Y, hf_stride, mp_stride = hf.shape[1], hf.shape[2], hf.shape[1] * hf.shape[2], 2
    points = torch.cat([p.to(device=dev, dtype=torch.float32) for p in body_points], dim=0).contiguous()
    return _MotionLossFn.apply(tgt_root_pos, tgt_root_rot, tgt_joint_dof, char_model, cfg,
                               (src_root_pos, src_root_rot_quat, src_joint_rot, src_body_vels, src_body_rot_vels, contacts),
                               (hf, hf_stride, X, Y, min_point, mp_stride, float(dxdy[0]), float(dxdy[1])), points)


def motion_terrain_contact_loss_localized(tgt_root_pos, tgt_root_rot, tgt_joint_dof, src_root_pos, src_root_rot_quat, src_joint_rot,
                                          src_body_vels, src_body_rot_vels, contacts, terrain, body_points: List[torch.Tensor],
                                          char_model: KinCharModel, w_root_pos: float, w_root_rot: float, w_joint_rot: float,
                                          w_smoothness: float, w_penetration: float, w_contact: float, w_sliding: float,
                                          w_body_constraints: float, w_jerk: float, body_constraints: list, max_jerk: float):
    """motion_optimization.py:426-656, one motion of T frames. Returns (loss, {LossType: float})."""
    if body_constraints is not None:
        raise NotImplementedError("body_constraints are not supported by the CUDA loss (pass None)")
    u = lambda t: t.unsqueeze(0)
    loss, terms = motion_loss_batched(u(tgt_root_pos), u(tgt_root_rot), u(tgt_joint_dof), u(src_root_pos), u(src_root_rot_quat),
                                      u(src_joint_rot), u(src_body_vels), u(src_body_rot_vels), u(contacts), terrain.hf,
                                      terrain.min_point, terrain.dxdy, body_points, char_model, w_root_pos, w_root_rot, w_joint_rot,
                                      w_smoothness, w_penetration, w_contact, w_sliding, w_body_constraints, w_jerk, max_jerk)
    vals = terms[0].tolist()       # the reference also reads every term back with .item() (:626-643)
    losses = {t: vals[t.value] for t in list(LossType)[:9]}
    return loss[0], losses
