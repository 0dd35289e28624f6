"""Ranking loss of generated motion batches -- mirror of `compute_motion_loss`
(PARC/motion_synthesis/procgen/mdm_path.py:31-127), which the autoregressive
generation loop calls on every MDM batch (`:263`) to keep the top-k motions.

Same signature and return dict. Forward kinematics runs in the CUDA `KinCharModel.forward_kinematics`; the per-body
`points_hf_sdf` loops (O(frames x points x cells) with a `[B, T*points, cells, 3]` intermediate upstream) are one
launch of the optimizer's pruned box-SDF kernel over the whole terrain (`parc_b200_path_motion_loss`,
include/parc_b200_opt.h). There is no CPU fallback.
"""
from typing import List

import torch

from . import _lib
from .kin_char_model import KinCharModel
from .motion_optimization import build_opt_config


class MDMPathSettings:                    # mdm_path.py:18-29
    next_node_lookahead = 7
    rewind_num_frames = 5
    end_of_path_buffer = 2
    max_motion_length = 10.0
    path_batch_size = 16
    mdm_batch_size = 32
    top_k = 4
    w_target = 2.0
    w_contact = 0.1
    w_pen = 0.1


def compute_motion_loss(motion_frames, path_nodes, terrain, char_model: KinCharModel, body_points: List[torch.Tensor], w_contact: float,
                        w_pen: float, w_path: float = 0.0, verbose: bool = True):
    """motion_frames: .root_pos [B,T,3], .root_rot [B,T,4], .joint_rot [B,T,J-1,4], .contacts [B,T,J] (MotionFrames);
    terrain: SubTerrain; body_points: geom_util.get_char_point_samples(char_model). `path_nodes` / `w_path` are unused
    upstream as well (:45-61). Returns {"total_loss", "contact_loss", "pen_loss"}, each [B]."""
    root_pos, root_rot, joint_rot, contacts = motion_frames.root_pos, motion_frames.root_rot, motion_frames.joint_rot, motion_frames.contacts
    assert root_pos.dim() == 3 and root_rot.dim() == 3 and joint_rot.dim() == 4 and contacts.dim() == 3
    B, T = int(root_pos.shape[0]), int(root_pos.shape[1])
    J = char_model.get_num_joints()
    dev, f32 = root_pos.device, torch.float32
    body_pos, body_rot = char_model.forward_kinematics(root_pos.detach(), root_rot.detach(), joint_rot.detach())      # :76
    body_pos, body_rot = body_pos.contiguous().float(), body_rot.contiguous().float()
    contacts = contacts.detach().to(dtype=f32).contiguous()
    assert contacts.shape[-1] >= J, "compute_motion_loss reads contacts[..., b] for every body b (:109)"
    hf = terrain.hf.to(dtype=f32).contiguous()
    min_z = torch.min(hf).item()                                                                                   # :78
    cfg = build_opt_config(char_model, body_points, int(contacts.shape[-1]), 0, 0, 0, 0, w_pen, w_contact, 0, 0, 0, 0, base_z=min_z - 10.0)
    for b in range(J):
        cfg.contact_col[b] = b
    pts = torch.cat([p.to(device=dev, dtype=f32) for p in body_points], dim=0).contiguous()
    L = _lib.lib()
    ws_bytes = int(L.parc_b200_path_motion_loss_workspace_bytes(B, T))
    ws = torch.empty(max(ws_bytes, 4) // 4, dtype=f32, device=dev)
    pen, con, total = (torch.empty(B, dtype=f32, device=dev) for _ in range(3))
    mp = terrain.min_point.to(device=dev, dtype=f32).contiguous()
    dd = terrain.dxdy.detach().cpu()
    p = _lib.ptr
    _lib.check(L.parc_b200_path_motion_loss(char_model._cstruct(), cfg, B, T, p(body_pos), p(body_rot), p(contacts), p(hf), int(hf.shape[0]),
                                            int(hf.shape[1]), p(mp), float(dd[0]), float(dd[1]), p(pts), float(w_pen), float(w_contact),
                                            p(ws), ws_bytes, p(pen), p(con), p(total), _lib.stream_ptr(dev)))
    return {"total_loss": total, "contact_loss": con, "pen_loss": pen}
