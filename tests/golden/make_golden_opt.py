"""Golden fixtures for the kinematic-optimizer rows, produced by the
UNMODIFIED PARC reference on CPU (oracle/ref_import.py; run in the authoring container only):

    python tests/golden/make_golden_opt.py

  ref_body_points.npz   geom_util.get_char_point_samples(humanoid) with the kin_gen defaults
                        (data/configs/parc_2_kin_gen_default.yaml:35-42): 304 points, per-body counts.
                        The icosahedron vertex order comes from oracle/ref_shim/trimesh.py (trimesh is
                        not installed here); sums and per-body minima do not depend on point order.
  ref_hf_sdf.npz        terrain_util.points_hf_sdf on seeded points over a 9x7 blocky patch, both
                        `inverted` settings
  ref_opt_loss.npz      motion_terrain_contact_loss_localized on 24 frames of the bundled teaser clip
                        (its own terrain), targets = source + seeded perturbation, first half pushed 6 cm into the
                        ground, second half lifted 5 cm: total loss, the 9 loss terms, and autograd gradients w.r.t. the three
                        parameter tensors. Two weight sets (kin_gen defaults; contact/sliding off).
                        Plus two body-constraint cases (motion_optimization.py:581-609) with the kin_gen weights:
                        `cons`  = the lists compute_approx_body_constraints (:70-205) derives from the clip's contact labels
                                  and terrain (its inputs are the src_* tensors above; its outputs are stored flattened as
                                  cons_body / cons_start / cons_end / cons_point),
                        `cons2` = hand-made lists that make every branch active (hinge on for the feet, both signs of the
                                  sphere distance for the hands, overlapping ranges, a range that runs past the last frame,
                                  a constraint on a capsule body, which the reference ignores).
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from oracle import ref_import  # noqa: E402

ref_import.setup()
torch.set_num_threads(1)

import parc.anim.kin_char_model as rkin  # noqa: E402
import parc.motion_synthesis.motion_opt.motion_optimization as ropt  # noqa: E402
import parc.util.file_io_helper as rfio  # noqa: E402
import parc.util.geom_util as rgeom  # noqa: E402
import parc.util.terrain_util as rterr  # noqa: E402
import parc.util.torch_util as rtu  # noqa: E402

KIN_GEN_WEIGHTS = dict(w_root_pos=1.0, w_root_rot=10.0, w_joint_rot=1.0, w_smoothness=10.0, w_penetration=1000.0,
                       w_contact=1000.0, w_sliding=10.0, w_body_constraints=1000.0, w_jerk=1000.0, max_jerk=1000.0)
POINT_SAMPLES = dict(sphere_num_subdivisions=0, box_num_slices=2, box_dim_x=3, box_dim_y=6, capsule_num_circle_points=4,
                     capsule_num_sphere_subdivisons=0, capsule_num_cylinder_slices=4)


def npz(name, **arrs):
    out = {k: (v.detach().cpu().numpy() if isinstance(v, torch.Tensor) else np.asarray(v)) for k, v in arrs.items()}
    np.savez_compressed(os.path.join(HERE, name), **out)
    print("wrote", name, {k: v.shape for k, v in out.items()})


def main():
    model = rkin.KinCharModel("cpu")
    model.load_char_file(ref_import.humanoid_xml())
    body_points = rgeom.get_char_point_samples(model, **POINT_SAMPLES)
    counts = np.array([p.shape[0] for p in body_points], dtype=np.int64)
    npz("ref_body_points.npz", points=torch.cat(body_points, dim=0), counts=counts)

    # ---- points_hf_sdf ---------------------------------------------------------------------------
    g = torch.Generator().manual_seed(5)
    B, X, Y, Np = 3, 9, 7, 64
    hf = 0.2 * torch.randint(-5, 6, (B, X, Y), generator=g).float()
    min_c = torch.tensor([[-1.0, 0.5], [0.3, -2.0], [10.0, 10.0]])
    dxdy = torch.tensor([0.4, 0.4])
    pts = torch.rand(B, Np, 3, generator=g)
    pts[..., 0] = min_c[:, None, 0] - 0.6 + pts[..., 0] * (X * 0.4 + 0.8)
    pts[..., 1] = min_c[:, None, 1] - 0.6 + pts[..., 1] * (Y * 0.4 + 0.8)
    pts[..., 2] = -1.5 + 3.0 * pts[..., 2]
    sd_inv = rterr.points_hf_sdf(pts, hf, min_c, dxdy, base_z=-10.0, inverted=True)
    sd_pos = rterr.points_hf_sdf(pts, hf, min_c, dxdy, base_z=-10.0, inverted=False)
    sd_inv_r = rterr.points_hf_sdf(pts, hf, min_c, dxdy, base_z=-10.0, inverted=True, radius=0.07)      # rounded columns (:1782-1785)
    sd_pos_r = rterr.points_hf_sdf(pts, hf, min_c, dxdy, base_z=-10.0, inverted=False, radius=0.07)
    npz("ref_hf_sdf.npz", points=pts, hf=hf, min_box_center=min_c, dxdy=dxdy, sd_inverted=sd_inv, sd_solid=sd_pos,
        sd_inverted_r=sd_inv_r, sd_solid_r=sd_pos_r, radius=np.float64(0.07))

    # ---- motion_terrain_contact_loss_localized ----------------------------------------------------
    d = rfio.load_ms_file(ref_import.clip("dec2024_teaser_717_1_opt_dm.pkl"), device="cpu")
    md, td = d.motion_data, d.terrain_data
    T = 24
    src_root_pos = torch.as_tensor(md.root_pos, dtype=torch.float32)[:T].clone()
    src_root_rot_quat = torch.as_tensor(md.root_rot, dtype=torch.float32)[:T].clone()
    src_joint_rot = torch.as_tensor(md.joint_rot, dtype=torch.float32)[:T].clone()
    contacts = torch.as_tensor(md.body_contacts, dtype=torch.float32)[:T].clone()
    terrain = rterr.SubTerrain.from_ms_terrain_data(td, device="cpu") if hasattr(rterr.SubTerrain, "from_ms_terrain_data") else None
    if terrain is None:
        terrain = rterr.SubTerrain("t", x_dim=td.hf.shape[0], y_dim=td.hf.shape[1], dx=float(td.dx), dy=float(td.dy),
                                   min_x=float(td.min_point[0]), min_y=float(td.min_point[1]), device="cpu")
        terrain.hf = torch.as_tensor(td.hf, dtype=torch.float32)
    src_body_pos, src_body_rot = model.forward_kinematics(src_root_pos, src_root_rot_quat, src_joint_rot)
    src_body_vels = src_body_pos[1:] - src_body_pos[:-1]
    src_body_rot_vels = rtu.quat_diff_angle(src_body_rot[1:], src_body_rot[:-1])
    g = torch.Generator().manual_seed(17)
    tgt_root_pos = src_root_pos + 0.02 * torch.randn(T, 3, generator=g)
    tgt_root_pos[:12, 2] -= 0.06                                    # into the ground: penetration term active
    tgt_root_pos[12:, 2] += 0.05                                    # floating: contact term active
    tgt_root_rot = rtu.quat_to_exp_map(src_root_rot_quat) + 0.03 * torch.randn(T, 3, generator=g)
    tgt_joint_dof = model.rot_to_dof(src_joint_rot) + 0.05 * torch.randn(T, model.get_dof_size(), generator=g)
    out = dict(src_root_pos=src_root_pos, src_root_rot_quat=src_root_rot_quat, src_joint_rot=src_joint_rot,
               src_body_vels=src_body_vels, src_body_rot_vels=src_body_rot_vels, contacts=contacts,
               hf=terrain.hf, min_point=terrain.min_point, dxdy=terrain.dxdy,
               tgt_root_pos=tgt_root_pos, tgt_root_rot=tgt_root_rot, tgt_joint_dof=tgt_joint_dof)
    for tag, wts in (("full", KIN_GEN_WEIGHTS), ("nocontact", dict(KIN_GEN_WEIGHTS, w_contact=0.0, w_sliding=0.0))):
        p = [t.clone().requires_grad_(True) for t in (tgt_root_pos, tgt_root_rot, tgt_joint_dof)]
        loss, losses = ropt.motion_terrain_contact_loss_localized(
            p[0], p[1], p[2], src_root_pos, src_root_rot_quat, src_joint_rot, src_body_vels, src_body_rot_vels, contacts, terrain,
            body_points, model, body_constraints=None, **wts)
        loss.backward()
        out[tag + "_loss"] = loss.detach()
        out[tag + "_terms"] = np.array([float(losses.get(t, 0.0)) for t in list(ropt.LossType)[:9]], dtype=np.float64)
        out[tag + "_weights"] = np.array([wts[k] for k in ("w_root_pos", "w_root_rot", "w_joint_rot", "w_smoothness", "w_penetration",
                                                           "w_contact", "w_sliding", "w_body_constraints", "w_jerk", "max_jerk")])
        out[tag + "_g_root_pos"], out[tag + "_g_root_rot"], out[tag + "_g_joint_dof"] = p[0].grad, p[1].grad, p[2].grad
        print(tag, float(loss), {k.name: round(v, 6) if isinstance(v, float) else v for k, v in losses.items()})
    # ---- motion_terrain_contact_loss (:207-423): every cell of the terrain instead of per-frame windows; the root is
    # raised so that for many points a neighbouring column is nearer than the ground below
    p = [t.clone().requires_grad_(True) for t in (tgt_root_pos + torch.tensor([0.0, 0.0, 0.9]), tgt_root_rot, tgt_joint_dof)]
    loss, losses = ropt.motion_terrain_contact_loss(
        p[0], p[1], p[2], src_root_pos, src_root_rot_quat, src_joint_rot, src_body_vels, src_body_rot_vels, contacts, terrain,
        body_points, model, body_constraints=None, **KIN_GEN_WEIGHTS)
    loss.backward()
    loc, loc_losses = ropt.motion_terrain_contact_loss_localized(
        p[0].detach(), p[1].detach(), p[2].detach(), src_root_pos, src_root_rot_quat, src_joint_rot, src_body_vels, src_body_rot_vels,
        contacts, terrain, body_points, model, body_constraints=None, **KIN_GEN_WEIGHTS)
    out["whole_tgt_root_pos"] = p[0].detach()
    out["whole_loss"] = loss.detach()
    out["whole_terms"] = np.array([float(losses.get(t, 0.0)) for t in list(ropt.LossType)[:9]], dtype=np.float64)
    out["whole_localized_terms"] = np.array([float(loc_losses.get(t, 0.0)) for t in list(ropt.LossType)[:9]], dtype=np.float64)
    out["whole_weights"] = out["full_weights"]
    out["whole_g_root_pos"], out["whole_g_root_rot"], out["whole_g_joint_dof"] = p[0].grad, p[1].grad, p[2].grad
    print("whole", float(loss), "localized on the same input", float(loc), out["whole_terms"][4:6], out["whole_localized_terms"][4:6])

    # ---- body constraints ------------------------------------------------------------------------
    def flatten(bc):
        rows = [(b, c.start_frame_idx, c.end_frame_idx, c.constraint_point) for b, lst in enumerate(bc) for c in lst]
        if not rows:
            return np.zeros(0, np.int64), np.zeros(0, np.int64), np.zeros(0, np.int64), np.zeros((0, 3), np.float32)
        return (np.array([r[0] for r in rows], np.int64), np.array([r[1] for r in rows], np.int64), np.array([r[2] for r in rows], np.int64),
                torch.stack([r[3] for r in rows]).numpy())

    J = model.get_num_joints()
    kinds, dims, offs = [], np.zeros((J, 3), np.float32), np.zeros((J, 3), np.float32)
    for b in range(J):                                  # first geom of every body: all the constraint term looks at (:590, :598)
        g0 = model.get_geoms(b)[0]
        kinds.append(g0._shape_type.name)
        dv = g0._dims.reshape(-1).numpy()
        dims[b, :dv.shape[0]] = dv
        offs[b] = g0._offset.numpy()
    out["geom_kind"], out["geom_dims"], out["geom_offset"] = np.array(kinds), dims, offs
    torch.manual_seed(0)
    auto = ropt.compute_approx_body_constraints(src_root_pos, src_root_rot_quat, src_joint_rot, contacts, model, terrain)
    tgt_body_pos, _ = model.forward_kinematics(tgt_root_pos, rtu.exp_map_to_quat(tgt_root_rot), model.dof_to_rot(tgt_joint_dof))

    def mk(b, s, e, pt):
        c = ropt.BodyConstraint()
        c.start_frame_idx, c.end_frame_idx, c.constraint_point = s, e, pt.clone()
        return b, c

    lf, rf, lh, rh = (model.get_body_id(n) for n in ("left_foot", "right_foot", "left_hand", "right_hand"))
    capsule_body = model.get_body_id("left_shin")
    hand = [[] for _ in range(model.get_num_joints())]
    for b, c in (mk(lf, 2, 9, tgt_body_pos[5, lf] + torch.tensor([0.25, -0.1, -0.05])),        # > 1.25 |dims| away: hinge active
                 mk(lf, 7, 15, tgt_body_pos[11, lf] + torch.tensor([0.02, 0.01, -0.03])),      # overlaps the first; mostly inside the ball
                 mk(rf, 0, 0, tgt_body_pos[0, rf] + torch.tensor([-0.2, 0.15, 0.1])),          # single frame at the start
                 mk(rf, 18, 40, tgt_body_pos[20, rf] + torch.tensor([0.1, 0.2, -0.1])),        # runs past the last frame (Python slicing clips it)
                 mk(lh, 3, 12, tgt_body_pos[8, lh] + torch.tensor([0.15, 0.0, 0.05])),         # outside the hand sphere
                 mk(lh, 14, 20, tgt_body_pos[17, lh] + torch.tensor([0.005, 0.0, 0.0])),       # inside it
                 mk(rh, 22, 23, tgt_body_pos[23, rh] + torch.tensor([0.0, -0.3, 0.0])),        # includes the last frame
                 mk(capsule_body, 4, 10, tgt_body_pos[6, capsule_body])):                      # capsule geom: ignored (:603-604)
        hand[b].append(c)
    for tag, bc in (("cons", auto), ("cons2", hand)):
        p = [t.clone().requires_grad_(True) for t in (tgt_root_pos, tgt_root_rot, tgt_joint_dof)]
        loss, losses = ropt.motion_terrain_contact_loss_localized(
            p[0], p[1], p[2], src_root_pos, src_root_rot_quat, src_joint_rot, src_body_vels, src_body_rot_vels, contacts, terrain,
            body_points, model, body_constraints=bc, **KIN_GEN_WEIGHTS)
        loss.backward()
        out[tag + "_loss"] = loss.detach()
        out[tag + "_terms"] = np.array([float(losses.get(t, 0.0)) for t in list(ropt.LossType)[:9]], dtype=np.float64)
        out[tag + "_weights"] = out["full_weights"]
        out[tag + "_g_root_pos"], out[tag + "_g_root_rot"], out[tag + "_g_joint_dof"] = p[0].grad, p[1].grad, p[2].grad
        out[tag + "_body"], out[tag + "_start"], out[tag + "_end"], out[tag + "_point"] = flatten(bc)
        print(tag, float(loss), {k.name: round(v, 6) if isinstance(v, float) else v for k, v in losses.items()},
              list(zip(out[tag + "_body"].tolist(), out[tag + "_start"].tolist(), out[tag + "_end"].tolist())))
    # ---- motion_contact_optimization (:721-823): 40 Adam iterations (kin_gen step size) on the 24 frames, the source pushed
    # 4 cm into the ground so that the penetration / contact terms drive the update, with the auto-derived constraints
    import parc.util.motion_util as rmu
    src = rmu.MotionFrames(root_pos=src_root_pos + torch.tensor([0.0, 0.0, -0.04]), root_rot=src_root_rot_quat.clone(),
                           joint_rot=src_joint_rot.clone(), contacts=contacts.clone())
    ret = ropt.motion_contact_optimization(src, body_points, terrain, model, num_iters=40, step_size=0.001, body_constraints=auto,
                                           exp_name=None, use_wandb=False, log_file=None, **KIN_GEN_WEIGHTS)
    out["optim_src_root_pos"] = src.root_pos
    out["optim_root_pos"], out["optim_root_rot"], out["optim_joint_rot"] = ret.root_pos, ret.root_rot, ret.joint_rot
    print("optim: root moved by", float((ret.root_pos - src.root_pos).abs().max()), "joints by", float((ret.joint_rot - src.joint_rot).abs().max()))
    npz("ref_opt_loss.npz", **out)


if __name__ == "__main__":
    main()
