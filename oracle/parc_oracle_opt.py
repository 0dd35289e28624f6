"""TEST INFRASTRUCTURE ONLY -- CPU (torch, autograd) restatement of the kinematic-optimizer rows of
SURVEY.md section 8 (a20-a22). Never imported by the product path (`parc_b200/`).

Parity pinned: tests/test_oracle_golden_opt.py checks every function here against
tests/golden/ref_body_points.npz, ref_hf_sdf.npz and ref_opt_loss.npz, which were produced by the
unmodified reference (tests/golden/make_golden_opt.py).

Reference files: PARC/util/geom_util.py (sdBox :124-145, get_char_point_samples :790-872),
PARC/util/terrain_util.py (points_boxes_sdf :1678-1705, points_hf_sdf :1736-1794),
PARC/motion_synthesis/motion_opt/motion_optimization.py (motion_terrain_contact_loss_localized :426-656).
"""
import torch

from oracle import parc_oracle as O

LOSS_NAMES = ("ROOT_POS_LOSS", "ROOT_ROT_LOSS", "JOINT_ROT_LOSS", "SMOOTHNESS_LOSS", "PENETRATION_LOSS", "CONTACT_LOSS",
              "SLIDING_LOSS", "BODY_CONSTRAINT_LOSS", "JERK_LOSS")      # motion_optimization.py:15-25 (enum order)


def sd_box(p, half):
    """geom_util.py:124-145: signed distance of p to an origin-centred box with half extents `half`."""
    q = p.abs() - half
    outside = torch.clamp(q, min=0.0).norm(dim=-1)
    inside = torch.clamp(q.max(dim=-1)[0], max=0.0)
    return outside + inside


def points_hf_sdf(points, hf, min_box_center, dxdy, base_z=-10.0, inverted=True, radius=None):
    """terrain_util.py:1736-1794. points [B,N,3], hf [B,X,Y], min_box_center [B,2], dxdy [2] -> [B,N].
    One axis-aligned box per cell, from base_z up to the cell height (solid terrain) or, inverted,
    from the cell height up to -base_z (the air above it; the result is negated), min over cells."""
    B, X, Y = hf.shape
    xs = torch.linspace(0.0, (X - 1.0) * float(dxdy[0]), X)
    ys = torch.linspace(0.0, (Y - 1.0) * float(dxdy[1]), Y)
    gx, gy = torch.meshgrid(xs, ys, indexing="ij")
    cx = gx[None] + min_box_center[:, 0, None, None]
    cy = gy[None] + min_box_center[:, 1, None, None]
    if inverted:
        top = -base_z
        cz, hz = (hf + top) / 2.0, (top - hf) / 2.0
    else:
        cz, hz = (hf + base_z) / 2.0, (hf - base_z) / 2.0
    centers = torch.stack([cx, cy, cz], dim=-1).reshape(B, X * Y, 3)
    half = torch.stack([torch.full_like(hz, float(dxdy[0]) / 2.0), torch.full_like(hz, float(dxdy[1]) / 2.0), hz], dim=-1).reshape(B, X * Y, 3)
    rel = points[:, :, None, :] - centers[:, None, :, :]
    sd = sd_box(rel, half[:, None, :, :])
    if radius is not None:
        sd = sd - radius                    # geom_util.sdRoundBox (geom_util.py:115-122): columns with rounded edges
    sd = sd.min(dim=-1)[0]
    return -sd if inverted else sd


def frame_patches(points_xy, hf, min_point, dxdy):
    """motion_optimization.py:507-551: per-frame heightfield window that covers the body points of the
    frame +- 2 cells, all frames sharing the largest window size. -> (start_idx [T,2] int64, (px, py))."""
    dims = torch.tensor(hf.shape, dtype=torch.int64)
    pad = dxdy * 2.0
    lo = points_xy.min(dim=1)[0] - pad
    hi = points_xy.max(dim=1)[0] + pad
    inv = 1.0 / dxdy
    gmin = torch.floor((lo - min_point) * inv).to(torch.int64)
    gmax = torch.ceil((hi - min_point) * inv).to(torch.int64)
    gmin = torch.maximum(gmin, torch.zeros(2, dtype=torch.int64))
    gmax = torch.minimum(gmax, dims - 1)
    size = torch.maximum((gmax - gmin + 1).max(dim=0).values, torch.ones_like(dims))
    size = torch.minimum(size, dims)
    start = torch.minimum(gmin, torch.maximum(dims - size, torch.zeros_like(dims)))
    start = torch.maximum(start, torch.zeros(2, dtype=torch.int64))
    return start, (int(size[0]), int(size[1]))


def motion_loss(model, body_points, body_point_counts, hf, min_point, dxdy, tgt_root_pos, tgt_root_rot, tgt_joint_dof,
                src_root_pos, src_root_rot_quat, src_joint_rot, src_body_vels, src_body_rot_vels, contacts, weights,
                contact_body_of_body=None, body_constraints=None, geoms=None, localized=True):
    """motion_terrain_contact_loss_localized (:426-656). `body_constraints` (:581-609): None or a list of records
    (body, start_frame, end_frame, point[3]); `geoms[b]` = ("SPHERE", radius, offset[3]) | ("BOX", dims[3]) | anything else
    (ignored), the first geom of body b. `localized=False` is motion_terrain_contact_loss (:207-423): the whole terrain for
    every frame instead of the per-frame windows. `weights` =
    [w_root_pos, w_root_rot, w_joint_rot, w_smoothness, w_penetration, w_contact, w_sliding,
    w_body_constraints, w_jerk, max_jerk]. body_points [Np,3] concatenated in body order with
    body_point_counts [J]. -> (loss, terms [9] in LOSS_NAMES order); differentiable in the tgt_* tensors."""
    (w_root_pos, w_root_rot, w_joint_rot, w_smooth, w_pen, w_contact, w_slide, w_bc, w_jerk, max_jerk) = [float(x) for x in weights]
    T, J = tgt_root_pos.shape[0], model.J
    root_pos_loss = (tgt_root_pos - src_root_pos).square().sum()
    tq = O.exp_map_to_quat(tgt_root_rot)
    root_rot_loss = O.quat_diff_angle(tq, src_root_rot_quat).square().sum()
    jq = model.dof_to_rot(tgt_joint_dof)
    joint_rot_loss = O.quat_diff_angle(jq, src_joint_rot).square().sum()
    body_pos, body_rot = model.forward_kinematics(tgt_root_pos, tq, jq)
    vel = body_pos[1:] - body_pos[:-1]
    vel_err_sq = (vel - src_body_vels).square()
    rvel = O.quat_diff_angle(body_rot[1:], body_rot[:-1])
    rvel_err_sq = (rvel - src_body_rot_vels).square()
    smooth_loss = vel_err_sq.sum() + rvel_err_sq.sum()
    both = torch.clamp(torch.minimum(contacts[1:], contacts[:-1]), min=0.0)

    body_of_point = torch.repeat_interleave(torch.arange(J), torch.as_tensor(body_point_counts))
    world = O.quat_rotate(body_rot[:, body_of_point], body_points[None].expand(T, -1, -1)) + body_pos[:, body_of_point]
    start, (px, py) = frame_patches(world[..., :2].detach(), hf, min_point, dxdy)
    if not localized:
        start, (px, py) = torch.zeros_like(start), (int(hf.shape[0]), int(hf.shape[1]))
    ix = start[:, 0, None] + torch.arange(px)[None]
    iy = start[:, 1, None] + torch.arange(py)[None]
    patches = hf[ix[:, :, None], iy[:, None, :]]
    patch_min = min_point[None] + start.float() * dxdy[None]
    neg = torch.clamp(points_hf_sdf(world, patches, patch_min, dxdy, -10.0, True), max=0.0)
    pen_loss = (-neg).sum()
    contact_loss = torch.zeros(())
    if w_contact != 0.0:
        pos = torch.clamp(points_hf_sdf(world, patches, patch_min, dxdy, -10.0, False), min=0.0)
        off = 0
        for b in range(J):
            n = int(body_point_counts[b])
            closest = pos[:, off:off + n].min(dim=-1)[0]
            off += n
            cb = b if contact_body_of_body is None else contact_body_of_body[b]
            if cb is not None and cb >= 0:
                contact_loss = contact_loss + (closest * contacts[:, cb]).sum()
    bc_loss = torch.zeros(())
    if body_constraints is not None:
        starts = [0]
        for n in body_point_counts:
            starts.append(starts[-1] + int(n))
        vel_keep = torch.ones(T - 1, J)
        for (b, s0, e0, pt) in body_constraints:
            b, s0, e0 = int(b), int(s0), int(e0)
            pt = torch.as_tensor(pt, dtype=torch.float32)
            g = geoms[b]
            if g[0] == "SPHERE":                               # :590-596: | |point - centre| - radius | per frame
                centre = O.quat_rotate(body_rot[:, b], torch.as_tensor(g[2], dtype=torch.float32)[None].expand(T, -1)) + body_pos[:, b]
                bc_loss = bc_loss + ((pt[None] - centre[s0:e0 + 1]).norm(dim=-1) - float(g[1])).abs().sum()
            elif g[0] == "BOX":                                # :598-602: hinge on the first 18 sample points, ball of 1.25 |dims|
                radius = torch.as_tensor(g[1], dtype=torch.float32).norm() * 1.25
                sole = world[s0:e0 + 1, starts[b]:starts[b + 1]][:, 0:18].reshape(-1, 3)
                bc_loss = bc_loss + torch.clamp((pt[None] - sole).norm(dim=-1) - radius, min=0.0).sum()
            else:
                continue
            vel_keep[s0:e0 + 1, b] = 0.0                       # :606-609, after the smoothness sum: the sliding term only
        vel_err_sq = vel_err_sq * vel_keep[..., None]
        rvel_err_sq = rvel_err_sq * vel_keep
    slide_loss = torch.zeros(())
    if w_slide != 0.0 and both.shape[-1] == J:
        c, c2 = 0.03, 0.0009
        slide_loss = ((torch.sqrt(vel_err_sq.sum(dim=-1) + c2) - c) * both).sum() + ((torch.sqrt(rvel_err_sq + c2) - c) * both).sum()
    acc = vel[1:] - vel[:-1]
    jerk = (acc[1:] - acc[:-1]).norm(dim=-1)
    jerk_loss = torch.clamp(jerk - max_jerk * (1.0 / 30.0) ** 3, min=0.0).sum()
    terms = [root_pos_loss, root_rot_loss, joint_rot_loss, smooth_loss, pen_loss, contact_loss, slide_loss, bc_loss, jerk_loss]
    ws = [w_root_pos, w_root_rot, w_joint_rot, w_smooth, w_pen, w_contact, w_slide, w_bc, w_jerk]
    loss = sum(w * t for w, t in zip(ws, terms))
    return loss, torch.stack([t.detach().double() for t in terms])


def approx_body_constraints(model, root_pos, root_rot, joint_rot, contacts, hf, min_point, dxdy, bodies, box_offsets, iters=1000):
    """compute_approx_body_constraints (:70-205). `bodies` = ids of (left_foot, right_foot, left_hand, right_hand);
    `box_offsets[k]` = offset of the k-th body's box geom or None (hands). Returns records (body, start, end, point[3]) in
    `bodies` order. Every run of consecutive frames with contact > 0.9 gives one record (a trailing run of ONE frame is
    dropped, :94-95), its point = mean body position over the run, moved by `iters` SGD steps (lr 0.01) on the squared
    solid-column distance to the whole terrain (base_z = min(hf) - 10)."""
    body_pos, body_rot = model.forward_kinematics(root_pos, root_rot, joint_rot)
    T = root_pos.shape[0]
    base_z = float(hf.min()) - 10.0
    records = []
    for k, b in enumerate(bodies):
        pos = body_pos[:, b]
        if box_offsets[k] is not None:
            pos = pos + O.quat_rotate(body_rot[:, b], torch.as_tensor(box_offsets[k], dtype=torch.float32)[None].expand(T, -1))
        idx = torch.nonzero(contacts[:, b] > 0.9).flatten().tolist()
        runs, cur = [], []
        for i in idx:
            if cur and i != cur[-1] + 1:
                runs.append(cur)
                cur = []
            cur.append(i)
        if len(cur) >= 2:                                      # the last run survives only with two or more frames
            runs.append(cur)
        for run in runs:
            pt = pos[run].mean(dim=0).detach().clone().requires_grad_(True)
            for _ in range(iters):
                sd = points_hf_sdf(pt[None, None], hf[None], min_point[None], dxdy, base_z, False)
                (g,) = torch.autograd.grad(sd.square().sum(), pt)
                pt = (pt - 0.01 * g).detach().requires_grad_(True)
            records.append((int(b), int(run[0]), int(run[-1]), pt.detach()))
    return records


# ------------------------------------------------------------------------------------------------
# body surface sample points (geom_util.py:745-872), host setup shared by oracle and kernels
# ------------------------------------------------------------------------------------------------
def icosahedron(radius):
    """trimesh.creation.icosphere(subdivisions=0): the 12 icosahedron vertices scaled to `radius`
    (vertex order of oracle/ref_shim/trimesh.py)."""
    p = (1.0 + 5.0 ** 0.5) / 2.0
    v = torch.tensor([(-1, p, 0), (1, p, 0), (-1, -p, 0), (1, -p, 0), (0, -1, p), (0, 1, p), (0, -1, -p), (0, 1, -p),
                      (p, 0, -1), (p, 0, 1), (-p, 0, -1), (-p, 0, 1)], dtype=torch.float64)
    return (v / v.norm(dim=-1, keepdim=True) * radius).float()


def compute_motion_loss(model, body_points, body_point_counts, hf, min_point, dxdy, root_pos, root_rot, joint_rot, contacts, w_contact,
                        w_pen):
    """mdm_path.compute_motion_loss (PARC/motion_synthesis/procgen/mdm_path.py:31-127):
    forward-only penetration + contact terms of a batch of generated motions [B,T,...] against ONE whole terrain.
    Per body (:83-113): penetration = w_pen * sum over frames and points of max(air-column sdf, 0); contact = w_contact *
    sum over frames of (closest solid-column distance of the body's points, clamped at 0) * contact label.
    Returns (total [B], contact [B], pen [B]). Pinned by tests/golden/ref_path_loss.npz."""
    B, T = root_pos.shape[0], root_pos.shape[1]
    body_pos, body_rot = model.forward_kinematics(root_pos, root_rot, joint_rot)
    base_z = float(hf.min()) - 10.0                                                         # :78, :97
    hfb = hf.unsqueeze(0).expand(B, -1, -1)
    mpb = min_point.unsqueeze(0).expand(B, -1)
    pen = torch.zeros(B)
    con = torch.zeros(B)
    start = 0
    for b, cnt in enumerate(body_point_counts):
        cnt = int(cnt)
        pts = body_points[start:start + cnt]
        start += cnt
        world = O.quat_rotate(body_rot[..., b, :].unsqueeze(-2).expand(-1, -1, cnt, -1), pts.expand(B, T, cnt, 3)) + body_pos[..., b, :].unsqueeze(-2)
        flat = world.reshape(B, T * cnt, 3)
        neg = torch.clamp(points_hf_sdf(flat, hfb, mpb, dxdy, base_z, True), max=0.0)        # :92-99
        pen = pen + torch.sum(-neg, dim=-1) * w_pen
        pos = torch.clamp(points_hf_sdf(flat, hfb, mpb, dxdy, base_z, False), min=0.0).view(B, T, cnt)   # :101-108
        con = con + torch.sum(pos.min(dim=-1)[0] * contacts[..., b], dim=-1) * w_contact     # :109-113
    return con + pen, con, pen
