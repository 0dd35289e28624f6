"""GPU parity of the kinematic-optimizer rows: CUDA kernels through the
C ABI against the golden vectors of the unmodified reference and against the pinned oracle."""
import numpy as np
import pytest
import torch

from oracle import parc_oracle_opt as OO
from tests import helpers as H

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
KIN_GEN = dict(w_root_pos=1.0, w_root_rot=10.0, w_joint_rot=1.0, w_smoothness=10.0, w_penetration=1000.0, w_contact=1000.0,
               w_sliding=10.0, w_body_constraints=1000.0, w_jerk=1000.0, max_jerk=1000.0)


def _t(x, dev="cpu"):
    return torch.as_tensor(np.asarray(x)).to(dev)


def _body_points(dev):
    bp = H.load_npz("ref_body_points.npz")
    pts, out, o = _t(bp["points"]), [], 0
    for c in bp["counts"]:
        out.append(pts[o:o + int(c)].to(dev))
        o += int(c)
    return out, bp


def test_points_hf_sdf_matches_reference_values_and_oracle_gradient():
    from parc_b200.terrain import points_hf_sdf
    d = H.load_npz("ref_hf_sdf.npz")
    for inverted, key in ((True, "sd_inverted"), (False, "sd_solid")):
        p = _t(d["points"], DEV).requires_grad_(True)
        sd = points_hf_sdf(p, _t(d["hf"], DEV), _t(d["min_box_center"], DEV), _t(d["dxdy"]), -10.0, inverted)
        torch.testing.assert_close(sd.detach().cpu(), _t(d[key]), rtol=1e-5, atol=2e-6)
        w = torch.linspace(0.5, 1.5, sd.numel(), device=DEV).reshape(sd.shape)
        (sd * w).sum().backward()
        po = _t(d["points"]).requires_grad_(True)
        so = OO.points_hf_sdf(po, _t(d["hf"]), _t(d["min_box_center"]), _t(d["dxdy"]), -10.0, inverted)
        (so * w.cpu()).sum().backward()
        torch.testing.assert_close(p.grad.cpu(), po.grad, rtol=1e-4, atol=1e-5)
        sd_r = points_hf_sdf(_t(d["points"], DEV), _t(d["hf"], DEV), _t(d["min_box_center"], DEV), _t(d["dxdy"]), -10.0, inverted,
                             radius=float(d["radius"]))                       # rounded columns (terrain_util.py:1782-1785)
        torch.testing.assert_close(sd_r.cpu(), _t(d[key + "_r"]), rtol=1e-5, atol=2e-6)


def _constraint_lists(d, tag, J, dev):
    """golden records -> the reference's structure: one list of BodyConstraint per body"""
    from parc_b200.motion_optimization import BodyConstraint
    out = [[] for _ in range(J)]
    for b, s0, e0, pt in zip(d[tag + "_body"], d[tag + "_start"], d[tag + "_end"], d[tag + "_point"]):
        c = BodyConstraint()
        c.start_frame_idx, c.end_frame_idx, c.constraint_point = int(s0), int(e0), _t(pt, dev)
        out[int(b)].append(c)
    return out


def _geoms(d):
    return [(str(k), (float(dm[0]) if str(k) == "SPHERE" else dm), off) for k, dm, off in zip(d["geom_kind"], d["geom_dims"], d["geom_offset"])]


@pytest.mark.parametrize("tag", ["full", "nocontact", "cons", "cons2"])
def test_motion_loss_matches_reference_golden(tag):
    """`cons` / `cons2`: with body constraints (motion_optimization.py:581-609)."""
    from parc_b200.motion_optimization import LossType, motion_terrain_contact_loss_localized
    d = H.load_npz("ref_opt_loss.npz")
    model = H.product_model(DEV)
    bc = _constraint_lists(d, tag, model.get_num_joints(), DEV) if tag.startswith("cons") else None
    pts, _ = _body_points(DEV)
    terrain = H.product_terrain(_t(d["hf"]), d["min_point"], float(d["dxdy"][0]), DEV)
    p = [_t(d[k], DEV).clone().requires_grad_(True) for k in ("tgt_root_pos", "tgt_root_rot", "tgt_joint_dof")]
    w = d[tag + "_weights"]
    kw = dict(zip(("w_root_pos", "w_root_rot", "w_joint_rot", "w_smoothness", "w_penetration", "w_contact", "w_sliding",
                   "w_body_constraints", "w_jerk", "max_jerk"), [float(x) for x in w]))
    loss, losses = motion_terrain_contact_loss_localized(
        p[0], p[1], p[2], _t(d["src_root_pos"], DEV), _t(d["src_root_rot_quat"], DEV), _t(d["src_joint_rot"], DEV),
        _t(d["src_body_vels"], DEV), _t(d["src_body_rot_vels"], DEV), _t(d["contacts"], DEV), terrain, pts, model,
        body_constraints=bc, **kw)
    loss.backward()
    got = np.array([losses[t] for t in list(LossType)[:9]])
    assert (got[7] > 1.0) == tag.startswith("cons")
    np.testing.assert_allclose(got, d[tag + "_terms"], rtol=2e-5, atol=2e-6)
    np.testing.assert_allclose(float(loss.detach()), float(d[tag + "_loss"]), rtol=2e-5)
    for g, k in zip(p, ("_g_root_pos", "_g_root_rot", "_g_joint_dof")):
        ref = d[tag + k]
        np.testing.assert_allclose(g.grad.cpu().numpy(), ref, rtol=1e-4, atol=1e-4 * float(np.abs(ref).max()))   # gradients <= 1e-4 relative (of the tensor scale)


def test_compute_approx_body_constraints_matches_reference():
    """motion_optimization.py:70-205 on the golden clip: the same contact runs per foot / hand, and constraint points
    (1,000 SGD steps through parc_b200_points_hf_sdf and its gradient) within 1e-5 m of the reference's."""
    from parc_b200.motion_optimization import compute_approx_body_constraints
    d = H.load_npz("ref_opt_loss.npz")
    model = H.product_model(DEV)
    terrain = H.product_terrain(_t(d["hf"]), d["min_point"], float(d["dxdy"][0]), DEV)
    bc = compute_approx_body_constraints(_t(d["src_root_pos"], DEV), _t(d["src_root_rot_quat"], DEV), _t(d["src_joint_rot"], DEV),
                                         _t(d["contacts"], DEV), model, terrain)
    assert len(bc) == model.get_num_joints()
    got = [(b, c.start_frame_idx, c.end_frame_idx) for b, lst in enumerate(bc) for c in lst]
    want = list(zip(d["cons_body"].tolist(), d["cons_start"].tolist(), d["cons_end"].tolist()))
    assert got == want and len(want) >= 3
    pts = torch.stack([c.constraint_point for lst in bc for c in lst]).cpu().numpy()
    np.testing.assert_allclose(pts, d["cons_point"], rtol=0, atol=1e-5)


def test_whole_terrain_motion_loss_matches_reference_golden():
    """motion_terrain_contact_loss (motion_optimization.py:207-423) through the mirror of the same name."""
    from parc_b200.motion_optimization import LossType, motion_terrain_contact_loss
    d = H.load_npz("ref_opt_loss.npz")
    model = H.product_model(DEV)
    pts, _ = _body_points(DEV)
    terrain = H.product_terrain(_t(d["hf"]), d["min_point"], float(d["dxdy"][0]), DEV)
    p = [_t(d[k], DEV).clone().requires_grad_(True) for k in ("whole_tgt_root_pos", "tgt_root_rot", "tgt_joint_dof")]
    loss, losses = motion_terrain_contact_loss(
        p[0], p[1], p[2], _t(d["src_root_pos"], DEV), _t(d["src_root_rot_quat"], DEV), _t(d["src_joint_rot"], DEV),
        _t(d["src_body_vels"], DEV), _t(d["src_body_rot_vels"], DEV), _t(d["contacts"], DEV), terrain, pts, model,
        body_constraints=None, **KIN_GEN)
    loss.backward()
    got = np.array([losses[t] for t in list(LossType)[:9]])
    np.testing.assert_allclose(got, d["whole_terms"], rtol=2e-5, atol=2e-6)
    np.testing.assert_allclose(float(loss.detach()), float(d["whole_loss"]), rtol=2e-5)
    for g, k in zip(p, ("_g_root_pos", "_g_root_rot", "_g_joint_dof")):
        ref = d["whole" + k]
        np.testing.assert_allclose(g.grad.cpu().numpy(), ref, rtol=1e-4, atol=1e-4 * float(np.abs(ref).max()))


def test_whole_terrain_and_windowed_losses_match_their_oracles():
    """32x32 terrains with 6 m walls every 7th row and the characters lifted 1.6 m, windowed and whole-terrain evaluation,
    each against the oracle in the same mode (on these inputs the two agree to rounding: a motion's frames share its
    largest window, which already holds every column that can be nearest)."""
    from parc_b200.motion_optimization import motion_loss_batched
    B, T = 2, 30
    d, dx = _synthetic_batch(B, T, seed=13)
    d["tgt_root_pos"] = d["tgt_root_pos"] + torch.tensor([0.0, 0.0, 1.6])
    d["hf"][:, ::7, :] += 6.0
    model = H.product_model(DEV)
    pts, bp = _body_points(DEV)
    om = H.oracle_model()
    w = [KIN_GEN[k] for k in ("w_root_pos", "w_root_rot", "w_joint_rot", "w_smoothness", "w_penetration", "w_contact", "w_sliding",
                              "w_body_constraints", "w_jerk", "max_jerk")]
    dxdy = torch.tensor([dx, dx])
    names = ("src_root_pos", "src_root_rot_quat", "src_joint_rot", "src_body_vels", "src_body_rot_vels", "contacts")
    got = {}
    for localized in (True, False):
        p = [d[k].to(DEV).clone().requires_grad_(True) for k in ("tgt_root_pos", "tgt_root_rot", "tgt_joint_dof")]
        loss, terms = motion_loss_batched(p[0], p[1], p[2], *[d[k].to(DEV) for k in names], d["hf"].to(DEV), d["min_point"].to(DEV), dxdy, pts,
                                          model, localized=localized, **KIN_GEN)
        loss.sum().backward()
        got[localized] = terms.cpu()
        for b in range(B):
            po = [d[k][b].clone().requires_grad_(True) for k in ("tgt_root_pos", "tgt_root_rot", "tgt_joint_dof")]
            lo, to = OO.motion_loss(om, _t(bp["points"]), bp["counts"], d["hf"][b], d["min_point"][b], dxdy, po[0], po[1], po[2],
                                    *[d[k][b] for k in names], w, localized=localized)
            lo.backward()
            np.testing.assert_allclose(terms[b].cpu().numpy(), to.numpy(), rtol=3e-5, atol=3e-6)
            for g, go in zip(p, po):
                ref = go.grad.numpy()
                np.testing.assert_allclose(g.grad[b].cpu().numpy(), ref, rtol=1e-4, atol=1e-4 * float(np.abs(ref).max()))
    assert float(got[False][:, 4].min()) > 1.0 and float(got[False][:, 5].min()) > 1.0         # both terrain terms active


def _synthetic_batch(B, T, seed):
    """B clips of the synthetic generator with their own 32x32 blocky terrains (BASELINE config 4 shape)."""
    from parc_b200 import synth
    spec = synth.humanoid_spec()
    clips = synth.make_clips(spec, B, seed=seed, min_frames=T, max_frames=T, wrap_fraction=0.0)
    om = H.oracle_model()
    g = torch.Generator().manual_seed(seed + 1)
    out = dict(src_root_pos=[], src_root_rot_quat=[], src_joint_rot=[], src_body_vels=[], src_body_rot_vels=[], contacts=[], hf=[],
               min_point=[], tgt_root_pos=[], tgt_root_rot=[], tgt_joint_dof=[])
    from oracle import parc_oracle as O
    for c in clips:
        rp, rr, jr = c["root_pos"], c["root_rot"], c["joint_rot"]
        bp, br = om.forward_kinematics(rp, rr, jr)
        out["src_root_pos"].append(rp); out["src_root_rot_quat"].append(rr); out["src_joint_rot"].append(jr)
        out["src_body_vels"].append(bp[1:] - bp[:-1]); out["src_body_rot_vels"].append(O.quat_diff_angle(br[1:], br[:-1]))
        out["contacts"].append(c["contacts"])
        out["hf"].append(torch.as_tensor(c["hf"], dtype=torch.float32)); out["min_point"].append(torch.as_tensor(c["min_point"], dtype=torch.float32))
        out["tgt_root_pos"].append(rp + 0.02 * torch.randn(T, 3, generator=g) + torch.tensor([0.0, 0.0, -0.04]))
        out["tgt_root_rot"].append(O.quat_to_exp_map(rr) + 0.03 * torch.randn(T, 3, generator=g))
        out["tgt_joint_dof"].append(om.rot_to_dof(jr) + 0.05 * torch.randn(T, 28, generator=g))
    return {k: torch.stack(v) for k, v in out.items()}, float(clips[0]["dx"])


def test_batched_motion_loss_vs_oracle_per_motion():
    from parc_b200.motion_optimization import motion_loss_batched
    B, T = 3, 40
    d, dx = _synthetic_batch(B, T, seed=31)
    model = H.product_model(DEV)
    pts, bp = _body_points(DEV)
    om = H.oracle_model()
    w = [KIN_GEN[k] for k in ("w_root_pos", "w_root_rot", "w_joint_rot", "w_smoothness", "w_penetration", "w_contact", "w_sliding",
                              "w_body_constraints", "w_jerk", "max_jerk")]
    p = [d[k].to(DEV).clone().requires_grad_(True) for k in ("tgt_root_pos", "tgt_root_rot", "tgt_joint_dof")]
    dxdy = torch.tensor([dx, dx])
    loss, terms = motion_loss_batched(p[0], p[1], p[2], *[d[k].to(DEV) for k in ("src_root_pos", "src_root_rot_quat", "src_joint_rot",
                                                                                 "src_body_vels", "src_body_rot_vels", "contacts")],
                                      d["hf"].to(DEV), d["min_point"].to(DEV), dxdy, pts, model, **KIN_GEN)
    loss.sum().backward()
    for b in range(B):
        po = [d[k][b].clone().requires_grad_(True) for k in ("tgt_root_pos", "tgt_root_rot", "tgt_joint_dof")]
        lo, to = OO.motion_loss(om, _t(bp["points"]), bp["counts"], d["hf"][b], d["min_point"][b], dxdy, po[0], po[1], po[2],
                                d["src_root_pos"][b], d["src_root_rot_quat"][b], d["src_joint_rot"][b], d["src_body_vels"][b],
                                d["src_body_rot_vels"][b], d["contacts"][b], w)
        lo.backward()
        np.testing.assert_allclose(terms[b].cpu().numpy(), to.numpy(), rtol=3e-5, atol=3e-6)
        np.testing.assert_allclose(float(loss[b]), float(lo), rtol=3e-5)
        for g, go in zip(p, po):
            ref = go.grad.numpy()
            np.testing.assert_allclose(g.grad[b].cpu().numpy(), ref, rtol=1e-4, atol=1e-4 * float(np.abs(ref).max()))   # gradients <= 1e-4 relative (of the tensor scale)


def test_batched_motion_loss_with_body_constraints_vs_oracle_per_motion():
    """Constraint records of a batch: a different number per motion (one motion has none, one passes None), feet and hands,
    overlapping and clipped ranges; every motion against the pinned oracle evaluated alone."""
    from parc_b200.motion_optimization import BodyConstraint, PackedBodyConstraints, motion_loss_batched
    B, T = 4, 36
    d, dx = _synthetic_batch(B, T, seed=57)
    g4 = H.load_npz("ref_opt_loss.npz")
    geoms = _geoms(g4)
    model = H.product_model(DEV)
    J = model.get_num_joints()
    pts, bp = _body_points(DEV)
    om = H.oracle_model()
    w = [KIN_GEN[k] for k in ("w_root_pos", "w_root_rot", "w_joint_rot", "w_smoothness", "w_penetration", "w_contact", "w_sliding",
                              "w_body_constraints", "w_jerk", "max_jerk")]
    tq = torch.stack([H.O.exp_map_to_quat(d["tgt_root_rot"][b]) for b in range(B)])
    gen = torch.Generator().manual_seed(3)
    ends = [om.body_names.index(n) for n in ("left_foot", "right_foot", "left_hand", "right_hand", "head")]   # head: a sphere WITH an offset
    records, lists = [], []
    for b in range(B):
        bpos, _ = om.forward_kinematics(d["tgt_root_pos"][b], tq[b], om.dof_to_rot(d["tgt_joint_dof"][b]))
        rec = []
        for k in range([5, 0, 9, 3][b]):
            body = ends[int(torch.randint(0, 5, (1,), generator=gen))]
            s0 = int(torch.randint(0, T - 2, (1,), generator=gen))
            e0 = s0 + int(torch.randint(0, 12, (1,), generator=gen))          # may run past T-1
            scale = 0.02 if k % 2 else 0.25
            rec.append((body, s0, e0, bpos[min(s0 + 1, T - 1), body] + scale * torch.randn(3, generator=gen)))
        records.append(rec)
        if b == 1:
            lists.append(None)
            continue
        per_body = [[] for _ in range(J)]
        for body, s0, e0, pt in rec:
            c = BodyConstraint()
            c.start_frame_idx, c.end_frame_idx, c.constraint_point = s0, e0, pt.to(DEV)
            per_body[body].append(c)
        lists.append(per_body)
    packed = PackedBodyConstraints(lists, model, torch.device(DEV))
    assert packed.num == 17
    p = [d[k].to(DEV).clone().requires_grad_(True) for k in ("tgt_root_pos", "tgt_root_rot", "tgt_joint_dof")]
    dxdy = torch.tensor([dx, dx])
    loss, terms = motion_loss_batched(p[0], p[1], p[2], *[d[k].to(DEV) for k in ("src_root_pos", "src_root_rot_quat", "src_joint_rot",
                                                                                 "src_body_vels", "src_body_rot_vels", "contacts")],
                                      d["hf"].to(DEV), d["min_point"].to(DEV), dxdy, pts, model, body_constraints=packed, **KIN_GEN)
    loss.sum().backward()
    assert float(terms[1, 7]) == 0.0 and float(terms[0, 7]) > 0.1
    for b in range(B):
        po = [d[k][b].clone().requires_grad_(True) for k in ("tgt_root_pos", "tgt_root_rot", "tgt_joint_dof")]
        # the reference keeps per-body lists: records sorted by body, stable
        rec = sorted(records[b], key=lambda r: r[0])
        lo, to = OO.motion_loss(om, _t(bp["points"]), bp["counts"], d["hf"][b], d["min_point"][b], dxdy, po[0], po[1], po[2],
                                d["src_root_pos"][b], d["src_root_rot_quat"][b], d["src_joint_rot"][b], d["src_body_vels"][b],
                                d["src_body_rot_vels"][b], d["contacts"][b], w, body_constraints=rec, geoms=geoms)
        lo.backward()
        np.testing.assert_allclose(terms[b].cpu().numpy(), to.numpy(), rtol=3e-5, atol=3e-6)
        np.testing.assert_allclose(float(loss[b]), float(lo), rtol=3e-5)
        for g, go in zip(p, po):
            ref = go.grad.numpy()
            np.testing.assert_allclose(g.grad[b].cpu().numpy(), ref, rtol=1e-4, atol=1e-4 * float(np.abs(ref).max()))


def test_config4_size_properties():
    """1,024 motions x 300 frames: finite, deterministic, and batch-invariant (a motion evaluated alone
    gives bit-identical terms and gradients to its row in the batch)."""
    from parc_b200.motion_optimization import motion_loss_batched
    B, T = 64, 300
    d, dx = _synthetic_batch(B, T, seed=77)
    rep = 16                                        # tile the 64 generated motions to 1,024
    dd = {k: v.repeat(rep, *([1] * (v.dim() - 1))).to(DEV) for k, v in d.items()}
    model = H.product_model(DEV)
    pts, _ = _body_points(DEV)
    dxdy = torch.tensor([dx, dx])
    names = ("src_root_pos", "src_root_rot_quat", "src_joint_rot", "src_body_vels", "src_body_rot_vels", "contacts")

    def run(sl):
        p = [dd[k][sl].clone().requires_grad_(True) for k in ("tgt_root_pos", "tgt_root_rot", "tgt_joint_dof")]
        loss, terms = motion_loss_batched(p[0], p[1], p[2], *[dd[k][sl] for k in names], dd["hf"][sl], dd["min_point"][sl], dxdy, pts,
                                          model, **KIN_GEN)
        loss.sum().backward()
        return loss.detach(), terms, [x.grad for x in p]

    loss, terms, grads = run(slice(0, B * rep))
    assert loss.shape == (B * rep,) and bool(torch.isfinite(loss).all()) and all(bool(torch.isfinite(g).all()) for g in grads)
    assert bool((terms[:, 4] > 0).any()) and bool((terms >= 0).all())
    loss2, terms2, grads2 = run(slice(0, B * rep))
    assert torch.equal(loss, loss2) and torch.equal(terms, terms2) and all(torch.equal(a, b) for a, b in zip(grads, grads2))
    assert torch.equal(terms[:B], terms[B:2 * B])   # tiled copies
    l1, t1, g1 = run(slice(5, 6))
    assert torch.equal(t1[0], terms[5]) and torch.equal(l1[0], loss[5]) and all(torch.equal(a[0], b[5]) for a, b in zip(g1, grads))


@pytest.mark.parametrize("tag", ["teaser", "civ"])
def test_path_motion_loss_matches_reference_golden(tag):
    """parc_b200.mdm_path.compute_motion_loss against the reference's outputs; `civ` has a 50x50 terrain
    (2,500 cells: the kernel reads it from global memory instead of staging it)."""
    from parc_b200 import geom_util, mdm_path
    from parc_b200.file_io import MotionFrames
    d = H.load_npz("ref_path_loss.npz")
    model = H.product_model(DEV)
    pts = geom_util.get_char_point_samples(model)
    g = lambda k: torch.as_tensor(d[tag + "_" + k]).to(DEV)
    terrain = H.product_terrain(torch.as_tensor(d[tag + "_hf"]), d[tag + "_min_point"], float(d[tag + "_dxdy"][0]), DEV)
    frames = MotionFrames(root_pos=g("root_pos"), root_rot=g("root_rot"), joint_rot=g("joint_rot"), contacts=g("contacts"))
    out = mdm_path.compute_motion_loss(frames, None, terrain, model, pts, w_contact=0.1, w_pen=0.1, w_path=0.0, verbose=False)
    torch.cuda.synchronize()
    for k in ("pen_loss", "contact_loss", "total_loss"):
        np.testing.assert_allclose(out[k].cpu().numpy(), d[tag + "_" + k], rtol=2e-5, atol=3e-6, err_msg=tag + " " + k)


def test_path_motion_loss_matches_oracle_on_synthetic_batch():
    """32 x 45 frames (the generation loop's batch, mdm_path.py:25) of synthetic clips on one blocky terrain."""
    from oracle import parc_oracle_opt as OO
    from parc_b200 import geom_util, mdm_path, synth
    from parc_b200.file_io import MotionFrames
    model = H.product_model(DEV)
    pts = geom_util.get_char_point_samples(model)
    B, T = 32, 45
    clips = synth.make_clips(synth.humanoid_spec(), 1, seed=5, min_frames=B * T, max_frames=B * T, wrap_fraction=0.0)
    c = clips[0]
    hf = torch.as_tensor(c["hf"], dtype=torch.float32)
    rp = c["root_pos"].reshape(B, T, 3).clone()
    ij = torch.round((rp[..., :2] - torch.as_tensor(c["min_point"])) / float(c["dx"])).long()
    ij[..., 0].clamp_(0, hf.shape[0] - 1); ij[..., 1].clamp_(0, hf.shape[1] - 1)
    rp[..., 2] = hf[ij[..., 0], ij[..., 1]] + 0.85 + 0.1 * torch.randn(B, 1, generator=torch.Generator().manual_seed(1))
    rr, jr, con = c["root_rot"].reshape(B, T, 4), c["joint_rot"].reshape(B, T, 14, 4), c["contacts"].reshape(B, T, 15)
    terrain = H.product_terrain(hf, c["min_point"], float(c["dx"]), DEV)
    out = mdm_path.compute_motion_loss(MotionFrames(root_pos=rp.to(DEV), root_rot=rr.to(DEV), joint_rot=jr.to(DEV), contacts=con.to(DEV)),
                                       None, terrain, model, pts, w_contact=0.1, w_pen=0.1)
    torch.cuda.synchronize()
    om = H.oracle_model()
    counts = [int(p.shape[0]) for p in pts]
    total, oc, op = OO.compute_motion_loss(om, torch.cat([p.cpu() for p in pts]), counts, hf, torch.as_tensor(c["min_point"], dtype=torch.float32),
                                           torch.tensor([float(c["dx"])] * 2), rp, rr, jr, con, 0.1, 0.1)
    assert float(op.max()) > 0 and float(oc.max()) > 0
    np.testing.assert_allclose(out["pen_loss"].cpu().numpy(), op.numpy(), rtol=3e-5, atol=1e-5)
    np.testing.assert_allclose(out["contact_loss"].cpu().numpy(), oc.numpy(), rtol=3e-5, atol=1e-5)
    np.testing.assert_allclose(out["total_loss"].cpu().numpy(), total.numpy(), rtol=3e-5, atol=1e-5)


def test_motion_contact_optimization_follows_the_reference_run():
    """motion_contact_optimization (motion_optimization.py:721-823): 40 Adam iterations with the kin_gen weights and the
    auto-derived body constraints, source 4 cm in the ground -- the optimised frames against the reference's own run.
    Adam divides every gradient component by its running magnitude, so components whose gradient is rounding noise still move
    by ~lr per step: the CPU oracle (the same torch ops in another order) ends up to 1.6e-3 m / 3.3e-3 (quaternions) away from
    the reference's own run after 40 iterations. The test therefore asks for 5e-3 on moves of up to 3.6 cm, plus a lower
    loss, not for trajectory identity."""
    from parc_b200.file_io import MotionFrames
    from parc_b200.motion_optimization import compute_approx_body_constraints, motion_contact_optimization
    d = H.load_npz("ref_opt_loss.npz")
    model = H.product_model(DEV)
    pts, _ = _body_points(DEV)
    terrain = H.product_terrain(_t(d["hf"]), d["min_point"], float(d["dxdy"][0]), DEV)
    bc = _constraint_lists(d, "cons", model.get_num_joints(), DEV)
    src = MotionFrames(root_pos=_t(d["optim_src_root_pos"], DEV), root_rot=_t(d["src_root_rot_quat"], DEV), joint_rot=_t(d["src_joint_rot"], DEV),
                       contacts=_t(d["contacts"], DEV))
    seen = []
    ret = motion_contact_optimization(src, pts, terrain, model, num_iters=40, step_size=0.001, body_constraints=bc, use_wandb=False,
                                      log=lambda it, loss, terms: seen.append((it, loss)), **KIN_GEN)
    assert [s[0] for s in seen] == [0, 25] and seen[1][1] < seen[0][1]          # logged every 25 iterations, loss going down
    moved = float((ret.root_pos.cpu() - _t(d["optim_src_root_pos"])).abs().max())
    assert moved > 0.02
    np.testing.assert_allclose(ret.root_pos.cpu().numpy(), d["optim_root_pos"], rtol=0, atol=5e-3)
    def quat_gap(a, b):          # q and -q are the same rotation (the root's exp-map angle sits next to pi here, where the sign flips)
        return np.minimum(np.abs(a - b).max(axis=-1), np.abs(a + b).max(axis=-1)).max()
    assert quat_gap(ret.root_rot.cpu().numpy(), d["optim_root_rot"]) < 5e-3
    assert quat_gap(ret.joint_rot.cpu().numpy(), d["optim_joint_rot"]) < 8e-3
    assert torch.equal(ret.contacts.cpu(), _t(d["contacts"]))
