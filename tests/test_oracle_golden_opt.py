"""CPU: the optimizer-row oracle (oracle/parc_oracle_opt.py) against the vectors the unmodified
reference produced (tests/golden/make_golden_opt.py)."""
import numpy as np
import torch

from oracle import parc_oracle_opt as OO
from tests import helpers as H


def _t(x):
    return torch.as_tensor(np.asarray(x))


def test_points_hf_sdf_matches_reference():
    d = H.load_npz("ref_hf_sdf.npz")
    for inverted, key in ((True, "sd_inverted"), (False, "sd_solid")):
        got = OO.points_hf_sdf(_t(d["points"]), _t(d["hf"]), _t(d["min_box_center"]), _t(d["dxdy"]), -10.0, inverted)
        torch.testing.assert_close(got, _t(d[key]), rtol=1e-6, atol=1e-6)
        got = OO.points_hf_sdf(_t(d["points"]), _t(d["hf"]), _t(d["min_box_center"]), _t(d["dxdy"]), -10.0, inverted, radius=float(d["radius"]))
        torch.testing.assert_close(got, _t(d[key + "_r"]), rtol=1e-6, atol=1e-6)


def test_motion_loss_and_gradients_match_reference():
    d = H.load_npz("ref_opt_loss.npz")
    bp = H.load_npz("ref_body_points.npz")
    model = H.oracle_model()
    for tag in ("full", "nocontact"):
        p = [_t(d[k]).clone().requires_grad_(True) for k in ("tgt_root_pos", "tgt_root_rot", "tgt_joint_dof")]
        loss, terms = OO.motion_loss(model, _t(bp["points"]), bp["counts"], _t(d["hf"]), _t(d["min_point"]), _t(d["dxdy"]), p[0], p[1], p[2],
                                     _t(d["src_root_pos"]), _t(d["src_root_rot_quat"]), _t(d["src_joint_rot"]), _t(d["src_body_vels"]),
                                     _t(d["src_body_rot_vels"]), _t(d["contacts"]), d[tag + "_weights"])
        loss.backward()
        np.testing.assert_allclose(terms.numpy(), d[tag + "_terms"], rtol=2e-5, atol=1e-6)
        np.testing.assert_allclose(float(loss.detach()), float(d[tag + "_loss"]), rtol=2e-5)
        for g, k in zip(p, ("_g_root_pos", "_g_root_rot", "_g_joint_dof")):
            ref = d[tag + k]
            np.testing.assert_allclose(g.grad.numpy(), ref, rtol=1e-3, atol=1e-3 * float(np.abs(ref).max()))


def test_whole_terrain_motion_loss_matches_reference():
    """motion_terrain_contact_loss (motion_optimization.py:207-423): no per-frame windows."""
    d = H.load_npz("ref_opt_loss.npz")
    bp = H.load_npz("ref_body_points.npz")
    model = H.oracle_model()
    p = [_t(d[k]).clone().requires_grad_(True) for k in ("whole_tgt_root_pos", "tgt_root_rot", "tgt_joint_dof")]
    loss, terms = OO.motion_loss(model, _t(bp["points"]), bp["counts"], _t(d["hf"]), _t(d["min_point"]), _t(d["dxdy"]), p[0], p[1], p[2],
                                 _t(d["src_root_pos"]), _t(d["src_root_rot_quat"]), _t(d["src_joint_rot"]), _t(d["src_body_vels"]),
                                 _t(d["src_body_rot_vels"]), _t(d["contacts"]), d["whole_weights"], localized=False)
    loss.backward()
    np.testing.assert_allclose(terms.numpy(), d["whole_terms"], rtol=2e-5, atol=1e-6)
    for g, k in zip(p, ("_g_root_pos", "_g_root_rot", "_g_joint_dof")):
        ref = d["whole" + k]
        np.testing.assert_allclose(g.grad.numpy(), ref, rtol=1e-3, atol=1e-3 * float(np.abs(ref).max()))


def _geoms(d):
    return [(str(k), (float(dm[0]) if str(k) == "SPHERE" else dm), off) for k, dm, off in zip(d["geom_kind"], d["geom_dims"], d["geom_offset"])]


def _records(d, tag):
    return [(int(b), int(s), int(e), p) for b, s, e, p in zip(d[tag + "_body"], d[tag + "_start"], d[tag + "_end"], d[tag + "_point"])]


def test_motion_loss_with_body_constraints_matches_reference():
    """motion_optimization.py:581-609: `cons` = the reference's own compute_approx_body_constraints lists, `cons2` = hand-made
    lists exercising every branch (tests/golden/make_golden_opt.py)."""
    d = H.load_npz("ref_opt_loss.npz")
    bp = H.load_npz("ref_body_points.npz")
    model = H.oracle_model()
    for tag in ("cons", "cons2"):
        p = [_t(d[k]).clone().requires_grad_(True) for k in ("tgt_root_pos", "tgt_root_rot", "tgt_joint_dof")]
        loss, terms = OO.motion_loss(model, _t(bp["points"]), bp["counts"], _t(d["hf"]), _t(d["min_point"]), _t(d["dxdy"]), p[0], p[1], p[2],
                                     _t(d["src_root_pos"]), _t(d["src_root_rot_quat"]), _t(d["src_joint_rot"]), _t(d["src_body_vels"]),
                                     _t(d["src_body_rot_vels"]), _t(d["contacts"]), d[tag + "_weights"],
                                     body_constraints=_records(d, tag), geoms=_geoms(d))
        loss.backward()
        assert d[tag + "_terms"][7] > 1.0
        np.testing.assert_allclose(terms.numpy(), d[tag + "_terms"], rtol=2e-5, atol=1e-6)
        np.testing.assert_allclose(float(loss.detach()), float(d[tag + "_loss"]), rtol=2e-5)
        for g, k in zip(p, ("_g_root_pos", "_g_root_rot", "_g_joint_dof")):
            ref = d[tag + k]
            np.testing.assert_allclose(g.grad.numpy(), ref, rtol=1e-3, atol=1e-3 * float(np.abs(ref).max()))


def test_approx_body_constraints_match_reference():
    """compute_approx_body_constraints (:70-205) on the golden clip: same runs, surface-projected points within 1e-5 m."""
    d = H.load_npz("ref_opt_loss.npz")
    model = H.oracle_model()
    bodies = [model.body_names.index(n) for n in ("left_foot", "right_foot", "left_hand", "right_hand")]
    offs = [d["geom_offset"][bodies[0]], d["geom_offset"][bodies[1]], None, None]
    rec = OO.approx_body_constraints(model, _t(d["src_root_pos"]), _t(d["src_root_rot_quat"]), _t(d["src_joint_rot"]), _t(d["contacts"]),
                                     _t(d["hf"]), _t(d["min_point"]), _t(d["dxdy"]), bodies, offs)
    # the reference returns per-body lists, flattened in body order; the oracle emits feet and hands in its own order
    rec.sort(key=lambda r: (r[0], r[1]))
    want = _records(d, "cons")
    assert [r[:3] for r in rec] == [w[:3] for w in want] and len(want) >= 3
    np.testing.assert_allclose(np.stack([r[3].numpy() for r in rec]), np.stack([w[3] for w in want]), rtol=0, atol=1e-5)


def test_path_motion_loss_matches_reference():
    """mdm_path.compute_motion_loss against tests/golden/ref_path_loss.npz."""
    d = H.load_npz("ref_path_loss.npz")
    bp = H.load_npz("ref_body_points.npz")
    model = H.oracle_model()
    for tag in ("teaser", "civ"):
        total, con, pen = OO.compute_motion_loss(model, _t(bp["points"]), bp["counts"], _t(d[tag + "_hf"]), _t(d[tag + "_min_point"]),
                                                 _t(d[tag + "_dxdy"]), _t(d[tag + "_root_pos"]), _t(d[tag + "_root_rot"]),
                                                 _t(d[tag + "_joint_rot"]), _t(d[tag + "_contacts"]), 0.1, 0.1)
        np.testing.assert_allclose(pen.numpy(), d[tag + "_pen_loss"], rtol=2e-5, atol=2e-6)
        np.testing.assert_allclose(con.numpy(), d[tag + "_contact_loss"], rtol=2e-5, atol=2e-6)
        np.testing.assert_allclose(total.numpy(), d[tag + "_total_loss"], rtol=2e-5, atol=2e-6)
