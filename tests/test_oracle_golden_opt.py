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
        np.testing.assert_allclose(float(loss), float(d[tag + "_loss"]), rtol=2e-5)
        for g, k in zip(p, ("_g_root_pos", "_g_root_rot", "_g_joint_dof")):
            ref = d[tag + k]
            np.testing.assert_allclose(g.grad.numpy(), ref, rtol=1e-3, atol=1e-3 * float(np.abs(ref).max()))


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
