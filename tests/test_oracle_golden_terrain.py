"""CPU: the terrain post-processing oracle (oracle/parc_oracle_terrain.py) against the unmodified reference's outputs
(tests/golden/ref_hf_mask.npz). SURVEY.md section 8(f) rank 4."""
import numpy as np
import torch

from oracle import parc_oracle as O
from oracle import parc_oracle_terrain as OT
from tests import helpers as H
from tests.helpers import load_npz, t


def golden_inputs():
    g = load_npz("ref_hf_mask.npz")
    bp = load_npz("ref_body_points.npz")
    terrain = O.Terrain(t(g["hf"]), t(g["min_point"]), t(g["dxdy"]))
    return g, bp, terrain


def test_hf_mask_inds_and_extra_vals_match_reference():
    g, bp, terrain = golden_inputs()
    m = H.oracle_model()
    inds, mm = OT.compute_hf_extra_vals(m, t(bp["points"]), bp["counts"], terrain, t(g["root_pos"]), t(g["root_rot"]), t(g["joint_rot"]))
    offs = g["offsets"]
    assert len(inds) == len(offs) - 1
    for k, fr in enumerate(inds):
        assert np.array_equal(fr.numpy(), g["inds"][offs[k]:offs[k + 1]]), "frame %d" % k
    _, mins = OT.compute_hf_mask_inds(m, t(bp["points"]), bp["counts"], terrain, t(g["root_pos"]), t(g["root_rot"]), t(g["joint_rot"]))
    np.testing.assert_allclose(mins.numpy(), g["min_body_heights"], rtol=1e-6, atol=2e-6)
    assert np.array_equal(OT.compute_hf_mask_from_inds(terrain, inds).numpy(), g["hf_mask"])
    np.testing.assert_allclose(mm.numpy(), g["hf_maxmin"], rtol=1e-6, atol=2e-6)
