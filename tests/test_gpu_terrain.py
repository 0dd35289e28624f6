"""GPU: motion-aware terrain post-processing (parc_b200/terrain.py -> parc_b200_hf_mask_inds / parc_b200_hf_extra_vals)
against the reference's outputs (tests/golden/ref_hf_mask.npz) and the CPU oracle. SURVEY.md section 8(f) rank 4.
The index lists are integer results: bit-exact."""
import numpy as np
import pytest
import torch

from oracle import parc_oracle as O
from oracle import parc_oracle_terrain as OT
from parc_b200 import synth
from tests import helpers as H
from tests.helpers import load_npz, t

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def test_hf_mask_inds_and_extra_vals_match_reference_golden():
    from parc_b200 import geom_util, terrain as T
    from parc_b200.file_io import MotionFrames
    g = load_npz("ref_hf_mask.npz")
    model = H.product_model(DEV)
    pts = geom_util.get_char_point_samples(model)
    ter = H.product_terrain(t(g["hf"]), g["min_point"], float(g["dxdy"][0]), DEV)
    frames = MotionFrames(root_pos=t(g["root_pos"]).to(DEV), root_rot=t(g["root_rot"]).to(DEV), joint_rot=t(g["joint_rot"]).to(DEV))
    inds, mins = T.compute_hf_mask_inds(frames, ter, model, pts)
    offs = g["offsets"]
    assert len(inds) == len(offs) - 1
    for k, fr in enumerate(inds):
        assert np.array_equal(fr.cpu().numpy(), g["inds"][offs[k]:offs[k + 1]]), "frame %d" % k
    np.testing.assert_allclose(mins.cpu().numpy(), g["min_body_heights"], rtol=1e-6, atol=3e-6)
    assert np.array_equal(T.compute_hf_mask_from_inds(ter, inds).cpu().numpy(), g["hf_mask"])
    assert np.array_equal(T.compute_hf_mask(frames, ter, model, pts).cpu().numpy(), g["hf_mask"])
    inds2, ret = T.compute_hf_extra_vals(frames, ter, model, pts)
    assert all(torch.equal(a, b) for a, b in zip(inds, inds2))
    np.testing.assert_allclose(ret.hf_maxmin.cpu().numpy(), g["hf_maxmin"], rtol=1e-6, atol=3e-6)
    assert float(ter.hf_maxmin[0, 0, 0]) == 1.0             # the input terrain is left alone


def test_hf_mask_inds_match_oracle_on_a_long_synthetic_motion():
    """2,000 frames over a 64x64 terrain (4,096 cells = 128 bitmap words) with the character leaving the grid (clamped cells)."""
    from parc_b200 import geom_util, terrain as T
    from parc_b200.file_io import MotionFrames
    model = H.product_model(DEV)
    pts = geom_util.get_char_point_samples(model)
    c = synth.make_clips(synth.humanoid_spec(), 1, seed=8, min_frames=2000, max_frames=2000, terrain_dim=64, dx=0.25)[0]
    rp = c["root_pos"].clone()
    rp[:, 0] += torch.linspace(0.0, 6.0, 2000)              # walks off the +x edge
    hf = torch.as_tensor(c["hf"], dtype=torch.float32)
    ter = H.product_terrain(hf, c["min_point"], float(c["dx"]), DEV)
    inds, mins = T.compute_hf_mask_inds(MotionFrames(root_pos=rp.to(DEV), root_rot=c["root_rot"].to(DEV), joint_rot=c["joint_rot"].to(DEV)),
                                        ter, model, pts)
    oter = O.Terrain(hf, torch.as_tensor(c["min_point"], dtype=torch.float32), [float(c["dx"])] * 2)
    counts = [int(p.shape[0]) for p in pts]
    oinds, omins = OT.compute_hf_mask_inds(H.oracle_model(), torch.cat([p.cpu() for p in pts]), counts, oter, rp, c["root_rot"], c["joint_rot"])
    # a sample point within float rounding of a cell boundary may land in the neighbouring cell on one side only
    same = sum(int(a.shape == b.shape and torch.equal(a.cpu(), b)) for a, b in zip(inds, oinds))
    assert same >= len(oinds) - 2, "%d of %d frames differ" % (len(oinds) - same, len(oinds))
    close = torch.isclose(mins.cpu(), omins, rtol=1e-5, atol=1e-5)
    assert int((~close).sum()) <= 2
