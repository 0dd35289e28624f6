"""GPU: library construction in one launch and the packed on-disk library."""
import pytest
import torch

from oracle import parc_oracle as O
from parc_b200 import synth
from tests import helpers as H
from tests.helpers import assert_close

pytestmark = pytest.mark.gpu
DEV = "cuda:0"

FRAME_TENSORS = ("_frame_root_pos", "_frame_root_rot", "_frame_root_vel", "_frame_root_ang_vel", "_frame_joint_rot", "_frame_dof_vel",
                 "_frame_contacts", "_motion_num_frames", "_motion_start_idx", "_motion_lengths", "_motion_loop_modes",
                 "_motion_root_pos_delta", "_motion_weights", "_motion_fps", "_motion_dt", "_blob_interp", "_blob_vel", "_blob_meta")


def mixed_clips():
    clips = synth.make_clips(synth.humanoid_spec(), 9, seed=31, min_frames=2, max_frames=70, wrap_fraction=0.3)
    for k, c in enumerate(clips):
        c["fps"] = (30, 60, 24)[k % 3]                      # per-clip frame rates
    one = dict(clips[0])                                     # a single-frame clip: velocities are zero
    for key in ("root_pos", "root_rot", "joint_rot", "contacts"):
        one[key] = one[key][:1].clone()
    clips.insert(4, one)
    return clips


def test_library_velocities_equal_per_clip_calls_and_oracle():
    model = H.product_model(DEV)
    clips = mixed_clips()
    lib = H.product_library(clips, model, DEV)
    rv, rav, dv = [], [], []
    for c in clips:
        a, b, d = model.frame_velocities(c["root_pos"].to(DEV), c["root_rot"].to(DEV), c["joint_rot"].to(DEV), c["fps"], 1.0 / c["fps"])
        rv.append(a); rav.append(b); dv.append(d)
    torch.cuda.synchronize()
    assert torch.equal(lib._frame_root_vel, torch.cat(rv)) and torch.equal(lib._frame_root_ang_vel, torch.cat(rav))
    assert torch.equal(lib._frame_dof_vel, torch.cat(dv))
    olib = O.Library(H.oracle_model(), [H.oracle_clip(c) for c in clips if c["root_pos"].shape[0] > 1])
    keep = torch.cat([torch.full((c["root_pos"].shape[0],), c["root_pos"].shape[0] > 1) for c in clips])
    assert_close(lib._frame_root_vel.cpu()[keep], olib.frame_root_vel, atol=2e-5, what="root_vel")
    assert_close(lib._frame_dof_vel.cpu()[keep], olib.frame_dof_vel, atol=2e-4, what="dof_vel")
    assert float(lib._frame_root_vel.cpu()[~keep].abs().max()) == 0.0


def test_packed_library_round_trip(tmp_path):
    model = H.product_model(DEV)
    clips = [H.golden_clip("clip_teaser_dm.npz"), H.golden_clip("clip_civilization.npz")] + mixed_clips()
    lib = H.product_library(clips, model, DEV)
    g = torch.Generator().manual_seed(3)
    lib._hf_mask_inds[0] = [torch.randint(0, 15, (int(torch.randint(1, 9, (1,), generator=g)), 2), generator=g).to(DEV)
                            for _ in range(int(lib._motion_num_frames[0]))]
    path = lib.save_packed(str(tmp_path / "lib.parcpack"))
    from parc_b200 import MotionLib
    lib2 = MotionLib.from_packed(path, model, DEV)
    torch.cuda.synchronize()
    for name in FRAME_TENSORS:
        assert torch.equal(getattr(lib, name), getattr(lib2, name)), name
    assert lib2.get_motion_names() == lib.get_motion_names() and lib2.num_motions() == len(clips)
    for a, b in zip(lib._terrains, lib2._terrains):
        assert torch.equal(a.hf, b.hf) and torch.equal(a.hf_maxmin, b.hf_maxmin) and torch.equal(a.min_point, b.min_point)
        assert torch.equal(a.dxdy, b.dxdy)
    assert lib2._hf_mask_inds[1] is None and len(lib2._hf_mask_inds[0]) == int(lib._motion_num_frames[0])
    assert all(torch.equal(x, y) for x, y in zip(lib._hf_mask_inds[0], lib2._hf_mask_inds[0]))
    ids = torch.randint(0, len(clips), (4096,), generator=g).to(DEV)
    t = torch.rand(4096, generator=g).to(DEV) * lib.get_motion_length(ids) * 1.5
    for x, y in zip(lib.calc_motion_frame(ids, t), lib2.calc_motion_frame(ids, t)):
        assert torch.equal(x, y)
    with pytest.raises(ValueError):
        bad = tmp_path / "bad.parcpack"
        bad.write_bytes(b"not a library")
        MotionLib.from_packed(str(bad), model, DEV)
