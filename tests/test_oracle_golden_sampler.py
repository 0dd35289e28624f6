"""CPU: the sampler oracle (oracle/parc_oracle_sampler.py) against the outputs of the unmodified reference
MDMHeightfieldContactMotionSampler.sample_motion_data (tests/golden/ref_mdm_sampler.npz, made by
tests/golden/make_golden_sampler.py). SURVEY.md section 8(f) rank 1."""
import numpy as np
import pytest
import torch

from oracle import parc_oracle as O
from oracle import parc_oracle_sampler as OS
from tests import helpers as H
from tests.helpers import assert_close, load_npz, t

SAMPLER_CLIPS = ["clip_teaser_dm.npz", "clip_teaser_modified.npz"]


def sampler_world():
    m = H.oracle_model()
    clips = [H.oracle_clip(H.golden_clip(n)) for n in SAMPLER_CLIPS]
    return m, O.Library(m, clips, weights=[1.0, 2.0])


def test_sampler_constants_match_reference():
    g = load_npz("ref_mdm_sampler.npz")
    cfg = OS.SamplerConfig()
    assert torch.equal(cfg.motion_times, t(g["motion_times"])) and cfg.seq_len == 15
    assert torch.equal(cfg.grid, t(g["grid"]))


@pytest.mark.parametrize("tag,style", [("root", OS.RELATIVE_TO_ROOT), ("floor", OS.RELATIVE_TO_ROOT_FLOOR)])
def test_sampler_oracle_matches_reference(tag, style):
    g = load_npz("ref_mdm_sampler.npz")
    m, lib = sampler_world()
    cfg = OS.SamplerConfig(relative_z_style=style)
    ids, starts = t(g[tag + "_ids"]), t(g[tag + "_starts"])
    ft = OS.future_times(lib, cfg, ids, starts, t(g[tag + "_u"]))
    out = OS.sample_motion_data(m, lib, cfg, ids, starts, future_t=ft, future_noise=t(g[tag + "_z"]))
    # heights come out of an integer cell lookup: a different cell would be a 0.1+ m jump, so the tolerance is tight
    assert_close(out["hfs"], t(g[tag + "_hfs"]), rtol=1e-6, atol=1e-6, what="hfs")
    for k in ("root_pos", "root_rot", "joint_pos", "joint_rot", "contacts", "future_pos", "future_rot"):
        assert_close(out[k], t(g[tag + "_" + k]), atol=2e-6, what=tag + " " + k)
    nohf = OS.sample_motion_data(m, lib, cfg, ids, starts, ret_hf_obs=False)
    assert_close(nohf["root_pos"], t(g[tag + "_nohf_root_pos"]), atol=2e-6, what="root_pos without heightmaps")
