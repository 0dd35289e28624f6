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


AUG_CASES = [("aug_default", OS.RELATIVE_TO_ROOT, "MAXPOOL_AND_BOXES"), ("aug_boxes", OS.RELATIVE_TO_ROOT_FLOOR, "MAXPOOL_AND_BOXES"),
             ("aug_noise", OS.RELATIVE_TO_ROOT, "NOISE")]


@pytest.mark.parametrize("tag,style,mode", AUG_CASES)
def test_augmented_heightmaps_match_reference(tag, style, mode):
    """The heightfield augmenter (:304-364, :390-400) and FLOOR_HEIGHTS (:267-278) with the reference's own draws."""
    g = load_npz("ref_mdm_sampler.npz")
    m, lib = sampler_world()
    mms, masks = H.aug_extras()
    cfg = OS.SamplerConfig(relative_z_style=style)
    ids, starts = t(g[tag + "_ids"]), t(g[tag + "_starts"])
    aug = dict(mode=mode, hf_maxmins=mms, mask_inds=masks)
    if mode == "NOISE":
        aug["noise"] = t(g[tag + "_noise"])
    else:
        aug["rec_i"], aug["rec_f"] = t(g[tag + "_rec_i"]), t(g[tag + "_rec_f"])
    ft = OS.future_times(lib, cfg, ids, starts, t(g[tag + "_u"]))
    out = OS.sample_motion_data(m, lib, cfg, ids, starts, future_t=ft, future_noise=t(g[tag + "_z"]), aug=aug,
                                want_floor_heights=(tag + "_floor_heights") in g)
    plain = OS.sample_motion_data(m, lib, cfg, ids, starts)["hfs"]
    assert float((plain - t(g[tag + "_hfs"])).abs().max()) > 0.1, "the augmenter must have changed something"
    assert_close(out["hfs"], t(g[tag + "_hfs"]), rtol=1e-6, atol=1e-6, what=tag + " hfs")
    assert_close(out["root_pos"], t(g[tag + "_root_pos"]), atol=2e-6, what=tag + " root_pos")
    assert_close(out["future_pos"], t(g[tag + "_future_pos"]), atol=2e-6, what=tag + " future_pos")
    if (tag + "_floor_heights") in g:
        assert out["floor_heights"].shape == (24, 24, 15, 1)
        assert_close(out["floor_heights"], t(g[tag + "_floor_heights"]), rtol=1e-6, atol=1e-6, what=tag + " floor heights")


def test_box_aug_draws_replay_the_reference_order():
    """draw_box_aug_records consumes `random` and torch's generator as the reference does: re-seeding reproduces the
    stored records (which the golden heightmaps above were checked with)."""
    import random
    import yaml
    g = load_npz("ref_mdm_sampler.npz")
    for tag in ("aug_default", "aug_boxes"):
        c = yaml.safe_load(str(g[tag + "_cfg"]))
        seeds = g[tag + "_seeds"]
        random.seed(int(seeds[0])); torch.manual_seed(int(seeds[1]))
        rec_i, rec_f = OS.draw_box_aug_records(24, (31, 31), c["heightmap"]["max_h"], c["max_num_boxes"], c["box_min_len"], c["box_max_len"],
                                               c["hf_maxpool_chance"], c["hf_max_maxpool_size"], c["hf_change_height_chance"])
        assert torch.equal(rec_i, t(g[tag + "_rec_i"])) and torch.equal(rec_f, t(g[tag + "_rec_f"]))
