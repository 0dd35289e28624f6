"""GPU: the CUDA MDM dataset sampler (parc_b200/motion_sampler.py -> parc_b200_mdm_sample_motion_data) against the
reference's own outputs (tests/golden/ref_mdm_sampler.npz) and against the CPU oracle on a larger seeded batch.
SURVEY.md section 8(f) rank 1. Tolerances: fp32 1e-5 relative (north_star); heights are integer cell lookups."""
import pytest
import torch

from oracle import parc_oracle as O
from oracle import parc_oracle_sampler as OS
from parc_b200 import synth
from tests import helpers as H
from tests.helpers import assert_close, load_npz, t
from tests.test_oracle_golden_sampler import SAMPLER_CLIPS, sampler_world

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def make_sampler(style="RELATIVE_TO_ROOT", clips=None, **over):
    from parc_b200.motion_sampler import MDMHeightfieldContactMotionSampler
    model = H.product_model(DEV)
    clips = [H.golden_clip(n) for n in SAMPLER_CLIPS] if clips is None else clips
    lib = H.product_library(clips, model, DEV)
    cfg = {"device": DEV, "char_file": synth.HUMANOID_JSON, "motion_lib_file": lib,
           "features": {"canonicalize_samples": True, "frame_components": ["ROOT_POS", "ROOT_ROT", "JOINT_POS", "JOINT_ROT", "CONTACTS"]},
           "sequence_fps": 30, "sequence_duration": 0.5, "autoregressive": True, "num_prev_states": 2, "future_pos_noise_scale": 0.05,
           "relative_z_style": style, "hf_augmentation_mode": "NONE",
           "heightmap": {"horizontal_scale": 0.2, "local_grid": {"num_x_neg": 10, "num_x_pos": 20, "num_y_neg": 15, "num_y_pos": 15},
                         "max_h": 3.0},
           "use_hf_augmentation": False, "angle_noise_scale": 0.01, "pos_noise_scale": 0.01, "future_window_min": 0.4,
           "future_window_max": 1.5}
    cfg.update(over)
    if cfg["use_hf_augmentation"]:
        # what the augmenter reads beyond the clip files: terrain hf_maxmin and the per-frame hf_mask_inds (clip 2 = WRAP copy of clip 0)
        mms, masks = H.aug_extras()
        for k, ter in enumerate(lib._terrains):
            ter.hf_maxmin = mms[k % 2].to(DEV)
        lib._hf_mask_inds = [masks[k % 2] for k in range(len(lib._terrains))]
    return MDMHeightfieldContactMotionSampler(cfg)


AUG_KEYS = dict(max_num_boxes=4, box_min_len=2, box_max_len=12, hf_maxpool_chance=0.15, hf_max_maxpool_size=10, hf_change_height_chance=0.1)


def compare(ret, hfs, tgt, ref, what, pos_atol=3e-6):
    """pos_atol: canonical positions are sums of metre-sized terms rotated by the inverse heading; the heading is
    atan2 of the rotated x axis and loses digits when the root pitches far out of the plane (its xy projection gets
    short), so over thousands of random frames a few positions move by ~1e-5 of the skeleton's 1-2 m extent."""
    from parc_b200.motion_sampler import MDMFrameType as F
    # a height is hf[cell] - base: the cell must be the reference's (a wrong one is off by decimetres)
    assert_close(hfs.cpu(), ref["hfs"], rtol=1e-6, atol=2e-6, what=what + " hfs")
    for key, ft in (("root_pos", F.ROOT_POS), ("root_rot", F.ROOT_ROT), ("joint_pos", F.JOINT_POS), ("joint_rot", F.JOINT_ROT),
                    ("contacts", F.CONTACTS)):
        assert_close(ret[ft].cpu(), ref[key], atol=pos_atol if key.endswith("pos") else 3e-6, what=what + " " + key)
    assert_close(tgt.future_pos.cpu(), ref["future_pos"], atol=pos_atol, what=what + " future_pos")
    assert_close(tgt.future_rot.cpu(), ref["future_rot"], atol=3e-6, what=what + " future_rot")


@pytest.mark.parametrize("tag,style", [("root", "RELATIVE_TO_ROOT"), ("floor", "RELATIVE_TO_ROOT_FLOOR")])
def test_sampler_matches_reference_golden(tag, style):
    g = load_npz("ref_mdm_sampler.npz")
    s = make_sampler(style)
    assert torch.equal(s._motion_times.cpu(), t(g["motion_times"])) and torch.equal(s._generic_heightmap.cpu(), t(g["grid"]))
    m, lib = sampler_world()
    ocfg = OS.SamplerConfig(relative_z_style=OS.RELATIVE_TO_ROOT if tag == "root" else OS.RELATIVE_TO_ROOT_FLOOR)
    ids, starts = t(g[tag + "_ids"]), t(g[tag + "_starts"])
    ft = OS.future_times(lib, ocfg, ids, starts, t(g[tag + "_u"]))
    ret, hfs, tgt = s.sample_motion_data(motion_ids=ids.to(DEV), motion_start_times=starts.to(DEV), future_times=ft.to(DEV),
                                         future_pos_noise=t(g[tag + "_z"]).to(DEV))
    torch.cuda.synchronize()
    compare(ret, hfs, tgt, {k: t(g[tag + "_" + k]) for k in ("hfs", "root_pos", "root_rot", "joint_pos", "joint_rot", "contacts",
                                                             "future_pos", "future_rot")}, "reference golden " + tag)
    from parc_b200.motion_sampler import MDMFrameType as F
    only = s.sample_motion_data(motion_ids=ids.to(DEV), motion_start_times=starts.to(DEV), ret_hf_obs=False, ret_target_info=False)
    assert isinstance(only, dict)
    assert_close(only[F.ROOT_POS].cpu(), t(g[tag + "_nohf_root_pos"]), atol=3e-6, what="root_pos without heightmaps")


@pytest.mark.parametrize("style,ostyle", [("RELATIVE_TO_ROOT", OS.RELATIVE_TO_ROOT), ("RELATIVE_TO_ROOT_FLOOR", OS.RELATIVE_TO_ROOT_FLOOR)])
def test_sampler_matches_oracle_on_a_seeded_batch(style, ostyle):
    """4,099 samples (not a multiple of the 16 groups per block) over the two clips + a WRAP copy of the first."""
    clips = [H.golden_clip(n) for n in SAMPLER_CLIPS]
    wrap = dict(clips[0]); wrap["loop_mode"] = O.WRAP
    clips.append(wrap)
    s = make_sampler(style, clips=clips)
    m = H.oracle_model()
    lib = O.Library(m, [H.oracle_clip(c) for c in clips])
    ocfg = OS.SamplerConfig(relative_z_style=ostyle)
    g = torch.Generator().manual_seed(77)
    n = 4099
    ids = torch.randint(0, 3, (n,), generator=g)
    lengths = lib.lengths[ids]
    starts = torch.rand(n, generator=g) * torch.where(ids == 2, lengths * 2.5, lengths - 0.5)     # WRAP clips may start anywhere
    ft = OS.future_times(lib, ocfg, ids, starts, torch.rand(n, generator=g))
    z = torch.randn(n, 3, generator=g)
    ref = OS.sample_motion_data(m, lib, ocfg, ids, starts, future_t=ft, future_noise=z)
    ret, hfs, tgt = s.sample_motion_data(motion_ids=ids.to(DEV), motion_start_times=starts.to(DEV), future_times=ft.to(DEV),
                                         future_pos_noise=z.to(DEV))
    torch.cuda.synchronize()
    compare(ret, hfs, tgt, ref, "oracle " + style, pos_atol=2e-5)
    assert_close(s._last_canon[:, 0:3].cpu(), ref["canon_pos"], what="canon pos")
    assert_close(s._last_canon[:, 7:11].cpu(), ref["canon_heading_quat_inv"], atol=2e-6, what="canon heading quat inv")


def test_sampler_public_call_draws_on_the_device_and_covers_every_window():
    """sample_motion_data(num_samples=...) with its own draws: shapes, finiteness, invariants of the canonical frame."""
    from parc_b200.motion_sampler import MDMFrameType as F
    s = make_sampler("RELATIVE_TO_ROOT_FLOOR")
    torch.manual_seed(3)
    ret, hfs, tgt = s.sample_motion_data(num_samples=1000)
    torch.cuda.synchronize()
    T, J = s.get_seq_len(), 15
    assert ret[F.ROOT_POS].shape == (1000, T, 3) and ret[F.JOINT_POS].shape == (1000, T, J - 1, 3) and hfs.shape == (1000, 31, 31)
    for v in list(ret.values()) + [hfs, tgt.future_pos, tgt.future_rot]:
        assert bool(torch.isfinite(v).all())
    # the reference frame sits at the origin of the canonical frame with zero heading; the cell under it is height 0
    assert float(ret[F.ROOT_POS][:, s._ref_frame_idx, 0:2].abs().max()) == 0.0
    assert float(hfs[:, s._num_x_neg, s._num_y_neg].abs().max()) == 0.0
    assert float(hfs.max()) <= 3.0 and float(hfs.min()) >= -3.0
    q = ret[F.ROOT_ROT][:, s._ref_frame_idx]
    heading = torch.atan2(2 * (q[:, 3] * q[:, 2] + q[:, 0] * q[:, 1]), 1 - 2 * (q[:, 1] ** 2 + q[:, 2] ** 2))
    assert float(heading.abs().median()) < 1e-5 and float(heading.abs().max()) < 2e-3    # ill-conditioned for pitched roots
    seqs = s.get_motion_sequences_for_id(1)
    assert seqs[F.JOINT_ROT].shape == (142 - T, T, J - 1, 4)


def _aug_sampler_from_golden_cfg(g, tag, **over):
    import yaml
    c = yaml.safe_load(str(g[tag + "_cfg"]))          # the reference-side config of the case (aug_default = parc_1_train_gen_default.yaml)
    keys = {k: c[k] for k in ("use_hf_augmentation", "hf_augmentation_mode", "max_num_boxes", "box_min_len", "box_max_len",
                              "hf_maxpool_chance", "hf_max_maxpool_size", "hf_change_height_chance", "features")}
    keys.update(over)
    return make_sampler(c["relative_z_style"], **keys), c


@pytest.mark.parametrize("tag", ["aug_default", "aug_boxes", "aug_noise"])
def test_augmented_heightmaps_match_reference_golden(tag):
    """parc_b200_mdm_augment_hfs / parc_b200_mdm_floor_heights with the reference's own draws as records."""
    from parc_b200.motion_sampler import MDMFrameType as F
    g = load_npz("ref_mdm_sampler.npz")
    s, c = _aug_sampler_from_golden_cfg(g, tag)
    ids, starts = t(g[tag + "_ids"]).to(DEV), t(g[tag + "_starts"]).to(DEV)
    kw = {}
    if c["hf_augmentation_mode"] == "NOISE":
        kw["hf_noise"] = t(g[tag + "_noise"]).to(DEV)
    else:
        kw["hf_aug_records"] = (t(g[tag + "_rec_i"]), t(g[tag + "_rec_f"]))
    ret, hfs, tgt = s.sample_motion_data(motion_ids=ids, motion_start_times=starts, **kw)
    torch.cuda.synchronize()
    assert_close(hfs.cpu(), t(g[tag + "_hfs"]), rtol=1e-6, atol=2e-6, what=tag + " hfs")
    assert_close(ret[F.ROOT_POS].cpu(), t(g[tag + "_root_pos"]), atol=3e-6, what=tag + " root_pos")
    if (tag + "_floor_heights") in g:
        assert ret[F.FLOOR_HEIGHTS].shape == (24, 24, 15, 1)
        assert_close(ret[F.FLOOR_HEIGHTS].cpu(), t(g[tag + "_floor_heights"]), rtol=1e-6, atol=2e-6, what=tag + " floor heights")


@pytest.mark.parametrize("tag", ["aug_default", "aug_boxes"])
def test_seeded_public_call_reproduces_the_reference_augmentation(tag):
    """No draws passed in: with Python's `random` and torch seeded as the golden run was, the mirror's own draws
    (_draw_box_aug_records) are the reference's, so the heightmaps come out the same. aug_default runs the values of
    data/configs/parc_1_train_gen_default.yaml unmodified."""
    import random
    g = load_npz("ref_mdm_sampler.npz")
    s, _ = _aug_sampler_from_golden_cfg(g, tag)
    seeds = g[tag + "_seeds"]
    random.seed(int(seeds[0])); torch.manual_seed(int(seeds[1]))
    ret, hfs, tgt = s.sample_motion_data(motion_ids=t(g[tag + "_ids"]).to(DEV), motion_start_times=t(g[tag + "_starts"]).to(DEV))
    torch.cuda.synchronize()
    assert_close(hfs.cpu(), t(g[tag + "_hfs"]), rtol=1e-6, atol=2e-6, what=tag + " hfs (own draws)")


@pytest.mark.parametrize("mode,style,ostyle", [("MAXPOOL_AND_BOXES", "RELATIVE_TO_ROOT_FLOOR", OS.RELATIVE_TO_ROOT_FLOOR),
                                               ("MAXPOOL_AND_BOXES", "RELATIVE_TO_ROOT", OS.RELATIVE_TO_ROOT),
                                               ("NOISE", "RELATIVE_TO_ROOT_FLOOR", OS.RELATIVE_TO_ROOT_FLOOR)])
def test_augmenter_matches_oracle_on_a_seeded_batch(mode, style, ostyle):
    """1,031 windows over the two clips + a WRAP copy (windows that start past the last frame have no mask rows), every
    pool kind / size up to 10, up to 6 boxes; FLOOR_HEIGHTS on a 37-window slice (the output is quadratic in the batch)."""
    import random
    from parc_b200.motion_sampler import MDMFrameType as F
    clips = [H.golden_clip(n) for n in SAMPLER_CLIPS]
    wrap = dict(clips[0]); wrap["loop_mode"] = O.WRAP
    clips.append(wrap)
    keys = dict(AUG_KEYS, max_num_boxes=6, hf_maxpool_chance=0.5, hf_change_height_chance=0.2)
    s = make_sampler(style, clips=clips, use_hf_augmentation=True, hf_augmentation_mode=mode, **keys)
    m = H.oracle_model()
    lib = O.Library(m, [H.oracle_clip(c) for c in clips])
    mms, masks = H.aug_extras()
    ocfg = OS.SamplerConfig(relative_z_style=ostyle)
    g = torch.Generator().manual_seed(91)
    n = 1031
    ids = torch.randint(0, 3, (n,), generator=g)
    lengths = lib.lengths[ids]
    starts = torch.rand(n, generator=g) * torch.where(ids == 2, lengths * 2.5, lengths - 0.5)
    aug = dict(mode=mode, hf_maxmins=[mms[0], mms[1], mms[0]], mask_inds=[masks[0], masks[1], masks[0]])
    kw = {}
    if mode == "NOISE":
        aug["noise"] = torch.rand(n, 31, 31, generator=g)
        kw["hf_noise"] = aug["noise"].to(DEV)
    else:
        random.seed(5); torch.manual_seed(6)
        aug["rec_i"], aug["rec_f"] = OS.draw_box_aug_records(n, (31, 31), 3.0, 6, 2, 12, 0.5, 10, 0.2)
        kw["hf_aug_records"] = (aug["rec_i"], aug["rec_f"])
    ref = OS.sample_motion_data(m, lib, ocfg, ids, starts, aug=aug)
    ret, hfs = s.sample_motion_data(motion_ids=ids.to(DEV), motion_start_times=starts.to(DEV), ret_target_info=False, **kw)
    torch.cuda.synchronize()
    # A grid point within an ulp of a terrain-cell border can land in the neighbouring cell (sinf / cosf of the heading differ in
    # the last bit between CUDA and the host's libm); the two cells may share a height and differ in their bounds, and a
    # max-pool spreads a single difference. Windows with a grid point that close to a border are left out.
    pts = OS.grid_points_world(ocfg, ref["canon_pos"], ref["canon_rot"])
    frac = torch.stack([(pts[i] - lib.terrains[int(ids[i])].min_point) / lib.terrains[int(ids[i])].dxdy for i in range(n)])
    same = (((frac - torch.floor(frac)) - 0.5).abs() > 1e-5).flatten(1).all(dim=1)          # 27 of the 1,031 windows
    assert int((~same).sum()) <= n // 20, "%d windows sit on a cell border" % int((~same).sum())
    plain_ref = OS.sample_motion_data(m, lib, ocfg, ids, starts)["hfs"]
    assert_close(hfs.cpu()[same], ref["hfs"][same], rtol=1e-6, atol=2e-6, what="augmented hfs vs oracle " + mode)
    assert float((hfs.cpu()[same] - plain_ref[same]).abs().max()) > 0.5, "the augmenter must have changed the heightmaps"
    # FLOOR_HEIGHTS on a slice
    s2 = make_sampler(style, clips=clips, use_hf_augmentation=True, hf_augmentation_mode=mode,
                      features={"canonicalize_samples": True, "frame_components": ["ROOT_POS", "FLOOR_HEIGHTS", "JOINT_VEL"]}, **keys)
    k = 37
    kw2 = {a: (b[:k] if not isinstance(b, tuple) else (b[0][:k], b[1][:k])) for a, b in kw.items()}
    ret2, hfs2 = s2.sample_motion_data(motion_ids=ids[:k].to(DEV), motion_start_times=starts[:k].to(DEV), ret_target_info=False, **kw2)
    torch.cuda.synchronize()
    assert set(ret2.keys()) == {F.ROOT_POS, F.FLOOR_HEIGHTS}            # JOINT_VEL: accepted, no entry (as upstream)
    want = OS.floor_heights(ocfg, ref["hfs"][:k], ref["root_pos"][:k])
    assert ret2[F.FLOOR_HEIGHTS].shape == (k, k, 15, 1)
    # a root within 1e-6 of a cell border may fall in the neighbouring cell: compare all but a handful
    bad = (ret2[F.FLOOR_HEIGHTS].cpu() - want).abs() > 2e-6
    assert int(bad.sum()) <= max(2, want.numel() // 5000), "floor heights: %d of %d differ" % (int(bad.sum()), want.numel())
