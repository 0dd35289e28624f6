"""Golden vectors for the MDM dataset sampler: runs the UNMODIFIED reference
MDMHeightfieldContactMotionSampler.sample_motion_data (PARC/motion_generator/
mdm_heightfield_contact_motion_sampler.py:197-302) on CPU over a two-motion library built from the bundled
clips that carry hf_mask_inds, for both relative_z_style settings, with the augmenter off.

Augmenter cases (:304-330, :332-364, :390-400; terrain_util.py:861-914, :1509-1536) and FLOOR_HEIGHTS (:267-278):
  aug_default  data/configs/parc_1_train_gen_default.yaml as it is (only device and the two file paths replaced)
  aug_boxes    MAXPOOL_AND_BOXES with the chances raised so that every branch fires, RELATIVE_TO_ROOT_FLOOR, plus the
               FLOOR_HEIGHTS frame component
  aug_noise    NOISE mode
Python's `random` and torch's generator are seeded before each call; the draws the reference makes are re-derived
afterwards by re-seeding and replaying them in the reference's order (oracle/parc_oracle_sampler.draw_box_aug_records),
and stored next to the outputs. clip_aug_extras.npz holds what the augmenter reads from the two clips beyond
clip_teaser_*.npz: terrain hf_maxmin and the per-frame hf_mask_inds (CSR).

Run here (authoring container) only:  python tests/golden/make_golden_sampler.py  -> ref_mdm_sampler.npz
The random draws the reference makes inside the call (future times, future-position noise) are recorded by
re-seeding torch's generator and repeating the draws in the reference's order.
"""
import os
import random
import sys
import tempfile

import numpy as np
import torch
import yaml

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from oracle import ref_import  # noqa: E402

ref_import.setup()
torch.set_num_threads(1)

import parc.motion_generator.mdm_heightfield_contact_motion_sampler as rs  # noqa: E402
from parc.motion_generator.diffusion_util import MDMFrameType  # noqa: E402

CLIPS = ["dec2024_teaser_717_1_opt_dm.pkl", "dec2024_teaser_717_1_modified_opt.pkl"]


def make_cfg(motion_yaml, style):
    return {
        "device": "cpu", "char_file": ref_import.humanoid_xml(), "motion_lib_file": motion_yaml,
        "features": {"canonicalize_samples": True, "frame_components": ["ROOT_POS", "ROOT_ROT", "JOINT_POS", "JOINT_ROT", "CONTACTS"]},
        "sequence_fps": 30, "sequence_duration": 0.5, "autoregressive": True, "num_prev_states": 2,
        "future_pos_noise_scale": 0.05, "relative_z_style": style, "hf_augmentation_mode": "NONE",
        "heightmap": {"horizontal_scale": 0.2, "local_grid": {"num_x_neg": 10, "num_x_pos": 20, "num_y_neg": 15, "num_y_pos": 15},
                      "max_h": 3.0},
        "use_hf_augmentation": False, "angle_noise_scale": 0.01, "pos_noise_scale": 0.01,
        "future_window_min": 0.4, "future_window_max": 1.5,
    }


def main():
    tmp = tempfile.mkdtemp()
    motion_yaml = os.path.join(tmp, "motions.yaml")
    with open(motion_yaml, "w") as f:
        yaml.safe_dump({"motions": [{"file": ref_import.clip(c), "weight": 1.0 + i} for i, c in enumerate(CLIPS)]}, f)
    out = {}
    g = torch.Generator().manual_seed(5)
    n = 24
    for style in ("RELATIVE_TO_ROOT", "RELATIVE_TO_ROOT_FLOOR"):
        s = rs.MDMHeightfieldContactMotionSampler(make_cfg(motion_yaml, style))
        mlib = s._mlib
        ids = torch.randint(0, mlib.num_motions(), (n,), generator=g)
        ids[0], ids[1] = 0, 1
        lengths = mlib.get_motion_length(ids)
        starts = torch.rand(n, generator=g) * (lengths - s._sample_seq_time)      # motion_sampler.py:47 (CLAMP clips)
        starts[0] = 0.0                                                           # first frames
        starts[1] = lengths[1] - s._sample_seq_time                               # the last admissible start
        starts[2] = 7.0 / 30.0                                                    # an exact frame time
        torch.manual_seed(1234)
        motion_ret, hfs, tgt = s.sample_motion_data(motion_ids=ids, motion_start_times=starts)
        # the draws made inside the call, in order: rand_like(start_times) (:143), randn_like(future_pos) (:154)
        torch.manual_seed(1234)
        u = torch.rand_like(starts)
        z = torch.randn(n, 3)
        tag = "root" if style == "RELATIVE_TO_ROOT" else "floor"
        out.update({tag + "_ids": ids, tag + "_starts": starts, tag + "_u": u, tag + "_z": z, tag + "_hfs": hfs,
                    tag + "_root_pos": motion_ret[MDMFrameType.ROOT_POS], tag + "_root_rot": motion_ret[MDMFrameType.ROOT_ROT],
                    tag + "_joint_pos": motion_ret[MDMFrameType.JOINT_POS], tag + "_joint_rot": motion_ret[MDMFrameType.JOINT_ROT],
                    tag + "_contacts": motion_ret[MDMFrameType.CONTACTS], tag + "_future_pos": tgt.future_pos,
                    tag + "_future_rot": tgt.future_rot})
        # without heightmaps / target: a plain dict comes back (:293-302)
        only = s.sample_motion_data(motion_ids=ids, motion_start_times=starts, ret_hf_obs=False, ret_target_info=False)
        assert torch.equal(only[MDMFrameType.JOINT_ROT], motion_ret[MDMFrameType.JOINT_ROT])
        out[tag + "_nohf_root_pos"] = only[MDMFrameType.ROOT_POS]
    # ---- heightfield augmenter + FLOOR_HEIGHTS ----------------------------------------------------------------
    from oracle import parc_oracle_sampler as OS
    with open(os.path.join(ref_import.REF_ROOT, "data", "configs", "parc_1_train_gen_default.yaml")) as f:
        default_cfg = yaml.safe_load(f)
    default_cfg.update(device="cpu", char_file=ref_import.humanoid_xml(), motion_lib_file=motion_yaml)
    boxes_cfg = make_cfg(motion_yaml, "RELATIVE_TO_ROOT_FLOOR")
    boxes_cfg.update(use_hf_augmentation=True, hf_augmentation_mode="MAXPOOL_AND_BOXES", max_num_boxes=5, box_min_len=2, box_max_len=12,
                     hf_maxpool_chance=0.6, hf_max_maxpool_size=6, hf_change_height_chance=0.3)
    boxes_cfg["features"]["frame_components"] = ["ROOT_POS", "ROOT_ROT", "JOINT_POS", "JOINT_ROT", "CONTACTS", "FLOOR_HEIGHTS"]
    noise_cfg = make_cfg(motion_yaml, "RELATIVE_TO_ROOT")
    noise_cfg.update(use_hf_augmentation=True, hf_augmentation_mode="NOISE", max_num_boxes=4, box_min_len=2, box_max_len=12,
                     hf_maxpool_chance=0.15, hf_max_maxpool_size=10, hf_change_height_chance=0.1)     # read by the constructor (:91-97)
    for tag, cfg, seeds in (("aug_default", default_cfg, (11, 12)), ("aug_boxes", boxes_cfg, (21, 22)), ("aug_noise", noise_cfg, (31, 32))):
        s = rs.MDMHeightfieldContactMotionSampler(cfg)
        mlib = s._mlib
        ids = torch.randint(0, mlib.num_motions(), (n,), generator=g)
        lengths = mlib.get_motion_length(ids)
        starts = torch.rand(n, generator=g) * (lengths - s._sample_seq_time)
        starts[0], starts[1] = 0.0, lengths[1] - s._sample_seq_time
        random.seed(seeds[0]); torch.manual_seed(seeds[1])
        motion_ret, hfs, tgt = s.sample_motion_data(motion_ids=ids, motion_start_times=starts)
        # replay the draws: the augmenter's (get_hfs_from_data runs first), then the target info's
        random.seed(seeds[0]); torch.manual_seed(seeds[1])
        if cfg["hf_augmentation_mode"] == "NOISE":
            out[tag + "_noise"] = torch.stack([torch.rand(s._grid_dim_x, s._grid_dim_y) for _ in range(n)])
        else:
            rec_i, rec_f = OS.draw_box_aug_records(n, (s._grid_dim_x, s._grid_dim_y), s._max_h, cfg["max_num_boxes"], cfg["box_min_len"],
                                                   cfg["box_max_len"], cfg["hf_maxpool_chance"], cfg["hf_max_maxpool_size"],
                                                   cfg["hf_change_height_chance"])
            out[tag + "_rec_i"], out[tag + "_rec_f"] = rec_i, rec_f
        u = torch.rand_like(starts)
        z = torch.randn(n, 3)
        out.update({tag + "_ids": ids, tag + "_starts": starts, tag + "_u": u, tag + "_z": z, tag + "_hfs": hfs,
                    tag + "_root_pos": motion_ret[MDMFrameType.ROOT_POS], tag + "_future_pos": tgt.future_pos,
                    tag + "_seeds": np.array(seeds), tag + "_cfg": np.array(yaml.safe_dump({k: v for k, v in cfg.items() if k not in ("device", "char_file", "motion_lib_file")}))})
        if MDMFrameType.FLOOR_HEIGHTS in motion_ret:
            out[tag + "_floor_heights"] = motion_ret[MDMFrameType.FLOOR_HEIGHTS]
        print(tag, "hfs range", float(hfs.min()), float(hfs.max()), "changed cells vs plain:", None)
    extras = {}
    for k in range(mlib.num_motions()):
        extras["c%d_hf_maxmin" % k] = mlib._terrains[k].hf_maxmin.numpy()
        inds = mlib._hf_mask_inds[k]
        extras["c%d_mask_start" % k] = np.cumsum([0] + [int(x.shape[0]) for x in inds]).astype(np.int64)
        extras["c%d_mask_cells" % k] = torch.cat([torch.as_tensor(x).reshape(-1, 2) for x in inds]).numpy().astype(np.int32)
    np.savez_compressed(os.path.join(HERE, "clip_aug_extras.npz"), **extras)
    out["motion_times"] = s._motion_times
    out["grid"] = s._generic_heightmap
    out["clip_names"] = np.array(CLIPS)
    # the second clip as plain arrays (the first is already tests/golden/clip_teaser_dm.npz)
    import parc.util.file_io_helper as rfio
    import parc.anim.motion_lib as rmlib
    d = rfio.load_ms_file(ref_import.clip(CLIPS[1]), device="cpu")
    md, td = d.motion_data, d.terrain_data
    np.savez_compressed(os.path.join(HERE, "clip_teaser_modified.npz"), root_pos=md.root_pos.numpy(), root_rot=md.root_rot.numpy(),
                        joint_rot=md.joint_rot.numpy(), contacts=md.body_contacts.numpy(), fps=np.array(md.fps),
                        loop_mode=np.array(rmlib.LoopMode[md.loop_mode].value), hf=td.hf.numpy(), min_point=np.asarray(td.min_point),
                        dx=np.array(float(td.dx), dtype=np.float32))
    np.savez_compressed(os.path.join(HERE, "ref_mdm_sampler.npz"),
                        **{k: (v.detach().numpy() if isinstance(v, torch.Tensor) else v) for k, v in out.items()})
    print({k: getattr(v, "shape", None) for k, v in out.items()})


if __name__ == "__main__":
    main()
