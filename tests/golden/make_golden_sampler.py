"""Golden vectors for the MDM dataset sampler: runs the UNMODIFIED reference
MDMHeightfieldContactMotionSampler.sample_motion_data (PARC/motion_generator/
mdm_heightfield_contact_motion_sampler.py:197-302) on CPU over a two-motion library built from the bundled
clips that carry hf_mask_inds, for both relative_z_style settings, with the augmenter off.

Run here (authoring container) only:  python tests/golden/make_golden_sampler.py  -> ref_mdm_sampler.npz
The random draws the reference makes inside the call (future times, future-position noise) are recorded by
re-seeding torch's generator and repeating the draws in the reference's order.
"""
import os
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
