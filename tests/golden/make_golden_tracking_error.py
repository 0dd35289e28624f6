"""Golden vector for compute_tracking_error: the UNMODIFIED reference function
(mgdm_dm_util.py:520-553) driven as IGParkourEnv._update_info drives it (ig_parkour_env.py:1066-1095)
on the recorded inputs/outputs of ref_env_step.npz -- simulated character state and the reference
motion state of the same step, two FK passes of the reference KinCharModel, then the jit function.

Run here only:  python tests/golden/make_golden_tracking_error.py   -> ref_tracking_error.npz
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from oracle import ref_import  # noqa: E402

ref_import.setup()
torch.set_num_threads(1)

import parc.anim.kin_char_model as rkin  # noqa: E402
import parc.motion_tracker.envs.ig_parkour.mgdm_dm_util as rutil  # noqa: E402


def main():
    g = np.load(os.path.join(HERE, "ref_env_step.npz"))
    t = lambda k: torch.from_numpy(g[k])
    model = rkin.KinCharModel("cpu")
    model.load_char_file(ref_import.humanoid_xml())
    char_joint_rot = model.dof_to_rot(t("in_char_dof_pos"))
    char_body_pos, char_body_rot = model.forward_kinematics(t("in_char_root_pos"), t("in_char_root_rot"), char_joint_rot)
    ref_body_pos, ref_body_rot = model.forward_kinematics(t("out_ref_root_pos"), t("out_ref_root_rot"), t("out_ref_joint_rot"))
    err = rutil.compute_tracking_error(
        root_pos=t("in_char_root_pos"), root_rot=t("in_char_root_rot"), body_rot=char_body_rot, body_pos=char_body_pos,
        tar_root_pos=t("out_ref_root_pos"), tar_root_rot=t("out_ref_root_rot"), tar_body_rot=ref_body_rot, tar_body_pos=ref_body_pos,
        root_vel=t("in_char_root_vel"), root_ang_vel=t("in_char_root_ang_vel"), dof_vel=t("in_char_dof_vel"),
        tar_root_vel=t("out_ref_root_vel"), tar_root_ang_vel=t("out_ref_root_ang_vel"), tar_dof_vel=t("out_ref_dof_vel"))
    print("tracking error mean per channel:", err.mean(dim=0).tolist())
    np.savez_compressed(os.path.join(HERE, "ref_tracking_error.npz"), tracking_error=err.numpy(), char_body_rot=char_body_rot.numpy(),
                        char_body_pos=char_body_pos.numpy(), ref_body_rot=ref_body_rot.numpy(), ref_body_pos=ref_body_pos.numpy())


if __name__ == "__main__":
    main()
