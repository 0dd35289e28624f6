"""Golden vectors for SURVEY.md section 8(f) rank 2: mdm_path.compute_motion_loss
(PARC/motion_synthesis/procgen/mdm_path.py:31-127) -- the forward-only penetration/contact ranking of generated motion
batches against a WHOLE terrain -- run with the UNMODIFIED reference on CPU:

    python tests/golden/make_golden_path_loss.py   -> ref_path_loss.npz

Two cases: windows of the bundled teaser clip on its 18x15 terrain, and windows of civilization.pkl on its 50x50
terrain (more cells than the CUDA kernel stages in shared memory). Half of each batch is pushed into / lifted off the
ground so that both terms are non-zero."""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from oracle import ref_import  # noqa: E402

ref_import.setup()
torch.set_num_threads(4)

import parc.anim.kin_char_model as rkin  # noqa: E402
import parc.motion_synthesis.procgen.mdm_path as rpath  # noqa: E402
import parc.util.file_io_helper as rfio  # noqa: E402
import parc.util.geom_util as rgeom  # noqa: E402
import parc.util.terrain_util as rterr  # noqa: E402
from parc.util.motion_util import MotionFrames  # noqa: E402

POINT_SAMPLES = dict(sphere_num_subdivisions=0, box_num_slices=2, box_dim_x=3, box_dim_y=6, capsule_num_circle_points=4,
                     capsule_num_sphere_subdivisons=0, capsule_num_cylinder_slices=4)


def main():
    model = rkin.KinCharModel("cpu")
    model.load_char_file(ref_import.humanoid_xml())
    body_points = rgeom.get_char_point_samples(model, **POINT_SAMPLES)
    out = {}
    g = torch.Generator().manual_seed(21)
    for tag, pkl, B, T in (("teaser", "dec2024_teaser_717_1_opt_dm.pkl", 6, 20), ("civ", "civilization.pkl", 4, 12)):
        d = rfio.load_ms_file(ref_import.clip(pkl), device="cpu")
        md, td = d.motion_data, d.terrain_data
        terrain = rterr.SubTerrain(x_dim=td.hf.shape[0], y_dim=td.hf.shape[1], dx=float(td.dx), dy=float(td.dx),
                                   min_x=float(td.min_point[0]), min_y=float(td.min_point[1]), device="cpu")
        terrain.hf[...] = td.hf
        nf = md.root_pos.shape[0]
        starts = torch.randint(0, nf - T, (B,), generator=g)
        idx = starts[:, None] + torch.arange(T)[None, :]
        root_pos = md.root_pos[idx].clone()
        root_pos[: B // 2, :, 2] -= 0.08                       # into the ground
        root_pos[B // 2:, :, 2] += 0.05                        # off the ground
        root_pos[..., 0:2] += 0.02 * torch.randn(B, T, 2, generator=g)
        frames = MotionFrames(root_pos=root_pos, root_rot=md.root_rot[idx].clone(), joint_rot=md.joint_rot[idx].clone(),
                              contacts=md.body_contacts[idx].clone())
        losses = rpath.compute_motion_loss(frames, None, terrain, model, body_points, w_contact=0.1, w_pen=0.1, w_path=0.0, verbose=False)
        out.update({tag + "_root_pos": frames.root_pos, tag + "_root_rot": frames.root_rot, tag + "_joint_rot": frames.joint_rot,
                    tag + "_contacts": frames.contacts, tag + "_hf": terrain.hf, tag + "_min_point": terrain.min_point,
                    tag + "_dxdy": terrain.dxdy, tag + "_total_loss": losses["total_loss"], tag + "_contact_loss": losses["contact_loss"],
                    tag + "_pen_loss": losses["pen_loss"]})
        print(tag, {k: v.tolist() for k, v in losses.items()})
    np.savez_compressed(os.path.join(HERE, "ref_path_loss.npz"), **{k: v.detach().numpy() for k, v in out.items()})


if __name__ == "__main__":
    main()
