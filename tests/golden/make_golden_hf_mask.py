"""Golden vectors for SURVEY.md section 8(f) rank 4 (terrain post-processing): terrain_util.compute_hf_mask_inds
(PARC/util/terrain_util.py:1851-1897), compute_hf_mask_from_inds (:1899-1906) and compute_hf_extra_vals (:1918-1947)
run with the UNMODIFIED reference on CPU over the bundled teaser clip (60 frames) and its terrain:

    python tests/golden/make_golden_hf_mask.py   -> ref_hf_mask.npz

The ragged per-frame index lists are stored flat with offsets."""
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
    d = rfio.load_ms_file(ref_import.clip("dec2024_teaser_717_1_opt_dm.pkl"), device="cpu")
    md, td = d.motion_data, d.terrain_data
    terrain = rterr.SubTerrain(x_dim=td.hf.shape[0], y_dim=td.hf.shape[1], dx=float(td.dx), dy=float(td.dx),
                               min_x=float(td.min_point[0]), min_y=float(td.min_point[1]), device="cpu")
    terrain.hf[...] = td.hf
    sl = slice(40, 100)
    frames = MotionFrames(root_pos=md.root_pos[sl].clone(), root_rot=md.root_rot[sl].clone(), joint_rot=md.joint_rot[sl].clone(),
                          contacts=md.body_contacts[sl].clone())
    frames.root_pos[20:30, 2] += 1.2          # a "jump": bodies far above the cells they cross (the jump_buf branch)
    inds, min_heights = rterr.compute_hf_mask_inds(frames, terrain, model, body_points)
    mask = rterr.compute_hf_mask_from_inds(terrain, inds)
    inds2, ret_terrain = rterr.compute_hf_extra_vals(frames, terrain, model, body_points)
    assert all(torch.equal(a, b) for a, b in zip(inds, inds2))
    offs = np.cumsum([0] + [int(x.shape[0]) for x in inds])
    np.savez_compressed(os.path.join(HERE, "ref_hf_mask.npz"), root_pos=frames.root_pos.numpy(), root_rot=frames.root_rot.numpy(),
                        joint_rot=frames.joint_rot.numpy(), hf=terrain.hf.numpy(), min_point=terrain.min_point.numpy(),
                        dxdy=terrain.dxdy.numpy(), inds=torch.cat(inds).numpy(), offsets=offs, min_body_heights=min_heights.numpy(),
                        hf_mask=mask.numpy(), hf_maxmin=ret_terrain.hf_maxmin.numpy())
    print("frames", len(inds), "cells per frame", [int(x.shape[0]) for x in inds][:8], "mask cells", int(mask.sum()),
          "jump cells", int((ret_terrain.hf_maxmin[..., 0] != ret_terrain.hf_maxmin[..., 1]).logical_and(mask).sum()))


if __name__ == "__main__":
    main()
