"""Golden vectors for DeepMimicEnv.build_terrain_square (PARC/motion_tracker/envs/ig_parkour/dm_env.py:157-316) and
geom_util.get_xy_grid_points (PARC/util/geom_util.py:212-224), from the UNMODIFIED reference on CPU.

Run here (the authoring container) only:  python tests/golden/make_golden_terrain_square.py
Seven synthetic clips with terrains of unequal dims are written as the reference's own .pkl files + a motions yaml, loaded
with MotionLib.from_file, and tiled by the reference; inputs and outputs go to ref_terrain_square.npz.
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

import parc.anim.kin_char_model as rkin  # noqa: E402
import parc.motion_tracker.envs.ig_parkour.dm_env as rdm  # noqa: E402
import parc.util.file_io as rfile  # noqa: E402
import parc.util.geom_util as rgeom  # noqa: E402

import parc.util.terrain_util as rterrain  # noqa: E402

from parc_b200 import synth  # noqa: E402  (input generation only)

# The reference targets torch 2.2 / NumPy 1.x, where `python_float - np.float32_scalar` (dm_env.py:222-223) is evaluated in
# float64 and the result is rounded when it is stored into the fp32 offsets tensor. NumPy 2 (this container) would keep the
# subtraction in float32 -- and torch then refuses the np.float32 assignment outright. The harness restores the NumPy 1.x
# behaviour WITHOUT touching the reference: numpy_copy() hands out min_point as float64 (same values), so the subtraction is
# a float64 one again. Nothing else of build_terrain_square depends on that dtype.
_numpy_copy = rterrain.SubTerrain.numpy_copy


def _numpy_copy_np1(self):
    t = _numpy_copy(self)
    t.min_point = np.asarray(t.min_point, dtype=np.float64)
    return t


rterrain.SubTerrain.numpy_copy = _numpy_copy_np1

DIMS = [(20, 32), (32, 24), (18, 15), (32, 32), (27, 33), (25, 25), (9, 40)]


def make_inputs():
    clips = []
    for k, (dx_, dy_) in enumerate(DIMS):
        c = synth.make_clips(synth.humanoid_spec(), 1, seed=300 + k, min_frames=20, max_frames=40, terrain_dim=max(dx_, dy_))[0]
        c["hf"] = c["hf"][:dx_, :dy_].contiguous()
        c["min_point"] = c["min_point"] + 0.4 * torch.tensor([k % 3, k // 3], dtype=torch.float32)   # unequal origins too
        clips.append(c)
    return clips


def main():
    clips = make_inputs()
    tmp = tempfile.mkdtemp(prefix="parc_square_")
    entries = []
    for k, c in enumerate(clips):
        md = rfile.MSMotionData(root_pos=c["root_pos"].numpy(), root_rot=c["root_rot"].numpy(), joint_rot=c["joint_rot"].numpy(),
                                body_contacts=c["contacts"].numpy(), fps=int(c["fps"]), loop_mode="CLAMP")
        hf = c["hf"].numpy()
        td = rfile.MSTerrainData(hf=hf, hf_maxmin=np.stack([hf, hf], axis=-1), min_point=c["min_point"].numpy(), dx=float(c["dx"]))
        path = os.path.join(tmp, "clip_%d.pkl" % k)
        rfile.save_ms_file(rfile.MSFileData(motion_data=md, terrain_data=td, misc_data=None), path)
        entries.append({"file": path, "weight": 1.0})
    motions_yaml = os.path.join(tmp, "motions.yaml")
    with open(motions_yaml, "w") as f:
        yaml.safe_dump({"motions": entries}, f)

    with open(ref_import.clip("dm_env_civilization.yaml"), "r") as f:
        cfg = yaml.safe_load(f)
    cfg["env"]["dm"]["motion_file"] = motions_yaml
    cfg["env"]["dm"]["terrain_build_mode"] = "square"
    cfg["env"]["dm"]["terrains_per_motion"] = 1
    model = rkin.KinCharModel("cpu")
    model.load_char_file(ref_import.humanoid_xml())
    out = {}
    for tag, padding, xo, yo in (("a", 0.8, 0.0, 0.0), ("b", 0.4, 1.2, -2.0), ("c", 1.2, -0.4, 0.8)):     # padding 0 crashes the reference itself (SubTerrain.pad slices [0:-0])
        cfg["env"]["dm"]["heightmap"] = {"horizontal_scale": 0.4, "padding": padding}
        env = rdm.DeepMimicEnv(cfg, 4, "cpu", False, model)       # fresh library: build_terrain_square pads the clips' terrains in place
        env.build_terrain(cfg["env"], os.path.join(tmp, "terrain_%s.pkl" % tag), x_offset=xo, y_offset=yo)
        t = env._terrain
        out.update({tag + "_padding": np.float64(padding), tag + "_x_offset": np.float64(xo), tag + "_y_offset": np.float64(yo),
                    tag + "_hf": t.hf.numpy(), tag + "_min_point": t.min_point.numpy(), tag + "_dxdy": t.dxdy.numpy(),
                    tag + "_dims": t.dims.numpy(), tag + "_dm_motion_offsets": env._dm_motion_offsets.numpy()})
    for k, c in enumerate(clips):
        out["clip%d_hf" % k] = c["hf"].numpy()
        out["clip%d_min_point" % k] = c["min_point"].numpy()
    out["num_clips"] = np.int64(len(clips))
    out["dx"] = np.float64(0.4)          # horizontal_scale: a Python double in the reference (dm_env.py:159)
    # ray lattices: the 32x32 config-2 lattice and an asymmetric one
    out["grid_32"] = rgeom.get_xy_grid_points(center=torch.zeros(2), dx=0.1, dy=0.1, num_x_neg=15, num_x_pos=16, num_y_neg=15,
                                              num_y_pos=16).reshape(-1, 2).numpy()
    out["grid_asym"] = rgeom.get_xy_grid_points(center=torch.zeros(2), dx=0.15, dy=0.1, num_x_neg=3, num_x_pos=9, num_y_neg=7,
                                                num_y_pos=2).reshape(-1, 2).numpy()
    np.savez_compressed(os.path.join(HERE, "ref_terrain_square.npz"), **out)
    print("wrote ref_terrain_square.npz", {k: np.asarray(v).shape for k, v in out.items() if not k.startswith("clip")})


if __name__ == "__main__":
    main()
