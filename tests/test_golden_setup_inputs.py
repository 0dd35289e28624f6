"""CPU: the host-side setup restatements that BOTH the oracle and the CUDA env consume as inputs, pinned bit for bit to
outputs of the unmodified reference (tests/golden/make_golden_terrain_square.py, make_golden_opt.py): an error in one of
them would be invisible to every oracle-vs-kernel comparison.
  synth.build_terrain_square   DeepMimicEnv.build_terrain_square, PARC/motion_tracker/envs/ig_parkour/dm_env.py:157-316
  synth.grid_ray_points        geom_util.get_xy_grid_points, PARC/util/geom_util.py:212-224
  geom_util.get_char_point_samples   PARC/util/geom_util.py:790-872"""
import numpy as np
import torch

from parc_b200 import synth
from tests import helpers as H
from tests.helpers import load_npz


def _golden_clips(g):
    return [dict(hf=torch.from_numpy(g["clip%d_hf" % k]), min_point=torch.from_numpy(g["clip%d_min_point" % k]), dx=float(g["dx"]))
            for k in range(int(g["num_clips"]))]


def test_build_terrain_square_reproduces_the_reference_bit_for_bit():
    g = load_npz("ref_terrain_square.npz")
    for tag in ("a", "b", "c"):
        clips = _golden_clips(g)
        pad_cells = synth.num_padding_cells(float(g[tag + "_padding"]), float(g["dx"]))
        hf, min_point, dx, offs = synth.build_terrain_square(clips, padding_cells=pad_cells, x_offset=float(g[tag + "_x_offset"]),
                                                             y_offset=float(g[tag + "_y_offset"]))
        assert tuple(hf.shape) == tuple(int(x) for x in g[tag + "_dims"]), tag
        assert np.array_equal(hf.numpy(), g[tag + "_hf"]), tag
        assert np.array_equal(min_point.numpy(), g[tag + "_min_point"]), tag
        assert np.array_equal(offs.numpy(), g[tag + "_dm_motion_offsets"]), tag
        assert np.float32(dx) == g[tag + "_dxdy"][0] == g[tag + "_dxdy"][1]
    # the reference truncates padding / dx (dm_env.py:162-166): 1.2 / 0.4 = 2.9999999999999996 -> 2 cells, as case "c" shows
    assert synth.num_padding_cells(1.2, 0.4) == 2 and synth.num_padding_cells(0.8, 0.4) == 2 and synth.num_padding_cells(0.4, 0.4) == 1
    assert g["c_hf"].shape == g["a_hf"].shape


def test_grid_ray_points_equal_get_xy_grid_points():
    g = load_npz("ref_terrain_square.npz")
    assert np.array_equal(synth.grid_ray_points().numpy(), g["grid_32"])
    assert np.array_equal(synth.xy_grid_points(dx=0.15, dy=0.1, num_x_neg=3, num_x_pos=9, num_y_neg=7, num_y_pos=2).numpy(), g["grid_asym"])


def test_char_point_samples_equal_the_reference():
    from parc_b200 import geom_util
    bp = load_npz("ref_body_points.npz")
    pts = geom_util.get_char_point_samples(H.product_model("cpu"))
    assert [int(p.shape[0]) for p in pts] == [int(c) for c in bp["counts"]]
    assert np.array_equal(torch.cat(pts, dim=0).numpy(), bp["points"])
