"""GPU: parc_b200.dm_env.DeepMimicEnv -- the reference's DeepMimicEnv / RefCharEnv surface -- replayed
method for method in the call order of tests/golden/make_golden.py (which drives the UNMODIFIED reference class the same way)
against the same golden file. The simulator state is injected the way Isaac Gym hands it out: root state as column slices
of an [N * actors, 13] tensor, dof state as stride-2 slices of [N * dofs, 2] (ig_char_env.py:172-196)."""
import numpy as np
import pytest
import torch
import yaml

from tests import helpers as H
from tests.helpers import assert_close, load_npz, t

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


@pytest.fixture(scope="module")
def model():
    return H.product_model(DEV)


def _env(model, g, actors=2):
    from parc_b200 import dm_util
    from parc_b200.dm_env import DeepMimicEnv
    clip = H.golden_clip()
    lib = H.product_library([clip], model, DEV)
    lib._terrains = [H.product_terrain(clip["hf"], clip["min_point"], clip["dx"], DEV)]
    ec = yaml.safe_load(str(g["cfg_yaml"]))["env"]
    ec["dm"] = {"motion_file": lib, "terrains_per_motion": 1, "terrain_build_mode": "file", "min_motion_weight": 0.01,
                "fail_rate_quantiles": [0.25, 0.5, 0.75], "motion_classes": []}
    cfg = {"env": ec}
    N = g["in_timestep_buf"].shape[0]
    J, D = model.get_num_joints(), model.get_dof_size()
    env = DeepMimicEnv(cfg, N, DEV, False, model)          # the reference's constructor signature (dm_env.py:20)
    env.build_terrain(ec, None)
    z = lambda *s, dt=torch.float32: torch.zeros(*s, dtype=dt, device=DEV)
    # Isaac Gym layout: two actors per env, the character is actor 0; dofs interleaved (pos, vel)
    root_all = z(N * actors, 13)
    root = root_all.view(N, actors, 13)[:, 0]
    dof = z(N * D, 2).view(N, D, 2)
    sim = dict(char_root_pos=root[:, 0:3], char_root_rot=root[:, 3:7], char_root_vel=root[:, 7:10], char_root_ang_vel=root[:, 10:13],
               char_dof_pos=dof[..., 0], char_dof_vel=dof[..., 1], char_contact_forces=z(N, J, 3), char_rigid_body_pos=z(N, J, 3),
               char_rigid_body_vel=z(N, J, 3), char_rigid_body_ang_vel=z(N, J, 3))
    ref = dict(ref_root_pos=z(N, 3), ref_root_rot=z(N, 4), ref_root_vel=z(N, 3), ref_root_ang_vel=z(N, 3), ref_body_pos=z(N, J, 3),
               ref_joint_rot=z(N, J - 1, 4), ref_dof_pos=z(N, D), ref_dof_vel=z(N, D), ref_contacts=z(N, J))
    assert not sim["char_root_pos"].is_contiguous() and sim["char_dof_pos"].stride(1) == 2
    env.get_sim_tensor_views(**ref, **sim)
    bufs = dict(reward_buf=z(N), done_buf=z(N, dt=torch.int32), time_buf=z(N), timestep_buf=z(N, dt=torch.int32),
                actors_need_reset=torch.zeros(N, actors, dtype=torch.bool, device=DEV), env_offsets=t(g["env_offsets"]).to(DEV),
                key_body_ids=torch.tensor([model.get_body_id(b) for b in ec["key_bodies"]], dtype=torch.int64, device=DEV),
                ray_xy_points=t(g["ray_xy"]).to(DEV), ray_hfs=z(N, g["ray_xy"].shape[0]))
    env.get_data_buffer_views(**bufs)
    return env, ec, sim, ref, bufs, dm_util


def test_reset_replays_the_reference(model):
    g = load_npz("ref_env_step.npz")
    env, ec, sim, ref, bufs, _ = _env(model, g)
    N = env._num_envs
    sim["char_rigid_body_vel"].fill_(1.0)
    env.reset(torch.arange(N, device=DEV), draws=(t(g["reset_motion_ids"]), t(g["reset_terrain_ids"]), t(g["reset_times"]), t(g["reset_noise"])))
    torch.cuda.synchronize()
    assert np.array_equal(env._motion_ids.cpu().numpy(), g["reset_motion_ids"])
    assert np.array_equal(env._motion_time_offsets.cpu().numpy(), g["reset_times"])
    W = float(np.abs(g["reset_ref_root_pos"]).max())
    for k, v in list(ref.items()) + list(sim.items()):
        if "reset_" + k in g:
            H.assert_close_world(v, g["reset_" + k], W, what=k)
    assert np.array_equal(bufs["actors_need_reset"].cpu().numpy(), g["reset_actors_need_reset"])
    assert int(bufs["timestep_buf"].abs().sum()) == 0 and int(bufs["done_buf"].abs().sum()) == 0


def test_post_physics_sequence_replays_the_reference(model):
    g = load_npz("ref_env_step.npz")
    env, ec, sim, ref, bufs, dm_util = _env(model, g)
    N, J, D = env._num_envs, model.get_num_joints(), model.get_dof_size()
    cu = lambda k: t(g[k]).to(DEV)
    # state as make_golden.py left it before the step
    bufs["timestep_buf"].copy_(cu("in_timestep_buf"))
    env._motion_ids.copy_(cu("in_motion_ids"))
    env._motion_terrain_ids.copy_(cu("in_motion_terrain_ids"))
    env._motion_time_offsets.copy_(cu("in_motion_time_offsets"))
    env._motion_id_fail_rates.copy_(cu("in_fail_rates"))
    for k in sim:
        if "in_" + k in g:
            sim[k].copy_(cu("in_" + k))
    timestep = 1.0 / ec["control_freq"]
    # ---- IGEnv._post_physics_step order (ig_env.py:368-377), as make_golden.py replays it ----
    from parc_b200 import _lib
    heading = torch.empty(N, device=DEV)
    _lib.check(_lib.lib().parc_b200_calc_heading(_lib.ptr(sim["char_root_rot"].contiguous()), N, _lib.ptr(heading), _lib.stream_ptr(torch.device(DEV))))
    env._refresh_obs_hfs(sim["char_root_pos"] + bufs["env_offsets"], heading)
    assert_close(bufs["ray_hfs"], g["ray_hfs"], what="_refresh_obs_hfs")
    bufs["timestep_buf"] += 1
    bufs["time_buf"][:] = timestep * bufs["timestep_buf"]
    assert np.array_equal(bufs["time_buf"].cpu().numpy(), g["time_buf"])
    env._update_ref_motion()
    W = float(np.abs(g["out_ref_root_pos"]).max())
    for k, v in ref.items():
        H.assert_close_world(v, g["out_" + k], W, what="_update_ref_motion " + k)
    tar = env.compute_tar_obs(torch.tensor(ec["tar_obs_steps"], dtype=torch.int), None)
    for name, a in zip(("tar_root_pos", "tar_root_rot", "tar_joint_rot", "tar_key_pos", "tar_contacts"), tar):
        H.assert_close_world(a, g[name], W, what="compute_tar_obs " + name)
    # subset variant: rows land in call order
    ids = torch.tensor([5, 0, 17, 3], device=DEV)
    sub = env.compute_tar_obs(torch.tensor(ec["tar_obs_steps"], dtype=torch.int), ids)
    for a, b in zip(sub, tar):
        assert torch.equal(a, b[ids])
    # observation and reward through the reference's free functions (their mirrors in dm_util), as IGParkourEnv calls them
    joint_rot = model.dof_to_rot(sim["char_dof_pos"])
    body_pos, _ = model.forward_kinematics(sim["char_root_pos"], sim["char_root_rot"], joint_rot)
    kb = bufs["key_body_ids"]
    od = dm_util.compute_deepmimic_obs(root_pos=sim["char_root_pos"], root_rot=sim["char_root_rot"], root_vel=sim["char_root_vel"],
                                       root_ang_vel=sim["char_root_ang_vel"], joint_rot=joint_rot, dof_vel=sim["char_dof_vel"],
                                       key_pos=body_pos[..., kb, :], global_obs=ec["global_obs"], root_height_obs=ec["global_root_height_obs"],
                                       enable_tar_obs=True, tar_root_pos=tar[0], tar_root_rot=tar[1], tar_joint_rot=tar[2], tar_key_pos=tar[3])
    char_contacts = (torch.norm(sim["char_contact_forces"], dim=-1) > 1e-5).to(torch.float32)
    obs = torch.cat([od["char_obs"], od["tar_obs"].reshape(N, -1), tar[4].reshape(N, -1), char_contacts, bufs["ray_hfs"]], dim=-1)
    H.assert_close_world(obs, g["obs"], W, what="obs row")
    # update_done with the reference's ten arguments
    env.update_done(termination_height=torch.tensor(ec["termination_height"], dtype=torch.float32), episode_length=ec["episode_length"],
                    contact_body_ids=torch.tensor([model.get_body_id(b) for b in ec["contact_bodies"]], dtype=torch.int64),
                    pose_termination=ec["pose_termination"], pose_termination_dist=torch.tensor(ec["pose_termination_dist"], dtype=torch.float32),
                    global_obs=ec["global_obs"], enable_early_termination=ec["enable_early_termination"], track_root=ec["track_root"],
                    root_pos_termination_dist=ec["root_pos_termination_dist"], root_rot_termination_angle=ec["root_rot_termination_angle"])
    torch.cuda.synchronize()
    assert np.array_equal(bufs["done_buf"].cpu().numpy(), g["done"]), "done flags must be bit-exact"
    assert_close(env._motion_id_fail_rates, g["out_fail_rates"], rtol=1e-6, atol=0, what="fail-rate EMA")
    # both attribute names the reference's callers use
    assert env._mlib is env._motion_lib


def test_setters_and_terrain_square(model, tmp_path):
    from parc_b200 import synth
    from parc_b200.dm_env import DeepMimicEnv
    g = load_npz("ref_terrain_square.npz")
    clips = synth.make_clips(synth.humanoid_spec(), int(g["num_clips"]), seed=3, min_frames=20, max_frames=30)
    for k, c in enumerate(clips):
        c["hf"], c["min_point"] = t(g["clip%d_hf" % k]), t(g["clip%d_min_point" % k])
    lib = H.product_library(clips, model, DEV)
    lib._terrains = [H.product_terrain(c["hf"], c["min_point"], 0.4, DEV) for c in clips]
    ec = dict(control_freq=30, rand_root_pos_offset_scale=0.0, max_obs_h=3.0, min_obs_h=-3.0, demo_mode=False, rand_reset=True,
              dm={"motion_file": lib, "terrains_per_motion": 1, "terrain_build_mode": "square", "fail_rate_quantiles": [0.5],
                  "heightmap": {"horizontal_scale": 0.4, "padding": float(g["b_padding"])}})
    env = DeepMimicEnv({"env": ec}, 8, DEV, False, model)
    path = str(tmp_path / "terrain.pkl")
    env.build_terrain(ec, path, x_offset=float(g["b_x_offset"]), y_offset=float(g["b_y_offset"]))
    assert np.array_equal(env._terrain.hf.cpu().numpy(), g["b_hf"])
    assert np.array_equal(env._terrain.min_point.cpu().numpy(), g["b_min_point"])
    assert np.array_equal(env._dm_motion_offsets.cpu().numpy(), g["b_dm_motion_offsets"])
    assert tuple(lib._terrains[2].hf.shape) == (g["clip2_hf"].shape[0] + 2, g["clip2_hf"].shape[1] + 2)    # padded in place, as upstream
    env2 = DeepMimicEnv({"env": dict(ec, dm=dict(ec["dm"], motion_file=lib))}, 8, DEV, False, model)
    env2.load_terrain(path)
    assert torch.equal(env2._terrain.hf, env._terrain.hf) and torch.equal(env2._dm_motion_offsets, env._dm_motion_offsets)
    assert env.set_demo_mode() is True and env.set_demo_mode(False) is False
    with pytest.raises(AssertionError):
        env.set_root_pos_offset(torch.zeros(7, 3))
    env.set_root_rot_offset(torch.zeros(8, 4))
    env.set_root_rot_offset(None)
