"""Pins the CPU oracle (oracle/parc_oracle.py) against golden vectors produced by the UNMODIFIED
reference (tests/golden/make_golden.py) and against the one golden vector the reference ships
(the recorded observation tape of dec2024_teaser_717_1_opt_dm.pkl). CPU only."""
import numpy as np
import torch
import yaml

from oracle import parc_oracle as O
from tests.helpers import assert_close, golden_clip, load_npz, oracle_clip, oracle_model, t

torch.set_num_threads(1)


def test_kin_char_model_matches_reference():
    g = load_npz("ref_kin.npz")
    m = oracle_model()
    rot = m.dof_to_rot(t(g["dof"]))
    assert_close(rot, g["rot"], atol=1e-6, what="dof_to_rot")
    pos, brot = m.forward_kinematics(t(g["root_pos"]), t(g["root_rot"]), t(g["rot"]))
    assert_close(pos, g["body_pos"], atol=1e-6, what="fk pos")
    assert_close(brot, g["body_rot"], atol=1e-6, what="fk rot")
    assert_close(m.rot_to_dof(t(g["rot"])), g["dof_back"], atol=1e-6, what="rot_to_dof")
    assert_close(m.rot_to_dof(t(g["rnd_rot"])), g["dof_rnd"], atol=1e-6, what="rot_to_dof random")


def test_motion_lib_matches_reference():
    g = load_npz("ref_motionlib.npz")
    m = oracle_model()
    clip = oracle_clip(golden_clip())
    lib = O.Library(m, [clip])
    assert_close(lib.frame_root_vel, g["frame_root_vel"], what="root_vel")
    assert_close(lib.frame_root_ang_vel, g["frame_root_ang_vel"], what="root_ang_vel")
    assert_close(lib.frame_dof_vel, g["frame_dof_vel"], atol=1e-5, what="dof_vel")
    assert float(lib.lengths[0]) == float(g["motion_length"][0])
    assert_close(lib.root_pos_delta, g["root_pos_delta"], what="delta")
    names = ["root_pos", "root_rot", "root_vel", "root_ang_vel", "joint_rot", "dof_vel", "contacts"]
    for prefix, loop in (("", O.CLAMP), ("w_", O.WRAP)):
        lib.loop_modes[:] = loop
        times = t(g["times_w" if prefix else "times"])
        ids = torch.zeros(times.shape[0], dtype=torch.int64)
        i0, i1, bl = lib.frame_blend(ids, times)
        assert np.array_equal(i0.numpy(), g[prefix + "idx0"]), "frame_idx0 must be bit-exact"
        assert np.array_equal(i1.numpy(), g[prefix + "idx1"]), "frame_idx1 must be bit-exact"
        assert np.array_equal(bl.numpy(), g[prefix + "blend"]), "blend is the same fp32 expression"
        out = lib.calc_motion_frame(ids, times)
        for k, v in zip(names, out):
            assert_close(v, g[prefix + k], atol=1e-6, what=prefix + k)


def _env_from_golden(g):
    m = oracle_model()
    clip = oracle_clip(golden_clip())
    lib = O.Library(m, [clip])
    ec = yaml.safe_load(str(g["cfg_yaml"]))["env"]
    cfg = O.EnvConfig(m, control_freq=ec["control_freq"], episode_length=ec["episode_length"],
                      tar_obs_steps=ec["tar_obs_steps"], key_bodies=ec["key_bodies"], global_obs=ec["global_obs"],
                      root_height_obs=ec["global_root_height_obs"], use_contact_info=ec["use_contact_info"],
                      track_root=ec["track_root"], track_root_h=ec["track_root_h"],
                      enable_early_termination=ec["enable_early_termination"], pose_termination=ec["pose_termination"],
                      contact_bodies=ec["contact_bodies"], termination_height=ec["termination_height"],
                      root_pos_termination_dist=ec["root_pos_termination_dist"],
                      root_rot_termination_angle=ec["root_rot_termination_angle"],
                      pose_termination_dist=ec["pose_termination_dist"], joint_err_w=ec["joint_err_w"],
                      contact_weights=ec["contact_weights"], pose_w=ec["pose_w"], vel_w=ec["vel_w"],
                      root_pos_w=ec["root_pos_w"], root_vel_w=ec["root_vel_w"], key_pos_w=ec["key_pos_w"],
                      min_obs_h=ec["min_obs_h"], max_obs_h=ec["max_obs_h"],
                      rand_root_pos_offset_scale=ec["rand_root_pos_offset_scale"])
    N = g["in_timestep_buf"].shape[0]
    env = O.TrackingEnv(m, lib, clip["terrain"], t(g["dm_motion_offsets"]), cfg, N, t(g["env_offsets"]), t(g["ray_xy"]))
    return env


def load_golden_step_inputs(env, g):
    env.timestep_buf[:] = t(g["in_timestep_buf"])
    env.time_buf[:] = env.timestep * env.timestep_buf
    env.motion_ids[:] = t(g["in_motion_ids"])
    env.motion_terrain_ids[:] = t(g["in_motion_terrain_ids"])
    env.motion_time_offsets[:] = t(g["in_motion_time_offsets"])
    env.fail_rates[:] = t(g["in_fail_rates"])
    for k in ("root_pos", "root_rot", "root_vel", "root_ang_vel", "dof_pos", "dof_vel", "contact_forces", "rigid_body_pos"):
        getattr(env, "char_" + k)[:] = t(g["in_char_" + k])


def test_env_step_matches_reference():
    g = load_npz("ref_env_step.npz")
    env = _env_from_golden(g)
    load_golden_step_inputs(env, g)
    env.post_physics_step()
    assert np.array_equal(env.done_buf.numpy(), g["done"]), "done flags must be bit-exact"
    assert np.array_equal(env.time_buf.numpy(), g["time_buf"])
    assert_close(env.ray_hfs, g["ray_hfs"], what="ray_hfs")
    for k in ("root_pos", "root_rot", "root_vel", "root_ang_vel", "body_pos", "joint_rot", "dof_pos", "dof_vel", "contacts"):
        assert_close(getattr(env, "ref_" + k), g["out_ref_" + k], atol=2e-6, what="ref_" + k)
    assert_close(env.obs_buf, g["obs"], atol=2e-6, what="obs")
    assert_close(env.reward_terms[:, 0:5], g["reward5"], what="reward5")
    assert_close(env.reward_terms[:, 5], g["contact_penalty"], what="contact penalty")
    assert_close(env.reward_buf, g["reward"], what="reward")
    assert_close(env.fail_rates, g["out_fail_rates"], rtol=1e-6, atol=0, what="fail-rate EMA")
    tar = env.tar_obs_data()
    for k, v in zip(("tar_root_pos", "tar_root_rot", "tar_joint_rot", "tar_key_pos", "tar_contacts"), tar):
        assert_close(v, g[k], atol=2e-6, what=k)


def test_env_reset_matches_reference():
    g = load_npz("ref_env_step.npz")
    env = _env_from_golden(g)
    N = env.N
    env.char_rigid_body_vel[:] = 1.0
    env.reset(torch.arange(N), t(g["reset_motion_ids"]), t(g["reset_terrain_ids"]), t(g["reset_times"]), t(g["reset_noise"]))
    assert np.array_equal(env.motion_ids.numpy(), g["reset_motion_ids"])
    for k in ("ref_root_pos", "ref_root_rot", "ref_root_vel", "ref_root_ang_vel", "ref_joint_rot", "ref_dof_pos", "ref_dof_vel",
              "ref_contacts", "char_root_pos", "char_root_rot", "char_root_vel", "char_root_ang_vel", "char_dof_pos",
              "char_dof_vel", "char_rigid_body_vel"):
        assert_close(getattr(env, k), g["reset_" + k], atol=2e-6, what=k)
    assert bool(env.actors_need_reset[:, 0].all())


def test_observation_tape_shipped_with_reference():
    """The recorded obs tape (write_agent_states, ig_parkour_env.py:759-774) next to the recorded
    frames: the columns that are a function of the stored frames alone must reproduce."""
    g = load_npz("clip_teaser_dm.npz")
    m = oracle_model()
    tape = t(g["obs_tape"])
    root_pos, root_rot, joint_rot = t(g["root_pos"]), t(g["root_rot"]), t(g["joint_rot"])
    n = root_pos.shape[0]
    body_pos = m.forward_kinematics(root_pos, root_rot, joint_rot)[0]
    key = [m.body_id(b) for b in ("right_hand", "left_hand", "right_foot", "left_foot")]
    obs = O.compute_char_obs(root_pos, root_rot, torch.zeros(n, 3), torch.zeros(n, 3), joint_rot, torch.zeros(n, m.D),
                             body_pos[:, key], False, False)
    assert obs.shape[1] == 136
    assert_close(obs[:, 0:6], tape[:, 0:6], rtol=1e-5, atol=5e-6, what="root tan-norm")
    pinned_joints = [0, 1, 8, 9, 10, 11, 12, 13]          # abdomen, neck, hips, knees, ankles (+ fixed hands below)
    for j in pinned_joints + [4, 7]:
        assert_close(obs[:, 12 + 6 * j:18 + 6 * j], tape[:, 12 + 6 * j:18 + 6 * j], rtol=1e-5, atol=5e-6, what="joint %d tan-norm" % j)
    assert_close(obs[:, 124:136], tape[:, 124:136], rtol=1e-4, atol=5e-5, what="key body positions")
    # heights: the sample points that fall inside the stored 18x15 terrain slice reproduce
    terrain = O.Terrain(t(g["hf"]), t(g["min_point"]), [float(g["dx"])] * 2)
    from parc_b200 import synth
    ray = synth.cone_ray_points()
    heading = O.calc_heading(root_rot)
    pts = O.rotate_2d_vec(ray.unsqueeze(0), heading.unsqueeze(-1)) + root_pos[:, None, 0:2]
    raw = torch.round((pts - terrain.min_point) / terrain.dxdy)
    inside = ((raw >= 0) & (raw <= (terrain.dims - 1))).all(dim=-1)
    h = torch.clamp(terrain.height_at(pts) - root_pos[:, 2:3], -3.0, 3.0)
    got, want = h[inside], tape[:, 871:][inside]
    assert inside.float().mean() > 0.8
    frac_ok = ((got - want).abs() <= 1e-5).float().mean().item()
    assert frac_ok > 0.999, "heightmap columns inside the stored slice: %.4f reproduce" % frac_ok


def test_tracking_error_matches_reference():
    """a19: oracle compute_tracking_error on the FK outputs the reference recorded (ref_tracking_error.npz)."""
    g, ge = load_npz("ref_env_step.npz"), load_npz("ref_tracking_error.npz")
    m = oracle_model()
    cbp, cbr = m.forward_kinematics(t(g["in_char_root_pos"]), t(g["in_char_root_rot"]), m.dof_to_rot(t(g["in_char_dof_pos"])))
    rbp, rbr = m.forward_kinematics(t(g["out_ref_root_pos"]), t(g["out_ref_root_rot"]), t(g["out_ref_joint_rot"]))
    assert_close(cbp, ge["char_body_pos"], atol=2e-6, what="char body pos")
    assert_close(rbr, ge["ref_body_rot"], atol=2e-6, what="ref body rot")
    err = O.compute_tracking_error(t(g["in_char_root_pos"]), t(g["in_char_root_rot"]), cbr, cbp, t(g["out_ref_root_pos"]),
                                   t(g["out_ref_root_rot"]), rbr, rbp, t(g["in_char_root_vel"]), t(g["in_char_root_ang_vel"]),
                                   t(g["in_char_dof_vel"]), t(g["out_ref_root_vel"]), t(g["out_ref_root_ang_vel"]), t(g["out_ref_dof_vel"]))
    assert_close(err, ge["tracking_error"], atol=2e-6, what="tracking error")
