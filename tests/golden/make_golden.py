"""Generates the golden fixtures under tests/golden/ by running the UNMODIFIED PARC reference
(imported from /root/reference through oracle/ref_import.py) on CPU.

Run here (the authoring container) only:  python tests/golden/make_golden.py
The GPU box has no /root/reference; tests read the committed .npz files instead.

Fixtures
  clip_civilization.npz   the bundled clip data/motion_terrains/civilization.pkl as plain arrays
  clip_teaser_dm.npz      dec2024_teaser_717_1_opt_dm.pkl + the recorded observation tape
                          misc_data["obs"] [142,1312] -- the only golden vector the reference ships
  ref_kin.npz             KinCharModel.dof_to_rot / forward_kinematics / rot_to_dof on seeded poses
  ref_motionlib.npz       MotionLib load-time tensors + calc_motion_frame/_calc_frame_blend (CLAMP and WRAP)
  ref_env_step.npz        one _post_physics_step of a standalone DeepMimicEnv (N=96) driven exactly
                          like IGParkourEnv drives it (heights, time, ref motion, obs, reward, done,
                          fail-rate EMA) and one RSI reset with recorded random draws
"""
import os
import sys
from collections import OrderedDict

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
import parc.anim.motion_lib as rmlib  # noqa: E402
import parc.motion_tracker.envs.ig_parkour.dm_env as rdm  # noqa: E402
import parc.motion_tracker.envs.ig_parkour.mgdm_dm_util as rutil  # noqa: E402
import parc.util.geom_util as rgeom  # noqa: E402
import parc.util.torch_util as rtu  # noqa: E402
import parc.util.file_io_helper as rfio  # noqa: E402

from parc_b200 import synth  # noqa: E402  (input generation only: seeded noise)


def npz(name, **arrs):
    out = {}
    for k, v in arrs.items():
        out[k] = v.detach().cpu().numpy() if isinstance(v, torch.Tensor) else np.asarray(v)
    np.savez_compressed(os.path.join(HERE, name), **out)
    print("wrote", name, {k: v.shape for k, v in out.items()})


def dump_clip(pkl, out_name, with_tape=False):
    d = rfio.load_ms_file(ref_import.clip(pkl), device="cpu")
    md, td = d.motion_data, d.terrain_data
    arrs = dict(root_pos=md.root_pos, root_rot=md.root_rot, joint_rot=md.joint_rot, contacts=md.body_contacts,
                fps=np.array(md.fps), loop_mode=np.array(rmlib.LoopMode[md.loop_mode].value),
                hf=td.hf, min_point=td.min_point, dx=np.array(float(td.dx), dtype=np.float32))
    if with_tape:
        arrs["obs_tape"] = np.asarray(d.misc_data["obs"], dtype=np.float32)
        arrs["min_point_offset"] = np.asarray(d.misc_data["min_point_offset"], dtype=np.float32)
    npz(out_name, **arrs)


def main():
    model = rkin.KinCharModel("cpu")
    model.load_char_file(ref_import.humanoid_xml())
    J, D = model.get_num_joints(), model.get_dof_size()

    dump_clip("civilization.pkl", "clip_civilization.npz")
    dump_clip("dec2024_teaser_717_1_opt_dm.pkl", "clip_teaser_dm.npz", with_tape=True)

    # ---- KinCharModel ---------------------------------------------------------------------------
    g = torch.Generator().manual_seed(11)
    n = 96
    lo, hi = model._lower_dof_limits, model._upper_dof_limits
    dof = lo + torch.rand(n, D, generator=g) * (hi - lo)
    dof[0] = 0.0                                   # identity pose: exercises the min_theta branches
    dof[1, 0:3] = torch.tensor([1e-6, 0.0, 0.0])   # below exp-map min_theta
    dof[2, 0:3] = torch.tensor([3.3, 0.0, 0.0])    # angle > pi: normalize_angle wraps
    root_pos = torch.randn(n, 3, generator=g)
    root_rot = rtu.quat_normalize(torch.randn(n, 4, generator=g))
    rot = model.dof_to_rot(dof)
    body_pos, body_rot = model.forward_kinematics(root_pos, root_rot, rot)
    dof_back = model.rot_to_dof(rot)
    # rot_to_dof on arbitrary unit quaternions (w < 0 included -> quat_pos branch, hinge sign flip)
    rnd_rot = rtu.normalize(torch.randn(n, J - 1, 4, generator=g))
    dof_rnd = model.rot_to_dof(rnd_rot)
    npz("ref_kin.npz", dof=dof, root_pos=root_pos, root_rot=root_rot, rot=rot, body_pos=body_pos, body_rot=body_rot,
        dof_back=dof_back, rnd_rot=rnd_rot, dof_rnd=dof_rnd)

    # ---- MotionLib ------------------------------------------------------------------------------
    lib = rmlib.MotionLib.from_file(ref_import.clip("civilization_motions.yaml") if False else ref_import.clip("civilization.pkl"),
                                    model, "cpu", contact_info=True)
    L = lib._motion_lengths[0].item()
    nq = 512
    times = torch.rand(nq, generator=g) * (L * 1.3) - 0.1 * L           # before 0 and past the end too
    times[:64] = (1.0 / 30.0) * torch.arange(64, dtype=torch.float32)   # exactly on frames (fp32 dt * k)
    times[64:128] = torch.tensor(1.0 / 30.0, dtype=torch.float32) * torch.arange(190, 254, dtype=torch.int32)
    ids = torch.zeros(nq, dtype=torch.int64)
    i0, i1, bl = lib._calc_frame_blend(ids, times)
    out = lib.calc_motion_frame(ids, times)
    # WRAP variant of the same clip (all bundled clips are CLAMP): flip the loop mode in place
    lib_w = rmlib.MotionLib.from_file(ref_import.clip("civilization.pkl"), model, "cpu", contact_info=True)
    lib_w._motion_loop_modes[:] = rmlib.LoopMode.WRAP.value
    times_w = torch.rand(nq, generator=g) * (L * 4.0) - L
    wi0, wi1, wbl = lib_w._calc_frame_blend(ids, times_w)
    out_w = lib_w.calc_motion_frame(ids, times_w)
    names = ["root_pos", "root_rot", "root_vel", "root_ang_vel", "joint_rot", "dof_vel", "contacts"]
    arrs = dict(times=times, idx0=i0, idx1=i1, blend=bl, times_w=times_w, w_idx0=wi0, w_idx1=wi1, w_blend=wbl,
                frame_root_vel=lib._frame_root_vel, frame_root_ang_vel=lib._frame_root_ang_vel,
                frame_dof_vel=lib._frame_dof_vel, motion_length=lib._motion_lengths,
                root_pos_delta=lib._motion_root_pos_delta)
    for k, v in zip(names, out):
        arrs[k] = v
    for k, v in zip(names, out_w):
        arrs["w_" + k] = v
    npz("ref_motionlib.npz", **arrs)

    # ---- standalone DeepMimicEnv step ------------------------------------------------------------
    with open(ref_import.clip("dm_env_civilization.yaml"), "r") as f:
        cfg = yaml.safe_load(f)
    cfg["env"]["dm"]["motion_file"] = ref_import.clip("civilization.pkl")
    cfg["env"]["dm"]["terrain_build_mode"] = "file"
    ec = cfg["env"]
    N = 96
    env = rdm.DeepMimicEnv(cfg, N, "cpu", False, model)
    env.build_terrain(ec, None)
    z = lambda *s: torch.zeros(*s, dtype=torch.float32)
    ref = OrderedDict(ref_root_pos=z(N, 3), ref_root_rot=z(N, 4), ref_root_vel=z(N, 3), ref_root_ang_vel=z(N, 3),
                      ref_body_pos=z(N, J, 3), ref_joint_rot=z(N, J - 1, 4), ref_dof_pos=z(N, D), ref_dof_vel=z(N, D),
                      ref_contacts=z(N, J))
    char = OrderedDict(char_root_pos=z(N, 3), char_root_rot=z(N, 4), char_root_vel=z(N, 3), char_root_ang_vel=z(N, 3),
                       char_dof_pos=z(N, D), char_dof_vel=z(N, D), char_contact_forces=z(N, J, 3),
                       char_rigid_body_pos=z(N, J, 3), char_rigid_body_vel=z(N, J, 3), char_rigid_body_ang_vel=z(N, J, 3))
    env.get_sim_tensor_views(**ref, **char)
    reward_buf, time_buf = z(N), z(N)
    done_buf = torch.zeros(N, dtype=torch.int32)
    timestep_buf = torch.zeros(N, dtype=torch.int32)
    actors_need_reset = torch.zeros(N, 2, dtype=torch.bool)
    env_offsets = synth.env_offsets(N, env_spacing=ec.get("env_spacing", 2.0))
    key_body_ids = torch.tensor([model.get_body_id(b) for b in ec["key_bodies"]], dtype=torch.int64)
    ray_xy = rgeom.get_xy_points_cone(center=torch.zeros(2), dx=ec["ray_dx"], num_neg=ec["ray_points_behind"],
                                      num_pos=ec["ray_points_ahead"], num_rays_neg=ec["ray_num_left"],
                                      num_rays_pos=ec["ray_num_right"], angle_between_rays=ec["ray_angle"])
    P = ray_xy.shape[0]
    ray_hfs = z(N, P)
    env.get_data_buffer_views(reward_buf, done_buf, time_buf, timestep_buf, actors_need_reset, env_offsets, key_body_ids,
                              ray_xy, ray_hfs)
    tar_obs_steps = torch.tensor(ec["tar_obs_steps"], dtype=torch.int)
    timestep = 1.0 / ec["control_freq"]

    # reset with recorded draws: replay torch's generator in the reference's call order
    all_ids = torch.arange(N, dtype=torch.int64)
    seed = 7
    torch.manual_seed(seed)
    w = torch.clamp(env._motion_id_fail_rates, min=env._min_motion_weight) * env._mlib._motion_weights
    d_mids = torch.multinomial(w, num_samples=N, replacement=True)
    d_tids = torch.randint(high=env._terrains_per_motion, size=(N,), dtype=torch.int64)
    d_phase = torch.rand(d_mids.shape)
    d_times = d_phase * env._mlib._motion_lengths[d_mids]
    d_noise = torch.rand(size=[N, 2], dtype=torch.float32) * 2.0 - 1.0
    torch.manual_seed(seed)
    env.reset(all_ids)
    assert torch.equal(env._motion_ids, d_mids) and torch.equal(env._motion_time_offsets, d_times)
    reset_out = dict(reset_motion_ids=d_mids, reset_terrain_ids=d_tids, reset_times=d_times, reset_noise=d_noise,
                     reset_scale=np.array(ec["rand_root_pos_offset_scale"], dtype=np.float32))
    for k, v in list(ref.items()) + list(char.items()):
        if k != "ref_body_pos":          # not written by reset
            reset_out["reset_" + k] = v.clone()
    reset_out["reset_actors_need_reset"] = actors_need_reset.clone()

    # put the envs at assorted times, make a simulated state around the reference state
    g2 = torch.Generator().manual_seed(21)
    timestep_buf[:] = torch.randint(0, 200, (N,), generator=g2, dtype=torch.int32)
    timestep_buf[0:4] = 0                       # first step: failures suppressed by `time > 1e-5`? (time = dt after update)
    timestep_buf[4:8] = 299                     # -> time >= episode_length after the update (TIME)
    env._motion_time_offsets[8:12] = env._mlib._motion_lengths[0] - 0.01   # motion end
    time_buf[:] = timestep * timestep_buf
    # state of the reference one step ahead, perturbed
    ref_next = env._mlib.calc_motion_frame(env._motion_ids, time_buf + timestep + env._motion_time_offsets)
    rp = ref_next[0].clone()
    rp[:, 0:2] = env._move_to_motion_terrain(rp[:, 0:2])
    ref_state = dict(root_pos=rp, root_rot=ref_next[1], root_vel=ref_next[2], root_ang_vel=ref_next[3],
                     dof_pos=model.rot_to_dof(ref_next[4]), dof_vel=ref_next[5])
    noise = synth.sim_noise(N, J, D, seed=5678)
    noise["root_pos"][12:20] += 0.7             # root position failures
    noise["root_rot_exp"][20:26, 2] += 1.6      # root rotation failures
    noise["dof_pos"][26:34] += 1.2              # pose failures
    sim = synth.perturb_state(ref_state, noise, model.forward_kinematics, model.dof_to_rot)
    char["char_root_pos"][:], char["char_root_rot"][:] = sim["root_pos"], sim["root_rot"]
    char["char_root_vel"][:], char["char_root_ang_vel"][:] = sim["root_vel"], sim["root_ang_vel"]
    char["char_dof_pos"][:], char["char_dof_vel"][:] = sim["dof_pos"], sim["dof_vel"]
    char["char_rigid_body_pos"][:], char["char_contact_forces"][:] = sim["body_pos"], sim["contact_forces"]
    env._motion_id_fail_rates[:] = 0.5
    inputs = dict(in_timestep_buf=timestep_buf.clone(), in_motion_ids=env._motion_ids.clone(),
                  in_motion_terrain_ids=env._motion_terrain_ids.clone(), in_motion_time_offsets=env._motion_time_offsets.clone(),
                  in_fail_rates=env._motion_id_fail_rates.clone(), env_offsets=env_offsets, ray_xy=ray_xy,
                  dm_motion_offsets=env._dm_motion_offsets[0:1].clone())
    for k, v in char.items():
        inputs["in_" + k] = v.clone()

    # ---- IGEnv._post_physics_step order (ig_env.py:368-377) -------------------------------------
    # _refresh_sim_tensors -> _refresh_obs_hfs (ig_parkour_env.py:518-533)
    char_root_pos_xyz = char["char_root_pos"] + env_offsets
    env._refresh_obs_hfs(char_root_pos_xyz, rtu.calc_heading(char["char_root_rot"]))
    # _update_time (ig_env.py:391-394)
    timestep_buf += 1
    time_buf[:] = timestep * timestep_buf
    # _update_misc -> _update_ref_motion
    env._update_ref_motion()
    # _update_observations -> _compute_obs (ig_parkour_env.py:842-965)
    joint_rot = model.dof_to_rot(char["char_dof_pos"])
    body_pos, _ = model.forward_kinematics(char["char_root_pos"], char["char_root_rot"], joint_rot)
    key_pos = body_pos[..., key_body_ids, :]
    tar = env.compute_tar_obs(tar_obs_steps, None)
    od = rutil.compute_deepmimic_obs(root_pos=char["char_root_pos"], root_rot=char["char_root_rot"],
                                     root_vel=char["char_root_vel"], root_ang_vel=char["char_root_ang_vel"],
                                     joint_rot=joint_rot, dof_vel=char["char_dof_vel"], key_pos=key_pos,
                                     global_obs=ec["global_obs"], root_height_obs=ec["global_root_height_obs"],
                                     enable_tar_obs=True, tar_root_pos=tar[0], tar_root_rot=tar[1], tar_joint_rot=tar[2],
                                     tar_key_pos=tar[3])
    char_contacts = (torch.norm(char["char_contact_forces"], dim=-1) > 1e-5).to(torch.float32)
    obs = torch.cat([od["char_obs"], od["tar_obs"].reshape(N, -1), tar[4].reshape(N, -1), char_contacts, ray_hfs], dim=-1)
    # _update_reward (ig_parkour_env.py:984-1044)
    joint_err_w = torch.tensor(ec["joint_err_w"], dtype=torch.float32)
    dof_err_w = torch.zeros(D)
    for j in range(1, J):
        dd = model.get_joint_dof_dim(j)
        if dd > 0:
            di = model.get_joint_dof_idx(j)
            dof_err_w[di:di + dd] = joint_err_w[j - 1]
    comp = rutil.compute_deepmimic_reward(
        root_pos=char["char_root_pos"], root_rot=char["char_root_rot"], root_vel=char["char_root_vel"],
        root_ang_vel=char["char_root_ang_vel"], joint_rot=joint_rot, dof_vel=char["char_dof_vel"],
        key_pos=char["char_rigid_body_pos"][..., key_body_ids, :], tar_root_pos=ref["ref_root_pos"],
        tar_root_rot=ref["ref_root_rot"], tar_root_vel=ref["ref_root_vel"], tar_root_ang_vel=ref["ref_root_ang_vel"],
        tar_joint_rot=ref["ref_joint_rot"], tar_dof_vel=ref["ref_dof_vel"],
        tar_key_pos=ref["ref_body_pos"][..., key_body_ids, :], joint_rot_err_w=joint_err_w, dof_err_w=dof_err_w,
        track_root_h=ec["track_root_h"], track_root=ec["track_root"])
    tw = ec["pose_w"] + ec["vel_w"] + ec["root_pos_w"] + ec["root_vel_w"] + ec["key_pos_w"]
    dm_r = (ec["pose_w"] / tw) * comp[:, 0] + (ec["vel_w"] / tw) * comp[:, 1] + (ec["root_pos_w"] / tw) * comp[:, 2] \
        + (ec["root_vel_w"] / tw) * comp[:, 3] + (ec["key_pos_w"] / tw) * comp[:, 4]
    cw = torch.tensor(ec["contact_weights"], dtype=torch.float32)
    pen = torch.mean(rutil.compute_contact_reward(tar_contacts=ref["ref_contacts"],
                                                  contact_forces=char["char_contact_forces"][:, model.get_contact_body_ids()],
                                                  contact_weights=cw), dim=-1)
    dm_r = dm_r + pen
    # _update_done (ig_parkour_env.py:967-982)
    contact_body_ids = torch.tensor([model.get_body_id(b) for b in ec["contact_bodies"]], dtype=torch.int64)
    env.update_done(termination_height=torch.tensor(ec["termination_height"], dtype=torch.float32),
                    episode_length=ec["episode_length"], contact_body_ids=contact_body_ids,
                    pose_termination=ec["pose_termination"],
                    pose_termination_dist=torch.tensor(ec["pose_termination_dist"], dtype=torch.float32),
                    global_obs=ec["global_obs"], enable_early_termination=ec["enable_early_termination"],
                    track_root=ec["track_root"], root_pos_termination_dist=ec["root_pos_termination_dist"],
                    root_rot_termination_angle=ec["root_rot_termination_angle"])
    print("done histogram:", torch.bincount(done_buf.long(), minlength=4).tolist())
    outs = dict(obs=obs, reward5=comp, contact_penalty=pen, reward=dm_r, done=done_buf, time_buf=time_buf,
                ray_hfs=ray_hfs, out_fail_rates=env._motion_id_fail_rates,
                tar_root_pos=tar[0], tar_root_rot=tar[1], tar_joint_rot=tar[2], tar_key_pos=tar[3], tar_contacts=tar[4])
    for k, v in ref.items():
        outs["out_" + k] = v
    cfg_arr = dict(cfg_yaml=np.array(yaml.safe_dump({"env": {k: v for k, v in ec.items() if k not in ("dm", "mgdm", "plane")}})))
    npz("ref_env_step.npz", **inputs, **outs, **reset_out, **cfg_arr)


if __name__ == "__main__":
    main()
