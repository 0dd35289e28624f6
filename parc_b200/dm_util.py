"""The TorchScript free functions of the tracking path, backed by the stand-alone CUDA operators of
include/parc_b200_env.h. Same names, argument order and meaning as the reference:

  compute_char_obs            PARC/motion_tracker/envs/ig_char_env.py:581-625
  compute_tar_obs             PARC/motion_tracker/envs/ig_parkour/mgdm_dm_util.py:404-461
  compute_deepmimic_reward    mgdm_dm_util.py:269-332
  compute_contact_reward      mgdm_dm_util.py:497-518
  compute_done                mgdm_dm_util.py:334-402
  compute_tracking_error      mgdm_dm_util.py:520-553
  calc_heading                PARC/util/torch_util.py:501-510

The reference functions are independent of the character (they take tensors only); the kernels need
the body count, which is read off the tensor shapes. Inputs must be CUDA tensors; outputs are freshly
allocated and owned by the caller. For a whole control step use the fused `parc_b200_env_step`
(parc_b200.tracking_env.TrackingEnv) instead: these exist for callers that use the pieces one at a time."""
import torch

from . import _lib


def _c(t, dtype=torch.float32):
    return t.detach().to(dtype=dtype).contiguous()


def _shape_model(num_bodies, dof_size=0):
    """A parc_char_model_t that only carries the sizes the element-wise operators read."""
    m = _lib.CharModelStruct()
    m.num_bodies, m.dof_size = int(num_bodies), int(dof_size)
    return m


def calc_heading(q):
    q = _c(q)
    n = q.numel() // 4
    out = torch.empty(q.shape[:-1], dtype=torch.float32, device=q.device)
    _lib.check(_lib.lib().parc_b200_calc_heading(_lib.ptr(q), n, _lib.ptr(out), _lib.stream_ptr(q.device)))
    return out


def compute_char_obs(root_pos, root_rot, root_vel, root_ang_vel, joint_rot, dof_vel, key_pos, global_obs: bool, root_height_obs: bool):
    n, J1, D = root_pos.shape[0], joint_rot.shape[-2], dof_vel.shape[-1]
    K = key_pos.shape[-2] if key_pos.dim() >= 2 else 0
    obs = torch.empty(n, (1 if root_height_obs else 0) + 12 + 6 * J1 + D + 3 * K, dtype=torch.float32, device=root_pos.device)
    if n == 0:
        return obs
    m = _shape_model(J1 + 1, D)
    a = [_c(x) for x in (root_pos, root_rot, root_vel, root_ang_vel, joint_rot, dof_vel)]
    kp = _c(key_pos) if K > 0 else None
    _lib.check(_lib.lib().parc_b200_compute_char_obs(m, *[_lib.ptr(x) for x in a], _lib.ptr(kp), K, int(bool(global_obs)),
                                                     int(bool(root_height_obs)), n, _lib.ptr(obs), _lib.stream_ptr(obs.device)))
    return obs


def compute_tar_obs(ref_root_pos, ref_root_rot, tar_root_pos, tar_root_rot, joint_rot, tar_key_pos, global_obs: bool,
                    global_tar_root_h_obs: bool):
    n, S, J1 = tar_root_pos.shape[0], tar_root_pos.shape[1], joint_rot.shape[-2]
    K = tar_key_pos.shape[-2] if tar_key_pos.dim() >= 2 else 0
    obs = torch.empty(n, S, 9 + 6 * J1 + 3 * K, dtype=torch.float32, device=tar_root_pos.device)
    if n == 0 or S == 0:
        return obs
    m = _shape_model(J1 + 1)
    a = [_c(x) for x in (ref_root_pos, ref_root_rot, tar_root_pos, tar_root_rot, joint_rot)]
    kp = _c(tar_key_pos) if K > 0 else None
    _lib.check(_lib.lib().parc_b200_compute_tar_obs(m, *[_lib.ptr(x) for x in a], _lib.ptr(kp), S, K, int(bool(global_obs)),
                                                    int(bool(global_tar_root_h_obs)), n, _lib.ptr(obs), _lib.stream_ptr(obs.device)))
    return obs


def compute_deepmimic_obs(root_pos, root_rot, root_vel, root_ang_vel, joint_rot, dof_vel, key_pos, global_obs, root_height_obs,
                          enable_tar_obs, tar_root_pos, tar_root_rot, tar_joint_rot, tar_key_pos):
    """mgdm_dm_util.py:463-495 -> OrderedDict(char_obs [n, .], tar_obs [n, S, .])."""
    from collections import OrderedDict
    obs = OrderedDict()
    obs["char_obs"] = compute_char_obs(root_pos=root_pos, root_rot=root_rot, root_vel=root_vel, root_ang_vel=root_ang_vel, joint_rot=joint_rot,
                                       dof_vel=dof_vel, key_pos=key_pos, global_obs=global_obs, root_height_obs=root_height_obs)
    if enable_tar_obs:
        obs["tar_obs"] = compute_tar_obs(ref_root_pos=root_pos, ref_root_rot=root_rot, tar_root_pos=tar_root_pos, tar_root_rot=tar_root_rot,
                                         joint_rot=tar_joint_rot, tar_key_pos=tar_key_pos, global_obs=global_obs, global_tar_root_h_obs=False)
    return obs


def compute_deepmimic_reward(root_pos, root_rot, root_vel, root_ang_vel, joint_rot, dof_vel, key_pos, tar_root_pos, tar_root_rot,
                             tar_root_vel, tar_root_ang_vel, tar_joint_rot, tar_dof_vel, tar_key_pos, joint_rot_err_w, dof_err_w,
                             track_root_h: bool, track_root: bool):
    n, J1, D = root_pos.shape[0], joint_rot.shape[-2], dof_vel.shape[-1]
    K = key_pos.shape[-2] if key_pos.dim() >= 2 else 0
    out = torch.empty(n, 5, dtype=torch.float32, device=root_pos.device)
    if n == 0:
        return out
    m = _shape_model(J1 + 1, D)
    a = [_c(x) for x in (root_pos, root_rot, root_vel, root_ang_vel, joint_rot, dof_vel)]
    b = [_c(x) for x in (tar_root_pos, tar_root_rot, tar_root_vel, tar_root_ang_vel, tar_joint_rot, tar_dof_vel)]
    kp, tkp = (_c(key_pos), _c(tar_key_pos)) if K > 0 else (None, None)
    w = [_c(joint_rot_err_w), _c(dof_err_w)]
    _lib.check(_lib.lib().parc_b200_compute_deepmimic_reward(
        m, *[_lib.ptr(x) for x in a], _lib.ptr(kp), *[_lib.ptr(x) for x in b], _lib.ptr(tkp), _lib.ptr(w[0]), _lib.ptr(w[1]), K,
        int(bool(track_root_h)), int(bool(track_root)), n, _lib.ptr(out), _lib.stream_ptr(out.device)))
    return out


def compute_contact_reward(tar_contacts, contact_forces, contact_weights):
    n, C = tar_contacts.shape[0], tar_contacts.shape[-1]
    out = torch.empty(n, C, dtype=torch.float32, device=tar_contacts.device)
    if n == 0:
        return out
    a = [_c(tar_contacts), _c(contact_forces), _c(contact_weights)]
    _lib.check(_lib.lib().parc_b200_compute_contact_reward(*[_lib.ptr(x) for x in a], C, n, _lib.ptr(out), _lib.stream_ptr(out.device)))
    return out


def compute_done(done_buf, time, ep_len: float, root_rot, body_pos, char_root_pos, tar_root_rot, tar_body_pos, contact_force,
                 contact_body_ids, termination_heights, pose_termination: bool, pose_termination_dist, global_obs: bool,
                 enable_early_termination: bool, track_root: bool, root_pos_termination_dist: float,
                 root_rot_termination_angle: float):
    """`char_root_pos` and `global_obs` are accepted and unused, as in the reference body."""
    n, J = time.shape[0], body_pos.shape[-2]
    out = torch.empty(n, dtype=torch.int32, device=time.device)
    if n == 0:
        return out
    f = [_c(x) for x in (time, root_rot, body_pos, tar_root_rot, tar_body_pos, contact_force)]
    ids = _c(contact_body_ids, torch.int64) if contact_body_ids is not None and contact_body_ids.numel() > 0 else None
    th = _c(termination_heights) if termination_heights is not None else None
    ptd = _c(pose_termination_dist).reshape(-1)
    if ptd.numel() == 1:                      # a scalar threshold broadcasts over the bodies, as in the reference
        ptd = ptd.expand(J - 1).contiguous()
    if pose_termination and ptd.numel() != J - 1:
        raise ValueError("pose_termination_dist must have 1 or %d entries, got %d" % (J - 1, ptd.numel()))
    done_in = _c(done_buf, torch.int32)
    _lib.check(_lib.lib().parc_b200_compute_done(
        _lib.ptr(done_in), _lib.ptr(f[0]), float(ep_len), _lib.ptr(f[1]), _lib.ptr(f[2]), _lib.ptr(f[3]), _lib.ptr(f[4]), _lib.ptr(f[5]),
        _lib.ptr(ids), 0 if ids is None else int(ids.numel()), _lib.ptr(th), int(bool(pose_termination)), _lib.ptr(ptd),
        int(bool(enable_early_termination)), int(bool(track_root)), float(root_pos_termination_dist),
        float(root_rot_termination_angle), J, n, _lib.ptr(out), _lib.stream_ptr(out.device)))
    return out


def compute_tracking_error(root_pos, root_rot, body_rot, body_pos, tar_root_pos, tar_root_rot, tar_body_rot, tar_body_pos, root_vel,
                           root_ang_vel, dof_vel, tar_root_vel, tar_root_ang_vel, tar_dof_vel):
    """-> [n,7]: root pos L2, |root rot angle|, mean root-relative body pos L2, mean |body rot angle|, mean |dof vel diff|,
    mean |root vel diff|, mean |root ang vel diff|"""
    n, J, D = root_pos.shape[0], body_rot.shape[-2], dof_vel.shape[-1]
    out = torch.empty(n, 7, dtype=torch.float32, device=root_pos.device)
    if n == 0:
        return out
    a = [_c(x) for x in (root_pos, root_rot, body_rot, body_pos, tar_root_pos, tar_root_rot, tar_body_rot, tar_body_pos, root_vel,
                         root_ang_vel, dof_vel, tar_root_vel, tar_root_ang_vel, tar_dof_vel)]
    _lib.check(_lib.lib().parc_b200_compute_tracking_error(*[_lib.ptr(x) for x in a], J, D, n, _lib.ptr(out), _lib.stream_ptr(out.device)))
    return out
