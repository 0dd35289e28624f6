"""Shared test plumbing: load fixtures, build the same scenario for the CPU oracle and the CUDA path."""
import json
import os

import numpy as np
import torch

from oracle import parc_oracle as O
from parc_b200 import synth

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load_npz(name):
    with np.load(os.path.join(GOLDEN, name), allow_pickle=False) as z:
        return {k: z[k] for k in z.files}


def t(x):
    return torch.from_numpy(np.ascontiguousarray(x))


def golden_clip(name="clip_civilization.npz"):
    g = load_npz(name)
    return dict(root_pos=t(g["root_pos"]), root_rot=t(g["root_rot"]), joint_rot=t(g["joint_rot"]), contacts=t(g["contacts"]),
                fps=int(g["fps"]), loop_mode=int(g["loop_mode"]), hf=t(g["hf"]), min_point=t(g["min_point"]), dx=float(g["dx"]))


def aug_extras(num_clips=2):
    """What the heightfield augmenter reads beyond the clip files (tests/golden/clip_aug_extras.npz, written by
    make_golden_sampler.py): -> (hf_maxmins [list of [X,Y,2]], mask_inds [list over clips of list over frames of [k,2] i64])"""
    g = load_npz("clip_aug_extras.npz")
    mms, masks = [], []
    for k in range(num_clips):
        mms.append(t(g["c%d_hf_maxmin" % k]))
        st, cells = g["c%d_mask_start" % k], t(g["c%d_mask_cells" % k]).long()
        masks.append([cells[int(st[f]):int(st[f + 1])] for f in range(len(st) - 1)])
    return mms, masks


def oracle_model():
    return O.CharModel(synth.humanoid_spec())


def oracle_clip(c):
    d = dict(c)
    d["terrain"] = O.Terrain(c["hf"], c["min_point"], [c["dx"], c["dx"]])
    return d


def assert_close(a, b, rtol=1e-5, atol=1e-6, what=""):
    """|a-b| <= atol + rtol*|b| (north_star tolerance: 1e-5 relative, 1e-6 absolute floor)."""
    a = torch.as_tensor(a).detach().cpu().double()
    b = torch.as_tensor(b).detach().cpu().double()
    assert a.shape == b.shape, "%s: shape %s vs %s" % (what, tuple(a.shape), tuple(b.shape))
    err = (a - b).abs()
    tol = atol + rtol * b.abs()
    bad = err > tol
    if bad.any():
        i = torch.argmax(err - tol)
        raise AssertionError("%s: %d/%d elements out of tolerance; worst |%g - %g| = %g at flat index %d" %
                             (what, int(bad.sum()), bad.numel(), a.flatten()[i], b.flatten()[i], err.flatten()[i], int(i)))


def assert_quat_close(a, b, atol=2e-6, what=""):
    """Unit quaternions are compared component-wise with an absolute tolerance (components are <= 1)."""
    assert_close(a, b, rtol=1e-5, atol=atol, what=what)


# ---------------------------------------------------------------------------------------------
# CUDA-side construction (imported lazily so CPU-only tests never touch the CUDA library)
# ---------------------------------------------------------------------------------------------
def product_model(device="cuda:0"):
    from parc_b200 import KinCharModel
    m = KinCharModel(device)
    m.load_char_file(synth.HUMANOID_JSON)
    return m


def product_terrain(hf, min_point, dx, device="cuda:0"):
    from parc_b200 import SubTerrain
    s = SubTerrain(x_dim=hf.shape[0], y_dim=hf.shape[1], dx=dx, dy=dx, min_x=float(min_point[0]), min_y=float(min_point[1]),
                   device=device)
    s.hf = hf.to(device=device, dtype=torch.float32).contiguous().clone()
    return s


def product_library(clips, model, device="cuda:0"):
    from parc_b200 import MotionLib
    pc = []
    for c in clips:
        pc.append(dict(root_pos=c["root_pos"], root_rot=c["root_rot"], joint_rot=c["joint_rot"], contacts=c["contacts"],
                       fps=c["fps"], loop_mode=c["loop_mode"], terrain=product_terrain(c["hf"], c["min_point"], c["dx"], device)))
    return MotionLib.from_clips(pc, model, device, contact_info=True)


def oracle_env_config_from_dict(m, ec):
    return O.EnvConfig(m, control_freq=ec["control_freq"], episode_length=ec["episode_length"],
                       tar_obs_steps=ec["tar_obs_steps"], key_bodies=ec["key_bodies"], global_obs=ec["global_obs"],
                       root_height_obs=ec["global_root_height_obs"], enable_tar_obs=ec.get("enable_tar_obs", True),
                       use_contact_info=ec["use_contact_info"], track_root=ec["track_root"], track_root_h=ec["track_root_h"],
                       enable_early_termination=ec["enable_early_termination"], pose_termination=ec["pose_termination"],
                       contact_bodies=ec["contact_bodies"], never_done=ec.get("never_done", False),
                       termination_height=ec["termination_height"],
                       root_pos_termination_dist=ec["root_pos_termination_dist"],
                       root_rot_termination_angle=ec["root_rot_termination_angle"],
                       pose_termination_dist=ec["pose_termination_dist"], joint_err_w=ec["joint_err_w"],
                       contact_weights=ec["contact_weights"], pose_w=ec["pose_w"], vel_w=ec["vel_w"],
                       root_pos_w=ec["root_pos_w"], root_vel_w=ec["root_vel_w"], key_pos_w=ec["key_pos_w"],
                       min_obs_h=ec["min_obs_h"], max_obs_h=ec["max_obs_h"],
                       rand_root_pos_offset_scale=ec["rand_root_pos_offset_scale"],
                       min_motion_weight=ec.get("dm", {}).get("min_motion_weight", 0.01),
                       terrains_per_motion=ec.get("dm", {}).get("terrains_per_motion", 1))


class Scenario:
    """The same synthetic world for the oracle (CPU) and, on demand, the CUDA TrackingEnv."""

    def __init__(self, num_envs, num_motions, ray="grid", seed=1234, wrap_fraction=0.1, env_config=None, min_frames=60,
                 max_frames=300):
        from parc_b200.tracking_env import default_env_config
        self.N, self.M = num_envs, num_motions
        self.ec = default_env_config() if env_config is None else env_config
        self.clips = synth.make_clips(synth.humanoid_spec(), num_motions, seed=seed, wrap_fraction=wrap_fraction,
                                      min_frames=min_frames, max_frames=max_frames)
        self.hf, self.min_point, self.dx, self.dm_offsets = synth.build_terrain_square(self.clips, padding_cells=1)
        self.ray = synth.grid_ray_points() if ray == "grid" else synth.cone_ray_points()
        self.env_offsets = synth.env_offsets(num_envs)
        self._oracle = None

    def oracle_env(self):
        m = oracle_model()
        lib = O.Library(m, [oracle_clip(c) for c in self.clips])
        terrain = O.Terrain(self.hf, self.min_point, [self.dx, self.dx])
        cfg = oracle_env_config_from_dict(m, self.ec)
        return O.TrackingEnv(m, lib, terrain, self.dm_offsets, cfg, self.N, self.env_offsets, self.ray)

    def product_env(self, device="cuda:0", **kw):
        from parc_b200.tracking_env import TrackingEnv
        m = product_model(device)
        lib = product_library(self.clips, m, device)
        terrain = product_terrain(self.hf, self.min_point, self.dx, device)
        return TrackingEnv(self.ec, self.N, device, m, lib, terrain, self.dm_offsets, self.ray, env_offsets=self.env_offsets, **kw)

    def draws(self, n, weights, lengths, seed):
        return synth.reset_draws(n, self.M, 1, lengths, weights, seed)


def copy_oracle_state_to_product(oenv, penv):
    """Uploads every input buffer of the oracle env into the CUDA env (same scenario)."""
    dev = penv._device
    for k in ("root_pos", "root_rot", "root_vel", "root_ang_vel", "dof_pos", "dof_vel", "contact_forces", "rigid_body_pos"):
        getattr(penv, "_char_" + k).copy_(getattr(oenv, "char_" + k).to(dev))
    for k in ("root_pos", "root_rot", "root_vel", "root_ang_vel", "body_pos", "joint_rot", "dof_pos", "dof_vel", "contacts"):
        getattr(penv, "_ref_" + k).copy_(getattr(oenv, "ref_" + k).to(dev))
    penv._timestep_buf.copy_(oenv.timestep_buf.to(dev))
    penv._time_buf.copy_(oenv.time_buf.to(dev))
    penv._done_buf.copy_(oenv.done_buf.to(dev))
    penv._motion_ids.copy_(oenv.motion_ids.to(dev))
    penv._motion_terrain_ids.copy_(oenv.motion_terrain_ids.to(dev))
    penv._motion_time_offsets.copy_(oenv.motion_time_offsets.to(dev))
    penv._motion_id_fail_rates.copy_(oenv.fail_rates.to(dev))


def randomize_oracle_state(oenv, seed=99, fail_fraction=0.15):
    """Reset every env with seeded draws, advance to random step counts and install a perturbed
    'simulated' state around the next reference pose (some envs pushed far enough to fail)."""
    N, m, lib = oenv.N, oenv.model, oenv.lib
    mids, tids, times, noise = synth.reset_draws(N, lib.num_motions(), oenv.cfg.terrains_per_motion, lib.lengths,
                                                 oenv.sampling_weights(), seed)
    oenv.reset(torch.arange(N), mids, tids, times, noise)
    g = torch.Generator().manual_seed(seed + 1)
    oenv.timestep_buf[:] = torch.randint(0, 120, (N,), generator=g, dtype=torch.int32)
    oenv.timestep_buf[: max(N // 64, 1)] = 299
    oenv.time_buf[:] = oenv.timestep * oenv.timestep_buf
    nxt = lib.calc_motion_frame(oenv.motion_ids, oenv.time_buf + oenv.timestep + oenv.motion_time_offsets)
    rp = nxt[0].clone()
    rp[:, 0:2] += oenv._terrain_offset()
    ref = dict(root_pos=rp, root_rot=nxt[1], root_vel=nxt[2], root_ang_vel=nxt[3], dof_pos=m.rot_to_dof(nxt[4]), dof_vel=nxt[5])
    nz = synth.sim_noise(N, m.J, m.D, seed=seed + 2)
    k = int(N * fail_fraction)
    if k > 0:
        nz["root_pos"][N - k:] += 0.5
        nz["dof_pos"][N - 2 * k:N - k] += 0.9
    sim = synth.perturb_state(ref, nz, m.forward_kinematics, m.dof_to_rot)
    oenv.char_root_pos[:], oenv.char_root_rot[:] = sim["root_pos"], sim["root_rot"]
    oenv.char_root_vel[:], oenv.char_root_ang_vel[:] = sim["root_vel"], sim["root_ang_vel"]
    oenv.char_dof_pos[:], oenv.char_dof_vel[:] = sim["dof_pos"], sim["dof_vel"]
    oenv.char_rigid_body_pos[:], oenv.char_contact_forces[:] = sim["body_pos"], sim["contact_forces"]
    oenv.fail_rates[:] = 0.25 + 0.5 * torch.rand(lib.num_motions(), generator=g)


def assert_close_world(a, b, world_mag, rtol=1e-5, atol=3e-6, strict_fraction=0.999, ulps=2.0, what=""):
    """For values derived from WORLD coordinates (positions, and observations that are differences
    of positions): env offsets and terrain-tile offsets put world coordinates at |x| ~ world_mag, so
    one fp32 rounding step of such a coordinate is ulp(world_mag) in absolute terms -- in the
    reference too. Every element must be within 2 ulp(world_mag) (+ the usual tolerance), and at
    least `strict_fraction` of them within the usual tolerance alone."""
    a = torch.as_tensor(a).detach().cpu().double()
    b = torch.as_tensor(b).detach().cpu().double()
    assert a.shape == b.shape, "%s: shape %s vs %s" % (what, tuple(a.shape), tuple(b.shape))
    ulp = float(np.spacing(np.float32(world_mag)))
    err = (a - b).abs()
    strict = err <= atol + rtol * b.abs()
    frac = float(strict.double().mean())
    assert frac >= strict_fraction, "%s: only %.5f of elements within rtol=%g atol=%g" % (what, frac, rtol, atol)
    loose = err <= atol + ulps * ulp + rtol * b.abs()
    if not bool(loose.all()):
        i = torch.argmax(err)
        raise AssertionError("%s: worst |%g - %g| = %g exceeds %g ulp(%g) = %g (strict fraction %.5f)" %
                             (what, a.flatten()[i], b.flatten()[i], err.flatten()[i], ulps, world_mag, ulps * ulp, frac))
