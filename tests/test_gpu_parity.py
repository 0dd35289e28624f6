"""Parity of the sm_100a kernels (through the C ABI / ctypes) against the CPU oracle and the
reference golden vectors. Tolerances are BASELINE.json's: bit-exact frame indices, done flags,
grid indices and reset selection; 1e-5 relative (1e-6..2e-6 absolute floor) for fp32 values."""
import copy

import numpy as np
import pytest
import torch
import yaml

from oracle import parc_oracle as O
from parc_b200 import synth
from parc_b200._lib import ParcB200Error
from tests import helpers as H
from tests.helpers import assert_close, load_npz, t

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


@pytest.fixture(scope="module")
def model():
    return H.product_model(DEV)


# ------------------------------------------------------------------------------------- KinCharModel
def test_kin_char_model_vs_reference_golden(model):
    g = load_npz("ref_kin.npz")
    cu = lambda k: t(g[k]).to(DEV)
    assert_close(model.dof_to_rot(cu("dof")), g["rot"], atol=1e-6, what="dof_to_rot")
    pos, rot = model.forward_kinematics(cu("root_pos"), cu("root_rot"), cu("rot"))
    assert_close(pos, g["body_pos"], atol=2e-6, what="fk pos")
    assert_close(rot, g["body_rot"], atol=2e-6, what="fk rot")
    assert_close(model.rot_to_dof(cu("rot")), g["dof_back"], atol=2e-6, what="rot_to_dof")
    assert_close(model.rot_to_dof(cu("rnd_rot")), g["dof_rnd"], atol=2e-6, what="rot_to_dof random")


def test_fk_leading_dims_and_empty(model):
    om = H.oracle_model()
    g = torch.Generator().manual_seed(3)
    rp = torch.randn(5, 7, 3, generator=g)
    rr = O.quat_normalize(torch.randn(5, 7, 4, generator=g))
    jr = O.normalize(torch.randn(5, 7, 14, 4, generator=g))
    pos, rot = model.forward_kinematics(rp.to(DEV), rr.to(DEV), jr.to(DEV))
    opos, orot = om.forward_kinematics(rp, rr, jr)
    assert pos.shape == (5, 7, 15, 3) and rot.shape == (5, 7, 15, 4)
    assert_close(pos, opos, atol=2e-6, what="fk pos [B,T]")
    assert_close(rot, orot, atol=2e-6, what="fk rot [B,T]")
    e = torch.zeros(0, 3, device=DEV)
    pos, rot = model.forward_kinematics(e, torch.zeros(0, 4, device=DEV), torch.zeros(0, 14, 4, device=DEV))
    assert pos.shape == (0, 15, 3) and rot.shape == (0, 15, 4)


def test_fk_backward_matches_autograd_of_oracle(model):
    om = H.oracle_model()
    g = torch.Generator().manual_seed(4)
    n = 33
    rp = torch.randn(n, 3, generator=g)
    rr = O.quat_normalize(torch.randn(n, 4, generator=g))
    jr = O.normalize(torch.randn(n, 14, 4, generator=g))
    gp, gr = torch.randn(n, 15, 3, generator=g), torch.randn(n, 15, 4, generator=g)
    a = [x.clone().requires_grad_(True) for x in (rp, rr, jr)]
    p, r = om.forward_kinematics(*a)
    ((p * gp).sum() + (r * gr).sum()).backward()
    b = [x.clone().to(DEV).requires_grad_(True) for x in (rp, rr, jr)]
    p2, r2 = model.forward_kinematics(*b)
    ((p2 * gp.to(DEV)).sum() + (r2 * gr.to(DEV)).sum()).backward()
    for x, y, nm in zip(b, a, ("root_pos", "root_rot", "joint_rot")):
        assert_close(x.grad, y.grad, rtol=1e-4, atol=1e-5, what="d/d" + nm)


def test_cpu_tensor_is_refused(model):
    with pytest.raises(ParcB200Error):
        model.dof_to_rot(torch.zeros(4, 28))


# ------------------------------------------------------------------------------------- MotionLib
def test_motion_lib_vs_reference_golden(model):
    g = load_npz("ref_motionlib.npz")
    clip = H.golden_clip()
    lib = H.product_library([clip], model, DEV)
    assert_close(lib._frame_root_vel, g["frame_root_vel"], what="root_vel")
    assert_close(lib._frame_root_ang_vel, g["frame_root_ang_vel"], atol=2e-6, what="root_ang_vel")
    assert_close(lib._frame_dof_vel, g["frame_dof_vel"], atol=2e-5, what="dof_vel")
    assert float(lib._motion_lengths[0]) == float(g["motion_length"][0])
    names = ["root_pos", "root_rot", "root_vel", "root_ang_vel", "joint_rot", "dof_vel", "contacts"]
    for prefix, loop in (("", 0), ("w_", 1)):
        lib._motion_loop_modes[:] = loop
        lib.repack()
        times = t(g["times_w" if prefix else "times"]).to(DEV)
        ids = torch.zeros(times.shape[0], dtype=torch.int64, device=DEV)
        i0, i1, bl = lib._calc_frame_blend(ids, times)
        assert np.array_equal(i0.cpu().numpy(), g[prefix + "idx0"]), "frame_idx0 must be bit-exact"
        assert np.array_equal(i1.cpu().numpy(), g[prefix + "idx1"]), "frame_idx1 must be bit-exact"
        assert np.array_equal(bl.cpu().numpy(), g[prefix + "blend"]), "blend must be bit-exact"
        out = lib.calc_motion_frame(ids, times)
        assert len(out) == 7
        for k, v in zip(names, out):
            assert_close(v, g[prefix + k], atol=2e-6, what=prefix + k)


def test_motion_lib_synthetic_multi_clip_vs_oracle(model):
    clips = synth.make_clips(synth.humanoid_spec(), 37, seed=5, wrap_fraction=0.3, min_frames=2, max_frames=90)
    lib = H.product_library(clips, model, DEV)
    om = H.oracle_model()
    olib = O.Library(om, [H.oracle_clip(c) for c in clips])
    assert np.array_equal(lib._motion_start_idx.cpu().numpy(), olib.start_idx.numpy())
    assert_close(lib._frame_dof_vel, olib.frame_dof_vel, atol=5e-5, what="dof_vel")
    g = torch.Generator().manual_seed(6)
    n = 20000
    ids = torch.randint(0, 37, (n,), generator=g)
    times = (torch.rand(n, generator=g) * 1.5 - 0.2) * olib.lengths[ids]
    # exact frame times: dt * k in fp32, the truncation boundary of frame_idx0
    k = torch.randint(0, 60, (n // 2,), generator=g, dtype=torch.int32)
    times[: n // 2] = torch.tensor(1.0 / 30.0, dtype=torch.float32) * k
    i0, i1, bl = lib._calc_frame_blend(ids.to(DEV), times.to(DEV))
    oi0, oi1, obl = olib.frame_blend(ids, times)
    assert torch.equal(i0.cpu(), oi0) and torch.equal(i1.cpu(), oi1), "frame indices must be bit-exact"
    assert torch.equal(bl.cpu(), obl)
    out = lib.calc_motion_frame(ids.to(DEV), times.to(DEV))
    oout = olib.calc_motion_frame(ids, times)
    for a, b, nm in zip(out, oout, ["root_pos", "root_rot", "root_vel", "root_ang_vel", "joint_rot", "dof_vel", "contacts"]):
        assert_close(a, b, atol=2e-6, what=nm)
    # empty and shaped queries
    e = lib.calc_motion_frame(torch.zeros(0, dtype=torch.int64, device=DEV), torch.zeros(0, device=DEV))
    assert e[0].shape == (0, 3) and e[4].shape == (0, 14, 4)
    sh = lib.calc_motion_frame(ids[:12].reshape(3, 4).to(DEV), times[:12].reshape(3, 4).to(DEV))
    assert sh[4].shape == (3, 4, 14, 4)


# ------------------------------------------------------------------------------------- terrain
def test_ray_heights_and_grid_index_bit_exact(model):
    g = torch.Generator().manual_seed(8)
    hf = torch.randn(40, 57, generator=g)
    terr = H.product_terrain(hf, torch.tensor([-3.2, 1.2]), 0.4, DEV)
    oterr = O.Terrain(hf, [-3.2, 1.2], [0.4, 0.4])
    pts = torch.rand(5000, 2, generator=g) * 30 - 8
    # points exactly on cell boundaries: (k + 0.5) * dx from the origin -> round-half-even decides
    kk = torch.arange(0, 30, dtype=torch.float32)
    pts[:30, 0] = torch.tensor(-3.2) + (kk + 0.5) * torch.tensor(0.4)
    idx = terr.get_grid_index(pts.to(DEV))
    assert torch.equal(idx.cpu(), oterr.grid_index(pts)), "grid indices must be bit-exact"
    assert torch.equal(terr.get_hf_val_from_points(pts.to(DEV)).cpu(), oterr.height_at(pts))


def test_cell_index_division_is_correctly_rounded():
    """The fused step's divide-by-cell-size (reciprocal + one fma correction) must equal div.rn bit for bit."""
    from parc_b200 import _lib
    g = torch.Generator().manual_seed(10)
    n = 1 << 22
    for d in (0.4, 0.1, 0.25, 1.0, 0.3):
        x = (torch.rand(n, generator=g) * 900.0 - 450.0)
        k = torch.arange(0, 4096, dtype=torch.float32)
        x[:4096] = (k + 0.5) * torch.tensor(d, dtype=torch.float32)          # exact cell boundaries
        x[4096:8192] = (k - 2048.0) * torch.tensor(d, dtype=torch.float32)
        x = x.to(DEV)
        qf, qi = torch.empty_like(x), torch.empty_like(x)
        _lib.check(_lib.lib().parc_b200_selftest_div(_lib.ptr(x), n, d, _lib.ptr(qf), _lib.ptr(qi), _lib.stream_ptr(torch.device(DEV))))
        assert torch.equal(qf, qi), "d=%g: %d quotients differ from div.rn" % (d, int((qf != qi).sum()))
        assert torch.equal(qi.cpu(), x.cpu() / torch.tensor(d, dtype=torch.float32))


def test_packed_cell_arithmetic_equals_the_separately_rounded_one():
    """The fused step computes the grid cell of a ray sample point with packed fp32x2 instructions; every product and sum
    must still be rounded on its own (torch's mul / sub / add), or points next to a cell boundary land in the wrong
    cell. ptxas contracts mul.rn.f32x2 + add.rn.f32x2 into FFMA2, which this test guards against; 16 M points per
    setting, a third of them placed within a few ulps of a cell boundary."""
    from parc_b200 import _lib
    g = torch.Generator().manual_seed(12)
    n = 1 << 24
    dev = torch.device(DEV)
    for dx, dim, heading in ((0.4, 1088, 0.7), (0.4, 1088, -2.4), (0.25, 64, 1.3), (0.1, 500, 3.0)):
        c, s = float(torch.cos(torch.tensor(heading))), float(torch.sin(torch.tensor(heading)))
        mn = -0.5 * dx * dim
        gx, gy = float(torch.rand(1, generator=g)) * 100.0 - 50.0, float(torch.rand(1, generator=g)) * 100.0 - 50.0
        pts = torch.rand(n, 2, generator=g) * 8.0 - 4.0
        # world points on cell boundaries, pulled back into the local frame (rounding puts them a few ulps off either way)
        k = n // 3
        bx = (torch.randint(0, dim, (k,), generator=g).float() + 0.5) * dx + mn
        by = (torch.randint(0, dim, (k,), generator=g).float() + 0.5) * dx + mn
        lx, ly = bx - gx, by - gy
        pts[:k, 0] = lx * c + ly * s
        pts[:k, 1] = -lx * s + ly * c
        pts = pts.to(dev).contiguous()
        a = torch.empty(n, dtype=torch.int32, device=dev)
        b = torch.empty(n, dtype=torch.int32, device=dev)
        _lib.check(_lib.lib().parc_b200_selftest_cells(_lib.ptr(pts), n, c, s, gx, gy, mn, mn, dx, dx, dim, dim, _lib.ptr(a), _lib.ptr(b),
                                                       _lib.stream_ptr(dev)))
        assert torch.equal(a, b), "dx=%g: %d of %d cells differ" % (dx, int((a != b).sum()), n)


# ------------------------------------------------------------------------------------- fused env step
def _golden_env(model):
    from parc_b200.tracking_env import TrackingEnv, default_env_config
    g = load_npz("ref_env_step.npz")
    clip = H.golden_clip()
    lib = H.product_library([clip], model, DEV)
    ec = default_env_config()
    ec.update(yaml.safe_load(str(g["cfg_yaml"]))["env"])
    terrain = H.product_terrain(clip["hf"], clip["min_point"], clip["dx"], DEV)
    N = g["in_timestep_buf"].shape[0]
    env = TrackingEnv(ec, N, DEV, model, lib, terrain, t(g["dm_motion_offsets"]), t(g["ray_xy"]), env_offsets=t(g["env_offsets"]),
                      keep_ray_hfs=True)
    return env, g


def test_fused_step_vs_reference_golden(model):
    env, g = _golden_env(model)
    cu = lambda k: t(g[k]).to(DEV)
    env._timestep_buf.copy_(cu("in_timestep_buf"))
    env._motion_ids.copy_(cu("in_motion_ids"))
    env._motion_terrain_ids.copy_(cu("in_motion_terrain_ids"))
    env._motion_time_offsets.copy_(cu("in_motion_time_offsets"))
    env._motion_id_fail_rates.copy_(cu("in_fail_rates"))
    for k in ("root_pos", "root_rot", "root_vel", "root_ang_vel", "dof_pos", "dof_vel", "contact_forces", "rigid_body_pos"):
        getattr(env, "_char_" + k).copy_(cu("in_char_" + k))
    env.post_physics_step()
    torch.cuda.synchronize()
    assert env.get_obs_dim() == 1312
    assert np.array_equal(env._done_buf.cpu().numpy(), g["done"]), "done flags must be bit-exact"
    assert np.array_equal(env._time_buf.cpu().numpy(), g["time_buf"])
    assert_close(env._ray_hfs, g["ray_hfs"], what="ray_hfs")
    W = float(np.abs(g["out_ref_root_pos"]).max())
    for k in ("root_pos", "root_rot", "root_vel", "root_ang_vel", "body_pos", "joint_rot", "dof_pos", "dof_vel", "contacts"):
        H.assert_close_world(getattr(env, "_ref_" + k), g["out_ref_" + k], W, what="ref_" + k)
    H.assert_close_world(env._obs_buf, g["obs"], W, what="obs")
    assert_close(env._reward_terms[:, 0:5], g["reward5"], atol=2e-6, what="reward5")
    assert_close(env._reward_terms[:, 5], g["contact_penalty"], atol=2e-6, what="contact penalty")
    assert_close(env._reward_buf, g["reward"], atol=2e-6, what="reward")
    assert_close(env._motion_id_fail_rates, g["out_fail_rates"], rtol=1e-6, atol=0, what="fail-rate EMA")


def test_reset_vs_reference_golden(model):
    env, g = _golden_env(model)
    N = env.get_num_envs()
    env._char_rigid_body_vel.fill_(1.0)
    ids = torch.arange(N, device=DEV)
    env.reset(ids, draws=(t(g["reset_motion_ids"]), t(g["reset_terrain_ids"]), t(g["reset_times"]), t(g["reset_noise"])))
    torch.cuda.synchronize()
    assert np.array_equal(env._motion_ids.cpu().numpy(), g["reset_motion_ids"])
    assert np.array_equal(env._motion_time_offsets.cpu().numpy(), g["reset_times"])
    for k in ("ref_root_pos", "ref_root_rot", "ref_root_vel", "ref_root_ang_vel", "ref_joint_rot", "ref_dof_pos", "ref_dof_vel",
              "ref_contacts", "char_root_pos", "char_root_rot", "char_root_vel", "char_root_ang_vel", "char_dof_pos",
              "char_dof_vel", "char_rigid_body_vel"):
        H.assert_close_world(getattr(env, "_" + k), g["reset_" + k], float(np.abs(g["reset_ref_root_pos"]).max()), what=k)
    assert bool(env._actors_need_reset[:, 0].all())
    assert int(env._timestep_buf.abs().sum()) == 0 and int(env._done_buf.abs().sum()) == 0


def _compare_step(oenv, penv, what=""):
    assert torch.equal(penv._done_buf.cpu(), oenv.done_buf), what + ": done flags must be bit-exact"
    assert torch.equal(penv._time_buf.cpu(), oenv.time_buf), what + ": time"
    W = float(oenv.ref_root_pos.abs().max())
    for k in ("root_pos", "body_pos"):
        H.assert_close_world(getattr(penv, "_ref_" + k), getattr(oenv, "ref_" + k), W, what=what + " ref_" + k)
    for k in ("root_rot", "root_vel", "root_ang_vel", "joint_rot", "dof_pos", "dof_vel", "contacts"):
        assert_close(getattr(penv, "_ref_" + k), getattr(oenv, "ref_" + k), atol=3e-6, what=what + " ref_" + k)
    H.assert_close_world(penv._obs_buf, oenv.obs_buf, W, what=what + " obs")
    # the root / key-position reward terms are exp(-k * |difference of world positions|^2): a 1-ulp(W) change of a
    # world coordinate moves them by up to ~10 ulp(W); everything else is held to the plain tolerance
    assert_close(penv._reward_terms[:, 0:2], oenv.reward_terms[:, 0:2], atol=2e-6, what=what + " pose/vel reward")
    assert_close(penv._reward_terms[:, 3], oenv.reward_terms[:, 3], atol=2e-6, what=what + " root vel reward")
    H.assert_close_world(penv._reward_terms[:, 0:7], oenv.reward_terms, W, strict_fraction=0.99, ulps=16.0, atol=1e-5, what=what + " reward terms")
    H.assert_close_world(penv._reward_buf, oenv.reward_buf, W, strict_fraction=0.99, ulps=16.0, atol=1e-5, what=what + " reward")
    assert_close(penv._motion_id_fail_rates, oenv.fail_rates, rtol=2e-6, atol=0, what=what + " fail rates")


VARIANTS = {
    "default_grid32": dict(),
    "global_obs_root_h": dict(global_obs=True, global_root_height_obs=True),
    "no_track_root": dict(track_root=False, track_root_h=False),
    "fall_test": dict(contact_bodies=["right_foot", "left_foot"]),
    "no_contacts_no_tar": dict(use_contact_info=False, enable_tar_obs=False),
    "never_done": dict(never_done=True),
    "no_early_termination": dict(enable_early_termination=False),
}


@pytest.mark.parametrize("variant", sorted(VARIANTS))
def test_fused_step_vs_oracle_synthetic(variant):
    from parc_b200.tracking_env import default_env_config
    ec = default_env_config()
    ec.update(VARIANTS[variant])
    sc = H.Scenario(num_envs=1531, num_motions=23, ray="grid" if variant != "fall_test" else "cone", seed=77,
                    wrap_fraction=0.2, env_config=ec, max_frames=120)
    oenv = sc.oracle_env()
    H.randomize_oracle_state(oenv, seed=31)
    penv = sc.product_env(DEV)
    H.copy_oracle_state_to_product(oenv, penv)
    for it in range(2):
        oenv.post_physics_step()
        penv.post_physics_step()
        torch.cuda.synchronize()
        _compare_step(oenv, penv, what="%s step %d" % (variant, it))


def _compare_step_strict(oenv, penv, what=""):
    """north_star tolerances with no allowance: flags and time bit-exact, EVERY fp32 element inside rtol 1e-5 / atol 1e-6."""
    assert torch.equal(penv._done_buf.cpu(), oenv.done_buf), what + ": done flags must be bit-exact"
    assert torch.equal(penv._time_buf.cpu(), oenv.time_buf), what + ": time"
    for k in ("root_pos", "body_pos", "root_rot", "root_vel", "root_ang_vel", "joint_rot", "dof_pos", "dof_vel", "contacts"):
        assert_close(getattr(penv, "_ref_" + k), getattr(oenv, "ref_" + k), rtol=1e-5, atol=1e-6, what=what + " ref_" + k)
    assert_close(penv._obs_buf, oenv.obs_buf, rtol=1e-5, atol=1e-6, what=what + " obs")
    assert_close(penv._reward_terms[:, 0:7], oenv.reward_terms, rtol=1e-5, atol=1e-6, what=what + " reward terms")
    assert_close(penv._reward_buf, oenv.reward_buf, rtol=1e-5, atol=1e-6, what=what + " reward")
    assert_close(penv._motion_id_fail_rates, oenv.fail_rates, rtol=2e-6, atol=0, what=what + " fail rates")


@pytest.mark.parametrize("variant", sorted(VARIANTS))
def test_fused_step_strict_tolerance_near_the_origin(variant):
    """The seven configurations of test_fused_step_vs_oracle_synthetic with env_offsets = 0 and the tile grid centred on the
    origin (one clip: world coordinates stay within ~ +-7 m, one fp32 ulp <= 4.8e-7): without the 200 m world offsets the plain north_star
    tolerance holds for 100 % of the elements, reward included -- the evidence behind assert_close_world's
    "ulp of the world magnitude" allowance in the large-world tests."""
    from parc_b200.tracking_env import default_env_config
    ec = default_env_config()
    ec.update(VARIANTS[variant])
    sc = H.Scenario(num_envs=1531, num_motions=1, ray="grid" if variant != "fall_test" else "cone", seed=78,
                    wrap_fraction=0.0, env_config=ec, max_frames=120)
    sc.env_offsets = torch.zeros_like(sc.env_offsets)
    assert float(sc.min_point.abs().max()) < 8.0
    oenv = sc.oracle_env()
    H.randomize_oracle_state(oenv, seed=31)
    penv = sc.product_env(DEV)
    H.copy_oracle_state_to_product(oenv, penv)
    for it in range(2):
        oenv.post_physics_step()
        penv.post_physics_step()
        torch.cuda.synchronize()
        _compare_step_strict(oenv, penv, what="%s strict step %d" % (variant, it))


def test_reset_subset_and_obs_refresh_vs_oracle():
    sc = H.Scenario(num_envs=700, num_motions=11, ray="cone", seed=9, wrap_fraction=0.2, max_frames=100)
    oenv = sc.oracle_env()
    H.randomize_oracle_state(oenv, seed=12)
    penv = sc.product_env(DEV, keep_ray_hfs=True)
    H.copy_oracle_state_to_product(oenv, penv)
    oenv.post_physics_step()
    penv.post_physics_step()
    ids = torch.nonzero(oenv.done_buf != 0).flatten()
    assert 0 < ids.numel() < sc.N
    draws = synth.reset_draws(ids.numel(), sc.M, 1, oenv.lib.lengths, oenv.sampling_weights(), seed=5)
    oenv.cfg.rand_root_pos_offset_scale = 0.075
    penv.set_rand_root_pos_offset_scale(0.075)
    oenv.reset(ids, *draws)
    oenv.refresh_obs_after_reset(ids)
    penv.reset(ids.to(DEV), draws=draws)
    torch.cuda.synchronize()
    assert torch.equal(penv._motion_ids.cpu(), oenv.motion_ids), "reset selection must be bit-exact"
    assert torch.equal(penv._timestep_buf.cpu(), oenv.timestep_buf)
    assert torch.equal(penv._done_buf.cpu(), oenv.done_buf)
    W = float(oenv.ref_root_pos.abs().max())
    for k in ("char_root_pos", "char_root_rot", "char_root_vel", "char_root_ang_vel", "char_dof_pos", "char_dof_vel",
              "ref_root_pos", "ref_joint_rot", "ref_dof_pos", "ref_dof_vel", "ref_contacts"):
        H.assert_close_world(getattr(penv, "_" + k), getattr(oenv, k), W, what=k)
    assert_close(penv._ray_hfs, oenv.ray_hfs, what="ray_hfs after reset")
    H.assert_close_world(penv._obs_buf, oenv.obs_buf, W, what="obs after reset")
    # empty reset is a no-op
    penv.reset(torch.zeros(0, dtype=torch.int64, device=DEV))


def test_reset_done_honours_offsets_start_fractions_demo_and_ignore_fail_rates():
    """TrackingEnv.reset_done (device-side draws) against TrackingEnv.reset(draws=...) -- the reference-order path that is
    itself pinned to DeepMimicEnv.reset -- with every non-random knob of the reference's reset set: root position / rotation
    offsets (mgdm_dm_util.py:108-115), rand_reset False with per-env start fractions (dm_env.py:500-502), then demo mode
    (:479-480) and dm.ignore_fail_rates (:485-486). Also: the sampling CDF is valid before the first EMA."""
    def twins(**ecu):
        sc = H.Scenario(num_envs=384, num_motions=11, ray="grid", seed=29, wrap_fraction=0.1, max_frames=80)
        sc.ec.update(ecu)
        oenv = sc.oracle_env()
        H.randomize_oracle_state(oenv, seed=4, fail_fraction=0.25)
        out = []
        for _ in range(2):
            pe = sc.product_env(DEV)
            H.copy_oracle_state_to_product(oenv, pe)
            out.append(pe)
        return sc, oenv, out[0], out[1]

    g = torch.Generator().manual_seed(8)
    # --- offsets + start-time fractions, rand_reset off ---
    sc, oenv, a, b = twins(rand_reset=False)
    N = sc.N
    assert float(a._cdf[-1]) > 0.99 and torch.all(a._cdf[1:] >= a._cdf[:-1])        # valid before any EMA has run
    pos_off = (torch.rand(N, 3, generator=g) - 0.5).to(DEV)
    rot_off = torch.nn.functional.normalize(torch.randn(N, 4, generator=g), dim=-1).to(DEV)
    frac = torch.rand(N, generator=g).to(DEV)
    for e in (a, b):
        e.set_root_pos_offset(pos_off); e.set_root_rot_offset(rot_off); e.set_motion_start_time_fraction(frac)
        e.post_physics_step(update_fail_rates=False)          # the CDF must still be usable without the EMA
    torch.cuda.synchronize()
    n = int(a._done_count.item())
    done_ids = torch.sort(a._done_list[:n].long()).values
    assert n > 20
    step = a._step_count
    a.reset_done()
    cdf = a._cdf.cpu().numpy()
    lengths = oenv.lib.lengths.numpy()
    d = [O.rsi_draws(int(e), a._seed, step, cdf, 1, lengths, False) for e in done_ids.tolist()]
    mids = torch.tensor([x[0] for x in d], dtype=torch.int64)
    times = b._mlib.get_motion_length(mids.to(DEV)) * frac[done_ids]
    noise = torch.tensor(np.array([[x[3], x[4]] for x in d], dtype=np.float32))
    assert torch.equal(a._motion_ids[done_ids].cpu(), mids)
    assert torch.equal(a._motion_time_offsets[done_ids], times), "start time = length * fraction, rounded as the reference does"
    b.reset(done_ids, draws=(mids, torch.tensor([x[1] for x in d], dtype=torch.int64), times, noise))
    torch.cuda.synchronize()
    for k in ("_root_state", "_dof_state", "_ref_root_pos", "_ref_root_rot", "_ref_joint_rot", "_ref_dof_pos", "_ref_dof_vel", "_ref_contacts",
              "_obs_buf", "_motion_ids", "_motion_terrain_ids", "_motion_time_offsets", "_timestep_buf", "_done_buf"):
        assert torch.equal(getattr(a, k), getattr(b, k)), k
    assert float((a._char_root_pos[done_ids] - a._ref_root_pos[done_ids]).abs().max()) > 0.05      # the offsets were applied
    # --- demo mode: motion id = env id % M ---
    sc, oenv, a, _ = twins()
    a.set_demo_mode(True)
    a.post_physics_step()
    n = int(a._done_count.item())
    done_ids = a._done_list[:n].long()
    a.reset_done()
    assert torch.equal(a._motion_ids[done_ids], done_ids % sc.M)
    # --- ignore_fail_rates: motion weights alone, whatever the fail rates are ---
    sc, oenv, a, _ = twins(dm={"min_motion_weight": 0.01, "terrains_per_motion": 1, "ignore_fail_rates": True})
    a._motion_id_fail_rates.copy_(torch.linspace(0.0, 1.0, sc.M))
    a.post_physics_step()
    n = int(a._done_count.item())
    done_ids = torch.sort(a._done_list[:n].long()).values
    step = a._step_count
    a.reset_done()
    cdf_w = np.cumsum(a._mlib._motion_weights.cpu().numpy(), dtype=np.float32)
    assert np.array_equal(a._cdf_weights_only.cpu().numpy(), torch.cumsum(a._mlib._motion_weights, 0).cpu().numpy())
    d = [O.rsi_draws(int(e), a._seed, step, a._cdf_weights_only.cpu().numpy(), 1, oenv.lib.lengths.numpy(), True) for e in done_ids.tolist()]
    assert torch.equal(a._motion_ids[done_ids].cpu(), torch.tensor([x[0] for x in d], dtype=torch.int64))
    del cdf_w


@pytest.mark.parametrize("M,n_done", [(1, 4096), (3, 6000), (7, 300), (1000, 28000), (40, 40)])
def test_fail_rate_ema_is_sequential_in_env_order_for_any_segment_size(M, n_done):
    """parc_b200_fail_rate_ema against the reference's host loop (dm_env.py:648-660: finished envs in ascending env id,
    r <- r * (1 - w) [+ w on FAIL] in fp32) on forced done lists: one clip with 4,096 finishes (every entry in one segment),
    a few clips with thousands each, the BASELINE configs[4] shape (28 k finishes over 1,000 clips), and tiny cases."""
    from parc_b200 import _lib
    sc = H.Scenario(num_envs=max(n_done + 100, 512), num_motions=M, ray="cone", seed=31, wrap_fraction=0.0, max_frames=70)
    penv = sc.product_env(DEV)
    N = sc.N
    g = torch.Generator().manual_seed(M * 7 + n_done)
    mids = torch.randint(0, M, (N,), generator=g)
    done_envs = torch.randperm(N, generator=g)[:n_done]                 # unordered, as the step kernel appends them
    code = torch.zeros(N, dtype=torch.uint8)
    code[done_envs] = torch.where(torch.rand(n_done, generator=g) < 0.4, torch.tensor(1, dtype=torch.uint8), torch.tensor(2, dtype=torch.uint8))
    rates0 = torch.rand(M, generator=g)
    penv._motion_ids.copy_(mids.to(DEV))
    penv._done_code.copy_(code.to(DEV))
    penv._done_list[:n_done].copy_(done_envs.to(torch.int32).to(DEV))
    penv._done_count.fill_(n_done)
    penv._motion_id_fail_rates.copy_(rates0.to(DEV))
    _lib.check(_lib.lib().parc_b200_fail_rate_ema(penv._cfg, penv._bufs, M, _lib.ptr(penv._cdf), _lib.ptr(penv._ema_scratch), penv._stream()))
    torch.cuda.synchronize()
    want = rates0.numpy().astype(np.float32).copy()
    keep, w = np.float32(1.0 - 0.01), np.float32(0.01)
    order = torch.sort(done_envs).values.tolist()
    mi, co = mids.numpy(), code.numpy()
    for e in order:
        r = np.float32(want[mi[e]] * keep)
        want[mi[e]] = np.float32(r + w) if co[e] == 1 else r
    assert np.array_equal(penv._motion_id_fail_rates.cpu().numpy(), want), "the EMA must equal the sequential fp32 recurrence bit for bit"
    cdf = np.cumsum(np.maximum(want, np.float32(0.01)) * penv._mlib._motion_weights.cpu().numpy(), dtype=np.float32)
    assert_close(penv._cdf, cdf, rtol=1e-5, atol=1e-7, what="RSI CDF")


def test_device_sampled_reset_matches_restated_draws():
    """parc_b200_env_reset_sampled: Philox + inverse-CDF draws are bit-exact against the CPU
    restatement (oracle.rsi_draws), and the resulting state equals an explicit reset with them."""
    sc = H.Scenario(num_envs=512, num_motions=17, ray="cone", seed=19, wrap_fraction=0.0, max_frames=90)
    oenv = sc.oracle_env()
    H.randomize_oracle_state(oenv, seed=2, fail_fraction=0.2)
    penv = sc.product_env(DEV)
    H.copy_oracle_state_to_product(oenv, penv)
    oenv.post_physics_step()
    penv.post_physics_step()
    torch.cuda.synchronize()
    cdf = penv._cdf.cpu().numpy()
    want = np.cumsum(oenv.sampling_weights().numpy().astype(np.float32), dtype=np.float32)
    assert_close(cdf, want, rtol=1e-5, atol=1e-7, what="RSI CDF")
    done_ids = torch.nonzero(oenv.done_buf != 0).flatten()
    step = penv._step_count
    penv.reset_done()
    torch.cuda.synchronize()
    lengths = oenv.lib.lengths.numpy()
    d = [O.rsi_draws(int(e), penv._seed, step, cdf, 1, lengths, True) for e in done_ids.tolist()]
    mids = torch.tensor([x[0] for x in d], dtype=torch.int64)
    assert torch.equal(penv._motion_ids.cpu()[done_ids], mids), "device RSI motion selection must be bit-exact"
    times = torch.tensor(np.array([x[2] for x in d], dtype=np.float32))
    assert torch.equal(penv._motion_time_offsets.cpu()[done_ids], times)
    oenv.cfg.rand_root_pos_offset_scale = sc.ec["rand_root_pos_offset_scale"]
    noise = torch.tensor(np.array([[x[3], x[4]] for x in d], dtype=np.float32))
    oenv.reset(done_ids, mids, torch.tensor([x[1] for x in d], dtype=torch.int64), times, noise)
    oenv.refresh_obs_after_reset(done_ids)
    W = float(oenv.ref_root_pos.abs().max())
    for k in ("char_root_pos", "char_root_rot", "char_dof_pos", "char_dof_vel", "ref_root_pos", "ref_dof_pos"):
        H.assert_close_world(getattr(penv, "_" + k), getattr(oenv, k), W, what=k)
    H.assert_close_world(penv._obs_buf[done_ids.to(DEV)], oenv.obs_buf[done_ids], W, what="obs of reset envs")
    assert int(penv._done_count.item()) == 0
    # chi^2 of the inverse-CDF sampler against the weights
    u = torch.rand(200000, device=DEV)
    ids = torch.empty(200000, dtype=torch.int64, device=DEV)
    from parc_b200 import _lib
    _lib.check(_lib.lib().parc_b200_sample_motions_cdf(_lib.ptr(penv._cdf), sc.M, _lib.ptr(u), 200000, _lib.ptr(ids),
                                                      _lib.stream_ptr(torch.device(DEV))))
    counts = torch.bincount(ids.cpu(), minlength=sc.M).double()
    p = oenv.sampling_weights().double()
    p = p / p.sum()
    chi2 = float((((counts - 200000 * p) ** 2) / (200000 * p)).sum())
    assert chi2 < 60.0, "chi^2 = %.1f for %d bins" % (chi2, sc.M)


# ------------------------------------------------------------------------------------- full size
def test_full_size_properties_65536_envs():
    """BASELINE config 2 shape (65,536 envs, 1,000 clips, 32x32 heights): size-independent checks --
    shard invariance (a rank owning envs [a,b) computes exactly the rows of the full run),
    determinism, flag domain, and spot parity of 256 random envs against the oracle."""
    N, M = 65536, 1000
    sc = H.Scenario(num_envs=N, num_motions=M, ray="grid", seed=1234, wrap_fraction=0.05)
    penv = sc.product_env(DEV)
    lib = penv._mlib
    draws = synth.reset_draws(N, M, 1, lib._motion_lengths, lib._motion_weights, seed=4321)
    penv.set_rand_root_pos_offset_scale(0.0)
    penv.reset(torch.arange(N, device=DEV), draws=draws, refresh_obs=False)
    g = torch.Generator().manual_seed(1)
    penv._timestep_buf.copy_(torch.randint(0, 100, (N,), generator=g, dtype=torch.int32).to(DEV))
    penv._time_buf.copy_(penv._timestep_buf.float() * np.float32(1.0 / 30.0))
    nz = synth.sim_noise(N, 15, 28, seed=5678)
    ref = dict(root_pos=penv._ref_root_pos, root_rot=penv._ref_root_rot, root_vel=penv._ref_root_vel,
               root_ang_vel=penv._ref_root_ang_vel, dof_pos=penv._ref_dof_pos, dof_vel=penv._ref_dof_vel)
    sim = synth.perturb_state(ref, nz, penv._kin_char_model.forward_kinematics, penv._kin_char_model.dof_to_rot)
    for k, dst in (("root_pos", "_char_root_pos"), ("root_rot", "_char_root_rot"), ("root_vel", "_char_root_vel"),
                   ("root_ang_vel", "_char_root_ang_vel"), ("dof_pos", "_char_dof_pos"), ("dof_vel", "_char_dof_vel"),
                   ("body_pos", "_char_rigid_body_pos"), ("contact_forces", "_char_contact_forces")):
        getattr(penv, dst).copy_(sim[k])
    snap = {k: getattr(penv, k).clone() for k in ("_root_state", "_dof_state", "_char_rigid_body_pos", "_char_contact_forces",
                                                  "_timestep_buf", "_time_buf", "_motion_ids", "_motion_terrain_ids",
                                                  "_motion_time_offsets")}
    penv.post_physics_step()
    torch.cuda.synchronize()
    obs, rew, done = penv._obs_buf.clone(), penv._reward_buf.clone(), penv._done_buf.clone()
    assert obs.shape == (N, 1895)
    assert bool(torch.isfinite(obs).all()) and bool(torch.isfinite(rew).all())
    assert set(torch.unique(done).tolist()) <= {0, 1, 3}
    assert float(obs[:, 871:].abs().max()) <= 3.0
    # determinism: same inputs -> bit-identical outputs
    for k, v in snap.items():
        getattr(penv, k).copy_(v)
    penv._motion_id_fail_rates.fill_(1.0)
    penv.post_physics_step()
    torch.cuda.synchronize()
    assert torch.equal(penv._obs_buf, obs) and torch.equal(penv._reward_buf, rew) and torch.equal(penv._done_buf, done)
    # spot parity against the oracle on 256 random envs (the oracle finishes that in seconds)
    sel = torch.randperm(N, generator=g)[:256].sort().values
    sub = H.Scenario.__new__(H.Scenario)
    sub.__dict__.update(sc.__dict__)
    sub.N, sub.env_offsets = 256, sc.env_offsets[sel]
    oenv = sub.oracle_env()
    cpu = lambda x: x.cpu()
    oenv.motion_ids[:], oenv.motion_terrain_ids[:] = cpu(snap["_motion_ids"])[sel], cpu(snap["_motion_terrain_ids"])[sel]
    oenv.motion_time_offsets[:] = cpu(snap["_motion_time_offsets"])[sel]
    oenv.timestep_buf[:], oenv.time_buf[:] = cpu(snap["_timestep_buf"])[sel], cpu(snap["_time_buf"])[sel]
    rs, ds = cpu(snap["_root_state"])[sel], cpu(snap["_dof_state"])[sel]
    oenv.char_root_pos[:], oenv.char_root_rot[:], oenv.char_root_vel[:], oenv.char_root_ang_vel[:] = rs[:, 0:3], rs[:, 3:7], rs[:, 7:10], rs[:, 10:13]
    oenv.char_dof_pos[:], oenv.char_dof_vel[:] = ds[:, :, 0], ds[:, :, 1]
    oenv.char_rigid_body_pos[:] = cpu(snap["_char_rigid_body_pos"])[sel]
    oenv.char_contact_forces[:] = cpu(snap["_char_contact_forces"])[sel]
    oenv.post_physics_step()
    assert torch.equal(done.cpu()[sel], oenv.done_buf), "done flags must be bit-exact"
    H.assert_close_world(obs.cpu()[sel], oenv.obs_buf, float(oenv.ref_root_pos.abs().max()), what="obs (65536-env run, 256 sampled rows)")
    assert_close(rew.cpu()[sel], oenv.reward_buf, atol=1e-5, what="reward")
    # shard invariance: a rank that owns global envs [a, b) -- not aligned to the kernel's 8-env tiles -- computes
    # bit-identical rows (bench.py --gpus N relies on this; there is no exchange on the data path)
    a, b = 16384 + 3, 16384 + 3 + 4099
    shard = H.Scenario.__new__(H.Scenario)
    shard.__dict__.update(sc.__dict__)
    shard.N, shard.env_offsets = b - a, sc.env_offsets[a:b]
    senv = shard.product_env(DEV)
    for k, v in snap.items():
        getattr(senv, k).copy_(v[a:b])
    senv._motion_id_fail_rates.fill_(1.0)
    senv.post_physics_step()
    torch.cuda.synchronize()
    assert torch.equal(senv._obs_buf, obs[a:b]) and torch.equal(senv._reward_buf, rew[a:b]) and torch.equal(senv._done_buf, done[a:b])


def test_simulator_standin_matches_torch_formulas():
    """parc_b200_sim_standin (bench utility) == synth.apply_standin_torch on the oracle env."""
    sc = H.Scenario(num_envs=333, num_motions=5, ray="cone", seed=4, max_frames=80)
    oenv = sc.oracle_env()
    H.randomize_oracle_state(oenv, seed=6)
    penv = sc.product_env(DEV)
    H.copy_oracle_state_to_product(oenv, penv)
    noise = synth.standin_noise(sc.N, 15, 28, seed=9)
    synth.apply_standin_torch(oenv, noise, 15, 28)
    penv.sim_standin(noise.to(DEV))
    torch.cuda.synchronize()
    W = float(oenv.ref_root_pos.abs().max())
    for k in ("root_pos", "rigid_body_pos"):
        H.assert_close_world(getattr(penv, "_char_" + k).cpu(), getattr(oenv, "char_" + k), W, what="standin " + k)
    for k in ("root_rot", "root_vel", "root_ang_vel", "dof_pos", "dof_vel"):
        assert_close(getattr(penv, "_char_" + k).cpu(), getattr(oenv, "char_" + k), what="standin " + k)


@pytest.mark.parametrize("num_chunks,use_standin", [(1, False), (5, False), (8, True), (64, False)])
def test_host_fed_step_equals_upload_then_step(num_chunks, use_standin):
    """parc_b200_env_step_host (chunked upload overlapped with the step) == upload_sim_state + post_physics_step,
    bit for bit, including the done list, for an env count that is not a multiple of the chunk / tile size."""
    sc = H.Scenario(num_envs=1003, num_motions=6, ray="grid", seed=11, wrap_fraction=0.2, max_frames=90)
    oenv = sc.oracle_env()
    H.randomize_oracle_state(oenv, seed=12)
    a, b = sc.product_env(DEV), sc.product_env(DEV)
    H.copy_oracle_state_to_product(oenv, a)
    H.copy_oracle_state_to_product(oenv, b)
    noise = synth.standin_noise(sc.N, 15, 28, seed=9).to(DEV) if use_standin else None
    host = a.make_host_state()
    for k, v in a.sim_state_tensors().items():
        host[k].copy_(v)
    torch.cuda.synchronize()
    # the device copies of `b` start out as garbage: everything must come from the host buffers
    for v in b.sim_state_tensors().values():
        v.fill_(float("nan"))
    if use_standin:
        a.post_physics_step(parts=2, update_fail_rates=False)     # the stand-in reads ref_*: same ref state on both
        b.post_physics_step_host(host, num_chunks=1, parts=2, update_fail_rates=False)
        a.sim_standin(noise)
    a.post_physics_step()
    b.post_physics_step_host(host, num_chunks=num_chunks, standin_noise=noise)
    torch.cuda.synchronize()
    for name in ("_obs_buf", "_reward_buf", "_done_buf", "_time_buf", "_timestep_buf", "_ref_body_pos", "_ref_dof_pos", "_done_code",
                 "_motion_id_fail_rates", "_char_rigid_body_pos"):
        assert torch.equal(getattr(a, name), getattr(b, name)), name
    na, nb = int(a._done_count.item()), int(b._done_count.item())
    assert na == nb and na > 0
    assert torch.equal(torch.sort(a._done_list[:na]).values, torch.sort(b._done_list[:nb]).values)
    # and through the host-to-host call: reward / done of the step are read BEFORE the reset clears them
    res = b.make_host_results()
    c = sc.product_env(DEV)
    H.copy_oracle_state_to_product(oenv, c)
    if use_standin:
        c.post_physics_step(parts=2, update_fail_rates=False)
    c.step_host(host, res, num_chunks=num_chunks, standin_noise=noise)
    torch.cuda.synchronize()
    assert torch.equal(res["reward"], a._reward_buf.cpu()) and torch.equal(res["done"], a._done_buf.cpu())
    assert int((res["done"] != 0).sum()) > 0 and int(c._done_buf.abs().sum()) == 0
