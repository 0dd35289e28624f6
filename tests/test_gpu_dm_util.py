"""The stand-alone operators of parc_b200.dm_util (the reference's TorchScript free functions) against
the CPU oracle on seeded inputs and against the reference golden step. Flags bit-exact, fp32 values
to 1e-5 relative."""
import numpy as np
import pytest
import torch
import yaml

from oracle import parc_oracle as O
from parc_b200 import dm_util
from tests import helpers as H
from tests.helpers import assert_close, load_npz, t

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
J, D, K, S = 15, 28, 4, 6


def _quats(g, *shape):
    q = torch.randn(*shape, 4, generator=g)
    return q / q.norm(dim=-1, keepdim=True)


def _state(seed, n):
    g = torch.Generator().manual_seed(seed)
    r = lambda *s: torch.randn(*s, generator=g)
    return dict(root_pos=r(n, 3) * 3, root_rot=_quats(g, n), root_vel=r(n, 3), root_ang_vel=r(n, 3), joint_rot=_quats(g, n, J - 1),
                dof_vel=r(n, D) * 2, key_pos=r(n, K, 3) * 3)


def _cu(*xs):
    return [x.to(DEV) if torch.is_tensor(x) else x for x in xs]


@pytest.mark.parametrize("n", [1, 37, 4096])
@pytest.mark.parametrize("global_obs,root_h", [(False, False), (True, True), (False, True)])
def test_compute_char_obs(n, global_obs, root_h):
    s = _state(1, n)
    args = [s[k] for k in ("root_pos", "root_rot", "root_vel", "root_ang_vel", "joint_rot", "dof_vel", "key_pos")]
    want = O.compute_char_obs(*args, global_obs, root_h)
    got = dm_util.compute_char_obs(*_cu(*args), global_obs, root_h)
    assert got.shape == want.shape
    assert_close(got, want, atol=5e-6, what="char obs")


def test_compute_char_obs_no_key_bodies_and_empty():
    s = _state(2, 9)
    args = [s[k] for k in ("root_pos", "root_rot", "root_vel", "root_ang_vel", "joint_rot", "dof_vel")]
    want = O.compute_char_obs(*args, torch.zeros(9, 0, 3), False, True)
    got = dm_util.compute_char_obs(*_cu(*args), torch.zeros(9, 0, 3, device=DEV), False, True)
    assert_close(got, want, atol=5e-6, what="char obs without key bodies")
    e = [x[:0] for x in _cu(*args)]
    assert dm_util.compute_char_obs(*e, torch.zeros(0, K, 3, device=DEV), False, True).shape == (0, 1 + 12 + 6 * (J - 1) + D + 3 * K)


@pytest.mark.parametrize("n", [1, 61, 2048])
@pytest.mark.parametrize("global_obs,global_h", [(False, False), (True, False), (False, True)])
def test_compute_tar_obs(n, global_obs, global_h):
    g = torch.Generator().manual_seed(3)
    ref_pos, ref_rot = torch.randn(n, 3, generator=g) * 3, _quats(g, n)
    tp = ref_pos.unsqueeze(1) + torch.randn(n, S, 3, generator=g)
    tr, jr = _quats(g, n, S), _quats(g, n, S, J - 1)
    kp = tp.unsqueeze(2) + torch.randn(n, S, K, 3, generator=g)
    want = O.compute_tar_obs(ref_pos, ref_rot, tp, tr, jr, kp, global_obs, global_h)
    got = dm_util.compute_tar_obs(*_cu(ref_pos, ref_rot, tp, tr, jr, kp), global_obs, global_h)
    assert got.shape == want.shape
    assert_close(got, want, atol=5e-6, what="tar obs")


@pytest.mark.parametrize("n", [1, 100, 4096])
@pytest.mark.parametrize("track_root_h,track_root", [(True, True), (False, True), (True, False)])
def test_compute_deepmimic_reward(n, track_root_h, track_root):
    a, b = _state(4, n), _state(5, n)
    a["key_pos"] = a["root_pos"].unsqueeze(1) + a["key_pos"] / 6      # key bodies within ~1 m of the root, as on a character
    for k in a:                                   # the tracked character stays near its target so that no term saturates at 0
        b[k] = a[k] + (0.01 if k in ("key_pos", "root_pos") else 0.05) * b[k]
    b["root_rot"] = b["root_rot"] / b["root_rot"].norm(dim=-1, keepdim=True)
    b["joint_rot"] = b["joint_rot"] / b["joint_rot"].norm(dim=-1, keepdim=True)
    g = torch.Generator().manual_seed(6)
    jw = torch.rand(J - 1, generator=g)
    dw = torch.rand(D, generator=g)
    names = ("root_pos", "root_rot", "root_vel", "root_ang_vel", "joint_rot", "dof_vel", "key_pos")
    args = [a[k] for k in names] + [b[k] for k in names] + [jw, dw]
    want = O.compute_deepmimic_reward(*args, track_root_h, track_root)
    got = dm_util.compute_deepmimic_reward(*_cu(*args), track_root_h, track_root)
    assert got.shape == (n, 5)
    assert n == 1 or float(want.median(dim=0).values.min()) > 1e-2, "the seeded case must exercise every reward term away from 0"
    assert_close(got[:, :4], want[:, :4], atol=2e-6, what="reward5[0:4]")
    # without root tracking the key bodies are rotated into the heading frame first: the last-bit differences between the CPU's and
    # the GPU's sin/cos of the heading enter the exponent times key_pos_scale * |key offset| * |key error|, ~3e-5 of the term here
    assert_close(got[:, 4], want[:, 4], rtol=1e-5 if track_root else 5e-5, atol=2e-6, what="reward5[4] (key bodies)")


def test_compute_contact_reward():
    g = torch.Generator().manual_seed(7)
    n = 513
    tar = (torch.rand(n, J, generator=g) > 0.5).float()
    f = torch.randn(n, J, 3, generator=g) * (torch.rand(n, J, 1, generator=g) > 0.4)
    f[::7] = 0
    w = torch.rand(J, generator=g) * 5
    want = O.compute_contact_reward(tar, f, w)
    got = dm_util.compute_contact_reward(*_cu(tar, f, w))
    assert_close(got, want, atol=1e-6, what="contact reward")


def _done_case(seed, n):
    g = torch.Generator().manual_seed(seed)
    body = torch.randn(n, J, 3, generator=g)
    body[..., 2] = body[..., 2].abs()
    tar_body = body + torch.randn(n, J, 3, generator=g) * 0.25
    rot = _quats(g, n)
    tar_rot = O.quat_mul(O.axis_angle_to_quat(torch.randn(n, 3, generator=g), torch.randn(n, generator=g) * 0.8), rot)
    time = torch.rand(n, generator=g) * 12
    time[::11] = 0.0
    force = torch.randn(n, J, 3, generator=g) * (torch.rand(n, J, 1, generator=g) > 0.8)
    heights = torch.rand(n, J, generator=g) * 0.5
    ptd = torch.rand(J - 1, generator=g) * 0.6 + 0.5
    return time, rot, body, tar_rot, tar_body, force, heights, ptd


@pytest.mark.parametrize("contact_ids", [[], [3, 7], [0, 5, 8, 11, 14]])
@pytest.mark.parametrize("pose_termination,track_root,early", [(True, True, True), (True, False, True), (False, False, True), (True, True, False)])
def test_compute_done(contact_ids, pose_termination, track_root, early):
    n = 3000
    time, rot, body, tar_rot, tar_body, force, heights, ptd = _done_case(8, n)
    ids = torch.tensor(contact_ids, dtype=torch.int64)
    want = O.compute_done(time, 10.0, rot, body, tar_rot, tar_body, force, ids, heights, pose_termination, ptd, early, track_root, 0.6, 1.309)
    done_buf = torch.zeros(n, dtype=torch.int32, device=DEV)
    got = dm_util.compute_done(done_buf, *_cu(time), 10.0, *_cu(rot, body, body[:, 0], tar_rot, tar_body, force, ids, heights), pose_termination,
                               ptd.to(DEV), False, early, track_root, 0.6, 1.309)
    assert got.dtype == torch.int32
    assert np.array_equal(got.cpu().numpy(), want.numpy()), "done flags must be bit-exact"
    if early and (pose_termination or contact_ids):
        assert set(np.unique(want.numpy()).tolist()) == {O.DONE_NULL, O.DONE_FAIL, O.DONE_TIME}, "the case must exercise all three outcomes"


def test_compute_done_scalar_pose_distance():
    n = 500
    time, rot, body, tar_rot, tar_body, force, heights, _ = _done_case(9, n)
    ids = torch.tensor([3, 7], dtype=torch.int64)
    want = O.compute_done(time, 10.0, rot, body, tar_rot, tar_body, force, ids, heights, True, torch.tensor(0.7), True, True, 0.6, 1.309)
    got = dm_util.compute_done(torch.zeros(n, dtype=torch.int32, device=DEV), *_cu(time), 10.0,
                               *_cu(rot, body, body[:, 0], tar_rot, tar_body, force, ids, heights), True, torch.tensor(0.7, device=DEV), False, True,
                               True, 0.6, 1.309)
    assert np.array_equal(got.cpu().numpy(), want.numpy())


def test_calc_heading():
    g = torch.Generator().manual_seed(10)
    q = _quats(g, 4, 250)
    assert_close(dm_util.calc_heading(q.to(DEV)), O.calc_heading(q), atol=2e-6, what="heading")


def test_operators_vs_reference_golden_step():
    """obs / reward5 / done of the recorded reference step, rebuilt from its recorded inputs with the
    stand-alone operators in the order IGParkourEnv._compute_obs and DeepMimicEnv use them."""
    g = load_npz("ref_env_step.npz")
    cfg = yaml.safe_load(str(g["cfg_yaml"]))["env"]
    model = H.product_model(DEV)
    cu = lambda k: t(g[k]).to(DEV)
    key = [model.get_body_id(b) for b in cfg["key_bodies"]]
    root_pos, root_rot = cu("in_char_root_pos"), cu("in_char_root_rot")
    joint_rot = model.dof_to_rot(cu("in_char_dof_pos"))
    body_pos = model.forward_kinematics(root_pos, root_rot, joint_rot)[0]
    W = float(np.abs(g["out_ref_root_pos"]).max())
    char_obs = dm_util.compute_char_obs(root_pos, root_rot, cu("in_char_root_vel"), cu("in_char_root_ang_vel"), joint_rot, cu("in_char_dof_vel"),
                                        body_pos[:, key], cfg["global_obs"], cfg["root_height_obs"])
    assert char_obs.shape[1] == 136
    H.assert_close_world(char_obs, g["obs"][:, :136], W, what="char obs vs golden")
    tar_obs = dm_util.compute_tar_obs(root_pos, root_rot, cu("tar_root_pos"), cu("tar_root_rot"), cu("tar_joint_rot"), cu("tar_key_pos"),
                                      cfg["global_obs"], False)
    H.assert_close_world(tar_obs.reshape(tar_obs.shape[0], -1), g["obs"][:, 136:136 + 630], W, what="tar obs vs golden")

    om = H.oracle_model()
    jw = torch.tensor(cfg["joint_err_w"], dtype=torch.float32)
    dw = O.make_dof_err_w(om, jw)
    rigid = cu("in_char_rigid_body_pos")
    ref_body = cu("out_ref_body_pos")
    r5 = dm_util.compute_deepmimic_reward(root_pos, root_rot, cu("in_char_root_vel"), cu("in_char_root_ang_vel"), joint_rot, cu("in_char_dof_vel"),
                                          rigid[:, key], cu("out_ref_root_pos"), cu("out_ref_root_rot"), cu("out_ref_root_vel"),
                                          cu("out_ref_root_ang_vel"), cu("out_ref_joint_rot"), cu("out_ref_dof_vel"), ref_body[:, key],
                                          jw.to(DEV), dw.to(DEV), cfg["track_root_h"], cfg["track_root"])
    assert_close(r5, g["reward5"], atol=2e-6, what="reward5 vs golden")
    pen = dm_util.compute_contact_reward(cu("out_ref_contacts"), cu("in_char_contact_forces"),
                                         torch.tensor(cfg["contact_weights"], dtype=torch.float32, device=DEV))
    assert_close(pen.mean(dim=-1), g["contact_penalty"], atol=2e-6, what="contact penalty vs golden")


def test_compute_tracking_error_vs_oracle():
    g = torch.Generator().manual_seed(11)
    n = 1000
    a, b = _state(12, n), _state(13, n)
    for k in a:
        b[k] = a[k] + 0.1 * b[k]
    b["root_rot"] = b["root_rot"] / b["root_rot"].norm(dim=-1, keepdim=True)
    body_rot, tar_body_rot = _quats(g, n, J), _quats(g, n, J)
    body_pos = a["root_pos"].unsqueeze(1) + torch.randn(n, J, 3, generator=g) * 0.4
    tar_body_pos = b["root_pos"].unsqueeze(1) + (body_pos - a["root_pos"].unsqueeze(1)) + torch.randn(n, J, 3, generator=g) * 0.1
    args = [a["root_pos"], a["root_rot"], body_rot, body_pos, b["root_pos"], b["root_rot"], tar_body_rot, tar_body_pos, a["root_vel"],
            a["root_ang_vel"], a["dof_vel"], b["root_vel"], b["root_ang_vel"], b["dof_vel"]]
    want = O.compute_tracking_error(*args)
    got = dm_util.compute_tracking_error(*_cu(*args))
    assert got.shape == (n, 7)
    assert_close(got, want, atol=2e-6, what="tracking error")


def test_tracking_error_vs_reference_golden():
    """a19: the env-level call (dof_to_rot + two FK passes + the error kernel) on the recorded step."""
    from tests.test_gpu_parity import _golden_env
    model = H.product_model(DEV)
    env, g = _golden_env(model)
    ge = load_npz("ref_tracking_error.npz")
    cu = lambda k: t(g[k]).to(DEV)
    for k in ("root_pos", "root_rot", "root_vel", "root_ang_vel", "dof_pos", "dof_vel"):
        getattr(env, "_char_" + k).copy_(cu("in_char_" + k))
    for k in ("root_pos", "root_rot", "root_vel", "root_ang_vel", "joint_rot", "dof_vel"):
        getattr(env, "_ref_" + k).copy_(cu("out_ref_" + k))
    got = env.compute_tracking_error()
    W = float(np.abs(g["out_ref_root_pos"]).max())
    H.assert_close_world(got, ge["tracking_error"], W, what="tracking error vs golden")
