"""GPU: the recording tape (TrackingEnv.build_agent_states_dict / write_agent_states -> parc_b200_record_agent_states)
against the unmodified reference methods' outputs (tests/golden/ref_recording.npz). SURVEY.md section 8(f) rank 4.
Counts, flags and contact labels are bit-exact; states within 1e-5."""
import numpy as np
import pytest
import torch

from tests import helpers as H
from tests.helpers import load_npz, t

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


@pytest.mark.parametrize("tag,record_ref,bypass,obs", [("sim", False, False, True), ("ref", True, True, False)])
def test_recording_tape_matches_reference(tag, record_ref, bypass, obs):
    g = load_npz("ref_recording.npz")
    K, N = g[tag + "_in_done"].shape
    # three clips whose lengths are the golden run's motion lengths (1.0, 2.5, 4.0 s at 30 fps)
    base = H.golden_clip("clip_teaser_dm.npz")
    clips = []
    for L in g[tag + "_motion_lengths"]:
        c = dict(base)
        nf = int(round(float(L) * 30)) + 1
        for key in ("root_pos", "root_rot", "joint_rot", "contacts"):
            c[key] = base[key][:nf].clone()
        clips.append(c)
    sc = H.Scenario(num_envs=N, num_motions=3, ray="cone", seed=2, max_frames=60)
    sc.clips = clips                                 # Scenario builds its library from .clips
    env = sc.product_env(DEV)
    assert torch.allclose(env._mlib._motion_lengths.cpu(), t(g[tag + "_motion_lengths"]))
    env._motion_ids.copy_(t(g[tag + "_motion_ids"]).to(DEV))
    env._motion_time_offsets.zero_()
    env.build_agent_states_dict(name_suffix="_dm", record_obs=obs, max_frames=K, record_ref=record_ref, bypass_record_fail=bypass)
    saved = []
    for k in range(K):
        gi = lambda name: t(g[tag + "_in_" + name][k]).to(DEV)
        env._char_root_pos.copy_(gi("root_pos")); env._char_root_rot.copy_(gi("root_rot")); env._char_dof_pos.copy_(gi("dof_pos"))
        env._char_contact_forces.copy_(gi("contact_forces"))
        env._ref_root_pos.copy_(gi("ref_root_pos")); env._ref_root_rot.copy_(gi("ref_root_rot")); env._ref_dof_pos.copy_(gi("ref_dof_pos"))
        env._ref_contacts.copy_(gi("ref_contacts"))
        env._done_buf.copy_(gi("done")); env._time_buf.copy_(gi("motion_times"))
        if obs:
            env._obs_buf.zero_()
            env._obs_buf[:, :g[tag + "_in_obs"].shape[-1]].copy_(gi("obs"))
        env.write_agent_states()
        saved += env.envs_finished_this_step().cpu().tolist()
    torch.cuda.synchronize()
    tape = env._tape
    assert np.array_equal(tape["cursor"].cpu().numpy(), g[tag + "_counts"])
    np.testing.assert_allclose(tape["frames"].cpu().numpy(), g[tag + "_frames"], rtol=1e-5, atol=3e-6)
    assert np.array_equal(tape["contacts"].cpu().numpy(), g[tag + "_contacts"])
    if obs:
        W = g[tag + "_in_obs"].shape[-1]
        assert np.array_equal(tape["obs"][:, :, :W].cpu().numpy(), g[tag + "_obs"])
    assert np.array_equal(env.get_env_success_states().cpu().numpy(), g[tag + "_success"])
    assert np.array_equal(tape["writing"].bool().cpu().numpy(), g[tag + "_writing"])
    assert saved == g[tag + "_saved_envs"].tolist()
    assert env.is_writing_agent_states() == bool(g[tag + "_still_recording"]) and int(tape["overflow"].item()) == 0
    m = env.get_agent_motion(3)
    assert m["frames"].shape == (int(g[tag + "_counts"][3]), 34) and m["loop_mode"] == "CLAMP" and m["fps"] == 30
    np.testing.assert_allclose((m["frames"][:, 0:2] - env._env_offsets[3, 0:2]).cpu().numpy(), g[tag + "_frames"][3, :m["frames"].shape[0], 0:2],
                               rtol=1e-5, atol=3e-6)
    # a tape that is too short drops the extra frames and says so
    env.build_agent_states_dict(max_frames=2, record_ref=record_ref, bypass_record_fail=bypass)
    env._done_buf.zero_()
    for _ in range(3):
        env.write_agent_states()
    assert int(env._tape["overflow"].item()) == 1 and int(env._tape["cursor"].max().item()) == 2
