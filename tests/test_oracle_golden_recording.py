"""CPU: the recording-path oracle against the unmodified reference methods' outputs (tests/golden/ref_recording.npz)."""
import numpy as np
import pytest
import torch

from oracle import parc_oracle_recording as OR
from tests.helpers import load_npz, t


@pytest.mark.parametrize("tag,record_ref,bypass,obs", [("sim", False, False, True), ("ref", True, True, False)])
def test_recording_oracle_matches_reference(tag, record_ref, bypass, obs):
    g = load_npz("ref_recording.npz")
    K, N = g[tag + "_in_done"].shape
    rec = OR.Recorder(N, K, 34, 15, obs_dim=g[tag + "_in_obs"].shape[-1] if obs else 0, record_ref=record_ref, bypass_record_fail=bypass)
    lens = t(g[tag + "_motion_lengths"])[t(g[tag + "_motion_ids"])]
    pre = "ref_" if record_ref else ""
    saved = []
    for k in range(K):
        saved += rec.step(t(g[tag + "_in_" + pre + "root_pos"][k]), t(g[tag + "_in_" + pre + "root_rot"][k]), t(g[tag + "_in_" + pre + "dof_pos"][k]),
                          t(g[tag + "_in_ref_contacts"][k]) if record_ref else t(g[tag + "_in_contact_forces"][k]), t(g[tag + "_in_done"][k]),
                          t(g[tag + "_in_motion_times"][k]), lens, obs=t(g[tag + "_in_obs"][k]))
    assert np.array_equal(rec.cursor.numpy(), g[tag + "_counts"])
    np.testing.assert_allclose(rec.frames.numpy(), g[tag + "_frames"], rtol=1e-5, atol=2e-6)
    assert np.array_equal(rec.contacts.numpy(), g[tag + "_contacts"])
    if obs:
        assert np.array_equal(rec.obs.numpy(), g[tag + "_obs"])
    assert np.array_equal(rec.success.numpy(), g[tag + "_success"]) and np.array_equal(rec.writing.numpy(), g[tag + "_writing"])
    assert saved == g[tag + "_saved_envs"].tolist()
    assert rec.flag == bool(g[tag + "_still_recording"])
