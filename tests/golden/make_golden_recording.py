"""Golden vectors for the recording path: the UNMODIFIED reference methods
IGParkourEnv.write_agent_states / _get_char_state / build_agent_states_dict
(PARC/motion_tracker/envs/ig_parkour/ig_parkour_env.py:664-687, 759-796, 1153-1182) driven for K control steps over
N envs through a stand-in `self` that carries only the attributes those methods touch (no simulator):

    python tests/golden/make_golden_recording.py   -> ref_recording.npz

Per step the inputs are seeded random character / reference states, contact forces, done flags and motion times;
the outputs are every env's recorded frames / contacts (/ observations), the success flags, and which envs were
handed to save_agent_states_to_file with which name."""
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from oracle import ref_import  # noqa: E402

ref_import.setup()
torch.set_num_threads(1)

import parc.motion_tracker.envs.ig_parkour.ig_parkour_env as renv  # noqa: E402

N, K, J, D, OBS = 12, 9, 15, 28, 20
FAIL = 1


class DmStub:
    def __init__(self, owner):
        self.o = owner

    def get_env_motion_length(self, env_id):
        return self.o.motion_lengths[self.o.motion_ids[env_id]]

    def get_env_motion_time(self, env_id):
        return self.o.motion_times[env_id]

    def get_env_motion_name(self, env_id):
        return "clip%d" % int(self.o.motion_ids[env_id])


def make_self(record_ref, bypass):
    s = types.SimpleNamespace()
    for name in ("write_agent_states", "_get_char_state", "is_writing_agent_states", "set_write_agent_states_flag", "is_writing_env_state",
                 "set_writing_env_state", "set_env_success_state", "get_env_success_states", "build_agent_states_dict"):
        setattr(s, name, types.MethodType(getattr(renv.IGParkourEnv, name), s))
    s._num_envs, s._num_dm_envs, s._control_freq, s._timestep, s._device = N, N, 30, 1.0 / 30.0, "cpu"
    s._bypass_record_fail = bypass
    if record_ref:
        s._record_ref = True
    s.saved = []
    s.save_agent_states_to_file = lambda env_id, name=None: s.saved.append((env_id, name))
    s.has_dm_envs = lambda: True
    s.get_dm_env = lambda: DmStub(s)
    s._compute_obs = lambda env_ids, ret_obs_shapes=False: {"obs": OBS}
    return s


def run(tag, record_ref, bypass, record_obs, out, g):
    s = make_self(record_ref, bypass)
    s.motion_lengths = torch.tensor([1.0, 2.5, 4.0])
    s.motion_ids = torch.randint(0, 3, (N,), generator=g)
    s.build_agent_states_dict(name_suffix="_dm", record_obs=record_obs)
    steps = {k: [] for k in ("root_pos", "root_rot", "dof_pos", "contact_forces", "ref_root_pos", "ref_root_rot", "ref_dof_pos",
                             "ref_contacts", "done", "motion_times", "obs")}
    for k in range(K):
        q = torch.randn(N, 4, generator=g); q = q / q.norm(dim=-1, keepdim=True)
        rq = torch.randn(N, 4, generator=g); rq = rq / rq.norm(dim=-1, keepdim=True)
        s._char_root_pos, s._char_root_rot = torch.randn(N, 3, generator=g), q
        s._char_dof_pos = torch.randn(N, D, generator=g)
        f = torch.randn(N, J, 3, generator=g) * (torch.rand(N, J, 1, generator=g) < 0.4)
        f[0, 0] = torch.tensor([6e-6, 0.0, 0.0])                       # below eps = 1e-5
        s._char_contact_forces = f
        s._ref_root_pos, s._ref_root_rot = torch.randn(N, 3, generator=g), rq
        s._ref_dof_pos, s._ref_contacts = torch.randn(N, D, generator=g), torch.rand(N, J, generator=g)
        s._obs_buf = torch.randn(N, OBS, generator=g)
        done = torch.zeros(N, dtype=torch.int32)
        done[torch.rand(N, generator=g) < 0.18] = FAIL
        done[torch.rand(N, generator=g) < 0.05] = 3                    # TIME: keeps recording (:776)
        s._done_buf = done
        # half of the envs sit near the end of their clip (a FAIL there counts as a tracked motion, :786-789)
        near = torch.rand(N, generator=g) < 0.5
        lens = s.motion_lengths[s.motion_ids]
        s.motion_times = torch.where(near, lens - torch.rand(N, generator=g) * 0.1, lens * torch.rand(N, generator=g) * 0.8)
        for name, v in (("root_pos", s._char_root_pos), ("root_rot", q), ("dof_pos", s._char_dof_pos), ("contact_forces", f),
                        ("ref_root_pos", s._ref_root_pos), ("ref_root_rot", rq), ("ref_dof_pos", s._ref_dof_pos),
                        ("ref_contacts", s._ref_contacts), ("done", done), ("motion_times", s.motion_times), ("obs", s._obs_buf)):
            steps[name].append(v.clone())
        s.write_agent_states()
    for name, v in steps.items():
        out[tag + "_in_" + name] = torch.stack(v).numpy()
    out[tag + "_motion_ids"] = s.motion_ids.numpy()
    out[tag + "_motion_lengths"] = s.motion_lengths.numpy()
    counts = [len(s._dm_agent_motion[e]["frames"]) for e in range(N)]
    out[tag + "_counts"] = np.array(counts)
    frames = np.zeros((N, K, 6 + D), dtype=np.float32); contacts = np.zeros((N, K, J), dtype=np.float32)
    obs = np.zeros((N, K, OBS), dtype=np.float32)
    for e in range(N):
        if counts[e]:
            frames[e, :counts[e]] = torch.stack(s._dm_agent_motion[e]["frames"]).numpy()
            contacts[e, :counts[e]] = torch.stack(s._dm_agent_motion[e]["contacts"]).numpy()
            if record_obs:
                obs[e, :counts[e]] = torch.stack(s._dm_agent_motion[e]["obs"]).numpy()
    out[tag + "_frames"], out[tag + "_contacts"], out[tag + "_obs"] = frames, contacts, obs
    out[tag + "_success"] = np.array(s.get_env_success_states())
    out[tag + "_writing"] = np.array([s.is_writing_env_state(e) for e in range(N)])
    out[tag + "_still_recording"] = np.array(s.is_writing_agent_states())
    out[tag + "_saved_envs"] = np.array([e for e, _ in s.saved], dtype=np.int64)
    out[tag + "_saved_names"] = np.array([n for _, n in s.saved])
    print(tag, "counts", counts, "success", int(np.sum(out[tag + "_success"])), "saved", s.saved)


def main():
    out = {}
    g = torch.Generator().manual_seed(17)
    run("sim", record_ref=False, bypass=False, record_obs=True, out=out, g=g)
    run("ref", record_ref=True, bypass=True, record_obs=False, out=out, g=g)
    np.savez_compressed(os.path.join(HERE, "ref_recording.npz"), **out)


if __name__ == "__main__":
    main()
