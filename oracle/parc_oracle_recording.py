"""TEST INFRASTRUCTURE ONLY -- CPU restatement of the recording path IGParkourEnv.write_agent_states /
_get_char_state (PARC/motion_tracker/envs/ig_parkour/ig_parkour_env.py:664-687, 759-796), vectorised over envs.
SURVEY.md section 8(f) rank 4. Imported by tests only.

Parity pin: tests/golden/ref_recording.npz (tests/golden/make_golden_recording.py drives the unmodified reference
methods); tests/test_oracle_golden_recording.py checks this file against it."""
import torch

from . import parc_oracle as O


class Recorder:
    def __init__(self, num_envs, max_frames, state_dim, num_contacts, obs_dim=0, record_ref=False, bypass_record_fail=False,
                 timestep=1.0 / 30.0):
        self.frames = torch.zeros(num_envs, max_frames, state_dim)
        self.contacts = torch.zeros(num_envs, max_frames, num_contacts)
        self.obs = torch.zeros(num_envs, max_frames, obs_dim) if obs_dim else None
        self.cursor = torch.zeros(num_envs, dtype=torch.int64)
        self.writing = torch.ones(num_envs, dtype=torch.bool)          # :1177
        self.success = torch.zeros(num_envs, dtype=torch.bool)         # :1178
        self.flag = True                                               # set_write_agent_states_flag(True), :1176
        self.record_ref, self.bypass, self.timestep = record_ref, bypass_record_fail, timestep

    def step(self, root_pos, root_rot, dof_pos, contacts_or_forces, done, motion_time, motion_length, obs=None, eps=1e-5):
        """One write_agent_states call; returns the env ids handed to save_agent_states_to_file, in env order."""
        if not self.flag:                                              # :760
            return []
        w = self.writing.clone()
        self.flag = bool(w.any())                                      # :761-766
        state = torch.cat([root_pos, O.quat_to_exp_map(root_rot), dof_pos], dim=-1)      # :664-680
        con = contacts_or_forces if self.record_ref else (torch.norm(contacts_or_forces, dim=-1) > eps).float()
        ids = torch.nonzero(w).flatten()
        self.frames[ids, self.cursor[ids]] = state[ids]
        self.contacts[ids, self.cursor[ids]] = con[ids]
        if self.obs is not None:
            self.obs[ids, self.cursor[ids]] = obs[ids]
        self.cursor[ids] += 1
        fail = w & (done == O.DONE_FAIL)                               # :776
        self.writing[fail] = False
        early = motion_time.double() < motion_length.double() - self.timestep * 2.0       # :787 (python floats)
        ok = fail & (torch.ones_like(early) if self.bypass else ~early)
        self.success[ok] = True
        return torch.nonzero(ok).flatten().tolist()
