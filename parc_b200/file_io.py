"""On-disk motion/terrain ("MS") file format and the MotionFrames container -- boundary types.

The format is the reference's (PARC/util/file_io.py:1-137): a pickled dict with three
individually pickled sub-dicts `motion_data`, `terrain_data`, `misc_data`. Only the fields the hot
path consumes are interpreted; `misc_data` is passed through untouched.
"""
import os
import pickle
from dataclasses import asdict, dataclass
from typing import Any, MutableMapping, Optional

import numpy as np
import torch

MOTION_DATA_KEY = "motion_data"
TERRAIN_DATA_KEY = "terrain_data"
MISC_DATA_KEY = "misc_data"
HF_MASK_INDS_KEY = "hf_mask_inds"


@dataclass
class MSMotionData:
    root_pos: Any            # [n, 3]
    root_rot: Any            # [n, 4] quat (x, y, z, w)
    joint_rot: Any           # [n, J-1, 4]
    body_contacts: Any       # [n, n_body] or None
    fps: int
    loop_mode: str           # "CLAMP" | "WRAP"


@dataclass
class MSTerrainData:
    hf: Any                  # [X, Y]
    hf_maxmin: Any           # [X, Y, 2]
    min_point: Any           # [2]
    dx: float


@dataclass
class MSFileData:
    motion_data: Optional[MSMotionData]
    terrain_data: Optional[MSTerrainData]
    misc_data: Optional[MutableMapping[str, Any]]


def _loads(blob):
    if blob is None:
        return None
    try:
        return pickle.loads(blob)
    except Exception:
        return None


def load_ms_file(filepath, device="cpu") -> MSFileData:
    """Reads an MS .pkl and moves the arrays the hot path needs onto `device` as fp32 tensors
    (reference: file_io.py:111-137 + file_io_helper.py:168-220). misc_data entries that need
    classes of the reference's optimizer module are left as stored."""
    with open(filepath, "rb") as f:
        container = pickle.load(f)
    motion = _loads(container.get(MOTION_DATA_KEY))
    terrain = _loads(container.get(TERRAIN_DATA_KEY))
    misc = _loads(container.get(MISC_DATA_KEY))

    def to_t(x):
        if x is None:
            return None
        if isinstance(x, np.ndarray):
            x = torch.from_numpy(x)
        return x.to(device=device, dtype=torch.float32)

    if motion is not None:
        motion = MSMotionData(**motion)
        motion.root_pos = to_t(motion.root_pos)
        motion.root_rot = to_t(motion.root_rot)
        motion.joint_rot = to_t(motion.joint_rot)
        motion.body_contacts = to_t(motion.body_contacts)
    if terrain is not None:
        terrain = MSTerrainData(**terrain)
        terrain.hf = to_t(terrain.hf)
        terrain.hf_maxmin = to_t(terrain.hf_maxmin)
    if misc is not None and HF_MASK_INDS_KEY in misc:
        misc[HF_MASK_INDS_KEY] = [torch.as_tensor(x).to(dtype=torch.int64, device=device) for x in misc[HF_MASK_INDS_KEY]]
    return MSFileData(motion_data=motion, terrain_data=terrain, misc_data=misc)


def save_ms_file(data: MSFileData, filepath):
    assert os.path.splitext(filepath)[1] == ".pkl", "unknown extension"

    def np_dict(d):
        if d is None:
            return None
        out = {}
        for k, v in asdict(d).items():
            out[k] = v.detach().cpu().numpy() if isinstance(v, torch.Tensor) else v
        return pickle.dumps(out)

    container = {
        MOTION_DATA_KEY: np_dict(data.motion_data),
        TERRAIN_DATA_KEY: np_dict(data.terrain_data),
        MISC_DATA_KEY: None if data.misc_data is None else pickle.dumps(data.misc_data),
    }
    with open(filepath, "wb") as f:
        pickle.dump(container, f)


class MotionFrames:
    """Bag of per-frame tensors (reference: PARC/util/motion_util.py:11-213)."""

    _FIELDS = ("root_pos", "root_rot", "joint_rot", "body_pos", "body_rot", "contacts")

    def __init__(self, root_pos=None, root_rot=None, joint_rot=None, body_pos=None, body_rot=None, contacts=None):
        self.root_pos = root_pos
        self.root_rot = root_rot
        self.joint_rot = joint_rot
        self.body_pos = body_pos
        self.body_rot = body_rot
        self.contacts = contacts

    def _map(self, fn):
        return MotionFrames(**{k: (None if getattr(self, k) is None else fn(getattr(self, k))) for k in self._FIELDS})

    def get_copy(self, device):
        return self._map(lambda t: t.clone().to(device=device))

    def get_slice(self, in_slice):
        nd = self.root_pos.dim()
        assert nd in (2, 3), "tensors have unsupported shape: " + str(self.root_pos.shape)
        return self._map((lambda t: t[:, in_slice]) if nd == 3 else (lambda t: t[in_slice]))

    def unsqueeze(self, dim):
        return self._map(lambda t: t.unsqueeze(dim))

    def squeeze(self, dim):
        return self._map(lambda t: t.squeeze(dim))

    def assert_num_dims(self, n):
        assert self.root_pos.dim() == n
