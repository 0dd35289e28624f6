"""parc_b200 -- B200-native (sm_100a) kernels for the data-parallel hot path of PARC's
DeepMimic-style motion-tracking environment, behind the reference's own Python method signatures.

Only what the path needs lives here: the CUDA kernels + C ABI (`csrc/`, `include/parc_b200.h`), the
ctypes binding (`_lib`), and host-side mirrors of the reference classes the path goes through
(`KinCharModel`, `MotionLib`, `SubTerrain`, `DeepMimicEnv`, `TrackingEnv`). There is no CPU
fallback: every hot method raises if the CUDA library is missing or a tensor is not on a GPU.
"""
from .kin_char_model import KinCharModel, Joint, JointType, Geom, GeomType   # noqa: F401
from .motion_lib import MotionLib, LoopMode                                   # noqa: F401
from .terrain import SubTerrain                                               # noqa: F401
from .file_io import MotionFrames, load_ms_file, save_ms_file                 # noqa: F401
from . import dm_util                                                         # noqa: F401
from .dm_env import DeepMimicEnv, DeepMimicEnvB200, RefCharEnv                # noqa: F401

__all__ = ["KinCharModel", "Joint", "JointType", "Geom", "GeomType", "MotionLib", "LoopMode", "SubTerrain",
           "MotionFrames", "load_ms_file", "save_ms_file", "dm_util", "DeepMimicEnv", "DeepMimicEnvB200", "RefCharEnv"]
