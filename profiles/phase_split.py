#!/usr/bin/env python
"""Times env_step_kernel on the bench workload (65,536 envs, 1,000 clips, P=1024) for several `parts` masks, to see how the
launch time splits over heights / target+character poses / reference pose + reward + termination. CUDA events, 20 launches each."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from parc_b200 import _lib  # noqa: E402

dev = torch.device("cuda", 0)
world = bench.build_world(1000)
env = bench.build_gpu_env(world, 65536, 0, 65536, dev)
bench.init_gpu_state(env, 0)
env._ray_hfs = torch.zeros(65536, 1024, device=dev)      # so that HEIGHTS and OBS can run apart
env._build_buffers_struct()
L = _lib.lib()
km, lib_cs = env._kin_char_model._cstruct(), env._mlib.cstruct()
T, R, H, O, W, D = 1, 2, 4, 8, 16, 32
for name, parts in [("all", 63), ("heights only -> ray_hfs", H), ("obs only (7 poses + copy-out; heights copied from ray_hfs)", O),
                    ("heights+obs", H | O), ("ref only", R), ("ref+reward+done", R | W | D), ("ref+reward+done+heights", R | W | D | H), ("all", 63)]:
    for _ in range(3):
        L.parc_b200_env_step(km, lib_cs, env._terrain_cs, env._cfg, env._bufs, parts, None, 0, env._stream())
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(20):
        env._done_count.zero_()
        _lib.check(L.parc_b200_env_step(km, lib_cs, env._terrain_cs, env._cfg, env._bufs, parts, None, 0, env._stream()))
    b.record()
    torch.cuda.synchronize()
    print("%-70s %.3f ms" % (name, a.elapsed_time(b) / 20))
