#!/usr/bin/env python
"""Marginal cost of the parts of env_step_kernel at BASELINE config 2: launch time for several `parts` masks (CUDA events, 20 launches each).
Parts masks other than 63 take the general (run-time layout) instantiation, so the 63 line is given for both."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from parc_b200 import _lib  # noqa: E402

dev = torch.device("cuda", 0)
N = 65536
world = bench.build_world(1000)
env = bench.build_gpu_env(world, N, 0, N, dev)
bench.init_gpu_state(env, 0)
L = _lib.lib()
km, lib_cs = env._kin_char_model._cstruct(), env._mlib.cstruct()
T, R, H, O, W, D = _lib.STEP_TIME, _lib.STEP_REF, _lib.STEP_HEIGHTS, _lib.STEP_OBS, _lib.STEP_REWARD, _lib.STEP_DONE
ids = torch.arange(N, device=dev, dtype=torch.int64)
for name, parts, rows in (("all (fast instantiation)", 63, None), ("all, row list (general instantiation)", 63, ids), ("no obs (heights to ray_hfs)", 63 - O, ids),
                          ("no obs, no heights", 63 - O - H, ids), ("ref only", R, ids), ("ref + reward + done", R + W + D, ids), ("heights only", H, ids),
                          ("obs + heights only", O + H, ids), ("time only", T, ids)):
    ts = []
    try:
        _lib.check(L.parc_b200_env_step(km, lib_cs, env._terrain_cs, env._cfg, env._bufs, parts, _lib.ptr(rows) if rows is not None else None,
                                        N if rows is not None else 0, env._stream()))
    except AssertionError as ex:
        print("%-44s refused: %s" % (name, ex))
        continue
    for k in range(25):
        env._done_count.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        _lib.check(L.parc_b200_env_step(km, lib_cs, env._terrain_cs, env._cfg, env._bufs, parts, _lib.ptr(rows) if rows is not None else None,
                                        N if rows is not None else 0, env._stream()))
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    ts = sorted(ts[5:])
    print("%-44s median %.4f ms  min %.4f ms" % (name, ts[len(ts) // 2], ts[0]))
