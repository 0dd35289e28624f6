#!/usr/bin/env python
"""Per-phase critical path of env_step_kernel from clock64() stamps (debug build, -DPARC_STEP_TIMING):
  nvcc <build.py flags> -DPARC_STEP_TIMING ... -o scratch/libparc_timing.so;  PARC_B200_LIB=scratch/libparc_timing.so python profiles/step_stamps.py
Stamps are taken by thread 0 of each block: 0 start | 1 after phase A barrier | 2 after lattice table + prefetch | 3 after phase B barrier |
4 after phase C barrier | 5 after bulk issue | 6 after phase D | 7 after heights | 8 end."""
import ctypes
import os
import sys

import numpy as np
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
names = ["A (uniforms, dof->quat)", "lattice table + prefetch", "B (poses)", "C (terms)", "E issue", "D (sums, done)", "H (heights)", "E general + wait"]
nb = N // 16
for parts in (63, 2):
    for _ in range(3):
        env._done_count.zero_()
        _lib.check(L.parc_b200_env_step(km, lib_cs, env._terrain_cs, env._cfg, env._bufs, parts, None, 0, env._stream()))
    torch.cuda.synchronize()
    out = np.zeros((nb, 16), dtype=np.int64)
    assert L.parc_b200_debug_step_stamps(out.ctypes.data_as(ctypes.POINTER(ctypes.c_longlong)), nb) == 0
    d = np.diff(out[:, :9], axis=1)
    print("parts = %d: mean cycles per phase over %d blocks (thread 0's view)" % (parts, nb))
    for k, nm in enumerate(names):
        print("  %-28s mean %8.0f  median %8.0f  p90 %8.0f" % (nm, d[:, k].mean(), np.median(d[:, k]), np.percentile(d[:, k], 90)))
    for k, nm in ((13, "A: thread 0 after the table staging"), (14, "A: root chain done (last warp)")):
        dd = out[:, k] - out[:, 0]
        print("  %-36s mean %8.0f  median %8.0f  p90 %8.0f   (cycles since block start)" % (nm, dd.mean(), np.median(dd), np.percentile(dd, 90)))
    for k, nm in ((9, "B: reference pose done (warp 0)"), (10, "B: target pose 1 done (warp 1)"), (11, "B: target pose 6 done (warp 6)"), (12, "B: character pose done (warp 7)")):
        dd = out[:, k] - out[:, 2]
        print("  %-36s mean %8.0f  median %8.0f  p90 %8.0f   (cycles since the start of phase B)" % (nm, dd.mean(), np.median(dd), np.percentile(dd, 90)))
    tot = out[:, 8] - out[:, 0]
    print("  %-28s mean %8.0f  median %8.0f  p90 %8.0f" % ("block total", tot.mean(), np.median(tot), np.percentile(tot, 90)))
    # per SM: busy span vs sum of block times -> resident blocks
    sm = out[:, 15]
    # phase offset between the two blocks that run side by side on an SM: start-time gap to the nearest other start, in units of a block time
    gaps = []
    for s_ in np.unique(sm):
        st = np.sort(out[sm == s_, 0])
        gaps.extend(np.diff(st))
    print("  start-to-start gaps on an SM (cycles): p10 %.0f median %.0f p90 %.0f" % (np.percentile(gaps, 10), np.median(gaps), np.percentile(gaps, 90)))
    span = [out[sm == s, 8].max() - out[sm == s, 0].min() for s in np.unique(sm)]
    print("  SMs %d, mean span %.0f cycles, blocks per SM %.1f, mean concurrency %.2f" % (len(span), np.mean(span), nb / len(span), tot.sum() / np.sum(span)))
