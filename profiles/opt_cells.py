#!/usr/bin/env python
"""Mean heightfield cells evaluated per body sample point in opt_sdf_kernel at BASELINE config 4 (SURVEY.md 8d asks for the
fp32 roofline of the optimizer loss against 304 x C x 2 x 20 x 3 FLOP per frame with the MEASURED C). Needs the counting build:
  bash profiles/build_variant.sh cells - -DPARC_OPT_COUNT_CELLS;  PARC_B200_LIB=scratch/lib_cells.so python profiles/opt_cells.py
Writes profiles/opt_sdf_cells.json (read by bench_optimizer.py)."""
import ctypes
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench_optimizer  # noqa: E402
from parc_b200 import _lib  # noqa: E402

L = _lib.lib()
out = (ctypes.c_ulonglong * 2)()
B, T = 1024, 300
bench_optimizer.run(B, T, steps=1, warmup=0, cpu_motions=0, dev=torch.device("cuda", 0), peak_gbs=None)     # builds the workload, 1 evaluation
torch.cuda.synchronize()
assert L.parc_b200_debug_opt_cells(out, 1) == 0
evals = bench_optimizer.LAST_EVALS if hasattr(bench_optimizer, "LAST_EVALS") else 1
cells, calls = int(out[0]), int(out[1])
res = {"motions": B, "frames": T, "evaluations_counted": evals, "cells_evaluated": cells, "scan_calls": calls,
       "cells_per_scan": cells / max(calls, 1), "scans_per_frame": calls / (B * T * evals), "cells_per_frame": cells / (B * T * evals),
       "note": "a scan = one sample point against its window (the gradient pass of the contact term re-scans the closest point of a body); "
               "counted with -DPARC_OPT_COUNT_CELLS, need_grad = 1"}
print(json.dumps(res, indent=1))
json.dump(res, open(os.path.join(ROOT, "profiles", "opt_sdf_cells.json"), "w"), indent=1)
