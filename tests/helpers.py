"""Shared test plumbing: load fixtures, build the same scenario for the CPU oracle and the CUDA path."""
import json
import os

import numpy as np
import torch

from oracle import parc_oracle as O
from parc_b200 import synth

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load_npz(name):
    with np.load(os.path.join(GOLDEN, name), allow_pickle=False) as z:
        return {k: z[k] for k in z.files}


def t(x):
    return torch.from_numpy(np.ascontiguousarray(x))


def golden_clip(name="clip_civilization.npz"):
    g = load_npz(name)
    return dict(root_pos=t(g["root_pos"]), root_rot=t(g["root_rot"]), joint_rot=t(g["joint_rot"]), contacts=t(g["contacts"]),
                fps=int(g["fps"]), loop_mode=int(g["loop_mode"]), hf=t(g["hf"]), min_point=t(g["min_point"]), dx=float(g["dx"]))


def oracle_model():
    return O.CharModel(synth.humanoid_spec())


def oracle_clip(c):
    d = dict(c)
    d["terrain"] = O.Terrain(c["hf"], c["min_point"], [c["dx"], c["dx"]])
    return d


def assert_close(a, b, rtol=1e-5, atol=1e-6, what=""):
    """|a-b| <= atol + rtol*|b| (north_star tolerance: 1e-5 relative, 1e-6 absolute floor)."""
    a = torch.as_tensor(a).detach().cpu().double()
    b = torch.as_tensor(b).detach().cpu().double()
    assert a.shape == b.shape, "%s: shape %s vs %s" % (what, tuple(a.shape), tuple(b.shape))
    err = (a - b).abs()
    tol = atol + rtol * b.abs()
    bad = err > tol
    if bad.any():
        i = torch.argmax(err - tol)
        raise AssertionError("%s: %d/%d elements out of tolerance; worst |%g - %g| = %g at flat index %d" %
                             (what, int(bad.sum()), bad.numel(), a.flatten()[i], b.flatten()[i], err.flatten()[i], int(i)))


def assert_quat_close(a, b, atol=2e-6, what=""):
    """Unit quaternions are compared component-wise with an absolute tolerance (components are <= 1)."""
    assert_close(a, b, rtol=1e-5, atol=atol, what=what)
