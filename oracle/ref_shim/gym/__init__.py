"""Stub of `gym` (PARC/motion_tracker/envs/base_env.py:4 imports it); test infrastructure only."""


class _Box:
    def __init__(self, low=None, high=None, shape=None, dtype=None):
        self.low, self.high, self.shape, self.dtype = low, high, shape, dtype


class _Spaces:
    Box = _Box


spaces = _Spaces()
