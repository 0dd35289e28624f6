"""Stub of `trimesh` for importing the PARC reference on a box without it (test infrastructure only).

Only what PARC/anim/kin_char_model.py:9 and PARC/util/geom_util.py:749 touch: `load` (never reached for
the bundled humanoid, which has no mesh assets) and `creation.icosphere(subdivisions=0, radius)`.
The icosahedron vertex order is ours (trimesh is unpinned in the reference's requirements.txt:8), so
body sample points are treated as kernel *inputs*, shared by oracle and kernels.
"""
import numpy as np


def load(*args, **kwargs):
    raise RuntimeError("trimesh stub: mesh loading unavailable")


class _Mesh:
    def __init__(self, vertices):
        self.vertices = vertices


class _Creation:
    @staticmethod
    def icosphere(subdivisions=0, radius=1.0):
        assert subdivisions == 0, "stub supports only the raw icosahedron"
        p = (1.0 + 5.0 ** 0.5) / 2.0
        v = np.array([(-1, p, 0), (1, p, 0), (-1, -p, 0), (1, -p, 0),
                      (0, -1, p), (0, 1, p), (0, -1, -p), (0, 1, -p),
                      (p, 0, -1), (p, 0, 1), (-p, 0, -1), (-p, 0, 1)], dtype=np.float64)
        v = v / np.linalg.norm(v, axis=-1, keepdims=True) * radius
        return _Mesh(v)


creation = _Creation()
