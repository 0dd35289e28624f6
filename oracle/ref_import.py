"""TEST INFRASTRUCTURE ONLY -- imports the unmodified PARC reference from /root/reference on CPU.

Only `tests/golden/make_golden.py` and the "reference present" branch of the CPU tests may use
this; nothing under `parc_b200/` imports it. `/root/reference` does not exist on the GPU box, so
everything that must run there reads committed fixtures under `tests/golden/` instead.

Recipe: the package directory is `PARC/` but every import says `parc.`
(PARC/anim/motion_lib.py:8), so a `parc -> /root/reference/PARC` symlink is put on sys.path
together with stubs for trimesh / gym / isaacgym (oracle/ref_shim/), and
`path_loader._load_data_dir` (PARC/util/path_loader.py:19-31) is pointed at the reference root.
"""
import os
import sys
import tempfile
from pathlib import Path

REF_ROOT = Path(os.environ.get("PARC_REFERENCE_ROOT", "/root/reference"))
_SHIM = Path(__file__).resolve().parent / "ref_shim"
_state = {"done": False}


def available() -> bool:
    return (REF_ROOT / "PARC" / "anim" / "motion_lib.py").is_file()


def setup():
    """Make `import parc.*` resolve to the reference. Idempotent."""
    if _state["done"]:
        return
    if not available():
        raise RuntimeError("PARC reference not present at %s" % REF_ROOT)
    link_dir = Path(tempfile.mkdtemp(prefix="parc_ref_link_"))
    os.symlink(REF_ROOT / "PARC", link_dir / "parc")
    sys.path.insert(0, str(link_dir))
    sys.path.insert(0, str(_SHIM))
    import parc.util.path_loader as pl
    pl._load_data_dir = lambda: REF_ROOT
    _state["done"] = True


def humanoid_xml() -> str:
    return str(REF_ROOT / "data" / "assets" / "humanoid.xml")


def clip(name: str) -> str:
    return str(REF_ROOT / "data" / "motion_terrains" / name)
