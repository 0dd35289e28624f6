"""CPU-side tests (no GPU, no compute calls): the C-ABI library loads and exports every symbol that
include/*.h declares, the ctypes structs have the layout of the C structs, the host-side planning
exposed through the ABI (tree parts, obs layout) is consistent, and the product classes refuse to
run without CUDA instead of falling back to anything."""
import ctypes
import glob
import os
import re
import subprocess

import numpy as np
import pytest
import torch

from tests import helpers as H

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _header_functions():
    names = set()
    for path in sorted(glob.glob(os.path.join(ROOT, "include", "*.h"))):
        src = open(path).read()
        src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)        # comments mention symbols too
        src = re.sub(r"//[^\n]*", "", src)
        for m in re.finditer(r"\b(parc_b200_[a-z0-9_]+)\s*\(", src):
            names.add(m.group(1))
    return sorted(names)


def test_library_exports_every_declared_symbol():
    from parc_b200 import _lib
    handle = _lib.lib()                       # builds with nvcc if the .so is missing (cross-compiles on CPU)
    declared = _header_functions()
    assert len(declared) >= 25
    missing = [n for n in declared if not hasattr(handle, n)]
    assert not missing, "declared in include/*.h but not exported: %s" % missing
    # and the Python binding table names exactly the header's functions
    assert sorted(_lib.EXPORTED_SYMBOLS) == declared
    assert handle.parc_b200_version().decode().startswith("parc_b200")


def test_library_is_sm100a_only():
    from parc_b200 import build
    out = subprocess.run(["cuobjdump", "-lelf", str(build.LIB)], capture_output=True, text=True).stdout
    archs = set(re.findall(r"sm_(\d+a?)", out))
    assert archs == {"100a"}, archs


def test_ctypes_struct_sizes_match_the_c_structs(tmp_path):
    """sizeof() of each struct passed by pointer, compiled from the headers with gcc."""
    from parc_b200 import _lib
    prog = tmp_path / "sizes.c"
    prog.write_text('#include <stdio.h>\n#include "parc_b200.h"\n#include "parc_b200_env.h"\n#include "parc_b200_sampler.h"\n'
                    'int main(void){printf("%zu %zu %zu %zu %zu %zu %zu ", sizeof(parc_char_model_t), sizeof(parc_motion_lib_t),'
                    'sizeof(parc_terrain_t), sizeof(parc_env_config_t), sizeof(parc_env_buffers_t), sizeof(parc_host_rows_t),'
                    'sizeof(parc_sampler_config_t));printf("%zu\\n", sizeof(parc_record_tape_t));return 0;}\n')
    exe = tmp_path / "sizes"
    subprocess.run(["gcc", "-I", os.path.join(ROOT, "include"), str(prog), "-o", str(exe)], check=True)
    sizes = [int(x) for x in subprocess.run([str(exe)], capture_output=True, text=True, check=True).stdout.split()]
    mine = [ctypes.sizeof(s) for s in (_lib.CharModelStruct, _lib.MotionLibStruct, _lib.TerrainStruct, _lib.EnvConfigStruct,
                                       _lib.EnvBuffersStruct, _lib.HostRowsStruct, _lib.SamplerConfigStruct, _lib.RecordTapeStruct)]
    assert mine == sizes


def _tree_parts(model):
    from parc_b200 import _lib
    L = _lib.lib()
    ent = np.full((4, 32), -1, dtype=np.int32)
    max_len, n_slots, root_slot = ctypes.c_int32(), ctypes.c_int32(), ctypes.c_int32()
    _lib.check(L.parc_b200_tree_parts(model._cstruct(), ent.ctypes.data_as(ctypes.POINTER(ctypes.c_int32)), ctypes.byref(max_len),
                                      ctypes.byref(n_slots), ctypes.byref(root_slot)))
    return ent, max_len.value, n_slots.value, root_slot.value


def _replay_schedule(parents, ent, n_iter, n_slots, root_slot):
    """Replays the joint schedule the way the kernel does: per lane a register holding the transform of the body it
    evaluated last (the root at first), per 4-lane group `n_slots` shared slots; reads of an iteration happen before its
    writes become visible to other lanes."""
    J = len(parents)
    reg = [0, 0, 0, 0]
    slots = {root_slot: 0} if root_slot >= 0 else {}
    done_at = {0: -1}
    for it in range(n_iter):
        writes = []
        for lane in range(4):
            e = int(ent[lane][it])
            if e < 0:
                continue
            j, from_slot, rslot = e & 0xFF, (e >> 16) & 1, (e >> 17) & 31
            publish, wslot, has_kids = (e >> 22) & 1, (e >> 23) & 31, (e >> 28) & 1
            assert 1 <= j < J and j not in done_at, "joint %d evaluated twice" % j
            par = parents[j]
            assert par in done_at and done_at[par] < it, "joint %d runs before its parent is available" % j
            if from_slot:
                assert rslot < n_slots and slots.get(rslot) == par, "joint %d would read a stale slot" % j
            else:
                assert reg[lane] == par, "joint %d: the lane's registers hold body %d, not the parent %d" % (j, reg[lane], par)
            assert has_kids == int(j in parents)
            reg[lane] = j
            done_at[j] = it
            if publish:
                assert has_kids and wslot < n_slots
                writes.append((wslot, j))
        for w, j in writes:
            slots[w] = j
    return done_at


def test_joint_schedule_of_the_humanoid():
    """The fused step evaluates the joint tree on 4 lanes per (env, pose) along a host-made schedule: every joint exactly
    once, after its parent, with the parent transform where the entry says it is; as many iterations as the tree is deep."""
    model = H.product_model("cpu")
    parents = [int(p) for p in model._parent_indices.tolist()]
    J = len(parents)
    ent, n_iter, n_slots, root_slot = _tree_parts(model)
    assert root_slot == -1
    done_at = _replay_schedule(parents, ent, n_iter, n_slots, root_slot)
    assert sorted(done_at) == list(range(J)), "each joint must be produced exactly once"
    depth = [0] * J
    for j in range(1, J):
        depth[j] = depth[parents[j]] + 1
    assert n_iter == max(max(depth), -(-(J - 1) // 4))          # humanoid: 4 = tree depth = ceil(14 / 4)
    assert n_slots == 1                                          # only the torso is needed by other lanes / later


def test_joint_schedule_of_random_trees():
    from parc_b200 import _lib
    rng = np.random.default_rng(5)
    for J in (1, 2, 5, 9, 17, 32):
        for _ in range(6):
            cs = _lib.CharModelStruct()
            cs.num_bodies, cs.dof_size = J, 0
            parents = [-1] + [int(rng.integers(0, j)) for j in range(1, J)]
            for j in range(J):
                cs.parent[j] = parents[j]
                cs.local_rot[j][3] = 1.0
            ent = np.full((4, 32), -1, dtype=np.int32)
            n_iter, n_slots, root_slot = ctypes.c_int32(), ctypes.c_int32(), ctypes.c_int32()
            _lib.check(_lib.lib().parc_b200_tree_parts(cs, ent.ctypes.data_as(ctypes.POINTER(ctypes.c_int32)), ctypes.byref(n_iter),
                                                       ctypes.byref(n_slots), ctypes.byref(root_slot)))
            done_at = _replay_schedule([0] + parents[1:], ent, n_iter.value, n_slots.value, root_slot.value)
            assert sorted(done_at) == list(range(J))
            assert 1 <= n_slots.value <= 32


def test_ray_points_lattice_detection():
    from parc_b200 import _lib, synth
    assert _lib.ray_points_lattice(synth.grid_ray_points()) == (32, 32)
    assert _lib.ray_points_lattice(synth.cone_ray_points()) == (0, 0)
    pts = synth.grid_ray_points().clone()
    pts[37, 1] += 1e-7                                           # one point off the lattice by an ulp
    assert _lib.ray_points_lattice(pts) == (0, 0)
    xs, ys = torch.linspace(-1, 1, 5), torch.linspace(0, 3, 7)
    x, y = torch.meshgrid(xs, ys, indexing="ij")
    assert _lib.ray_points_lattice(torch.stack([x, y], -1).reshape(-1, 2)) == (5, 7)


def test_obs_dim_matches_reference_layout():
    from parc_b200 import _lib
    from parc_b200.tracking_env import build_config_struct, default_env_config
    model = H.product_model("cpu")
    for P, want in ((441, 1312), (1024, 1895), (0, 871)):
        cfg = build_config_struct(default_env_config(), model, P)
        assert _lib.lib().parc_b200_obs_dim(model._cstruct(), cfg) == want
    ec = default_env_config()
    ec.update(global_obs=True, global_root_height_obs=True)
    assert _lib.lib().parc_b200_obs_dim(model._cstruct(), build_config_struct(ec, model, 441)) == 1313
    ec = default_env_config()
    ec.update(enable_tar_obs=False, use_contact_info=False)
    assert _lib.lib().parc_b200_obs_dim(model._cstruct(), build_config_struct(ec, model, 441)) == 136 + 441


def test_config_struct_rounding_follows_the_reference():
    from parc_b200.tracking_env import build_config_struct, default_env_config
    model = H.product_model("cpu")
    cfg = build_config_struct(default_env_config(), model, 441)
    assert np.float32(cfg.root_pos_termination_dist) == np.float32(0.6) * np.float32(0.6) or \
        abs(cfg.root_pos_termination_dist - 0.36) < 1e-7
    assert np.float32(cfg.ema_keep) == np.float32(1.0 - 0.01)
    assert cfg.num_tar_steps == 6 and cfg.num_key_bodies == 4
    steps = torch.tensor([1, 2, 3, 10, 20, 30], dtype=torch.int32)
    want = (np.float32(cfg.timestep) * steps).float()
    assert [np.float32(cfg.tar_dt[i]) for i in range(6)] == [np.float32(x) for x in want.tolist()]


def test_product_path_refuses_to_run_without_cuda():
    from parc_b200 import _lib
    with pytest.raises(_lib.ParcB200Error):
        _lib.ptr(torch.zeros(4))
    model = H.product_model("cpu")
    if not torch.cuda.is_available():
        with pytest.raises((_lib.ParcB200Error, AssertionError, RuntimeError)):
            model.forward_kinematics(torch.zeros(2, 3), torch.tensor([[0., 0, 0, 1]] * 2), torch.zeros(2, 14, 4))


def test_null_arguments_are_rejected_with_the_reference_error_type():
    """Invalid arguments come back as PARC_ERR_INVALID_ARG -> AssertionError (the reference asserts)."""
    from parc_b200 import _lib
    L = _lib.lib()
    model = H.product_model("cpu")
    rc = L.parc_b200_forward_kinematics(model._cstruct(), None, None, None, 4, None, None, None)
    assert rc != 0
    with pytest.raises(AssertionError):
        _lib.check(rc)
    assert b"NULL" in L.parc_b200_last_error() or len(L.parc_b200_last_error()) > 0
    # n == 0 is a no-op, as an empty env_ids tensor is in the reference (dm_env.py:567-570)
    assert L.parc_b200_forward_kinematics(model._cstruct(), None, None, None, 0, None, None, None) == 0


def test_synthetic_library_is_deterministic_and_in_spec():
    from parc_b200 import synth
    spec = synth.humanoid_spec()
    a = synth.make_clips(spec, 5, seed=11, min_frames=60, max_frames=80)
    b = synth.make_clips(spec, 5, seed=11, min_frames=60, max_frames=80)
    for x, y in zip(a, b):
        assert torch.equal(x["root_pos"], y["root_pos"]) and torch.equal(x["joint_rot"], y["joint_rot"])
        assert 60 <= x["root_pos"].shape[0] <= 80 and x["joint_rot"].shape[1:] == (14, 4)
        n = x["joint_rot"].norm(dim=-1)
        assert float((n - 1).abs().max()) < 1e-5
    hf, min_point, dx, offs = synth.build_terrain_square(a, padding_cells=1)
    tiles = int(np.ceil(np.sqrt(5)))
    assert hf.shape[0] == hf.shape[1] and hf.shape[0] % tiles == 0 and offs.shape == (5, 1, 2)
    assert synth.grid_ray_points().shape == (1024, 2) and synth.cone_ray_points().shape == (441, 2)
