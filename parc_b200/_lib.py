"""ctypes binding of the C ABI declared in include/parc_b200.h / parc_b200_env.h.

There is no CPU fallback: if the CUDA library cannot be loaded, or a tensor is not a CUDA tensor,
every call fails loudly. torch is used for device memory and streams only.
"""
import ctypes as C
import threading

import torch

from . import build as _build

MAX_BODIES = 32
MAX_TAR_STEPS = 16
MAX_KEY_BODIES = 8
MAX_DOFS = 96
EMA_NOALLOC_MAX_INTS = 1 << 20      # PARC_EMA_NOALLOC_MAX_INTS, include/parc_b200_env.h

PARC_OK = 0
PARC_ERR_INVALID_ARG = -1
PARC_ERR_UNSUPPORTED = -2
PARC_ERR_CUDA = -3

STEP_TIME, STEP_REF, STEP_HEIGHTS, STEP_OBS, STEP_REWARD, STEP_DONE, STEP_ALL = 1, 2, 4, 8, 16, 32, 63


class CharModelStruct(C.Structure):
    _fields_ = [
        ("num_bodies", C.c_int32),
        ("dof_size", C.c_int32),
        ("parent", C.c_int32 * MAX_BODIES),
        ("joint_type", C.c_int32 * MAX_BODIES),
        ("dof_idx", C.c_int32 * MAX_BODIES),
        ("local_trans", (C.c_float * 3) * MAX_BODIES),
        ("local_rot", (C.c_float * 4) * MAX_BODIES),
        ("axis", (C.c_float * 3) * MAX_BODIES),
    ]


class MotionLibStruct(C.Structure):
    _fields_ = [
        ("interp", C.c_void_p),
        ("vel", C.c_void_p),
        ("meta", C.c_void_p),
        ("total_frames", C.c_int64),
        ("num_motions", C.c_int32),
        ("irec", C.c_int32),
        ("vrec", C.c_int32),
        ("num_contacts", C.c_int32),
    ]


class SamplerConfigStruct(C.Structure):     # parc_sampler_config_t, include/parc_b200_sampler.h
    _fields_ = [("seq_len", C.c_int32), ("ref_frame_idx", C.c_int32), ("relative_z_style", C.c_int32), ("grid_dim_x", C.c_int32),
                ("grid_dim_y", C.c_int32), ("center_ix", C.c_int32), ("center_iy", C.c_int32), ("min_h", C.c_float),
                ("max_h", C.c_float), ("future_pos_noise_scale", C.c_float)]


class HfAugStruct(C.Structure):              # parc_hf_aug_t, include/parc_b200_sampler.h
    _fields_ = [("mode", C.c_int32), ("max_boxes", C.c_int32), ("max_terrain_cells", C.c_int32), ("timestep", C.c_float),
                ("hf_maxmin", C.c_void_p), ("maxmin_offset", C.c_void_p), ("mask_frame_off", C.c_void_p), ("mask_start", C.c_void_p),
                ("mask_cells", C.c_void_p), ("rec_i", C.c_void_p), ("rec_f", C.c_void_p), ("noise", C.c_void_p)]


AUG_REC_INTS, AUG_BOX_FLOATS = 8, 7           # PARC_AUG_REC_INTS, PARC_AUG_BOX_FLOATS


class RecordTapeStruct(C.Structure):         # parc_record_tape_t, include/parc_b200_env.h
    _fields_ = [("frames", C.c_void_p), ("contacts", C.c_void_p), ("obs", C.c_void_p), ("cursor", C.c_void_p), ("writing", C.c_void_p),
                ("success", C.c_void_p), ("saved_list", C.c_void_p), ("saved_count", C.c_void_p), ("still_recording", C.c_void_p),
                ("overflow", C.c_void_p), ("max_frames", C.c_int32), ("obs_dim", C.c_int32)]


class ResetOptionsStruct(C.Structure):       # parc_reset_options_t, include/parc_b200_env.h
    _fields_ = [("root_pos_offset", C.c_void_p), ("root_rot_offset", C.c_void_p), ("motion_start_time_fraction", C.c_void_p),
                ("demo_mode", C.c_int32)]


class HostRowsStruct(C.Structure):
    _fields_ = [("dst_device", C.c_void_p), ("src_host", C.c_void_p), ("row_bytes", C.c_int64)]


class TerrainStruct(C.Structure):
    _fields_ = [
        ("hf", C.c_void_p),
        ("dim_x", C.c_int32),
        ("dim_y", C.c_int32),
        ("min_x", C.c_float),
        ("min_y", C.c_float),
        ("dx", C.c_float),
        ("dy", C.c_float),
    ]


class View3(C.Structure):
    _fields_ = [
        ("ptr", C.c_void_p),
        ("row_stride", C.c_int64),
        ("mid_stride", C.c_int64),
        ("elem_stride", C.c_int64),
    ]


class EnvConfigStruct(C.Structure):
    _fields_ = [
        ("num_tar_steps", C.c_int32),
        ("tar_dt", C.c_float * MAX_TAR_STEPS),
        ("num_key_bodies", C.c_int32),
        ("key_body_ids", C.c_int32 * MAX_KEY_BODIES),
        ("num_ray_points", C.c_int32),
        ("global_obs", C.c_int32),
        ("root_height_obs", C.c_int32),
        ("enable_tar_obs", C.c_int32),
        ("use_contact_info", C.c_int32),
        ("track_root", C.c_int32),
        ("track_root_h", C.c_int32),
        ("enable_early_termination", C.c_int32),
        ("pose_termination", C.c_int32),
        ("num_fall_contact_bodies", C.c_int32),
        ("fall_contact_body_mask", C.c_uint32),
        ("never_done", C.c_int32),
        ("timestep", C.c_float),
        ("episode_length", C.c_float),
        ("termination_height", C.c_float),
        ("root_pos_termination_dist", C.c_float),
        ("root_rot_termination_angle", C.c_float),
        ("pose_termination_dist", C.c_float * MAX_BODIES),
        ("joint_err_w", C.c_float * MAX_BODIES),
        ("dof_err_w", C.c_float * MAX_DOFS),
        ("contact_weights", C.c_float * MAX_BODIES),
        ("pose_w", C.c_float),
        ("vel_w", C.c_float),
        ("root_pos_w", C.c_float),
        ("root_vel_w", C.c_float),
        ("key_pos_w", C.c_float),
        ("contact_eps", C.c_float),
        ("min_obs_h", C.c_float),
        ("max_obs_h", C.c_float),
        ("ema_weight", C.c_float),
        ("ema_keep", C.c_float),
        ("min_motion_weight", C.c_float),
        ("rand_root_pos_offset_scale", C.c_float),
        ("ray_grid_nx", C.c_int32),
        ("ray_grid_ny", C.c_int32),
    ]


_VIEW_NAMES = [
    "char_root_pos", "char_root_rot", "char_root_vel", "char_root_ang_vel", "char_dof_pos",
    "char_dof_vel", "char_contact_forces", "char_rigid_body_pos", "char_rigid_body_vel",
    "char_rigid_body_ang_vel", "ref_root_pos", "ref_root_rot", "ref_root_vel", "ref_root_ang_vel",
    "ref_body_pos", "ref_joint_rot", "ref_dof_pos", "ref_dof_vel", "ref_contacts",
]


class EnvBuffersStruct(C.Structure):
    _fields_ = (
        [("num_envs", C.c_int64)]
        + [(name, View3) for name in _VIEW_NAMES]
        + [
            ("reward_buf", C.c_void_p),
            ("reward_terms", C.c_void_p),
            ("done_buf", C.c_void_p),
            ("time_buf", C.c_void_p),
            ("timestep_buf", C.c_void_p),
            ("actors_need_reset", C.c_void_p),
            ("actors_per_env", C.c_int64),
            ("env_offsets", C.c_void_p),
            ("ray_xy_points", C.c_void_p),
            ("ray_hfs", C.c_void_p),
            ("obs_buf", C.c_void_p),
            ("obs_row_stride", C.c_int64),
            ("motion_ids", C.c_void_p),
            ("motion_terrain_ids", C.c_void_p),
            ("motion_time_offsets", C.c_void_p),
            ("motion_id_fail_rates", C.c_void_p),
            ("motion_weights", C.c_void_p),
            ("dm_motion_offsets", C.c_void_p),
            ("terrains_per_motion", C.c_int32),
            ("done_count", C.c_void_p),
            ("done_list", C.c_void_p),
            ("done_code", C.c_void_p),
        ]
    )


PARC_NUM_LOSS_TERMS = 9


class OptConfigStruct(C.Structure):
    """parc_opt_config_t (include/parc_b200_opt.h)."""
    _fields_ = [
        ("w", C.c_float * PARC_NUM_LOSS_TERMS),
        ("max_jerk", C.c_float),
        ("base_z", C.c_float),
        ("num_points", C.c_int32),
        ("point_start", C.c_int32 * (MAX_BODIES + 1)),
        ("contact_col", C.c_int32 * MAX_BODIES),
        ("num_contact_cols", C.c_int32),
        ("whole_terrain", C.c_int32),
    ]


GEOM_NONE, GEOM_SPHERE, GEOM_BOX = 0, 1, 2


class OptConstraintsStruct(C.Structure):
    """parc_opt_constraints_t (include/parc_b200_opt.h)."""
    _fields_ = [
        ("num", C.c_int32),
        ("motion_start", C.c_void_p),
        ("body", C.c_void_p),
        ("start_frame", C.c_void_p),
        ("end_frame", C.c_void_p),
        ("point", C.c_void_p),
        ("geom_kind", C.c_int32 * MAX_BODIES),
        ("geom_radius", C.c_float * MAX_BODIES),
        ("geom_offset", (C.c_float * 3) * MAX_BODIES),
        ("sole_points", C.c_int32),
    ]


class ParcB200Error(RuntimeError):
    pass


_lock = threading.Lock()
_lib = None


def _declare(lib):
    vp, i64, i32, f32, u32, u64, f64 = C.c_void_p, C.c_int64, C.c_int32, C.c_float, C.c_uint32, C.c_uint64, C.c_double
    P = C.POINTER
    lib.parc_b200_version.restype = C.c_char_p
    lib.parc_b200_version.argtypes = []
    lib.parc_b200_last_error.restype = C.c_char_p
    lib.parc_b200_last_error.argtypes = []
    sigs = {
        "parc_b200_forward_kinematics": [P(CharModelStruct), vp, vp, vp, i64, vp, vp, vp],
        "parc_b200_forward_kinematics_bwd": [P(CharModelStruct), vp, vp, vp, i64, vp, vp, vp, vp, vp, vp],
        "parc_b200_dof_to_rot": [P(CharModelStruct), vp, i64, vp, vp],
        "parc_b200_rot_to_dof": [P(CharModelStruct), vp, i64, vp, vp],
        "parc_b200_frame_velocities": [P(CharModelStruct), vp, vp, vp, i64, f32, f32, vp, vp, vp, vp],
        "parc_b200_motionlib_pack": [P(CharModelStruct), i64, i32, i32] + [vp] * 15 + [vp],
        "parc_b200_calc_motion_frame": [P(CharModelStruct), P(MotionLibStruct), vp, vp, i64] + [vp] * 10 + [vp],
        "parc_b200_ray_heights": [P(TerrainStruct), vp, vp, vp, i32, i64, f32, f32, vp, i64, vp, vp],
        "parc_b200_calc_heading": [vp, i64, vp, vp],
        "parc_b200_obs_dim": [P(CharModelStruct), P(EnvConfigStruct)],
        "parc_b200_env_step": [P(CharModelStruct), P(MotionLibStruct), P(TerrainStruct), P(EnvConfigStruct),
                               P(EnvBuffersStruct), u32, vp, i64, vp],
        "parc_b200_env_step_host": [P(CharModelStruct), P(MotionLibStruct), P(TerrainStruct), P(EnvConfigStruct),
                                    P(EnvBuffersStruct), u32, P(HostRowsStruct), i32, i32, vp, vp],
        "parc_b200_compute_tar_obs_data": [P(CharModelStruct), P(MotionLibStruct), P(EnvConfigStruct),
                                           P(EnvBuffersStruct), vp, i64, vp, vp, vp, vp, vp, vp],
        "parc_b200_compute_char_obs": [P(CharModelStruct)] + [vp] * 7 + [i32, i32, i32, i64, vp, vp],
        "parc_b200_compute_tar_obs": [P(CharModelStruct)] + [vp] * 6 + [i32, i32, i32, i32, i64, vp, vp],
        "parc_b200_compute_deepmimic_reward": [P(CharModelStruct)] + [vp] * 16 + [i32, i32, i32, i64, vp, vp],
        "parc_b200_compute_contact_reward": [vp, vp, vp, i32, i64, vp, vp],
        "parc_b200_compute_done": [vp, vp, f32, vp, vp, vp, vp, vp, vp, i32, vp, i32, vp, i32, i32, f64, f32,
                                   i32, i64, vp, vp],
        "parc_b200_compute_tracking_error": [vp] * 14 + [i32, i32, i64, vp, vp],
        "parc_b200_fail_rate_ema": [P(EnvConfigStruct), P(EnvBuffersStruct), i32, vp, vp, vp],
        "parc_b200_fail_rate_ema_noalloc": [P(EnvConfigStruct), P(EnvBuffersStruct), i32, vp, vp],
        "parc_b200_env_reset": [P(CharModelStruct), P(MotionLibStruct), P(EnvConfigStruct), P(EnvBuffersStruct),
                                vp, i64, vp, vp, vp, vp, vp, vp, vp],
        "parc_b200_env_reset_sampled": [P(CharModelStruct), P(MotionLibStruct), P(TerrainStruct),
                                        P(EnvConfigStruct), P(EnvBuffersStruct), vp, i32, u64, u64, i32, P(ResetOptionsStruct), vp],
        "parc_b200_sample_motions_cdf": [vp, i32, vp, i64, vp, vp],
        "parc_b200_sim_standin": [P(CharModelStruct), P(EnvBuffersStruct), vp, vp],
        "parc_b200_points_hf_sdf": [vp, i64, i64, vp, i32, i32, vp, f32, f32, f32, i32, vp, vp, vp],
        "parc_b200_motion_loss": [P(CharModelStruct), P(OptConfigStruct), i64, i64] + [vp] * 9 + [vp, i64, i32, i32, vp, i64, f32, f32,
                                                                                                 vp, vp, i64, vp, vp, vp, vp, vp, vp],
        "parc_b200_motion_loss_constrained": [P(CharModelStruct), P(OptConfigStruct), P(OptConstraintsStruct), i64, i64] + [vp] * 9 +
                                             [vp, i64, i32, i32, vp, i64, f32, f32, vp, vp, i64, vp, vp, vp, vp, vp, vp],
        "parc_b200_selftest_div": [vp, i64, f32, vp, vp, vp],
        "parc_b200_selftest_cells": [vp, i64, f32, f32, f32, f32, f32, f32, f32, f32, i32, i32, vp, vp, vp],
        "parc_b200_record_agent_states": [P(CharModelStruct), P(MotionLibStruct), P(EnvConfigStruct), P(EnvBuffersStruct),
                                          P(RecordTapeStruct), i32, i32, f64, vp],
        "parc_b200_hf_mask_inds": [P(CharModelStruct), P(TerrainStruct), vp, vp, i64, vp, vp, i32, vp, vp, vp, vp, vp],
        "parc_b200_hf_extra_vals": [P(TerrainStruct), vp, vp, f32, f32, f32, f32, vp, vp],
        "parc_b200_library_velocities": [P(CharModelStruct), vp, vp, vp, i64, i32, vp, vp, vp, vp, vp, vp, vp, vp],
        "parc_b200_path_motion_loss": [P(CharModelStruct), P(OptConfigStruct), i64, i64, vp, vp, vp, vp, i32, i32, vp, f32, f32, vp,
                                       f32, f32, vp, i64, vp, vp, vp, vp],
        "parc_b200_mdm_sample_motion_data": [P(CharModelStruct), P(MotionLibStruct), vp, P(SamplerConfigStruct), vp, vp, i64] + [vp] * 14,
        "parc_b200_mdm_augment_hfs": [vp, P(SamplerConfigStruct), P(HfAugStruct), vp, vp, i64, vp, vp, vp, vp],
        "parc_b200_mdm_floor_heights": [P(SamplerConfigStruct), f32, vp, vp, i64, vp, vp],
        "parc_b200_outcome_counts": [P(EnvBuffersStruct), i32, vp, vp],
        "parc_b200_tree_parts": [P(CharModelStruct), P(i32), P(i32), P(i32), P(i32)],
        "parc_b200_ray_points_lattice": [P(f32), i32, P(i32), P(i32)],
    }
    for name, argtypes in sigs.items():
        fn = getattr(lib, name)   # AttributeError here = header/library mismatch: fail loudly
        fn.restype = C.c_int
        fn.argtypes = argtypes
    lib.parc_b200_path_motion_loss_workspace_bytes.restype = i64
    lib.parc_b200_path_motion_loss_workspace_bytes.argtypes = [i64, i64]
    sigs["parc_b200_path_motion_loss_workspace_bytes"] = lib.parc_b200_path_motion_loss_workspace_bytes.argtypes
    lib.parc_b200_motion_loss_workspace_bytes.restype = i64
    lib.parc_b200_motion_loss_workspace_bytes.argtypes = [P(CharModelStruct), i64, i64]
    sigs["parc_b200_motion_loss_workspace_bytes"] = lib.parc_b200_motion_loss_workspace_bytes.argtypes
    return sigs


EXPORTED_SYMBOLS = None


def lib():
    """Load (building if needed) the CUDA library. Raises if that is impossible -- no fallback."""
    global _lib, EXPORTED_SYMBOLS
    with _lock:
        if _lib is None:
            import os
            path = os.environ.get("PARC_B200_LIB") or _build.build()      # PARC_B200_LIB: a debug build of the same sources (profiles/)
            handle = C.CDLL(str(path))
            EXPORTED_SYMBOLS = sorted(_declare(handle).keys()) + ["parc_b200_version", "parc_b200_last_error"]
            _lib = handle
    return _lib


def check(rc: int):
    if rc == PARC_OK:
        return
    msg = lib().parc_b200_last_error().decode()
    if rc == PARC_ERR_INVALID_ARG:
        raise AssertionError(msg)
    if rc == PARC_ERR_UNSUPPORTED:
        raise NotImplementedError(msg)
    raise ParcB200Error(msg)


def ray_points_lattice(ray_xy_points):
    """(nx, ny) if the [P,2] points are meshgrid(xs, ys, 'ij') bit for bit (geom_util.get_xy_grid_points), else (0, 0)."""
    import numpy as np
    pts = np.ascontiguousarray(ray_xy_points.detach().cpu().numpy(), dtype=np.float32).reshape(-1, 2)
    nx, ny = C.c_int32(), C.c_int32()
    check(lib().parc_b200_ray_points_lattice(pts.ctypes.data_as(C.POINTER(C.c_float)), int(pts.shape[0]), C.byref(nx), C.byref(ny)))
    return nx.value, ny.value


def stream_ptr(device=None) -> int:
    return torch.cuda.current_stream(device).cuda_stream


def ptr(t):
    """Device pointer of a CUDA tensor (None -> NULL). Refuses CPU tensors: there is no CPU path."""
    if t is None:
        return None
    if not t.is_cuda:
        raise ParcB200Error("parc_b200 kernels need CUDA tensors (got a %s tensor); there is no CPU fallback" % t.device)
    return t.data_ptr()


def cptr(t, dtype):
    """Pointer of a contiguous CUDA tensor of the given dtype (asserts both)."""
    if t is None:
        return None
    assert t.dtype == dtype, "expected %s, got %s" % (dtype, t.dtype)
    assert t.is_contiguous(), "tensor must be contiguous"
    return ptr(t)


def view3(t) -> View3:
    """Strided [N, A(, B)] fp32 view descriptor of a CUDA tensor (no copy)."""
    v = View3()
    if t is None:
        v.ptr, v.row_stride, v.mid_stride, v.elem_stride = None, 0, 0, 0
        return v
    assert t.dtype == torch.float32
    v.ptr = ptr(t)
    if t.dim() == 2:
        v.row_stride, v.mid_stride, v.elem_stride = t.stride(0), 0, t.stride(1)
    elif t.dim() == 3:
        v.row_stride, v.mid_stride, v.elem_stride = t.stride(0), t.stride(1), t.stride(2)
    else:
        raise AssertionError("view must be 2-D or 3-D, got %d-D" % t.dim())
    return v
