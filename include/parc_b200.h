/* parc_b200 -- C ABI of the B200-native kernels for PARC's motion-tracking hot path.
 *
 * The reference (CloudEngineHub/PARC) has no FFI layer: the seam is a set of Python method
 * signatures. Each entry point below replaces the torch-op sequence of one
 * such method and is what the Python classes in parc_b200/ (same names and signatures as the
 * reference classes) bind through ctypes. Citations are file:line into the reference.
 *
 * Conventions
 *   - every pointer argument is a DEVICE pointer unless its name starts with `h_` or it is one of
 *     the descriptor structs below (host structs holding device pointers);
 *   - all floating data is fp32, ids are int64, done / timestep / loop mode are int32;
 *   - quaternions are (x, y, z, w);
 *   - `stream` is a cudaStream_t passed as void* (0 = legacy default stream);
 *   - return value: 0 = ok, negative = parc_status_t error; parc_b200_last_error() gives the text.
 *     The Python wrappers re-raise AssertionError / ValueError like the reference does.
 *   - no entry point synchronises the device or allocates device memory.
 */
#ifndef PARC_B200_H
#define PARC_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PARC_MAX_BODIES 32
#define PARC_MAX_TAR_STEPS 16
#define PARC_MAX_KEY_BODIES 8

typedef enum {
    PARC_OK = 0,
    PARC_ERR_INVALID_ARG = -1,   /* shape / config error: AssertionError in the reference */
    PARC_ERR_UNSUPPORTED = -2,   /* valid in the reference, not built here (documented in DESIGN.md) */
    PARC_ERR_CUDA = -3           /* launch failed; text holds cudaGetErrorString */
} parc_status_t;

/* JointType, PARC/anim/kin_char_model.py:15-19 */
enum { PARC_JOINT_ROOT = 0, PARC_JOINT_HINGE = 1, PARC_JOINT_SPHERICAL = 2, PARC_JOINT_FIXED = 3 };
/* LoopMode, PARC/anim/motion_lib.py:14-16 */
enum { PARC_LOOP_CLAMP = 0, PARC_LOOP_WRAP = 1 };
/* DoneFlags, PARC/motion_tracker/envs/base_env.py:14-18 */
enum { PARC_DONE_NULL = 0, PARC_DONE_FAIL = 1, PARC_DONE_SUCC = 2, PARC_DONE_TIME = 3 };

/* Kinematic tree of KinCharModel (PARC/anim/kin_char_model.py:154-197): body 0 is the root,
 * parents precede children. Passed by value to the kernels (constant bank). */
typedef struct {
    int32_t num_bodies;                      /* J <= PARC_MAX_BODIES */
    int32_t dof_size;                        /* D = sum of joint dof dims */
    int32_t parent[PARC_MAX_BODIES];         /* _parent_indices, -1 for the root */
    int32_t joint_type[PARC_MAX_BODIES];     /* PARC_JOINT_* */
    int32_t dof_idx[PARC_MAX_BODIES];        /* Joint.dof_idx */
    float local_trans[PARC_MAX_BODIES][3];   /* _local_translation */
    float local_rot[PARC_MAX_BODIES][4];     /* _local_rotation */
    float axis[PARC_MAX_BODIES][3];          /* hinge axis (unused otherwise) */
} parc_char_model_t;

/* Device-resident motion library: the reference's flat frame tensors
 * (PARC/anim/motion_lib.py:374-396) repacked into two 16-byte-aligned frame records so that one
 * pose evaluation reads two ADJACENT records of one array.
 *   interp record (irec floats): quats[J][4] (root rotation first, then the J-1 joints) |
 *                                root_pos[3] | pad | contacts[C] | pad to even | pair[J][2] | pad to a multiple of 4
 *                                pair[j] = the blend-independent half of slerp (PARC/util/torch_util.py:474-499) between
 *                                quaternion j of this frame and of the clip's next frame (of this frame itself on the last
 *                                one): (acos|c|, sign(c) / sqrt(1 - c^2)), or (-1, w) for the reference's two special cases
 *                                (|c| >= 1: w = 0, result q0; sin < 0.001: w = +-0.5, the un-normalised midpoint)
 *   vel record (vrec floats):    root_vel[3] | root_ang_vel[3] | dof_vel[D] | pad to multiple of 4
 *   meta record (8 x 4 bytes):   int32 start_idx | int32 num_frames | float length |
 *                                int32 loop_mode | float root_pos_delta[3] | pad */
typedef struct {
    const float* interp;
    const float* vel;
    const int32_t* meta;
    int64_t total_frames;    /* F */
    int32_t num_motions;     /* M */
    int32_t irec;
    int32_t vrec;
    int32_t num_contacts;    /* C (0 when the library was built with contact_info=False) */
} parc_motion_lib_t;

/* SubTerrain, PARC/util/terrain_util.py:18-31: hf is [dim_x, dim_y] row-major. */
typedef struct {
    const float* hf;
    int32_t dim_x, dim_y;
    float min_x, min_y;
    float dx, dy;
} parc_terrain_t;

/* A [N, width] fp32 view with explicit strides in ELEMENTS: Isaac Gym hands the env column slices
 * of its [N*actors, 13] root state and stride-2 slices of its [N*dofs, 2] dof state
 * (PARC/motion_tracker/envs/ig_char_env.py:172-196); the kernels read and write through them in
 * place instead of copying. */
typedef struct {
    float* ptr;
    int64_t row_stride;
    int64_t elem_stride;
} parc_view_t;

/* ------------------------------------------------------------------------------------------
 * Library identity / errors
 * ---------------------------------------------------------------------------------------- */
const char* parc_b200_version(void);
const char* parc_b200_last_error(void);

/* ------------------------------------------------------------------------------------------
 * KinCharModel
 * ---------------------------------------------------------------------------------------- */

/* KinCharModel.forward_kinematics (PARC/anim/kin_char_model.py:617-649).
 * root_pos [n,3], root_rot [n,4], joint_rot [n,J-1,4] -> body_pos [n,J,3], body_rot [n,J,4]
 * (either output may be NULL). */
int parc_b200_forward_kinematics(const parc_char_model_t* h_model, const float* root_pos,
                                 const float* root_rot, const float* joint_rot, int64_t n,
                                 float* body_pos, float* body_rot, void* stream);

/* Vector-Jacobian product of forward_kinematics (the reference relies on torch autograd through
 * kin_char_model.py:617-649, e.g. PARC/motion_synthesis/motion_opt/motion_optimization.py:463).
 * g_body_pos / g_body_rot may be NULL (treated as zero). */
int parc_b200_forward_kinematics_bwd(const parc_char_model_t* h_model, const float* root_pos,
                                     const float* root_rot, const float* joint_rot, int64_t n,
                                     const float* g_body_pos, const float* g_body_rot,
                                     float* g_root_pos, float* g_root_rot, float* g_joint_rot,
                                     void* stream);

/* KinCharModel.dof_to_rot (kin_char_model.py:586-599, Joint.dof_to_rot :61-81): dof [n,D] -> rot [n,J-1,4] */
int parc_b200_dof_to_rot(const parc_char_model_t* h_model, const float* dof, int64_t n, float* rot,
                         void* stream);
/* KinCharModel.rot_to_dof (kin_char_model.py:601-615, Joint.rot_to_dof :83-104): rot [n,J-1,4] -> dof [n,D] */
int parc_b200_rot_to_dof(const parc_char_model_t* h_model, const float* rot, int64_t n, float* dof,
                         void* stream);

/* KinCharModel.compute_frame_dof_vel + the root velocity finite differences of
 * MotionLib._load_motion_file (kin_char_model.py:651-690, motion_lib.py:318-327) for ONE clip of
 * nf frames: root_vel/root_ang_vel [nf,3], dof_vel [nf,D]; the last frame repeats the previous. */
int parc_b200_frame_velocities(const parc_char_model_t* h_model, const float* root_pos,
                               const float* root_rot, const float* joint_rot, int64_t nf, float fps,
                               float dt, float* root_vel, float* root_ang_vel, float* dof_vel,
                               void* stream);

/* The same for ALL clips of a library in one launch (the per-file loop of MotionLib._load_motion_file,
 * motion_lib.py:285-372): the frames of the clips are concatenated
 * (total_frames rows); clip c owns rows [clip_start[c], clip_start[c] + clip_frames[c]) and has its own
 * fps / dt (= 1/fps as the host computed it). Bit-identical to one parc_b200_frame_velocities per clip. */
int parc_b200_library_velocities(const parc_char_model_t* h_model, const float* root_pos, const float* root_rot,
                                 const float* joint_rot, int64_t total_frames, int32_t num_clips,
                                 const int64_t* clip_start, const int64_t* clip_frames, const float* clip_fps,
                                 const float* clip_dt, float* root_vel, float* root_ang_vel, float* dof_vel,
                                 void* stream);

/* ------------------------------------------------------------------------------------------
 * MotionLib
 * ---------------------------------------------------------------------------------------- */

/* Repack the reference's flat tensors (motion_lib.py:374-396) into the records described at
 * parc_motion_lib_t. contacts may be NULL (num_contacts 0). Output buffers are caller-allocated:
 * interp [F, irec], vel [F, vrec], meta [M, 8] int32. */
int parc_b200_motionlib_pack(const parc_char_model_t* h_model, int64_t total_frames,
                             int32_t num_motions, int32_t num_contacts, const float* frame_root_pos,
                             const float* frame_root_rot, const float* frame_root_vel,
                             const float* frame_root_ang_vel, const float* frame_joint_rot,
                             const float* frame_dof_vel, const float* frame_contacts,
                             const int64_t* motion_num_frames, const int64_t* motion_start_idx,
                             const float* motion_lengths, const int32_t* motion_loop_modes,
                             const float* motion_root_pos_delta, float* interp, float* vel,
                             int32_t* meta, void* stream);

/* MotionLib.calc_motion_frame (motion_lib.py:94-126) with _calc_frame_blend (:425-438),
 * calc_phase (:519-530) and _calc_loop_offset (:440-457).
 * motion_ids [n] i64, motion_times [n] f32 -> root_pos [n,3], root_rot [n,4], root_vel [n,3],
 * root_ang_vel [n,3], joint_rot [n,J-1,4], dof_vel [n,D], contacts [n,C] (NULL if C == 0).
 * Optional (may be NULL) frame_idx0/1 [n] i64 and blend [n] f32 expose _calc_frame_blend for the
 * bit-exact index tests. */
int parc_b200_calc_motion_frame(const parc_char_model_t* h_model, const parc_motion_lib_t* h_lib,
                                const int64_t* motion_ids, const float* motion_times, int64_t n,
                                float* root_pos, float* root_rot, float* root_vel,
                                float* root_ang_vel, float* joint_rot, float* dof_vel,
                                float* contacts, int64_t* frame_idx0, int64_t* frame_idx1,
                                float* blend, void* stream);

/* ------------------------------------------------------------------------------------------
 * Terrain
 * ---------------------------------------------------------------------------------------- */

/* RefCharEnv._refresh_ray_obs_hfs (PARC/motion_tracker/envs/ig_parkour/mgdm_dm_util.py:128-145)
 * with SubTerrain.get_hf_val_from_points (PARC/util/terrain_util.py:143-156).
 * root_xyz [n,3] (global), heading [n], ray_xy [P,2] -> ray_hfs [n,P] written with row stride
 * out_row_stride (elements); optional grid_idx [n,P,2] i64 (NULL to skip). */
int parc_b200_ray_heights(const parc_terrain_t* h_terrain, const float* root_xyz,
                          const float* heading, const float* ray_xy, int32_t num_points, int64_t n,
                          float min_obs_h, float max_obs_h, float* ray_hfs, int64_t out_row_stride,
                          int64_t* grid_idx, void* stream);

/* torch_util.calc_heading (PARC/util/torch_util.py:501-510): root_rot [n,4] -> heading [n] */
int parc_b200_calc_heading(const float* root_rot, int64_t n, float* heading, void* stream);

#ifdef __cplusplus
}
#endif

#include "parc_b200_env.h"

#endif /* PARC_B200_H */
