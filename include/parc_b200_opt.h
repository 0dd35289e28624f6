/* parc_b200_opt.h -- C ABI of the kinematic motion optimizer's loss.
 *
 * Replaces, for the FK + terrain penetration / contact loss of PARC's motion_synthesis optimizer:
 *   terrain_util.points_hf_sdf                      PARC/util/terrain_util.py:1736-1794
 *   terrain_util.points_boxes_sdf / geom_util.sdBox PARC/util/terrain_util.py:1678-1705, PARC/util/geom_util.py:124-145
 *   motion_terrain_contact_loss_localized           PARC/motion_synthesis/motion_opt/motion_optimization.py:426-656
 *   (and the torch autograd pass the optimizer runs through it, motion_optimization.py:772-797)
 *
 * Conventions as in parc_b200.h: device pointers, row-major fp32, quaternions xyzw, `stream` is a
 * cudaStream_t, return 0 = ok / negative PARC_ERR_*. */
#ifndef PARC_B200_OPT_H
#define PARC_B200_OPT_H

#include "parc_b200.h"

#ifdef __cplusplus
extern "C" {
#endif

/* LossType order of motion_optimization.py:15-25 (first nine members). */
enum {
    PARC_LOSS_ROOT_POS = 0,
    PARC_LOSS_ROOT_ROT = 1,
    PARC_LOSS_JOINT_ROT = 2,
    PARC_LOSS_SMOOTHNESS = 3,
    PARC_LOSS_PENETRATION = 4,
    PARC_LOSS_CONTACT = 5,
    PARC_LOSS_SLIDING = 6,
    PARC_LOSS_BODY_CONSTRAINT = 7, /* 0 unless parc_b200_motion_loss_constrained is given constraint records */
    PARC_LOSS_JERK = 8,
    PARC_NUM_LOSS_TERMS = 9
};

typedef struct {
    /* w_root_pos, w_root_rot, w_joint_rot, w_smoothness, w_penetration, w_contact, w_sliding,
     * w_body_constraints, w_jerk (indexed by the enum above) */
    float w[PARC_NUM_LOSS_TERMS];
    float max_jerk;                                 /* multiplied by (1/30)^3 inside, :621-623 */
    float base_z;                                   /* -10.0 in the reference (:553, :559) */
    int32_t num_points;                             /* Np = sum of the per-body sample counts */
    int32_t point_start[PARC_MAX_BODIES + 1];       /* body b owns points [point_start[b], point_start[b+1]) */
    int32_t contact_col[PARC_MAX_BODIES];           /* column of `contacts` for body b, or -1 (:575-579) */
    int32_t num_contact_cols;                       /* trailing dim of `contacts`; sliding needs == num_bodies (:608) */
    int32_t whole_terrain;                          /* 0: motion_terrain_contact_loss_localized, per-frame windows (:507-551);
                                                     * 1: motion_terrain_contact_loss (:207-423), every point against every cell */
} parc_opt_config_t;

/* terrain_util.points_hf_sdf without `radius`: points [B,N,3], hf [B,X,Y], min_box_center [B,2]
 * -> sd [B,N]. One axis-aligned box per heightfield cell (solid column from base_z up to the cell
 * height, or, inverted, the air column from the cell height up to -base_z with the result negated),
 * minimum over all X*Y cells. grad (may be NULL) receives d sd / d point [B,N,3] -- what autograd
 * would propagate through the argmin cell. */
int parc_b200_points_hf_sdf(const float* points, int64_t B, int64_t N, const float* hf, int32_t X, int32_t Y,
                            const float* min_box_center, float dx, float dy, float base_z, int32_t inverted,
                            float* sd, float* grad, void* stream);

/* Bytes of scratch parc_b200_motion_loss needs for B motions of T frames. */
int64_t parc_b200_motion_loss_workspace_bytes(const parc_char_model_t* h_model, int64_t B, int64_t T);

/* motion_terrain_contact_loss_localized for B independent motions of T frames each (the reference
 * takes one motion per call; BASELINE config 4 is 1,024 x 300), value and gradient in one pass:
 *   tgt_root_pos [B,T,3], tgt_root_rot [B,T,3] (exp map), tgt_joint_dof [B,T,D]      -- the parameters
 *   src_root_pos [B,T,3], src_root_rot_quat [B,T,4], src_joint_rot [B,T,J-1,4],
 *   src_body_vels [B,T-1,J,3], src_body_rot_vels [B,T-1,J], contacts [B,T,C]
 *   terrain: hf [B,X,Y] with hf_motion_stride = X*Y, or one shared [X,Y] with stride 0;
 *            min_point [B,2] / [2] likewise (min_point_motion_stride 2 or 0)
 *   body_points [Np,3] in body order (geom_util.get_char_point_samples, geom_util.py:790-872)
 * Outputs: terms [B,9] unweighted loss terms, loss [B] weighted total; if the g_* pointers are
 * non-NULL, d loss / d parameter in g_root_pos [B,T,3], g_root_rot [B,T,3], g_joint_dof [B,T,D]
 * (all three or none). Heightfield windows follow the reference: per frame the cells covering the
 * frame's body points +- 2 cells, all frames of a motion sharing the largest window (:507-551). */
int parc_b200_motion_loss(const parc_char_model_t* h_model, const parc_opt_config_t* h_cfg, int64_t B, int64_t T,
                          const float* tgt_root_pos, const float* tgt_root_rot, const float* tgt_joint_dof,
                          const float* src_root_pos, const float* src_root_rot_quat, const float* src_joint_rot,
                          const float* src_body_vels, const float* src_body_rot_vels, const float* contacts,
                          const float* hf, int64_t hf_motion_stride, int32_t X, int32_t Y,
                          const float* min_point, int64_t min_point_motion_stride, float dx, float dy,
                          const float* body_points, void* workspace, int64_t workspace_bytes,
                          float* terms, float* loss, float* g_root_pos, float* g_root_rot, float* g_joint_dof,
                          void* stream);

/* body_constraints of motion_terrain_contact_loss_localized (motion_optimization.py:581-609), flattened. The reference
 * passes, per motion, one Python list of BodyConstraint{start_frame_idx, end_frame_idx, constraint_point} per body
 * (:28-31, built by compute_approx_body_constraints :70-205); here every constraint is one record, the records of
 * motion b being [motion_start[b], motion_start[b+1]). What a record does depends on the FIRST geom of its body:
 *   sphere (:590-596)  term += sum over frames start..end of | |point - (rot * geom_offset + pos)| - radius |
 *   box    (:598-602)  term += sum over frames start..end and the body's first `sole_points` sample points P of
 *                      max(|point - P| - 1.25 * |box dims|, 0)
 *   other              the record is ignored
 * and, for sphere and box, the body's velocity / angular-velocity errors of velocity indices start..end drop out of the
 * SLIDING term (:606-609; the smoothness term was summed before and keeps them). */
enum { PARC_GEOM_NONE = 0, PARC_GEOM_SPHERE = 1, PARC_GEOM_BOX = 2 };
typedef struct {
    int32_t num;                                   /* records over all motions */
    const int32_t* motion_start;                   /* [B+1] device */
    const int32_t* body;                           /* [num] device */
    const int32_t* start_frame;                    /* [num] device */
    const int32_t* end_frame;                      /* [num] device, inclusive; frames past T-1 do not exist (Python slicing) */
    const float* point;                            /* [num,3] device */
    int32_t geom_kind[PARC_MAX_BODIES];            /* PARC_GEOM_* of body b's first geom */
    float geom_radius[PARC_MAX_BODIES];            /* sphere: its radius; box: 1.25 * |dims| (:599) */
    float geom_offset[PARC_MAX_BODIES][3];         /* sphere: the geom's offset in the body frame */
    int32_t sole_points;                           /* 18 in the reference (:600) */
} parc_opt_constraints_t;

/* parc_b200_motion_loss with body constraints (h_constraints may be NULL = none; then identical to it). Fills
 * terms[:, PARC_LOSS_BODY_CONSTRAINT], adds w[PARC_LOSS_BODY_CONSTRAINT] x term to the loss and its gradient to the g_*. */
int parc_b200_motion_loss_constrained(const parc_char_model_t* h_model, const parc_opt_config_t* h_cfg,
                                      const parc_opt_constraints_t* h_constraints, int64_t B, int64_t T,
                                      const float* tgt_root_pos, const float* tgt_root_rot, const float* tgt_joint_dof,
                                      const float* src_root_pos, const float* src_root_rot_quat, const float* src_joint_rot,
                                      const float* src_body_vels, const float* src_body_rot_vels, const float* contacts,
                                      const float* hf, int64_t hf_motion_stride, int32_t X, int32_t Y,
                                      const float* min_point, int64_t min_point_motion_stride, float dx, float dy,
                                      const float* body_points, void* workspace, int64_t workspace_bytes,
                                      float* terms, float* loss, float* g_root_pos, float* g_root_rot, float* g_joint_dof,
                                      void* stream);

/* mdm_path.compute_motion_loss (PARC/motion_synthesis/procgen/mdm_path.py:31-127): the
 * forward-only penetration / contact score that ranks a batch of generated motions, against ONE terrain taken whole
 * (no per-frame window). Inputs are the forward-kinematics outputs of the batch (KinCharModel.forward_kinematics, :76):
 *   body_pos [B,T,J,3], body_rot [B,T,J,4], contacts [B,T,C] (column h_cfg->contact_col[b] labels body b; upstream
 *   uses column b for every body, :109), hf [X,Y], min_point [2], body_points [Np,3] in body order.
 * h_cfg supplies point_start / contact_col / num_points / num_contact_cols and base_z (upstream: min(hf) - 10, :78);
 * its weights are ignored in favour of w_pen / w_contact. Outputs (any may be NULL): pen_loss [B] = w_pen * sum over
 * frames and points of max(air-column sdf, 0); contact_loss [B] = w_contact * sum over frames and bodies of the
 * body's closest solid-column distance (>= 0) times its contact label; total_loss = their sum. */
int64_t parc_b200_path_motion_loss_workspace_bytes(int64_t B, int64_t T);
int parc_b200_path_motion_loss(const parc_char_model_t* h_model, const parc_opt_config_t* h_cfg, int64_t B, int64_t T,
                               const float* body_pos, const float* body_rot, const float* contacts,
                               const float* hf, int32_t X, int32_t Y, const float* min_point, float dx, float dy,
                               const float* body_points, float w_pen, float w_contact,
                               void* workspace, int64_t workspace_bytes,
                               float* pen_loss, float* contact_loss, float* total_loss, void* stream);

#ifdef __cplusplus
}
#endif
#endif
