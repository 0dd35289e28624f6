/* parc-b200: C ABI of the MDM dataset sampler -- the second caller of the
 * motion-evaluation / forward-kinematics / local-heightmap kernels, at batch num_samples x seq_len.
 *
 * Replaces the body of MDMHeightfieldContactMotionSampler.sample_motion_data
 * (PARC/motion_generator/mdm_heightfield_contact_motion_sampler.py:197-302):
 *   _extract_motion_data (:162-190)  -> MotionLib.calc_motion_frame on num_samples x seq_len times
 *   get_hfs_from_data (:366-401) + get_hfs_from_data_helper (:332-364, height lookup) -> local heightmaps
 *   canonicalisation by the heading of the reference frame (:234-239), KinCharModel.forward_kinematics (:241)
 *   _sample_target_info (:146-160) -> future root position / rotation in the canonical frame
 * The random draws (motion ids, start times, future times, position noise) are inputs: the host side makes
 * them in the reference's order. The heightfield augmenter (:304-330, Python `random` + per-sample loops)
 * is outside this path.
 *
 * Same conventions as parc_b200.h: device pointers, fp32, quaternions (x,y,z,w), `stream` = cudaStream_t,
 * return 0 or PARC_ERR_*.
 */
#ifndef PARC_B200_SAMPLER_H
#define PARC_B200_SAMPLER_H

#include "parc_b200.h"

#ifdef __cplusplus
extern "C" {
#endif

enum { PARC_RELATIVE_TO_ROOT = 0, PARC_RELATIVE_TO_ROOT_FLOOR = 1 };   /* RelativeZStyle, diffusion_util.py:9-11 */

typedef struct parc_sampler_config_t {
    int32_t seq_len;            /* T = len(arange(0, sequence_duration, 1/fps)), :50-52 */
    int32_t ref_frame_idx;      /* num_prev_states - 1, :99 */
    int32_t relative_z_style;   /* PARC_RELATIVE_TO_* , :61 */
    int32_t grid_dim_x;         /* num_x_neg + 1 + num_x_pos, :77 */
    int32_t grid_dim_y;
    int32_t center_ix;          /* num_x_neg: the grid cell under the root, :378 */
    int32_t center_iy;          /* num_y_neg */
    float min_h;                /* -max_h, :86-87 */
    float max_h;
    float future_pos_noise_scale;   /* :59 */
} parc_sampler_config_t;

/* sample_motion_data for n samples.
 *   terrains        [num_motions] device array of parc_terrain_t: MotionLib._terrains[motion id] (motion_lib.py:355-360)
 *   motion_ids      [n] i64, start_times [n]
 *   motion_times    [T] the sampler's _motion_times (:50-51), grid_xy [gx*gy*2] its _generic_heightmap (:83)
 *   future_times    [n] or NULL (no target info); future_noise [n,3] standard-normal draws or NULL
 * Outputs (any may be NULL except as noted):
 *   root_pos [n,T,3], root_rot [n,T,4], joint_pos [n,T,J-1,3], joint_rot [n,T,J-1,4], contacts [n,T,C]
 *   hfs [n,gx,gy] (NULL = ret_hf_obs False: then no floor-relative z shift is applied either, as upstream :226-232)
 *   future_pos [n,3], future_rot [n,4] (required when future_times is given)
 *   canon [n,12] scratch, REQUIRED: canon_root_pos[3] | canon_root_rot[4] | heading_quat_inv[4] | center_h */
int parc_b200_mdm_sample_motion_data(const parc_char_model_t* h_model, const parc_motion_lib_t* h_lib,
                                     const parc_terrain_t* terrains, const parc_sampler_config_t* h_cfg,
                                     const int64_t* motion_ids, const float* start_times, int64_t n,
                                     const float* motion_times, const float* grid_xy,
                                     const float* future_times, const float* future_noise,
                                     float* root_pos, float* root_rot, float* joint_pos, float* joint_rot,
                                     float* contacts, float* hfs, float* future_pos, float* future_rot,
                                     float* canon, void* stream);

#ifdef __cplusplus
}
#endif
#endif
