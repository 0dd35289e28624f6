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
 * them in the reference's order. The heightfield augmenter (:304-364, :390-400) is a third launch,
 * parc_b200_mdm_augment_hfs, fed with the draws of Python's `random` / torch's generator as per-window records;
 * the FLOOR_HEIGHTS frame component (:267-278) is parc_b200_mdm_floor_heights.
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

/* ---- heightfield augmenter (use_hf_augmentation, :390-400) -------------------------------------------------------
 * Per window, on its gx x gy local heightmap (after the relative-z shift, before the [min_h, max_h] clamp of :228):
 *   bounds   get_hfs_from_data_helper (:332-364): every grid point gets (max, min) = the terrain's hf_maxmin of its cell
 *            if some body passes over that cell during the window's frames (MotionLib._hf_mask_inds), else
 *            (2 max_h, 2 min_h); shifted like the heights.
 *   NOISE    _noise_hf_augmentation (:327-330): clamp(u (max_h - min_h) + min_h, min = bound max, max = bound min) with
 *            torch.clamp's rule that max wins -- as upstream, whose two bounds arrive swapped.
 *   MAXPOOL_AND_BOXES  _box_hf_augmentation (:304-325): optional constant height, up to three max-pools (2-D / along x /
 *            along y, window 2 size + 1, each followed by a clamp into the bounds; terrain_util.py:1509-1536), then
 *            rotated boxes of constant height and a final clamp (add_boxes_to_hf2, terrain_util.py:861-914).
 * The draws are inputs, one record per window:
 *   rec_i [n, 8] int32   change_height | (pool kind, pool size) x 3 in applied order, kind -1 = skipped | num_boxes
 *   rec_f [n, 1 + 7 max_boxes] fp32   new_height | per box: centre x, y (cells) | side x, y | cos, sin(angle) | height
 *   noise [n, gx, gy] uniform [0,1) (NOISE mode) */
enum { PARC_HF_AUG_NOISE = 0, PARC_HF_AUG_MAXPOOL_AND_BOXES = 1, PARC_HF_AUG_NONE = 2 };   /* HFAugmentationMode, :19-22 */
enum { PARC_POOL_2D = 0, PARC_POOL_X = 1, PARC_POOL_Y = 2 };
#define PARC_AUG_REC_INTS 8
#define PARC_AUG_BOX_FLOATS 7

typedef struct parc_hf_aug_t {
    int32_t mode;                     /* PARC_HF_AUG_* */
    int32_t max_boxes;                /* boxes per record of rec_f */
    int32_t max_terrain_cells;        /* largest dim_x * dim_y over the library's terrains (sizes the cell bitmap) */
    float timestep;                   /* 1 / sequence_fps: frame index of a window = round(start_time / timestep), :213 */
    const float* hf_maxmin;           /* device: every motion's terrain hf_maxmin [X,Y,2] = (max, min), concatenated */
    const int64_t* maxmin_offset;     /* device [M]: float offset of motion m's block */
    const int32_t* mask_frame_off;    /* device [M+1]: motion m owns rows [off[m], off[m+1]) of mask_start, one per frame */
    const int32_t* mask_start;        /* device [rows+1]: CSR over mask_cells */
    const int32_t* mask_cells;        /* device: ix * dim_y + iy of every masked cell (MotionLib._hf_mask_inds) */
    const int32_t* rec_i;             /* device, MAXPOOL_AND_BOXES */
    const float* rec_f;
    const float* noise;               /* device, NOISE */
} parc_hf_aug_t;

/* hfs [n,gx,gy] is overwritten with the augmented, clamped heightmaps. `canon` is the scratch
 * parc_b200_mdm_sample_motion_data filled for the same windows (root pose of the reference frame + centre height). */
int parc_b200_mdm_augment_hfs(const parc_terrain_t* terrains, const parc_sampler_config_t* h_cfg, const parc_hf_aug_t* h_aug,
                              const int64_t* motion_ids, const float* start_times, int64_t n, const float* grid_xy,
                              const float* canon, float* hfs, void* stream);

/* FLOOR_HEIGHTS (:267-278): out[i, j, t] = hfs[i][cell under root_pos[j, t]] for ALL pairs of windows (i, j) -- upstream
 * indexes the whole batch inside its per-window loop, giving [n, n, T, 1]. cell = clamp(round((xy - grid_min) / dx)),
 * grid_min = (-num_x_neg, -num_y_neg) * dx (:72). root_pos [n,T,3] canonical, hfs [n,gx,gy], out [n,n,T]. */
int parc_b200_mdm_floor_heights(const parc_sampler_config_t* h_cfg, float dx, const float* hfs, const float* root_pos,
                                int64_t n, float* out, void* stream);

#ifdef __cplusplus
}
#endif
#endif
