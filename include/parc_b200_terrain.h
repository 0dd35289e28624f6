/* parc-b200: C ABI of the motion-aware terrain post-processing.
 *
 * Replaces terrain_util.compute_hf_mask_inds (PARC/util/terrain_util.py:1851-1897: a Python loop over frames x
 * bodies x points), compute_hf_mask_from_inds (:1899-1906) and the grid arithmetic of compute_hf_extra_vals
 * (:1918-1947), which stage 2 runs on every generated motion (scripts/parc_2_kin_gen.py:472-475) to tell the
 * heightfield augmenter which cells a motion touches.
 * Conventions as in parc_b200.h. */
#ifndef PARC_B200_TERRAIN_H
#define PARC_B200_TERRAIN_H

#include "parc_b200.h"

#ifdef __cplusplus
extern "C" {
#endif

/* For T frames with forward-kinematics outputs body_pos [T,J,3], body_rot [T,J,4] and the body sample points
 * body_points [Np,3] (body b owns points [point_start[b], point_start[b+1]), Np <= 2048):
 *   inds   [T, Np, 2] i64: frame t's UNIQUE grid cells (SubTerrain.get_grid_index of the points' xy), in the
 *          lexicographic order torch.unique(dim=0) gives, rows [0, counts[t]) valid
 *   counts [T] i32
 *   min_body_heights [X,Y]: lowest sample-point z seen over each cell, 99999.9999 where none (:1873-1896)
 *   hf_mask [X,Y] u8 (may be NULL): union of all frames' cells = compute_hf_mask_from_inds of the result */
int parc_b200_hf_mask_inds(const parc_char_model_t* h_model, const parc_terrain_t* h_terrain, const float* body_pos,
                           const float* body_rot, int64_t T, const float* body_points, const int32_t* point_start,
                           int32_t num_points, int64_t* inds, int32_t* counts, float* min_body_heights,
                           uint8_t* hf_mask, void* stream);

/* compute_hf_extra_vals (:1918-1947) given the mask and min_body_heights: hf_maxmin [X,Y,2] =
 * (max_h + z_buf, min_h - z_buf) everywhere, (hf, hf) on masked cells, and (min_body_height - jump_buf,
 * min_h - z_buf) on masked cells the character passes at least jump_buf above. max_h = highest root z of the
 * motion, min_h = lowest cell of the terrain (the caller's two reductions). */
int parc_b200_hf_extra_vals(const parc_terrain_t* h_terrain, const uint8_t* hf_mask, const float* min_body_heights,
                            float max_h, float min_h, float z_buf, float jump_buf, float* hf_maxmin, void* stream);

#ifdef __cplusplus
}
#endif
#endif
