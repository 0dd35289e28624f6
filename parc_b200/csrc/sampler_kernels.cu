// MDM dataset sampler: MDMHeightfieldContactMotionSampler.sample_motion_data
// (PARC/motion_generator/mdm_heightfield_contact_motion_sampler.py:197-302) in two launches.
//   mdm_canon_hf_kernel   warp per sample: the reference frame's root pose (canonical frame), its heading,
//                         the local heightmap looked up in the motion's own terrain, the target info
//   mdm_frames_kernel<G>  G lanes per (sample, frame): frame evaluation, canonicalisation, forward kinematics
// Both reuse the lane building blocks of the tracking path (parc_lanes.cuh). HBM-bound: per sample the output is
// T x (7 + 7 (J-1) + C) floats + gx*gy heights (7.6 KB at T=15, 31x31) against two 320 B record reads per frame.
#include "parc_common.cuh"
#include "parc_lanes.cuh"
#include "../../include/parc_b200_sampler.h"

namespace parc {

constexpr int CANON_W = 12;   // canon_root_pos[3] | canon_root_rot[4] | heading_quat_inv[4] | center_h

// Root pose of MotionLib.calc_motion_frame (motion_lib.py:94-126) alone: what _extract_root_pos_and_rot (:192-195) keeps.
__device__ __forceinline__ void root_pose_at(const parc_motion_lib_t& lib, int J, long long mid, float t, Vec3& pos, Quat& rot) {
    const MotionMeta mm = load_motion_meta(lib.meta, mid);
    const FrameBlend fb = calc_frame_blend(mm, t);
    const float* r0 = lib.interp + fb.idx0 * lib.irec;
    const float* r1 = lib.interp + fb.idx1 * lib.irec;
    const float b = fb.blend, omb = __fsub_rn(1.0f, b);
    rot = slerp(quat_from_f4(__ldg(reinterpret_cast<const float4*>(r0))), quat_from_f4(__ldg(reinterpret_cast<const float4*>(r1))), b);
    const float4 p0 = __ldg(reinterpret_cast<const float4*>(r0 + interp_pos_off(J)));
    const float4 p1 = __ldg(reinterpret_cast<const float4*>(r1 + interp_pos_off(J)));
    pos = make_vec3(__fadd_rn(lerp_exact(omb, p0.x, b, p1.x), fb.loop_offset.x), __fadd_rn(lerp_exact(omb, p0.y, b, p1.y), fb.loop_offset.y),
                    __fadd_rn(lerp_exact(omb, p0.z, b, p1.z), fb.loop_offset.z));
}

// One WARP per sample: lane 0 evaluates the canonical frame (and the target info) and broadcasts the five numbers the
// height lookups need; the 32 lanes then stride over the gx*gy grid points (coalesced stores into hfs[i]).
__global__ void __launch_bounds__(256) mdm_canon_hf_kernel(
    const __grid_constant__ parc_motion_lib_t lib, const __grid_constant__ parc_sampler_config_t cfg, const int J,
    const parc_terrain_t* __restrict__ terrains, const long long* __restrict__ motion_ids, const float* __restrict__ start_times,
    const long long n, const float* __restrict__ motion_times, const float* __restrict__ grid_xy,
    const float* __restrict__ future_times, const float* __restrict__ future_noise, float* __restrict__ hfs,
    float* __restrict__ future_pos, float* __restrict__ future_rot, float* __restrict__ canon) {
    const int lane = threadIdx.x & 31;
    const long long wpb = blockDim.x / 32;
    for (long long i = (long long)blockIdx.x * wpb + threadIdx.x / 32; i < n; i += (long long)gridDim.x * wpb) {
        const long long mid = __ldg(motion_ids + i);
        float rx = 0.f, ry = 0.f, rz = 0.f, sh = 0.f, ch = 1.f;
        if (lane == 0) {
            // canon_root_pos / canon_root_rot = frame ref_frame_idx of the sample (:222-224); times are start + _motion_times (:168)
            Vec3 cp; Quat cq;
            root_pose_at(lib, J, mid, __fadd_rn(__ldg(start_times + i), __ldg(motion_times + cfg.ref_frame_idx)), cp, cq);
            const float heading = calc_heading(cq);                                                // :370
            const Quat hinv = axis_angle_to_quat(make_vec3(0.f, 0.f, 1.f), -heading);             // calc_heading_quat_inv, :236
            float* c = canon + i * CANON_W;
            c[0] = cp.x; c[1] = cp.y; c[2] = cp.z; c[3] = cq.x; c[4] = cq.y; c[5] = cq.z; c[6] = cq.w;
            c[7] = hinv.x; c[8] = hinv.y; c[9] = hinv.z; c[10] = hinv.w;
            rx = cp.x; ry = cp.y; rz = cp.z; sh = sinf(heading); ch = cosf(heading);
            if (future_times) {
                // _sample_target_info (:146-160): root pose at the future time, noise, canonical frame
                Vec3 fp; Quat fq;
                root_pose_at(lib, J, mid, __ldg(future_times + i), fp, fq);
                if (future_noise)
                    fp = make_vec3(__fadd_rn(fp.x, __fmul_rn(cfg.future_pos_noise_scale, __ldg(future_noise + i * 3))),
                                   __fadd_rn(fp.y, __fmul_rn(cfg.future_pos_noise_scale, __ldg(future_noise + i * 3 + 1))),
                                   __fadd_rn(fp.z, __fmul_rn(cfg.future_pos_noise_scale, __ldg(future_noise + i * 3 + 2))));
                fp = quat_rotate(hinv, fp - cp);
                fq = quat_multiply(hinv, fq);
                future_pos[i * 3] = fp.x; future_pos[i * 3 + 1] = fp.y; future_pos[i * 3 + 2] = fp.z;
                reinterpret_cast<float4*>(future_rot)[i] = quat_to_f4(fq);
            }
        }
        if (hfs == nullptr) { if (lane == 0) canon[i * CANON_W + 11] = 0.f; continue; }
        rx = __shfl_sync(FULL_MASK, rx, 0); ry = __shfl_sync(FULL_MASK, ry, 0); rz = __shfl_sync(FULL_MASK, rz, 0);
        sh = __shfl_sync(FULL_MASK, sh, 0); ch = __shfl_sync(FULL_MASK, ch, 0);
        // get_hfs_from_data (:366-401): the generic grid rotated by the heading, moved to the root, looked up in the
        // MOTION'S OWN terrain (get_hfs_from_data_helper :336-339, SubTerrain.get_grid_index terrain_util.py:143-152)
        const parc_terrain_t tr = terrains[mid];
        auto height_at = [&](int p) {
            const float2 pt = __ldg(reinterpret_cast<const float2*>(grid_xy) + p);
            const float wx = __fadd_rn(__fsub_rn(__fmul_rn(pt.x, ch), __fmul_rn(pt.y, sh)), rx);   // rotate_2d_vec, torch_util.py:650-662
            const float wy = __fadd_rn(__fadd_rn(__fmul_rn(pt.x, sh), __fmul_rn(pt.y, ch)), ry);
            const int ix = grid_index(wx, tr.min_x, tr.dx, tr.dim_x), iy = grid_index(wy, tr.min_y, tr.dy, tr.dim_y);
            return __ldg(tr.hf + (long long)ix * tr.dim_y + iy);
        };
        const float center_h = height_at(cfg.center_ix * cfg.grid_dim_y + cfg.center_iy);          // :378
        const float base = cfg.relative_z_style == PARC_RELATIVE_TO_ROOT_FLOOR ? center_h : rz;     // :379-388
        if (lane == 0) canon[i * CANON_W + 11] = center_h;
        const int P = cfg.grid_dim_x * cfg.grid_dim_y;
        float* o = hfs + i * P;
#pragma unroll 4
        for (int p = lane; p < P; p += 32) o[p] = fminf(fmaxf(__fsub_rn(height_at(p), base), cfg.min_h), cfg.max_h);   // clamp :228
    }
}

template <int G>
#ifndef PARC_MDM_THREADS
#define PARC_MDM_THREADS 128
#endif
#ifndef PARC_MDM_MINB
#define PARC_MDM_MINB 8
#endif
__global__ void __launch_bounds__(PARC_MDM_THREADS, PARC_MDM_MINB) mdm_frames_kernel(
    const __grid_constant__ parc_char_model_t cm, const __grid_constant__ parc_motion_lib_t lib,
    const __grid_constant__ parc_sampler_config_t cfg, const long long* __restrict__ motion_ids, const float* __restrict__ start_times,
    const long long n, const float* __restrict__ motion_times, const float* __restrict__ canon, const int floor_shift,
    float* __restrict__ root_pos, float* __restrict__ root_rot, float* __restrict__ joint_pos, float* __restrict__ joint_rot,
    float* __restrict__ contacts) {
    const LaneBody lb = load_lane_body<G>(cm);
    const int J = cm.num_bodies, C = lib.num_contacts, T = cfg.seq_len;
    const int sub = threadIdx.x & (G - 1);
    const long long gpb = blockDim.x / G, total = n * T;
    const long long first = (long long)blockIdx.x * gpb + threadIdx.x / G;
    const long long warp_first = first - ((threadIdx.x & 31) / G);
    for (long long base = warp_first; base < total; base += (long long)gridDim.x * gpb) {
        const long long k = base + ((threadIdx.x & 31) / G);
        const bool live = k < total;
        const long long kk = live ? k : 0;
        const long long i = kk / T;
        const int f = (int)(kk - i * T);
        const long long mid = __ldg(motion_ids + i);
        const float t = __fadd_rn(__ldg(start_times + i), __ldg(motion_times + f));                // :168
        const MotionMeta mm = load_motion_meta(lib.meta, mid);
        const FrameBlend fb = calc_frame_blend(mm, t);
        const PoseSample ps = eval_pose_lane(lib, lb, fb, J);
        const float* c = canon + i * CANON_W;
        const Vec3 cp = make_vec3(__ldg(c), __ldg(c + 1), __ldg(c + 2));
        const Quat hinv = make_quat(__ldg(c + 7), __ldg(c + 8), __ldg(c + 9), __ldg(c + 10));
        Vec3 rp = ps.root_pos;
        if (floor_shift) rp.z = __fsub_rn(rp.z, __ldg(c + 11));                                    // :230-232
        rp = quat_rotate(hinv, rp - cp);                                                           // :235, :239
        const Quat rr = quat_multiply(hinv, group_bcast<G>(ps.q, 0));                              // :237
        Vec3 pos; Quat rot;
        fk_lanes<G>(lb, ps.q, rp, rr, pos, rot);                                                   // :241
        if (!live) continue;
        if (sub == 0) {
            if (root_pos) { root_pos[k * 3] = rp.x; root_pos[k * 3 + 1] = rp.y; root_pos[k * 3 + 2] = rp.z; }
            if (root_rot) reinterpret_cast<float4*>(root_rot)[k] = quat_to_f4(rr);
        } else if (lb.active) {
            if (joint_rot) reinterpret_cast<float4*>(joint_rot)[k * (J - 1) + (sub - 1)] = quat_to_f4(ps.q);
            if (joint_pos) { float* o = joint_pos + (k * (J - 1) + (sub - 1)) * 3; o[0] = pos.x; o[1] = pos.y; o[2] = pos.z; }   // body_pos[..., 1:, :], :243
        }
        if (contacts && sub < C) contacts[k * C + sub] = ps.contact;
    }
}

// ---- heightfield augmenter (:304-364, :390-400): one block per window, the gx x gy patch and its bounds in shared memory ----
// s_a / s_b ping-pong for the separable max-pools; the motion's masked terrain cells of the window's frames are OR-ed into a
// bitmap first (CSR rows of consecutive frames are contiguous, so the union is one flat range of mask_cells).
__global__ void __launch_bounds__(256) mdm_hf_augment_kernel(const __grid_constant__ parc_sampler_config_t cfg,
                                                             const __grid_constant__ parc_hf_aug_t aug,
                                                             const parc_terrain_t* __restrict__ terrains,
                                                             const long long* __restrict__ motion_ids, const float* __restrict__ start_times,
                                                             const long long n, const float* __restrict__ grid_xy,
                                                             const float* __restrict__ canon, float* __restrict__ hfs) {
    extern __shared__ float smem[];
    const int gx = cfg.grid_dim_x, gy = cfg.grid_dim_y, P = gx * gy;
    float* s_a = smem;
    float* s_b = s_a + P;
    float* s_max = s_b + P;
    float* s_min = s_max + P;
    unsigned* s_bits = reinterpret_cast<unsigned*>(s_min + P);
    const int tid = threadIdx.x, nt = blockDim.x;
    const float span = __fsub_rn(cfg.max_h, cfg.min_h);                      // (self._max_h - self._min_h)
    const float open_max = __fmul_rn(cfg.max_h, 2.0f), open_min = __fmul_rn(cfg.min_h, 2.0f);   // :354-355
    for (long long i = blockIdx.x; i < n; i += gridDim.x) {
        const long long mid = __ldg(motion_ids + i);
        const parc_terrain_t tr = terrains[mid];
        const int cells = tr.dim_x * tr.dim_y, words = (cells + 31) >> 5;
        for (int w = tid; w < words; w += nt) s_bits[w] = 0u;
        __syncthreads();
        // frames of the window: round(start / timestep) + [0, T) (:213-216), sliced like a Python list (:348-349)
        const long long f0 = (long long)rintf(__fdiv_rn(__ldg(start_times + i), aug.timestep));
        const int row0 = __ldg(aug.mask_frame_off + mid), rows = __ldg(aug.mask_frame_off + mid + 1) - row0;
        const long long lo = f0 < 0 ? 0 : (f0 > rows ? rows : f0);
        long long hi = f0 + cfg.seq_len;
        hi = hi < lo ? lo : (hi > rows ? rows : hi);
        const int c0 = __ldg(aug.mask_start + row0 + lo), c1 = __ldg(aug.mask_start + row0 + hi);
        for (int c = c0 + tid; c < c1; c += nt) {
            const int cell = __ldg(aug.mask_cells + c);
            if (cell >= 0 && cell < cells) atomicOr(&s_bits[cell >> 5], 1u << (cell & 31));
        }
        __syncthreads();
        const float* cn = canon + i * CANON_W;
        const float rx = __ldg(cn), ry = __ldg(cn + 1);
        // heading of the reference frame: once per window (the same calls as mdm_canon_hf_kernel), shared through s_trig
        __shared__ float s_trig[2];
        if (tid == 0) {
            const float heading = calc_heading(make_quat(__ldg(cn + 3), __ldg(cn + 4), __ldg(cn + 5), __ldg(cn + 6)));
            s_trig[0] = sinf(heading); s_trig[1] = cosf(heading);
        }
        __syncthreads();
        const float sh = s_trig[0], ch = s_trig[1];
        const float base = cfg.relative_z_style == PARC_RELATIVE_TO_ROOT_FLOOR ? __ldg(cn + 11) : __ldg(cn + 2);    // :379-388
        const float* mm = aug.hf_maxmin + __ldg(aug.maxmin_offset + mid);
        for (int p = tid; p < P; p += nt) {
            const float2 pt = __ldg(reinterpret_cast<const float2*>(grid_xy) + p);
            const float wx = __fadd_rn(__fsub_rn(__fmul_rn(pt.x, ch), __fmul_rn(pt.y, sh)), rx);
            const float wy = __fadd_rn(__fadd_rn(__fmul_rn(pt.x, sh), __fmul_rn(pt.y, ch)), ry);
            const int ix = grid_index(wx, tr.min_x, tr.dx, tr.dim_x), iy = grid_index(wy, tr.min_y, tr.dy, tr.dim_y);
            const int cell = ix * tr.dim_y + iy;
            const bool masked = (s_bits[cell >> 5] >> (cell & 31)) & 1u;
            s_a[p] = __fsub_rn(__ldg(tr.hf + cell), base);
            s_max[p] = __fsub_rn(masked ? __ldg(mm + 2 * (long long)cell) : open_max, base);
            s_min[p] = __fsub_rn(masked ? __ldg(mm + 2 * (long long)cell + 1) : open_min, base);
        }
        __syncthreads();
        float* cur = s_a;
        if (aug.mode == PARC_HF_AUG_NOISE) {
            const float* u = aug.noise + i * P;
            // torch.clamp(x, min = bound max, max = bound min) = min(max(x, bound max), bound min)  (:329)
            for (int p = tid; p < P; p += nt)
                s_a[p] = fminf(fmaxf(__fadd_rn(__fmul_rn(__ldg(u + p), span), cfg.min_h), s_max[p]), s_min[p]);
        } else if (aug.mode == PARC_HF_AUG_MAXPOOL_AND_BOXES) {
            const int* ri = aug.rec_i + i * PARC_AUG_REC_INTS;
            const float* rf = aug.rec_f + i * (1 + PARC_AUG_BOX_FLOATS * aug.max_boxes);
            if (__ldg(ri)) {                                                                         // :305-309
                const float h = __ldg(rf);
                for (int p = tid; p < P; p += nt) s_a[p] = h;
            }
            __syncthreads();
            for (int k = 0; k < 3; ++k) {                                                            // :322-332
                const int kind = __ldg(ri + 1 + 2 * k), size = __ldg(ri + 2 + 2 * k);
                if (kind < 0) continue;
                float* other = cur == s_a ? s_b : s_a;
                if (kind == PARC_POOL_2D || kind == PARC_POOL_X) {                                   // along x (rows of the patch)
                    const bool last = kind == PARC_POOL_X;
                    for (int p = tid; p < P; p += nt) {
                        const int x = p / gy, y = p - x * gy;
                        const int x0 = max(x - size, 0), x1 = min(x + size, gx - 1);
                        float v = cur[x0 * gy + y];
                        for (int q = x0 + 1; q <= x1; ++q) v = fmaxf(v, cur[q * gy + y]);
                        other[p] = last ? fminf(fmaxf(v, s_min[p]), s_max[p]) : v;                   // clamp(new, min, max), terrain_util.py:1513
                    }
                    __syncthreads();
                    float* t = cur; cur = other; other = t;
                }
                if (kind == PARC_POOL_2D || kind == PARC_POOL_Y) {                                   // along y
                    for (int p = tid; p < P; p += nt) {
                        const int x = p / gy, y = p - x * gy;
                        const int y0 = max(y - size, 0), y1 = min(y + size, gy - 1);
                        float v = cur[x * gy + y0];
                        for (int q = y0 + 1; q <= y1; ++q) v = fmaxf(v, cur[x * gy + q]);
                        other[p] = fminf(fmaxf(v, s_min[p]), s_max[p]);
                    }
                    __syncthreads();
                    float* t = cur; cur = other; other = t;
                }
            }
            // add_boxes_to_hf2 (terrain_util.py:861-914): cells are independent, the boxes apply in order per cell
            const int nb = min(__ldg(ri + 7), aug.max_boxes);
            for (int p = tid; p < P; p += nt) {
                const int x = p / gy, y = p - x * gy;
                float v = cur[p];
                for (int b = 0; b < nb; ++b) {
                    const float* bx = rf + 1 + PARC_AUG_BOX_FLOATS * b;
                    const float cx = __ldg(bx), cy = __ldg(bx + 1), hx = __fmul_rn(__ldg(bx + 2), 0.5f), hy = __fmul_rn(__ldg(bx + 3), 0.5f);
                    const float c = __ldg(bx + 4), s = __ldg(bx + 5);
                    const float dxv = __fsub_rn((float)x, cx), dyv = __fsub_rn((float)y, cy);
                    const float qx = __fadd_rn(__fsub_rn(__fmul_rn(dxv, c), __fmul_rn(dyv, s)), cx);  // rotate_2d_vec(xy - centre) + centre
                    const float qy = __fadd_rn(__fadd_rn(__fmul_rn(dxv, s), __fmul_rn(dyv, c)), cy);
                    if (qx < __fadd_rn(cx, hx) && qx > __fsub_rn(cx, hx) && qy < __fadd_rn(cy, hy) && qy > __fsub_rn(cy, hy)) v = __ldg(bx + 6);
                }
                cur[p] = fminf(fmaxf(v, s_min[p]), s_max[p]);                                        // terrain_util.py:911-912
            }
        }
        __syncthreads();
        float* o = hfs + i * P;
        for (int p = tid; p < P; p += nt) o[p] = fminf(fmaxf(cur[p], cfg.min_h), cfg.max_h);         // :228
        __syncthreads();
    }
}

// FLOOR_HEIGHTS (:267-278): thread per (i, j, t)
__global__ void __launch_bounds__(256) mdm_floor_heights_kernel(const __grid_constant__ parc_sampler_config_t cfg, const float dx,
                                                                const float* __restrict__ hfs, const float* __restrict__ root_pos,
                                                                const long long n, float* __restrict__ out) {
    const long long T = cfg.seq_len, total = n * n * T;
    const float gmx = __fmul_rn((float)(-cfg.center_ix), dx), gmy = __fmul_rn((float)(-cfg.center_iy), dx);   // _grid_min_point, :72
    const long long P = (long long)cfg.grid_dim_x * cfg.grid_dim_y;
    for (long long k = (long long)blockIdx.x * blockDim.x + threadIdx.x; k < total; k += (long long)gridDim.x * blockDim.x) {
        const long long i = k / (n * T), jt = k - i * (n * T);
        const float x = __ldg(root_pos + jt * 3), y = __ldg(root_pos + jt * 3 + 1);
        int ix = (int)rintf(__fdiv_rn(__fsub_rn(x, gmx), dx)), iy = (int)rintf(__fdiv_rn(__fsub_rn(y, gmy), dx));
        ix = min(max(ix, 0), cfg.grid_dim_x - 1); iy = min(max(iy, 0), cfg.grid_dim_y - 1);
        out[k] = __ldg(hfs + i * P + (long long)ix * cfg.grid_dim_y + iy);
    }
}

}  // namespace parc

using namespace parc;

extern "C" int parc_b200_mdm_augment_hfs(const parc_terrain_t* terrains, const parc_sampler_config_t* cfg, const parc_hf_aug_t* aug,
                                         const int64_t* motion_ids, const float* start_times, int64_t n, const float* grid_xy,
                                         const float* canon, float* hfs, void* stream) {
    PARC_REQUIRE(cfg && aug, "mdm_augment_hfs: NULL config");
    PARC_REQUIRE(n >= 0, "mdm_augment_hfs: negative window count");
    if (n == 0 || aug->mode == PARC_HF_AUG_NONE) return PARC_OK;
    PARC_REQUIRE(terrains && motion_ids && start_times && grid_xy && canon && hfs, "mdm_augment_hfs: NULL input");
    PARC_REQUIRE(aug->mode == PARC_HF_AUG_NOISE || aug->mode == PARC_HF_AUG_MAXPOOL_AND_BOXES, "mdm_augment_hfs: unknown mode %d", aug->mode);
    PARC_REQUIRE(cfg->grid_dim_x > 0 && cfg->grid_dim_y > 0 && cfg->seq_len > 0, "mdm_augment_hfs: bad local grid / sequence length");
    PARC_REQUIRE(aug->timestep > 0.f, "mdm_augment_hfs: timestep must be positive");
    PARC_REQUIRE(aug->hf_maxmin && aug->maxmin_offset && aug->mask_frame_off && aug->mask_start && aug->mask_cells,
                 "mdm_augment_hfs: NULL terrain bounds / mask tables");
    if (aug->mode == PARC_HF_AUG_NOISE) PARC_REQUIRE(aug->noise != nullptr, "mdm_augment_hfs: NOISE mode without noise draws");
    else PARC_REQUIRE(aug->rec_i && aug->rec_f && aug->max_boxes >= 0, "mdm_augment_hfs: MAXPOOL_AND_BOXES mode without records");
    const size_t P = (size_t)cfg->grid_dim_x * cfg->grid_dim_y;
    PARC_REQUIRE(aug->max_terrain_cells > 0, "mdm_augment_hfs: max_terrain_cells not set");
    const size_t bytes = 4 * P * sizeof(float) + (((size_t)aug->max_terrain_cells + 31) / 32) * sizeof(unsigned);
    PARC_REQUIRE(bytes <= 200 * 1024, "mdm_augment_hfs: %zu B of shared memory (local grid %zu points, terrains up to %d cells) exceed 200 KB",
                 bytes, P, aug->max_terrain_cells);
    static bool attr_set[64] = {};
    int dev = 0;
    cudaGetDevice(&dev);
    if (bytes > 48 * 1024 && dev >= 0 && dev < 64 && !attr_set[dev]) {
        cudaError_t ce = cudaFuncSetAttribute(mdm_hf_augment_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
        if (ce != cudaSuccess) { set_error("mdm_augment_hfs: %s", cudaGetErrorString(ce)); return PARC_ERR_CUDA; }
        attr_set[dev] = true;
    }
    mdm_hf_augment_kernel<<<(unsigned)std::min<long long>(n, (long long)kNumSMs * 8), 256, bytes, as_stream(stream)>>>(
        *cfg, *aug, terrains, (const long long*)motion_ids, start_times, n, grid_xy, canon, hfs);
    return check_launch("mdm_augment_hfs");
}

extern "C" int parc_b200_mdm_floor_heights(const parc_sampler_config_t* cfg, float dx, const float* hfs, const float* root_pos, int64_t n,
                                           float* out, void* stream) {
    PARC_REQUIRE(cfg != nullptr, "mdm_floor_heights: NULL config");
    PARC_REQUIRE(n >= 0, "mdm_floor_heights: negative window count");
    if (n == 0) return PARC_OK;
    PARC_REQUIRE(hfs && root_pos && out, "mdm_floor_heights: NULL pointer");
    PARC_REQUIRE(dx > 0.f && cfg->grid_dim_x > 0 && cfg->grid_dim_y > 0 && cfg->seq_len > 0, "mdm_floor_heights: bad grid");
    const long long total = n * n * cfg->seq_len;
    mdm_floor_heights_kernel<<<(unsigned)std::min<long long>((total + 255) / 256, (long long)kNumSMs * 16), 256, 0, as_stream(stream)>>>(
        *cfg, dx, hfs, root_pos, n, out);
    return check_launch("mdm_floor_heights");
}

extern "C" int parc_b200_mdm_sample_motion_data(const parc_char_model_t* m, const parc_motion_lib_t* lib, const parc_terrain_t* terrains,
                                                const parc_sampler_config_t* cfg, const int64_t* motion_ids, const float* start_times,
                                                int64_t n, const float* motion_times, const float* grid_xy, const float* future_times,
                                                const float* future_noise, float* root_pos, float* root_rot, float* joint_pos,
                                                float* joint_rot, float* contacts, float* hfs, float* future_pos, float* future_rot,
                                                float* canon, void* stream) {
    if (!model_ok(m, "mdm_sample_motion_data")) return PARC_ERR_INVALID_ARG;
    PARC_REQUIRE(lib && lib->interp && lib->vel && lib->meta, "mdm_sample_motion_data: motion library not packed");
    PARC_REQUIRE(cfg != nullptr, "mdm_sample_motion_data: NULL config");
    PARC_REQUIRE(n >= 0, "mdm_sample_motion_data: negative sample count");
    if (n == 0) return PARC_OK;
    PARC_REQUIRE(motion_ids && start_times && motion_times && canon, "mdm_sample_motion_data: NULL input");
    PARC_REQUIRE(cfg->seq_len > 0 && cfg->ref_frame_idx >= 0 && cfg->ref_frame_idx < cfg->seq_len,
                 "mdm_sample_motion_data: ref_frame_idx %d outside the %d-frame sequence", cfg->ref_frame_idx, cfg->seq_len);
    PARC_REQUIRE(cfg->relative_z_style == PARC_RELATIVE_TO_ROOT || cfg->relative_z_style == PARC_RELATIVE_TO_ROOT_FLOOR,
                 "mdm_sample_motion_data: unknown relative_z_style %d", cfg->relative_z_style);
    if (hfs) {
        PARC_REQUIRE(terrains && grid_xy, "mdm_sample_motion_data: heightmaps requested without terrains / grid points");
        PARC_REQUIRE(cfg->grid_dim_x > 0 && cfg->grid_dim_y > 0 && cfg->center_ix >= 0 && cfg->center_ix < cfg->grid_dim_x &&
                         cfg->center_iy >= 0 && cfg->center_iy < cfg->grid_dim_y, "mdm_sample_motion_data: bad local grid");
    }
    if (future_times) PARC_REQUIRE(future_pos && future_rot, "mdm_sample_motion_data: target info needs future_pos / future_rot");
    PARC_REQUIRE(m->num_bodies <= 32, "mdm_sample_motion_data: more than 32 bodies");
    mdm_canon_hf_kernel<<<(unsigned)std::min<long long>((n + 7) / 8, kNumSMs * 32), 256, 0, as_stream(stream)>>>(
        *lib, *cfg, m->num_bodies, terrains, (const long long*)motion_ids, start_times, n, motion_times, grid_xy, future_times,
        future_noise, hfs, future_pos, future_rot, canon);
    int rc = check_launch("mdm_canon_hf");
    if (rc) return rc;
    if (root_pos || root_rot || joint_pos || joint_rot || contacts) {
        const int G = m->num_bodies <= 16 ? 16 : 32;
        const long long groups = n * cfg->seq_len, gpb = PARC_MDM_THREADS / G;
        const unsigned blocks = (unsigned)std::min<long long>((groups + gpb - 1) / gpb, kNumSMs * 32);
        const int floor_shift = hfs != nullptr && cfg->relative_z_style == PARC_RELATIVE_TO_ROOT_FLOOR;
        if (G == 16)
            mdm_frames_kernel<16><<<blocks, PARC_MDM_THREADS, 0, as_stream(stream)>>>(*m, *lib, *cfg, (const long long*)motion_ids, start_times, n,
                                                                        motion_times, canon, floor_shift, root_pos, root_rot, joint_pos,
                                                                        joint_rot, contacts);
        else
            mdm_frames_kernel<32><<<blocks, PARC_MDM_THREADS, 0, as_stream(stream)>>>(*m, *lib, *cfg, (const long long*)motion_ids, start_times, n,
                                                                        motion_times, canon, floor_shift, root_pos, root_rot, joint_pos,
                                                                        joint_rot, contacts);
        rc = check_launch("mdm_frames");
    }
    return rc;
}
