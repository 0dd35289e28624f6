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

}  // namespace parc

using namespace parc;

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
