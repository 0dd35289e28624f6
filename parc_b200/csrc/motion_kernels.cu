// MotionLib kernels: record packing and calc_motion_frame (PARC/anim/motion_lib.py:94-126, 425-457).
// Also the standalone terrain kernels (ray heights, heading).
#include "parc_common.cuh"
#include "parc_lanes.cuh"

namespace parc {

// One thread per output float of the interp / vel records; reads are coalesced on the source
// tensors' trailing dims, writes on the record. The pair block of frame f holds slerp_pair(q_j(f), q_j(f')) for the
// quaternions j = 0..J-1 (0 = root rotation), f' = the next frame of the same clip (f itself on a clip's last frame:
// _calc_frame_blend, motion_lib.py:425-438, clamps idx1 to the last frame).
__global__ void __launch_bounds__(256) pack_frames_kernel(int J, int D, int C, long long F, int irec, int vrec, int M,
                                                          const long long* __restrict__ start_idx,
                                                          const long long* __restrict__ num_frames,
                                                          const float* __restrict__ root_pos,
                                                          const float* __restrict__ root_rot,
                                                          const float* __restrict__ root_vel,
                                                          const float* __restrict__ root_ang_vel,
                                                          const float* __restrict__ joint_rot,
                                                          const float* __restrict__ dof_vel,
                                                          const float* __restrict__ contacts,
                                                          float* __restrict__ interp, float* __restrict__ vel) {
    const long long total_i = F * irec;
    const long long total = total_i + F * vrec;
    const int pair0 = interp_pair_off(J, C);
    for (long long k = (long long)blockIdx.x * blockDim.x + threadIdx.x; k < total; k += (long long)gridDim.x * blockDim.x) {
        if (k < total_i) {
            const long long f = k / irec;
            const int e = (int)(k - f * irec);
            float v = 0.f;
            if (e < 4) v = root_rot[f * 4 + e];
            else if (e < 4 * J) v = joint_rot[f * (J - 1) * 4 + (e - 4)];
            else if (e < 4 * J + 3) v = root_pos[f * 3 + (e - 4 * J)];
            else if (e >= 4 * J + 4 && e < 4 * J + 4 + C) v = contacts[f * C + (e - 4 * J - 4)];
            else if (e >= pair0 && e < pair0 + 2 * J) {
                // clip of frame f: last m with start_idx[m] <= f
                int lo = 0, hi = M - 1;
                while (lo < hi) { const int mid = (lo + hi + 1) >> 1; if (start_idx[mid] <= f) lo = mid; else hi = mid - 1; }
                const long long last = start_idx[lo] + num_frames[lo] - 1;
                const long long f1 = f < last ? f + 1 : f;
                const int j = (e - pair0) >> 1;
                const float* a = j == 0 ? root_rot + f * 4 : joint_rot + (f * (J - 1) + (j - 1)) * 4;
                const float* b = j == 0 ? root_rot + f1 * 4 : joint_rot + (f1 * (J - 1) + (j - 1)) * 4;
                const float2 pr = slerp_pair(make_quat(a[0], a[1], a[2], a[3]), make_quat(b[0], b[1], b[2], b[3]));
                v = ((e - pair0) & 1) ? pr.y : pr.x;
            }
            interp[k] = v;
        } else {
            const long long kk = k - total_i;
            const long long f = kk / vrec;
            const int e = (int)(kk - f * vrec);
            float v = 0.f;
            if (e < 3) v = root_vel[f * 3 + e];
            else if (e < 6) v = root_ang_vel[f * 3 + (e - 3)];
            else if (e < 6 + D) v = dof_vel[f * D + (e - 6)];
            vel[kk] = v;
        }
    }
}

__global__ void pack_meta_kernel(int M, const long long* __restrict__ num_frames, const long long* __restrict__ start_idx,
                                 const float* __restrict__ lengths, const int* __restrict__ loop_modes,
                                 const float* __restrict__ delta, int* __restrict__ meta) {
    for (int m = blockIdx.x * blockDim.x + threadIdx.x; m < M; m += gridDim.x * blockDim.x) {
        int* o = meta + 8 * m;
        o[0] = (int)start_idx[m];
        o[1] = (int)num_frames[m];
        o[2] = __float_as_int(lengths[m]);
        o[3] = loop_modes[m];
        o[4] = __float_as_int(delta[m * 3]);
        o[5] = __float_as_int(delta[m * 3 + 1]);
        o[6] = __float_as_int(delta[m * 3 + 2]);
        o[7] = 0;
    }
}

// calc_motion_frame: one group of G lanes per (motion id, time) sample.
template <int G>
__global__ void __launch_bounds__(256) calc_motion_frame_kernel(
    const __grid_constant__ parc_char_model_t cm, const __grid_constant__ parc_motion_lib_t lib,
    const long long* __restrict__ motion_ids, const float* __restrict__ motion_times, long long n,
    float* __restrict__ root_pos, float* __restrict__ root_rot, float* __restrict__ root_vel,
    float* __restrict__ root_ang_vel, float* __restrict__ joint_rot, float* __restrict__ dof_vel,
    float* __restrict__ contacts, long long* __restrict__ frame_idx0, long long* __restrict__ frame_idx1,
    float* __restrict__ blend) {
    const LaneBody lb = load_lane_body<G>(cm);
    const int J = cm.num_bodies, D = cm.dof_size, C = lib.num_contacts;
    const int sub = threadIdx.x & (G - 1);
    const long long gpb = blockDim.x / G;
    for (long long i = (long long)blockIdx.x * gpb + threadIdx.x / G; i < n; i += (long long)gridDim.x * gpb) {
        const long long mid = __ldg(motion_ids + i);
        const float t = __ldg(motion_times + i);
        const MotionMeta mm = load_motion_meta(lib.meta, mid);
        const FrameBlend fb = calc_frame_blend(mm, t);
        const PoseSample ps = eval_pose_lane(lib, lb, fb, J);
        if (sub == 0) {
            if (root_rot) reinterpret_cast<float4*>(root_rot)[i] = quat_to_f4(ps.q);
            if (root_pos) { root_pos[i * 3] = ps.root_pos.x; root_pos[i * 3 + 1] = ps.root_pos.y; root_pos[i * 3 + 2] = ps.root_pos.z; }
            if (frame_idx0) frame_idx0[i] = fb.idx0;
            if (frame_idx1) frame_idx1[i] = fb.idx1;
            if (blend) blend[i] = fb.blend;
        } else if (lb.active && joint_rot) {
            reinterpret_cast<float4*>(joint_rot)[i * (J - 1) + (sub - 1)] = quat_to_f4(ps.q);
        }
        if (contacts && sub < C) contacts[i * C + sub] = ps.contact;
        // velocities come from frame idx0 without interpolation (motion_lib.py:102-110)
        const float* v = lib.vel + fb.idx0 * lib.vrec;
        for (int e = sub; e < 6 + D; e += G) {
            float x = __ldg(v + e);
            if (e < 3) { if (root_vel) root_vel[i * 3 + e] = x; }
            else if (e < 6) { if (root_ang_vel) root_ang_vel[i * 3 + (e - 3)] = x; }
            else if (dof_vel) dof_vel[i * D + (e - 6)] = x;
        }
    }
}

// ---- terrain ---------------------------------------------------------------------------------

// rotate_2d_vec (PARC/util/torch_util.py:650-662) + root xy, each product rounded separately.
__device__ __forceinline__ void ray_point_world(float px, float py, float c, float s, float rx, float ry, float& wx, float& wy) {
    wx = __fadd_rn(__fsub_rn(__fmul_rn(px, c), __fmul_rn(py, s)), rx);
    wy = __fadd_rn(__fadd_rn(__fmul_rn(px, s), __fmul_rn(py, c)), ry);
}

__global__ void __launch_bounds__(256) ray_heights_kernel(const __grid_constant__ parc_terrain_t tr,
                                                          const float* __restrict__ root_xyz,
                                                          const float* __restrict__ heading,
                                                          const float* __restrict__ ray_xy, int P, long long n,
                                                          float lo, float hi, float* __restrict__ out,
                                                          long long out_stride, long long* __restrict__ grid_idx) {
    // one warp per env, lanes over the P sample points
    const int lane = threadIdx.x & 31;
    const long long wpb = blockDim.x / 32;
    for (long long i = (long long)blockIdx.x * wpb + threadIdx.x / 32; i < n; i += (long long)gridDim.x * wpb) {
        const float rx = __ldg(root_xyz + i * 3), ry = __ldg(root_xyz + i * 3 + 1), rz = __ldg(root_xyz + i * 3 + 2);
        float s, c;
        const float h = __ldg(heading + i);
        s = sinf(h); c = cosf(h);
        for (int p = lane; p < P; p += 32) {
            const float2 pt = __ldg(reinterpret_cast<const float2*>(ray_xy) + p);
            float wx, wy;
            ray_point_world(pt.x, pt.y, c, s, rx, ry, wx, wy);
            const int ix = grid_index(wx, tr.min_x, tr.dx, tr.dim_x);
            const int iy = grid_index(wy, tr.min_y, tr.dy, tr.dim_y);
            const float hval = __ldg(tr.hf + (long long)ix * tr.dim_y + iy);
            out[i * out_stride + p] = fminf(fmaxf(__fsub_rn(hval, rz), lo), hi);   // mgdm_dm_util.py:139-140
            if (grid_idx) { grid_idx[(i * P + p) * 2] = ix; grid_idx[(i * P + p) * 2 + 1] = iy; }
        }
    }
}

__global__ void calc_heading_kernel(const float* __restrict__ root_rot, long long n, float* __restrict__ heading) {
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
        heading[i] = calc_heading(quat_from_f4(__ldg(reinterpret_cast<const float4*>(root_rot) + i)));
}

static int grid_cap(long long blocks) {
    long long cap = (long long)kNumSMs * 16;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    return (int)blocks;
}

}  // namespace parc

using namespace parc;

extern "C" int parc_b200_motionlib_pack(const parc_char_model_t* m, int64_t F, int32_t M, int32_t C,
                                        const float* root_pos, const float* root_rot, const float* root_vel,
                                        const float* root_ang_vel, const float* joint_rot, const float* dof_vel,
                                        const float* contacts, const int64_t* num_frames, const int64_t* start_idx,
                                        const float* lengths, const int32_t* loop_modes, const float* delta,
                                        float* interp, float* vel, int32_t* meta, void* stream) {
    if (!model_ok(m, "motionlib_pack")) return PARC_ERR_INVALID_ARG;
    PARC_REQUIRE(F > 0 && M > 0, "motionlib_pack: empty library (F=%lld, M=%d)", (long long)F, M);
    PARC_REQUIRE(F < (1ll << 31), "motionlib_pack: more than 2^31 frames");
    PARC_REQUIRE(C >= 0 && C <= m->num_bodies, "motionlib_pack: num_contacts %d must be in [0, num_bodies]", C);
    PARC_REQUIRE(C == 0 || contacts, "motionlib_pack: contacts is NULL");
    PARC_REQUIRE(root_pos && root_rot && root_vel && root_ang_vel && dof_vel && interp && vel && meta && num_frames && start_idx,
                 "motionlib_pack: NULL pointer");
    const int J = m->num_bodies, D = m->dof_size;
    const int irec = interp_record_floats(J, C), vrec = vel_record_floats(D);
    const long long total = F * (irec + vrec);
    pack_frames_kernel<<<grid_cap((total + 255) / 256), 256, 0, as_stream(stream)>>>(
        J, D, C, F, irec, vrec, M, (const long long*)start_idx, (const long long*)num_frames, root_pos, root_rot, root_vel, root_ang_vel,
        joint_rot, dof_vel, contacts, interp, vel);
    int rc = check_launch("motionlib_pack(frames)");
    if (rc) return rc;
    pack_meta_kernel<<<grid_cap((M + 255) / 256), 256, 0, as_stream(stream)>>>(
        M, (const long long*)num_frames, (const long long*)start_idx, lengths, loop_modes, delta, meta);
    return check_launch("motionlib_pack(meta)");
}

extern "C" int parc_b200_calc_motion_frame(const parc_char_model_t* m, const parc_motion_lib_t* lib,
                                           const int64_t* motion_ids, const float* motion_times, int64_t n,
                                           float* root_pos, float* root_rot, float* root_vel, float* root_ang_vel,
                                           float* joint_rot, float* dof_vel, float* contacts, int64_t* frame_idx0,
                                           int64_t* frame_idx1, float* blend, void* stream) {
    if (!model_ok(m, "calc_motion_frame")) return PARC_ERR_INVALID_ARG;
    PARC_REQUIRE(lib && lib->interp && lib->vel && lib->meta, "calc_motion_frame: library not packed");
    PARC_REQUIRE(n >= 0, "calc_motion_frame: n < 0");
    if (n == 0) return PARC_OK;
    PARC_REQUIRE(motion_ids && motion_times, "calc_motion_frame: NULL ids/times");
    PARC_REQUIRE(contacts == nullptr || lib->num_contacts > 0, "calc_motion_frame: library has no contact info");
    if (m->num_bodies <= 16)
        calc_motion_frame_kernel<16><<<grid_cap((n + 15) / 16), 256, 0, as_stream(stream)>>>(
            *m, *lib, (const long long*)motion_ids, motion_times, n, root_pos, root_rot, root_vel, root_ang_vel,
            joint_rot, dof_vel, contacts, (long long*)frame_idx0, (long long*)frame_idx1, blend);
    else
        calc_motion_frame_kernel<32><<<grid_cap((n + 7) / 8), 256, 0, as_stream(stream)>>>(
            *m, *lib, (const long long*)motion_ids, motion_times, n, root_pos, root_rot, root_vel, root_ang_vel,
            joint_rot, dof_vel, contacts, (long long*)frame_idx0, (long long*)frame_idx1, blend);
    return check_launch("calc_motion_frame");
}

extern "C" int parc_b200_ray_heights(const parc_terrain_t* tr, const float* root_xyz, const float* heading,
                                     const float* ray_xy, int32_t P, int64_t n, float lo, float hi, float* out,
                                     int64_t out_stride, int64_t* grid_idx, void* stream) {
    PARC_REQUIRE(tr && tr->hf && tr->dim_x > 0 && tr->dim_y > 0, "ray_heights: empty terrain");
    PARC_REQUIRE(P >= 0 && n >= 0, "ray_heights: negative size");
    if (n == 0 || P == 0) return PARC_OK;
    PARC_REQUIRE(root_xyz && heading && ray_xy && out, "ray_heights: NULL pointer");
    PARC_REQUIRE(out_stride >= P, "ray_heights: out_row_stride %lld < num_points %d", (long long)out_stride, P);
    ray_heights_kernel<<<grid_cap((n + 7) / 8), 256, 0, as_stream(stream)>>>(*tr, root_xyz, heading, ray_xy, P, n, lo, hi,
                                                                          out, out_stride, (long long*)grid_idx);
    return check_launch("ray_heights");
}

extern "C" int parc_b200_calc_heading(const float* root_rot, int64_t n, float* heading, void* stream) {
    if (n == 0) return PARC_OK;
    PARC_REQUIRE(root_rot && heading, "calc_heading: NULL pointer");
    calc_heading_kernel<<<grid_cap((n + 255) / 256), 256, 0, as_stream(stream)>>>(root_rot, n, heading);
    return check_launch("calc_heading");
}
