// Stand-alone forms of the reference's per-call functions (same inputs/outputs as the Python
// methods they replace) and the reset / fail-rate kernels. The fused step lives in env_kernels.cu.
#include "parc_common.cuh"
#include "parc_lanes.cuh"

namespace parc {

static int blocks_for(long long items, int per_block) {
    long long b = (items + per_block - 1) / per_block;
    const long long cap = (long long)kNumSMs * 32;
    if (b > cap) b = cap;
    if (b < 1) b = 1;
    return (int)b;
}

// ---- DeepMimicEnv.compute_tar_obs (dm_env.py:594-626) + fetch_tar_obs_data (mgdm_dm_util.py:221-244)
// One group of G lanes per (env, target step) sample, env-major like the reference's flatten.
template <int G>
__global__ void __launch_bounds__(256) tar_obs_data_kernel(
    const __grid_constant__ parc_char_model_t cm, const __grid_constant__ parc_motion_lib_t lib,
    const __grid_constant__ parc_env_config_t cfg, const __grid_constant__ parc_env_buffers_t env,
    const long long* __restrict__ env_ids, long long n, float* __restrict__ tar_root_pos, float* __restrict__ tar_root_rot,
    float* __restrict__ tar_joint_rot, float* __restrict__ tar_key_pos, float* __restrict__ tar_contacts) {
    const LaneBody lb = load_lane_body<G>(cm);
    const int J = cm.num_bodies, C = lib.num_contacts, S = cfg.num_tar_steps, K = cfg.num_key_bodies;
    const int sub = threadIdx.x & (G - 1);
    const long long gpb = blockDim.x / G;
    const long long total = n * S;
    const long long g0 = (long long)blockIdx.x * gpb + threadIdx.x / G;
    const long long warp_first = g0 - ((threadIdx.x & 31) / G);
    for (long long base = warp_first; base < total; base += (long long)gridDim.x * gpb) {
        const long long i = base + ((threadIdx.x & 31) / G);
        const bool live = i < total;
        const long long ii = live ? i : 0;
        const long long r = ii / S;
        const int s = (int)(ii - r * S);
        const long long e = env_ids ? env_ids[r] : r;
        const long long mid = env.motion_ids[e];
        const MotionMeta mm = load_motion_meta(lib.meta, mid);
        const float t0 = __fadd_rn(env.time_buf[e], env.motion_time_offsets[e]);
        const float t = __fadd_rn(t0, cfg.tar_dt[s]);
        const FrameBlend fb = calc_frame_blend(mm, t);
        const PoseSample ps = eval_pose_lane(lib, lb, fb, J);
        const float* dmo = env.dm_motion_offsets + (mid * env.terrains_per_motion + env.motion_terrain_ids[e]) * 2;
        const float tox = __fsub_rn(__ldg(dmo), __ldg(env.env_offsets + e * 3));
        const float toy = __fsub_rn(__ldg(dmo + 1), __ldg(env.env_offsets + e * 3 + 1));
        const Vec3 rpos = make_vec3(__fadd_rn(ps.root_pos.x, tox), __fadd_rn(ps.root_pos.y, toy), ps.root_pos.z);
        const Quat rrot = group_bcast<G>(ps.q, 0);
        Vec3 bpos; Quat brot;
        fk_lanes<G>(lb, ps.q, rpos, rrot, bpos, brot);
        if (!live) continue;
        if (sub == 0) {
            tar_root_pos[ii * 3] = rpos.x; tar_root_pos[ii * 3 + 1] = rpos.y; tar_root_pos[ii * 3 + 2] = rpos.z;
            reinterpret_cast<float4*>(tar_root_rot)[ii] = quat_to_f4(ps.q);
        } else if (lb.active) {
            reinterpret_cast<float4*>(tar_joint_rot)[ii * (J - 1) + sub - 1] = quat_to_f4(ps.q);
        }
        if (lb.active && tar_key_pos)
            for (int k = 0; k < K; ++k)
                if (cfg.key_body_ids[k] == lb.body) {
                    float* o = tar_key_pos + (ii * K + k) * 3;
                    o[0] = bpos.x; o[1] = bpos.y; o[2] = bpos.z;
                }
        if (tar_contacts && sub < C) tar_contacts[ii * C + sub] = ps.contact;
    }
}

// ---- compute_char_obs (ig_char_env.py:581-625): one group per env --------------------------------
template <int G>
__global__ void __launch_bounds__(256) char_obs_kernel(const __grid_constant__ parc_char_model_t cm,
                                                       const float* __restrict__ root_pos, const float* __restrict__ root_rot,
                                                       const float* __restrict__ root_vel, const float* __restrict__ root_ang_vel,
                                                       const float* __restrict__ joint_rot, const float* __restrict__ dof_vel,
                                                       const float* __restrict__ key_pos, int K, int global_obs, int root_h,
                                                       long long n, float* __restrict__ obs) {
    const int J = cm.num_bodies, D = cm.dof_size;
    const int sub = threadIdx.x & (G - 1);
    const int dim = root_h + 12 + 6 * (J - 1) + D + 3 * K;
    const long long gpb = blockDim.x / G;
    for (long long i = (long long)blockIdx.x * gpb + threadIdx.x / G; i < n; i += (long long)gridDim.x * gpb) {
        const Quat rr = quat_from_f4(__ldg(reinterpret_cast<const float4*>(root_rot) + i));
        const Vec3 rp = make_vec3(root_pos[i * 3], root_pos[i * 3 + 1], root_pos[i * 3 + 2]);
        const Quat hinv = calc_heading_quat_inv(rr);
        float* o = obs + i * dim + root_h;
        if (sub == 0) {
            if (root_h) obs[i * dim] = rp.z;
            float tn[6];
            quat_to_tan_norm(global_obs ? rr : quat_mul(hinv, rr), tn);
            for (int k = 0; k < 6; ++k) o[k] = tn[k];
            Vec3 v = make_vec3(root_vel[i * 3], root_vel[i * 3 + 1], root_vel[i * 3 + 2]);
            Vec3 w = make_vec3(root_ang_vel[i * 3], root_ang_vel[i * 3 + 1], root_ang_vel[i * 3 + 2]);
            if (!global_obs) { v = quat_rotate(hinv, v); w = quat_rotate(hinv, w); }
            o[6] = v.x; o[7] = v.y; o[8] = v.z; o[9] = w.x; o[10] = w.y; o[11] = w.z;
        } else if (sub < J) {
            float tn[6];
            quat_to_tan_norm(quat_from_f4(__ldg(reinterpret_cast<const float4*>(joint_rot) + i * (J - 1) + sub - 1)), tn);
            for (int k = 0; k < 6; ++k) o[12 + 6 * (sub - 1) + k] = tn[k];
        }
        for (int k = sub; k < D; k += G) o[12 + 6 * (J - 1) + k] = dof_vel[i * D + k];
        for (int k = sub; k < K; k += G) {
            Vec3 kp = make_vec3(key_pos[(i * K + k) * 3] - rp.x, key_pos[(i * K + k) * 3 + 1] - rp.y, key_pos[(i * K + k) * 3 + 2] - rp.z);
            if (!global_obs) kp = quat_rotate(hinv, kp);
            float* kd = o + 12 + 6 * (J - 1) + D + 3 * k;
            kd[0] = kp.x; kd[1] = kp.y; kd[2] = kp.z;
        }
    }
}

// ---- compute_tar_obs (mgdm_dm_util.py:404-461): one group per (env, step) -------------------------
template <int G>
__global__ void __launch_bounds__(256) tar_obs_kernel(const __grid_constant__ parc_char_model_t cm,
                                                      const float* __restrict__ ref_root_pos, const float* __restrict__ ref_root_rot,
                                                      const float* __restrict__ tar_root_pos, const float* __restrict__ tar_root_rot,
                                                      const float* __restrict__ tar_joint_rot, const float* __restrict__ tar_key_pos,
                                                      int S, int K, int global_obs, int global_h, long long n, float* __restrict__ obs) {
    const int J = cm.num_bodies;
    const int sub = threadIdx.x & (G - 1);
    const int dim = 9 + 6 * (J - 1) + 3 * K;
    const long long gpb = blockDim.x / G;
    const long long total = n * S;
    for (long long i = (long long)blockIdx.x * gpb + threadIdx.x / G; i < total; i += (long long)gridDim.x * gpb) {
        const long long r = i / S;
        const Quat rr = quat_from_f4(__ldg(reinterpret_cast<const float4*>(ref_root_rot) + r));
        const Quat hinv = calc_heading_quat_inv(rr);
        const Vec3 trp = make_vec3(tar_root_pos[i * 3], tar_root_pos[i * 3 + 1], tar_root_pos[i * 3 + 2]);
        Vec3 rp_obs = make_vec3(trp.x - ref_root_pos[r * 3], trp.y - ref_root_pos[r * 3 + 1], trp.z - ref_root_pos[r * 3 + 2]);
        if (!global_obs) rp_obs = quat_rotate(hinv, rp_obs);
        float* o = obs + i * dim;
        if (sub == 0) {
            o[0] = rp_obs.x; o[1] = rp_obs.y; o[2] = global_h ? trp.z : rp_obs.z;
            Quat q = quat_from_f4(__ldg(reinterpret_cast<const float4*>(tar_root_rot) + i));
            if (!global_obs) q = quat_mul(hinv, q);
            float tn[6];
            quat_to_tan_norm(q, tn);
            for (int k = 0; k < 6; ++k) o[3 + k] = tn[k];
        } else if (sub < J) {
            float tn[6];
            quat_to_tan_norm(quat_from_f4(__ldg(reinterpret_cast<const float4*>(tar_joint_rot) + i * (J - 1) + sub - 1)), tn);
            for (int k = 0; k < 6; ++k) o[9 + 6 * (sub - 1) + k] = tn[k];
        }
        for (int k = sub; k < K; k += G) {
            Vec3 kp = make_vec3(tar_key_pos[(i * K + k) * 3] - trp.x, tar_key_pos[(i * K + k) * 3 + 1] - trp.y, tar_key_pos[(i * K + k) * 3 + 2] - trp.z);
            // the key offsets use root_pos_obs BEFORE the optional global-height overwrite (:437-446)
            if (!global_obs) kp = quat_rotate(hinv, kp) + rp_obs;
            float* kd = o + 9 + 6 * (J - 1) + 3 * k;
            kd[0] = kp.x; kd[1] = kp.y; kd[2] = kp.z;
        }
    }
}

// ---- compute_deepmimic_reward (mgdm_dm_util.py:269-332): one warp per env --------------------------
__global__ void __launch_bounds__(256) dm_reward_kernel(int J, int D, int K, const float* __restrict__ root_pos,
                                                        const float* __restrict__ root_rot, const float* __restrict__ root_vel,
                                                        const float* __restrict__ root_ang_vel, const float* __restrict__ joint_rot,
                                                        const float* __restrict__ dof_vel, const float* __restrict__ key_pos,
                                                        const float* __restrict__ t_root_pos, const float* __restrict__ t_root_rot,
                                                        const float* __restrict__ t_root_vel, const float* __restrict__ t_root_ang_vel,
                                                        const float* __restrict__ t_joint_rot, const float* __restrict__ t_dof_vel,
                                                        const float* __restrict__ t_key_pos, const float* __restrict__ jw,
                                                        const float* __restrict__ dw, int track_root_h, int track_root, long long n,
                                                        float* __restrict__ out) {
    const int lane = threadIdx.x & 31;
    const long long wpb = blockDim.x / 32;
    for (long long i = (long long)blockIdx.x * wpb + threadIdx.x / 32; i < n; i += (long long)gridDim.x * wpb) {
        float pose_t = 0.f, vel_t = 0.f, key_t = 0.f;
        for (int j = lane; j < J - 1; j += 32) {
            const Quat a = quat_from_f4(__ldg(reinterpret_cast<const float4*>(joint_rot) + i * (J - 1) + j));
            const Quat b = quat_from_f4(__ldg(reinterpret_cast<const float4*>(t_joint_rot) + i * (J - 1) + j));
            const float ang = quat_diff_angle(a, b);
            pose_t += jw[j] * ang * ang;
        }
        for (int k = lane; k < D; k += 32) {
            const float d = t_dof_vel[i * D + k] - dof_vel[i * D + k];
            vel_t += dw[k] * d * d;
        }
        const Vec3 rp = make_vec3(root_pos[i * 3], root_pos[i * 3 + 1], root_pos[i * 3 + 2]);
        const Vec3 trp = make_vec3(t_root_pos[i * 3], t_root_pos[i * 3 + 1], t_root_pos[i * 3 + 2]);
        Quat a_rot = quat_from_f4(__ldg(reinterpret_cast<const float4*>(root_rot) + i));
        Quat b_rot = quat_from_f4(__ldg(reinterpret_cast<const float4*>(t_root_rot) + i));
        const Quat ha = calc_heading_quat_inv(a_rot), hb = calc_heading_quat_inv(b_rot);
        for (int k = lane; k < K; k += 32) {
            Vec3 kp = make_vec3(key_pos[(i * K + k) * 3] - rp.x, key_pos[(i * K + k) * 3 + 1] - rp.y, key_pos[(i * K + k) * 3 + 2] - rp.z);
            Vec3 tk = make_vec3(t_key_pos[(i * K + k) * 3] - trp.x, t_key_pos[(i * K + k) * 3 + 1] - trp.y, t_key_pos[(i * K + k) * 3 + 2] - trp.z);
            if (!track_root) { kp = quat_rotate(ha, kp); tk = quat_rotate(hb, tk); }
            const Vec3 d = tk - kp;
            key_t += d.x * d.x + d.y * d.y + d.z * d.z;
        }
        for (int o = 16; o > 0; o >>= 1) {
            pose_t += __shfl_xor_sync(FULL_MASK, pose_t, o);
            vel_t += __shfl_xor_sync(FULL_MASK, vel_t, o);
            key_t += __shfl_xor_sync(FULL_MASK, key_t, o);
        }
        if (lane == 0) {
            Vec3 rpd = trp - rp;
            if (!track_root) { rpd.x = 0.f; rpd.y = 0.f; }
            if (!track_root_h) rpd.z = 0.f;
            const float root_pos_err = rpd.x * rpd.x + rpd.y * rpd.y + rpd.z * rpd.z;
            Vec3 av = make_vec3(root_vel[i * 3], root_vel[i * 3 + 1], root_vel[i * 3 + 2]);
            Vec3 bv = make_vec3(t_root_vel[i * 3], t_root_vel[i * 3 + 1], t_root_vel[i * 3 + 2]);
            Vec3 aw = make_vec3(root_ang_vel[i * 3], root_ang_vel[i * 3 + 1], root_ang_vel[i * 3 + 2]);
            Vec3 bw = make_vec3(t_root_ang_vel[i * 3], t_root_ang_vel[i * 3 + 1], t_root_ang_vel[i * 3 + 2]);
            if (!track_root) {
                av = quat_rotate(ha, av); aw = quat_rotate(ha, aw); bv = quat_rotate(hb, bv); bw = quat_rotate(hb, bw);
                a_rot = quat_mul(ha, a_rot); b_rot = quat_mul(hb, b_rot);
            }
            const float rra = quat_diff_angle(a_rot, b_rot);
            const Vec3 dv = bv - av, dwv = bw - aw;
            const float rve = dv.x * dv.x + dv.y * dv.y + dv.z * dv.z;
            const float rae = dwv.x * dwv.x + dwv.y * dwv.y + dwv.z * dwv.z;
            float* o = out + i * 5;
            o[0] = expf(-0.25f * pose_t);
            o[1] = expf(-0.01f * vel_t);
            o[2] = expf(-5.0f * (root_pos_err + 0.1f * rra * rra));
            o[3] = expf(-1.0f * (rve + 0.1f * rae));
            o[4] = expf(-10.0f * key_t);
        }
    }
}

// ---- compute_contact_reward (mgdm_dm_util.py:497-518) ------------------------------------------------
__global__ void contact_reward_kernel(const float* __restrict__ tar, const float* __restrict__ f, const float* __restrict__ w, int C,
                                      long long n, float* __restrict__ out) {
    for (long long k = (long long)blockIdx.x * blockDim.x + threadIdx.x; k < n * C; k += (long long)gridDim.x * blockDim.x) {
        const float fx = f[k * 3], fy = f[k * 3 + 1], fz = f[k * 3 + 2];
        const float m = fminf(sqrtf(fx * fx + fy * fy + fz * fz), 1.0f);
        const float c = tar[k];
        float r = -(1.0f - c) * m;
        r += c * m;
        out[k] = w[k % C] * r;
    }
}

// ---- compute_tracking_error (mgdm_dm_util.py:520-553): one warp per env, lanes over bodies / dofs -----
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(FULL_MASK, v, o);
    return v;
}

__global__ void __launch_bounds__(256) tracking_error_kernel(
    int J, int D, const float* __restrict__ root_pos, const float* __restrict__ root_rot, const float* __restrict__ body_rot,
    const float* __restrict__ body_pos, const float* __restrict__ tar_root_pos, const float* __restrict__ tar_root_rot,
    const float* __restrict__ tar_body_rot, const float* __restrict__ tar_body_pos, const float* __restrict__ root_vel,
    const float* __restrict__ root_ang_vel, const float* __restrict__ dof_vel, const float* __restrict__ tar_root_vel,
    const float* __restrict__ tar_root_ang_vel, const float* __restrict__ tar_dof_vel, long long n, float* __restrict__ out) {
    const int lane = threadIdx.x & 31;
    const long long wpb = blockDim.x / 32;
    for (long long i = (long long)blockIdx.x * wpb + threadIdx.x / 32; i < n; i += (long long)gridDim.x * wpb) {
        const Vec3 rp = make_vec3(root_pos[i * 3], root_pos[i * 3 + 1], root_pos[i * 3 + 2]);
        const Vec3 trp = make_vec3(tar_root_pos[i * 3], tar_root_pos[i * 3 + 1], tar_root_pos[i * 3 + 2]);
        float pose = 0.f, bpos = 0.f, dv = 0.f;
        for (int b = lane; b < J; b += 32) {
            const Quat q = quat_from_f4(__ldg(reinterpret_cast<const float4*>(body_rot) + i * J + b));
            const Quat tq = quat_from_f4(__ldg(reinterpret_cast<const float4*>(tar_body_rot) + i * J + b));
            pose += fabsf(quat_diff_angle(q, tq));
            const float* p = body_pos + (i * J + b) * 3;
            const float* tp = tar_body_pos + (i * J + b) * 3;
            const Vec3 d = (make_vec3(tp[0], tp[1], tp[2]) - trp) - (make_vec3(p[0], p[1], p[2]) - rp);
            bpos += sqrtf(sqnorm3_seq(d));
        }
        for (int d = lane; d < D; d += 32) dv += fabsf(__fsub_rn(tar_dof_vel[i * D + d], dof_vel[i * D + d]));
        pose = warp_sum(pose);
        bpos = warp_sum(bpos);
        dv = warp_sum(dv);
        if (lane == 0) {
            float* o = out + i * 7;
            o[0] = sqrtf(sqnorm3_seq(trp - rp));
            const Quat a = quat_from_f4(__ldg(reinterpret_cast<const float4*>(root_rot) + i));
            const Quat b = quat_from_f4(__ldg(reinterpret_cast<const float4*>(tar_root_rot) + i));
            o[1] = fabsf(quat_diff_angle(a, b));
            o[2] = bpos / (float)J;
            o[3] = pose / (float)J;
            o[4] = D > 0 ? dv / (float)D : nanf("");     // torch.mean over an empty dim is NaN
            float rv = 0.f, rav = 0.f;
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                rv += fabsf(__fsub_rn(tar_root_vel[i * 3 + k], root_vel[i * 3 + k]));
                rav += fabsf(__fsub_rn(tar_root_ang_vel[i * 3 + k], root_ang_vel[i * 3 + k]));
            }
            o[5] = rv / 3.0f;
            o[6] = rav / 3.0f;
        }
    }
}

// ---- compute_done (mgdm_dm_util.py:334-402): one warp per env, lanes over bodies ---------------------
__global__ void __launch_bounds__(256) compute_done_kernel(const float* __restrict__ time, float ep_len, const float* __restrict__ root_rot,
                                                           const float* __restrict__ body_pos, const float* __restrict__ tar_root_rot,
                                                           const float* __restrict__ tar_body_pos, const float* __restrict__ contact_force,
                                                           const long long* __restrict__ cb_ids, int n_cb, const float* __restrict__ term_h,
                                                           int pose_termination, const float* __restrict__ pose_dist, int early, int track_root,
                                                           float root_pos_dist_sq, float root_rot_angle, int J, long long n, int* __restrict__ done_out) {
    const int lane = threadIdx.x & 31;
    const long long wpb = blockDim.x / 32;
    for (long long i = (long long)blockIdx.x * wpb + threadIdx.x / 32; i < n; i += (long long)gridDim.x * wpb) {
        int done = PARC_DONE_NULL;
        const float t = time[i];
        if (t >= ep_len) done = PARC_DONE_TIME;
        if (early) {
            bool failed = false;
            if (n_cb > 0) {
                bool fc = false, fh = false;
                for (int b = lane; b < J; b += 32) {
                    bool is_cb = false;
                    for (int k = 0; k < n_cb; ++k) is_cb = is_cb || (cb_ids[k] == b);
                    if (is_cb) continue;
                    const float* f = contact_force + (i * J + b) * 3;
                    fc = fc || fabsf(f[0]) > 0.1f || fabsf(f[1]) > 0.1f || fabsf(f[2]) > 0.1f;
                    fh = fh || (body_pos[(i * J + b) * 3 + 2] < term_h[i * J + b]);
                }
                failed = __any_sync(FULL_MASK, fc) && __any_sync(FULL_MASK, fh);
            }
            if (pose_termination) {
                bool pf = false;
                const float* b0 = body_pos + i * J * 3;
                const float* t0 = tar_body_pos + i * J * 3;
                for (int b = lane + 1; b < J; b += 32) {
                    const Vec3 a = make_vec3(__fsub_rn(b0[b * 3], b0[0]), __fsub_rn(b0[b * 3 + 1], b0[1]), __fsub_rn(b0[b * 3 + 2], b0[2]));
                    const Vec3 c = make_vec3(__fsub_rn(t0[b * 3], t0[0]), __fsub_rn(t0[b * 3 + 1], t0[1]), __fsub_rn(t0[b * 3 + 2], t0[2]));
                    const Vec3 d = make_vec3(__fsub_rn(c.x, a.x), __fsub_rn(c.y, a.y), __fsub_rn(c.z, a.z));
                    const float td = pose_dist[b - 1];
                    pf = pf || (sqnorm3_seq(d) > __fmul_rn(td, td));
                }
                pf = __any_sync(FULL_MASK, pf);
                if (track_root) {
                    const Vec3 d = make_vec3(__fsub_rn(b0[0], t0[0]), __fsub_rn(b0[1], t0[1]), __fsub_rn(b0[2], t0[2]));
                    pf = pf || (sqnorm3_seq(d) > root_pos_dist_sq);
                    const Quat a = quat_from_f4(__ldg(reinterpret_cast<const float4*>(root_rot) + i));
                    const Quat b = quat_from_f4(__ldg(reinterpret_cast<const float4*>(tar_root_rot) + i));
                    pf = pf || (fabsf(quat_diff_angle(a, b)) > root_rot_angle);
                }
                failed = failed || pf;
            }
            failed = failed && (t > 1e-5f);
            if (failed) done = PARC_DONE_FAIL;
        }
        if (lane == 0) done_out[i] = done;
    }
}

}  // namespace parc

namespace parc {

// ---- DeepMimicEnv.reset (dm_env.py:567-592) --------------------------------------------------------
// Shared by the explicit-draw entry point and the device-sampled one: one group of G lanes per
// reset env writes the ids, evaluates the clip at the start time and copies ref -> char state.
template <int G>
__device__ __forceinline__ void reset_one_env(const parc_char_model_t& cm, const parc_motion_lib_t& lib, const parc_env_config_t& cfg,
                                              const parc_env_buffers_t& env, const LaneBody& lb, long long e, long long mid, long long tid,
                                              float mtime, float nx, float ny, const float* root_pos_offset, const float* root_rot_offset) {
    const int J = cm.num_bodies, D = cm.dof_size, C = lib.num_contacts;
    const int sub = threadIdx.x & (G - 1);
    const MotionMeta mm = load_motion_meta(lib.meta, mid);
    const FrameBlend fb = calc_frame_blend(mm, mtime);
    const PoseSample ps = eval_pose_lane(lib, lb, fb, J);
    // _move_to_motion_terrain (dm_env.py:554-565)
    const float* dmo = env.dm_motion_offsets + (mid * env.terrains_per_motion + tid) * 2;
    const float tox = __fsub_rn(__ldg(dmo), __ldg(env.env_offsets + e * 3));
    const float toy = __fsub_rn(__ldg(dmo + 1), __ldg(env.env_offsets + e * 3 + 1));
    const Vec3 rpos = make_vec3(__fadd_rn(ps.root_pos.x, tox), __fadd_rn(ps.root_pos.y, toy), ps.root_pos.z);
    if (sub == 0) {
        env.motion_ids[e] = mid;                                   // _reset_ref_motion, dm_env.py:504-506
        env.motion_terrain_ids[e] = tid;
        env.motion_time_offsets[e] = mtime;
        view_st_vec3(env.ref_root_pos, e, rpos);
        view_st_quat(env.ref_root_rot, e, ps.q);
        // _char_state_init_from_ref (mgdm_dm_util.py:89-100) + add_noise_to_char_state (:102-106)
        Vec3 cp = rpos;
        cp.x = __fadd_rn(cp.x, __fmul_rn(cfg.rand_root_pos_offset_scale, nx));
        cp.y = __fadd_rn(cp.y, __fmul_rn(cfg.rand_root_pos_offset_scale, ny));
        Quat cr = ps.q;
        if (root_pos_offset) { cp.x += root_pos_offset[e * 3]; cp.y += root_pos_offset[e * 3 + 1]; cp.z += root_pos_offset[e * 3 + 2]; }
        if (root_rot_offset) cr = quat_multiply(quat_from_f4(reinterpret_cast<const float4*>(root_rot_offset)[e]), cr);   // :112-113
        view_st_vec3(env.char_root_pos, e, cp);
        view_st_quat(env.char_root_rot, e, cr);
        env.timestep_buf[e] = 0;                                   // dm_env.py:586-590
        env.time_buf[e] = 0.f;
        env.done_buf[e] = PARC_DONE_NULL;
        if (env.actors_need_reset) env.actors_need_reset[e * env.actors_per_env] = 1;
    } else if (lb.active) {
        float* jr = env.ref_joint_rot.ptr + e * env.ref_joint_rot.row_stride + (sub - 1) * env.ref_joint_rot.mid_stride;
        const long long es = env.ref_joint_rot.elem_stride;
        jr[0] = ps.q.x; jr[es] = ps.q.y; jr[2 * es] = ps.q.z; jr[3 * es] = ps.q.w;
        float dd[3];
        const int nd = rot_to_dof_lane(lb, ps.q, dd);
        for (int k = 0; k < nd; ++k) {
            view_st(env.ref_dof_pos, e, lb.dof_idx + k, dd[k]);
            view_st(env.char_dof_pos, e, lb.dof_idx + k, dd[k]);
        }
    }
    if (sub < C && env.ref_contacts.ptr) view_st(env.ref_contacts, e, sub, ps.contact);
    const float* v = lib.vel + fb.idx0 * lib.vrec;
    for (int k = sub; k < 6 + D; k += G) {
        const float x = __ldg(v + k);
        if (k < 3) { view_st(env.ref_root_vel, e, k, x); view_st(env.char_root_vel, e, k, x); }
        else if (k < 6) { view_st(env.ref_root_ang_vel, e, k - 3, x); view_st(env.char_root_ang_vel, e, k - 3, x); }
        else { view_st(env.ref_dof_vel, e, k - 6, x); view_st(env.char_dof_vel, e, k - 6, x); }
    }
    if (lb.active) {
        for (int k = 0; k < 3; ++k) {
            if (env.char_rigid_body_vel.ptr) view_st3(env.char_rigid_body_vel, e, sub, k, 0.f);
            if (env.char_rigid_body_ang_vel.ptr) view_st3(env.char_rigid_body_ang_vel, e, sub, k, 0.f);
        }
    }
}

template <int G>
__global__ void __launch_bounds__(256) env_reset_kernel(const __grid_constant__ parc_char_model_t cm, const __grid_constant__ parc_motion_lib_t lib,
                                                        const __grid_constant__ parc_env_config_t cfg, const __grid_constant__ parc_env_buffers_t env,
                                                        const long long* __restrict__ env_ids, long long n, const long long* __restrict__ motion_ids,
                                                        const long long* __restrict__ terrain_ids, const float* __restrict__ motion_times,
                                                        const float* __restrict__ xy_noise, const float* __restrict__ root_pos_offset,
                                                        const float* __restrict__ root_rot_offset) {
    const LaneBody lb = load_lane_body<G>(cm);
    const long long gpb = blockDim.x / G;
    for (long long i = (long long)blockIdx.x * gpb + threadIdx.x / G; i < n; i += (long long)gridDim.x * gpb) {
        const float nx = xy_noise ? xy_noise[i * 2] : 0.f, ny = xy_noise ? xy_noise[i * 2 + 1] : 0.f;
        reset_one_env<G>(cm, lib, cfg, env, lb, env_ids[i], motion_ids[i], terrain_ids[i], motion_times[i], nx, ny, root_pos_offset, root_rot_offset);
    }
}

// ---- device-side RSI draws: Philox4x32-10 keyed by (seed, env id, step) -----------------------------
__device__ __forceinline__ uint4 philox4x32_10(uint4 ctr, uint2 key) {
    const unsigned M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        const unsigned hi0 = __umulhi(M0, ctr.x), lo0 = M0 * ctr.x;
        const unsigned hi1 = __umulhi(M1, ctr.z), lo1 = M1 * ctr.z;
        ctr = make_uint4(hi1 ^ ctr.y ^ key.x, lo1, hi0 ^ ctr.w ^ key.y, lo0);
        key.x += W0; key.y += W1;
    }
    return ctr;
}
__device__ __forceinline__ float u01(unsigned x) { return (float)(x >> 8) * (1.0f / 16777216.0f); }

// first m with cdf[m] > u * cdf[M-1]  (inverse-CDF form of a weighted draw with replacement)
__device__ __forceinline__ int sample_cdf(const float* __restrict__ cdf, int M, float u) {
    const float target = u * __ldg(cdf + M - 1);
    int lo = 0, hi = M - 1;
    while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (__ldg(cdf + mid) > target) hi = mid; else lo = mid + 1;
    }
    return lo;
}

template <int G>
__global__ void __launch_bounds__(256) env_reset_sampled_kernel(const __grid_constant__ parc_char_model_t cm, const __grid_constant__ parc_motion_lib_t lib,
                                                                const __grid_constant__ parc_env_config_t cfg, const __grid_constant__ parc_env_buffers_t env,
                                                                const float* __restrict__ cdf, int M, unsigned long long seed, unsigned long long step, int rand_reset,
                                                                const parc_reset_options_t opt) {
    const LaneBody lb = load_lane_body<G>(cm);
    const long long gpb = blockDim.x / G;
    const long long n = min((long long)*env.done_count, (long long)env.num_envs);
    for (long long i = (long long)blockIdx.x * gpb + threadIdx.x / G; i < n; i += (long long)gridDim.x * gpb) {
        const long long e = env.done_list[i];
        if (env.done_buf[e] == PARC_DONE_NULL) continue;   // never_done: counted for the EMA, not reset
        const uint2 key = make_uint2((unsigned)seed, (unsigned)(seed >> 32));
        const uint4 r0 = philox4x32_10(make_uint4((unsigned)e, (unsigned)(e >> 32), (unsigned)step, (unsigned)(step >> 32)), key);
        // demo mode: env_id % num_motions (dm_env.py:479-480); else sample_motions over the CDF (motion_lib.py:56-60)
        const long long mid = opt.demo_mode ? e % M : sample_cdf(cdf, M, u01(r0.x));
        const long long tid = (long long)(r0.y % (unsigned)env.terrains_per_motion);               // randint, dm_env.py:496
        const float len = __int_as_float(__ldg(lib.meta + 8 * mid + 2));
        // sample_time (motion_lib.py:62-71), or length * _motion_start_time_fraction[env] (dm_env.py:500-502; zeros when not set)
        const float mtime = rand_reset ? u01(r0.z) * len : (opt.motion_start_time_fraction ? __fmul_rn(len, opt.motion_start_time_fraction[e]) : 0.f);
        const uint4 r1 = philox4x32_10(make_uint4((unsigned)e, (unsigned)(e >> 32), (unsigned)step, (unsigned)(step >> 32) ^ 0x80000000u), key);
        const float nx = u01(r1.x) * 2.0f - 1.0f, ny = u01(r1.y) * 2.0f - 1.0f;                    // mgdm_dm_util.py:103
        reset_one_env<G>(cm, lib, cfg, env, lb, e, mid, tid, mtime, nx, ny, opt.root_pos_offset, opt.root_rot_offset);
    }
}

__global__ void sample_cdf_kernel(const float* __restrict__ cdf, int M, const float* __restrict__ u, long long n, long long* __restrict__ ids) {
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
        ids[i] = sample_cdf(cdf, M, u[i]);
}

}  // namespace parc

namespace parc {

// ---- fail-rate EMA (dm_env.py:648-660) --------------------------------------------------------------
// The reference walks the finished envs on the host IN ENV-ID ORDER and applies, per motion,
//   rate <- rate*(1-w) + w   (FAIL)      or      rate <- rate*(1-w)   (TIME / SUCC / motion end)
// as separately rounded fp32 operations. Order matters (w^2 = 1e-4 between a fail-then-pass and a
// pass-then-fail), so the updates of one motion are applied sequentially in ascending env id by
// one thread; different motions proceed in parallel. One block:
//   histogram per motion -> exclusive scan -> bucket scatter -> per-motion sort + apply -> RSI CDF.
constexpr int EMA_THREADS = 1024;
constexpr int EMA_ILP = 8;              // done-list entries a thread keeps in flight

// Block-wide exclusive scan of one value per thread (warp shuffles + one pass over the 32 warp totals).
template <typename T>
__device__ __forceinline__ T block_exclusive_scan(T v, T* warp_tot /* [32] smem */, T* total) {
    const int t = threadIdx.x, lane = t & 31, w = t >> 5;
    T inc = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const T o = __shfl_up_sync(0xffffffffu, inc, d);
        if (lane >= d) inc += o;
    }
    if (lane == 31) warp_tot[w] = inc;
    __syncthreads();
    if (w == 0) {
        const T x = lane < (int)(blockDim.x >> 5) ? warp_tot[lane] : T(0);
        T xi = x;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const T o = __shfl_up_sync(0xffffffffu, xi, d);
            if (lane >= d) xi += o;
        }
        warp_tot[lane] = xi - x;                   // exclusive prefix of the warp totals
        if (lane == 31 && total) *total = xi;
    }
    __syncthreads();
    const T r = warp_tot[w] + (inc - v);
    __syncthreads();
    return r;
}

// Ascending in-place sort of a[0..n) by NT cooperating threads (t = 0..NT-1), any n: the normalised bitonic network (every
// merge starts with a mirror step, all comparisons ascending), in which the virtual +inf elements past n never move.
// sync() separates the steps: __syncwarp for a warp, __syncthreads for the block.
template <typename SyncF>
__device__ __forceinline__ void bitonic_sort_ints(int* __restrict__ a, int n, int t, int NT, SyncF sync) {
    int np2 = 1;
    while (np2 < n) np2 <<= 1;
    const int half = np2 >> 1;
    for (int size = 2; size <= np2; size <<= 1) {
        const int hs = size >> 1;
        for (int idx = t; idx < half; idx += NT) {
            const int i = (idx / hs) * size + (idx % hs), p = i ^ (size - 1);
            if (p < n) { const int x = a[i], y = a[p]; if (x > y) { a[i] = y; a[p] = x; } }
        }
        sync();
        for (int stride = size >> 2; stride >= 1; stride >>= 1) {
            for (int idx = t; idx < half; idx += NT) {
                const int i = (idx / stride) * (2 * stride) + (idx % stride), p = i + stride;
                if (p < n) { const int x = a[i], y = a[p]; if (x > y) { a[i] = y; a[p] = x; } }
            }
            sync();
        }
    }
}
constexpr int EMA_THREAD_MAX = 48;      // segments up to this many finished envs: one thread sorts (insertion) and applies
constexpr int EMA_WARP_MAX = 2048;      // up to this many: one warp sorts, lane 0 applies; larger ones: the whole block sorts
constexpr int EMA_BIG_LIST = 128;       // block-sorted segments per launch (more than that fall back to the warp tier)

// Dynamic shared memory: [hist M | offs M] when `tables_in_smem`, then a bucket of `bucket_cap` entries; the global
// scratch (2M + N ints) takes whatever does not fit. A bucket entry is env_id << 1 | failed, so that sorting the
// entries sorts by env id and the apply loop needs no further global reads.
__global__ void __launch_bounds__(EMA_THREADS) fail_rate_ema_kernel(const __grid_constant__ parc_env_config_t cfg,
                                                                    const __grid_constant__ parc_env_buffers_t env, int M,
                                                                    float* __restrict__ cdf, int* __restrict__ scratch,
                                                                    int tables_in_smem, int bucket_cap) {
    extern __shared__ int ema_smem[];
    __shared__ int warp_i[32];
    __shared__ float warp_f[32];
    __shared__ int big_list[EMA_BIG_LIST];
    __shared__ int big_count;
    const int t = threadIdx.x, T = blockDim.x;
    if (t == 0) big_count = 0;
    const int n = min(*env.done_count, (int)env.num_envs);
    int* hist = tables_in_smem ? ema_smem : scratch;                 // [M]
    int* offs = hist + M;                                            // [M]
    int* bucket = n <= bucket_cap ? ema_smem + (tables_in_smem ? 2 * M : 0) : scratch + 2 * M;   // [n]
    if (n > 0) {
        for (int m = t; m < M; m += T) hist[m] = 0;
        __syncthreads();
        // pass 1: histogram of the finished envs' motions; EMA_ILP independent (done_list -> motion id, code) chains
        for (int i0 = t; i0 < n; i0 += T * EMA_ILP) {
            int e[EMA_ILP], m[EMA_ILP];
#pragma unroll
            for (int k = 0; k < EMA_ILP; ++k) e[k] = env.done_list[min(i0 + k * T, n - 1)];
#pragma unroll
            for (int k = 0; k < EMA_ILP; ++k) m[k] = (int)env.motion_ids[e[k]];
#pragma unroll
            for (int k = 0; k < EMA_ILP; ++k)
                if (i0 + k * T < n) atomicAdd(&hist[m[k]], 1);
        }
        __syncthreads();
        // exclusive scan over motions -> segment starts; hist becomes the per-motion cursor
        {
            const int chunk = (M + T - 1) / T;
            const int b = min(t * chunk, M), e = min(b + chunk, M);
            int sum = 0;
            for (int k = b; k < e; ++k) sum += hist[k];
            int run = block_exclusive_scan<int>(sum, warp_i, nullptr);
            for (int k = b; k < e; ++k) {
                const int v = hist[k];
                offs[k] = run; run += v; hist[k] = 0;
                if (v > EMA_WARP_MAX) { const int slot = atomicAdd(&big_count, 1); if (slot < EMA_BIG_LIST) big_list[slot] = k; }
            }
        }
        __syncthreads();
        // pass 2: scatter env_id << 1 | failed into the motion's segment
        for (int i0 = t; i0 < n; i0 += T * EMA_ILP) {
            int e[EMA_ILP], m[EMA_ILP], c[EMA_ILP];
#pragma unroll
            for (int k = 0; k < EMA_ILP; ++k) e[k] = env.done_list[min(i0 + k * T, n - 1)];
#pragma unroll
            for (int k = 0; k < EMA_ILP; ++k) { m[k] = (int)env.motion_ids[e[k]]; c[k] = env.done_code[e[k]] == 1; }
#pragma unroll
            for (int k = 0; k < EMA_ILP; ++k)
                if (i0 + k * T < n) bucket[offs[m[k]] + atomicAdd(&hist[m[k]], 1)] = (e[k] << 1) | c[k];
        }
        __syncthreads();
        const float w = cfg.ema_weight;
        const float keep = cfg.ema_keep;   // Python double (1.0 - 0.01) cast to fp32 when it meets the fp32 tensor
        // the reference's rule, applied to one motion's finished envs in ascending env id (dm_env.py:648-660)
        auto apply = [&](int m, const int* seg, int cnt) {
            float r = env.motion_id_fail_rates[m];
            for (int a = 0; a < cnt; ++a) {
                r = __fmul_rn(r, keep);
                if (seg[a] & 1) r = __fadd_rn(r, w);
            }
            env.motion_id_fail_rates[m] = r;
        };
        // tier 1: the many short segments, one thread per motion
        for (int m = t; m < M; m += T) {
            const int cnt = hist[m];
            if (cnt == 0 || cnt > EMA_THREAD_MAX) continue;
            int* seg = bucket + offs[m];
            for (int a = 1; a < cnt; ++a) {   // insertion sort by env id
                const int v = seg[a];
                int b = a - 1;
                while (b >= 0 && seg[b] > v) { seg[b + 1] = seg[b]; --b; }
                seg[b + 1] = v;
            }
            apply(m, seg, cnt);
        }
        // tier 3: giant segments (few clips, many envs), the whole block sorts one after the other
        const int n_big = min(big_count, EMA_BIG_LIST);
        for (int k = 0; k < n_big; ++k) {
            const int m = big_list[k];
            bitonic_sort_ints(bucket + offs[m], hist[m], t, T, [] { __syncthreads(); });
            if (t == 0) apply(m, bucket + offs[m], hist[m]);
        }
        // tier 2: one warp per medium segment
        {
            const int warp = t >> 5, lane = t & 31, nwarp = T >> 5;
            for (int m = warp; m < M; m += nwarp) {
                const int cnt = hist[m];
                if (cnt <= EMA_THREAD_MAX) continue;
                if (cnt > EMA_WARP_MAX) {
                    bool listed = false;
                    for (int k = 0; k < n_big; ++k) listed = listed || big_list[k] == m;
                    if (listed) continue;          // already done by the block
                }
                bitonic_sort_ints(bucket + offs[m], cnt, lane, 32, [] { __syncwarp(); });
                if (lane == 0) apply(m, bucket + offs[m], cnt);
            }
        }
        __syncthreads();
    }
    // RSI sampling CDF (dm_env.py:487-490): clamp(fail_rate, min_motion_weight) * motion_weight
    if (cdf) {
        const int chunk = (M + T - 1) / T;
        const int b = min(t * chunk, M), e = min(b + chunk, M);
        float s = 0.f;
        for (int k = b; k < e; ++k) s += fmaxf(env.motion_id_fail_rates[k], cfg.min_motion_weight) * env.motion_weights[k];
        float run = block_exclusive_scan<float>(s, warp_f, nullptr);
        for (int k = b; k < e; ++k) {
            run += fmaxf(env.motion_id_fail_rates[k], cfg.min_motion_weight) * env.motion_weights[k];
            cdf[k] = run;
        }
    }
}

// ---- simulator stand-in for benchmarks (see parc_b200_sim_standin) -----------------------------------
// G lanes per env: the lanes first write the root / dof columns (ref + noise), then run dof_to_rot + forward
// kinematics on that state, in registers, for the rigid body positions the simulator would report.
#ifndef PARC_STANDIN_THREADS
#define PARC_STANDIN_THREADS 128
#endif
#ifndef PARC_STANDIN_MINB
#define PARC_STANDIN_MINB 12
#endif
template <int G>
__global__ void __launch_bounds__(PARC_STANDIN_THREADS, PARC_STANDIN_MINB) sim_standin_kernel(const __grid_constant__ parc_char_model_t cm,
                                                          const __grid_constant__ parc_env_buffers_t env,
                                                          const float* __restrict__ noise, const long long first_env,
                                                          const long long n_rows) {
    const int J = cm.num_bodies, D = cm.dof_size;
    const int W = 13 + 2 * D + 3 * J, W1 = 13 + 2 * D;
    const LaneBody lb = load_lane_body<G>(cm);
    const long long gpb = blockDim.x / G;
    const long long gid0 = (long long)blockIdx.x * gpb + threadIdx.x / G;
    const long long warp_first = gid0 - ((threadIdx.x & 31) / G);
    for (long long base = warp_first; base < n_rows; base += (long long)gridDim.x * gpb) {
        const long long i = base + ((threadIdx.x & 31) / G);
        const bool live = i < n_rows;
        const long long e = first_env + (live ? i : 0);
        const float* n = noise + e * W;
        const Vec3 rp = view_ld_vec3(env.ref_root_pos, e) + make_vec3(n[0], n[1], n[2]);
        const Quat rr = quat_unit(quat_mul(view_ld_quat(env.ref_root_rot, e), make_quat(n[3], n[4], n[5], n[6])));
        if (live) {
            // root / dof columns: lane j of the env's group takes columns j, j + G, ...
            for (int c = threadIdx.x & (G - 1); c < W1; c += G) {
                if (c < 3) view_st(env.char_root_pos, e, c, c == 0 ? rp.x : (c == 1 ? rp.y : rp.z));
                else if (c < 7) { if (c == 3) view_st_quat(env.char_root_rot, e, rr); }
                else if (c < 10) view_st(env.char_root_vel, e, c - 7, view_ld(env.ref_root_vel, e, c - 7) + n[c]);
                else if (c < 13) view_st(env.char_root_ang_vel, e, c - 10, view_ld(env.ref_root_ang_vel, e, c - 10) + n[c]);
                else if (c < 13 + D) view_st(env.char_dof_pos, e, c - 13, view_ld(env.ref_dof_pos, e, c - 13) + n[c]);
                else view_st(env.char_dof_vel, e, c - 13 - D, view_ld(env.ref_dof_vel, e, c - 13 - D) + n[c]);
            }
        }
        Quat q = quat_identity();
        if (lb.active && lb.body > 0) q = dof_to_rot_lane(lb, [&](int k) { return view_ld(env.ref_dof_pos, e, k) + n[13 + k]; });
        Vec3 pos; Quat rot;
        fk_lanes<G>(lb, q, rp, rr, pos, rot);
        if (live && lb.active) {
            const float* nb = n + W1 + 3 * lb.body;
            view_st3(env.char_rigid_body_pos, e, lb.body, 0, pos.x + nb[0]);
            view_st3(env.char_rigid_body_pos, e, lb.body, 1, pos.y + nb[1]);
            view_st3(env.char_rigid_body_pos, e, lb.body, 2, pos.z + nb[2]);
        }
    }
}

__global__ void zero_int_kernel(int* p) { *p = 0; }

// Cross-rank fail rates: per motion, how many of this rank's envs finished the last step as FAIL (row 0) / otherwise
// (row 1), ADDED to counts [2, M]. Reads the step kernel's done list, so it costs a handful of atomics per finished env.
__global__ void outcome_counts_kernel(const __grid_constant__ parc_env_buffers_t env, const int num_motions, unsigned long long* __restrict__ counts) {
    const int n = *env.done_count;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const long long e = env.done_list[i];
        const int code = env.done_code[e];
        const long long m = env.motion_ids[e];
        if (code != 0 && m >= 0 && m < num_motions) atomicAdd(counts + (code == 1 ? 0 : num_motions) + m, 1ull);
    }
}

}  // namespace parc

#include "env_aux_api.cuh"

// ---- allocation-free variant of the fail-rate EMA (no counterpart in the reference) ---------------
// The scratch of fail_rate_ema_kernel lives in a file-scope device array instead of a caller buffer, so a caller that
// cannot allocate (CUDA graph capture, serving loops, harnesses that forbid cudaMalloc) needs nothing but its inputs and
// outputs. Every GPU that loads the library holds PARC_EMA_NOALLOC_MAX_INTS * 4 bytes (4 MiB) until exit. In host code
// the array's NAME is a host-side stand-in, not a device address: the address is looked up once per device with
// cudaGetSymbolAddress (that device current) and passed to the kernel.
static __device__ __align__(256) int32_t g_ema_noalloc_scratch[PARC_EMA_NOALLOC_MAX_INTS];

extern "C" int parc_b200_fail_rate_ema_noalloc(const parc_env_config_t* cfg, const parc_env_buffers_t* env, int32_t num_motions,
                                               float* cdf, void* stream) {
    PARC_REQUIRE(cfg && env && num_motions > 0, "fail_rate_ema_noalloc: bad arguments");
    PARC_REQUIRE(env->done_count && env->done_list && env->done_code && env->motion_ids && env->motion_id_fail_rates,
                 "fail_rate_ema_noalloc: done/fail-rate buffers not set");
    PARC_REQUIRE(cdf == nullptr || env->motion_weights, "fail_rate_ema_noalloc: motion_weights not set");
    PARC_REQUIRE(env->num_envs >= 0, "fail_rate_ema_noalloc: num_envs < 0");
    const long long need = 2ll * num_motions + env->num_envs;      // hist[M] | offs[M] | bucket[<= N]
    PARC_REQUIRE(need <= (long long)PARC_EMA_NOALLOC_MAX_INTS,
                 "fail_rate_ema_noalloc: 2 * num_motions + num_envs = %lld exceeds the built-in scratch of %d int32 "
                 "(PARC_EMA_NOALLOC_MAX_INTS); use parc_b200_fail_rate_ema with a caller-provided scratch",
                 need, (int)PARC_EMA_NOALLOC_MAX_INTS);
    // per-device set-up, done once with that device current: address of the scratch, opt-in shared memory of the kernel
    static int32_t* scratch_of_device[64] = {};
    int dev = -1;
    cudaError_t ce = cudaGetDevice(&dev);
    if (ce != cudaSuccess) { set_error("fail_rate_ema_noalloc: cudaGetDevice: %s", cudaGetErrorString(ce)); return PARC_ERR_CUDA; }
    PARC_REQUIRE(dev >= 0 && dev < 64, "fail_rate_ema_noalloc: device ordinal %d outside [0, 64)", dev);
    if (scratch_of_device[dev] == nullptr) {
        int32_t* p = nullptr;
        ce = cudaGetSymbolAddress((void**)&p, g_ema_noalloc_scratch);
        if (ce != cudaSuccess || p == nullptr) {
            set_error("fail_rate_ema_noalloc: cudaGetSymbolAddress: %s", cudaGetErrorString(ce));
            return PARC_ERR_CUDA;
        }
        ce = cudaFuncSetAttribute(fail_rate_ema_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(200 * 1024));
        if (ce != cudaSuccess) { set_error("fail_rate_ema_noalloc: %s", cudaGetErrorString(ce)); return PARC_ERR_CUDA; }
        scratch_of_device[dev] = p;
    }
    // same staging rule as parc_b200_fail_rate_ema: per-motion tables in shared memory when they fit in 64 KB, then as much
    // bucket as is left of 200 KB. The kernel zeroes the histogram it uses before counting, so the array's stale contents
    // (every call shares it and leaves its last state behind) are never read.
    const int tables = (size_t)num_motions * 8 <= 64 * 1024;
    const size_t table_bytes = tables ? (size_t)num_motions * 8 : 0;
    const long long cap = std::min<long long>(env->num_envs, (long long)((200 * 1024 - table_bytes) / 4));
    const size_t smem = table_bytes + (size_t)cap * 4;
    fail_rate_ema_kernel<<<1, EMA_THREADS, smem, as_stream(stream)>>>(*cfg, *env, num_motions, cdf, scratch_of_device[dev], tables, (int)cap);
    return check_launch("fail_rate_ema_noalloc");
}
