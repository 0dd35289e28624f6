// Fused post-physics step of the DeepMimic-style tracking env:
//   heights -> time -> reference motion -> observation -> reward -> termination
// i.e. IGEnv._post_physics_step (PARC/motion_tracker/envs/ig_env.py:368-377) minus PhysX, in ONE
// launch. One warp per env; a block owns a tile of TILE consecutive envs whose observation rows
// are assembled in shared memory and leave the SM as one bulk (TMA) shared->global copy.
#include "parc_common.cuh"
#include "parc_lanes.cuh"

namespace parc {

constexpr int TILE = 8;                 // envs (= warps) per block
constexpr int STEP_THREADS = TILE * 32;

// Column layout of one observation row (IGParkourEnv._compute_obs, ig_parkour_env.py:842-965;
// compute_char_obs ig_char_env.py:581-625; compute_tar_obs mgdm_dm_util.py:404-461).
struct ObsLayout {
    int root_h;        // 0/1 leading root height
    int char_dim;      // root_h + 12 + 6(J-1) + D + 3K
    int tar_step_dim;  // 9 + 6(J-1) + 3K
    int o_tar, o_tarc, o_charc, o_hf;
    int obs_dim;
};

__host__ __device__ inline ObsLayout make_layout(int J, int D, int C, const parc_env_config_t& cfg) {
    ObsLayout L;
    const int K = cfg.num_key_bodies, S = cfg.num_tar_steps;
    L.root_h = cfg.root_height_obs ? 1 : 0;
    L.char_dim = L.root_h + 12 + 6 * (J - 1) + D + 3 * K;
    L.tar_step_dim = 9 + 6 * (J - 1) + 3 * K;
    L.o_tar = L.char_dim;
    L.o_tarc = L.o_tar + (cfg.enable_tar_obs ? S * L.tar_step_dim : 0);
    L.o_charc = L.o_tarc + ((cfg.use_contact_info && cfg.enable_tar_obs) ? S * C : 0);
    L.o_hf = L.o_charc + (cfg.use_contact_info ? J : 0);
    L.obs_dim = L.o_hf + cfg.num_ray_points;
    return L;
}

// bulk shared->global copy (TMA, SASS UBLKCP); bytes % 16 == 0, both addresses 16-byte aligned
__device__ __forceinline__ void bulk_store_s2g(void* gdst, const void* ssrc, unsigned bytes) {
    unsigned s = (unsigned)__cvta_generic_to_shared(ssrc);
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst), "r"(s), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_commit_wait_read() {
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
}
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

__device__ __forceinline__ int find_key_slot(const parc_env_config_t& cfg, int body) {
    int slot = -1;
    for (int k = 0; k < cfg.num_key_bodies; ++k)
        if (cfg.key_body_ids[k] == body) slot = k;   // a body listed twice keeps the last slot (and see write loop)
    return slot;
}

// rotate_2d_vec (torch_util.py:650-662) + root xy, products rounded separately (bit-exact cells)
__device__ __forceinline__ void ray_world(float px, float py, float c, float s, float rx, float ry, float& wx, float& wy) {
    wx = __fadd_rn(__fsub_rn(__fmul_rn(px, c), __fmul_rn(py, s)), rx);
    wy = __fadd_rn(__fadd_rn(__fmul_rn(px, s), __fmul_rn(py, c)), ry);
}

template <int G>
__global__ void __launch_bounds__(STEP_THREADS) env_step_kernel(
    const __grid_constant__ parc_char_model_t cm, const __grid_constant__ parc_motion_lib_t lib,
    const __grid_constant__ parc_terrain_t terrain, const __grid_constant__ parc_env_config_t cfg,
    const __grid_constant__ parc_env_buffers_t env, const unsigned parts, const long long* __restrict__ env_ids,
    const long long n_rows_host, const int bulk_ok, const int* __restrict__ env_ids32, const int* __restrict__ n_rows_dev) {
    extern __shared__ __align__(16) float smem[];
    // row count either from the host or (graph-captured reset path) from the device-side done counter
    const long long n_rows = n_rows_dev ? min((long long)*n_rows_dev, n_rows_host) : n_rows_host;
    if ((long long)blockIdx.x * TILE >= n_rows) return;
    constexpr int HALVES = 32 / G;
    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    const int h = lane / G;          // which pose of the round this lane works on
    const int sub = lane & (G - 1);  // body owned by this lane
    const int J = cm.num_bodies, D = cm.dof_size, C = lib.num_contacts;
    const int S = cfg.enable_tar_obs ? cfg.num_tar_steps : 0;
    const int K = cfg.num_key_bodies, P = cfg.num_ray_points;
    const ObsLayout L = make_layout(J, D, C, cfg);
    const LaneBody lb = load_lane_body<G>(cm);
    const int my_key = lb.active ? find_key_slot(cfg, lb.body) : -1;

    const long long row = (long long)blockIdx.x * TILE + warp;
    const bool live = row < n_rows;
    const long long e = live ? (env_ids ? env_ids[row] : (env_ids32 ? (long long)env_ids32[row] : row)) : 0;
    float* srow = smem + warp * L.obs_dim;

    const bool do_obs = parts & PARC_STEP_OBS;
    const bool do_ref = parts & PARC_STEP_REF;
    const bool do_reward = parts & PARC_STEP_REWARD;
    const bool do_done = parts & PARC_STEP_DONE;

    if (live) {
        // ---- uniform per-env state (broadcast loads) ---------------------------------------
        const Vec3 c_pos = view_ld_vec3(env.char_root_pos, e);
        const Quat c_rot = view_ld_quat(env.char_root_rot, e);
        const Vec3 c_vel = view_ld_vec3(env.char_root_vel, e);
        const Vec3 c_avel = view_ld_vec3(env.char_root_ang_vel, e);
        const float eox = __ldg(env.env_offsets + e * 3), eoy = __ldg(env.env_offsets + e * 3 + 1), eoz = __ldg(env.env_offsets + e * 3 + 2);

        // ---- _refresh_obs_hfs (ig_parkour_env.py:518-533, mgdm_dm_util.py:128-145) ----------
        if (parts & PARC_STEP_HEIGHTS) {
            const float gx = __fadd_rn(c_pos.x, eox), gy = __fadd_rn(c_pos.y, eoy), gz = __fadd_rn(c_pos.z, eoz);
            const float heading = calc_heading(c_rot);
            const float sh = sinf(heading), ch = cosf(heading);
            for (int p = lane; p < P; p += 32) {
                const float2 pt = __ldg(reinterpret_cast<const float2*>(env.ray_xy_points) + p);
                float wx, wy;
                ray_world(pt.x, pt.y, ch, sh, gx, gy, wx, wy);
                const int ix = grid_index(wx, terrain.min_x, terrain.dx, terrain.dim_x);
                const int iy = grid_index(wy, terrain.min_y, terrain.dy, terrain.dim_y);
                const float hv = __ldg(terrain.hf + (long long)ix * terrain.dim_y + iy);
                const float v = fminf(fmaxf(__fsub_rn(hv, gz), cfg.min_obs_h), cfg.max_obs_h);
                if (do_obs) srow[L.o_hf + p] = v;
                if (env.ray_hfs) env.ray_hfs[e * P + p] = v;
            }
        } else if (do_obs) {
            for (int p = lane; p < P; p += 32) srow[L.o_hf + p] = env.ray_hfs[e * P + p];
        }

        // ---- _update_time (ig_env.py:391-394) -----------------------------------------------
        float time;
        if (parts & PARC_STEP_TIME) {
            const int ts = env.timestep_buf[e] + 1;
            time = __fmul_rn(cfg.timestep, (float)ts);
            __syncwarp();   // every lane has read the old step count before lane 0 replaces it
            if (lane == 0) { env.timestep_buf[e] = ts; env.time_buf[e] = time; }
        } else {
            time = env.time_buf[e];
        }

        const long long mid = env.motion_ids[e];
        const long long tid = env.motion_terrain_ids[e];
        const MotionMeta mm = load_motion_meta(lib.meta, mid);
        const float t0 = __fadd_rn(time, env.motion_time_offsets[e]);              // dm_env.py:547-552
        // _move_to_motion_terrain (dm_env.py:554-565): offset = dm_offsets[m, t] - env_offsets[:, 0:2]
        const float* dmo = env.dm_motion_offsets + (mid * env.terrains_per_motion + tid) * 2;
        const float tox = __fsub_rn(__ldg(dmo), eox), toy = __fsub_rn(__ldg(dmo + 1), eoy);

        const Quat hinv = calc_heading_quat_inv(c_rot);   // heading frame of the CHARACTER root

        // per-lane values that outlive the pose loop
        Quat ref_q = quat_identity(), char_q = quat_identity();
        Vec3 ref_bpos = make_vec3(0.f, 0.f, 0.f);
        Vec3 ref_rpos = make_vec3(0.f, 0.f, 0.f);
        Quat ref_rrot = quat_identity();
        float ref_contact = 0.f;
        long long ref_idx0 = 0;

        const int n_poses = S + 2;   // 0 = reference pose, 1..S = target steps, S+1 = simulated character
        const bool need_poses = do_obs || do_ref || do_reward || do_done;
        const int n_rounds = need_poses ? (n_poses + HALVES - 1) / HALVES : 0;
        for (int r = 0; r < n_rounds; ++r) {
            const int pose = r * HALVES + h;
            const bool is_ref = pose == 0;
            const bool is_tar = pose >= 1 && pose <= S;
            const bool is_char = pose == S + 1;
            Quat q = quat_identity();
            Vec3 rpos = c_pos;
            float contact = 0.f;
            long long idx0 = 0;
            if (is_ref || is_tar) {
                // fetch_tar_obs_data (mgdm_dm_util.py:221-244): motion_times + timestep * tar_obs_steps
                const float t = is_ref ? t0 : __fadd_rn(t0, cfg.tar_dt[pose - 1]);
                const FrameBlend fb = calc_frame_blend(mm, t);
                const PoseSample ps = eval_pose_lane(lib, lb, fb, J);
                q = ps.q;
                rpos = make_vec3(__fadd_rn(ps.root_pos.x, tox), __fadd_rn(ps.root_pos.y, toy), ps.root_pos.z);
                contact = ps.contact;
                idx0 = fb.idx0;
            } else if (is_char) {
                if (sub == 0) q = c_rot;
                else if (lb.active) q = dof_to_rot_lane(lb, [&](int k) { return view_ld(env.char_dof_pos, e, k); });
            }
            // forward kinematics of this pose (root rotation lives in the group's lane 0)
            const Quat rrot = group_bcast<G>(q, 0);
            Vec3 bpos; Quat brot;
            fk_lanes<G>(lb, q, rpos, rrot, bpos, brot);

            if (is_ref) {
                ref_q = q; ref_bpos = bpos; ref_rpos = rpos; ref_rrot = rrot; ref_contact = contact; ref_idx0 = idx0;
                if (do_ref) {
                    // DeepMimicEnv._update_ref_motion (dm_env.py:523-545)
                    if (sub == 0) {
                        view_st_vec3(env.ref_root_pos, e, rpos);
                        view_st_quat(env.ref_root_rot, e, q);
                    } else if (lb.active) {
                        float* jr = env.ref_joint_rot.ptr + e * env.ref_joint_rot.row_stride + (sub - 1) * env.ref_joint_rot.mid_stride;
                        const long long es = env.ref_joint_rot.elem_stride;
                        jr[0] = q.x; jr[es] = q.y; jr[2 * es] = q.z; jr[3 * es] = q.w;
                        float dd[3];
                        const int nd = rot_to_dof_lane(lb, q, dd);
                        for (int k = 0; k < nd; ++k) view_st(env.ref_dof_pos, e, lb.dof_idx + k, dd[k]);
                    }
                    if (lb.active) {
                        view_st3(env.ref_body_pos, e, sub, 0, bpos.x);
                        view_st3(env.ref_body_pos, e, sub, 1, bpos.y);
                        view_st3(env.ref_body_pos, e, sub, 2, bpos.z);
                    }
                    if (sub < C && env.ref_contacts.ptr) view_st(env.ref_contacts, e, sub, contact);
                    const float* v = lib.vel + idx0 * lib.vrec;
                    for (int k = sub; k < 6 + D; k += G) {
                        const float x = __ldg(v + k);
                        if (k < 3) view_st(env.ref_root_vel, e, k, x);
                        else if (k < 6) view_st(env.ref_root_ang_vel, e, k - 3, x);
                        else view_st(env.ref_dof_vel, e, k - 6, x);
                    }
                }
            } else if (is_tar) {
                if (do_obs) {
                    // compute_tar_obs (mgdm_dm_util.py:404-461) for target step `pose - 1`
                    float* o = srow + L.o_tar + (pose - 1) * L.tar_step_dim;
                    Vec3 rp_obs = rpos - c_pos;
                    if (!cfg.global_obs) rp_obs = quat_rotate(hinv, rp_obs);
                    Quat oq = q;
                    if (sub == 0) {
                        o[0] = rp_obs.x; o[1] = rp_obs.y; o[2] = rp_obs.z;
                        if (!cfg.global_obs) oq = quat_mul(hinv, q);
                    }
                    if (lb.active) {
                        float tn[6];
                        quat_to_tan_norm(oq, tn);
                        float* d = o + 3 + 6 * sub;
#pragma unroll
                        for (int k = 0; k < 6; ++k) d[k] = tn[k];
                        if (my_key >= 0) {
                            Vec3 kp = bpos - rpos;
                            if (!cfg.global_obs) kp = quat_rotate(hinv, kp) + rp_obs;
                            for (int k = 0; k < K; ++k)
                                if (cfg.key_body_ids[k] == lb.body) {
                                    float* kd = o + 9 + 6 * (J - 1) + 3 * k;
                                    kd[0] = kp.x; kd[1] = kp.y; kd[2] = kp.z;
                                }
                        }
                    }
                    if (cfg.use_contact_info && sub < C) srow[L.o_tarc + (pose - 1) * C + sub] = contact;
                }
            } else if (is_char) {
                char_q = q;
                if (do_obs) {
                    // compute_char_obs (ig_char_env.py:581-625)
                    float* o = srow + L.root_h;
                    if (sub == 0) {
                        if (L.root_h) srow[0] = c_pos.z;
                        Quat oq = cfg.global_obs ? c_rot : quat_mul(hinv, c_rot);
                        float tn[6];
                        quat_to_tan_norm(oq, tn);
#pragma unroll
                        for (int k = 0; k < 6; ++k) o[k] = tn[k];
                        Vec3 v = cfg.global_obs ? c_vel : quat_rotate(hinv, c_vel);
                        Vec3 w = cfg.global_obs ? c_avel : quat_rotate(hinv, c_avel);
                        o[6] = v.x; o[7] = v.y; o[8] = v.z; o[9] = w.x; o[10] = w.y; o[11] = w.z;
                    } else if (lb.active) {
                        float tn[6];
                        quat_to_tan_norm(q, tn);
                        float* d = o + 12 + 6 * (sub - 1);
#pragma unroll
                        for (int k = 0; k < 6; ++k) d[k] = tn[k];
                    }
                    for (int k = sub; k < D; k += G) o[12 + 6 * (J - 1) + k] = view_ld(env.char_dof_vel, e, k);
                    if (my_key >= 0) {
                        Vec3 kp = bpos - c_pos;
                        if (!cfg.global_obs) kp = quat_rotate(hinv, kp);
                        for (int k = 0; k < K; ++k)
                            if (cfg.key_body_ids[k] == lb.body) {
                                float* kd = o + 12 + 6 * (J - 1) + D + 3 * k;
                                kd[0] = kp.x; kd[1] = kp.y; kd[2] = kp.z;
                            }
                    }
                    if (cfg.use_contact_info && lb.active) {
                        // _get_char_contact_state (ig_parkour_env.py:655-662): ||f|| > eps -> 0/1
                        const float fx = view_ld3(env.char_contact_forces, e, sub, 0), fy = view_ld3(env.char_contact_forces, e, sub, 1),
                                    fz = view_ld3(env.char_contact_forces, e, sub, 2);
                        srow[L.o_charc + sub] = (sqrtf(fx * fx + fy * fy + fz * fz) > cfg.contact_eps) ? 1.f : 0.f;
                    }
                }
            }
        }

        // Bring the reference pose (group 0, round 0) and the character pose (group hc) to every lane.
        constexpr int ref_lane_base = 0;
        const int char_lane_base = ((S + 1) % HALVES) * G;
        if (do_reward || do_done) {
            ref_q = make_quat(__shfl_sync(FULL_MASK, ref_q.x, ref_lane_base + sub), __shfl_sync(FULL_MASK, ref_q.y, ref_lane_base + sub),
                              __shfl_sync(FULL_MASK, ref_q.z, ref_lane_base + sub), __shfl_sync(FULL_MASK, ref_q.w, ref_lane_base + sub));
            char_q = make_quat(__shfl_sync(FULL_MASK, char_q.x, char_lane_base + sub), __shfl_sync(FULL_MASK, char_q.y, char_lane_base + sub),
                               __shfl_sync(FULL_MASK, char_q.z, char_lane_base + sub), __shfl_sync(FULL_MASK, char_q.w, char_lane_base + sub));
            ref_bpos = make_vec3(__shfl_sync(FULL_MASK, ref_bpos.x, ref_lane_base + sub), __shfl_sync(FULL_MASK, ref_bpos.y, ref_lane_base + sub),
                                 __shfl_sync(FULL_MASK, ref_bpos.z, ref_lane_base + sub));
            ref_rpos = make_vec3(__shfl_sync(FULL_MASK, ref_rpos.x, 0), __shfl_sync(FULL_MASK, ref_rpos.y, 0), __shfl_sync(FULL_MASK, ref_rpos.z, 0));
            ref_rrot = make_quat(__shfl_sync(FULL_MASK, ref_rrot.x, 0), __shfl_sync(FULL_MASK, ref_rrot.y, 0),
                                 __shfl_sync(FULL_MASK, ref_rrot.z, 0), __shfl_sync(FULL_MASK, ref_rrot.w, 0));
            ref_contact = __shfl_sync(FULL_MASK, ref_contact, ref_lane_base + sub);
            ref_idx0 = __shfl_sync(FULL_MASK, ref_idx0, 0);
        }
        const bool first_group = h == 0;   // per-body terms are summed over one group only

        // simulator rigid-body position of the lane's body (reward key bodies + termination)
        Vec3 sim_bpos = make_vec3(0.f, 0.f, 0.f);
        Vec3 sim_b0 = make_vec3(0.f, 0.f, 0.f);
        if (do_reward || do_done) {
            if (lb.active)
                sim_bpos = make_vec3(view_ld3(env.char_rigid_body_pos, e, sub, 0), view_ld3(env.char_rigid_body_pos, e, sub, 1),
                                     view_ld3(env.char_rigid_body_pos, e, sub, 2));
            sim_b0 = make_vec3(view_ld3(env.char_rigid_body_pos, e, 0, 0), view_ld3(env.char_rigid_body_pos, e, 0, 1),
                               view_ld3(env.char_rigid_body_pos, e, 0, 2));
        }

        // ---- _update_reward (ig_parkour_env.py:984-1044) ------------------------------------
        if (do_reward) {
            const float* rv = lib.vel + ref_idx0 * lib.vrec;
            const Vec3 r_vel = make_vec3(__ldg(rv), __ldg(rv + 1), __ldg(rv + 2));
            const Vec3 r_avel = make_vec3(__ldg(rv + 3), __ldg(rv + 4), __ldg(rv + 5));
            // compute_deepmimic_reward (mgdm_dm_util.py:269-332)
            float pose_t = 0.f;
            if (first_group && lb.active && sub > 0) {
                const float a = quat_diff_angle(char_q, ref_q);
                pose_t = cfg.joint_err_w[sub - 1] * a * a;
            }
            float vel_t = 0.f;
            for (int k = lane; k < D; k += 32) {
                const float dv = __ldg(rv + 6 + k) - view_ld(env.char_dof_vel, e, k);
                vel_t += cfg.dof_err_w[k] * dv * dv;
            }
            Vec3 rpd = ref_rpos - c_pos;
            if (!cfg.track_root) { rpd.x = 0.f; rpd.y = 0.f; }
            if (!cfg.track_root_h) rpd.z = 0.f;
            const float root_pos_err = rpd.x * rpd.x + rpd.y * rpd.y + rpd.z * rpd.z;
            Quat a_rot = c_rot, b_rot = ref_rrot;
            Vec3 a_vel = c_vel, b_vel = r_vel, a_av = c_avel, b_av = r_avel;
            Vec3 kp = sim_bpos - c_pos, tkp = ref_bpos - ref_rpos;
            if (!cfg.track_root) {   // convert_to_local (mgdm_dm_util.py:246-267)
                const Quat hr = calc_heading_quat_inv(ref_rrot);
                a_rot = quat_mul(hinv, c_rot); b_rot = quat_mul(hr, ref_rrot);
                a_vel = quat_rotate(hinv, c_vel); b_vel = quat_rotate(hr, r_vel);
                a_av = quat_rotate(hinv, c_avel); b_av = quat_rotate(hr, r_avel);
                kp = quat_rotate(hinv, kp); tkp = quat_rotate(hr, tkp);
            }
            float key_t = 0.f;
            if (first_group && my_key >= 0) {
                const Vec3 d = tkp - kp;
                int mult = 0;
                for (int k = 0; k < K; ++k) mult += (cfg.key_body_ids[k] == lb.body);
                key_t = mult * (d.x * d.x + d.y * d.y + d.z * d.z);
            }
            float con_t = 0.f;
            if (cfg.use_contact_info && first_group && sub < C) {
                // compute_contact_reward (mgdm_dm_util.py:497-518); contact bodies == all bodies
                const float fx = view_ld3(env.char_contact_forces, e, sub, 0), fy = view_ld3(env.char_contact_forces, e, sub, 1),
                            fz = view_ld3(env.char_contact_forces, e, sub, 2);
                const float f = fminf(sqrtf(fx * fx + fy * fy + fz * fz), 1.0f);
                float cr = -(1.0f - ref_contact) * f;
                cr += ref_contact * f;
                con_t = cfg.contact_weights[sub] * cr;
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                pose_t += __shfl_xor_sync(FULL_MASK, pose_t, o);
                vel_t += __shfl_xor_sync(FULL_MASK, vel_t, o);
                key_t += __shfl_xor_sync(FULL_MASK, key_t, o);
                con_t += __shfl_xor_sync(FULL_MASK, con_t, o);
            }
            if (lane == 0) {
                float rra = quat_diff_angle(a_rot, b_rot);
                const float root_rot_err = rra * rra;
                const Vec3 dv = b_vel - a_vel, dw = b_av - a_av;
                const float root_vel_err = dv.x * dv.x + dv.y * dv.y + dv.z * dv.z;
                const float root_av_err = dw.x * dw.x + dw.y * dw.y + dw.z * dw.z;
                const float pose_r = expf(-0.25f * pose_t);
                const float vel_r = expf(-0.01f * vel_t);
                const float root_pose_r = expf(-5.0f * (root_pos_err + 0.1f * root_rot_err));
                const float root_vel_r = expf(-1.0f * (root_vel_err + 0.1f * root_av_err));
                const float key_r = expf(-10.0f * key_t);
                float total = cfg.pose_w * pose_r + cfg.vel_w * vel_r + cfg.root_pos_w * root_pose_r + cfg.root_vel_w * root_vel_r + cfg.key_pos_w * key_r;
                float pen = 0.f;
                if (cfg.use_contact_info) { pen = con_t / (float)C; total += pen; }
                env.reward_buf[e] = total;
                if (env.reward_terms) {
                    float4* t4 = reinterpret_cast<float4*>(env.reward_terms + e * 8);
                    t4[0] = make_float4(pose_r, vel_r, root_pose_r, root_vel_r);
                    t4[1] = make_float4(key_r, pen, total, 0.f);
                }
            }
        }

        // ---- _update_done (ig_parkour_env.py:967-982; mgdm_dm_util.py:147-172, 334-402; dm_env.py:628-665)
        if (do_done) {
            int done = PARC_DONE_NULL;
            if (time >= cfg.episode_length) done = PARC_DONE_TIME;
            if (cfg.enable_early_termination) {
                bool failed = false;
                if (cfg.num_fall_contact_bodies > 0) {
                    bool fc = false, fh = false;
                    if (first_group && lb.active && !((cfg.fall_contact_body_mask >> sub) & 1u)) {
                        const float fx = view_ld3(env.char_contact_forces, e, sub, 0), fy = view_ld3(env.char_contact_forces, e, sub, 1),
                                    fz = view_ld3(env.char_contact_forces, e, sub, 2);
                        fc = fabsf(fx) > 0.1f || fabsf(fy) > 0.1f || fabsf(fz) > 0.1f;
                        const int ix = grid_index(__fadd_rn(sim_bpos.x, eox), terrain.min_x, terrain.dx, terrain.dim_x);
                        const int iy = grid_index(__fadd_rn(sim_bpos.y, eoy), terrain.min_y, terrain.dy, terrain.dim_y);
                        const float th = __fadd_rn(__ldg(terrain.hf + (long long)ix * terrain.dim_y + iy), cfg.termination_height);
                        fh = sim_bpos.z < th;
                    }
                    failed = __any_sync(FULL_MASK, fc) && __any_sync(FULL_MASK, fh);
                }
                if (cfg.pose_termination) {
                    bool pf = false;
                    if (first_group && lb.active && sub > 0) {
                        // (tar_body - tar_root) - (body - root), squared norm summed left to right
                        const Vec3 a = make_vec3(__fsub_rn(sim_bpos.x, sim_b0.x), __fsub_rn(sim_bpos.y, sim_b0.y), __fsub_rn(sim_bpos.z, sim_b0.z));
                        const Vec3 b = make_vec3(__fsub_rn(ref_bpos.x, ref_rpos.x), __fsub_rn(ref_bpos.y, ref_rpos.y), __fsub_rn(ref_bpos.z, ref_rpos.z));
                        const Vec3 d = make_vec3(__fsub_rn(b.x, a.x), __fsub_rn(b.y, a.y), __fsub_rn(b.z, a.z));
                        const float td = cfg.pose_termination_dist[sub - 1];
                        pf = sqnorm3_seq(d) > __fmul_rn(td, td);
                    }
                    pf = __any_sync(FULL_MASK, pf);
                    if (cfg.track_root) {
                        const Vec3 d = make_vec3(__fsub_rn(sim_b0.x, ref_rpos.x), __fsub_rn(sim_b0.y, ref_rpos.y), __fsub_rn(sim_b0.z, ref_rpos.z));
                        pf = pf || (sqnorm3_seq(d) > cfg.root_pos_termination_dist);   // host passes fp32(dist*dist)
                        pf = pf || (fabsf(quat_diff_angle(c_rot, ref_rrot)) > cfg.root_rot_termination_angle);
                    }
                    failed = failed || pf;
                }
                failed = failed && (time > 1e-5f);
                if (failed) done = PARC_DONE_FAIL;
            }
            // DeepMimicEnv.update_done tail (dm_env.py:638-665)
            const bool motion_end = (t0 >= mm.length) && (mm.loop_mode != PARC_LOOP_WRAP);
            int code = 0;
            if (done == PARC_DONE_FAIL) code = 1;
            else if (done != PARC_DONE_NULL || motion_end) code = 2;
            if (motion_end) done = PARC_DONE_FAIL;
            if (cfg.never_done) done = PARC_DONE_NULL;
            if (lane == 0) {
                env.done_buf[e] = done;
                if (env.done_code) env.done_code[e] = (unsigned char)code;
                if (code != 0 && env.done_list) {
                    const int slot = atomicAdd(env.done_count, 1);
                    if (slot < env.num_envs) env.done_list[slot] = (int)e;   // a stale (un-zeroed) count must not overrun the list
                }
            }
        }
    }

    // ---- observation rows leave the SM -------------------------------------------------------
    if (do_obs) {
        const long long row0 = (long long)blockIdx.x * TILE;
        const long long rows = min((long long)TILE, n_rows - row0);
        const long long total = rows * L.obs_dim;
        if (bulk_ok && ((total & 3) == 0)) {
            fence_async_smem();
            __syncthreads();
            if (threadIdx.x == 0) {
                float* g = env.obs_buf + row0 * env.obs_row_stride;
                unsigned bytes = (unsigned)(total * 4);
                unsigned off = 0;
                while (off < bytes) {
                    unsigned chunk = min(bytes - off, 32768u);
                    bulk_store_s2g(reinterpret_cast<char*>(g) + off, reinterpret_cast<const char*>(smem) + off, chunk);
                    off += chunk;
                }
                bulk_commit_wait_read();
            }
        } else {
            __syncwarp();
            if (live) {
                float* g = env.obs_buf + e * env.obs_row_stride;
                for (int k = lane; k < L.obs_dim; k += 32) g[k] = srow[k];
            }
        }
    }
}

}  // namespace parc

using namespace parc;

static int validate_env(const parc_char_model_t* m, const parc_motion_lib_t* lib, const parc_env_config_t* cfg,
                        const parc_env_buffers_t* env, const char* who) {
    if (!model_ok(m, who)) return PARC_ERR_INVALID_ARG;
    PARC_REQUIRE(lib && lib->interp && lib->vel && lib->meta, "%s: motion library not packed", who);
    PARC_REQUIRE(cfg && env, "%s: NULL config/buffers", who);
    PARC_REQUIRE(cfg->num_tar_steps >= 0 && cfg->num_tar_steps <= PARC_MAX_TAR_STEPS, "%s: num_tar_steps %d outside [0,%d]", who,
                 cfg->num_tar_steps, PARC_MAX_TAR_STEPS);
    PARC_REQUIRE(cfg->num_key_bodies >= 0 && cfg->num_key_bodies <= PARC_MAX_KEY_BODIES, "%s: num_key_bodies %d outside [0,%d]", who,
                 cfg->num_key_bodies, PARC_MAX_KEY_BODIES);
    for (int k = 0; k < cfg->num_key_bodies; ++k)
        PARC_REQUIRE(cfg->key_body_ids[k] >= 0 && cfg->key_body_ids[k] < m->num_bodies, "%s: key body id %d out of range", who, cfg->key_body_ids[k]);
    PARC_REQUIRE(!cfg->use_contact_info || lib->num_contacts == m->num_bodies,
                 "%s: use_contact_info needs a library with one contact label per body (got %d for %d bodies)", who,
                 lib->num_contacts, m->num_bodies);
    PARC_REQUIRE(env->num_envs >= 0, "%s: num_envs < 0", who);
    return PARC_OK;
}

extern "C" int parc_b200_obs_dim(const parc_char_model_t* m, const parc_env_config_t* cfg) {
    if (!model_ok(m, "obs_dim") || cfg == nullptr) return PARC_ERR_INVALID_ARG;
    // contact labels: one per body (validated at step time)
    return make_layout(m->num_bodies, m->dof_size, m->num_bodies, *cfg).obs_dim;
}

extern "C" int parc_b200_env_step(const parc_char_model_t* m, const parc_motion_lib_t* lib, const parc_terrain_t* terrain,
                                  const parc_env_config_t* cfg, const parc_env_buffers_t* env, uint32_t parts,
                                  const int64_t* env_ids, int64_t n_sub, void* stream) {
    int rc = validate_env(m, lib, cfg, env, "env_step");
    if (rc) return rc;
    const long long n_rows = env_ids ? n_sub : env->num_envs;
    PARC_REQUIRE(n_rows >= 0, "env_step: negative row count");
    if (n_rows == 0 || parts == 0) return PARC_OK;
    PARC_REQUIRE(terrain && terrain->hf && terrain->dim_x > 0 && terrain->dim_y > 0, "env_step: empty terrain");
    PARC_REQUIRE(env->char_root_pos.ptr && env->char_root_rot.ptr && env->char_root_vel.ptr && env->char_root_ang_vel.ptr &&
                     env->char_dof_pos.ptr && env->char_dof_vel.ptr,
                 "env_step: simulator state views not set");
    PARC_REQUIRE(env->motion_ids && env->motion_terrain_ids && env->motion_time_offsets && env->time_buf && env->env_offsets &&
                     env->dm_motion_offsets,
                 "env_step: bookkeeping buffers not set");
    if (parts & PARC_STEP_TIME) PARC_REQUIRE(env->timestep_buf, "env_step: timestep_buf not set");
    if (parts & (PARC_STEP_HEIGHTS | PARC_STEP_OBS)) PARC_REQUIRE(cfg->num_ray_points == 0 || env->ray_xy_points, "env_step: ray_xy_points not set");
    if ((parts & PARC_STEP_HEIGHTS) && !(parts & PARC_STEP_OBS)) PARC_REQUIRE(env->ray_hfs, "env_step: heights requested without obs need ray_hfs");
    if ((parts & PARC_STEP_OBS) && !(parts & PARC_STEP_HEIGHTS)) PARC_REQUIRE(env->ray_hfs || cfg->num_ray_points == 0, "env_step: obs without heights reads ray_hfs");
    if (parts & PARC_STEP_OBS) PARC_REQUIRE(env->obs_buf, "env_step: obs_buf not set");
    if (parts & PARC_STEP_REF)
        PARC_REQUIRE(env->ref_root_pos.ptr && env->ref_root_rot.ptr && env->ref_root_vel.ptr && env->ref_root_ang_vel.ptr &&
                         env->ref_body_pos.ptr && env->ref_joint_rot.ptr && env->ref_dof_pos.ptr && env->ref_dof_vel.ptr,
                     "env_step: ref_* views not set");
    if (parts & (PARC_STEP_REWARD | PARC_STEP_DONE))
        PARC_REQUIRE(env->char_rigid_body_pos.ptr && env->char_contact_forces.ptr, "env_step: rigid body / contact views not set");
    if (parts & PARC_STEP_REWARD) PARC_REQUIRE(env->reward_buf, "env_step: reward_buf not set");
    if (parts & PARC_STEP_DONE) PARC_REQUIRE(env->done_buf && (!env->done_list || env->done_count), "env_step: done buffers not set");
    if (cfg->use_contact_info && (parts & PARC_STEP_REF)) PARC_REQUIRE(env->ref_contacts.ptr, "env_step: ref_contacts not set");

    const ObsLayout L = make_layout(m->num_bodies, m->dof_size, lib->num_contacts, *cfg);
    PARC_REQUIRE(!(parts & PARC_STEP_OBS) || env->obs_row_stride >= L.obs_dim, "env_step: obs_row_stride %lld < obs_dim %d",
                 (long long)env->obs_row_stride, L.obs_dim);
    const size_t smem = (parts & PARC_STEP_OBS) ? (size_t)TILE * L.obs_dim * sizeof(float) : 0;
    PARC_REQUIRE(smem <= 227 * 1024, "env_step: observation tile of %zu bytes exceeds shared memory", smem);
    return parc::launch_env_step(m, lib, terrain, cfg, env, parts, env_ids, n_rows, nullptr, nullptr, stream);
}

namespace parc {
// Shared launcher: rows come from env_ids (i64, host count), from env_ids32 + a device-side count
// (done_list/done_count of the previous kernels, grid sized for max_rows), or are simply 0..N-1.
int launch_env_step(const parc_char_model_t* m, const parc_motion_lib_t* lib, const parc_terrain_t* terrain,
                    const parc_env_config_t* cfg, const parc_env_buffers_t* env, uint32_t parts, const int64_t* env_ids,
                    long long max_rows, const int* env_ids32, const int* n_rows_dev, void* stream) {
    const ObsLayout L = make_layout(m->num_bodies, m->dof_size, lib->num_contacts, *cfg);
    const size_t smem = (parts & PARC_STEP_OBS) ? (size_t)TILE * L.obs_dim * sizeof(float) : 0;
    const int bulk_ok = (env_ids == nullptr) && (env_ids32 == nullptr) && (env->obs_row_stride == L.obs_dim) &&
                        ((reinterpret_cast<uintptr_t>(env->obs_buf) & 15) == 0);
    const long long blocks = (max_rows + TILE - 1) / TILE;
    cudaError_t ce;
    if (m->num_bodies <= 16) {
        ce = cudaFuncSetAttribute(env_step_kernel<16>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (ce == cudaSuccess)
            env_step_kernel<16><<<(unsigned)blocks, STEP_THREADS, smem, as_stream(stream)>>>(
                *m, *lib, *terrain, *cfg, *env, parts, (const long long*)env_ids, max_rows, bulk_ok, env_ids32, n_rows_dev);
    } else {
        ce = cudaFuncSetAttribute(env_step_kernel<32>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (ce == cudaSuccess)
            env_step_kernel<32><<<(unsigned)blocks, STEP_THREADS, smem, as_stream(stream)>>>(
                *m, *lib, *terrain, *cfg, *env, parts, (const long long*)env_ids, max_rows, bulk_ok, env_ids32, n_rows_dev);
    }
    if (ce != cudaSuccess) { set_error("env_step: %s", cudaGetErrorString(ce)); return PARC_ERR_CUDA; }
    return check_launch("env_step");
}
}  // namespace parc
