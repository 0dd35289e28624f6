// Included by opt_kernels_loss.cuh: temporal terms, parameter backward, reduction and the C ABI.
#pragma once

namespace parc {

// ---- 3. smoothness, sliding, jerk (:460-471, 608-624) and their gradients ----------------------------
// Lane = body j of frame t. The lane owns velocity index t (t <= T-2) and jerk index t (t <= T-4) for the
// VALUES, and gathers every contribution to d loss / d body_pos[t,j], d loss / d body_rot[t,j].
// CONS: body-constraint records are present (:581-609). The lane scans its motion's records for its own body: frames inside a
// record add the sphere / sole-point hinge (value + gradient w.r.t. this frame's body transform), and velocity indices
// inside a record are masked out of the sliding term.
template <int G, bool CONS>
__global__ void __launch_bounds__(256, PARC_OPT_LANES_MINB) opt_temporal_kernel(const __grid_constant__ parc_char_model_t cm, const __grid_constant__ parc_opt_config_t cfg,
                                                           const __grid_constant__ parc_opt_constraints_t cons,
                                                           long long n, long long T, const float* __restrict__ src_body_vels,
                                                           const float* __restrict__ src_body_rot_vels, const float* __restrict__ contacts,
                                                           const float* __restrict__ body_points, OptWs ws, int need_grad) {
    const int J = cm.num_bodies;
    const int j = threadIdx.x & (G - 1);
    const bool active = j < J;
    const long long gpb = blockDim.x / G;
    const long long gid0 = (long long)blockIdx.x * gpb + threadIdx.x / G;
    const long long warp_first = gid0 - ((threadIdx.x & 31) / G);
    const float w_s = cfg.w[PARC_LOSS_SMOOTHNESS], w_j = cfg.w[PARC_LOSS_JERK];
    const bool sliding = cfg.w[PARC_LOSS_SLIDING] != 0.f && cfg.num_contact_cols == J;
    const float w_sl = sliding ? cfg.w[PARC_LOSS_SLIDING] : 0.f;
    const float c0 = 0.03f, c2 = 0.0009f;
    const float dt = 1.0f / 30.0f;
    const float mj = cfg.max_jerk * (dt * dt * dt);
    for (long long base = warp_first; base < n; base += (long long)gridDim.x * gpb) {
        const long long i0 = base + ((threadIdx.x & 31) / G);
        const bool live = i0 < n && active;
        const long long i = i0 < n ? i0 : 0;
        const long long b = i / T;
        const int t = (int)(i - b * T);
        const int jj = active ? j : 0;
        float v_smooth = 0.f, v_slide = 0.f, v_jerk = 0.f, v_cons = 0.f;
        Vec3 gP = make_vec3(0.f, 0.f, 0.f);
        Quat gR = make_quat(0.f, 0.f, 0.f, 0.f);
        bool masked_t = false, masked_tm1 = false;      // velocity indices t and t-1 of this body sit inside a constraint
        if (CONS && live) {
            const int kind = cons.geom_kind[jj];
            if (kind != PARC_GEOM_NONE) {
                const float w_c = cfg.w[PARC_LOSS_BODY_CONSTRAINT];
                const float* bp = ws.body_pos + (i * J + jj) * 3;
                const Vec3 pos = make_vec3(bp[0], bp[1], bp[2]);
                const Quat rot = quat_from_f4(reinterpret_cast<const float4*>(ws.body_rot)[i * J + jj]);
                const float radius = cons.geom_radius[jj];
                const int r1 = cons.motion_start[b + 1];
                for (int r = cons.motion_start[b]; r < r1; ++r) {
                    if (cons.body[r] != jj) continue;
                    const int s0 = cons.start_frame[r], e0 = cons.end_frame[r];
                    masked_t |= (t >= s0 && t <= e0);
                    masked_tm1 |= (t - 1 >= s0 && t - 1 <= e0);
                    if (t < s0 || t > e0) continue;
                    const Vec3 cp = make_vec3(cons.point[r * 3], cons.point[r * 3 + 1], cons.point[r * 3 + 2]);
                    if (kind == PARC_GEOM_SPHERE) {
                        const Vec3 off = make_vec3(cons.geom_offset[jj][0], cons.geom_offset[jj][1], cons.geom_offset[jj][2]);
                        const Vec3 ctr = quat_rotate(rot, off) + pos;
                        const Vec3 dv = make_vec3(ctr.x - cp.x, ctr.y - cp.y, ctr.z - cp.z);
                        const float m = sqrtf(dv.x * dv.x + dv.y * dv.y + dv.z * dv.z);
                        const float d = m - radius;                                          // geom_util.sdSphere (:169-173)
                        v_cons += fabsf(d);
                        if (need_grad && m > 0.f && d != 0.f) {
                            const float k = (d > 0.f ? w_c : -w_c) / m;
                            const Vec3 g = make_vec3(k * dv.x, k * dv.y, k * dv.z);
                            const Quat gq = rotate_bwd_q(rot, off, g);
                            gP = gP + g;
                            gR.x += gq.x; gR.y += gq.y; gR.z += gq.z; gR.w += gq.w;
                        }
                    } else {
                        const int p0 = cfg.point_start[jj];
                        const int p1 = min(cfg.point_start[jj + 1], p0 + cons.sole_points);
                        for (int q = p0; q < p1; ++q) {
                            const Vec3 lp = make_vec3(__ldg(body_points + q * 3), __ldg(body_points + q * 3 + 1), __ldg(body_points + q * 3 + 2));
                            const Vec3 wp = quat_rotate(rot, lp) + pos;
                            const Vec3 dv = make_vec3(wp.x - cp.x, wp.y - cp.y, wp.z - cp.z);
                            const float m = sqrtf(dv.x * dv.x + dv.y * dv.y + dv.z * dv.z);
                            const float d = m - radius;
                            if (!(d > 0.f)) continue;
                            v_cons += d;
                            if (need_grad) {
                                const float k = w_c / m;
                                const Vec3 g = make_vec3(k * dv.x, k * dv.y, k * dv.z);
                                const Quat gq = rotate_bwd_q(rot, lp, g);
                                gP = gP + g;
                                gR.x += gq.x; gR.y += gq.y; gR.z += gq.z; gR.w += gq.w;
                            }
                        }
                    }
                }
            }
        }
        if (live) {
            // positions of frames t-3 .. t+3 of this body
            Vec3 Pw[7];
#pragma unroll
            for (int k = 0; k < 7; ++k) {
                const int tt = t - 3 + k;
                const bool ok = tt >= 0 && tt < T;
                const float* p = ws.body_pos + ((b * T + (ok ? tt : t)) * J + jj) * 3;
                Pw[k] = make_vec3(p[0], p[1], p[2]);
            }
            auto both_at = [&](int tt) {   // clamp(min(contacts[tt+1], contacts[tt]), 0) (:468-469)
                if (!sliding) return 0.f;
                return fmaxf(fminf(contacts[(b * T + tt + 1) * J + jj], contacts[(b * T + tt) * J + jj]), 0.f);
            };
            auto vel_err = [&](int tt, int k0) {   // k0 = index of P(tt) in Pw
                const float* sv = src_body_vels + ((b * (T - 1) + tt) * J + jj) * 3;
                return make_vec3(Pw[k0 + 1].x - Pw[k0].x - sv[0], Pw[k0 + 1].y - Pw[k0].y - sv[1], Pw[k0 + 1].z - Pw[k0].z - sv[2]);
            };
            auto jerk_vec = [&](int k0) {          // P(s+3) - 3 P(s+2) + 3 P(s+1) - P(s), k0 = index of P(s)
                const Vec3 v0 = make_vec3(Pw[k0 + 1].x - Pw[k0].x, Pw[k0 + 1].y - Pw[k0].y, Pw[k0 + 1].z - Pw[k0].z);
                const Vec3 v1 = make_vec3(Pw[k0 + 2].x - Pw[k0 + 1].x, Pw[k0 + 2].y - Pw[k0 + 1].y, Pw[k0 + 2].z - Pw[k0 + 1].z);
                const Vec3 v2 = make_vec3(Pw[k0 + 3].x - Pw[k0 + 2].x, Pw[k0 + 3].y - Pw[k0 + 2].y, Pw[k0 + 3].z - Pw[k0 + 2].z);
                return make_vec3((v2.x - v1.x) - (v1.x - v0.x), (v2.y - v1.y) - (v1.y - v0.y), (v2.z - v1.z) - (v1.z - v0.z));
            };
            auto rot_at = [&](int tt) { return quat_from_f4(reinterpret_cast<const float4*>(ws.body_rot)[(b * T + tt) * J + jj]); };
            // ---- values owned by (t, j)
            if (t <= T - 2) {
                const Vec3 ve = vel_err(t, 3);
                const float S = ve.x * ve.x + ve.y * ve.y + ve.z * ve.z;
                const float re = qda_value(rot_at(t + 1), rot_at(t)) - src_body_rot_vels[(b * (T - 1) + t) * J + jj];
                v_smooth = S + re * re;
                if (sliding) {
                    const float bo = both_at(t);
                    // inside a constraint the errors were zeroed before the sliding sum (:606-609)
                    v_slide = masked_t ? (sqrtf(c2) - c0) * bo + (sqrtf(c2) - c0) * bo
                                       : (sqrtf(S + c2) - c0) * bo + (sqrtf(re * re + c2) - c0) * bo;
                }
            }
            if (t <= T - 4) {
                const Vec3 jv = jerk_vec(3);
                v_jerk = fmaxf(sqrtf(jv.x * jv.x + jv.y * jv.y + jv.z * jv.z) - mj, 0.f);
            }
            if (need_grad) {
                // d loss / d vel[tt]
                auto gvel = [&](int tt, int k0) {
                    if (tt < 0 || tt > T - 2) return make_vec3(0.f, 0.f, 0.f);
                    const Vec3 ve = vel_err(tt, k0);
                    float k = 2.0f * w_s;
                    if (sliding && !(tt == t ? masked_t : masked_tm1)) k += w_sl * both_at(tt) / sqrtf(ve.x * ve.x + ve.y * ve.y + ve.z * ve.z + c2);
                    return make_vec3(k * ve.x, k * ve.y, k * ve.z);
                };
                // d loss / d jerk vector[s]
                auto gjerk = [&](int s, int k0) {
                    if (s < 0 || s > T - 4) return make_vec3(0.f, 0.f, 0.f);
                    const Vec3 jv = jerk_vec(k0);
                    const float m = sqrtf(jv.x * jv.x + jv.y * jv.y + jv.z * jv.z);
                    if (!(m > mj) || m == 0.f) return make_vec3(0.f, 0.f, 0.f);
                    const float k = w_j / m;
                    return make_vec3(k * jv.x, k * jv.y, k * jv.z);
                };
                const Vec3 a = gvel(t - 1, 2), c = gvel(t, 3);
                const Vec3 j3 = gjerk(t - 3, 0), j2 = gjerk(t - 2, 1), j1 = gjerk(t - 1, 2), j0 = gjerk(t, 3);
                gP = gP + make_vec3(a.x - c.x + j3.x - 3.f * j2.x + 3.f * j1.x - j0.x, a.y - c.y + j3.y - 3.f * j2.y + 3.f * j1.y - j0.y,
                               a.z - c.z + j3.z - 3.f * j2.z + 3.f * j1.z - j0.z);
                // rotational velocity: rvel[tt] = quat_diff_angle(rot[tt+1], rot[tt]) (:464)
                auto g_angle = [&](int tt, Quat q0, Quat q1) {
                    const float re = qda_value(q0, q1) - src_body_rot_vels[(b * (T - 1) + tt) * J + jj];
                    float k = 2.0f * w_s * re;
                    if (sliding && !(tt == t ? masked_t : masked_tm1)) k += w_sl * both_at(tt) * re / sqrtf(re * re + c2);
                    return k;
                };
                const Quat Rt = rot_at(t);
                Quat dump = make_quat(0.f, 0.f, 0.f, 0.f);
                if (t >= 1) { const Quat Rm = rot_at(t - 1); qda_bwd(Rt, Rm, g_angle(t - 1, Rt, Rm), gR, dump); }       // q0 = rot[t]
                if (t <= T - 2) { const Quat Rp = rot_at(t + 1); qda_bwd(Rp, Rt, g_angle(t, Rp, Rt), dump, gR); }       // q1 = rot[t]
            }
        }
        v_smooth = group_sum<G>(v_smooth); v_slide = group_sum<G>(v_slide); v_jerk = group_sum<G>(v_jerk);
        if (CONS) v_cons = group_sum<G>(v_cons);
        if (i0 < n && j == 0) {
            float* ft = ws.frame_terms + i * OPT_FT;
            ft[PARC_LOSS_SMOOTHNESS] = v_smooth; ft[PARC_LOSS_SLIDING] = v_slide; ft[PARC_LOSS_JERK] = v_jerk;
            if (CONS) ft[PARC_LOSS_BODY_CONSTRAINT] = v_cons;
        }
        if (live && need_grad) {
            float* gp = ws.g_body_pos + (i * J + j) * 3;
            gp[0] += gP.x; gp[1] += gP.y; gp[2] += gP.z;
            float* gr = ws.g_body_rot + (i * J + j) * 4;
            gr[0] += gR.x; gr[1] += gR.y; gr[2] += gR.z; gr[3] += gR.w;
        }
    }
}

// ---- 5. regulariser gradients + chain through dof_to_rot / exp_map_to_quat ----------------------------
template <int G>
__global__ void __launch_bounds__(256) opt_param_kernel(const __grid_constant__ parc_char_model_t cm, const __grid_constant__ parc_opt_config_t cfg,
                                                        long long n, const float* __restrict__ tgt_root_pos, const float* __restrict__ tgt_root_rot,
                                                        const float* __restrict__ tgt_joint_dof, const float* __restrict__ src_root_pos,
                                                        const float* __restrict__ src_root_rot, const float* __restrict__ src_joint_rot, OptWs ws,
                                                        float* __restrict__ g_root_pos, float* __restrict__ g_root_rot, float* __restrict__ g_joint_dof) {
    const LaneBody lb = load_lane_body<G>(cm);
    const int J = cm.num_bodies, D = cm.dof_size;
    const long long gpb = blockDim.x / G;
    for (long long i = (long long)blockIdx.x * gpb + threadIdx.x / G; i < n; i += (long long)gridDim.x * gpb) {
        if (!lb.active) continue;
        if (lb.body == 0) {
            const float w = 2.0f * cfg.w[PARC_LOSS_ROOT_POS];
#pragma unroll
            for (int c = 0; c < 3; ++c) g_root_pos[i * 3 + c] = ws.g_rp[i * 3 + c] + w * (tgt_root_pos[i * 3 + c] - src_root_pos[i * 3 + c]);
            const Quat tq = quat_from_f4(reinterpret_cast<const float4*>(ws.tq)[i]);
            const Quat sq = quat_from_f4(reinterpret_cast<const float4*>(src_root_rot)[i]);
            Quat gq = quat_from_f4(reinterpret_cast<const float4*>(ws.g_tq)[i]), dump = make_quat(0.f, 0.f, 0.f, 0.f);
            qda_bwd(tq, sq, 2.0f * cfg.w[PARC_LOSS_ROOT_ROT] * qda_value(tq, sq), gq, dump);
            const Vec3 ge = exp_map_bwd(make_vec3(tgt_root_rot[i * 3], tgt_root_rot[i * 3 + 1], tgt_root_rot[i * 3 + 2]), gq);
            g_root_rot[i * 3] = ge.x; g_root_rot[i * 3 + 1] = ge.y; g_root_rot[i * 3 + 2] = ge.z;
        } else {
            const long long k = i * (J - 1) + lb.body - 1;
            const Quat q = quat_from_f4(reinterpret_cast<const float4*>(ws.jq)[k]);
            const Quat sq = quat_from_f4(reinterpret_cast<const float4*>(src_joint_rot)[k]);
            Quat gq = quat_from_f4(reinterpret_cast<const float4*>(ws.g_jq)[k]), dump = make_quat(0.f, 0.f, 0.f, 0.f);
            qda_bwd(q, sq, 2.0f * cfg.w[PARC_LOSS_JOINT_ROT] * qda_value(q, sq), gq, dump);
            const float* dof = tgt_joint_dof + i * D + lb.dof_idx;
            float* gd = g_joint_dof + i * D + lb.dof_idx;
            if (lb.jtype == PARC_JOINT_SPHERICAL) {
                const Vec3 ge = exp_map_bwd(make_vec3(dof[0], dof[1], dof[2]), gq);
                gd[0] = ge.x; gd[1] = ge.y; gd[2] = ge.z;
            } else if (lb.jtype == PARC_JOINT_HINGE) {
                gd[0] = hinge_bwd(lb.axis, dof[0], gq);
            }
        }
    }
}

// ---- 6. per-motion sums (fp64) ----------------------------------------------------------------------
__global__ void __launch_bounds__(256) opt_reduce_kernel(const __grid_constant__ parc_opt_config_t cfg, long long T, OptWs ws,
                                                         float* __restrict__ terms, float* __restrict__ loss) {
    __shared__ double s[PARC_NUM_LOSS_TERMS][8];
    const long long b = blockIdx.x;
    double acc[PARC_NUM_LOSS_TERMS];
#pragma unroll
    for (int k = 0; k < PARC_NUM_LOSS_TERMS; ++k) acc[k] = 0.0;
    for (long long t = threadIdx.x; t < T; t += blockDim.x) {
        const float* ft = ws.frame_terms + (b * T + t) * OPT_FT;
#pragma unroll
        for (int k = 0; k < PARC_NUM_LOSS_TERMS; ++k) acc[k] += (double)ft[k];
    }
#pragma unroll
    for (int k = 0; k < PARC_NUM_LOSS_TERMS; ++k) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) acc[k] += __shfl_xor_sync(FULL_MASK, acc[k], o);
        if ((threadIdx.x & 31) == 0) s[k][threadIdx.x >> 5] = acc[k];
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        double total = 0.0;
        for (int k = 0; k < PARC_NUM_LOSS_TERMS; ++k) {
            double v = 0.0;
            for (int w = 0; w < (int)(blockDim.x >> 5); ++w) v += s[k][w];
            terms[b * PARC_NUM_LOSS_TERMS + k] = (float)v;
            total += (double)cfg.w[k] * v;
        }
        loss[b] = (float)total;
    }
}

}  // namespace parc

using namespace parc;

extern "C" int parc_b200_points_hf_sdf(const float* points, int64_t B, int64_t N, const float* hf, int32_t X, int32_t Y,
                                       const float* min_box_center, float dx, float dy, float base_z, int32_t inverted, float* sd,
                                       float* grad, void* stream) {
    if (B == 0 || N == 0) return PARC_OK;
    PARC_REQUIRE(points && hf && min_box_center && sd, "points_hf_sdf: NULL pointer");
    PARC_REQUIRE(B > 0 && N > 0 && X > 0 && Y > 0 && B <= 65535, "points_hf_sdf: bad sizes");
    dim3 grid((unsigned)((N + 255) / 256), (unsigned)B);
    points_hf_sdf_kernel<<<grid, 256, 0, as_stream(stream)>>>(points, N, hf, X, Y, min_box_center, dx, dy, base_z, inverted, sd, grad);
    return check_launch("points_hf_sdf");
}

#ifdef PARC_OPT_COUNT_CELLS
extern "C" int parc_b200_debug_opt_cells(unsigned long long* h_out2, int reset) {
    cudaError_t ce = cudaMemcpyFromSymbol(h_out2, g_opt_cells, sizeof(unsigned long long) * 2);
    if (ce == cudaSuccess && reset) { const unsigned long long z[2] = {0ull, 0ull}; ce = cudaMemcpyToSymbol(g_opt_cells, z, sizeof(z)); }
    return ce == cudaSuccess ? 0 : -3;
}
#endif

extern "C" int64_t parc_b200_motion_loss_workspace_bytes(const parc_char_model_t* m, int64_t B, int64_t T) {
    if (!m || B <= 0 || T <= 0) return 0;
    return (int64_t)(opt_ws_floats(B * T, B, m->num_bodies) * sizeof(float));
}

extern "C" int parc_b200_motion_loss_constrained(const parc_char_model_t* m, const parc_opt_config_t* cfg, const parc_opt_constraints_t* cons,
                                     int64_t B, int64_t T, const float* tgt_root_pos, const float* tgt_root_rot, const float* tgt_joint_dof,
                                     const float* src_root_pos, const float* src_root_rot_quat, const float* src_joint_rot,
                                     const float* src_body_vels, const float* src_body_rot_vels, const float* contacts,
                                     const float* hf, int64_t hf_motion_stride, int32_t X, int32_t Y, const float* min_point,
                                     int64_t min_point_motion_stride, float dx, float dy, const float* body_points, void* workspace,
                                     int64_t workspace_bytes, float* terms, float* loss, float* g_root_pos, float* g_root_rot,
                                     float* g_joint_dof, void* stream) {
    PARC_REQUIRE(m && cfg, "motion_loss: NULL model / config");
    if (B == 0) return PARC_OK;
    PARC_REQUIRE(B > 0 && T >= 4, "motion_loss: needs at least 4 frames per motion (jerk term)");
    PARC_REQUIRE(tgt_root_pos && tgt_root_rot && tgt_joint_dof && src_root_pos && src_root_rot_quat && src_joint_rot && src_body_vels &&
                     src_body_rot_vels && contacts && hf && min_point && body_points && workspace && terms && loss,
                 "motion_loss: NULL pointer");
    const int J = m->num_bodies;
    PARC_REQUIRE(J >= 1 && J <= PARC_MAX_BODIES, "motion_loss: bad body count");
    PARC_REQUIRE(cfg->num_points > 0 && cfg->point_start[0] == 0 && cfg->point_start[J] == cfg->num_points, "motion_loss: bad point table");
    PARC_REQUIRE(cfg->num_points <= 1184, "motion_loss: %d body sample points (at most 1,184 supported)", cfg->num_points);
    PARC_REQUIRE(X > 0 && Y > 0 && dx > 0.f && dy > 0.f, "motion_loss: bad terrain");
    const bool need_grad = g_root_pos || g_root_rot || g_joint_dof;
    PARC_REQUIRE(!need_grad || (g_root_pos && g_root_rot && g_joint_dof), "motion_loss: pass all three gradient outputs or none");
    const bool constrained = cons != nullptr && cons->num > 0;
    if (constrained) {
        PARC_REQUIRE(cons->motion_start && cons->body && cons->start_frame && cons->end_frame && cons->point, "motion_loss: NULL constraint array");
        PARC_REQUIRE(cons->sole_points >= 0, "motion_loss: negative sole point count");
    }
    const long long n = B * T;
    PARC_REQUIRE(workspace_bytes >= (int64_t)(opt_ws_floats(n, B, J) * sizeof(float)), "motion_loss: workspace too small");
    PARC_REQUIRE(n * (long long)J * 4 < (1ll << 40), "motion_loss: batch too large");
    cudaStream_t st = as_stream(stream);
    OptWs ws = carve_ws(workspace, n, B, J);
    OptTerrain ter{hf, hf_motion_stride, X, Y, min_point, min_point_motion_stride, dx, dy};
    cudaError_t ce = cudaMemsetAsync(ws.patch, 0, sizeof(int) * 2 * B, st);
    if (ce != cudaSuccess) { set_error("motion_loss: %s", cudaGetErrorString(ce)); return PARC_ERR_CUDA; }
    const int G = J <= 16 ? 16 : 32;
    const long long gpb = 256 / G;
    const unsigned blocks = (unsigned)std::min<long long>((n + gpb - 1) / gpb, (long long)kNumSMs * 32);
#define PARC_OPT_G(KERNEL, ...)                                                         \
    do {                                                                                \
        if (G == 16) KERNEL<16><<<blocks, 256, 0, st>>>(__VA_ARGS__);                   \
        else KERNEL<32><<<blocks, 256, 0, st>>>(__VA_ARGS__);                           \
    } while (0)
    PARC_OPT_G(opt_pose_kernel, *m, *cfg, n, T, tgt_root_pos, tgt_root_rot, tgt_joint_dof, src_root_pos, src_root_rot_quat, src_joint_rot,
               body_points, ter, ws);
    int rc = check_launch("motion_loss (pose)");
    if (rc) return rc;
    const size_t pg_bytes = need_grad ? sizeof(float) * 7 * (size_t)cfg->num_points : 0;
    PARC_REQUIRE(pg_bytes <= 32 * 1024, "motion_loss: more than 1,170 body sample points");
    opt_sdf_kernel<<<(unsigned)n, OPT_SDF_THREADS, pg_bytes, st>>>(*m, *cfg, T, contacts, body_points, ter, ws, need_grad ? 1 : 0);
    if ((rc = check_launch("motion_loss (sdf)"))) return rc;
    if (constrained) {
        if (G == 16) opt_temporal_kernel<16, true><<<blocks, 256, 0, st>>>(*m, *cfg, *cons, n, T, src_body_vels, src_body_rot_vels, contacts, body_points, ws, need_grad ? 1 : 0);
        else opt_temporal_kernel<32, true><<<blocks, 256, 0, st>>>(*m, *cfg, *cons, n, T, src_body_vels, src_body_rot_vels, contacts, body_points, ws, need_grad ? 1 : 0);
    } else {
        const parc_opt_constraints_t none = {};
        if (G == 16) opt_temporal_kernel<16, false><<<blocks, 256, 0, st>>>(*m, *cfg, none, n, T, src_body_vels, src_body_rot_vels, contacts, body_points, ws, need_grad ? 1 : 0);
        else opt_temporal_kernel<32, false><<<blocks, 256, 0, st>>>(*m, *cfg, none, n, T, src_body_vels, src_body_rot_vels, contacts, body_points, ws, need_grad ? 1 : 0);
    }
    if ((rc = check_launch("motion_loss (temporal)"))) return rc;
    if (need_grad) {
        rc = parc_b200_forward_kinematics_bwd(m, tgt_root_pos, ws.tq, ws.jq, n, ws.g_body_pos, ws.g_body_rot, ws.g_rp, ws.g_tq, ws.g_jq, stream);
        if (rc) return rc;
        ce = cudaMemsetAsync(g_joint_dof, 0, sizeof(float) * n * m->dof_size, st);
        if (ce != cudaSuccess) { set_error("motion_loss: %s", cudaGetErrorString(ce)); return PARC_ERR_CUDA; }
        PARC_OPT_G(opt_param_kernel, *m, *cfg, n, tgt_root_pos, tgt_root_rot, tgt_joint_dof, src_root_pos, src_root_rot_quat, src_joint_rot, ws,
                   g_root_pos, g_root_rot, g_joint_dof);
        if ((rc = check_launch("motion_loss (param)"))) return rc;
    }
#undef PARC_OPT_G
    opt_reduce_kernel<<<(unsigned)B, 256, 0, st>>>(*cfg, T, ws, terms, loss);
    return check_launch("motion_loss (reduce)");
}

extern "C" int parc_b200_motion_loss(const parc_char_model_t* m, const parc_opt_config_t* cfg, int64_t B, int64_t T,
                                     const float* tgt_root_pos, const float* tgt_root_rot, const float* tgt_joint_dof,
                                     const float* src_root_pos, const float* src_root_rot_quat, const float* src_joint_rot,
                                     const float* src_body_vels, const float* src_body_rot_vels, const float* contacts,
                                     const float* hf, int64_t hf_motion_stride, int32_t X, int32_t Y, const float* min_point,
                                     int64_t min_point_motion_stride, float dx, float dy, const float* body_points, void* workspace,
                                     int64_t workspace_bytes, float* terms, float* loss, float* g_root_pos, float* g_root_rot,
                                     float* g_joint_dof, void* stream) {
    return parc_b200_motion_loss_constrained(m, cfg, nullptr, B, T, tgt_root_pos, tgt_root_rot, tgt_joint_dof, src_root_pos, src_root_rot_quat,
                                             src_joint_rot, src_body_vels, src_body_rot_vels, contacts, hf, hf_motion_stride, X, Y, min_point,
                                             min_point_motion_stride, dx, dy, body_points, workspace, workspace_bytes, terms, loss, g_root_pos,
                                             g_root_rot, g_joint_dof, stream);
}

// ---- mdm_path.compute_motion_loss: forward-only penetration / contact against a whole terrain ----
namespace parc {
// every frame's window is the whole grid: first cell (0,0), size (X,Y); the per-frame terms start at zero
__global__ void path_loss_window_kernel(long long n, long long B, int X, int Y, OptWs ws) {
    for (long long k = (long long)blockIdx.x * blockDim.x + threadIdx.x; k < n * OPT_FT; k += (long long)gridDim.x * blockDim.x) {
        ws.frame_terms[k] = 0.f;
        if (k < n * 2) ws.gmin[k] = 0;
        if (k < B * 2) ws.patch[k] = (k & 1) ? Y : X;
    }
}
// one block per motion: frames in order, fp64 accumulation (the reference sums per body over frames x points in fp32;
// the order does not carry below the 1e-5 tolerance)
__global__ void __launch_bounds__(128) path_loss_reduce_kernel(long long T, float w_pen, float w_contact, OptWs ws, float* __restrict__ pen,
                                                               float* __restrict__ con, float* __restrict__ total) {
    __shared__ double s_p[128], s_c[128];
    const long long b = blockIdx.x;
    double p = 0.0, c = 0.0;
    for (long long t = threadIdx.x; t < T; t += blockDim.x) {
        p += ws.frame_terms[(b * T + t) * OPT_FT + PARC_LOSS_PENETRATION];
        c += ws.frame_terms[(b * T + t) * OPT_FT + PARC_LOSS_CONTACT];
    }
    s_p[threadIdx.x] = p; s_c[threadIdx.x] = c;
    __syncthreads();
    for (int o = 64; o > 0; o >>= 1) {
        if ((int)threadIdx.x < o) { s_p[threadIdx.x] += s_p[threadIdx.x + o]; s_c[threadIdx.x] += s_c[threadIdx.x + o]; }
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        const float pl = (float)s_p[0] * w_pen, cl = (float)s_c[0] * w_contact;
        if (pen) pen[b] = pl;
        if (con) con[b] = cl;
        if (total) total[b] = cl + pl;                                                     // mdm_path.py:115
    }
}
}  // namespace parc

extern "C" int64_t parc_b200_path_motion_loss_workspace_bytes(int64_t B, int64_t T) {
    if (B < 0 || T < 0) return 0;
    return (int64_t)sizeof(float) * ((B * T) * (OPT_FT + 2) + B * 2 + 64);
}

extern "C" int parc_b200_path_motion_loss(const parc_char_model_t* m, const parc_opt_config_t* cfg, int64_t B, int64_t T,
                                          const float* body_pos, const float* body_rot, const float* contacts, const float* hf, int32_t X,
                                          int32_t Y, const float* min_point, float dx, float dy, const float* body_points,
                                          float w_pen, float w_contact, void* workspace, int64_t workspace_bytes, float* pen_loss,
                                          float* contact_loss, float* total_loss, void* stream) {
    if (!model_ok(m, "path_motion_loss")) return PARC_ERR_INVALID_ARG;
    PARC_REQUIRE(cfg != nullptr, "path_motion_loss: NULL config");
    PARC_REQUIRE(B >= 0 && T >= 0, "path_motion_loss: negative batch / frame count");
    if (B == 0) return PARC_OK;
    PARC_REQUIRE(T > 0, "path_motion_loss: motions without frames");
    PARC_REQUIRE(body_pos && body_rot && contacts && hf && min_point && body_points, "path_motion_loss: NULL input");
    PARC_REQUIRE(X > 0 && Y > 0 && dx > 0.f && dy > 0.f, "path_motion_loss: empty terrain");
    PARC_REQUIRE(cfg->num_points > 0 && cfg->num_points <= 1184, "path_motion_loss: %d body sample points (1..1184 supported)", cfg->num_points);
    PARC_REQUIRE(workspace && workspace_bytes >= parc_b200_path_motion_loss_workspace_bytes(B, T), "path_motion_loss: workspace too small");
    PARC_REQUIRE(pen_loss || contact_loss || total_loss, "path_motion_loss: no output requested");
    const long long n = B * T;
    OptWs ws = {};
    float* p = reinterpret_cast<float*>(workspace);
    ws.frame_terms = p; p += n * OPT_FT;
    ws.gmin = reinterpret_cast<int*>(p); p += n * 2;
    ws.patch = reinterpret_cast<int*>(p);
    ws.body_pos = const_cast<float*>(body_pos);       // read-only in the forward pass
    ws.body_rot = const_cast<float*>(body_rot);
    parc_opt_config_t c = *cfg;
    c.w[PARC_LOSS_PENETRATION] = w_pen;
    c.w[PARC_LOSS_CONTACT] = w_contact;
    OptTerrain ter = {hf, 0, X, Y, min_point, 0, dx, dy};    // one terrain shared by the batch (mdm_path.py:94-96)
    cudaStream_t st = as_stream(stream);
    path_loss_window_kernel<<<(unsigned)std::min<long long>((n * OPT_FT + 255) / 256, (long long)kNumSMs * 8), 256, 0, st>>>(n, B, X, Y, ws);
    int rc = check_launch("path_motion_loss (window)");
    if (rc) return rc;
    opt_sdf_kernel<<<(unsigned)n, OPT_SDF_THREADS, 0, st>>>(*m, c, T, contacts, body_points, ter, ws, 0);
    if ((rc = check_launch("path_motion_loss (sdf)"))) return rc;
    path_loss_reduce_kernel<<<(unsigned)B, 128, 0, st>>>(T, w_pen, w_contact, ws, pen_loss, contact_loss, total_loss);
    return check_launch("path_motion_loss (reduce)");
}
