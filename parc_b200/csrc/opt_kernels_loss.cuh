// Included by opt_kernels.cu: the motion-loss kernels and the C ABI of include/parc_b200_opt.h.
#pragma once

namespace parc {

constexpr int OPT_FT = 12;   // per-frame scratch terms: the 9 LossType slots (+3 spare)

// Scratch carved out of the caller's workspace; all [n = B*T, ...] row-major.
struct OptWs {
    float* tq;          // [n,4]      root quaternion = exp_map_to_quat(tgt_root_rot)
    float* jq;          // [n,J-1,4]  joint quaternions = dof_to_rot(tgt_joint_dof)
    float* body_pos;    // [n,J,3]
    float* body_rot;    // [n,J,4]
    float* g_body_pos;  // [n,J,3]    d loss / d body_pos
    float* g_body_rot;  // [n,J,4]
    float* g_tq;        // [n,4]
    float* g_jq;        // [n,J-1,4]
    float* g_rp;        // [n,3]      d loss / d root_pos through FK
    float* frame_terms; // [n,OPT_FT]
    int* gmin;          // [n,2]      first cell of the frame's own window (before the common size is applied)
    int* patch;         // [B,2]      largest per-frame window size of the motion
};

inline size_t opt_ws_floats(long long n, long long B, int J) {
    return (size_t)n * (4 + 4 * (J - 1) + 3 * J + 4 * J + 3 * J + 4 * J + 4 + 4 * (J - 1) + 3 + OPT_FT + 2) + (size_t)B * 2 + 64;
}

inline OptWs carve_ws(void* base, long long n, long long B, int J) {
    float* p = reinterpret_cast<float*>(base);
    OptWs w;
    auto take = [&](size_t k) { float* r = p; p += (k + 3) & ~size_t(3); return r; };
    w.tq = take(n * 4); w.jq = take(n * 4 * (J - 1)); w.body_pos = take(n * 3 * J); w.body_rot = take(n * 4 * J);
    w.g_body_pos = take(n * 3 * J); w.g_body_rot = take(n * 4 * J); w.g_tq = take(n * 4); w.g_jq = take(n * 4 * (J - 1));
    w.g_rp = take(n * 3); w.frame_terms = take(n * OPT_FT);
    w.gmin = reinterpret_cast<int*>(take(n * 2)); w.patch = reinterpret_cast<int*>(take(B * 2));
    return w;
}

struct OptTerrain {
    const float* hf; long long hf_stride; int X, Y;
    const float* min_point; long long mp_stride; float dx, dy;
};

// ---- 1. poses, regulariser terms, FK, heightfield window ---------------------------------------------
template <int G>
__global__ void __launch_bounds__(256) opt_pose_kernel(const __grid_constant__ parc_char_model_t cm, const __grid_constant__ parc_opt_config_t cfg,
                                                       long long n, long long T, const float* __restrict__ tgt_root_pos,
                                                       const float* __restrict__ tgt_root_rot, const float* __restrict__ tgt_joint_dof,
                                                       const float* __restrict__ src_root_pos, const float* __restrict__ src_root_rot,
                                                       const float* __restrict__ src_joint_rot, const float* __restrict__ body_points,
                                                       OptTerrain ter, OptWs ws) {
    const LaneBody lb = load_lane_body<G>(cm);
    const int J = cm.num_bodies, D = cm.dof_size;
    const long long gpb = blockDim.x / G;
    const long long gid0 = (long long)blockIdx.x * gpb + threadIdx.x / G;
    const long long warp_first = gid0 - ((threadIdx.x & 31) / G);
    for (long long base = warp_first; base < n; base += (long long)gridDim.x * gpb) {
        const long long i0 = base + ((threadIdx.x & 31) / G);
        const bool live = i0 < n;
        const long long i = live ? i0 : 0;
        const long long b = i / T;
        const Vec3 rp = make_vec3(tgt_root_pos[i * 3], tgt_root_pos[i * 3 + 1], tgt_root_pos[i * 3 + 2]);
        const Vec3 e = make_vec3(tgt_root_rot[i * 3], tgt_root_rot[i * 3 + 1], tgt_root_rot[i * 3 + 2]);
        const Quat tq = exp_map_to_quat(e);
        Quat q = quat_identity();
        const float* dof = tgt_joint_dof + i * D;
        const bool joint = lb.active && lb.body > 0;
        if (joint) q = dof_to_rot_lane(lb, [&](int k) { return dof[k]; });
        Vec3 pos; Quat rot;
        fk_lanes<G>(lb, q, rp, tq, pos, rot);
        // regulariser terms (:448-458)
        float t_pos = 0.f, t_rot = 0.f, t_joint = 0.f;
        if (lb.body == 0) {
            const float ex = rp.x - src_root_pos[i * 3], ey = rp.y - src_root_pos[i * 3 + 1], ez = rp.z - src_root_pos[i * 3 + 2];
            t_pos = ex * ex + ey * ey + ez * ez;
            const float a = qda_value(tq, quat_from_f4(reinterpret_cast<const float4*>(src_root_rot)[i]));
            t_rot = a * a;
        } else if (joint) {
            const float a = qda_value(q, quat_from_f4(reinterpret_cast<const float4*>(src_joint_rot)[i * (J - 1) + lb.body - 1]));
            t_joint = a * a;
        }
        t_joint = group_sum<G>(t_joint);
        // xy bounding box of the frame's body sample points (:497-505)
        float lox = 3.0e38f, loy = 3.0e38f, hix = -3.0e38f, hiy = -3.0e38f;
        if (lb.active) {
            for (int k = cfg.point_start[lb.body]; k < cfg.point_start[lb.body + 1]; ++k) {
                const Vec3 w = quat_rotate(rot, make_vec3(body_points[k * 3], body_points[k * 3 + 1], body_points[k * 3 + 2])) + pos;
                lox = fminf(lox, w.x); loy = fminf(loy, w.y); hix = fmaxf(hix, w.x); hiy = fmaxf(hiy, w.y);
            }
        }
        lox = group_min<G>(lox); loy = group_min<G>(loy); hix = group_max<G>(hix); hiy = group_max<G>(hiy);
        if (!live) continue;
        if (lb.active) {
            float* bp = ws.body_pos + (i * J + lb.body) * 3;
            bp[0] = pos.x; bp[1] = pos.y; bp[2] = pos.z;
            reinterpret_cast<float4*>(ws.body_rot)[i * J + lb.body] = quat_to_f4(rot);
            if (joint) reinterpret_cast<float4*>(ws.jq)[i * (J - 1) + lb.body - 1] = quat_to_f4(q);
        }
        if (lb.body == 0) {
            reinterpret_cast<float4*>(ws.tq)[i] = quat_to_f4(tq);
            float* ft = ws.frame_terms + i * OPT_FT;
#pragma unroll
            for (int k = 0; k < OPT_FT; ++k) ft[k] = 0.f;
            ft[PARC_LOSS_ROOT_POS] = t_pos; ft[PARC_LOSS_ROOT_ROT] = t_rot; ft[PARC_LOSS_JOINT_ROT] = t_joint;
            // window: floor / ceil of ((bbox -+ 2 cells) - min_point) / dxdy, clamped to the grid (:507-524)
            const float mpx = ter.min_point[b * ter.mp_stride], mpy = ter.min_point[b * ter.mp_stride + 1];
            const float idx = 1.0f / ter.dx, idy = 1.0f / ter.dy;
            int g0x = (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(lox, ter.dx * 2.0f), mpx), idx));
            int g0y = (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(loy, ter.dy * 2.0f), mpy), idy));
            int g1x = (int)ceilf(__fmul_rn(__fsub_rn(__fadd_rn(hix, ter.dx * 2.0f), mpx), idx));
            int g1y = (int)ceilf(__fmul_rn(__fsub_rn(__fadd_rn(hiy, ter.dy * 2.0f), mpy), idy));
            g0x = max(g0x, 0); g0y = max(g0y, 0); g1x = min(g1x, ter.X - 1); g1y = min(g1y, ter.Y - 1);
            ws.gmin[i * 2] = g0x; ws.gmin[i * 2 + 1] = g0y;
            atomicMax(&ws.patch[b * 2], g1x - g0x + 1);
            atomicMax(&ws.patch[b * 2 + 1], g1y - g0y + 1);
        }
    }
}

// ---- 2. penetration + contact against the frame's heightfield window ---------------------------------
// One block per frame. Thread k < Np owns body sample point k. Shared memory: the window's heights, the
// per-body closest (distance, point) key, the per-body gradient accumulators.
constexpr int OPT_MAX_PATCH = 1024;
__global__ void __launch_bounds__(320) opt_sdf_kernel(const __grid_constant__ parc_char_model_t cm, const __grid_constant__ parc_opt_config_t cfg,
                                                      long long T, const float* __restrict__ contacts, const float* __restrict__ body_points,
                                                      OptTerrain ter, OptWs ws, int need_grad) {
    __shared__ float s_h[OPT_MAX_PATCH];
    __shared__ unsigned long long s_key[PARC_MAX_BODIES];
    __shared__ float s_g[PARC_MAX_BODIES][8];
    __shared__ float s_red[16];
    extern __shared__ float s_pg[];          // [num_points][7] per-point gradient w.r.t. the body transform (need_grad)
    const long long i = blockIdx.x, b = i / T;
    const int J = cm.num_bodies, tid = threadIdx.x;
    const int psx = min(max(ws.patch[b * 2], 1), ter.X), psy = min(max(ws.patch[b * 2 + 1], 1), ter.Y);
    const int sx = max(min(ws.gmin[i * 2], max(ter.X - psx, 0)), 0), sy = max(min(ws.gmin[i * 2 + 1], max(ter.Y - psy, 0)), 0);   // :526-528
    const int cells = psx * psy;
    const bool staged = cells <= OPT_MAX_PATCH;
    const float* hf = ter.hf + b * ter.hf_stride;
    if (staged)
        for (int k = tid; k < cells; k += blockDim.x) { const int cx = k / psy, cy = k - cx * psy; s_h[k] = hf[(sx + cx) * ter.Y + sy + cy]; }
    for (int k = tid; k < J; k += blockDim.x) {
        s_key[k] = ~0ull;
#pragma unroll
        for (int c = 0; c < 8; ++c) s_g[k][c] = 0.f;
    }
    __syncthreads();
    // window origin (:549): min_point + start * dxdy
    const float ox = __fadd_rn(ter.min_point[b * ter.mp_stride], __fmul_rn((float)sx, ter.dx));
    const float oy = __fadd_rn(ter.min_point[b * ter.mp_stride + 1], __fmul_rn((float)sy, ter.dy));
    const float hx = ter.dx * 0.5f, hy = ter.dy * 0.5f, top = -cfg.base_z;
    const bool want_contact = cfg.w[PARC_LOSS_CONTACT] != 0.f;
    float pen_sum = 0.f;
    for (int k = tid; k < cfg.num_points; k += blockDim.x) {
        int body = 0;
        while (k >= cfg.point_start[body + 1]) ++body;
        const float* bp = ws.body_pos + (i * J + body) * 3;
        const Quat rot = quat_from_f4(reinterpret_cast<const float4*>(ws.body_rot)[i * J + body]);
        const Vec3 lp = make_vec3(body_points[k * 3], body_points[k * 3 + 1], body_points[k * 3 + 2]);
        const Vec3 P = quat_rotate(rot, lp) + make_vec3(bp[0], bp[1], bp[2]);
        float best_a = 3.0e38f, best_s = 3.0e38f;
        int arg_a = 0;
        for (int cx = 0; cx < psx; ++cx) {
            const float rx = P.x - (cx * ter.dx + ox);
            for (int cy = 0; cy < psy; ++cy) {
                const int c = cx * psy + cy;
                const float h = staged ? s_h[c] : hf[(sx + cx) * ter.Y + sy + cy];
                const CellSd r = cell_sd(rx, P.y - (cy * ter.dy + oy), P.z, h, hx, hy, cfg.base_z);
                if (r.air < best_a) { best_a = r.air; arg_a = c; }
                best_s = fminf(best_s, r.solid);
            }
        }
        // penetration (:553-556): -min(-min_air, 0) = max(min_air, 0)
        const float pen = fmaxf(best_a, 0.f);
        pen_sum += pen;
        Vec3 gP = make_vec3(0.f, 0.f, 0.f);
        if (need_grad && best_a > 0.f) {
            const int cx = arg_a / psy, cy = arg_a - cx * psy;
            const float h = staged ? s_h[arg_a] : hf[(sx + cx) * ter.Y + sy + cy];
            gP = cfg.w[PARC_LOSS_PENETRATION] * cell_sd_grad(P.x - (cx * ter.dx + ox), P.y - (cy * ter.dy + oy), P.z - (h + top) * 0.5f, hx, hy, (top - h) * 0.5f);
        }
        const int col = cfg.contact_col[body];
        const float pos_sd = fmaxf(best_s, 0.f);
        if (want_contact && col >= 0) {
            // closest point of the body (:570-572): min over its points, first index on ties
            const unsigned long long key = ((unsigned long long)__float_as_uint(pos_sd) << 32) | (unsigned)(k - cfg.point_start[body]);
            atomicMin(&s_key[body], key);
        }
        if (need_grad) {
            const Quat gq = rotate_bwd_q(rot, lp, gP);
            float* o = s_pg + k * 7;
            o[0] = gP.x; o[1] = gP.y; o[2] = gP.z; o[3] = gq.x; o[4] = gq.y; o[5] = gq.z; o[6] = gq.w;
        }
    }
    __syncthreads();
    if (need_grad) {
        // per-body sums in point order (deterministic): one thread per (body, component)
        for (int k = tid; k < J * 7; k += blockDim.x) {
            const int body = k / 7, c = k - body * 7;
            float a = 0.f;
            for (int q = cfg.point_start[body]; q < cfg.point_start[body + 1]; ++q) a += s_pg[q * 7 + c];
            s_g[body][c] = a;
        }
        __syncthreads();
    }
    // contact term and its gradient: the closest point of each contact body
    float con_sum = 0.f;
    if (want_contact) {
        for (int body = tid; body < J; body += blockDim.x) {
            const int col = cfg.contact_col[body];
            if (col < 0 || s_key[body] == ~0ull) continue;
            const float d = __uint_as_float((unsigned)(s_key[body] >> 32));
            const float cw = contacts[i * cfg.num_contact_cols + col];
            con_sum += d * cw;
            if (need_grad && d > 0.f && cw != 0.f) {
                const int k = cfg.point_start[body] + (int)(s_key[body] & 0xffffffffu);
                const float* bp = ws.body_pos + (i * J + body) * 3;
                const Quat rot = quat_from_f4(reinterpret_cast<const float4*>(ws.body_rot)[i * J + body]);
                const Vec3 lp = make_vec3(body_points[k * 3], body_points[k * 3 + 1], body_points[k * 3 + 2]);
                const Vec3 P = quat_rotate(rot, lp) + make_vec3(bp[0], bp[1], bp[2]);
                float best_s = 3.0e38f; int arg_s = 0;
                for (int c = 0; c < cells; ++c) {
                    const int cx = c / psy, cy = c - cx * psy;
                    const float h = staged ? s_h[c] : hf[(sx + cx) * ter.Y + sy + cy];
                    const CellSd r = cell_sd(P.x - (cx * ter.dx + ox), P.y - (cy * ter.dy + oy), P.z, h, hx, hy, cfg.base_z);
                    if (r.solid < best_s) { best_s = r.solid; arg_s = c; }
                }
                const int cx = arg_s / psy, cy = arg_s - cx * psy;
                const float h = staged ? s_h[arg_s] : hf[(sx + cx) * ter.Y + sy + cy];
                const Vec3 g = (cfg.w[PARC_LOSS_CONTACT] * cw) *
                               cell_sd_grad(P.x - (cx * ter.dx + ox), P.y - (cy * ter.dy + oy), P.z - (h + cfg.base_z) * 0.5f, hx, hy, (h - cfg.base_z) * 0.5f);
                const Quat gq = rotate_bwd_q(rot, lp, g);
                s_g[body][0] += g.x; s_g[body][1] += g.y; s_g[body][2] += g.z;          // one thread per body here
                s_g[body][3] += gq.x; s_g[body][4] += gq.y; s_g[body][5] += gq.z; s_g[body][6] += gq.w;
            }
        }
    }
    // block sums of the two terms
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { pen_sum += __shfl_xor_sync(FULL_MASK, pen_sum, o); con_sum += __shfl_xor_sync(FULL_MASK, con_sum, o); }
    if ((tid & 31) == 0) s_red[tid >> 5] = pen_sum;
    __syncthreads();
    if (tid == 0) {
        float s = 0.f;
        for (int w = 0; w < (int)(blockDim.x >> 5); ++w) s += s_red[w];
        ws.frame_terms[i * OPT_FT + PARC_LOSS_PENETRATION] = s;
    }
    __syncthreads();
    if ((tid & 31) == 0) s_red[tid >> 5] = con_sum;
    __syncthreads();
    if (tid == 0) {
        float s = 0.f;
        for (int w = 0; w < (int)(blockDim.x >> 5); ++w) s += s_red[w];
        ws.frame_terms[i * OPT_FT + PARC_LOSS_CONTACT] = s;
    }
    if (need_grad) {
        for (int k = tid; k < J * 7; k += blockDim.x) {
            const int body = k / 7, c = k - body * 7;
            if (c < 3) ws.g_body_pos[(i * J + body) * 3 + c] = s_g[body][c];
            else ws.g_body_rot[(i * J + body) * 4 + (c - 3)] = s_g[body][c];
        }
    }
}

}  // namespace parc

#include "opt_kernels_time.cuh"
