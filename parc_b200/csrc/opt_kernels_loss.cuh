// Included by opt_kernels.cu: the motion-loss kernels and the C ABI of include/parc_b200_opt.h.
#pragma once

#ifndef PARC_OPT_LANES_MINB
#define PARC_OPT_LANES_MINB 4
#endif

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
__global__ void __launch_bounds__(256, PARC_OPT_LANES_MINB) opt_pose_kernel(const __grid_constant__ parc_char_model_t cm, const __grid_constant__ parc_opt_config_t cfg,
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
            if (cfg.whole_terrain) { g0x = 0; g0y = 0; g1x = ter.X - 1; g1y = ter.Y - 1; }       // motion_terrain_contact_loss (:273-276)
            ws.gmin[i * 2] = g0x; ws.gmin[i * 2 + 1] = g0y;
            atomicMax(&ws.patch[b * 2], g1x - g0x + 1);
            atomicMax(&ws.patch[b * 2 + 1], g1y - g0y + 1);
        }
    }
}

// Minimum of the column SDFs over the window's cells for one point. The cell under the point is evaluated first; a
// cell k columns away is at least (k - 1) * dx from the point, so only cells within floor(bound / min(dx, dy)) + 1
// of it can still be closer than `bound`. What still needs a scan:
//   air columns   only while the running minimum is > 0 (<= 0 means "not inside the ground": the penetration term
//                 is 0 whatever the other cells say);
//   solid columns only for bodies with a non-zero contact label, and only below `solid_cap` (the body's closest
//                 distance found so far: a point that cannot beat it does not matter); `solid_zlb` = how far the
//                 point is above the highest column top of the window (0 if it is not).
// Within those limits the result equals the reference's minimum over the whole window.
// Debug build only (-DPARC_OPT_COUNT_CELLS, profiles/opt_cells.py): cells evaluated and scan_cells calls, for the mean cells
// visited per sample point that the config-4 roofline is stated against.
#ifdef PARC_OPT_COUNT_CELLS
__device__ unsigned long long g_opt_cells[2];
#define PARC_COUNT_CELLS(n) do { atomicAdd(&g_opt_cells[0], (unsigned long long)(n)); atomicAdd(&g_opt_cells[1], 1ull); } while (0)
#else
#define PARC_COUNT_CELLS(n) do { } while (0)
#endif
struct CellScan { float air, solid; int arg_air, arg_solid; };
// AIR / SOLID: the column(s) whose minimum is wanted (compile time); OWN_BOTH: the own cell is evaluated for both (pass 1 keeps
// the own-cell solid distance as the body's first bound). A member that is not scanned keeps its own-cell value.
template <bool AIR, bool SOLID, bool OWN_BOTH, typename HF>
__device__ __forceinline__ CellScan scan_cells(Vec3 P, int psx, int psy, float ox, float oy, float dx, float dy, float base_z,
                                               float solid_cap, float solid_zlb, HF height) {
    constexpr bool need_air = AIR, need_solid = SOLID;
    const float hx = dx * 0.5f, hy = dy * 0.5f;
    // the cell to start from (any cell next to the point would do: the scan below is exact), hence the approximate division
    const int ix0 = max(0, min(__float2int_rn(__fdividef(P.x - ox, dx)), psx - 1)), iy0 = max(0, min(__float2int_rn(__fdividef(P.y - oy, dy)), psy - 1));
    const int c0 = ix0 * psy + iy0;
    const CellSd r0 = cell_sd<AIR || OWN_BOTH, SOLID || OWN_BOTH>(P.x - (ix0 * dx + ox), P.y - (iy0 * dy + oy), P.z, height(ix0, iy0), hx, hy, base_z);
    CellScan out;
    out.air = r0.air; out.solid = r0.solid; out.arg_air = c0; out.arg_solid = c0;
    const float b_air = (need_air && out.air > 0.f) ? out.air : 0.f;
    const float b_solid = (need_solid && out.solid > 0.f) ? fminf(out.solid, solid_cap) : 0.f;
    if (!(fmaxf(b_air, b_solid) > 0.f)) { PARC_COUNT_CELLS(1); return out; }
    // a point that is at least solid_zlb above every column top is, k columns away, at least
    // sqrt(((k - 1) dx)^2 + zlb^2) from a column: over flat ground only the neighbouring cells remain
    const float b_solid_xy = sqrtf(fmaxf(b_solid * b_solid - solid_zlb * solid_zlb, 0.f));
    const int r = (int)fminf(fmaxf(b_air, b_solid_xy) / fminf(dx, dy), 1.0e6f) + 1;
    const int x0 = max(ix0 - r, 0), x1 = min(ix0 + r, psx - 1), y0 = max(iy0 - r, 0), y1 = min(iy0 + r, psy - 1);
    PARC_COUNT_CELLS(1 + (x1 - x0 + 1) * (y1 - y0 + 1));
    const float fy0 = (float)y0;
    for (int cx = x0; cx <= x1; ++cx) {
        const float rx = P.x - (cx * dx + ox);
        float fy = fy0;                                 // cy as a float, stepped exactly (no int -> float conversion per cell)
        int c = cx * psy + y0;
        for (int cy = y0; cy <= y1; ++cy, ++c, fy += 1.0f) {
            const CellSd v = cell_sd<AIR, SOLID>(rx, P.y - (fy * dy + oy), P.z, height(cx, cy), hx, hy, base_z);
            // the lower cell index wins ties, as a scan in index order would
            if (AIR && (v.air < out.air || (v.air == out.air && c < out.arg_air))) { out.air = v.air; out.arg_air = c; }
            if (SOLID && (v.solid < out.solid || (v.solid == out.solid && c < out.arg_solid))) { out.solid = v.solid; out.arg_solid = c; }
        }
    }
    return out;
}

// ---- 2. penetration + contact against the frame's heightfield window ---------------------------------
// One block per frame. Thread k < Np owns body sample point k. Shared memory: the window's heights, the
// per-body closest (distance, point) key, the per-body gradient accumulators.
constexpr int OPT_MAX_PATCH = 1024;
#ifndef PARC_OPT_SDF_THREADS
#define PARC_OPT_SDF_THREADS 96
#endif
constexpr int OPT_SDF_THREADS = PARC_OPT_SDF_THREADS;   // threads of the block that owns a frame; each takes ceil(Np / threads) sample points
#ifndef PARC_OPT_SDF_MINB
#define PARC_OPT_SDF_MINB 12
#endif
__global__ void __launch_bounds__(OPT_SDF_THREADS, PARC_OPT_SDF_MINB) opt_sdf_kernel(const __grid_constant__ parc_char_model_t cm, const __grid_constant__ parc_opt_config_t cfg,
                                                      long long T, const float* __restrict__ contacts, const float* __restrict__ body_points,
                                                      OptTerrain ter, OptWs ws, int need_grad) {
    __shared__ float s_h[OPT_MAX_PATCH];
    __shared__ unsigned long long s_key[PARC_MAX_BODIES];
    __shared__ unsigned s_bub[PARC_MAX_BODIES];     // per body: smallest own-cell solid distance of its points (float bits, >= 0)
    __shared__ float s_hmax;
    __shared__ float s_bpos[PARC_MAX_BODIES][3];     // the frame's body transforms and contact labels, fetched once
    __shared__ float4 s_brot[PARC_MAX_BODIES];
    __shared__ float s_cw[PARC_MAX_BODIES];
    __shared__ unsigned char s_pbody[1184];         // body of each sample point (num_points <= 1170, checked by the caller)
    __shared__ float s_g[PARC_MAX_BODIES][8];
    __shared__ float s_red[16];
    __shared__ int s_any[PARC_MAX_BODIES];          // body has at least one penetrating point (its per-point gradients are not all zero)
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
    for (int k = tid; k < J; k += blockDim.x)
        for (int q = cfg.point_start[k]; q < cfg.point_start[k + 1]; ++q) s_pbody[q] = (unsigned char)k;
    for (int k = tid; k < J * 3; k += blockDim.x) s_bpos[k / 3][k % 3] = ws.body_pos[i * J * 3 + k];
    for (int k = tid; k < J; k += blockDim.x) {
        s_brot[k] = reinterpret_cast<const float4*>(ws.body_rot)[i * J + k];
        const int col = cfg.contact_col[k];
        s_cw[k] = (cfg.w[PARC_LOSS_CONTACT] != 0.f && col >= 0) ? contacts[i * cfg.num_contact_cols + col] : 0.f;
    }
    for (int k = tid; k < J; k += blockDim.x) {
        s_key[k] = ~0ull;
        s_bub[k] = 0x7f800000u;
        s_any[k] = 0;
#pragma unroll
        for (int c = 0; c < 8; ++c) s_g[k][c] = 0.f;
    }
    __syncthreads();
    if (tid < 32) {   // highest cell of the window: a point above it is at least (z - hmax) from every solid column
        float m = -3.0e38f;
        for (int k = tid; k < cells; k += 32) { const int cx = k / psy, cy = k - cx * psy; m = fmaxf(m, staged ? s_h[k] : hf[(sx + cx) * ter.Y + sy + cy]); }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(FULL_MASK, m, o));
        if (tid == 0) s_hmax = m;
    }
    // window origin (:549): min_point + start * dxdy
    const float ox = __fadd_rn(ter.min_point[b * ter.mp_stride], __fmul_rn((float)sx, ter.dx));
    const float oy = __fadd_rn(ter.min_point[b * ter.mp_stride + 1], __fmul_rn((float)sy, ter.dy));
    const float hx = ter.dx * 0.5f, hy = ter.dy * 0.5f, top = -cfg.base_z;
    const bool want_contact = cfg.w[PARC_LOSS_CONTACT] != 0.f;
    // one base pointer + pitch for both homes of the window's heights (generic load): the staged copy or the terrain itself
    const float* hbase = staged ? s_h : hf + (long long)sx * ter.Y + sy;
    const int hpitch = staged ? psy : ter.Y;
    auto height = [&](int cx, int cy) { return hbase[cx * hpitch + cy]; };
    float pen_sum = 0.f;
    // ---- pass 1: penetration (air columns) of every point; own-cell solid distance of the contact bodies' points
    for (int k = tid; k < cfg.num_points; k += blockDim.x) {
        const int body = s_pbody[k];
        const float* bp = s_bpos[body];
        const Quat rot = quat_from_f4(s_brot[body]);
        const Vec3 lp = make_vec3(__ldg(body_points + k * 3), __ldg(body_points + k * 3 + 1), __ldg(body_points + k * 3 + 2));
        const Vec3 P = quat_rotate(rot, lp) + make_vec3(bp[0], bp[1], bp[2]);
        const bool need_solid = s_cw[body] != 0.f;
        const CellScan sc = scan_cells<true, false, true>(P, psx, psy, ox, oy, ter.dx, ter.dy, cfg.base_z, 0.f, 0.f, height);
        // penetration (:553-556): -min(-min_air, 0) = max(min_air, 0)
        pen_sum += fmaxf(sc.air, 0.f);
        if (need_solid) atomicMin(&s_bub[body], __float_as_uint(fmaxf(sc.solid, 0.f)));   // sc.solid here = the cell under the point
        if (need_grad) {
            float* o = s_pg + k * 7;
            if (sc.air > 0.f) {      // only a penetrating point has a gradient (most do not: no chain through the rotation for them)
                const int cx = sc.arg_air / psy, cy = sc.arg_air - cx * psy;
                const float h = height(cx, cy);
                const Vec3 gP = cfg.w[PARC_LOSS_PENETRATION] * cell_sd_grad(P.x - (cx * ter.dx + ox), P.y - (cy * ter.dy + oy), P.z - (h + top) * 0.5f, hx, hy, (top - h) * 0.5f);
                const Quat gq = rotate_bwd_q(rot, lp, gP);
                o[0] = gP.x; o[1] = gP.y; o[2] = gP.z; o[3] = gq.x; o[4] = gq.y; o[5] = gq.z; o[6] = gq.w;
                s_any[body] = 1;
            } else {
                o[0] = 0.f; o[1] = 0.f; o[2] = 0.f; o[3] = 0.f; o[4] = 0.f; o[5] = 0.f; o[6] = 0.f;
            }
        }
    }
    __syncthreads();
    // ---- pass 2: closest point of each contact body (:570-572): min over its points of the solid-column distance,
    //      first index on ties. A point is scanned only if it can still beat the body's best own-cell distance.
    if (want_contact) {
        const float hmax = s_hmax;
        for (int k = tid; k < cfg.num_points; k += blockDim.x) {
            const int body = s_pbody[k];
            if (s_cw[body] == 0.f) continue;
            const float* bp = s_bpos[body];
            const Vec3 P = quat_rotate(quat_from_f4(s_brot[body]), make_vec3(__ldg(body_points + k * 3), __ldg(body_points + k * 3 + 1), __ldg(body_points + k * 3 + 2))) +
                           make_vec3(bp[0], bp[1], bp[2]);
            const float cap = __uint_as_float(s_bub[body]);
            // above every column by more than the body's best (1e-4 m of slack: the column distance is computed as
            // |z - centre| - half height, which rounds differently from z - top)
            if (P.z - hmax > cap + 1e-4f) continue;
            const CellScan sc = scan_cells<false, true, false>(P, psx, psy, ox, oy, ter.dx, ter.dy, cfg.base_z, cap, fmaxf(P.z - hmax - 1e-4f, 0.f), height);
            const unsigned long long key = ((unsigned long long)__float_as_uint(fmaxf(sc.solid, 0.f)) << 32) | (unsigned)(k - cfg.point_start[body]);
            atomicMin(&s_key[body], key);
        }
    }
    __syncthreads();
    if (need_grad) {
        // per-body sums in point order (deterministic): one thread per (body, component)
        for (int k = tid; k < J * 7; k += blockDim.x) {
            const int body = k / 7, c = k - body * 7;
            float a = 0.f;
            if (s_any[body])     // a body without a penetrating point sums zeros
                for (int q = cfg.point_start[body]; q < cfg.point_start[body + 1]; ++q) a += s_pg[q * 7 + c];
            s_g[body][c] = a;
        }
        __syncthreads();
    }
    // contact term and its gradient: the closest point of each contact body
    float con_sum = 0.f;
    if (want_contact) {
        // one thread per body: the bodies' gradient scans run side by side in the lanes of one warp
        for (int body = tid; body < J; body += blockDim.x) {
            if (s_cw[body] == 0.f || s_key[body] == ~0ull) continue;
            const float d = __uint_as_float((unsigned)(s_key[body] >> 32));
            const float cw = s_cw[body];
            con_sum += d * cw;
            if (need_grad && d > 0.f && cw != 0.f) {
                const int k = cfg.point_start[body] + (int)(s_key[body] & 0xffffffffu);
                const float* bp = s_bpos[body];
                const Quat rot = quat_from_f4(s_brot[body]);
                const Vec3 lp = make_vec3(__ldg(body_points + k * 3), __ldg(body_points + k * 3 + 1), __ldg(body_points + k * 3 + 2));
                const Vec3 P = quat_rotate(rot, lp) + make_vec3(bp[0], bp[1], bp[2]);
                const int arg_s = scan_cells<false, true, false>(P, psx, psy, ox, oy, ter.dx, ter.dy, cfg.base_z, d * 1.00001f + 1e-6f, fmaxf(P.z - s_hmax - 1e-4f, 0.f), height).arg_solid;
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
