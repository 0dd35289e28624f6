// Kinematic-optimizer loss: FK + terrain penetration / contact SDF + temporal regularisers, value and
// gradient (PARC/motion_synthesis/motion_opt/motion_optimization.py:426-656, terrain_util.py:1736-1794,
// geom_util.py:124-145). See include/parc_b200_opt.h.
//
// Pipeline per call (B motions x T frames, n = B*T frames):
//   opt_pose_kernel      G lanes per frame (lane = body): exp map / dof -> quaternions, regulariser terms, FK,
//                        xy bounding box of the frame's body points -> heightfield window, per-motion window size
//   opt_sdf_kernel       one block per frame, thread per body sample point: box SDF against the window's cells
//                        (air columns -> penetration, solid columns -> contact), per-body closest point,
//                        gradient w.r.t. the body transform accumulated in shared memory
//   opt_temporal_kernel  G lanes per frame: smoothness, sliding, jerk and their gradients (gather form: the
//                        owner of (frame, body) sums every contribution to it, no atomics)
//   fk_bwd_kernel        (kin_kernels.cu) gradient through forward kinematics
//   opt_param_kernel     G lanes per frame: regulariser gradients + chain through dof_to_rot / exp_map_to_quat
//   opt_reduce_kernel    one block per motion: per-frame terms -> the 9 loss terms and the weighted total (fp64)
#include <algorithm>

#include "parc_b200_opt.h"
#include "parc_common.cuh"
#include "parc_lanes.cuh"

namespace parc {

// ---- derivative helpers ------------------------------------------------------------------------
// r = a (x) b (Hamilton product): ga = dL/da, gb = dL/db given g = dL/dr.
__device__ __forceinline__ void qmul_bwd(Quat a, Quat b, Quat g, Quat& ga, Quat& gb) {
    ga.x = g.x * b.w - g.y * b.z + g.z * b.y - g.w * b.x;
    ga.y = g.x * b.z + g.y * b.w - g.z * b.x - g.w * b.y;
    ga.z = -g.x * b.y + g.y * b.x + g.z * b.w - g.w * b.z;
    ga.w = g.x * b.x + g.y * b.y + g.z * b.z + g.w * b.w;
    gb.x = a.w * g.x + a.z * g.y - a.y * g.z - a.x * g.w;
    gb.y = -a.z * g.x + a.w * g.y + a.x * g.z - a.y * g.w;
    gb.z = a.y * g.x - a.x * g.y + a.w * g.z - a.z * g.w;
    gb.w = a.x * g.x + a.y * g.y + a.z * g.z + a.w * g.w;
}

// quat_diff_angle(q0, q1) = angle of quat_pos(q1 (x) conj(q0)) (torch_util.py:447-462, 69-91), value and
// gradient: g0 += gA * d angle / d q0, g1 += gA * d angle / d q1.
__device__ __forceinline__ float qda_value(Quat q0, Quat q1) {
    const Quat d = quat_pos(quat_mul(q1, quat_conj(q0)));
    const float len = sqrtf(d.x * d.x + d.y * d.y + d.z * d.z);
    return len > 1e-5f ? 2.0f * atan2f(len, d.w) : 0.f;
}
__device__ __forceinline__ void qda_bwd(Quat q0, Quat q1, float gA, Quat& g0, Quat& g1) {
    const Quat c0 = quat_conj(q0);
    Quat d = quat_mul(q1, c0);
    const float s = d.w < 0.f ? -1.f : 1.f;
    d = make_quat(s * d.x, s * d.y, s * d.z, s * d.w);
    const float len2 = d.x * d.x + d.y * d.y + d.z * d.z, len = sqrtf(len2);
    if (!(len > 1e-5f) || gA == 0.f) return;
    const float n2 = len2 + d.w * d.w;
    const float kx = gA * 2.0f * d.w / (n2 * len), kw = -gA * 2.0f * len / n2;
    const Quat gd = make_quat(s * kx * d.x, s * kx * d.y, s * kx * d.z, s * kw);
    Quat ga, gb;
    qmul_bwd(q1, c0, gd, ga, gb);
    g1.x += ga.x; g1.y += ga.y; g1.z += ga.z; g1.w += ga.w;
    g0.x -= gb.x; g0.y -= gb.y; g0.z -= gb.z; g0.w += gb.w;
}

// quat_unit backward at a unit quaternion: remove the radial component.
__device__ __forceinline__ Quat unit_bwd(Quat q, Quat g) {
    const float d = q.x * g.x + q.y * g.y + q.z * g.z + q.w * g.w;
    return make_quat(g.x - q.x * d, g.y - q.y * d, g.z - q.z * d, g.w - q.w * d);
}

// exp_map_to_quat (torch_util.py:425-451) backward: e -> dL/de given g = dL/dq. The masked branch
// (|normalize_angle(||e||)| <= 1e-5 -> identity) has zero gradient.
__device__ __forceinline__ Vec3 exp_map_bwd(Vec3 e, Quat g) {
    const float theta = sqrtf(e.x * e.x + e.y * e.y + e.z * e.z);
    if (!(theta > 0.f)) return make_vec3(0.f, 0.f, 0.f);
    const float phi = normalize_angle(theta);
    if (!(fabsf(phi) > 1e-5f)) return make_vec3(0.f, 0.f, 0.f);
    const float it = 1.0f / theta;
    const Vec3 a = make_vec3(e.x * it, e.y * it, e.z * it);
    float sh, ch;
    sincosf(0.5f * phi, &sh, &ch);
    const Quat q = make_quat(a.x * sh, a.y * sh, a.z * sh, ch);
    const Quat gp = unit_bwd(q, g);
    const float ga_dot = gp.x * a.x + gp.y * a.y + gp.z * a.z;
    const float gphi = 0.5f * (ga_dot * ch - gp.w * sh);
    // axis path: (I - a a^T) (g_xyz sin(phi/2)) / theta
    const float k = sh * it;
    return make_vec3(a.x * gphi + k * (gp.x - a.x * ga_dot), a.y * gphi + k * (gp.y - a.y * ga_dot), a.z * gphi + k * (gp.z - a.z * ga_dot));
}

// axis_angle_to_quat(axis const, angle) backward (hinge joints, kin_char_model.py:61-81).
__device__ __forceinline__ float hinge_bwd(Vec3 axis, float angle, Quat g) {
    const Vec3 n = normalize3(axis);
    float sh, ch;
    sincosf(0.5f * angle, &sh, &ch);
    const Quat q = make_quat(n.x * sh, n.y * sh, n.z * sh, ch);
    const Quat gp = unit_bwd(q, g);
    return 0.5f * ((gp.x * n.x + gp.y * n.y + gp.z * n.z) * ch - gp.w * sh);
}

// out = quat_rotate(q, v) (torch_util.py:61-67): dL/dq for constant v.
__device__ __forceinline__ Quat rotate_bwd_q(Quat q, Vec3 v, Vec3 g) {
    const Vec3 qv = make_vec3(q.x, q.y, q.z);
    const Vec3 t = 2.f * cross(qv, v);
    const float gw = g.x * t.x + g.y * t.y + g.z * t.z;
    const Vec3 gt = q.w * g + cross(g, qv);
    const Vec3 gq = cross(t, g) + 2.f * cross(v, gt);
    return make_quat(gq.x, gq.y, gq.z, gw);
}

template <int G>
__device__ __forceinline__ float group_min(float v) {
#pragma unroll
    for (int o = G / 2; o > 0; o >>= 1) v = fminf(v, __shfl_xor_sync(FULL_MASK, v, o, G));
    return v;
}
template <int G>
__device__ __forceinline__ float group_max(float v) {
#pragma unroll
    for (int o = G / 2; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(FULL_MASK, v, o, G));
    return v;
}

// ---- box SDF of one point against one heightfield cell --------------------------------------------
// sdBox (geom_util.py:124-145) for the air column [h, -base_z] and the solid column [base_z, h] of a cell
// whose xy half extents are (hx, hy); rx, ry = point - cell centre. The two share the xy part.
struct CellSd { float air, solid; };
// AIR / SOLID: which of the two columns the caller needs (the other member is left at 0). The distances only feed minima and
// clamped sums with a 2e-5 tolerance, so the square root is the approximate one (2 ulp).
template <bool AIR, bool SOLID>
__device__ __forceinline__ CellSd cell_sd(float rx, float ry, float pz, float h, float hx, float hy, float base_z) {
    const float qx = fabsf(rx) - hx, qy = fabsf(ry) - hy;
    const float ox = fmaxf(qx, 0.f), oy = fmaxf(qy, 0.f);
    const float oxy2 = ox * ox + oy * oy, mxy = fmaxf(qx, qy);
    CellSd r;
    r.air = 0.f; r.solid = 0.f;
    if (AIR) {
        const float top = -base_z;
        const float qa = fabsf(pz - (h + top) * 0.5f) - (top - h) * 0.5f;
        const float oa = fmaxf(qa, 0.f);
        r.air = fast_sqrt(oxy2 + oa * oa) + fminf(fmaxf(mxy, qa), 0.f);
    }
    if (SOLID) {
        const float qs = fabsf(pz - (h + base_z) * 0.5f) - (h - base_z) * 0.5f;
        const float os = fmaxf(qs, 0.f);
        r.solid = fast_sqrt(oxy2 + os * os) + fminf(fmaxf(mxy, qs), 0.f);
    }
    return r;
}
// d sdBox / d point for the column (centre z cz, half height hz)
__device__ __forceinline__ Vec3 cell_sd_grad(float rx, float ry, float rz, float hx, float hy, float hz) {
    const float qx = fabsf(rx) - hx, qy = fabsf(ry) - hy, qz = fabsf(rz) - hz;
    const float sx = rx > 0.f ? 1.f : (rx < 0.f ? -1.f : 0.f), sy = ry > 0.f ? 1.f : (ry < 0.f ? -1.f : 0.f),
                sz = rz > 0.f ? 1.f : (rz < 0.f ? -1.f : 0.f);
    const float ox = fmaxf(qx, 0.f), oy = fmaxf(qy, 0.f), oz = fmaxf(qz, 0.f);
    const float on = sqrtf(ox * ox + oy * oy + oz * oz);
    Vec3 g = make_vec3(0.f, 0.f, 0.f);
    if (on > 0.f) { const float r = 1.0f / on; g = make_vec3(ox * r * sx, oy * r * sy, oz * r * sz); }
    const float m = fmaxf(qx, fmaxf(qy, qz));
    if (m < 0.f) {   // inside: the closest face
        if (qx >= qy && qx >= qz) g.x += sx;
        else if (qy >= qz) g.y += sy;
        else g.z += sz;
    }
    return g;
}

// ---- points_hf_sdf (terrain_util.py:1736-1794) ----------------------------------------------------
// One thread per point; the batch's heightfield is staged in shared memory in tiles of cells.
__global__ void __launch_bounds__(256) points_hf_sdf_kernel(const float* __restrict__ points, long long N, const float* __restrict__ hf,
                                                            int X, int Y, const float* __restrict__ min_box_center, float dx, float dy,
                                                            float base_z, int inverted, float* __restrict__ sd, float* __restrict__ grad) {
    __shared__ float s_h[1024];
    const long long b = blockIdx.y;
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const bool live = i < N;
    const float* p = points + (b * N + (live ? i : 0)) * 3;
    const float px = p[0], py = p[1], pz = p[2];
    const float mx = min_box_center[b * 2], my = min_box_center[b * 2 + 1];
    const float hx = dx * 0.5f, hy = dy * 0.5f;
    const int cells = X * Y;
    float best = 3.0e38f;
    int best_c = 0;
    for (int c0 = 0; c0 < cells; c0 += 1024) {
        __syncthreads();
        for (int k = threadIdx.x; k < 1024 && c0 + k < cells; k += blockDim.x) s_h[k] = hf[b * cells + c0 + k];
        __syncthreads();
        const int lim = min(1024, cells - c0);
        for (int k = 0; k < lim; ++k) {
            const int c = c0 + k, cx = c / Y, cy = c - cx * Y;
            const CellSd r = cell_sd<true, true>(px - (cx * dx + mx), py - (cy * dy + my), pz, s_h[k], hx, hy, base_z);
            const float v = inverted ? r.air : r.solid;
            if (v < best) { best = v; best_c = c; }
        }
    }
    if (!live) return;
    sd[b * N + i] = inverted ? -best : best;
    if (grad) {
        const int cx = best_c / Y, cy = best_c - cx * Y;
        const float h = hf[b * cells + best_c];
        const float top = -base_z;
        const float cz = inverted ? (h + top) * 0.5f : (h + base_z) * 0.5f, hz = inverted ? (top - h) * 0.5f : (h - base_z) * 0.5f;
        Vec3 g = cell_sd_grad(px - (cx * dx + mx), py - (cy * dy + my), pz - cz, hx, hy, hz);
        const float s = inverted ? -1.f : 1.f;
        float* o = grad + (b * N + i) * 3;
        o[0] = s * g.x; o[1] = s * g.y; o[2] = s * g.z;
    }
}

}  // namespace parc

#include "opt_kernels_loss.cuh"
