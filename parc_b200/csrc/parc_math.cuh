// Quaternion / exp-map / heading arithmetic for the PARC tracking hot path (sm_100a).
// Quaternions are (x, y, z, w) everywhere, as in the reference (PARC/util/torch_util.py:46-47).
//
// Every function states the reference lines whose *arithmetic and branch thresholds* it follows.
// Index- and flag-critical expressions use __fmul_rn/__fadd_rn/__fdiv_rn so nvcc cannot contract
// them into FMAs or approximate divisions; the remaining fp32 math is allowed to contract
// (tolerance 1e-5 relative, BASELINE.json north_star).
#pragma once
#include <cuda_runtime.h>
#include <math.h>

namespace parc {

struct Quat { float x, y, z, w; };
struct Vec3 { float x, y, z; };

__device__ __forceinline__ Quat make_quat(float x, float y, float z, float w) { Quat q; q.x = x; q.y = y; q.z = z; q.w = w; return q; }
__device__ __forceinline__ Vec3 make_vec3(float x, float y, float z) { Vec3 v; v.x = x; v.y = y; v.z = z; return v; }
__device__ __forceinline__ Quat quat_from_f4(float4 v) { return make_quat(v.x, v.y, v.z, v.w); }
__device__ __forceinline__ float4 quat_to_f4(Quat q) { return make_float4(q.x, q.y, q.z, q.w); }
__device__ __forceinline__ Quat quat_identity() { return make_quat(0.f, 0.f, 0.f, 1.f); }

__device__ __forceinline__ Vec3 operator+(Vec3 a, Vec3 b) { return make_vec3(__fadd_rn(a.x, b.x), __fadd_rn(a.y, b.y), __fadd_rn(a.z, b.z)); }
__device__ __forceinline__ Vec3 operator-(Vec3 a, Vec3 b) { return make_vec3(__fsub_rn(a.x, b.x), __fsub_rn(a.y, b.y), __fsub_rn(a.z, b.z)); }
__device__ __forceinline__ Vec3 operator*(float s, Vec3 a) { return make_vec3(s * a.x, s * a.y, s * a.z); }
// torch.cross on CPU (ATen, compiled with FMA contraction) evaluates each component as
// fma(a1, b2, -round(a2 * b1)); measured bit-identical on 1e5 random vectors, the separately
// rounded form only on 75 % of them. Reproducing it keeps quat_rotate -- and with it every FK
// position, which lives at world-coordinate magnitude -- bit-exact against the reference.
__device__ __forceinline__ Vec3 cross(Vec3 a, Vec3 b) {
    return make_vec3(__fmaf_rn(a.y, b.z, -__fmul_rn(a.z, b.y)), __fmaf_rn(a.z, b.x, -__fmul_rn(a.x, b.z)),
                     __fmaf_rn(a.x, b.y, -__fmul_rn(a.y, b.x)));
}

// torch.sum(a*b, dim=-1) over 3 / 4 trailing elements: separately rounded products added left to
// right (measured against torch 2.11 CPU: ((p0+p1)+p2)+p3 is bit-identical, pairwise is not).
__device__ __forceinline__ float dot3_seq(Vec3 a, Vec3 b) {
    return __fadd_rn(__fadd_rn(__fmul_rn(a.x, b.x), __fmul_rn(a.y, b.y)), __fmul_rn(a.z, b.z));
}
__device__ __forceinline__ float dot4_seq(Quat a, Quat b) {
    return __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(a.x, b.x), __fmul_rn(a.y, b.y)), __fmul_rn(a.z, b.z)),
                     __fmul_rn(a.w, b.w));
}
__device__ __forceinline__ float sqnorm3_seq(Vec3 a) { return dot3_seq(a, a); }

// Approximate reciprocal / square root (MUFU, ~1 ulp, no slow path) for values that are only held to
// the 1e-5 tolerance. Index- and flag-critical expressions keep the IEEE forms.
__device__ __forceinline__ float fast_rcp(float x) { float r; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
__device__ __forceinline__ float fast_sqrt(float x) { float r; asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }

// quat_conjugate, torch_util.py:30-32
__device__ __forceinline__ Quat quat_conj(Quat q) { return make_quat(-q.x, -q.y, -q.z, q.w); }

// quat_pos, torch_util.py:34-39: q * (1 - 2*[w < 0])
__device__ __forceinline__ Quat quat_pos(Quat q) {
    float s = (q.w < 0.f) ? -1.f : 1.f;
    return make_quat(s * q.x, s * q.y, s * q.z, s * q.w);
}

// quat_mul, torch_util.py:41-59: the reference's 8-multiplication factoring, every operation
// rounded on its own in the reference's order (elementwise torch ops), so products of
// near-identical rotations (quat_diff_angle of small pose errors, finite-difference velocities)
// cancel exactly as they do upstream.
__device__ __forceinline__ Quat quat_mul(Quat a, Quat b) {
    const float ww = __fmul_rn(__fadd_rn(a.z, a.x), __fadd_rn(b.x, b.y));
    const float yy = __fmul_rn(__fsub_rn(a.w, a.y), __fadd_rn(b.w, b.z));
    const float zz = __fmul_rn(__fadd_rn(a.w, a.y), __fsub_rn(b.w, b.z));
    const float xx = __fadd_rn(__fadd_rn(ww, yy), zz);
    const float qq = __fmul_rn(0.5f, __fadd_rn(xx, __fmul_rn(__fsub_rn(a.z, a.x), __fsub_rn(b.x, b.y))));
    Quat r;
    r.w = __fadd_rn(__fsub_rn(qq, ww), __fmul_rn(__fsub_rn(a.z, a.y), __fsub_rn(b.y, b.z)));
    r.x = __fadd_rn(__fsub_rn(qq, xx), __fmul_rn(__fadd_rn(a.x, a.w), __fadd_rn(b.x, b.w)));
    r.y = __fadd_rn(__fsub_rn(qq, yy), __fmul_rn(__fsub_rn(a.w, a.x), __fadd_rn(b.y, b.z)));
    r.z = __fadd_rn(__fsub_rn(qq, zz), __fmul_rn(__fadd_rn(a.z, a.y), __fsub_rn(b.w, b.x)));
    return r;
}

// quat_rotate, torch_util.py:61-67: t = 2 q_v x v; (v + w t) + q_v x t, each op rounded separately
__device__ __forceinline__ Vec3 quat_rotate(Quat q, Vec3 v) {
    const Vec3 qv = make_vec3(q.x, q.y, q.z);
    const Vec3 c0 = cross(qv, v);
    const Vec3 t = make_vec3(__fmul_rn(2.f, c0.x), __fmul_rn(2.f, c0.y), __fmul_rn(2.f, c0.z));
    const Vec3 c1 = cross(qv, t);
    return make_vec3(__fadd_rn(__fadd_rn(v.x, __fmul_rn(q.w, t.x)), c1.x), __fadd_rn(__fadd_rn(v.y, __fmul_rn(q.w, t.y)), c1.y),
                     __fadd_rn(__fadd_rn(v.z, __fmul_rn(q.w, t.z)), c1.z));
}

// normalize, torch_util.py:10-13: x / max(||x||, 1e-9)
__device__ __forceinline__ Vec3 normalize3(Vec3 v) {
    const float r = fast_rcp(fmaxf(fast_sqrt(v.x * v.x + v.y * v.y + v.z * v.z), 1e-9f));
    return make_vec3(v.x * r, v.y * r, v.z * r);
}
__device__ __forceinline__ Quat quat_unit(Quat q) {
    const float r = fast_rcp(fmaxf(fast_sqrt(q.x * q.x + q.y * q.y + q.z * q.z + q.w * q.w), 1e-9f));
    return make_quat(q.x * r, q.y * r, q.z * r, q.w * r);
}
// quat_normalize, torch_util.py:469-472: unit(pos(q))
__device__ __forceinline__ Quat quat_normalize(Quat q) { return quat_unit(quat_pos(q)); }

// quat_to_axis_angle, torch_util.py:69-91: eps 1e-5 on ||xyz||, safe length 1e-6, default axis z.
__device__ __forceinline__ void quat_to_axis_angle(Quat q, Vec3& axis, float& angle) {
    q = quat_pos(q);
    const float len = fast_sqrt(q.x * q.x + q.y * q.y + q.z * q.z);
    const float a = 2.0f * atan2f(len, q.w);
    const float rs = fast_rcp(fmaxf(len, 1e-6f));
    const bool mask = len > 1e-5f;
    angle = mask ? a : 0.f;
    axis = mask ? make_vec3(q.x * rs, q.y * rs, q.z * rs) : make_vec3(0.f, 0.f, 1.f);
}

// quat_to_exp_map, torch_util.py:371-376
__device__ __forceinline__ Vec3 quat_to_exp_map(Quat q) {
    Vec3 axis; float angle;
    quat_to_axis_angle(q, axis, angle);
    return angle * axis;
}

// normalize_angle, torch_util.py:5-8
__device__ __forceinline__ float normalize_angle(float x) { return atan2f(sinf(x), cosf(x)); }
// the same for x >= 0 (a vector norm): atan2(sin x, cos x) is x itself below pi, up to rounding
__device__ __forceinline__ float normalize_angle_nonneg(float x) { return x < 3.1415925f ? x : normalize_angle(x); }

// sin / cos of a half angle |x| <= pi/2 + eps (the argument of axis_angle_to_quat after
// normalize_angle): Taylor polynomials, truncation error < 1e-7
__device__ __forceinline__ float cos_first_quadrant(float x) {
    const float x2 = x * x;
    float p = fmaf(x2, 2.0876757e-9f, -2.7557319e-7f);
    p = fmaf(p, x2, 2.4801587e-5f);
    p = fmaf(p, x2, -1.3888889e-3f);
    p = fmaf(p, x2, 4.1666668e-2f);
    p = fmaf(p, x2, -0.5f);
    return fmaf(p, x2, 1.0f);
}

// axis_angle_to_quat, torch_util.py:336-342: normalises the axis AND the resulting quaternion.
__device__ __forceinline__ Quat axis_angle_to_quat(Vec3 axis, float angle) {
    float s, c;
    const float h = angle * 0.5f;
    if (fabsf(h) <= 1.5708f) {
        const float ah = fabsf(h);
        const float x2 = ah * ah;
        float p = fmaf(x2, -2.5052108e-8f, 2.7557319e-6f);
        p = fmaf(p, x2, -1.9841270e-4f);
        p = fmaf(p, x2, 8.3333338e-3f);
        p = fmaf(p, x2, -1.6666667e-1f);
        s = copysignf(fmaf(ah * x2, p, ah), h);
        c = cos_first_quadrant(ah);
    } else {
        sincosf(h, &s, &c);
    }
    Vec3 n = normalize3(axis);
    return quat_unit(make_quat(n.x * s, n.y * s, n.z * s, c));
}

// exp_map_to_axis_angle + exp_map_to_quat, torch_util.py:425-450: min_theta 1e-5, default axis z.
__device__ __forceinline__ Quat exp_map_to_quat(Vec3 e) {
    const float ang = fast_sqrt(e.x * e.x + e.y * e.y + e.z * e.z);
    const float nang = normalize_angle_nonneg(ang);
    const bool mask = fabsf(nang) > 1e-5f;
    const float ra = fast_rcp(fmaxf(ang, 1e-30f));
    Vec3 axis = mask ? make_vec3(e.x * ra, e.y * ra, e.z * ra) : make_vec3(0.f, 0.f, 1.f);
    float angle = mask ? nang : 0.f;
    return axis_angle_to_quat(axis, angle);
}

// quat_diff_angle, torch_util.py:453-462: angle of q1 (x) conj(q0)
__device__ __forceinline__ float quat_diff_angle(Quat q0, Quat q1) {
    Quat dq = quat_mul(q1, quat_conj(q0));
    Vec3 axis; float angle;
    quat_to_axis_angle(dq, axis, angle);
    return angle;
}

// sin(x) for 0 <= x <= pi/2 (slerp arguments never leave that range): odd Taylor polynomial to
// x^11, truncation error < 6e-8 absolute, no range reduction and no slow path.
__device__ __forceinline__ float sin_first_quadrant(float x) {
    const float x2 = x * x;
    float p = fmaf(x2, -2.5052108e-8f, 2.7557319e-6f);
    p = fmaf(p, x2, -1.9841270e-4f);
    p = fmaf(p, x2, 8.3333338e-3f);
    p = fmaf(p, x2, -1.6666667e-1f);
    return fmaf(x * x2, p, x);
}

// slerp, torch_util.py:474-499. Branches are taken on the SAME fp32 values as the reference
// (dot product summed left to right, 1 - c*c unfused, correctly rounded sqrt); the result is not
// re-normalised and the |sin| < 0.001 fallback is the un-normalised midpoint, exactly as upstream.
// The two special cases are tested first (the reference evaluates everything and selects): static
// joints have c == 1 and would otherwise send sqrt / divide down their zero-operand slow paths.
__device__ __forceinline__ Quat slerp(Quat q0, Quat q1, float t) {
    float c = dot4_seq(q0, q1);
    if (c < 0.f) q1 = make_quat(-q1.x, -q1.y, -q1.z, -q1.w);
    c = fabsf(c);
    if (c >= 1.f) return q0;
    const float s = __fsqrt_rn(__fsub_rn(1.0f, __fmul_rn(c, c)));
    if (s < 0.001f)
        return make_quat(0.5f * q0.x + 0.5f * q1.x, 0.5f * q0.y + 0.5f * q1.y, 0.5f * q0.z + 0.5f * q1.z, 0.5f * q0.w + 0.5f * q1.w);
    const float half_theta = acosf(c);
    const float inv_s = __fdividef(1.0f, s);
    const float ra = sin_first_quadrant((1.f - t) * half_theta) * inv_s;
    const float rb = sin_first_quadrant(t * half_theta) * inv_s;
    return make_quat(ra * q0.x + rb * q1.x, ra * q0.y + rb * q1.y, ra * q0.z + rb * q1.z, ra * q0.w + rb * q1.w);
}

// The blend-independent half of slerp (torch_util.py:474-499) for one quaternion pair, computed once when the library is
// packed: (h, w) such that slerp(q0, q1, t) = ra q0 + rb q1 with
//   h >= 0 (general case):  ra = sin((1 - t) h) |w|,  rb = sin(t h) w      h = acos|c|, w = sign(c) / sqrt(1 - c c)
//   h <  0 (special cases): ra = 1 - |w|,             rb = w               |c| >= 1 -> w = 0 (q0);  sin < 0.001 -> w = +-0.5
// c = sum(q0 q1) left to right, 1 - c c unfused, correctly rounded sqrt: the branch decisions are the reference's.
__device__ __forceinline__ float2 slerp_pair(Quat q0, Quat q1) {
    const float c0 = dot4_seq(q0, q1);
    const float sgn = (c0 < 0.f) ? -1.f : 1.f;
    const float c = fabsf(c0);
    if (c >= 1.f) return make_float2(-1.f, 0.f);
    const float s = __fsqrt_rn(__fsub_rn(1.0f, __fmul_rn(c, c)));
    if (s < 0.001f) return make_float2(-1.f, 0.5f * sgn);
    return make_float2(acosf(c), sgn * __fdiv_rn(1.0f, s));
}
// slerp from the packed pair constants; omt = fp32(1 - t) as the reference forms it
__device__ __forceinline__ Quat slerp_pre(Quat q0, Quat q1, float2 pr, float t, float omt) {
    const float aw = fabsf(pr.y);
    float ra = sin_first_quadrant(omt * pr.x) * aw;
    float rb = sin_first_quadrant(t * pr.x) * pr.y;
    if (pr.x < 0.f) { ra = 1.f - aw; rb = pr.y; }
    return make_quat(ra * q0.x + rb * q1.x, ra * q0.y + rb * q1.y, ra * q0.z + rb * q1.z, ra * q0.w + rb * q1.w);
}

// calc_heading, torch_util.py:501-510: atan2 of the rotated x axis.
__device__ __forceinline__ float calc_heading(Quat q) {
    Vec3 d = quat_rotate(q, make_vec3(1.f, 0.f, 0.f));
    return atan2f(d.y, d.x);
}
// calc_heading_quat_inv, torch_util.py:522-530: axis_angle_to_quat(z, -heading)
__device__ __forceinline__ Quat calc_heading_quat_inv(Quat q) {
    return axis_angle_to_quat(make_vec3(0.f, 0.f, 1.f), -calc_heading(q));
}

// quat_to_tan_norm, torch_util.py:392-404: [R x_hat, R z_hat]
__device__ __forceinline__ void quat_to_tan_norm(Quat q, float out[6]) {
    Vec3 t = quat_rotate(q, make_vec3(1.f, 0.f, 0.f));
    Vec3 n = quat_rotate(q, make_vec3(0.f, 0.f, 1.f));
    out[0] = t.x; out[1] = t.y; out[2] = t.z; out[3] = n.x; out[4] = n.y; out[5] = n.z;
}

// quat_multiply, torch_util.py:607-625 (Hamilton form used only for the reset rotation offset)
__device__ __forceinline__ Quat quat_multiply(Quat a, Quat b) {
    Quat r;
    r.w = a.w * b.w - a.x * b.x - a.y * b.y - a.z * b.z;
    r.x = a.w * b.x + a.x * b.w + a.y * b.z - a.z * b.y;
    r.y = a.w * b.y - a.x * b.z + a.y * b.w + a.z * b.x;
    r.z = a.w * b.z + a.x * b.y - a.y * b.x + a.z * b.w;
    return r;
}

// Nearest-cell index of SubTerrain.get_grid_index (PARC/util/terrain_util.py:143-152):
// clamp(round_half_even((p - min) / d), 0, dim - 1); rintf is half-to-even like torch.round.
__device__ __forceinline__ int grid_index(float p, float mn, float d, int dim) {
    float u = rintf(__fdiv_rn(__fsub_rn(p, mn), d));
    // (long) conversion of values beyond int range: clamp in float first (dims << 2^24)
    u = fminf(fmaxf(u, 0.f), (float)(dim - 1));
    return (int)u;
}

}  // namespace parc
