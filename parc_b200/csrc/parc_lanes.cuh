// Lane-parallel building blocks shared by every kernel of the tracking hot path.
//
// Mapping: one GROUP of G lanes (G = 16 when the character has <= 16 bodies, else 32) evaluates one
// pose; lane `sub` of the group owns body `sub` (lane 0 = root, lanes 1..J-1 = joints). A 15-body
// humanoid therefore keeps 30 of 32 lanes busy with two poses per warp. Forward kinematics walks
// the tree level by level: each level is 7 width-G shuffles that fetch the parent's world
// transform from the parent's lane (KinCharModel.forward_kinematics,
// PARC/anim/kin_char_model.py:617-649).
#pragma once
#include "parc_math.cuh"
#include "../../include/parc_b200.h"

namespace parc {

constexpr unsigned FULL_MASK = 0xffffffffu;

// Per-lane copy of one body's row of the kinematic tree (registers, loaded once per warp).
struct LaneBody {
    int body;        // body index owned by this lane (== sub-lane); >= num_bodies -> inactive
    int parent;      // parent body index (sub-lane to shuffle from); 0 for root/inactive
    int jtype;
    int dof_idx;
    int depth;       // root 0
    int max_depth;   // uniform
    bool active;
    bool lrot_identity;   // uniform: every _local_rotation is (0,0,0,1)
    Vec3 trans;
    Quat lrot;
    Vec3 axis;
};

__device__ __forceinline__ int joint_dof_dim(int jtype) {
    return jtype == PARC_JOINT_SPHERICAL ? 3 : (jtype == PARC_JOINT_HINGE ? 1 : 0);
}

template <int G>
__device__ __forceinline__ LaneBody load_lane_body(const parc_char_model_t& cm) {
    LaneBody lb;
    const int sub = threadIdx.x & (G - 1);
    const int J = cm.num_bodies;
    lb.body = sub;
    lb.active = sub < J;
    const int b = lb.active ? sub : 0;
    lb.parent = cm.parent[b] < 0 ? 0 : cm.parent[b];
    lb.jtype = lb.active ? cm.joint_type[b] : PARC_JOINT_FIXED;
    lb.dof_idx = cm.dof_idx[b];
    lb.trans = make_vec3(cm.local_trans[b][0], cm.local_trans[b][1], cm.local_trans[b][2]);
    lb.lrot = make_quat(cm.local_rot[b][0], cm.local_rot[b][1], cm.local_rot[b][2], cm.local_rot[b][3]);
    lb.axis = make_vec3(cm.axis[b][0], cm.axis[b][1], cm.axis[b][2]);
    // depth by walking up (J <= 32, depth small)
    int d = 0;
    int p = lb.active ? cm.parent[b] : -1;
    while (p >= 0) { ++d; p = cm.parent[p]; }
    lb.depth = d;
    int md = d;
    bool ident = !lb.active || (lb.lrot.x == 0.f && lb.lrot.y == 0.f && lb.lrot.z == 0.f && lb.lrot.w == 1.f);
#pragma unroll
    for (int o = G / 2; o > 0; o >>= 1) md = max(md, __shfl_xor_sync(FULL_MASK, md, o, G));
    lb.max_depth = md;
    lb.lrot_identity = __all_sync(FULL_MASK, ident);
    return lb;
}

template <int G>
__device__ __forceinline__ float group_bcast(float v, int src) { return __shfl_sync(FULL_MASK, v, src, G); }
template <int G>
__device__ __forceinline__ Vec3 group_bcast(Vec3 v, int src) {
    return make_vec3(__shfl_sync(FULL_MASK, v.x, src, G), __shfl_sync(FULL_MASK, v.y, src, G),
                     __shfl_sync(FULL_MASK, v.z, src, G));
}
template <int G>
__device__ __forceinline__ Quat group_bcast(Quat q, int src) {
    return make_quat(__shfl_sync(FULL_MASK, q.x, src, G), __shfl_sync(FULL_MASK, q.y, src, G),
                     __shfl_sync(FULL_MASK, q.z, src, G), __shfl_sync(FULL_MASK, q.w, src, G));
}
template <int G>
__device__ __forceinline__ float group_sum(float v) {
#pragma unroll
    for (int o = G / 2; o > 0; o >>= 1) v += __shfl_xor_sync(FULL_MASK, v, o, G);
    return v;
}

// Level-by-level forward kinematics. `q` is this lane's LOCAL joint rotation (ignored on the root
// lane), root_pos/root_rot are group-uniform. On return pos/rot hold the WORLD transform of the
// lane's body. Must be called by all 32 lanes of the warp.
template <int G>
__device__ __forceinline__ void fk_lanes(const LaneBody& lb, Quat q, Vec3 root_pos, Quat root_rot, Vec3& pos, Quat& rot) {
    Quat cq = lb.lrot_identity ? q : quat_mul(lb.lrot, q);   // local_rot (x) j_rot, :640
    pos = root_pos;
    rot = root_rot;
    for (int d = 1; d <= lb.max_depth; ++d) {
        Vec3 ppos = group_bcast<G>(pos, lb.parent);
        Quat prot = group_bcast<G>(rot, lb.parent);
        if (lb.depth == d) {
            pos = ppos + quat_rotate(prot, lb.trans);   // :637-639
            rot = quat_mul(prot, cq);                   // :641
        }
    }
}

// Joint.dof_to_rot (kin_char_model.py:61-81) for the lane's joint; dof points at the pose's D dofs.
template <typename LoadF>
__device__ __forceinline__ Quat dof_to_rot_lane(const LaneBody& lb, LoadF load_dof) {
    if (lb.jtype == PARC_JOINT_SPHERICAL) {
        Vec3 e = make_vec3(load_dof(lb.dof_idx), load_dof(lb.dof_idx + 1), load_dof(lb.dof_idx + 2));
        return exp_map_to_quat(e);
    } else if (lb.jtype == PARC_JOINT_HINGE) {
        return axis_angle_to_quat(lb.axis, load_dof(lb.dof_idx));
    }
    return quat_identity();
}

// Joint.rot_to_dof (kin_char_model.py:83-104); returns the number of dofs written into out[0..2].
__device__ __forceinline__ int rot_to_dof_lane(const LaneBody& lb, Quat q, float out[3]) {
    if (lb.jtype == PARC_JOINT_SPHERICAL) {
        Vec3 e = quat_to_exp_map(q);
        out[0] = e.x; out[1] = e.y; out[2] = e.z;
        return 3;
    } else if (lb.jtype == PARC_JOINT_HINGE) {
        Vec3 axis; float angle;
        quat_to_axis_angle(q, axis, angle);
        float d = dot3_seq(lb.axis, axis);
        out[0] = (d < 0.f) ? -angle : angle;
        return 1;
    }
    return 0;
}

// ---- motion library -------------------------------------------------------------------------

struct MotionMeta {
    int start_idx;
    int num_frames;
    float length;
    int loop_mode;
    Vec3 delta;
};

__device__ __forceinline__ MotionMeta load_motion_meta(const int32_t* __restrict__ meta, long long motion_id) {
    const int4* m4 = reinterpret_cast<const int4*>(meta) + 2 * motion_id;
    int4 a = __ldg(m4);
    int4 b = __ldg(m4 + 1);
    MotionMeta mm;
    mm.start_idx = a.x;
    mm.num_frames = a.y;
    mm.length = __int_as_float(a.z);
    mm.loop_mode = a.w;
    mm.delta = make_vec3(__int_as_float(b.x), __int_as_float(b.y), __int_as_float(b.z));
    return mm;
}

struct FrameBlend {
    long long idx0, idx1;   // global frame rows (start_idx already added)
    float blend;
    Vec3 loop_offset;
};

// MotionLib._calc_frame_blend (PARC/anim/motion_lib.py:425-438) + calc_phase (:519-530) +
// _calc_loop_offset (:440-457). Every fp32 operation is rounded separately, in the reference's
// order, so idx0/idx1 are bit-exact against torch (IEEE divide, no FMA contraction).
__device__ __forceinline__ FrameBlend calc_frame_blend(const MotionMeta& mm, float t) {
    FrameBlend fb;
    float phase = __fdiv_rn(t, mm.length);                 // times / motion_len
    fb.loop_offset = make_vec3(0.f, 0.f, 0.f);
    if (mm.loop_mode == PARC_LOOP_WRAP) {
        float fl = floorf(phase);
        fb.loop_offset = make_vec3(__fmul_rn(fl, mm.delta.x), __fmul_rn(fl, mm.delta.y), __fmul_rn(fl, mm.delta.z));
        phase = __fsub_rn(phase, fl);                      // phase - floor(phase)
    }
    // torch.clip(phase, 0, 1): NaN (0/0 for a 1-frame clip) propagates in torch; (long)NaN is
    // implementation-defined there, here it maps to frame 0.
    phase = fminf(fmaxf(phase, 0.f), 1.f);
    const float nfm1 = (float)(mm.num_frames - 1);
    const float scaled = __fmul_rn(phase, nfm1);           // phase * (num_frames - 1)
    int i0 = (int)scaled;                                  // .long(): truncation
    int i1 = min(i0 + 1, mm.num_frames - 1);
    fb.blend = __fsub_rn(scaled, (float)i0);
    fb.idx0 = (long long)i0 + mm.start_idx;
    fb.idx1 = (long long)i1 + mm.start_idx;
    return fb;
}

// Record offsets inside the interp / vel records (see parc_motion_lib_t in include/parc_b200.h).
__host__ __device__ __forceinline__ int interp_pos_off(int J) { return 4 * J; }
__host__ __device__ __forceinline__ int interp_contact_off(int J) { return 4 * J + 4; }
// slerp constants of the pair (this frame, next frame of the clip): [J][2] floats, 8-byte aligned
__host__ __device__ __forceinline__ int interp_pair_off(int J, int C) { return (4 * J + 4 + C + 1) & ~1; }
__host__ __device__ __forceinline__ int interp_record_floats(int J, int C) { return (interp_pair_off(J, C) + 2 * J + 3) & ~3; }
__host__ __device__ __forceinline__ int vel_record_floats(int D) { return (6 + D + 3) & ~3; }

// One pose sample of MotionLib.calc_motion_frame (motion_lib.py:94-126), lane-parallel:
// lane j slerps quaternion j of the two adjacent records (j = 0 is the root rotation), every lane
// gets the lerped root position (+ loop offset), lane c the lerped contact label c.
struct PoseSample {
    Quat q;          // lane j: root rot (j == 0) or local joint rot j
    Vec3 root_pos;   // group-uniform
    float contact;   // lane c < C
};

// (1 - b) * x0 + b * x1 with each product and the sum rounded on its own (torch elementwise ops)
__device__ __forceinline__ float lerp_exact(float omb, float x0, float b, float x1) {
    return __fadd_rn(__fmul_rn(omb, x0), __fmul_rn(b, x1));
}

__device__ __forceinline__ PoseSample eval_pose_lane(const parc_motion_lib_t& lib, const LaneBody& lb,
                                                     const FrameBlend& fb, int J) {
    PoseSample ps;
    const float* r0 = lib.interp + fb.idx0 * lib.irec;
    const float* r1 = lib.interp + fb.idx1 * lib.irec;
    const float b = fb.blend;
    const float omb = __fsub_rn(1.0f, b);
    Quat q0 = quat_identity(), q1 = quat_identity();
    if (lb.active) {
        q0 = quat_from_f4(__ldg(reinterpret_cast<const float4*>(r0) + lb.body));
        q1 = quat_from_f4(__ldg(reinterpret_cast<const float4*>(r1) + lb.body));
    }
    float4 p0 = __ldg(reinterpret_cast<const float4*>(r0 + interp_pos_off(J)));
    float4 p1 = __ldg(reinterpret_cast<const float4*>(r1 + interp_pos_off(J)));
    ps.contact = 0.f;
    if (lb.body < lib.num_contacts) {
        float c0 = __ldg(r0 + interp_contact_off(J) + lb.body);
        float c1 = __ldg(r1 + interp_contact_off(J) + lb.body);
        ps.contact = lerp_exact(omb, c0, b, c1);           // :121
    }
    ps.q = slerp(q0, q1, b);                               // :113-115
    // :112, :117-118 -- root positions live at world-coordinate magnitude: same rounding steps as torch
    ps.root_pos = make_vec3(__fadd_rn(lerp_exact(omb, p0.x, b, p1.x), fb.loop_offset.x),
                            __fadd_rn(lerp_exact(omb, p0.y, b, p1.y), fb.loop_offset.y),
                            __fadd_rn(lerp_exact(omb, p0.z, b, p1.z), fb.loop_offset.z));
    return ps;
}

// ---- strided views --------------------------------------------------------------------------
__device__ __forceinline__ float view_ld(const parc_view3_t& v, long long row, long long e) {
    return v.ptr[row * v.row_stride + e * v.elem_stride];
}
__device__ __forceinline__ float view_ld3(const parc_view3_t& v, long long row, long long mid, long long e) {
    return v.ptr[row * v.row_stride + mid * v.mid_stride + e * v.elem_stride];
}
__device__ __forceinline__ void view_st(const parc_view3_t& v, long long row, long long e, float x) {
    v.ptr[row * v.row_stride + e * v.elem_stride] = x;
}
__device__ __forceinline__ void view_st3(const parc_view3_t& v, long long row, long long mid, long long e, float x) {
    v.ptr[row * v.row_stride + mid * v.mid_stride + e * v.elem_stride] = x;
}
__device__ __forceinline__ Vec3 view_ld_vec3(const parc_view3_t& v, long long row) {
    return make_vec3(view_ld(v, row, 0), view_ld(v, row, 1), view_ld(v, row, 2));
}
__device__ __forceinline__ Quat view_ld_quat(const parc_view3_t& v, long long row) {
    return make_quat(view_ld(v, row, 0), view_ld(v, row, 1), view_ld(v, row, 2), view_ld(v, row, 3));
}
__device__ __forceinline__ void view_st_vec3(const parc_view3_t& v, long long row, Vec3 a) {
    view_st(v, row, 0, a.x); view_st(v, row, 1, a.y); view_st(v, row, 2, a.z);
}
__device__ __forceinline__ void view_st_quat(const parc_view3_t& v, long long row, Quat q) {
    view_st(v, row, 0, q.x); view_st(v, row, 1, q.y); view_st(v, row, 2, q.z); view_st(v, row, 3, q.w);
}

}  // namespace parc
