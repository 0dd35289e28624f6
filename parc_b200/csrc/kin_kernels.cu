// KinCharModel kernels: forward kinematics, dof <-> rotation, frame velocities.
// Replaces the per-joint Python loops of PARC/anim/kin_char_model.py:586-690.
#include <stdarg.h>
#include <string.h>

#include "parc_common.cuh"
#include "parc_lanes.cuh"

namespace parc {

static thread_local char g_error[512] = "";

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_error, sizeof(g_error), fmt, ap);
    va_end(ap);
}

int check_launch(const char* what) {
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        set_error("%s: %s", what, cudaGetErrorString(e));
        return PARC_ERR_CUDA;
    }
    return PARC_OK;
}

// ---- forward kinematics ---------------------------------------------------------------------
// One group of G lanes per pose; lane j loads joint quaternion j-1 as one 128-bit load (the J-1
// quaternions of a pose are contiguous, so a group reads a coalesced 16*(J-1) byte run).
template <int G>
__global__ void __launch_bounds__(256) fk_kernel(const __grid_constant__ parc_char_model_t cm,
                                                 const float* __restrict__ root_pos,
                                                 const float* __restrict__ root_rot,
                                                 const float* __restrict__ joint_rot, long long n,
                                                 float* __restrict__ body_pos, float* __restrict__ body_rot) {
    const LaneBody lb = load_lane_body<G>(cm);
    const int J = cm.num_bodies;
    const long long groups_per_block = blockDim.x / G;
    const long long gid0 = (long long)blockIdx.x * groups_per_block + threadIdx.x / G;
    const long long stride = (long long)gridDim.x * groups_per_block;
    // all lanes of a warp must run the same number of iterations (shuffles inside): iterate on the
    // warp's first group and mask per group.
    const long long warp_first = gid0 - ((threadIdx.x & 31) / G);
    for (long long base = warp_first; base < n; base += stride) {
        const long long i = base + ((threadIdx.x & 31) / G);
        const bool live = i < n;
        const long long ii = live ? i : 0;
        Vec3 rp = make_vec3(__ldg(root_pos + ii * 3), __ldg(root_pos + ii * 3 + 1), __ldg(root_pos + ii * 3 + 2));
        Quat rr = quat_from_f4(__ldg(reinterpret_cast<const float4*>(root_rot) + ii));
        Quat q = quat_identity();
        if (lb.active && lb.body > 0)
            q = quat_from_f4(__ldg(reinterpret_cast<const float4*>(joint_rot) + ii * (J - 1) + (lb.body - 1)));
        Vec3 pos; Quat rot;
        fk_lanes<G>(lb, q, rp, rr, pos, rot);
        if (live && lb.active) {
            if (body_pos) {
                float* o = body_pos + (ii * J + lb.body) * 3;
                o[0] = pos.x; o[1] = pos.y; o[2] = pos.z;
            }
            if (body_rot) reinterpret_cast<float4*>(body_rot)[ii * J + lb.body] = quat_to_f4(rot);
        }
    }
}

// ---- forward kinematics, vector-Jacobian product ---------------------------------------------
// One thread per pose: recompute the forward pass into local arrays, then sweep children ->
// parents accumulating d/d(body_pos), d/d(body_rot). Poses in a motion-optimizer batch number
// ~3e5 (config 4), so thread-per-pose already fills the machine; J <= 32 keeps the arrays small.
__device__ __forceinline__ void quat_mul_bwd(Quat a, Quat b, Quat g, Quat& ga, Quat& gb) {
    // r = a (x) b (Hamilton). ga = g (x) conj(b), gb = conj(a) (x) g  (as 4-vectors, no conj sign on w)
    ga.x = g.x * b.w - g.y * b.z + g.z * b.y - g.w * b.x;
    ga.y = g.x * b.z + g.y * b.w - g.z * b.x - g.w * b.y;
    ga.z = -g.x * b.y + g.y * b.x + g.z * b.w - g.w * b.z;
    ga.w = g.x * b.x + g.y * b.y + g.z * b.z + g.w * b.w;
    gb.x = a.w * g.x + a.z * g.y - a.y * g.z - a.x * g.w;
    gb.y = -a.z * g.x + a.w * g.y + a.x * g.z - a.y * g.w;
    gb.z = a.y * g.x - a.x * g.y + a.w * g.z - a.z * g.w;
    gb.w = a.x * g.x + a.y * g.y + a.z * g.z + a.w * g.w;
}

// out = v + w t + qv x t, t = 2 qv x v (quat_rotate, torch_util.py:61-67); gradient wrt q (v const)
__device__ __forceinline__ Quat quat_rotate_bwd_q(Quat q, Vec3 v, Vec3 g) {
    Vec3 qv = make_vec3(q.x, q.y, q.z);
    Vec3 t = 2.f * cross(qv, v);
    // out = v + w*t + cross(qv, t)
    float gw = g.x * t.x + g.y * t.y + g.z * t.z;
    // d/dt: w*g + cross(g, qv)   [since d(qv x t)/dt^T g = g x qv]
    Vec3 gt = q.w * g + cross(g, qv);
    // d/dqv from cross(qv, t): t x g ; from t = 2 qv x v: 2 * (v x gt)
    Vec3 gq = cross(t, g) + 2.f * cross(v, gt);
    return make_quat(gq.x, gq.y, gq.z, gw);
}

__global__ void __launch_bounds__(128) fk_bwd_kernel(const __grid_constant__ parc_char_model_t cm,
                                                     const float* __restrict__ root_pos,
                                                     const float* __restrict__ root_rot,
                                                     const float* __restrict__ joint_rot, long long n,
                                                     const float* __restrict__ g_body_pos,
                                                     const float* __restrict__ g_body_rot,
                                                     float* __restrict__ g_root_pos,
                                                     float* __restrict__ g_root_rot,
                                                     float* __restrict__ g_joint_rot) {
    const int J = cm.num_bodies;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        Quat rot[PARC_MAX_BODIES];
        Vec3 gp[PARC_MAX_BODIES];
        Quat gr[PARC_MAX_BODIES];
        rot[0] = quat_from_f4(reinterpret_cast<const float4*>(root_rot)[i]);
        for (int j = 1; j < J; ++j) {
            Quat q = quat_from_f4(reinterpret_cast<const float4*>(joint_rot)[i * (J - 1) + (j - 1)]);
            Quat lr = make_quat(cm.local_rot[j][0], cm.local_rot[j][1], cm.local_rot[j][2], cm.local_rot[j][3]);
            rot[j] = quat_mul(rot[cm.parent[j]], quat_mul(lr, q));
        }
        for (int j = 0; j < J; ++j) {
            gp[j] = g_body_pos ? make_vec3(g_body_pos[(i * J + j) * 3], g_body_pos[(i * J + j) * 3 + 1], g_body_pos[(i * J + j) * 3 + 2])
                               : make_vec3(0.f, 0.f, 0.f);
            gr[j] = g_body_rot ? quat_from_f4(reinterpret_cast<const float4*>(g_body_rot)[i * J + j]) : make_quat(0.f, 0.f, 0.f, 0.f);
        }
        for (int j = J - 1; j >= 1; --j) {
            const int p = cm.parent[j];
            Quat q = quat_from_f4(reinterpret_cast<const float4*>(joint_rot)[i * (J - 1) + (j - 1)]);
            Quat lr = make_quat(cm.local_rot[j][0], cm.local_rot[j][1], cm.local_rot[j][2], cm.local_rot[j][3]);
            Vec3 tr = make_vec3(cm.local_trans[j][0], cm.local_trans[j][1], cm.local_trans[j][2]);
            Quat cq = quat_mul(lr, q);
            // rot_j = rot_p (x) cq
            Quat g_rp, g_cq;
            quat_mul_bwd(rot[p], cq, gr[j], g_rp, g_cq);
            // cq = lr (x) q
            Quat g_lr, g_q;
            quat_mul_bwd(lr, q, g_cq, g_lr, g_q);
            if (g_joint_rot) reinterpret_cast<float4*>(g_joint_rot)[i * (J - 1) + (j - 1)] = quat_to_f4(g_q);
            // pos_j = pos_p + rotate(rot_p, tr)
            Quat g_rot_from_pos = quat_rotate_bwd_q(rot[p], tr, gp[j]);
            gp[p] = gp[p] + gp[j];
            gr[p] = make_quat(gr[p].x + g_rp.x + g_rot_from_pos.x, gr[p].y + g_rp.y + g_rot_from_pos.y,
                              gr[p].z + g_rp.z + g_rot_from_pos.z, gr[p].w + g_rp.w + g_rot_from_pos.w);
        }
        if (g_root_pos) { g_root_pos[i * 3] = gp[0].x; g_root_pos[i * 3 + 1] = gp[0].y; g_root_pos[i * 3 + 2] = gp[0].z; }
        if (g_root_rot) reinterpret_cast<float4*>(g_root_rot)[i] = quat_to_f4(gr[0]);
    }
}

// ---- dof <-> rot ------------------------------------------------------------------------------
template <int G>
__global__ void __launch_bounds__(256) dof_to_rot_kernel(const __grid_constant__ parc_char_model_t cm,
                                                         const float* __restrict__ dof, long long n,
                                                         float* __restrict__ rot) {
    const LaneBody lb = load_lane_body<G>(cm);
    const int J = cm.num_bodies, D = cm.dof_size;
    const long long gpb = blockDim.x / G;
    for (long long i = (long long)blockIdx.x * gpb + threadIdx.x / G; i < n; i += (long long)gridDim.x * gpb) {
        if (lb.active && lb.body > 0) {
            const float* d = dof + i * D;
            Quat q = dof_to_rot_lane(lb, [&](int k) { return __ldg(d + k); });
            reinterpret_cast<float4*>(rot)[i * (J - 1) + (lb.body - 1)] = quat_to_f4(q);
        }
    }
}

template <int G>
__global__ void __launch_bounds__(256) rot_to_dof_kernel(const __grid_constant__ parc_char_model_t cm,
                                                         const float* __restrict__ rot, long long n,
                                                         float* __restrict__ dof) {
    const LaneBody lb = load_lane_body<G>(cm);
    const int J = cm.num_bodies, D = cm.dof_size;
    const long long gpb = blockDim.x / G;
    for (long long i = (long long)blockIdx.x * gpb + threadIdx.x / G; i < n; i += (long long)gridDim.x * gpb) {
        if (lb.active && lb.body > 0) {
            Quat q = quat_from_f4(__ldg(reinterpret_cast<const float4*>(rot) + i * (J - 1) + (lb.body - 1)));
            float out[3];
            int nd = rot_to_dof_lane(lb, q, out);
            for (int k = 0; k < nd; ++k) dof[i * D + lb.dof_idx + k] = out[k];
        }
    }
}

// ---- frame velocities (load-time precompute) ---------------------------------------------------
// motion_lib.py:318-327: root_vel = fps*(p[t+1]-p[t]); root_ang_vel = fps*exp_map(q[t+1] (x) conj(q[t]));
// kin_char_model.py:661-690: drot = normalize(conj(q0) (x) q1); spherical: exp_map(drot)/dt;
// hinge: dot(axis, exp_map(drot)/dt). Last frame copies the one before.
// One group of G lanes per frame. With clip tables (num_clips > 0) the frames of ALL clips of a library are done in one
// launch: the group finds its clip by binary search over the exclusive frame offsets and uses that clip's fps / dt and
// last-frame rule; without them the nf frames are one clip.
template <int G>
__global__ void __launch_bounds__(256) frame_vel_kernel(const __grid_constant__ parc_char_model_t cm,
                                                        const float* __restrict__ root_pos,
                                                        const float* __restrict__ root_rot,
                                                        const float* __restrict__ joint_rot, long long nf,
                                                        float fps1, float dt1, int num_clips, const long long* __restrict__ clip_start,
                                                        const long long* __restrict__ clip_frames, const float* __restrict__ clip_fps,
                                                        const float* __restrict__ clip_dt, float* __restrict__ root_vel,
                                                        float* __restrict__ root_ang_vel,
                                                        float* __restrict__ dof_vel) {
    const LaneBody lb = load_lane_body<G>(cm);
    const int J = cm.num_bodies, D = cm.dof_size;
    const long long gpb = blockDim.x / G;
    for (long long t = (long long)blockIdx.x * gpb + threadIdx.x / G; t < nf; t += (long long)gridDim.x * gpb) {
        long long first = 0, count = nf;
        float fps = fps1, dt = dt1;
        if (num_clips > 0) {
            int lo = 0, hi = num_clips - 1;          // last clip whose first frame is <= t
            while (lo < hi) { const int mid = (lo + hi + 1) >> 1; if (__ldg(clip_start + mid) <= t) lo = mid; else hi = mid - 1; }
            first = __ldg(clip_start + lo); count = __ldg(clip_frames + lo); fps = __ldg(clip_fps + lo); dt = __ldg(clip_dt + lo);
        }
        // frame t of the output reads the pair (s, s+1) with s = min(t, last-1) inside its clip
        const long long s = (t < first + count - 1) ? t : first + count - 2;
        if (count < 2) {
            if (lb.body == 0) for (int k = 0; k < 3; ++k) { root_vel[t * 3 + k] = 0.f; root_ang_vel[t * 3 + k] = 0.f; }
            if (lb.active && lb.body > 0) for (int k = 0; k < joint_dof_dim(lb.jtype); ++k) dof_vel[t * D + lb.dof_idx + k] = 0.f;
            continue;
        }
        if (lb.body == 0) {
            for (int k = 0; k < 3; ++k) root_vel[t * 3 + k] = fps * (root_pos[(s + 1) * 3 + k] - root_pos[s * 3 + k]);
            Quat q0 = quat_from_f4(reinterpret_cast<const float4*>(root_rot)[s]);
            Quat q1 = quat_from_f4(reinterpret_cast<const float4*>(root_rot)[s + 1]);
            Vec3 e = quat_to_exp_map(quat_mul(q1, quat_conj(q0)));   // quat_diff, torch_util.py:453-456
            root_ang_vel[t * 3] = fps * e.x; root_ang_vel[t * 3 + 1] = fps * e.y; root_ang_vel[t * 3 + 2] = fps * e.z;
        } else if (lb.active) {
            Quat q0 = quat_from_f4(reinterpret_cast<const float4*>(joint_rot)[s * (J - 1) + lb.body - 1]);
            Quat q1 = quat_from_f4(reinterpret_cast<const float4*>(joint_rot)[(s + 1) * (J - 1) + lb.body - 1]);
            Quat d = quat_normalize(quat_mul(quat_conj(q0), q1));
            Vec3 e = quat_to_exp_map(d);
            Vec3 v = make_vec3(e.x / dt, e.y / dt, e.z / dt);
            if (lb.jtype == PARC_JOINT_SPHERICAL) {
                dof_vel[t * D + lb.dof_idx] = v.x; dof_vel[t * D + lb.dof_idx + 1] = v.y; dof_vel[t * D + lb.dof_idx + 2] = v.z;
            } else if (lb.jtype == PARC_JOINT_HINGE) {
                dof_vel[t * D + lb.dof_idx] = dot3_seq(lb.axis, v);
            }
        }
    }
}

static int grid_for(long long work_items, int per_block) {
    long long blocks = (work_items + per_block - 1) / per_block;
    long long cap = (long long)kNumSMs * 16;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    return (int)blocks;
}

}  // namespace parc

using namespace parc;

extern "C" const char* parc_b200_version(void) { return "parc_b200 0.1 (sm_100a)"; }
extern "C" const char* parc_b200_last_error(void) { return g_error; }

extern "C" int parc_b200_forward_kinematics(const parc_char_model_t* m, const float* root_pos,
                                            const float* root_rot, const float* joint_rot, int64_t n,
                                            float* body_pos, float* body_rot, void* stream) {
    if (!model_ok(m, "forward_kinematics")) return PARC_ERR_INVALID_ARG;
    PARC_REQUIRE(n >= 0, "forward_kinematics: n < 0");
    if (n == 0) return PARC_OK;
    PARC_REQUIRE(root_pos && root_rot && (joint_rot || m->num_bodies == 1), "forward_kinematics: NULL input");
    if (m->num_bodies <= 16) {
        fk_kernel<16><<<grid_for(n, 16), 256, 0, as_stream(stream)>>>(*m, root_pos, root_rot, joint_rot, n, body_pos, body_rot);
    } else {
        fk_kernel<32><<<grid_for(n, 8), 256, 0, as_stream(stream)>>>(*m, root_pos, root_rot, joint_rot, n, body_pos, body_rot);
    }
    return check_launch("forward_kinematics");
}

extern "C" int parc_b200_forward_kinematics_bwd(const parc_char_model_t* m, const float* root_pos,
                                                const float* root_rot, const float* joint_rot, int64_t n,
                                                const float* g_body_pos, const float* g_body_rot,
                                                float* g_root_pos, float* g_root_rot, float* g_joint_rot,
                                                void* stream) {
    if (!model_ok(m, "forward_kinematics_bwd")) return PARC_ERR_INVALID_ARG;
    if (n == 0) return PARC_OK;
    PARC_REQUIRE(root_pos && root_rot, "forward_kinematics_bwd: NULL input");
    fk_bwd_kernel<<<grid_for(n, 128), 128, 0, as_stream(stream)>>>(*m, root_pos, root_rot, joint_rot, n, g_body_pos,
                                                                  g_body_rot, g_root_pos, g_root_rot, g_joint_rot);
    return check_launch("forward_kinematics_bwd");
}

extern "C" int parc_b200_dof_to_rot(const parc_char_model_t* m, const float* dof, int64_t n, float* rot, void* stream) {
    if (!model_ok(m, "dof_to_rot")) return PARC_ERR_INVALID_ARG;
    if (n == 0 || m->num_bodies == 1) return PARC_OK;
    PARC_REQUIRE(dof && rot, "dof_to_rot: NULL pointer");
    if (m->num_bodies <= 16) dof_to_rot_kernel<16><<<grid_for(n, 16), 256, 0, as_stream(stream)>>>(*m, dof, n, rot);
    else dof_to_rot_kernel<32><<<grid_for(n, 8), 256, 0, as_stream(stream)>>>(*m, dof, n, rot);
    return check_launch("dof_to_rot");
}

extern "C" int parc_b200_rot_to_dof(const parc_char_model_t* m, const float* rot, int64_t n, float* dof, void* stream) {
    if (!model_ok(m, "rot_to_dof")) return PARC_ERR_INVALID_ARG;
    if (n == 0 || m->num_bodies == 1) return PARC_OK;
    PARC_REQUIRE(dof && rot, "rot_to_dof: NULL pointer");
    if (m->num_bodies <= 16) rot_to_dof_kernel<16><<<grid_for(n, 16), 256, 0, as_stream(stream)>>>(*m, rot, n, dof);
    else rot_to_dof_kernel<32><<<grid_for(n, 8), 256, 0, as_stream(stream)>>>(*m, rot, n, dof);
    return check_launch("rot_to_dof");
}

extern "C" int parc_b200_frame_velocities(const parc_char_model_t* m, const float* root_pos, const float* root_rot,
                                          const float* joint_rot, int64_t nf, float fps, float dt, float* root_vel,
                                          float* root_ang_vel, float* dof_vel, void* stream) {
    if (!model_ok(m, "frame_velocities")) return PARC_ERR_INVALID_ARG;
    if (nf == 0) return PARC_OK;
    PARC_REQUIRE(root_pos && root_rot && root_vel && root_ang_vel && dof_vel, "frame_velocities: NULL pointer");
    if (m->num_bodies <= 16)
        frame_vel_kernel<16><<<grid_for(nf, 16), 256, 0, as_stream(stream)>>>(*m, root_pos, root_rot, joint_rot, nf, fps, dt, 0, nullptr, nullptr, nullptr, nullptr, root_vel, root_ang_vel, dof_vel);
    else
        frame_vel_kernel<32><<<grid_for(nf, 8), 256, 0, as_stream(stream)>>>(*m, root_pos, root_rot, joint_rot, nf, fps, dt, 0, nullptr, nullptr, nullptr, nullptr, root_vel, root_ang_vel, dof_vel);
    return check_launch("frame_velocities");
}

extern "C" int parc_b200_library_velocities(const parc_char_model_t* m, const float* root_pos, const float* root_rot,
                                            const float* joint_rot, int64_t total_frames, int32_t num_clips,
                                            const int64_t* clip_start, const int64_t* clip_frames, const float* clip_fps,
                                            const float* clip_dt, float* root_vel, float* root_ang_vel, float* dof_vel, void* stream) {
    if (!model_ok(m, "library_velocities")) return PARC_ERR_INVALID_ARG;
    PARC_REQUIRE(total_frames >= 0 && num_clips >= 0, "library_velocities: negative size");
    if (total_frames == 0) return PARC_OK;
    PARC_REQUIRE(num_clips > 0 && clip_start && clip_frames && clip_fps && clip_dt, "library_velocities: clip tables missing");
    PARC_REQUIRE(root_pos && root_rot && root_vel && root_ang_vel && dof_vel && (joint_rot || m->num_bodies == 1), "library_velocities: NULL pointer");
    const long long* cs = reinterpret_cast<const long long*>(clip_start);
    const long long* cf = reinterpret_cast<const long long*>(clip_frames);
    if (m->num_bodies <= 16)
        frame_vel_kernel<16><<<grid_for(total_frames, 16), 256, 0, as_stream(stream)>>>(*m, root_pos, root_rot, joint_rot, total_frames, 0.f, 0.f, num_clips, cs, cf, clip_fps, clip_dt, root_vel, root_ang_vel, dof_vel);
    else
        frame_vel_kernel<32><<<grid_for(total_frames, 8), 256, 0, as_stream(stream)>>>(*m, root_pos, root_rot, joint_rot, total_frames, 0.f, 0.f, num_clips, cs, cf, clip_fps, clip_dt, root_vel, root_ang_vel, dof_vel);
    return check_launch("library_velocities");
}
