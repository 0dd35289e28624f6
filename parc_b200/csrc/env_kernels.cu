// Fused post-physics step of the DeepMimic-style tracking env:
//   time -> reference motion -> observation -> reward -> termination -> heights
// i.e. IGEnv._post_physics_step (PARC/motion_tracker/envs/ig_env.py:368-377) minus PhysX, in ONE
// launch that reads every input once and writes every output once.
//
// Work decomposition (a block owns TILE = 16 consecutive rows, 512 threads = two groups of 8 envs x 8 warps):
//   A  per-env uniforms (root state, heading, time, clip meta) as two dependent-load chains in the last two warps;
//      character dof -> quaternion, one thread per (env, joint); plan tables and ray lattice into shared memory
//   B  warp w evaluates pose (w mod 8) -- 0 = reference, 1..S = target steps, S+1 = simulated character -- for 8 of the
//      tile's envs; the NPART = 4 lanes of an (env, pose) follow a host-made schedule of the joint tree: every joint is
//      evaluated once, a lane hands its transform to its next joint in registers, and a transform that another lane (or
//      a later iteration) needs goes through a shared slot of the 4-lane group. Slerp uses the pair constants packed
//      with the library (parc_math.cuh: slerp_pair / slerp_pre). Outputs go to the staged observation row or to the
//      tile's ref_* block in shared memory
//   C  reward / termination terms, flattened over the tile into warp-sized tasks
//   D  16 lanes: ordered sums, exponentials, reward, done flag, done-list append -- while
//   H  all warps sample the heightmaps (straight to global memory); rows are handed out by a shared counter
//   E  copy-out: cp.async.bulk (TMA) of each staged observation row and of the ref_* blocks, issued right after phase C
//      so that the copies run under D / H; plain stores for partial tiles, row subsets and strided ref views
#include <string.h>
#include <type_traits>

#include "parc_common.cuh"
#include "parc_lanes.cuh"

// The kernel is instruction-issue / instruction-fetch bound, so loops that run a handful of iterations per block are kept
// rolled: nvcc's default 4x unrolling of them only adds code.
#define PARC_COLD_LOOP _Pragma("unroll 1")
// Tuning constants, each set by an A/B measurement on a B200:
#ifndef PARC_HEIGHT_GATHERS
#define PARC_HEIGHT_GATHERS 2                 // general ray sets: independent heightfield gathers per lane and loop iteration
#endif
#ifndef PARC_LATTICE_ROWS
#define PARC_LATTICE_ROWS 8                   // lattice ray sets: lattice rows (= independent gathers) per lane and loop iteration
#endif
#ifndef PARC_HEIGHT_SPLIT
#define PARC_HEIGHT_SPLIT 2                   // an env's sample points are handed out in this many pieces (load balance across the warps)
#endif
#ifndef PARC_PREFETCH_FRAMES
#define PARC_PREFETCH_FRAMES 0                // 1: a separate stage prefetches the frame records before the pose loop (measured slower: the
#endif                                        //    root part of a pose touches nearly every line of its two records anyway)
#ifndef PARC_SPLIT_JOINT_TASKS
#define PARC_SPLIT_JOINT_TASKS 1              // 1: phase C deals rot_to_dof / pose error / pose termination as separate tasks
#endif
#ifndef PARC_TAIL_ROLES
#define PARC_TAIL_ROLES 1                     // 1: after phase C reward, termination and the bulk hand-over run in different warps
#endif
#ifndef PARC_HEIGHT_PATCH
#define PARC_HEIGHT_PATCH 0                   // 1 (lattice ray sets): the env's heightfield neighbourhood is staged in shared memory
#endif
#ifndef PARC_PREFETCH_NEXT_JOINT
#define PARC_PREFETCH_NEXT_JOINT 0            // 1: a pose lane asks for its next joint's record lines one iteration ahead (measured 2 % slower)
#endif
#ifndef PARC_STREAM_HINTS
#define PARC_STREAM_HINTS 1                   // 1: height rows are stored with the streaming hint (st.global.cs): phase B's frame records
#endif                                        //    want every byte of the L1 the block's shared memory leaves
#ifndef PARC_STEP_FAST
#define PARC_STEP_FAST 1                      // 1: full steps of the default humanoid layout run the compile-time-specialised instantiation
#endif
#ifndef PARC_STAGGER_NS
#define PARC_STAGGER_NS 0                     // > 0: the second block of every SM's first wave starts this much later, so that the
#endif                                        //      latency-bound phases of an SM's two resident blocks do not coincide
#ifndef PARC_TAR_FAST_MATH
#define PARC_TAR_FAST_MATH 0                  // 1 (measured: no gain, +0.3 %): the target poses (observation only, 1e-5 tolerance) use FMA-contracted quaternion products / rotations;
#endif                                        //    the reference and character poses, which feed the termination flags, keep the reference's rounding
#ifndef PARC_CELL_MAGIC
#define PARC_CELL_MAGIC 1                     // 1: a sample's cell index is rounded by a magic-number addition (FMA pipe) instead of F2I (XU pipe)
#endif
#ifndef PARC_POSE_TAIL_COPIES
#define PARC_POSE_TAIL_COPIES 1               // 1: the copies only the reference pose (velocity record) and the character pose (dof velocities, contact
#endif                                        //    flags) make are asked for at the top of the pose and done after its joint loop with their loads batched;
                                              //    0: done up front in rolled loops (each load's latency then sits on the block's critical path: those two warps
                                              //    finish phase B last)
#ifndef PARC_REF_TAIL_BY_CHAR
#define PARC_REF_TAIL_BY_CHAR 0               // 1 (measured 0.5 % slower: the target poses, not the reference pose, then end phase B): when both run, the character-pose warp (no slerps: it finishes phase B first) makes the reference pose's
#endif                                        //    velocity-record copy and contact-label lerp, and asks L1 for the rigid-body rows phase C reads
#ifndef PARC_PHASE_A_WIDE
#define PARC_PHASE_A_WIDE 0                   // 1 (measured: phase A 5.2k -> 4.9k cycles, kernel time unchanged): phase A's tables and joint constants arrive as 16-byte parameter loads (a divergent parameter read costs one pass per
#endif                                        //    distinct address), and the root chain's cos / inverse heading quaternion run in two further warps
#ifndef PARC_EARLY_ROW_PREFETCH
#define PARC_EARLY_ROW_PREFETCH 0             // 1 (measured 0.5 % slower): the L2 prefetch of the simulator rows phase C reads is issued by warps that idle in phase A (it sat between the
#endif                                        //    phase-A barrier and the poses, on every warp's path); the character-pose warps, first to finish phase B, then ask
                                              //    L1 for the rigid-body rows
#ifndef PARC_REWARD_TWO_LANES
#define PARC_REWARD_TWO_LANES 0               // 1 (measured 0.8 % slower): an env's reward is summed by two lanes of warp 0 (pose / key / root terms | dof-velocity and contact terms), the
#endif                                        //    order inside every sum unchanged: warp 0 reaches the height samples sooner
#ifndef PARC_BULK_ROWS
#define PARC_BULK_ROWS 1                      // 1: the tile's ref_* blocks leave through cp.async.bulk where alignment allows
#endif
#ifndef PARC_BULK_OBS
#define PARC_BULK_OBS 1                       // 1: so do the staged observation rows
#endif

namespace parc {

// Debug build only (-DPARC_STEP_TIMING, profiles/step_stamps.py): per-block clock stamps at the phase boundaries.
#ifdef PARC_STEP_TIMING
__device__ long long g_step_stamps[16 * 16384];
#define PARC_STAMP(k) do { if (threadIdx.x == 0 && blockIdx.x < 16384) g_step_stamps[blockIdx.x * 16 + (k)] = clock64(); } while (0)
#define PARC_STAMP_T(k, t) do { if (threadIdx.x == (t) && blockIdx.x < 16384) g_step_stamps[blockIdx.x * 16 + (k)] = clock64(); } while (0)
#else
#define PARC_STAMP(k) do { } while (0)
#define PARC_STAMP_T(k, t) do { } while (0)
#endif

#ifndef PARC_STEP_TILE
#define PARC_STEP_TILE 16                      // envs per block; a multiple of 8
#endif
constexpr int TILE = PARC_STEP_TILE;      // envs per block (== warps per block)
#ifndef PARC_STEP_MIN_BLOCKS
#define PARC_STEP_MIN_BLOCKS (32 / PARC_STEP_TILE)   // launch bound: 1,024 threads per SM at 64 registers
#endif
constexpr int STEP_THREADS = 32 * TILE;
constexpr int NPART = 4;                  // lanes cooperating on one (env, pose)
constexpr int NGROUP = STEP_THREADS / NPART;
constexpr int LATTICE_NY = 32;            // lattice ray sets take the shared-product height path when they have 32 columns

// Column layout of one observation row (IGParkourEnv._compute_obs, ig_parkour_env.py:842-965;
// compute_char_obs ig_char_env.py:581-625; compute_tar_obs mgdm_dm_util.py:404-461).
struct ObsLayout {
    int root_h;        // 0/1 leading root height
    int char_dim;      // root_h + 12 + 6(J-1) + D + 3K
    int tar_step_dim;  // 9 + 6(J-1) + 3K
    int o_tar, o_tarc, o_charc, o_hf;
    int obs_dim;
};

__host__ __device__ constexpr ObsLayout layout_of(int J, int D, int C, int K, int S, int P, bool root_h, bool tar_obs, bool contact_info) {
    ObsLayout L{};
    L.root_h = root_h ? 1 : 0;
    L.char_dim = L.root_h + 12 + 6 * (J - 1) + D + 3 * K;
    L.tar_step_dim = 9 + 6 * (J - 1) + 3 * K;
    L.o_tar = L.char_dim;
    L.o_tarc = L.o_tar + (tar_obs ? S * L.tar_step_dim : 0);
    L.o_charc = L.o_tarc + ((contact_info && tar_obs) ? S * C : 0);
    L.o_hf = L.o_charc + (contact_info ? J : 0);
    L.obs_dim = L.o_hf + P;
    return L;
}
__host__ __device__ inline ObsLayout make_layout(int J, int D, int C, const parc_env_config_t& cfg) {
    return layout_of(J, D, C, cfg.num_key_bodies, cfg.num_tar_steps, cfg.num_ray_points, cfg.root_height_obs != 0, cfg.enable_tar_obs != 0,
                     cfg.use_contact_info != 0);
}

// Reference-motion outputs (DeepMimicEnv._update_ref_motion, dm_env.py:523-545). The tile's ref_* values are staged
// TENSOR-MAJOR: field f of width w_f occupies floats [off_f * TILE, (off_f + w_f) * TILE) as [TILE][w_f] -- which is
// exactly the memory image of rows e0 .. e0 + TILE - 1 of a contiguous [N, w_f] tensor, so a full tile leaves with one
// bulk copy per tensor.
struct RefLayout {
    int root_pos, root_rot, root_vel, root_ang_vel, body_pos, joint_rot, dof_pos, dof_vel, contacts, width;
};
__host__ __device__ constexpr RefLayout make_ref_layout(int J, int D, int C) {
    RefLayout R{};
    R.root_pos = 0; R.root_rot = 3; R.root_vel = 7; R.root_ang_vel = 10; R.body_pos = 13;
    R.joint_rot = R.body_pos + 3 * J;
    R.dof_pos = R.joint_rot + 4 * (J - 1);
    R.dof_vel = R.dof_pos + D;
    R.contacts = R.dof_vel + D;
    R.width = R.contacts + C;
    return R;
}
constexpr int NREF = 9;   // ref_* tensors

// Per-env uniforms computed once (phase A) and read by every later phase.
struct EnvU {
    long long e;             // env index of this row
    Vec3 c_pos; Quat c_rot; Vec3 c_vel; Vec3 c_avel;
    Quat hinv;               // inverse heading rotation of the CHARACTER root
    Vec3 eo;                 // env offset
    Vec3 g;                  // global root position (c_pos + env offset)
    float sh, ch;            // sin / cos of the character heading
    float time, t0, tox, toy;
    MotionMeta mm;
    int live;
    int row_shift;           // floats the staged row is shifted by (16-byte phase of the env's obs_buf row)
    int flags;               // phase C: bit0 fall contact, bit1 fall height, bit2 pose failure
};

// Shared-memory plan (dynamic): sized on the host.
struct SmemPlan {
    int row_pitch;           // floats per staged row: o_hf (+ up to 3 floats of alignment shift), = 4 mod 32
    int n_pose;              // poses per env: S + 2
    int n_gslots;            // transforms a 4-lane group keeps in shared memory
    int term_pitch;          // floats of reward terms per env: (J-1) + D + J + K
    size_t off_env, off_row, off_ref, off_charq, off_gx, off_terms, off_refmap, off_tab, off_pts, off_lat, off_patch, off_ctr, total;
};

// One entry of the joint schedule: what lane `part` of an (env, pose) group does in iteration `it`.
//   bits 0-7 joint | 8-15 mask of key-body slots that are this joint | 16 parent transform comes from a group slot (else:
//   from this lane's registers = its previous joint, or the root) | 17-21 slot to read | 22 publish to a slot |
//   23-27 slot to write | 28 body has children (its rotation is needed);   -1 = idle
constexpr int SCHED_FROM_SLOT = 1 << 16, SCHED_RSLOT_SHIFT = 17, SCHED_PUBLISH = 1 << 22, SCHED_WSLOT_SHIFT = 23, SCHED_NEEDS_ROT = 1 << 28;
constexpr int SCHED_MAX_SLOTS = 32;

// Everything the kernel would otherwise re-derive per thread from the model / config, computed once
// on the host per launch and passed by value.
struct StepPlan {
    ObsLayout L;
    RefLayout R;
    SmemPlan S;
    unsigned key_mask;                    // bit j: body j is listed in key_body_ids
    int lrot_ident;                       // every _local_rotation is the identity
    int n_iter;                           // iterations of the joint schedule
    int root_slot;                        // >= 0: the root transform is published to this group slot before the joint loop
    alignas(16) int sched[PARC_MAX_BODIES * NPART];   // [it][part]
    alignas(16) float jc[PARC_MAX_BODIES * 8];        // per joint: local translation[3] | local rotation[4] | 0 (copied to shared memory in phase A)
    alignas(16) int jd[PARC_MAX_BODIES * 4];          // per joint: joint type | dof index << 8, hinge axis x, y, z (float bits): one 16-byte read in phase A
    float rdx, rdy;                       // correctly rounded 1/dx, 1/dy
    int lat_nx, lat_ny;                   // > 0: ray_xy_points is the lattice xs[i] x ys[j], point i * ny + j (cfg.ray_grid_*)
    int ref_fast;                         // every ref_* view is a contiguous, 16-byte aligned [N, w] tensor
};

// List scheduling of the joints 1..J-1 on NPART lanes. A joint may run once its parent ran in an EARLIER iteration
// (transforms change hands between iterations). A lane first continues below the joint it has just finished (the parent
// is then in its registers); free lanes take the ready joint with the longest chain below it. A parent that is not
// "the same lane, previous iteration" is read from a group slot; slots are reused once their last reader is done.
inline void plan_schedule(const parc_char_model_t& cm, const parc_env_config_t& cfg, StepPlan& sp) {
    const int J = cm.num_bodies;
    int height[PARC_MAX_BODIES], iter_of[PARC_MAX_BODIES], lane_of[PARC_MAX_BODIES], nkids[PARC_MAX_BODIES];
    for (int j = 0; j < J; ++j) { height[j] = 1; iter_of[j] = -1; lane_of[j] = -1; nkids[j] = 0; }
    for (int j = J - 1; j >= 1; --j) {
        const int p = cm.parent[j];
        if (height[j] + 1 > height[p]) height[p] = height[j] + 1;
        ++nkids[p];
    }
    iter_of[0] = -1;   // the root is available before iteration 0
    int prev[NPART];
    for (int p = 0; p < NPART; ++p) prev[p] = -1;
    for (int k = 0; k < PARC_MAX_BODIES * NPART; ++k) sp.sched[k] = -1;
    int remaining = J - 1, it = 0;
    int joint_at[PARC_MAX_BODIES][NPART];
    while (remaining > 0) {
        int chosen[NPART];
        for (int p = 0; p < NPART; ++p) {
            chosen[p] = -1;
            if (prev[p] < 0) continue;
            for (int c = prev[p] + 1; c < J; ++c)
                if (cm.parent[c] == prev[p] && iter_of[c] < 0 && (chosen[p] < 0 || height[c] > height[chosen[p]])) chosen[p] = c;
            if (chosen[p] >= 0) iter_of[chosen[p]] = it;
        }
        for (int p = 0; p < NPART; ++p) {
            if (chosen[p] >= 0) continue;
            int best = -1;
            for (int c = 1; c < J; ++c) {
                if (iter_of[c] >= 0) continue;
                const int par = cm.parent[c];
                const bool ready = par == 0 || (iter_of[par] >= 0 && iter_of[par] < it);
                if (ready && (best < 0 || height[c] > height[best])) best = c;
            }
            if (best >= 0) { chosen[p] = best; iter_of[best] = it; }
        }
        for (int p = 0; p < NPART; ++p) {
            joint_at[it][p] = chosen[p];
            if (chosen[p] >= 0) { lane_of[chosen[p]] = p; --remaining; }
            prev[p] = chosen[p];
        }
        ++it;
    }
    sp.n_iter = it;
    // which parents go through a slot, and until when they are needed
    int last_reader[PARC_MAX_BODIES];
    bool from_slot[PARC_MAX_BODIES];
    for (int j = 0; j < J; ++j) { last_reader[j] = -1; from_slot[j] = false; }
    for (int c = 1; c < J; ++c) {
        const int par = cm.parent[c];
        bool in_regs;
        if (par == 0) {
            // root: in registers as long as the lane has not evaluated any joint yet
            in_regs = true;
            for (int t = 0; t < iter_of[c]; ++t) if (joint_at[t][lane_of[c]] >= 0) in_regs = false;
        } else {
            in_regs = lane_of[par] == lane_of[c] && iter_of[par] == iter_of[c] - 1;
        }
        from_slot[c] = !in_regs;
        if (!in_regs && iter_of[c] > last_reader[par]) last_reader[par] = iter_of[c];
    }
    // interval colouring: body j occupies a slot over [iter_of[j], last_reader[j]] (the root from -1)
    int slot_of[PARC_MAX_BODIES], slot_free_at[SCHED_MAX_SLOTS], ns = 0;
    for (int s = 0; s < SCHED_MAX_SLOTS; ++s) slot_free_at[s] = -2;
    for (int j = 0; j < J; ++j) slot_of[j] = -1;
    for (int t = -1; t < it; ++t)
        for (int j = 0; j < J; ++j) {
            if (last_reader[j] < 0 || iter_of[j] != t) continue;
            int s = 0;
            while (slot_free_at[s] >= t) ++s;               // free once the last reader's iteration is over; J <= 32 bodies never need more than 32 slots
            slot_of[j] = s; slot_free_at[s] = last_reader[j];
            if (s + 1 > ns) ns = s + 1;
        }
    sp.root_slot = slot_of[0];
    for (int t = 0; t < it; ++t)
        for (int p = 0; p < NPART; ++p) {
            const int j = joint_at[t][p];
            if (j < 0) continue;
            int e = j;
            for (int k = 0; k < cfg.num_key_bodies; ++k) if (cfg.key_body_ids[k] == j) e |= 1 << (8 + k);
            if (from_slot[j]) e |= SCHED_FROM_SLOT | (slot_of[cm.parent[j]] << SCHED_RSLOT_SHIFT);
            if (slot_of[j] >= 0) e |= SCHED_PUBLISH | (slot_of[j] << SCHED_WSLOT_SHIFT);
            if (nkids[j] > 0) e |= SCHED_NEEDS_ROT;
            sp.sched[t * NPART + p] = e;
        }
    sp.S.n_gslots = ns > 0 ? ns : 1;
}

__host__ __device__ constexpr size_t plan_take(size_t& o, size_t bytes) { const size_t at = o; o = (o + bytes + 15) & ~size_t(15); return at; }
__host__ __device__ constexpr size_t plan_max(size_t a, size_t b) { return a > b ? a : b; }
// o_hf = staged observation columns, ref_w = RefLayout::width, n_pose = S + 2, lat_nx / lat_ny > 0 for a lattice ray set
__host__ __device__ constexpr SmemPlan smem_plan_of(int J, int D, int K, int P, int o_hf, int ref_w, int n_pose, int n_gslots, int lat_nx, int lat_ny) {
    SmemPlan p{};
    // A row starts at a multiple of 4 floats plus the env's shift (0..3), so that column k of the staged row and of the
    // obs_buf row have the same 16-byte phase. Pitch = 4 x odd: the 8 envs a warp touches at one column land in banks
    // 4 (el x odd mod 8) + shift -- all different.
    p.row_pitch = (o_hf + 3 + 3) & ~3;
    if (((p.row_pitch >> 2) & 1) == 0) p.row_pitch += 4;
    p.n_pose = n_pose;
    p.n_gslots = n_gslots;
    p.term_pitch = (J - 1) + D + J + K;
    size_t o = 0;
    p.off_env = plan_take(o, sizeof(EnvU) * TILE);
    p.off_row = plan_take(o, sizeof(float) * TILE * p.row_pitch);
    p.off_ref = plan_take(o, sizeof(float) * TILE * ref_w);
    // character joint quaternions (phases A..C); the heights (after phase C) keep one lattice-row table per warp in the same bytes
    p.off_charq = p.off_lat = plan_take(o, plan_max(sizeof(float4) * TILE * J, lat_nx > 0 ? sizeof(float2) * TILE * lat_nx : 0));
    // The group slots are only live in phase B; the reward terms (phase C/D) and the ref_* scatter map (built in phase C,
    // read in phase E) reuse the same bytes.
    const size_t gx_bytes = sizeof(float) * 7 * n_gslots * NGROUP;
    const size_t terms_bytes = (sizeof(float) * TILE * p.term_pitch + 15) & ~size_t(15);
    p.off_gx = p.off_terms = o;
    p.off_refmap = o + terms_bytes;
    plan_take(o, plan_max(gx_bytes, terms_bytes + 16 * (size_t)ref_w));
    p.off_tab = plan_take(o, sizeof(int) * NPART * PARC_MAX_BODIES + sizeof(float) * 8 * J);
    if (lat_nx > 0) {
        p.off_pts = plan_take(o, sizeof(float) * (lat_nx + lat_ny));            // lattice axes xs | ys
        p.off_patch = PARC_HEIGHT_PATCH ? plan_take(o, sizeof(float) * 16 * 17 * TILE) : o;   // one heightfield patch per warp (PATCH x PATCH_PITCH)
    } else {
        p.off_pts = plan_take(o, sizeof(float2) * (size_t)P);                   // ray_xy_points, shared by the tile's envs
        p.off_patch = o;
    }
    p.off_ctr = plan_take(o, 16);
    p.total = o;
    return p;
}
inline void make_smem_plan(const parc_char_model_t& cm, int C, const parc_env_config_t& cfg, StepPlan& sp) {
    const int n_gslots = sp.S.n_gslots;      // set by plan_schedule
    sp.S = smem_plan_of(cm.num_bodies, cm.dof_size, cfg.num_key_bodies, cfg.num_ray_points, sp.L.o_hf, sp.R.width,
                        (cfg.enable_tar_obs ? cfg.num_tar_steps : 0) + 2, n_gslots, sp.lat_nx, sp.lat_ny);
}

// The layout the specialised instantiation of the step kernel is compiled for: the reference's shipped humanoid
// (data/assets/humanoid.xml: 15 bodies, 28 dofs) with the shipped observation config (dm_env_default.yaml: 6 target steps,
// 4 key bodies, contact labels, local observations) on a 32 x 32 lattice of ray points. Every size, column offset and
// shared-memory offset is then a compile-time constant; anything else runs the generic instantiation.
struct FastDims {
    static constexpr int J = 15, D = 28, C = 15, S = 6, K = 4, NX = 32, NY = 32, P = NX * NY, N_ITER = 4;
};
constexpr ObsLayout kFastL = layout_of(FastDims::J, FastDims::D, FastDims::C, FastDims::K, FastDims::S, FastDims::P, false, true, true);
constexpr RefLayout kFastR = make_ref_layout(FastDims::J, FastDims::D, FastDims::C);
constexpr SmemPlan kFastS = smem_plan_of(FastDims::J, FastDims::D, FastDims::K, FastDims::P, kFastL.o_hf, kFastR.width, FastDims::S + 2, 1, FastDims::NX, FastDims::NY);

static bool view_is_rows(const parc_view3_t& v, int a, int b) {
    // [N, a, b] (b = 0: [N, a]) row-major without gaps, base 16-byte aligned
    if (!v.ptr || (reinterpret_cast<uintptr_t>(v.ptr) & 15)) return false;
    if (b == 0) return v.row_stride == a && v.elem_stride == 1;
    return v.row_stride == (long long)a * b && v.mid_stride == b && v.elem_stride == 1;
}

inline StepPlan make_step_plan(const parc_char_model_t& cm, int C, const parc_env_config_t& cfg, const parc_terrain_t& tr,
                               const parc_env_buffers_t* env = nullptr) {
    StepPlan sp;
    const int J = cm.num_bodies, D = cm.dof_size;
    sp.L = make_layout(J, D, C, cfg);
    sp.R = make_ref_layout(J, D, C);
    sp.key_mask = 0;
    for (int k = 0; k < cfg.num_key_bodies; ++k) sp.key_mask |= 1u << cfg.key_body_ids[k];
    sp.lrot_ident = 1;
    for (int j = 1; j < J; ++j)
        if (!(cm.local_rot[j][0] == 0.f && cm.local_rot[j][1] == 0.f && cm.local_rot[j][2] == 0.f && cm.local_rot[j][3] == 1.f)) sp.lrot_ident = 0;
    plan_schedule(cm, cfg, sp);
    for (int j = 0; j < PARC_MAX_BODIES; ++j) {
        for (int c = 0; c < 8; ++c) sp.jc[j * 8 + c] = j < J ? (c < 3 ? cm.local_trans[j][c] : (c < 7 ? cm.local_rot[j][c - 3] : 0.f)) : 0.f;
        sp.jd[j * 4] = j < J ? (cm.joint_type[j] | (cm.dof_idx[j] << 8)) : 0;
        for (int c = 0; c < 3; ++c) { const float a = j < J ? cm.axis[j][c] : 0.f; memcpy(&sp.jd[j * 4 + 1 + c], &a, sizeof(float)); }
    }
    // the lattice sampler gives lane l the lattice column l (32 columns) and walks PARC_LATTICE_ROWS rows per iteration
    const bool lat = cfg.ray_grid_nx > 0 && cfg.ray_grid_ny > 0 && cfg.ray_grid_nx * cfg.ray_grid_ny == cfg.num_ray_points &&
                     cfg.ray_grid_ny == LATTICE_NY && cfg.ray_grid_nx % PARC_LATTICE_ROWS == 0;
    sp.lat_nx = lat ? cfg.ray_grid_nx : 0;
    sp.lat_ny = lat ? cfg.ray_grid_ny : 0;
    make_smem_plan(cm, C, cfg, sp);
    sp.rdx = 1.0f / tr.dx; sp.rdy = 1.0f / tr.dy;    // IEEE single division on the host == RN(1/d)
    sp.ref_fast = 0;
    if (env)
        sp.ref_fast = view_is_rows(env->ref_root_pos, 3, 0) && view_is_rows(env->ref_root_rot, 4, 0) && view_is_rows(env->ref_root_vel, 3, 0) &&
                      view_is_rows(env->ref_root_ang_vel, 3, 0) && view_is_rows(env->ref_body_pos, J, 3) &&
                      (J == 1 || view_is_rows(env->ref_joint_rot, J - 1, 4)) && (D == 0 || (view_is_rows(env->ref_dof_pos, D, 0) &&
                      view_is_rows(env->ref_dof_vel, D, 0))) && (C == 0 || view_is_rows(env->ref_contacts, C, 0));
    return sp;
}

__device__ __forceinline__ void st_stream(float* p, float x) {
#if PARC_STREAM_HINTS
    __stcs(p, x);
#else
    *p = x;
#endif
}
__device__ __forceinline__ float4 ldg4(const float* p) { return __ldg(reinterpret_cast<const float4*>(p)); }
__device__ __forceinline__ float2 ldg2(const float* p) { return __ldg(reinterpret_cast<const float2*>(p)); }

// rotate_2d_vec (torch_util.py:650-662) + root xy, products rounded separately (bit-exact cells)
__device__ __forceinline__ void ray_world(float px, float py, float c, float s, float rx, float ry, float& wx, float& wy) {
    wx = __fadd_rn(__fsub_rn(__fmul_rn(px, c), __fmul_rn(py, s)), rx);
    wy = __fadd_rn(__fadd_rn(__fmul_rn(px, s), __fmul_rn(py, c)), ry);
}

// quat_to_tan_norm (torch_util.py:392-404) = [rotate(q, x), rotate(q, z)] with the multiplications by
// the 0 / 1 components of the unit vectors folded away. Only exact operations are removed
// (a*0, a*1, a+0), so the result is bit-identical to the general quat_rotate.
__device__ __forceinline__ void tan_norm_exact(Quat q, float* __restrict__ o) {
    {   // v = (1,0,0): c0 = (0, qz, -qy), t = (0, 2qz, -2qy)
        const float ty = __fmul_rn(2.f, q.z), tz = __fmul_rn(2.f, -q.y);
        const float c1x = __fmaf_rn(q.y, tz, -__fmul_rn(q.z, ty));
        const float c1y = -__fmul_rn(q.x, tz);
        const float c1z = __fmul_rn(q.x, ty);
        o[0] = __fadd_rn(1.f, c1x);
        o[1] = __fadd_rn(__fmul_rn(q.w, ty), c1y);
        o[2] = __fadd_rn(__fmul_rn(q.w, tz), c1z);
    }
    {   // v = (0,0,1): c0 = (qy, -qx, 0), t = (2qy, -2qx, 0)
        const float tx = __fmul_rn(2.f, q.y), ty = __fmul_rn(2.f, -q.x);
        const float c1x = -__fmul_rn(q.z, ty);
        const float c1y = __fmul_rn(q.z, tx);
        const float c1z = __fmaf_rn(q.x, ty, -__fmul_rn(q.y, tx));
        o[3] = __fadd_rn(__fmul_rn(q.w, tx), c1x);
        o[4] = __fadd_rn(__fmul_rn(q.w, ty), c1y);
        o[5] = __fadd_rn(1.f, c1z);
    }
}

// FMA-contracted versions for values that only feed the observation (tolerance 1e-5, no flag or index depends on them)
__device__ __forceinline__ Quat quat_mul_fast(Quat a, Quat b) {
    return make_quat(fmaf(a.w, b.x, fmaf(a.x, b.w, fmaf(a.y, b.z, -a.z * b.y))), fmaf(a.w, b.y, fmaf(a.y, b.w, fmaf(a.z, b.x, -a.x * b.z))),
                     fmaf(a.w, b.z, fmaf(a.z, b.w, fmaf(a.x, b.y, -a.y * b.x))), fmaf(a.w, b.w, -fmaf(a.x, b.x, fmaf(a.y, b.y, a.z * b.z))));
}
__device__ __forceinline__ Vec3 quat_rotate_fast(Quat q, Vec3 v) {      // v + 2 (w c + q_v x c), c = q_v x v
    const float cx = fmaf(q.y, v.z, -q.z * v.y), cy = fmaf(q.z, v.x, -q.x * v.z), cz = fmaf(q.x, v.y, -q.y * v.x);
    const float dx = fmaf(q.w, cx, fmaf(q.y, cz, -q.z * cy)), dy = fmaf(q.w, cy, fmaf(q.z, cx, -q.x * cz)), dz = fmaf(q.w, cz, fmaf(q.x, cy, -q.y * cx));
    return make_vec3(fmaf(2.f, dx, v.x), fmaf(2.f, dy, v.y), fmaf(2.f, dz, v.z));
}
__device__ __forceinline__ void tan_norm_fast(Quat q, float* __restrict__ o) {      // R(q) x_hat, R(q) z_hat
    const float x2 = q.x + q.x, y2 = q.y + q.y, z2 = q.z + q.z;
    o[0] = 1.f - fmaf(q.y, y2, q.z * z2);
    o[1] = fmaf(q.x, y2, q.w * z2);
    o[2] = fmaf(q.x, z2, -q.w * y2);
    o[3] = fmaf(q.x, z2, q.w * y2);
    o[4] = fmaf(q.y, z2, -q.w * x2);
    o[5] = 1.f - fmaf(q.x, x2, q.y * y2);
}

// (x - mn) / d for the nearest-cell index with d, 1/d uniform: q = x * rc; r = fma(-q, d, x);
// q' = fma(r, rc, q) (Markstein) is the correctly rounded quotient for normal operands when rc is
// the correctly rounded reciprocal; checked against div.rn by parc_b200_selftest_div.
__device__ __forceinline__ float div_const(float x, float d, float rc) {
    const float q = __fmul_rn(x, rc);
    const float r = __fmaf_rn(-q, d, x);
    return __fmaf_rn(r, rc, q);
}

// Grid cell of one ray sample point: rotate_2d_vec + root xy, (px c - py s + gx, px s + py c + gy), then
// clamp(round_half_even((w - min) / dxdy), 0, dim - 1) (terrain_util.py:139-152). The caller supplies the four products
// a = (px c, px s) and b = (-py s, py c), each a scalar mul.rn (for a lattice of sample points they are shared by a
// whole lattice row / column). The sums and the Markstein division (div_const) are packed fp32x2 (FADD2 / FMUL2 / FFMA2,
// sm_100): x and y of the point share those instructions. Every product and sum must be rounded on its own, as torch's
// separate mul / sub / add are. ptxas (CUDA 12.9) contracts mul.rn.f32x2 + add.rn.f32x2 into FFMA2 despite the .rn
// (seen in the SASS; an FFMA2 with a -0 addend is folded back and contracted too), so the products stay scalar -- which
// it does leave alone -- and only the sums and the division are packed. parc_b200_selftest_cells checks the result
// against the scalar path on the GPU.
__device__ __forceinline__ int packed_cell_ab(float2 a, float2 b, float2 g2, float2 nmn, float2 rd2, float2 nd2, int imx, int imy, int dimy) {
    const float2 w = __fadd2_rn(__fadd2_rn(a, b), g2);
    const float2 x = __fadd2_rn(w, nmn);
    float2 q = __fmul2_rn(x, rd2);
    const float2 r = __ffma2_rn(q, nd2, x);
    q = __ffma2_rn(r, rd2, q);
#if PARC_CELL_MAGIC
    // round-half-even without the conversion unit (F2I runs on the XU pipe at a quarter of the FP32 rate, and the height samples
    // are the only heavy user of it): q + 1.5 * 2^23 has its integer value in the low mantissa bits, rounded to nearest even by the
    // addition itself -- exact for |q| < 2^22; beyond that the bit pattern still lands on the correct side of [0, dim - 1]
    // (larger exponent -> large positive, smaller exponent -> negative), and NaN goes to dim - 1
    const float2 m = __fadd2_rn(q, make_float2(12582912.0f, 12582912.0f));
    const int ix = max(0, min(__float_as_int(m.x) - 0x4B400000, imx));
    const int iy = max(0, min(__float_as_int(m.y) - 0x4B400000, imy));
#else
    const int ix = max(0, min(__float2int_rn(q.x), imx));
    const int iy = max(0, min(__float2int_rn(q.y), imy));
#endif
    return ix * dimy + iy;
}
__device__ __forceinline__ int packed_cell(float2 pt, float2 cs, float2 nsc, float2 g2, float2 nmn, float2 rd2, float2 nd2, int imx, int imy,
                                           int dimy) {
    const float2 a = make_float2(__fmul_rn(pt.x, cs.x), __fmul_rn(pt.x, cs.y));
    const float2 b = make_float2(__fmul_rn(pt.y, nsc.x), __fmul_rn(pt.y, nsc.y));
    return packed_cell_ab(a, b, g2, nmn, rd2, nd2, imx, imy, dimy);
}

struct HeightArgs {
    const float* hf;
    int dimy, imx, imy;
    float2 g2, nmn, rd2, nd2;
    float gz, lo, hi;
};

// One env's heightmap samples by one warp, general ray set (_refresh_obs_hfs, ig_parkour_env.py:518-533; get_hf_val,
// terrain_util.py:139-156). Points come from shared memory, PARC_HEIGHT_GATHERS per lane and iteration; the two
// destination rows (either may be absent) are running pointers with immediate offsets. FULL: no tail predicates.
template <bool FULL>
__device__ __forceinline__ void heights_of_env(const float2* __restrict__ s_pts, int P, int lane, const HeightArgs& h, float2 cs, float2 nsc,
                                               float* __restrict__ hrow, float* __restrict__ orow) {
    constexpr int U = PARC_HEIGHT_GATHERS;
    const bool HAS_H = hrow != nullptr, HAS_O = orow != nullptr;   // warp-uniform: predicated stores, one code copy
    const float2* sp = s_pts + lane;
    float* ph = hrow + lane;
    float* po = orow + lane;
    for (int p = lane; p < P; p += 32 * U, sp += 32 * U, ph += 32 * U, po += 32 * U) {
        float2 pt[U];
        int cell[U];
        float v[U];
#pragma unroll
        for (int k = 0; k < U; ++k) pt[k] = (FULL || p + 32 * k < P) ? sp[32 * k] : make_float2(0.f, 0.f);
#pragma unroll
        for (int k = 0; k < U; ++k) cell[k] = packed_cell(pt[k], cs, nsc, h.g2, h.nmn, h.rd2, h.nd2, h.imx, h.imy, h.dimy);
#pragma unroll
        for (int k = 0; k < U; ++k) v[k] = __ldg(h.hf + cell[k]);
#pragma unroll
        for (int k = 0; k < U; ++k) {
            const float x = fminf(fmaxf(__fsub_rn(v[k], h.gz), h.lo), h.hi);
            if (FULL || p + 32 * k < P) {
                if (HAS_H) st_stream(ph + 32 * k, x);
                if (HAS_O) st_stream(po + 32 * k, x);
            }
        }
    }
}

// The same for a lattice ray set xs[i] x ys[j] (point i * 32 + j) with 32 columns: lane l owns lattice column l; its two
// products with ys[l] are formed once per env, the two with xs[i] come from the env's table s_a (phase A), so a sample
// costs three packed adds, the packed division and the gather. A warp store covers one lattice row = 128 contiguous bytes;
// the U rows of an iteration are immediate offsets from one running pointer.
template <bool HAS_H, bool HAS_O>
__device__ __forceinline__ void heights_of_env_lattice(const float2* __restrict__ s_a, float y, int nx, int lane, const HeightArgs& h, float c,
                                                       float s, float* __restrict__ hrow, float* __restrict__ orow) {
    constexpr int U = PARC_LATTICE_ROWS;
    const float2 b = make_float2(__fmul_rn(y, -s), __fmul_rn(y, c));
    float* ph = HAS_H ? hrow + lane : nullptr;
    float* po = HAS_O ? orow + lane : nullptr;
    const float2* a = s_a;
    PARC_COLD_LOOP
    for (int i = 0; i < nx; i += U, a += U) {
        int cell[U];
        float v[U];
#pragma unroll
        for (int k = 0; k < U; ++k) cell[k] = packed_cell_ab(a[k], b, h.g2, h.nmn, h.rd2, h.nd2, h.imx, h.imy, h.dimy);
#pragma unroll
        for (int k = 0; k < U; ++k) v[k] = __ldg(h.hf + cell[k]);
#pragma unroll
        for (int k = 0; k < U; ++k) {
            const float x = fminf(fmaxf(__fsub_rn(v[k], h.gz), h.lo), h.hi);
            if (HAS_H) st_stream(ph + k * LATTICE_NY, x);
            if (HAS_O) st_stream(po + k * LATTICE_NY, x);
        }
        if (HAS_H) ph += U * LATTICE_NY;
        if (HAS_O) po += U * LATTICE_NY;
    }
}

// The lattice sampler with the env's heightfield neighbourhood staged in shared memory. The cell of a lattice point is
// monotone in its row i and in its column j (every step -- the rounded products, the three sums, the correctly rounded
// division, round-half-even -- is monotone), so the cells of all points lie between the cells of the four lattice corners.
// If that range spans at most PATCH x PATCH cells (3.1 m lattices on 0.4 m cells: <= 12), the warp copies hf[clamp(range)]
// into its patch with one round of independent loads -- the clamps of get_grid_index (terrain_util.py:139-141) are applied
// there -- and every sample is an LDS with no clamp and no global latency. Returns false (nothing written) if the range
// is too large: the caller then takes the gather path.
constexpr int PATCH = 16, PATCH_PITCH = 17, PATCH_FLOATS = PATCH * PATCH_PITCH;
__device__ __forceinline__ int2 packed_cell_raw(float2 a, float2 b, float2 g2, float2 nmn, float2 rd2, float2 nd2) {
    const float2 w = __fadd2_rn(__fadd2_rn(a, b), g2);
    const float2 x = __fadd2_rn(w, nmn);
    float2 q = __fmul2_rn(x, rd2);
    const float2 r = __ffma2_rn(q, nd2, x);
    q = __ffma2_rn(r, rd2, q);
    return make_int2(__float2int_rn(q.x), __float2int_rn(q.y));
}
template <bool HAS_H, bool HAS_O>
__device__ __forceinline__ bool heights_of_env_patch(const float2* __restrict__ s_a, float y, int i0, int i1, int nx_all, int lane, const HeightArgs& h,
                                                     float c, float s, float* __restrict__ hrow, float* __restrict__ orow, float* __restrict__ s_patch) {
    constexpr int U = PARC_LATTICE_ROWS;
    const float2 b = make_float2(__fmul_rn(y, -s), __fmul_rn(y, c));
    // cell range of the env's whole lattice: the first and the last lattice row contain the four corners
    const int2 ca = packed_cell_raw(s_a[0], b, h.g2, h.nmn, h.rd2, h.nd2), cb = packed_cell_raw(s_a[nx_all - 1], b, h.g2, h.nmn, h.rd2, h.nd2);
    const int lo_x = __reduce_min_sync(FULL_MASK, min(ca.x, cb.x)), hi_x = __reduce_max_sync(FULL_MASK, max(ca.x, cb.x));
    const int lo_y = __reduce_min_sync(FULL_MASK, min(ca.y, cb.y)), hi_y = __reduce_max_sync(FULL_MASK, max(ca.y, cb.y));
    constexpr int LIM = 1 << 28;
    if (lo_x < -LIM || lo_y < -LIM || hi_x > LIM || hi_y > LIM || hi_x - lo_x >= PATCH || hi_y - lo_y >= PATCH) return false;
    __syncwarp();                                       // the previous unit's samples have been read
#pragma unroll
    for (int k = 0; k < PATCH * PATCH / 32; ++k) {
        const int idx = lane + 32 * k, pa = idx / PATCH, pb = idx % PATCH;
        const int gx = max(0, min(lo_x + pa, h.imx)), gy = max(0, min(lo_y + pb, h.imy));
        s_patch[pa * PATCH_PITCH + pb] = __ldg(h.hf + gx * h.dimy + gy);
    }
    __syncwarp();
    const float* pz = s_patch - (lo_x * PATCH_PITCH + lo_y);
    float* ph = HAS_H ? hrow + lane : nullptr;
    float* po = HAS_O ? orow + lane : nullptr;
    const float2* a = s_a + i0;
    PARC_COLD_LOOP
    for (int i = i0; i < i1; i += U, a += U) {
        float v[U];
#pragma unroll
        for (int k = 0; k < U; ++k) {
            const int2 cc = packed_cell_raw(a[k], b, h.g2, h.nmn, h.rd2, h.nd2);
            v[k] = pz[cc.x * PATCH_PITCH + cc.y];
        }
#pragma unroll
        for (int k = 0; k < U; ++k) {
            const float x = fminf(fmaxf(__fsub_rn(v[k], h.gz), h.lo), h.hi);
            if (HAS_H) st_stream(ph + k * LATTICE_NY, x);
            if (HAS_O) st_stream(po + k * LATTICE_NY, x);
        }
        if (HAS_H) ph += U * LATTICE_NY;
        if (HAS_O) po += U * LATTICE_NY;
    }
    return true;
}

__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
// shared -> global bulk copy (TMA, no tensor map): both addresses 16-byte aligned, bytes a multiple of 16
__device__ __forceinline__ void bulk_store(void* gdst, const void* ssrc, unsigned bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst), "r"(smem_u32(ssrc)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// FAST: the layout of FastDims / kFastL / kFastR / kFastS with the shipped default flags, all rows (no env_ids), every part --
// the launcher checks all of that; sizes, offsets and flags are then compile-time constants.
template <int THREADS, bool LATTICE, bool FAST>
__global__ void __launch_bounds__(THREADS, PARC_STEP_MIN_BLOCKS) env_step_kernel(
    const __grid_constant__ parc_char_model_t cm, const __grid_constant__ parc_motion_lib_t lib,
    const __grid_constant__ parc_terrain_t terrain, const __grid_constant__ parc_env_config_t cfg,
    const __grid_constant__ parc_env_buffers_t env, const __grid_constant__ StepPlan sp, const unsigned parts,
    const long long* __restrict__ env_ids, const long long n_rows_host, const int* __restrict__ env_ids32,
    const int* __restrict__ n_rows_dev, const long long first_env) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    static_assert(!FAST || LATTICE, "the specialised instantiation is a lattice one");
    const int* n_rows_dev_ = FAST ? nullptr : n_rows_dev;
    const long long* env_ids_ = FAST ? nullptr : env_ids;
    const int* env_ids32_ = FAST ? nullptr : env_ids32;
    const long long n_rows = n_rows_dev_ ? min((long long)*n_rows_dev_, n_rows_host) : n_rows_host;
    // A block owns one tile -- except in launches whose row count is only known on the device (the refresh of the reset rows):
    // those use a small grid instead of one mostly empty block per possible tile, and a block walks its tiles here. The
    // specialised instantiation never loops (`if (FAST) break` below), so its code is that of a single pass.
    for (long long row0 = (long long)blockIdx.x * TILE; row0 < n_rows; row0 += (long long)gridDim.x * TILE) {
    const int tid = threadIdx.x;
    const int J = FAST ? FastDims::J : cm.num_bodies, D = FAST ? FastDims::D : cm.dof_size, C = FAST ? FastDims::C : lib.num_contacts;
    const int S = FAST ? FastDims::S : (cfg.enable_tar_obs ? cfg.num_tar_steps : 0);
    const int K = FAST ? FastDims::K : cfg.num_key_bodies, P = FAST ? FastDims::P : cfg.num_ray_points;
    constexpr ObsLayout FL = layout_of(FastDims::J, FastDims::D, FastDims::C, FastDims::K, FastDims::S, FastDims::P, false, true, true);
    constexpr RefLayout FR = make_ref_layout(FastDims::J, FastDims::D, FastDims::C);
    constexpr SmemPlan FS = smem_plan_of(FastDims::J, FastDims::D, FastDims::K, FastDims::P, FL.o_hf, FR.width, FastDims::S + 2, 1, FastDims::NX, FastDims::NY);
    const ObsLayout L = FAST ? FL : sp.L;
    const RefLayout R = FAST ? FR : sp.R;
    const SmemPlan plan = FAST ? FS : sp.S;
    const int lat_nx = FAST ? FastDims::NX : sp.lat_nx, lat_ny = FAST ? FastDims::NY : sp.lat_ny;
    const int irec_f = FAST ? interp_record_floats(FastDims::J, FastDims::C) : lib.irec, vrec_f = FAST ? vel_record_floats(FastDims::D) : lib.vrec;
    // config flags (dm_env_default.yaml values in the specialised instantiation)
    const bool c_global_obs = FAST ? false : cfg.global_obs != 0, c_contact_info = FAST ? true : cfg.use_contact_info != 0;
    const bool c_track_root = FAST ? true : cfg.track_root != 0, c_track_root_h = FAST ? true : cfg.track_root_h != 0;
    const bool c_early_term = FAST ? true : cfg.enable_early_termination != 0, c_pose_term = FAST ? true : cfg.pose_termination != 0;
    const int c_n_fall = FAST ? 0 : cfg.num_fall_contact_bodies;
    const bool c_never_done = FAST ? false : cfg.never_done != 0;
    const bool lrot_ident_ = FAST ? true : sp.lrot_ident != 0;
    const int n_iter = FAST ? FastDims::N_ITER : sp.n_iter, root_slot = FAST ? -1 : sp.root_slot;
    EnvU* s_env = reinterpret_cast<EnvU*>(smem_raw + plan.off_env);
    float* s_row = reinterpret_cast<float*>(smem_raw + plan.off_row);
    float* s_ref = reinterpret_cast<float*>(smem_raw + plan.off_ref);
    float4* s_charq = reinterpret_cast<float4*>(smem_raw + plan.off_charq);
    float* s_gx = reinterpret_cast<float*>(smem_raw + plan.off_gx);
    float* s_terms = reinterpret_cast<float*>(smem_raw + plan.off_terms);
    uint4* s_refmap = reinterpret_cast<uint4*>(smem_raw + plan.off_refmap);
    int* s_sched = reinterpret_cast<int*>(smem_raw + plan.off_tab);
    float* s_jc = reinterpret_cast<float*>(smem_raw + plan.off_tab + sizeof(int) * NPART * PARC_MAX_BODIES);
    float2* s_pts = reinterpret_cast<float2*>(smem_raw + plan.off_pts);     // general ray set
    float* s_axes = reinterpret_cast<float*>(smem_raw + plan.off_pts);      // lattice: xs[nx] | ys[ny]
    float2* s_lat = reinterpret_cast<float2*>(smem_raw + plan.off_lat);     // lattice, after phase C: per warp [nx] (x c, x s) of the env it samples
    float* s_patch = reinterpret_cast<float*>(smem_raw + plan.off_patch);   // lattice: one heightfield patch per warp
    int* s_ctr = reinterpret_cast<int*>(smem_raw + plan.off_ctr);          // height-unit counter
    const int ROWP = plan.row_pitch;

    const bool do_heights = FAST || (parts & PARC_STEP_HEIGHTS);
    const bool do_obs = FAST || (parts & PARC_STEP_OBS);
    const bool do_ref = FAST || (parts & PARC_STEP_REF);
    const bool do_reward = FAST || (parts & PARC_STEP_REWARD);
    const bool do_done = FAST || (parts & PARC_STEP_DONE);
    const bool do_time = FAST || (parts & PARC_STEP_TIME);
    const bool need_ref_pose = do_ref || do_reward || do_done;
    const bool need_char_q = do_obs || do_reward;
    const int tile_rows = (int)min((long long)TILE, n_rows - row0);
    // the tile's ref_* block of field f (see RefLayout): [TILE][w] floats
    auto refp = [&](int field_off, int w, int el) -> float* { return s_ref + field_off * TILE + el * w; };

    // ---- height units -------------------------------------------------------------------------------------------------
    // _refresh_obs_hfs (ig_parkour_env.py:518-533, mgdm_dm_util.py:128-145): a unit = one piece of one env's sample points, done
    // by one warp; units are handed out by a shared counter. The samples depend only on the root / heading uniforms of
    // phase A and go straight to global memory. (Handing units to warps that wait at the barriers behind phases A / B was
    // tried and measured 10 % slower: a unit once taken delays the barrier it was meant to fill.)
    auto height_units = [&]() {
        if (!(P > 0 && (do_heights || do_obs))) return;
        const int lane = tid & 31;
        HeightArgs ha;
        ha.hf = terrain.hf; ha.dimy = terrain.dim_y; ha.imx = terrain.dim_x - 1; ha.imy = terrain.dim_y - 1;
        ha.nmn = make_float2(-terrain.min_x, -terrain.min_y); ha.rd2 = make_float2(sp.rdx, sp.rdy); ha.nd2 = make_float2(-terrain.dx, -terrain.dy);
        ha.lo = cfg.min_obs_h; ha.hi = cfg.max_obs_h;
        // an env's sample points in PARC_HEIGHT_SPLIT pieces, so that the warps finish together
        constexpr int HS = PARC_HEIGHT_SPLIT;
        const bool split = LATTICE ? (lat_nx % (HS * PARC_LATTICE_ROWS) == 0) : (P % (HS * 32 * PARC_HEIGHT_GATHERS) == 0);
        const int n_piece = split ? HS : 1;
        const int ppp = P / n_piece;                 // points per piece
        for (;;) {
            int unit = 0x7fffffff;
            if (lane == 0) unit = atomicAdd(s_ctr, 1);
            unit = __shfl_sync(FULL_MASK, unit, 0);
            if (unit >= tile_rows * n_piece) break;
            const int el = unit / n_piece, p0 = (unit - el * n_piece) * ppp;
            const EnvU& u = s_env[el];
            float* orow = do_obs ? env.obs_buf + u.e * env.obs_row_stride + L.o_hf + p0 : nullptr;
            float* hrow = env.ray_hfs ? env.ray_hfs + u.e * P + p0 : nullptr;
            if (do_heights) {
                ha.g2 = make_float2(u.g.x, u.g.y); ha.gz = u.g.z;
                if (LATTICE) {
                    // (xs[i] c, xs[i] s): the two products of rotate_2d_vec that a whole lattice row shares
                    const int nxp = lat_nx / n_piece, i0 = (unit - el * n_piece) * nxp;
                    float2* a = s_lat + (tid >> 5) * lat_nx;
                    __syncwarp();
                    for (int i = lane; i < lat_nx; i += 32) a[i] = make_float2(__fmul_rn(s_axes[i], u.ch), __fmul_rn(s_axes[i], u.sh));
                    __syncwarp();
                    const float y = s_axes[lat_nx + lane];
                    bool done = false;
                    if (PARC_HEIGHT_PATCH) {
                        float* patch = s_patch + (tid >> 5) * PATCH_FLOATS;
                        if (orow && hrow) done = heights_of_env_patch<true, true>(a, y, i0, i0 + nxp, lat_nx, lane, ha, u.ch, u.sh, hrow, orow, patch);
                        else if (orow) done = heights_of_env_patch<false, true>(a, y, i0, i0 + nxp, lat_nx, lane, ha, u.ch, u.sh, hrow, orow, patch);
                        else done = heights_of_env_patch<true, false>(a, y, i0, i0 + nxp, lat_nx, lane, ha, u.ch, u.sh, hrow, orow, patch);
                    }
                    if (!done) {
                        if (orow && hrow) heights_of_env_lattice<true, true>(a + i0, y, nxp, lane, ha, u.ch, u.sh, hrow, orow);
                        else if (orow) heights_of_env_lattice<false, true>(a + i0, y, nxp, lane, ha, u.ch, u.sh, hrow, orow);
                        else heights_of_env_lattice<true, false>(a + i0, y, nxp, lane, ha, u.ch, u.sh, hrow, orow);
                    }
                } else {
                    const float2 cs = make_float2(u.ch, u.sh), nsc = make_float2(-u.sh, u.ch);
                    if (ppp % (32 * PARC_HEIGHT_GATHERS) == 0) heights_of_env<true>(s_pts + p0, ppp, lane, ha, cs, nsc, hrow, orow);
                    else heights_of_env<false>(s_pts + p0, ppp, lane, ha, cs, nsc, hrow, orow);
                }
            } else if (orow) {
                for (int p = lane; p < ppp; p += 32) orow[p] = hrow[p];
            }
        }
    };

    if (PARC_STAGGER_NS > 0 && blockIdx.x >= kNumSMs && blockIdx.x < 2 * kNumSMs) __nanosleep(PARC_STAGGER_NS);
    // ================================ phase A ====================================================
    PARC_STAMP(0);
    if (PARC_PHASE_A_WIDE) {
        if (tid < NPART * PARC_MAX_BODIES / 4) reinterpret_cast<int4*>(s_sched)[tid] = reinterpret_cast<const int4*>(sp.sched)[tid];
        else if (tid >= 32 && tid < 32 + 2 * J) reinterpret_cast<float4*>(s_jc)[tid - 32] = reinterpret_cast<const float4*>(sp.jc)[tid - 32];
    } else {
        PARC_COLD_LOOP
        for (int k = tid; k < NPART * PARC_MAX_BODIES; k += THREADS) s_sched[k] = sp.sched[k];
        PARC_COLD_LOOP
        for (int k = tid; k < 8 * J; k += THREADS) {
            const int j = k >> 3, c = k & 7;
            s_jc[k] = c < 3 ? cm.local_trans[j][c] : (c < 7 ? cm.local_rot[j][c - 3] : 0.f);
        }
    }
    if (tid == 0) *s_ctr = 0;
    PARC_STAMP(13);
    if (do_heights) {
        if (LATTICE) {
            // xs[i] = x of point i * ny, ys[j] = y of point j
            PARC_COLD_LOOP
            for (int k = tid; k < lat_nx + lat_ny; k += THREADS)
                s_axes[k] = k < lat_nx ? __ldg(env.ray_xy_points + 2 * (long long)k * lat_ny) : __ldg(env.ray_xy_points + 2 * (k - lat_nx) + 1);
        } else {
            PARC_COLD_LOOP
            for (int k = tid; k < P; k += THREADS) s_pts[k] = __ldg(reinterpret_cast<const float2*>(env.ray_xy_points) + k);
        }
    }
    // A.3 per-env uniforms as two chains in two otherwise idle warps: root state -> heading -> sin/cos/h^-1 (last warp), and
    //     time -> motion id -> clip meta -> terrain offset (the warp before it); each writes its own fields of s_env[el].
    //     Rows past the end of a partial tile are filled with the tile's first row: every lane of phases B / C then works on
    //     valid data, and only stores to global memory look at `live`.
    if (tid >= THREADS - 32) {
      if (tid < THREADS - 32 + TILE) {
        const int el = tid - (THREADS - 32);
        EnvU& u = s_env[el];
        const int live = el < tile_rows;
        const long long row = row0 + (live ? el : 0);
        const long long e = env_ids_ ? env_ids_[row] : (env_ids32_ ? (long long)env_ids32_[row] : row + first_env);
        u.live = live;
        u.e = e;
        const Vec3 c_pos = view_ld_vec3(env.char_root_pos, e);
        const Quat c_rot = view_ld_quat(env.char_root_rot, e);
        u.c_pos = c_pos; u.c_rot = c_rot;
        u.c_vel = view_ld_vec3(env.char_root_vel, e);
        u.c_avel = view_ld_vec3(env.char_root_ang_vel, e);
        const Vec3 eo = make_vec3(__ldg(env.env_offsets + e * 3), __ldg(env.env_offsets + e * 3 + 1), __ldg(env.env_offsets + e * 3 + 2));
        u.eo = eo;
        u.g = c_pos + eo;                                              // ig_parkour_env.py:522
        const float heading = calc_heading(c_rot);
        u.sh = sinf(heading);
        if (!PARC_PHASE_A_WIDE) {
            u.ch = cosf(heading);
            u.hinv = axis_angle_to_quat(make_vec3(0.f, 0.f, 1.f), -heading);   // calc_heading_quat_inv
        }
        u.flags = 0;
        u.row_shift = do_obs ? (int)(((reinterpret_cast<unsigned long long>(env.obs_buf) >> 2) + (unsigned long long)(e * env.obs_row_stride)) & 3ull) : 0;
        if (need_char_q) s_charq[el * J] = quat_to_f4(c_rot);
        PARC_STAMP_T(14, THREADS - 32);
      }
    } else if (PARC_PHASE_A_WIDE && tid >= THREADS - 128 && tid < THREADS - 64) {
      // the other two ends of the root chain, each in its own warp (the heading is recomputed: the same operations on the same
      // quaternion): warp THREADS / 32 - 4 -> cos, warp THREADS / 32 - 3 -> inverse heading quaternion
      const int el = tid & 31;
      if (el < TILE) {
        const long long row = row0 + (el < tile_rows ? el : 0);
        const long long e = env_ids_ ? env_ids_[row] : (env_ids32_ ? (long long)env_ids32_[row] : row + first_env);
        const float heading = calc_heading(view_ld_quat(env.char_root_rot, e));
        if (tid < THREADS - 96) s_env[el].ch = cosf(heading);
        else s_env[el].hinv = axis_angle_to_quat(make_vec3(0.f, 0.f, 1.f), -heading);   // calc_heading_quat_inv
      }
    } else if (tid >= THREADS - 64) {
      if (tid < THREADS - 64 + TILE) {
        const int el = tid - (THREADS - 64);
        EnvU& u = s_env[el];
        const int live = el < tile_rows;
        const long long row = row0 + (live ? el : 0);
        const long long e = env_ids_ ? env_ids_[row] : (env_ids32_ ? (long long)env_ids32_[row] : row + first_env);
        float time;
        if (do_time) {                                   // _update_time, ig_env.py:391-394
            const int ts = env.timestep_buf[e] + 1;
            time = __fmul_rn(cfg.timestep, (float)ts);
            if (live) { env.timestep_buf[e] = ts; env.time_buf[e] = time; }
        } else {
            time = env.time_buf[e];
        }
        u.time = time;
        const long long mid = env.motion_ids[e];
        const long long tidx = env.motion_terrain_ids[e];
        u.mm = load_motion_meta(lib.meta, mid);
        u.t0 = __fadd_rn(time, env.motion_time_offsets[e]);            // dm_env.py:547-552
        const float* dmo = env.dm_motion_offsets + (mid * env.terrains_per_motion + tidx) * 2;   // dm_env.py:554-565
        u.tox = __fsub_rn(__ldg(dmo), __ldg(env.env_offsets + e * 3)); u.toy = __fsub_rn(__ldg(dmo + 1), __ldg(env.env_offsets + e * 3 + 1));
      }
    }
    if (PARC_EARLY_ROW_PREFETCH && (do_reward || do_done) && tid >= THREADS / 2 && tid < THREADS / 2 + TILE * 3 * 4) {
        // simulator state read in phase C (rigid body positions, contact forces, dof velocities): 3 rows per env, asked for by
        // threads that have nothing else to do in phase A
        const int it = tid - THREADS / 2;
        const int el = it / 12, w = (it >> 2) % 3, ln = it & 3;
        if (el < tile_rows) {
            const long long row = row0 + el;
            const long long e = env_ids_ ? env_ids_[row] : (env_ids32_ ? (long long)env_ids32_[row] : row + first_env);
            const parc_view3_t& v = w == 0 ? env.char_rigid_body_pos : (w == 1 ? env.char_contact_forces : env.char_dof_vel);
            const long long span = (w == 2 ? (long long)D * v.elem_stride : (long long)J * v.mid_stride) * (long long)sizeof(float);
            if (v.ptr && ln * 128 < span) asm volatile("prefetch.global.L2 [%0];" ::"l"(reinterpret_cast<const char*>(v.ptr + e * v.row_stride) + ln * 128));
        }
    }
    // A.1 character joint rotations: KinCharModel.dof_to_rot (kin_char_model.py:586-599), one thread per (env, joint)
    if (need_char_q) {
        PARC_COLD_LOOP
        for (int it = tid; it < TILE * (J - 1); it += THREADS) {
            const int el = it / (J - 1), j = 1 + (it - el * (J - 1));
            const long long row = row0 + (el < tile_rows ? el : 0);
            const long long e = env_ids_ ? env_ids_[row] : (env_ids32_ ? (long long)env_ids32_[row] : row + first_env);
            int jt, di; Vec3 ax;
            if (PARC_PHASE_A_WIDE) {
                const int4 d = reinterpret_cast<const int4*>(sp.jd)[j];
                jt = d.x & 0xff; di = d.x >> 8; ax = make_vec3(__int_as_float(d.y), __int_as_float(d.z), __int_as_float(d.w));
            } else {
                jt = cm.joint_type[j]; di = cm.dof_idx[j]; ax = make_vec3(cm.axis[j][0], cm.axis[j][1], cm.axis[j][2]);
            }
            Quat q = quat_identity();
            if (jt == PARC_JOINT_SPHERICAL)
                q = exp_map_to_quat(make_vec3(view_ld(env.char_dof_pos, e, di), view_ld(env.char_dof_pos, e, di + 1), view_ld(env.char_dof_pos, e, di + 2)));
            else if (jt == PARC_JOINT_HINGE)
                q = axis_angle_to_quat(ax, view_ld(env.char_dof_pos, e, di));
            s_charq[el * J + j] = quat_to_f4(q);
        }
    }
    __syncthreads();
    PARC_STAMP(1);

    // A.0b warm the frame records this tile is about to interpolate: one thread per (env, library pose) asks for
    //      the two adjacent interp records (and the velocity record of the reference pose)
    if (!PARC_EARLY_ROW_PREFETCH && (do_reward || do_done)) {
        // simulator state read in phase C (rigid body positions, contact forces, dof velocities): 3 rows per env
        PARC_COLD_LOOP
        for (int it = tid; it < TILE * 3 * 4; it += THREADS) {
            const int el = it / 12, w = (it >> 2) % 3, ln = it & 3;
            if (el >= tile_rows) continue;
            const parc_view3_t& v = w == 0 ? env.char_rigid_body_pos : (w == 1 ? env.char_contact_forces : env.char_dof_vel);
            const long long span = (w == 2 ? (long long)D * v.elem_stride : (long long)J * v.mid_stride) * (long long)sizeof(float);
            if (v.ptr && ln * 128 < span)
                asm volatile("prefetch.global.L2 [%0];" ::"l"(reinterpret_cast<const char*>(v.ptr + s_env[el].e * v.row_stride) + ln * 128));
        }
    }
    if (PARC_PREFETCH_FRAMES && (do_obs || need_ref_pose)) {
        PARC_COLD_LOOP
        for (int it = tid; it < TILE * (S + 1); it += THREADS) {
            const int el = it / (S + 1), pose = it - el * (S + 1);
            if (el >= tile_rows) continue;
            const EnvU& u = s_env[el];
            const float t = pose == 0 ? u.t0 : __fadd_rn(u.t0, cfg.tar_dt[pose - 1]);
            const FrameBlend fb = calc_frame_blend(u.mm, t);
            const char* r0 = reinterpret_cast<const char*>(lib.interp + fb.idx0 * irec_f);
            const int bytes = (int)((fb.idx1 - fb.idx0 + 1) * irec_f * sizeof(float));
            PARC_COLD_LOOP
            for (int off = 0; off < bytes; off += 128) asm volatile("prefetch.global.L1 [%0];" ::"l"(r0 + off));
            if (pose == 0) {
                const char* v = reinterpret_cast<const char*>(lib.vel + fb.idx0 * vrec_f);
                PARC_COLD_LOOP
                for (int off = 0; off < (int)(vrec_f * sizeof(float)); off += 128) asm volatile("prefetch.global.L1 [%0];" ::"l"(v + off));
            }
        }
    }

    // ================================ phase B ====================================================
    PARC_STAMP(2);
    static_assert(TILE % 8 == 0 && NPART == 4, "one warp = one pose of 8 of the tile's envs x NPART lanes");
    const bool pose_work = do_obs || need_ref_pose;
    const bool ref_tail_by_char = PARC_REF_TAIL_BY_CHAR && PARC_POSE_TAIL_COPIES && do_obs && need_ref_pose;
    // warp w takes poses w, w + 8, ...; lane = (env, part). Rows past a partial tile hold a copy of the tile's first row
    // (phase A), so no lane leaves the loop early: the __syncwarp() of the joint schedule sees the whole warp.
    for (int pose = (tid >> 5) & 7; pose_work && pose < plan.n_pose; pose += 8) {
        const int lane = tid & 31, el = (tid >> 8) * 8 + lane / NPART, part = lane & (NPART - 1);
        const bool is_ref = pose == 0, is_char = pose == S + 1, is_tar = pose >= 1 && pose <= S;
        const bool wanted = (is_ref && need_ref_pose) || (is_tar && do_obs) || (is_char && do_obs);
        if (!wanted) continue;
        const EnvU& u = s_env[el];
        float* row = s_row + el * ROWP + u.row_shift;
        const float *r0 = nullptr, *r1 = nullptr, *pr0 = nullptr;
        float b = 0.f, omb = 1.f;
        Quat rq = u.c_rot;
        Vec3 rpos = u.c_pos;
        const float* vrec = nullptr;
        if (!is_char) {
            // MotionLib.calc_motion_frame (motion_lib.py:94-126); fetch_tar_obs_data (mgdm_dm_util.py:221-244)
            const float t = is_ref ? u.t0 : __fadd_rn(u.t0, cfg.tar_dt[pose - 1]);
            const FrameBlend fb = calc_frame_blend(u.mm, t);
            r0 = lib.interp + fb.idx0 * irec_f;
            r1 = lib.interp + fb.idx1 * irec_f;
            pr0 = r0 + interp_pair_off(J, C);
            b = fb.blend; omb = __fsub_rn(1.0f, b);
            rq = slerp_pre(quat_from_f4(ldg4(r0)), quat_from_f4(ldg4(r1)), ldg2(pr0), b, omb);
            const float4 p0 = ldg4(r0 + 4 * J), p1 = ldg4(r1 + 4 * J);
            rpos = make_vec3(__fadd_rn(__fadd_rn(lerp_exact(omb, p0.x, b, p1.x), fb.loop_offset.x), u.tox),
                             __fadd_rn(__fadd_rn(lerp_exact(omb, p0.y, b, p1.y), fb.loop_offset.y), u.toy),
                             __fadd_rn(lerp_exact(omb, p0.z, b, p1.z), fb.loop_offset.z));
            vrec = lib.vel + fb.idx0 * vrec_f;
        }
        // ---- root-level outputs (scalar pieces by part 0, the per-dof / per-body lists strided over the parts) ----
        float* o_pose = nullptr;        // start of this pose's joint tan-norm block
        float* o_key = nullptr;         // start of this pose's key-body block
        Vec3 rp_obs = make_vec3(0.f, 0.f, 0.f);
        if (is_ref) {
            if (part == 0) {
                float* o = refp(R.root_pos, 3, el); o[0] = rpos.x; o[1] = rpos.y; o[2] = rpos.z;
                *reinterpret_cast<float4*>(refp(R.root_rot, 4, el)) = quat_to_f4(rq);
                o = refp(R.body_pos, 3 * J, el); o[0] = rpos.x; o[1] = rpos.y; o[2] = rpos.z;
            }
            // velocities come from frame idx0 without interpolation (motion_lib.py:102-110)
            if (PARC_POSE_TAIL_COPIES) {
                if (!ref_tail_by_char && part * 32 < 6 + D) asm volatile("prefetch.global.L1 [%0];" ::"l"(vrec + part * 32));
            } else {
                float* o_rv = refp(R.root_vel, 3, el); float* o_rav = refp(R.root_ang_vel, 3, el); float* o_dv = refp(R.dof_vel, D, el);
                PARC_COLD_LOOP
                for (int k = 4 * part; k < 6 + D; k += 4 * NPART) {
                    const float4 x = ldg4(vrec + k);
                    const float xs[4] = {x.x, x.y, x.z, x.w};
#pragma unroll
                    for (int c = 0; c < 4; ++c) {
                        const int kk = k + c;
                        if (kk < 3) o_rv[kk] = xs[c];
                        else if (kk < 6) o_rav[kk - 3] = xs[c];
                        else if (kk < 6 + D) o_dv[kk - 6] = xs[c];
                    }
                }
            }
        } else if (is_tar) {
            // compute_tar_obs (mgdm_dm_util.py:404-461) for target step pose-1
            float* o = row + L.o_tar + (pose - 1) * L.tar_step_dim;
            rp_obs = rpos - u.c_pos;
            Quat oq = rq;
            if (!c_global_obs) { rp_obs = quat_rotate(u.hinv, rp_obs); oq = quat_mul(u.hinv, rq); }
            if (part == 0) {
                o[0] = rp_obs.x; o[1] = rp_obs.y; o[2] = rp_obs.z;
                tan_norm_exact(oq, o + 3);
            }
            o_pose = o + 9; o_key = o + 9 + 6 * (J - 1);
        } else {
            // compute_char_obs (ig_char_env.py:581-625)
            float* o = row + L.root_h;
            if (part == 0) {
                if (L.root_h) row[0] = u.c_pos.z;
                Quat oq = u.c_rot; Vec3 v = u.c_vel, w = u.c_avel;
                if (!c_global_obs) { oq = quat_mul(u.hinv, u.c_rot); v = quat_rotate(u.hinv, u.c_vel); w = quat_rotate(u.hinv, u.c_avel); }
                tan_norm_exact(oq, o);
                o[6] = v.x; o[7] = v.y; o[8] = v.z; o[9] = w.x; o[10] = w.y; o[11] = w.z;
            }
            o_pose = o + 12; o_key = o + 12 + 6 * (J - 1) + D;
            if (PARC_POSE_TAIL_COPIES) {
                // the rows read after the joint loop: 128-byte pieces of the dof-velocity span and of the contact-force row
                const float* dv = env.char_dof_vel.ptr + u.e * env.char_dof_vel.row_stride;
                if ((long long)part * 32 < (long long)D * env.char_dof_vel.elem_stride) asm volatile("prefetch.global.L1 [%0];" ::"l"(dv + part * 32));
                if (c_contact_info && part * 32 < 3 * J)
                    asm volatile("prefetch.global.L1 [%0];" ::"l"(env.char_contact_forces.ptr + u.e * env.char_contact_forces.row_stride + part * 32));
                if (ref_tail_by_char) {
                    // the reference pose's velocity record, copied by this warp after its joint loop
                    const FrameBlend fb = calc_frame_blend(u.mm, u.t0);
                    if (part * 32 < 6 + D) asm volatile("prefetch.global.L1 [%0];" ::"l"(lib.vel + fb.idx0 * vrec_f + part * 32));
                }
            } else {
                PARC_COLD_LOOP
                for (int k = part; k < D; k += NPART) o[12 + 6 * (J - 1) + k] = view_ld(env.char_dof_vel, u.e, k);
                if (c_contact_info) {
                    // _get_char_contact_state (ig_parkour_env.py:655-662): ||f|| > eps -> 0/1
                    PARC_COLD_LOOP
                    for (int c = part; c < J; c += NPART) {
                        const float fx = view_ld3(env.char_contact_forces, u.e, c, 0), fy = view_ld3(env.char_contact_forces, u.e, c, 1),
                                    fz = view_ld3(env.char_contact_forces, u.e, c, 2);
                        row[L.o_charc + c] = (sqrtf(fx * fx + fy * fy + fz * fz) > cfg.contact_eps) ? 1.f : 0.f;
                    }
                }
            }
        }
        // contact labels: lerp of the two records (motion_lib.py:121)
        if (!is_char && C > 0 && (is_ref ? !ref_tail_by_char : c_contact_info)) {
            float* oc = is_ref ? refp(R.contacts, C, el) : (row + L.o_tarc + (pose - 1) * C);
            const float* c0 = r0 + 4 * J + 4; const float* c1 = r1 + 4 * J + 4;
            PARC_COLD_LOOP
            for (int c = 4 * part; c < C; c += 4 * NPART) {
                const float4 a = ldg4(c0 + c), d = ldg4(c1 + c);
                oc[c] = lerp_exact(omb, a.x, b, d.x);
                if (c + 1 < C) oc[c + 1] = lerp_exact(omb, a.y, b, d.y);
                if (c + 2 < C) oc[c + 2] = lerp_exact(omb, a.z, b, d.z);
                if (c + 3 < C) oc[c + 3] = lerp_exact(omb, a.w, b, d.w);
            }
        }
        // a key body may also be the root itself
        if (!is_ref && part == 0 && (sp.key_mask & 1u))
            for (int k = 0; k < K; ++k)
                if (cfg.key_body_ids[k] == 0) {
                    Vec3 kp = make_vec3(0.f, 0.f, 0.f);
                    if (!c_global_obs && is_tar) kp = kp + rp_obs;
                    float* kd = o_key + 3 * k;
                    kd[0] = kp.x; kd[1] = kp.y; kd[2] = kp.z;
                }
        // ---- joints: slerp + forward kinematics (kin_char_model.py:617-649) along the schedule ----
        const bool lrot_ident = lrot_ident_;
        float* gx = s_gx + (tid >> 2);                  // [slot][7][NGROUP], this 4-lane group's column
        Vec3 cpos = rpos; Quat crot = rq;               // transform of the body this lane evaluated last (the root at first)
        if (root_slot >= 0 && part == 0) {
            float* x = gx + root_slot * 7 * NGROUP;
            x[0] = rpos.x; x[NGROUP] = rpos.y; x[2 * NGROUP] = rpos.z;
            x[3 * NGROUP] = rq.x; x[4 * NGROUP] = rq.y; x[5 * NGROUP] = rq.z; x[6 * NGROUP] = rq.w;
        }
        float* o_jr = is_ref ? refp(R.joint_rot, 4 * (J - 1), el) : nullptr;
        float* o_bp = is_ref ? refp(R.body_pos, 3 * J, el) : nullptr;
        __syncwarp();
        auto joint_loop = [&](auto fast_tag) {
        constexpr bool FASTM = decltype(fast_tag)::value;
        PARC_COLD_LOOP
        for (int it = 0; it < n_iter; ++it) {
            const int entry = s_sched[it * NPART + part];
#if PARC_PREFETCH_NEXT_JOINT
            // the next iteration's quaternions and pair constants: asked for now, they arrive while this joint is evaluated
            if (!is_char && it + 1 < n_iter) {
                const int en = s_sched[(it + 1) * NPART + part];
                if (en >= 0) {
                    const int jn = en & 0xff;
                    asm volatile("prefetch.global.L1 [%0];" ::"l"(r0 + 4 * jn));
                    asm volatile("prefetch.global.L1 [%0];" ::"l"(r1 + 4 * jn));
                    asm volatile("prefetch.global.L1 [%0];" ::"l"(pr0 + 2 * jn));
                }
            }
#endif
            if (entry >= 0) {
                const int j = entry & 0xff;
                Quat q;
                if (is_char) q = quat_from_f4(s_charq[el * J + j]);
                else q = slerp_pre(quat_from_f4(ldg4(r0 + 4 * j)), quat_from_f4(ldg4(r1 + 4 * j)), ldg2(pr0 + 2 * j), b, omb);
                if (entry & SCHED_FROM_SLOT) {
                    const float* x = gx + ((entry >> SCHED_RSLOT_SHIFT) & 31) * 7 * NGROUP;
                    cpos = make_vec3(x[0], x[NGROUP], x[2 * NGROUP]);
                    crot = make_quat(x[3 * NGROUP], x[4 * NGROUP], x[5 * NGROUP], x[6 * NGROUP]);
                }
                const float* jc = s_jc + 8 * j;
                const Vec3 bpos = cpos + (FASTM ? quat_rotate_fast(crot, make_vec3(jc[0], jc[1], jc[2])) : quat_rotate(crot, make_vec3(jc[0], jc[1], jc[2])));
                cpos = bpos;
                if (entry & SCHED_NEEDS_ROT) {
                    const Quat lq = make_quat(jc[3], jc[4], jc[5], jc[6]);
                    const Quat cq = lrot_ident ? q : (FASTM ? quat_mul_fast(lq, q) : quat_mul(lq, q));
                    crot = FASTM ? quat_mul_fast(crot, cq) : quat_mul(crot, cq);
                    if (entry & SCHED_PUBLISH) {
                        float* x = gx + ((entry >> SCHED_WSLOT_SHIFT) & 31) * 7 * NGROUP;
                        x[0] = bpos.x; x[NGROUP] = bpos.y; x[2 * NGROUP] = bpos.z;
                        x[3 * NGROUP] = crot.x; x[4 * NGROUP] = crot.y; x[5 * NGROUP] = crot.z; x[6 * NGROUP] = crot.w;
                    }
                }
                if (is_ref) {
                    *reinterpret_cast<float4*>(o_jr + 4 * (j - 1)) = quat_to_f4(q);
                    float* bp = o_bp + 3 * j;
                    bp[0] = bpos.x; bp[1] = bpos.y; bp[2] = bpos.z;
                } else {
                    if (FASTM) tan_norm_fast(q, o_pose + 6 * (j - 1));
                    else tan_norm_exact(q, o_pose + 6 * (j - 1));
                    PARC_COLD_LOOP
                    for (unsigned km = (entry >> 8) & 0xffu; km; km &= km - 1) {
                        Vec3 kp = bpos - rpos;
                        if (!c_global_obs) { kp = FASTM ? quat_rotate_fast(u.hinv, kp) : quat_rotate(u.hinv, kp); if (is_tar) kp = kp + rp_obs; }
                        float* kd = o_key + 3 * (__ffs(km) - 1);
                        kd[0] = kp.x; kd[1] = kp.y; kd[2] = kp.z;
                    }
                }
            }
            __syncwarp();
        }
        };
        if (PARC_TAR_FAST_MATH && is_tar) joint_loop(std::true_type{});
        else joint_loop(std::false_type{});
        if (PARC_POSE_TAIL_COPIES && (is_ref ? !ref_tail_by_char : (is_char && ref_tail_by_char))) {
            if (is_char) {
                // the reference pose's frame pair again (same operations on the same uniforms as in that pose's warp)
                const FrameBlend fb = calc_frame_blend(u.mm, u.t0);
                vrec = lib.vel + fb.idx0 * vrec_f;
                if (C > 0) {
                    // contact labels of the reference pose: lerp of the two records (motion_lib.py:121)
                    const float* c0 = lib.interp + fb.idx0 * irec_f + 4 * J + 4; const float* c1 = lib.interp + fb.idx1 * irec_f + 4 * J + 4;
                    const float bb = fb.blend, ob = __fsub_rn(1.0f, bb);
                    float* oc = refp(R.contacts, C, el);
                    PARC_COLD_LOOP
                    for (int c = 4 * part; c < C; c += 4 * NPART) {
                        const float4 a = ldg4(c0 + c), d = ldg4(c1 + c);
                        oc[c] = lerp_exact(ob, a.x, bb, d.x);
                        if (c + 1 < C) oc[c + 1] = lerp_exact(ob, a.y, bb, d.y);
                        if (c + 2 < C) oc[c + 2] = lerp_exact(ob, a.z, bb, d.z);
                        if (c + 3 < C) oc[c + 3] = lerp_exact(ob, a.w, bb, d.w);
                    }
                }
            }
            // velocity record -> ref_root_vel | ref_root_ang_vel | ref_dof_vel: all of this lane's loads first, then the scatter
            float* o_rv = refp(R.root_vel, 3, el); float* o_rav = refp(R.root_ang_vel, 3, el); float* o_dv = refp(R.dof_vel, D, el);
            constexpr int NX = FAST ? 3 : 2;                                       // pieces of 4 floats per lane and batch (humanoid: 34 floats = 3 / 2 / 2 / 2)
            for (int k0 = 4 * part; k0 < 6 + D; k0 += 4 * NPART * NX) {
                float4 x[NX];
#pragma unroll
                for (int i = 0; i < NX; ++i) if (k0 + 4 * NPART * i < 6 + D) x[i] = ldg4(vrec + k0 + 4 * NPART * i);
#pragma unroll
                for (int i = 0; i < NX; ++i) {
                    if (k0 + 4 * NPART * i < 6 + D) {
                        const float xs[4] = {x[i].x, x[i].y, x[i].z, x[i].w};
#pragma unroll
                        for (int c = 0; c < 4; ++c) {
                            const int kk = k0 + 4 * NPART * i + c;
                            if (kk < 3) o_rv[kk] = xs[c];
                            else if (kk < 6) o_rav[kk - 3] = xs[c];
                            else if (kk < 6 + D) o_dv[kk - 6] = xs[c];
                        }
                    }
                }
            }
        }
        if (PARC_POSE_TAIL_COPIES && is_char) {
            if ((PARC_REF_TAIL_BY_CHAR || PARC_EARLY_ROW_PREFETCH) && (do_reward || do_done) && env.char_rigid_body_pos.ptr && (long long)part * 32 < (long long)J * env.char_rigid_body_pos.mid_stride)
                asm volatile("prefetch.global.L1 [%0];" ::"l"(env.char_rigid_body_pos.ptr + u.e * env.char_rigid_body_pos.row_stride + part * 32));   // read in phase C
            float* o = row + L.root_h + 12 + 6 * (J - 1);
            constexpr int ND = FAST ? (FastDims::D + NPART - 1) / NPART : 2;       // dof velocities of this lane per batch
            for (int k0 = part; k0 < D; k0 += NPART * ND) {
                float v[ND];
#pragma unroll
                for (int i = 0; i < ND; ++i) if (k0 + NPART * i < D) v[i] = view_ld(env.char_dof_vel, u.e, k0 + NPART * i);
#pragma unroll
                for (int i = 0; i < ND; ++i) if (k0 + NPART * i < D) o[k0 + NPART * i] = v[i];
            }
            if (c_contact_info) {
                // _get_char_contact_state (ig_parkour_env.py:655-662): ||f|| > eps -> 0/1
                constexpr int NC = FAST ? (FastDims::J + NPART - 1) / NPART : 1;
                for (int c0 = part; c0 < J; c0 += NPART * NC) {
                    float f[NC][3];
#pragma unroll
                    for (int i = 0; i < NC; ++i)
                        if (c0 + NPART * i < J) {
#pragma unroll
                            for (int a = 0; a < 3; ++a) f[i][a] = view_ld3(env.char_contact_forces, u.e, c0 + NPART * i, a);
                        }
#pragma unroll
                    for (int i = 0; i < NC; ++i)
                        if (c0 + NPART * i < J)
                            row[L.o_charc + c0 + NPART * i] = (sqrtf(f[i][0] * f[i][0] + f[i][1] * f[i][1] + f[i][2] * f[i][2]) > cfg.contact_eps) ? 1.f : 0.f;
                }
            }
        }
    }
    PARC_STAMP_T(9, 0); PARC_STAMP_T(10, 32); PARC_STAMP_T(11, 6 * 32); PARC_STAMP_T(12, 7 * 32);     // poses 0, 1, S, S + 1 of the first env group
    __syncthreads();
    PARC_STAMP(3);

    // ================================ phase C ====================================================
    // per-(env, item) terms; item lists are flattened over the tile so that every lane has work
    const int o_pose_t = 0, o_vel_t = J - 1, o_con_t = o_vel_t + D, o_key_t = o_con_t + J;
    const int TP = plan.term_pitch;
    // The full-tile fast path of phase E: rows row0 .. row0 + TILE - 1 are consecutive env ids, so the staged ref_* blocks
    // are the memory image of those rows of every (contiguous) ref_* tensor.
    const long long e_first = row0 + first_env;
    const bool ref_bulk = PARC_BULK_ROWS && do_ref && (FAST || sp.ref_fast) && tile_rows == TILE && env_ids_ == nullptr && env_ids32_ == nullptr &&
                          (e_first & 3) == 0;
    const bool obs_bulk = PARC_BULK_OBS && do_obs;
    // C.0 scatter map of the ref_* values for the general copy-out of phase E: column k of an env's ref_* values ->
    //     (global base pointer, row stride in bytes, position in the tile's staged block). Lives in the bytes the group
    //     slots used during phase B; phase E reads it behind the barrier that ends phase C.
    if (do_ref && !ref_bulk) {
        PARC_COLD_LOOP
        for (int k = tid; k < R.width; k += THREADS) {
            const parc_view3_t* v; int off, fo, w;
            if (k < R.root_rot) { v = &env.ref_root_pos; fo = R.root_pos; w = 3; }
            else if (k < R.root_vel) { v = &env.ref_root_rot; fo = R.root_rot; w = 4; }
            else if (k < R.root_ang_vel) { v = &env.ref_root_vel; fo = R.root_vel; w = 3; }
            else if (k < R.body_pos) { v = &env.ref_root_ang_vel; fo = R.root_ang_vel; w = 3; }
            else if (k < R.joint_rot) { v = &env.ref_body_pos; fo = R.body_pos; w = 3 * J; }
            else if (k < R.dof_pos) { v = &env.ref_joint_rot; fo = R.joint_rot; w = 4 * (J - 1); }
            else if (k < R.dof_vel) { v = &env.ref_dof_pos; fo = R.dof_pos; w = D; }
            else if (k < R.contacts) { v = &env.ref_dof_vel; fo = R.dof_vel; w = D; }
            else { v = &env.ref_contacts; fo = R.contacts; w = C; }
            off = k - fo;
            long long eoff;
            if (k >= R.body_pos && k < R.joint_rot) eoff = (off / 3) * v->mid_stride + (off % 3) * v->elem_stride;
            else if (k >= R.joint_rot && k < R.dof_pos) eoff = (off / 4) * v->mid_stride + (off % 4) * v->elem_stride;
            else eoff = off * v->elem_stride;
            const unsigned long long base = v->ptr ? reinterpret_cast<unsigned long long>(v->ptr + eoff) : 0ull;
            s_refmap[k] = make_uint4((unsigned)base, (unsigned)(base >> 32), (unsigned)(v->row_stride * (long long)sizeof(float)),
                                     (unsigned)(fo * TILE + off) | ((unsigned)w << 20));
        }
    }
    {
        // Four item kinds (joints, contacts, key bodies, dof velocities) cut into warp-sized tasks that are dealt
        // round-robin to the warps: every warp runs about the same number of (prefetched) global reads.
        const bool want_c = (do_reward && c_contact_info) || (do_done && c_early_term && c_n_fall > 0);
        const int nt_j1 = need_ref_pose ? (TILE * (J - 1) + 31) >> 5 : 0;      // tasks per joint sub-kind
        const int nt_j = PARC_SPLIT_JOINT_TASKS ? 3 * nt_j1 : nt_j1;
        const int nt_c = want_c ? (TILE * J + 31) >> 5 : 0;
        const int nt_k = do_reward ? (TILE * K + 31) >> 5 : 0;
        const int nt_d = do_reward ? (TILE * D + 31) >> 5 : 0;
        const int b_c = nt_j, b_k = b_c + nt_c, b_d = b_k + nt_k, n_task = b_d + nt_d;
        const int lane = tid & 31;
        PARC_COLD_LOOP
        for (int task = tid >> 5; task < n_task; task += THREADS / 32) {
            if (task < b_c) {
                // joints: rot_to_dof of the reference pose (kin_char_model.py:601-615), pose error term, pose termination test
                // -- one task does all three (sub < 0), or they are three tasks of similar length (sub = 0, 1, 2)
                const int sub = PARC_SPLIT_JOINT_TASKS ? task / nt_j1 : -1;
                const int it = (PARC_SPLIT_JOINT_TASKS ? task - sub * nt_j1 : task) * 32 + lane;
                if (it >= TILE * (J - 1)) continue;
                const int el = it / (J - 1), j = 1 + (it - el * (J - 1));
                if (el >= tile_rows) continue;
                const Quat q = quat_from_f4(*reinterpret_cast<const float4*>(refp(R.joint_rot, 4 * (J - 1), el) + 4 * (j - 1)));
                if (do_ref && sub <= 0) {
                    const int jt = cm.joint_type[j], di = cm.dof_idx[j];
                    float* dp = refp(R.dof_pos, D, el) + di;
                    if (jt == PARC_JOINT_SPHERICAL) {
                        const Vec3 ex = quat_to_exp_map(q);
                        dp[0] = ex.x; dp[1] = ex.y; dp[2] = ex.z;
                    } else if (jt == PARC_JOINT_HINGE) {
                        Vec3 axis; float angle;
                        quat_to_axis_angle(q, axis, angle);
                        const float dd = dot3_seq(make_vec3(cm.axis[j][0], cm.axis[j][1], cm.axis[j][2]), axis);
                        dp[0] = (dd < 0.f) ? -angle : angle;
                    }
                }
                if (do_reward && (sub < 0 || sub == 1)) {   // compute_deepmimic_reward pose term (mgdm_dm_util.py:285-287)
                    const float a = quat_diff_angle(quat_from_f4(s_charq[el * J + j]), q);
                    s_terms[el * TP + o_pose_t + (j - 1)] = cfg.joint_err_w[j - 1] * a * a;
                }
                if (do_done && c_early_term && c_pose_term && (sub < 0 || sub == 2)) {
                    // compute_done pose test (mgdm_dm_util.py:372-384): squared norm summed left to right
                    const EnvU& u = s_env[el];
                    const Vec3 sb = make_vec3(view_ld3(env.char_rigid_body_pos, u.e, j, 0), view_ld3(env.char_rigid_body_pos, u.e, j, 1), view_ld3(env.char_rigid_body_pos, u.e, j, 2));
                    const Vec3 s0 = make_vec3(view_ld3(env.char_rigid_body_pos, u.e, 0, 0), view_ld3(env.char_rigid_body_pos, u.e, 0, 1), view_ld3(env.char_rigid_body_pos, u.e, 0, 2));
                    const float* bp = refp(R.body_pos, 3 * J, el);
                    const Vec3 a = sb - s0;
                    const Vec3 bb = make_vec3(bp[3 * j], bp[3 * j + 1], bp[3 * j + 2]) - make_vec3(bp[0], bp[1], bp[2]);
                    const Vec3 d = bb - a;
                    const float td = cfg.pose_termination_dist[j - 1];
                    if (sqnorm3_seq(d) > __fmul_rn(td, td)) atomicOr(&s_env[el].flags, 4);
                }
            } else if (task < b_k) {
                const int it = (task - b_c) * 32 + lane;
                if (it >= TILE * J) continue;
                const int el = it / J, c = it - el * J;
                if (el >= tile_rows) continue;
                const EnvU& u = s_env[el];
                const float fx = view_ld3(env.char_contact_forces, u.e, c, 0), fy = view_ld3(env.char_contact_forces, u.e, c, 1),
                            fz = view_ld3(env.char_contact_forces, u.e, c, 2);
                if (do_reward && c_contact_info) {
                    // compute_contact_reward (mgdm_dm_util.py:497-518); contact bodies == all bodies
                    const float f = fminf(sqrtf(fx * fx + fy * fy + fz * fz), 1.0f);
                    const float rc = refp(R.contacts, C, el)[c];
                    float cr = -(1.0f - rc) * f;
                    cr += rc * f;
                    s_terms[el * TP + o_con_t + c] = cfg.contact_weights[c] * cr;
                }
                if (do_done && c_early_term && c_n_fall > 0 && !((cfg.fall_contact_body_mask >> c) & 1u)) {
                    // fall test (mgdm_dm_util.py:147-150, 349-362)
                    int fl = 0;
                    if (fabsf(fx) > 0.1f || fabsf(fy) > 0.1f || fabsf(fz) > 0.1f) fl |= 1;
                    const float bx = view_ld3(env.char_rigid_body_pos, u.e, c, 0), by = view_ld3(env.char_rigid_body_pos, u.e, c, 1),
                                bz = view_ld3(env.char_rigid_body_pos, u.e, c, 2);
                    const int ix = grid_index(__fadd_rn(bx, u.eo.x), terrain.min_x, terrain.dx, terrain.dim_x);
                    const int iy = grid_index(__fadd_rn(by, u.eo.y), terrain.min_y, terrain.dy, terrain.dim_y);
                    const float th = __fadd_rn(__ldg(terrain.hf + (long long)ix * terrain.dim_y + iy), cfg.termination_height);
                    if (bz < th) fl |= 2;
                    if (fl) atomicOr(&s_env[el].flags, fl);
                }
            } else if (task < b_d) {
                // key body term (mgdm_dm_util.py:303-323); key positions of the CHARACTER come from the simulator's rigid bodies
                const int it = (task - b_k) * 32 + lane;
                if (it >= TILE * K) continue;
                const int el = it / K, k = it - el * K;
                if (el >= tile_rows) continue;
                const EnvU& u = s_env[el];
                const int body = cfg.key_body_ids[k];
                const float* bp = refp(R.body_pos, 3 * J, el);
                const Vec3 ref_rpos = make_vec3(bp[0], bp[1], bp[2]);
                Vec3 kp = make_vec3(view_ld3(env.char_rigid_body_pos, u.e, body, 0), view_ld3(env.char_rigid_body_pos, u.e, body, 1),
                                    view_ld3(env.char_rigid_body_pos, u.e, body, 2)) - u.c_pos;
                Vec3 tkp = make_vec3(bp[3 * body], bp[3 * body + 1], bp[3 * body + 2]) - ref_rpos;
                if (!c_track_root) {   // convert_to_local (mgdm_dm_util.py:246-267)
                    const Quat rr = quat_from_f4(*reinterpret_cast<const float4*>(refp(R.root_rot, 4, el)));
                    kp = quat_rotate(u.hinv, kp);
                    tkp = quat_rotate(calc_heading_quat_inv(rr), tkp);
                }
                const Vec3 d = tkp - kp;
                s_terms[el * TP + o_key_t + k] = d.x * d.x + d.y * d.y + d.z * d.z;
            } else {
                // dof velocity term (mgdm_dm_util.py:289-291)
                const int it = (task - b_d) * 32 + lane;
                if (it >= TILE * D) continue;
                const int el = it / D, k = it - el * D;
                if (el >= tile_rows) continue;
                const EnvU& u = s_env[el];
                const float dv = refp(R.dof_vel, D, el)[k] - view_ld(env.char_dof_vel, u.e, k);
                s_terms[el * TP + o_vel_t + k] = cfg.dof_err_w[k] * dv * dv;
            }
        }
    }
    if (obs_bulk || ref_bulk) fence_async_smem();   // the staged rows / blocks are about to be read by the bulk-copy engine
    __syncthreads();
    PARC_STAMP(4);

    // ================================ phase E (issue) =============================================
    // Everything staged is final: hand the aligned body of each observation row and the ref_* blocks to the bulk-copy
    // engine now, so that the copies run while phase D and the height samples execute. The few floats in front of /
    // behind a row's 16-byte aligned body are stored by the same lane.
    bool bulk_pending = false;
    {
        // Issuing lanes: the first tile_rows lanes of one warp hand over an observation row each, NREF further lanes of it
        // one ref_* tensor each. (UBLKCP takes uniform operands, so the lanes of a warp issue one after the other; one lane
        // per warp instead measured slower.) With PARC_TAIL_ROLES that warp is warp 1, so that it delays neither the reward
        // (warp 0) nor the termination lanes (warp 2).
        const int warp = tid >> 5, lane = tid & 31;
        constexpr int W_ISSUE = PARC_TAIL_ROLES ? 1 : 0;
        if (obs_bulk && warp == W_ISSUE && lane < tile_rows) {
            const EnvU& u = s_env[lane];
            const float* srow = s_row + lane * ROWP + u.row_shift;
            float* g = env.obs_buf + u.e * env.obs_row_stride;
            const int n = L.o_hf, head = min((4 - u.row_shift) & 3, n), nvec = (n - head) >> 2, t0 = head + 4 * nvec;
            if (nvec > 0) { bulk_store(g + head, srow + head, 16u * nvec); bulk_pending = true; }
            for (int k = 0; k < head; ++k) g[k] = srow[k];
            for (int k = t0; k < n; ++k) g[k] = srow[k];
        } else if (ref_bulk && warp == W_ISSUE && lane >= TILE && lane < TILE + NREF) {
            const int f = lane - TILE;
            const parc_view3_t& v = f == 0 ? env.ref_root_pos : f == 1 ? env.ref_root_rot : f == 2 ? env.ref_root_vel : f == 3 ? env.ref_root_ang_vel
                                  : f == 4 ? env.ref_body_pos : f == 5 ? env.ref_joint_rot : f == 6 ? env.ref_dof_pos : f == 7 ? env.ref_dof_vel : env.ref_contacts;
            const int fo = f == 0 ? R.root_pos : f == 1 ? R.root_rot : f == 2 ? R.root_vel : f == 3 ? R.root_ang_vel : f == 4 ? R.body_pos
                         : f == 5 ? R.joint_rot : f == 6 ? R.dof_pos : f == 7 ? R.dof_vel : R.contacts;
            const int fe = f == 0 ? R.root_rot : f == 1 ? R.root_vel : f == 2 ? R.root_ang_vel : f == 3 ? R.body_pos : f == 4 ? R.joint_rot
                         : f == 5 ? R.dof_pos : f == 6 ? R.dof_vel : f == 7 ? R.contacts : R.width;
            const int w = fe - fo;
            if (w > 0 && v.ptr) { bulk_store(v.ptr + e_first * w, s_ref + fo * TILE, (unsigned)(w * TILE * sizeof(float))); bulk_pending = true; }
        }
    }
    if (bulk_pending) bulk_commit();

    // ================================ phase D ====================================================
    PARC_STAMP(5);
    static_assert(TILE + NREF <= 32 && THREADS >= 96, "issuing lanes fit one warp; three warps take the tail roles");
    // 16 lanes each: the reward in warp 0, the termination flags in warp 2 (PARC_TAIL_ROLES; else both in warp 0)
    constexpr bool RW2 = PARC_REWARD_TWO_LANES && 2 * TILE <= 32;
    // second reward lane of an env (lanes TILE .. 2 TILE - 1 of warp 0): the two long sums, handed over by a shuffle
    float vel_hi = 0.f, con_hi = 0.f;
    if (RW2 && do_reward && (tid >> 5) == 0) {
        const int lane = tid & 31;
        if (lane >= TILE && lane - TILE < tile_rows) {
            const float* tm = s_terms + (lane - TILE) * TP;
#pragma unroll 4
            for (int k = 0; k < D; ++k) vel_hi += tm[o_vel_t + k];
            if (c_contact_info) {
#pragma unroll 4
                for (int k = 0; k < J; ++k) con_hi += tm[o_con_t + k];
            }
        }
        vel_hi = __shfl_down_sync(FULL_MASK, vel_hi, TILE);
        con_hi = __shfl_down_sync(FULL_MASK, con_hi, TILE);
    }
    const bool run_reward = do_reward && (tid >> 5) == 0 && (tid & 31) < tile_rows;
    const bool run_done = do_done && (tid >> 5) == (PARC_TAIL_ROLES ? 2 : 0) && (tid & 31) < tile_rows;
    if (run_reward || run_done) {
        const int el = tid & 31;
        const EnvU& u = s_env[el];
        const long long e = u.e;
        const float* rp = refp(R.root_pos, 3, el);
        const Vec3 ref_rpos = make_vec3(rp[0], rp[1], rp[2]);
        const Quat ref_rrot = quat_from_f4(*reinterpret_cast<const float4*>(refp(R.root_rot, 4, el)));
        if (run_reward) {
            // _update_reward (ig_parkour_env.py:984-1044) + compute_deepmimic_reward (mgdm_dm_util.py:269-332)
            const float* tm = s_terms + el * TP;
            float pose_t = 0.f, vel_t = 0.f, key_t = 0.f, con_t = 0.f;
#pragma unroll 4
            for (int k = 0; k < J - 1; ++k) pose_t += tm[o_pose_t + k];
            if (RW2) vel_t = vel_hi;
            else {
#pragma unroll 4
                for (int k = 0; k < D; ++k) vel_t += tm[o_vel_t + k];
            }
            for (int k = 0; k < K; ++k) key_t += tm[o_key_t + k];
            Vec3 rpd = ref_rpos - u.c_pos;
            if (!c_track_root) { rpd.x = 0.f; rpd.y = 0.f; }
            if (!c_track_root_h) rpd.z = 0.f;
            const float root_pos_err = rpd.x * rpd.x + rpd.y * rpd.y + rpd.z * rpd.z;
            Quat a_rot = u.c_rot, b_rot = ref_rrot;
            Vec3 a_vel = u.c_vel, a_av = u.c_avel;
            const float* rv = refp(R.root_vel, 3, el); const float* rav = refp(R.root_ang_vel, 3, el);
            Vec3 b_vel = make_vec3(rv[0], rv[1], rv[2]);
            Vec3 b_av = make_vec3(rav[0], rav[1], rav[2]);
            if (!c_track_root) {   // convert_to_local (mgdm_dm_util.py:246-267)
                const Quat hr = calc_heading_quat_inv(ref_rrot);
                a_rot = quat_mul(u.hinv, u.c_rot); b_rot = quat_mul(hr, ref_rrot);
                a_vel = quat_rotate(u.hinv, a_vel); b_vel = quat_rotate(hr, b_vel);
                a_av = quat_rotate(u.hinv, a_av); b_av = quat_rotate(hr, b_av);
            }
            const float rra = quat_diff_angle(a_rot, b_rot);
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
            if (c_contact_info) {
                if (RW2) con_t = con_hi;
                else {
#pragma unroll 4
                    for (int k = 0; k < J; ++k) con_t += tm[o_con_t + k];
                }
                pen = con_t / (float)C;
                total += pen;
            }
            env.reward_buf[e] = total;
            if (env.reward_terms) {
                float4* t4 = reinterpret_cast<float4*>(env.reward_terms + e * 8);
                t4[0] = make_float4(pose_r, vel_r, root_pose_r, root_vel_r);
                t4[1] = make_float4(key_r, pen, total, 0.f);
            }
        }
        if (run_done) {
            // _update_done (ig_parkour_env.py:967-982; mgdm_dm_util.py:147-172, 334-402; dm_env.py:628-665)
            int done = PARC_DONE_NULL;
            if (u.time >= cfg.episode_length) done = PARC_DONE_TIME;
            if (c_early_term) {
                bool failed = false;
                if (c_n_fall > 0) failed = (u.flags & 1) && (u.flags & 2);
                if (c_pose_term) {
                    bool pf = (u.flags & 4) != 0;
                    if (c_track_root) {
                        const Vec3 s0 = make_vec3(view_ld3(env.char_rigid_body_pos, e, 0, 0), view_ld3(env.char_rigid_body_pos, e, 0, 1),
                                                  view_ld3(env.char_rigid_body_pos, e, 0, 2));
                        const Vec3 d = s0 - ref_rpos;
                        pf = pf || (sqnorm3_seq(d) > cfg.root_pos_termination_dist);   // host passes fp32(dist*dist)
                        pf = pf || (fabsf(quat_diff_angle(u.c_rot, ref_rrot)) > cfg.root_rot_termination_angle);
                    }
                    failed = failed || pf;
                }
                failed = failed && (u.time > 1e-5f);
                if (failed) done = PARC_DONE_FAIL;
            }
            // DeepMimicEnv.update_done tail (dm_env.py:638-665)
            const bool motion_end = (u.t0 >= u.mm.length) && (u.mm.loop_mode != PARC_LOOP_WRAP);
            int code = 0;
            if (done == PARC_DONE_FAIL) code = 1;
            else if (done != PARC_DONE_NULL || motion_end) code = 2;
            if (motion_end) done = PARC_DONE_FAIL;
            if (c_never_done) done = PARC_DONE_NULL;
            env.done_buf[e] = done;
            if (env.done_code) env.done_code[e] = (unsigned char)code;
            if (code != 0 && env.done_list) {
                const int slot = atomicAdd(env.done_count, 1);
                if (slot < env.num_envs) env.done_list[slot] = (int)e;   // a stale (un-zeroed) count must not overrun the list
            }
        }
    }

    PARC_STAMP(6);
    // ================================ phase H ====================================================
    // the height units nobody has taken yet (all of them unless the heights were filler work)
    height_units();

    PARC_STAMP(7);
    // ================================ phase E (general) ===========================================
    // staged values -> global with plain stores where the bulk path does not apply: warp w streams env w
    {
        const int warp = tid >> 5, lane = tid & 31;
        if ((do_obs && !obs_bulk) || (do_ref && !ref_bulk)) {
            for (int el = warp; el < tile_rows; el += THREADS / 32) {
                const EnvU& u = s_env[el];
                if (do_obs && !obs_bulk) {
                    const float* row = s_row + el * ROWP + u.row_shift;
                    float* g = env.obs_buf + u.e * env.obs_row_stride;
                    // column k is 16-byte aligned in both the staged row and the obs_buf row when (row_shift + k) % 4 == 0
                    const int n = L.o_hf, head = min((4 - u.row_shift) & 3, n), nvec = (n - head) >> 2, t0 = head + 4 * nvec;
                    if (lane < head) g[lane] = row[lane];
                    const float4* r4 = reinterpret_cast<const float4*>(row + head);
                    float4* g4 = reinterpret_cast<float4*>(g + head);
#pragma unroll 2
                    for (int k = lane; k < nvec; k += 32) g4[k] = r4[k];
                    if (t0 + lane < n) g[t0 + lane] = row[t0 + lane];
                }
                if (do_ref && !ref_bulk) {
#pragma unroll 2
                    for (int k = lane; k < R.width; k += 32) {
                        const uint4 m = s_refmap[k];
                        const unsigned long long base = ((unsigned long long)m.y << 32) | m.x;
                        if (base) *reinterpret_cast<float*>(base + (unsigned long long)u.e * m.z) = s_ref[(m.w & 0xfffffu) + el * (m.w >> 20)];
                    }
                }
            }
        }
    }
    if (bulk_pending) bulk_wait_read();   // the shared-memory sources must outlive the copies
    if (FAST) break;
    __syncthreads();                      // the block's next tile reuses the staging
    }
    PARC_STAMP(8);
#ifdef PARC_STEP_TIMING
    if (threadIdx.x == 0 && blockIdx.x < 16384) { unsigned smid; asm("mov.u32 %0, %%smid;" : "=r"(smid)); g_step_stamps[blockIdx.x * 16 + 15] = smid; }
#endif
}

}  // namespace parc

namespace parc {
__global__ void selftest_div_kernel(const float* __restrict__ x, long long n, float d, float* __restrict__ q_fast, float* __restrict__ q_ieee) {
    const float rc = __frcp_rn(d);
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        q_fast[i] = div_const(x[i], d, rc);
        q_ieee[i] = __fdiv_rn(x[i], d);
    }
}
}  // namespace parc

using namespace parc;

#ifdef PARC_STEP_TIMING
extern "C" int parc_b200_debug_step_stamps(long long* h_out, int n_blocks) {
    cudaError_t ce = cudaMemcpyFromSymbol(h_out, g_step_stamps, sizeof(long long) * 16 * (size_t)n_blocks);
    return ce == cudaSuccess ? 0 : -3;
}
#endif

namespace parc {
// packed_cell against the scalar, separately rounded ray_world + grid_index for n points of one (heading, root) pair
__global__ void selftest_cells_kernel(const float2* __restrict__ pts, long long n, float c, float s, float gx, float gy, float mnx, float mny,
                                      float dx, float dy, int dimx, int dimy, int* __restrict__ cell_packed, int* __restrict__ cell_scalar) {
    const float2 cs = make_float2(c, s), nsc = make_float2(-s, c), g2 = make_float2(gx, gy), nmn = make_float2(-mnx, -mny);
    const float2 rd2 = make_float2(__frcp_rn(dx), __frcp_rn(dy)), nd2 = make_float2(-dx, -dy);
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const float2 pt = pts[i];
        cell_packed[i] = packed_cell(pt, cs, nsc, g2, nmn, rd2, nd2, dimx - 1, dimy - 1, dimy);
        float wx, wy;
        ray_world(pt.x, pt.y, c, s, gx, gy, wx, wy);
        cell_scalar[i] = grid_index(wx, mnx, dx, dimx) * dimy + grid_index(wy, mny, dy, dimy);
    }
}
}  // namespace parc

extern "C" int parc_b200_selftest_cells(const float* pts_xy, int64_t n, float c, float s, float gx, float gy, float min_x, float min_y,
                                        float dx, float dy, int32_t dim_x, int32_t dim_y, int32_t* cell_packed, int32_t* cell_scalar, void* stream) {
    if (n == 0) return PARC_OK;
    PARC_REQUIRE(pts_xy && cell_packed && cell_scalar && dim_x > 0 && dim_y > 0, "selftest_cells: bad arguments");
    selftest_cells_kernel<<<kNumSMs * 8, 256, 0, as_stream(stream)>>>(reinterpret_cast<const float2*>(pts_xy), n, c, s, gx, gy, min_x, min_y, dx, dy,
                                                                     dim_x, dim_y, cell_packed, cell_scalar);
    return check_launch("selftest_cells");
}

extern "C" int parc_b200_selftest_div(const float* x, int64_t n, float d, float* q_fast, float* q_ieee, void* stream) {
    if (n == 0) return PARC_OK;
    PARC_REQUIRE(x && q_fast && q_ieee, "selftest_div: NULL pointer");
    selftest_div_kernel<<<kNumSMs * 8, 256, 0, as_stream(stream)>>>(x, n, d, q_fast, q_ieee);
    return check_launch("selftest_div");
}

extern "C" int parc_b200_tree_parts(const parc_char_model_t* m, int32_t* entries, int32_t* max_len, int32_t* num_slots, int32_t* root_slot) {
    if (!model_ok(m, "tree_parts")) return PARC_ERR_INVALID_ARG;
    PARC_REQUIRE(entries && max_len && num_slots && root_slot, "tree_parts: NULL output");
    parc_env_config_t cfg = {};
    parc_terrain_t tr = {};
    tr.dx = tr.dy = 1.f; tr.dim_x = tr.dim_y = 1;
    const StepPlan sp = make_step_plan(*m, 0, cfg, tr);
    for (int p = 0; p < NPART; ++p)
        for (int i = 0; i < PARC_MAX_BODIES; ++i) entries[p * PARC_MAX_BODIES + i] = i < sp.n_iter ? sp.sched[i * NPART + p] : -1;
    *max_len = sp.n_iter;
    *num_slots = sp.S.n_gslots;
    *root_slot = sp.root_slot;
    return PARC_OK;
}

extern "C" int parc_b200_ray_points_lattice(const float* h_pts_xy, int32_t num_points, int32_t* nx, int32_t* ny) {
    PARC_REQUIRE(nx && ny && num_points >= 0 && (num_points == 0 || h_pts_xy), "ray_points_lattice: bad arguments");
    *nx = *ny = 0;
    if (num_points < 2) return PARC_OK;
    // ny = length of the leading run of points that share x; then every point must be (xs[i], ys[j]) bit for bit
    int n_y = 1;
    while (n_y < num_points && h_pts_xy[2 * n_y] == h_pts_xy[0]) ++n_y;
    if (num_points % n_y != 0) return PARC_OK;
    const int n_x = num_points / n_y;
    for (int i = 0; i < n_x; ++i)
        for (int j = 0; j < n_y; ++j) {
            const float* p = h_pts_xy + 2 * ((long long)i * n_y + j);
            if (p[0] != h_pts_xy[2 * (long long)i * n_y] || p[1] != h_pts_xy[2 * j + 1]) return PARC_OK;
        }
    *nx = n_x; *ny = n_y;
    return PARC_OK;
}

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

namespace parc {
// Argument checks shared by parc_b200_env_step and parc_b200_env_step_host.
int validate_env_step(const parc_char_model_t* m, const parc_motion_lib_t* lib, const parc_terrain_t* terrain,
                      const parc_env_config_t* cfg, const parc_env_buffers_t* env, uint32_t parts) {
    int rc = validate_env(m, lib, cfg, env, "env_step");
    if (rc) return rc;
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
    PARC_REQUIRE((long long)terrain->dim_x * terrain->dim_y < (1ll << 31), "env_step: terrain has more than 2^31 cells");
    PARC_REQUIRE(cfg->ray_grid_nx >= 0 && cfg->ray_grid_ny >= 0 && (cfg->ray_grid_nx == 0 || (long long)cfg->ray_grid_nx * cfg->ray_grid_ny == cfg->num_ray_points),
                 "env_step: ray_grid %d x %d does not cover the %d ray points", cfg->ray_grid_nx, cfg->ray_grid_ny, cfg->num_ray_points);
    const StepPlan sp_check = make_step_plan(*m, lib->num_contacts, *cfg, *terrain);
    PARC_REQUIRE(sp_check.S.total <= 227 * 1024, "env_step: staging of %zu bytes exceeds shared memory", sp_check.S.total);
    return PARC_OK;
}
}  // namespace parc

extern "C" int parc_b200_env_step(const parc_char_model_t* m, const parc_motion_lib_t* lib, const parc_terrain_t* terrain,
                                  const parc_env_config_t* cfg, const parc_env_buffers_t* env, uint32_t parts,
                                  const int64_t* env_ids, int64_t n_sub, void* stream) {
    int rc = validate_env(m, lib, cfg, env, "env_step");
    if (rc) return rc;
    const long long n_rows = env_ids ? n_sub : env->num_envs;
    PARC_REQUIRE(n_rows >= 0, "env_step: negative row count");
    if (n_rows == 0 || parts == 0) return PARC_OK;
    rc = parc::validate_env_step(m, lib, terrain, cfg, env, parts);
    if (rc) return rc;
    return parc::launch_env_step(m, lib, terrain, cfg, env, parts, env_ids, n_rows, nullptr, nullptr, stream);
}

namespace parc {
// Shared launcher: rows come from env_ids (i64, host count), from env_ids32 + a device-side count
// (done_list/done_count of the previous kernels, grid sized for max_rows), or are simply 0..N-1.
int launch_env_step(const parc_char_model_t* m, const parc_motion_lib_t* lib, const parc_terrain_t* terrain,
                    const parc_env_config_t* cfg, const parc_env_buffers_t* env, uint32_t parts, const int64_t* env_ids,
                    long long max_rows, const int* env_ids32, const int* n_rows_dev, void* stream, long long first_env) {
    const StepPlan sp = make_step_plan(*m, lib->num_contacts, *cfg, *terrain, env);
    const SmemPlan& plan = sp.S;
    const bool lattice = sp.lat_nx > 0;
    // the specialised instantiation: exactly the layout and flags it was compiled for (see FastDims), every part, all rows
    const bool fast = PARC_STEP_FAST && lattice && parts == PARC_STEP_ALL && env_ids == nullptr && env_ids32 == nullptr && n_rows_dev == nullptr &&
                      m->num_bodies == FastDims::J && m->dof_size == FastDims::D && lib->num_contacts == FastDims::C &&
                      cfg->enable_tar_obs && cfg->num_tar_steps == FastDims::S && cfg->num_key_bodies == FastDims::K &&
                      cfg->num_ray_points == FastDims::P && sp.lat_nx == FastDims::NX && sp.lat_ny == FastDims::NY &&
                      !cfg->global_obs && !cfg->root_height_obs && cfg->use_contact_info && cfg->track_root && cfg->track_root_h &&
                      cfg->enable_early_termination && cfg->pose_termination && cfg->num_fall_contact_bodies == 0 && !cfg->never_done &&
                      sp.lrot_ident && sp.n_iter == FastDims::N_ITER && sp.root_slot == -1 && sp.ref_fast &&
                      lib->irec == interp_record_floats(FastDims::J, FastDims::C) && lib->vrec == vel_record_floats(FastDims::D) &&
                      memcmp(&sp.L, &kFastL, sizeof(ObsLayout)) == 0 && memcmp(&sp.R, &kFastR, sizeof(RefLayout)) == 0 &&
                      sp.S.row_pitch == kFastS.row_pitch && sp.S.n_pose == kFastS.n_pose && sp.S.n_gslots == kFastS.n_gslots &&
                      sp.S.term_pitch == kFastS.term_pitch && sp.S.off_env == kFastS.off_env && sp.S.off_row == kFastS.off_row &&
                      sp.S.off_ref == kFastS.off_ref && sp.S.off_charq == kFastS.off_charq && sp.S.off_gx == kFastS.off_gx &&
                      sp.S.off_refmap == kFastS.off_refmap && sp.S.off_tab == kFastS.off_tab && sp.S.off_pts == kFastS.off_pts &&
                      sp.S.off_lat == kFastS.off_lat && sp.S.off_patch == kFastS.off_patch && sp.S.off_ctr == kFastS.off_ctr && sp.S.total == kFastS.total;
    const int variant = fast ? 2 : (lattice ? 1 : 0);
    // opt-in dynamic shared memory: a per-device function attribute, raised only when a larger plan shows up
    static size_t smem_attr[64][3] = {};
    int dev = 0;
    cudaGetDevice(&dev);
    size_t& have = smem_attr[dev & 63][variant];
    if (plan.total > have) {
        cudaError_t ce = variant == 2 ? cudaFuncSetAttribute(env_step_kernel<STEP_THREADS, true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)plan.total)
                       : variant == 1 ? cudaFuncSetAttribute(env_step_kernel<STEP_THREADS, true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)plan.total)
                                      : cudaFuncSetAttribute(env_step_kernel<STEP_THREADS, false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)plan.total);
        if (ce != cudaSuccess) { set_error("env_step: %s", cudaGetErrorString(ce)); return PARC_ERR_CUDA; }
#ifdef PARC_STEP_CARVEOUT
        if (variant == 2) cudaFuncSetAttribute(env_step_kernel<STEP_THREADS, true, true>, cudaFuncAttributePreferredSharedMemoryCarveout, PARC_STEP_CARVEOUT);
#endif
        have = plan.total;
    }
    long long blocks = (max_rows + TILE - 1) / TILE;
    if (n_rows_dev) blocks = std::min<long long>(blocks, (long long)kNumSMs * 4);      // row count on the device: blocks loop over the tiles
    if (variant == 2)
        env_step_kernel<STEP_THREADS, true, true><<<(unsigned)blocks, STEP_THREADS, plan.total, as_stream(stream)>>>(
            *m, *lib, *terrain, *cfg, *env, sp, parts, (const long long*)env_ids, max_rows, env_ids32, n_rows_dev, first_env);
    else if (variant == 1)
        env_step_kernel<STEP_THREADS, true, false><<<(unsigned)blocks, STEP_THREADS, plan.total, as_stream(stream)>>>(
            *m, *lib, *terrain, *cfg, *env, sp, parts, (const long long*)env_ids, max_rows, env_ids32, n_rows_dev, first_env);
    else
        env_step_kernel<STEP_THREADS, false, false><<<(unsigned)blocks, STEP_THREADS, plan.total, as_stream(stream)>>>(
            *m, *lib, *terrain, *cfg, *env, sp, parts, (const long long*)env_ids, max_rows, env_ids32, n_rows_dev, first_env);
    return check_launch("env_step");
}
}  // namespace parc
