// Fused post-physics step of the DeepMimic-style tracking env:
//   heights -> time -> reference motion -> observation -> reward -> termination
// i.e. IGEnv._post_physics_step (PARC/motion_tracker/envs/ig_env.py:368-377) minus PhysX, in ONE
// launch that reads every input once and writes every output once.
//
// Work decomposition (a block owns TILE = 16 consecutive envs, 512 threads = two groups of 8 envs x 8 warps):
//   A  per-env uniforms (heading, time, clip meta) by 8 lanes; character dof -> quaternion by
//      one thread per (env, joint)
//   B  warps 0..1: one thread per (env, pose) -- pose 0 = reference, 1..S = target steps,
//      S+1 = simulated character -- walks the joint list once: slerp of the two frame records,
//      forward kinematics against the parent transform kept in shared memory, observation
//      features straight into the staged row;   warps 2..7 meanwhile: heightmap samples, one
//      thread per sample point, written straight into the observation rows (coalesced)
//   C  one thread per (env, joint | dof | body | key body): rot_to_dof, reward / termination terms
//   D  8 lanes: ordered sums, exponentials, reward, done flag, done-list append
//   E  all threads: staged rows (observation minus heights, ref_* buffers) -> global, coalesced
// The v1 kernel (one warp per env, level-by-level FK with warp shuffles) executed 9.8k warp
// instructions per env and sat at 10 % of the HBM roofline, issue-bound with 22 % of the lanes
// doing FK work; this layout keeps every lane on its own pose and needs ~2k.
#include "parc_common.cuh"
#include "parc_lanes.cuh"

// The kernel is instruction-fetch bound (ncu: stall_no_inst is the largest stall class; the SASS is ~110 KB against a
// 32 KB L1.5 instruction cache), so loops that run a handful of iterations per block are kept rolled: nvcc's default
// 4x unrolling of them cost 5 % of the launch time.
#define PARC_COLD_LOOP _Pragma("unroll 1")
// Tuning constants, each set by an A/B measurement on a B200:
#ifndef PARC_STEP_BLOCKS_PER_SM_AT_TILE8
#define PARC_STEP_BLOCKS_PER_SM_AT_TILE8 4   // launch bound: 4 blocks per SM at 8 envs per block = 2 at 16 (64 registers)
#endif
#ifndef PARC_HEIGHT_GATHERS
#define PARC_HEIGHT_GATHERS 2                 // independent heightfield gathers per lane and loop iteration
#endif
#ifndef PARC_HEIGHTS_LATE
#define PARC_HEIGHTS_LATE 1   // 1: height samples after phases C/D (nothing waits for them at a barrier; they cover phase D)
#endif
#ifndef PARC_SPLIT_A
#define PARC_SPLIT_A 1      // 1: phase A uniforms as two dependent-load chains in two warps
#endif
#ifndef PARC_VEC_ROWS
#define PARC_VEC_ROWS 1      // 1: staged rows are shifted to the 16-byte phase of their obs_buf row, copy-out in 128-bit pieces
#endif

namespace parc {

#ifndef PARC_STEP_TILE
#define PARC_STEP_TILE 16                      // envs per block; a multiple of 8
#endif
constexpr int TILE = PARC_STEP_TILE;      // envs per block (== warps per block)
constexpr int STEP_THREADS = 32 * TILE;

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

// Reference-motion outputs staged behind the observation columns of a row (DeepMimicEnv.
// _update_ref_motion, dm_env.py:523-545): offsets relative to the start of the section.
struct RefLayout {
    int root_pos, root_rot, root_vel, root_ang_vel, body_pos, joint_rot, dof_pos, dof_vel, contacts, width;
};
__host__ __device__ inline RefLayout make_ref_layout(int J, int D, int C) {
    RefLayout R;
    R.root_pos = 0; R.root_rot = 3; R.root_vel = 7; R.root_ang_vel = 10; R.body_pos = 13;
    R.joint_rot = R.body_pos + 3 * J;
    R.dof_pos = R.joint_rot + 4 * (J - 1);
    R.dof_vel = R.dof_pos + D;
    R.contacts = R.dof_vel + D;
    R.width = R.contacts + C;
    return R;
}

// Per-env uniforms computed once (phase A) and read by every later phase.
struct EnvU {
    long long e;             // env index of this row
    long long mid;           // motion id
    const float* vel_rec;    // velocity record of the reference pose's frame idx0 (set by phase B)
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

// Shared-memory plan (dynamic): sized on the host by step_smem_bytes().
struct SmemPlan {
    int row_pitch;           // floats per staged row: o_hf + ref width (+ up to 3 floats of alignment shift), = 4 mod 32
    int n_pose;              // poses per env: S + 2
    int n_slots;             // FK transforms a thread keeps: max over parts of bodies with children inside the part
    int term_pitch;          // floats of reward terms per env: (J-1) + D + J + K
    size_t off_env, off_row, off_charq, off_xf, off_terms, off_refmap, off_tab, off_pts, total;
};

constexpr int NPART = 4;                // threads cooperating on one (env, pose): each walks one part of the tree

inline SmemPlan make_plan(const parc_char_model_t& cm, int C, const parc_env_config_t& cfg, const ObsLayout& L, int n_slots) {
    SmemPlan p;
    const int J = cm.num_bodies, D = cm.dof_size;
    const RefLayout R = make_ref_layout(J, D, C);
#if PARC_VEC_ROWS
    // A row starts at a multiple of 4 floats plus the env's shift (0..3), so that column k of the staged row and of the
    // obs_buf row have the same 16-byte phase. Pitch = 4 (mod 32): the 8 envs a warp touches at one column land in
    // banks 4 el + shift -- all different.
    p.row_pitch = (L.o_hf + R.width + 3 + 3) & ~3;
    while ((p.row_pitch & 31) != 4) p.row_pitch += 4;
#else
    p.row_pitch = (L.o_hf + R.width) | 1;
#endif
    p.n_pose = (cfg.enable_tar_obs ? cfg.num_tar_steps : 0) + 2;
    p.n_slots = n_slots;
    p.term_pitch = (J - 1) + D + J + cfg.num_key_bodies;
    size_t o = 0;
    p.off_env = o;    o += sizeof(EnvU) * TILE;                          o = (o + 15) & ~size_t(15);
    p.off_row = o;    o += sizeof(float) * TILE * p.row_pitch;           o = (o + 15) & ~size_t(15);
    p.off_charq = o;  o += sizeof(float4) * TILE * J;                    o = (o + 15) & ~size_t(15);
    // The FK transform columns are only live in phase B; the reward terms (phase C/D) and the ref_* scatter map
    // (built in phase C, read in phase E) reuse the same bytes.
    const size_t xf_bytes = sizeof(float) * 7 * p.n_slots * STEP_THREADS;
    const size_t terms_bytes = (sizeof(float) * TILE * p.term_pitch + 15) & ~size_t(15);
    p.off_xf = o;
    p.off_terms = o;
    p.off_refmap = o + terms_bytes;
    o += std::max(xf_bytes, terms_bytes + 16 * (size_t)R.width);      o = (o + 15) & ~size_t(15);
    p.off_tab = o;    o += sizeof(int) * NPART * PARC_MAX_BODIES + sizeof(float) * 8 * J;   o = (o + 15) & ~size_t(15);
    p.off_pts = o;    o += sizeof(float2) * (size_t)cfg.num_ray_points;   // ray_xy_points, shared by the tile's envs
    p.total = o;
    return p;
}

// Everything the kernel would otherwise re-derive per thread from the model / config, computed once
// on the host per launch and passed by value.
struct StepPlan {
    ObsLayout L;
    RefLayout R;
    SmemPlan S;
    unsigned has_child;                   // bit j: body j has children (root excluded from slots)
    unsigned key_mask;                    // bit j: body j is listed in key_body_ids
    int lrot_ident;                       // every _local_rotation is the identity
    int max_part_len;                     // longest joint list of a part
    // part_entry[p][i]: joint j | owned << 8 | (store slot + 1) << 12 | (parent slot + 1) << 16; -1 past the end.
    // A part's list is closed under ancestors (shared ancestors are recomputed, only one part owns = outputs them).
    int part_entry[NPART][PARC_MAX_BODIES];
    float rdx, rdy;                       // correctly rounded 1/dx, 1/dy
    float max_ix, max_iy;                 // dim - 1 as floats
};

// Splits the joints 1..J-1 into NPART ancestor-closed lists of similar length: start from the
// subtrees below the root, keep splitting the largest one at its first branching body (the chain
// above the branch is recomputed by every piece, owned by the first), then pack the pieces into
// NPART bins, largest first.
struct TreeUnit { int prefix[PARC_MAX_BODIES]; unsigned char prefix_owned[PARC_MAX_BODIES]; int n_prefix; int top; int size; };

static int subtree_size(const parc_char_model_t& cm, int r) {
    int n = 1;
    for (int j = r + 1; j < cm.num_bodies; ++j) {
        int a = cm.parent[j];
        while (a > r) a = cm.parent[a];
        if (a == r) ++n;
    }
    return n;
}
static bool in_subtree(const parc_char_model_t& cm, int r, int j) {
    while (j > r) j = cm.parent[j];
    return j == r;
}

inline StepPlan make_step_plan(const parc_char_model_t& cm, int C, const parc_env_config_t& cfg, const parc_terrain_t& tr) {
    StepPlan sp;
    const int J = cm.num_bodies;
    sp.L = make_layout(J, cm.dof_size, C, cfg);
    sp.R = make_ref_layout(J, cm.dof_size, C);
    sp.has_child = 0;
    for (int j = 1; j < J; ++j) sp.has_child |= 1u << cm.parent[j];
    sp.key_mask = 0;
    for (int k = 0; k < cfg.num_key_bodies; ++k) sp.key_mask |= 1u << cfg.key_body_ids[k];
    sp.lrot_ident = 1;
    for (int j = 1; j < J; ++j)
        if (!(cm.local_rot[j][0] == 0.f && cm.local_rot[j][1] == 0.f && cm.local_rot[j][2] == 0.f && cm.local_rot[j][3] == 1.f)) sp.lrot_ident = 0;
    // ---- units ----
    TreeUnit units[2 * PARC_MAX_BODIES];
    int nu = 0;
    for (int j = 1; j < J; ++j)
        if (cm.parent[j] == 0) { TreeUnit u; u.n_prefix = 0; u.top = j; u.size = subtree_size(cm, j); units[nu++] = u; }
    const int target = (J - 1 + NPART - 1) / NPART;
    for (int iter = 0; iter < 4 * PARC_MAX_BODIES; ++iter) {
        int best = -1;
        for (int a = 0; a < nu; ++a)
            if (units[a].size > target && (best < 0 || units[a].size > units[best].size)) {
                // splittable only if some body on the single-child chain below `top` branches
                int node = units[a].top, nch;
                for (;;) { nch = 0; int only = -1; for (int j = node + 1; j < J; ++j) if (cm.parent[j] == node) { ++nch; only = j; } if (nch != 1) break; node = only; }
                if (nch >= 2) best = a;
            }
        if (best < 0 || (nu >= NPART && units[best].size <= target + 1) || nu + PARC_MAX_BODIES / 2 > 2 * PARC_MAX_BODIES) break;
        TreeUnit u = units[best];
        int node = u.top;
        for (;;) {
            u.prefix[u.n_prefix] = node; u.prefix_owned[u.n_prefix] = 1; ++u.n_prefix;
            int nch = 0, only = -1;
            for (int j = node + 1; j < J; ++j) if (cm.parent[j] == node) { ++nch; only = j; }
            if (nch != 1) break;
            node = only;
        }
        units[best] = units[--nu];
        bool first = true;
        for (int j = node + 1; j < J; ++j)
            if (cm.parent[j] == node) {
                TreeUnit c = u;
                if (!first) for (int k = 0; k < c.n_prefix; ++k) c.prefix_owned[k] = 0;
                first = false;
                c.top = j; c.size = c.n_prefix + subtree_size(cm, j);
                units[nu++] = c;
            }
    }
    // ---- pack into NPART bins, largest unit first ----
    unsigned bin_has[NPART] = {0, 0, 0, 0}, bin_owned[NPART] = {0, 0, 0, 0};
    int load[NPART] = {0, 0, 0, 0};
    bool used[2 * PARC_MAX_BODIES] = {false};
    for (int n = 0; n < nu; ++n) {
        int a = -1;
        for (int k = 0; k < nu; ++k) if (!used[k] && (a < 0 || units[k].size > units[a].size)) a = k;
        used[a] = true;
        int b = 0;
        for (int k = 1; k < NPART; ++k) if (load[k] < load[b]) b = k;
        const TreeUnit& u = units[a];
        for (int k = 0; k < u.n_prefix; ++k) { bin_has[b] |= 1u << u.prefix[k]; if (u.prefix_owned[k]) bin_owned[b] |= 1u << u.prefix[k]; }
        for (int j = u.top; j < J; ++j) if (in_subtree(cm, u.top, j)) { bin_has[b] |= 1u << j; bin_owned[b] |= 1u << j; }
        load[b] = __builtin_popcount(bin_has[b]);
    }
    // a prefix body owned by a unit packed next to another copy must be owned exactly once overall
    unsigned owned_seen = 0;
    for (int b = 0; b < NPART; ++b) { bin_owned[b] &= ~owned_seen; owned_seen |= bin_owned[b]; }
    sp.max_part_len = 0;
    int max_slots = 1;
    for (int b = 0; b < NPART; ++b) {
        // Slots are handed out by liveness: a body's transform is kept only until its last child inside the part has
        // read it, and a child may take over its parent's slot (the walk reads the parent before storing the child).
        int slot_of[PARC_MAX_BODIES], kids_left[PARC_MAX_BODIES];
        bool slot_busy[PARC_MAX_BODIES] = {false};
        int ns = 0, len = 0;
        for (int j = 1; j < J; ++j) {
            slot_of[j] = -1; kids_left[j] = 0;
            if (!((bin_has[b] >> j) & 1u)) continue;
            for (int c = j + 1; c < J; ++c) if (cm.parent[c] == j && ((bin_has[b] >> c) & 1u)) ++kids_left[j];
        }
        for (int j = 1; j < J; ++j) {
            if (!((bin_has[b] >> j) & 1u)) continue;
            const int par = cm.parent[j];
            const int pslot = par == 0 ? -1 : slot_of[par];
            if (par != 0 && --kids_left[par] == 0) slot_busy[pslot] = false;
            if (kids_left[j] > 0) {
                int f = 0;
                while (slot_busy[f]) ++f;
                slot_busy[f] = true; slot_of[j] = f;
                if (f + 1 > ns) ns = f + 1;
            }
            sp.part_entry[b][len++] = j | ((int)((bin_owned[b] >> j) & 1u) << 8) | ((slot_of[j] + 1) << 12) | ((pslot + 1) << 16);
        }
        for (int k = len; k < PARC_MAX_BODIES; ++k) sp.part_entry[b][k] = -1;
        if (len > sp.max_part_len) sp.max_part_len = len;
        if (ns > max_slots) max_slots = ns;
    }
    sp.S = make_plan(cm, C, cfg, sp.L, max_slots);
    sp.rdx = 1.0f / tr.dx; sp.rdy = 1.0f / tr.dy;    // IEEE single division on the host == RN(1/d)
    sp.max_ix = (float)(tr.dim_x - 1); sp.max_iy = (float)(tr.dim_y - 1);
    return sp;
}

__device__ __forceinline__ float4 ldg4(const float* p) { return __ldg(reinterpret_cast<const float4*>(p)); }

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

// (x - mn) / d for the nearest-cell index with d, 1/d uniform: q = x * rc; r = fma(-q, d, x);
// q' = fma(r, rc, q) (Markstein) is the correctly rounded quotient for normal operands when rc is
// the correctly rounded reciprocal; checked against div.rn by parc_b200_selftest_div.
__device__ __forceinline__ float div_const(float x, float d, float rc) {
    const float q = __fmul_rn(x, rc);
    const float r = __fmaf_rn(-q, d, x);
    return __fmaf_rn(r, rc, q);
}
struct GridParams { float min_x, min_y, dx, dy, rdx, rdy, max_ix, max_iy; };
__device__ __forceinline__ int grid_index_fast(float p, float mn, float d, float rd, float mx) {
    float u = rintf(div_const(__fsub_rn(p, mn), d, rd));
    u = fminf(fmaxf(u, 0.f), mx);
    return (int)u;
}

// Grid cell of one ray sample point: rotate_2d_vec + root xy, (px c - py s + gx, px s + py c + gy), then
// clamp(round_half_even((w - min) / dxdy), 0, dim - 1) (terrain_util.py:139-152). The sums and the Markstein division
// (div_const) are packed fp32x2 (FADD2 / FMUL2 / FFMA2, sm_100): x and y of the point share those instructions.
// Every product and sum must be rounded on its own, as torch's separate mul / sub / add are. ptxas (CUDA 12.9)
// contracts mul.rn.f32x2 + add.rn.f32x2 into FFMA2 despite the .rn (seen in the SASS; an FFMA2 with a -0 addend is
// folded back and contracted too), so the four products are scalar mul.rn -- which it does leave alone -- and only
// the sums and the division are packed. parc_b200_selftest_cells checks the result against the scalar path on the GPU.
__device__ __forceinline__ int packed_cell(float2 pt, float2 cs, float2 nsc, float2 g2, float2 nmn, float2 rd2, float2 nd2, int imx, int imy,
                                           int dimy) {
    const float2 a = make_float2(__fmul_rn(pt.x, cs.x), __fmul_rn(pt.x, cs.y));
    const float2 b = make_float2(__fmul_rn(pt.y, nsc.x), __fmul_rn(pt.y, nsc.y));
    const float2 w = __fadd2_rn(__fadd2_rn(a, b), g2);
    const float2 x = __fadd2_rn(w, nmn);
    float2 q = __fmul2_rn(x, rd2);
    const float2 r = __ffma2_rn(q, nd2, x);
    q = __ffma2_rn(r, rd2, q);
    const int ix = max(0, min(__float2int_rn(q.x), imx));
    const int iy = max(0, min(__float2int_rn(q.y), imy));
    return ix * dimy + iy;
}

// One env's heightmap samples by one warp (_refresh_obs_hfs, ig_parkour_env.py:518-533; get_hf_val, terrain_util.py:139-156).
// Points come from shared memory, four per lane and iteration so that four independent gathers are in flight; the two
// destination rows (either may be absent) are running pointers with immediate offsets. FULL: P is a multiple of 128
// (no tail predicates).
template <bool FULL>
__device__ __forceinline__ void heights_of_env(const float2* __restrict__ s_pts, int P, int lane, const float* __restrict__ hf, int dimy,
                                               int imx, int imy, float2 cs, float2 nsc, float2 g2, float gz, float2 nmn, float2 rd2,
                                               float2 nd2, float lo, float hi, float* __restrict__ hrow, float* __restrict__ orow) {
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
        for (int k = 0; k < U; ++k) {
            cell[k] = packed_cell(pt[k], cs, nsc, g2, nmn, rd2, nd2, imx, imy, dimy);
        }
#pragma unroll
        for (int k = 0; k < U; ++k) v[k] = __ldg(hf + cell[k]);
#pragma unroll
        for (int k = 0; k < U; ++k) {
            const float h = fminf(fmaxf(__fsub_rn(v[k], gz), lo), hi);
            if (FULL || p + 32 * k < P) {
                if (HAS_H) ph[32 * k] = h;
                if (HAS_O) po[32 * k] = h;
            }
        }
    }
}

template <int THREADS>
__global__ void __launch_bounds__(THREADS, PARC_STEP_BLOCKS_PER_SM_AT_TILE8 * 8 / PARC_STEP_TILE) env_step_kernel(
    const __grid_constant__ parc_char_model_t cm, const __grid_constant__ parc_motion_lib_t lib,
    const __grid_constant__ parc_terrain_t terrain, const __grid_constant__ parc_env_config_t cfg,
    const __grid_constant__ parc_env_buffers_t env, const __grid_constant__ StepPlan sp, const unsigned parts,
    const long long* __restrict__ env_ids, const long long n_rows_host, const int* __restrict__ env_ids32,
    const int* __restrict__ n_rows_dev, const long long first_env) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const long long n_rows = n_rows_dev ? min((long long)*n_rows_dev, n_rows_host) : n_rows_host;
    const long long row0 = (long long)blockIdx.x * TILE;
    if (row0 >= n_rows) return;
    const int tid = threadIdx.x;
    const int J = cm.num_bodies, D = cm.dof_size, C = lib.num_contacts;
    const int S = cfg.enable_tar_obs ? cfg.num_tar_steps : 0;
    const int K = cfg.num_key_bodies, P = cfg.num_ray_points;
    const ObsLayout& L = sp.L;
    const RefLayout& R = sp.R;
    const SmemPlan& plan = sp.S;
    EnvU* s_env = reinterpret_cast<EnvU*>(smem_raw + plan.off_env);
    float* s_row = reinterpret_cast<float*>(smem_raw + plan.off_row);
    float4* s_charq = reinterpret_cast<float4*>(smem_raw + plan.off_charq);
    float* s_xf = reinterpret_cast<float*>(smem_raw + plan.off_xf);
    float* s_terms = reinterpret_cast<float*>(smem_raw + plan.off_terms);
    unsigned long long* s_refmap = reinterpret_cast<unsigned long long*>(smem_raw + plan.off_refmap);
    int* s_part = reinterpret_cast<int*>(smem_raw + plan.off_tab);
    float* s_jc = reinterpret_cast<float*>(smem_raw + plan.off_tab + sizeof(int) * NPART * PARC_MAX_BODIES);
    float2* s_pts = reinterpret_cast<float2*>(smem_raw + plan.off_pts);
    const int ROWP = plan.row_pitch;

    const bool do_heights = parts & PARC_STEP_HEIGHTS;
    const bool do_obs = parts & PARC_STEP_OBS;
    const bool do_ref = parts & PARC_STEP_REF;
    const bool do_reward = parts & PARC_STEP_REWARD;
    const bool do_done = parts & PARC_STEP_DONE;
    const bool need_ref_pose = do_ref || do_reward || do_done;
    const bool need_char_q = do_obs || do_reward;
    const int tile_rows = (int)min((long long)TILE, n_rows - row0);

    // ================================ phase A ====================================================
    PARC_COLD_LOOP
    for (int k = tid; k < NPART * PARC_MAX_BODIES; k += THREADS) s_part[k] = sp.part_entry[k / PARC_MAX_BODIES][k % PARC_MAX_BODIES];
    PARC_COLD_LOOP
    for (int k = tid; k < 8 * J; k += THREADS) {
        const int j = k >> 3, c = k & 7;
        s_jc[k] = c < 3 ? cm.local_trans[j][c] : (c < 7 ? cm.local_rot[j][c - 3] : 0.f);
    }
    if (do_heights)
        PARC_COLD_LOOP
        for (int k = tid; k < P; k += THREADS) s_pts[k] = __ldg(reinterpret_cast<const float2*>(env.ray_xy_points) + k);
    // A.3 per-env uniforms: the first lanes of the LAST warp, so that the dependent loads (env id -> motion id -> clip
    //     meta) overlap the dof -> quaternion work of the leading warps
#if PARC_SPLIT_A
    // two chains in two otherwise idle warps: root state -> heading -> sin/cos/h^-1 (last warp), and time -> motion id ->
    // clip meta -> terrain offset (the warp before it); each writes its own fields of s_env[el]
    if (tid >= THREADS - 32 && tid < THREADS - 32 + TILE) {
        const int el = tid - (THREADS - 32);
        EnvU& u = s_env[el];
        const int live = el < tile_rows;
        const long long row = row0 + (live ? el : 0);
        const long long e = env_ids ? env_ids[row] : (env_ids32 ? (long long)env_ids32[row] : row + first_env);
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
        u.sh = sinf(heading); u.ch = cosf(heading);
        u.hinv = axis_angle_to_quat(make_vec3(0.f, 0.f, 1.f), -heading);   // calc_heading_quat_inv
        u.vel_rec = lib.vel;
        u.flags = 0;
#if PARC_VEC_ROWS
        u.row_shift = do_obs ? (int)(((reinterpret_cast<unsigned long long>(env.obs_buf) >> 2) + (unsigned long long)(e * env.obs_row_stride)) & 3ull) : 0;
#else
        u.row_shift = 0;
#endif
        if (need_char_q) s_charq[el * J] = quat_to_f4(c_rot);
    } else if (tid >= THREADS - 64 && tid < THREADS - 64 + TILE) {
        const int el = tid - (THREADS - 64);
        EnvU& u = s_env[el];
        const int live = el < tile_rows;
        const long long row = row0 + (live ? el : 0);
        const long long e = env_ids ? env_ids[row] : (env_ids32 ? (long long)env_ids32[row] : row + first_env);
        float time;
        if (parts & PARC_STEP_TIME) {                                   // _update_time, ig_env.py:391-394
            const int ts = env.timestep_buf[e] + 1;
            time = __fmul_rn(cfg.timestep, (float)ts);
            if (live) { env.timestep_buf[e] = ts; env.time_buf[e] = time; }
        } else {
            time = env.time_buf[e];
        }
        u.time = time;
        const long long mid = env.motion_ids[e];
        u.mid = mid;
        const long long tidx = env.motion_terrain_ids[e];
        u.mm = load_motion_meta(lib.meta, mid);
        u.t0 = __fadd_rn(time, env.motion_time_offsets[e]);            // dm_env.py:547-552
        const float* dmo = env.dm_motion_offsets + (mid * env.terrains_per_motion + tidx) * 2;   // dm_env.py:554-565
        u.tox = __fsub_rn(__ldg(dmo), __ldg(env.env_offsets + e * 3)); u.toy = __fsub_rn(__ldg(dmo + 1), __ldg(env.env_offsets + e * 3 + 1));
    }
#else
    if (tid >= THREADS - 32 && tid < THREADS - 32 + TILE) {
        EnvU u;
        const int el = tid - (THREADS - 32);
        u.live = el < tile_rows;
        const long long row = row0 + (u.live ? el : 0);
        const long long e = env_ids ? env_ids[row] : (env_ids32 ? (long long)env_ids32[row] : row + first_env);
        u.e = e;
        u.c_pos = view_ld_vec3(env.char_root_pos, e);
        u.c_rot = view_ld_quat(env.char_root_rot, e);
        u.c_vel = view_ld_vec3(env.char_root_vel, e);
        u.c_avel = view_ld_vec3(env.char_root_ang_vel, e);
        u.eo = make_vec3(__ldg(env.env_offsets + e * 3), __ldg(env.env_offsets + e * 3 + 1), __ldg(env.env_offsets + e * 3 + 2));
        u.g = u.c_pos + u.eo;                                          // ig_parkour_env.py:522
        const float heading = calc_heading(u.c_rot);
        u.sh = sinf(heading); u.ch = cosf(heading);
        u.hinv = axis_angle_to_quat(make_vec3(0.f, 0.f, 1.f), -heading);   // calc_heading_quat_inv
        if (parts & PARC_STEP_TIME) {                                   // _update_time, ig_env.py:391-394
            const int ts = env.timestep_buf[e] + 1;
            u.time = __fmul_rn(cfg.timestep, (float)ts);
            if (u.live) { env.timestep_buf[e] = ts; env.time_buf[e] = u.time; }
        } else {
            u.time = env.time_buf[e];
        }
        u.mid = env.motion_ids[e];
        const long long tidx = env.motion_terrain_ids[e];
        u.mm = load_motion_meta(lib.meta, u.mid);
        u.t0 = __fadd_rn(u.time, env.motion_time_offsets[e]);          // dm_env.py:547-552
        const float* dmo = env.dm_motion_offsets + (u.mid * env.terrains_per_motion + tidx) * 2;   // dm_env.py:554-565
        u.tox = __fsub_rn(__ldg(dmo), u.eo.x); u.toy = __fsub_rn(__ldg(dmo + 1), u.eo.y);
        u.vel_rec = lib.vel;
        u.flags = 0;
#if PARC_VEC_ROWS
        u.row_shift = do_obs ? (int)(((reinterpret_cast<unsigned long long>(env.obs_buf) >> 2) + (unsigned long long)(e * env.obs_row_stride)) & 3ull) : 0;
#else
        u.row_shift = 0;
#endif
        s_env[el] = u;
        if (need_char_q) s_charq[el * J] = quat_to_f4(u.c_rot);
    }
#endif
    // A.1 character joint rotations: KinCharModel.dof_to_rot (kin_char_model.py:586-599), one thread per (env, joint)
    if (need_char_q) {
        PARC_COLD_LOOP
        for (int it = tid; it < TILE * (J - 1); it += THREADS) {
            const int el = it / (J - 1), j = 1 + (it - el * (J - 1));
            if (el >= tile_rows) continue;
            const long long row = row0 + el;
            const long long e = env_ids ? env_ids[row] : (env_ids32 ? (long long)env_ids32[row] : row + first_env);
            const int jt = cm.joint_type[j], di = cm.dof_idx[j];
            Quat q = quat_identity();
            if (jt == PARC_JOINT_SPHERICAL)
                q = exp_map_to_quat(make_vec3(view_ld(env.char_dof_pos, e, di), view_ld(env.char_dof_pos, e, di + 1), view_ld(env.char_dof_pos, e, di + 2)));
            else if (jt == PARC_JOINT_HINGE)
                q = axis_angle_to_quat(make_vec3(cm.axis[j][0], cm.axis[j][1], cm.axis[j][2]), view_ld(env.char_dof_pos, e, di));
            s_charq[el * J + j] = quat_to_f4(q);
        }
    }
    __syncthreads();

    // A.0b warm the frame records this tile is about to interpolate: one thread per (env, library pose) asks for
    //      the two adjacent interp records (and the velocity record of the reference pose); the dof -> quaternion
    //      work below runs while they arrive
    if (do_reward || do_done) {
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
    if (do_obs || need_ref_pose) {
        PARC_COLD_LOOP
        for (int it = tid; it < TILE * (S + 1); it += THREADS) {
            const int el = it / (S + 1), pose = it - el * (S + 1);
            if (el >= tile_rows) continue;
            const EnvU& u = s_env[el];
            const float t = pose == 0 ? u.t0 : __fadd_rn(u.t0, cfg.tar_dt[pose - 1]);
            const FrameBlend fb = calc_frame_blend(u.mm, t);
            const char* r0 = reinterpret_cast<const char*>(lib.interp + fb.idx0 * lib.irec);
            const int bytes = (int)((fb.idx1 - fb.idx0 + 1) * lib.irec * sizeof(float));
            PARC_COLD_LOOP
            for (int off = 0; off < bytes; off += 128) asm volatile("prefetch.global.L1 [%0];" ::"l"(r0 + off));
            if (pose == 0) {
                const char* v = reinterpret_cast<const char*>(lib.vel + fb.idx0 * lib.vrec);
                PARC_COLD_LOOP
                for (int off = 0; off < (int)(lib.vrec * sizeof(float)); off += 128) asm volatile("prefetch.global.L1 [%0];" ::"l"(v + off));
            }
        }
    }

    // ================================ phase B ====================================================
    static_assert(TILE % 8 == 0 && NPART == 4, "one warp = one pose of 8 of the tile's envs x NPART tree parts");
    const bool pose_work = do_obs || need_ref_pose;
    // ---- B.1 heights: _refresh_obs_hfs (ig_parkour_env.py:518-533, mgdm_dm_util.py:128-145) for the sample points [p0, p1):
    //      warp w samples the points of env w. Where this runs is a scheduling choice (PARC_HEIGHTS_LATE): the
    //      samples depend only on the phase A uniforms and write straight to global memory.
    auto sample_heights = [&](const int p0, const int p1) {
        if (!(P > 0 && (do_heights || do_obs)) || p1 <= p0) return;
        const int lane = tid & 31, cnt = p1 - p0;
        const float2 nmn = make_float2(-terrain.min_x, -terrain.min_y), rd2 = make_float2(sp.rdx, sp.rdy), nd2 = make_float2(-terrain.dx, -terrain.dy);
        const int imx = terrain.dim_x - 1, imy = terrain.dim_y - 1, dimy = terrain.dim_y;
        const float lo = cfg.min_obs_h, hi = cfg.max_obs_h;
        for (int el = tid >> 5; el < tile_rows; el += THREADS / 32) {
            const EnvU& u = s_env[el];
            float* orow = do_obs ? env.obs_buf + u.e * env.obs_row_stride + L.o_hf + p0 : nullptr;
            float* hrow = env.ray_hfs ? env.ray_hfs + u.e * P + p0 : nullptr;
            if (do_heights) {
                const float2 cs = make_float2(u.ch, u.sh), nsc = make_float2(-u.sh, u.ch), g2 = make_float2(u.g.x, u.g.y);
                if (cnt % (32 * PARC_HEIGHT_GATHERS) == 0)
                    heights_of_env<true>(s_pts + p0, cnt, lane, terrain.hf, dimy, imx, imy, cs, nsc, g2, u.g.z, nmn, rd2, nd2, lo, hi, hrow, orow);
                else
                    heights_of_env<false>(s_pts + p0, cnt, lane, terrain.hf, dimy, imx, imy, cs, nsc, g2, u.g.z, nmn, rd2, nd2, lo, hi, hrow, orow);
            } else if (orow) {
                for (int p = lane; p < cnt; p += 32) orow[p] = hrow[p];
            }
        }
    };
#if !PARC_HEIGHTS_LATE
    sample_heights(0, P);
#endif
    // ---- B.2 warp w takes poses w, w + 8, ...; lane = (env, part): NPART lanes share one (env, pose) and each
    //      walks its own ancestor-closed part of the joint tree, so all lanes of a warp run the same code ----
    for (int pose = (tid >> 5) & 7; pose_work && pose < plan.n_pose; pose += 8) {
        const int lane = tid & 31, el = (tid >> 8) * 8 + lane / NPART, part = lane & (NPART - 1);
        const bool is_ref = pose == 0, is_char = pose == S + 1, is_tar = pose >= 1 && pose <= S;
        const bool wanted = (is_ref && need_ref_pose) || (is_tar && do_obs) || (is_char && do_obs);
        if (!wanted || !s_env[el].live) continue;
        const EnvU& u = s_env[el];
        float* row = s_row + el * ROWP + u.row_shift;
        float* rsec = row + L.o_hf;                    // ref_* section
        const float *r0 = nullptr, *r1 = nullptr;
        float b = 0.f, omb = 1.f;
        Quat rq = u.c_rot;
        Vec3 rpos = u.c_pos;
        const float* vrec = nullptr;
        if (!is_char) {
            // MotionLib.calc_motion_frame (motion_lib.py:94-126); fetch_tar_obs_data (mgdm_dm_util.py:221-244)
            const float t = is_ref ? u.t0 : __fadd_rn(u.t0, cfg.tar_dt[pose - 1]);
            const FrameBlend fb = calc_frame_blend(u.mm, t);
            r0 = lib.interp + fb.idx0 * lib.irec;
            r1 = lib.interp + fb.idx1 * lib.irec;
            b = fb.blend; omb = __fsub_rn(1.0f, b);
            rq = slerp(quat_from_f4(ldg4(r0)), quat_from_f4(ldg4(r1)), b);
            const float4 p0 = ldg4(r0 + 4 * J), p1 = ldg4(r1 + 4 * J);
            rpos = make_vec3(__fadd_rn(__fadd_rn(lerp_exact(omb, p0.x, b, p1.x), fb.loop_offset.x), u.tox),
                             __fadd_rn(__fadd_rn(lerp_exact(omb, p0.y, b, p1.y), fb.loop_offset.y), u.toy),
                             __fadd_rn(lerp_exact(omb, p0.z, b, p1.z), fb.loop_offset.z));
            vrec = lib.vel + fb.idx0 * lib.vrec;
        }
        // ---- root-level outputs (scalar pieces by part 0, the per-dof / per-body lists strided over the parts) ----
        float* o_pose = nullptr;        // start of this pose's joint tan-norm block
        float* o_key = nullptr;         // start of this pose's key-body block
        Vec3 rp_obs = make_vec3(0.f, 0.f, 0.f);
        if (is_ref) {
            if (part == 0) {
                rsec[R.root_pos] = rpos.x; rsec[R.root_pos + 1] = rpos.y; rsec[R.root_pos + 2] = rpos.z;
                rsec[R.root_rot] = rq.x; rsec[R.root_rot + 1] = rq.y; rsec[R.root_rot + 2] = rq.z; rsec[R.root_rot + 3] = rq.w;
                rsec[R.body_pos] = rpos.x; rsec[R.body_pos + 1] = rpos.y; rsec[R.body_pos + 2] = rpos.z;
            }
            // velocities come from frame idx0 without interpolation (motion_lib.py:102-110)
            PARC_COLD_LOOP
            for (int k = 4 * part; k < 6 + D; k += 4 * NPART) {
                const float4 x = ldg4(vrec + k);
                const float xs[4] = {x.x, x.y, x.z, x.w};
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    const int kk = k + c;
                    if (kk < 3) rsec[R.root_vel + kk] = xs[c];
                    else if (kk < 6) rsec[R.root_ang_vel + kk - 3] = xs[c];
                    else if (kk < 6 + D) rsec[R.dof_vel + kk - 6] = xs[c];
                }
            }
        } else if (is_tar) {
            // compute_tar_obs (mgdm_dm_util.py:404-461) for target step pose-1
            float* o = row + L.o_tar + (pose - 1) * L.tar_step_dim;
            rp_obs = rpos - u.c_pos;
            Quat oq = rq;
            if (!cfg.global_obs) { rp_obs = quat_rotate(u.hinv, rp_obs); oq = quat_mul(u.hinv, rq); }
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
                if (!cfg.global_obs) { oq = quat_mul(u.hinv, u.c_rot); v = quat_rotate(u.hinv, u.c_vel); w = quat_rotate(u.hinv, u.c_avel); }
                tan_norm_exact(oq, o);
                o[6] = v.x; o[7] = v.y; o[8] = v.z; o[9] = w.x; o[10] = w.y; o[11] = w.z;
            }
            o_pose = o + 12; o_key = o + 12 + 6 * (J - 1) + D;
            PARC_COLD_LOOP
            for (int k = part; k < D; k += NPART) o[12 + 6 * (J - 1) + k] = view_ld(env.char_dof_vel, u.e, k);
            if (cfg.use_contact_info) {
                // _get_char_contact_state (ig_parkour_env.py:655-662): ||f|| > eps -> 0/1
                PARC_COLD_LOOP
                for (int c = part; c < J; c += NPART) {
                    const float fx = view_ld3(env.char_contact_forces, u.e, c, 0), fy = view_ld3(env.char_contact_forces, u.e, c, 1),
                                fz = view_ld3(env.char_contact_forces, u.e, c, 2);
                    row[L.o_charc + c] = (sqrtf(fx * fx + fy * fy + fz * fz) > cfg.contact_eps) ? 1.f : 0.f;
                }
            }
        }
        // contact labels: lerp of the two records (motion_lib.py:121)
        if (!is_char && C > 0 && (is_ref || cfg.use_contact_info)) {
            float* oc = is_ref ? (rsec + R.contacts) : (row + L.o_tarc + (pose - 1) * C);
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
        // ---- joints of this part: slerp + forward kinematics (kin_char_model.py:617-649) ----
        const bool lrot_ident = sp.lrot_ident != 0;
        float* xf = s_xf + tid;                         // [slot][7][THREADS], this thread's column
        const int* my_part = s_part + part * PARC_MAX_BODIES;
        PARC_COLD_LOOP
        for (int i = 0; i < sp.max_part_len; ++i) {
            const int entry = my_part[i];
            if (entry < 0) continue;
            const int j = entry & 0xff, owned = (entry >> 8) & 1, st_slot = ((entry >> 12) & 0xf) - 1, p_slot = ((entry >> 16) & 0xf) - 1;
            Quat q;
            if (is_char) q = quat_from_f4(s_charq[el * J + j]);
            else q = slerp(quat_from_f4(ldg4(r0 + 4 * j)), quat_from_f4(ldg4(r1 + 4 * j)), b);
            Vec3 ppos = rpos; Quat prot = rq;
            if (p_slot >= 0) {
                const float* x = xf + p_slot * 7 * THREADS;
                ppos = make_vec3(x[0], x[THREADS], x[2 * THREADS]);
                prot = make_quat(x[3 * THREADS], x[4 * THREADS], x[5 * THREADS], x[6 * THREADS]);
            }
            const float* jc = s_jc + 8 * j;
            const Vec3 bpos = ppos + quat_rotate(prot, make_vec3(jc[0], jc[1], jc[2]));
            if (st_slot >= 0) {
                const Quat cq = lrot_ident ? q : quat_mul(make_quat(jc[3], jc[4], jc[5], jc[6]), q);
                const Quat brot = quat_mul(prot, cq);
                float* x = xf + st_slot * 7 * THREADS;
                x[0] = bpos.x; x[THREADS] = bpos.y; x[2 * THREADS] = bpos.z;
                x[3 * THREADS] = brot.x; x[4 * THREADS] = brot.y; x[5 * THREADS] = brot.z; x[6 * THREADS] = brot.w;
            }
            if (!owned) continue;
            if (is_ref) {
                float* jr = rsec + R.joint_rot + 4 * (j - 1);
                jr[0] = q.x; jr[1] = q.y; jr[2] = q.z; jr[3] = q.w;
                float* bp = rsec + R.body_pos + 3 * j;
                bp[0] = bpos.x; bp[1] = bpos.y; bp[2] = bpos.z;
            } else {
                tan_norm_exact(q, o_pose + 6 * (j - 1));
                if ((sp.key_mask >> j) & 1u)
                PARC_COLD_LOOP
                for (int k = 0; k < K; ++k) {
                    if (cfg.key_body_ids[k] != j) continue;
                    Vec3 kp = bpos - rpos;
                    if (!cfg.global_obs) { kp = quat_rotate(u.hinv, kp); if (is_tar) kp = kp + rp_obs; }
                    float* kd = o_key + 3 * k;
                    kd[0] = kp.x; kd[1] = kp.y; kd[2] = kp.z;
                }
            }
        }
        // a key body may also be the root itself
        if (!is_ref && part == 0 && (sp.key_mask & 1u))
            for (int k = 0; k < K; ++k)
                if (cfg.key_body_ids[k] == 0) {
                    Vec3 kp = make_vec3(0.f, 0.f, 0.f);
                    if (!cfg.global_obs && is_tar) kp = kp + rp_obs;
                    float* kd = o_key + 3 * k;
                    kd[0] = kp.x; kd[1] = kp.y; kd[2] = kp.z;
                }
    }
    __syncthreads();

    // ================================ phase C ====================================================
    // per-(env, item) terms; item lists are flattened over the tile so that every lane has work
    const int o_pose_t = 0, o_vel_t = J - 1, o_con_t = o_vel_t + D, o_key_t = o_con_t + J;
    const int TP = plan.term_pitch;
    // C.0 scatter map of the ref_* section for phase E: column k -> (base pointer, row stride in bytes). Lives in the
    //     bytes the FK columns used during phase B; phase E starts behind the barrier that ends phase C.
    if (do_ref) {
        PARC_COLD_LOOP
        for (int k = tid; k < R.width; k += THREADS) {
            const parc_view3_t* v; int off;
            if (k < R.root_rot) { v = &env.ref_root_pos; off = k - R.root_pos; }
            else if (k < R.root_vel) { v = &env.ref_root_rot; off = k - R.root_rot; }
            else if (k < R.root_ang_vel) { v = &env.ref_root_vel; off = k - R.root_vel; }
            else if (k < R.body_pos) { v = &env.ref_root_ang_vel; off = k - R.root_ang_vel; }
            else if (k < R.joint_rot) { v = &env.ref_body_pos; off = k - R.body_pos; }
            else if (k < R.dof_pos) { v = &env.ref_joint_rot; off = k - R.joint_rot; }
            else if (k < R.dof_vel) { v = &env.ref_dof_pos; off = k - R.dof_pos; }
            else if (k < R.contacts) { v = &env.ref_dof_vel; off = k - R.dof_vel; }
            else { v = &env.ref_contacts; off = k - R.contacts; }
            long long eoff;
            if (k >= R.body_pos && k < R.joint_rot) eoff = (off / 3) * v->mid_stride + (off % 3) * v->elem_stride;
            else if (k >= R.joint_rot && k < R.dof_pos) eoff = (off / 4) * v->mid_stride + (off % 4) * v->elem_stride;
            else eoff = off * v->elem_stride;
            s_refmap[2 * k] = v->ptr ? reinterpret_cast<unsigned long long>(v->ptr + eoff) : 0ull;
            s_refmap[2 * k + 1] = (unsigned long long)(v->row_stride * (long long)sizeof(float));
        }
    }
    {
        // Four item kinds (joints, contacts, key bodies, dof velocities) cut into warp-sized tasks that are dealt
        // round-robin to the 8 warps: every warp runs about the same number of (prefetched) global reads.
        const bool want_c = (do_reward && cfg.use_contact_info) || (do_done && cfg.enable_early_termination && cfg.num_fall_contact_bodies > 0);
        const int nt_j = need_ref_pose ? (TILE * (J - 1) + 31) >> 5 : 0;
        const int nt_c = want_c ? (TILE * J + 31) >> 5 : 0;
        const int nt_k = do_reward ? (TILE * K + 31) >> 5 : 0;
        const int nt_d = do_reward ? (TILE * D + 31) >> 5 : 0;
        const int b_c = nt_j, b_k = b_c + nt_c, b_d = b_k + nt_k, n_task = b_d + nt_d;
        const int lane = tid & 31;
        PARC_COLD_LOOP
        for (int task = tid >> 5; task < n_task; task += THREADS / 32) {
            if (task < b_c) {
                // joints: rot_to_dof of the reference pose (kin_char_model.py:601-615) + pose error term
                const int it = task * 32 + lane;
                if (it >= TILE * (J - 1)) continue;
                const int el = it / (J - 1), j = 1 + (it - el * (J - 1));
                if (el >= tile_rows) continue;
                float* rsec = s_row + el * ROWP + s_env[el].row_shift + L.o_hf;
                const float* jr = rsec + R.joint_rot + 4 * (j - 1);
                const Quat q = make_quat(jr[0], jr[1], jr[2], jr[3]);
                if (do_ref) {
                    const int jt = cm.joint_type[j], di = cm.dof_idx[j];
                    if (jt == PARC_JOINT_SPHERICAL) {
                        const Vec3 ex = quat_to_exp_map(q);
                        rsec[R.dof_pos + di] = ex.x; rsec[R.dof_pos + di + 1] = ex.y; rsec[R.dof_pos + di + 2] = ex.z;
                    } else if (jt == PARC_JOINT_HINGE) {
                        Vec3 axis; float angle;
                        quat_to_axis_angle(q, axis, angle);
                        const float dd = dot3_seq(make_vec3(cm.axis[j][0], cm.axis[j][1], cm.axis[j][2]), axis);
                        rsec[R.dof_pos + di] = (dd < 0.f) ? -angle : angle;
                    }
                }
                if (do_reward) {   // compute_deepmimic_reward pose term (mgdm_dm_util.py:285-287)
                    const float a = quat_diff_angle(quat_from_f4(s_charq[el * J + j]), q);
                    s_terms[el * TP + o_pose_t + (j - 1)] = cfg.joint_err_w[j - 1] * a * a;
                }
                if (do_done && cfg.enable_early_termination && cfg.pose_termination) {
                    // compute_done pose test (mgdm_dm_util.py:372-384): squared norm summed left to right
                    const EnvU& u = s_env[el];
                    const Vec3 sb = make_vec3(view_ld3(env.char_rigid_body_pos, u.e, j, 0), view_ld3(env.char_rigid_body_pos, u.e, j, 1), view_ld3(env.char_rigid_body_pos, u.e, j, 2));
                    const Vec3 s0 = make_vec3(view_ld3(env.char_rigid_body_pos, u.e, 0, 0), view_ld3(env.char_rigid_body_pos, u.e, 0, 1), view_ld3(env.char_rigid_body_pos, u.e, 0, 2));
                    const float* bp = rsec + R.body_pos;
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
                if (do_reward && cfg.use_contact_info) {
                    // compute_contact_reward (mgdm_dm_util.py:497-518); contact bodies == all bodies
                    const float f = fminf(sqrtf(fx * fx + fy * fy + fz * fz), 1.0f);
                    const float rc = s_row[el * ROWP + u.row_shift + L.o_hf + R.contacts + c];
                    float cr = -(1.0f - rc) * f;
                    cr += rc * f;
                    s_terms[el * TP + o_con_t + c] = cfg.contact_weights[c] * cr;
                }
                if (do_done && cfg.enable_early_termination && cfg.num_fall_contact_bodies > 0 && !((cfg.fall_contact_body_mask >> c) & 1u)) {
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
                const float* rsec = s_row + el * ROWP + u.row_shift + L.o_hf;
                const float* bp = rsec + R.body_pos;
                const Vec3 ref_rpos = make_vec3(bp[0], bp[1], bp[2]);
                Vec3 kp = make_vec3(view_ld3(env.char_rigid_body_pos, u.e, body, 0), view_ld3(env.char_rigid_body_pos, u.e, body, 1),
                                    view_ld3(env.char_rigid_body_pos, u.e, body, 2)) - u.c_pos;
                Vec3 tkp = make_vec3(bp[3 * body], bp[3 * body + 1], bp[3 * body + 2]) - ref_rpos;
                if (!cfg.track_root) {   // convert_to_local (mgdm_dm_util.py:246-267)
                    const Quat rr = make_quat(rsec[R.root_rot], rsec[R.root_rot + 1], rsec[R.root_rot + 2], rsec[R.root_rot + 3]);
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
                const float dv = s_row[el * ROWP + u.row_shift + L.o_hf + R.dof_vel + k] - view_ld(env.char_dof_vel, u.e, k);
                s_terms[el * TP + o_vel_t + k] = cfg.dof_err_w[k] * dv * dv;
            }
        }
    }
    __syncthreads();

    // ================================ phase D ====================================================
    if (tid < tile_rows && (do_reward || do_done)) {
        const int el = tid;
        const EnvU& u = s_env[el];
        const long long e = u.e;
        const float* rsec = s_row + el * ROWP + u.row_shift + L.o_hf;
        const Vec3 ref_rpos = make_vec3(rsec[R.root_pos], rsec[R.root_pos + 1], rsec[R.root_pos + 2]);
        const Quat ref_rrot = make_quat(rsec[R.root_rot], rsec[R.root_rot + 1], rsec[R.root_rot + 2], rsec[R.root_rot + 3]);
        if (do_reward) {
            // _update_reward (ig_parkour_env.py:984-1044) + compute_deepmimic_reward (mgdm_dm_util.py:269-332)
            const float* tm = s_terms + el * TP;
            float pose_t = 0.f, vel_t = 0.f, key_t = 0.f, con_t = 0.f;
#pragma unroll 4
            for (int k = 0; k < J - 1; ++k) pose_t += tm[o_pose_t + k];
#pragma unroll 4
            for (int k = 0; k < D; ++k) vel_t += tm[o_vel_t + k];
            for (int k = 0; k < K; ++k) key_t += tm[o_key_t + k];
            Vec3 rpd = ref_rpos - u.c_pos;
            if (!cfg.track_root) { rpd.x = 0.f; rpd.y = 0.f; }
            if (!cfg.track_root_h) rpd.z = 0.f;
            const float root_pos_err = rpd.x * rpd.x + rpd.y * rpd.y + rpd.z * rpd.z;
            Quat a_rot = u.c_rot, b_rot = ref_rrot;
            Vec3 a_vel = u.c_vel, a_av = u.c_avel;
            Vec3 b_vel = make_vec3(rsec[R.root_vel], rsec[R.root_vel + 1], rsec[R.root_vel + 2]);
            Vec3 b_av = make_vec3(rsec[R.root_ang_vel], rsec[R.root_ang_vel + 1], rsec[R.root_ang_vel + 2]);
            if (!cfg.track_root) {   // convert_to_local (mgdm_dm_util.py:246-267)
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
            if (cfg.use_contact_info) {
#pragma unroll 4
                for (int k = 0; k < J; ++k) con_t += tm[o_con_t + k];
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
        if (do_done) {
            // _update_done (ig_parkour_env.py:967-982; mgdm_dm_util.py:147-172, 334-402; dm_env.py:628-665)
            int done = PARC_DONE_NULL;
            if (u.time >= cfg.episode_length) done = PARC_DONE_TIME;
            if (cfg.enable_early_termination) {
                bool failed = false;
                if (cfg.num_fall_contact_bodies > 0) failed = (u.flags & 1) && (u.flags & 2);
                if (cfg.pose_termination) {
                    bool pf = (u.flags & 4) != 0;
                    if (cfg.track_root) {
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
            if (cfg.never_done) done = PARC_DONE_NULL;
            env.done_buf[e] = done;
            if (env.done_code) env.done_code[e] = (unsigned char)code;
            if (code != 0 && env.done_list) {
                const int slot = atomicAdd(env.done_count, 1);
                if (slot < env.num_envs) env.done_list[slot] = (int)e;   // a stale (un-zeroed) count must not overrun the list
            }
        }
    }

#if PARC_HEIGHTS_LATE
    // heights after the reward / termination phases: nothing waits for them at a barrier, and they fill the issue slots
    // while the 16 lanes of phase D run their serial sums
    sample_heights(0, P);
#endif
    // ================================ phase E ====================================================
    // staged rows -> global: warp w streams env w; consecutive lanes write consecutive columns
    {
        const int warp = tid >> 5, lane = tid & 31;
        for (int el = warp; el < tile_rows; el += THREADS / 32) {
            const EnvU& u = s_env[el];
            const float* row = s_row + el * ROWP + u.row_shift;
            if (do_obs) {
                float* g = env.obs_buf + u.e * env.obs_row_stride;
#if PARC_VEC_ROWS
                // column k is 16-byte aligned in both the staged row and the obs_buf row when (row_shift + k) % 4 == 0
                const int n = L.o_hf, head = min((4 - u.row_shift) & 3, n), nvec = (n - head) >> 2, t0 = head + 4 * nvec;
                if (lane < head) g[lane] = row[lane];
                const float4* r4 = reinterpret_cast<const float4*>(row + head);
                float4* g4 = reinterpret_cast<float4*>(g + head);
#pragma unroll 2
                for (int k = lane; k < nvec; k += 32) g4[k] = r4[k];
                if (t0 + lane < n) g[t0 + lane] = row[t0 + lane];
#else
#pragma unroll 4
                for (int k = lane; k < L.o_hf; k += 32) g[k] = row[k];
#endif
            }
            if (do_ref) {
                const float* rsec = row + L.o_hf;
#pragma unroll 2
                for (int k = lane; k < R.width; k += 32) {
                    const unsigned long long base = s_refmap[2 * k];
                    if (base) *reinterpret_cast<float*>(base + (unsigned long long)u.e * s_refmap[2 * k + 1]) = rsec[k];
                }
            }
        }
    }
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

extern "C" int parc_b200_tree_parts(const parc_char_model_t* m, int32_t* entries, int32_t* max_len, int32_t* num_slots) {
    if (!model_ok(m, "tree_parts")) return PARC_ERR_INVALID_ARG;
    PARC_REQUIRE(entries && max_len && num_slots, "tree_parts: NULL output");
    parc_env_config_t cfg = {};
    parc_terrain_t tr = {};
    tr.dx = tr.dy = 1.f; tr.dim_x = tr.dim_y = 1;
    const StepPlan sp = make_step_plan(*m, 0, cfg, tr);
    for (int p = 0; p < NPART; ++p)
        for (int i = 0; i < PARC_MAX_BODIES; ++i) entries[p * PARC_MAX_BODIES + i] = sp.part_entry[p][i];
    *max_len = sp.max_part_len;
    *num_slots = sp.S.n_slots;
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
    const StepPlan sp = make_step_plan(*m, lib->num_contacts, *cfg, *terrain);
    const SmemPlan& plan = sp.S;
    static size_t smem_attr = 0;   // opt-in dynamic shared memory, raised only when a larger plan shows up
    if (plan.total > smem_attr) {
        cudaError_t ce = cudaFuncSetAttribute(env_step_kernel<STEP_THREADS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)plan.total);
        if (ce != cudaSuccess) { set_error("env_step: %s", cudaGetErrorString(ce)); return PARC_ERR_CUDA; }
        smem_attr = plan.total;
    }
    const long long blocks = (max_rows + TILE - 1) / TILE;
    env_step_kernel<STEP_THREADS><<<(unsigned)blocks, STEP_THREADS, plan.total, as_stream(stream)>>>(
        *m, *lib, *terrain, *cfg, *env, sp, parts, (const long long*)env_ids, max_rows, env_ids32, n_rows_dev, first_env);
    return check_launch("env_step");
}
}  // namespace parc
