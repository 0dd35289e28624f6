// extern "C" entry points of env_aux_kernels.cu (included at its end; kept separate for size).
#pragma once

using namespace parc;

#define G_DISPATCH(kernel, n_groups, ...)                                                              \
    do {                                                                                               \
        if (m->num_bodies <= 16) kernel<16><<<blocks_for((n_groups), 16), 256, 0, as_stream(stream)>>>(__VA_ARGS__); \
        else kernel<32><<<blocks_for((n_groups), 8), 256, 0, as_stream(stream)>>>(__VA_ARGS__);         \
    } while (0)

extern "C" int parc_b200_compute_tar_obs_data(const parc_char_model_t* m, const parc_motion_lib_t* lib, const parc_env_config_t* cfg,
                                              const parc_env_buffers_t* env, const int64_t* env_ids, int64_t n, float* tar_root_pos,
                                              float* tar_root_rot, float* tar_joint_rot, float* tar_key_pos, float* tar_contacts, void* stream) {
    if (!model_ok(m, "compute_tar_obs_data")) return PARC_ERR_INVALID_ARG;
    PARC_REQUIRE(lib && lib->interp && cfg && env, "compute_tar_obs_data: NULL descriptor");
    PARC_REQUIRE(cfg->num_tar_steps > 0 && cfg->num_tar_steps <= PARC_MAX_TAR_STEPS, "compute_tar_obs_data: num_steps must be in [1,%d]", PARC_MAX_TAR_STEPS);
    PARC_REQUIRE(n >= 0, "compute_tar_obs_data: n < 0");
    if (n == 0) return PARC_OK;
    PARC_REQUIRE(tar_root_pos && tar_root_rot && (tar_joint_rot || m->num_bodies == 1), "compute_tar_obs_data: NULL output");
    PARC_REQUIRE(tar_contacts == nullptr || lib->num_contacts > 0, "compute_tar_obs_data: library has no contact info");
    G_DISPATCH(tar_obs_data_kernel, n * cfg->num_tar_steps, *m, *lib, *cfg, *env, (const long long*)env_ids, n, tar_root_pos, tar_root_rot,
               tar_joint_rot, tar_key_pos, tar_contacts);
    return check_launch("compute_tar_obs_data");
}

extern "C" int parc_b200_compute_char_obs(const parc_char_model_t* m, const float* root_pos, const float* root_rot, const float* root_vel,
                                          const float* root_ang_vel, const float* joint_rot, const float* dof_vel, const float* key_pos,
                                          int32_t num_key, int32_t global_obs, int32_t root_height_obs, int64_t n, float* obs, void* stream) {
    if (!model_ok(m, "compute_char_obs")) return PARC_ERR_INVALID_ARG;
    if (n == 0) return PARC_OK;
    PARC_REQUIRE(root_pos && root_rot && root_vel && root_ang_vel && dof_vel && obs, "compute_char_obs: NULL pointer");
    PARC_REQUIRE(num_key == 0 || key_pos, "compute_char_obs: key_pos is NULL");
    G_DISPATCH(char_obs_kernel, n, *m, root_pos, root_rot, root_vel, root_ang_vel, joint_rot, dof_vel, key_pos, num_key, global_obs,
               root_height_obs ? 1 : 0, n, obs);
    return check_launch("compute_char_obs");
}

extern "C" int parc_b200_compute_tar_obs(const parc_char_model_t* m, const float* ref_root_pos, const float* ref_root_rot,
                                         const float* tar_root_pos, const float* tar_root_rot, const float* tar_joint_rot,
                                         const float* tar_key_pos, int32_t num_steps, int32_t num_key, int32_t global_obs,
                                         int32_t global_tar_root_h_obs, int64_t n, float* obs, void* stream) {
    if (!model_ok(m, "compute_tar_obs")) return PARC_ERR_INVALID_ARG;
    if (n == 0 || num_steps == 0) return PARC_OK;
    PARC_REQUIRE(ref_root_pos && ref_root_rot && tar_root_pos && tar_root_rot && obs, "compute_tar_obs: NULL pointer");
    PARC_REQUIRE(num_key == 0 || tar_key_pos, "compute_tar_obs: tar_key_pos is NULL");
    G_DISPATCH(tar_obs_kernel, n * num_steps, *m, ref_root_pos, ref_root_rot, tar_root_pos, tar_root_rot, tar_joint_rot, tar_key_pos,
               num_steps, num_key, global_obs, global_tar_root_h_obs, n, obs);
    return check_launch("compute_tar_obs");
}

extern "C" int parc_b200_compute_deepmimic_reward(const parc_char_model_t* m, const float* root_pos, const float* root_rot,
                                                  const float* root_vel, const float* root_ang_vel, const float* joint_rot,
                                                  const float* dof_vel, const float* key_pos, const float* tar_root_pos,
                                                  const float* tar_root_rot, const float* tar_root_vel, const float* tar_root_ang_vel,
                                                  const float* tar_joint_rot, const float* tar_dof_vel, const float* tar_key_pos,
                                                  const float* joint_rot_err_w, const float* dof_err_w, int32_t num_key,
                                                  int32_t track_root_h, int32_t track_root, int64_t n, float* reward5, void* stream) {
    if (!model_ok(m, "compute_deepmimic_reward")) return PARC_ERR_INVALID_ARG;
    if (n == 0) return PARC_OK;
    PARC_REQUIRE(root_pos && root_rot && root_vel && root_ang_vel && dof_vel && tar_root_pos && tar_root_rot && tar_root_vel &&
                     tar_root_ang_vel && tar_dof_vel && joint_rot_err_w && dof_err_w && reward5,
                 "compute_deepmimic_reward: NULL pointer");
    dm_reward_kernel<<<blocks_for(n, 8), 256, 0, as_stream(stream)>>>(m->num_bodies, m->dof_size, num_key, root_pos, root_rot, root_vel,
                                                                      root_ang_vel, joint_rot, dof_vel, key_pos, tar_root_pos, tar_root_rot,
                                                                      tar_root_vel, tar_root_ang_vel, tar_joint_rot, tar_dof_vel, tar_key_pos,
                                                                      joint_rot_err_w, dof_err_w, track_root_h, track_root, n, reward5);
    return check_launch("compute_deepmimic_reward");
}

extern "C" int parc_b200_compute_contact_reward(const float* tar_contacts, const float* contact_forces, const float* contact_weights,
                                                int32_t num_contacts, int64_t n, float* out, void* stream) {
    if (n == 0 || num_contacts == 0) return PARC_OK;
    PARC_REQUIRE(tar_contacts && contact_forces && contact_weights && out, "compute_contact_reward: NULL pointer");
    contact_reward_kernel<<<blocks_for(n * num_contacts, 256), 256, 0, as_stream(stream)>>>(tar_contacts, contact_forces, contact_weights,
                                                                                          num_contacts, n, out);
    return check_launch("compute_contact_reward");
}

extern "C" int parc_b200_compute_done(const int32_t* done_buf_in, const float* time, float ep_len, const float* root_rot,
                                      const float* body_pos, const float* tar_root_rot, const float* tar_body_pos,
                                      const float* contact_force, const int64_t* contact_body_ids, int32_t n_cb,
                                      const float* termination_heights, int32_t pose_termination, const float* pose_termination_dist,
                                      int32_t enable_early_termination, int32_t track_root, double root_pos_termination_dist,
                                      float root_rot_termination_angle, int32_t num_bodies, int64_t n, int32_t* done_out, void* stream) {
    (void)done_buf_in;   // the reference only uses it for shape/dtype (torch.full_like)
    if (n == 0) return PARC_OK;
    PARC_REQUIRE(time && done_out, "compute_done: NULL pointer");
    PARC_REQUIRE(num_bodies >= 1 && num_bodies <= PARC_MAX_BODIES, "compute_done: num_bodies out of range");
    if (enable_early_termination) {
        PARC_REQUIRE(n_cb == 0 || (contact_force && contact_body_ids && termination_heights && body_pos), "compute_done: fall test inputs missing");
        PARC_REQUIRE(!pose_termination || (body_pos && tar_body_pos && pose_termination_dist), "compute_done: pose termination inputs missing");
        PARC_REQUIRE(!(pose_termination && track_root) || (root_rot && tar_root_rot), "compute_done: root rotations missing");
    }
    // the reference squares the python float in double and compares the fp32 tensor with the rounded product
    const float root_pos_dist_sq = (float)(root_pos_termination_dist * root_pos_termination_dist);
    compute_done_kernel<<<blocks_for(n, 8), 256, 0, as_stream(stream)>>>(time, ep_len, root_rot, body_pos, tar_root_rot, tar_body_pos, contact_force,
                                                                         (const long long*)contact_body_ids, n_cb, termination_heights,
                                                                         pose_termination, pose_termination_dist, enable_early_termination,
                                                                         track_root, root_pos_dist_sq, root_rot_termination_angle,
                                                                         num_bodies, n, done_out);
    return check_launch("compute_done");
}

extern "C" int parc_b200_compute_tracking_error(const float* root_pos, const float* root_rot, const float* body_rot, const float* body_pos,
                                                const float* tar_root_pos, const float* tar_root_rot, const float* tar_body_rot,
                                                const float* tar_body_pos, const float* root_vel, const float* root_ang_vel,
                                                const float* dof_vel, const float* tar_root_vel, const float* tar_root_ang_vel,
                                                const float* tar_dof_vel, int32_t num_bodies, int32_t dof_size, int64_t n, float* out,
                                                void* stream) {
    if (n == 0) return PARC_OK;
    PARC_REQUIRE(num_bodies >= 1 && num_bodies <= PARC_MAX_BODIES && dof_size >= 0 && dof_size <= PARC_MAX_DOFS,
                 "compute_tracking_error: sizes out of range");
    PARC_REQUIRE(root_pos && root_rot && body_rot && body_pos && tar_root_pos && tar_root_rot && tar_body_rot && tar_body_pos && root_vel &&
                     root_ang_vel && tar_root_vel && tar_root_ang_vel && out && (dof_size == 0 || (dof_vel && tar_dof_vel)),
                 "compute_tracking_error: NULL pointer");
    tracking_error_kernel<<<blocks_for(n, 8), 256, 0, as_stream(stream)>>>(num_bodies, dof_size, root_pos, root_rot, body_rot, body_pos,
                                                                           tar_root_pos, tar_root_rot, tar_body_rot, tar_body_pos, root_vel,
                                                                           root_ang_vel, dof_vel, tar_root_vel, tar_root_ang_vel,
                                                                           tar_dof_vel, n, out);
    return check_launch("compute_tracking_error");
}

extern "C" int parc_b200_fail_rate_ema(const parc_env_config_t* cfg, const parc_env_buffers_t* env, int32_t num_motions, float* cdf,
                                       int32_t* scratch, void* stream) {
    PARC_REQUIRE(cfg && env && num_motions > 0, "fail_rate_ema: bad arguments");
    PARC_REQUIRE(env->done_count && env->done_list && env->done_code && env->motion_ids && env->motion_id_fail_rates && scratch,
                 "fail_rate_ema: done/fail-rate buffers not set");
    PARC_REQUIRE(cdf == nullptr || env->motion_weights, "fail_rate_ema: motion_weights not set");
    PARC_REQUIRE(env->num_envs < (1ll << 30), "fail_rate_ema: more than 2^30 envs");
    // shared-memory staging: the two per-motion tables when they fit in 64 KB, then as much bucket as is left of 200 KB
    const int tables = (size_t)num_motions * 8 <= 64 * 1024;
    const size_t table_bytes = tables ? (size_t)num_motions * 8 : 0;
    const long long cap = std::min<long long>(env->num_envs, (long long)((200 * 1024 - table_bytes) / 4));
    const size_t smem = table_bytes + (size_t)cap * 4;
    static size_t smem_set_of_device[64] = {};             // the opt-in is a per-device function attribute
    int dev = 0;
    cudaGetDevice(&dev);
    size_t& smem_set = smem_set_of_device[dev & 63];
    if (smem > smem_set) {
        cudaError_t ce = cudaFuncSetAttribute(fail_rate_ema_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(200 * 1024));
        if (ce != cudaSuccess) { set_error("fail_rate_ema: %s", cudaGetErrorString(ce)); return PARC_ERR_CUDA; }
        smem_set = 200 * 1024;
    }
    fail_rate_ema_kernel<<<1, EMA_THREADS, smem, as_stream(stream)>>>(*cfg, *env, num_motions, cdf, scratch, tables, (int)cap);
    return check_launch("fail_rate_ema");
}

extern "C" int parc_b200_outcome_counts(const parc_env_buffers_t* env, int32_t num_motions, int64_t* counts, void* stream) {
    PARC_REQUIRE(env && num_motions > 0 && counts, "outcome_counts: bad arguments");
    PARC_REQUIRE(env->done_count && env->done_list && env->done_code && env->motion_ids, "outcome_counts: done buffers not set");
    const long long blocks = std::min<long long>((env->num_envs + 255) / 256, (long long)kNumSMs * 4);
    outcome_counts_kernel<<<(unsigned)std::max<long long>(blocks, 1), 256, 0, as_stream(stream)>>>(*env, num_motions,
                                                                                                   reinterpret_cast<unsigned long long*>(counts));
    return check_launch("outcome_counts");
}

static int check_reset_views(const parc_env_buffers_t* env, const char* who) {
    PARC_REQUIRE(env->char_root_pos.ptr && env->char_root_rot.ptr && env->char_root_vel.ptr && env->char_root_ang_vel.ptr &&
                     env->char_dof_pos.ptr && env->char_dof_vel.ptr && env->ref_root_pos.ptr && env->ref_root_rot.ptr &&
                     env->ref_root_vel.ptr && env->ref_root_ang_vel.ptr && env->ref_joint_rot.ptr && env->ref_dof_pos.ptr &&
                     env->ref_dof_vel.ptr,
                 "%s: state views not set", who);
    PARC_REQUIRE(env->motion_ids && env->motion_terrain_ids && env->motion_time_offsets && env->timestep_buf && env->time_buf &&
                     env->done_buf && env->env_offsets && env->dm_motion_offsets,
                 "%s: bookkeeping buffers not set", who);
    return PARC_OK;
}

extern "C" int parc_b200_env_reset(const parc_char_model_t* m, const parc_motion_lib_t* lib, const parc_env_config_t* cfg,
                                   const parc_env_buffers_t* env, const int64_t* env_ids, int64_t n, const int64_t* motion_ids,
                                   const int64_t* terrain_ids, const float* motion_times, const float* xy_noise,
                                   const float* root_pos_offset, const float* root_rot_offset, void* stream) {
    if (!model_ok(m, "env_reset")) return PARC_ERR_INVALID_ARG;
    PARC_REQUIRE(lib && lib->interp && cfg && env, "env_reset: NULL descriptor");
    PARC_REQUIRE(n >= 0, "env_reset: n < 0");
    if (n == 0) return PARC_OK;
    PARC_REQUIRE(env_ids && motion_ids && terrain_ids && motion_times, "env_reset: NULL draw arrays");
    int rc = check_reset_views(env, "env_reset");
    if (rc) return rc;
    G_DISPATCH(env_reset_kernel, n, *m, *lib, *cfg, *env, (const long long*)env_ids, n, (const long long*)motion_ids,
               (const long long*)terrain_ids, motion_times, xy_noise, root_pos_offset, root_rot_offset);
    return check_launch("env_reset");
}

extern "C" int parc_b200_env_reset_sampled(const parc_char_model_t* m, const parc_motion_lib_t* lib, const parc_terrain_t* terrain,
                                           const parc_env_config_t* cfg, const parc_env_buffers_t* env, const float* cdf,
                                           int32_t num_motions, uint64_t seed, uint64_t step, int32_t rand_reset,
                                           const parc_reset_options_t* options, void* stream) {
    if (!model_ok(m, "env_reset_sampled")) return PARC_ERR_INVALID_ARG;
    parc_reset_options_t opt = {};
    if (options) opt = *options;
    PARC_REQUIRE(lib && lib->interp && cfg && env && terrain && cdf && num_motions > 0, "env_reset_sampled: NULL descriptor");
    PARC_REQUIRE(env->done_count && env->done_list, "env_reset_sampled: done list not set");
    int rc = check_reset_views(env, "env_reset_sampled");
    if (rc) return rc;
    if (env->num_envs == 0) return PARC_OK;
    // sized for the worst case (every env finished); groups beyond the device-side count exit at once
    G_DISPATCH(env_reset_sampled_kernel, env->num_envs, *m, *lib, *cfg, *env, cdf, num_motions, (unsigned long long)seed,
               (unsigned long long)step, rand_reset, opt);
    rc = check_launch("env_reset_sampled(reset)");
    if (rc) return rc;
    // IGParkourEnv.reset (ig_parkour_env.py:809-829): heights + observation of the reset envs
    if (env->obs_buf) {
        rc = launch_env_step(m, lib, terrain, cfg, env, PARC_STEP_HEIGHTS | PARC_STEP_OBS, nullptr, env->num_envs, env->done_list,
                             env->done_count, stream);
        if (rc) return rc;
    }
    zero_int_kernel<<<1, 1, 0, as_stream(stream)>>>(env->done_count);
    return check_launch("env_reset_sampled(clear)");
}

extern "C" int parc_b200_sample_motions_cdf(const float* cdf, int32_t num_motions, const float* u, int64_t n, int64_t* ids, void* stream) {
    if (n == 0) return PARC_OK;
    PARC_REQUIRE(cdf && u && ids && num_motions > 0, "sample_motions_cdf: bad arguments");
    sample_cdf_kernel<<<blocks_for(n, 256), 256, 0, as_stream(stream)>>>(cdf, num_motions, u, n, (long long*)ids);
    return check_launch("sample_motions_cdf");
}

namespace parc {
int launch_sim_standin(const parc_char_model_t* m, const parc_env_buffers_t* env, const float* noise, long long first_env,
                       long long n_rows, void* stream) {
    const long long groups_per_block = PARC_STANDIN_THREADS / (m->num_bodies <= 16 ? 16 : 32);
    const unsigned blocks = (unsigned)std::min<long long>((n_rows + groups_per_block - 1) / groups_per_block, kNumSMs * 32);
    if (m->num_bodies <= 16) sim_standin_kernel<16><<<blocks, PARC_STANDIN_THREADS, 0, as_stream(stream)>>>(*m, *env, noise, first_env, n_rows);
    else sim_standin_kernel<32><<<blocks, PARC_STANDIN_THREADS, 0, as_stream(stream)>>>(*m, *env, noise, first_env, n_rows);
    return check_launch("sim_standin");
}
}  // namespace parc

static int validate_standin(const parc_char_model_t* m, const parc_env_buffers_t* env, const float* noise) {
    PARC_REQUIRE(m && env && noise, "sim_standin: NULL argument");
    PARC_REQUIRE(env->char_root_pos.ptr && env->char_root_rot.ptr && env->char_root_vel.ptr && env->char_root_ang_vel.ptr &&
                     env->char_dof_pos.ptr && env->char_dof_vel.ptr && env->char_rigid_body_pos.ptr && env->ref_root_pos.ptr &&
                     env->ref_root_rot.ptr && env->ref_root_vel.ptr && env->ref_root_ang_vel.ptr && env->ref_dof_pos.ptr &&
                     env->ref_dof_vel.ptr && env->ref_body_pos.ptr, "sim_standin: state views not set");
    PARC_REQUIRE(m->num_bodies <= 32, "sim_standin: more than 32 bodies");
    return PARC_OK;
}

extern "C" int parc_b200_sim_standin(const parc_char_model_t* m, const parc_env_buffers_t* env, const float* noise, void* stream) {
    const int rc = validate_standin(m, env, noise);
    if (rc) return rc;
    if (env->num_envs == 0) return PARC_OK;
    return parc::launch_sim_standin(m, env, noise, 0, env->num_envs, stream);
}

// ---- parc_b200_env_step_host: the fused step fed from HOST simulator state ------------------------
namespace parc {
constexpr int kMaxHostChunks = 64;
struct HostPipe {            // one per device: the upload stream and the per-chunk "rows have arrived" events
    cudaStream_t copy = nullptr;
    cudaEvent_t arrived[kMaxHostChunks] = {};
    cudaEvent_t free_to_overwrite = nullptr;
};
static HostPipe* host_pipe() {
    static HostPipe pipes[64];
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return nullptr;
    HostPipe& p = pipes[dev];
    if (p.copy == nullptr) {
        if (cudaStreamCreateWithFlags(&p.copy, cudaStreamNonBlocking) != cudaSuccess) return nullptr;
        for (int c = 0; c < kMaxHostChunks; ++c)
            if (cudaEventCreateWithFlags(&p.arrived[c], cudaEventDisableTiming) != cudaSuccess) return nullptr;
        if (cudaEventCreateWithFlags(&p.free_to_overwrite, cudaEventDisableTiming) != cudaSuccess) return nullptr;
    }
    return &p;
}
}  // namespace parc

extern "C" int parc_b200_env_step_host(const parc_char_model_t* m, const parc_motion_lib_t* lib, const parc_terrain_t* terrain,
                                       const parc_env_config_t* cfg, const parc_env_buffers_t* env, uint32_t parts,
                                       const parc_host_rows_t* uploads, int32_t num_uploads, int32_t num_chunks,
                                       const float* standin_noise, void* stream) {
    PARC_REQUIRE(env != nullptr, "env_step_host: NULL buffers");
    const long long N = env->num_envs;
    if (N == 0 || parts == 0) return PARC_OK;
    int rc = validate_env_step(m, lib, terrain, cfg, env, parts);
    if (rc) return rc;
    PARC_REQUIRE(num_uploads >= 0 && (num_uploads == 0 || uploads), "env_step_host: NULL upload list");
    PARC_REQUIRE(num_chunks >= 1 && num_chunks <= kMaxHostChunks, "env_step_host: num_chunks %d outside [1, %d]", num_chunks, kMaxHostChunks);
    for (int k = 0; k < num_uploads; ++k)
        PARC_REQUIRE(uploads[k].dst_device && uploads[k].src_host && uploads[k].row_bytes > 0, "env_step_host: upload %d is incomplete", k);
    if (standin_noise) { rc = validate_standin(m, env, standin_noise); if (rc) return rc; }
    HostPipe* pipe = host_pipe();
    if (pipe == nullptr) { set_error("env_step_host: cannot create the upload stream"); return PARC_ERR_CUDA; }
    cudaStream_t cs = as_stream(stream);
    // the previous step's kernels (enqueued on `stream`) must be done reading the state before it is overwritten
    cudaEventRecord(pipe->free_to_overwrite, cs);
    cudaStreamWaitEvent(pipe->copy, pipe->free_to_overwrite, 0);
    // Range boundaries on multiples of 16 envs (the step kernel's tile). The ranges shrink geometrically (x 0.8): what is
    // left after the last upload has landed is the step of the LAST range only, so that one should be small, while a
    // larger first range costs nothing (its step overlaps the following uploads, and the step is faster per env than PCIe).
    const long long tiles = (N + 15) / 16;
    double cum[kMaxHostChunks + 1];
    cum[0] = 0.0;
    for (int c = 0; c < num_chunks; ++c) cum[c + 1] = cum[c] + pow(0.8, (double)c);
    for (int c = 0; c < num_chunks; ++c) {
        const long long a = std::min(N, (long long)(tiles * (cum[c] / cum[num_chunks])) * 16);
        const long long b = c + 1 == num_chunks ? N : std::min(N, (long long)(tiles * (cum[c + 1] / cum[num_chunks])) * 16);
        if (b <= a) continue;
        for (int k = 0; k < num_uploads; ++k) {
            const long long rb = uploads[k].row_bytes;
            cudaError_t ce = cudaMemcpyAsync(static_cast<char*>(uploads[k].dst_device) + a * rb, static_cast<const char*>(uploads[k].src_host) + a * rb,
                                             (size_t)((b - a) * rb), cudaMemcpyHostToDevice, pipe->copy);
            if (ce != cudaSuccess) { set_error("env_step_host: upload %d: %s", k, cudaGetErrorString(ce)); return PARC_ERR_CUDA; }
        }
        cudaEventRecord(pipe->arrived[c], pipe->copy);
        cudaStreamWaitEvent(cs, pipe->arrived[c], 0);
        if (standin_noise) { rc = launch_sim_standin(m, env, standin_noise, a, b - a, stream); if (rc) return rc; }
        rc = launch_env_step(m, lib, terrain, cfg, env, parts, nullptr, b - a, nullptr, nullptr, stream, a);
        if (rc) return rc;
    }
    return PARC_OK;
}

// ---- recording path: IGParkourEnv.write_agent_states (ig_parkour_env.py:759-796) ---------------------------------
namespace parc {
__global__ void __launch_bounds__(256) record_agent_states_kernel(const int J, const int D, const int C, const __grid_constant__ parc_motion_lib_t lib,
                                                                  const __grid_constant__ parc_env_config_t cfg, const __grid_constant__ parc_env_buffers_t env,
                                                                  const __grid_constant__ parc_record_tape_t tape, const int record_ref,
                                                                  const int bypass, const double timestep) {
    // one warp per env; lanes stride over the row's columns
    const int lane = threadIdx.x & 31;
    const long long wpb = blockDim.x / 32;
    const int W = 6 + D;
    for (long long e = (long long)blockIdx.x * wpb + threadIdx.x / 32; e < env.num_envs; e += (long long)gridDim.x * wpb) {
        if (!tape.writing[e]) continue;                                              // :764-765 (warp-uniform)
        if (lane == 0) *tape.still_recording = 1;                                    // :766
        const int cur = tape.cursor[e];
        const bool room = cur < tape.max_frames;
        if (room) {
            float* fr = tape.frames + ((long long)e * tape.max_frames + cur) * W;
            float* co = tape.contacts + ((long long)e * tape.max_frames + cur) * C;
            const parc_view3_t& vp = record_ref ? env.ref_root_pos : env.char_root_pos;
            const parc_view3_t& vr = record_ref ? env.ref_root_rot : env.char_root_rot;
            const parc_view3_t& vd = record_ref ? env.ref_dof_pos : env.char_dof_pos;
            if (lane < 3) fr[lane] = view_ld(vp, e, lane);
            if (lane == 3) {
                const Vec3 em = quat_to_exp_map(view_ld_quat(vr, e));                // _get_char_state, :664-676
                fr[3] = em.x; fr[4] = em.y; fr[5] = em.z;
            }
            for (int k = lane; k < D; k += 32) fr[6 + k] = view_ld(vd, e, k);
            for (int c = lane; c < C; c += 32) {
                float v;
                if (record_ref) v = view_ld(env.ref_contacts, e, c);
                else {
                    const float fx = view_ld3(env.char_contact_forces, e, c, 0), fy = view_ld3(env.char_contact_forces, e, c, 1),
                                fz = view_ld3(env.char_contact_forces, e, c, 2);
                    v = sqrtf(fx * fx + fy * fy + fz * fz) > cfg.contact_eps ? 1.f : 0.f;   // :674-676
                }
                co[c] = v;
            }
            if (tape.obs) {
                const float* src = env.obs_buf + e * env.obs_row_stride;
                float* dst = tape.obs + ((long long)e * tape.max_frames + cur) * tape.obs_dim;
                for (int k = lane; k < tape.obs_dim; k += 32) dst[k] = src[k];       // :773-774
            }
        }
        if (lane == 0) {
            if (room) tape.cursor[e] = cur + 1; else *tape.overflow = 1;
            if (env.done_buf[e] == PARC_DONE_FAIL) {                                 // :776-794
                tape.writing[e] = 0;
                const MotionMeta mm = load_motion_meta(lib.meta, env.motion_ids[e]);
                const float mt = __fadd_rn(env.time_buf[e], env.motion_time_offsets[e]);
                const bool failed_early = !bypass && ((double)mt < (double)mm.length - timestep * 2.0);
                if (!failed_early) {
                    tape.success[e] = 1;
                    tape.saved_list[atomicAdd(tape.saved_count, 1)] = (int)e;
                }
            }
        }
    }
}
}  // namespace parc

extern "C" int parc_b200_record_agent_states(const parc_char_model_t* m, const parc_motion_lib_t* lib, const parc_env_config_t* cfg,
                                             const parc_env_buffers_t* env, const parc_record_tape_t* tape, int32_t record_ref,
                                             int32_t bypass_record_fail, double timestep, void* stream) {
    if (!model_ok(m, "record_agent_states")) return PARC_ERR_INVALID_ARG;
    PARC_REQUIRE(lib && lib->meta && cfg && env && tape, "record_agent_states: NULL argument");
    PARC_REQUIRE(tape->frames && tape->contacts && tape->cursor && tape->writing && tape->success && tape->saved_list && tape->saved_count &&
                     tape->still_recording && tape->overflow && tape->max_frames > 0, "record_agent_states: incomplete tape");
    PARC_REQUIRE(env->done_buf && env->time_buf && env->motion_time_offsets && env->motion_ids, "record_agent_states: env buffers not set");
    if (record_ref) PARC_REQUIRE(env->ref_root_pos.ptr && env->ref_root_rot.ptr && env->ref_dof_pos.ptr && env->ref_contacts.ptr, "record_agent_states: ref_* views not set");
    else PARC_REQUIRE(env->char_root_pos.ptr && env->char_root_rot.ptr && env->char_dof_pos.ptr && env->char_contact_forces.ptr, "record_agent_states: simulator views not set");
    if (tape->obs) PARC_REQUIRE(env->obs_buf && tape->obs_dim > 0 && tape->obs_dim <= env->obs_row_stride, "record_agent_states: observation tape without a matching obs_buf");
    cudaStream_t st = as_stream(stream);
    zero_int_kernel<<<1, 1, 0, st>>>(tape->saved_count);
    zero_int_kernel<<<1, 1, 0, st>>>(tape->still_recording);
    if (env->num_envs == 0) return check_launch("record_agent_states (clear)");
    const int C = record_ref ? lib->num_contacts : m->num_bodies;
    const unsigned blocks = (unsigned)std::min<long long>((env->num_envs + 7) / 8, kNumSMs * 16);
    record_agent_states_kernel<<<blocks, 256, 0, st>>>(m->num_bodies, m->dof_size, C, *lib, *cfg, *env, *tape, record_ref, bypass_record_fail, timestep);
    return check_launch("record_agent_states");
}
