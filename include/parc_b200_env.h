/* parc_b200 -- C ABI, tracking-env part (observation, reward, termination, RSI reset).
 * Included by parc_b200.h. Citations are file:line into the reference (CloudEngineHub/PARC). */
#ifndef PARC_B200_ENV_H
#define PARC_B200_ENV_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PARC_MAX_DOFS 96

/* A [N, A, B] fp32 view with strides in ELEMENTS (mid_stride unused for 2-D buffers). */
typedef struct {
    float* ptr;
    int64_t row_stride;
    int64_t mid_stride;
    int64_t elem_stride;
} parc_view3_t;

/* Static configuration of the tracking env: the knobs of
 * data/configs/tracker_config/dm_env_default.yaml that change hot-path arithmetic, as parsed by
 * IGParkourEnv.__init__ (PARC/motion_tracker/envs/ig_parkour/ig_parkour_env.py:43-149) and
 * RefCharEnv.__init__ (ig_parkour/mgdm_dm_util.py:15-35). */
typedef struct {
    int32_t num_tar_steps;                       /* S */
    float tar_dt[PARC_MAX_TAR_STEPS];            /* fp32(timestep * tar_obs_steps[s]), mgdm_dm_util.py:232 */
    int32_t num_key_bodies;                      /* K */
    int32_t key_body_ids[PARC_MAX_KEY_BODIES];
    int32_t num_ray_points;                      /* P */
    int32_t global_obs;
    int32_t root_height_obs;
    int32_t enable_tar_obs;
    int32_t use_contact_info;
    int32_t track_root;
    int32_t track_root_h;
    int32_t enable_early_termination;
    int32_t pose_termination;
    int32_t num_fall_contact_bodies;             /* len(contact_body_ids); 0 disables the fall test */
    uint32_t fall_contact_body_mask;             /* bit b set <=> body b in contact_body_ids */
    int32_t never_done;
    float timestep;                              /* fp32(1 / control_freq) */
    float episode_length;
    float termination_height;
    float root_pos_termination_dist;
    float root_rot_termination_angle;
    float pose_termination_dist[PARC_MAX_BODIES];/* [j-1] for body j = 1..J-1 */
    float joint_err_w[PARC_MAX_BODIES];          /* [j-1] */
    float dof_err_w[PARC_MAX_DOFS];
    float contact_weights[PARC_MAX_BODIES];
    float pose_w, vel_w, root_pos_w, root_vel_w, key_pos_w;   /* already normalised (:91-101) */
    float contact_eps;                           /* _get_char_contact_state eps, 1e-5 (:655) */
    float min_obs_h, max_obs_h;
    float ema_weight;                            /* DeepMimicEnv._ema_weight, dm_env.py:90 */
    float ema_keep;                              /* fp32(1.0 - ema_weight) evaluated in double like the Python expression, dm_env.py:654 */
    float min_motion_weight;                     /* dm_env.py:26 */
    float rand_root_pos_offset_scale;            /* mgdm_dm_util.py:174 */
    /* ray_xy_points as a lattice (geom_util.get_xy_grid_points, PARC/util/geom_util.py:212-224: meshgrid(xs, ys, 'ij')):
     * point i * ray_grid_ny + j == (xs[i], ys[j]) bit for bit. 0 x 0 = not a lattice (e.g. get_xy_points_cone, :251-272).
     * Fill with parc_b200_ray_points_lattice; a lattice lets the step share the rotated xs / ys products per row / column. */
    int32_t ray_grid_nx, ray_grid_ny;
} parc_env_config_t;

/* All per-env buffers the reference injects through RefCharEnv.get_sim_tensor_views /
 * get_data_buffer_views (mgdm_dm_util.py:37-84) plus the ones DeepMimicEnv owns
 * (dm_env.py:80-92). The env does not own the views: every write is in place. */
typedef struct {
    int64_t num_envs;                            /* N */
    /* simulator state (inputs; written only by reset) */
    parc_view3_t char_root_pos;                  /* [N,3]  */
    parc_view3_t char_root_rot;                  /* [N,4]  */
    parc_view3_t char_root_vel;                  /* [N,3]  */
    parc_view3_t char_root_ang_vel;              /* [N,3]  */
    parc_view3_t char_dof_pos;                   /* [N,D]  */
    parc_view3_t char_dof_vel;                   /* [N,D]  */
    parc_view3_t char_contact_forces;            /* [N,J,3] */
    parc_view3_t char_rigid_body_pos;            /* [N,J,3] */
    parc_view3_t char_rigid_body_vel;            /* [N,J,3] (zeroed by reset; may be NULL) */
    parc_view3_t char_rigid_body_ang_vel;        /* [N,J,3] (zeroed by reset; may be NULL) */
    /* reference-motion state (outputs of _update_ref_motion / reset) */
    parc_view3_t ref_root_pos, ref_root_rot, ref_root_vel, ref_root_ang_vel;
    parc_view3_t ref_body_pos;                   /* [N,J,3] */
    parc_view3_t ref_joint_rot;                  /* [N,J-1,4] */
    parc_view3_t ref_dof_pos, ref_dof_vel;       /* [N,D] */
    parc_view3_t ref_contacts;                   /* [N,C] */
    /* data buffers */
    float* reward_buf;                           /* [N] */
    float* reward_terms;                         /* [N,8]: pose_r, vel_r, root_pos_r, root_vel_r, key_pos_r, contact_penalty, total_r, pad */
    int32_t* done_buf;                           /* [N] */
    float* time_buf;                             /* [N] */
    int32_t* timestep_buf;                       /* [N] */
    uint8_t* actors_need_reset;                  /* [N, actors] bool; column 0 set by reset */
    int64_t actors_per_env;
    const float* env_offsets;                    /* [N,3] */
    const float* ray_xy_points;                  /* [P,2] */
    float* ray_hfs;                              /* [N,P] or NULL (heights then live only in obs_buf) */
    float* obs_buf;                              /* [N, obs_dim] */
    int64_t obs_row_stride;
    /* DeepMimicEnv-owned bookkeeping */
    int64_t* motion_ids;                         /* [N] */
    int64_t* motion_terrain_ids;                 /* [N] */
    float* motion_time_offsets;                  /* [N] */
    float* motion_id_fail_rates;                 /* [M] */
    const float* motion_weights;                 /* [M] MotionLib._motion_weights */
    const float* dm_motion_offsets;              /* [M, terrains_per_motion, 2] */
    int32_t terrains_per_motion;
    /* per-step scratch written by the step kernel, consumed by fail_rate_ema / reset */
    int32_t* done_count;                         /* [1] */
    int32_t* done_list;                          /* [N] env ids with done != NULL (unordered) */
    uint8_t* done_code;                          /* [N] 0 = running, 1 = FAIL before the motion-end override, 2 = TIME/SUCC/motion end */
} parc_env_buffers_t;

/* Which parts of IGEnv._post_physics_step (PARC/motion_tracker/envs/ig_env.py:368-377) one call
 * of parc_b200_env_step performs. */
enum {
    PARC_STEP_TIME = 1,        /* _update_time, ig_env.py:391-394 */
    PARC_STEP_REF = 2,         /* DeepMimicEnv._update_ref_motion, dm_env.py:523-545 */
    PARC_STEP_HEIGHTS = 4,     /* _refresh_obs_hfs, ig_parkour_env.py:518-533 */
    PARC_STEP_OBS = 8,         /* IGParkourEnv._compute_obs, ig_parkour_env.py:842-965 */
    PARC_STEP_REWARD = 16,     /* _update_reward, ig_parkour_env.py:984-1044 */
    PARC_STEP_DONE = 32,       /* _update_done + DeepMimicEnv.update_done flags, ig_parkour_env.py:967-982, dm_env.py:628-665 */
    PARC_STEP_ALL = 63
};

/* Observation row width for a model/config (ig_parkour_env.py:842-965). */
int parc_b200_obs_dim(const parc_char_model_t* h_model, const parc_env_config_t* h_cfg);

/* One fused pass of the post-physics step over all N envs (or over the subset env_ids [n_sub],
 * i64, when env_ids != NULL -- the reset-time call IGParkourEnv._update_observations(env_ids),
 * whose rows land at obs_buf[env_ids]). `parts` is a mask of PARC_STEP_*.
 * With PARC_STEP_DONE the kernel also appends finished envs to done_list/done_count (the count
 * must be zeroed by the caller, e.g. by parc_b200_fail_rate_ema of the previous step) and writes
 * done_code for parc_b200_fail_rate_ema. */
int parc_b200_env_step(const parc_char_model_t* h_model, const parc_motion_lib_t* h_lib,
                       const parc_terrain_t* h_terrain, const parc_env_config_t* h_cfg,
                       const parc_env_buffers_t* h_env, uint32_t parts, const int64_t* env_ids,
                       int64_t n_sub, void* stream);

/* One [N, row_bytes] array of simulator state that lives in (pinned) HOST memory and its device twin. */
typedef struct parc_host_rows_t {
    void* dst_device;        /* device array the env views (char_root_pos, ...) point into */
    const void* src_host;    /* the same rows in host memory; pinned, or the copy is not asynchronous */
    int64_t row_bytes;       /* bytes per env */
} parc_host_rows_t;

/* parc_b200_env_step for a simulator whose state tensors live on the HOST (the CPU pipeline of
 * IGEnv._post_physics_step, ig_env.py:368-377, where _refresh_sim_tensors hands over host arrays):
 * the N envs are cut into num_chunks row ranges (multiples of 8 envs); the rows of range c are
 * uploaded on an internal copy stream while the step kernel of range c-1 runs on `stream`, so the
 * PCIe transfer hides behind the computation instead of preceding it. Results are those of
 * parc_b200_env_step over all envs (each env depends on its own rows only). The caller zeroes
 * done_count first, as for parc_b200_env_step. standin_noise (NULL in production) applies the
 * benchmark stand-in of parc_b200_sim_standin to each range after its upload. */
int parc_b200_env_step_host(const parc_char_model_t* h_model, const parc_motion_lib_t* h_lib,
                            const parc_terrain_t* h_terrain, const parc_env_config_t* h_cfg,
                            const parc_env_buffers_t* h_env, uint32_t parts,
                            const parc_host_rows_t* uploads, int32_t num_uploads, int32_t num_chunks,
                            const float* standin_noise, void* stream);

/* DeepMimicEnv.compute_tar_obs (dm_env.py:594-626) + fetch_tar_obs_data (mgdm_dm_util.py:221-244):
 * -> tar_root_pos [n,S,3], tar_root_rot [n,S,4], tar_joint_rot [n,S,J-1,4], tar_key_pos [n,S,K,3],
 * tar_contacts [n,S,C]; env_ids may be NULL (all envs, n = N). */
int parc_b200_compute_tar_obs_data(const parc_char_model_t* h_model, const parc_motion_lib_t* h_lib,
                                   const parc_env_config_t* h_cfg, const parc_env_buffers_t* h_env,
                                   const int64_t* env_ids, int64_t n, float* tar_root_pos,
                                   float* tar_root_rot, float* tar_joint_rot, float* tar_key_pos,
                                   float* tar_contacts, void* stream);

/* The free jit functions of the reference, on caller-provided contiguous tensors. */

/* compute_char_obs (PARC/motion_tracker/envs/ig_char_env.py:581-625) -> obs [n, char_obs_dim] */
int parc_b200_compute_char_obs(const parc_char_model_t* h_model, const float* root_pos,
                               const float* root_rot, const float* root_vel,
                               const float* root_ang_vel, const float* joint_rot,
                               const float* dof_vel, const float* key_pos, int32_t num_key,
                               int32_t global_obs, int32_t root_height_obs, int64_t n, float* obs,
                               void* stream);

/* compute_tar_obs (mgdm_dm_util.py:404-461) -> obs [n, S, 3+6+6(J-1)+3K] */
int parc_b200_compute_tar_obs(const parc_char_model_t* h_model, const float* ref_root_pos,
                              const float* ref_root_rot, const float* tar_root_pos,
                              const float* tar_root_rot, const float* tar_joint_rot,
                              const float* tar_key_pos, int32_t num_steps, int32_t num_key,
                              int32_t global_obs, int32_t global_tar_root_h_obs, int64_t n,
                              float* obs, void* stream);

/* compute_deepmimic_reward (mgdm_dm_util.py:269-332) -> [n,5] */
int parc_b200_compute_deepmimic_reward(const parc_char_model_t* h_model, const float* root_pos,
                                       const float* root_rot, const float* root_vel,
                                       const float* root_ang_vel, const float* joint_rot,
                                       const float* dof_vel, const float* key_pos,
                                       const float* tar_root_pos, const float* tar_root_rot,
                                       const float* tar_root_vel, const float* tar_root_ang_vel,
                                       const float* tar_joint_rot, const float* tar_dof_vel,
                                       const float* tar_key_pos, const float* joint_rot_err_w,
                                       const float* dof_err_w, int32_t num_key, int32_t track_root_h,
                                       int32_t track_root, int64_t n, float* reward5, void* stream);

/* compute_contact_reward (mgdm_dm_util.py:497-518): [n,C],[n,C,3],[C] -> [n,C] */
int parc_b200_compute_contact_reward(const float* tar_contacts, const float* contact_forces,
                                     const float* contact_weights, int32_t num_contacts, int64_t n,
                                     float* out, void* stream);

/* compute_done (mgdm_dm_util.py:334-402); termination_heights [n,J] precomputed by the caller as
 * in RefCharEnv.update_done (:147-172), contact_body_ids i64 [n_cb], pose_termination_dist [J-1].
 * root_pos_termination_dist is the python float (a double): it is squared in double and the product
 * rounded to fp32 for the comparison, as TorchScript does. */
int parc_b200_compute_done(const int32_t* done_buf_in, const float* time, float ep_len,
                           const float* root_rot, const float* body_pos, const float* tar_root_rot,
                           const float* tar_body_pos, const float* contact_force,
                           const int64_t* contact_body_ids, int32_t n_cb,
                           const float* termination_heights, int32_t pose_termination,
                           const float* pose_termination_dist, int32_t enable_early_termination,
                           int32_t track_root, double root_pos_termination_dist,
                           float root_rot_termination_angle, int32_t num_bodies, int64_t n,
                           int32_t* done_out, void* stream);

/* compute_tracking_error (mgdm_dm_util.py:520-553) -> [n,7] = root pos L2, |root rot angle|, mean body
 * pos L2 (root-relative), mean |body rot angle|, mean |dof vel diff|, mean |root vel diff|, mean |root
 * ang vel diff|. body_rot / body_pos are the FK outputs of both characters, as the caller
 * (ig_parkour_env.py:1066-1095) computes them. */
int parc_b200_compute_tracking_error(const float* root_pos, const float* root_rot, const float* body_rot,
                                     const float* body_pos, const float* tar_root_pos,
                                     const float* tar_root_rot, const float* tar_body_rot,
                                     const float* tar_body_pos, const float* root_vel,
                                     const float* root_ang_vel, const float* dof_vel,
                                     const float* tar_root_vel, const float* tar_root_ang_vel,
                                     const float* tar_dof_vel, int32_t num_bodies, int32_t dof_size,
                                     int64_t n, float* out, void* stream);

/* Fail-rate EMA of DeepMimicEnv.update_done (dm_env.py:648-660): for every finished env IN
 * ASCENDING ENV-ID ORDER per motion, rate <- rate*(1-w) + w (FAIL) or rate*(1-w) (otherwise).
 * Consumes done_list/done_code written by parc_b200_env_step; also refreshes the RSI sampling CDF
 * cdf [M] = inclusive prefix sum of clamp(rate, min_motion_weight) * motion_weight
 * (dm_env.py:487-490). scratch must hold 2*M + done capacity int32. Does NOT reset done_count. */
int parc_b200_fail_rate_ema(const parc_env_config_t* h_cfg, const parc_env_buffers_t* h_env,
                            int32_t num_motions, float* cdf, int32_t* scratch, void* stream);

/* Env-sharded runs (no counterpart in the single-process reference): ADD to counts [2, num_motions]
 * (int64, device) how many envs of the done list finished the last step as FAIL (row 0) / as time-out, success or motion end
 * (row 1), per motion. One rank's input to the cross-rank fail-rate rule: the ranks sum their tables with one all-reduce
 * and apply the closed form of dm_env.py:648-660 (parc_b200/sharding.py). Call after the step, before the reset. */
int parc_b200_outcome_counts(const parc_env_buffers_t* h_env, int32_t num_motions, int64_t* counts, void* stream);

/* Allocation-free variant of parc_b200_fail_rate_ema. It has NO counterpart in the reference: it was added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving
 * loops, harnesses that forbid it). Same inputs, outputs and results as parc_b200_fail_rate_ema, but the scratch is a
 * file-scope device array inside the library (PARC_EMA_NOALLOC_MAX_INTS int32 = 4 MiB, held by every GPU that loads the
 * library until exit); the call allocates nothing (no cudaMalloc / cudaFree / allocating library routine).
 * Limit: 2 * num_motions + num_envs <= PARC_EMA_NOALLOC_MAX_INTS, otherwise PARC_ERR_INVALID_ARG with a message naming
 * the limit. All calls on a device share the one array, which keeps the last call's contents (the kernel clears what it
 * needs at zero before use): calls that overlap on the GPU -- different streams without an ordering between them -- race;
 * serialise them or use parc_b200_fail_rate_ema with separate scratch buffers. */
#define PARC_EMA_NOALLOC_MAX_INTS (1 << 20)
int parc_b200_fail_rate_ema_noalloc(const parc_env_config_t* h_cfg, const parc_env_buffers_t* h_env,
                                    int32_t num_motions, float* cdf, void* stream);

/* DeepMimicEnv.reset (dm_env.py:567-592) for env_ids [n] i64 given the random draws the reference
 * takes from torch's generator: motion_ids [n] i64 (sample_motions / demo-mode ids),
 * terrain_ids [n] i64 (randint), motion_times [n] f32 (sample_time or len*start_fraction),
 * xy_noise [n,2] f32 in [-1,1) (add_noise_to_char_state, mgdm_dm_util.py:102-106).
 * Optional per-env offsets (apply_offsets_to_char_state, :108-126) may be NULL. */
int parc_b200_env_reset(const parc_char_model_t* h_model, const parc_motion_lib_t* h_lib,
                        const parc_env_config_t* h_cfg, const parc_env_buffers_t* h_env,
                        const int64_t* env_ids, int64_t n, const int64_t* motion_ids,
                        const int64_t* terrain_ids, const float* motion_times,
                        const float* xy_noise, const float* root_pos_offset,
                        const float* root_rot_offset, void* stream);

/* The per-call knobs of DeepMimicEnv.reset that are not random draws (all optional; a NULL struct = all defaults). */
typedef struct parc_reset_options_t {
    const float* root_pos_offset;              /* [N,3] RefCharEnv._root_pos_offset (mgdm_dm_util.py:108-112), indexed by env id, or NULL */
    const float* root_rot_offset;              /* [N,4] RefCharEnv._root_rot_offset (:113-115), or NULL */
    const float* motion_start_time_fraction;   /* [N] DeepMimicEnv._motion_start_time_fraction (dm_env.py:92, 500-502): used when
                                                  rand_reset == 0; NULL = zeros */
    int32_t demo_mode;                         /* dm_env.py:479-480: motion id = env id % num_motions instead of a draw */
} parc_reset_options_t;

/* Device-side RSI for the graph-captured loop: for each entry of done_list (count read on the
 * device) draw motion id (inverse CDF over `cdf`, binary search), terrain id and phase from a
 * Philox4x32-10 counter keyed by (seed, env id, step), then do everything parc_b200_env_reset does
 * and re-evaluate heights + observation of the reset envs (IGParkourEnv.reset,
 * ig_parkour_env.py:809-829). Zeroes done_count when finished.
 * `cdf` [M] is the inclusive prefix sum of the sampling weights: clamp(fail_rate, min_motion_weight) * motion_weight as
 * parc_b200_fail_rate_ema leaves it (dm_env.py:487-490), or of motion_weight alone for dm.ignore_fail_rates (:485-486);
 * it must be positive at the end (an all-zero CDF would send every env to the last motion). */
int parc_b200_env_reset_sampled(const parc_char_model_t* h_model, const parc_motion_lib_t* h_lib,
                                const parc_terrain_t* h_terrain, const parc_env_config_t* h_cfg,
                                const parc_env_buffers_t* h_env, const float* cdf,
                                int32_t num_motions, uint64_t seed, uint64_t step,
                                int32_t rand_reset, const parc_reset_options_t* h_options, void* stream);

/* Inverse-CDF sampler alone (MotionLib.sample_motions semantics given uniforms u [n] in [0,1)):
 * ids[i] = first m with cdf[m] > u[i] * cdf[M-1]. */
int parc_b200_sample_motions_cdf(const float* cdf, int32_t num_motions, const float* u, int64_t n,
                                 int64_t* ids, void* stream);

/* ---- recording path (stage 4 `record`) ---------------------------------------------
 * IGParkourEnv.write_agent_states (PARC/motion_tracker/envs/ig_parkour/ig_parkour_env.py:759-796) appends, per control
 * step and per env in a host loop, the character state [root_pos | exp_map(root_rot) | dof_pos] and the contact flags
 * (_get_char_state, :664-687; the reference character instead when record_ref), optionally the observation row, and
 * stops an env at its first FAIL. Here the tape lives in HBM and one launch does a step for all envs. */
typedef struct parc_record_tape_t {
    float* frames;            /* [N, max_frames, 6 + D] */
    float* contacts;          /* [N, max_frames, C] */
    float* obs;               /* [N, max_frames, obs_dim] or NULL (record_obs False, :1157-1163) */
    int32_t* cursor;          /* [N] frames recorded so far */
    uint8_t* writing;         /* [N] _writing_env_state (:1177) */
    uint8_t* success;         /* [N] _env_success_state (:1178) */
    int32_t* saved_list;      /* [N] envs the reference would hand to save_agent_states_to_file in THIS step (unordered) */
    int32_t* saved_count;     /* [1] zeroed by the call */
    int32_t* still_recording; /* [1] set to 1 iff some env was writing at this step = the _write_agent_states_flag the loop leaves (:762-767) */
    int32_t* overflow;        /* [1] set to 1 if an env ran past max_frames (its later frames are dropped) */
    int32_t max_frames;
    int32_t obs_dim;
} parc_record_tape_t;

/* One write_agent_states step. An env that is writing gets its row appended; if done_buf == FAIL it stops, and it
 * counts as a success unless (!bypass_record_fail and motion_time < motion_length - 2 * timestep, compared in double
 * like the Python floats, :786-789), motion_time = time_buf + motion_time_offsets (dm_env.py:547-552). */
int parc_b200_record_agent_states(const parc_char_model_t* h_model, const parc_motion_lib_t* h_lib,
                                  const parc_env_config_t* h_cfg, const parc_env_buffers_t* h_env,
                                  const parc_record_tape_t* h_tape, int32_t record_ref, int32_t bypass_record_fail,
                                  double timestep, void* stream);

/* Benchmark / test utility, NOT a reference function: a stand-in for the physics simulator between two
 * control steps, so that a device-resident loop sees a stationary workload. For every env it rewrites
 * the simulated character state as the env's current reference state plus a fixed per-env
 * perturbation: char_root_pos = ref_root_pos + n[0:3]; char_root_rot = normalize(ref_root_rot (x)
 * n[3:7]) (n[3:7] a unit quaternion); char_root_vel / ang_vel / dof_pos / dof_vel = ref_* + noise;
 * char_rigid_body_pos = forward kinematics of that root + dof state + body noise (what the simulator
 * would report; ref_body_pos is stale for envs that were just reset). noise is [N, 13 + 2 D + 3 J]
 * row-major in that column order. Contact forces are left as they are. */
int parc_b200_sim_standin(const parc_char_model_t* h_model, const parc_env_buffers_t* h_env, const float* noise,
                          void* stream);

/* Host-side plan of the fused step's forward kinematics (no GPU needed): a schedule of the joints 1..J-1 on the 4
 * lanes that share one (env, pose). entries [4 * PARC_MAX_BODIES]: entries[lane * PARC_MAX_BODIES + it] = what the lane
 * evaluates in iteration it (-1 = idle): joint | key-slot mask << 8 | parent-from-slot << 16 | read slot << 17 |
 * publish << 22 | write slot << 23 | has-children << 28. A joint runs in a later iteration than its parent; without the
 * from-slot bit the parent is what the same lane evaluated in the previous iteration (or the root, if it has not
 * evaluated anything yet). max_len = iterations, num_slots = shared transform slots per lane group, root_slot = the slot
 * the root transform is published to before the first iteration (-1: no entry reads the root from a slot). */
int parc_b200_tree_parts(const parc_char_model_t* h_model, int32_t* entries, int32_t* max_len, int32_t* num_slots,
                         int32_t* root_slot);

/* Host-side test (no GPU): are the num_points HOST points h_pts_xy [P,2] the lattice meshgrid(xs, ys, 'ij') of
 * geom_util.get_xy_grid_points (PARC/util/geom_util.py:212-224), i.e. point i * ny + j == (xs[i], ys[j]) exactly?
 * nx, ny receive the lattice dims, or 0, 0. The values belong in parc_env_config_t.ray_grid_nx / ray_grid_ny. */
int parc_b200_ray_points_lattice(const float* h_pts_xy, int32_t num_points, int32_t* nx, int32_t* ny);

/* Self-test: the packed fp32x2 cell computation of the fused step against the scalar, separately rounded one
 * (rotate_2d_vec + SubTerrain.get_grid_index) for n sample points [n,2] under one heading (c = cos, s = sin) and
 * root xy (gx, gy): both cell ids (ix * dim_y + iy) are returned and must be equal everywhere. */
int parc_b200_selftest_cells(const float* pts_xy, int64_t n, float c, float s, float gx, float gy, float min_x,
                             float min_y, float dx, float dy, int32_t dim_x, int32_t dim_y,
                             int32_t* cell_packed, int32_t* cell_scalar, void* stream);

/* Self-test of the heightmap index arithmetic: the fused step divides by the (uniform) cell size
 * with q = x*rc; r = fma(-q, d, x); q' = fma(r, rc, q), rc = RN(1/d), instead of div.rn. This entry
 * evaluates both on x [n] so that a test can demand bit-identical quotients (the nearest-cell
 * index of SubTerrain.get_grid_index, PARC/util/terrain_util.py:143-152, must be bit-exact). */
int parc_b200_selftest_div(const float* x, int64_t n, float d, float* q_fast, float* q_ieee, void* stream);

#ifdef __cplusplus
}
#endif

#endif /* PARC_B200_ENV_H */
