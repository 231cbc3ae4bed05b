/* pointfoot_b200 -- C ABI of the B200-native hot path of peachvegetable/pointfoot (legged_gym fork).
 *
 * Plain pointers and sizes only; every pointer is a DEVICE pointer into caller-owned memory (torch tensors,
 * or PhysX tensors wrapped zero-copy) unless the name says host.  No allocation, no ownership transfer;
 * the only retained state is the opaque pf_ctx (constants + small scratch).  Every entry point returns 0
 * on success; on failure a non-zero code, with pf_last_error() giving the message.  All launches go to the
 * cudaStream_t passed as `void* stream` (NULL = legacy default stream).  One ctx per device; thread-compatible.
 *
 * Citations: PF = legged_gym/envs/pointfoot/point_foot.py, LR = legged_gym/envs/base/legged_robot.py,
 * rsl_rl = rsl_rl v1.0.2 (third-party).
 */
#ifndef POINTFOOT_B200_H
#define POINTFOOT_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PF_ABI_VERSION 2
#define PF_MAX_DOF 6
#define PF_MAX_BODIES 9
#define PF_NUM_FEET 2
#define PF_MAX_REWARD_TERMS 24
#define PF_MAX_HEIGHT_POINTS 256
#define PF_MAX_OBS 512

/* reward term ids: the 24 _reward_* methods, PF:1132-1253 */
enum pf_reward_id {
    PF_REW_LIN_VEL_Z = 0, PF_REW_ANG_VEL_XY, PF_REW_ORIENTATION, PF_REW_BASE_HEIGHT, PF_REW_TORQUES,
    PF_REW_DOF_VEL, PF_REW_DOF_ACC, PF_REW_ACTION_RATE, PF_REW_COLLISION, PF_REW_TERMINATION,
    PF_REW_DOF_POS_LIMITS, PF_REW_DOF_VEL_LIMITS, PF_REW_TORQUE_LIMITS, PF_REW_TRACKING_LIN_VEL,
    PF_REW_TRACKING_ANG_VEL, PF_REW_FEET_AIR_TIME, PF_REW_NO_FLY, PF_REW_UNBALANCE_FEET_AIR_TIME,
    PF_REW_UNBALANCE_FEET_HEIGHT, PF_REW_STUMBLE, PF_REW_STAND_STILL, PF_REW_FEET_CONTACT_FORCES,
    PF_REW_FEET_DISTANCE, PF_REW_SURVIVAL, PF_REW_COUNT
};

enum pf_control_type { PF_CONTROL_P = 0, PF_CONTROL_V = 1, PF_CONTROL_T = 2 };   /* PF:537-546 */
enum pf_terrain_kind { PF_TERRAIN_PLANE = 0, PF_TERRAIN_HEIGHTFIELD = 1 };     /* 'plane' vs 'heightfield'/'trimesh' */
/* (x / horizontal_scale).long() at PF:1102: CPU eager divides, CUDA eager multiplies by fp32 1/scale. */
enum pf_index_div_mode { PF_DIV_TRUE = 0, PF_DIV_RECIPROCAL = 1 };

/* Constants of one task, flattened from the nested cfg classes (pointfoot_rough_config.py, PF:999-1009,
 * PF:698-820).  POD, copied into the ctx at pf_ctx_create. */
typedef struct pf_params {
    int32_t abi_version;              /* = PF_ABI_VERSION */
    int32_t num_envs;                 /* envs on THIS device */
    int64_t env_id_offset;            /* global id of local env 0 (multi-GPU block sharding) */
    int32_t num_dof, num_bodies;      /* 6, 9 */
    int32_t num_commands;             /* 3 flat / 4 rough */
    int32_t num_obs, num_priv_obs;    /* 27, 27/148; num_priv_obs = 0 => no privileged buffer */
    uint32_t seed;                    /* RNG contract seed (train cfg seed) */

    /* control, PF:523-547 */
    int32_t control_type, decimation;
    float action_scale, sim_dt, dt;   /* dt = decimation * sim_dt, PF:1000 */
    float clip_actions, clip_obs;     /* PF:176, 189 */
    float p_gains[PF_MAX_DOF], d_gains[PF_MAX_DOF], torque_limits[PF_MAX_DOF];
    float default_dof_pos[PF_MAX_DOF];
    float dof_pos_limits[PF_MAX_DOF][2], dof_vel_limits[PF_MAX_DOF];   /* soft limits, PF:449-463 */

    /* body index lists, PF:951-968 */
    int32_t feet_indices[PF_NUM_FEET];
    int32_t num_termination, termination_indices[PF_MAX_BODIES];
    int32_t num_penalised, penalised_indices[PF_MAX_BODIES];

    /* episode / commands, PF:479-493, 1006-1009 */
    float max_episode_length;         /* np.ceil(episode_length_s/dt), compared as float: PF:246 */
    float max_episode_length_s;
    int32_t resample_period;          /* int(resampling_time/dt), PF:484 */
    int32_t heading_command;          /* PF:487 */
    int32_t resample_commands;        /* 0 = PF (no-op, PF:498-521), 1 = LR:337-351 */
    float cmd_range[4][2];            /* lin_vel_x, lin_vel_y, ang_vel_yaw, heading */

    /* observations, PF:320-382, 652-695 */
    float obs_scale_ang_vel, obs_scale_dof_pos, obs_scale_dof_vel, obs_scale_height;
    float commands_scale[3];
    int32_t add_noise;
    int32_t measure_heights_actor, measure_heights_critic;

    /* terrain + height scan, PF:1032-1112 */
    int32_t terrain_kind, index_div_mode;
    int32_t hf_rows, hf_cols;         /* height_samples.shape */
    float border_size, horizontal_scale, vertical_scale;
    int32_t num_height_points;        /* P; 0 when nothing measures heights */

    /* resets + curriculum, PF:249-299, 549-588, 613-636 */
    int32_t custom_origins, terrain_curriculum;
    int32_t max_terrain_level, num_terrain_types;    /* terrain_origins is [max_terrain_level][num_terrain_types][3] */
    float terrain_env_length;
    float base_init_state[13];
    int32_t send_timeouts;

    /* rewards, PF:301-318, 796-820: terms in summation order (alphabetical), scales already * dt */
    int32_t num_reward_terms;                          /* excludes 'termination' */
    int32_t reward_ids[PF_MAX_REWARD_TERMS];
    float reward_scales[PF_MAX_REWARD_TERMS];
    int32_t has_termination_term; float termination_scale;   /* added after the clip, PF:315-318 */
    int32_t only_positive_rewards;
    int32_t include_feet_height_rewards;               /* PF:232-237 */
    float base_height_target, soft_dof_vel_limit, soft_torque_limit, tracking_sigma, max_contact_force;
    float min_feet_air_time, max_feet_air_time, min_feet_distance;

    /* push, PF:590-611 */
    float max_push_vel_xy;
} pf_params;

/* Device pointers of one env instance.  Shapes in [] with N = num_envs, D = num_dof, B = num_bodies,
 * F = 2 feet, C = num_commands, P = num_height_points, R = num_reward_terms (+1 if termination term). */
typedef struct pf_env_buffers {
    /* PhysX-owned, AoS, written in place for reset rows (PF:713-724) */
    float* root_states;               /* [N,13] pos3 quat4(xyzw) linvel3 angvel3 */
    float* dof_state;                 /* [N*D,2] (pos, vel) */
    const float* contact_forces;      /* [N*B,3] */
    const float* rigid_body_states;   /* [N*B,13] */
    /* env-owned */
    float* actions;                   /* [N,D] clipped actions (PF:177) */
    float* torques;                   /* [N,D] */
    float* commands;                  /* [N,C] */
    int64_t* episode_length_buf;      /* [N] */
    float* base_lin_vel; float* base_ang_vel; float* projected_gravity;   /* [N,3] each */
    float* last_actions; float* last_dof_vel;    /* [N,D] */
    float* last_root_vel;             /* [N,6] */
    float* feet_air_time; float* last_feet_air_time;                    /* [N,F] */
    float* current_max_feet_height; float* last_max_feet_height; float* feet_height;   /* [N,F] */
    uint8_t* contact_filt; uint8_t* first_contact; uint8_t* last_contacts;      /* [N,F] torch.bool */
    float* episode_sums;              /* [R,N] one row per term, rows in reward order (termination last) */
    float* rew_buf;                   /* [N] */
    uint8_t* reset_buf; uint8_t* time_out_buf; uint8_t* terminate;      /* [N] torch.bool */
    float* obs_buf;                   /* [N,num_obs] */
    float* priv_obs_buf;              /* [N,num_priv_obs] or NULL */
    float* measured_heights;          /* [N,P] or NULL (not materialised) */
    /* terrain (rough only, else NULL) */
    const int16_t* height_samples;    /* [hf_rows,hf_cols] */
    const float* terrain_origins;     /* [max_terrain_level,num_terrain_types,3] */
    int64_t* terrain_levels;          /* [N] */
    const int64_t* terrain_types;     /* [N] */
    float* env_origins;               /* [N,3] */
    /* per-step outputs for the caller / gym setters */
    int32_t* reset_ids;               /* [N] ascending ids of reset envs (PF:221, 561, 585) */
    int32_t* reset_count;             /* [1] device count */
    float* episode_stats;             /* [R+2] extras["episode"]: R term means /max_episode_length_s,
                                         terrain_level mean, reset count as float; rewritten only when >=1 reset (PF:259) */
    uint8_t* extras_time_outs;        /* [N] or NULL: extras["time_outs"] -- a copy of time_out_buf taken only on
                                         steps with >=1 reset (PF:259-260, 287-288: the dict entry is set inside
                                         reset_idx, after its early return) */
    int64_t* step_counter;            /* [1] or NULL: device copy of common_step_counter.  When non-NULL,
                                         pf_post_physics_step / pf_reset_all ignore their host argument: the step
                                         uses *step_counter + 1 and stores it back (PF:205), so that a whole step
                                         can be replayed from a CUDA graph */
} pf_env_buffers;

typedef struct pf_ctx pf_ctx;

const char* pf_last_error(void);
int pf_abi_version(void);
int pf_sizeof_params(void);        /* layout checks for FFI bindings */
int pf_sizeof_env_buffers(void);

/* Create / destroy.  noise_scale_vec: HOST float[num_priv_obs or num_obs] (PF:652-695 output, obs part then
 * privileged-extra part); height_points_xy: HOST float[P][2] grid in the base frame (PF:1032-1046). */
int pf_ctx_create(const pf_params* params, const float* noise_scale_vec_host,
                  const float* height_points_xy_host, pf_ctx** out);
int pf_ctx_destroy(pf_ctx* ctx);
/* update_command_curriculum (PF:638-650, LR:443-453) widens command_ranges["lin_vel_x"] on the host; ranges[4][2] =
 * lin_vel_x, lin_vel_y, ang_vel_yaw, heading replaces pf_params.cmd_range for every later launch of this ctx (a step
 * captured in a CUDA graph keeps the old values: re-capture). */
int pf_set_command_ranges(pf_ctx* ctx, const float* ranges_host);

/* PF:176-177 + PF:523-547 (LR:353-375).  If actions_in != NULL it is clipped to +-clip_actions into
 * bufs->actions first (the step() prologue); torques written to bufs->torques. */
int pf_pd_torques(pf_ctx* ctx, const pf_env_buffers* bufs, const float* actions_in, void* stream);

/* PF:195-227 + the obs clip of PF:189-192: everything post_physics_step does between the gym refresh calls
 * and the return of step(); `common_step_counter` is the value AFTER the increment at PF:205.  Also fills
 * reset_ids / reset_count / episode_stats (PF:221, 276-288). */
int pf_post_physics_step(pf_ctx* ctx, const pf_env_buffers* bufs, int64_t common_step_counter, void* stream);

/* pf_post_physics_step in two halves.  _begin leaves the step's tail (reset_ids, reset_count, episode_stats,
 * extras_time_outs, step_counter: PF:221, 259-288) running on a context-owned stream; obs / priv / rew / reset flags and
 * the reset rows are complete in `stream` order, so ActorCritic.act queued next overlaps it.  _join makes `stream` wait
 * for the tail: call it before reading those outputs, before the next step of this ctx and before a capture ends. */
int pf_post_physics_step_begin(pf_ctx* ctx, const pf_env_buffers* bufs, int64_t common_step_counter, void* stream);
int pf_post_physics_step_join(pf_ctx* ctx, void* stream);

/* Diagnostic (tools/step_timeline.py): install a DEVICE log of 4096 uint64 (log[0] = record count, records =
 * tag << 48 | globaltimer ns) that thread 0 of CTA 0 of the env kernels / the fused act kernel append to; NULL removes it.
 * The marks are compiled in only with -DPF_TIMELINE (PF_BUILD_TIMELINE=1 at build time); otherwise installing a log fails. */
int pf_debug_timeline(unsigned long long* device_log);
/* Diagnostic: number of kernel launches this library has issued in this process so far (a launch recorded during a stream
 * capture counts once; replays are the caller's to multiply).  bench.py's `gpu_launches` is measured with it. */
unsigned long long pf_debug_launch_count(void);
int pf_rl_debug_timeline(unsigned long long* device_log);
int pf_tt_debug_timeline(unsigned long long* device_log);   /* the same for the training GEMM's translation unit (tools/train_trace.py) */

/* PF:590-611: random push forces into forces[N,B,3] (body 0 rows; everything else zeroed).
 * max_push_force = base_mass.mean() * max_push_vel_xy / sim_dt, evaluated by the host in double (PF:592-596). */
int pf_push_forces(pf_ctx* ctx, const pf_env_buffers* bufs, float* forces, float max_push_force,
                   int64_t common_step_counter, void* stream);

/* PF:137-139 reset(): reset_idx over ALL envs (before the zero-action step). */
int pf_reset_all(pf_ctx* ctx, const pf_env_buffers* bufs, int64_t common_step_counter, void* stream);

/* Host-buffer step: the call a host without a device runtime of its own binds (cgo / ctypes drop-in), and the
 * reference's actual deployment (the fork runs PhysX on the CPU pipeline, PF:41,50: state tensors live in host
 * memory).  If `state` != NULL its non-NULL members (HOST, same AoS layouts as pf_env_buffers) are copied into the
 * device tensors first; then actions_host[N,D] -> H2D, decimation x pd_torques, post_physics_step,
 * obs / priv / rew / reset (+ torques if torques_host != NULL) -> D2H, stream sync.
 * PINNED host buffers (cudaHostAlloc / torch pin_memory) are read and written by the device directly: one gather
 * kernel in, one scatter kernel out; pageable buffers go through cudaMemcpyAsync. */
typedef struct pf_host_state {
    const float* root_states;         /* [N,13] */
    const float* dof_state;           /* [N*D,2] */
    const float* contact_forces;      /* [N*B,3] */
    const float* rigid_body_states;   /* [N*B,13]; only the feet position columns are transferred (all the path reads) */
} pf_host_state;

int pf_step_host(pf_ctx* ctx, const pf_env_buffers* bufs, const pf_host_state* state, const float* actions_host,
                 float* obs_host, float* priv_obs_host, float* rew_host, uint8_t* reset_host, float* torques_host,
                 int64_t common_step_counter, void* stream);
/* Same, but everything is only ENQUEUED on `stream` (pinned host buffers required): the caller adds what follows
 * the env step -- e.g. pf_rl_act on the device observations and the copy of the sampled actions back to the host --
 * and synchronises once.  The host buffers are valid after that synchronisation. */
int pf_step_host_async(pf_ctx* ctx, const pf_env_buffers* bufs, const pf_host_state* state, const float* actions_host,
                       float* obs_host, float* priv_obs_host, float* rew_host, uint8_t* reset_host, float* torques_host,
                       int64_t common_step_counter, void* stream);
/* pf_step_host_async in two halves.  _begin enqueues the step but leaves the results' transfer to the (pinned) host
 * buffers running on a context-owned stream, so that what the caller queues next on `stream` (pf_rl_act on the device
 * observations) overlaps it; _join makes `stream` wait for that transfer.  _join must come before the host reads the
 * outputs, before the next step and before a stream capture ends.  (Inside, the state tensors the torque law does
 * not read cross the bus on the same side stream while the torque sub-steps run.) */
int pf_step_host_begin(pf_ctx* ctx, const pf_env_buffers* bufs, const pf_host_state* state, const float* actions_host,
                       float* obs_host, float* priv_obs_host, float* rew_host, uint8_t* reset_host, float* torques_host,
                       int64_t common_step_counter, void* stream);
int pf_step_host_join(pf_ctx* ctx, void* stream);

/* ================================================================================================
 * rsl_rl v1.0.2 side of the path (third-party dependency of the reference; call sites
 * legged_gym/utils/task_registry.py:147-154, legged_gym/scripts/train.py:44, play.py:59).
 * ActorCritic.act / evaluate / act_inference, PPO.act / process_env_step / compute_returns / update,
 * RolloutStorage.add_transitions / compute_returns / mini_batch_generator.
 * ============================================================================================== */
#define PF_MAX_MLP_LAYERS 6

enum pf_rl_precision {
    PF_RL_FP32 = 0,      /* hand-written fp32 SIMT kernels: the parity path (rtol 1e-4 against torch fp32) */
    PF_RL_BF16_TC = 1    /* tensor-core paths, fp32 accumulation (rtol 2e-2): 16-bit operands -- fp16 in the inference kernels, bf16 in the PPO.update GEMMs */
};

/* Hyper-parameters and shapes (PointFootRoughCfgPPO, pointfoot_rough_config.py:216-259).  POD. */
typedef struct pf_rl_params {
    int32_t abi_version;
    int32_t num_envs;                 /* envs on THIS device */
    int64_t env_id_offset;            /* global id of local env 0 */
    int32_t num_steps;                /* num_steps_per_env (24) */
    int32_t num_obs, num_critic_obs, num_actions;
    /* MLPs: Linear(dims[i] -> dims[i+1]) for i < layers, ELU between (actor_critic.py).  dims[0] is the
     * observation width, dims[layers] the output width (num_actions / 1). */
    int32_t actor_layers, actor_dims[PF_MAX_MLP_LAYERS + 1];
    int32_t critic_layers, critic_dims[PF_MAX_MLP_LAYERS + 1];
    uint32_t seed;
    float gamma, lam, clip_param, value_loss_coef, entropy_coef, max_grad_norm;
    int32_t use_clipped_value_loss, adaptive_schedule;
    int32_t num_mini_batches, num_learning_epochs;
    int32_t precision;                /* enum pf_rl_precision */
    double desired_kl;                /* python floats upstream: kept in double so the schedule and Adam's */
    double adam_beta1, adam_beta2, adam_eps;   /* bias corrections round like torch.optim.Adam's host arithmetic */
} pf_rl_params;

/* RolloutStorage tensors (rollout_storage.py), all [T, N, .] contiguous, caller-owned device memory. */
typedef struct pf_rl_storage {
    float* observations;              /* [T,N,num_obs] */
    float* privileged_observations;   /* [T,N,num_critic_obs] or NULL (critic sees observations) */
    float* actions;                   /* [T,N,A] */
    float* rewards;                   /* [T,N,1] */
    uint8_t* dones;                   /* [T,N,1] */
    float* values;                    /* [T,N,1] */
    float* actions_log_prob;          /* [T,N,1] */
    float* mu; float* sigma;          /* [T,N,A] */
    float* returns; float* advantages;/* [T,N,1] */
} pf_rl_storage;

/* Learner state, caller-owned device memory.  `params` is ONE flat fp32 buffer in ActorCritic.parameters()
 * order: std[A], actor.{0,2,..}.weight[out,in] / .bias[out], critic.{0,2,..}.weight / .bias. */
typedef struct pf_rl_learner {
    float* params; float* grads;      /* [pf_rl_param_count] */
    float* adam_m; float* adam_v;     /* [pf_rl_param_count] exp_avg / exp_avg_sq */
    double* lr;                       /* [1] current learning rate (adaptive-KL schedule runs on the device) */
    int64_t* adam_step;               /* [1] */
    double* loss_stats;               /* [8] per minibatch: surrogate sum, value-loss sum, entropy sum, kl sum,
                                         sample count, grad-norm^2, (2 spare); written by pf_rl_minibatch_grad
                                         (grad-norm^2 by pf_rl_adam_step) */
} pf_rl_learner;

typedef struct pf_rl_ctx pf_rl_ctx;

int pf_sizeof_rl_params(void);
int pf_rl_create(const pf_rl_params* params, pf_rl_ctx** out);   /* allocates the activation workspace */
int pf_rl_destroy(pf_rl_ctx* ctx);
int64_t pf_rl_param_count(const pf_rl_ctx* ctx);
/* offset (in floats) of a tensor inside the flat parameter buffer: net 0 = actor, 1 = critic, -1 = std;
 * is_bias 0 = weight [out,in], 1 = bias [out] */
int64_t pf_rl_param_offset(const pf_rl_ctx* ctx, int net, int layer, int is_bias);
/* must be called after the caller changes `params` by other means than pf_rl_adam_step (init, load):
 * refreshes the packed 16-bit copies the tensor-core path reads */
int pf_rl_params_changed(pf_rl_ctx* ctx, const float* params, void* stream);
/* Host-only diagnostic, no device needed: can the fused tensor-core kernel (PF_RL_BF16_TC) run an MLP of this shape
 * (dims[0..layers], the ctor args of actor_critic.py ActorCritic)?  0 yes, 6 no -- the code pf_rl_create would return.
 * If `text` is non-NULL the kernel's static schedule is printed into it. */
int pf_rl_tc_plan(const int32_t* dims, int32_t layers, char* text, int32_t cap);
/* Diagnostic: CTA 0 of every following tensor-core launch writes its role timelines (4 x 512 uint64, tag << 48 |
 * SM clock) into `device_buf`; NULL switches it off (tools/tc_trace.py). */
int pf_rl_tc_trace(pf_rl_ctx* ctx, unsigned long long* device_buf);
/* Opt-in (PF_RL_BF16_TC): the fused kernel processes one DRY item (zeros for the input rows, no outputs) before its
 * programmatic-dependency wait, so that the code of all its roles has been executed once while the predecessor kernel
 * (the env step) is still running.  For chains launched back to back (a CUDA-graphed step + act); off by default. */
int pf_rl_tc_rehearse(pf_rl_ctx* ctx, int on);
/* The inference kernels of the tensor-core paths read PACKED copies of the parameters, refreshed by a packing kernel after
 * every optimizer step / pf_rl_params_changed.  Launched as programmatic dependents they stage those copies BEFORE their
 * dependency wait, under the tail of the kernel in front of them -- unless that kernel may be the packing kernel itself.
 * The context keeps track of that for what it launches; two calls cover what it cannot see:
 *   pf_rl_weights_written  after replaying a CUDA graph that contains an optimizer step / pf_rl_params_changed
 *                          (rl.PPO.update does): the next inference launch waits first;
 *   pf_rl_early_weights    (ctx, 1): the caller vouches that inference launches it CAPTURES into a graph are never the
 *                          first kernel after a repack when replayed (a step graph: env kernels, then act); without it
 *                          captured launches always wait first.
 * PF_RL_EARLY_WEIGHTS=0 in the environment disables the early staging altogether (every launch waits first). */
int pf_rl_weights_written(pf_rl_ctx* ctx);
int pf_rl_early_weights(pf_rl_ctx* ctx, int on);
/* Which kernels serve act / act_inference / evaluate of this context: 0 = fp32 SIMT chain (PF_RL_FP32), 1 = fused tcgen05
 * kernel (pf_mlp_tc.cu), 2 = register-resident warp-MMA kernel for small nets (pf_mlp_wm.cu: every hidden width a multiple of
 * 16 and <= 128, inputs <= 128, outputs <= 8 -- the pointfoot_flat policy; PF_RL_WM=0 in the environment disables it). */
int pf_rl_inference_path(const pf_rl_ctx* ctx);

/* PPO.act (ppo.py act) fused with RolloutStorage.add_transitions' copies: actor forward -> Normal(mean, std)
 * sample (RNG contract, purpose POLICY, step = policy_step) -> log-prob; critic forward; writes row `t` of the
 * storage (observations, privileged_observations, actions, values, actions_log_prob, mu, sigma) and
 * actions_out[N,A].  storage may be NULL (pure ActorCritic.act + evaluate; values_out / logp_out optional). */
int pf_rl_act(pf_rl_ctx* ctx, const float* params, const float* obs, const float* critic_obs, int32_t t,
              const pf_rl_storage* storage, int64_t policy_step, float* actions_out, float* values_out,
              float* logp_out, void* stream);
/* ActorCritic.act_inference: actions_out[rows, A] = actor(obs[rows, num_obs]); any row count (upstream accepts any
 * batch: one row in play / export checks), processed in chunks of the workspace size */
int pf_rl_act_inference(pf_rl_ctx* ctx, const float* params, const float* obs, int32_t rows, float* actions_out, void* stream);
/* ActorCritic.evaluate: values_out[rows] = critic(critic_obs[rows, num_critic_obs]) */
int pf_rl_evaluate(pf_rl_ctx* ctx, const float* params, const float* critic_obs, int32_t rows, float* values_out, void* stream);
/* PPO.process_env_step: rewards[t] = rew + gamma * values[t] * time_outs (time_outs may be NULL), dones[t] = dones */
int pf_rl_process_env_step(pf_rl_ctx* ctx, const pf_rl_storage* storage, int32_t t, const float* rew,
                           const uint8_t* dones, const uint8_t* time_outs, void* stream);
/* RolloutStorage.compute_returns, first half: GAE reverse scan (one thread per env) -> returns, raw advantages;
 * stats[3] (device, double) = sum, sum of squares, count of the raw advantages on THIS device.  Multi-GPU:
 * all-reduce `stats` (sum) between the two calls. */
int pf_rl_gae(pf_rl_ctx* ctx, const pf_rl_storage* storage, const float* last_values, double* stats, void* stream);
/* pf_rl_gae for callers that cannot allocate device memory (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * Needs no pf_rl_ctx and calls no cudaMalloc / cudaFree / allocating library routine: the per-CTA partials of the reduction
 * live in a static 64 KB device array inside the library (held by every GPU that loads it until the process exits).
 * Inputs / outputs are the rollout_storage.py tensors, [T,N,1] contiguous DEVICE memory owned by the caller: rewards, dones
 * (uint8), values in; returns and raw advantages (returns - values) out; last_values[N]; stats[3] (device, double) = sum, sum
 * of squares, count of the raw advantages, exactly what pf_rl_gae writes (feed it to pf_rl_normalize_advantages or to the
 * caller's own normalisation).  Same kernels, same bits as pf_rl_gae.
 * num_envs > PF_GAE_NOALLOC_MAX_ENVS is rejected (code 8, message in pf_last_error()).  All calls on a device share the
 * array: it keeps the previous call's contents (none of which a later call reads), and calls that OVERLAP on the GPU --
 * different streams with no ordering between them -- race; order them on one stream or with events. */
#define PF_GAE_NOALLOC_MAX_ENVS 1048576
int pf_rl_gae_noalloc(const float* rewards, const uint8_t* dones, const float* values, const float* last_values,
                      int32_t num_envs, int32_t num_steps, float gamma, float lam, float* returns, float* advantages,
                      double* stats, void* stream);
/* second half: advantages = (adv - mean) / (std_unbiased + 1e-8) from stats */
int pf_rl_normalize_advantages(pf_rl_ctx* ctx, const pf_rl_storage* storage, const double* stats, void* stream);
/* mini_batch_generator's shuffle: the Feistel permutation contract (oracle/rsl_rl_ref.py feistel_permutation)
 * for this `iteration`; kept inside the ctx.  Once per PPO.update. */
int pf_rl_make_permutation(pf_rl_ctx* ctx, int64_t iteration, void* stream);
/* copies the current permutation (num_mini_batches * minibatch_size int32 indices into the flattened [T*N]
 * storage) to out: what mini_batch_generator's `indices` tensor would hold */
int pf_rl_get_permutation(pf_rl_ctx* ctx, int32_t* out, void* stream);
/* One minibatch of PPO.update up to (not including) the optimizer step: gather by permutation, actor + critic
 * forward, adaptive-KL statistic, clipped surrogate + clipped value + entropy loss, backward.  Writes
 * learner->grads (d loss / d params, mean over THIS device's minibatch) and learner->loss_stats.
 * Multi-GPU: all-reduce grads (mean) and loss_stats (sum) before pf_rl_adam_step. */
int pf_rl_minibatch_grad(pf_rl_ctx* ctx, const pf_rl_storage* storage, const pf_rl_learner* learner,
                         int32_t minibatch, void* stream);
/* adaptive-KL learning-rate update (ppo.py update), clip_grad_norm_(max_grad_norm), Adam step; also refreshes
 * the packed 16-bit parameter copies. */
int pf_rl_adam_step(pf_rl_ctx* ctx, const pf_rl_learner* learner, void* stream);

/* ================================================================================================
 * Actuator-network torques: legged_gym/envs/anymal_c/anymal.py:62-78, the only other
 * _compute_torques of the reference (cfg.control.use_actuator_network, anymal_c_rough_config.py:68-69).
 * ============================================================================================== */
#define PF_ACTUATOR_MAX_DOF 12
#define PF_ACTUATOR_NET_WEIGHTS 969     /* floats: the module's state_dict without in_scale / out_scale, in its order:
                                            lstm.weight_ih_l0[32,2] weight_hh_l0[32,8] bias_ih_l0[32] bias_hh_l0[32]
                                            weight_ih_l1[32,8] weight_hh_l1[32,8] bias_ih_l1[32] bias_hh_l1[32]
                                            linear.weight[1,8] linear.bias[1] */
typedef struct pf_actuator_net_params {
    int32_t num_envs, num_dof;
    float action_scale;                          /* cfg.control.action_scale (anymal.py:66) */
    float in_scale[2], out_scale;                /* the module's buffers (anydrive_v3_lstm.pt: (2.0, 0.25), 20.0) */
    float default_dof_pos[PF_ACTUATOR_MAX_DOF];
} pf_actuator_net_params;
/* anymal.py:62-72: torques[N,A] = out_scale * Linear(LSTM((action * action_scale + default - q, qd) * in_scale, (h, c))),
 * h / c [2, N*A, 8] updated in place (anymal.py:48-49 sea_hidden_state / sea_cell_state).  weights: DEVICE float[969].
 * Context-free: allocates nothing, retains nothing. */
int pf_actuator_net_torques(const pf_actuator_net_params* params, const float* weights, const float* actions,
                            const float* dof_state, float* hidden, float* cell, float* torques, void* stream);
/* anymal.py:52-56 (reset_idx): zero the recurrent state of the envs in reset_ids[0 .. *reset_count) (both DEVICE, as
 * pf_post_physics_step leaves them: no host sync). */
int pf_actuator_net_reset(int32_t num_envs, int32_t num_dof, const int32_t* reset_ids, const int32_t* reset_count,
                          float* hidden, float* cell, void* stream);

/* ================================================================================================
 * LeggedRobot base class: legged_gym/envs/base/legged_robot.py (LR), the env of the a1 /
 * anymal / cassie tasks.  Differences from PointFoot that matter on the step path: observations lead with base_lin_vel and
 * put the commands in front of the joints (48 columns + the height scan in ONE buffer, LR:209-226); commands are really
 * resampled (LR:337-351, from the callback and again in reset_idx); the push overwrites the root xy velocity (LR:414-419); no
 * feet state machine in the step -- feet_air_time / last_contacts are updated inside _reward_feet_air_time (LR:882-893);
 * _reward_stand_still is the joint-pose variant (LR:900-902); reset_idx zeroes fewer buffers (LR:168-173).
 * Context-free entry points: parameters and buffers are passed on every call, nothing is allocated or retained.
 * ============================================================================================== */
#define PF_LR_MAX_DOF 12
#define PF_LR_MAX_BODIES 24
#define PF_LR_MAX_FEET 4

typedef struct pf_lr_params {
    int32_t abi_version, num_envs;
    int64_t env_id_offset;
    int32_t num_dof, num_bodies, num_feet, num_commands, num_obs;
    uint32_t seed;
    int32_t control_type, decimation;                                 /* LR:353-375 */
    float action_scale, sim_dt, dt, clip_actions, clip_obs;
    float p_gains[PF_LR_MAX_DOF], d_gains[PF_LR_MAX_DOF], torque_limits[PF_LR_MAX_DOF], default_dof_pos[PF_LR_MAX_DOF];
    float dof_pos_limits[PF_LR_MAX_DOF][2], dof_vel_limits[PF_LR_MAX_DOF];   /* soft limits, LR:291-305 */
    int32_t feet_indices[PF_LR_MAX_FEET];
    int32_t num_termination, termination_indices[PF_LR_MAX_BODIES];
    int32_t num_penalised, penalised_indices[PF_LR_MAX_BODIES];
    float max_episode_length, max_episode_length_s;                  /* LR:735-736 */
    int32_t resample_period, heading_command;                         /* LR:325-330 */
    float cmd_range[4][2];                                            /* command_ranges: lin_vel_x, lin_vel_y, ang_vel_yaw, heading */
    float obs_scale_lin_vel, obs_scale_ang_vel, obs_scale_dof_pos, obs_scale_dof_vel, obs_scale_height;
    float commands_scale[3];
    int32_t add_noise, measure_heights;
    int32_t terrain_kind, index_div_mode, hf_rows, hf_cols;
    float border_size, horizontal_scale, vertical_scale;
    int32_t num_height_points;
    int32_t custom_origins, terrain_curriculum, max_terrain_level, num_terrain_types;
    float terrain_env_length;
    float base_init_state[13];
    int32_t send_timeouts;
    int32_t num_reward_terms, reward_ids[PF_MAX_REWARD_TERMS];       /* enum pf_reward_id, in summation (dir()) order */
    float reward_scales[PF_MAX_REWARD_TERMS];                        /* already * dt (LR:553) */
    int32_t has_termination_term; float termination_scale;
    int32_t only_positive_rewards;
    float base_height_target, soft_dof_vel_limit, soft_torque_limit, tracking_sigma, max_contact_force;
    float max_push_vel_xy;                                            /* LR:417 */
} pf_lr_params;

/* Device pointers; N = num_envs, D = num_dof, B = num_bodies, F = num_feet, P = num_height_points, R = reward rows. */
typedef struct pf_lr_buffers {
    float* root_states;               /* [N,13]  PhysX-owned: reset rows and the push are written in place */
    float* dof_state;                 /* [N*D,2] */
    const float* contact_forces;      /* [N*B,3] */
    float* actions; float* torques;   /* [N,D] */
    float* commands;                  /* [N,4] */
    int64_t* episode_length_buf;      /* [N] */
    float* base_lin_vel; float* base_ang_vel; float* projected_gravity;   /* [N,3] */
    float* last_actions; float* last_dof_vel;                             /* [N,D] */
    float* last_root_vel;             /* [N,6] */
    float* feet_air_time;             /* [N,F] */
    uint8_t* last_contacts;           /* [N,F] torch.bool */
    float* episode_sums;              /* [R,N], termination last */
    float* rew_buf;                   /* [N] */
    uint8_t* reset_buf; uint8_t* time_out_buf;                            /* [N] torch.bool */
    float* obs_buf;                   /* [N,num_obs] */
    float* measured_heights;          /* [N,P]; required when measure_heights (LR:331-332 keeps it as an attribute) */
    const int16_t* height_samples; const float* terrain_origins;
    int64_t* terrain_levels; const int64_t* terrain_types;
    float* env_origins;               /* [N,3] */
    int32_t* reset_ids; int32_t* reset_count;                             /* ascending ids (LR:128), device count */
    float* episode_stats;             /* [R+2]: term means / max_episode_length_s, terrain level mean, reset count (LR:175-184) */
    uint8_t* extras_time_outs;        /* [N] or NULL (LR:186-187) */
    const float* noise_scale_vec;     /* [num_obs] DEVICE: _get_noise_scale_vec, LR:455-478 */
    const float* height_points_xy;    /* [P][2] DEVICE: _init_height_points, LR:761-775 */
} pf_lr_buffers;

int pf_lr_sizeof_params(void);
int pf_lr_sizeof_buffers(void);
/* LR:86-87 (action clip, when actions_in != NULL) + LR:353-375 _compute_torques */
int pf_lr_pd_torques(const pf_lr_params* params, const pf_lr_buffers* bufs, const float* actions_in, void* stream);
/* LR:105-136 post_physics_step between the gym refresh calls and the return of step(), + the obs clip of LR:99-102.
 * `push` != 0: this is a step on which _push_robots fires (LR:333: common_step_counter % push_interval == 0; the host decides). */
int pf_lr_post_physics_step(const pf_lr_params* params, const pf_lr_buffers* bufs, int64_t common_step_counter,
                            int32_t push, void* stream);
/* base_task.py:105-109 reset(): reset_idx over every env (LR:145-188; time_out_buf is read as it stands for
 * extras["time_outs"]).  The zero-action step that follows is the caller's. */
int pf_lr_reset_all(const pf_lr_params* params, const pf_lr_buffers* bufs, int64_t common_step_counter, void* stream);

/* ================================================================================================
 * Terrain generation: legged_gym/utils/terrain.py::Terrain (terrain.py:38-187) + the sub-terrain
 * primitives of isaacgym.terrain_utils it calls, as ONE raster kernel over the whole int16 height field (the field the
 * height scan of the step path gathers from), a per-sub-terrain kernel for env_origins (terrain.py:160-167) and the
 * height-field -> triangle-mesh conversion handed to PhysX for mesh_type 'trimesh' (terrain.py:66-70).
 * The host reduces every sub-terrain to integers exactly as terrain_utils does (int(step_height / vertical_scale) ...,
 * pointfoot_b200/terrain.py) and passes one pf_subterrain per map cell; random draws are evaluated per pixel from the
 * counter-based contract (purpose PF_RNG_TERRAIN, "env" = row * num_cols + col, lanes as listed below).
 * ============================================================================================== */
enum pf_subterrain_kind {
    PF_SUB_SLOPE = 0,        /* pyramid_sloped_terrain: p0 max_height (signed counts), p1 platform half size [px] */
    PF_SUB_NOISY_SLOPE = 1,  /* + random_uniform_terrain: p2 lowest height, p3 number of heights, p4 step [counts], p5 / p6 coarse grid
                                    nx / ny; draw lane a * ny + b for grid node (a, b) */
    PF_SUB_STAIRS = 2,       /* pyramid_stairs_terrain: p0 step width [px], p1 step height (signed counts), p2 number of steps */
    PF_SUB_OBSTACLES = 3,    /* discrete_obstacles_terrain: p0 max height, p1 / p2 min / max size [px], p3 number of rectangles,
                                    p4..p7 platform x1 x2 y1 y2; lanes 5 r + {0 width, 1 length, 2 start_i, 3 start_j, 4 height} */
    PF_SUB_STONES = 4,       /* stepping_stones_terrain: p0 stone size, p1 stone distance [px], p2 max height, p3 depth (counts),
                                    p4..p7 platform x1 x2 y1 y2; lanes 128 band + {0 offset, 1 first piece, 2 + s stone s} */
    PF_SUB_GAP = 5,          /* gap_terrain (terrain.py:169-181): p0..p3 outer x1 x2 y1 y2 (-1000), p4..p7 inner (0) */
    PF_SUB_PIT = 6           /* pit_terrain (terrain.py:183-190): p0 height (= -depth counts), p4..p7 x1 x2 y1 y2 */
};

typedef struct pf_subterrain { int32_t kind; int32_t p[15]; } pf_subterrain;

typedef struct pf_terrain_params {
    int32_t abi_version;
    int32_t num_rows, num_cols;        /* sub-terrains: levels x types */
    int32_t cell_px;                   /* width_per_env_pixels = length_per_env_pixels (the reference builds square sub-terrains, terrain.py:110-114) */
    int32_t border_px, tot_rows, tot_cols;
    uint32_t seed;
    int32_t origin_patch[4];           /* x1 x2 y1 y2 of the patch whose max is env_origin_z (terrain.py:163-166) */
    double env_length, env_width, vertical_scale;
} pf_terrain_params;

int pf_terrain_sizeof_params(void);
/* subs: DEVICE pf_subterrain[num_rows * num_cols] (row-major); height_field: DEVICE int16[tot_rows, tot_cols];
 * env_origins: DEVICE double[num_rows, num_cols, 3] */
int pf_terrain_generate(const pf_terrain_params* params, const pf_subterrain* subs, int16_t* height_field, double* env_origins,
                        void* stream);
/* isaacgym.terrain_utils.convert_heightfield_to_trimesh: vertices fp32[rows * cols, 3], triangles uint32[2 (rows-1)(cols-1), 3];
 * slope_threshold < 0: no correction */
int pf_terrain_trimesh(const int16_t* height_field, int32_t rows, int32_t cols, double horizontal_scale, double vertical_scale,
                       double slope_threshold, float* vertices, uint32_t* triangles, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* POINTFOOT_B200_H */
