/* ppo_b200.h — C ABI of libppo_b200.so, the B200-native rollout-and-update engine for ppo.cpp's MuJoCo path.
 *
 * The reference has no FFI today: the seam this library sits behind is the C++ class interface of gymcpp plus
 * the two mains.  Each entry point below names the reference interface it replaces (paths relative to the
 * reference repo).  Plain pointers and sizes only; every call returns 0 on success or a negative pb_status and
 * leaves a message retrievable with pb_last_error().  A handle is not thread-safe; all work of a handle is
 * enqueued on its own CUDA stream (pb_env_stream) and the *_host variants synchronise before returning.
 */
#ifndef PPO_B200_H
#define PPO_B200_H

#ifdef __cplusplus
extern "C" {
#endif

typedef struct pb_env pb_env;

enum pb_status {
  PB_OK = 0,
  PB_ERR_BAD_ENV = -1,      /* unknown env id            (reference: exit code 1, src/ppo_continuous_action.cpp:330-335) */
  PB_ERR_BAD_DEVICE = -2,   /* no such CUDA device        (reference: exit codes 2/3, src/ac_ppo_continuous_action.cpp:452,464) */
  PB_ERR_CUDA = -3,         /* CUDA runtime error         (reference: CHECK_CUDA throws, include/defines.h:37-48) */
  PB_ERR_ARG = -4,          /* bad argument */
  PB_ERR_MODEL = -5,        /* compiled model does not match the kernels' compile-time dimensions (task header asserts) */
  PB_ERR_COMM = -6          /* NCCL failure               (reference: exit code 6, src/ac_ppo_continuous_action.cpp:471-478) */
};

/* wrapper chain selector */
enum { PB_CHAIN_STATS_ONLY = 0, /* ac_ppo make_env: RecordEpisodeStatistics only (src/ac_ppo_continuous_action.cpp:50-53) */
       PB_CHAIN_PPO = 1 };      /* ppo make_env: + NormalizeObservation, clamp, NormalizeReward, clamp (src/ppo_continuous_action.cpp:41-49) */

const char *pb_last_error(void);

/* ---- vector env: replaces ParVectorEnv / SeqVectorEnv (libs/gymcpp/gym.h:75-164, 276-366) -------------------- */

/* Builds num_envs copies of task `env_id` ("HalfCheetah-v5", "Hopper-v5", "Ant-v5", "Humanoid-v4") with the chosen
 * wrapper chain on CUDA device `device`.  Replaces the env_array construction + `ParVectorEnv(env_array, clip_actions)`
 * (src/ppo_continuous_action.cpp:298-354) and MujocoEnv's constructor (libs/gymcpp/mujoco/mujoco_env.h:33-48). */
int pb_env_create(const char *env_id, int num_envs, int device, int chain, float gamma, int clip_actions, pb_env **out);
int pb_env_destroy(pb_env *e);

/* get_num_envs / get_observation_space / get_action_space / get_action_space_min / _max (gym.h:104-129,320-322) */
int pb_env_num_envs(const pb_env *e);
int pb_env_obs_dim(const pb_env *e);
int pb_env_act_dim(const pb_env *e);
float pb_env_act_min(const pb_env *e);
float pb_env_act_max(const pb_env *e);
void *pb_env_stream(const pb_env *e); /* cudaStream_t the handle enqueues on */

/* Tensor reset(const int seed)  (gym.h:310-318): env i is reset with seed+i (reseeded only if seed+i > 0), autoreset
 * flags cleared.  obs_out: [num_envs, obs_dim] float32, DEVICE pointer; asynchronous on the handle's stream. */
int pb_env_reset(pb_env *e, int seed, float *obs_out);

/* tuple<obs, rewards, terminations, truncations, infos> step(const Tensor& actions)  (gym.h:324-365): clamps the
 * actions if clip_actions, next-step autoreset, flags as 0.0/1.0 floats.  `infos` (vector<optional<env_info>>,
 * gym.h:19-24) becomes three arrays: has_info[i] != 0 iff env i finished an episode, with info_r/info_l its return
 * and length (the wall-clock field t is not reproduced).  All pointers are DEVICE pointers of num_envs rows;
 * asynchronous on the handle's stream.  has_info/info_r/info_l may be NULL. */
int pb_env_step(pb_env *e, const float *actions, float *obs_out, float *rew_out, float *term_out, float *trunc_out,
                int *has_info, float *info_r, int *info_l);

/* Same two calls with HOST buffers: what the reference mains do today (actions arrive on the CPU,
 * src/ppo_continuous_action.cpp:402; results are CPU tensors).  Copies in/out and synchronises. */
int pb_env_reset_host(pb_env *e, int seed, float *obs_out);
int pb_env_step_host(pb_env *e, const float *actions, float *obs_out, float *rew_out, float *term_out, float *trunc_out,
                     int *has_info, float *info_r, int *info_l);
int pb_env_sync(pb_env *e);

/* ---- state access (testing / checkpointing): the mjData fields the tasks read directly -------------------------- */
/* qpos[nq], qvel[nv], qacc_warmstart[nv] of env i as doubles (absolute root translation included). */
int pb_env_get_state(pb_env *e, int i, double *qpos, double *qvel, double *qacc_warmstart);
int pb_env_set_state(pb_env *e, int i, const double *qpos, const double *qvel, const double *qacc_warmstart);
int pb_env_model_dims(const pb_env *e, int *nq, int *nv, int *nu);
/* mj_forward on every env at its stored state (the `mj_forward` of mujocoReset, e.g. half_cheetah_v5.h:71-75, without
 * the noise): refreshes qacc_warmstart and the stale x reference.  raw_obs_out (DEVICE, may be NULL) receives the
 * task's unwrapped observation. */
int pb_env_forward(pb_env *e, float *raw_obs_out);
/* Per-env NormalizeObservation / NormalizeReward statistics (libs/gymcpp/wrappers/stateful_observation.h:25-28,64-84,
 * stateful_reward.h:27-33,72-91), which the reference keeps inside the wrapper objects and never saves: read env i's, or
 * overwrite env i's (i >= 0) / every env's (i < 0).  HOST pointers; ob_mean/ob_var are [obs_dim], rw4 = {mean, var, acc, count}
 * (may be NULL).  Used by checkpoints of a trained policy (the policy is only meaningful together with its normaliser). */
int pb_env_get_norm_stats(pb_env *e, int i, float *ob_mean, float *ob_var, float *ob_count, float *rw4);
int pb_env_set_norm_stats(pb_env *e, int i, const float *ob_mean, const float *ob_var, float ob_count, const float *rw4);
/* counters[0] = forward-dynamics evaluations so far, counters[1] = env steps whose contact/row caps overflowed */
int pb_env_counters(pb_env *e, unsigned long long *counters2);
/* solver diagnostics (PGS models): counters[0] = Gauss-Seidel sweeps so far, counters[1] = row steps (sweeps x rows); divide
 * by pb_env_counters()[0] for the per-evaluation means that tools/ and DESIGN.md quote next to the oracle's. */
int pb_env_solver_counters(pb_env *e, unsigned long long *counters2);


/* ---- fused rollout-and-update engine: replaces the body of main()'s iteration loop ------------------------------ */
/* The reference keeps this loop in main (src/ppo_continuous_action.cpp:375-586); the entry points below are what the
 * new ppo_continuous_action CLI calls.  All device work is enqueued on the trainer's stream. */
typedef struct pb_trainer pb_trainer;

typedef struct pb_ppo_config {       /* GlobalConfig of src/ppo_continuous_action.cpp:51-81 (same names, same defaults) */
  const char *env_id;
  int num_envs;                      /* envs on THIS device (num_envs_per_device, src/ac_ppo_continuous_action.cpp:398-407) */
  int num_steps;
  int device;
  int seed;
  float gamma, gae_lambda;
  int num_minibatches, update_epochs;
  int norm_adv;
  float clip_coef;
  int clip_vloss;
  float ent_coef, vf_coef, max_grad_norm, adam_eps;
  int clip_actions;
  int rank, world_size;              /* data-parallel replica id / count (1 = single GPU) */
} pb_ppo_config;

typedef struct pb_ppo_stats {        /* the scalars main() logs (src/ppo_continuous_action.cpp:426-432,575-583) */
  float value_loss, policy_loss, entropy, old_approx_kl, approx_kl, clipfrac; /* last minibatch, except clipfrac = mean */
  float episodic_return_sum, episodic_length_sum; int episodes;  /* finished episodes since the last call */
  float rollout_ms, gae_ms, update_ms;                           /* CUDA-event times of the last iteration */
  unsigned long long forward_evals, overflow_envs;               /* simulator counters (cumulative) */
} pb_ppo_stats;

void pb_ppo_config_default(pb_ppo_config *cfg);
int pb_trainer_create(const pb_ppo_config *cfg, pb_trainer **out);
int pb_trainer_destroy(pb_trainer *t);
int pb_trainer_num_params(const pb_trainer *t);
/* flat fp32 parameters, module by module in declaration order: critic.{0,2,4}.{weight,bias}, actor_mean.{0,2,4}.{weight,bias},
 * actor_logstd (src/ppo_continuous_action.cpp:124-137); HOST pointers. */
int pb_trainer_get_params(pb_trainer *t, float *host_out);
int pb_trainer_set_params(pb_trainer *t, const float *host_in);
int pb_trainer_reset(pb_trainer *t, int seed);               /* envs->reset(seed), :369 */
int pb_trainer_rollout(pb_trainer *t);                       /* the num_steps loop, :387-408 */
int pb_trainer_gae(pb_trainer *t);                           /* :447-467 */
int pb_trainer_update(pb_trainer *t, float lr);              /* all epochs x minibatches, :489-542 */
int pb_trainer_iteration(pb_trainer *t, float lr);           /* rollout + gae + update */
int pb_trainer_stats(pb_trainer *t, pb_ppo_stats *out);      /* synchronises */
int pb_trainer_sync(pb_trainer *t);
void *pb_trainer_stream(const pb_trainer *t);
pb_env *pb_trainer_env(pb_trainer *t);
/* One optimiser step on an explicit minibatch (HOST int32 indices into the flattened [T*N] rollout), for parity
 * tests against torch autograd: same math as one pass of the loop body :492-541. */
int pb_trainer_update_minibatch(pb_trainer *t, const int *idx_host, int n, float lr);
/* Device pointer and element count of a named rollout/update buffer ("obs","actions","logprobs","rewards","dones",
 * "values","advantages","returns","next_obs","grad"); for tests and checkpointing. */
int pb_trainer_buffer(pb_trainer *t, const char *name, void **dev_ptr, long long *count);
/* Copy host data into a named buffer (float32). */
int pb_trainer_write_buffer(pb_trainer *t, const char *name, const float *host, long long count);
int pb_trainer_read_buffer(pb_trainer *t, const char *name, float *host, long long count);

/* Final evaluation of src/ppo_continuous_action.cpp:589-626: envs->reset(eval_seed) on the TRAINING envs (their normaliser
 * statistics are kept), sampled actions, until `num_eval_runs` episodes have finished (all episodes finishing in the same
 * step are taken, in env order, like the reference's loop over infos).  returns_out: HOST array of capacity `cap`;
 * *n_out = episodes written (>= num_eval_runs unless cap is smaller). */
int pb_trainer_eval(pb_trainer *t, int eval_seed, int num_eval_runs, float *returns_out, int cap, int *n_out);
/* Checkpoint: parameters, Adam moments + step count, per-env normaliser statistics (which the reference loses,
 * src/ppo_continuous_action.cpp:595-597) in one flat binary file; replaces save_state (:173-180).  Load requires the same
 * env id / num_envs. */
int pb_trainer_save(pb_trainer *t, const char *path);
int pb_trainer_load(pb_trainer *t, const char *path);

/* ---- ac_ppo_continuous_action engine: replaces the body of src/ac_ppo_continuous_action.cpp's iteration loop (:624-950) ------ */
/* Agent of :150-266 (fixed obs mean/std, Linear-LayerNorm-ReLU x2 critic and actor encoder, Beta policy); env chain of :50-53
 * (episode statistics only); per-rank permutation generator seed*1500+rank (:618); cross-rank advantage moments (:830-849)
 * and gradient average (:878-885) over NCCL.  Collection is one batched device step for all envs of the rank instead of one
 * host thread per env (:641-698). */
typedef struct pb_ac_trainer pb_ac_trainer;
typedef struct pb_ac_config {        /* GlobalConfig of src/ac_ppo_continuous_action.cpp:55-100 (same names, same defaults) */
  const char *env_id;
  int num_envs;                      /* num_envs_per_device (:398-407) */
  int num_steps;
  int device;
  int seed;
  float gamma, gae_lambda;
  int num_minibatches, update_epochs;
  int norm_adv;
  float clip_coef;
  int clip_vloss;
  float ent_coef, vf_coef, max_grad_norm, adam_eps;
  int clip_actions;
  int rank, world_size;
  const float *obs_mean, *obs_std;   /* [obs_dim] HOST, fixed observation statistics (:496-534); NULL = zeros / ones */
} pb_ac_config;
void pb_ac_config_default(pb_ac_config *cfg);
int pb_ac_trainer_create(const pb_ac_config *cfg, pb_ac_trainer **out);
int pb_ac_trainer_destroy(pb_ac_trainer *t);
int pb_ac_trainer_num_params(const pb_ac_trainer *t);
/* flat fp32 parameters in torch registration order: critic.{0,1,3,4,6}, actor_mean.{0,1,3,4}, dist_alpha.0, dist_beta.0
 * (weight then bias each; :159-186); HOST pointers. */
int pb_ac_trainer_get_params(pb_ac_trainer *t, float *host_out);
int pb_ac_trainer_set_params(pb_ac_trainer *t, const float *host_in);
int pb_ac_trainer_reset(pb_ac_trainer *t, int seed);         /* envs[i]->reset(seed + i), :605-606 */
int pb_ac_trainer_rollout(pb_ac_trainer *t);                 /* collection, :641-715 */
int pb_ac_trainer_gae(pb_ac_trainer *t);                     /* :759-779 */
int pb_ac_trainer_update(pb_ac_trainer *t, float lr);        /* :800-890 */
int pb_ac_trainer_iteration(pb_ac_trainer *t, float lr);
int pb_ac_trainer_stats(pb_ac_trainer *t, pb_ppo_stats *out);
int pb_ac_trainer_sync(pb_ac_trainer *t);
void *pb_ac_trainer_stream(const pb_ac_trainer *t);
pb_env *pb_ac_trainer_env(pb_ac_trainer *t);
int pb_ac_trainer_update_minibatch(pb_ac_trainer *t, const int *idx_host, int n, float lr);
int pb_ac_trainer_forward(pb_ac_trainer *t);                 /* logits ("zab") and values ("vbuf") of the current next_obs */
long long pb_ac_trainer_buffer_count(pb_ac_trainer *t, const char *name);
int pb_ac_trainer_write_buffer(pb_ac_trainer *t, const char *name, const float *host, long long count);
int pb_ac_trainer_read_buffer(pb_ac_trainer *t, const char *name, float *host, long long count);
/* final evaluation with action = Beta mean (:965-1001) */
int pb_ac_trainer_eval(pb_ac_trainer *t, int eval_seed, int num_eval_runs, float *returns_out, int cap, int *n_out);
/* --estimate_mean_std (:663-665, 956-963): per-column sum x / sum x^2 (sums_host[2*o], [2*o+1]) over the rollout buffer's
 * T*N observation rows of this rank (the reference samples env 0 only); the caller merges iterations and ranks. */
int pb_ac_trainer_obs_sums(pb_ac_trainer *t, double *sums_host, long long *rows_out);
/* replace the fixed observation statistics of the agent (:496-534) */
int pb_ac_trainer_set_obs_stats(pb_ac_trainer *t, const float *mean_host, const float *std_host);
int pb_ac_trainer_obs_dim(const pb_ac_trainer *t);
int pb_ac_trainer_act_dim(const pb_ac_trainer *t);
int pb_ac_trainer_comm_init(pb_ac_trainer *t, const void *unique_id128, int rank, int world_size);
int pb_ac_trainer_broadcast_params(pb_ac_trainer *t);

/* torch::randperm(n) on a CPU mt19937 generator (src/ppo_continuous_action.cpp:490; src/ac_ppo_continuous_action.cpp:618,804):
 * forward Fisher-Yates with `random() % (n - i)`; bit-exact with torch for the same seed. */
typedef struct pb_rng pb_rng;
pb_rng *pb_rng_create(unsigned long long seed);
void pb_rng_destroy(pb_rng *r);
int pb_rng_randperm(pb_rng *r, int n, long long *out);

/* ---- DD-PPO pre-emption (src/ac_ppo_continuous_action.cpp:629-631,680-693,703-722,804-810; include/tcp_store.h) -------------
 * The reference lets every collector thread stop early once a fraction of all collectors has finished, trains on the first
 * min_collected_steps steps of every env and fills the batch by repeating the permutation.  Here collection is one batched
 * device step per rank, so the host loop decides per step: collect_begin, collect_step(s) for s = 0, 1, ... until
 * pb_dd_ppo_should_preempt(step = s) says stop (the counter it is fed replaces the ZeroMQ store; the CLI keeps it in a shared
 * file), collect_end(preempted); pb_ac_trainer_gae / _update then use the first num_train_steps_per_env steps only.  pb_ac_trainer_rollout = all num_steps. */
int pb_ac_trainer_collect_begin(pb_ac_trainer *t);
int pb_ac_trainer_collect_step(pb_ac_trainer *t, int s);
int pb_ac_trainer_collect_end(pb_ac_trainer *t, int preempted); /* preempted: the last collected step is kept as GAE's bootstrap only */
int pb_ac_trainer_barrier(pb_ac_trainer *t);                /* comm->allreduce(sychronize) of :632 + stream sync */
int pb_ac_trainer_train_steps(const pb_ac_trainer *t);      /* num_train_steps_per_env of the last collection (:722) */
/* (num_done / num_envs > threshold) && step > lround(min_perc * num_steps)   (:681-685); returns 0 / 1 */
int pb_dd_ppo_should_preempt(int num_done, int num_envs, float threshold, int step, float min_perc, int num_steps);
/* b_inds of one epoch (:804-810): randperm(num_collected) on `r`, repeated ceil(batch / num_collected) times and cut to
 * `batch` entries; bit-exact with torch for the same generator state.  out: HOST int64[batch]. */
int pb_dd_ppo_batch_indices(pb_rng *r, int num_collected, int batch, long long *out);

/* ---- data-parallel gradient exchange: replaces torchfort::Comm (include/distributed.h:41-60, src/distributed.cpp) */
/* 128-byte ncclUniqueId made on rank 0 and handed to the other ranks by the launcher (the reference broadcasts it
 * with MPI_Bcast, src/distributed.cpp:66-79). */
int pb_comm_unique_id(void *out128);
int pb_trainer_comm_init(pb_trainer *t, const void *unique_id128, int rank, int world_size);
int pb_trainer_broadcast_params(pb_trainer *t);  /* rank 0 -> all (src/ac_ppo_continuous_action.cpp:551-553) */
/* Which exchange step pb_trainer_update uses after pb_trainer_comm_init: 0 = single rank, 1 = ncclAllReduce(avg)
 * (src/distributed.cpp:117-133), 2 = the fused peer-memory all-reduce kernel over NVLink (csrc/pb_peer.cuh; chosen when
 * every rank could map every other rank's exchange buffer with CUDA IPC; PB_PEER_AR=0 forces 1). */
int pb_trainer_exchange_mode(const pb_trainer *t);
int pb_ac_trainer_exchange_mode(const pb_ac_trainer *t);   /* same codes; the ac_ppo trainer exchanges its gradient and the two advantage-moment scalars */

/* ---- allocation-free entry point.  NO COUNTERPART IN THE REFERENCE: added for callers
 * that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * Same result as pb_ac_trainer_obs_sums on one rank - sums_out[2*o] = sum x, sums_out[2*o+1] = sum x^2 over `rows` rows of the
 * row-major fp32 matrix obs_dev[rows][obs_dim], fp64 accumulation - but inputs and outputs are the caller's DEVICE pointers,
 * the work is enqueued on `stream` (a cudaStream_t, may be NULL) of CUDA device `device` and nothing is allocated or freed: the
 * block partials live in a static __device__ array of the library (512 KiB per GPU that loads it, held until exit).
 * Limit: obs_dim <= pb_obs_moments_noalloc_max_obs() = 512, larger sizes are rejected with PB_ERR_ARG.  Every call on a device
 * shares that array, which keeps the last call's contents (each call overwrites all it reads, nothing needs clearing); calls
 * that overlap on the GPU race on it - serialise them (one stream, or events between streams).  Asynchronous: synchronise
 * `stream` before reading sums_out_dev. */
int pb_obs_moments_noalloc(int device, void *stream, const float *obs_dev, long long rows, int obs_dim, double *sums_out_dev);
int pb_obs_moments_noalloc_max_obs(void);

/* ---- measurement helper: FP32 FMA throughput of the device (the roof SURVEY.md 8d names for the rigid-body kernel;
 * MEASURED_PEAKS.json only holds HBM and bf16 tensor peaks).  Runs a register-resident FFMA kernel on every SM for
 * ~`ms_target` ms and returns 2 * FMAs / time in TFLOP/s. */
int pb_fp32_fma_peak(int device, float ms_target, float *tflops_out);

#ifdef __cplusplus
}
#endif
#endif
