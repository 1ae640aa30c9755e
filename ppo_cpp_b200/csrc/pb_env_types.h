// pb_env_types.h — argument blocks shared by the env kernels and the host code (vector env + trainer).
#pragma once
#include <stdint.h>

// Per-env persistent state in HBM (env-major rows so a warp's lanes touch consecutive addresses).
struct EnvBuffers {
  int num_envs;
  float *qpos, *qvel, *warm;  // [N][NQ], [N][NV], [N][NV]; root x (and y for free-joint models) kept relative
  double *xy_base;            // [N][2] absolute root translation removed from qpos
  float *xref;                // [N] stale x reference of the last forward (Ant torso x / Humanoid com x), relative
  int *elapsed, *autoreset, *ep_length;
  float *ep_return;
  float *ob_mean, *ob_var, *ob_count;  // NormalizeObservation, per env
  float *rw_mean, *rw_var, *rw_acc, *rw_count;  // NormalizeReward, per env
  uint32_t *mt;               // [N][624] std::mt19937 state
  int *mt_idx, *nrm_ok;       // [N]
  double *nrm_saved;          // [N] cached second value of std::normal_distribution
  int *work;                  // [2] env work counters of the persistent step kernel (launch parity selects one)
  int *cost, *order;          // [N] 64-cycle units the env's warp spent in the last step's forward evaluations; next step's queue order
  unsigned long long *counters;  // [0] forward evaluations, [1] envs whose contact/row caps overflowed, [2] PGS sweeps, [3] PGS row steps
};

struct StepArgs {
  const float *actions;  // [N][NU]
  float *obs, *rew, *term, *trunc;  // [N][OBS], [N]
  int *has_info, *info_l;
  float *info_r;
  int chain;          // 0: RecordEpisodeStatistics only (ac_ppo), 1: ppo chain with normalisers and clamps
  int clip_actions;
  float act_min, act_max, gamma;
  int mode;           // 0: step (with autoreset), 1: reset(seed), 2: mj_forward at the stored state (raw obs, no wrappers)
  int seed;
  int parity;         // which of the two work counters this launch consumes (the other is zeroed for the next launch)
  int ordered;        // 1: the work queue is eb.order (set by the launcher after env_order_kernel), 0: env index order
  float *ep_stat;     // optional [3]: sum of finished-episode returns, lengths, count (atomicAdd)
};

