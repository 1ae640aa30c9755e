// pb_internal.h — host-side internals shared by the translation units of libppo_b200.so (not part of the ABI).
#pragma once
#include <cuda_runtime.h>

#include <string>
#include <vector>

#include "../../include/ppo_b200.h"
#include "pb_devmodel.h"
#include "pb_env_types.h"

int pb_fail(int code, const std::string &msg);
#define fail pb_fail
#define CK(call)                                                                                              \
  do {                                                                                                        \
    cudaError_t e_ = (call);                                                                                  \
    if (e_ != cudaSuccess) return pb_fail(PB_ERR_CUDA, std::string(#call) + ": " + cudaGetErrorString(e_));   \
  } while (0)

struct pb_env {
  int task, num_envs, device, chain, clip_actions, nq, nv, nu, obs_dim;
  float gamma, act_min, act_max;
  cudaStream_t stream;
  mutable int launch_parity = 0;  // which work counter the next env-kernel launch consumes
  // the next step's queue order is computed on a side stream right after an env step, beside whatever the caller launches next
  cudaStream_t order_stream = nullptr;
  cudaEvent_t step_done = nullptr, order_done = nullptr;
  mutable bool order_pending = false;
  DevModel *d_model;
  EnvBuffers eb;
  std::vector<void *> allocs;
  // staging for the *_host calls (pinned)
  float *h_pin = nullptr, *d_act = nullptr, *d_obs = nullptr, *d_rew = nullptr, *d_term = nullptr, *d_trunc = nullptr,
        *d_info_r = nullptr;
  int *d_has = nullptr, *d_info_l = nullptr;
};

// launch the fused env kernel with a caller-filled argument block (used by the trainer's rollout loop)
StepArgs pb_env_make_args(const pb_env *e, int mode, int seed);
cudaError_t pb_env_dispatch(const pb_env *e, const StepArgs &a);

#include <random>
// torch::randperm on a CPU mt19937 and the DD-PPO repeat / truncate rule on top of it (pb_trainer.cu)
void pb_randperm_i64(std::mt19937 &gen, int n, long long *out);
void pb_dd_ppo_indices(std::mt19937 &gen, int num_collected, int batch, long long *out);
