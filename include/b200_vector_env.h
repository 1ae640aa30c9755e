// b200_vector_env.h — header-only C++ mirror of gymcpp's vector env (libs/gymcpp/gym.h) over the C ABI of libppo_b200.so.
//
// Same method names and argument meaning as ParVectorEnv / SeqVectorEnv (gym.h:75-164, 276-366):
//   reset(seed)     -> observations  [num_envs, obs_dim]                               (gym.h:310-318)
//   step(actions)   -> {obs, rewards, terminations, truncations, infos}                (gym.h:324-365)
//   get_num_envs(), get_observation_space(), get_action_space(), get_action_space_min/max()   (gym.h:104-129,320-322)
// Tensors become plain float32 host vectors owned by the env (the reference's tensors are CPU tensors owned by the env
// too, gym.h:301-304); flags are 0.0/1.0 floats; infos is vector<optional<env_info>> with the same env_info{r,l,t}
// (gym.h:19-24; t is filled with wall-clock seconds since construction like wrappers/common.h:57-60).
// Errors surface as std::runtime_error, the reference's convention (include/defines.h:37-72).
// No torch dependency; a libtorch adapter is a from_blob() around these vectors.
#pragma once
#include <chrono>
#include <optional>
#include <stdexcept>
#include <string>
#include <tuple>
#include <vector>

#include "ppo_b200.h"

namespace b200 {

struct env_info {  // gym.h:19-24
  float r;  // cumulative reward
  int l;    // episode length
  float t;  // elapsed time since instantiation
};

class B200VectorEnv {
 public:
  using StepResult = std::tuple<const std::vector<float> &, const std::vector<float> &, const std::vector<float> &,
                                const std::vector<float> &, const std::vector<std::optional<env_info>> &>;

  // wrapper_chain: PB_CHAIN_PPO reproduces make_env of src/ppo_continuous_action.cpp:41-49, PB_CHAIN_STATS_ONLY that of
  // src/ac_ppo_continuous_action.cpp:50-53.  clip_actions as ParVectorEnv's constructor argument (gym.h:293).
  B200VectorEnv(const std::string &env_id, int num_envs, bool clip_actions = true, int wrapper_chain = PB_CHAIN_PPO,
                float gamma = 0.99f, int device = 0)
      : num_envs_(num_envs), start_(std::chrono::steady_clock::now()) {
    check(pb_env_create(env_id.c_str(), num_envs, device, wrapper_chain, gamma, clip_actions ? 1 : 0, &env_));
    obs_dim_ = pb_env_obs_dim(env_);
    act_dim_ = pb_env_act_dim(env_);
    obs_.assign((size_t)num_envs * obs_dim_, 0.f);
    rewards_.assign(num_envs, 0.f);
    terminations_.assign(num_envs, 0.f);
    truncations_.assign(num_envs, 0.f);
    has_.assign(num_envs, 0);
    info_l_.assign(num_envs, 0);
    info_r_.assign(num_envs, 0.f);
    infos_.assign(num_envs, std::nullopt);
  }
  ~B200VectorEnv() { pb_env_destroy(env_); }
  B200VectorEnv(const B200VectorEnv &) = delete;
  B200VectorEnv &operator=(const B200VectorEnv &) = delete;

  const std::vector<float> &reset(const int seed) {
    check(pb_env_reset_host(env_, seed, obs_.data()));
    return obs_;
  }

  // actions: [num_envs, act_dim] float32, row-major
  StepResult step(const std::vector<float> &actions) {
    if (actions.size() != (size_t)num_envs_ * act_dim_) throw std::runtime_error("B200VectorEnv::step: actions must be [num_envs, act_dim]");
    return step(actions.data());
  }
  StepResult step(const float *actions) {
    check(pb_env_step_host(env_, actions, obs_.data(), rewards_.data(), terminations_.data(), truncations_.data(), has_.data(),
                           info_r_.data(), info_l_.data()));
    const float t = std::chrono::duration<float>(std::chrono::steady_clock::now() - start_).count();
    for (int i = 0; i < num_envs_; i++) {
      if (has_[i]) infos_[i] = env_info{info_r_[i], info_l_[i], t}; else infos_[i] = std::nullopt;
    }
    return StepResult(obs_, rewards_, terminations_, truncations_, infos_);
  }

  [[nodiscard]] int get_num_envs() const { return num_envs_; }
  [[nodiscard]] int get_observation_space() const { return obs_dim_; }
  [[nodiscard]] int get_action_space() const { return act_dim_; }
  [[nodiscard]] float get_action_space_min() const { return pb_env_act_min(env_); }
  [[nodiscard]] float get_action_space_max() const { return pb_env_act_max(env_); }
  pb_env *handle() { return env_; }

 private:
  static void check(int rc) {
    if (rc != PB_OK) throw std::runtime_error(std::string("libppo_b200: ") + pb_last_error());
  }
  pb_env *env_ = nullptr;
  int num_envs_, obs_dim_ = 0, act_dim_ = 0;
  std::chrono::steady_clock::time_point start_;
  std::vector<float> obs_, rewards_, terminations_, truncations_, info_r_;
  std::vector<int> has_, info_l_;
  std::vector<std::optional<env_info>> infos_;
};

}  // namespace b200
