// b200_gym_torch.h — the gymcpp vector-env interface with its literal libtorch signatures (libs/gymcpp/gym.h:276-366) over
// libppo_b200.so, for a main() that is written against gymcpp and links libtorch.  Compiled only where a
// libtorch is present (host/Makefile probes the Python torch package's headers; the reference pins libtorch 2.4.1).
//
//   Tensor reset(const int seed);                                                                   gym.h:310-318
//   tuple<Tensor, Tensor, Tensor, Tensor, vector<optional<env_info>>> step(const Tensor& actions);  gym.h:324-365
//   get_num_envs() plus SeqVectorEnv's get_observation_space / get_action_space / _min / _max       gym.h:104-129,320-322
//
// Like the reference's class the env owns its output tensors (kFloat32 CPU tensors: obs [N, O], rewards / terminations /
// truncations [N], flags as 0.0 / 1.0) and returns them by value sharing storage - the caller copies what it keeps before the
// next step (the mains do, src/ppo_continuous_action.cpp:391,406-408).  Construction differs by necessity: gymcpp takes a
// vector of per-env wrapper objects, here the task id and the wrapper chain select the batched device env
// (b200::make_ppo_envs / make_ac_ppo_envs mirror the two make_env functions).  Errors are std::runtime_error.
#pragma once
#include <torch/torch.h>

#include <memory>
#include <optional>
#include <string>
#include <tuple>
#include <vector>

#include "b200_vector_env.h"

namespace b200 {

class ParVectorEnv {
 public:
  const unsigned int num_envs_;

  ParVectorEnv(const std::string &env_id, int num_envs, const bool clip_actions, int wrapper_chain = PB_CHAIN_PPO, float gamma = 0.99f,
               int device = 0)
      : num_envs_((unsigned)num_envs), env_(env_id, num_envs, clip_actions, wrapper_chain, gamma, device) {
    obs_ = torch::zeros({(long)num_envs_, (long)env_.get_observation_space()});
    rewards_ = torch::zeros({(long)num_envs_});
    terminations_ = torch::zeros({(long)num_envs_});
    truncations_ = torch::zeros({(long)num_envs_});
    infos_.assign(num_envs_, std::nullopt);
    has_.assign(num_envs_, 0);
    info_l_.assign(num_envs_, 0);
    info_r_.assign(num_envs_, 0.f);
  }

  torch::Tensor reset(const int seed) {
    check(pb_env_reset_host(env_.handle(), seed, obs_.data_ptr<float>()));
    return obs_;
  }

  std::tuple<torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor, std::vector<std::optional<env_info>>> step(const torch::Tensor &actions) {
    const torch::Tensor a = actions.to(torch::kCPU, torch::kFloat32).contiguous();
    if (a.numel() != (long)num_envs_ * env_.get_action_space()) throw std::runtime_error("b200::ParVectorEnv::step: actions must be [num_envs, act_dim]");
    check(pb_env_step_host(env_.handle(), a.data_ptr<float>(), obs_.data_ptr<float>(), rewards_.data_ptr<float>(), terminations_.data_ptr<float>(),
                           truncations_.data_ptr<float>(), has_.data(), info_r_.data(), info_l_.data()));
    const float t = std::chrono::duration<float>(std::chrono::steady_clock::now() - start_).count();
    for (unsigned i = 0; i < num_envs_; i++) {
      if (has_[i]) infos_[i] = env_info{info_r_[i], info_l_[i], t}; else infos_[i] = std::nullopt;
    }
    return {obs_, rewards_, terminations_, truncations_, infos_};
  }

  [[nodiscard]] unsigned int get_num_envs() const { return num_envs_; }
  [[nodiscard]] int get_observation_space() { return env_.get_observation_space(); }
  [[nodiscard]] int get_action_space() { return env_.get_action_space(); }
  [[nodiscard]] float get_action_space_min() { return env_.get_action_space_min(); }
  [[nodiscard]] float get_action_space_max() { return env_.get_action_space_max(); }

 private:
  static void check(int rc) {
    if (rc != PB_OK) throw std::runtime_error(std::string("libppo_b200: ") + pb_last_error());
  }
  B200VectorEnv env_;
  torch::Tensor obs_, rewards_, terminations_, truncations_;
  std::vector<std::optional<env_info>> infos_;
  std::vector<int> has_, info_l_;
  std::vector<float> info_r_;
  std::chrono::steady_clock::time_point start_ = std::chrono::steady_clock::now();
};
using SeqVectorEnv = ParVectorEnv;   // same interface, gym.h:75-164

// make_env of src/ppo_continuous_action.cpp:41-49 / src/ac_ppo_continuous_action.cpp:50-53 for a whole vector env
inline std::shared_ptr<ParVectorEnv> make_ppo_envs(const std::string &env_id, int num_envs, float gamma, bool clip_actions, int device = 0) {
  return std::make_shared<ParVectorEnv>(env_id, num_envs, clip_actions, PB_CHAIN_PPO, gamma, device);
}
inline std::shared_ptr<ParVectorEnv> make_ac_ppo_envs(const std::string &env_id, int num_envs, bool clip_actions, int device = 0) {
  return std::make_shared<ParVectorEnv>(env_id, num_envs, clip_actions, PB_CHAIN_STATS_ONLY, 0.99f, device);
}

}  // namespace b200
