// vector_env_selftest — exercises include/b200_vector_env.h (and, when built with -DWITH_TORCH, include/b200_gym_torch.h)
// the way a gymcpp caller would: reset(seed), then steps with a deterministic action pattern; prints one line per check that
// tests/test_cli.py compares with the Python mirror driven by the same actions.  Usage: vector_env_selftest <env_id> <num_envs> <steps>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <string>
#include <vector>

#ifdef WITH_TORCH
#include "b200_gym_torch.h"
#else
#include "b200_vector_env.h"
#endif

static double checksum(const float *p, size_t n) {
  double s = 0;
  for (size_t i = 0; i < n; i++) s += (double)p[i] * (double)((i % 7) + 1);
  return s;
}

int main(int argc, char **argv) {
  const std::string env_id = argc > 1 ? argv[1] : "Hopper-v5";
  const int n = argc > 2 ? atoi(argv[2]) : 64, steps = argc > 3 ? atoi(argv[3]) : 40;
  try {
#ifdef WITH_TORCH
    auto envs = b200::make_ppo_envs(env_id, n, 0.99f, true);
    const int O = envs->get_observation_space(), A = envs->get_action_space();
    torch::Tensor obs = envs->reset(1);
    printf("reset %.9e\n", checksum(obs.data_ptr<float>(), (size_t)n * O));
    long episodes = 0; double ret_sum = 0;
    for (int t = 0; t < steps; t++) {
      torch::Tensor act = torch::zeros({n, A});
      float *a = act.data_ptr<float>();
      for (int i = 0; i < n * A; i++) a[i] = 2.0f * std::sin(0.37f * (float)(t + 1) + 0.11f * (float)i);
      auto [o, r, te, tr, infos] = envs->step(act);
      for (const auto &inf : infos) if (inf) { episodes++; ret_sum += inf->r; }
      printf("step %d %.9e %.9e %.0f %.0f\n", t, checksum(o.data_ptr<float>(), (size_t)n * O), checksum(r.data_ptr<float>(), n),
             te.sum().item<float>(), tr.sum().item<float>());
    }
#else
    b200::B200VectorEnv envs(env_id, n, true, PB_CHAIN_PPO, 0.99f);
    const int O = envs.get_observation_space(), A = envs.get_action_space();
    const std::vector<float> &obs = envs.reset(1);
    printf("reset %.9e\n", checksum(obs.data(), (size_t)n * O));
    long episodes = 0; double ret_sum = 0;
    std::vector<float> act((size_t)n * A);
    for (int t = 0; t < steps; t++) {
      for (int i = 0; i < n * A; i++) act[i] = 2.0f * std::sin(0.37f * (float)(t + 1) + 0.11f * (float)i);
      auto [o, r, te, tr, infos] = envs.step(act);
      for (const auto &inf : infos) if (inf) { episodes++; ret_sum += inf->r; }
      double ts = 0, us = 0;
      for (int i = 0; i < n; i++) { ts += te[i]; us += tr[i]; }
      printf("step %d %.9e %.9e %.0f %.0f\n", t, checksum(o.data(), (size_t)n * O), checksum(r.data(), n), ts, us);
    }
#endif
    printf("episodes %ld %.6e\n", episodes, ret_sum);
    printf("spaces %d %d %.3f %.3f\n", O, A, (double)(
#ifdef WITH_TORCH
        envs->get_action_space_min()), (double)envs->get_action_space_max());
#else
        envs.get_action_space_min()), (double)envs.get_action_space_max());
#endif
  } catch (const std::exception &e) {
    fprintf(stderr, "vector_env_selftest: %s\n", e.what());
    return 2;
  }
  return 0;
}
