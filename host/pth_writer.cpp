// pth_writer — writes what the reference's save_state() writes (src/ppo_continuous_action.cpp:173-180,546-563,587;
// src/ac_ppo_continuous_action.cpp:268-275): torch::save(agent, model.pth) and torch::save(optimizer, optimizer.pth), from a raw dump
// of the engine's parameters and Adam state that the CLIs write next to it.  Linked against libtorch, kept OUT of the CLIs so the
// training binaries carry no libtorch (north_star); built only where a libtorch is found (host/Makefile).
//   pth_writer write  <dump.raw> <model.pth> <optimizer.pth>
//   pth_writer verify <dump.raw> <model.pth> <optimizer.pth>     loads both files back with torch::load and compares with the dump
// The two module skeletons below register the same names, in the same order, as the reference's AgentImpl classes
// (src/ppo_continuous_action.cpp:120-137, src/ac_ppo_continuous_action.cpp:150-186): that is what makes the archives interchangeable.
#include <torch/torch.h>

#include <cstdint>
#include <cstdio>
#include <fstream>
#include <iostream>
#include <string>
#include <vector>

using torch::Tensor;
namespace nn = torch::nn;

struct Dump {   // written by host/cli_common.h write_raw_dump()
  int32_t magic = 0, kind = 0, O = 0, A = 0, P = 0;
  int64_t step = 0;
  float lr = 0, eps = 0, act_high = 0, act_low = 0;
  std::vector<float> params, m, v, mean, std_;
};
static bool read_dump(const std::string &path, Dump &d) {
  std::ifstream f(path, std::ios::binary);
  if (!f) return false;
  f.read((char *)&d.magic, 4); f.read((char *)&d.kind, 4); f.read((char *)&d.O, 4); f.read((char *)&d.A, 4); f.read((char *)&d.P, 4);
  f.read((char *)&d.step, 8); f.read((char *)&d.lr, 4); f.read((char *)&d.eps, 4); f.read((char *)&d.act_high, 4); f.read((char *)&d.act_low, 4);
  if (d.magic != 0x50425257 || d.P <= 0) return false;
  d.params.resize(d.P); d.m.resize(d.P); d.v.resize(d.P);
  f.read((char *)d.params.data(), 4l * d.P); f.read((char *)d.m.data(), 4l * d.P); f.read((char *)d.v.data(), 4l * d.P);
  if (d.kind == 1) { d.mean.resize(d.O); d.std_.resize(d.O); f.read((char *)d.mean.data(), 4l * d.O); f.read((char *)d.std_.data(), 4l * d.O); }
  return (bool)f;
}

struct PpoAgentImpl : nn::Module {
  nn::Sequential critic{nullptr}, actor_mean{nullptr};
  Tensor actor_logstd;
  PpoAgentImpl(int O, int A) {
    critic = nn::Sequential(nn::Linear(O, 64), nn::Tanh(), nn::Linear(64, 64), nn::Tanh(), nn::Linear(64, 1));
    register_module("critic", critic);
    actor_mean = nn::Sequential(nn::Linear(O, 64), nn::Tanh(), nn::Linear(64, 64), nn::Tanh(), nn::Linear(64, A));
    register_module("actor_mean", actor_mean);
    actor_logstd = register_parameter("actor_logstd", torch::zeros({1, A}, torch::kFloat32));
  }
};
TORCH_MODULE(PpoAgent);

struct AcAgentImpl : nn::Module {
  Tensor action_space_high, action_space_low, mean_, std_;
  nn::Sequential critic{nullptr}, actor_encoder{nullptr}, dist_alpha{nullptr}, dist_beta{nullptr};
  AcAgentImpl(int O, int A, float hi, float lo, Tensor mean, Tensor sd) {
    action_space_high = register_parameter("action_space_high", torch::tensor(hi), false);
    action_space_low = register_parameter("action_space_low", torch::tensor(lo), false);
    mean_ = register_parameter("mean_", mean.unsqueeze(0), false);
    std_ = register_parameter("std_", sd.unsqueeze(0), false);
    critic = nn::Sequential(nn::Linear(O, 256), nn::LayerNorm(nn::LayerNormOptions({256})), nn::ReLU(), nn::Linear(256, 256),
                            nn::LayerNorm(nn::LayerNormOptions({256})), nn::ReLU(), nn::Linear(256, 1));
    register_module("critic", critic);
    actor_encoder = nn::Sequential(nn::Linear(O, 256), nn::LayerNorm(nn::LayerNormOptions({256})), nn::ReLU(), nn::Linear(256, 256),
                                   nn::LayerNorm(nn::LayerNormOptions({256})), nn::ReLU());
    register_module("actor_mean", actor_encoder);
    dist_alpha = nn::Sequential(nn::Linear(256, A));
    register_module("dist_alpha", dist_alpha);
    dist_beta = nn::Sequential(nn::Linear(256, A));
    register_module("dist_beta", dist_beta);
  }
};
TORCH_MODULE(AcAgent);

// the engine's flat order (include/ppo_b200.h: pb_trainer_get_params / pb_ac_trainer_get_params)
static std::vector<std::string> flat_names(int kind) {
  std::vector<std::string> n;
  auto wb = [&](const std::string &p) { n.push_back(p + ".weight"); n.push_back(p + ".bias"); };
  if (kind == 0) {
    for (const char *p : {"critic.0", "critic.2", "critic.4", "actor_mean.0", "actor_mean.2", "actor_mean.4"}) wb(p);
    n.push_back("actor_logstd");
  } else {
    for (const char *p : {"critic.0", "critic.1", "critic.3", "critic.4", "critic.6", "actor_mean.0", "actor_mean.1", "actor_mean.3", "actor_mean.4",
                          "dist_alpha.0", "dist_beta.0"}) wb(p);
  }
  return n;
}

template <class Agent>
static int run(const std::string &mode, const Dump &d, Agent agent, const std::string &model_path, const std::string &opt_path) {
  torch::NoGradGuard ng;
  auto named = agent->named_parameters(true);
  torch::optim::Adam optimizer(agent->parameters(), torch::optim::AdamOptions(d.lr).eps(d.eps));   // :341-342
  const auto names = flat_names(d.kind);
  if (mode == "write") {
    long off = 0;
    for (const auto &nm : names) {
      Tensor *p = named.find(nm);
      if (!p) { std::cerr << "pth_writer: no parameter " << nm << std::endl; return 3; }
      const long n = p->numel();
      p->copy_(torch::from_blob((void *)(d.params.data() + off), p->sizes(), torch::kFloat32));
      if (d.step > 0) {
        auto st = std::make_unique<torch::optim::AdamParamState>();
        st->step(d.step);
        st->exp_avg(torch::from_blob((void *)(d.m.data() + off), p->sizes(), torch::kFloat32).clone());
        st->exp_avg_sq(torch::from_blob((void *)(d.v.data() + off), p->sizes(), torch::kFloat32).clone());
        optimizer.state()[p->unsafeGetTensorImpl()] = std::move(st);
      }
      off += n;
    }
    if (off != d.P) { std::cerr << "pth_writer: parameter count mismatch " << off << " vs " << d.P << std::endl; return 3; }
    torch::save(agent, model_path);
    torch::save(optimizer, opt_path);
    return 0;
  }
  // verify: what a reference-side tool does - torch::load into a freshly constructed agent / optimizer
  torch::load(agent, model_path);
  torch::load(optimizer, opt_path);
  named = agent->named_parameters(true);
  long off = 0;
  double max_err = 0, max_m = 0;
  long steps_ok = 0, with_state = 0;
  for (const auto &nm : names) {
    Tensor *p = named.find(nm);
    const long n = p->numel();
    max_err = std::max(max_err, (p->flatten() - torch::from_blob((void *)(d.params.data() + off), {n}, torch::kFloat32)).abs().max().item<double>());
    auto it = optimizer.state().find(p->unsafeGetTensorImpl());
    if (it != optimizer.state().end()) {
      auto &st = static_cast<torch::optim::AdamParamState &>(*it->second);
      with_state++;
      steps_ok += st.step() == d.step;
      max_m = std::max(max_m, (st.exp_avg().flatten() - torch::from_blob((void *)(d.m.data() + off), {n}, torch::kFloat32)).abs().max().item<double>());
      max_m = std::max(max_m, (st.exp_avg_sq().flatten() - torch::from_blob((void *)(d.v.data() + off), {n}, torch::kFloat32)).abs().max().item<double>());
    }
    off += n;
  }
  const auto &opts = static_cast<torch::optim::AdamOptions &>(optimizer.param_groups()[0].options());
  printf("verify params_max_err %.3e adam_max_err %.3e tensors %zu with_state %ld step_ok %ld lr %.6g eps %.6g\n", max_err, max_m, names.size(), with_state,
         steps_ok, opts.lr(), opts.eps());
  return (max_err == 0 && max_m == 0 && (d.step == 0 || (with_state == (long)names.size() && steps_ok == with_state))) ? 0 : 4;
}

int main(int argc, char **argv) {
  if (argc != 5) { std::cerr << "usage: pth_writer write|verify <dump.raw> <model.pth> <optimizer.pth>" << std::endl; return 1; }
  Dump d;
  if (!read_dump(argv[2], d)) { std::cerr << "pth_writer: cannot read " << argv[2] << std::endl; return 2; }
  try {
    if (d.kind == 0) return run(argv[1], d, PpoAgent(d.O, d.A), argv[3], argv[4]);
    Tensor mean = torch::from_blob(d.mean.data(), {d.O}, torch::kFloat32).clone(), sd = torch::from_blob(d.std_.data(), {d.O}, torch::kFloat32).clone();
    return run(argv[1], d, AcAgent(d.O, d.A, d.act_high, d.act_low, mean, sd), argv[3], argv[4]);
  } catch (const std::exception &e) {
    std::cerr << "pth_writer: " << e.what() << std::endl;
    return 5;
  }
}
