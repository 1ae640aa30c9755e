// ac_ppo_continuous_action — command-line driver with the flags, defaults, stdout lines and scalar tags of the reference's
// src/ac_ppo_continuous_action.cpp, on top of libppo_b200.so (pb_ac_trainer_* C ABI).
//
// What changed shape (the reference lines are :line of src/ac_ppo_continuous_action.cpp):
//  * the reference runs one host thread per env on the collect device and a second model copy on the train device
//    (:581-603, 641-698); here every rank steps all of its envs in one batched device kernel per step, so
//    --collect_device / --train_device are accepted and ignored (both are the rank's GPU) and the DD-PPO pre-emption
//    flags (:86-88, 680-693) act between ranks (a counter file replaces the TCPStore); within a rank the envs step together;
//  * the MPI + TCPStore rendezvous (:360-372) is replaced by PB_RANK / PB_WORLD / PB_LOCAL_RANK (or torchrun's RANK /
//    WORLD_SIZE / LOCAL_RANK) plus PB_NCCL_ID_FILE; --rdzv_addr / --tcp_store_port are accepted for compatibility;
//  * the observation mean / std tables (:496-534) are the default (host/ac_obs_stats.h, data extracted by tools/extract_ac_obs_stats.py);
//    zero-std entries standardise with std 1 (the reference produces inf / NaN there); --obs_stats_file <file> replaces them
//    (two lines: mean, std); --obs_stats_file estimate takes them from one warm-up rollout of the initial policy instead
//    (all ranks agree through an allreduce).
#include <cmath>
#include <filesystem>
#include <fstream>
#include <iomanip>
#include <numeric>
#include <sstream>

#include "ac_obs_stats.h"
#include "cli_common.h"

namespace fs = std::filesystem;

struct GlobalConfig {  // :55-100, same names and defaults
  int seed = 1;
  int eval_seed = 2;
  long total_timesteps = 10000000;
  float learning_rate = 2.5e-4f;
  int num_envs = 8;
  int num_steps = 128;
  float gamma = 0.99f;
  float gae_lambda = 0.95f;
  int num_minibatches = 4;
  int update_epochs = 4;
  bool norm_adv = true;
  float clip_coef = 0.1f;
  bool clip_vloss = true;
  float ent_coef = 0.01f;
  float vf_coef = 0.5f;
  float max_grad_norm = 0.5f;
  float adam_eps = 1e-5f;
  bool anneal_lr = true;
  int num_eval_runs = 128;
  bool clip_actions = true;
  bool torch_deterministic = true;
  std::string exp_name_stem = "Ant-v5_AC_PPO_Atari";
  std::string env_id = "Ant-v5";
  std::string render = "rgb_array";
  std::string gpu_ids = "0";
  std::string collect_device = "gpu";
  std::string train_device = "gpu";
  std::string rdzv_addr = "localhost";
  int tcp_store_port = 29500;
  int use_dd_ppo_preempt = 0;
  float dd_ppo_min_perc = 0.25f;
  float dd_ppo_preempt_threshold = 0.6f;
  bool estimate_mean_std = false;
  std::string obs_stats_file;  // new: "mean..." line, "std..." line
  std::string exp_name;
  long batch_size = 0, minibatch_size = 0, num_iterations = 0;
  int num_devices = 1, num_envs_per_device = 0;
  long batch_size_per_device = 0, minibatch_per_device = 0;

  std::string to_string() const {
    std::ostringstream o;
    o << "|param|value|\n|-|-|\n"
      << "|seed|" << seed << "|\n|eval_seed|" << eval_seed << "|\n|total_timesteps|" << total_timesteps << "|\n|learning_rate|"
      << learning_rate << "|\n|num_envs|" << num_envs << "|\n|num_steps|" << num_steps << "|\n|gamma|" << gamma << "|\n|gae_lambda|"
      << gae_lambda << "|\n|num_minibatches|" << num_minibatches << "|\n|update_epochs|" << update_epochs << "|\n|norm_adv|" << norm_adv
      << "|\n|clip_coef|" << clip_coef << "|\n|clip_vloss|" << clip_vloss << "|\n|ent_coef|" << ent_coef << "|\n|vf_coef|" << vf_coef
      << "|\n|max_grad_norm|" << max_grad_norm << "|\n|adam_eps|" << adam_eps << "|\n|anneal_lr|" << anneal_lr << "|\n|num_eval_runs|"
      << num_eval_runs << "|\n|clip_actions|" << clip_actions << "|\n|exp_name|" << exp_name << "|\n|batch_size|" << batch_size
      << "|\n|minibatch_size|" << minibatch_size << "|\n|num_iterations|" << num_iterations << "|\n|torch_deterministic|"
      << torch_deterministic << "|\n|exp_name_stem|" << exp_name_stem << "|\n|env_id|" << env_id << "|\n|collect_device|" << collect_device
      << "|\n|train_device|" << train_device << "|\n|num_envs_per_device|" << num_envs_per_device << "|\n|batch_size_per_device|"
      << batch_size_per_device << "|\n|use_dd_ppo_preempt|" << use_dd_ppo_preempt << "|\n|dd_ppo_min_perc|" << dd_ppo_min_perc
      << "|\n|dd_ppo_preempt_threshold|" << dd_ppo_preempt_threshold << "|\n|minibatch_per_device|" << minibatch_per_device << "|\n";
    return o.str();
  }
};

static int die(int code, const std::string &what) {
  std::cerr << what << ": " << pb_last_error() << std::endl;
  return code;
}

static std::vector<int> parse_ids(const std::string &s) {  // "0,1,2" or "0 1 2"
  std::vector<int> ids;
  std::string tok;
  for (char c : s + ",") {
    if (c == ',' || c == ' ') {
      if (!tok.empty()) ids.push_back(std::atoi(tok.c_str()));
      tok.clear();
    } else tok.push_back(c);
  }
  return ids;
}

static bool read_stats_file(const std::string &path, int O, std::vector<float> &mean, std::vector<float> &sd) {
  std::ifstream f(path);
  if (!f) return false;
  auto row = [&](std::vector<float> &v) {
    std::string line;
    while (std::getline(f, line)) {
      for (char &c : line) if (c == ',' || c == '[' || c == ']' || c == '{' || c == '}') c = ' ';
      std::istringstream is(line);
      v.clear();
      float x;
      while (is >> x) v.push_back(x);
      if (!v.empty()) return (int)v.size() == O;
    }
    return false;
  };
  return row(mean) && row(sd);
}

// running sum x / sum x^2 -> population mean / std (:957-959); constant columns get std 1 so that (x - mean) / std stays finite
struct ObsMoments {
  std::vector<double> sums;
  long long rows = 0;
  void add(const std::vector<double> &s, long long r) {
    if (sums.empty()) sums.assign(s.size(), 0.0);
    for (size_t i = 0; i < s.size(); i++) sums[i] += s[i];
    rows += r;
  }
  void finish(std::vector<float> &mean, std::vector<float> &sd, bool fix_zero) const {
    const size_t O = sums.size() / 2;
    mean.resize(O); sd.resize(O);
    for (size_t o = 0; o < O; o++) {
      const double m = sums[2 * o] / (double)rows, var = std::max(0.0, sums[2 * o + 1] / (double)rows - m * m);
      mean[o] = (float)m;
      sd[o] = (float)std::sqrt(var);
      if (fix_zero && sd[o] < 1e-6f) sd[o] = 1.f;
    }
  }
};

static void print_vec(const char *title, const std::vector<float> &v) {
  std::cout << title << "\n";
  for (size_t i = 0; i < v.size(); i++) std::cout << std::fixed << std::setprecision(4) << v[i] << (i + 1 < v.size() ? ", " : "\n");
  std::cout << std::defaultfloat << std::setprecision(6) << std::endl;
}

int main(int argc, char **argv) {
  GlobalConfig config;
  cli::Flags f;
  f.add("seed", &config.seed, "Training seed");
  f.add("eval_seed", &config.eval_seed, "Seed of final evaluation run");
  f.add("total_timesteps", &config.total_timesteps, "Number of environment steps");
  f.add("learning_rate", &config.learning_rate, "Adam learning rate");
  f.add("num_envs", &config.num_envs, "Number of environments to be used (all devices together).");
  f.add("num_steps", &config.num_steps, "Num environment steps per iteration");
  f.add("gamma", &config.gamma, "Discount factor");
  f.add("gae_lambda", &config.gae_lambda, "Lambda of generalized advantage estimation");
  f.add("num_minibatches", &config.num_minibatches, "Number of training iterations per epoch");
  f.add("update_epochs", &config.update_epochs, "Number of training epochs per iteration");
  f.add("norm_adv", &config.norm_adv, "Whether to normalize the advantage");
  f.add("clip_coef", &config.clip_coef, "PPO clip coefficient");
  f.add("clip_vloss", &config.clip_vloss, "Whether to apply clipping to the value loss");
  f.add("ent_coef", &config.ent_coef, "Weigth of entropy loss.");
  f.add("vf_coef", &config.vf_coef, "Weigth of value loss.");
  f.add("max_grad_norm", &config.max_grad_norm, "Factor for gradient clipping.");
  f.add("adam_eps", &config.adam_eps, "Epsilon of adam.");
  f.add("anneal_lr", &config.anneal_lr, "Whether to anneal the learning rate linearly.");
  f.add("num_eval_runs", &config.num_eval_runs, "How many environments to evaluate");
  f.add("clip_actions", &config.clip_actions, "Whether to clip action into the valid range.");
  f.add("torch_deterministic", &config.torch_deterministic, "Accepted for compatibility");
  f.add("exp_name_stem", &config.exp_name_stem, "Name of the experiment.");
  f.add("env_id", &config.env_id, "Name of the mujoco env to be executed.");
  f.add("render", &config.render, "rgb_array only: the GLFW/OpenGL 'human' mode is unavailable on this box");
  f.add("gpu_ids", &config.gpu_ids, "CUDA device of each local rank, e.g. 0,1,2,3");
  f.add("collect_device", &config.collect_device, "Accepted for compatibility: collection runs on the rank's GPU");
  f.add("train_device", &config.train_device, "Accepted for compatibility: training runs on the rank's GPU");
  f.add("rdzv_addr", &config.rdzv_addr, "Accepted for compatibility (rendezvous is PB_NCCL_ID_FILE)");
  f.add("tcp_store_port", &config.tcp_store_port, "Accepted for compatibility (rendezvous is PB_NCCL_ID_FILE)");
  f.add("use_dd_ppo_preempt", &config.use_dd_ppo_preempt, "Flag to toggle the dd_ppo pre-emption trick (between ranks: the envs of one rank step together)");
  f.add("dd_ppo_min_perc", &config.dd_ppo_min_perc, "Minimum percentage of env steps that need to finish before preemtion.");
  f.add("dd_ppo_preempt_threshold", &config.dd_ppo_preempt_threshold, "Minimum number of envs that need to finish before preemtion.");
  f.add("estimate_mean_std", &config.estimate_mean_std, "Estimates mean and std of the observations of the run and prints them at the end.");
  f.add("obs_stats_file", &config.obs_stats_file, "Text file with the observation mean (line 1) and std (line 2) the agent standardises with (default: the built-in tables; \"estimate\": from a warm-up rollout)");
  int pr = f.parse(argc, argv);
  if (pr == 1) return 0;
  if (pr == 2) return 1;

  const cli::Dist dist = cli::Dist::from_env();
  config.num_devices = dist.world;
  if (config.num_envs % config.num_devices != 0) { std::cerr << "num_envs must be a multiple of num_devices." << std::endl; return 1; }  // :400
  config.num_envs_per_device = config.num_envs / config.num_devices;
  config.exp_name = config.exp_name_stem + "_" + std::to_string(config.seed);
  config.batch_size = (long)config.num_steps * config.num_envs;
  config.minibatch_size = config.batch_size / config.num_minibatches;
  config.num_iterations = config.total_timesteps / config.batch_size;
  config.batch_size_per_device = config.batch_size / config.num_devices;
  config.minibatch_per_device = config.minibatch_size / config.num_devices;
  if (config.render != "rgb_array") std::cerr << "--render " << config.render << ": unavailable on this box, continuing headless" << std::endl;
  const std::vector<int> ids = parse_ids(config.gpu_ids);
  const int device = dist.local_rank < (int)ids.size() ? ids[dist.local_rank] : dist.local_rank;

  fs::path exe = fs::canonical(argv[0]);
  fs::path exp_folder = exe.parent_path() / ".." / "models" / config.exp_name;
  if (dist.rank == 0) fs::create_directories(exp_folder);
  cli::ScalarLogger logger(dist.rank == 0 ? (exp_folder / "scalars.jsonl").string() : std::string("/dev/null"),
                           dist.rank == 0 ? (exp_folder / "tfevents_logs.pb").string() : std::string());
  logger.add_text("hyperparameters", 0, config.to_string());

  pb_ac_config pc;
  pb_ac_config_default(&pc);
  pc.env_id = config.env_id.c_str(); pc.num_envs = config.num_envs_per_device; pc.num_steps = config.num_steps; pc.device = device;
  pc.seed = config.seed; pc.gamma = config.gamma; pc.gae_lambda = config.gae_lambda; pc.num_minibatches = config.num_minibatches;
  pc.update_epochs = config.update_epochs; pc.norm_adv = config.norm_adv; pc.clip_coef = config.clip_coef; pc.clip_vloss = config.clip_vloss;
  pc.ent_coef = config.ent_coef; pc.vf_coef = config.vf_coef; pc.max_grad_norm = config.max_grad_norm; pc.adam_eps = config.adam_eps;
  pc.clip_actions = config.clip_actions; pc.rank = dist.rank; pc.world_size = dist.world;
  pb_ac_trainer *t = nullptr;
  int rc = pb_ac_trainer_create(&pc, &t);
  if (rc == PB_ERR_BAD_ENV) { std::cerr << "env_id: " << config.env_id << " is not implemented." << std::endl; return 1; }  // :538-541
  if (rc == PB_ERR_BAD_DEVICE) return die(2, "device");
  if (rc) return die(1, "pb_ac_trainer_create");
  const int O = pb_ac_trainer_obs_dim(t);
  if (dist.world > 1) {
    unsigned char id[128];
    if (dist.exchange_unique_id(id) != 0 || pb_ac_trainer_comm_init(t, id, dist.rank, dist.world) != PB_OK) return die(6, "comm init");
    dist.remove_unique_id();   // every rank has read it once the collective init returned
    if (pb_ac_trainer_broadcast_params(t) != PB_OK) return die(6, "broadcast");  // :551-553
  }

  // env i of rank r is seeded seed + r*envs_per_device + i (:605-606 seeds env i of every rank with seed + i)
  const int reset_seed = config.seed + dist.rank * config.num_envs_per_device;
  std::vector<double> sums(2 * (size_t)O);
  std::vector<float> mean, sd;   // the agent's mean_ / std_ (also part of its .pth, :153-156)
  {  // observation statistics of the agent (:496-534)
    if (!config.obs_stats_file.empty() && config.obs_stats_file != "estimate") {
      if (!read_stats_file(config.obs_stats_file, O, mean, sd)) { std::cerr << "cannot read " << O << " means and stds from " << config.obs_stats_file << std::endl; return 1; }
      if (pb_ac_trainer_set_obs_stats(t, mean.data(), sd.data()) != PB_OK) return die(1, "set_obs_stats");
    } else if (config.obs_stats_file == "estimate") {   // not in the reference: statistics of one warm-up rollout
      long long rows = 0;
      ObsMoments m;
      if (pb_ac_trainer_reset(t, reset_seed) != PB_OK || pb_ac_trainer_rollout(t) != PB_OK) return die(1, "warm-up rollout");
      if (pb_ac_trainer_obs_sums(t, sums.data(), &rows) != PB_OK) return die(1, "obs_sums");
      m.add(sums, rows);
      m.finish(mean, sd, true);
      if (pb_ac_trainer_set_obs_stats(t, mean.data(), sd.data()) != PB_OK) return die(1, "set_obs_stats");
      if (dist.rank == 0) std::cout << "Observation statistics estimated from " << rows << " warm-up observations" << std::endl;
    } else {
      // default: the reference's fixed tables (:496-497 Humanoid-v4, :506-508 HalfCheetah-v5 = 0 / 1, :521-522 Ant-v5, :533-534
      // Hopper-v5; host/ac_obs_stats.h).  The Humanoid table holds std = 0 for the 19 constant features (body masses in cinert,
      // the world body's cvel): the reference divides by them, (x - mean) / 0 = +-inf for a mass that differs from its 4-digit
      // table value, and the network output is NaN.  Here those entries standardise with std 1 (the constant feature becomes
      // a ~1e-5 offset) - the one deviation from the table, printed below.
      const AcObsStats *tab = ac_obs_stats_for(config.env_id.c_str());
      if (!tab || tab->n != O) return die(1, "no observation statistics table for this env");
      mean.assign(tab->mean, tab->mean + O);
      sd.assign(tab->std, tab->std + O);
      int zeros = 0;
      for (auto &v : sd) if (v == 0.f) { v = 1.f; zeros++; }
      if (pb_ac_trainer_set_obs_stats(t, mean.data(), sd.data()) != PB_OK) return die(1, "set_obs_stats");
      if (dist.rank == 0 && zeros) std::cout << "Observation statistics: reference table, " << zeros << " zero-std entries standardised with std 1" << std::endl;
    }
  }

  long global_step = 0;
  const auto start_time = std::chrono::steady_clock::now();
  if (pb_ac_trainer_reset(t, reset_seed) != PB_OK) return die(1, "reset");
  std::string last_ckpt;
  cli::PthSaver pth(argv[0]);
  const int n_params = pb_ac_trainer_num_params(t);
  std::vector<float> pth_p((size_t)n_params), pth_m((size_t)n_params), pth_v((size_t)n_params);
  if (mean.empty()) { mean.assign((size_t)O, 0.f); sd.assign((size_t)O, 1.f); }
  auto save_pth = [&](const std::string &model_name, const std::string &opt_name, float lr_now, bool rotate) {
    float sc[2] = {0.f, 0.f};   // scalars: lr, optimiser step count
    if (pb_ac_trainer_get_params(t, pth_p.data()) != PB_OK || pb_ac_trainer_read_buffer(t, "adam_m", pth_m.data(), n_params) != PB_OK ||
        pb_ac_trainer_read_buffer(t, "adam_v", pth_v.data(), n_params) != PB_OK || pb_ac_trainer_read_buffer(t, "scalars", sc, 2) != PB_OK) return;
    pb_env *ev = pb_ac_trainer_env(t);
    pth.save(exp_folder, model_name, opt_name, 1, O, pb_ac_trainer_act_dim(t), (long long)sc[1], lr_now, config.adam_eps, pth_p, pth_m, pth_v,
             pb_env_act_max(ev), pb_env_act_min(ev), &mean, &sd, rotate);
  };
  pb_ppo_stats st;
  ObsMoments run_moments;
  std::vector<float> params((size_t)pb_ac_trainer_num_params(t));
  auto save_params = [&](const std::string &path) {  // model_*.bin: flat fp32 parameters in torch registration order
    if (pb_ac_trainer_get_params(t, params.data()) != PB_OK) return false;
    std::ofstream o(path, std::ios::binary);
    o.write((const char *)params.data(), (std::streamsize)(params.size() * sizeof(float)));
    return (bool)o;
  };
  // the TCPStore of :360-372 becomes a counter file next to the NCCL id file (one node); single rank: nothing to pre-empt
  const bool preempt = config.use_dd_ppo_preempt && dist.world > 1;
  cli::SharedCounter ddppo;
  if (preempt) {
    const char *idf = std::getenv("PB_NCCL_ID_FILE");
    const std::string path = std::string(idf ? idf : "/tmp/pb_ddppo") + ".ddppo";
    if (dist.rank == 0 && !ddppo.open(path, true)) return die(6, "dd_ppo counter file");
    if (pb_ac_trainer_barrier(t) != PB_OK) return die(6, "barrier");
    if (dist.rank != 0 && !ddppo.open(path, false)) return die(6, "dd_ppo counter file");
  }
  for (long iteration = 0; iteration < config.num_iterations; iteration++) {
    float lr = config.learning_rate;
    if (config.anneal_lr) lr = (1.0f - (float)iteration / (float)config.num_iterations) * config.learning_rate;  // :626-631
    if (preempt) {
      // DD-PPO pre-emption (:629-631,680-693): rank 0 resets the shared counter, everybody meets, then every rank collects step by
      // step and stops once enough collectors of OTHER ranks have finished (the envs of one rank step together, so within a
      // rank there is no straggler); a finished rank adds its envs to the counter.
      if (dist.rank == 0) ddppo.reset();
      if (pb_ac_trainer_barrier(t) != PB_OK) return die(6, "barrier");
      if (pb_ac_trainer_collect_begin(t) != PB_OK) return die(1, "collect");
      int stopped = 0;
      for (int step = 0; step < config.num_steps; step++) {
        if (pb_ac_trainer_collect_step(t, step) != PB_OK || pb_ac_trainer_sync(t) != PB_OK) return die(1, "collect");
        if (pb_dd_ppo_should_preempt((int)ddppo.get(), config.num_envs, config.dd_ppo_preempt_threshold, step, config.dd_ppo_min_perc, config.num_steps)) { stopped = 1; break; }
      }
      ddppo.add(config.num_envs_per_device);
      if (pb_ac_trainer_collect_end(t, stopped) != PB_OK) return die(1, "collect");
    } else if (pb_ac_trainer_rollout(t) != PB_OK) return die(1, "rollout");
    if (config.estimate_mean_std) {
      long long rows = 0;
      if (pb_ac_trainer_obs_sums(t, sums.data(), &rows) != PB_OK) return die(1, "obs_sums");
      run_moments.add(sums, rows);
    }
    if (pb_ac_trainer_gae(t) != PB_OK) return die(1, "gae");
    if (pb_ac_trainer_update(t, lr) != PB_OK) return die(1, "update");
    if (pb_ac_trainer_stats(t, &st) != PB_OK) return die(1, "stats");
    global_step += config.batch_size;
    if (dist.rank != 0) continue;
    std::cout << "Time to collect data: " << st.rollout_ms * 1e-3 << " s\nTime to compute advantage: " << st.gae_ms * 1e-3
              << " s\nTime to train " << st.update_ms * 1e-3 << " s" << std::endl;
    if (st.episodes > 0) {
      const float r = st.episodic_return_sum / st.episodes, l = st.episodic_length_sum / st.episodes;
      std::cout << "global_step=" << global_step << ", episodic_return=" << std::fixed << std::setprecision(6) << r
                << std::defaultfloat << " (mean of " << st.episodes << " episodes)" << std::endl;
      logger.add_scalar("charts/episodic_return", global_step, r);
      logger.add_scalar("charts/episodic_length", global_step, l);
    }
    {  // checkpoint rotation (:892-909): keep only the latest
      char name[64];
      std::snprintf(name, sizeof(name), "model_latest_%09ld.bin", iteration);
      const std::string path = (exp_folder / name).string();
      if (save_params(path)) {
        if (!last_ckpt.empty()) fs::remove(last_ckpt);
        last_ckpt = path;
      }
      char mname[64], oname[64];   // the reference's own files (:892-909), through build/pth_writer
      std::snprintf(mname, sizeof(mname), "model_latest_%09ld.pth", iteration);
      std::snprintf(oname, sizeof(oname), "optimizer_latest_%09ld.pth", iteration);
      save_pth(mname, oname, lr, true);
    }
    const float passed = std::chrono::duration<float>(std::chrono::steady_clock::now() - start_time).count();
    float sps = 0;
    if (passed > 0) {
      sps = (float)global_step / passed;
      std::cout << std::fixed << std::setprecision(0) << "SPS: " << sps << std::defaultfloat << std::setprecision(6) << std::endl;
    }
    logger.add_scalar("charts/learning_rate", global_step, lr);
    logger.add_scalar("losses/value_loss", global_step, st.value_loss);
    logger.add_scalar("losses/policy_loss", global_step, st.policy_loss);
    logger.add_scalar("losses/entropy", global_step, st.entropy);
    logger.add_scalar("losses/old_approx_kl", global_step, st.old_approx_kl);
    logger.add_scalar("losses/approx_kl", global_step, st.approx_kl);
    logger.add_scalar("losses/clipfrac", global_step, st.clipfrac);
    logger.add_scalar("charts/SPS", global_step, sps);
    logger.flush();
  }
  if (dist.rank == 0) {
    save_params((exp_folder / "model_final.bin").string());
    save_pth("model_final.pth", "optimizer_final.pth", config.learning_rate, false);
    pth.wait();
  }

  if (config.estimate_mean_std && run_moments.rows > 0 && dist.rank == 0) {  // :956-963
    std::vector<float> mean, sd;
    run_moments.finish(mean, sd, false);
    print_vec("Mean obs:", mean);
    print_vec("Std obs:", sd);
  }

  if (dist.rank == 0 && config.num_eval_runs > 0) {  // final evaluation with the Beta mean (:965-1001)
    std::vector<float> returns((size_t)config.num_eval_runs + config.num_envs_per_device);
    int n = 0;
    if (pb_ac_trainer_eval(t, config.eval_seed, config.num_eval_runs, returns.data(), (int)returns.size(), &n) != PB_OK) return die(1, "eval");
    for (int i = 0; i < n; i++) {
      std::printf("Evaluation result: episode=%d episodic_return=%.2f \n", i, returns[i]);
      logger.add_scalar("eval/episodic_return", i, returns[i]);
    }
    const float avg = n ? std::accumulate(returns.begin(), returns.begin() + n, 0.0f) / n : 0.f;
    logger.add_scalar("eval/avg_return", n, avg);
    std::printf("Average evaluation return=%.2f over %d episodes\n", avg, n);
  }
  pb_ac_trainer_destroy(t);
  return 0;
}
