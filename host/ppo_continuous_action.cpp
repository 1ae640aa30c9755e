// ppo_continuous_action — command-line driver with the flags, defaults, stdout lines and scalar tags of the reference's
// src/ppo_continuous_action.cpp, on top of libppo_b200.so (pb_trainer_* C ABI).  The iteration loop of the reference
// (:375-586) maps to: lr anneal (:379-384) -> pb_trainer_rollout (:387-434) -> pb_trainer_gae (:447-467) ->
// pb_trainer_update (:489-542) -> checkpoint rotation (:546-563) -> SPS + scalars (:565-583); the final evaluation
// (:589-626) is pb_trainer_eval.  Data parallel over several GPUs: one process per GPU, PB_RANK / PB_WORLD /
// PB_NCCL_ID_FILE (or torchrun's RANK / WORLD_SIZE / LOCAL_RANK), --num_envs is the global env count.
#include <cmath>
#include <filesystem>
#include <iomanip>
#include <numeric>
#include <sstream>

#include "cli_common.h"

namespace fs = std::filesystem;

struct GlobalConfig {  // src/ppo_continuous_action.cpp:51-81, same names and defaults
  int seed = 1;
  int eval_seed = 2;
  int total_timesteps = 1000000;
  float learning_rate = 3e-4f;
  int num_envs = 1;
  int num_steps = 2048;
  float gamma = 0.99f;
  float gae_lambda = 0.95f;
  int num_minibatches = 32;
  int update_epochs = 10;
  bool norm_adv = true;
  float clip_coef = 0.2f;
  bool clip_vloss = true;
  float ent_coef = 0.0f;
  float vf_coef = 0.5f;
  float max_grad_norm = 0.5f;
  float adam_eps = 1e-5f;
  bool anneal_lr = true;
  int num_eval_runs = 10;
  bool clip_actions = true;
  bool torch_deterministic = true;  // accepted for compatibility; the kernels are deterministic per env already
  std::string exp_name_stem = "PPO_002";
  std::string env_id = "Humanoid-v4";
  std::string render = "rgb_array";  // "human" (GLFW/OpenGL) is unavailable on this box
  std::string load_checkpoint;       // new: resume from a pb_trainer_save file
  std::string exp_name;
  int batch_size = 0, minibatch_size = 0, num_iterations = 0;

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
      << torch_deterministic << "|\n|exp_name_stem|" << exp_name_stem << "|\n|env_id|" << env_id << "|\n";
    return o.str();
  }
};

static int die(int code, const std::string &what) {
  std::cerr << what << ": " << pb_last_error() << std::endl;
  return code;
}

int main(int argc, char **argv) {
  GlobalConfig config;
  cli::Flags f;
  f.add("seed", &config.seed, "Training seed");
  f.add("eval_seed", &config.eval_seed, "Seed of final evaluation run");
  f.add("total_timesteps", &config.total_timesteps, "Number of environment steps");
  f.add("learning_rate", &config.learning_rate, "Adam learning rate");
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
  f.add("num_envs", &config.num_envs, "Number of environments to be used (all GPUs together).");
  f.add("render", &config.render, "rgb_array only: the GLFW/OpenGL 'human' mode is unavailable on this box");
  f.add("load_checkpoint", &config.load_checkpoint, "Resume from a checkpoint written by this program");
  int pr = f.parse(argc, argv);
  if (pr == 1) return 0;
  if (pr == 2) return 1;

  const cli::Dist dist = cli::Dist::from_env();
  if (config.num_envs % dist.world != 0) { std::cerr << "num_envs must be divisible by the number of devices" << std::endl; return 1; }
  const int envs_per_device = config.num_envs / dist.world;
  config.exp_name = config.exp_name_stem + "_" + std::to_string(config.seed);
  config.batch_size = config.num_steps * config.num_envs;
  config.minibatch_size = config.batch_size / config.num_minibatches;
  config.num_iterations = config.total_timesteps / config.batch_size;
  if (config.render != "rgb_array") std::cerr << "--render " << config.render << ": unavailable on this box, continuing headless" << std::endl;

  fs::path exe = fs::canonical(argv[0]);
  fs::path exp_folder = exe.parent_path() / ".." / "models" / config.exp_name;
  if (dist.rank == 0) fs::create_directories(exp_folder);
  cli::ScalarLogger logger(dist.rank == 0 ? (exp_folder / "scalars.jsonl").string() : std::string("/dev/null"),
                           dist.rank == 0 ? (exp_folder / "tfevents_logs.pb").string() : std::string());
  logger.add_text("hyperparameters", 0, config.to_string());

  pb_ppo_config pc;
  pb_ppo_config_default(&pc);
  pc.env_id = config.env_id.c_str(); pc.num_envs = envs_per_device; pc.num_steps = config.num_steps; pc.device = dist.local_rank;
  pc.seed = config.seed; pc.gamma = config.gamma; pc.gae_lambda = config.gae_lambda; pc.num_minibatches = config.num_minibatches;
  pc.update_epochs = config.update_epochs; pc.norm_adv = config.norm_adv; pc.clip_coef = config.clip_coef; pc.clip_vloss = config.clip_vloss;
  pc.ent_coef = config.ent_coef; pc.vf_coef = config.vf_coef; pc.max_grad_norm = config.max_grad_norm; pc.adam_eps = config.adam_eps;
  pc.clip_actions = config.clip_actions; pc.rank = dist.rank; pc.world_size = dist.world;
  pb_trainer *t = nullptr;
  int rc = pb_trainer_create(&pc, &t);
  if (rc == PB_ERR_BAD_ENV) { std::cerr << "Env not implemented. Use one of: Humanoid-v4, HalfCheetah-v5, Hopper-v5, Ant-v5" << std::endl; return 1; }
  if (rc == PB_ERR_BAD_DEVICE) return die(2, "device");
  if (rc) return die(1, "pb_trainer_create");
  if (dist.world > 1) {
    unsigned char id[128];
    if (dist.exchange_unique_id(id) != 0 || pb_trainer_comm_init(t, id, dist.rank, dist.world) != PB_OK) return die(6, "comm init");
    dist.remove_unique_id();   // every rank has read it once the collective init returned
    if (pb_trainer_broadcast_params(t) != PB_OK) return die(6, "broadcast");
  }
  if (!config.load_checkpoint.empty() && pb_trainer_load(t, config.load_checkpoint.c_str()) != PB_OK) return die(1, "load_checkpoint");

  long global_step = 0;
  const auto start_time = std::chrono::steady_clock::now();
  // env i of rank r is seeded seed + r*envs_per_device + i (the reference seeds env i with seed + i, gym.h:313)
  if (pb_trainer_reset(t, config.seed + dist.rank * envs_per_device) != PB_OK) return die(1, "reset");
  std::string last_ckpt;
  cli::PthSaver pth(argv[0]);
  const int n_params = pb_trainer_num_params(t);
  std::vector<float> pth_p((size_t)n_params), pth_m((size_t)n_params), pth_v((size_t)n_params);
  auto save_pth = [&](const std::string &model_name, const std::string &opt_name, float lr_now, bool rotate) {
    float sc[2] = {0.f, 0.f};   // scalars: lr, optimiser step count
    if (pb_trainer_get_params(t, pth_p.data()) != PB_OK || pb_trainer_read_buffer(t, "adam_m", pth_m.data(), n_params) != PB_OK ||
        pb_trainer_read_buffer(t, "adam_v", pth_v.data(), n_params) != PB_OK || pb_trainer_read_buffer(t, "scalars", sc, 2) != PB_OK) return;
    pth.save(exp_folder, model_name, opt_name, 0, pb_env_obs_dim(pb_trainer_env(t)), pb_env_act_dim(pb_trainer_env(t)), (long long)sc[1], lr_now,
             config.adam_eps, pth_p, pth_m, pth_v, 0.f, 0.f, nullptr, nullptr, rotate);
  };
  pb_ppo_stats st;
  cli::TicToc tt;
  for (int iteration = 0; iteration < config.num_iterations; iteration++) {
    float lr = config.learning_rate;
    if (config.anneal_lr) lr = (1.0f - (float)iteration / (float)config.num_iterations) * config.learning_rate;  // :379-384
    tt.tic();
    if (pb_trainer_rollout(t) != PB_OK) return die(1, "rollout");
    if (pb_trainer_gae(t) != PB_OK) return die(1, "gae");
    if (pb_trainer_update(t, lr) != PB_OK) return die(1, "update");
    if (pb_trainer_stats(t, &st) != PB_OK) return die(1, "stats");
    global_step += config.batch_size;
    if (dist.rank != 0) continue;
    std::cout << "Time to collect data: " << st.rollout_ms * 1e-3 << " s\nTime to compute advantage: " << st.gae_ms * 1e-3
              << " s\nTime to train " << st.update_ms * 1e-3 << " s" << std::endl;
    if (st.episodes > 0) {  // the reference prints one line per finished episode (:413-433); here: the iteration's mean
      const float r = st.episodic_return_sum / st.episodes, l = st.episodic_length_sum / st.episodes;
      std::cout << "global_step=" << global_step << ", episodic_return=" << std::fixed << std::setprecision(6) << r
                << std::defaultfloat << " (mean of " << st.episodes << " episodes)" << std::endl;
      logger.add_scalar("charts/episodic_return", global_step, r);
      logger.add_scalar("charts/episodic_length", global_step, l);
    }
    {  // checkpoint rotation (:546-563): keep only the latest
      char name[64];
      std::snprintf(name, sizeof(name), "model_latest_%09d.bin", iteration);
      const std::string path = (exp_folder / name).string();
      if (pb_trainer_save(t, path.c_str()) == PB_OK) {
        if (!last_ckpt.empty()) fs::remove(last_ckpt);
        last_ckpt = path;
      }
      char mname[64], oname[64];   // the reference's own files, through build/pth_writer
      std::snprintf(mname, sizeof(mname), "model_latest_%09d.pth", iteration);
      std::snprintf(oname, sizeof(oname), "optimizer_latest_%09d.pth", iteration);
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
    pb_trainer_save(t, (exp_folder / "model_final.bin").string().c_str());
    save_pth("model_final.pth", "optimizer_final.pth", config.learning_rate, false);   // :587 (the reference saves the model only)
    pth.wait();
  }

  if (dist.rank == 0 && config.num_eval_runs > 0) {  // final evaluation on the training envs (:589-626)
    std::vector<float> returns(config.num_eval_runs + envs_per_device);
    int n = 0;
    if (pb_trainer_eval(t, config.eval_seed, config.num_eval_runs, returns.data(), (int)returns.size(), &n) != PB_OK) return die(1, "eval");
    for (int i = 0; i < n; i++) {
      std::printf("Evaluation result: episode=%d episodic_return=%.2f \n", i, returns[i]);
      logger.add_scalar("eval/episodic_return", i, returns[i]);
    }
    const float avg = n ? std::accumulate(returns.begin(), returns.begin() + n, 0.0f) / n : 0.f;
    logger.add_scalar("eval/avg_return", n, avg);
    std::printf("Average evaluation return=%.2f over %d episodes\n", avg, n);
  }
  pb_trainer_destroy(t);
  return 0;
}
