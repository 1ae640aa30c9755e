"""The C++ command-line driver (host/ppo_continuous_action.cpp): flag surface and exit codes of the reference's main
(src/ppo_continuous_action.cpp:199-272, 330-335) on the CPU; a short training run, checkpoint rotation, resume and the
final evaluation on the GPU."""
import json
import os
import re
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "build", "ppo_continuous_action")

REFERENCE_FLAGS = ["seed", "eval_seed", "total_timesteps", "learning_rate", "num_steps", "gamma", "gae_lambda", "num_minibatches",
                   "update_epochs", "norm_adv", "clip_coef", "clip_vloss", "ent_coef", "vf_coef", "max_grad_norm", "adam_eps",
                   "anneal_lr", "num_eval_runs", "clip_actions", "torch_deterministic", "exp_name_stem", "env_id", "num_envs", "render"]


@pytest.fixture(scope="module")
def binary():
    if not os.path.exists(BIN):
        subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "ppo_cpp_b200", "csrc")])
        subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "host")])
    return BIN


def test_help_lists_every_reference_flag_with_its_default(binary):
    out = subprocess.run([binary, "--help"], capture_output=True, text=True)
    assert out.returncode == 0
    for f in REFERENCE_FLAGS:
        assert f"--{f} " in out.stdout, f
    # defaults of GlobalConfig (src/ppo_continuous_action.cpp:53-76)
    for frag in ["--num_steps  Num environment steps per iteration (default 2048)", "(default Humanoid-v4)", "(default PPO_002)",
                 "--update_epochs  Number of training epochs per iteration (default 10)", "--clip_coef  PPO clip coefficient (default 0.2"]:
        assert frag in out.stdout, frag


def test_unknown_flag_and_unknown_env_exit_1(binary):
    assert subprocess.run([binary, "--nonsense", "1"], capture_output=True).returncode == 1
    r = subprocess.run([binary, "--env_id", "Walker2d-v5", "--total_timesteps", "64"], capture_output=True, text=True)
    assert r.returncode == 1 and "Env not implemented" in r.stderr


@pytest.mark.gpu
def test_short_training_run_checkpoints_and_evaluates(binary, tmp_path):
    args = [binary, "--env_id", "Hopper-v5", "--num_envs", "256", "--num_steps", "16", "--total_timesteps", str(256 * 16 * 3),
            "--num_minibatches", "4", "--update_epochs", "2", "--num_eval_runs", "8", "--exp_name_stem", "pytest_cli", "--seed", "3"]
    r = subprocess.run(args, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]
    assert len(re.findall(r"^SPS: \d+$", r.stdout, re.M)) == 3
    assert "Time to collect data:" in r.stdout and "Time to train" in r.stdout
    assert re.search(r"global_step=\d+, episodic_return=", r.stdout)
    assert len(re.findall(r"Evaluation result: episode=\d+ episodic_return=", r.stdout)) >= 8
    assert "Average evaluation return=" in r.stdout
    folder = os.path.join(ROOT, "models", "pytest_cli_3")
    files = sorted(os.listdir(folder))
    assert "model_final.bin" in files and "model_latest_000000002.bin" in files
    assert not any(f.startswith("model_latest_00000000") and f.endswith(".bin") and f != "model_latest_000000002.bin" for f in files)  # rotation
    if os.path.exists(os.path.join(ROOT, "build", "pth_writer")):
        # the reference's own files (src/ppo_continuous_action.cpp:546-563,587), written through libtorch's torch::save
        import torch
        from ppo_cpp_b200 import pth_compat
        pth = [f for f in files if f.endswith(".pth")]
        assert sorted(pth) == ["model_final.pth", "model_latest_000000002.pth", "optimizer_final.pth", "optimizer_latest_000000002.pth"], pth
        named = dict(torch.jit.load(os.path.join(folder, "model_final.pth")).named_parameters())
        flat = np.concatenate([named[n].detach().numpy().ravel() for n in pth_compat.NAMES])
        assert np.array_equal(flat, pth_compat.read_checkpoint(os.path.join(folder, "model_final.bin"))["params"])
    tags = {json.loads(line)["tag"] for line in open(os.path.join(folder, "scalars.jsonl"))}
    for t in ["hyperparameters", "charts/learning_rate", "charts/SPS", "losses/value_loss", "losses/policy_loss", "losses/entropy",
              "losses/old_approx_kl", "losses/approx_kl", "losses/clipfrac", "eval/episodic_return", "eval/avg_return"]:
        assert t in tags, t
    # the same scalars in the reference's TensorBoard event file (src/ppo_continuous_action.cpp:281)
    loader = pytest.importorskip("tensorboard.backend.event_processing.event_file_loader")
    tb_tags = {v.tag for e in loader.LegacyEventFileLoader(os.path.join(folder, "tfevents_logs.pb")).Load() for v in e.summary.value}
    assert tags <= tb_tags, tags - tb_tags
    # resume: the checkpoint loads into a fresh process
    r2 = subprocess.run(args[:-2] + ["--seed", "3", "--load_checkpoint", os.path.join(folder, "model_final.bin"),
                                     "--total_timesteps", str(256 * 16)], capture_output=True, text=True, timeout=300)
    assert r2.returncode == 0, r2.stderr[-2000:]


@pytest.mark.gpu
def test_checkpoint_roundtrip_restores_params_moments_and_normalisers(tmp_path):
    from ppo_cpp_b200 import trainer
    kw = dict(env_id="Hopper-v5", num_envs=64, num_steps=8, num_minibatches=2, update_epochs=2)
    a = trainer.Trainer(**kw)
    a.reset(1)
    a.iteration(3e-4)
    path = str(tmp_path / "ck.bin")
    a.save(path)
    b = trainer.Trainer(**kw)
    b.load(path)
    assert np.array_equal(a.get_params(), b.get_params())
    assert np.array_equal(a.read("adam_m"), b.read("adam_m")) and np.array_equal(a.read("adam_v"), b.read("adam_v"))
    # same minibatch from the same state -> identical next parameters (step count and moments restored)
    for name in ("obs", "actions", "logprobs", "advantages", "returns", "values"):
        b.write(name, a.read(name))
    idx = np.arange(64 * 8 // 2, dtype=np.int32)
    a.update_minibatch(idx, 3e-4); b.update_minibatch(idx, 3e-4)
    assert np.array_equal(a.get_params(), b.get_params())
    c = trainer.Trainer(env_id="Hopper-v5", num_envs=32, num_steps=8, num_minibatches=2, update_epochs=2)
    with pytest.raises(Exception):
        c.load(path)
    for t in (a, b, c):
        t.close()


# ---------------------------------------------------------------- ac_ppo_continuous_action (src/ac_ppo_continuous_action.cpp)
AC_BIN = os.path.join(ROOT, "build", "ac_ppo_continuous_action")
AC_REFERENCE_FLAGS = REFERENCE_FLAGS + ["gpu_ids", "collect_device", "train_device", "rdzv_addr", "tcp_store_port", "use_dd_ppo_preempt",
                                        "dd_ppo_min_perc", "dd_ppo_preempt_threshold", "estimate_mean_std"]


def test_ac_help_lists_every_reference_flag_with_its_default(binary):
    out = subprocess.run([AC_BIN, "--help"], capture_output=True, text=True)
    assert out.returncode == 0
    for f in AC_REFERENCE_FLAGS:
        assert f"--{f} " in out.stdout, f
    # defaults of GlobalConfig (src/ac_ppo_continuous_action.cpp:58-90)
    for frag in ["--total_timesteps  Number of environment steps (default 10000000)", "(default Ant-v5)", "(default Ant-v5_AC_PPO_Atari)",
                 "--num_envs  Number of environments to be used (all devices together). (default 8)",
                 "--num_steps  Num environment steps per iteration (default 128)", "--clip_coef  PPO clip coefficient (default 0.1",
                 "--ent_coef  Weigth of entropy loss. (default 0.01", "--num_eval_runs  How many environments to evaluate (default 128)",
                 "--learning_rate  Adam learning rate (default 0.00025"]:
        assert frag in out.stdout, frag


def test_ac_unknown_flag_and_indivisible_envs_exit_1(binary):
    assert subprocess.run([AC_BIN, "--nonsense", "1"], capture_output=True).returncode == 1
    env = dict(os.environ, PB_WORLD="3", PB_RANK="0")
    r = subprocess.run([AC_BIN, "--num_envs", "8"], capture_output=True, text=True, env=env)
    assert r.returncode == 1 and "multiple of num_devices" in r.stderr


@pytest.mark.gpu
def test_ac_short_training_run(binary, tmp_path):
    args = [AC_BIN, "--env_id", "Hopper-v5", "--num_envs", "256", "--num_steps", "16", "--total_timesteps", str(256 * 16 * 3),
            "--num_eval_runs", "8", "--exp_name_stem", "pytest_ac_cli", "--seed", "3", "--estimate_mean_std", "1", "--obs_stats_file", "estimate"]
    r = subprocess.run(args, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]
    assert "Observation statistics estimated from 4096 warm-up observations" in r.stdout
    if os.path.exists(os.path.join(ROOT, "build", "pth_writer")):
        files = os.listdir(os.path.join(ROOT, "models", "pytest_ac_cli_3"))
        assert {"model_final.pth", "optimizer_final.pth"} <= set(files) and sum(f.startswith("model_latest_") and f.endswith(".pth") for f in files) == 1
    assert len(re.findall(r"^SPS: \d+$", r.stdout, re.M)) == 3
    assert "Mean obs:" in r.stdout and "Std obs:" in r.stdout
    assert len(re.findall(r"Evaluation result: episode=\d+ episodic_return=", r.stdout)) >= 8
    folder = os.path.join(ROOT, "models", "pytest_ac_cli_3")
    files = sorted(os.listdir(folder))
    assert "model_final.bin" in files and "model_latest_000000002.bin" in files
    assert not any(f.startswith("model_latest_00000000") and f.endswith(".bin") and f != "model_latest_000000002.bin" for f in files)
    # 2 encoders (11->256->256 with LayerNorm) + value head + two 3-wide Beta heads
    n = 2 * (11 * 256 + 256 + 512 + 256 * 256 + 256 + 512) + 257 + 2 * (256 * 3 + 3)
    assert os.path.getsize(os.path.join(folder, "model_final.bin")) == 4 * n
    # the statistics file path: mean line + std line
    stats = tmp_path / "stats.txt"
    stats.write_text(", ".join(["0.1"] * 11) + "\n" + ", ".join(["2.0"] * 11) + "\n")
    r2 = subprocess.run(args[:-2] + ["--obs_stats_file", str(stats), "--total_timesteps", str(256 * 16)], capture_output=True, text=True,
                        timeout=300)
    assert r2.returncode == 0 and "warm-up" not in r2.stdout, r2.stderr[-2000:]


@pytest.mark.gpu
def test_ac_cli_defaults_to_the_reference_observation_tables():
    """src/ac_ppo_continuous_action.cpp:496-497: Humanoid-v4 standardises with the fixed tables; their 19 zero-std entries would
    give inf / NaN in the reference and standardise with std 1 here (printed)."""
    args = [AC_BIN, "--env_id", "Humanoid-v4", "--num_envs", "64", "--num_steps", "16", "--total_timesteps", str(64 * 16 * 2),
            "--num_eval_runs", "2", "--exp_name_stem", "pytest_ac_tab", "--seed", "3"]
    r = subprocess.run(args, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]
    assert "Observation statistics: reference table, 19 zero-std entries standardised with std 1" in r.stdout
    assert "nan" not in r.stdout.lower()


def _selftest_reference(env_id, n, steps):
    """What host/vector_env_selftest.cpp prints, computed through the Python mirror of the same C ABI."""
    from ppo_cpp_b200 import capi

    def checksum(a):
        a = np.asarray(a, np.float64).ravel()
        return float((a * ((np.arange(a.size) % 7) + 1)).sum())
    v = capi.ParVectorEnv(env_id, n, clip_actions=True, chain=capi.PB_CHAIN_PPO)
    out = [("reset", checksum(v.reset(1)))]
    episodes, ret_sum = 0, 0.0
    for t in range(steps):
        i = np.arange(n * v.act_dim, dtype=np.float32)
        act = (np.float32(2.0) * np.sin(np.float32(0.37) * np.float32(t + 1) + np.float32(0.11) * i)).astype(np.float32).reshape(n, v.act_dim)
        o, r, te, tr, infos = v.step(act)
        for inf in infos:
            if inf is not None:
                episodes += 1
                ret_sum += inf[0]
        out.append(("step", checksum(o), checksum(r), float(te.sum()), float(tr.sum())))
    v.close()
    return out, episodes, ret_sum


@pytest.mark.gpu
@pytest.mark.parametrize("binary", ["vector_env_selftest", "gym_torch_selftest"])
def test_cpp_vector_env_mirrors_match_the_python_mirror(binary):
    """include/b200_vector_env.h (plain C++) and include/b200_gym_torch.h (the literal gymcpp libtorch signatures, gym.h:310,324)
    driven like a gymcpp caller: every reset / step output and the episode infos equal the Python mirror's for the same actions."""
    exe = os.path.join(ROOT, "build", binary)
    if not os.path.exists(exe):
        pytest.skip(f"{binary} not built (no libtorch at build time)")
    env_id, n, steps = "Hopper-v5", 64, 60
    r = subprocess.run([exe, env_id, str(n), str(steps)], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [ln.split() for ln in r.stdout.strip().splitlines()]
    ref, episodes, ret_sum = _selftest_reference(env_id, n, steps)
    # sinf in C++ and numpy's float32 sin may differ in the last ulp of an action; one contact-rich step amplifies that to ~1e-4
    assert lines[0][0] == "reset" and abs(float(lines[0][1]) - ref[0][1]) <= 1e-6 * abs(ref[0][1])
    for t in range(steps):
        got, exp = lines[1 + t], ref[1 + t]
        assert got[0] == "step" and int(got[1]) == t
        assert abs(float(got[2]) - exp[1]) <= 2e-3 * max(1.0, abs(exp[1])), (t, got, exp)
        assert abs(float(got[3]) - exp[2]) <= 2e-3 * max(1.0, abs(exp[2])), (t, got, exp)
        if t < 5:
            assert float(got[4]) == exp[3] and float(got[5]) == exp[4]
    ep = [ln for ln in lines if ln[0] == "episodes"][0]
    assert int(ep[1]) > 0 and abs(int(ep[1]) - episodes) <= max(2, episodes // 10)
    sp = [ln for ln in lines if ln[0] == "spaces"][0]
    assert [int(sp[1]), int(sp[2]), float(sp[3]), float(sp[4])] == [11, 3, -1.0, 1.0]
