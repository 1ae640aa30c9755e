#!/usr/bin/env python
"""bench.py — env-steps/sec (SPS) of the PPO rollout-and-update hot path, Humanoid-v4 (BASELINE.json metric).

One "step" = one complete PPO iteration (rollout of num_steps x num_envs env steps with the fused policy, GAE, and
update_epochs x num_minibatches clipped-loss minibatch steps with Adam), same SPS definition as
src/ppo_continuous_action.cpp:567-571 (env steps / elapsed).  Weak scaling: 8192 envs per GPU (65536 at 8 GPUs, the
named config), one process per GPU, gradients averaged with one flat NCCL allreduce per minibatch.

  python bench.py [--gpus N --steps K --warmup W]            our arm (needs the CUDA library and a B200)
  python bench.py --impl reference [...]                       CPU arm: the oracle port on the host cores

Prints ONE JSON line on rank 0.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

ENV_ID = "Humanoid-v4"
OBS, ACT = 376, 17
ALG_BYTES_PER_ENV_STEP = 8172  # state r/w + per-env normaliser r/w + obs/action/scalars, Humanoid
# dram__bytes_read.sum + dram__bytes_write.sum of one env_step_kernel launch (8192 envs) from the ncu --set full capture
# summarised in profiles/r01_envstep_v3_ncu.txt; below the algorithmic bytes because the normaliser state stays in the 126 MB L2
NCU_TRAFFIC_BYTES = 29.2e6  # dram read + write of one env_step_kernel launch, profiles/r01_envstep_v9_ncu.txt
# FP32 thread-level operations one env_step_kernel launch (8192 Humanoid envs, bench state) executes, counted by ncu
# (2 x FFMA + FADD + FMUL, smsp__sass_thread_inst_executed_op_*_pred_on.sum; profiles/r01_envstep_v9_ncu.txt).  Counted
# from the kernel itself, so it includes whatever the kernel recomputes: it is the FP32-pipe utilisation figure
# north_star asks for, not an algorithmic minimum.  2 * 1.670e9 + 1.127e9 + 0.791e9 (FMA pipe active 13.6 % of cycles).
NCU_FP32_FLOP_PER_LAUNCH = 5.257e9


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return d.get("hbm_gbs", 6650.0), "measured"
    return 6650.0, "fallback"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled during the timed region."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.stop = index, [], False
        self.th = threading.Thread(target=self.run, daemon=True)

    def run(self):
        while not self.stop:
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits"],
                                     capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.rows.append([x.strip() for x in out.split(",")])
            except Exception:
                pass
            time.sleep(0.2)

    def __enter__(self):
        self.th.start()
        return self

    def __exit__(self, *a):
        self.stop = True
        self.th.join(timeout=6)

    def summary(self):
        if not self.rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["unsampled"]}
        sm = sorted(int(r[0]) for r in self.rows if r[0].isdigit())
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(r[2 + i].lower().startswith("active") for r in self.rows if len(r) > 2 + i)]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": int(self.rows[0][1]) if self.rows[0][1].isdigit() else None,
                "reasons": reasons}


# ------------------------------------------------------------------ CPU arm (oracle port on the host cores)
def cpu_iteration_sps(num_envs, num_steps, nthreads, iters=1, update=True):
    """One PPO iteration of the same workload on the CPU: oracle C physics/wrappers on `nthreads` host threads +
    torch-CPU agent forward, GAE and minibatch update (the reference's own execution model, gym.h:333-362)."""
    import torch

    from oracle import oracle_nn, pbo
    torch.set_num_threads(1)  # the 64-64 MLP on tens of rows is fastest single-threaded (measured)
    env = pbo.VecEnv(ENV_ID, num_envs, chain=1, nthreads=nthreads)
    agent = oracle_nn.Agent(OBS, ACT)
    opt = torch.optim.Adam(agent.parameters(), lr=3e-4, eps=1e-5)
    next_obs = torch.from_numpy(env.reset(1).copy())
    next_done = torch.zeros(num_envs)
    T, N = num_steps, num_envs
    obs = torch.zeros(T, N, OBS); actions = torch.zeros(T, N, ACT); logprobs = torch.zeros(T, N)
    rewards = torch.zeros(T, N); dones = torch.zeros(T, N); values = torch.zeros(T, N)
    t0 = time.perf_counter()
    for _ in range(iters):
        with torch.no_grad():
            for s in range(T):
                obs[s], dones[s] = next_obs, next_done
                a, lp, _, v = agent.get_action_and_value(next_obs)
                values[s], actions[s], logprobs[s] = v.flatten(), a, lp
                o, r, te, tr, _ = env.step(a.numpy())
                rewards[s] = torch.from_numpy(r.copy())
                next_obs = torch.from_numpy(o.copy())
                next_done = torch.from_numpy(np.maximum(te, tr).copy())
            adv, ret = oracle_nn.gae(rewards, values, dones, agent.get_value(next_obs).flatten(), next_done, 0.99, 0.95)
        if update:
            B = T * N
            b = [x.reshape(B, -1) if x.dim() == 3 else x.reshape(B) for x in (obs, actions, logprobs, adv, ret, values)]
            mbs = max(1, B // 32)
            for _e in range(10):
                perm = torch.randperm(B)
                for st in range(0, B - mbs + 1, mbs):
                    oracle_nn.ppo_minibatch_step(agent, opt, *b, perm[st:st + mbs])
    dt = time.perf_counter() - t0
    return iters * T * N / dt, dt


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    n_envs, T = 2048, 32  # bounded sample of the 8192-env workload: ~3-4 s per step on 16 threads, same minibatch / epoch counts
    for _ in range(min(args.warmup, 1)):
        cpu_iteration_sps(n_envs, 4, cores)
    vals, t_tot = [], 0.0
    for _ in range(args.steps):
        sps, dt = cpu_iteration_sps(n_envs, T, cores)
        vals.append(sps); t_tot += dt
        if t_tot > 150:
            break
    v = float(np.mean(vals))
    sample = f"{ENV_ID}, {n_envs} envs x {T} steps per iteration, oracle C physics on {cores} threads + torch-CPU agent/GAE/update"
    print(json.dumps({
        "impl": "reference", "metric": "env-steps/sec (SPS) Humanoid-v4 PPO iteration (rollout+GAE+update)", "value": v,
        "unit": "env-steps/s", "n_gpus": args.gpus, "steps": len(vals), "warmup": args.warmup,
        "ms_per_step": 1e3 * t_tot / len(vals), "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic",
        "config": {"workload": f"{ENV_ID} PPO, bounded CPU sample ({n_envs} envs, T={T}, 32 minibatches x 10 epochs)"},
        "cpu_baseline": {"value": v, "unit": "env-steps/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": v, "unit": "env-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }))


# ------------------------------------------------------------------ our arm
def run_ours(args):
    import torch
    import torch.distributed as dist

    from ppo_cpp_b200 import capi, trainer

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product path has no CPU fallback (use --impl reference)")
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    N, T, NMB, EP = args.num_envs, args.num_steps, 32, 10
    lr = 3e-4

    def make_warm_trainer():
        """Trainer at the start of iteration W+1 of training from reset(seed): the cost of a rollout drifts upward as the
        policy changes (more contacts per step), so both timed legs start from this same state."""
        tr = trainer.Trainer(env_id=ENV_ID, num_envs=N, num_steps=T, num_minibatches=NMB, update_epochs=EP, device=local,
                             seed=1, rank=rank, world_size=world)
        if world > 1:
            uid = [trainer.comm_unique_id() if rank == 0 else None]
            dist.broadcast_object_list(uid, src=0)
            tr.comm_init(uid[0], rank, world)
            tr.broadcast_params()
        tr.reset(1 + rank * N)
        for _ in range(args.warmup):
            tr.iteration(lr)
        tr.sync()
        return tr

    t = make_warm_trainer()
    stream = torch.cuda.ExternalStream(capi.lib().pb_trainer_stream(t._h))
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    roll_ms, upd_ms = [], []
    with ClockSampler(local) as cs:
        e0.record(stream)
        for _ in range(args.steps):
            t.iteration(lr)
        e1.record(stream)
        t.sync()
        torch.cuda.synchronize()
    st = t.stats()
    ms = e0.elapsed_time(e1)
    if world > 1:
        dist.barrier()
        tm = torch.tensor([ms], device="cuda")
        dist.all_reduce(tm, op=dist.ReduceOp.MAX)
        ms = float(tm.item())
    value = world * N * T * args.steps / (ms * 1e-3)

    # e2e: the public call a user makes (pb_trainer_iteration + pb_trainer_stats), wall clock around it; every step
    # uploads the iteration's host inputs (epoch permutations from pinned memory, lr) and reads the stats back.  Measured
    # on a fresh trainer brought to the same training state (iterations W+1..W+K) as the device-timed leg.
    t.close()
    if world > 1:
        dist.barrier()
    t = make_warm_trainer()
    stream = torch.cuda.ExternalStream(capi.lib().pb_trainer_stream(t._h))
    if world > 1:
        dist.barrier()
    w0 = time.perf_counter()
    e2e_iter_ms = []
    for _ in range(args.steps):
        wi = time.perf_counter()
        t.iteration(lr)
        s2 = t.stats()
        e2e_iter_ms.append(round((time.perf_counter() - wi) * 1e3, 2))
    w1 = time.perf_counter() - w0
    if world > 1:
        tw = torch.tensor([w1], device="cuda")
        dist.all_reduce(tw, op=dist.ReduceOp.MAX)
        w1 = float(tw.item())
    e2e = world * N * T * args.steps / w1
    h2d = EP * N * T * 4 + 4
    d2h = 14 * 4 + 16 + 16

    # dominant kernel (fused env step) timed alone on its launch stream, device pointers, warm state
    n_launch = 16
    act_ptr, _ = t.buffer_ptr("actions")
    obs_ptr, _ = t.buffer_ptr("next_obs")
    rew_ptr, _ = t.buffer_ptr("rewards")
    term_ptr, _ = t.buffer_ptr("term")
    trunc_ptr, _ = t.buffer_ptr("trunc")
    L = capi.lib()
    eh = L.pb_trainer_env(t._h)
    import ctypes as C
    k0, k1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    k0.record(stream)
    for i in range(n_launch):
        L.pb_env_step(eh, C.c_void_p(act_ptr + (i % T) * N * ACT * 4), C.c_void_p(obs_ptr), C.c_void_p(rew_ptr),
                      C.c_void_p(term_ptr), C.c_void_p(trunc_ptr), None, None, None)
    k1.record(stream)
    t.sync()
    kms = k0.elapsed_time(k1) / n_launch
    hbm_peak, peak_kind = peaks()
    achieved = ALG_BYTES_PER_ENV_STEP * N / (kms * 1e-3) / 1e9
    # the roof that actually bounds the rigid-body kernel: FP32 FMA throughput, measured here by a
    # register-resident FFMA kernel (pb_fp32_fma_peak) because MEASURED_PEAKS.json holds only HBM and bf16 tensor peaks
    fp32 = None
    if rank == 0:
        pk = C.c_float()
        if L.pb_fp32_fma_peak(local, C.c_float(20.0), C.byref(pk)) == 0 and pk.value > 0:
            fl = NCU_FP32_FLOP_PER_LAUNCH * (N / 8192) if NCU_FP32_FLOP_PER_LAUNCH else None
            fp32 = {"bound": "fp32_fma", "kernel": "env_step_kernel<CfgHumanoid>", "peak": pk.value, "peak_kind": "measured (pb_fp32_fma_peak)",
                    "unit": "TFLOP/s", "flop_per_launch": fl, "achieved": fl / (kms * 1e-3) / 1e12 if fl else None,
                    "frac": fl / (kms * 1e-3) / 1e12 / pk.value if fl else None,
                    "flop_source": "ncu thread-level 2*FFMA+FADD+FMUL of one launch at 8192 envs, scaled by N/8192"}
    exchange = {0: "none (single rank)", 1: "ncclAllReduce(avg)", 2: "fused peer-memory all-reduce kernel (CUDA IPC over NVLink)"}[
        L.pb_trainer_exchange_mode(t._h)]

    if rank == 0:
        cpu = None
        if world == 1 and not args.no_cpu_baseline:
            cores = os.cpu_count() or 1
            n_cpu = N if cores >= 8 else 2048  # the full per-GPU workload (~14 s on 16 threads), else a quarter of it
            sps, dt = cpu_iteration_sps(n_cpu, 32, cores)
            cpu = {"value": sps, "unit": "env-steps/s", "cores": cores, "kind": "port",
                   "sample": f"{ENV_ID}, {n_cpu} envs x 32 steps + GAE + 320 minibatch updates on {cores} host threads ({dt:.1f} s): "
                             "oracle C physics + torch-CPU agent"}
        # per iteration: T x (policy, queue order, env step) + bootstrap policy + GAE + 2 weight transposes (+ 2 TF32 weight
        # splits on the tcgen05 path) + per epoch (gather, advantage statistics) + per minibatch: forward/backward, weight
        # gradients, reduce, Adam (+ tcgen05 first-layer forward, + tcgen05 first-layer weight gradient, + the squared-norm
        # partials after the allreduce when N > 1; NCCL's own kernel is not counted)
        tc_path = os.environ.get("PB_UPDATE_MMA", "2") == "2"
        # N > 1: one more kernel of ours per minibatch either way — peer_avg_kernel (the exchange itself), or
        # sqnorm_partials_kernel after ncclAllReduce
        xmode = L.pb_trainer_exchange_mode(t._h)
        fused_x = xmode == 2 and os.environ.get("PB_PEER_FUSED", "0") != "0"   # slab reduction + exchange are one kernel
        per_mb = 4 + (1 if tc_path else 0) + (1 if tc_path and os.environ.get("PB_DW1_TC", "1") != "0" else 0) + (1 if world > 1 and not fused_x else 0)
        launches_per_iter = 3 * T + 2 + 2 + (2 if tc_path else 0) + EP * 2 + EP * NMB * per_mb
        print(json.dumps({
            "metric": "env-steps/sec (SPS) Humanoid-v4 PPO iteration (rollout+GAE+update)", "value": value, "unit": "env-steps/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"{ENV_ID} PPO: {N} envs/GPU x {T} steps, {NMB} minibatches x {EP} epochs, 64-64 tanh MLP, "
                                   f"random-init policy, reset(seed) states; both timed legs cover training iterations "
                                   f"{args.warmup + 1}..{args.warmup + args.steps} (rollout cost rises with training as contacts increase)",
                       "num_envs_per_gpu": N, "num_steps": T,
                       "l2": "working set per iteration (rollout obs 394 MB + per-env normalisers 25 MB) exceeds the 126 MB L2",
                       "parallelism": f"dp{world}"},
            "phase_ms": {"rollout": st["rollout_ms"], "gae": st["gae_ms"], "update": st["update_ms"]},
            "update_ms_per_iteration": st["update_ms"],
            "roofline": {"bound": "hbm", "kernel": "env_step_kernel<CfgHumanoid>", "achieved": achieved, "peak": hbm_peak,
                         "peak_kind": peak_kind, "unit": "GB/s", "frac": achieved / hbm_peak, "traffic": NCU_TRAFFIC_BYTES,
                         "alg_bytes_per_launch": ALG_BYTES_PER_ENV_STEP * N,
                         "kernel_ms": kms, "env_steps_per_s_kernel_only": N / (kms * 1e-3),
                         "note": "instruction-issue/latency-bound rigid-body kernel (ncu: IPC 1.55 of 4, ~8.7k warp-instructions per forward "
                                 "evaluation, 20 evaluations per env step); HBM is not the binding roof (SURVEY.md 8d, DESIGN.md 4.1)"},
            "roofline_fp32": fp32,
            "gradient_exchange": exchange,
            "cpu_baseline": cpu,
            "e2e": {"value": e2e, "unit": "env-steps/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "call": "pb_trainer_iteration + pb_trainer_stats (wall clock, host permutations uploaded from pinned memory)",
                    "iter_ms": e2e_iter_ms, "phase_ms_last": {"rollout": s2["rollout_ms"], "gae": s2["gae_ms"], "update": s2["update_ms"]}},
            "gpu_launches": launches_per_iter * args.steps,
            "clocks": cs.summary(),
            "sim_counters": {"forward_evals": st["forward_evals"], "overflow_envs": st["overflow_envs"], "episodes_last_iter": s2["episodes"]},
        }))
    t.close()
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--num-envs", type=int, default=8192, help="envs per GPU")
    ap.add_argument("--num-steps", type=int, default=32)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
