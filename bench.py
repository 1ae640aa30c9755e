#!/usr/bin/env python
"""bench.py — env-steps/sec (SPS) of the PPO rollout-and-update hot path, Humanoid-v4 (BASELINE.json metric).

One "step" = one complete PPO iteration (rollout of num_steps x num_envs env steps with the fused policy, GAE, and
update_epochs x num_minibatches clipped-loss minibatch steps with Adam), same SPS definition as
src/ppo_continuous_action.cpp:567-571 (env steps / elapsed).  Weak scaling: 8192 envs per GPU (65536 at 8 GPUs, the
named config), one process per GPU, gradients averaged once per minibatch.

PINNED STATE (SURVEY 8d "walking" workload): every leg starts from the committed policy snapshot after 50 training iterations
(tests/golden/walking_Humanoid-v4.npz: parameters + observation / reward normaliser) and runs its iterations with lr = 0, so
the policy - and with it the contact load of the rigid-body kernel - is the same in this file, under ncu and in the driver's
run, whatever --warmup / --steps are.  An iteration with lr = 0 executes exactly the kernels and FLOPs of one with lr > 0
(forward/backward, weight gradients, exchange, clip, Adam moments); only the parameter delta is zero.

  python bench.py [--gpus N --steps K --warmup W]            our arm (needs the CUDA library and a B200)
  python bench.py --impl reference [...]                       CPU arm: the oracle port on the host cores, same workload
  python bench.py --workload ac [--impl reference] [...]      BASELINE config 4: HalfCheetah-v5 ac_ppo_continuous_action, 8192 envs x 128 steps per GPU
  python bench.py --num-envs E --gpus N                        strong scaling: E = 65536 / N envs per GPU

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
NMB, EPOCHS = 32, 10
SNAPSHOT = os.path.join(ROOT, "tests", "golden", "walking_Humanoid-v4.npz")
# ncu figures of one env_step_kernel launch AT THE PINNED STATE (8192 envs): DRAM traffic, executed FP32 thread operations;
# written by tools/ncu_to_json.py from the committed capture, next to the oracle-counted algorithmic FLOPs per env step
KERNEL_PROFILE = os.path.join(ROOT, "profiles", "r02_envstep_walk.json")
ALG_BYTES_PER_ENV_STEP = 8172  # state r/w + per-env normaliser r/w + obs/action/scalars, Humanoid
METRIC = "env-steps/sec (SPS) Humanoid-v4 PPO iteration (rollout+GAE+update)"


def workload(num_envs, num_steps):
    """The same string on both arms (the driver compares the arms' configs)."""
    return (f"{ENV_ID} PPO at the pinned walking state (policy + normaliser after 50 iterations, lr = 0): {num_envs} envs/GPU x "
            f"{num_steps} steps, {NMB} minibatches x {EPOCHS} epochs, 64-64 tanh MLP")


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return d.get("hbm_gbs", 6650.0), "measured"
    return 6650.0, "fallback"


def kernel_profile():
    if os.path.exists(KERNEL_PROFILE):
        return json.load(open(KERNEL_PROFILE))
    return {}


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
    torch-CPU agent forward, GAE and minibatch update (the reference's own execution model, gym.h:333-362), from the same
    pinned state: snapshot policy + normaliser, env states = the snapshot's harvested walking states tiled over the envs."""
    import torch

    from oracle import oracle_nn, pbo
    torch.set_num_threads(1)  # the reference pins torch to one thread (src/ppo_continuous_action.cpp:187-188)
    snap = np.load(SNAPSHOT)
    env = pbo.VecEnv(ENV_ID, num_envs, chain=1, nthreads=nthreads)
    agent = oracle_nn.Agent(OBS, ACT)
    agent.load_flat(snap["params"])
    opt = torch.optim.Adam(agent.parameters(), lr=0.0, eps=1e-5)
    env.reset(1)
    env.set_norm_stats(snap["ob_mean"], snap["ob_var"], float(snap["ob_count"]), snap["rw4"])
    k = len(snap["qpos"])
    for i in range(num_envs):
        s = env.sim(i)
        s.qpos[:] = snap["qpos"][i % k]
        s.qvel[:] = snap["qvel"][i % k]
        s.qacc_warmstart[:] = snap["warm"][i % k]
    # one untimed step so next_obs / stale kinematics belong to the walking states
    o, _, te, tr, _ = env.step(np.zeros((num_envs, ACT), np.float32))
    next_obs = torch.from_numpy(o.copy())
    next_done = torch.from_numpy(np.maximum(te, tr).copy())
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
            mbs = max(1, B // NMB)
            for _e in range(EPOCHS):
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
    n_envs, T = args.num_envs, args.num_steps   # the same per-GPU workload as our arm; the run is bounded in time, not in size
    for _ in range(min(args.warmup, 1)):
        cpu_iteration_sps(min(n_envs, 512), 4, cores)
    vals, t_tot = [], 0.0
    for _ in range(args.steps):
        sps, dt = cpu_iteration_sps(n_envs, T, cores)
        vals.append(sps); t_tot += dt
        if t_tot > 120:
            break
    v = float(np.mean(vals))
    sample = (f"{len(vals)} iteration(s) of {n_envs} envs x {T} steps + GAE + {NMB * EPOCHS} minibatch updates ({t_tot:.0f} s): oracle C physics on "
              f"{cores} threads + single-thread torch-CPU agent/GAE/update, as the reference runs them")
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": "env-steps/s", "n_gpus": args.gpus, "steps": len(vals),
        "warmup": args.warmup, "ms_per_step": 1e3 * t_tot / len(vals), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": workload(n_envs, T), "num_envs_per_gpu": n_envs, "num_steps": T},
        "cpu_baseline": {"value": v, "unit": "env-steps/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": v, "unit": "env-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }))


# ------------------------------------------------------------------ our arm
def exchange_check(trainer, dist, torch, rank, world, local):
    """Multi-GPU correctness signal inside the bench run (the 2-GPU pytest is skipped on 1-GPU boxes): one minibatch step on
    rank-specific synthetic rows through the fused peer-memory all-reduce and through ncclAllReduce(avg); the averaged
    gradient must be bit-identical on all ranks and within 1e-6 (relative to the largest entry) of NCCL's."""
    N, T, nmb = 64, 8, 4
    B = N * T

    def one(peer):
        os.environ["PB_PEER_AR"] = "1" if peer else "0"
        t = trainer.Trainer(env_id=ENV_ID, num_envs=N, num_steps=T, num_minibatches=nmb, update_epochs=1, device=local, rank=rank,
                            world_size=world, seed=3)
        rng = np.random.default_rng(100 + rank)
        for k, v in dict(obs=rng.normal(size=(B, OBS)), actions=rng.normal(size=(B, ACT)) * 0.5, logprobs=rng.normal(size=B) * 0.1 - 0.5 * ACT,
                         advantages=rng.normal(size=B), returns=rng.normal(size=B), values=rng.normal(size=B) * 0.5).items():
            t.write(k, v.astype(np.float32))
        uid = [trainer.comm_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(uid, src=0)
        t.comm_init(uid[0], rank, world)
        t.broadcast_params()
        mode = capi_lib().pb_trainer_exchange_mode(t._h)
        t.update_minibatch(np.random.default_rng(7).permutation(B)[: B // nmb].astype(np.int32), 3e-4)
        g = torch.from_numpy(t.read("grad")).cuda()
        t.close()
        return g, mode

    saved = os.environ.get("PB_PEER_AR")
    g_peer, mode_peer = one(True)
    g_nccl, _ = one(False)
    if saved is None:
        os.environ.pop("PB_PEER_AR", None)
    else:
        os.environ["PB_PEER_AR"] = saved
    allg = [torch.empty_like(g_peer) for _ in range(world)]
    dist.all_gather(allg, g_peer)
    identical = all(torch.equal(allg[0], x) for x in allg)
    rel = float((g_peer - g_nccl).abs().max() / g_nccl.abs().max())
    ok = identical and rel <= 1e-6 and bool(torch.isfinite(g_peer).all())
    return {"status": "ok" if ok else "FAILED", "replicas_bit_identical": identical, "rel_vs_nccl": rel,
            "peer_path": {0: "none", 1: "ncclAllReduce (peer mapping unavailable)", 2: "fused peer-memory all-reduce"}[mode_peer]}


def capi_lib():
    from ppo_cpp_b200 import capi
    return capi.lib()


def run_ours(args):
    import ctypes as C

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
    N, T = args.num_envs, args.num_steps
    lr = 0.0   # pinned state, see the module docstring
    L = capi.lib()
    xcheck = exchange_check(trainer, dist, torch, rank, world, local) if world > 1 else None

    def make_warm_trainer():
        """Trainer at the pinned walking state: snapshot policy + normaliser, then untimed iterations until the episode /
        contact mix under that policy is stationary (at least 8 rollouts = 256 env steps, mean episode length ~130)."""
        tr = trainer.Trainer(env_id=ENV_ID, num_envs=N, num_steps=T, num_minibatches=NMB, update_epochs=EPOCHS, device=local,
                             seed=1, rank=rank, world_size=world)
        if world > 1:
            uid = [trainer.comm_unique_id() if rank == 0 else None]
            dist.broadcast_object_list(uid, src=0)
            tr.comm_init(uid[0], rank, world)
        tr.reset(1 + rank * N)
        tr.load_snapshot(SNAPSHOT)
        if world > 1:
            tr.broadcast_params()
        for _ in range(max(0, 8 - args.warmup)):
            tr.rollout()
        for _ in range(args.warmup):
            tr.iteration(lr)
        tr.sync()
        return tr

    t = make_warm_trainer()
    stream = torch.cuda.ExternalStream(L.pb_trainer_stream(t._h))
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with ClockSampler(local) as cs:
        e0.record(stream)
        for i in range(args.steps):
            if args.profile_iteration and i == args.steps - 1:   # ncu --profile-from-start off captures exactly this iteration
                t.sync()
                torch.cuda.profiler.start()
            t.iteration(lr)
        e1.record(stream)
        t.sync()
        torch.cuda.synchronize()
        if args.profile_iteration:
            torch.cuda.profiler.stop()
    st = t.stats()
    ms = e0.elapsed_time(e1)
    if world > 1:
        dist.barrier()
        tm = torch.tensor([ms], device="cuda")
        dist.all_reduce(tm, op=dist.ReduceOp.MAX)
        ms = float(tm.item())
    value = world * N * T * args.steps / (ms * 1e-3)

    # e2e: the public call a user makes (pb_trainer_iteration + pb_trainer_stats), wall clock around it; every step
    # uploads the iteration's host inputs (epoch permutations from pinned memory, lr) and reads the stats back.
    t.close()
    if world > 1:
        dist.barrier()
    t = make_warm_trainer()
    stream = torch.cuda.ExternalStream(L.pb_trainer_stream(t._h))
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
    h2d = EPOCHS * N * T * 4 + 4
    d2h = 14 * 4 + 16 + 16

    # dominant kernel (fused env step) timed alone on its launch stream, device pointers, at the pinned state: the rollout's own
    # sampled actions are replayed, so the envs keep walking under the same action distribution
    n_launch = 16
    act_ptr, _ = t.buffer_ptr("actions")
    obs_ptr, _ = t.buffer_ptr("next_obs")
    rew_ptr, _ = t.buffer_ptr("rewards")
    term_ptr, _ = t.buffer_ptr("term")
    trunc_ptr, _ = t.buffer_ptr("trunc")
    eh = L.pb_trainer_env(t._h)
    k0, k1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    k0.record(stream)
    for i in range(n_launch):
        L.pb_env_step(eh, C.c_void_p(act_ptr + (i % T) * N * ACT * 4), C.c_void_p(obs_ptr), C.c_void_p(rew_ptr),
                      C.c_void_p(term_ptr), C.c_void_p(trunc_ptr), None, None, None)
    k1.record(stream)
    t.sync()
    kms = k0.elapsed_time(k1) / n_launch
    # the literal gymcpp seam with HOST buffers (pb_env_step_host: actions H2D, step, obs/reward/flags/infos D2H), wall clock
    host_step = None
    if rank == 0:
        acts = t.read("actions").reshape(T, N, ACT)
        ho = [np.zeros((N, OBS), np.float32)] + [np.zeros(N, np.float32) for _ in range(3)]
        hi, hr, hl = np.zeros(N, np.int32), np.zeros(N, np.float32), np.zeros(N, np.int32)
        fp, ip = capi._fp, capi._ip
        w0h = time.perf_counter()
        nh = 8
        for i in range(nh):
            a = np.ascontiguousarray(acts[i % T])
            capi._check(L.pb_env_step_host(eh, a.ctypes.data_as(fp), ho[0].ctypes.data_as(fp), ho[1].ctypes.data_as(fp), ho[2].ctypes.data_as(fp),
                                           ho[3].ctypes.data_as(fp), hi.ctypes.data_as(ip), hr.ctypes.data_as(fp), hl.ctypes.data_as(ip)))
        dth = (time.perf_counter() - w0h) / nh
        host_step = {"ms_per_step": dth * 1e3, "env_steps_per_s": N / dth, "h2d_bytes": N * ACT * 4, "d2h_bytes": N * (OBS + 6) * 4,
                     "call": "pb_env_step_host (gymcpp ParVectorEnv::step seam, host buffers, wall clock)"}
    hbm_peak, peak_kind = peaks()
    achieved = ALG_BYTES_PER_ENV_STEP * N / (kms * 1e-3) / 1e9
    prof = kernel_profile()
    scale = N / float(prof.get("num_envs", 8192))
    traffic = prof.get("dram_bytes_per_launch")
    traffic = traffic * scale if traffic else None
    # the roof that actually bounds the rigid-body kernel: FP32 FMA throughput, measured here by a
    # register-resident FFMA kernel (pb_fp32_fma_peak) because MEASURED_PEAKS.json holds only HBM and bf16 tensor peaks.
    # Two numerators: the oracle-counted ALGORITHMIC flops per env step at this state (tools/count_oracle_flops.py) and the FP32
    # operations the kernel executes (ncu, same state): their ratio is the redundant / recomputed work.
    fp32 = None
    if rank == 0:
        pk = C.c_float()
        if L.pb_fp32_fma_peak(local, C.c_float(20.0), C.byref(pk)) == 0 and pk.value > 0:
            alg = prof.get("oracle_flop_per_env_step")
            exe = prof.get("executed_fp32_flop_per_launch")
            exe = exe * scale if exe else None
            fl = alg * N if alg else None
            fp32 = {"bound": "fp32_fma", "kernel": "env_step_kernel<CfgHumanoid>", "peak": pk.value, "peak_kind": "measured (pb_fp32_fma_peak)",
                    "unit": "TFLOP/s", "alg_flop_per_env_step": alg, "alg_flop_per_launch": fl,
                    "achieved": fl / (kms * 1e-3) / 1e12 if fl else None, "frac": fl / (kms * 1e-3) / 1e12 / pk.value if fl else None,
                    "executed_flop_per_launch": exe, "executed_frac": exe / (kms * 1e-3) / 1e12 / pk.value if exe else None,
                    "flop_source": prof.get("source")}
    # ... and the resource ncu shows it is closest to: warp-instruction issue (4 schedulers per SM, one instruction per cycle each)
    issue = None
    wi = prof.get("executed_warp_instructions")
    if rank == 0 and wi:
        props = torch.cuda.get_device_properties(local)
        mhz = cs.summary().get("sm_mhz") or 1965
        peak_issue = 4.0 * props.multi_processor_count * mhz * 1e6 / 1e12
        ach_issue = wi * scale / (kms * 1e-3) / 1e12
        issue = {"bound": "warp_issue", "kernel": "env_step_kernel<CfgHumanoid>", "achieved": ach_issue, "peak": peak_issue, "unit": "T warp-instructions/s",
                 "frac": ach_issue / peak_issue, "warp_instructions_per_launch": wi * scale,
                 "note": "IPC / 4: the rest is dependent-issue latency (one warp per env, 16 warps per SM) and the 23 % lockstep wait"}
    xmode = L.pb_trainer_exchange_mode(t._h)
    exchange = {0: "none (single rank)", 1: "ncclAllReduce(avg)", 2: "fused peer-memory all-reduce kernel (CUDA IPC over NVLink)"}[xmode]

    if rank == 0:
        cpu = None
        if world == 1 and not args.no_cpu_baseline:
            cores = os.cpu_count() or 1
            n_cpu = N if cores >= 8 else 2048  # the full per-GPU workload (~12 s on 16 threads), else a quarter of it
            sps, dt = cpu_iteration_sps(n_cpu, T, cores)
            cpu = {"value": sps, "unit": "env-steps/s", "cores": cores, "kind": "port",
                   "sample": f"{ENV_ID} at the pinned walking state, {n_cpu} envs x {T} steps + GAE + {NMB * EPOCHS} minibatch updates on {cores} host "
                             f"threads ({dt:.1f} s): oracle C physics + torch-CPU agent"}
        # launches per iteration: T x (policy, queue order, env step) + bootstrap policy + GAE + 2 weight transposes (+ 2 TF32
        # weight splits on the tcgen05 path) + per epoch (gather, advantage statistics) + per minibatch: forward/backward, weight
        # gradients, reduce, Adam (+ tcgen05 first-layer forward, + tcgen05 first-layer weight gradient, + one exchange kernel
        # of ours when N > 1; NCCL's own kernel is not counted)
        tc_path = os.environ.get("PB_UPDATE_MMA", "2") == "2"
        fused_x = xmode == 2 and os.environ.get("PB_PEER_FUSED", "0") != "0"
        ll_x = xmode == 2 and not fused_x and os.environ.get("PB_PEER_LL", "1") != "0"   # LL exchange: inside the reduction kernel
        extra_x = 1 if world > 1 and not fused_x and not ll_x else 0   # peer_avg_kernel, or sqnorm_partials_kernel after ncclAllReduce
        per_mb = 4 + (1 if tc_path else 0) + (1 if tc_path and os.environ.get("PB_DW1_TC", "1") != "0" else 0) + extra_x
        pol_tc = tc_path and os.environ.get("PB_POLICY_TC", "1") != "0" and N >= 128   # tcgen05 first layer in the rollout forward
        launches_per_iter = 3 * T + 2 + 2 + (2 if tc_path else 0) + ((T + 1) if pol_tc else 0) + EPOCHS * 2 + EPOCHS * NMB * per_mb
        print(json.dumps({
            "metric": METRIC, "value": value, "unit": "env-steps/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": workload(N, T), "num_envs_per_gpu": N, "num_steps": T,
                       "state": "tests/golden/walking_Humanoid-v4.npz (policy after 50 iterations), lr = 0 in every leg; "
                                f"{max(8, args.warmup)} untimed rollouts before the timed region",
                       "l2": "working set per iteration (rollout obs 394 MB + per-env normalisers 25 MB) exceeds the 126 MB L2",
                       "parallelism": f"dp{world}"},
            "phase_ms": {"rollout": st["rollout_ms"], "gae": st["gae_ms"], "update": st["update_ms"]},
            "update_ms_per_iteration": st["update_ms"],
            "rollout_only_sps": world * N * T / (st["rollout_ms"] * 1e-3),
            "roofline": {"bound": "hbm", "kernel": "env_step_kernel<CfgHumanoid>", "achieved": achieved, "peak": hbm_peak,
                         "peak_kind": peak_kind, "unit": "GB/s", "frac": achieved / hbm_peak, "traffic": traffic,
                         "alg_bytes_per_launch": ALG_BYTES_PER_ENV_STEP * N,
                         "kernel_ms": kms, "kernel_ms_ncu": prof.get("kernel_ms_ncu"), "env_steps_per_s_kernel_only": N / (kms * 1e-3),
                         "note": "instruction-issue/latency-bound rigid-body kernel; HBM is not the binding roof (SURVEY.md 8d, DESIGN.md 4.1): "
                                 "see roofline_fp32"},
            "roofline_fp32": fp32, "roofline_issue": issue,
            "gradient_exchange": exchange, "exchange_check": xcheck,
            "cpu_baseline": cpu,
            "e2e": {"value": e2e, "unit": "env-steps/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "call": "pb_trainer_iteration + pb_trainer_stats (wall clock, host permutations uploaded from pinned memory)",
                    "iter_ms": e2e_iter_ms, "phase_ms_last": {"rollout": s2["rollout_ms"], "gae": s2["gae_ms"], "update": s2["update_ms"]}},
            "e2e_env_step_host": host_step,
            "gpu_launches": launches_per_iter * args.steps,
            "clocks": cs.summary(),
            "sim_counters": {"forward_evals": st["forward_evals"], "overflow_envs": st["overflow_envs"], "episodes_last_iter": s2["episodes"]},
        }))
    t.close()
    if world > 1:
        dist.destroy_process_group()


# ------------------------------------------------------------------ C5: ac_ppo_continuous_action workload (BASELINE.json configs[4])
AC_ENV, AC_OBS, AC_ACT, AC_T, AC_NMB, AC_EPOCHS = "HalfCheetah-v5", 17, 6, 128, 4, 4
AC_METRIC = "env-steps/sec (SPS) HalfCheetah-v5 ac_ppo iteration (collection+GAE+update)"
# algorithmic FLOPs of the update per sample and epoch = 6 x MACs of the agent (859 k for HalfCheetah ac_ppo);
# elementwise work (LayerNorm, softplus, lgamma / digamma) excluded
# algorithmic bytes of one HalfCheetah env step (SURVEY.md 8d accounting: qpos / qvel / warm start read + written, obs + reward + flags written,
# action read, episode statistics): (9 + 9 + 9) * 2 * 4 + 17 * 4 + 6 * 4 + 6 * 4
AC_ENV_BYTES_PER_STEP = (9 + 9 + 9) * 2 * 4 + 17 * 4 + 6 * 4 + 6 * 4
AC_UPDATE_FLOP_PER_SAMPLE_EPOCH = 6 * (2 * (AC_OBS * 256 + 256 * 256) + 256 + 2 * 256 * AC_ACT)


def ac_workload(num_envs):
    return (f"{AC_ENV} ac_ppo at the seeded initial policy (lr = 0, Beta policy, 256-256 LayerNorm MLPs): {num_envs} envs/GPU x {AC_T} steps, "
            f"{AC_NMB} minibatches x {AC_EPOCHS} epochs")


def cpu_ac_iteration_sps(num_envs, nthreads):
    """One ac_ppo iteration on the CPU: oracle C physics on `nthreads` threads, torch-CPU agent (batched forward - the reference
    runs it per env, :655), GAE and the 4 x 4 minibatch update of src/ac_ppo_continuous_action.cpp:800-890."""
    import torch

    from oracle import oracle_nn, oracle_nn_ac, pbo
    torch.set_num_threads(nthreads)
    env = pbo.VecEnv(AC_ENV, num_envs, chain=0, nthreads=nthreads)
    agent = oracle_nn_ac.AcAgent(AC_OBS, AC_ACT, 1.0, -1.0)
    opt = torch.optim.Adam(agent.parameters(), lr=0.0, eps=1e-5)
    next_obs = torch.from_numpy(env.reset(1).copy())
    next_done = torch.zeros(num_envs)
    T, N = AC_T, num_envs
    obs = torch.zeros(T, N, AC_OBS); actions = torch.zeros(T, N, AC_ACT); logprobs = torch.zeros(T, N)
    rewards = torch.zeros(T, N); dones = torch.zeros(T, N); values = torch.zeros(T, N)
    t0 = time.perf_counter()
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
    B = T * N
    b = [x.reshape(B, -1) if x.dim() == 3 else x.reshape(B) for x in (obs, actions, logprobs, adv, ret, values)]
    mbs = B // AC_NMB
    for _e in range(AC_EPOCHS):
        perm = torch.randperm(B)
        for st in range(0, B, mbs):
            oracle_nn_ac.ac_minibatch_step(agent, opt, *b, perm[st:st + mbs])
    dt = time.perf_counter() - t0
    return T * N / dt, dt


def run_reference_ac(args):
    if int(os.environ.get("RANK", "0")) != 0:
        return
    cores = os.cpu_count() or 1
    n_envs = min(args.num_envs, 1024)     # bounded sample: one eighth of the per-GPU workload, same T / minibatch / epoch counts
    vals, t_tot = [], 0.0
    for _ in range(max(1, min(args.steps, 4))):
        sps, dt = cpu_ac_iteration_sps(n_envs, cores)
        vals.append(sps); t_tot += dt
        if t_tot > 100:
            break
    v = float(np.mean(vals))
    sample = (f"{len(vals)} iteration(s) of {n_envs} envs x {AC_T} steps + GAE + {AC_NMB * AC_EPOCHS} minibatch updates ({t_tot:.0f} s): oracle C physics "
              f"and torch-CPU agent on {cores} threads")
    print(json.dumps({
        "impl": "reference", "metric": AC_METRIC, "value": v, "unit": "env-steps/s", "n_gpus": args.gpus, "steps": len(vals), "warmup": args.warmup,
        "ms_per_step": 1e3 * t_tot / len(vals), "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": ac_workload(args.num_envs), "num_envs_per_gpu": args.num_envs, "num_steps": AC_T, "cpu_sample_envs": n_envs},
        "cpu_baseline": {"value": v, "unit": "env-steps/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": v, "unit": "env-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}))


def run_ours_ac(args):
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
    N, T = args.num_envs, AC_T
    L = capi.lib()

    def make():
        tr = trainer.AcTrainer(env_id=AC_ENV, num_envs=N, num_steps=T, num_minibatches=AC_NMB, update_epochs=AC_EPOCHS, device=local, seed=1,
                               rank=rank, world_size=world)
        if world > 1:
            uid = [trainer.comm_unique_id() if rank == 0 else None]
            dist.broadcast_object_list(uid, src=0)
            tr.comm_init(uid[0], rank, world)
            tr.broadcast_params()
        tr.reset(1 + rank * N)
        for _ in range(args.warmup):
            tr.iteration(0.0)
        tr.sync()
        return tr

    t = make()
    stream = torch.cuda.ExternalStream(L.pb_ac_trainer_stream(t._h))
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with ClockSampler(local) as cs:
        e0.record(stream)
        for i in range(args.steps):
            if args.profile_iteration and i == args.steps - 1:
                t.sync()
                torch.cuda.profiler.start()
            t.iteration(0.0)
        e1.record(stream)
        t.sync()
        torch.cuda.synchronize()
        if args.profile_iteration:
            torch.cuda.profiler.stop()
    st = t.stats()
    ms = e0.elapsed_time(e1)
    if world > 1:
        dist.barrier()
        tm = torch.tensor([ms], device="cuda")
        dist.all_reduce(tm, op=dist.ReduceOp.MAX)
        ms = float(tm.item())
    value = world * N * T * args.steps / (ms * 1e-3)
    # e2e: pb_ac_trainer_iteration + pb_ac_trainer_stats per iteration, wall clock (epoch permutations uploaded from pinned memory)
    w0 = time.perf_counter()
    for _ in range(args.steps):
        t.iteration(0.0)
        s2 = t.stats()
    w1 = time.perf_counter() - w0
    if world > 1:
        tw = torch.tensor([w1], device="cuda")
        dist.all_reduce(tw, op=dist.ReduceOp.MAX)
        w1 = float(tw.item())
    e2e = world * N * T * args.steps / w1
    # replicas must stay bit-identical under the data-parallel update (gradient average + cross-rank advantage moments, :830-885)
    replicas = None
    if world > 1:
        p = torch.from_numpy(t.get_params()).cuda()
        allp = [torch.empty_like(p) for _ in range(world)]
        dist.all_gather(allp, p)
        replicas = all(torch.equal(allp[0], x) for x in allp)
    # dominant kernel of the iteration (the HalfCheetah env step) timed alone on its launch stream; actions replayed from the rollout
    import ctypes as C
    eh = L.pb_ac_trainer_env(t._h)
    acts = torch.from_numpy(t.read("actions").reshape(T, N, AC_ACT)).cuda()
    o_d = torch.empty(N, AC_OBS, device="cuda")
    r_d, te_d, tr_d = (torch.empty(N, device="cuda") for _ in range(3))
    torch.cuda.synchronize()
    k0, k1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n_launch = 32
    k0.record(stream)
    for i in range(n_launch):
        L.pb_env_step(eh, C.c_void_p(acts[i % T].data_ptr()), C.c_void_p(o_d.data_ptr()), C.c_void_p(r_d.data_ptr()), C.c_void_p(te_d.data_ptr()),
                      C.c_void_p(tr_d.data_ptr()), None, None, None)
    k1.record(stream)
    t.sync()
    kms = k0.elapsed_time(k1) / n_launch
    if rank == 0:
        hbm_peak, peak_kind = peaks()
        peaks_json = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else {}
        tensor_peak = peaks_json.get("bf16_tflops_sustained", 1378.5)
        flop = AC_UPDATE_FLOP_PER_SAMPLE_EPOCH * N * T * AC_EPOCHS
        ach = flop / (st["update_ms"] * 1e-3) / 1e12
        env_bytes = AC_ENV_BYTES_PER_STEP * N
        hc_prof = {}
        hc_path = os.path.join(ROOT, "profiles", "r02_envstep_halfcheetah.json")
        if os.path.exists(hc_path):
            hc_prof = json.load(open(hc_path))
            if hc_prof.get("num_envs") != N:
                hc_prof = {}
        # update: every 256-wide activation matrix of a minibatch is rows x 1 KB.  Algorithmic minimum per net: store x_hat1, x_hat2 in the
        # forward pass, read them in the backward pass (4 passes); the implementation makes 12.
        mb_rows = N * T // AC_NMB
        upd_alg = 2 * 4 * mb_rows * 1024 * AC_NMB * AC_EPOCHS
        upd_impl = 2 * 12 * mb_rows * 1024 * AC_NMB * AC_EPOCHS
        upd_ach = upd_alg / (st["update_ms"] * 1e-3) / 1e9
        cpu = None
        if world == 1 and not args.no_cpu_baseline:
            cores = os.cpu_count() or 1
            sps, dt = cpu_ac_iteration_sps(min(N, 1024), cores)
            cpu = {"value": sps, "unit": "env-steps/s", "cores": cores, "kind": "port",
                   "sample": f"{min(N, 1024)} envs x {T} steps + GAE + {AC_NMB * AC_EPOCHS} minibatch updates on {cores} threads ({dt:.1f} s): oracle C physics + torch-CPU agent"}
        fused = os.environ.get("PB_AC_FUSED", "1") != "0" and os.environ.get("PB_AC_TC", "1") != "0"
        launches = (T * 12 + 6 + AC_NMB * AC_EPOCHS * (25 + (2 if world > 1 else 0)) + AC_EPOCHS * 2) if fused else None
        print(json.dumps({
            "metric": AC_METRIC, "value": value, "unit": "env-steps/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": ac_workload(N), "num_envs_per_gpu": N, "num_steps": T, "parallelism": f"dp{world}",
                       "l2": "per-iteration working set (rollout obs 71 MB, activations of a 262144-row minibatch 1.3 GB) exceeds the 126 MB L2"},
            "phase_ms": {"rollout": st["rollout_ms"], "gae": st["gae_ms"], "update": st["update_ms"]},
            "update_ms_per_iteration": st["update_ms"],
            "roofline": {"bound": "hbm", "kernel": "env_step_kernel<CfgHalfCheetah> (dominant kernel: 128 launches per iteration)",
                         "achieved": env_bytes / (kms * 1e-3) / 1e9, "peak": hbm_peak, "peak_kind": peak_kind, "unit": "GB/s",
                         "frac": env_bytes / (kms * 1e-3) / 1e9 / hbm_peak, "traffic": hc_prof.get("dram_bytes_per_launch"),
                         "alg_bytes_per_launch": env_bytes, "kernel_ms": kms, "kernel_ms_ncu": hc_prof.get("kernel_ms_ncu"),
                         "env_steps_per_s_kernel_only": N / (kms * 1e-3),
                         "note": "instruction-issue/latency-bound rigid-body kernel, as for the ppo workload; HBM is not its binding roof"},
            "roofline_update": {"bound": "hbm", "kernel": "ac_ppo update (16 minibatches: tcgen05 GEMMs with fused LayerNorm + row kernels)",
                                "achieved": upd_ach, "peak": hbm_peak, "unit": "GB/s", "frac": upd_ach / hbm_peak,
                                "alg_bytes_per_update": upd_alg, "implementation_bytes_per_update": upd_impl,
                                "implementation_frac": upd_impl / (st["update_ms"] * 1e-3) / 1e9 / hbm_peak,
                                "tensor_tflops_3xtf32_counted_once": ach, "tensor_peak_bf16_tflops": tensor_peak,
                                "alg_flop_per_iteration": flop},
            "replicas_bit_identical": replicas, "cpu_baseline": cpu,
            "e2e": {"value": e2e, "unit": "env-steps/s", "h2d_bytes_per_step": AC_EPOCHS * N * T * 4 + 4, "d2h_bytes_per_step": 88,
                    "call": "pb_ac_trainer_iteration + pb_ac_trainer_stats (wall clock)",
                    "phase_ms_last": {"rollout": s2["rollout_ms"], "gae": s2["gae_ms"], "update": s2["update_ms"]}},
            "gpu_launches": launches * args.steps if launches else None, "clocks": cs.summary()}))
    t.close()
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--num-envs", type=int, default=8192, help="envs per GPU")
    ap.add_argument("--num-steps", type=int, default=32)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--profile-iteration", action="store_true",
                    help="cudaProfilerStart/Stop around the last timed iteration (for ncu --profile-from-start off; not a bench run)")
    ap.add_argument("--workload", default="ppo", choices=["ppo", "ac"],
                    help="ppo: Humanoid-v4 ppo_continuous_action (the headline); ac: HalfCheetah-v5 ac_ppo_continuous_action (BASELINE config 4)")
    args = ap.parse_args()
    if args.workload == "ac":
        (run_reference_ac if args.impl == "reference" else run_ours_ac)(args)
    elif args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
