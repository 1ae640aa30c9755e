"""Device-side timing of the fused env-step kernel: python tools/time_env_step.py [env_id] [num_envs] [warm] [timed]
Random N(0,1) actions (clamped by the env), device buffers, CUDA events on the env's own stream."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ppo_cpp_b200 import capi  # noqa: E402

env_id = sys.argv[1] if len(sys.argv) > 1 else "Humanoid-v4"
n = int(sys.argv[2]) if len(sys.argv) > 2 else 8192
warm = int(sys.argv[3]) if len(sys.argv) > 3 else 20
timed = int(sys.argv[4]) if len(sys.argv) > 4 else 20
v = capi.ParVectorEnv(env_id, n, chain=capi.PB_CHAIN_PPO)
st = torch.cuda.ExternalStream(v.stream())
dev = torch.device("cuda", 0)
g = torch.Generator(device=dev).manual_seed(0)
acts = torch.randn(warm + timed, n, v.act_dim, device=dev, generator=g)
obs = torch.zeros(n, v.obs_dim, device=dev)
rew, term, trunc = (torch.zeros(n, device=dev) for _ in range(3))
torch.cuda.synchronize()
v.reset_device(1, obs.data_ptr())
for i in range(warm):
    v.step_device(acts[i].data_ptr(), obs.data_ptr(), rew.data_ptr(), term.data_ptr(), trunc.data_ptr())
v.sync()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(st)
for i in range(warm, warm + timed):
    v.step_device(acts[i].data_ptr(), obs.data_ptr(), rew.data_ptr(), term.data_ptr(), trunc.data_ptr())
e1.record(st)
v.sync()
ms = e0.elapsed_time(e1) / timed
print(f"{env_id} N={n} variant={os.environ.get('PB_HUM_VARIANT', '0')}: {ms:.3f} ms/step -> {n / ms * 1e3:.3e} env-steps/s | "
      f"obs checksum {float(obs.double().sum()):.6f} rew {float(rew.double().sum()):.6f} counters {v.counters()}")
v.close()
