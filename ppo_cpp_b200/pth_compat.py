"""`.pth` compatibility for the ppo_continuous_action agent.

The reference writes `model_latest_%09d.pth` / `model_final.pth` with `torch::save(agent, path)`
(src/ppo_continuous_action.cpp:173-180,546-563,587): a TorchScript zip archive whose parameters carry the names that
`AgentImpl` registers (:120-137): `critic.{0,2,4}.{weight,bias}`, `actor_mean.{0,2,4}.{weight,bias}`, `actor_logstd`.
Such a file is read by `torch.jit.load` in Python and by `torch::load(agent, path)` in C++, and a file written by
`torch.jit.save` of a scripted module with the same parameter names loads into the reference's `Agent` the same way.

This module converts between that archive and the engine's own checkpoint (`pb_trainer_save`: 40-byte header, then
params | adam_m | adam_v | normaliser statistics as flat fp32 — csrc/pb_trainer.cu), so weights trained here open in
reference-side tools and weights trained by the reference can be loaded with `Trainer.set_params`.  The flat parameter
order is AgentImpl's registration order (trainer.split_params).  Host-side tool: needs torch, never touches the GPU.

    python -m ppo_cpp_b200.pth_compat export models/run/model_final.bin 376 17 model_final.pth
    python -m ppo_cpp_b200.pth_compat import model_final.pth 376 17 params.f32
"""
from __future__ import annotations

import struct
import sys

import numpy as np

HEADER = struct.Struct("<8siiiif3i")   # CkptHeader in csrc/pb_trainer.cu: magic, version, P, N, O, step, reserved[3]

NAMES = ["critic.0.weight", "critic.0.bias", "critic.2.weight", "critic.2.bias", "critic.4.weight", "critic.4.bias",
         "actor_mean.0.weight", "actor_mean.0.bias", "actor_mean.2.weight", "actor_mean.2.bias", "actor_mean.4.weight",
         "actor_mean.4.bias", "actor_logstd"]


def shapes(obs, act):
    return [(64, obs), (64,), (64, 64), (64,), (1, 64), (1,), (64, obs), (64,), (64, 64), (64,), (act, 64), (act,), (1, act)]


def num_params(obs, act):
    return int(sum(np.prod(s) for s in shapes(obs, act)))


def read_checkpoint(path):
    """Engine checkpoint -> dict(step, P, N, O, params, adam_m, adam_v)."""
    with open(path, "rb") as f:
        magic, version, P, N, O, step, *_ = HEADER.unpack(f.read(HEADER.size))
        if not magic.startswith(b"PB200CK") or version != 1:
            raise ValueError(f"{path}: not a ppo_b200 checkpoint")
        flat = np.fromfile(f, dtype=np.float32, count=3 * P)
    if flat.size != 3 * P:
        raise ValueError(f"{path}: truncated")
    return dict(step=step, P=P, N=N, O=O, params=flat[:P].copy(), adam_m=flat[P:2 * P].copy(), adam_v=flat[2 * P:].copy())


def _agent_module(obs, act):
    import torch
    from torch import nn

    class Agent(nn.Module):   # parameter names = AgentImpl's (src/ppo_continuous_action.cpp:120-137)
        def __init__(self, o: int, a: int):
            super().__init__()
            self.critic = nn.Sequential(nn.Linear(o, 64), nn.Tanh(), nn.Linear(64, 64), nn.Tanh(), nn.Linear(64, 1))
            self.actor_mean = nn.Sequential(nn.Linear(o, 64), nn.Tanh(), nn.Linear(64, 64), nn.Tanh(), nn.Linear(64, a))
            self.actor_logstd = nn.Parameter(torch.zeros(1, a))

    return Agent(obs, act)


def state_dict_from_flat(flat, obs, act):
    import torch
    flat = np.asarray(flat, dtype=np.float32)
    if flat.size != num_params(obs, act):
        raise ValueError(f"expected {num_params(obs, act)} parameters for obs={obs} act={act}, got {flat.size}")
    out, o = {}, 0
    for name, shp in zip(NAMES, shapes(obs, act)):
        n = int(np.prod(shp))
        out[name] = torch.from_numpy(flat[o:o + n].reshape(shp).copy())
        o += n
    return out


def flat_from_named(named):
    """{name: tensor} (any superset of NAMES) -> flat fp32 vector in registration order."""
    missing = [n for n in NAMES if n not in named]
    if missing:
        raise KeyError(f"missing parameters: {missing}")
    return np.concatenate([named[n].detach().cpu().numpy().astype(np.float32).ravel() for n in NAMES])


def save_pth(path, flat, obs, act):
    """Write a TorchScript archive with AgentImpl's parameter names (what `torch::save(agent, path)` produces)."""
    import torch
    agent = _agent_module(obs, act)
    agent.load_state_dict(state_dict_from_flat(flat, obs, act))
    torch.jit.save(torch.jit.script(agent), str(path))


def load_pth(path):
    """Archive written by the reference (`torch::save(agent)`), by save_pth, or a plain state_dict -> flat fp32 vector."""
    import torch
    try:
        mod = torch.jit.load(str(path), map_location="cpu")
        named = dict(mod.named_parameters())
        named.update(dict(mod.named_buffers()))
    except RuntimeError:
        named = torch.load(str(path), map_location="cpu", weights_only=True)
    return flat_from_named(named)


def main(argv):
    if len(argv) != 5 or argv[0] not in ("export", "import"):
        print(__doc__)
        return 2
    obs, act = int(argv[2]), int(argv[3])
    if argv[0] == "export":
        ck = read_checkpoint(argv[1])
        if ck["O"] != obs:
            raise SystemExit(f"checkpoint has obs dim {ck['O']}, not {obs}")
        save_pth(argv[4], ck["params"], obs, act)
    else:
        load_pth(argv[1]).tofile(argv[4])
    return 0


if __name__ == "__main__":
    sys.exit(main(sys.argv[1:]))
