"""CPU ORACLE (test infrastructure) for the agent / loss / optimiser math: a torch-fp32 restatement of
src/ppo_continuous_action.cpp:120-169 (AgentImpl), :447-467 (GAE), :496-541 (clipped PPO loss, clip_grad_norm_,
Adam) and include/rl_utils.h:21-46 (Normal).  torch 2.11 CPU implements the same operators the reference calls
through libtorch 2.4.1, so this part of the oracle is pinned by construction.  Never imported by the product."""
import math

import numpy as np
import torch
from torch import nn


class Normal:
    def __init__(self, mean, std):
        self.mean, self.std = mean, std

    def log_prob(self, value):  # rl_utils.h:35-39
        var = self.std * self.std
        return -(value - self.mean) ** 2 / (2 * var) - torch.log(self.std) - math.log(math.sqrt(2 * math.pi))

    def entropy(self):  # rl_utils.h:41-43
        return 0.5 + 0.5 * math.log(2 * math.pi) + torch.log(self.std)


class Agent(nn.Module):
    def __init__(self, obs, act):
        super().__init__()
        self.critic = nn.Sequential(nn.Linear(obs, 64), nn.Tanh(), nn.Linear(64, 64), nn.Tanh(), nn.Linear(64, 1))
        self.actor_mean = nn.Sequential(nn.Linear(obs, 64), nn.Tanh(), nn.Linear(64, 64), nn.Tanh(), nn.Linear(64, act))
        self.actor_logstd = nn.Parameter(torch.zeros(1, act))
        for seq, last in ((self.critic, 1.0), (self.actor_mean, 0.01)):
            for i, gain in ((0, math.sqrt(2)), (2, math.sqrt(2)), (4, last)):
                nn.init.orthogonal_(seq[i].weight, gain)
                nn.init.constant_(seq[i].bias, 0.0)

    def ordered(self):
        """Flat-buffer order of the engine (AgentLayout): critic.*, actor_mean.*, actor_logstd.  (torch's own
        parameters() yields the directly registered actor_logstd first; order does not affect any arithmetic.)"""
        return list(self.critic.parameters()) + list(self.actor_mean.parameters()) + [self.actor_logstd]

    def flat(self):
        return torch.cat([p.detach().reshape(-1) for p in self.ordered()]).numpy().copy()

    def flat_grad(self):
        return torch.cat([p.grad.reshape(-1) for p in self.ordered()]).numpy().copy()

    def load_flat(self, flat):
        o = 0
        with torch.no_grad():
            for p in self.ordered():
                n = p.numel()
                p.copy_(torch.from_numpy(np.asarray(flat[o:o + n], dtype=np.float32)).reshape(p.shape))
                o += n

    def get_value(self, x):
        return self.critic(x)

    def get_action_and_value(self, x, action=None):
        mean = self.actor_mean(x)
        std = torch.exp(self.actor_logstd.expand_as(mean))
        probs = Normal(mean, std)
        if action is None:
            action = torch.normal(mean, std)
        return action, probs.log_prob(action).sum(1), probs.entropy().sum(1), self.critic(x)


def gae(rewards, values, dones, next_value, next_done, gamma, lam):
    T = rewards.shape[0]
    adv = torch.zeros_like(rewards)
    last = torch.zeros_like(next_value)
    for t in reversed(range(T)):
        if t == T - 1:
            nnt, nv = 1.0 - next_done, next_value
        else:
            nnt, nv = 1.0 - dones[t + 1], values[t + 1]
        delta = rewards[t] + gamma * nv * nnt - values[t]
        adv[t] = last = delta + gamma * lam * nnt * last
    return adv, adv + values


def ppo_minibatch_step(agent, opt, b_obs, b_actions, b_logprobs, b_adv, b_returns, b_values, mb_inds, clip_coef=0.2,
                       norm_adv=True, clip_vloss=True, ent_coef=0.0, vf_coef=0.5, max_grad_norm=0.5):
    _, newlogprob, entropy, newvalue = agent.get_action_and_value(b_obs[mb_inds], b_actions[mb_inds])
    logratio = newlogprob - b_logprobs[mb_inds]
    ratio = logratio.exp()
    with torch.no_grad():
        old_approx_kl = (-logratio).mean()
        approx_kl = ((ratio - 1) - logratio).mean()
        clipfrac = ((ratio - 1.0).abs() > clip_coef).float().mean()
    mb_adv = b_adv[mb_inds]
    if norm_adv:
        mb_adv = (mb_adv - mb_adv.mean()) / (mb_adv.std() + 1e-8)
    pg_loss = torch.max(-mb_adv * ratio, -mb_adv * torch.clamp(ratio, 1 - clip_coef, 1 + clip_coef)).mean()
    newvalue = newvalue.view(-1)
    if clip_vloss:
        v_unclipped = (newvalue - b_returns[mb_inds]) ** 2
        v_clipped = b_values[mb_inds] + torch.clamp(newvalue - b_values[mb_inds], -clip_coef, clip_coef)
        v_loss = 0.5 * torch.max(v_unclipped, (v_clipped - b_returns[mb_inds]) ** 2).mean()
    else:
        v_loss = 0.5 * ((newvalue - b_returns[mb_inds]) ** 2).mean()
    entropy_loss = entropy.mean()
    loss = pg_loss - ent_coef * entropy_loss + v_loss * vf_coef
    opt.zero_grad()
    loss.backward()
    raw_grad = agent.flat_grad()
    nn.utils.clip_grad_norm_(agent.parameters(), max_grad_norm)
    opt.step()
    return dict(policy_loss=pg_loss.item(), value_loss=v_loss.item(), entropy=entropy_loss.item(),
                old_approx_kl=old_approx_kl.item(), approx_kl=approx_kl.item(), clipfrac=clipfrac.item(), grad=raw_grad)
