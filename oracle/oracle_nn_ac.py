"""CPU ORACLE (test infrastructure, NOT the product path) for the ac_ppo_continuous_action agent and update.

torch-CPU restatement of
  * AgentImpl of src/ac_ppo_continuous_action.cpp:150-266 (fixed obs mean/std, Linear-LayerNorm-ReLU x2 critic and actor
    encoder, 256 wide; alpha, beta = softplus(Linear(256, A)) + 1; default torch init, no orthogonal call),
  * Beta via Dirichlet of include/rl_utils.h:48-110 (log_prob with xlogy, entropy with digamma, mean),
  * scale/unscale of actions with the clamp [1e-7, 1 + 1e-7] (:194-210),
  * the minibatch step of :816-888 with the advantage normalisation over world*mb samples (:830-849).
Pinned by construction: torch in this container implements the same operators the reference calls in libtorch.
"""
import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F

HID = 256


class AcAgent(nn.Module):
    def __init__(self, obs, act, act_high, act_low, mean=None, std=None):
        super().__init__()
        self.high, self.low = float(act_high), float(act_low)
        self.register_buffer("mean_", torch.zeros(1, obs) if mean is None else torch.as_tensor(mean, dtype=torch.float32).view(1, -1))
        self.register_buffer("std_", torch.ones(1, obs) if std is None else torch.as_tensor(std, dtype=torch.float32).view(1, -1))
        self.critic = nn.Sequential(nn.Linear(obs, HID), nn.LayerNorm(HID), nn.ReLU(), nn.Linear(HID, HID), nn.LayerNorm(HID), nn.ReLU(),
                                    nn.Linear(HID, 1))
        self.actor_mean = nn.Sequential(nn.Linear(obs, HID), nn.LayerNorm(HID), nn.ReLU(), nn.Linear(HID, HID), nn.LayerNorm(HID), nn.ReLU())
        self.dist_alpha = nn.Sequential(nn.Linear(HID, act))
        self.dist_beta = nn.Sequential(nn.Linear(HID, act))

    # registration order of the reference: critic.{0,1,3,4,6}, actor_mean.{0,1,3,4}, dist_alpha.0, dist_beta.0 (weight, bias each)
    def ordered(self):
        return list(self.critic.parameters()) + list(self.actor_mean.parameters()) + list(self.dist_alpha.parameters()) + \
            list(self.dist_beta.parameters())

    def flat(self):
        return np.concatenate([p.detach().numpy().ravel() for p in self.ordered()]).astype(np.float32)

    def flat_grad(self):
        return np.concatenate([p.grad.detach().numpy().ravel() for p in self.ordered()]).astype(np.float32)

    def load_flat(self, flat):
        off = 0
        with torch.no_grad():
            for p in self.ordered():
                n = p.numel()
                p.copy_(torch.from_numpy(np.asarray(flat[off:off + n], dtype=np.float32)).view_as(p))
                off += n

    def get_value(self, x):
        return self.critic((x - self.mean_) / self.std_)

    def scale_action(self, a):
        s = (a - self.low) / (self.high - self.low)
        return torch.clamp(s, 0.0 + 1e-7, 1.0 + 1e-7)

    def unscale_action(self, a):
        return a * (self.high - self.low) + self.low

    def alpha_beta(self, x):
        f = self.actor_mean((x - self.mean_) / self.std_)
        return F.softplus(self.dist_alpha(f)) + 1.0, F.softplus(self.dist_beta(f)) + 1.0

    def get_action_and_value(self, x, action=None, sample_type="sample", generator=None):
        alpha, beta = self.alpha_beta(x)
        conc = torch.stack([alpha, beta], -1)
        if action is None:
            if sample_type == "sample":
                with torch.no_grad():
                    a = torch._sample_dirichlet(conc, generator)[..., 0]
            else:
                a = alpha / (alpha + beta)
        else:
            a = self.scale_action(action)
        ht = torch.stack([a, 1.0 - a], -1)
        logprob = (torch.special.xlogy(conc - 1.0, ht).sum(-1) + torch.lgamma(conc.sum(-1)) - torch.lgamma(conc).sum(-1)).sum(1)
        a0 = conc.sum(-1)
        ent = (torch.lgamma(conc).sum(-1) - torch.lgamma(a0) - (2 - a0) * torch.digamma(a0) - ((conc - 1.0) * torch.digamma(conc)).sum(-1)).sum(1)
        value = self.critic((x - self.mean_) / self.std_)
        return self.unscale_action(a), logprob, ent, value


def ac_minibatch_step(agent, opt, b_obs, b_actions, b_logprobs, b_adv, b_returns, b_values, mb_inds, clip_coef=0.1, norm_adv=True,
                      clip_vloss=True, ent_coef=0.01, vf_coef=0.5, max_grad_norm=0.5, adv_moments=None):
    """One optimiser step.  adv_moments=(mean, std) overrides the local statistics with the cross-rank ones the reference
    computes with two allreduces (:830-849); with one rank they equal the local mean / unbiased std."""
    _, newlogprob, entropy, newvalue = agent.get_action_and_value(b_obs[mb_inds], b_actions[mb_inds])
    logratio = newlogprob - b_logprobs[mb_inds]
    ratio = logratio.exp()
    with torch.no_grad():
        old_approx_kl = (-logratio).mean()
        approx_kl = ((ratio - 1) - logratio).mean()
        clipfrac = ((ratio - 1.0).abs() > clip_coef).float().mean()
    mb_adv = b_adv[mb_inds]
    if norm_adv:
        if adv_moments is None:
            mean, std = mb_adv.mean(), torch.sqrt(((mb_adv - mb_adv.mean()) ** 2).sum() / (mb_adv.numel() - 1))
        else:
            mean, std = adv_moments
        mb_adv = (mb_adv - mean) / (std + 1e-8)
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
    return dict(policy_loss=pg_loss.item(), value_loss=v_loss.item(), entropy=entropy_loss.item(), old_approx_kl=old_approx_kl.item(),
                approx_kl=approx_kl.item(), clipfrac=clipfrac.item(), grad=raw_grad)
