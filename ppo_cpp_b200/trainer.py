"""Host-side mirror of the reference's training loop objects over the pb_trainer_* C ABI (include/ppo_b200.h).

`Trainer` owns what `main()` owns in src/ppo_continuous_action.cpp:338-364 (agent, Adam, vector env, rollout storage)
and exposes the loop phases under the reference's names: rollout (:387-408), GAE (:447-467), update (:489-542).
`randperm` is torch::randperm on a CPU mt19937 generator, bit-exact (:490).
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import capi
from .capi import _check, lib


class PpoConfig(C.Structure):
    _fields_ = [("env_id", C.c_char_p), ("num_envs", C.c_int), ("num_steps", C.c_int), ("device", C.c_int),
                ("seed", C.c_int), ("gamma", C.c_float), ("gae_lambda", C.c_float), ("num_minibatches", C.c_int),
                ("update_epochs", C.c_int), ("norm_adv", C.c_int), ("clip_coef", C.c_float), ("clip_vloss", C.c_int),
                ("ent_coef", C.c_float), ("vf_coef", C.c_float), ("max_grad_norm", C.c_float), ("adam_eps", C.c_float),
                ("clip_actions", C.c_int), ("rank", C.c_int), ("world_size", C.c_int)]


class PpoStats(C.Structure):
    _fields_ = [("value_loss", C.c_float), ("policy_loss", C.c_float), ("entropy", C.c_float),
                ("old_approx_kl", C.c_float), ("approx_kl", C.c_float), ("clipfrac", C.c_float),
                ("episodic_return_sum", C.c_float), ("episodic_length_sum", C.c_float), ("episodes", C.c_int),
                ("rollout_ms", C.c_float), ("gae_ms", C.c_float), ("update_ms", C.c_float),
                ("forward_evals", C.c_ulonglong), ("overflow_envs", C.c_ulonglong)]


def randperm(seed_or_rng, n):
    """torch.randperm(n) for a CPU generator seeded with `seed` (one-shot), or continue an RNG handle."""
    L = lib()
    own = not isinstance(seed_or_rng, C.c_void_p)
    r = C.c_void_p(L.pb_rng_create(int(seed_or_rng))) if own else seed_or_rng
    out = np.zeros(n, np.int64)
    _check(L.pb_rng_randperm(r, n, out.ctypes.data_as(C.POINTER(C.c_longlong))))
    if own:
        L.pb_rng_destroy(r)
    return out


class Trainer:
    def __init__(self, **kw):
        L = lib()
        self.cfg = PpoConfig()
        L.pb_ppo_config_default(C.byref(self.cfg))
        self._env_id = kw.pop("env_id", "Humanoid-v4").encode()
        self.cfg.env_id = self._env_id
        for k, v in kw.items():
            if not hasattr(self.cfg, k):
                raise TypeError(f"unknown config field {k}")
            setattr(self.cfg, k, v)
        self._h = C.c_void_p()
        _check(L.pb_trainer_create(C.byref(self.cfg), C.byref(self._h)))
        self.num_params = L.pb_trainer_num_params(self._h)

    def close(self):
        if self._h:
            lib().pb_trainer_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def reset(self, seed):
        _check(lib().pb_trainer_reset(self._h, int(seed)))

    def rollout(self):
        _check(lib().pb_trainer_rollout(self._h))

    def gae(self):
        _check(lib().pb_trainer_gae(self._h))

    def update(self, lr):
        _check(lib().pb_trainer_update(self._h, float(lr)))

    def iteration(self, lr):
        _check(lib().pb_trainer_iteration(self._h, float(lr)))

    def update_minibatch(self, idx, lr):
        idx = np.ascontiguousarray(idx, dtype=np.int32)
        _check(lib().pb_trainer_update_minibatch(self._h, idx.ctypes.data_as(capi._ip), len(idx), float(lr)))

    def sync(self):
        _check(lib().pb_trainer_sync(self._h))

    def stats(self):
        s = PpoStats()
        _check(lib().pb_trainer_stats(self._h, C.byref(s)))
        return {k: getattr(s, k) for k, _ in PpoStats._fields_}

    def get_params(self):
        p = np.zeros(self.num_params, np.float32)
        _check(lib().pb_trainer_get_params(self._h, p.ctypes.data_as(capi._fp)))
        return p

    def set_params(self, p):
        p = np.ascontiguousarray(p, dtype=np.float32)
        assert p.size == self.num_params
        _check(lib().pb_trainer_set_params(self._h, p.ctypes.data_as(capi._fp)))

    def buffer_ptr(self, name):
        ptr, n = C.c_void_p(), C.c_longlong()
        _check(lib().pb_trainer_buffer(self._h, name.encode(), C.byref(ptr), C.byref(n)))
        return ptr.value, n.value

    def read(self, name):
        _, n = self.buffer_ptr(name)
        out = np.zeros(n, np.float32)
        _check(lib().pb_trainer_read_buffer(self._h, name.encode(), out.ctypes.data_as(capi._fp), n))
        return out

    def write(self, name, arr):
        a = np.ascontiguousarray(arr, dtype=np.float32).ravel()
        _check(lib().pb_trainer_write_buffer(self._h, name.encode(), a.ctypes.data_as(capi._fp), a.size))

    def evaluate(self, eval_seed, num_eval_runs):
        """Final evaluation (src/ppo_continuous_action.cpp:589-626): episodic returns of the first finished episodes."""
        cap = num_eval_runs + self.cfg.num_envs
        out, n = np.zeros(cap, np.float32), C.c_int()
        _check(lib().pb_trainer_eval(self._h, int(eval_seed), int(num_eval_runs), out.ctypes.data_as(capi._fp), cap, C.byref(n)))
        return out[: n.value]

    def save(self, path):
        _check(lib().pb_trainer_save(self._h, str(path).encode()))

    def load(self, path):
        _check(lib().pb_trainer_load(self._h, str(path).encode()))

    def set_norm_stats(self, ob_mean, ob_var, ob_count, rw4=None):
        """Give every env the same NormalizeObservation / NormalizeReward statistics (a trained policy's normaliser)."""
        m = np.ascontiguousarray(ob_mean, dtype=np.float32)
        v = np.ascontiguousarray(ob_var, dtype=np.float32)
        r = None if rw4 is None else np.ascontiguousarray(rw4, dtype=np.float32)
        _check(lib().pb_env_set_norm_stats(lib().pb_trainer_env(self._h), -1, m.ctypes.data_as(capi._fp), v.ctypes.data_as(capi._fp),
                                           float(ob_count), None if r is None else r.ctypes.data_as(capi._fp)))

    def load_snapshot(self, npz):
        """Policy snapshot written by tools/make_walking_checkpoint.py: parameters + the normaliser they were trained with."""
        d = np.load(npz) if isinstance(npz, (str, bytes)) or hasattr(npz, "__fspath__") else npz
        self.set_params(d["params"])
        self.set_norm_stats(d["ob_mean"], d["ob_var"], float(d["ob_count"]), d["rw4"])

    def comm_init(self, unique_id: bytes, rank, world):
        buf = C.create_string_buffer(unique_id, 128)
        _check(lib().pb_trainer_comm_init(self._h, buf, rank, world))

    def broadcast_params(self):
        _check(lib().pb_trainer_broadcast_params(self._h))


def comm_unique_id() -> bytes:
    buf = C.create_string_buffer(128)
    _check(lib().pb_comm_unique_id(buf))
    return buf.raw


def dd_ppo_batch_indices(rng, num_collected, batch):
    """b_inds of one ac_ppo epoch under DD-PPO pre-emption (src/ac_ppo_continuous_action.cpp:804-810) on an RNG handle."""
    out = np.zeros(batch, np.int64)
    _check(lib().pb_dd_ppo_batch_indices(rng, int(num_collected), int(batch), out.ctypes.data_as(C.POINTER(C.c_longlong))))
    return out


def split_params(flat, obs, act):
    """Flat parameter vector -> dict of torch-layout arrays (registration order of AgentImpl)."""
    o, out = 0, {}
    for name, shape in [("critic.0.weight", (64, obs)), ("critic.0.bias", (64,)), ("critic.2.weight", (64, 64)),
                        ("critic.2.bias", (64,)), ("critic.4.weight", (1, 64)), ("critic.4.bias", (1,)),
                        ("actor_mean.0.weight", (64, obs)), ("actor_mean.0.bias", (64,)),
                        ("actor_mean.2.weight", (64, 64)), ("actor_mean.2.bias", (64,)),
                        ("actor_mean.4.weight", (act, 64)), ("actor_mean.4.bias", (act,)), ("actor_logstd", (1, act))]:
        n = int(np.prod(shape))
        out[name] = np.asarray(flat[o:o + n]).reshape(shape)
        o += n
    assert o == len(flat)
    return out


class AcConfig(C.Structure):
    """pb_ac_config (include/ppo_b200.h); GlobalConfig of src/ac_ppo_continuous_action.cpp:55-100."""
    _fields_ = PpoConfig._fields_ + [("obs_mean", C.POINTER(C.c_float)), ("obs_std", C.POINTER(C.c_float))]


class AcTrainer:
    """The ac_ppo_continuous_action loop objects (agent :150-266, Adam, envs, rollout storage :581-603) over pb_ac_trainer_*."""

    def __init__(self, obs_mean=None, obs_std=None, **kw):
        L = lib()
        self.cfg = AcConfig()
        L.pb_ac_config_default(C.byref(self.cfg))
        self._env_id = kw.pop("env_id", "Ant-v5").encode()
        self.cfg.env_id = self._env_id
        for k, v in kw.items():
            if not hasattr(self.cfg, k):
                raise TypeError(f"unknown config field {k}")
            setattr(self.cfg, k, v)
        self._mean = None if obs_mean is None else np.ascontiguousarray(obs_mean, dtype=np.float32)
        self._std = None if obs_std is None else np.ascontiguousarray(obs_std, dtype=np.float32)
        if self._mean is not None:
            self.cfg.obs_mean = self._mean.ctypes.data_as(capi._fp)
        if self._std is not None:
            self.cfg.obs_std = self._std.ctypes.data_as(capi._fp)
        self._h = C.c_void_p()
        _check(L.pb_ac_trainer_create(C.byref(self.cfg), C.byref(self._h)))
        self.num_params = L.pb_ac_trainer_num_params(self._h)

    def close(self):
        if self._h:
            lib().pb_ac_trainer_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def reset(self, seed):
        _check(lib().pb_ac_trainer_reset(self._h, int(seed)))

    def rollout(self):
        _check(lib().pb_ac_trainer_rollout(self._h))

    def gae(self):
        _check(lib().pb_ac_trainer_gae(self._h))

    def update(self, lr):
        _check(lib().pb_ac_trainer_update(self._h, float(lr)))

    def iteration(self, lr):
        _check(lib().pb_ac_trainer_iteration(self._h, float(lr)))

    def collect(self, steps, preempted=False):
        """Collect `steps` steps per env; preempted: the collector broke out after the last of them (:680-693), which then only
        serves as GAE's bootstrap - steps - 1 are trained on."""
        _check(lib().pb_ac_trainer_collect_begin(self._h))
        for s in range(steps):
            _check(lib().pb_ac_trainer_collect_step(self._h, s))
        _check(lib().pb_ac_trainer_collect_end(self._h, int(preempted)))

    def train_steps(self):
        return lib().pb_ac_trainer_train_steps(self._h)

    def update_minibatch(self, idx, lr):
        idx = np.ascontiguousarray(idx, dtype=np.int32)
        _check(lib().pb_ac_trainer_update_minibatch(self._h, idx.ctypes.data_as(capi._ip), len(idx), float(lr)))

    def forward(self):
        """alpha/beta logits [N, 2A] and values [N] of the current next_obs (forward of :212-265 without the heads' softplus)."""
        _check(lib().pb_ac_trainer_forward(self._h))
        return self.read("zab"), self.read("vbuf")

    def sync(self):
        _check(lib().pb_ac_trainer_sync(self._h))

    def stats(self):
        s = PpoStats()
        _check(lib().pb_ac_trainer_stats(self._h, C.byref(s)))
        return {k: getattr(s, k) for k, _ in PpoStats._fields_}

    def get_params(self):
        p = np.zeros(self.num_params, np.float32)
        _check(lib().pb_ac_trainer_get_params(self._h, p.ctypes.data_as(capi._fp)))
        return p

    def set_params(self, p):
        p = np.ascontiguousarray(p, dtype=np.float32)
        assert p.size == self.num_params
        _check(lib().pb_ac_trainer_set_params(self._h, p.ctypes.data_as(capi._fp)))

    def read(self, name, count=None):
        n = lib().pb_ac_trainer_buffer_count(self._h, name.encode())
        if n < 0:
            raise KeyError(name)
        n = n if count is None else min(n, count)
        out = np.zeros(n, np.float32)
        _check(lib().pb_ac_trainer_read_buffer(self._h, name.encode(), out.ctypes.data_as(capi._fp), n))
        return out

    def write(self, name, arr):
        a = np.ascontiguousarray(arr, dtype=np.float32).ravel()
        _check(lib().pb_ac_trainer_write_buffer(self._h, name.encode(), a.ctypes.data_as(capi._fp), a.size))

    def evaluate(self, eval_seed, num_eval_runs):
        cap = num_eval_runs + self.cfg.num_envs
        out, n = np.zeros(cap, np.float32), C.c_int()
        _check(lib().pb_ac_trainer_eval(self._h, int(eval_seed), int(num_eval_runs), out.ctypes.data_as(capi._fp), cap, C.byref(n)))
        return out[: n.value]

    def comm_init(self, unique_id: bytes, rank, world):
        buf = C.create_string_buffer(unique_id, 128)
        _check(lib().pb_ac_trainer_comm_init(self._h, buf, rank, world))

    def broadcast_params(self):
        _check(lib().pb_ac_trainer_broadcast_params(self._h))
