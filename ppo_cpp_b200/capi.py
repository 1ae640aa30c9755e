"""ctypes binding of libppo_b200.so (include/ppo_b200.h) and a host-side mirror of gymcpp's vector env.

``ParVectorEnv`` below keeps the reference's method names and argument meaning (libs/gymcpp/gym.h:276-366):
``reset(seed) -> obs[N, O]`` and ``step(actions[N, A]) -> (obs, rewards, terminations, truncations, infos)`` with
flags as 0.0/1.0 float32 and ``infos[i]`` either ``None`` or ``(r, l)`` like ``optional<env_info>``.
There is no CPU fallback: importing works without a GPU, every compute call needs the CUDA library and a device.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libppo_b200.so")
_lib = None

PB_CHAIN_STATS_ONLY, PB_CHAIN_PPO = 0, 1
_fp, _ip, _dp = C.POINTER(C.c_float), C.POINTER(C.c_int), C.POINTER(C.c_double)

# name -> (restype, argtypes); every symbol include/ppo_b200.h declares
SIGNATURES = {
    "pb_last_error": (C.c_char_p, []),
    "pb_env_create": (C.c_int, [C.c_char_p, C.c_int, C.c_int, C.c_int, C.c_float, C.c_int, C.POINTER(C.c_void_p)]),
    "pb_env_destroy": (C.c_int, [C.c_void_p]),
    "pb_env_num_envs": (C.c_int, [C.c_void_p]),
    "pb_env_obs_dim": (C.c_int, [C.c_void_p]),
    "pb_env_act_dim": (C.c_int, [C.c_void_p]),
    "pb_env_act_min": (C.c_float, [C.c_void_p]),
    "pb_env_act_max": (C.c_float, [C.c_void_p]),
    "pb_env_stream": (C.c_void_p, [C.c_void_p]),
    "pb_env_reset": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p]),
    "pb_env_step": (C.c_int, [C.c_void_p] + [C.c_void_p] * 8),
    "pb_env_reset_host": (C.c_int, [C.c_void_p, C.c_int, _fp]),
    "pb_env_step_host": (C.c_int, [C.c_void_p, _fp, _fp, _fp, _fp, _fp, _ip, _fp, _ip]),
    "pb_env_sync": (C.c_int, [C.c_void_p]),
    "pb_env_get_state": (C.c_int, [C.c_void_p, C.c_int, _dp, _dp, _dp]),
    "pb_env_set_state": (C.c_int, [C.c_void_p, C.c_int, _dp, _dp, _dp]),
    "pb_env_model_dims": (C.c_int, [C.c_void_p, _ip, _ip, _ip]),
    "pb_env_forward": (C.c_int, [C.c_void_p, C.c_void_p]),
    "pb_env_get_norm_stats": (C.c_int, [C.c_void_p, C.c_int, _fp, _fp, _fp, _fp]),
    "pb_env_set_norm_stats": (C.c_int, [C.c_void_p, C.c_int, _fp, _fp, C.c_float, _fp]),
    "pb_env_solver_counters": (C.c_int, [C.c_void_p, C.POINTER(C.c_ulonglong)]),
    "pb_env_counters": (C.c_int, [C.c_void_p, C.POINTER(C.c_ulonglong)]),
    "pb_ppo_config_default": (None, [C.c_void_p]),
    "pb_trainer_create": (C.c_int, [C.c_void_p, C.POINTER(C.c_void_p)]),
    "pb_trainer_destroy": (C.c_int, [C.c_void_p]),
    "pb_trainer_num_params": (C.c_int, [C.c_void_p]),
    "pb_trainer_get_params": (C.c_int, [C.c_void_p, _fp]),
    "pb_trainer_set_params": (C.c_int, [C.c_void_p, _fp]),
    "pb_trainer_reset": (C.c_int, [C.c_void_p, C.c_int]),
    "pb_trainer_rollout": (C.c_int, [C.c_void_p]),
    "pb_trainer_gae": (C.c_int, [C.c_void_p]),
    "pb_trainer_update": (C.c_int, [C.c_void_p, C.c_float]),
    "pb_trainer_iteration": (C.c_int, [C.c_void_p, C.c_float]),
    "pb_trainer_stats": (C.c_int, [C.c_void_p, C.c_void_p]),
    "pb_trainer_sync": (C.c_int, [C.c_void_p]),
    "pb_trainer_stream": (C.c_void_p, [C.c_void_p]),
    "pb_trainer_env": (C.c_void_p, [C.c_void_p]),
    "pb_trainer_update_minibatch": (C.c_int, [C.c_void_p, _ip, C.c_int, C.c_float]),
    "pb_trainer_buffer": (C.c_int, [C.c_void_p, C.c_char_p, C.POINTER(C.c_void_p), C.POINTER(C.c_longlong)]),
    "pb_trainer_write_buffer": (C.c_int, [C.c_void_p, C.c_char_p, _fp, C.c_longlong]),
    "pb_trainer_read_buffer": (C.c_int, [C.c_void_p, C.c_char_p, _fp, C.c_longlong]),
    "pb_trainer_eval": (C.c_int, [C.c_void_p, C.c_int, C.c_int, _fp, C.c_int, _ip]),
    "pb_trainer_save": (C.c_int, [C.c_void_p, C.c_char_p]),
    "pb_trainer_load": (C.c_int, [C.c_void_p, C.c_char_p]),
    "pb_ac_config_default": (None, [C.c_void_p]),
    "pb_ac_trainer_create": (C.c_int, [C.c_void_p, C.POINTER(C.c_void_p)]),
    "pb_ac_trainer_destroy": (C.c_int, [C.c_void_p]),
    "pb_ac_trainer_num_params": (C.c_int, [C.c_void_p]),
    "pb_ac_trainer_get_params": (C.c_int, [C.c_void_p, _fp]),
    "pb_ac_trainer_set_params": (C.c_int, [C.c_void_p, _fp]),
    "pb_ac_trainer_reset": (C.c_int, [C.c_void_p, C.c_int]),
    "pb_ac_trainer_rollout": (C.c_int, [C.c_void_p]),
    "pb_ac_trainer_gae": (C.c_int, [C.c_void_p]),
    "pb_ac_trainer_update": (C.c_int, [C.c_void_p, C.c_float]),
    "pb_ac_trainer_iteration": (C.c_int, [C.c_void_p, C.c_float]),
    "pb_ac_trainer_stats": (C.c_int, [C.c_void_p, C.c_void_p]),
    "pb_ac_trainer_sync": (C.c_int, [C.c_void_p]),
    "pb_ac_trainer_stream": (C.c_void_p, [C.c_void_p]),
    "pb_ac_trainer_env": (C.c_void_p, [C.c_void_p]),
    "pb_ac_trainer_update_minibatch": (C.c_int, [C.c_void_p, _ip, C.c_int, C.c_float]),
    "pb_ac_trainer_forward": (C.c_int, [C.c_void_p]),
    "pb_ac_trainer_buffer_count": (C.c_longlong, [C.c_void_p, C.c_char_p]),
    "pb_ac_trainer_write_buffer": (C.c_int, [C.c_void_p, C.c_char_p, _fp, C.c_longlong]),
    "pb_ac_trainer_read_buffer": (C.c_int, [C.c_void_p, C.c_char_p, _fp, C.c_longlong]),
    "pb_ac_trainer_eval": (C.c_int, [C.c_void_p, C.c_int, C.c_int, _fp, C.c_int, _ip]),
    "pb_ac_trainer_obs_sums": (C.c_int, [C.c_void_p, C.POINTER(C.c_double), C.POINTER(C.c_longlong)]),
    "pb_ac_trainer_set_obs_stats": (C.c_int, [C.c_void_p, _fp, _fp]),
    "pb_ac_trainer_obs_dim": (C.c_int, [C.c_void_p]),
    "pb_ac_trainer_act_dim": (C.c_int, [C.c_void_p]),
    "pb_ac_trainer_comm_init": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int]),
    "pb_ac_trainer_broadcast_params": (C.c_int, [C.c_void_p]),
    "pb_ac_trainer_collect_begin": (C.c_int, [C.c_void_p]),
    "pb_ac_trainer_collect_step": (C.c_int, [C.c_void_p, C.c_int]),
    "pb_ac_trainer_collect_end": (C.c_int, [C.c_void_p, C.c_int]),
    "pb_ac_trainer_barrier": (C.c_int, [C.c_void_p]),
    "pb_ac_trainer_train_steps": (C.c_int, [C.c_void_p]),
    "pb_dd_ppo_should_preempt": (C.c_int, [C.c_int, C.c_int, C.c_float, C.c_int, C.c_float, C.c_int]),
    "pb_dd_ppo_batch_indices": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.POINTER(C.c_longlong)]),
    "pb_rng_create": (C.c_void_p, [C.c_ulonglong]),
    "pb_rng_destroy": (None, [C.c_void_p]),
    "pb_rng_randperm": (C.c_int, [C.c_void_p, C.c_int, C.POINTER(C.c_longlong)]),
    "pb_comm_unique_id": (C.c_int, [C.c_void_p]),
    "pb_trainer_comm_init": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int]),
    "pb_trainer_broadcast_params": (C.c_int, [C.c_void_p]),
    "pb_trainer_exchange_mode": (C.c_int, [C.c_void_p]),
    "pb_ac_trainer_exchange_mode": (C.c_int, [C.c_void_p]),
    "pb_obs_moments_noalloc": (C.c_int, [C.c_int, C.c_void_p, C.c_void_p, C.c_longlong, C.c_int, C.c_void_p]),
    "pb_obs_moments_noalloc_max_obs": (C.c_int, []),
    "pb_fp32_fma_peak": (C.c_int, [C.c_int, C.c_float, _fp]),
}


def lib():
    """Load the CUDA library; fails loudly if it has not been built (there is no CPU path)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'`")
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            if not hasattr(L, name):
                continue
            f = getattr(L, name)
            f.restype, f.argtypes = res, args
        _lib = L
    return _lib


class PbError(RuntimeError):
    pass


def _check(rc):
    if rc != 0:
        raise PbError(f"pb status {rc}: {lib().pb_last_error().decode()}")


class ParVectorEnv:
    """Drop-in for gymcpp's ParVectorEnv/SeqVectorEnv on host buffers (reset/step signatures of gym.h:310,324)."""

    def __init__(self, env_id, num_envs, clip_actions=True, chain=PB_CHAIN_PPO, gamma=0.99, device=0):
        self._h = C.c_void_p()
        _check(lib().pb_env_create(env_id.encode(), num_envs, device, chain, gamma, int(clip_actions), C.byref(self._h)))
        L = lib()
        self.num_envs_ = num_envs
        self.obs_dim, self.act_dim = L.pb_env_obs_dim(self._h), L.pb_env_act_dim(self._h)
        nq, nv, nu = C.c_int(), C.c_int(), C.c_int()
        L.pb_env_model_dims(self._h, C.byref(nq), C.byref(nv), C.byref(nu))
        self.nq, self.nv, self.nu = nq.value, nv.value, nu.value
        self._obs = np.zeros((num_envs, self.obs_dim), np.float32)
        self._rew = np.zeros(num_envs, np.float32)
        self._term = np.zeros(num_envs, np.float32)
        self._trunc = np.zeros(num_envs, np.float32)
        self._has = np.zeros(num_envs, np.int32)
        self._ir = np.zeros(num_envs, np.float32)
        self._il = np.zeros(num_envs, np.int32)

    def close(self):
        if self._h:
            lib().pb_env_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- gymcpp interface
    def get_num_envs(self):
        return self.num_envs_

    def get_observation_space(self):
        return self.obs_dim

    def get_action_space(self):
        return self.act_dim

    def get_action_space_min(self):
        return lib().pb_env_act_min(self._h)

    def get_action_space_max(self):
        return lib().pb_env_act_max(self._h)

    def reset(self, seed):
        _check(lib().pb_env_reset_host(self._h, int(seed), self._obs.ctypes.data_as(_fp)))
        return self._obs

    def step(self, actions):
        a = np.ascontiguousarray(actions, dtype=np.float32)
        if a.shape != (self.num_envs_, self.act_dim):
            raise ValueError(f"actions must be [{self.num_envs_}, {self.act_dim}], got {a.shape}")
        _check(lib().pb_env_step_host(self._h, a.ctypes.data_as(_fp), self._obs.ctypes.data_as(_fp),
                                      self._rew.ctypes.data_as(_fp), self._term.ctypes.data_as(_fp),
                                      self._trunc.ctypes.data_as(_fp), self._has.ctypes.data_as(_ip),
                                      self._ir.ctypes.data_as(_fp), self._il.ctypes.data_as(_ip)))
        infos = [(float(self._ir[i]), int(self._il[i])) if self._has[i] else None for i in range(self.num_envs_)]
        return self._obs, self._rew, self._term, self._trunc, infos

    # ---- device-pointer variants (asynchronous on the handle's stream)
    def reset_device(self, seed, obs_ptr):
        _check(lib().pb_env_reset(self._h, int(seed), C.c_void_p(obs_ptr)))

    def step_device(self, act_ptr, obs_ptr, rew_ptr, term_ptr, trunc_ptr, has_ptr=None, ir_ptr=None, il_ptr=None):
        _check(lib().pb_env_step(self._h, *[C.c_void_p(p) if p else None for p in
                                            (act_ptr, obs_ptr, rew_ptr, term_ptr, trunc_ptr, has_ptr, ir_ptr, il_ptr)]))

    def sync(self):
        _check(lib().pb_env_sync(self._h))

    def stream(self):
        return lib().pb_env_stream(self._h)

    # ---- state access for parity tests
    def get_state(self, i):
        q, v, w = np.zeros(self.nq), np.zeros(self.nv), np.zeros(self.nv)
        _check(lib().pb_env_get_state(self._h, i, q.ctypes.data_as(_dp), v.ctypes.data_as(_dp), w.ctypes.data_as(_dp)))
        return q, v, w

    def set_state(self, i, qpos, qvel, warm=None):
        q = np.ascontiguousarray(qpos, dtype=np.float64)
        v = np.ascontiguousarray(qvel, dtype=np.float64)
        w = np.ascontiguousarray(warm if warm is not None else np.zeros(self.nv), dtype=np.float64)
        _check(lib().pb_env_set_state(self._h, i, q.ctypes.data_as(_dp), v.ctypes.data_as(_dp), w.ctypes.data_as(_dp)))

    def forward(self):
        _check(lib().pb_env_forward(self._h, None))
        self.sync()

    def solver_counters(self):
        c = (C.c_ulonglong * 2)()
        _check(lib().pb_env_solver_counters(self._h, c))
        return int(c[0]), int(c[1])

    def counters(self):
        c = (C.c_ulonglong * 2)()
        _check(lib().pb_env_counters(self._h, c))
        return int(c[0]), int(c[1])
