"""ctypes loader for the CPU oracle (test infrastructure; never imported by the product package)."""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

NO_GRAVITY, NO_CONTACT, NO_LIMIT, NO_PASSIVE, NO_ARMATURE = 1, 2, 4, 8, 16


def build():
    subprocess.check_call(["make", "-s", "-C", _HERE, "libpbo.so"])


def lib():
    global _LIB
    if _LIB is None:
        path = os.environ.get("PBO_LIB") or os.path.join(_HERE, "libpbo.so")   # PBO_LIB: the op-counting build (oracle/count)
        if not os.path.exists(path):
            build()
        L = C.CDLL(path)
        L.pbo_model_by_name.restype = C.c_void_p
        L.pbo_model_by_name.argtypes = [C.c_char_p]
        L.pbo_model_variant.restype = C.c_void_p
        L.pbo_model_variant.argtypes = [C.c_char_p, C.c_int, C.c_int, C.c_double]
        L.pbo_model_free.argtypes = [C.c_void_p]
        L.pbo_model_dims.argtypes = [C.c_void_p, C.POINTER(C.c_int)]
        L.pbo_data_new.restype = C.c_void_p
        L.pbo_data_new.argtypes = [C.c_void_p]
        for f in ("pbo_data_free", "pbo_data_reset", "pbo_forward", "pbo_step", "pbo_rne_post_constraint"):
            getattr(L, f).argtypes = [C.c_void_p]
            getattr(L, f).restype = None
        L.pbo_field.argtypes = [C.c_void_p, C.c_char_p, C.POINTER(C.POINTER(C.c_double))]
        L.pbo_counts.argtypes = [C.c_void_p, C.POINTER(C.c_int)]
        L.pbo_set_caps.argtypes = [C.c_void_p, C.c_int, C.c_int]
        L.pbo_overflow.argtypes = [C.c_void_p]
        L.pbo_trace_reset.argtypes = [C.c_void_p]
        L.pbo_trace_get.argtypes = [C.c_void_p, C.POINTER(C.c_short), C.POINTER(C.c_short)]
        L.pbo_contact_get.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_int)]
        L.pbo_energy.restype = C.c_double
        L.pbo_energy.argtypes = [C.c_void_p, C.POINTER(C.c_double), C.POINTER(C.c_double)]
        _LIB = L
    return _LIB


class Sim:
    """One oracle mjData-equivalent. ``flags``/``integrator``/``timestep`` build a private model variant."""

    def __init__(self, env_id, flags=0, integrator=-1, timestep=-1.0):
        L = lib()
        self.env_id = env_id
        self._own = bool(flags or integrator >= 0 or timestep > 0)
        if self._own:
            self.model = L.pbo_model_variant(env_id.encode(), flags, integrator, timestep)
        else:
            self.model = L.pbo_model_by_name(env_id.encode())
        assert self.model, env_id
        dims = (C.c_int * 7)()
        L.pbo_model_dims(self.model, dims)
        self.nq, self.nv, self.nu, self.nbody, self.njnt, self.ngeom, self.npair = list(dims)
        self.d = L.pbo_data_new(self.model)

    @classmethod
    def borrowed(cls, env_id, data_ptr, keep):
        s = cls.__new__(cls)
        L = lib()
        s.env_id, s._own, s._borrowed, s._keep = env_id, False, True, keep
        s.model = L.pbo_model_by_name(env_id.encode())
        dims = (C.c_int * 7)()
        L.pbo_model_dims(s.model, dims)
        s.nq, s.nv, s.nu, s.nbody, s.njnt, s.ngeom, s.npair = list(dims)
        s.d = data_ptr
        return s

    def __del__(self):
        try:
            if getattr(self, "_borrowed", False):
                return
            lib().pbo_data_free(self.d)
            if self._own:
                lib().pbo_model_free(self.model)
        except Exception:
            pass

    def field(self, name):
        """numpy view (no copy) into the oracle's data."""
        p = C.POINTER(C.c_double)()
        n = lib().pbo_field(self.d, name.encode(), C.byref(p))
        if n < 0:
            raise KeyError(name)
        if n == 0:
            return np.zeros(0)
        return np.ctypeslib.as_array(p, shape=(n,))

    def __getattr__(self, name):
        if name.startswith("_") or name in ("d", "model", "env_id"):
            raise AttributeError(name)
        return self.field(name)

    def counts(self):
        out = (C.c_int * 4)()
        lib().pbo_counts(self.d, out)
        return dict(ncon=out[0], nefc=out[1], niter=out[2], nforward=out[3])

    def model_field(self, name):
        """Copy of a model constant array (doubles, else ints) of this Sim's model."""
        L = lib()
        L.pbo_model_field.argtypes = [C.c_void_p, C.c_char_p, C.POINTER(C.POINTER(C.c_double))]
        L.pbo_model_ifield.argtypes = [C.c_void_p, C.c_char_p, C.POINTER(C.POINTER(C.c_int))]
        p = C.POINTER(C.c_double)()
        n = L.pbo_model_field(self.model, name.encode(), C.byref(p))
        if n >= 0:
            return np.ctypeslib.as_array(p, shape=(n,)).copy()
        q = C.POINTER(C.c_int)()
        n = L.pbo_model_ifield(self.model, name.encode(), C.byref(q))
        if n < 0:
            raise KeyError(name)
        return np.ctypeslib.as_array(q, shape=(n,)).copy()

    def contact_params(self, i):
        out = (C.c_double * 11)()
        lib().pbo_contact_params.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_double)]
        assert lib().pbo_contact_params(self.d, i, out) == 0
        o = np.array(out)
        return dict(includemargin=o[0], friction=o[1:4], solref=o[4:6], solimp=o[6:11])

    def set_caps(self, cap_con, cap_efc):
        """Storage caps of the engine under test (csrc/pb_sim_common.cuh MAXCON / MAXEFC); same drop rule as the kernels."""
        lib().pbo_set_caps(self.d, cap_con, cap_efc)

    def overflow(self):
        return bool(lib().pbo_overflow(self.d))

    def trace_reset(self):
        lib().pbo_trace_reset(self.d)

    def trace(self):
        """(nefc, solver iterations) of every forward evaluation since trace_reset (at most 64)."""
        a, b = (C.c_short * 64)(), (C.c_short * 64)()
        n = lib().pbo_trace_get(self.d, a, b)
        return np.array(a[:n]), np.array(b[:n])

    def contact(self, i):
        out = (C.c_double * 13)()
        io = (C.c_int * 4)()
        assert lib().pbo_contact_get(self.d, i, out, io) == 0
        o = np.array(out)
        return dict(dist=o[0], pos=o[1:4], frame=o[4:13].reshape(3, 3), dim=io[0], g1=io[1], g2=io[2], efc=io[3])

    def reset(self):
        lib().pbo_data_reset(self.d)

    def forward(self):
        lib().pbo_forward(self.d)

    def step(self, n=1):
        for _ in range(n):
            lib().pbo_step(self.d)

    def rne_post(self):
        lib().pbo_rne_post_constraint(self.d)

    def energy(self):
        ke, pe = C.c_double(), C.c_double()
        e = lib().pbo_energy(self.d, C.byref(ke), C.byref(pe))
        return e, ke.value, pe.value

    def mass_matrix(self):
        return self.field("M").reshape(24, 24)[: self.nv, : self.nv].copy()


class VecEnv:
    """Oracle vector env: the reference's ParVectorEnv + wrapper chain (gym.h:276-366) on the CPU."""

    def __init__(self, env_id, num_envs, chain=1, gamma=0.99, clip_actions=True, nthreads=1):
        L = lib()
        L.pbo_vec_create.restype = C.c_void_p
        L.pbo_vec_create.argtypes = [C.c_char_p, C.c_int, C.c_int, C.c_float, C.c_int, C.c_int]
        L.pbo_vec_destroy.argtypes = [C.c_void_p]
        L.pbo_vec_obs_dim.argtypes = [C.c_void_p]
        L.pbo_vec_act_dim.argtypes = [C.c_void_p]
        fp, ip = C.POINTER(C.c_float), C.POINTER(C.c_int)
        L.pbo_vec_reset.argtypes = [C.c_void_p, C.c_int, fp]
        L.pbo_vec_step.argtypes = [C.c_void_p, fp, fp, fp, fp, fp, ip, fp, ip]
        L.pbo_vec_data.restype = C.c_void_p
        L.pbo_vec_data.argtypes = [C.c_void_p, C.c_int]
        L.pbo_vec_total_forwards.restype = C.c_long
        L.pbo_vec_total_forwards.argtypes = [C.c_void_p]
        self.v = L.pbo_vec_create(env_id.encode(), num_envs, chain, gamma, int(clip_actions), nthreads)
        assert self.v, env_id
        self.n = num_envs
        self.obs_dim = L.pbo_vec_obs_dim(self.v)
        self.act_dim = L.pbo_vec_act_dim(self.v)
        self.obs = np.zeros((num_envs, self.obs_dim), np.float32)
        self.rew = np.zeros(num_envs, np.float32)
        self.term = np.zeros(num_envs, np.float32)
        self.trunc = np.zeros(num_envs, np.float32)
        self.has_info = np.zeros(num_envs, np.int32)
        self.info_r = np.zeros(num_envs, np.float32)
        self.info_l = np.zeros(num_envs, np.int32)

    def __del__(self):
        try:
            lib().pbo_vec_destroy(self.v)
        except Exception:
            pass

    @staticmethod
    def _fp(a):
        return a.ctypes.data_as(C.POINTER(C.c_float))

    def reset(self, seed):
        lib().pbo_vec_reset(self.v, seed, self._fp(self.obs))
        return self.obs

    def step(self, actions):
        a = np.ascontiguousarray(actions, dtype=np.float32).reshape(self.n, self.act_dim)
        ip = C.POINTER(C.c_int)
        lib().pbo_vec_step(self.v, self._fp(a), self._fp(self.obs), self._fp(self.rew), self._fp(self.term),
                           self._fp(self.trunc), self.has_info.ctypes.data_as(ip), self._fp(self.info_r),
                           self.info_l.ctypes.data_as(ip))
        return self.obs, self.rew, self.term, self.trunc, (self.has_info, self.info_r, self.info_l)

    def set_norm_stats(self, ob_mean, ob_var, ob_count, rw4=None):
        L = lib()
        fp = C.POINTER(C.c_float)
        L.pbo_vec_set_norm_stats.argtypes = [C.c_void_p, fp, fp, C.c_float, fp]
        m = np.ascontiguousarray(ob_mean, dtype=np.float32)
        v = np.ascontiguousarray(ob_var, dtype=np.float32)
        r = None if rw4 is None else np.ascontiguousarray(rw4, dtype=np.float32)
        L.pbo_vec_set_norm_stats(self.v, self._fp(m), self._fp(v), float(ob_count), None if r is None else self._fp(r))

    def sim(self, i):
        """Sim view onto env i's physics data (borrowed pointer; keeps the vector env alive)."""
        return Sim.borrowed(ENV_OF_OBS[self.obs_dim], lib().pbo_vec_data(self.v, i), self)

    def total_forwards(self):
        return lib().pbo_vec_total_forwards(self.v)


ENV_OF_OBS = {17: "HalfCheetah-v5", 11: "Hopper-v5", 105: "Ant-v5", 376: "Humanoid-v4"}
