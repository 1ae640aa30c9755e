// pb_sim_common.cuh — per-task compile-time configuration, the per-env shared-memory record and small math
// helpers for the warp-per-env rigid-body kernels (one warp simulates one env; all env state lives in shared
// memory for the whole env step; lanes split bodies / dofs / pairs / constraint rows).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "pb_devmodel.h"

#define PB_FULL 0xffffffffu
#define PB_MINVAL 1e-15f

enum { PB_TASK_HALFCHEETAH = 0, PB_TASK_HOPPER = 1, PB_TASK_ANT = 2, PB_TASK_HUMANOID = 3 };

// Compile-time dimensions per task (checked against the compiled model at create time).
struct CfgHalfCheetah {
  static constexpr int TASK = PB_TASK_HALFCHEETAH, NQ = 9, NV = 9, NU = 6, NB = 8, NJ = 9, NG = 9, NPAIR = 8, NM = 36;
  static constexpr int MAXCON = 16, MAXEFC = 72, SOLVER = PB_SOL_NEWTON, INTEGRATOR = PB_INT_EULER;
  static constexpr int OBS = 17, FRAME_SKIP = 5;
  static constexpr bool NEED_CFRC = false;
  static constexpr int DOF_PARENT[NV] = {-1, 0, 1, 2, 3, 4, 2, 6, 7};
};
struct CfgHopper {
  static constexpr int TASK = PB_TASK_HOPPER, NQ = 6, NV = 6, NU = 3, NB = 5, NJ = 6, NG = 5, NPAIR = 7, NM = 21;
  static constexpr int MAXCON = 12, MAXEFC = 40, SOLVER = PB_SOL_NEWTON, INTEGRATOR = PB_INT_RK4;
  static constexpr int OBS = 11, FRAME_SKIP = 4;
  static constexpr bool NEED_CFRC = false;
  static constexpr int DOF_PARENT[NV] = {-1, 0, 1, 2, 3, 4};
};
struct CfgAnt {
  static constexpr int TASK = PB_TASK_ANT, NQ = 15, NV = 14, NU = 8, NB = 14, NJ = 9, NG = 14, NPAIR = 13, NM = 81;
  static constexpr int MAXCON = 20, MAXEFC = 64, SOLVER = PB_SOL_NEWTON, INTEGRATOR = PB_INT_RK4;
  static constexpr int OBS = 105, FRAME_SKIP = 5;
  static constexpr bool NEED_CFRC = true;
  static constexpr int DOF_PARENT[NV] = {-1, 0, 1, 2, 3, 4, 5, 6, 5, 8, 5, 10, 5, 12};
};
struct CfgHumanoid {
  static constexpr int TASK = PB_TASK_HUMANOID, NQ = 24, NV = 23, NU = 17, NB = 14, NJ = 18, NG = 18, NPAIR = 126, NM = 185;
  static constexpr int MAXCON = 16, MAXEFC = 32, SOLVER = PB_SOL_PGS, INTEGRATOR = PB_INT_RK4;
  static constexpr int OBS = 376, FRAME_SKIP = 5;
  static constexpr bool NEED_CFRC = true;
  static constexpr int DOF_PARENT[NV] = {-1, 0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11, 8, 13, 14, 15, 5, 17, 18, 5, 20, 21};
};

// Compile-time view of the dof tree (the same numbers pb_build_devmodel derives at run time; pb_env_create checks that
// the compiled model agrees): chain length of dof k including itself, and the address of its row in the sparse M / LD
// layout (row k = diagonal, then its ancestors going up).
template <class C>
struct Tree {
  static constexpr int depth(int k) { int d = 0; for (int j = k; j >= 0; j = C::DOF_PARENT[j]) d++; return d; }
  static constexpr int madr(int k) { int a = 0; for (int i = 0; i < k; i++) a += depth(i); return a; }
  static constexpr int nanc_pairs(int k) { int c = depth(k) - 1; return c * (c + 1) / 2; }
  static constexpr int fac_start(int step) { int n = 0; for (int s = 0; s < step; s++) n += nanc_pairs(C::NV - 1 - s); return n; }
  static constexpr unsigned ancmask(int k) { unsigned mk = 0; for (int j = C::DOF_PARENT[k]; j >= 0; j = C::DOF_PARENT[j]) mk |= 1u << j; return mk; }  // strict ancestors
  static constexpr int maxdepth() { int d = 0; for (int k = 0; k < C::NV; k++) d = depth(k) > d ? depth(k) : d; return d; }
  static constexpr unsigned descmask(int k) { unsigned mk = 0; for (int i = k + 1; i < C::NV; i++) if ((ancmask(i) >> k) & 1u) mk |= 1u << i; return mk; }  // strict descendants
};

// Contact record kept for the constraint build and for cfrc_ext.
struct Contact {
  float dist, pos[3], n[3], t1[3];  // frame x (normal) and y; z = n x t1
  int pair, efc;                    // pair index, first constraint row (-1: excluded)
};

// Per-env shared-memory record.  Shared memory per env bounds the warps an SM can hold (the kernel is latency-bound,
// so resident warps are throughput), hence arrays whose lifetimes inside one forward evaluation do not overlap
// share storage:
//   xanchor/xaxis (kinematics -> com_pos)            with tmpA (crb in crb_factor, cacc in the velocity stage)
//   gxpos/gz      (kinematics -> collision)          with tmpB (crb*cdof in crb_factor, cdof_dot in the velocity stage)
//   xmat/xipos    (kinematics -> com_pos)            with the constraint Jacobian (written by make_constraint)
//   cfrc          (velocity stage; cfrc_ext at the end) with the head of the solver matrices
// Stage order in sim_forward: kinematics, com_pos, collision, crb_factor, make_constraint, velocity, solver.
// Row stride of J is odd so lane-per-row accesses are bank-conflict free.
template <class C>
struct EnvS {
  static constexpr int LDJ = (C::NV | 1);
  static constexpr bool PGS = (C::SOLVER == PB_SOL_PGS);
  static constexpr int NTRI = C::MAXEFC * (C::MAXEFC + 1) / 2;
  // state
  float qpos[C::NQ], qvel[C::NV], qacc[C::NV], warm[C::NV], ctrl[C::NU];
  // kinematics that outlives the position stage
  float xpos[C::NB][3];
  float com[3];
  int ncon, nefc, overflow;
  int pgs_sweeps, pgs_rowsteps;
#ifdef PB_IMBALANCE
  unsigned dbg_t[4];   // solver set-up / sweeps / tail cycles of this env step
#endif
   // diagnostics: PGS sweeps / row steps of this env step (pb_env_solver_counters)
  // spatial quantities (c-frame)
  float cinert[C::NB][10], cdof[C::NV][6], cvel[C::NB][6];
  union {
    float tmpA[C::NB][10];   // crb (crb_factor) | cacc (velocity stage, 6 used)
    struct { float xanchor[C::NJ][3], xaxis[C::NJ][3]; };
  };
  union {
    float tmpB[C::NV][6];    // crb*cdof buffer (crb_factor) | cdof_dot (velocity stage)
    struct { float gxpos[C::NG][3], gz[C::NG][3]; };
  };
  // inertia: sparse M is kept only where a primal solver / implicit damping needs it
  float qM[PGS ? 1 : C::NM], qLD[C::NM], qLDinv[C::NV];
  float qfrc_smooth[C::NV], qacc_smooth[C::NV], qfrc_actuator[C::NV], qfrc_constraint[PGS ? 1 : C::NV];
  float vecA[C::NV], vecB[PGS ? 1 : C::NV], vecC[PGS ? 1 : C::NV], vecD[PGS ? 1 : C::NV];  // solver / integrator scratch
  // RK4 bookkeeping
  float q0[C::NQ], v0[C::NV], accv[C::NV], acca[C::NV];
  // contacts and constraint rows
  Contact con[C::MAXCON];
  union {
    float J[C::MAXEFC][LDJ];
    struct { float xmat[C::NB][9], xipos[C::NB][3]; };
  };
  float efc_R[C::MAXEFC], efc_aref[PGS ? 1 : C::MAXEFC], efc_b[PGS ? 1 : C::MAXEFC], efc_force[C::MAXEFC], efc_kip[C::MAXEFC],
      efc_bd[C::MAXEFC], efc_jar[PGS ? 1 : C::MAXEFC], efc_jv[PGS ? 1 : C::MAXEFC];
  // solver matrices: PGS -> packed lower triangle of AR; Newton -> dense M and Hessian/Cholesky [NV][NV+1] each
  union {
    float solv[PGS ? NTRI : 2 * C::NV * (C::NV + 1)];
    float cfrc[C::NB][6];    // RNE body forces / subtree sums; cfrc_ext at the end of the env step
  };
};

// ------------------------------------------------------------------ math helpers
__device__ __forceinline__ float dot3(const float *a, const float *b) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }
__device__ __forceinline__ void cross3(float *r, const float *a, const float *b) {
  float x = a[1] * b[2] - a[2] * b[1], y = a[2] * b[0] - a[0] * b[2], z = a[0] * b[1] - a[1] * b[0];
  r[0] = x; r[1] = y; r[2] = z;
}
__device__ __forceinline__ void mat3vec(float *r, const float *m, const float *v) {
  float x = m[0] * v[0] + m[1] * v[1] + m[2] * v[2];
  float y = m[3] * v[0] + m[4] * v[1] + m[5] * v[2];
  float z = m[6] * v[0] + m[7] * v[1] + m[8] * v[2];
  r[0] = x; r[1] = y; r[2] = z;
}
__device__ __forceinline__ void quat_mul(float *r, const float *a, const float *b) {
  float w = a[0] * b[0] - a[1] * b[1] - a[2] * b[2] - a[3] * b[3];
  float x = a[0] * b[1] + a[1] * b[0] + a[2] * b[3] - a[3] * b[2];
  float y = a[0] * b[2] - a[1] * b[3] + a[2] * b[0] + a[3] * b[1];
  float z = a[0] * b[3] + a[1] * b[2] - a[2] * b[1] + a[3] * b[0];
  r[0] = w; r[1] = x; r[2] = y; r[3] = z;
}
__device__ __forceinline__ void quat_normalize(float *q) {
  float n2 = q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3];
  if (n2 < 1e-30f) { q[0] = 1.f; q[1] = q[2] = q[3] = 0.f; return; }
  float inv = rsqrtf(n2);
  // one Newton step on rsqrt keeps the norm error at the fp32 rounding level
  inv = inv * (1.5f - 0.5f * n2 * inv * inv);
  q[0] *= inv; q[1] *= inv; q[2] *= inv; q[3] *= inv;
}
__device__ __forceinline__ void quat2mat(float *m, const float *q) {
  float w = q[0], x = q[1], y = q[2], z = q[3];
  m[0] = w * w + x * x - y * y - z * z; m[1] = 2.f * (x * y - w * z); m[2] = 2.f * (x * z + w * y);
  m[3] = 2.f * (x * y + w * z); m[4] = w * w - x * x + y * y - z * z; m[5] = 2.f * (y * z - w * x);
  m[6] = 2.f * (x * z - w * y); m[7] = 2.f * (y * z + w * x); m[8] = w * w - x * x - y * y + z * z;
}
__device__ __forceinline__ float normalize3(float *a) {
  float n = sqrtf(dot3(a, a));
  if (n < PB_MINVAL) { a[0] = 1.f; a[1] = 0.f; a[2] = 0.f; } else { float i = 1.f / n; a[0] *= i; a[1] *= i; a[2] *= i; }
  return n;
}
// spatial algebra, [angular, linear] about the tree com
__device__ __forceinline__ void mul_inert_vec(float *r, const float *i, const float *v) {
  r[0] = i[0] * v[0] + i[3] * v[1] + i[4] * v[2] - i[8] * v[4] + i[7] * v[5];
  r[1] = i[3] * v[0] + i[1] * v[1] + i[5] * v[2] + i[8] * v[3] - i[6] * v[5];
  r[2] = i[4] * v[0] + i[5] * v[1] + i[2] * v[2] - i[7] * v[3] + i[6] * v[4];
  r[3] = i[8] * v[1] - i[7] * v[2] + i[9] * v[3];
  r[4] = i[6] * v[2] - i[8] * v[0] + i[9] * v[4];
  r[5] = i[7] * v[0] - i[6] * v[1] + i[9] * v[5];
}
__device__ __forceinline__ void cross_motion(float *r, const float *vel, const float *v) {
  float a[3], b[3], c[3];
  cross3(a, vel, v); cross3(b, vel, v + 3); cross3(c, vel + 3, v);
  r[0] = a[0]; r[1] = a[1]; r[2] = a[2]; r[3] = b[0] + c[0]; r[4] = b[1] + c[1]; r[5] = b[2] + c[2];
}
__device__ __forceinline__ void cross_force(float *r, const float *vel, const float *f) {
  float a[3], b[3], c[3];
  cross3(a, vel, f); cross3(b, vel + 3, f + 3); cross3(c, vel, f + 3);
  r[0] = a[0] + b[0]; r[1] = a[1] + b[1]; r[2] = a[2] + b[2]; r[3] = c[0]; r[4] = c[1]; r[5] = c[2];
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(PB_FULL, v, o);
  return v;
}
__device__ __forceinline__ int warp_excl_scan(int v, int lane, int *total) {
  int x = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    int y = __shfl_up_sync(PB_FULL, x, o);
    if (lane >= o) x += y;
  }
  *total = __shfl_sync(PB_FULL, x, 31);
  return x - v;
}
