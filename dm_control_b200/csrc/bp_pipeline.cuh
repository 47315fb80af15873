// The mj_step pipeline as ONE kernel, templated on the execution shape:
//   G = 32, image staged in shared memory across all sub-steps of a launch
//       (one warp per environment), or
//   G < 32, image addressed in global memory through L1/L2 (32/G environments per
//       warp; G = 1 is one thread per environment with warp-interleaved images).
//
// Replaces the three ctypes calls of the reference's Physics.step()
// (/root/reference/dm_control/mujoco/engine.py:149-166: mj_step2 + mj_step1, or
// mj_step + mj_step1 for RK4 models), Physics.forward() (engine.py:297-305) and the
// sub-step loop of control.Environment.step (rl/control.py:100-106).
#ifndef BP_PIPELINE_CUH_
#define BP_PIPELINE_CUH_
#include <cuda_runtime.h>
#include <cstdlib>

#include "bp_collide.cuh"
#include "bp_constraint.cuh"
#include "bp_sensor.cuh"
#include "bp_solver.cuh"
#include "bp_task.cuh"

// factor M and, for the implicit-damping Euler step, M + dt diag(damping) together
template <class C>
__device__ __forceinline__ void bp_factor_stage(const C& c) {
    BP_MODEL(m);
    auto qLD = DF(qLD), qM = DF(qM), MM = DF(s_MM);
    // (not in the shared-memory shape: its scratch does not survive the launch)
    const bool implicit = C::G != 32 && m.any_damping && m.opt_integrator == 0;
    bp_copy(c, qLD, qM, m.nM);
    if (implicit) bp_copy(c, MM, qM, m.nM);
    SYNC();
    if (implicit) {
      for (int i = c.lane; i < m.nv; i += C::G)
        MM[m.dof_Madr[i]] += m.opt_timestep * m.dof_damping[i];
      SYNC();
    }
    bp_factor_sparse(c, qLD, DF(qLDiagInv), implicit ? MM : c.null(),
                     implicit ? DF(s_MMdinv) : c.null());
}

// The composite functions below run for EVERY thread of a block, also for groups
// without an environment (c.act false), which only take part in the BP_LOCK barriers.
template <class C>
__device__ __forceinline__ void bp_fwd_position(const C& c) {
  if (c.act) {
    bp_kinematics(c);
  }
  BP_LOCK(9);
  if (c.act) {
    bp_com_pos(c);
    bp_transmission(c);
    bp_crb(c);
  }
  BP_LOCK(10);
  if (c.act) bp_factor_stage(c);
  BP_LOCK(0);
  RUN(bp_collision(c));
  BP_LOCK(1);
  RUN(bp_make_constraint(c));
  BP_LOCK(2);
}

// position sensors / energy and mj_fwdVelocity, in two lock-step halves
template <class C>
__device__ __forceinline__ void bp_fwd_velocity(const C& c, bool skipsensor) {
  if (c.act) {
    if (!skipsensor) bp_sensors(c, STAGE_POS);
    bp_energy_pos(c);
    bp_actuator_velocity(c);
    bp_com_vel(c);
    bp_passive(c);
  }
  BP_LOCK(11);
  if (c.act) {
    bp_reference_constraint(c);
    bp_rne(c, false, DF(s_cacc), DF(s_cfrc), c.null(), DF(qfrc_bias), 0);
    if (!skipsensor) bp_sensors(c, STAGE_VEL);
    bp_energy_vel(c);
  }
}

template <class C>
__device__ __forceinline__ void bp_forward(const C& c, bool skipsensor) {
  bp_fwd_position(c);
  bp_fwd_velocity(c, skipsensor);
  BP_LOCK(3);
  if (c.act) {
    bp_fwd_actuation(c);
    bp_fwd_acceleration(c);
  }
  BP_LOCK(4);
  bp_fwd_constraint(c);
  BP_LOCK(5);
  if (c.act && !skipsensor) bp_sensors(c, STAGE_ACC);
}

// zero the persistent image, qpos = qpos0 (mj_resetData, engine.py:268-289)
BP_STAGE void bp_reset_image(const C& c) {
  BP_MODEL(m);
  typename C::DP img{c.sm};
  for (int w = c.lane; w < bp_lv(c).persist; w += C::G) img[w] = 0;
  SYNC();
  for (int i = c.lane; i < m.nq; i += C::G) DF(qpos)[i] = m.qpos0[i];
  SYNC();
}

// MuJoCo resets the data and raises a warning when the state is not finite
BP_STAGE bool bp_check(const C& c, typename C::DP x, int n, int w) {
  bool bad = false;
  for (int i = c.lane; i < n; i += C::G) bad |= is_bad(x[i]);
  if (!group_any(c, bad)) return false;
  bp_reset_image(c);
  if (c.lane == 0) IF(warning)[w] = 1;
  SYNC();
  return true;
}

BP_STAGE void bp_integrate_pos(const C& c, typename C::DP qpos, typename C::DP qvel,
                               double dt) {
  BP_MODEL(m);
#pragma unroll 1
  for (int j = c.lane; j < m.njnt; j += C::G) {
    int qa = m.jnt_qposadr[j], da = m.jnt_dofadr[j];
    switch (m.jnt_type[j]) {
      case JNT_FREE:
        for (int k = 0; k < 3; k++) qpos[qa + k] += dt * qvel[da + k];
        quat_integrate(qpos + (qa + 3), qvel + (da + 3), dt);
        break;
      case JNT_BALL: quat_integrate(qpos + qa, qvel + da, dt); break;
      default: qpos[qa] += dt * qvel[da];
    }
  }
  SYNC();
}

BP_STAGE void bp_advance(const C& c, typename C::DP act_dot, typename C::DP qacc,
                         typename C::DP qvel_for_pos) {
  BP_MODEL(m);
  double dt = m.opt_timestep;
  for (int i = c.lane; i < m.na; i += C::G) DF(act)[i] += dt * act_dot[i];
  for (int i = c.lane; i < m.nv; i += C::G) DF(qvel)[i] += dt * qacc[i];
  SYNC();
  bp_integrate_pos(c, DF(qpos), qvel_for_pos.valid() ? qvel_for_pos : DF(qvel), dt);
  if (c.lane == 0) DF(time)[0] += dt;
  SYNC();
}

BP_STAGE void bp_euler(const C& c) {
  BP_MODEL(m);
  auto qa = DF(s_tmp);
  if (!m.any_damping) {
    for (int i = c.lane; i < m.nv; i += C::G) qa[i] = DF(qacc)[i];
    SYNC();
  } else {
    // implicit joint damping: (M + dt diag(damping)) a = f_smooth + f_constraint, with the
    // factor of that matrix left in s_MM by the position stage (bp_fwd_position)
    if (C::G == 32) {
      auto MM = DF(s_MM);
      bp_copy(c, MM, DF(qM), m.nM);
      SYNC();
      for (int i = c.lane; i < m.nv; i += C::G)
        MM[m.dof_Madr[i]] += m.opt_timestep * m.dof_damping[i];
      SYNC();
      bp_factor_sparse(c, MM, DF(s_MMdinv), c.null(), c.null());
    }
    for (int i = c.lane; i < m.nv; i += C::G)
      qa[i] = DF(qfrc_smooth)[i] + DF(qfrc_constraint)[i];
    SYNC();
    bp_solve_sparse(c, DF(s_MM), DF(s_MMdinv), qa);
  }
  bp_advance(c, DF(act_dot), qa, c.null());
}

// ---- RK4 (mj_RungeKutta(4)): state saved in s_rk = [q0 | v0 | a0 | F[4][nf] | dX[nf]] ----
template <class C>
struct RkView {
  typename C::DP q0, v0, a0, F, dX;
  int nq, nv, na, nf;
  __device__ __forceinline__ RkView(const C& c) {
    BP_MODEL(m);
    nq = m.nq; nv = m.nv; na = m.na; nf = 2 * nv + na;
    q0 = DF(s_rk); v0 = q0 + nq; a0 = v0 + nv; F = a0 + na; dX = F + 4 * nf;
  }
};

BP_STAGE void bp_rk4_begin(const C& c) {
  RkView<C> r(c);
  SYNC();
  for (int i = c.lane; i < r.nq; i += C::G) r.q0[i] = DF(qpos)[i];
  for (int i = c.lane; i < r.nv; i += C::G) {
    double v = DF(qvel)[i];
    r.v0[i] = v; r.F[i] = v; r.F[r.nv + i] = DF(qacc)[i];
  }
  for (int i = c.lane; i < r.na; i += C::G) {
    r.a0[i] = DF(act)[i]; r.F[2 * r.nv + i] = DF(act_dot)[i];
  }
  if (c.lane == 0) r.dX[r.nf] = DF(time)[0];        // t0 rides behind dX
  SYNC();
}

// state <- x0 + h * sum_j A[s-1][j] F_j, time <- t0 + h * C[s-1]
BP_STAGE void bp_rk4_stage(const C& c, int s) {
  RkView<C> r(c);
  const double A[3][3] = {{0.5, 0, 0}, {0, 0.5, 0}, {0, 0, 1}};
  const double Ct[3] = {0.5, 0.5, 1.0};
  const double h = bp_mv(c).opt_timestep, t0 = r.dX[r.nf];
  for (int k = c.lane; k < r.nf; k += C::G) {
    double x = 0;
    for (int j = 0; j < s; j++) x += A[s - 1][j] * r.F[j * r.nf + k];
    r.dX[k] = x;
  }
  for (int i = c.lane; i < r.nq; i += C::G) DF(qpos)[i] = r.q0[i];
  SYNC();
  bp_integrate_pos(c, DF(qpos), r.dX, h);
  for (int k = c.lane; k < r.nv; k += C::G) DF(qvel)[k] = r.v0[k] + h * r.dX[r.nv + k];
  for (int k = c.lane; k < r.na; k += C::G) DF(act)[k] = r.a0[k] + h * r.dX[2 * r.nv + k];
  if (c.lane == 0) DF(time)[0] = t0 + h * Ct[s - 1];
  SYNC();
}

BP_STAGE void bp_rk4_collect(const C& c, int s) {
  RkView<C> r(c);
  auto Fs = r.F + s * r.nf;
  for (int i = c.lane; i < r.nv; i += C::G) { Fs[i] = DF(qvel)[i]; Fs[r.nv + i] = DF(qacc)[i]; }
  for (int i = c.lane; i < r.na; i += C::G) Fs[2 * r.nv + i] = DF(act_dot)[i];
  SYNC();
}

BP_STAGE void bp_rk4_finish(const C& c) {
  RkView<C> r(c);
  const double B[4] = {1.0 / 6, 1.0 / 3, 1.0 / 3, 1.0 / 6};
  const double t0 = r.dX[r.nf];
  SYNC();
  for (int k = c.lane; k < r.nf; k += C::G) {
    double x = 0;
    for (int j = 0; j < 4; j++) x += B[j] * r.F[j * r.nf + k];
    r.dX[k] = x;
  }
  for (int i = c.lane; i < r.nq; i += C::G) DF(qpos)[i] = r.q0[i];
  for (int i = c.lane; i < r.nv; i += C::G) DF(qvel)[i] = r.v0[i];
  for (int i = c.lane; i < r.na; i += C::G) DF(act)[i] = r.a0[i];
  if (c.lane == 0) DF(time)[0] = t0;
  SYNC();
  bp_advance(c, r.dX + 2 * r.nv, r.dX + r.nv, r.dX);
}

template <class C>
__device__ __forceinline__ void bp_rk4(const C& c) {
  RUN(bp_rk4_begin(c));
#pragma unroll 1
  for (int s = 1; s < 4; s++) {
    RUN(bp_rk4_stage(c, s));
    bp_forward(c, true);
    RUN(bp_rk4_collect(c, s));
  }
  RUN(bp_rk4_finish(c));
}

// mj_checkAcc: a bad qacc resets the env and recomputes forward.  Data dependent, so
// the recomputation runs without stage barriers.
BP_STAGE void bp_check_acc(const C& c) {
  if (bp_check(c, DF(qacc), bp_mv(c).nv, W_BADQACC)) {
    C solo = c;
    solo.lock = false;
    bp_forward(solo, false);
  }
}

BP_STAGE void bp_check_state(const C& c) {
  BP_MODEL(m);
  bp_check(c, DF(qpos), m.nq, W_BADQPOS);
  bp_check(c, DF(qvel), m.nv, W_BADQVEL);
}

// `skipsensor`: sensors and energy are outputs only; inside a multi-sub-step launch they
// are evaluated on the last sub-step alone (the values every earlier one would have
// written are overwritten before anyone can read them).
template <class C>
__device__ __forceinline__ void bp_mj_step(const C& c, bool skipsensor) {
  RUN(bp_check_state(c));
  bp_forward(c, skipsensor);
  RUN(bp_check_acc(c));
  if (bp_mv(c).opt_integrator == 1) bp_rk4(c); else RUN(bp_euler(c));
  BP_LOCK(6);
}

template <class C>
__device__ __forceinline__ void bp_step1(const C& c, bool skipsensor) {
  RUN(bp_check_state(c));
  bp_fwd_position(c);
  bp_fwd_velocity(c, skipsensor);
  BP_LOCK(7);
}

template <class C>
__device__ __forceinline__ void bp_step2(const C& c, bool skipsensor) {
  if (c.act) {
    bp_fwd_actuation(c);
    bp_fwd_acceleration(c);
  }
  BP_LOCK(4);
  bp_fwd_constraint(c);
  BP_LOCK(5);
  if (c.act) {
    if (!skipsensor) bp_sensors(c, STAGE_ACC);
    bp_check_acc(c);
    bp_euler(c);
  }
  BP_LOCK(6);
}

// prologue of the fused Environment.step: this env's row of the dense control array
template <class C>
__device__ __forceinline__ void bp_load_ctrl(const C& c, long long env) {
  const BpParams& P = c.P;
  if (!P.ctrl_in || !c.act) return;
  const double* src = P.ctrl_in + env * P.m.nu;
  for (int i = c.lane; i < P.m.nu; i += C::G) DF(ctrl)[i] = src[i];
  SYNC();
}

template <class C>
__device__ __forceinline__ void bp_run_mode(const C& c) {
  const BpParams& P = c.P;
  switch (P.mode) {
    case BP_MODE_STEP:
#pragma unroll 1
      for (int s = 0; s < P.nsub; s++) {
        const bool skip = s + 1 < P.nsub;
        if (P.m.opt_integrator == 0) bp_step2(c, skip); else bp_mj_step(c, skip);
        bp_step1(c, skip);
      }
      break;
    case BP_MODE_FORWARD: bp_forward(c, false); break;
    case BP_MODE_STEP1: bp_step1(c, false); break;
    case BP_MODE_STEP2: bp_step2(c, false); break;
    case BP_MODE_MJSTEP:
#pragma unroll 1
      for (int s = 0; s < P.nsub; s++) bp_mj_step(c, s + 1 < P.nsub);
      break;
  }
}

// ---- G < 32: image in global memory, 32/G environments per warp -----------------
// Upper bound of threads per block = register budget of a shape: 512 threads -> 128
// registers, 896 -> 72, 1024 -> 64.  The wide shapes (G >= 8) take the 72-register build: a
// one-block-per-SM launch of 16384 x 8 or 8192 x 16 threads is 28 warps = 896 threads per SM
// (humanoid G = 16 at 64 registers: 1.04 M vs 0.96 M env-steps/s for G = 8 at 128 registers;
// 72 instead of 64 registers: cheetah 28.3 -> 29.6 M, humanoid 1.152 -> 1.189 M, quadruped-
// escape 1.39 -> 1.42 M, profiles/r1_tuning.md r1w).
#ifndef BP_TPB
#define BP_TPB(G) ((G) >= 8 ? 896 : 512)
#endif
#ifndef BP_LOCKSTEP
#define BP_LOCKSTEP 1    // 1: warps of a block re-align at stage boundaries (BP_LOCK)
#endif
#define BP_BOUNDS __launch_bounds__(BP_TPB(G))
template <int G, int S, int SHAPE, class MT = void>
__global__ void BP_BOUNDS
bp_step_kernel_gmem(const __grid_constant__ BpParams P) {
  const long long tid = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  long long env = tid / G;
  const bool act = env < P.nenv && (!P.mask || P.mask[env]);
  const bool lock = BP_LOCKSTEP != 0;
  if (!act) {
    if (!lock) return;
    env = 0;                 // pointers stay valid; nothing is touched when !act
  }
  const int wl = threadIdx.x & 31;
  double* base;
  int* ib;
  if (S == 1) {
    base = P.image + env * P.stride;
    ib = (int*)(base + P.L.ints);
  } else {          // S == 32: [warp][element][32 envs]
    double* wb = P.image + (env / 32) * P.stride * 32;
    base = wb + (env & 31);
    ib = (int*)(wb + (long long)P.L.ints * 32) + (env & 31);
  }
  unsigned mask = G == 32 ? 0xffffffffu : (((1u << G) - 1u) << (wl & ~(G - 1)));
  extern __shared__ double bp_arena[];
  double* hot = (S == 1 && P.hot_doubles > 0)
      ? bp_arena + (long long)(threadIdx.x / G) * P.hot_doubles : nullptr;
  const float* hf = P.hfield ? P.hfield + env * (long long)P.m.nhfielddata : nullptr;
  const double* gp = P.env_geom_pos ? P.env_geom_pos + env * 3LL * P.m.ngeom : (const double*)P.m.geom_pos;
  const double* sp = P.env_site_pos ? P.env_site_pos + env * 3LL * P.m.nsite : (const double*)P.m.site_pos;
  const double* bp = P.env_body_pos ? P.env_body_pos + env * 3LL * P.m.nbody : (const double*)P.m.body_pos;
  const double* bq = P.env_body_quat ? P.env_body_quat + env * 4LL * P.m.nbody : (const double*)P.m.body_quat;
  const double* wp = P.env_wrap_prm ? P.env_wrap_prm + env * (long long)P.env_wrap_n : (const double*)P.m.wrap_prm;
  Ctx<G, S, SHAPE, MT> c{P, base, ib, wl & (G - 1), mask, act, lock, hot, hf, gp, sp, bp, bq, wp};
  bp_load_ctrl(c, env);
  bp_run_mode(c);
  if (P.rew_out && act) {           // epilogue: the task's observation and reward
    if (G > 1) __syncwarp(mask);
    if (c.lane == 0) bp_task_eval(P, P.task, env, P.task_nobs, P.obs_out, P.rew_out);
  }
  // results straight to pinned host memory, block by block (BpParams.obs_host); every thread
  // of a lock-step block is still here
  if (lock && (P.obs_host || P.rew_host)) {
    __syncthreads();
    const long long e0 = blockIdx.x * (long long)blockDim.x / G;
    long long cnt = blockDim.x / G;
    if (e0 + cnt > P.nenv) cnt = P.nenv - e0;
    if (cnt > 0) {
      if (P.obs_host && P.obs_out) {
        const double* src = P.obs_out + e0 * P.task_nobs;
        double* dst = P.obs_host + e0 * P.task_nobs;
        for (long long i = threadIdx.x; i < cnt * P.task_nobs; i += blockDim.x) dst[i] = __ldcg(src + i);
      }
      if (P.rew_host) {
        for (long long i = threadIdx.x; i < cnt; i += blockDim.x)
          P.rew_host[e0 + i] = __ldcg(P.rew_out + e0 + i);
      }
    }
  }
}


// ---- split pipeline: one kernel per stage ---------------------------------------------
// The same stage functions, cut at the BP_LOCK points of Physics.step() for Euler models
// (mj_step2 then mj_step1, engine.py:149-166) into separate launches on the batch's stream.
// Each launch runs one stage's code only, so the instruction cache holds it without the
// block-wide re-alignment barriers of the fused kernel, and the grid no longer has to be
// one block per SM.  The image lives in global memory in these shapes anyway.  The solver
// segment keeps its block-uniform Newton loop.
enum { SEG_ACT = 0, SEG_SOLVE, SEG_INTEGRATE, SEG_COM, SEG_FACTOR, SEG_COLLIDE, SEG_CONSTRAINT,
       SEG_VEL1, SEG_VEL2, SEG_COUNT };

template <int G, int SEG>
__global__ void BP_BOUNDS
bp_seg_kernel(const __grid_constant__ BpParams P, int first, int last) {
  const long long tid = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  long long env = tid / G;
  const bool act = env < P.nenv && (!P.mask || P.mask[env]);
  const bool lock = SEG == SEG_SOLVE;
  if (!act) {
    if (!lock) return;
    env = 0;
  }
  const int wl = threadIdx.x & 31;
  double* base = P.image + env * P.stride;
  int* ib = (int*)(base + P.L.ints);
  unsigned mask = G == 32 ? 0xffffffffu : (((1u << G) - 1u) << (wl & ~(G - 1)));
  const float* hf = P.hfield ? P.hfield + env * (long long)P.m.nhfielddata : nullptr;
  const double* gp = P.env_geom_pos ? P.env_geom_pos + env * 3LL * P.m.ngeom : (const double*)P.m.geom_pos;
  const double* sp = P.env_site_pos ? P.env_site_pos + env * 3LL * P.m.nsite : (const double*)P.m.site_pos;
  const double* bp = P.env_body_pos ? P.env_body_pos + env * 3LL * P.m.nbody : (const double*)P.m.body_pos;
  const double* bq = P.env_body_quat ? P.env_body_quat + env * 4LL * P.m.nbody : (const double*)P.m.body_quat;
  const double* wp = P.env_wrap_prm ? P.env_wrap_prm + env * (long long)P.env_wrap_n : (const double*)P.m.wrap_prm;
  using C = Ctx<G, 1>;
  C c{P, base, ib, wl & (G - 1), mask, act, lock, nullptr, hf, gp, sp, bp, bq, wp};
  const bool skip = !last;          // sensors / energy on the last sub-step only
  if (SEG == SEG_ACT) {
    if (first) bp_load_ctrl(c, env);
    bp_fwd_actuation(c);
    bp_fwd_acceleration(c);
  } else if (SEG == SEG_SOLVE) {
    bp_fwd_constraint(c);
  } else if (SEG == SEG_INTEGRATE) {
    if (!skip) bp_sensors(c, STAGE_ACC);
    bp_check_acc(c);
    bp_euler(c);
    bp_check_state(c);
    bp_kinematics(c);
  } else if (SEG == SEG_COM) {
    bp_com_pos(c);
    bp_transmission(c);
    bp_crb(c);
  } else if (SEG == SEG_FACTOR) {
    bp_factor_stage(c);
  } else if (SEG == SEG_COLLIDE) {
    bp_collision(c);
  } else if (SEG == SEG_CONSTRAINT) {
    bp_make_constraint(c);
  } else if (SEG == SEG_VEL1) {
    if (!skip) bp_sensors(c, STAGE_POS);
    bp_energy_pos(c);
    bp_actuator_velocity(c);
    bp_com_vel(c);
    bp_passive(c);
  } else if (SEG == SEG_VEL2) {
    bp_reference_constraint(c);
    bp_rne(c, false, DF(s_cacc), DF(s_cfrc), c.null(), DF(qfrc_bias), 0);
    if (!skip) bp_sensors(c, STAGE_VEL);
    bp_energy_vel(c);
    if (last && P.rew_out) {        // epilogue: the task's observation and reward
      if (G > 1) __syncwarp(mask);
      __threadfence_block();
      if (c.lane == 0) bp_task_eval(P, P.task, env, P.task_nobs, P.obs_out, P.rew_out);
    }
  }
}

template <int G, int SEG>
static inline void bp_launch_seg(const BpParams& Q, int threads, int first, int last,
                                 cudaStream_t stream) {
  const long long total = Q.nenv * G;
  const unsigned blocks = (unsigned)((total + threads - 1) / threads);
  bp_seg_kernel<G, SEG><<<blocks, threads, 0, stream>>>(Q, first, last);
}

// one segment at this TU's lane width (called across TUs: a stage may run at its own G)
template <int G>
static int bp_launch_seg_any(const BpParams* P, int seg, int threads, int first, int last,
                             cudaStream_t stream) {
  BpParams Q = *P;
  Q.hot_doubles = 0;
  switch (seg) {
    case SEG_ACT: bp_launch_seg<G, SEG_ACT>(Q, threads, first, last, stream); break;
    case SEG_SOLVE: bp_launch_seg<G, SEG_SOLVE>(Q, threads, first, last, stream); break;
    case SEG_INTEGRATE: bp_launch_seg<G, SEG_INTEGRATE>(Q, threads, first, last, stream); break;
    case SEG_COM: bp_launch_seg<G, SEG_COM>(Q, threads, first, last, stream); break;
    case SEG_FACTOR: bp_launch_seg<G, SEG_FACTOR>(Q, threads, first, last, stream); break;
    case SEG_COLLIDE: bp_launch_seg<G, SEG_COLLIDE>(Q, threads, first, last, stream); break;
    case SEG_CONSTRAINT: bp_launch_seg<G, SEG_CONSTRAINT>(Q, threads, first, last, stream); break;
    case SEG_VEL1: bp_launch_seg<G, SEG_VEL1>(Q, threads, first, last, stream); break;
    case SEG_VEL2: bp_launch_seg<G, SEG_VEL2>(Q, threads, first, last, stream); break;
  }
  return 0;
}

extern "C" int bp_launch_seg_g4(const BpParams*, int, int, int, int, cudaStream_t);
extern "C" int bp_launch_seg_g8(const BpParams*, int, int, int, int, cudaStream_t);
extern "C" int bp_launch_seg_g16(const BpParams*, int, int, int, int, cudaStream_t);

static inline int bp_block_threads(long long threads_total, int tpb_max);

// nsub x 9 launches.  BPHYS_SEG_THREADS: block size of the non-solver segments;
// BPHYS_SEG_G: lane width per segment, e.g. "8,8,4,4,8,16,8,8,8" (4, 8 or 16; the image
// layout does not depend on it), default the batch's.
template <int G>
static int bp_launch_split(const BpParams* P, cudaStream_t stream, int solver_threads) {
  int t = 256;
  if (const char* e = getenv("BPHYS_SEG_THREADS")) {
    int v = atoi(e);
    if (v >= 32 && v <= 512 && v % 32 == 0) t = v;
  }
  int g[SEG_COUNT];
  for (int k = 0; k < SEG_COUNT; k++) g[k] = G;
  if (const char* e = getenv("BPHYS_SEG_G")) {
    int k = 0;
    for (const char* q = e; *q && k < SEG_COUNT; k++) {
      int v = atoi(q);
      if (v == 4 || v == 8 || v == 16) g[k] = v;
      while (*q && *q != ',') q++;
      if (*q == ',') q++;
    }
  }
  for (int s = 0; s < P->nsub; s++) {
    const int first = s == 0, last = s + 1 == P->nsub;
    for (int k = 0; k < SEG_COUNT; k++) {
      int th = t;
      if (k == SEG_SOLVE)
        th = g[k] == G ? solver_threads
                       : bp_block_threads(P->nenv * g[k], BP_TPB(g[k]));
      if (g[k] == 4) bp_launch_seg_g4(P, k, th, first, last, stream);
      else if (g[k] == 8) bp_launch_seg_g8(P, k, th, first, last, stream);
      else bp_launch_seg_g16(P, k, th, first, last, stream);
    }
  }
  return (int)cudaGetLastError();
}

// Block size of a launch: as many warps per block as possible (they share instruction
// cache lines between BP_LOCK barriers: +50..160 % measured, profiles/r1d_*) while the
// grid still covers the 148 SMs.
static inline int bp_block_threads(long long threads_total, int tpb_max) {
  if (const char* e = getenv("BPHYS_BLOCK_THREADS")) {       // tuning knob (multiple of 32)
    int v = atoi(e);
    if (v >= 32 && v <= tpb_max && v % 32 == 0) return v;
  }
  const long long warps = (threads_total + 31) / 32;
  const long long per_sm = (warps + 147) / 148;        // 148 SMs
  if (per_sm * 32 <= tpb_max) return per_sm < 2 ? 64 : (int)per_sm * 32;  // one block per SM
  // larger batches: blocks that fit twice per SM at the shape's register budget
  // (2 x 448 x 72 registers for the wide shapes)
  return tpb_max > 512 ? 448 : tpb_max;
}

// Launch of one global-memory shape.  BPHYS_ARENA=1 moves the per-env solver arena
// (packed Hessian + pivots + Mgrad) to shared memory when it fits; it is off by default
// because the L1 capacity it takes costs more than the shorter latency gains
// (humanoid-run: 710 k vs 790 k env-steps/s, profiles/r1_tuning.md).
template <int G, int S, int SHAPE = 0>
static int bp_launch_gmem(const BpParams* P, cudaStream_t stream) {
  const long long total = P->nenv * G;
  const int threads = bp_block_threads(total, BP_TPB(G));
  const unsigned blocks = (unsigned)((total + threads - 1) / threads);
  if constexpr (S == 1 && G >= 4 && SHAPE == 0) {
    if (P->mode == BP_MODE_STEP && P->m.opt_integrator == 0 && getenv("BPHYS_SPLIT"))
      return bp_launch_split<G>(P, stream, threads);
  }
  BpParams Q = *P;
  Q.hot_doubles = 0;
  size_t bytes = 0;
  if (S == 1 && getenv("BPHYS_ARENA")) {
    const int hot = P->m.ntri + 3 * P->m.nv;
    const size_t need = (size_t)(threads / G) * hot * sizeof(double);
    if (need <= 200 * 1024) { Q.hot_doubles = hot; bytes = need; }
  }
  static bool carved = false;
  if (!carved) {            // no shared memory in the default shape: ask for all of it as L1
    if (const char* e = getenv("BPHYS_CARVEOUT"))
      cudaFuncSetAttribute(bp_step_kernel_gmem<G, S, SHAPE>,
                           cudaFuncAttributePreferredSharedMemoryCarveout, atoi(e));
    carved = true;
  }
  static size_t configured = 48 * 1024;
  if (bytes > configured) {
    cudaError_t e = cudaFuncSetAttribute(bp_step_kernel_gmem<G, S, SHAPE>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)bytes);
    if (e != cudaSuccess) return (int)e;
    configured = bytes;
  }
  bp_step_kernel_gmem<G, S, SHAPE><<<blocks, threads, bytes, stream>>>(Q);
  return (int)cudaGetLastError();
}

#endif  // BP_PIPELINE_CUH_
