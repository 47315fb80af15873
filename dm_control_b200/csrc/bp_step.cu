// The mj_step pipeline as ONE kernel: a warp per environment, the env's data image
// resident in shared memory across all sub-steps of a launch.
//
// Replaces the three ctypes calls of the reference's Physics.step()
// (/root/reference/dm_control/mujoco/engine.py:149-166: mj_step2 + mj_step1, or
// mj_step + mj_step1 for RK4 models), Physics.forward() (engine.py:297-305) and the
// sub-step loop of control.Environment.step (rl/control.py:100-106).
#include <cuda_runtime.h>

#include "bp_collide.cuh"
#include "bp_constraint.cuh"
#include "bp_sensor.cuh"
#include "bp_solver.cuh"

__device__ void bp_fwd_position(const Ctx& c) {
  bp_kinematics(c);
  bp_com_pos(c);
  bp_transmission(c);
  bp_crb(c);
  const BpModel& m = c.P.m;
  double *qLD = DF(qLD), *qM = DF(qM);
  for (int i = c.lane; i < m.nM; i += 32) qLD[i] = qM[i];
  SYNC();
  bp_factor_sparse(c, qLD, DF(qLDiagInv));
  bp_collision(c);
  bp_make_constraint(c);
}

__device__ void bp_fwd_velocity(const Ctx& c) {
  bp_actuator_velocity(c);
  bp_com_vel(c);
  bp_passive(c);
  bp_reference_constraint(c);
  bp_rne(c, false, DF(s_cacc), DF(s_cfrc), nullptr, DF(qfrc_bias), 0);
}

__device__ void bp_forward(const Ctx& c, bool skipsensor) {
  bp_fwd_position(c);
  if (!skipsensor) bp_sensors(c, STAGE_POS);
  bp_energy_pos(c);
  bp_fwd_velocity(c);
  if (!skipsensor) bp_sensors(c, STAGE_VEL);
  bp_energy_vel(c);
  bp_fwd_actuation(c);
  bp_fwd_acceleration(c);
  bp_fwd_constraint(c);
  if (!skipsensor) bp_sensors(c, STAGE_ACC);
}

// zero the persistent image, qpos = qpos0 (mj_resetData, engine.py:268-289)
__device__ void bp_reset_image(const Ctx& c) {
  const BpModel& m = c.P.m;
  for (int w = c.lane; w < c.P.L.persist; w += 32) c.sm[w] = 0;
  SYNC();
  for (int i = c.lane; i < m.nq; i += 32) DF(qpos)[i] = m.qpos0[i];
  SYNC();
}

// MuJoCo resets the data and raises a warning when the state is not finite
__device__ bool bp_check(const Ctx& c, const double* x, int n, int w) {
  bool bad = false;
  for (int i = c.lane; i < n; i += 32) bad |= is_bad(x[i]);
  if (!__any_sync(FULL, bad)) return false;
  bp_reset_image(c);
  if (c.lane == 0) IF(warning)[w] = 1;
  SYNC();
  return true;
}

__device__ void bp_integrate_pos(const Ctx& c, double* qpos, const double* qvel, double dt) {
  const BpModel& m = c.P.m;
  for (int j = c.lane; j < m.njnt; j += 32) {
    int qa = m.jnt_qposadr[j], da = m.jnt_dofadr[j];
    switch (m.jnt_type[j]) {
      case JNT_FREE:
        for (int k = 0; k < 3; k++) qpos[qa + k] += dt * qvel[da + k];
        quat_integrate(qpos + qa + 3, qvel + da + 3, dt);
        break;
      case JNT_BALL: quat_integrate(qpos + qa, qvel + da, dt); break;
      default: qpos[qa] += dt * qvel[da];
    }
  }
  SYNC();
}

__device__ void bp_advance(const Ctx& c, const double* act_dot, const double* qacc,
                           const double* qvel_for_pos) {
  const BpModel& m = c.P.m;
  double dt = m.opt_timestep;
  for (int i = c.lane; i < m.na; i += 32) DF(act)[i] += dt * act_dot[i];
  for (int i = c.lane; i < m.nv; i += 32) DF(qvel)[i] += dt * qacc[i];
  SYNC();
  bp_integrate_pos(c, DF(qpos), qvel_for_pos ? qvel_for_pos : DF(qvel), dt);
  if (c.lane == 0) DF(time)[0] += dt;
  SYNC();
}

__device__ void bp_euler(const Ctx& c) {
  const BpModel& m = c.P.m;
  double* qa = DF(s_tmp);
  if (!m.any_damping) {
    for (int i = c.lane; i < m.nv; i += 32) qa[i] = DF(qacc)[i];
    SYNC();
  } else {
    // implicit joint damping: (M + dt diag(damping)) a = f_smooth + f_constraint
    double *MM = DF(s_MM), *qM = DF(qM);
    for (int i = c.lane; i < m.nM; i += 32) MM[i] = qM[i];
    SYNC();
    for (int i = c.lane; i < m.nv; i += 32) {
      MM[m.dof_Madr[i]] += m.opt_timestep * m.dof_damping[i];
      qa[i] = DF(qfrc_smooth)[i] + DF(qfrc_constraint)[i];
    }
    SYNC();
    bp_factor_sparse(c, MM, DF(s_MMdinv));
    bp_solve_sparse(c, MM, DF(s_MMdinv), qa);
  }
  bp_advance(c, DF(act_dot), qa, nullptr);
}

__device__ void bp_rk4(const Ctx& c) {
  const BpModel& m = c.P.m;
  const double A[3][3] = {{0.5, 0, 0}, {0, 0.5, 0}, {0, 0, 1}};
  const double B[4] = {1.0 / 6, 1.0 / 3, 1.0 / 3, 1.0 / 6};
  const double C[3] = {0.5, 0.5, 1.0};
  const int nq = m.nq, nv = m.nv, na = m.na, nf = 2 * nv + na;
  const double h = m.opt_timestep, t0 = DF(time)[0];
  double *q0 = DF(s_rk), *v0 = q0 + nq, *a0 = v0 + nv, *F = a0 + na, *dX = F + 4 * nf;
  for (int i = c.lane; i < nq; i += 32) q0[i] = DF(qpos)[i];
  for (int i = c.lane; i < nv; i += 32) { v0[i] = DF(qvel)[i]; F[i] = v0[i]; F[nv + i] = DF(qacc)[i]; }
  for (int i = c.lane; i < na; i += 32) { a0[i] = DF(act)[i]; F[2 * nv + i] = DF(act_dot)[i]; }
  SYNC();
  for (int s = 1; s < 4; s++) {
    for (int k = c.lane; k < nf; k += 32) {
      double x = 0;
      for (int j = 0; j < s; j++) x += A[s - 1][j] * F[j * nf + k];
      dX[k] = x;
    }
    for (int i = c.lane; i < nq; i += 32) DF(qpos)[i] = q0[i];
    SYNC();
    bp_integrate_pos(c, DF(qpos), dX, h);
    for (int k = c.lane; k < nv; k += 32) DF(qvel)[k] = v0[k] + h * dX[nv + k];
    for (int k = c.lane; k < na; k += 32) DF(act)[k] = a0[k] + h * dX[2 * nv + k];
    if (c.lane == 0) DF(time)[0] = t0 + h * C[s - 1];
    SYNC();
    bp_forward(c, true);
    double* Fs = F + s * nf;
    for (int i = c.lane; i < nv; i += 32) { Fs[i] = DF(qvel)[i]; Fs[nv + i] = DF(qacc)[i]; }
    for (int i = c.lane; i < na; i += 32) Fs[2 * nv + i] = DF(act_dot)[i];
    SYNC();
  }
  for (int k = c.lane; k < nf; k += 32) {
    double x = 0;
    for (int j = 0; j < 4; j++) x += B[j] * F[j * nf + k];
    dX[k] = x;
  }
  for (int i = c.lane; i < nq; i += 32) DF(qpos)[i] = q0[i];
  for (int i = c.lane; i < nv; i += 32) DF(qvel)[i] = v0[i];
  for (int i = c.lane; i < na; i += 32) DF(act)[i] = a0[i];
  if (c.lane == 0) DF(time)[0] = t0;
  SYNC();
  bp_advance(c, dX + 2 * nv, dX + nv, dX);
}

__device__ void bp_check_acc(const Ctx& c) {
  if (bp_check(c, DF(qacc), c.P.m.nv, W_BADQACC)) bp_forward(c, false);
}

__device__ void bp_mj_step(const Ctx& c) {
  const BpModel& m = c.P.m;
  bp_check(c, DF(qpos), m.nq, W_BADQPOS);
  bp_check(c, DF(qvel), m.nv, W_BADQVEL);
  bp_forward(c, false);
  bp_check_acc(c);
  if (m.opt_integrator == 1) bp_rk4(c); else bp_euler(c);
}

__device__ void bp_step1(const Ctx& c) {
  const BpModel& m = c.P.m;
  bp_check(c, DF(qpos), m.nq, W_BADQPOS);
  bp_check(c, DF(qvel), m.nv, W_BADQVEL);
  bp_fwd_position(c);
  bp_sensors(c, STAGE_POS);
  bp_energy_pos(c);
  bp_fwd_velocity(c);
  bp_sensors(c, STAGE_VEL);
  bp_energy_vel(c);
}

__device__ void bp_step2(const Ctx& c) {
  bp_fwd_actuation(c);
  bp_fwd_acceleration(c);
  bp_fwd_constraint(c);
  bp_sensors(c, STAGE_ACC);
  bp_check_acc(c);
  bp_euler(c);
}

extern "C" __global__ void __launch_bounds__(32)
bp_step_kernel(const __grid_constant__ BpParams P) {
  extern __shared__ double smem[];
  const int lane = threadIdx.x;
  const long long env = blockIdx.x;
  if (env >= P.nenv || (P.mask && !P.mask[env])) return;
  double* g = P.image + env * (long long)P.L.persist;
  for (int w = lane; w < P.L.persist; w += 32) smem[w] = g[w];
  __syncwarp();
  Ctx c{P, smem, (int*)(smem + P.L.ints), lane};
  switch (P.mode) {
    case BP_MODE_STEP:
      for (int s = 0; s < P.nsub; s++) {
        if (P.m.opt_integrator == 0) bp_step2(c); else bp_mj_step(c);
        bp_step1(c);
      }
      break;
    case BP_MODE_FORWARD: bp_forward(c, false); break;
    case BP_MODE_STEP1: bp_step1(c); break;
    case BP_MODE_STEP2: bp_step2(c); break;
    case BP_MODE_MJSTEP: for (int s = 0; s < P.nsub; s++) bp_mj_step(c); break;
  }
  __syncwarp();
  for (int w = lane; w < P.L.persist; w += 32) g[w] = smem[w];
}

extern "C" int bp_launch_step(const BpParams* P, cudaStream_t stream) {
  size_t bytes = (size_t)P->L.total * sizeof(double);
  static size_t configured = 0;
  if (bytes > configured) {
    cudaError_t e = cudaFuncSetAttribute(bp_step_kernel,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)bytes);
    if (e != cudaSuccess) return (int)e;
    configured = bytes;
  }
  bp_step_kernel<<<(unsigned)P->nenv, 32, bytes, stream>>>(*P);
  return (int)cudaGetLastError();
}
