// Velocity stage, actuation and unconstrained acceleration (SURVEY.md §8a rows a2/a3:
// mj_fwdVelocity, mj_fwdActuation, mj_fwdAcceleration behind engine.py:162-166).
#ifndef BP_DYN_CUH_
#define BP_DYN_CUH_
#include "bp_smooth.cuh"

__device__ void bp_com_vel(const Ctx& c) {
  const BpModel& m = c.P.m;
  double *cvel = DF(cvel), *cdof = DF(cdof), *cdof_dot = DF(cdof_dot), *qvel = DF(qvel);
  if (c.lane < 6) cvel[c.lane] = 0;
  SYNC();
  for (int lev = 1; lev < m.nlevel; lev++) {
    for (int idx = m.body_level_adr[lev] + c.lane; idx < m.body_level_adr[lev + 1];
         idx += 32) {
      int b = m.body_level_list[idx];
      double cv[6];
      for (int i = 0; i < 6; i++) cv[i] = cvel[6 * m.body_parentid[b] + i];
      int jn = m.body_jntnum[b], j0 = m.body_jntadr[b];
      for (int k = 0; k < jn; k++) {
        int j = j0 + k, da = m.jnt_dofadr[j], jt = m.jnt_type[j];
        if (jt == JNT_FREE) {
          for (int i = 0; i < 3; i++) {
            for (int q = 0; q < 6; q++) {
              cdof_dot[6 * (da + i) + q] = 0;
              cv[q] += cdof[6 * (da + i) + q] * qvel[da + i];
            }
          }
          da += 3;
        }
        int n = (jt == JNT_FREE || jt == JNT_BALL) ? 3 : 1;
        for (int i = 0; i < n; i++) {
          double t[6];
          cross_motion(t, cv, cdof + 6 * (da + i));
          for (int q = 0; q < 6; q++) cdof_dot[6 * (da + i) + q] = t[q];
        }
        for (int i = 0; i < n; i++)
          for (int q = 0; q < 6; q++) cv[q] += cdof[6 * (da + i) + q] * qvel[da + i];
      }
      for (int i = 0; i < 6; i++) cvel[6 * b + i] = cv[i];
    }
    SYNC();
  }
}

__device__ void bp_passive(const Ctx& c) {
  const BpModel& m = c.P.m;
  double *qp = DF(qfrc_passive), *qpos = DF(qpos), *qvel = DF(qvel);
  bool off = m.opt_disableflags & DSBL_PASSIVE;
  for (int i = c.lane; i < m.nv; i += 32) {
    double f = 0;
    if (!off) {
      int j = m.dof_jntid[i], jt = m.jnt_type[j];
      double k = m.jnt_stiffness[j];
      if (k != 0 && (jt == JNT_SLIDE || jt == JNT_HINGE)) {
        int qa = m.jnt_qposadr[j];
        f -= k * (qpos[qa] - m.qpos_spring[qa]);
      }
      f -= m.dof_damping[i] * qvel[i];
    }
    qp[i] = f;
  }
  SYNC();
}

// Recursive Newton-Euler.  with_acc adds cdof*qacc; cfrc_ext (may be null) is
// subtracted from the body forces (post-constraint variant used by sensors).
// Writes cacc / cfrc into the given buffers; result (may be null) = J^T cfrc.
__device__ void bp_rne(const Ctx& c, bool with_acc, double* cacc, double* cfrc,
                       const double* cfrc_ext, double* result, int stop_level) {
  const BpModel& m = c.P.m;
  double *cdof = DF(cdof), *cdof_dot = DF(cdof_dot), *qvel = DF(qvel), *qacc = DF(qacc);
  double *cinert = DF(cinert), *cvel = DF(cvel);
  bool grav = !(m.opt_disableflags & DSBL_GRAVITY);
  if (c.lane < 6) {
    cacc[c.lane] = (c.lane >= 3 && grav) ? -m.opt_gravity[c.lane - 3] : 0.0;
    cfrc[c.lane] = 0;
  }
  SYNC();
  for (int lev = 1; lev < m.nlevel; lev++) {
    for (int idx = m.body_level_adr[lev] + c.lane; idx < m.body_level_adr[lev + 1];
         idx += 32) {
      int b = m.body_level_list[idx];
      double a[6];
      for (int i = 0; i < 6; i++) a[i] = cacc[6 * m.body_parentid[b] + i];
      int da = m.body_dofadr[b], dn = m.body_dofnum[b];
      for (int k = 0; k < dn; k++) {
        double v = qvel[da + k], qa = with_acc ? qacc[da + k] : 0.0;
        for (int i = 0; i < 6; i++) {
          if (with_acc)
            a[i] += cdof_dot[6 * (da + k) + i] * v + cdof[6 * (da + k) + i] * qa;
          else
            a[i] += cdof_dot[6 * (da + k) + i] * v;
        }
      }
      for (int i = 0; i < 6; i++) cacc[6 * b + i] = a[i];
      double Ia[6], Iv[6], vxIv[6];
      inert_mul(Ia, cinert + 10 * b, a);
      inert_mul(Iv, cinert + 10 * b, cvel + 6 * b);
      cross_force(vxIv, cvel + 6 * b, Iv);
      for (int i = 0; i < 6; i++)
        cfrc[6 * b + i] = Ia[i] + vxIv[i] - (cfrc_ext ? cfrc_ext[6 * b + i] : 0.0);
    }
    SYNC();
  }
  bp_tree_gather(c, cfrc, 6, stop_level);
  if (result) {
    for (int i = c.lane; i < m.nv; i += 32)
      result[i] = dot6(cdof + 6 * i, cfrc + 6 * m.dof_bodyid[i]);
    SYNC();
  }
}

// joint transmission with scalar joints (the only kind the compiled subset emits)
__device__ void bp_transmission(const Ctx& c) {
  const BpModel& m = c.P.m;
  double *len = DF(actuator_length), *qpos = DF(qpos);
  for (int i = c.lane; i < m.nu; i += 32)
    len[i] = m.actuator_gear[6 * i] * qpos[m.jnt_qposadr[m.actuator_trnid[2 * i]]];
  SYNC();
}

__device__ void bp_actuator_velocity(const Ctx& c) {
  const BpModel& m = c.P.m;
  double *vel = DF(actuator_velocity), *qvel = DF(qvel);
  for (int i = c.lane; i < m.nu; i += 32)
    vel[i] = m.actuator_gear[6 * i] * qvel[m.jnt_dofadr[m.actuator_trnid[2 * i]]];
  SYNC();
}

__device__ void bp_fwd_actuation(const Ctx& c) {
  const BpModel& m = c.P.m;
  int nu = m.nu, nv = m.nv;
  double *ctrl = DF(ctrl), *force = DF(actuator_force), *qa = DF(qfrc_actuator);
  double *act = DF(act), *act_dot = DF(act_dot);
  bool off = nu == 0 || (m.opt_disableflags & DSBL_ACTUATION) || c.P.disable_actuation;
  for (int i = c.lane; i < m.na; i += 32) act_dot[i] = 0;
  if (off) {
    for (int i = c.lane; i < nu; i += 32) force[i] = 0;
    for (int i = c.lane; i < nv; i += 32) qa[i] = 0;
    SYNC();
    return;
  }
  bool bad = false;
  for (int i = c.lane; i < nu; i += 32) {
    double x = ctrl[i];
    bad |= isnan(x) || fabs(x) > BP_MAXVAL;
  }
  if (__any_sync(FULL, bad)) {
    if (c.lane == 0) IF(warning)[W_BADCTRL]++;
    for (int i = c.lane; i < nu; i += 32) ctrl[i] = 0;
  }
  SYNC();
  for (int i = c.lane; i < nu; i += 32) {
    double x = ctrl[i];
    if (m.actuator_ctrllimited[i] && !(m.opt_disableflags & DSBL_CLAMPCTRL)) {
      double lo = m.actuator_ctrlrange[2 * i], hi = m.actuator_ctrlrange[2 * i + 1];
      x = x < lo ? lo : (x > hi ? hi : x);
    }
    double in = x;
    int dyn = m.actuator_dyntype[i];
    if (dyn != 0) {
      int ia = 0;   // activation index: count of stateful actuators before i
      for (int k = 0; k < i; k++) ia += m.actuator_dyntype[k] != 0;
      double a = act[ia];
      act_dot[ia] = dyn == 1 ? x : (x - a) / fmax(BP_MINVAL, m.actuator_dynprm[3 * i]);
      in = a;
    }
    double f = m.actuator_gainprm[3 * i] * in;
    if (m.actuator_biastype[i] == 1)
      f += m.actuator_biasprm[3 * i] + m.actuator_biasprm[3 * i + 1] * DF(actuator_length)[i] +
           m.actuator_biasprm[3 * i + 2] * DF(actuator_velocity)[i];
    if (m.actuator_forcelimited[i]) {
      double lo = m.actuator_forcerange[2 * i], hi = m.actuator_forcerange[2 * i + 1];
      f = f < lo ? lo : (f > hi ? hi : f);
    }
    force[i] = f;
  }
  SYNC();
  // qfrc_actuator = moment^T force: each dof sums its own actuators in id order
  for (int d = c.lane; d < nv; d += 32) {
    double s = 0;
    for (int k = m.dof_actadr[d]; k < m.dof_actadr[d + 1]; k++) {
      int u = m.dof_actlist[k];
      s += m.actuator_gear[6 * u] * force[u];
    }
    qa[d] = s;
  }
  SYNC();
}

__device__ void bp_fwd_acceleration(const Ctx& c) {
  const BpModel& m = c.P.m;
  double *fs = DF(qfrc_smooth), *as = DF(qacc_smooth);
  for (int i = c.lane; i < m.nv; i += 32) {
    double f = DF(qfrc_passive)[i] - DF(qfrc_bias)[i] + DF(qfrc_actuator)[i] +
               DF(qfrc_applied)[i];
    fs[i] = f; as[i] = f;
  }
  SYNC();
  bp_solve_sparse(c, DF(qLD), DF(qLDiagInv), as);
}

#endif  // BP_DYN_CUH_
