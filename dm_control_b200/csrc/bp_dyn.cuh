// Velocity stage, actuation and unconstrained acceleration (SURVEY.md §8a rows a2/a3:
// mj_fwdVelocity, mj_fwdActuation, mj_fwdAcceleration behind engine.py:162-166).
#ifndef BP_DYN_CUH_
#define BP_DYN_CUH_
#include "bp_smooth.cuh"

BP_STAGE void bp_com_vel(const C& c) {
  BP_MODEL(m);
  auto cvel = DF(cvel), cdof = DF(cdof), cdof_dot = DF(cdof_dot), qvel = DF(qvel);
  for (int i = c.lane; i < 6; i += C::G) cvel[i] = 0;
  SYNC();
#pragma unroll (C::UF)
  for (int lev = 1; lev < m.nlevel; lev++) {
#pragma unroll (C::UF)
    for (int idx = m.body_level_adr[lev] + c.lane; idx < m.body_level_adr[lev + 1];
         idx += C::G) {
      int b = m.body_level_list[idx];
      double cv[6];
      for (int i = 0; i < 6; i++) cv[i] = cvel[6 * m.body_parentid[b] + i];
      int jn = m.body_jntnum[b], j0 = m.body_jntadr[b];
#pragma unroll (C::UF)
      for (int k = 0; k < jn; k++) {
        int j = j0 + k, da = m.jnt_dofadr[j], jt = m.jnt_type[j];
        if (jt == JNT_FREE) {
          for (int i = 0; i < 3; i++) {
            double v = qvel[da + i];
            for (int q = 0; q < 6; q++) {
              cdof_dot[6 * (da + i) + q] = 0;
              cv[q] += cdof[6 * (da + i) + q] * v;
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
        for (int i = 0; i < n; i++) {
          double v = qvel[da + i];
          for (int q = 0; q < 6; q++) cv[q] += cdof[6 * (da + i) + q] * v;
        }
      }
      for (int i = 0; i < 6; i++) cvel[6 * b + i] = cv[i];
    }
    SYNC();
  }
}

BP_STAGE void bp_passive(const C& c) {
  BP_MODEL(m);
  auto qp = DF(qfrc_passive), qpos = DF(qpos), qvel = DF(qvel);
  bool off = m.opt_disableflags & DSBL_PASSIVE;
  for (int i = c.lane; i < m.nv; i += C::G) {
    double f = 0;
    if (!off) {
      int j = m.dof_jntid[i], jt = m.jnt_type[j];
      double k = m.jnt_stiffness[j];
      if (k != 0 && (jt == JNT_SLIDE || jt == JNT_HINGE)) {
        int qa = m.jnt_qposadr[j];
        f -= k * (qpos[qa] - m.qpos_spring[qa]);
      }
      f -= m.dof_damping[i] * qvel[i];
      // tendon spring-dampers: each dof gathers the tendons that wrap its joint
      if constexpr (C::FULL) for (int t = 0; t < m.ntendon; t++) {
        double kt = m.tendon_stiffness[t], bt = m.tendon_damping[t];
        if (kt == 0 && bt == 0) continue;
        double frc = -kt * (DF(ten_length)[t] - m.tendon_length0[t]) - bt * DF(ten_velocity)[t];
        for (int w = m.tendon_adr[t]; w < m.tendon_adr[t] + m.tendon_num[t]; w++)
          if (m.wrap_objid[w] == j) f += c.wprm[w] * frc;
      }
    }
    qp[i] = f;
  }
  SYNC();
  // fluid forces [MJ-doc: inertia-box model of mj_passive]: every body is replaced by the
  // box with its mass and inertia; viscous force / torque of the equivalent sphere plus
  // quadratic drag per box face.  Bodies on lanes build their wrench (moved to the subtree
  // centre the cdofs refer to) in scratch, then dofs on lanes gather over their subtree.
  if constexpr (C::FULL) if (!off && (m.opt_density > 0 || m.opt_viscosity > 0)) {
    const double PI = 3.14159265358979323846;
    auto wr = DF(s_cfrc), cvel = DF(cvel), xipos = DF(xipos), ximat = DF(ximat), scom = DF(subtree_com);
#pragma unroll (C::UF)
    for (int b = 1 + c.lane; b < m.nbody; b += C::G) {
      double w[6] = {0, 0, 0, 0, 0, 0};
      const double mass = m.body_mass[b];
      if (!(mass < BP_MINVAL)) {
        const double* I = m.body_inertia + 3 * b;
        double box[3] = {sqrt(fmax(BP_MINVAL, I[1] + I[2] - I[0]) / mass * 6.0),
                         sqrt(fmax(BP_MINVAL, I[0] + I[2] - I[1]) / mass * 6.0),
                         sqrt(fmax(BP_MINVAL, I[0] + I[1] - I[2]) / mass * 6.0)};
        auto com = scom + 3 * m.body_rootid[b];
        double cv[6], R[9], dif[3], t3[3], vlin[3], lw[3], lv[3], fa[3] = {0, 0, 0}, fl[3] = {0, 0, 0};
        for (int k = 0; k < 6; k++) cv[k] = cvel[6 * b + k];
        for (int k = 0; k < 9; k++) R[k] = ximat[9 * b + k];
        for (int k = 0; k < 3; k++) dif[k] = xipos[3 * b + k] - com[k];
        cross3(t3, cv, dif);
        for (int k = 0; k < 3; k++) vlin[k] = cv[3 + k] + t3[k];
        matT_vec3(lw, R, cv);
        matT_vec3(lv, R, vlin);
        if (m.opt_viscosity > 0) {
          double diam = (box[0] + box[1] + box[2]) / 3.0;
          for (int k = 0; k < 3; k++) {
            fa[k] = -PI * diam * diam * diam * m.opt_viscosity * lw[k];
            fl[k] = -3.0 * PI * diam * m.opt_viscosity * lv[k];
          }
        }
        if (m.opt_density > 0) {
          const double rho = m.opt_density;
          fl[0] -= 0.5 * rho * box[1] * box[2] * fabs(lv[0]) * lv[0];
          fl[1] -= 0.5 * rho * box[0] * box[2] * fabs(lv[1]) * lv[1];
          fl[2] -= 0.5 * rho * box[0] * box[1] * fabs(lv[2]) * lv[2];
          double b0 = box[0] * box[0], b1 = box[1] * box[1], b2 = box[2] * box[2];
          fa[0] -= rho * box[0] * (b1 * b1 + b2 * b2) * fabs(lw[0]) * lw[0] / 64.0;
          fa[1] -= rho * box[1] * (b0 * b0 + b2 * b2) * fabs(lw[1]) * lw[1] / 64.0;
          fa[2] -= rho * box[2] * (b0 * b0 + b1 * b1) * fabs(lw[2]) * lw[2] / 64.0;
        }
        double ga[3], gl[3];
        mat_vec3(ga, R, fa);
        mat_vec3(gl, R, fl);
        cross3(t3, dif, gl);
        for (int k = 0; k < 3; k++) { w[k] = ga[k] + t3[k]; w[3 + k] = gl[k]; }
      }
      for (int k = 0; k < 6; k++) wr[6 * b + k] = w[k];
    }
    SYNC();
    auto cdof = DF(cdof);
#pragma unroll (C::UF)
    for (int i = c.lane; i < m.nv; i += C::G) {
      double acc = 0;
      for (int b = 1; b < m.nbody; b++)
        if (m.body_dofmask[b] >> i & 1ull) acc += dot6(cdof + 6 * i, wr + 6 * b);
      qp[i] += acc;
    }
    SYNC();
  }
}

// Recursive Newton-Euler.  with_acc adds cdof*qacc; cfrc_ext (may be null) is
// subtracted from the body forces (post-constraint variant used by sensors).
// Writes cacc / cfrc into the given buffers; result (may be null) = J^T cfrc.
BP_STAGE void bp_rne(const C& c, bool with_acc, typename C::DP cacc, typename C::DP cfrc,
                     typename C::DP cfrc_ext, typename C::DP result, int stop_level) {
  BP_MODEL(m);
  auto cdof = DF(cdof), cdof_dot = DF(cdof_dot), qvel = DF(qvel), qacc = DF(qacc);
  auto cinert = DF(cinert), cvel = DF(cvel);
  bool grav = !(m.opt_disableflags & DSBL_GRAVITY);
  for (int i = c.lane; i < 6; i += C::G) {
    cacc[i] = (i >= 3 && grav) ? -m.opt_gravity[i - 3] : 0.0;
    cfrc[i] = 0;
  }
  SYNC();
#pragma unroll (C::UF)
  for (int lev = 1; lev < m.nlevel; lev++) {
#pragma unroll (C::UF)
    for (int idx = m.body_level_adr[lev] + c.lane; idx < m.body_level_adr[lev + 1];
         idx += C::G) {
      int b = m.body_level_list[idx];
      double a[6];
      for (int i = 0; i < 6; i++) a[i] = cacc[6 * m.body_parentid[b] + i];
      int da = m.body_dofadr[b], dn = m.body_dofnum[b];
#pragma unroll (C::UF)
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
      double Ia[6], Iv[6], vxIv[6], cv[6];
      for (int i = 0; i < 6; i++) cv[i] = cvel[6 * b + i];
      inert_mul(Ia, cinert + 10 * b, a);
      inert_mul(Iv, cinert + 10 * b, cv);
      cross_force(vxIv, cv, Iv);
      for (int i = 0; i < 6; i++)
        cfrc[6 * b + i] = Ia[i] + vxIv[i] - (cfrc_ext.valid() ? cfrc_ext[6 * b + i] : 0.0);
    }
    SYNC();
  }
  bp_tree_gather<6>(c, cfrc, stop_level);
  if (result.valid()) {
    for (int i = c.lane; i < m.nv; i += C::G)
      result[i] = dot6(cdof + 6 * i, cfrc + 6 * m.dof_bodyid[i]);
    SYNC();
  }
}

// joint transmission with scalar joints (the only kind the compiled subset emits)
// mj_tendon (fixed tendons: length = sum coef * qpos) + mj_transmission
BP_STAGE void bp_transmission(const C& c) {
  BP_MODEL(m);
  auto len = DF(actuator_length), qpos = DF(qpos);
  if constexpr (!C::LEAN) if (m.ntendon > 0) {
    auto tl = DF(ten_length);
    for (int t = c.lane; t < m.ntendon; t += C::G) {
      if (C::FULL && m.tendon_Jadr[t] >= 0) continue;
      double s = 0;
      for (int w = m.tendon_adr[t]; w < m.tendon_adr[t] + m.tendon_num[t]; w++)
        s += c.wprm[w] * qpos[m.jnt_qposadr[m.wrap_objid[w]]];
      tl[t] = s;
    }
    // spatial tendons (straight segments between sites): dofs on lanes build the dense
    // Jacobian row  sum_seg (jacp(p1) - jacp(p0))^T dir;  lane 0 keeps the length
    if constexpr (C::FULL) if (m.nspatial > 0) {
      auto sx = DF(site_xpos), cdof = DF(cdof), com = DF(subtree_com), tJ = DF(ten_J);
#pragma unroll (C::UF)
      for (int t = 0; t < m.ntendon; t++) {
        const int ja = m.tendon_Jadr[t];
        if (ja < 0) continue;
        double length = 0;
        for (int i0 = 0; i0 < m.nv; i0 += C::G) {
          const int i = i0 + c.lane;
          double acc = 0;
          length = 0;
#pragma unroll (C::UF)
          for (int w = m.tendon_adr[t]; w + 1 < m.tendon_adr[t] + m.tendon_num[t]; w++) {
            const int s0 = m.wrap_objid[w], s1 = m.wrap_objid[w + 1];
            double p0[3], p1[3], dif[3];
            for (int k = 0; k < 3; k++) { p0[k] = sx[3 * s0 + k]; p1[k] = sx[3 * s1 + k]; dif[k] = p1[k] - p0[k]; }
            const double dist = sqrt(dif[0] * dif[0] + dif[1] * dif[1] + dif[2] * dif[2]);
            if (dist < BP_MINVAL) { dif[0] = 1; dif[1] = 0; dif[2] = 0; }
            else for (int k = 0; k < 3; k++) dif[k] /= dist;
            length += dist;
            if (i >= m.nv) continue;
            const int b0 = m.site_bodyid[s0], b1 = m.site_bodyid[s1];
            double cd[6], j0[3] = {0, 0, 0}, j1[3] = {0, 0, 0};
            for (int k = 0; k < 6; k++) cd[k] = cdof[6 * i + k];
            if ((m.body_dofmask[b0] >> i) & 1ull) {
              auto cm = com + 3 * m.body_rootid[b0];
              double off[3] = {p0[0] - cm[0], p0[1] - cm[1], p0[2] - cm[2]}, x[3];
              cross3(x, cd, off);
              for (int k = 0; k < 3; k++) j0[k] = cd[3 + k] + x[k];
            }
            if ((m.body_dofmask[b1] >> i) & 1ull) {
              auto cm = com + 3 * m.body_rootid[b1];
              double off[3] = {p1[0] - cm[0], p1[1] - cm[1], p1[2] - cm[2]}, x[3];
              cross3(x, cd, off);
              for (int k = 0; k < 3; k++) j1[k] = cd[3 + k] + x[k];
            }
            for (int k = 0; k < 3; k++) acc += (j1[k] - j0[k]) * dif[k];
          }
          if (i < m.nv) tJ[ja * m.nv + i] = acc;
        }
        if (c.lane == 0) tl[t] = length;
      }
    }
    SYNC();
  }
  for (int i = c.lane; i < m.nu; i += C::G) {
    int id = m.actuator_trnid[2 * i];
    len[i] = m.actuator_gear[6 * i] *
             (!C::LEAN && m.actuator_trntype[i] == 3 ? DF(ten_length)[id]
                                                     : qpos[m.jnt_qposadr[id]]);
  }
  SYNC();
}

BP_STAGE void bp_actuator_velocity(const C& c) {
  BP_MODEL(m);
  auto vel = DF(actuator_velocity), qvel = DF(qvel);
  if constexpr (!C::LEAN) for (int t = c.lane; t < m.ntendon; t += C::G) {
    double s = 0;
    if (C::FULL && m.tendon_Jadr[t] >= 0)
      for (int i = 0; i < m.nv; i++) s += DF(ten_J)[m.tendon_Jadr[t] * m.nv + i] * qvel[i];
    else
      for (int w = m.tendon_adr[t]; w < m.tendon_adr[t] + m.tendon_num[t]; w++)
        s += c.wprm[w] * qvel[m.jnt_dofadr[m.wrap_objid[w]]];
    DF(ten_velocity)[t] = s;
  }
  for (int i = c.lane; i < m.nu; i += C::G) {
    int id = m.actuator_trnid[2 * i];
    const double g = m.actuator_gear[6 * i];
    if (!C::LEAN && m.actuator_trntype[i] == 3) {   // moment row = gear * tendon coefficients
      double s = 0;
      for (int w = m.tendon_adr[id]; w < m.tendon_adr[id] + m.tendon_num[id]; w++)
        s += g * c.wprm[w] * qvel[m.jnt_dofadr[m.wrap_objid[w]]];
      vel[i] = s;
    } else {
      vel[i] = g * qvel[m.jnt_dofadr[id]];
    }
  }
  SYNC();
}

BP_STAGE void bp_fwd_actuation(const C& c) {
  BP_MODEL(m);
  int nu = m.nu, nv = m.nv;
  auto ctrl = DF(ctrl), force = DF(actuator_force), qa = DF(qfrc_actuator);
  auto act = DF(act), act_dot = DF(act_dot);
  bool off = nu == 0 || (m.opt_disableflags & DSBL_ACTUATION) || c.P.disable_actuation;
  for (int i = c.lane; i < m.na; i += C::G) act_dot[i] = 0;
  if (off) {
    for (int i = c.lane; i < nu; i += C::G) force[i] = 0;
    for (int i = c.lane; i < nv; i += C::G) qa[i] = 0;
    SYNC();
    return;
  }
  bool bad = false;
  for (int i = c.lane; i < nu; i += C::G) {
    double x = ctrl[i];
    bad |= isnan(x) || fabs(x) > BP_MAXVAL;
  }
  if (group_any(c, bad)) {
    if (c.lane == 0) IF(warning)[W_BADCTRL]++;
    for (int i = c.lane; i < nu; i += C::G) ctrl[i] = 0;
  }
  SYNC();
#pragma unroll (C::UF)
  for (int i = c.lane; i < nu; i += C::G) {
    double x = ctrl[i];
    if (m.actuator_ctrllimited[i] && !(m.opt_disableflags & DSBL_CLAMPCTRL))
      x = clampd(x, m.actuator_ctrlrange[2 * i], m.actuator_ctrlrange[2 * i + 1]);
    double in = x;
    int dyn = C::LEAN ? 0 : m.actuator_dyntype[i];
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
    if (m.actuator_forcelimited[i])
      f = clampd(f, m.actuator_forcerange[2 * i], m.actuator_forcerange[2 * i + 1]);
    force[i] = f;
  }
  SYNC();
  // qfrc_actuator = moment^T force: each dof sums its own actuators in id order
#pragma unroll (C::UF)
  for (int d = c.lane; d < nv; d += C::G) {
    double s = 0;
    for (int k = m.dof_actadr[d]; k < m.dof_actadr[d + 1]; k++) {
      int u = m.dof_actlist[k];
      double cf = m.dof_actcoef[k];
      if constexpr (C::FULL) {
        // per-env tendon coefficients (point_mass.py:85-101): the host-derived coefficient
        // was built from the model's wrap_prm, rebuild it from this env's in the same order
        if (c.P.env_wrap_prm && m.actuator_trntype[u] == 3) {
          const int t = m.actuator_trnid[2 * u];
          cf = 0;
          for (int w = m.tendon_adr[t]; w < m.tendon_adr[t] + m.tendon_num[t]; w++)
            if (m.jnt_dofadr[m.wrap_objid[w]] == d) cf += m.actuator_gear[6 * u] * c.wprm[w];
        }
      }
      s += cf * force[u];
    }
    qa[d] = s;
  }
  SYNC();
}

BP_STAGE void bp_fwd_acceleration(const C& c) {
  BP_MODEL(m);
  auto fs = DF(qfrc_smooth), as = DF(qacc_smooth);
  for (int i = c.lane; i < m.nv; i += C::G) {
    double f = DF(qfrc_passive)[i] - DF(qfrc_bias)[i] + DF(qfrc_actuator)[i] +
               DF(qfrc_applied)[i];
    fs[i] = f; as[i] = f;
  }
  SYNC();
  bp_solve_sparse(c, DF(qLD), DF(qLDiagInv), as);
}

#endif  // BP_DYN_CUH_
