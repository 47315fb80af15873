// mj_makeConstraint + reference acceleration: rows in the order equality, friction loss,
// joint limits, contacts.  Row slots come from
// group prefix sums so row ids are reproducible; Jacobian entries are produced one
// (contact, dof) pair per lane using per-body dof-ancestry bit masks.
#ifndef BP_CONSTRAINT_CUH_
#define BP_CONSTRAINT_CUH_
#include "bp_math.cuh"

// one shared copy of pow(): the inlined expansion is ~200 instructions per call site
static __device__ __noinline__ double bp_pow(double x, double p) { return pow(x, p); }

static __device__ BP_STAGE_NOINLINE void bp_impedance(const double* si, double pos, double margin,
                                                 double* imp, double* impP) {
  if (si[0] == si[1] || si[2] <= BP_MINVAL) { *imp = 0.5 * (si[0] + si[1]); *impP = 0; return; }
  double x = (pos - margin) / si[2], sgn = 1;
  if (x < 0) { x = -x; sgn = -1; }
  if (x >= 1 || x <= 0) { *imp = x >= 1 ? si[1] : si[0]; *impP = 0; return; }
  double y, yP, p = si[4], mid = si[3];
  if (p == 1) { y = x; yP = 1; }
  else if (p == 2) {          // the default power: pow(x, 2) = x * x and pow(x, 1) = x exactly
    if (x <= mid) { double a = 1 / mid; y = a * (x * x); yP = p * a * x; }
    else { double b = 1 / (1 - mid), xm = 1 - x; y = 1 - b * (xm * xm); yP = p * b * xm; }
  } else if (x <= mid) {
    double a = 1 / bp_pow(mid, p - 1);
    y = a * bp_pow(x, p); yP = p * a * bp_pow(x, p - 1);
  } else {
    double b = 1 / bp_pow(1 - mid, p - 1);
    y = 1 - b * bp_pow(1 - x, p); yP = p * b * bp_pow(1 - x, p - 1);
  }
  *imp = si[0] + y * (si[1] - si[0]);
  *impP = yP * sgn * (si[1] - si[0]) / si[2];
}

template <class C>
__device__ __forceinline__ void bp_row_header(const C& c, int r, int type, int id,
                                              double pos, double margin, double floss) {
  IF(efc_type)[r] = type; IF(efc_id)[r] = id;
  DF(efc_pos)[r] = pos; DF(efc_margin)[r] = margin; DF(efc_frictionloss)[r] = floss;
}

BP_STAGE void bp_make_constraint(const C& c) {
  BP_MODEL(m);
  const int nv = m.nv, nvp = m.nvp, njmax = bp_lv(c).njmax;
  auto J = DF(efc_J);
  if (c.lane == 0) { IF(nefc)[0] = 0; IF(ne)[0] = 0; IF(nf)[0] = 0; }
  SYNC();
  if (m.opt_disableflags & DSBL_CONSTRAINT) return;
  int base = 0, nf = 0;
  // ---- pass 1: headers + slots -------------------------------------------------
  // equalities between joints / fixed tendons: (x1 - x1_0) - poly(x2 - x2_0)
  if constexpr (!C::LEAN) if (m.neq_active > 0 && !(m.opt_disableflags & DSBL_EQUALITY)) {
    for (int e = c.lane; e < m.neq; e += C::G) {
      if (!m.eq_active[e]) continue;
      int r = 0;
      for (int k = 0; k < e; k++) r += m.eq_active[k] != 0;
      const int o1 = m.eq_obj1id[e], o2 = m.eq_obj2id[e];
      const bool tend = m.eq_type[e] == 3;
      const double* cf = m.eq_data + 5 * e;
      double x1 = tend ? DF(ten_length)[o1] - m.tendon_length0[o1]
                       : DF(qpos)[m.jnt_qposadr[o1]] - m.qpos0[m.jnt_qposadr[o1]];
      double pos = x1 - cf[0];
      if (o2 >= 0) {
        double x2 = tend ? DF(ten_length)[o2] - m.tendon_length0[o2]
                         : DF(qpos)[m.jnt_qposadr[o2]] - m.qpos0[m.jnt_qposadr[o2]];
        pos = x1 - (cf[0] + x2 * (cf[1] + x2 * (cf[2] + x2 * (cf[3] + x2 * cf[4]))));
      }
      if (r < njmax) bp_row_header(c, r, CN_EQUALITY, e, pos, 0, 0);
    }
    base += m.neq_active;
  }
  const int eq_base = base;
  if (c.lane == 0) IF(ne)[0] = eq_base < njmax ? eq_base : njmax;
  if (!(m.opt_disableflags & DSBL_FRICTIONLOSS)) {
    for (int k = c.lane; k < m.nfric; k += C::G) {
      int r = base + k, i = m.fric_dof[k];
      if (r < njmax) bp_row_header(c, r, CN_FRICTION_DOF, i, 0, 0, m.dof_frictionloss[i]);
    }
    base += m.nfric;
    nf = (base < njmax ? base : njmax) - (eq_base < njmax ? eq_base : njmax);
  }
  int lim_base = base;
  if (!(m.opt_disableflags & DSBL_LIMIT)) {
#pragma unroll 1
    for (int k0 = 0; k0 < m.nlim; k0 += C::G) {
      int k = k0 + c.lane, n = 0, j = 0;
      double dist0 = 0, dist1 = 0, margin = 0;
      bool on0 = false, on1 = false;
      if (k < m.nlim) {
        j = m.lim_jnt[k];
        double q = DF(qpos)[m.jnt_qposadr[j]];
        margin = m.jnt_margin[j];
        dist0 = -(m.jnt_range[2 * j] - q);        // side = -1: lower bound
        dist1 = m.jnt_range[2 * j + 1] - q;       // side = +1: upper bound
        on0 = dist0 < margin; on1 = dist1 < margin;
        n = on0 + on1;
      }
      int tot, start = base + group_excl_scan(c, n, &tot);
      if (on0) {
        if (start < njmax) bp_row_header(c, start, CN_LIMIT_JOINT, j, dist0, margin, 0);
        start++;
      }
      if (on1 && start < njmax) bp_row_header(c, start, CN_LIMIT_JOINT, j, dist1, margin, 0);
      base += tot;
    }
  }
  // tendon limits, after the joint limits (tendons are few: one group pass each)
  const int ten_base = base;
  if constexpr (C::FULL) if (!(m.opt_disableflags & DSBL_LIMIT)) {
#pragma unroll 1
    for (int t0 = 0; t0 < m.ntendon; t0 += C::G) {
      int t = t0 + c.lane, n = 0;
      double dist0 = 0, dist1 = 0, margin = 0;
      bool on0 = false, on1 = false;
      if (t < m.ntendon && m.tendon_limited[t]) {
        double len = DF(ten_length)[t];
        margin = m.tendon_margin[t];
        dist0 = -(m.tendon_range[2 * t] - len);
        dist1 = m.tendon_range[2 * t + 1] - len;
        on0 = dist0 < margin; on1 = dist1 < margin;
        n = on0 + on1;
      }
      int tot, start = base + group_excl_scan(c, n, &tot);
      if (on0) {
        if (start < njmax) bp_row_header(c, start, CN_LIMIT_TENDON, t, dist0, margin, 0);
        start++;
      }
      if (on1 && start < njmax) bp_row_header(c, start, CN_LIMIT_TENDON, t, dist1, margin, 0);
      base += tot;
    }
  }
  int con_base = base, ncon = IF(ncon)[0], con_end = 0;
  const bool pyramidal = !C::FULL || m.opt_cone == 0;
  // a contact gets all of its rows or none (mj_addConstraint reserves the whole block);
  // the running sum is monotone, so nothing fits after the first contact that does not
#pragma unroll 1
  for (int c0 = 0; c0 < ncon; c0 += C::G) {
    int ci = c0 + c.lane, n = 0, dim = 0;
    if (ci < ncon && !IF(contact_exclude)[ci]) {
      dim = IF(contact_dim)[ci];
      n = dim == 1 ? 1 : (pyramidal ? 2 * (dim - 1) : dim);
    }
    int tot, start = base + group_excl_scan(c, n, &tot);
    if (ci < ncon) {
      const bool fits = n > 0 && start + n <= njmax;
      IF(contact_efc_address)[ci] = fits ? start : -1;
      int type = dim == 1 ? CN_CONTACT_FRICTIONLESS
                          : (pyramidal ? CN_CONTACT_PYRAMIDAL : CN_CONTACT_ELLIPTIC);
      double dist = DF(contact_dist)[ci], inc = DF(contact_includemargin)[ci];
      if (fits) {
        for (int k = 0; k < n; k++) bp_row_header(c, start + k, type, ci, dist, inc, 0);
        con_end = start + n;
      }
    }
    base += tot;
  }
#pragma unroll
  for (int o = C::G / 2; o > 0; o >>= 1) {
    int other = __shfl_xor_sync(c.mask, con_end, o, C::G);
    con_end = other > con_end ? other : con_end;
  }
  int nefc = con_base < njmax ? con_base : njmax;
  if (con_end > nefc) nefc = con_end;
  if (c.lane == 0) {
    IF(nefc)[0] = nefc; IF(nf)[0] = nf;
    if (base > njmax) IF(warning)[W_CNSTRFULL]++;
  }
  SYNC();
  // ---- pass 2: Jacobian ----------------------------------------------------------
  for (int w = c.lane; w < nefc * nvp; w += C::G) J[w] = 0;
  SYNC();
  if constexpr (!C::LEAN) for (int r = c.lane; r < eq_base && r < nefc; r += C::G) {
    const int e = IF(efc_id)[r], o1 = m.eq_obj1id[e], o2 = m.eq_obj2id[e];
    const bool tend = m.eq_type[e] == 3;
    const double* cf = m.eq_data + 5 * e;
    double deriv = 0;
    if (o2 >= 0) {
      double x2 = tend ? DF(ten_length)[o2] - m.tendon_length0[o2]
                       : DF(qpos)[m.jnt_qposadr[o2]] - m.qpos0[m.jnt_qposadr[o2]];
      deriv = cf[1] + x2 * (2 * cf[2] + x2 * (3 * cf[3] + x2 * 4 * cf[4]));
    }
    if (tend) {
      for (int w = m.tendon_adr[o1]; w < m.tendon_adr[o1] + m.tendon_num[o1]; w++)
        J[r * nvp + m.jnt_dofadr[m.wrap_objid[w]]] += c.wprm[w];
      if (o2 >= 0)
        for (int w = m.tendon_adr[o2]; w < m.tendon_adr[o2] + m.tendon_num[o2]; w++)
          J[r * nvp + m.jnt_dofadr[m.wrap_objid[w]]] -= deriv * c.wprm[w];
    } else {
      J[r * nvp + m.jnt_dofadr[o1]] += 1;
      if (o2 >= 0) J[r * nvp + m.jnt_dofadr[o2]] -= deriv;
    }
  }
  for (int r = eq_base + c.lane; r < lim_base && r < nefc; r += C::G)
    J[r * nvp + IF(efc_id)[r]] = 1;
  // limit rows: the Jacobian entry is -side; a joint with both sides active owns two
  // adjacent rows, lower bound first
  for (int r = lim_base + c.lane; r < ten_base && r < nefc; r += C::G) {
    int j = IF(efc_id)[r];
    double q = DF(qpos)[m.jnt_qposadr[j]];
    bool lower_on = -(m.jnt_range[2 * j] - q) < m.jnt_margin[j];
    bool first = (r == lim_base) || IF(efc_id)[r - 1] != j;
    int side = (lower_on && first) ? -1 : 1;
    J[r * nvp + m.jnt_dofadr[j]] = -side;
  }
  // tendon limit rows: -side * (tendon Jacobian)
  if constexpr (C::FULL) for (int r = ten_base + c.lane; r < con_base && r < nefc; r += C::G) {
    int t = IF(efc_id)[r];
    bool lower_on = -(m.tendon_range[2 * t] - DF(ten_length)[t]) < m.tendon_margin[t];
    bool first = (r == ten_base) || IF(efc_id)[r - 1] != t;
    double side = (lower_on && first) ? -1 : 1;
    if (m.tendon_Jadr[t] >= 0) {
      for (int i = 0; i < nv; i++) J[r * nvp + i] = -side * DF(ten_J)[m.tendon_Jadr[t] * nv + i];
    } else {
      for (int w = m.tendon_adr[t]; w < m.tendon_adr[t] + m.tendon_num[t]; w++)
        J[r * nvp + m.jnt_dofadr[m.wrap_objid[w]]] += -side * c.wprm[w];
    }
  }
  auto cdof = DF(cdof), com = DF(subtree_com);
#pragma unroll 1
  for (int w = c.lane; w < ncon * nv; w += C::G) {
    int ci = w / nv, i = w - ci * nv;
    int a = IF(contact_efc_address)[ci];
    if (a < 0) continue;
    int b1 = m.geom_bodyid[IF(contact_geom1)[ci]], b2 = m.geom_bodyid[IF(contact_geom2)[ci]];
    bool in1 = (m.body_dofmask[b1] >> i) & 1ull, in2 = (m.body_dofmask[b2] >> i) & 1ull;
    if (!in1 && !in2) continue;          // column stays zero
    double cd[6], pos[3];
    for (int k = 0; k < 6; k++) cd[k] = cdof[6 * i + k];
    for (int k = 0; k < 3; k++) pos[k] = DF(contact_pos)[3 * ci + k];
    double dp[3] = {0, 0, 0}, dr[3] = {0, 0, 0};
    if (in2) {
      auto cm = com + 3 * m.body_rootid[b2];
      double off[3] = {pos[0] - cm[0], pos[1] - cm[1], pos[2] - cm[2]}, t[3];
      cross3(t, cd, off);
      for (int k = 0; k < 3; k++) { dp[k] = cd[3 + k] + t[k]; dr[k] = cd[k]; }
    }
    if (in1) {
      auto cm = com + 3 * m.body_rootid[b1];
      double off[3] = {pos[0] - cm[0], pos[1] - cm[1], pos[2] - cm[2]}, t[3];
      cross3(t, cd, off);
      for (int k = 0; k < 3; k++) { dp[k] -= cd[3 + k] + t[k]; dr[k] -= cd[k]; }
    }
    auto fr = DF(contact_frame) + 9 * ci;
    double Jc[6];
    for (int k = 0; k < 3; k++) { Jc[k] = dot3(fr + 3 * k, dp); Jc[3 + k] = dot3(fr + 3 * k, dr); }
    int dim = IF(contact_dim)[ci];
    if (dim == 1) J[a * nvp + i] = Jc[0];
    else if (pyramidal) {
#pragma unroll 1
      for (int k = 1; k < dim; k++) {
        double mu = DF(contact_friction)[5 * ci + k - 1];
        int r = a + 2 * (k - 1);
        if (r < nefc) J[r * nvp + i] = Jc[0] + mu * Jc[k];
        if (r + 1 < nefc) J[(r + 1) * nvp + i] = Jc[0] - mu * Jc[k];
      }
    } else {
#pragma unroll 1
      for (int k = 0; k < dim; k++)
        if (a + k < nefc) J[(a + k) * nvp + i] = Jc[k];
    }
  }
  SYNC();
  // ---- pass 3: per-row solver parameters -----------------------------------------
  const double dt = m.opt_timestep;
#pragma unroll 1
  for (int r = c.lane; r < nefc; r += C::G) {
    int type = IF(efc_type)[r], id = IF(efc_id)[r];
    double solref[2], solimp[5], dA;
    if (!C::LEAN && type == CN_EQUALITY) {
      for (int k = 0; k < 2; k++) solref[k] = m.eq_solref[2 * id + k];
      for (int k = 0; k < 5; k++) solimp[k] = m.eq_solimp[5 * id + k];
      const int o1 = m.eq_obj1id[id], o2 = m.eq_obj2id[id];
      if (m.eq_type[id] == 3)
        dA = m.tendon_invweight0[o1] + (o2 >= 0 ? m.tendon_invweight0[o2] : 0.0);
      else
        dA = m.dof_invweight0[m.jnt_dofadr[o1]] +
             (o2 >= 0 ? m.dof_invweight0[m.jnt_dofadr[o2]] : 0.0);
    } else if (type == CN_FRICTION_DOF) {
      for (int k = 0; k < 2; k++) solref[k] = m.dof_solref[2 * id + k];
      for (int k = 0; k < 5; k++) solimp[k] = m.dof_solimp[5 * id + k];
      dA = m.dof_invweight0[id];
    } else if (type == CN_LIMIT_JOINT) {
      for (int k = 0; k < 2; k++) solref[k] = m.jnt_solref[2 * id + k];
      for (int k = 0; k < 5; k++) solimp[k] = m.jnt_solimp[5 * id + k];
      dA = m.dof_invweight0[m.jnt_dofadr[id]];
    } else if (C::FULL && type == CN_LIMIT_TENDON) {
      for (int k = 0; k < 2; k++) solref[k] = m.tendon_solref_lim[2 * id + k];
      for (int k = 0; k < 5; k++) solimp[k] = m.tendon_solimp_lim[5 * id + k];
      dA = m.tendon_invweight0[id];
    } else {
      for (int k = 0; k < 2; k++) solref[k] = DF(contact_solref)[2 * id + k];
      for (int k = 0; k < 5; k++) solimp[k] = DF(contact_solimp)[5 * id + k];
      int b1 = m.geom_bodyid[IF(contact_geom1)[id]], b2 = m.geom_bodyid[IF(contact_geom2)[id]];
      double tran = m.body_invweight0[2 * b1] + m.body_invweight0[2 * b2];
      double rot = m.body_invweight0[2 * b1 + 1] + m.body_invweight0[2 * b2 + 1];
      int k = r - IF(contact_efc_address)[id];
      if (type == CN_CONTACT_FRICTIONLESS) dA = tran;
      else if (type == CN_CONTACT_PYRAMIDAL) {
        double fri = DF(contact_friction)[5 * id + k / 2];
        dA = tran + fri * fri * (k < 4 ? tran : rot);
      } else dA = k < 3 ? tran : rot;
    }
    DF(efc_diagApprox)[r] = dA;
    double sr0 = solref[0], sr1 = solref[1];
    if (!(m.opt_disableflags & DSBL_REFSAFE) && sr0 > 0 && sr0 < 2 * dt) sr0 = 2 * dt;
    double imp, impP;
    bp_impedance(solimp, DF(efc_pos)[r], DF(efc_margin)[r], &imp, &impP);
    imp = clampd(imp, 0.0001, 0.9999);
    DF(efc_R)[r] = fmax(BP_MINVAL, (1 - imp) * dA / imp);
    double dmax = clampd(solimp[1], 0.0001, 0.9999), K, B;
    if (type == CN_FRICTION_DOF) { K = 0; B = 0; }
    else if (sr0 > 0) {
      K = 1 / fmax(BP_MINVAL, dmax * dmax * sr0 * sr0 * sr1 * sr1);
      B = 2 / fmax(BP_MINVAL, dmax * sr0);
    } else { K = -sr0 / fmax(BP_MINVAL, dmax * dmax); B = -sr1 / fmax(BP_MINVAL, dmax); }
    auto kb = DF(efc_KBIP) + 4 * r;
    kb[0] = K; kb[1] = B; kb[2] = imp; kb[3] = impP;
  }
  SYNC();
#pragma unroll 1
  for (int ci = c.lane; ci < ncon; ci += C::G) {
    int a = IF(contact_efc_address)[ci], dim = IF(contact_dim)[ci];
    if (a < 0 || dim == 1) continue;
    auto R = DF(efc_R);
    auto f = DF(contact_friction) + 5 * ci;
    if (pyramidal) {
      double mu = f[0] / sqrt(m.opt_impratio);
      DF(contact_mu)[ci] = mu;
      double Rpy = 2 * mu * mu * R[a];
      for (int k = 0; k < 2 * (dim - 1); k++) if (a + k < nefc) R[a + k] = Rpy;
    } else {
      if (a + 1 < nefc) R[a + 1] = R[a] / fmax(BP_MINVAL, m.opt_impratio);
      double R1 = a + 1 < nefc ? R[a + 1] : R[a];
      DF(contact_mu)[ci] = f[0] * sqrt(R1 / R[a]);
      for (int k = 2; k < dim; k++)
        if (a + k < nefc) R[a + k] = R1 * f[0] * f[0] / (f[k - 1] * f[k - 1]);
    }
  }
  SYNC();
  for (int r = c.lane; r < nefc; r += C::G) DF(efc_D)[r] = 1 / DF(efc_R)[r];
  SYNC();
}

// efc_vel = J qvel, efc_aref = -B vel - K imp (pos - margin)
BP_STAGE void bp_reference_constraint(const C& c) {
  BP_MODEL(m);
  const int nv = m.nv, nvp = m.nvp, nefc = IF(nefc)[0];
  auto J = DF(efc_J), qvel = DF(qvel);
#pragma unroll 1
  for (int r = c.lane; r < nefc; r += C::G) {
    double v = 0;
    for (int i = 0; i < nv; i++) v += J[r * nvp + i] * qvel[i];
    DF(efc_vel)[r] = v;
    auto kb = DF(efc_KBIP) + 4 * r;
    DF(efc_aref)[r] = -kb[1] * v - kb[0] * kb[2] * (DF(efc_pos)[r] - DF(efc_margin)[r]);
  }
  SYNC();
}

#endif  // BP_CONSTRAINT_CUH_
