// Position-stage smooth dynamics: kinematics, mj_comPos, mj_crb, sparse L^T D L
// factor/solve, for a group of G lanes per environment (bp_math.cuh).
//
// These are the stages the reference reaches through `mj_step1`
// (/root/reference/dm_control/mujoco/engine.py:166).  Tree
// recursions run level by level (bodies of equal depth on different lanes);
// backward accumulations gather children in descending id, i.e. the order of a
// serial loop over body ids, so sums are reproducible for every G.
#ifndef BP_SMOOTH_CUH_
#define BP_SMOOTH_CUH_
#include "bp_math.cuh"

BP_STAGE void bp_kinematics(const C& c) {
  BP_MODEL(m);
  auto qpos = DF(qpos), xpos = DF(xpos), xquat = DF(xquat), xmat = DF(xmat);
  auto xanchor = DF(xanchor), xaxis = DF(xaxis);
  for (int j = c.lane; j < m.njnt; j += C::G) {
    int jt = m.jnt_type[j];
    if (jt == JNT_FREE) normalize4(qpos + (m.jnt_qposadr[j] + 3));
    else if (jt == JNT_BALL) normalize4(qpos + m.jnt_qposadr[j]);
  }
  for (int i = c.lane; i < 9; i += C::G) {
    if (i < 3) xpos[i] = 0;
    if (i < 4) xquat[i] = i == 0 ? 1.0 : 0.0;
    xmat[i] = (i % 4 == 0) ? 1.0 : 0.0;
  }
  SYNC();
#pragma unroll (C::UF)
  for (int lev = 1; lev < m.nlevel; lev++) {
#pragma unroll (C::UF)
    for (int idx = m.body_level_adr[lev] + c.lane; idx < m.body_level_adr[lev + 1];
         idx += C::G) {
      int b = m.body_level_list[idx], p = m.body_parentid[b];
      double pos[3], quat[4], R[9], t[3];
      mat_vec3(t, xmat + 9 * p, (C::FULL ? c.bpos : (const double*)m.body_pos) + 3 * b);
      for (int k = 0; k < 3; k++) pos[k] = xpos[3 * p + k] + t[k];
      quat_mul(quat, xquat + 4 * p, (C::FULL ? c.bquat : (const double*)m.body_quat) + 4 * b);
      int jn = m.body_jntnum[b], j0 = m.body_jntadr[b];
#pragma unroll (C::UF)
      for (int k = 0; k < jn; k++) {
        int j = j0 + k, qa = m.jnt_qposadr[j], jt = m.jnt_type[j];
        if (jt == JNT_FREE) {
          for (int i = 0; i < 3; i++) { pos[i] = qpos[qa + i]; xanchor[3 * j + i] = pos[i]; }
          for (int i = 0; i < 4; i++) quat[i] = qpos[qa + 3 + i];
          xaxis[3 * j] = 0; xaxis[3 * j + 1] = 0; xaxis[3 * j + 2] = 1;
          continue;
        }
        quat2mat(R, quat);
        mat_vec3(t, R, m.jnt_pos + 3 * j);
        double anchor[3], axis[3];
        for (int i = 0; i < 3; i++) { anchor[i] = pos[i] + t[i]; xanchor[3 * j + i] = anchor[i]; }
        mat_vec3(axis, R, m.jnt_axis + 3 * j);
        for (int i = 0; i < 3; i++) xaxis[3 * j + i] = axis[i];
        if (jt == JNT_SLIDE) {
          double dq = qpos[qa] - m.qpos0[qa];
          for (int i = 0; i < 3; i++) pos[i] += axis[i] * dq;
        } else {
          double qr[4], qn[4];
          if (jt == JNT_HINGE) axis_angle_quat(qr, m.jnt_axis + 3 * j, qpos[qa] - m.qpos0[qa]);
          else for (int i = 0; i < 4; i++) qr[i] = qpos[qa + i];
          quat_mul(qn, quat, qr);
          for (int i = 0; i < 4; i++) quat[i] = qn[i];
          quat2mat(R, quat);
          mat_vec3(t, R, m.jnt_pos + 3 * j);
          for (int i = 0; i < 3; i++) pos[i] = anchor[i] - t[i];
        }
      }
      normalize4(quat);
      for (int i = 0; i < 3; i++) xpos[3 * b + i] = pos[i];
      for (int i = 0; i < 4; i++) xquat[4 * b + i] = quat[i];
      quat2mat(R, quat);
      for (int i = 0; i < 9; i++) xmat[9 * b + i] = R[i];
    }
    SYNC();
  }
  // inertial frames, geoms and sites hang off finished body frames: one flat loop
  int nb = m.nbody, ng = m.ngeom, ns = m.nsite;
#pragma unroll (C::UF)
  for (int w = c.lane; w < nb + ng + ns; w += C::G) {
    const double *lp, *lq;
    typename C::DP op, om;
    int b;
    if (w < nb) {
      b = w; lp = m.body_ipos + 3 * w; lq = m.body_iquat + 4 * w;
      op = DF(xipos) + 3 * w; om = DF(ximat) + 9 * w;
    } else if (w < nb + ng) {
      int g = w - nb;
      b = m.geom_bodyid[g]; lp = (C::FULL ? c.gpos : (const double*)m.geom_pos) + 3 * g; lq = m.geom_quat + 4 * g;
      op = DF(geom_xpos) + 3 * g; om = DF(geom_xmat) + 9 * g;
    } else {
      int s = w - nb - ng;
      b = m.site_bodyid[s]; lp = (C::FULL ? c.spos : (const double*)m.site_pos) + 3 * s; lq = m.site_quat + 4 * s;
      op = DF(site_xpos) + 3 * s; om = DF(site_xmat) + 9 * s;
    }
    double t[3], q[4], R[9];
    mat_vec3(t, xmat + 9 * b, lp);
    for (int i = 0; i < 3; i++) op[i] = xpos[3 * b + i] + t[i];
    quat_mul(q, xquat + 4 * b, lq);
    quat2mat(R, q);
    for (int i = 0; i < 9; i++) om[i] = R[i];
  }
  SYNC();
}

// vals[b*n .. b*n+n) += sum over children, from the deepest level up to `stop_level`
// (stop_level = 0 also accumulates into the world body, 1 leaves it untouched)
template <int n, class C>
__device__ BP_STAGE_NOINLINE void bp_tree_gather(const C& c, typename C::DP vals, int stop_level) {
  BP_MODEL(m);
#pragma unroll (C::UF)
  for (int lev = m.nlevel - 2; lev >= stop_level; lev--) {
    int a0 = m.body_level_adr[lev], cnt = m.body_level_adr[lev + 1] - a0;
#pragma unroll (C::UF)
    for (int w = c.lane; w < cnt * n; w += C::G) {
      int b = m.body_level_list[a0 + w / n], k = w % n;
      int c0 = m.body_child_adr[b], c1 = m.body_child_adr[b + 1];
      if (c1 > c0) {
        double s = vals[b * n + k];
#pragma unroll (C::UF)
        for (int ci = c0; ci < c1; ci++) s += vals[m.body_child_list[ci] * n + k];
        vals[b * n + k] = s;
      }
    }
    SYNC();
  }
}

BP_STAGE void bp_com_pos(const C& c) {
  BP_MODEL(m);
  int nb = m.nbody;
  auto com = DF(subtree_com), xipos = DF(xipos);
  for (int w = c.lane; w < 3 * nb; w += C::G) com[w] = m.body_mass[w / 3] * xipos[w];
  SYNC();
  bp_tree_gather<3>(c, com, 0);
  for (int w = c.lane; w < 3 * nb; w += C::G) {
    double sm = m.body_subtreemass[w / 3];
    com[w] = sm < BP_MINVAL ? xipos[w] : com[w] / sm;
  }
  SYNC();
  auto cinert = DF(cinert), ximat = DF(ximat);
#pragma unroll (C::UF)
  for (int b = 1 + c.lane; b < nb; b += C::G) {
    auto R = ximat + 9 * b;
    const double* I = m.body_inertia + 3 * b;
    auto cm = com + 3 * m.body_rootid[b];
    double mass = m.body_mass[b], dif[3], T[9], Rl[9];
    for (int k = 0; k < 9; k++) Rl[k] = R[k];
    for (int k = 0; k < 3; k++) dif[k] = xipos[3 * b + k] - cm[k];
    for (int r = 0; r < 3; r++)
      for (int k = r; k < 3; k++)
        T[3 * r + k] = Rl[3 * r] * I[0] * Rl[3 * k] + Rl[3 * r + 1] * I[1] * Rl[3 * k + 1] +
                       Rl[3 * r + 2] * I[2] * Rl[3 * k + 2];
    auto ci = cinert + 10 * b;
    ci[0] = T[0] + mass * (dif[1] * dif[1] + dif[2] * dif[2]);
    ci[1] = T[4] + mass * (dif[0] * dif[0] + dif[2] * dif[2]);
    ci[2] = T[8] + mass * (dif[0] * dif[0] + dif[1] * dif[1]);
    ci[3] = T[1] - mass * dif[0] * dif[1];
    ci[4] = T[2] - mass * dif[0] * dif[2];
    ci[5] = T[5] - mass * dif[1] * dif[2];
    ci[6] = mass * dif[0]; ci[7] = mass * dif[1]; ci[8] = mass * dif[2];
    ci[9] = mass;
  }
  for (int i = c.lane; i < 10; i += C::G) cinert[i] = 0;
  auto cdof = DF(cdof), xanchor = DF(xanchor), xaxis = DF(xaxis), xmat = DF(xmat);
#pragma unroll (C::UF)
  for (int j = c.lane; j < m.njnt; j += C::G) {
    int b = m.jnt_bodyid[j], da = m.jnt_dofadr[j], jt = m.jnt_type[j];
    auto cm = com + 3 * m.body_rootid[b];
    double off[3];
    for (int k = 0; k < 3; k++) off[k] = cm[k] - xanchor[3 * j + k];
    if (jt == JNT_SLIDE) {
      auto cd = cdof + 6 * da;
      for (int k = 0; k < 3; k++) { cd[k] = 0; cd[3 + k] = xaxis[3 * j + k]; }
    } else if (jt == JNT_HINGE) {
      auto cd = cdof + 6 * da;
      double ax[3] = {xaxis[3 * j], xaxis[3 * j + 1], xaxis[3 * j + 2]}, t[3];
      cross3(t, ax, off);
      for (int k = 0; k < 3; k++) { cd[k] = ax[k]; cd[3 + k] = t[k]; }
    } else {
      if (jt == JNT_FREE) {
        for (int k = 0; k < 3; k++) {
          auto cd = cdof + 6 * (da + k);
          for (int i = 0; i < 6; i++) cd[i] = (i == 3 + k) ? 1.0 : 0.0;
        }
        da += 3;
      }
      auto R = xmat + 9 * b;
      for (int k = 0; k < 3; k++) {
        auto cd = cdof + 6 * (da + k);
        double ax[3] = {R[k], R[3 + k], R[6 + k]}, t[3];
        cross3(t, ax, off);
        for (int i = 0; i < 3; i++) { cd[i] = ax[i]; cd[3 + i] = t[i]; }
      }
    }
  }
  SYNC();
}

BP_STAGE void bp_crb(const C& c) {
  BP_MODEL(m);
  auto crb = DF(crb), cinert = DF(cinert), cdof = DF(cdof), qM = DF(qM);
  for (int w = c.lane; w < 10 * m.nbody; w += C::G) crb[w] = cinert[w];
  SYNC();
  bp_tree_gather<10>(c, crb, 1);
#pragma unroll (C::UF)
  for (int i = c.lane; i < m.nv; i += C::G) {
    double buf[6];
    inert_mul(buf, crb + 10 * m.dof_bodyid[i], cdof + 6 * i);
    int adr = m.dof_Madr[i];
    qM[adr++] = m.dof_armature[i] + dot6(cdof + 6 * i, buf);
#pragma unroll (C::UF)
    for (int j = m.dof_parentid[i]; j >= 0; j = m.dof_parentid[j])
      qM[adr++] = dot6(cdof + 6 * j, buf);
  }
  SYNC();
}

// in-place sparse L^T D L of a matrix in qM layout; row k's ancestors on lanes.  A second
// matrix with the same layout (M + dt * damping for the implicit Euler step) can ride
// along: it shares every index load and barrier of the first.
BP_STAGE void bp_factor_sparse(const C& c, typename C::DP LD, typename C::DP diaginv,
                               typename C::DP LD2, typename C::DP diaginv2) {
  BP_MODEL(m);
  const bool two = LD2.valid();
#pragma unroll (C::UF)
  for (int k = m.nv - 1; k >= 0; k--) {
    int akk = m.dof_Madr[k], na = m.dof_depth[k];
    double dkk = LD[akk], dkk2 = two ? LD2[akk] : 1.0;
#pragma unroll (C::UF)
    for (int a = 1 + c.lane; a <= na; a += C::G) {
      double t = LD[akk + a] / dkk;
      int i = m.M_col[akk + a];
      int aij = m.dof_Madr[i];
      if (two) {
        double t2 = LD2[akk + a] / dkk2;
#pragma unroll (C::UF)
        for (int b = a; b <= na; b++, aij++) {
          LD[aij] -= t * LD[akk + b];
          LD2[aij] -= t2 * LD2[akk + b];
        }
      } else {
#pragma unroll (C::UF)
        for (int b = a; b <= na; b++) LD[aij++] -= t * LD[akk + b];
      }
    }
    SYNC();
    for (int a = 1 + c.lane; a <= na; a += C::G) {
      LD[akk + a] = LD[akk + a] / dkk;
      if (two) LD2[akk + a] = LD2[akk + a] / dkk2;
    }
    if (c.lane == 0) {
      diaginv[k] = 1.0 / dkk;
      if (two) diaginv2[k] = 1.0 / dkk2;
    }
    SYNC();
  }
}

// x <- M^-1 x with the factor above
BP_STAGE void bp_solve_sparse(const C& c, typename C::DP LD, typename C::DP diaginv,
                              typename C::DP x) {
  BP_MODEL(m);
  int nv = m.nv;
#pragma unroll (C::UF)
  for (int i = nv - 1; i >= 0; i--) {          // x <- L^-T x
    int a0 = m.dof_Madr[i] + 1, na = m.dof_depth[i];
    double xi = x[i];
    for (int a = c.lane; a < na; a += C::G) x[m.M_col[a0 + a]] -= LD[a0 + a] * xi;
    SYNC();
  }
  for (int i = c.lane; i < nv; i += C::G) x[i] *= diaginv[i];
  SYNC();
#pragma unroll (C::UF)
  for (int lev = 1; lev < m.ndoflevel; lev++) { // x <- L^-1 x, depth by depth
#pragma unroll (C::UF)
    for (int idx = m.dof_level_adr[lev] + c.lane; idx < m.dof_level_adr[lev + 1];
         idx += C::G) {
      int i = m.dof_level_list[idx], a = m.dof_Madr[i] + 1;
      double s = x[i];
      for (int j = m.dof_parentid[i]; j >= 0; j = m.dof_parentid[j]) s -= LD[a++] * x[j];
      x[i] = s;
    }
    SYNC();
  }
}

#endif  // BP_SMOOTH_CUH_
