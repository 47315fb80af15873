// Position-stage smooth dynamics, one warp per environment, state in shared memory:
// kinematics, mj_comPos, mj_crb, sparse L^T D L factor/solve.
//
// These are the stages the reference reaches through `mj_step1`
// (/root/reference/dm_control/mujoco/engine.py:166).  Tree
// recursions run level by level (bodies of equal depth on different lanes);
// backward accumulations gather children in descending id, i.e. the order of a
// serial loop over body ids, so sums are reproducible.
#ifndef BP_SMOOTH_CUH_
#define BP_SMOOTH_CUH_
#include "bp_math.cuh"

__device__ void bp_kinematics(const Ctx& c) {
  const BpModel& m = c.P.m;
  double *qpos = DF(qpos), *xpos = DF(xpos), *xquat = DF(xquat), *xmat = DF(xmat);
  double *xanchor = DF(xanchor), *xaxis = DF(xaxis);
  for (int j = c.lane; j < m.njnt; j += 32) {
    int jt = m.jnt_type[j];
    if (jt == JNT_FREE) normalize4(qpos + m.jnt_qposadr[j] + 3);
    else if (jt == JNT_BALL) normalize4(qpos + m.jnt_qposadr[j]);
  }
  if (c.lane < 3) xpos[c.lane] = 0;
  if (c.lane < 4) xquat[c.lane] = c.lane == 0 ? 1.0 : 0.0;
  if (c.lane < 9) xmat[c.lane] = (c.lane % 4 == 0) ? 1.0 : 0.0;
  SYNC();
  for (int lev = 1; lev < m.nlevel; lev++) {
    for (int idx = m.body_level_adr[lev] + c.lane; idx < m.body_level_adr[lev + 1];
         idx += 32) {
      int b = m.body_level_list[idx], p = m.body_parentid[b];
      double pos[3], quat[4], R[9], t[3];
      mat_vec3(t, xmat + 9 * p, m.body_pos + 3 * b);
      for (int k = 0; k < 3; k++) pos[k] = xpos[3 * p + k] + t[k];
      quat_mul(quat, xquat + 4 * p, m.body_quat + 4 * b);
      int jn = m.body_jntnum[b], j0 = m.body_jntadr[b];
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
  for (int w = c.lane; w < nb + ng + ns; w += 32) {
    const double *lp, *lq;
    double *op, *om;
    int b;
    if (w < nb) {
      b = w; lp = m.body_ipos + 3 * w; lq = m.body_iquat + 4 * w;
      op = DF(xipos) + 3 * w; om = DF(ximat) + 9 * w;
    } else if (w < nb + ng) {
      int g = w - nb;
      b = m.geom_bodyid[g]; lp = m.geom_pos + 3 * g; lq = m.geom_quat + 4 * g;
      op = DF(geom_xpos) + 3 * g; om = DF(geom_xmat) + 9 * g;
    } else {
      int s = w - nb - ng;
      b = m.site_bodyid[s]; lp = m.site_pos + 3 * s; lq = m.site_quat + 4 * s;
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
__device__ __forceinline__ void bp_tree_gather(const Ctx& c, double* vals, int n,
                                               int stop_level) {
  const BpModel& m = c.P.m;
  for (int lev = m.nlevel - 2; lev >= stop_level; lev--) {
    int a0 = m.body_level_adr[lev], cnt = m.body_level_adr[lev + 1] - a0;
    for (int w = c.lane; w < cnt * n; w += 32) {
      int b = m.body_level_list[a0 + w / n], k = w % n;
      int c0 = m.body_child_adr[b], c1 = m.body_child_adr[b + 1];
      if (c1 > c0) {
        double s = vals[b * n + k];
        for (int ci = c0; ci < c1; ci++) s += vals[m.body_child_list[ci] * n + k];
        vals[b * n + k] = s;
      }
    }
    SYNC();
  }
}

__device__ void bp_com_pos(const Ctx& c) {
  const BpModel& m = c.P.m;
  int nb = m.nbody;
  double *com = DF(subtree_com), *xipos = DF(xipos);
  for (int w = c.lane; w < 3 * nb; w += 32) com[w] = m.body_mass[w / 3] * xipos[w];
  SYNC();
  bp_tree_gather(c, com, 3, 0);
  for (int w = c.lane; w < 3 * nb; w += 32) {
    double sm = m.body_subtreemass[w / 3];
    com[w] = sm < BP_MINVAL ? xipos[w] : com[w] / sm;
  }
  SYNC();
  double *cinert = DF(cinert), *ximat = DF(ximat);
  for (int b = 1 + c.lane; b < nb; b += 32) {
    const double* R = ximat + 9 * b;
    const double* I = m.body_inertia + 3 * b;
    const double* cm = com + 3 * m.body_rootid[b];
    double mass = m.body_mass[b], dif[3], T[9];
    for (int k = 0; k < 3; k++) dif[k] = xipos[3 * b + k] - cm[k];
    for (int r = 0; r < 3; r++)
      for (int k = r; k < 3; k++)
        T[3 * r + k] = R[3 * r] * I[0] * R[3 * k] + R[3 * r + 1] * I[1] * R[3 * k + 1] +
                       R[3 * r + 2] * I[2] * R[3 * k + 2];
    double* ci = cinert + 10 * b;
    ci[0] = T[0] + mass * (dif[1] * dif[1] + dif[2] * dif[2]);
    ci[1] = T[4] + mass * (dif[0] * dif[0] + dif[2] * dif[2]);
    ci[2] = T[8] + mass * (dif[0] * dif[0] + dif[1] * dif[1]);
    ci[3] = T[1] - mass * dif[0] * dif[1];
    ci[4] = T[2] - mass * dif[0] * dif[2];
    ci[5] = T[5] - mass * dif[1] * dif[2];
    ci[6] = mass * dif[0]; ci[7] = mass * dif[1]; ci[8] = mass * dif[2];
    ci[9] = mass;
  }
  if (c.lane < 10) cinert[c.lane] = 0;
  double *cdof = DF(cdof), *xanchor = DF(xanchor), *xaxis = DF(xaxis), *xmat = DF(xmat);
  for (int j = c.lane; j < m.njnt; j += 32) {
    int b = m.jnt_bodyid[j], da = m.jnt_dofadr[j], jt = m.jnt_type[j];
    const double* cm = com + 3 * m.body_rootid[b];
    double off[3];
    for (int k = 0; k < 3; k++) off[k] = cm[k] - xanchor[3 * j + k];
    if (jt == JNT_SLIDE) {
      double* cd = cdof + 6 * da;
      for (int k = 0; k < 3; k++) { cd[k] = 0; cd[3 + k] = xaxis[3 * j + k]; }
    } else if (jt == JNT_HINGE) {
      double* cd = cdof + 6 * da;
      double ax[3] = {xaxis[3 * j], xaxis[3 * j + 1], xaxis[3 * j + 2]}, t[3];
      cross3(t, ax, off);
      for (int k = 0; k < 3; k++) { cd[k] = ax[k]; cd[3 + k] = t[k]; }
    } else {
      if (jt == JNT_FREE) {
        for (int k = 0; k < 3; k++) {
          double* cd = cdof + 6 * (da + k);
          for (int i = 0; i < 6; i++) cd[i] = (i == 3 + k) ? 1.0 : 0.0;
        }
        da += 3;
      }
      const double* R = xmat + 9 * b;
      for (int k = 0; k < 3; k++) {
        double* cd = cdof + 6 * (da + k);
        double ax[3] = {R[k], R[3 + k], R[6 + k]}, t[3];
        cross3(t, ax, off);
        for (int i = 0; i < 3; i++) { cd[i] = ax[i]; cd[3 + i] = t[i]; }
      }
    }
  }
  SYNC();
}

__device__ void bp_crb(const Ctx& c) {
  const BpModel& m = c.P.m;
  double *crb = DF(crb), *cinert = DF(cinert), *cdof = DF(cdof), *qM = DF(qM);
  for (int w = c.lane; w < 10 * m.nbody; w += 32) crb[w] = cinert[w];
  SYNC();
  bp_tree_gather(c, crb, 10, 1);
  for (int i = c.lane; i < m.nv; i += 32) {
    double buf[6];
    inert_mul(buf, crb + 10 * m.dof_bodyid[i], cdof + 6 * i);
    int adr = m.dof_Madr[i];
    qM[adr++] = m.dof_armature[i] + dot6(cdof + 6 * i, buf);
    for (int j = m.dof_parentid[i]; j >= 0; j = m.dof_parentid[j])
      qM[adr++] = dot6(cdof + 6 * j, buf);
  }
  SYNC();
}

// in-place sparse L^T D L of a matrix in qM layout; row k's ancestors on lanes
__device__ void bp_factor_sparse(const Ctx& c, double* LD, double* diaginv) {
  const BpModel& m = c.P.m;
  for (int k = m.nv - 1; k >= 0; k--) {
    int akk = m.dof_Madr[k], na = m.dof_depth[k];
    double dkk = LD[akk];
    // lane a handles ancestor a (1-based position in row k)
    double t = 0;
    for (int a = 1 + c.lane; a <= na; a += 32) {
      t = LD[akk + a] / dkk;
      int i = m.M_col[akk + a];
      int aij = m.dof_Madr[i];
      for (int b = a; b <= na; b++) LD[aij++] -= t * LD[akk + b];
    }
    SYNC();
    for (int a = 1 + c.lane; a <= na; a += 32) LD[akk + a] = LD[akk + a] / dkk;
    if (c.lane == 0) diaginv[k] = 1.0 / dkk;
    SYNC();
  }
}

// x <- M^-1 x with the factor above
__device__ void bp_solve_sparse(const Ctx& c, const double* LD, const double* diaginv,
                                double* x) {
  const BpModel& m = c.P.m;
  int nv = m.nv;
  for (int i = nv - 1; i >= 0; i--) {          // x <- L^-T x
    int a0 = m.dof_Madr[i] + 1, na = m.dof_depth[i];
    double xi = x[i];
    for (int a = c.lane; a < na; a += 32) x[m.M_col[a0 + a]] -= LD[a0 + a] * xi;
    SYNC();
  }
  for (int i = c.lane; i < nv; i += 32) x[i] *= diaginv[i];
  SYNC();
  for (int lev = 1; lev < m.ndoflevel; lev++) { // x <- L^-1 x, depth by depth
    for (int idx = m.dof_level_adr[lev] + c.lane; idx < m.dof_level_adr[lev + 1];
         idx += 32) {
      int i = m.dof_level_list[idx], a = m.dof_Madr[i] + 1;
      double s = x[i];
      for (int j = m.dof_parentid[i]; j >= 0; j = m.dof_parentid[j]) s -= LD[a++] * x[j];
      x[i] = s;
    }
    SYNC();
  }
}

#endif  // BP_SMOOTH_CUH_
