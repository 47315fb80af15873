// mj_fwdConstraint: primal Newton solver in qacc with exact line search
// (SURVEY.md §8a row a2, Appendix A.9).  One warp per environment:
//   rows of J on lanes for J*x products, cost/line-search sums by warp shuffles,
//   Hessian M + J^T D J one lower-triangle entry per lane, dense Cholesky with
//   rows on lanes.
#ifndef BP_SOLVER_CUH_
#define BP_SOLVER_CUH_
#include "bp_math.cuh"

// y = Md * x, dense symmetric nv x nv with row stride nvp (odd => conflict free)
__device__ __forceinline__ void bp_mul_md(const Ctx& c, const double* Md, double* y,
                                          const double* x) {
  const int nv = c.P.m.nv, nvp = c.P.m.nvp;
  for (int i = c.lane; i < nv; i += 32) {
    double s = 0;
    for (int j = 0; j < nv; j++) s += Md[i * nvp + j] * x[j];
    y[i] = s;
  }
  SYNC();
}

// out[r] = J[r,:] x - (sub ? sub[r] : 0)
__device__ __forceinline__ void bp_mul_j(const Ctx& c, double* out, const double* x,
                                         const double* sub, int nefc) {
  const int nv = c.P.m.nv, nvp = c.P.m.nvp;
  const double* J = DF(efc_J);
  for (int r = c.lane; r < nefc; r += 32) {
    double s = 0;
    for (int i = 0; i < nv; i++) s += J[r * nvp + i] * x[i];
    out[r] = sub ? s - sub[r] : s;
  }
  SYNC();
}

// constraint cost at residual jar; optionally writes force/state and J^T force
__device__ double bp_update_constraint(const Ctx& c, const double* jar, int nefc,
                                       bool write) {
  const int nv = c.P.m.nv, nvp = c.P.m.nvp;
  double cost = 0;
  double *force = DF(efc_force), *D = DF(efc_D), *R = DF(efc_R);
  int *type = IF(efc_type), *state = IF(efc_state);
  for (int r = c.lane; r < nefc; r += 32) {
    double d = D[r], x = jar[r], f, cr = 0;
    int t = type[r], st;
    if (t == CN_EQUALITY) { f = -d * x; st = ST_QUADRATIC; cr = 0.5 * d * x * x; }
    else if (t == CN_FRICTION_DOF || t == CN_FRICTION_TENDON) {
      double fl = DF(efc_frictionloss)[r], rr = R[r];
      if (x <= -rr * fl) { f = fl; st = ST_LINEARNEG; cr = -fl * (0.5 * rr * fl + x); }
      else if (x >= rr * fl) { f = -fl; st = ST_LINEARPOS; cr = -fl * (0.5 * rr * fl - x); }
      else { f = -d * x; st = ST_QUADRATIC; cr = 0.5 * d * x * x; }
    } else {
      if (x >= 0) { f = 0; st = ST_SATISFIED; }
      else { f = -d * x; st = ST_QUADRATIC; cr = 0.5 * d * x * x; }
    }
    cost += cr;
    if (write) { force[r] = f; state[r] = st; }
  }
  cost = warp_sum(cost);
  if (write) {
    SYNC();
    const double* J = DF(efc_J);
    double* qc = DF(qfrc_constraint);
    for (int i = c.lane; i < nv; i += 32) {
      double s = 0;
      for (int r = 0; r < nefc; r++) s += J[r * nvp + i] * force[r];
      qc[i] = s;
    }
    SYNC();
  }
  return cost;
}

__device__ __forceinline__ double bp_gauss_cost(const Ctx& c) {
  double g = 0;
  for (int i = c.lane; i < c.P.m.nv; i += 32)
    g += 0.5 * (DF(s_Ma)[i] - DF(qfrc_smooth)[i]) * (DF(qacc)[i] - DF(qacc_smooth)[i]);
  return warp_sum(g);
}

struct LSPoint { double cost, d0, d1; };

__device__ LSPoint bp_ls_eval(const Ctx& c, double alpha, double g0, double g1, double g2,
                              int nefc) {
  double c0 = 0, c1 = 0, c2 = 0;
  const double *D = DF(efc_D), *jar = DF(s_jar), *jvv = DF(s_jv);
  const int* type = IF(efc_type);
  for (int r = c.lane; r < nefc; r += 32) {
    double d = D[r], j0 = jar[r], jv = jvv[r], x = j0 + alpha * jv;
    int t = type[r];
    double q0 = 0.5 * d * j0 * j0, q1 = d * j0 * jv, q2 = 0.5 * d * jv * jv;
    if (t == CN_EQUALITY) { c0 += q0; c1 += q1; c2 += q2; }
    else if (t == CN_FRICTION_DOF || t == CN_FRICTION_TENDON) {
      double f = DF(efc_frictionloss)[r], rr = DF(efc_R)[r];
      if (x <= -rr * f) { c0 += -f * (0.5 * rr * f + j0); c1 += -f * jv; }
      else if (x >= rr * f) { c0 += -f * (0.5 * rr * f - j0); c1 += f * jv; }
      else { c0 += q0; c1 += q1; c2 += q2; }
    } else if (x < 0) { c0 += q0; c1 += q1; c2 += q2; }
  }
  c0 = g0 + warp_sum(c0); c1 = g1 + warp_sum(c1); c2 = g2 + warp_sum(c2);
  LSPoint p;
  p.cost = alpha * alpha * c2 + alpha * c1 + c0;
  p.d0 = 2 * alpha * c2 + c1;
  p.d1 = 2 * c2;
  return p;
}

// exact 1-D minimisation along s_search (safeguarded Newton on the derivative)
__device__ double bp_linesearch(const Ctx& c, double scale, int nefc) {
  const BpModel& m = c.P.m;
  double sn = 0, g0 = 0, g1 = 0, g2 = 0;
  for (int i = c.lane; i < m.nv; i += 32) {
    double s = DF(s_search)[i], r = DF(s_Ma)[i] - DF(qfrc_smooth)[i];
    sn += s * s;
    g0 += 0.5 * r * (DF(qacc)[i] - DF(qacc_smooth)[i]);
    g1 += s * r;
    g2 += 0.5 * s * DF(s_Mv)[i];
  }
  sn = sqrt(warp_sum(sn)); g0 = warp_sum(g0); g1 = warp_sum(g1); g2 = warp_sum(g2);
  if (sn < BP_MINVAL) return 0;
  double gtol = m.opt_tolerance * m.opt_ls_tolerance * sn / scale;
  LSPoint p = bp_ls_eval(c, 0, g0, g1, g2, nefc);
  if (p.d1 < BP_MINVAL) return 0;
  double lo = 0, hi = 0, a = -p.d0 / p.d1;
  bool have_hi = false;
  for (int it = 0; it < m.opt_ls_iterations; it++) {
    p = bp_ls_eval(c, a, g0, g1, g2, nefc);
    if (fabs(p.d0) < gtol) break;
    if (p.d0 < 0) lo = a; else { hi = a; have_hi = true; }
    double an = a - p.d0 / p.d1;
    if (have_hi && !(an > lo && an < hi)) an = 0.5 * (lo + hi);
    a = an;
  }
  return a;
}

// dense Cholesky H = L L^T in place (lower); rows on lanes, right-looking.
// Returns false when a pivot is not positive.
__device__ bool bp_chol_factor(const Ctx& c, double* H, double* dinv) {
  const int n = c.P.m.nv, nvp = c.P.m.nvp;
  for (int j = 0; j < n; j++) {
    double s = H[j * nvp + j];
    if (s < BP_MINVAL) return false;      // uniform across the warp
    s = sqrt(s);
    double inv = 1.0 / s;
    SYNC();
    if (c.lane == 0) { H[j * nvp + j] = s; dinv[j] = inv; }
    for (int i = j + 1 + c.lane; i < n; i += 32) H[i * nvp + j] /= s;
    SYNC();
    for (int i = j + 1 + c.lane; i < n; i += 32) {
      double lij = H[i * nvp + j];
      for (int k = j + 1; k <= i; k++) H[i * nvp + k] -= lij * H[k * nvp + j];
    }
    SYNC();
  }
  return true;
}

__device__ void bp_chol_solve(const Ctx& c, const double* L, const double* dinv, double* x) {
  const int n = c.P.m.nv, nvp = c.P.m.nvp;
  for (int k = 0; k < n; k++) {
    double xk = x[k] * dinv[k];
    SYNC();
    if (c.lane == 0) x[k] = xk;
    for (int i = k + 1 + c.lane; i < n; i += 32) x[i] -= L[i * nvp + k] * xk;
    SYNC();
  }
  for (int k = n - 1; k >= 0; k--) {
    double xk = x[k] * dinv[k];
    SYNC();
    if (c.lane == 0) x[k] = xk;
    for (int i = c.lane; i < k; i += 32) x[i] -= L[k * nvp + i] * xk;
    SYNC();
  }
}

// grad = Ma - qfrc_smooth - qfrc_constraint;  H = M + J^T D_active J;  Mgrad = H^-1 grad
__device__ void bp_newton_gradient(const Ctx& c, int nefc) {
  const BpModel& m = c.P.m;
  const int nv = m.nv, nvp = m.nvp;
  double *H = DF(s_H), *Md = DF(s_Md), *grad = DF(s_grad), *Mgrad = DF(s_Mgrad);
  const double *J = DF(efc_J), *D = DF(efc_D);
  const int* state = IF(efc_state);
  for (int i = c.lane; i < nv; i += 32) {
    double g = DF(s_Ma)[i] - DF(qfrc_smooth)[i] - DF(qfrc_constraint)[i];
    grad[i] = g; Mgrad[i] = g;
  }
  for (int e = c.lane; e < m.ntri; e += 32) {
    int i = m.tri_i[e], j = m.tri_j[e];
    double acc = Md[i * nvp + j];
    for (int r = 0; r < nefc; r++)
      if (state[r] == ST_QUADRATIC) acc += D[r] * J[r * nvp + i] * J[r * nvp + j];
    H[i * nvp + j] = acc;
  }
  SYNC();
  if (!bp_chol_factor(c, H, DF(s_Hdinv))) {
    SYNC();
    for (int e = c.lane; e < m.ntri; e += 32) {
      int i = m.tri_i[e], j = m.tri_j[e];
      H[i * nvp + j] = Md[i * nvp + j];
    }
    SYNC();
    bp_chol_factor(c, H, DF(s_Hdinv));
  }
  bp_chol_solve(c, H, DF(s_Hdinv), Mgrad);
}

__device__ void bp_fwd_constraint(const Ctx& c) {
  const BpModel& m = c.P.m;
  const int nv = m.nv, nvp = m.nvp, nefc = *IF(nefc);
  double *qacc = DF(qacc), *warm = DF(qacc_warmstart), *as = DF(qacc_smooth);
  if (nefc == 0) {
    for (int i = c.lane; i < nv; i += 32) {
      double a = as[i];
      qacc[i] = a; warm[i] = a; DF(qfrc_constraint)[i] = 0;
    }
    if (c.lane == 0) *IF(solver_iter) = 0;
    SYNC();
    return;
  }
  // dense copy of M
  double *Md = DF(s_Md), *qM = DF(qM);
  for (int w = c.lane; w < nv * nvp; w += 32) Md[w] = 0;
  SYNC();
  for (int i = c.lane; i < nv; i += 32) {
    int a = m.dof_Madr[i];
    for (int j = i; j >= 0; j = m.dof_parentid[j], a++) {
      Md[i * nvp + j] = qM[a]; Md[j * nvp + i] = qM[a];
    }
  }
  SYNC();
  double *Ma = DF(s_Ma), *jar = DF(s_jar), *search = DF(s_search);
  bool use_warm = false;
  if (!(m.opt_disableflags & DSBL_WARMSTART)) {
    for (int i = c.lane; i < nv; i += 32) qacc[i] = warm[i];
    SYNC();
    bp_mul_md(c, Md, Ma, qacc);
    bp_mul_j(c, jar, qacc, DF(efc_aref), nefc);
    double cw = bp_update_constraint(c, jar, nefc, false) + bp_gauss_cost(c);
    bp_mul_j(c, jar, as, DF(efc_aref), nefc);
    double cs = bp_update_constraint(c, jar, nefc, false);
    use_warm = cw < cs;
    SYNC();
  }
  for (int i = c.lane; i < nv; i += 32) qacc[i] = use_warm ? warm[i] : as[i];
  SYNC();
  bp_mul_md(c, Md, Ma, qacc);
  bp_mul_j(c, jar, qacc, DF(efc_aref), nefc);
  double cost = bp_update_constraint(c, jar, nefc, true) + bp_gauss_cost(c);
  bp_newton_gradient(c, nefc);
  for (int i = c.lane; i < nv; i += 32) search[i] = -DF(s_Mgrad)[i];
  SYNC();
  const double scale = 1.0 / (m.stat_meaninertia * (nv > 1 ? nv : 1));
  int iter = 0;
  double improvement = 0, gradient = 0;
  while (iter < m.opt_iterations) {
    bp_mul_md(c, Md, DF(s_Mv), search);
    bp_mul_j(c, DF(s_jv), search, nullptr, nefc);
    double alpha = bp_linesearch(c, scale, nefc);
    if (alpha == 0) break;
    for (int i = c.lane; i < nv; i += 32) {
      qacc[i] += alpha * search[i];
      Ma[i] += alpha * DF(s_Mv)[i];
    }
    for (int r = c.lane; r < nefc; r += 32) jar[r] += alpha * DF(s_jv)[r];
    SYNC();
    double old = cost;
    cost = bp_update_constraint(c, jar, nefc, true) + bp_gauss_cost(c);
    bp_newton_gradient(c, nefc);
    double gn = 0;
    for (int i = c.lane; i < nv; i += 32) gn += DF(s_grad)[i] * DF(s_grad)[i];
    gn = warp_sum(gn);
    improvement = scale * (old - cost); gradient = scale * sqrt(gn);
    iter++;
    if (improvement < m.opt_tolerance || gradient < m.opt_tolerance) break;
    for (int i = c.lane; i < nv; i += 32) search[i] = -DF(s_Mgrad)[i];
    SYNC();
  }
  if (c.lane == 0) {
    *IF(solver_iter) = iter;
    DF(solver_stat)[0] = improvement; DF(solver_stat)[1] = gradient;
  }
  for (int i = c.lane; i < nv; i += 32) warm[i] = qacc[i];
  SYNC();
}

#endif  // BP_SOLVER_CUH_
