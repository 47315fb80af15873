// mj_fwdConstraint: primal Newton solver in qacc with exact line search
// (SURVEY.md §8a row a2, Appendix A.9).  Per environment group of G lanes:
//   rows of J on lanes for J*x products, cost/line-search sums by group shuffles,
//   Hessian M + J^T D J one lower-triangle entry per lane, dense Cholesky with
//   rows on lanes.
#ifndef BP_SOLVER_CUH_
#define BP_SOLVER_CUH_
#include "bp_math.cuh"

// y = Md * x, dense symmetric nv x nv with row stride nvp (odd => conflict free)
BP_STAGE void bp_mul_md(const C& c, typename C::DP Md, typename C::DP y, typename C::DP x) {
  const int nv = c.P.m.nv, nvp = c.P.m.nvp;
#pragma unroll 1
  for (int i = c.lane; i < nv; i += C::G) {
    double s = 0;
    for (int j = 0; j < nv; j++) s += Md[i * nvp + j] * x[j];
    y[i] = s;
  }
  SYNC();
}

// out[r] = J[r,:] x - (sub ? sub[r] : 0)
BP_STAGE void bp_mul_j(const C& c, typename C::DP out, typename C::DP x, typename C::DP sub,
                       int nefc) {
  const int nv = c.P.m.nv, nvp = c.P.m.nvp;
  auto J = DF(efc_J);
#pragma unroll 1
  for (int r = c.lane; r < nefc; r += C::G) {
    double s = 0;
    for (int i = 0; i < nv; i++) s += J[r * nvp + i] * x[i];
    out[r] = sub.valid() ? s - sub[r] : s;
  }
  SYNC();
}

// constraint cost at residual jar; optionally writes force/state and J^T force
BP_STAGE double bp_update_constraint(const C& c, typename C::DP jar, int nefc, bool write) {
  const int nv = c.P.m.nv, nvp = c.P.m.nvp;
  double cost = 0;
  auto force = DF(efc_force), D = DF(efc_D), R = DF(efc_R);
  auto type = IF(efc_type), state = IF(efc_state);
#pragma unroll 1
  for (int r = c.lane; r < nefc; r += C::G) {
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
    if (write) {
      force[r] = f; state[r] = st;
      DF(s_rmask)[r] = st == ST_QUADRATIC ? DF(s_rdof)[r] : 0.0;   // bit patterns, see below
    }
  }
  cost = group_sum(c, cost);
  if (write) {
    SYNC();
    auto J = DF(efc_J);
    auto qc = DF(qfrc_constraint);
#pragma unroll 1
    for (int i = c.lane; i < nv; i += C::G) {
      double s = 0;
      for (int r = 0; r < nefc; r++) s += J[r * nvp + i] * force[r];
      qc[i] = s;
    }
    SYNC();
  }
  return cost;
}

template <class C>
__device__ __forceinline__ double bp_gauss_cost(const C& c) {
  double g = 0;
  for (int i = c.lane; i < c.P.m.nv; i += C::G)
    g += 0.5 * (DF(s_Ma)[i] - DF(qfrc_smooth)[i]) * (DF(qacc)[i] - DF(qacc_smooth)[i]);
  return group_sum(c, g);
}

struct LSPoint { double cost, d0, d1; };

BP_STAGE LSPoint bp_ls_eval(const C& c, double alpha, double g0, double g1, double g2,
                            int nefc) {
  double c0 = 0, c1 = 0, c2 = 0;
  auto D = DF(efc_D), jar = DF(s_jar), jvv = DF(s_jv);
  auto type = IF(efc_type);
#pragma unroll 1
  for (int r = c.lane; r < nefc; r += C::G) {
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
  c0 = g0 + group_sum(c, c0); c1 = g1 + group_sum(c, c1); c2 = g2 + group_sum(c, c2);
  LSPoint p;
  p.cost = alpha * alpha * c2 + alpha * c1 + c0;
  p.d0 = 2 * alpha * c2 + c1;
  p.d1 = 2 * c2;
  return p;
}

// exact 1-D minimisation along s_search (safeguarded Newton on the derivative)
BP_STAGE double bp_linesearch(const C& c, double scale, int nefc) {
  const BpModel& m = c.P.m;
  double sn = 0, g0 = 0, g1 = 0, g2 = 0;
  for (int i = c.lane; i < m.nv; i += C::G) {
    double s = DF(s_search)[i], r = DF(s_Ma)[i] - DF(qfrc_smooth)[i];
    sn += s * s;
    g0 += 0.5 * r * (DF(qacc)[i] - DF(qacc_smooth)[i]);
    g1 += s * r;
    g2 += 0.5 * s * DF(s_Mv)[i];
  }
  sn = sqrt(group_sum(c, sn)); g0 = group_sum(c, g0); g1 = group_sum(c, g1);
  g2 = group_sum(c, g2);
  if (sn < BP_MINVAL) return 0;
  double gtol = m.opt_tolerance * m.opt_ls_tolerance * sn / scale;
  LSPoint p = bp_ls_eval(c, 0, g0, g1, g2, nefc);
  if (p.d1 < BP_MINVAL) return 0;
  double lo = 0, hi = 0, a = -p.d0 / p.d1;
  bool have_hi = false;
#pragma unroll 1
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

// ---- dense Cholesky H = L L^T in place (lower triangle, row stride nvp) -------------
// Row i of the system belongs to lane i % G, register slot i / G.  When nv <= 4 G every
// lane owns at most BP_RS = 4 rows, so a column step keeps its 4 partial dot products in
// registers (left-looking: one broadcast load of L[j][k] feeds 4 independent FMAs) and
// the triangular solves keep the whole right-hand side in registers, exchanging pivots
// by group shuffles instead of memory + barriers.  Larger systems fall back to the
// generic right-looking code.
#define BP_RS 4

template <class C>
__device__ __forceinline__ double bp_slot(const double (&v)[BP_RS], int s) {
  return s == 0 ? v[0] : (s == 1 ? v[1] : (s == 2 ? v[2] : v[3]));
}

template <class C>
__device__ __forceinline__ double group_bcast(const C& c, double v, int src) {
  return C::G == 1 ? v : __shfl_sync(c.mask, v, src, C::G);
}

// Returns false when a pivot is not positive.
BP_STAGE bool bp_chol_factor_generic(const C& c, typename C::DP H, typename C::DP dinv) {
  const int n = c.P.m.nv, nvp = c.P.m.nvp;
#pragma unroll 1
  for (int j = 0; j < n; j++) {
    double s = H[j * nvp + j];
    if (s < BP_MINVAL) return false;      // uniform across the group
    s = sqrt(s);
    double inv = 1.0 / s;
    SYNC();
    if (c.lane == 0) { H[j * nvp + j] = s; dinv[j] = inv; }
    for (int i = j + 1 + c.lane; i < n; i += C::G) H[i * nvp + j] /= s;
    SYNC();
#pragma unroll 1
    for (int i = j + 1 + c.lane; i < n; i += C::G) {
      double lij = H[i * nvp + j];
      for (int k = j + 1; k <= i; k++) H[i * nvp + k] -= lij * H[k * nvp + j];
    }
    SYNC();
  }
  return true;
}

BP_STAGE bool bp_chol_factor(const C& c, typename C::DP H, typename C::DP dinv) {
  const int n = c.P.m.nv, nvp = c.P.m.nvp;
  if (n > BP_RS * C::G) return bp_chol_factor_generic(c, H, dinv);
  int row[BP_RS];
#pragma unroll
  for (int s = 0; s < BP_RS; s++) row[s] = (c.lane + C::G * s) * nvp;
#pragma unroll 1
  for (int j = 0; j < n; j++) {
    // acc[s] = H[i][j] - sum_{k<j} L[i][k] L[j][k] for this lane's rows i >= j
    double acc[BP_RS];
    bool on[BP_RS];
#pragma unroll
    for (int s = 0; s < BP_RS; s++) {
      int i = c.lane + C::G * s;
      on[s] = i >= j && i < n;
      acc[s] = on[s] ? H[row[s] + j] : 0.0;
    }
    const int rj = j * nvp;
#pragma unroll 2
    for (int k = 0; k < j; k++) {
      double ljk = H[rj + k];
#pragma unroll
      for (int s = 0; s < BP_RS; s++)
        if (on[s]) acc[s] -= H[row[s] + k] * ljk;
    }
    double piv = group_bcast(c, bp_slot<C>(acc, j / C::G), j % C::G);
    if (piv < BP_MINVAL) return false;     // uniform across the group
    double d = sqrt(piv), inv = 1.0 / d;
#pragma unroll
    for (int s = 0; s < BP_RS; s++) {
      int i = c.lane + C::G * s;
      if (on[s]) H[row[s] + j] = i == j ? d : acc[s] * inv;
    }
    if (c.lane == 0) dinv[j] = inv;
    SYNC();                                 // column j (and row j+1 up to it) visible
  }
  return true;
}

BP_STAGE void bp_chol_solve_generic(const C& c, typename C::DP L, typename C::DP dinv,
                                    typename C::DP x) {
  const int n = c.P.m.nv, nvp = c.P.m.nvp;
#pragma unroll 1
  for (int k = 0; k < n; k++) {
    double xk = x[k] * dinv[k];
    SYNC();
    if (c.lane == 0) x[k] = xk;
    for (int i = k + 1 + c.lane; i < n; i += C::G) x[i] -= L[i * nvp + k] * xk;
    SYNC();
  }
#pragma unroll 1
  for (int k = n - 1; k >= 0; k--) {
    double xk = x[k] * dinv[k];
    SYNC();
    if (c.lane == 0) x[k] = xk;
    for (int i = c.lane; i < k; i += C::G) x[i] -= L[k * nvp + i] * xk;
    SYNC();
  }
}

BP_STAGE void bp_chol_solve(const C& c, typename C::DP L, typename C::DP dinv,
                            typename C::DP x) {
  const int n = c.P.m.nv, nvp = c.P.m.nvp;
  if (n > BP_RS * C::G) { bp_chol_solve_generic(c, L, dinv, x); return; }
  double xs[BP_RS];
#pragma unroll
  for (int s = 0; s < BP_RS; s++) {
    int i = c.lane + C::G * s;
    xs[s] = i < n ? x[i] : 0.0;
  }
  // forward sweep L y = b by columns: pivot k is final on its owner, everyone below it
  // subtracts L[i][k] y_k
#pragma unroll 1
  for (int k = 0; k < n; k++) {
    double yk = group_bcast(c, bp_slot<C>(xs, k / C::G), k % C::G) * dinv[k];
#pragma unroll
    for (int s = 0; s < BP_RS; s++) {
      int i = c.lane + C::G * s;
      if (i == k) xs[s] = yk;
      else if (i > k && i < n) xs[s] -= L[i * nvp + k] * yk;
    }
  }
  // backward sweep L^T x = y by rows of L
#pragma unroll 1
  for (int k = n - 1; k >= 0; k--) {
    double xk = group_bcast(c, bp_slot<C>(xs, k / C::G), k % C::G) * dinv[k];
    const int rk = k * nvp;
#pragma unroll
    for (int s = 0; s < BP_RS; s++) {
      int i = c.lane + C::G * s;
      if (i == k) xs[s] = xk;
      else if (i < k) xs[s] -= L[rk + i] * xk;
    }
  }
#pragma unroll
  for (int s = 0; s < BP_RS; s++) {
    int i = c.lane + C::G * s;
    if (i < n) x[i] = xs[s];
  }
  SYNC();
}

// grad = Ma - qfrc_smooth - qfrc_constraint;  H = M + J^T D_active J;  Mgrad = H^-1 grad
BP_STAGE void bp_newton_gradient(const C& c, int nefc) {
  const BpModel& m = c.P.m;
  const int nv = m.nv, nvp = m.nvp;
  auto H = DF(s_H), Md = DF(s_Md), grad = DF(s_grad), Mgrad = DF(s_Mgrad);
  auto J = DF(efc_J), D = DF(efc_D);
  for (int i = c.lane; i < nv; i += C::G) {
    double g = DF(s_Ma)[i] - DF(qfrc_smooth)[i] - DF(qfrc_constraint)[i];
    grad[i] = g; Mgrad[i] = g;
  }
  // entry (i, j) only visits the active rows whose dof set holds both i and j
  // (s_rmask: dof bit mask of quadratic rows, 0 for the others): limit and friction
  // rows touch one diagonal entry, a contact row the ancestors of its two bodies.
  auto rmask = DF(s_rmask);
#pragma unroll 1
  for (int e = c.lane; e < m.ntri; e += C::G) {
    int i = m.tri_i[e], j = m.tri_j[e];
    const unsigned long long bits = (1ull << i) | (1ull << j);
    double acc = Md[i * nvp + j];
#pragma unroll 2
    for (int r = 0; r < nefc; r++) {
      unsigned long long mk = (unsigned long long)__double_as_longlong(rmask[r]);
      if ((mk & bits) == bits) acc += D[r] * J[r * nvp + i] * J[r * nvp + j];
    }
    H[i * nvp + j] = acc;
  }
  SYNC();
  if (!bp_chol_factor(c, H, DF(s_Hdinv))) {
    SYNC();
    for (int e = c.lane; e < m.ntri; e += C::G) {
      int i = m.tri_i[e], j = m.tri_j[e];
      H[i * nvp + j] = Md[i * nvp + j];
    }
    SYNC();
    bp_chol_factor(c, H, DF(s_Hdinv));
  }
  bp_chol_solve(c, H, DF(s_Hdinv), Mgrad);
}

BP_STAGE void bp_fwd_constraint(const C& c) {
  const BpModel& m = c.P.m;
  const int nv = m.nv, nvp = m.nvp, nefc = IF(nefc)[0];
  auto qacc = DF(qacc), warm = DF(qacc_warmstart), as = DF(qacc_smooth);
  if (nefc == 0) {
    for (int i = c.lane; i < nv; i += C::G) {
      double a = as[i];
      qacc[i] = a; warm[i] = a; DF(qfrc_constraint)[i] = 0;
    }
    if (c.lane == 0) IF(solver_iter)[0] = 0;
    SYNC();
    return;
  }
  // dof set of every row as a bit mask (nv <= 64), stored bit-for-bit in a double slot
#pragma unroll 1
  for (int r = c.lane; r < nefc; r += C::G) {
    int t = IF(efc_type)[r], id = IF(efc_id)[r];
    unsigned long long mk;
    if (t == CN_FRICTION_DOF) mk = 1ull << id;
    else if (t == CN_LIMIT_JOINT) mk = 1ull << m.jnt_dofadr[id];
    else mk = m.body_dofmask[m.geom_bodyid[IF(contact_geom1)[id]]] |
              m.body_dofmask[m.geom_bodyid[IF(contact_geom2)[id]]];
    DF(s_rdof)[r] = __longlong_as_double((long long)mk);
  }
  // dense copy of M
  auto Md = DF(s_Md), qM = DF(qM);
  for (int w = c.lane; w < nv * nvp; w += C::G) Md[w] = 0;
  SYNC();
#pragma unroll 1
  for (int i = c.lane; i < nv; i += C::G) {
    int a = m.dof_Madr[i];
    for (int j = i; j >= 0; j = m.dof_parentid[j], a++) {
      double v = qM[a];
      Md[i * nvp + j] = v; Md[j * nvp + i] = v;
    }
  }
  SYNC();
  auto Ma = DF(s_Ma), jar = DF(s_jar), search = DF(s_search);
  bool use_warm = false;
  if (!(m.opt_disableflags & DSBL_WARMSTART)) {
    for (int i = c.lane; i < nv; i += C::G) qacc[i] = warm[i];
    SYNC();
    bp_mul_md(c, Md, Ma, qacc);
    bp_mul_j(c, jar, qacc, DF(efc_aref), nefc);
    double cw = bp_update_constraint(c, jar, nefc, false) + bp_gauss_cost(c);
    bp_mul_j(c, jar, as, DF(efc_aref), nefc);
    double cs = bp_update_constraint(c, jar, nefc, false);
    use_warm = cw < cs;
    SYNC();
  }
  for (int i = c.lane; i < nv; i += C::G) qacc[i] = use_warm ? warm[i] : as[i];
  SYNC();
  bp_mul_md(c, Md, Ma, qacc);
  bp_mul_j(c, jar, qacc, DF(efc_aref), nefc);
  double cost = bp_update_constraint(c, jar, nefc, true) + bp_gauss_cost(c);
  bp_newton_gradient(c, nefc);
  for (int i = c.lane; i < nv; i += C::G) search[i] = -DF(s_Mgrad)[i];
  SYNC();
  const double scale = 1.0 / (m.stat_meaninertia * (nv > 1 ? nv : 1));
  int iter = 0;
  double improvement = 0, gradient = 0;
#pragma unroll 1
  while (iter < m.opt_iterations) {
    bp_mul_md(c, Md, DF(s_Mv), search);
    bp_mul_j(c, DF(s_jv), search, c.null(), nefc);
    double alpha = bp_linesearch(c, scale, nefc);
    if (alpha == 0) break;
    for (int i = c.lane; i < nv; i += C::G) {
      qacc[i] += alpha * search[i];
      Ma[i] += alpha * DF(s_Mv)[i];
    }
    for (int r = c.lane; r < nefc; r += C::G) jar[r] += alpha * DF(s_jv)[r];
    SYNC();
    double old = cost;
    cost = bp_update_constraint(c, jar, nefc, true) + bp_gauss_cost(c);
    bp_newton_gradient(c, nefc);
    double gn = 0;
    for (int i = c.lane; i < nv; i += C::G) gn += DF(s_grad)[i] * DF(s_grad)[i];
    gn = group_sum(c, gn);
    improvement = scale * (old - cost); gradient = scale * sqrt(gn);
    iter++;
    if (improvement < m.opt_tolerance || gradient < m.opt_tolerance) break;
    for (int i = c.lane; i < nv; i += C::G) search[i] = -DF(s_Mgrad)[i];
    SYNC();
  }
  if (c.lane == 0) {
    IF(solver_iter)[0] = iter;
    DF(solver_stat)[0] = improvement; DF(solver_stat)[1] = gradient;
  }
  for (int i = c.lane; i < nv; i += C::G) warm[i] = qacc[i];
  SYNC();
}

#endif  // BP_SOLVER_CUH_
