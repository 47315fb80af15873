// mj_fwdConstraint: primal Newton solver in qacc with exact line search
// (SURVEY.md §8a row a2, Appendix A.9).  Per environment group of G lanes:
//   rows of J on lanes for J*x products, cost/line-search sums by group shuffles,
//   Hessian M + J^T D J one lower-triangle entry per lane, dense Cholesky with
//   rows on lanes.
#ifndef BP_SOLVER_CUH_
#define BP_SOLVER_CUH_
#include "bp_math.cuh"

#ifndef BP_NEWTON_GROUP
#define BP_NEWTON_GROUP 0     // warps that share the per-iteration barrier (0: the block)
#endif
// barrier `id` over `nthreads` threads with an OR reduction of pred (bar.red.or)
__device__ __forceinline__ bool bp_bar_or(int id, int nthreads, bool pred) {
  int r;
  asm volatile("{ .reg .pred p, q; setp.ne.u32 q, %1, 0; bar.red.or.pred p, %2, %3, q; "
               "selp.u32 %0, 1, 0, p; }" : "=r"(r) : "r"((int)pred), "r"(id), "r"(nthreads) : "memory");
  return r != 0;
}

// the Hessian and its factor are stored as a packed lower triangle, row i at i (i + 1) / 2
#define PK(i, j) ((i) * ((i) + 1) / 2 + (j))

// dst[0..n) = src[0..n): four loads in flight per lane before the stores (the compiler
// cannot reorder them itself, the two ranges might alias)
template <class C>
__device__ __forceinline__ void bp_copy(const C& c, typename C::DP dst, typename C::DP src, int n) {
  int e = c.lane;
  for (; e + 3 * C::G < n; e += 4 * C::G) {
    double a0 = src[e], a1 = src[e + C::G], a2 = src[e + 2 * C::G], a3 = src[e + 3 * C::G];
    dst[e] = a0; dst[e + C::G] = a1; dst[e + 2 * C::G] = a2; dst[e + 3 * C::G] = a3;
  }
  for (; e < n; e += C::G) dst[e] = src[e];
}

// y = M x with M a packed lower triangle (row i at i (i + 1) / 2): row i reads its own
// packed row up to the diagonal, then walks down column i
BP_STAGE void bp_mul_md(const C& c, typename C::DP Mp, typename C::DP y, typename C::DP x) {
  const int nv = bp_mv(c).nv;
#pragma unroll (C::UF)
  for (int i = c.lane; i < nv; i += C::G) {
    double s = 0;
    const int ri = i * (i + 1) / 2;
#pragma unroll (C::UF > 4 ? C::UF : 4)
    for (int j = 0; j <= i; j++) s += Mp[ri + j] * x[j];
    int a = ri + 2 * i + 1;                       // (i + 1, i)
#pragma unroll (C::UF > 4 ? C::UF : 4)
    for (int j = i + 1; j < nv; j++) { s += Mp[a] * x[j]; a += j + 1; }
    y[i] = s;
  }
  SYNC();
}

// out[r] = J[r,:] x - (sub ? sub[r] : 0)
BP_STAGE void bp_mul_j(const C& c, typename C::DP out, typename C::DP x, typename C::DP sub,
                       int nefc) {
  const int nv = bp_mv(c).nv, nvp = bp_mv(c).nvp;
  auto J = DF(efc_J);
#pragma unroll 1
  for (int r = c.lane; r < nefc; r += C::G) {
    double s = 0;
#pragma unroll (C::UF > 4 ? C::UF : 4)
    for (int i = 0; i < nv; i++) s += J[r * nvp + i] * x[i];
    out[r] = sub.valid() ? s - sub[r] : s;
  }
  SYNC();
}

// ---- elliptic friction cones (opt_cone == 1; Appendix A.9) ---------------------------
// In the scaled coordinates U = (mu x0, fri_j x_j), N = U0, T = |U_1..|: N >= mu T is the
// top zone (no force), mu N + T <= 0 the bottom zone (every row plainly quadratic) and in
// between the cost is 0.5 Dm (N - mu T)^2, Dm = D0 / (mu^2 (1 + mu^2)).  Contacts on lanes.
// Kept out of line: the pyramidal models never execute any of it.
BP_STAGE double bp_cone_update(const C& c, typename C::DP jar, int nefc, bool write, bool* chg) {
  const int ncon = IF(ncon)[0];
  auto force = DF(efc_force), D = DF(efc_D), rw = DF(s_rw);
  auto state = IF(efc_state);
  double cost = 0;
#pragma unroll 1
  for (int ci = c.lane; ci < ncon; ci += C::G) {
    const int a = IF(contact_efc_address)[ci], dim = IF(contact_dim)[ci];
    if (a < 0 || dim == 1 || a + dim > nefc) continue;
    const double mu = DF(contact_mu)[ci];
    auto fri = DF(contact_friction) + 5 * ci;
    const double N = jar[a] * mu;
    double tt = 0;
    for (int j = 1; j < dim; j++) { double u = jar[a + j] * fri[j - 1]; tt += u * u; }
    const double T = sqrt(tt);
    if (N >= mu * T || (T <= 0 && N >= 0)) {
      if (write)
        for (int j = 0; j < dim; j++) {
          force[a + j] = 0; state[a + j] = ST_SATISFIED;
          *chg |= rw[a + j] != 0; rw[a + j] = 0;
        }
    } else if (mu * N + T <= 0 || (T <= 0 && N < 0)) {
      for (int j = 0; j < dim; j++) {
        double d = D[a + j], x = jar[a + j];
        cost += 0.5 * d * x * x;
        if (write) {
          force[a + j] = -d * x; state[a + j] = ST_QUADRATIC;
          *chg |= rw[a + j] != d; rw[a + j] = d;
        }
      }
    } else {
      const double Dm = D[a] / (mu * mu * (1 + mu * mu)), NmT = N - mu * T;
      cost += 0.5 * Dm * NmT * NmT;
      if (write) {
        const double f0 = -Dm * NmT * mu;
        force[a] = f0; state[a] = ST_CONE; rw[a] = 0;
        for (int j = 1; j < dim; j++) {
          force[a + j] = -f0 / T * (jar[a + j] * fri[j - 1]) * fri[j - 1];
          state[a + j] = ST_CONE; rw[a + j] = 0;
        }
        *chg = true;                 // the cone Hessian moves with jar
      }
    }
  }
  return cost;
}

// line-search sums of the elliptic contacts at step alpha: bottom-zone rows add to the
// quadratic coefficients q[0..2], middle-zone cones to the value / slope / curvature e[0..2]
BP_STAGE void bp_cone_ls(const C& c, double alpha, int nefc, double* q, double* e) {
  const int ncon = IF(ncon)[0];
  auto D = DF(efc_D), jar = DF(s_jar), jvv = DF(s_jv);
#pragma unroll 1
  for (int ci = c.lane; ci < ncon; ci += C::G) {
    const int a = IF(contact_efc_address)[ci], dim = IF(contact_dim)[ci];
    if (a < 0 || dim == 1 || a + dim > nefc) continue;
    const double mu = DF(contact_mu)[ci];
    auto fri = DF(contact_friction) + 5 * ci;
    const double U0 = jar[a] * mu, V0 = jvv[a] * mu;
    double UU = 0, UV = 0, VV = 0;
    for (int j = 1; j < dim; j++) {
      double u = jar[a + j] * fri[j - 1], v = jvv[a + j] * fri[j - 1];
      UU += u * u; UV += u * v; VV += v * v;
    }
    const double N = U0 + alpha * V0, Tsqr = UU + alpha * (2 * UV + alpha * VV);
    const double T = Tsqr > 0 ? sqrt(Tsqr) : 0;
    if (N >= mu * T || (T <= 0 && N >= 0)) continue;
    if (mu * N + T <= 0 || (T <= 0 && N < 0)) {
      for (int j = 0; j < dim; j++) {
        double d = D[a + j], j0 = jar[a + j], jv = jvv[a + j];
        q[0] += 0.5 * d * j0 * j0; q[1] += d * j0 * jv; q[2] += 0.5 * d * jv * jv;
      }
    } else {
      const double Dm = D[a] / (mu * mu * (1 + mu * mu));
      const double N1 = V0, T1 = (UV + alpha * VV) / T;
      const double T2 = VV / T - (UV + alpha * VV) * T1 / (T * T);
      const double NmT = N - mu * T, NmT1 = N1 - mu * T1;
      e[0] += 0.5 * Dm * NmT * NmT;
      e[1] += Dm * NmT * NmT1;
      e[2] += Dm * (NmT1 * NmT1 - NmT * mu * T2);
    }
  }
}

// H += Jc^T Hc Jc for the contacts in the middle zone, Hc = Dm S (g g^T + kappa P) S with
// S = diag(mu, fri), u = U_t / T, g = (1, -mu u), P = I - u u^T on the tangential block,
// kappa = mu^2 - mu N / T.  One contact at a time; its dof-set pairs over the lanes.
BP_STAGE void bp_cone_hessian(const C& c, typename C::DP H, int nefc) {
  BP_MODEL(m);
  const int ncon = IF(ncon)[0], nvp = m.nvp;
  auto J = DF(efc_J), jar = DF(s_jar), D = DF(efc_D);
#pragma unroll 1
  for (int ci = 0; ci < ncon; ci++) {
    const int a = IF(contact_efc_address)[ci], dim = IF(contact_dim)[ci];
    if (a < 0 || dim == 1 || a + dim > nefc || IF(efc_state)[a] != ST_CONE) continue;
    const double mu = DF(contact_mu)[ci];
    auto fri = DF(contact_friction) + 5 * ci;
    double S[6], g[6], u[6];
    const double N = jar[a] * mu;
    double tt = 0;
    S[0] = mu; g[0] = 1; u[0] = 0;
    for (int j = 1; j < dim; j++) { S[j] = fri[j - 1]; u[j] = jar[a + j] * S[j]; tt += u[j] * u[j]; }
    const double T = sqrt(tt);
    for (int j = 1; j < dim; j++) { u[j] /= T; g[j] = -mu * u[j]; }
    const double Dm = D[a] / (mu * mu * (1 + mu * mu)), kappa = mu * mu - mu * N / T;
    const unsigned long long mk = m.body_dofmask[m.geom_bodyid[IF(contact_geom1)[ci]]] |
                                  m.body_dofmask[m.geom_bodyid[IF(contact_geom2)[ci]]];
    unsigned long long remi = mk;
#pragma unroll 1
    for (int ap = 0; remi; ap++, remi &= remi - 1) {
      const int i = __ffsll((long long)remi) - 1;
      double y[6], gv = 0, ut = 0;
      for (int q = 0; q < dim; q++) {
        y[q] = S[q] * J[(a + q) * nvp + i];
        gv += g[q] * y[q];
        if (q > 0) ut += u[q] * y[q];
      }
      for (int q = 0; q < dim; q++)
        y[q] = Dm * S[q] * (g[q] * gv + (q > 0 ? kappa * (y[q] - u[q] * ut) : 0.0));
      unsigned long long remj = mk;
      for (int t = 0; t < c.lane; t++) remj &= remj - 1;
#pragma unroll 1
      for (int bp = c.lane; bp <= ap; bp += C::G) {
        const int j = __ffsll((long long)remj) - 1;
        for (int t = 0; t < C::G; t++) remj &= remj - 1;
        double sum = 0;
        for (int q = 0; q < dim; q++) sum += y[q] * J[(a + q) * nvp + j];
        H[PK(i, j)] += sum;
      }
    }
    SYNC();
  }
}

// constraint cost at residual jar; optionally writes force/state and J^T force
// *changed (write only): some row's Hessian weight differs from the previous call, i.e.
// the factorisation of H is stale
BP_STAGE double bp_update_constraint(const C& c, typename C::DP jar, int nefc, bool write,
                                     bool* changed) {
  const int nv = bp_mv(c).nv, nvp = bp_mv(c).nvp;
  double cost = 0;
  bool chg = false;
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
    } else if (t == CN_CONTACT_ELLIPTIC) {
      continue;                                    // whole contacts: bp_cone_update
    } else {
      if (x >= 0) { f = 0; st = ST_SATISFIED; }
      else { f = -d * x; st = ST_QUADRATIC; cr = 0.5 * d * x * x; }
    }
    cost += cr;
    if (write) {
      force[r] = f; state[r] = st;
      double w = st == ST_QUADRATIC ? d : 0.0;         // row weight in the Hessian
      chg |= DF(s_rw)[r] != w;
      DF(s_rw)[r] = w;
    }
  }
  if constexpr (C::FULL) if (bp_mv(c).opt_cone == 1) cost += bp_cone_update(c, jar, nefc, write, &chg);
  cost = group_sum(c, cost);
  if (write) {
    *changed = group_any(c, chg);
    SYNC();
    auto J = DF(efc_J);
    auto qc = DF(qfrc_constraint);
#pragma unroll (C::UF)
    for (int i = c.lane; i < nv; i += C::G) {
      double s = 0;
#pragma unroll 4
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
  for (int i = c.lane; i < bp_mv(c).nv; i += C::G)
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
    } else if (t == CN_CONTACT_ELLIPTIC) continue;
    else if (x < 0) { c0 += q0; c1 += q1; c2 += q2; }
  }
  double e[3] = {0, 0, 0};
  if constexpr (C::FULL) if (bp_mv(c).opt_cone == 1) {
    double q[3] = {0, 0, 0};
    bp_cone_ls(c, alpha, nefc, q, e);
    c0 += q[0]; c1 += q[1]; c2 += q[2];
    e[0] = group_sum(c, e[0]); e[1] = group_sum(c, e[1]); e[2] = group_sum(c, e[2]);
  }
  c0 = g0 + group_sum(c, c0); c1 = g1 + group_sum(c, c1); c2 = g2 + group_sum(c, c2);
  LSPoint p;
  p.cost = alpha * alpha * c2 + alpha * c1 + c0 + e[0];
  p.d0 = 2 * alpha * c2 + c1 + e[1];
  p.d1 = 2 * c2 + e[2];
  return p;
}

// exact 1-D minimisation along s_search (safeguarded Newton on the derivative)
BP_STAGE double bp_linesearch(const C& c, double scale, int nefc) {
  BP_MODEL(m);
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

// ---- dense Cholesky H = L L^T in place (packed lower triangle) ---------------------
// Row i of the system belongs to lane i % G, register slot i / G.  When the rows fit
// BP_RS = 4 slots per lane (nv <= 4 G) a column step keeps its partial dot products in
// registers (left-looking: one broadcast load of L[j][k] feeds NS independent FMAs) and
// the triangular solves keep the whole right-hand side in registers, exchanging pivots
// by group shuffles instead of memory + barriers.  The code is instantiated for the
// number of slots NS a model really uses (humanoid on 16 lanes: 2), so dead slots cost no
// instructions.  Larger systems fall back to the generic right-looking code.
#define BP_RS 4

template <int NS>
__device__ __forceinline__ double bp_slot(const double (&v)[NS], int s) {
  double r = v[0];
#pragma unroll
  for (int t = 1; t < NS; t++) r = s == t ? v[t] : r;
  return r;
}

template <class C>
__device__ __forceinline__ double group_bcast(const C& c, double v, int src) {
  return C::G == 1 ? v : __shfl_sync(c.mask, v, src, C::G);
}

// Returns false when a pivot is not positive.
BP_STAGE bool bp_chol_factor_generic(const C& c, typename C::DP H, typename C::DP dinv) {
  const int n = bp_mv(c).nv;
#pragma unroll (C::UF)
  for (int j = 0; j < n; j++) {
    double s = H[PK(j, j)];
    if (s < BP_MINVAL) return false;      // uniform across the group
    s = sqrt(s);
    double inv = 1.0 / s;
    SYNC();
    if (c.lane == 0) { H[PK(j, j)] = s; dinv[j] = inv; }
    for (int i = j + 1 + c.lane; i < n; i += C::G) H[PK(i, j)] /= s;
    SYNC();
#pragma unroll (C::UF)
    for (int i = j + 1 + c.lane; i < n; i += C::G) {
      double lij = H[PK(i, j)];
      for (int k = j + 1; k <= i; k++) H[PK(i, k)] -= lij * H[PK(k, j)];
    }
    SYNC();
  }
  return true;
}

// The right-hand side of the solve that follows rides along as row n of the packed
// matrix (H[ntri + k] = g[k]): factoring the bordered matrix leaves y = L^-1 g in that
// row, which replaces the forward sweep.  bp_chol_fused says whether that applies.
template <class C>
__device__ __forceinline__ bool bp_chol_fused(int n) { return n + 1 <= BP_RS * C::G; }

// Two columns per step: the dot products of columns j and j + 1 share the loads of the
// lane's rows, and the pair costs one barrier.
template <class C, int NS>
__device__ __forceinline__ bool bp_chol_factor_t(const C& c, typename C::DP H,
                                                 typename C::DP dinv, int n, int nrows) {
  int row[NS];
#pragma unroll
  for (int s = 0; s < NS; s++) { int i = c.lane + C::G * s; row[s] = i * (i + 1) / 2; }
#pragma unroll 1
  for (int j = 0; j < n; j += 2) {
    const bool pair = j + 1 < n;
    // acc0[s] = H[i][j] - sum_{k<j} L[i][k] L[j][k], acc1 the same for column j + 1
    double acc0[NS], acc1[NS];
    bool on0[NS], on1[NS];
#pragma unroll
    for (int s = 0; s < NS; s++) {
      int i = c.lane + C::G * s;
      on0[s] = i >= j && i < nrows;
      on1[s] = pair && i > j && i < nrows;
      acc0[s] = on0[s] ? H[row[s] + j] : 0.0;
      acc1[s] = on1[s] ? H[row[s] + j + 1] : 0.0;
    }
    const int rj = j * (j + 1) / 2, rj1 = rj + j + 1;
#pragma unroll 2
    for (int k = 0; k < j; k++) {
      double l0 = H[rj + k], l1 = pair ? H[rj1 + k] : 0.0;
#pragma unroll
      for (int s = 0; s < NS; s++)
        if (on0[s]) {
          double v = H[row[s] + k];
          acc0[s] -= v * l0;
          acc1[s] -= v * l1;
        }
    }
    double piv0 = group_bcast(c, bp_slot<NS>(acc0, j / C::G), j % C::G);
    if (piv0 < BP_MINVAL) return false;    // uniform across the group
    double d0 = sqrt(piv0), inv0 = 1.0 / d0;
    double lcol[NS];                       // column j of L on this lane's rows
#pragma unroll
    for (int s = 0; s < NS; s++) {
      int i = c.lane + C::G * s;
      lcol[s] = i == j ? d0 : acc0[s] * inv0;
      if (on0[s]) H[row[s] + j] = lcol[s];
    }
    if (c.lane == 0) dinv[j] = inv0;
    if (pair) {
      const int j1 = j + 1;
      double l10 = group_bcast(c, bp_slot<NS>(lcol, j1 / C::G), j1 % C::G);   // L[j+1][j]
#pragma unroll
      for (int s = 0; s < NS; s++)
        if (on1[s]) acc1[s] -= lcol[s] * l10;
      double piv1 = group_bcast(c, bp_slot<NS>(acc1, j1 / C::G), j1 % C::G);
      if (piv1 < BP_MINVAL) return false;
      double d1 = sqrt(piv1), inv1 = 1.0 / d1;
#pragma unroll
      for (int s = 0; s < NS; s++) {
        int i = c.lane + C::G * s;
        if (on1[s]) H[row[s] + j1] = i == j1 ? d1 : acc1[s] * inv1;
      }
      if (c.lane == 0) dinv[j1] = inv1;
    }
    SYNC();                                 // columns j, j + 1 visible
  }
  return true;
}

BP_STAGE bool bp_chol_factor(const C& c, typename C::DP H, typename C::DP dinv) {
  const int n = bp_mv(c).nv;
  if (n > BP_RS * C::G) return bp_chol_factor_generic(c, H, dinv);
  const int nrows = bp_chol_fused<C>(n) ? n + 1 : n;
  switch ((nrows + C::G - 1) / C::G) {
    case 1: return bp_chol_factor_t<C, 1>(c, H, dinv, n, nrows);
    case 2: return bp_chol_factor_t<C, 2>(c, H, dinv, n, nrows);
    case 3: return bp_chol_factor_t<C, 3>(c, H, dinv, n, nrows);
    default: return bp_chol_factor_t<C, 4>(c, H, dinv, n, nrows);
  }
}

BP_STAGE void bp_chol_solve_generic(const C& c, typename C::DP L, typename C::DP dinv,
                                    typename C::DP x) {
  const int n = bp_mv(c).nv;
#pragma unroll (C::UF)
  for (int k = 0; k < n; k++) {
    double xk = x[k] * dinv[k];
    SYNC();
    if (c.lane == 0) x[k] = xk;
    for (int i = k + 1 + c.lane; i < n; i += C::G) x[i] -= L[PK(i, k)] * xk;
    SYNC();
  }
#pragma unroll (C::UF)
  for (int k = n - 1; k >= 0; k--) {
    double xk = x[k] * dinv[k];
    SYNC();
    if (c.lane == 0) x[k] = xk;
    for (int i = c.lane; i < k; i += C::G) x[i] -= L[PK(k, i)] * xk;
    SYNC();
  }
}

template <class C, int NS>
__device__ __forceinline__ void bp_chol_solve_t(const C& c, typename C::DP L, typename C::DP dinv,
                                                typename C::DP x, bool fwd_done, int n) {
  double xs[NS];
  const int rn = n * (n + 1) / 2;
#pragma unroll
  for (int s = 0; s < NS; s++) {
    int i = c.lane + C::G * s;
    xs[s] = i < n ? (fwd_done ? L[rn + i] : x[i]) : 0.0;
  }
  // forward sweep L y = b by columns: pivot k is final on its owner, everyone below it
  // subtracts L[i][k] y_k
#pragma unroll 1
  for (int k = 0; k < (fwd_done ? 0 : n); k++) {
    double yk = group_bcast(c, bp_slot<NS>(xs, k / C::G), k % C::G) * dinv[k];
#pragma unroll
    for (int s = 0; s < NS; s++) {
      int i = c.lane + C::G * s;
      if (i == k) xs[s] = yk;
      else if (i > k && i < n) xs[s] -= L[PK(i, k)] * yk;
    }
  }
  // backward sweep L^T x = y by rows of L
#pragma unroll 2
  for (int k = n - 1; k >= 0; k--) {
    double xk = group_bcast(c, bp_slot<NS>(xs, k / C::G), k % C::G) * dinv[k];
    const int rk = PK(k, 0);
#pragma unroll
    for (int s = 0; s < NS; s++) {
      int i = c.lane + C::G * s;
      if (i == k) xs[s] = xk;
      else if (i < k) xs[s] -= L[rk + i] * xk;
    }
  }
#pragma unroll
  for (int s = 0; s < NS; s++) {
    int i = c.lane + C::G * s;
    if (i < n) x[i] = xs[s];
  }
  SYNC();
}

// x <- H^-1 x.  `fwd_done`: y = L^-1 x already sits in row n of L (see bp_chol_fused)
BP_STAGE void bp_chol_solve(const C& c, typename C::DP L, typename C::DP dinv,
                            typename C::DP x, bool fwd_done) {
  const int n = bp_mv(c).nv;
  if (n > BP_RS * C::G) { bp_chol_solve_generic(c, L, dinv, x); return; }
  switch ((n + C::G - 1) / C::G) {
    case 1: bp_chol_solve_t<C, 1>(c, L, dinv, x, fwd_done, n); break;
    case 2: bp_chol_solve_t<C, 2>(c, L, dinv, x, fwd_done, n); break;
    case 3: bp_chol_solve_t<C, 3>(c, L, dinv, x, fwd_done, n); break;
    default: bp_chol_solve_t<C, 4>(c, L, dinv, x, fwd_done, n); break;
  }
}

// grad = Ma - qfrc_smooth - qfrc_constraint; returns |grad|^2
template <class C>
__device__ __forceinline__ double bp_newton_grad(const C& c) {
  auto grad = DF(s_grad);
  double gn = 0;
  for (int i = c.lane; i < bp_mv(c).nv; i += C::G) {
    double g = DF(s_Ma)[i] - DF(qfrc_smooth)[i] - DF(qfrc_constraint)[i];
    grad[i] = g;
    gn += g * g;
  }
  gn = group_sum(c, gn);
  SYNC();
  return gn;
}

// H = M + J^T D_active J;  Mgrad = H^-1 grad  (grad from bp_newton_grad)
//
// The Hessian is assembled from the structure of J instead of a dense triple loop:
// friction-loss and limit rows (`nsingle` rows after the equalities) have one non-zero,
// +-1, and add their weight to one diagonal entry; the rows of a contact are non-zero only on
// the dofs that move its two bodies (body_dofmask), so a contact updates the
// k (k + 1) / 2 entries of that dof set, split over the lanes by outer dof.
BP_STAGE void bp_newton_gradient(const C& c, int nefc, int nsingle, int ntenlim,
                                  bool refactor) {
  BP_MODEL(m);
  const int nv = m.nv, nvp = m.nvp;
  const int ntri = m.ntri;
  auto H = HOT(s_H, 0), Hdinv = HOT(s_Hdinv, ntri + nv), Mgrad = HOT(s_Mgrad, ntri + 2 * nv);
  auto Md = DF(s_Md), grad = DF(s_grad);
  auto J = DF(efc_J), rw = DF(s_rw), rdof = DF(s_rdof);
  for (int i = c.lane; i < nv; i += C::G) Mgrad[i] = grad[i];
  if (!refactor) {            // same active set as last time: H and its factor stand
    SYNC();
    BP_TL(-1);
    bp_chol_solve(c, H, Hdinv, Mgrad, false);
    BP_TL(8);
    return;
  }
  BP_TL(-1);
  bp_copy(c, H, Md, ntri);
  for (int i = c.lane; i < nv; i += C::G) H[ntri + i] = grad[i];     // bordered rhs row
  SYNC();
  const int ne = IF(ne)[0];
#pragma unroll (C::UF)
  for (int i = c.lane; i < nv; i += C::G) {
    double acc = 0;
    for (int r = ne; r < ne + nsingle; r++)
      if ((int)rdof[r] == i) acc += rw[r];
    if (acc != 0) H[PK(i, i)] += acc;
    // equality rows (the first ne) and tendon-limit rows (after the single-entry rows): a
    // few non-zeros each, row i of the lower triangle
#pragma unroll 1
    if constexpr (!C::LEAN) for (int r = 0; r < ne + nsingle + ntenlim; r++) {
      if (r == ne) r += nsingle;
      if (r >= ne + nsingle + ntenlim) break;
      const double ji = J[r * nvp + i];
      if (ji == 0) continue;
      const double w = rw[r] * ji;
      for (int j = 0; j <= i; j++) {
        const double jj = J[r * nvp + j];
        if (jj != 0) H[PK(i, j)] += w * jj;
      }
    }
  }
  SYNC();
  BP_TL(4);
  const int ncon = IF(ncon)[0];
#pragma unroll 1
  for (int ci = 0; ci < ncon; ci++) {
    const int a = IF(contact_efc_address)[ci];
    if (a < 0) continue;
    const int dim = IF(contact_dim)[ci];
    int n = dim == 1 ? 1 : (m.opt_cone == 0 ? 2 * (dim - 1) : dim);
    if (a + n > nefc) n = nefc - a;
    bool any = false;
    for (int q = 0; q < n; q++) any |= rw[a + q] != 0;
    if (!any) continue;                                   // uniform across the group
    const unsigned long long mk = m.body_dofmask[m.geom_bodyid[IF(contact_geom1)[ci]]] |
                                  m.body_dofmask[m.geom_bodyid[IF(contact_geom2)[ci]]];
    const int k = __popcll(mk);
#pragma unroll 1
    for (int q0 = 0; q0 < n; q0 += 4) {                   // rows in chunks of 4
      const int nq = n - q0 < 4 ? n - q0 : 4;
      const int r0 = (a + q0) * nvp;
      // the k (k + 1) / 2 pairs (i >= j) of the dof set are tiled over a GA x GB grid of
      // lanes: lane (la, lb) takes the outer dofs la, la + GA, ... and, for each, the inner
      // dofs lb, lb + GB, ... up to the outer one
      constexpr int GB = C::G >= 16 ? 4 : (C::G >= 4 ? 2 : 1), GA = C::G / GB;
      const int la = c.lane / GB, lb = c.lane % GB;
      unsigned long long remi = mk;
      for (int t = 0; t < la; t++) remi &= remi - 1;
      unsigned long long remj0 = mk;
      for (int t = 0; t < lb; t++) remj0 &= remj0 - 1;
#pragma unroll 1
      for (int ap = la; ap < k; ap += GA) {
        const int i = __ffsll((long long)remi) - 1;
        double ji[4];
#pragma unroll
        for (int q = 0; q < 4; q++) ji[q] = q < nq ? rw[a + q0 + q] * J[r0 + q * nvp + i] : 0.0;
        unsigned long long remj = remj0;
#pragma unroll 1
        for (int bp = lb; bp <= ap; bp += GB) {
          const int j = __ffsll((long long)remj) - 1;
#pragma unroll
          for (int t = 0; t < GB; t++) remj &= remj - 1;
          double sum = 0;
#pragma unroll
          for (int q = 0; q < 4; q++)
            if (q < nq) sum += ji[q] * J[r0 + q * nvp + j];
          H[PK(i, j)] += sum;
        }
#pragma unroll
        for (int t = 0; t < GA; t++) remi &= remi - 1;
      }
      SYNC();
    }
  }
  if constexpr (C::FULL) if (m.opt_cone == 1) bp_cone_hessian(c, H, nefc);
  BP_TL(5);
  if (!bp_chol_factor(c, H, Hdinv)) {
    SYNC();
    bp_copy(c, H, Md, ntri);
    for (int i = c.lane; i < nv; i += C::G) H[ntri + i] = grad[i];
    SYNC();
    bp_chol_factor(c, H, Hdinv);
  }
  BP_TL(6);
  bp_chol_solve(c, H, Hdinv, Mgrad, bp_chol_fused<C>(nv));
  BP_TL(7);
}

// CG [MJ-doc A.9]: the primal gradient preconditioned by the sparse factor of M
BP_STAGE void bp_cg_gradient(const C& c) {
  BP_MODEL(m);
  auto Mg = HOT(s_Mgrad, m.ntri + 2 * m.nv), grad = DF(s_grad);
  for (int i = c.lane; i < m.nv; i += C::G) Mg[i] = grad[i];
  SYNC();
  bp_solve_sparse(c, DF(qLD), DF(qLDiagInv), Mg);
}

// PGS [MJ-doc A.9]: dual projected Gauss-Seidel on 0.5 f^T AR f + f^T b with
// AR = J M^-1 J^T + diag(R), b = J qacc_smooth - aref; rows in order, equality rows free,
// friction-loss rows clamped to +-frictionloss, limit / contact rows to f >= 0 (pyramidal
// cones only).  The warm start maps qacc_warmstart to forces through the primal update and is
// kept when its dual cost is negative (f = 0 costs 0).  Rows are sequential by nature; the
// row dot products run on the lanes.  oracle/mjstep_oracle.c pgs_solve is the restatement
// this is held to.
BP_STAGE void bp_solver_pgs(const C& c) {
  BP_MODEL(m);
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
  auto AR = DF(s_AR), b = DF(s_jv), x = DF(s_grad), f = DF(efc_force), J = DF(efc_J);
  auto R = DF(efc_R), jar = DF(s_jar);
  auto type = IF(efc_type), state = IF(efc_state);
#pragma unroll 1
  for (int i = 0; i < nefc; i++) {
    for (int k = c.lane; k < nv; k += C::G) x[k] = J[i * nvp + k];
    SYNC();
    bp_solve_sparse(c, DF(qLD), DF(qLDiagInv), x);
#pragma unroll 1
    for (int j = c.lane; j < nefc; j += C::G) {
      double s = 0;
      for (int k = 0; k < nv; k++) s += J[j * nvp + k] * x[k];
      AR[i * nefc + j] = j == i ? s + R[i] : s;
    }
    SYNC();
  }
  bp_mul_j(c, b, as, DF(efc_aref), nefc);
  bool use_warm = false;
  if (!(m.opt_disableflags & DSBL_WARMSTART)) {
    for (int i = c.lane; i < nv; i += C::G) qacc[i] = warm[i];
    SYNC();
    bp_mul_j(c, jar, qacc, DF(efc_aref), nefc);
    bool chg;
    bp_update_constraint(c, jar, nefc, true, &chg);
    double cost = 0;
#pragma unroll 1
    for (int i = c.lane; i < nefc; i += C::G) {
      double s = 0;
      for (int j = 0; j < nefc; j++) s += AR[i * nefc + j] * f[j];
      cost += f[i] * (0.5 * s + b[i]);
    }
    use_warm = group_sum(c, cost) < 0;
  }
  if (!use_warm) {
    for (int i = c.lane; i < nefc; i += C::G) f[i] = 0;
  }
  SYNC();
  const double scale = 1.0 / (m.stat_meaninertia * (nv > 1 ? nv : 1));
  int iter = 0;
  double improvement = 0;
#pragma unroll 1
  while (iter < m.opt_iterations) {
    improvement = 0;
#pragma unroll 1
    for (int i = 0; i < nefc; i++) {
      double res = 0;
      for (int j = c.lane; j < nefc; j += C::G) res += AR[i * nefc + j] * f[j];
      res = b[i] + group_sum(c, res);
      const double aii = AR[i * nefc + i], old = f[i];
      double fn = old - res / aii;
      const int t = type[i];
      if (t == CN_FRICTION_DOF || t == CN_FRICTION_TENDON) {
        const double fl = DF(efc_frictionloss)[i];
        fn = fn < -fl ? -fl : (fn > fl ? fl : fn);
      } else if (t != CN_EQUALITY) {
        if (fn < 0) fn = 0;
      }
      const double delta = fn - old;
      SYNC();                       // every lane has read f[i]
      if (c.lane == 0) f[i] = fn;
      SYNC();
      improvement -= 0.5 * delta * delta * aii + delta * res;
    }
    improvement *= scale;
    iter++;
    if (improvement < m.opt_tolerance) break;
  }
  for (int i = c.lane; i < nefc; i += C::G) {
    const int t = type[i];
    const double fi = f[i];
    int st;
    if (t == CN_EQUALITY) st = ST_QUADRATIC;
    else if (t == CN_FRICTION_DOF || t == CN_FRICTION_TENDON) {
      const double fl = DF(efc_frictionloss)[i];
      st = fi >= fl ? ST_LINEARNEG : (fi <= -fl ? ST_LINEARPOS : ST_QUADRATIC);
    } else st = fi > 0 ? ST_QUADRATIC : ST_SATISFIED;
    state[i] = st;
  }
  auto qc = DF(qfrc_constraint);
#pragma unroll (C::UF)
  for (int i = c.lane; i < nv; i += C::G) {
    double s = 0;
    for (int r = 0; r < nefc; r++) s += J[r * nvp + i] * f[r];
    qc[i] = s; x[i] = s;
  }
  SYNC();
  bp_solve_sparse(c, DF(qLD), DF(qLDiagInv), x);
  for (int i = c.lane; i < nv; i += C::G) {
    double a = as[i] + x[i];
    qacc[i] = a; warm[i] = a;
  }
  if (c.lane == 0) {
    IF(solver_iter)[0] = iter;
    DF(solver_stat)[0] = improvement; DF(solver_stat)[1] = 0;
  }
  SYNC();
}

// Solver state that lives across Newton iterations (registers of the calling kernel)
struct BpSolve {
  double cost, improvement, gradient;
  int iter, nefc, nsingle, ntenlim;
  bool done, stale;      // stale: the factorisation does not match the active set
  double cg_den;         // CG: grad . Mgrad of the previous iterate (FULL shape only)
};

// everything up to the first search direction; st.done when there is nothing to iterate
BP_STAGE void bp_solver_begin(const C& c, BpSolve& st) {
  BP_MODEL(m);
  const int nv = m.nv, nvp = m.nvp, nefc = IF(nefc)[0];
  auto qacc = DF(qacc), warm = DF(qacc_warmstart), as = DF(qacc_smooth);
  st.iter = 0; st.improvement = 0; st.gradient = 0; st.nefc = nefc; st.nsingle = 0; st.ntenlim = 0;
  st.stale = false;
  st.cost = 0;
  if (nefc == 0) {
    for (int i = c.lane; i < nv; i += C::G) {
      double a = as[i];
      qacc[i] = a; warm[i] = a; DF(qfrc_constraint)[i] = 0;
    }
    if (c.lane == 0) IF(solver_iter)[0] = 0;
    SYNC();
    st.done = true;
    return;
  }
  // friction-loss and limit rows come first and have a single +-1 entry: remember
  // its dof; `nsingle` = number of such rows
  int nsingle = 0, ntenlim = 0;
#pragma unroll 1
  for (int r = c.lane; r < nefc; r += C::G) {
    int t = IF(efc_type)[r], id = IF(efc_id)[r];
    if (t == CN_FRICTION_DOF) { DF(s_rdof)[r] = id; nsingle++; }
    else if (t == CN_LIMIT_JOINT) { DF(s_rdof)[r] = m.jnt_dofadr[id]; nsingle++; }
    else if (t == CN_LIMIT_TENDON) ntenlim++;
  }
  nsingle = (int)(group_sum(c, (double)nsingle) + 0.5);
  if (C::FULL && m.ntendon > 0) ntenlim = (int)(group_sum(c, (double)ntenlim) + 0.5);
  st.nsingle = nsingle; st.ntenlim = ntenlim;
  // dense copy of M as a packed lower triangle
  auto Md = DF(s_Md), qM = DF(qM);
  for (int w = c.lane; w < m.ntri; w += C::G) Md[w] = 0;
  SYNC();
#pragma unroll (C::UF)
  for (int i = c.lane; i < nv; i += C::G) {
    int a = m.dof_Madr[i];
    const int ri = i * (i + 1) / 2;
    for (int j = i; j >= 0; j = m.dof_parentid[j], a++) Md[ri + j] = qM[a];
  }
  SYNC();
  auto Ma = DF(s_Ma), jar = DF(s_jar), search = DF(s_search);
  bool use_warm = false;
  if (!(m.opt_disableflags & DSBL_WARMSTART)) {
    for (int i = c.lane; i < nv; i += C::G) qacc[i] = warm[i];
    SYNC();
    bp_mul_md(c, Md, Ma, qacc);
    bp_mul_j(c, jar, qacc, DF(efc_aref), nefc);
    double cw = bp_update_constraint(c, jar, nefc, false, nullptr) + bp_gauss_cost(c);
    // residual at qacc_smooth goes to s_jv so that Ma / jar of the warm start survive
    bp_mul_j(c, DF(s_jv), as, DF(efc_aref), nefc);
    double cs = bp_update_constraint(c, DF(s_jv), nefc, false, nullptr);
    use_warm = cw < cs;
    SYNC();
  }
  if (!use_warm) {
    for (int i = c.lane; i < nv; i += C::G) qacc[i] = as[i];
    SYNC();
    bp_mul_md(c, Md, Ma, qacc);
    bp_mul_j(c, jar, qacc, DF(efc_aref), nefc);
  }
  bool changed;
  st.cost = bp_update_constraint(c, jar, nefc, true, &changed) + bp_gauss_cost(c);
  bp_newton_grad(c);
  if (C::FULL && m.opt_solver == 1) bp_cg_gradient(c);
  else bp_newton_gradient(c, nefc, nsingle, ntenlim, true);
  for (int i = c.lane; i < nv; i += C::G) search[i] = -HOT(s_Mgrad, m.ntri + 2 * nv)[i];
  SYNC();
  st.done = m.opt_iterations <= 0;
}

// One Newton iteration in two halves (the block re-aligns between them):
//   step:      line search along s_search, state update, cost, gradient, convergence test
//   direction: new search direction -- only for environments that go on
BP_STAGE void bp_solver_step(const C& c, BpSolve& st) {
  BP_MODEL(m);
  const int nv = m.nv, nefc = st.nefc;
  auto qacc = DF(qacc), Ma = DF(s_Ma), jar = DF(s_jar), search = DF(s_search);
  const double scale = 1.0 / (m.stat_meaninertia * (nv > 1 ? nv : 1));
  BP_TL(-1);
  bp_mul_md(c, DF(s_Md), DF(s_Mv), search);
  bp_mul_j(c, DF(s_jv), search, c.null(), nefc);
  BP_TL(0);
  double alpha = bp_linesearch(c, scale, nefc);
  BP_TL(1);
  if (alpha == 0) { st.done = true; return; }
  for (int i = c.lane; i < nv; i += C::G) {
    qacc[i] += alpha * search[i];
    Ma[i] += alpha * DF(s_Mv)[i];
  }
  for (int r = c.lane; r < nefc; r += C::G) jar[r] += alpha * DF(s_jv)[r];
  SYNC();
  double old = st.cost;
  bool changed;
  if constexpr (C::FULL) if (m.opt_solver == 1) {   // Polak-Ribiere needs the previous pair
    double den = 0;
    auto Mg = HOT(s_Mgrad, m.ntri + 2 * nv), Mgold = DF(s_tmp);
    for (int i = c.lane; i < nv; i += C::G) { double v = Mg[i]; den += DF(s_grad)[i] * v; Mgold[i] = v; }
    st.cg_den = group_sum(c, den);
    SYNC();
  }
  st.cost = bp_update_constraint(c, jar, nefc, true, &changed) + bp_gauss_cost(c);
  BP_TL(2);
  double gn = bp_newton_grad(c);
  BP_TL(3);
  st.improvement = scale * (old - st.cost); st.gradient = scale * sqrt(gn);
  st.iter++;
  st.stale |= changed;
  // the search direction (assembly, factorisation, solve) is only needed to go on
  if (st.improvement < m.opt_tolerance || st.gradient < m.opt_tolerance ||
      st.iter >= m.opt_iterations) st.done = true;
}

BP_STAGE void bp_solver_direction(const C& c, BpSolve& st) {
  BP_MODEL(m);
  if constexpr (C::FULL) if (m.opt_solver == 1) {     // CG: search = -Mgrad + beta search
    bp_cg_gradient(c);
    auto Mg = HOT(s_Mgrad, m.ntri + 2 * m.nv), Mgold = DF(s_tmp), sr = DF(s_search);
    double num = 0;
    for (int i = c.lane; i < m.nv; i += C::G) num += DF(s_grad)[i] * (Mg[i] - Mgold[i]);
    num = group_sum(c, num);
    double beta = num / (st.cg_den > BP_MINVAL ? st.cg_den : BP_MINVAL);
    if (beta < 0) beta = 0;
    for (int i = c.lane; i < m.nv; i += C::G) sr[i] = -Mg[i] + beta * sr[i];
    SYNC();
    return;
  }
  bp_newton_gradient(c, st.nefc, st.nsingle, st.ntenlim, st.stale);
  st.stale = false;
  auto search = DF(s_search);
  for (int i = c.lane; i < m.nv; i += C::G) search[i] = -HOT(s_Mgrad, m.ntri + 2 * m.nv)[i];
  SYNC();
}

template <class C>
__device__ __forceinline__ void bp_solver_iterate(const C& c, BpSolve& st) {
  bp_solver_step(c, st);
  if (!st.done) bp_solver_direction(c, st);
}

BP_STAGE void bp_solver_end(const C& c, const BpSolve& st) {
  if (st.nefc == 0) return;
  const int nv = bp_mv(c).nv;
  if (c.lane == 0) {
    IF(solver_iter)[0] = st.iter;
    DF(solver_stat)[0] = st.improvement; DF(solver_stat)[1] = st.gradient;
  }
  for (int i = c.lane; i < nv; i += C::G) DF(qacc_warmstart)[i] = DF(qacc)[i];
  SYNC();
}

// mj_fwdConstraint.  In the lock-step shapes the Newton loop runs until every
// environment of the block has converged, and the block's warps re-align at every
// iteration (environments that are done skip the body).
template <class C>
__device__ __forceinline__ void bp_fwd_constraint(const C& c) {
  if constexpr (C::FULL) if (bp_mv(c).opt_solver == 0) {    // uniform per launch
    RUN(bp_solver_pgs(c));
    return;
  }
  BpSolve st;
  st.done = true; st.nefc = 0; st.cg_den = 0;
  RUN(bp_solver_begin(c, st));
  if (c.lock && (BP_LOCKMASK >> 8 & 1)) {
#if BP_NEWTON_GROUP > 0
    // the Newton loop re-aligns groups of BP_NEWTON_GROUP warps (named barriers 1..15), not
    // the whole block: a group iterates as long as ITS slowest environment
    const int warp = threadIdx.x >> 5, nwarp = (blockDim.x + 31) >> 5;
    const int grp = warp / BP_NEWTON_GROUP;
    const int gcount = min(BP_NEWTON_GROUP, nwarp - grp * BP_NEWTON_GROUP) * 32;
    while (bp_bar_or(1 + grp, gcount, !st.done)) {
      if (!st.done) bp_solver_iterate(c, st);
    }
  } else if (false) {
#endif
    while (__syncthreads_or(!st.done)) {
#ifdef BP_TIMELINE
      if (blockIdx.x == 0 && threadIdx.x == 0) bp_timeline[16 + 8]++;
#endif
      if (!st.done) bp_solver_iterate(c, st);
    }
  } else {
#pragma unroll 1
    while (!st.done) bp_solver_iterate(c, st);
  }
  RUN(bp_solver_end(c, st));
}

#endif  // BP_SOLVER_CUH_
