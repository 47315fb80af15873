// mj_collision: static candidate pairs on lanes, primitive narrowphase, contacts
// emitted in pair order through a group prefix sum so that contact ids are
// reproducible for every group size.
#ifndef BP_COLLIDE_CUH_
#define BP_COLLIDE_CUH_
#include "bp_math.cuh"
#include "bp_hfield.cuh"

__device__ __forceinline__ int col_plane_sphere(const double* ppos, const double* pn,
                                                const double* ctr, double r,
                                                double margin, RawCon& o) {
  double dif[3] = {ctr[0] - ppos[0], ctr[1] - ppos[1], ctr[2] - ppos[2]};
  double dist = dot3(dif, pn) - r;
  if (dist > margin) return 0;
  o.dist = dist;
  for (int k = 0; k < 3; k++) {
    o.pos[k] = ctr[k] - pn[k] * (r + 0.5 * dist);
    o.n[k] = pn[k];
    o.hint[k] = 0;
  }
  return 1;
}

__device__ __forceinline__ int col_sphere_sphere(const double* c1, double r1,
                                                 const double* c2, double r2,
                                                 double margin, RawCon& o) {
  double n[3] = {c2[0] - c1[0], c2[1] - c1[1], c2[2] - c1[2]};
  double len = norm3(n), dist = len - r1 - r2;
  if (dist > margin) return 0;
  if (len < BP_MINVAL) { n[0] = 1; n[1] = 0; n[2] = 0; }
  else { n[0] /= len; n[1] /= len; n[2] /= len; }
  o.dist = dist;
  for (int k = 0; k < 3; k++) {
    o.pos[k] = c1[k] + n[k] * (r1 + 0.5 * dist);
    o.n[k] = n[k];
    o.hint[k] = 0;
  }
  return 1;
}

// closest point of a solid cylinder (centre p, unit axis a, radius r, half length h) to q
__device__ __forceinline__ double cyl_closest(const double* p, const double* a, double r,
                                              double h, const double* q, double* cl) {
  double dq[3] = {q[0] - p[0], q[1] - p[1], q[2] - p[2]};
  double x = dot3(dq, a);
  double w[3] = {dq[0] - x * a[0], dq[1] - x * a[1], dq[2] - x * a[2]};
  double rho = norm3(w);
  double xc = clampd(x, -h, h);
  double sc = rho > r ? r / rho : 1.0;
  for (int k = 0; k < 3; k++) cl[k] = p[k] + xc * a[k] + sc * w[k];
  double e[3] = {q[0] - cl[0], q[1] - cl[1], q[2] - cl[2]};
  return norm3(e);
}

// sphere (geom 1) against a cylinder (geom 2): along the closest-point direction; a
// centre inside the cylinder leaves through the nearer of the side and the caps
__device__ __forceinline__ int col_sphere_cylinder(const double* q, double rs, const double* p,
                                                   const double* a, double r, double h,
                                                   double margin, RawCon& o) {
  double cl[3], nrm[3], dist;
  double dd = cyl_closest(p, a, r, h, q, cl);
  if (dd > BP_MINVAL) {
    for (int k = 0; k < 3; k++) nrm[k] = (cl[k] - q[k]) / dd;
    dist = dd - rs;
  } else {
    double dq[3] = {q[0] - p[0], q[1] - p[1], q[2] - p[2]};
    double x = dot3(dq, a);
    double w[3] = {dq[0] - x * a[0], dq[1] - x * a[1], dq[2] - x * a[2]};
    double rho = norm3(w);
    if (r - rho < h - fabs(x) && rho > BP_MINVAL) {
      for (int k = 0; k < 3; k++) nrm[k] = -w[k] / rho;
      dist = -(r - rho) - rs;
    } else {
      double sg = x >= 0 ? 1.0 : -1.0;
      for (int k = 0; k < 3; k++) nrm[k] = -sg * a[k];
      dist = -(h - fabs(x)) - rs;
    }
  }
  if (dist > margin) return 0;
  o.dist = dist;
  for (int k = 0; k < 3; k++) {
    o.pos[k] = q[k] + nrm[k] * (rs + 0.5 * dist);
    o.n[k] = nrm[k];
    o.hint[k] = 0;
  }
  return 1;
}

// Pairs whose second geom is a cylinder.  plane: analytic, up to four rim points.
// sphere: closest point.  capsule: the distance from a point of the capsule's segment to
// the cylinder is convex along the segment, so the closest point is bracketed by
// bisection on the sign of the derivative.  cylinder - cylinder: geom 1 is replaced by
// its enclosing capsule (exact while the two are apart).
static __device__ __noinline__ int bp_collide_cylinder(int t1, const double* p1, const double* z1,
                                                       const double* s1, const double* p2,
                                                       const double* R2, const double* s2,
                                                       double margin, RawCon* out) {
  double a2[3] = {R2[2], R2[5], R2[8]};
  if (t1 == GEOM_PLANE) {
    double axis[3] = {a2[0], a2[1], a2[2]}, vec[3];
    double prjaxis = dot3(z1, axis);
    if (prjaxis > 0) { for (int k = 0; k < 3; k++) axis[k] = -axis[k]; prjaxis = -prjaxis; }
    double dif[3] = {p2[0] - p1[0], p2[1] - p1[1], p2[2] - p1[2]};
    double dist0 = dot3(dif, z1);
    for (int k = 0; k < 3; k++) vec[k] = axis[k] * prjaxis - z1[k];
    double len = norm3(vec);
    if (len >= BP_MINVAL) { for (int k = 0; k < 3; k++) vec[k] *= s2[0] / len; }
    else { vec[0] = R2[0] * s2[0]; vec[1] = R2[3] * s2[0]; vec[2] = R2[6] * s2[0]; }
    double prjvec = dot3(vec, z1);
    for (int k = 0; k < 3; k++) axis[k] *= s2[1];
    prjaxis *= s2[1];
    if (dist0 + prjaxis + prjvec > margin) return 0;
    int cnt = 0;
    double dd = dist0 + prjaxis + prjvec;
    out[cnt].dist = dd;
    for (int k = 0; k < 3; k++) {
      out[cnt].pos[k] = p2[k] + vec[k] + axis[k] - z1[k] * dd * 0.5;
      out[cnt].n[k] = z1[k]; out[cnt].hint[k] = 0;
    }
    cnt++;
    dd = dist0 - prjaxis + prjvec;
    if (dd <= margin) {
      out[cnt].dist = dd;
      for (int k = 0; k < 3; k++) {
        out[cnt].pos[k] = p2[k] + vec[k] - axis[k] - z1[k] * dd * 0.5;
        out[cnt].n[k] = z1[k]; out[cnt].hint[k] = 0;
      }
      cnt++;
    }
    double prjvec1 = -prjvec * 0.5;
    dd = dist0 + prjaxis + prjvec1;
    if (dd <= margin) {
      double vec1[3];
      cross3(vec1, vec, axis);
      normalize3(vec1);
      for (int k = 0; k < 3; k++) vec1[k] *= s2[0] * sqrt(3.0) * 0.5;
      for (int sgn = 1; sgn >= -1; sgn -= 2) {
        out[cnt].dist = dd;
        for (int k = 0; k < 3; k++) {
          out[cnt].pos[k] = p2[k] + sgn * vec1[k] + axis[k] - vec[k] * 0.5 - z1[k] * dd * 0.5;
          out[cnt].n[k] = z1[k]; out[cnt].hint[k] = 0;
        }
        cnt++;
      }
    }
    return cnt;
  }
  if (t1 == GEOM_SPHERE)
    return col_sphere_cylinder(p1, s1[0], p2, a2, s2[0], s2[1], margin, out[0]);
  if (t1 == GEOM_CAPSULE || t1 == GEOM_CYLINDER) {
    double lo = -s1[1], hi = s1[1], q[3], cl[3];
#pragma unroll 1
    for (int it = 0; it < 64; it++) {
      double t = 0.5 * (lo + hi);
      for (int k = 0; k < 3; k++) q[k] = p1[k] + t * z1[k];
      cyl_closest(p2, a2, s2[0], s2[1], q, cl);
      double g = (q[0] - cl[0]) * z1[0] + (q[1] - cl[1]) * z1[1] + (q[2] - cl[2]) * z1[2];
      if (g >= 0) hi = t; else lo = t;
    }
    double t = 0.5 * (lo + hi);
    for (int k = 0; k < 3; k++) q[k] = p1[k] + t * z1[k];
    int n = col_sphere_cylinder(q, s1[0], p2, a2, s2[0], s2[1], margin, out[0]);
    if (n) for (int k = 0; k < 3; k++) out[0].hint[k] = z1[k];
    return n;
  }
  return 0;
}

// ---- ellipsoids (MuJoCo: general convex routine; here the closest point directly) --------
// Point q in the ellipsoid's frame, semi-axes a: the closest surface point is
// x_i = a_i^2 q_i / (t + a_i^2), t >= 0 the root of F(t) = sum (a_i q_i / (t + a_i^2))^2 - 1
// (convex and decreasing, so Newton from the left never overshoots).  t carries the
// multiplier between neighbouring queries.  Returns the distance, 0 inside.
__device__ __forceinline__ double ell_closest(const double* a, const double* q, double* x, double& t) {
  double f0 = 0;
  for (int k = 0; k < 3; k++) f0 += (q[k] / a[k]) * (q[k] / a[k]);
  if (f0 <= 1) { for (int k = 0; k < 3; k++) x[k] = q[k]; t = 0; return 0; }
  double tt = t > 0 ? t : 0;
#pragma unroll 1
  for (int it = 0; it < 100; it++) {
    double F = -1, dF = 0;
    for (int k = 0; k < 3; k++) {
      double den = tt + a[k] * a[k], r = a[k] * q[k] / den;
      F += r * r; dF -= 2 * r * r / den;
    }
    if (F < 0 && it == 0 && tt > 0) { tt = 0; continue; }
    double step = -F / dF;
    tt += step;
    if (fabs(step) <= 1e-16 * (tt + a[0] * a[0])) break;
  }
  double dd = 0;
  for (int k = 0; k < 3; k++) {
    x[k] = a[k] * a[k] * q[k] / (tt + a[k] * a[k]);
    dd += (q[k] - x[k]) * (q[k] - x[k]);
  }
  t = tt;
  return sqrt(dd);
}

__device__ __forceinline__ int col_sphere_ellipsoid(const double* qw, double rs, const double* p,
                                                    const double* R, const double* a,
                                                    double margin, double& tmul, RawCon& o) {
  double dq[3] = {qw[0] - p[0], qw[1] - p[1], qw[2] - p[2]}, q[3], x[3], nl[3], dist;
  matT_vec3(q, R, dq);
  double dd = ell_closest(a, q, x, tmul);
  if (dd > BP_MINVAL) {
    for (int k = 0; k < 3; k++) nl[k] = (x[k] - q[k]) / dd;
    dist = dd - rs;
  } else {
    double s = 0;
    for (int k = 0; k < 3; k++) s += (q[k] / a[k]) * (q[k] / a[k]);
    s = sqrt(s);
    double g[3], gl = 0, pen = 0;
    for (int k = 0; k < 3; k++) {
      double xs = s > BP_MINVAL ? q[k] / s : (k == 2 ? a[2] : 0.0);
      g[k] = xs / (a[k] * a[k]); gl += g[k] * g[k];
      pen += (xs - q[k]) * (xs - q[k]);
    }
    gl = sqrt(gl);
    for (int k = 0; k < 3; k++) nl[k] = -g[k] / gl;
    dist = -sqrt(pen) - rs;
  }
  if (dist > margin) return 0;
  double nw[3];
  mat_vec3(nw, R, nl);
  o.dist = dist;
  for (int k = 0; k < 3; k++) {
    o.pos[k] = qw[k] + nw[k] * (rs + 0.5 * dist);
    o.n[k] = nw[k];
    o.hint[k] = 0;
  }
  return 1;
}

// slope of the distance from the capsule's segment point to the ellipsoid
__device__ __forceinline__ double cap_ell_slope(const double* pc, const double* al, double t,
                                                const double* ac, const double* p, const double* R,
                                                const double* a, double& tmul) {
  double dq[3] = {pc[0] + t * ac[0] - p[0], pc[1] + t * ac[1] - p[1], pc[2] + t * ac[2] - p[2]};
  double q[3], x[3];
  matT_vec3(q, R, dq);
  double dd = ell_closest(a, q, x, tmul);
  if (dd <= BP_MINVAL) return 0;
  return ((q[0] - x[0]) * al[0] + (q[1] - x[1]) * al[1] + (q[2] - x[2]) * al[2]) / dd;
}

// Pairs whose second geom is an ellipsoid: plane (support point along -n), sphere
// (closest point), capsule (the distance along the segment is convex: Illinois regula
// falsi on its slope, then a sphere at the closest segment point).
static __device__ __noinline__ int bp_collide_ellipsoid(int t1, const double* p1, const double* z1,
                                                        const double* s1, const double* p2,
                                                        const double* R2, const double* a,
                                                        double margin, RawCon* out) {
  if (t1 == GEOM_PLANE) {
    double dl[3], x[3], xw[3], s = 0;
    matT_vec3(dl, R2, z1);
    for (int k = 0; k < 3; k++) s += a[k] * a[k] * dl[k] * dl[k];
    s = sqrt(s);
    for (int k = 0; k < 3; k++) x[k] = -a[k] * a[k] * dl[k] / s;
    mat_vec3(xw, R2, x);
    double dif[3] = {p2[0] + xw[0] - p1[0], p2[1] + xw[1] - p1[1], p2[2] + xw[2] - p1[2]};
    double dist = dot3(dif, z1);
    if (dist > margin) return 0;
    out[0].dist = dist;
    for (int k = 0; k < 3; k++) {
      out[0].pos[k] = p2[k] + xw[k] - z1[k] * dist * 0.5;
      out[0].n[k] = z1[k]; out[0].hint[k] = 0;
    }
    return 1;
  }
  double tmul = 0;
  if (t1 == GEOM_SPHERE) return col_sphere_ellipsoid(p1, s1[0], p2, R2, a, margin, tmul, out[0]);
  if (t1 == GEOM_CAPSULE) {
    const double l = s1[1];
    double al[3];
    matT_vec3(al, R2, z1);
    double lo = -l, hi = l;
    double glo = cap_ell_slope(p1, al, lo, z1, p2, R2, a, tmul);
    double ghi = cap_ell_slope(p1, al, hi, z1, p2, R2, a, tmul);
    double t;
    if (glo >= 0) t = lo;
    else if (ghi <= 0) t = hi;
    else {
      int side = 0;
      t = 0.5 * (lo + hi);
#pragma unroll 1
      for (int it = 0; it < 100; it++) {
        t = (lo * ghi - hi * glo) / (ghi - glo);
        double g = cap_ell_slope(p1, al, t, z1, p2, R2, a, tmul);
        if (fabs(g) < 1e-15 || hi - lo < 1e-15 * l) break;
        if (g > 0) { hi = t; ghi = g; if (side == 1) glo *= 0.5; side = 1; }
        else { lo = t; glo = g; if (side == -1) ghi *= 0.5; side = -1; }
      }
    }
    double q[3] = {p1[0] + t * z1[0], p1[1] + t * z1[1], p1[2] + t * z1[2]};
    int n = col_sphere_ellipsoid(q, s1[0], p2, R2, a, margin, tmul, out[0]);
    if (n) for (int k = 0; k < 3; k++) out[0].hint[k] = z1[k];
    return n;
  }
  return 0;
}

// ---- box narrowphase: sphere / capsule / box against a box (FULL shape only) ------------
// [MJ-doc] MuJoCo's mjc_SphereBox / mjc_CapsuleBox / mjc_BoxBox are not in the reference;
// the geometry is restated from first principles (oracle/mjstep_oracle.c holds the same
// restatement):
//   sphere - box   closest point of the solid box; a centre inside leaves through the
//                  nearest face
//   capsule - box  squared distance from the segment to the box is convex along the segment:
//                  the ends of its minimiser set (bisection on the derivative's sign) and the
//                  segment ends are candidate sphere centres; the deepest hit and the hit
//                  farthest from it along the axis are kept (<= 2, ascending along the axis)
//   box - box      separating-axis test (15 axes, faces preferred); face case: incident face
//                  clipped against the reference face (<= 8, clipping order); edge case:
//                  closest points of the two edges (1)
static __device__ __noinline__ int col_sphere_box(const double* c, double r, const double* pb,
                                                  const double* Rb, const double* h,
                                                  double margin, RawCon& o) {
  double dif[3] = {c[0] - pb[0], c[1] - pb[1], c[2] - pb[2]}, d[3], del[3], nl[3], pl[3];
  matT_vec3(d, Rb, dif);
  for (int k = 0; k < 3; k++) {
    double q = d[k] < -h[k] ? -h[k] : (d[k] > h[k] ? h[k] : d[k]);
    del[k] = d[k] - q;
  }
  double dd = norm3(del), dist;
  if (dd > 1e-12) {
    for (int k = 0; k < 3; k++) nl[k] = -del[k] / dd;
    dist = dd - r;
  } else {
    int best = 0;
    double depth = h[0] - fabs(d[0]);
    for (int k = 1; k < 3; k++) {
      double dk = h[k] - fabs(d[k]);
      if (dk < depth - 1e-10) { depth = dk; best = k; }
    }
    for (int k = 0; k < 3; k++) nl[k] = k == best ? (d[k] >= 0 ? -1.0 : 1.0) : 0.0;
    dist = -depth - r;
  }
  if (dist > margin) return 0;
  for (int k = 0; k < 3; k++) pl[k] = d[k] + nl[k] * (r + 0.5 * dist);
  o.dist = dist;
  mat_vec3(o.pos, Rb, pl);
  for (int k = 0; k < 3; k++) { o.pos[k] += pb[k]; o.hint[k] = 0; }
  mat_vec3(o.n, Rb, nl);
  return 1;
}

__device__ __forceinline__ double capbox_dg(const double* c0, const double* u, const double* h,
                                            double t) {
  double s = 0;
  for (int k = 0; k < 3; k++) {
    double v = c0[k] + t * u[k];
    double e = v > h[k] ? v - h[k] : (v < -h[k] ? v + h[k] : 0.0);
    s += e * u[k];
  }
  return s;
}
__device__ __forceinline__ double capbox_g(const double* c0, const double* u, const double* h,
                                           double t) {
  double s = 0;
  for (int k = 0; k < 3; k++) {
    double v = c0[k] + t * u[k];
    double e = v > h[k] ? v - h[k] : (v < -h[k] ? v + h[k] : 0.0);
    s += e * e;
  }
  return s;
}

static __device__ __noinline__ int col_capsule_box(const double* p1, const double* a1, double r,
                                                   double l, const double* pb, const double* Rb,
                                                   const double* h, double margin, RawCon* out) {
  double dif[3] = {p1[0] - pb[0], p1[1] - pb[1], p1[2] - pb[2]}, c0[3], u[3];
  matT_vec3(c0, Rb, dif);
  matT_vec3(u, Rb, a1);
  double tlo, thi;
  if (capbox_dg(c0, u, h, -l) >= 0) tlo = -l;
  else if (capbox_dg(c0, u, h, l) < 0) tlo = l;
  else {
    double lo = -l, hi = l;
#pragma unroll 1
    for (int it = 0; it < 60; it++) {
      double t = 0.5 * (lo + hi);
      if (capbox_dg(c0, u, h, t) >= 0) hi = t; else lo = t;
    }
    tlo = hi;
  }
  if (capbox_dg(c0, u, h, -l) > 0) thi = -l;
  else if (!(capbox_dg(c0, u, h, l) > 0)) thi = l;
  else {
    double lo = -l, hi = l;
#pragma unroll 1
    for (int it = 0; it < 60; it++) {
      double t = 0.5 * (lo + hi);
      if (capbox_dg(c0, u, h, t) > 0) hi = t; else lo = t;
    }
    thi = lo;
  }
  double cand[4] = {-l, l, tlo, thi};
  int nc = 4;
  if (!(capbox_g(c0, u, h, tlo) > 1e-24)) { cand[2] = 0.5 * (tlo + thi); nc = 3; }
  RawCon hit[4];
  double th[4];
  int nh = 0;
#pragma unroll 1
  for (int i = 0; i < nc; i++) {
    double q[3] = {p1[0] + cand[i] * a1[0], p1[1] + cand[i] * a1[1], p1[2] + cand[i] * a1[2]};
    if (col_sphere_box(q, r, pb, Rb, h, margin, hit[nh])) {
      for (int k = 0; k < 3; k++) hit[nh].hint[k] = a1[k];
      th[nh++] = cand[i];
    }
  }
  if (!nh) return 0;
  int first = 0, second = -1;
  for (int i = 1; i < nh; i++) if (hit[i].dist < hit[first].dist - 1e-10) first = i;
  double sep = 1e-6 * (l + r);
  for (int i = 0; i < nh; i++) {
    double si = fabs(th[i] - th[first]);
    if (i != first && si > sep) { sep = si; second = i; }
  }
  if (second < 0) { out[0] = hit[first]; return 1; }
  if (th[first] < th[second]) { out[0] = hit[first]; out[1] = hit[second]; }
  else { out[0] = hit[second]; out[1] = hit[first]; }
  return 2;
}

// clips the polygon (x, y, z)[n] against sgn * coord <= lim (coord = x or y)
static __device__ __noinline__ int boxbox_clip(const double* xi, const double* yi,
                                               const double* zi, int n, int usey, double sgn,
                                               double lim, double* xo, double* yo, double* zo) {
  int m = 0;
#pragma unroll 1
  for (int i = 0; i < n; i++) {
    int j = (i + 1) % n;
    double ci = sgn * (usey ? yi[i] : xi[i]) - lim, cj = sgn * (usey ? yi[j] : xi[j]) - lim;
    /* a vertex on the boundary (boxes of equal width side by side) counts as inside on both
     * sides of a rounding error, so that the polygon keeps its vertex order */
    const int in_i = ci <= 1e-10, in_j = cj <= 1e-10;
    if (in_i) { xo[m] = xi[i]; yo[m] = yi[i]; zo[m] = zi[i]; m++; }
    if (in_i != in_j) {
      double w = ci / (ci - cj);
      xo[m] = xi[i] + w * (xi[j] - xi[i]);
      yo[m] = yi[i] + w * (yi[j] - yi[i]);
      zo[m] = zi[i] + w * (zi[j] - zi[i]);
      m++;
    }
  }
  return m;
}

__device__ __forceinline__ void col_axis(const double* R, int col, double* ax) {
  ax[0] = R[col]; ax[1] = R[3 + col]; ax[2] = R[6 + col];
}

static __device__ __noinline__ int col_box_box(const double* pa, const double* Ra,
                                               const double* ha, const double* pb,
                                               const double* Rb, const double* hb, double margin,
                                               RawCon* out) {
  double R[9], Q[9], t[3], dw[3] = {pb[0] - pa[0], pb[1] - pa[1], pb[2] - pa[2]};
  matT_vec3(t, Ra, dw);
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 3; j++) {
      R[3 * i + j] = Ra[i] * Rb[j] + Ra[3 + i] * Rb[3 + j] + Ra[6 + i] * Rb[6 + j];
      Q[3 * i + j] = fabs(R[3 * i + j]);
    }
  double best = -1e300;
  int code = -1;
  for (int i = 0; i < 3; i++) {
    double s = fabs(t[i]) - (ha[i] + hb[0] * Q[3 * i] + hb[1] * Q[3 * i + 1] + hb[2] * Q[3 * i + 2]);
    if (s > margin) return 0;
    if (s > best + 1e-10) { best = s; code = i; }
  }
  for (int j = 0; j < 3; j++) {
    double tb = t[0] * R[j] + t[1] * R[3 + j] + t[2] * R[6 + j];
    double s = fabs(tb) - (hb[j] + ha[0] * Q[j] + ha[1] * Q[3 + j] + ha[2] * Q[6 + j]);
    if (s > margin) return 0;
    if (s > best + 1e-10) { best = s; code = 3 + j; }
  }
#pragma unroll 1
  for (int i = 0; i < 3; i++)
#pragma unroll 1
    for (int j = 0; j < 3; j++) {
      int i1 = (i + 1) % 3, i2 = (i + 2) % 3, j1 = (j + 1) % 3, j2 = (j + 2) % 3;
      double len2 = 1.0 - R[3 * i + j] * R[3 * i + j];
      if (len2 < 1e-12) continue;
      double len = sqrt(len2);
      double tl = fabs(t[i2] * R[3 * i1 + j] - t[i1] * R[3 * i2 + j]);
      double ra = ha[i1] * Q[3 * i2 + j] + ha[i2] * Q[3 * i1 + j];
      double rb = hb[j1] * Q[3 * i + j2] + hb[j2] * Q[3 * i + j1];
      double s = (tl - ra - rb) / len;
      if (s > margin) return 0;
      if (s > best + 0.05 * fabs(best) + 1e-10) { best = s; code = 6 + 3 * i + j; }
    }
  if (code < 6) {
    const bool refA = code < 3;
    const int ai = refA ? code : code - 3;
    const double *pr = refA ? pa : pb, *Rr = refA ? Ra : Rb, *hr = refA ? ha : hb;
    const double *pi = refA ? pb : pa, *Ri = refA ? Rb : Ra, *hi = refA ? hb : ha;
    double ax[3], nref[3], dc[3] = {pi[0] - pr[0], pi[1] - pr[1], pi[2] - pr[2]};
    col_axis(Rr, ai, ax);
    double sg = dot3(dc, ax) > 0 ? 1.0 : -1.0;
    for (int k = 0; k < 3; k++) nref[k] = sg * ax[k];
    int kf = 0;
    double bestdot = -1;
    for (int k = 0; k < 3; k++) {
      double ak[3];
      col_axis(Ri, k, ak);
      double dk = fabs(dot3(nref, ak));
      if (dk > bestdot + 1e-10) { bestdot = dk; kf = k; }
    }
    double af[3], a1[3], a2[3], fc[3];
    const int k1 = (kf + 1) % 3, k2 = (kf + 2) % 3, r1 = (ai + 1) % 3, r2 = (ai + 2) % 3;
    col_axis(Ri, kf, af); col_axis(Ri, k1, a1); col_axis(Ri, k2, a2);
    double sf = dot3(nref, af) > 0 ? -1.0 : 1.0;
    for (int k = 0; k < 3; k++) fc[k] = pi[k] + sf * hi[kf] * af[k] - pr[k];
    double t1[3], t2[3];
    col_axis(Rr, r1, t1); col_axis(Rr, r2, t2);
    double x[2][9], y[2][9], z[2][9];
#pragma unroll 1
    for (int v = 0; v < 4; v++) {
      const double S1 = (v == 0 || v == 3) ? 1.0 : -1.0, S2 = v < 2 ? 1.0 : -1.0;
      double w[3];
      for (int k = 0; k < 3; k++) w[k] = fc[k] + S1 * hi[k1] * a1[k] + S2 * hi[k2] * a2[k];
      x[0][v] = dot3(w, t1); y[0][v] = dot3(w, t2); z[0][v] = dot3(w, nref) - hr[ai];
    }
    int n = 4;
    n = boxbox_clip(x[0], y[0], z[0], n, 0, 1.0, hr[r1], x[1], y[1], z[1]);
    n = boxbox_clip(x[1], y[1], z[1], n, 0, -1.0, hr[r1], x[0], y[0], z[0]);
    n = boxbox_clip(x[0], y[0], z[0], n, 1, 1.0, hr[r2], x[1], y[1], z[1]);
    n = boxbox_clip(x[1], y[1], z[1], n, 1, -1.0, hr[r2], x[0], y[0], z[0]);
    int cnt = 0;
#pragma unroll 1
    for (int v = 0; v < n && cnt < 8; v++) {
      if (z[0][v] > margin) continue;
      out[cnt].dist = z[0][v];
      for (int k = 0; k < 3; k++) {
        out[cnt].pos[k] = pr[k] + x[0][v] * t1[k] + y[0][v] * t2[k] +
                          (hr[ai] + 0.5 * z[0][v]) * nref[k];
        out[cnt].n[k] = refA ? nref[k] : -nref[k];
        out[cnt].hint[k] = 0;
      }
      cnt++;
    }
    return cnt;
  }
  const int i = (code - 6) / 3, j = (code - 6) % 3;
  double ua[3], ub[3], L[3];
  col_axis(Ra, i, ua); col_axis(Rb, j, ub);
  cross3(L, ua, ub);
  normalize3(L);
  if (dot3(L, dw) < 0) { L[0] = -L[0]; L[1] = -L[1]; L[2] = -L[2]; }
  double ea[3] = {pa[0], pa[1], pa[2]}, eb[3] = {pb[0], pb[1], pb[2]};
  for (int k = 0; k < 3; k++) {
    double ak[3], bk[3];
    if (k != i) {
      col_axis(Ra, k, ak);
      double sgk = dot3(L, ak) > 0 ? 1.0 : -1.0;
      for (int q = 0; q < 3; q++) ea[q] += sgk * ha[k] * ak[q];
    }
    if (k != j) {
      col_axis(Rb, k, bk);
      double sgk = dot3(L, bk) > 0 ? -1.0 : 1.0;
      for (int q = 0; q < 3; q++) eb[q] += sgk * hb[k] * bk[q];
    }
  }
  double d[3] = {eb[0] - ea[0], eb[1] - ea[1], eb[2] - ea[2]};
  double uaub = dot3(ua, ub), q1 = dot3(ua, d), q2 = -dot3(ub, d), den = 1.0 - uaub * uaub;
  double al = (q1 + uaub * q2) / den, be = (uaub * q1 + q2) / den;
  al = clampd(al, -ha[i], ha[i]);
  be = clampd(be, -hb[j], hb[j]);
  double ca[3], cb[3], sepv[3];
  for (int k = 0; k < 3; k++) {
    ca[k] = ea[k] + al * ua[k]; cb[k] = eb[k] + be * ub[k]; sepv[k] = cb[k] - ca[k];
  }
  double dist = dot3(sepv, L);
  if (dist > margin) return 0;
  out[0].dist = dist;
  for (int k = 0; k < 3; k++) {
    out[0].pos[k] = 0.5 * (ca[k] + cb[k]); out[0].n[k] = L[k]; out[0].hint[k] = 0;
  }
  return 1;
}

// Narrowphase of ordered pair (type1 <= type2) on geom poses already copied to
// registers (p = centre, z = local z axis, R2 = full frame of geom 2 for boxes).
// Returns the number of raw contacts written to out[0..3].
template <int SHAPE>
static __device__ BP_STAGE_NOINLINE int bp_collide_prims(int t1, int t2, const double* p1,
                                                    const double* z1, const double* s1,
                                                    const double* p2, const double* z2,
                                                    const double* R2, const double* s2,
                                                    double margin, RawCon* out,
                                                    const double* R1) {
  if constexpr (SHAPE == 0) {
    if (t2 == GEOM_BOX && t1 != GEOM_PLANE) {
      if (t1 == GEOM_SPHERE) return col_sphere_box(p1, s1[0], p2, R2, s2, margin, out[0]);
      if (t1 == GEOM_CAPSULE) return col_capsule_box(p1, z1, s1[0], s1[1], p2, R2, s2, margin, out);
      if (t1 == GEOM_BOX) return col_box_box(p1, R1, s1, p2, R2, s2, margin, out);
      return 0;
    }
  }
  if constexpr (SHAPE != 1) {
    if (t2 == GEOM_CYLINDER) return bp_collide_cylinder(t1, p1, z1, s1, p2, R2, s2, margin, out);
    if (t2 == GEOM_ELLIPSOID) return bp_collide_ellipsoid(t1, p1, z1, s1, p2, R2, s2, margin, out);
  }
  if (t1 == GEOM_PLANE) {
    if (t2 == GEOM_SPHERE) return col_plane_sphere(p1, z1, p2, s2[0], margin, out[0]);
    if (t2 == GEOM_CAPSULE) {
      double e[3], hl = s2[1];
      int cnt = 0;
      for (int sgn = 1; sgn >= -1; sgn -= 2) {
        for (int k = 0; k < 3; k++) e[k] = p2[k] + sgn * z2[k] * hl;
        if (col_plane_sphere(p1, z1, e, s2[0], margin, out[cnt])) {
          for (int k = 0; k < 3; k++) out[cnt].hint[k] = z2[k];
          cnt++;
        }
      }
      return cnt;
    }
    if (SHAPE == 0 && t2 == GEOM_BOX) {
      double dif[3] = {p2[0] - p1[0], p2[1] - p1[1], p2[2] - p1[2]};
      double dist = dot3(dif, z1);
      int cnt = 0;
#pragma unroll 1
      for (int i = 0; i < 8 && cnt < 4; i++) {
        double v[3] = {(i & 1) ? s2[0] : -s2[0], (i & 2) ? s2[1] : -s2[1],
                       (i & 4) ? s2[2] : -s2[2]}, cr[3];
        mat_vec3(cr, R2, v);
        double ld = dot3(z1, cr);
        if (dist + ld > margin) continue;
        out[cnt].dist = dist + ld;
        for (int k = 0; k < 3; k++) {
          out[cnt].pos[k] = p2[k] + cr[k] - z1[k] * 0.5 * (dist + ld);
          out[cnt].n[k] = z1[k];
          out[cnt].hint[k] = 0;
        }
        cnt++;
      }
      return cnt;
    }
    return 0;
  }
  if (t1 == GEOM_SPHERE && t2 == GEOM_SPHERE)
    return col_sphere_sphere(p1, s1[0], p2, s2[0], margin, out[0]);
  if (t1 == GEOM_SPHERE && t2 == GEOM_CAPSULE) {
    double dif[3] = {p1[0] - p2[0], p1[1] - p2[1], p1[2] - p2[2]};
    double x = clampd(dot3(z2, dif), -s2[1], s2[1]);
    double cc[3] = {p2[0] + z2[0] * x, p2[1] + z2[1] * x, p2[2] + z2[2] * x};
    return col_sphere_sphere(p1, s1[0], cc, s2[0], margin, out[0]);
  }
  if (t1 == GEOM_CAPSULE && t2 == GEOM_CAPSULE) {
    // closest points of segments p1 + x1*z1, p2 + x2*z2 with |xi| <= li
    double dif[3] = {p1[0] - p2[0], p1[1] - p2[1], p1[2] - p2[2]};
    double l1 = s1[1], l2 = s2[1];
    double b = dot3(z1, z2), u = -dot3(z1, dif), v = dot3(z2, dif);
    double det = 1.0 - b * b, x1, x2;
    if (fabs(det) >= 1e-12) {
      x1 = (u + b * v) / det;
      x2 = (v + b * u) / det;
      if (x1 > l1) { x1 = l1; x2 = v + b * x1; }
      else if (x1 < -l1) { x1 = -l1; x2 = v + b * x1; }
      if (x2 > l2) { x2 = l2; x1 = clampd(u + b * x2, -l1, l1); }
      else if (x2 < -l2) { x2 = -l2; x1 = clampd(u + b * x2, -l1, l1); }
    } else {
      x1 = clampd(u, -l1, l1);
      x2 = clampd(v + b * x1, -l2, l2);
      x1 = clampd(u + b * x2, -l1, l1);
    }
    double c1[3], c2[3];
    for (int k = 0; k < 3; k++) { c1[k] = p1[k] + z1[k] * x1; c2[k] = p2[k] + z2[k] * x2; }
    return col_sphere_sphere(c1, s1[0], c2, s2[0], margin, out[0]);
  }
  return 0;
}

// geom (type gt, centre gp, frame gR row-major, sizes gs; all in the field's frame) against
// one triangle; returns the number of contacts written (at most 4), normals pointing from
// the terrain to the geom.  sphere: closest point of the triangle (face always, edge /
// vertex when owned); capsule: both end spheres + the segment's interior against the owned
// edges; ellipsoid, cylinder: support point(s) along -n against the face + the owned vertex
// against the solid.
static __device__ __noinline__ int hf_tri_geom(const HField& h, int tid, const double* a,
                                               const double* b, const double* c, int gt,
                                               const double* gp, const double* gR,
                                               const double* gs, double margin, RawCon* out) {
  const int half = tid & 1;
  double ab[3], ac[3], nt[3];
  for (int k = 0; k < 3; k++) { ab[k] = b[k] - a[k]; ac[k] = c[k] - a[k]; }
  cross3(nt, ab, ac);
  normalize3(nt);
  int n = 0;
  if (gt == GEOM_SPHERE)
    return hf_sphere_tri(h, tid, a, b, c, nt, gp, gs[0], margin, nullptr, out[0]);
  if (gt == GEOM_CAPSULE) {
    double ax[3] = {gR[2], gR[5], gR[8]}, e[3];
#pragma unroll 1
    for (int sgn = 1; sgn >= -1; sgn -= 2) {
      for (int k = 0; k < 3; k++) e[k] = gp[k] + sgn * ax[k] * gs[1];
      n += hf_sphere_tri(h, tid, a, b, c, nt, e, gs[0], margin, ax, out[n]);
    }
    if (half == 0) n += hf_capsule_edge(h, gp, ax, gs[1], gs[0], a, b, 1, margin, out[n]);
    n += hf_capsule_edge(h, gp, ax, gs[1], gs[0], c, a, 0, margin, out[n]);
    return n;
  }
  if (gt == GEOM_ELLIPSOID) {
    double dl[3], x[3], xw[3], s = 0;
    matT_vec3(dl, gR, nt);
    for (int k = 0; k < 3; k++) s += gs[k] * gs[k] * dl[k] * dl[k];
    s = sqrt(s);
    for (int k = 0; k < 3; k++) x[k] = -gs[k] * gs[k] * dl[k] / s;
    mat_vec3(xw, gR, x);
    for (int k = 0; k < 3; k++) xw[k] += gp[k];
    n += hf_point_face(a, b, c, nt, xw, margin, out[n]);
    if (half == 0) {
      double tmul = 0;
      n += col_sphere_ellipsoid(a, 0.0, gp, gR, gs, margin, tmul, out[n]);
    }
    return n;
  }
  if (gt == GEOM_CYLINDER) {
    double axis[3] = {gR[2], gR[5], gR[8]}, vec[3], x[3];
    double prj = dot3(nt, axis);
    for (int k = 0; k < 3; k++) vec[k] = axis[k] * prj - nt[k];
    double len = norm3(vec);
    if (len >= BP_MINVAL) { for (int k = 0; k < 3; k++) vec[k] *= gs[0] / len; }
    else { vec[0] = gR[0] * gs[0]; vec[1] = gR[3] * gs[0]; vec[2] = gR[6] * gs[0]; }
#pragma unroll 1
    for (int sgn = 1; sgn >= -1; sgn -= 2) {
      for (int k = 0; k < 3; k++) x[k] = gp[k] + vec[k] + sgn * axis[k] * gs[1];
      n += hf_point_face(a, b, c, nt, x, margin, out[n]);
    }
    if (half == 0) n += col_sphere_cylinder(a, 0.0, gp, axis, gs[0], gs[1], margin, out[n]);
    return n;
  }
  return 0;
}

// cells of the grid under [lo, hi] along one axis: first cell and one past the last
__device__ __forceinline__ void hf_range(double lo, double hi, double rad, double dd, int nn,
                                         int& i0, int& i1) {
  double f0 = floor((lo + rad) / dd), f1 = ceil((hi + rad) / dd);
  if (f0 < 0) f0 = 0;
  if (f1 > nn - 1) f1 = nn - 1;
  i0 = (int)f0; i1 = (int)f1;
}

// frame rows: normal, tangent 1 (from the hint or a default axis), tangent 2
__device__ __forceinline__ void make_frame(double* f) {
  normalize3(f);
  if (norm3(f + 3) < 0.5) {
    f[3] = f[4] = f[5] = 0;
    if (f[1] < 0.5 && f[1] > -0.5) f[4] = 1; else f[5] = 1;
  }
  double t = dot3(f, f + 3);
  for (int k = 0; k < 3; k++) f[3 + k] -= t * f[k];
  normalize3(f + 3);
  cross3(f + 6, f, f + 3);
}

BP_STAGE void bp_write_contact(const C& c, int id, int g1, int g2, double margin, double gap,
                               const RawCon& raw) {
  BP_MODEL(m);
  double f[9];
  for (int k = 0; k < 3; k++) { f[k] = raw.n[k]; f[3 + k] = raw.hint[k]; f[6 + k] = 0; }
  make_frame(f);
  DF(contact_dist)[id] = raw.dist;
  for (int k = 0; k < 3; k++) DF(contact_pos)[3 * id + k] = raw.pos[k];
  for (int k = 0; k < 9; k++) DF(contact_frame)[9 * id + k] = f[k];
  double inc = margin - gap;
  DF(contact_includemargin)[id] = inc;
  IF(contact_geom1)[id] = g1; IF(contact_geom2)[id] = g2;
  // equal priority: max condim, max friction, solmix-weighted solref / solimp; otherwise
  // everything from the geom with the higher priority
  int d1 = m.geom_condim[g1], d2 = m.geom_condim[g2];
  const int pr1 = C::FULL ? m.geom_priority[g1] : 0, pr2 = C::FULL ? m.geom_priority[g2] : 0;
  IF(contact_dim)[id] = pr1 > pr2 ? d1 : pr2 > pr1 ? d2 : (d1 > d2 ? d1 : d2);
  const double *f1 = m.geom_friction + 3 * g1, *f2 = m.geom_friction + 3 * g2;
  double fr[3];
  for (int k = 0; k < 3; k++) fr[k] = pr1 > pr2 ? f1[k] : pr2 > pr1 ? f2[k] : fmax(f1[k], f2[k]);
  auto cf = DF(contact_friction) + 5 * id;
  cf[0] = fr[0]; cf[1] = fr[0]; cf[2] = fr[1]; cf[3] = fr[2]; cf[4] = fr[2];
  double mix1 = m.geom_solmix[g1], mix2 = m.geom_solmix[g2], mix;
  if (mix1 >= BP_MINVAL && mix2 >= BP_MINVAL) mix = mix1 / (mix1 + mix2);
  else if (mix1 < BP_MINVAL && mix2 < BP_MINVAL) mix = 0.5;
  else mix = mix1 < BP_MINVAL ? 0.0 : 1.0;
  if (pr1 != pr2) mix = pr1 > pr2 ? 1.0 : 0.0;
  for (int k = 0; k < 2; k++)
    DF(contact_solref)[2 * id + k] =
        mix * m.geom_solref[2 * g1 + k] + (1 - mix) * m.geom_solref[2 * g2 + k];
  for (int k = 0; k < 5; k++)
    DF(contact_solimp)[5 * id + k] =
        mix * m.geom_solimp[5 * g1 + k] + (1 - mix) * m.geom_solimp[5 * g2 + k];
  IF(contact_exclude)[id] = raw.dist >= inc;
  IF(contact_efc_address)[id] = -1;
  DF(contact_mu)[id] = 0;
}

BP_STAGE void bp_collision(const C& c) {
  BP_MODEL(m);
  if (c.lane == 0) IF(ncon)[0] = 0;
  SYNC();
  if (m.opt_disableflags & (DSBL_CONTACT | DSBL_CONSTRAINT)) return;
  bool filterparent = !(m.opt_disableflags & DSBL_FILTERPARENT);
  int nconmax = bp_lv(c).nconmax, base = 0;
  bool full = false;
  auto gpos = DF(geom_xpos), gmat = DF(geom_xmat);
#pragma unroll 1
  const int nprim = C::LEAN ? m.npair : m.npair - m.nhfpair;
  for (int p0 = 0; p0 < nprim; p0 += C::G) {
    int p = p0 + c.lane, n = 0, g1 = 0, g2 = 0;
    double margin = 0, gap = 0;
    RawCon raw[C::FULL ? 8 : 4];           // box - box: up to 8
    if (p < nprim && !(filterparent && m.pair_parentchild[p])) {
      g1 = m.pair_geom1[p]; g2 = m.pair_geom2[p];
      margin = fmax(m.geom_margin[g1], m.geom_margin[g2]);
      gap = fmax(m.geom_gap[g1], m.geom_gap[g2]);
      double p1[3], p2[3], z1[3], dif[3];
      for (int k = 0; k < 3; k++) {
        p1[k] = gpos[3 * g1 + k]; p2[k] = gpos[3 * g2 + k];
        z1[k] = gmat[9 * g1 + 3 * k + 2];
        dif[k] = p2[k] - p1[k];
      }
      int t1 = m.geom_type[g1], t2 = m.geom_type[g2];
      bool keep;
      if (t1 == GEOM_PLANE) keep = !(dot3(dif, z1) > m.geom_rbound[g2] + margin);
      else {
        double bound = m.geom_rbound[g1] + m.geom_rbound[g2] + margin;
        keep = !(dot3(dif, dif) > bound * bound);
      }
      if (keep) {
        double R2[9], R1[C::FULL ? 9 : 1], z2[3], s1[3], s2[3];
        if constexpr (!C::LEAN)
          for (int k = 0; k < 9; k++) R2[k] = gmat[9 * g2 + k];
        if constexpr (C::FULL)
          if (t1 == GEOM_BOX) for (int k = 0; k < 9; k++) R1[k] = gmat[9 * g1 + k];
        for (int k = 0; k < 3; k++) {
          z2[k] = gmat[9 * g2 + 3 * k + 2];
          s1[k] = m.geom_size[3 * g1 + k]; s2[k] = m.geom_size[3 * g2 + k];
        }
        n = bp_collide_prims<(C::FULL ? 0 : C::LEAN ? 1 : 2)>(t1, t2, p1, z1, s1, p2, z2, R2, s2, margin, raw, R1);
      }
    }
    // exclusive prefix sum of n over the group -> slots in pair order
    int tot, start = base + group_excl_scan(c, n, &tot);
    base += tot;
#pragma unroll 1
    for (int i = 0; i < n; i++) {
      int id = start + i;
      if (id >= nconmax) { full = true; break; }
      bp_write_contact(c, id, g1, g2, margin, gap, raw[i]);
    }
    if (base > nconmax) break;    // uniform inside the group
  }
  // ---- height-field pairs (the last nhfpair of the list): one pair at a time, the
  // triangles under the geom's bounding sphere spread over the lanes, contacts in
  // (row, column, half) order
  if constexpr (!C::LEAN) if (m.nhfpair > 0 && c.hf != nullptr) {
#pragma unroll 1
    for (int p = nprim; p < m.npair && base <= nconmax; p++) {
      if (filterparent && m.pair_parentchild[p]) continue;
      const int g1 = m.pair_geom1[p], g2 = m.pair_geom2[p];
      const double margin = fmax(m.geom_margin[g1], m.geom_margin[g2]);
      const double gap = fmax(m.geom_gap[g1], m.geom_gap[g2]);
      HField h;
      hf_init(m, c.hf, m.geom_dataid[g1], h);
      double p1[3], R1[9], R2[9], dif[3], gp[3], gR[9], gs[3];
      for (int k = 0; k < 3; k++) { p1[k] = gpos[3 * g1 + k]; dif[k] = gpos[3 * g2 + k] - p1[k]; }
      for (int k = 0; k < 9; k++) { R1[k] = gmat[9 * g1 + k]; R2[k] = gmat[9 * g2 + k]; }
      for (int k = 0; k < 3; k++) gs[k] = m.geom_size[3 * g2 + k];
      matT_vec3(gp, R1, dif);
      for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++)
          gR[3 * i + j] = R1[i] * R2[j] + R1[3 + i] * R2[3 + j] + R1[6 + i] * R2[6 + j];
      const double rb = m.geom_rbound[g2] + margin;
      if (gp[0] - rb > h.rx || gp[0] + rb < -h.rx || gp[1] - rb > h.ry || gp[1] + rb < -h.ry ||
          gp[2] - rb > h.zs || gp[2] + rb < -h.base) continue;
      int c0, c1, r0, r1;
      hf_range(gp[0] - rb, gp[0] + rb, h.rx, h.dx, h.ncol, c0, c1);
      hf_range(gp[1] - rb, gp[1] + rb, h.ry, h.dy, h.nrow, r0, r1);
      const int ncol = c1 - c0, ntri = ncol > 0 && r1 > r0 ? 2 * ncol * (r1 - r0) : 0;
      const int gt = m.geom_type[g2];
#pragma unroll 1
      for (int t0 = 0; t0 < ntri; t0 += C::G) {
        const int t = t0 + c.lane;
        int n = 0;
        RawCon raw[4];
        if (t < ntri) {
          const int cell = t >> 1, half = t & 1;
          double a[3], b[3], cc[3];
          const int rr = r0 + cell / ncol, cl = c0 + cell % ncol;
          hf_triangle(h, rr, cl, half, a, b, cc);
          if (!(fmax(a[2], fmax(b[2], cc[2])) < gp[2] - rb))
            n = hf_tri_geom(h, ((rr * h.ncol + cl) << 1) | half, a, b, cc, gt, gp, gR, gs, margin,
                            raw);
        }
        int tot, start = base + group_excl_scan(c, n, &tot);
        base += tot;
#pragma unroll 1
        for (int i = 0; i < n; i++) {
          int id = start + i;
          if (id >= nconmax) { full = true; break; }
          RawCon w;
          w.dist = raw[i].dist;
          mat_vec3(w.pos, R1, raw[i].pos);
          for (int k = 0; k < 3; k++) w.pos[k] += p1[k];
          mat_vec3(w.n, R1, raw[i].n);
          mat_vec3(w.hint, R1, raw[i].hint);
          bp_write_contact(c, id, g1, g2, margin, gap, w);
        }
        if (base > nconmax) break;
      }
    }
  }
  full = group_any(c, full);
  if (c.lane == 0) {
    IF(ncon)[0] = base < nconmax ? base : nconmax;
    if (full) IF(warning)[W_CONTACTFULL]++;
  }
  SYNC();
}

#endif  // BP_COLLIDE_CUH_
