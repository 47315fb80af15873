// Height fields: narrowphase against the triangles of the grid and ray casting.
//
// MuJoCo walks the grid cells under the geom's bounding sphere, builds a triangular prism
// per half cell and sends prism-geom to its general convex routine (libccd MPR inside the
// closed binary).  Here the same cell walk feeds closed-form tests against the prism's top
// triangle, the ground below counting as solid (oracle
// mjstep_oracle.c "height fields").  Everything is in the height field's frame.
//   cell (r, c), half 0: a = v(r,c), b = v(r,c+1),   c = v(r+1,c+1)
//                half 1: a = v(r,c), b = v(r+1,c+1), c = v(r+1,c)      (both CCW from above)
// Features shared by neighbouring triangles are tested by ONE owner, so a contact is not
// reported twice: half 0 owns vertex a and edges ab, ca; half 1 owns edge ca.
#ifndef BP_HFIELD_CUH_
#define BP_HFIELD_CUH_
#include "bp_math.cuh"

struct RawCon { double dist, pos[3], n[3], hint[3]; };

struct HField {
  const float* data;
  int nrow, ncol;
  double rx, ry, zs, base, dx, dy;
};

template <class M_>
__device__ __forceinline__ void hf_init(const M_& m, const float* env_data, int hid, HField& h) {
  h.data = env_data + m.hfield_adr[hid];
  h.nrow = m.hfield_nrow[hid]; h.ncol = m.hfield_ncol[hid];
  h.rx = m.hfield_size[4 * hid]; h.ry = m.hfield_size[4 * hid + 1];
  h.zs = m.hfield_size[4 * hid + 2]; h.base = m.hfield_size[4 * hid + 3];
  h.dx = 2 * h.rx / (h.ncol - 1); h.dy = 2 * h.ry / (h.nrow - 1);
}

__device__ __forceinline__ void hf_vertex(const HField& h, int r, int c, double* v) {
  v[0] = c * h.dx - h.rx; v[1] = r * h.dy - h.ry;
  v[2] = (double)__ldg(h.data + r * h.ncol + c) * h.zs;
}

__device__ __forceinline__ void hf_triangle(const HField& h, int r, int c, int half, double* a,
                                            double* b, double* cc) {
  hf_vertex(h, r, c, a);
  if (half == 0) { hf_vertex(h, r, c + 1, b); hf_vertex(h, r + 1, c + 1, cc); }
  else { hf_vertex(h, r + 1, c + 1, b); hf_vertex(h, r + 1, c, cc); }
}

__device__ __forceinline__ int hf_cell(double u, double rad, double dd, int nn) {
  double f = floor((u + rad) / dd);
  if (f < 0) f = 0;
  if (f > nn - 2) f = nn - 2;
  return (int)f;
}

// elevation of the surface over (x, y); tid = id of the triangle that holds the point,
// (cell index << 1) | half
static __device__ __noinline__ double hf_surface(const HField& h, double x, double y, int& tid) {
  int c = hf_cell(x, h.rx, h.dx, h.ncol), r = hf_cell(y, h.ry, h.dy, h.nrow);
  double fx = (x + h.rx) / h.dx - c, fy = (y + h.ry) / h.dy - r;
  fx = clampd(fx, 0.0, 1.0); fy = clampd(fy, 0.0, 1.0);
  const float* dt = h.data;
  double z00 = __ldg(dt + r * h.ncol + c), z10 = __ldg(dt + r * h.ncol + c + 1),
         z01 = __ldg(dt + (r + 1) * h.ncol + c), z11 = __ldg(dt + (r + 1) * h.ncol + c + 1);
  int half = fx >= fy ? 0 : 1;
  double zsurf = half == 0 ? z00 + fx * (z10 - z00) + fy * (z11 - z10)
                           : z00 + fx * (z11 - z01) + fy * (z01 - z00);
  tid = ((r * h.ncol + c) << 1) | half;
  return zsurf * h.zs;
}

// closest point of triangle abc to p; returns the Voronoi region: 0 face, 1 / 2 / 3 vertex
// a / b / c, 4 edge ab, 5 edge bc, 6 edge ca
static __device__ __noinline__ int tri_closest(const double* p, const double* a, const double* b,
                                               const double* c, double* q) {
  double ab[3], ac[3], ap[3], bp[3], cp[3];
  for (int k = 0; k < 3; k++) {
    ab[k] = b[k] - a[k]; ac[k] = c[k] - a[k];
    ap[k] = p[k] - a[k]; bp[k] = p[k] - b[k]; cp[k] = p[k] - c[k];
  }
  double d1 = dot3(ab, ap), d2 = dot3(ac, ap);
  if (d1 <= 0 && d2 <= 0) { for (int k = 0; k < 3; k++) q[k] = a[k]; return 1; }
  double d3 = dot3(ab, bp), d4 = dot3(ac, bp);
  if (d3 >= 0 && d4 <= d3) { for (int k = 0; k < 3; k++) q[k] = b[k]; return 2; }
  double vc = d1 * d4 - d3 * d2;
  if (vc <= 0 && d1 >= 0 && d3 <= 0) {
    double v = d1 / (d1 - d3);
    for (int k = 0; k < 3; k++) q[k] = a[k] + v * ab[k];
    return 4;
  }
  double d5 = dot3(ab, cp), d6 = dot3(ac, cp);
  if (d6 >= 0 && d5 <= d6) { for (int k = 0; k < 3; k++) q[k] = c[k]; return 3; }
  double vb = d5 * d2 - d1 * d6;
  if (vb <= 0 && d2 >= 0 && d6 <= 0) {
    double w = d2 / (d2 - d6);
    for (int k = 0; k < 3; k++) q[k] = a[k] + w * ac[k];
    return 6;
  }
  double va = d3 * d6 - d5 * d4;
  if (va <= 0 && (d4 - d3) >= 0 && (d5 - d6) >= 0) {
    double w = (d4 - d3) / ((d4 - d3) + (d5 - d6));
    for (int k = 0; k < 3; k++) q[k] = b[k] + w * (c[k] - b[k]);
    return 5;
  }
  double den = 1.0 / (va + vb + vc), v = vb * den, w = vc * den;
  for (int k = 0; k < 3; k++) q[k] = a[k] + ab[k] * v + ac[k] * w;
  return 0;
}

#define HF_OWN0 ((1 << 0) | (1 << 1) | (1 << 4) | (1 << 6))
#define HF_OWN1 ((1 << 0) | (1 << 6))

__device__ __forceinline__ void raw_set(RawCon& o, double dist, const double* pos, const double* n,
                                        const double* hint) {
  o.dist = dist;
  for (int k = 0; k < 3; k++) { o.pos[k] = pos[k]; o.n[k] = n[k]; o.hint[k] = hint ? hint[k] : 0.0; }
}

// sphere (centre q, radius r) against triangle `tid` with unit normal nt.  A centre above
// the surface touches the closest point of the triangle (face always, edges / vertices when
// owned); a centre that has sunk below the surface is pushed out through the face of the
// one triangle that holds it, so no sideways normals from neighbouring edges
static __device__ __noinline__ int hf_sphere_tri(const HField& h, int tid, const double* a,
                                                 const double* b, const double* c, const double* nt,
                                                 const double* q, double r, double margin,
                                                 const double* hint, RawCon& out) {
  double aq[3] = {q[0] - a[0], q[1] - a[1], q[2] - a[2]}, cl[3], nrm[3], dist;
  double s = dot3(nt, aq);
  int holder;
  if (q[2] < hf_surface(h, q[0], q[1], holder)) {
    if (holder != tid) return 0;
    for (int k = 0; k < 3; k++) nrm[k] = nt[k];
    dist = s - r;
  } else {
    if (s < 0) return 0;
    int reg = tri_closest(q, a, b, c, cl);
    if (!(((tid & 1) ? HF_OWN1 : HF_OWN0) >> reg & 1)) return 0;
    double dv[3] = {q[0] - cl[0], q[1] - cl[1], q[2] - cl[2]};
    double dd = norm3(dv);
    if (dd < BP_MINVAL) { for (int k = 0; k < 3; k++) nrm[k] = nt[k]; }
    else { for (int k = 0; k < 3; k++) nrm[k] = dv[k] / dd; }
    dist = dd - r;
  }
  if (dist > margin) return 0;
  double pos[3];
  for (int k = 0; k < 3; k++) pos[k] = q[k] - nrm[k] * (r + 0.5 * dist);
  raw_set(out, dist, pos, nrm, hint);
  return 1;
}

// interior of the capsule's segment (p +- l ax) against the edge e0 -> e1; a hit at the
// start vertex counts when vertex0 is set (the edge that owns its first vertex)
static __device__ __noinline__ int hf_capsule_edge(const HField& h, const double* p,
                                                   const double* ax, double l, double r,
                                                   const double* e0, const double* e1,
                                                   int vertex0, double margin, RawCon& out) {
  double d1[3], d2[3], rr[3];
  for (int k = 0; k < 3; k++) {
    d1[k] = 2 * l * ax[k]; d2[k] = e1[k] - e0[k];
    rr[k] = p[k] - l * ax[k] - e0[k];
  }
  double A = dot3(d1, d1), E = dot3(d2, d2), F = dot3(d2, rr), C = dot3(d1, rr),
         B = dot3(d1, d2);
  double den = A * E - B * B, s, t;
  if (den > 1e-12 * A * E) s = clampd((B * F - C * E) / den, 0.0, 1.0);
  else s = 0.5;
  t = (B * s + F) / E;
  if (t < 0) { t = 0; s = clampd(-C / A, 0.0, 1.0); }
  else if (t > 1) { t = 1; s = clampd((B - C) / A, 0.0, 1.0); }
  if (!(s > 0 && s < 1)) return 0;                 // the ends are tested as spheres
  if (!(t < 1 && (t > 0 || vertex0))) return 0;
  double c1[3], c2[3], dv[3];
  for (int k = 0; k < 3; k++) {
    c1[k] = p[k] - l * ax[k] + s * d1[k]; c2[k] = e0[k] + t * d2[k];
    dv[k] = c1[k] - c2[k];
  }
  double dd = norm3(dv);
  if (dd < BP_MINVAL || dv[2] <= 0) return 0;      // the axis has sunk below the edge
  int holder;
  if (c1[2] < hf_surface(h, c1[0], c1[1], holder)) return 0;     // ... or the surface
  double dist = dd - r;
  if (dist > margin) return 0;
  double nrm[3], pos[3];
  for (int k = 0; k < 3; k++) { nrm[k] = dv[k] / dd; pos[k] = c1[k] - nrm[k] * (r + 0.5 * dist); }
  raw_set(out, dist, pos, nrm, ax);
  return 1;
}

// a surface point x of a convex geom against the face: inside the prism's footprint and
// below (or within margin of) the face
static __device__ __noinline__ int hf_point_face(const double* a, const double* b, const double* c,
                                                 const double* nt, const double* x, double margin,
                                                 RawCon& out) {
  double cl[3];
  if (tri_closest(x, a, b, c, cl) != 0) return 0;
  double ax[3] = {x[0] - a[0], x[1] - a[1], x[2] - a[2]};
  double dist = dot3(nt, ax);
  if (dist > margin) return 0;
  double pos[3];
  for (int k = 0; k < 3; k++) pos[k] = x[k] - nt[k] * 0.5 * dist;
  raw_set(out, dist, pos, nt, nullptr);
  return 1;
}

// ---- rays [mj_ray, mju_rayGeom]: distance along a unit ray to the nearest surface, -1
// without a hit.  Quadratics return the smallest non-negative root, so a ray that starts
// inside a geom reports its exit.
__device__ __forceinline__ double ray_quad(double a, double b, double c) {  // a x^2 + 2 b x + c
  double det = b * b - a * c;
  if (det < 0 || a < BP_MINVAL) return -1;
  det = sqrt(det);
  double x0 = (-b - det) / a, x1 = (-b + det) / a;
  if (x0 >= 0) return x0;
  if (x1 >= 0) return x1;
  return -1;
}
__device__ __forceinline__ double ray_min(double x, double y) {
  if (x < 0) return y;
  if (y < 0) return x;
  return x < y ? x : y;
}
// Moeller-Trumbore, both faces
__device__ __forceinline__ double ray_triangle(const double* o, const double* v, const double* a,
                                               const double* b, const double* c) {
  double e1[3], e2[3], pv[3], tv[3], qv[3];
  for (int k = 0; k < 3; k++) { e1[k] = b[k] - a[k]; e2[k] = c[k] - a[k]; tv[k] = o[k] - a[k]; }
  cross3(pv, v, e2);
  double det = dot3(e1, pv);
  if (fabs(det) < BP_MINVAL) return -1;
  double u = dot3(tv, pv) / det;
  if (u < 0 || u > 1) return -1;
  cross3(qv, tv, e1);
  double w = dot3(v, qv) / det;
  if (w < 0 || u + w > 1) return -1;
  double t = dot3(e2, qv) / det;
  return t >= 0 ? t : -1;
}

// ray (origin o, unit direction v, field frame): the box is entered through a wall or the
// bottom (solid), or the surface is hit further on; the grid is swept slab by slab along
// the axis the ray advances fastest on
static __device__ __noinline__ double ray_hfield(const HField& h, const double* o, const double* v) {
  double lo[3] = {-h.rx, -h.ry, -h.base}, hi[3] = {h.rx, h.ry, h.zs};
  double t0 = 0, t1 = 1e300;
  int axis0 = -1;
#pragma unroll
  for (int k = 0; k < 3; k++) {
    if (fabs(v[k]) < BP_MINVAL) {
      if (o[k] < lo[k] || o[k] > hi[k]) return -1;
      continue;
    }
    double ta = (lo[k] - o[k]) / v[k], tb = (hi[k] - o[k]) / v[k];
    if (ta > tb) { double s = ta; ta = tb; tb = s; }
    if (ta > t0) { t0 = ta; axis0 = k; }
    if (tb < t1) t1 = tb;
  }
  if (t0 > t1) return -1;
  double best = -1;
  if (axis0 >= 0 && !(axis0 == 2 && v[2] < 0)) {
    if (axis0 == 2) return t0;
    double x = o[0] + t0 * v[0], y = o[1] + t0 * v[1], z = o[2] + t0 * v[2];
    int holder;
    if (z <= hf_surface(h, x, y, holder)) return t0;
  }
  const int am = fabs(v[0]) >= fabs(v[1]) ? 0 : 1, an = 1 - am;
  const double radm = am == 0 ? h.rx : h.ry, radn = am == 0 ? h.ry : h.rx;
  const double ddm = am == 0 ? h.dx : h.dy, ddn = am == 0 ? h.dy : h.dx;
  const int nm = am == 0 ? h.ncol : h.nrow, nn = am == 0 ? h.nrow : h.ncol;
  const double om = am == 0 ? o[0] : o[1], vm = am == 0 ? v[0] : v[1];
  const double on = am == 0 ? o[1] : o[0], vn = am == 0 ? v[1] : v[0];
  int i0 = hf_cell(om + t0 * vm, radm, ddm, nm);
  int i1 = hf_cell(om + t1 * vm, radm, ddm, nm);
  int step = i1 >= i0 ? 1 : -1;
#pragma unroll 1
  for (int i = i0; i != i1 + step; i += step) {
    double ta = t0, tb = t1;
    if (fabs(vm) >= BP_MINVAL) {
      double sa = (i * ddm - radm - om) / vm, sb = ((i + 1) * ddm - radm - om) / vm;
      if (sa > sb) { double s = sa; sa = sb; sb = s; }
      if (sa > ta) ta = sa;
      if (sb < tb) tb = sb;
    }
    if (ta > tb) continue;
    double na = on + ta * vn, nb = on + tb * vn;
    int j0 = hf_cell(na < nb ? na : nb, radn, ddn, nn), j1 = hf_cell(na < nb ? nb : na, radn, ddn, nn);
#pragma unroll 1
    for (int j = j0; j <= j1; j++) {
      int r = am == 0 ? j : i, c = am == 0 ? i : j;
#pragma unroll 1
      for (int half = 0; half < 2; half++) {
        double a[3], b[3], cc[3];
        hf_triangle(h, r, c, half, a, b, cc);
        best = ray_min(best, ray_triangle(o, v, a, b, cc));
      }
    }
    if (best >= 0) return best;
  }
  return -1;
}

// ray in the geom's frame (origin o, direction v) against a primitive of type gt
static __device__ __noinline__ double ray_prim(int gt, const double* sz, const double* o,
                                               const double* v) {
  switch (gt) {
    case GEOM_PLANE: {
      if (v[2] > -BP_MINVAL) return -1;         // only the front face, from above
      double x = -o[2] / v[2];
      if (x < 0) return -1;
      double p0 = o[0] + x * v[0], p1 = o[1] + x * v[1];
      if ((sz[0] <= 0 || fabs(p0) <= sz[0]) && (sz[1] <= 0 || fabs(p1) <= sz[1])) return x;
      return -1;
    }
    case GEOM_SPHERE:
      return ray_quad(dot3(v, v), dot3(v, o), dot3(o, o) - sz[0] * sz[0]);
    case GEOM_CAPSULE: {
      double a = v[0] * v[0] + v[1] * v[1], b = v[0] * o[0] + v[1] * o[1],
             c = o[0] * o[0] + o[1] * o[1] - sz[0] * sz[0];
      double x = ray_quad(a, b, c);
      if (x >= 0 && fabs(o[2] + x * v[2]) > sz[1]) x = -1;
#pragma unroll 1
      for (int sgn = -1; sgn <= 1; sgn += 2) {
        double oz = o[2] - sgn * sz[1];
        double dd = dot3(v, v), db = v[0] * o[0] + v[1] * o[1] + v[2] * oz,
               dc = o[0] * o[0] + o[1] * o[1] + oz * oz - sz[0] * sz[0];
        double det = db * db - dd * dc;
        if (det < 0) continue;
        det = sqrt(det);
        for (int root = 0; root < 2; root++) {
          double xs = (root == 0 ? -db - det : -db + det) / dd;
          if (xs >= 0 && sgn * (o[2] + xs * v[2]) >= sz[1]) { x = ray_min(x, xs); break; }
        }
      }
      return x;
    }
    case GEOM_ELLIPSOID: {
      double os[3] = {o[0] / sz[0], o[1] / sz[1], o[2] / sz[2]};
      double vs[3] = {v[0] / sz[0], v[1] / sz[1], v[2] / sz[2]};
      return ray_quad(dot3(vs, vs), dot3(vs, os), dot3(os, os) - 1);
    }
    case GEOM_CYLINDER: {
      double a = v[0] * v[0] + v[1] * v[1], b = v[0] * o[0] + v[1] * o[1],
             c = o[0] * o[0] + o[1] * o[1] - sz[0] * sz[0];
      double x = ray_quad(a, b, c);
      if (x >= 0 && fabs(o[2] + x * v[2]) > sz[1]) x = -1;
      if (fabs(v[2]) >= BP_MINVAL)
        for (int sgn = -1; sgn <= 1; sgn += 2) {
          double xs = (sgn * sz[1] - o[2]) / v[2];
          if (xs < 0) continue;
          double p0 = o[0] + xs * v[0], p1 = o[1] + xs * v[1];
          if (p0 * p0 + p1 * p1 <= sz[0] * sz[0]) x = ray_min(x, xs);
        }
      return x;
    }
    case GEOM_BOX: {
      double x = -1;
#pragma unroll 1
      for (int k = 0; k < 3; k++) {
        if (fabs(v[k]) < BP_MINVAL) continue;
        for (int sgn = -1; sgn <= 1; sgn += 2) {
          double xs = (sgn * sz[k] - o[k]) / v[k];
          if (xs < 0) continue;
          int k1 = (k + 1) % 3, k2 = (k + 2) % 3;
          if (fabs(o[k1] + xs * v[k1]) <= sz[k1] && fabs(o[k2] + xs * v[k2]) <= sz[k2])
            x = ray_min(x, xs);
        }
      }
      return x;
    }
  }
  return -1;
}

#endif  // BP_HFIELD_CUH_
