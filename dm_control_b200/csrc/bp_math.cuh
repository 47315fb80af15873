// Small fixed-size math used by every stage, plus the execution context.
//
// An environment is advanced by a GROUP of G lanes (G = 1, 2, 4, 8, 16 or 32; 32/G
// environments share a warp).  Its data image is addressed through strided pointers:
// element i of a field lives at base[i * S].  S = 1 is a contiguous image (shared
// memory for G = 32, global memory for 1 < G < 32); S = 32 interleaves the images of
// the 32 environments of a warp so that the G = 1 (thread per environment) variant
// makes fully coalesced accesses.  All helpers are templates over "pointer-like"
// arguments so that they accept strided pointers, model pointers and local arrays.
#ifndef BP_MATH_CUH_
#define BP_MATH_CUH_
#include "bp_types.h"
#include <type_traits>

// GL: the pointer is known to address global memory (the images of a static instance: no
// shared-memory arena, no shared-memory image), which lets the compiler emit LDG / STG instead
// of generic LD / ST.
template <class T, int S, bool GL = false>
struct SP {
  T* p;
  __device__ __forceinline__ T& operator[](int i) const {
    T* q = p + (long long)i * S;
    if constexpr (GL) __builtin_assume(__isGlobal(q));
    return *q;
  }
  __device__ __forceinline__ SP operator+(int n) const { return SP{p + (long long)n * S}; }
  __device__ __forceinline__ bool valid() const { return p != nullptr; }
};

// LEAN: a second instantiation of every stage for models that use none of: height fields,
// rangefinders, tendons, equalities, elliptic cones, geoms other than plane / sphere /
// capsule, fluid forces, actuator dynamics, geom priorities, per-env model frames.  Those
// branches are compiled out of the hot functions (each of them cost the plain locomotion
// models ~2 % when it was added, profiles/r1_tuning.md r1u); bp_api.cu picks the shape.
// SHAPE 2 (MID) keeps height fields, rangefinders, fixed tendons, equalities, ellipsoid /
// cylinder geoms and actuator dynamics (the quadruped) and drops what only FULL has:
// elliptic cones, fluid forces, spatial tendons, tendon limits and springs, box geoms, geom
// priorities, per-env model frames.
// lane inside the group; for G = 1 it is the compile-time constant 0, also inside the
// out-of-line stage functions (which see the context by reference)
template <int G>
struct LaneT {
  int v;
  __device__ __forceinline__ LaneT(int x) : v(x) {}
  __device__ __forceinline__ operator int() const { return v; }
};
template <>
struct LaneT<1> {
  __device__ __forceinline__ LaneT(int) {}
  __device__ __forceinline__ operator int() const { return 0; }
};

// MT: static model traits (csrc/static/*.cuh, generated ahead of time from a compiled model
// by tools/gen_static_model.py) or void.  With traits, the model's structure -- sizes, tree,
// addresses, type tables, pair list, sensor list, the data layout -- is compile-time constant,
// the structural loops unroll (UF) and every image access has a constant offset; parameters
// that may be edited at run time (all float arrays, opt_*) are still read from BpModel.
template <int G_, int S_, int SHAPE_ = 0, class MT_ = void>
struct Ctx {
  static constexpr int G = G_, S = S_;
  static constexpr bool LEAN = SHAPE_ == 1, FULL = SHAPE_ == 0;
  using MT = MT_;
  static constexpr bool STATIC = !std::is_void<MT_>::value;
#ifndef BP_UF
#define BP_UF 1
#endif
  static constexpr int UF = STATIC ? BP_UF : 1;  // unroll factor of structural loops
  using DP = SP<double, S_, STATIC>;
  using IP = SP<int, S_, STATIC>;
  const BpParams& P;
  double* sm;      // element 0 of this env's image
  int* ip;         // element 0 of the int32 region
  LaneT<G_> lane;  // lane inside the group, 0 .. G-1
  unsigned mask;   // warp lanes of this group
  bool act;        // false: this group has no environment (ragged tail / masked out)
  bool lock;       // true: the block's warps re-align at stage boundaries (BP_LOCK)
  double* hot;     // this env's shared-memory arena (nullptr: those arrays stay global)
  const float* hf; // this env's height-field elevation data (nullptr: none)
  const double* gpos;  // this env's geom_pos / site_pos (the model's unless overridden)
  const double* spos;
  const double* bpos;  // this env's body_pos / body_quat
  const double* bquat;
  const double* wprm;  // this env's tendon coefficients (wrap_prm)
  __device__ __forceinline__ DP null() const { return DP{nullptr}; }
};

// model / layout as the stage functions see them: BpModel / BpLayout of the launch, or the
// static traits' views of them
template <class C>
__device__ __forceinline__ decltype(auto) bp_mv(const C& c) {
  if constexpr (C::STATIC) return typename C::MT::Model(c.P.m);
  else return (c.P.m);
}
template <class C>
__device__ __forceinline__ decltype(auto) bp_lv(const C& c) {
  if constexpr (C::STATIC) return typename C::MT::Layout();
  else return (c.P.L);
}
#define BP_MODEL(m) const auto& m = bp_mv(c)
#define DF(f) (typename C::DP{c.sm + (long long)bp_lv(c).f * C::S})
#define IF(f) (typename C::IP{c.ip + (long long)bp_lv(c).f * C::S})
// solver arrays that live in the shared-memory arena when there is one (S == 1 only)
#define HOT(f, off) ((C::S == 1 && c.hot) ? typename C::DP{c.hot + (off)} : DF(f))
#define SYNC() do { if (C::G > 1) __syncwarp(c.mask); } while (0)
#if defined(BP_STAGE_FORCE_INLINE)
#define BP_STAGE template <class C> __device__ __forceinline__
#define BP_STAGE_NOINLINE
#elif defined(BP_STAGE_AUTO_INLINE)  // static instances: inlining of the stage functions left to the compiler
#define BP_STAGE template <class C> __device__
#define BP_STAGE_NOINLINE
#else
#define BP_STAGE template <class C> __device__ __noinline__
#define BP_STAGE_NOINLINE __noinline__
#endif
// Stage boundary of the lock-step shapes: every thread of the block (active or not)
// arrives, so the block's warps execute the same stage -- and share its instruction
// cache lines -- at the same time.  Only in code whose trip count is uniform per launch.
// Barrier ids: 0 after kinematics..factorM, 1 after collision, 2 after makeConstraint,
// 3 after the velocity stage, 4 before the solver, 5 after the solver, 6 end of a step,
// 7 after mj_step1's velocity stage, 8 every Newton iteration, 9 after kinematics,
// 10 after comPos..crb, 11 inside the velocity stage.  BP_LOCKMASK selects the ones that are compiled in.
#ifndef BP_LOCKMASK
#define BP_LOCKMASK 0x1ff
#endif
#ifdef BP_TIMELINE   // debug builds: wall time of block 0 between consecutive barriers
static __device__ unsigned long long bp_timeline[32];   // one copy per translation unit
static __device__ long long bp_timeline_prev;
#define BP_LOCK(id) do { if (((BP_LOCKMASK) >> (id) & 1) && c.lock) { __syncthreads(); \
    if (blockIdx.x == 0 && threadIdx.x == 0) { long long t = clock64(); \
      bp_timeline[id] += (unsigned long long)(t - bp_timeline_prev); bp_timeline[16 + (id)]++; \
      bp_timeline_prev = t; } } } while (0)
// finer stamps inside the solver: cycles of block 0's first group between consecutive BP_TL
static __device__ unsigned long long bp_tl2[64];
static __device__ long long bp_tl2_prev;
#define BP_TL(slot) do { if (blockIdx.x == 0 && threadIdx.x == 0) { long long t_ = clock64(); \
    if ((slot) >= 0) { bp_tl2[slot] += (unsigned long long)(t_ - bp_tl2_prev); bp_tl2[32 + (slot)]++; } \
    bp_tl2_prev = t_; } } while (0)
#else
#define BP_LOCK(id) do { if (((BP_LOCKMASK) >> (id) & 1) && c.lock) __syncthreads(); } while (0)
#define BP_TL(slot) do { } while (0)
#endif
#define RUN(call) do { if (c.act) { call; } } while (0)

template <class A, class B>
__device__ __forceinline__ double dot3(A a, B b) {
  return a[0] * b[0] + a[1] * b[1] + a[2] * b[2];
}
template <class R, class A, class B>
__device__ __forceinline__ void cross3(R r, A a, B b) {
  double x = a[1] * b[2] - a[2] * b[1], y = a[2] * b[0] - a[0] * b[2],
         z = a[0] * b[1] - a[1] * b[0];
  r[0] = x; r[1] = y; r[2] = z;
}
template <class A>
__device__ __forceinline__ double norm3(A a) { return sqrt(dot3(a, a)); }
template <class A>
__device__ __forceinline__ double normalize3(A a) {
  double n = norm3(a);
  if (n < BP_MINVAL) { a[0] = 1; a[1] = 0; a[2] = 0; }
  else { a[0] /= n; a[1] /= n; a[2] /= n; }
  return n;
}
template <class A>
__device__ __forceinline__ void normalize4(A q) {
  double q0 = q[0], q1 = q[1], q2 = q[2], q3 = q[3];
  double n = sqrt(q0 * q0 + q1 * q1 + q2 * q2 + q3 * q3);
  if (n < BP_MINVAL) { q[0] = 1; q[1] = 0; q[2] = 0; q[3] = 0; }
  else { q[0] = q0 / n; q[1] = q1 / n; q[2] = q2 / n; q[3] = q3 / n; }
}
template <class R, class A, class B>
__device__ __forceinline__ void quat_mul(R r, A a, B b) {
  double a0 = a[0], a1 = a[1], a2 = a[2], a3 = a[3];
  double b0 = b[0], b1 = b[1], b2 = b[2], b3 = b[3];
  r[0] = a0 * b0 - a1 * b1 - a2 * b2 - a3 * b3;
  r[1] = a0 * b1 + a1 * b0 + a2 * b3 - a3 * b2;
  r[2] = a0 * b2 - a1 * b3 + a2 * b0 + a3 * b1;
  r[3] = a0 * b3 + a1 * b2 - a2 * b1 + a3 * b0;
}
template <class R, class A>
__device__ __forceinline__ void quat2mat(R R_, A q) {
  double w = q[0], x = q[1], y = q[2], z = q[3];
  double q00 = w * w, q01 = w * x, q02 = w * y, q03 = w * z;
  double q11 = x * x, q12 = x * y, q13 = x * z;
  double q22 = y * y, q23 = y * z, q33 = z * z;
  R_[0] = q00 + q11 - q22 - q33; R_[4] = q00 - q11 + q22 - q33;
  R_[8] = q00 - q11 - q22 + q33;
  R_[1] = 2 * (q12 - q03); R_[2] = 2 * (q13 + q02);
  R_[3] = 2 * (q12 + q03); R_[5] = 2 * (q23 - q01);
  R_[6] = 2 * (q13 - q02); R_[7] = 2 * (q23 + q01);
}
template <class R, class M, class V>
__device__ __forceinline__ void mat_vec3(R r, M R_, V v) {
  double v0 = v[0], v1 = v[1], v2 = v[2];
  r[0] = R_[0] * v0 + R_[1] * v1 + R_[2] * v2;
  r[1] = R_[3] * v0 + R_[4] * v1 + R_[5] * v2;
  r[2] = R_[6] * v0 + R_[7] * v1 + R_[8] * v2;
}
template <class R, class M, class V>
__device__ __forceinline__ void matT_vec3(R r, M R_, V v) {
  double v0 = v[0], v1 = v[1], v2 = v[2];
  r[0] = R_[0] * v0 + R_[3] * v1 + R_[6] * v2;
  r[1] = R_[1] * v0 + R_[4] * v1 + R_[7] * v2;
  r[2] = R_[2] * v0 + R_[5] * v1 + R_[8] * v2;
}
template <class Q, class A>
__device__ __forceinline__ void axis_angle_quat(Q q, A axis, double ang) {
  double s, co;
  sincos(0.5 * ang, &s, &co);
  q[0] = co; q[1] = axis[0] * s; q[2] = axis[1] * s; q[3] = axis[2] * s;
}
// q <- q * exp(w*dt/2), w in the body frame (free/ball joint velocity)
template <class Q, class W>
__device__ __forceinline__ void quat_integrate(Q q, W w, double dt) {
  double v[3] = {w[0], w[1], w[2]};
  double n = norm3(v);
  if (n < BP_MINVAL) return;
  v[0] /= n; v[1] /= n; v[2] /= n;
  double qr[4], qn[4];
  axis_angle_quat(qr, v, n * dt);
  quat_mul(qn, q, qr);
  normalize4(qn);
  q[0] = qn[0]; q[1] = qn[1]; q[2] = qn[2]; q[3] = qn[3];
}
// spatial vectors are (angular; linear).  r = v x s (motion), r = v x* f (force)
template <class R, class V, class S_>
__device__ __forceinline__ void cross_motion(R r, V v, S_ s) {
  double vv[6], ss[6];
  for (int i = 0; i < 6; i++) { vv[i] = v[i]; ss[i] = s[i]; }
  double a[3], b[3], w[3];
  cross3(w, vv, ss); cross3(a, vv, ss + 3); cross3(b, vv + 3, ss);
  r[0] = w[0]; r[1] = w[1]; r[2] = w[2];
  r[3] = a[0] + b[0]; r[4] = a[1] + b[1]; r[5] = a[2] + b[2];
}
template <class R, class V, class F>
__device__ __forceinline__ void cross_force(R r, V v, F f) {
  double vv[6], ff[6];
  for (int i = 0; i < 6; i++) { vv[i] = v[i]; ff[i] = f[i]; }
  double a[3], b[3], w[3];
  cross3(a, vv, ff); cross3(b, vv + 3, ff + 3); cross3(w, vv, ff + 3);
  r[0] = a[0] + b[0]; r[1] = a[1] + b[1]; r[2] = a[2] + b[2];
  r[3] = w[0]; r[4] = w[1]; r[5] = w[2];
}
// 10-vector inertia (Ixx Iyy Izz Ixy Ixz Iyz, m*c, m) times a motion vector
template <class R, class I_, class V>
__device__ __forceinline__ void inert_mul(R r, I_ Iin, V vin) {
  double I[10], v[6];
  for (int i = 0; i < 10; i++) I[i] = Iin[i];
  for (int i = 0; i < 6; i++) v[i] = vin[i];
  r[0] = I[0] * v[0] + I[3] * v[1] + I[4] * v[2] - I[8] * v[4] + I[7] * v[5];
  r[1] = I[3] * v[0] + I[1] * v[1] + I[5] * v[2] + I[8] * v[3] - I[6] * v[5];
  r[2] = I[4] * v[0] + I[5] * v[1] + I[2] * v[2] - I[7] * v[3] + I[6] * v[4];
  r[3] = I[8] * v[1] - I[7] * v[2] + I[9] * v[3];
  r[4] = I[6] * v[2] - I[8] * v[0] + I[9] * v[4];
  r[5] = I[7] * v[0] - I[6] * v[1] + I[9] * v[5];
}
template <class A, class B>
__device__ __forceinline__ double dot6(A a, B b) {
  return a[0] * b[0] + a[1] * b[1] + a[2] * b[2] + a[3] * b[3] + a[4] * b[4] +
         a[5] * b[5];
}
// group-wide reductions / scans (all lanes of the group must call them together)
template <class C>
__device__ __forceinline__ double group_sum(const C& c, double v) {
#pragma unroll
  for (int o = C::G / 2; o > 0; o >>= 1) v += __shfl_xor_sync(c.mask, v, o, C::G);
  return v;
}
template <class C>
__device__ __forceinline__ bool group_any(const C& c, bool x) {
  if (C::G == 1) return x;
  return (__ballot_sync(c.mask, x) & c.mask) != 0;
}
template <class C>
__device__ __forceinline__ int group_excl_scan(const C& c, int n, int* total) {
  int incl = n;
#pragma unroll
  for (int o = 1; o < C::G; o <<= 1) {
    int t = __shfl_up_sync(c.mask, incl, o, C::G);
    if (c.lane >= o) incl += t;
  }
  *total = __shfl_sync(c.mask, incl, C::G - 1, C::G);
  return incl - n;
}
__device__ __forceinline__ bool is_bad(double x) {
  return isnan(x) || x > BP_MAXVAL || x < -BP_MAXVAL;
}
__device__ __forceinline__ double clampd(double x, double lo, double hi) {
  return x < lo ? lo : (x > hi ? hi : x);
}
#endif  // BP_MATH_CUH_
