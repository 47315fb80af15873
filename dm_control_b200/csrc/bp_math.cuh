// Small fixed-size math used by every stage, plus the warp context.
#ifndef BP_MATH_CUH_
#define BP_MATH_CUH_
#include "bp_types.h"

struct Ctx {
  const BpParams& P;
  double* sm;   // this env's image in shared memory
  int* ip;      // int32 region of the image
  int lane;
};

#define DF(f) (c.sm + c.P.L.f)
#define IF(f) (c.ip + c.P.L.f)
#define SYNC() __syncwarp()
#define FULL 0xffffffffu

__device__ __forceinline__ double dot3(const double* a, const double* b) {
  return a[0] * b[0] + a[1] * b[1] + a[2] * b[2];
}
__device__ __forceinline__ void cross3(double* r, const double* a, const double* b) {
  double x = a[1] * b[2] - a[2] * b[1], y = a[2] * b[0] - a[0] * b[2],
         z = a[0] * b[1] - a[1] * b[0];
  r[0] = x; r[1] = y; r[2] = z;
}
__device__ __forceinline__ double norm3(const double* a) { return sqrt(dot3(a, a)); }
__device__ __forceinline__ double normalize3(double* a) {
  double n = norm3(a);
  if (n < BP_MINVAL) { a[0] = 1; a[1] = 0; a[2] = 0; }
  else { a[0] /= n; a[1] /= n; a[2] /= n; }
  return n;
}
__device__ __forceinline__ void normalize4(double* q) {
  double n = sqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
  if (n < BP_MINVAL) { q[0] = 1; q[1] = q[2] = q[3] = 0; }
  else { q[0] /= n; q[1] /= n; q[2] /= n; q[3] /= n; }
}
__device__ __forceinline__ void quat_mul(double* r, const double* a, const double* b) {
  r[0] = a[0] * b[0] - a[1] * b[1] - a[2] * b[2] - a[3] * b[3];
  r[1] = a[0] * b[1] + a[1] * b[0] + a[2] * b[3] - a[3] * b[2];
  r[2] = a[0] * b[2] - a[1] * b[3] + a[2] * b[0] + a[3] * b[1];
  r[3] = a[0] * b[3] + a[1] * b[2] - a[2] * b[1] + a[3] * b[0];
}
__device__ __forceinline__ void quat2mat(double* R, const double* q) {
  double q00 = q[0] * q[0], q01 = q[0] * q[1], q02 = q[0] * q[2], q03 = q[0] * q[3];
  double q11 = q[1] * q[1], q12 = q[1] * q[2], q13 = q[1] * q[3];
  double q22 = q[2] * q[2], q23 = q[2] * q[3], q33 = q[3] * q[3];
  R[0] = q00 + q11 - q22 - q33; R[4] = q00 - q11 + q22 - q33;
  R[8] = q00 - q11 - q22 + q33;
  R[1] = 2 * (q12 - q03); R[2] = 2 * (q13 + q02);
  R[3] = 2 * (q12 + q03); R[5] = 2 * (q23 - q01);
  R[6] = 2 * (q13 - q02); R[7] = 2 * (q23 + q01);
}
__device__ __forceinline__ void mat_vec3(double* r, const double* R, const double* v) {
  r[0] = R[0] * v[0] + R[1] * v[1] + R[2] * v[2];
  r[1] = R[3] * v[0] + R[4] * v[1] + R[5] * v[2];
  r[2] = R[6] * v[0] + R[7] * v[1] + R[8] * v[2];
}
__device__ __forceinline__ void matT_vec3(double* r, const double* R, const double* v) {
  r[0] = R[0] * v[0] + R[3] * v[1] + R[6] * v[2];
  r[1] = R[1] * v[0] + R[4] * v[1] + R[7] * v[2];
  r[2] = R[2] * v[0] + R[5] * v[1] + R[8] * v[2];
}
__device__ __forceinline__ void axis_angle_quat(double* q, const double* axis, double ang) {
  double s, co;
  sincos(0.5 * ang, &s, &co);
  q[0] = co; q[1] = axis[0] * s; q[2] = axis[1] * s; q[3] = axis[2] * s;
}
// q <- q * exp(w*dt/2), w in the body frame (free/ball joint velocity)
__device__ __forceinline__ void quat_integrate(double* q, const double* w, double dt) {
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
__device__ __forceinline__ void cross_motion(double* r, const double* v, const double* s) {
  cross3(r, v, s);
  double a[3], b[3];
  cross3(a, v, s + 3); cross3(b, v + 3, s);
  r[3] = a[0] + b[0]; r[4] = a[1] + b[1]; r[5] = a[2] + b[2];
}
__device__ __forceinline__ void cross_force(double* r, const double* v, const double* f) {
  double a[3], b[3];
  cross3(a, v, f); cross3(b, v + 3, f + 3);
  r[0] = a[0] + b[0]; r[1] = a[1] + b[1]; r[2] = a[2] + b[2];
  cross3(r + 3, v, f + 3);
}
// 10-vector inertia (Ixx Iyy Izz Ixy Ixz Iyz, m*c, m) times a motion vector
__device__ __forceinline__ void inert_mul(double* r, const double* I, const double* v) {
  r[0] = I[0] * v[0] + I[3] * v[1] + I[4] * v[2] - I[8] * v[4] + I[7] * v[5];
  r[1] = I[3] * v[0] + I[1] * v[1] + I[5] * v[2] + I[8] * v[3] - I[6] * v[5];
  r[2] = I[4] * v[0] + I[5] * v[1] + I[2] * v[2] - I[7] * v[3] + I[6] * v[4];
  r[3] = I[8] * v[1] - I[7] * v[2] + I[9] * v[3];
  r[4] = I[6] * v[2] - I[8] * v[0] + I[9] * v[4];
  r[5] = I[7] * v[0] - I[6] * v[1] + I[9] * v[5];
}
__device__ __forceinline__ double dot6(const double* a, const double* b) {
  return a[0] * b[0] + a[1] * b[1] + a[2] * b[2] + a[3] * b[3] + a[4] * b[4] +
         a[5] * b[5];
}
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(FULL, v, o);
  return v;
}
__device__ __forceinline__ bool is_bad(double x) {
  return isnan(x) || x > BP_MAXVAL || x < -BP_MAXVAL;
}
__device__ __forceinline__ double clampd(double x, double lo, double hi) {
  return x < lo ? lo : (x > hi ? hi : x);
}
#endif  // BP_MATH_CUH_
