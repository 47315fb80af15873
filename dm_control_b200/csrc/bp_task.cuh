// Suite observation / reward functions as one device function, shared by the stand-alone
// task kernel (bp_task.cu) and the fused epilogue of the step kernel (bp_pipeline.cuh).
//
// Each task mirrors get_observation()/get_reward() of the reference's task class
// (file:line cited per case) on the data image in HBM; the few fields it reads are
// addressed through the image layout.  `utils/rewards.py` tolerance()
// (/root/reference/dm_control/utils/rewards.py:93-135) is restated for the three sigmoids
// the configured tasks use.
#ifndef BP_TASK_CUH_
#define BP_TASK_CUH_
#include <cuda_runtime.h>
#include <math.h>

#include "bp_types.h"

enum { TASK_NONE = 0, TASK_CARTPOLE = 1, TASK_ACROBOT = 2, TASK_PENDULUM = 3,
       TASK_CHEETAH = 4, TASK_WALKER = 5, TASK_HUMANOID = 6, TASK_HOPPER = 7,
       TASK_REACHER = 8, TASK_FINGER = 9, TASK_QUADRUPED = 10, TASK_QUADRUPED_ESCAPE = 11,
       TASK_LQR = 12, TASK_POINT_MASS = 13, TASK_FINGER_TURN = 14,
       TASK_SWIMMER = 15, TASK_FISH_UPRIGHT = 16,
       TASK_FISH_SWIM = 17, TASK_BALL_IN_CUP = 18,
       TASK_MANIPULATOR = 19, TASK_QUADRUPED_FETCH = 20, TASK_STACKER = 21 };
enum { SIG_GAUSSIAN = 0, SIG_LINEAR = 1, SIG_QUADRATIC = 2, SIG_LONG_TAIL = 3 };

// read-only strided window of one environment's image (see bp_math.cuh for layouts)
struct ImgView {
  const double* p; int S;
  __device__ __forceinline__ double operator[](int i) const { return p[(long long)i * S]; }
  __device__ __forceinline__ ImgView at(int off) const { return ImgView{p + (long long)off * S, S}; }
};
__device__ __forceinline__ double* bp_env_base(const BpParams& P, long long env) {
  const int S = P.interleave;
  return P.image + (env / S) * P.stride * S + (env % S);
}

__device__ __forceinline__ double sigmoid_(double x, double v1, int kind) {
  if (kind == SIG_GAUSSIAN) {
    double scale = sqrt(-2 * log(v1));
    return exp(-0.5 * (x * scale) * (x * scale));
  } else if (kind == SIG_LINEAR) {
    double sx = x * (1 - v1);
    return fabs(sx) < 1 ? 1 - sx : 0.0;
  } else if (kind == SIG_LONG_TAIL) {      // rewards.py:64-66
    double scale = sqrt(1 / v1 - 1);
    return 1 / ((x * scale) * (x * scale) + 1);
  } else {
    double sx = x * sqrt(1 - v1);
    return fabs(sx) < 1 ? 1 - sx * sx : 0.0;
  }
}

// rewards.tolerance (utils/rewards.py:93-135)
__device__ __forceinline__ double tolerance(double x, double lower, double upper,
                                            double margin, int sig, double v1) {
  bool in = lower <= x && x <= upper;
  if (in) return 1.0;
  if (margin == 0) return 0.0;
  double d = (x < lower ? lower - x : x - upper) / margin;
  return sigmoid_(d, v1, sig);
}

// observation (obs_out + env * nobs, may be null) and reward (rew_out[env]) of one env
static __device__ __noinline__ void bp_task_eval(const BpParams& P, const BpTask& T, long long env,
                                                 int nobs, double* __restrict__ obs_out,
                                                 double* __restrict__ rew_out) {
  const BpModel& m = P.m;
  const BpLayout& L = P.L;
  const ImgView d{bp_env_base(P, env), P.interleave};
  const ImgView qpos = d.at(L.qpos), qvel = d.at(L.qvel), xmat = d.at(L.xmat), xpos = d.at(L.xpos);
  const ImgView ctrl = d.at(L.ctrl), sens = d.at(L.sensordata);
  double* o = obs_out ? obs_out + env * (long long)nobs : nullptr;
  double r = 0;
  int k = 0;
  const double INF = 1.0 / 0.0;
  switch (T.kind) {
    case TASK_CARTPOLE: {   // suite/cartpole.py:131-148,197-221
      double cart = qpos[T.ip[0]];
      if (o) {
        o[k++] = cart;
        for (int b = 2; b < m.nbody; b++) { o[k++] = xmat[9 * b + 8]; o[k++] = xmat[9 * b + 2]; }
        for (int i = 0; i < m.nv; i++) o[k++] = qvel[i];
      }
      int npole = m.nbody - 2;
      if (T.ip[1]) {        // sparse
        double ang = 1;
        for (int b = 2; b < m.nbody; b++) ang *= tolerance(xmat[9 * b + 8], .995, 1, 0, 0, .1);
        r = tolerance(cart, -.25, .25, 0, 0, .1) * ang;
      } else {
        double up = 0;
        for (int b = 2; b < m.nbody; b++) up += (xmat[9 * b + 8] + 1) / 2;
        up /= npole;
        double centered = (1 + tolerance(cart, 0, 0, 2, SIG_GAUSSIAN, .1)) / 2;
        double small_control = (4 + tolerance(ctrl[0], 0, 0, 1, SIG_QUADRATIC, 0)) / 5;
        double sv = 1;
        for (int i = 1; i < m.nv; i++) sv = fmin(sv, tolerance(qvel[i], 0, 0, 5, SIG_GAUSSIAN, .1));
        r = up * small_control * ((1 + sv) / 2) * centered;
      }
      break;
    }
    case TASK_ACROBOT: {    // suite/acrobot.py:59-78,108-119
      const ImgView sx = d.at(L.site_xpos);
      int ua = T.ip[0], la = T.ip[1], tgt = T.ip[2], tip = T.ip[3];
      if (o) {
        o[k++] = xmat[9 * ua + 2]; o[k++] = xmat[9 * la + 2];
        o[k++] = xmat[9 * ua + 8]; o[k++] = xmat[9 * la + 8];
        for (int i = 0; i < m.nv; i++) o[k++] = qvel[i];
      }
      double dx = sx[3 * tgt] - sx[3 * tip], dy = sx[3 * tgt + 1] - sx[3 * tip + 1],
             dz = sx[3 * tgt + 2] - sx[3 * tip + 2];
      r = tolerance(sqrt(dx * dx + dy * dy + dz * dz), 0, T.fp[0], T.fp[1], SIG_GAUSSIAN, .1);
      break;
    }
    case TASK_PENDULUM: {   // suite/pendulum.py:51-64,104-110
      int b = T.ip[0];
      if (o) { o[k++] = xmat[9 * b + 8]; o[k++] = xmat[9 * b + 2]; o[k++] = qvel[T.ip[1]]; }
      r = tolerance(xmat[9 * b + 8], T.fp[0], 1, 0, 0, .1);
      break;
    }
    case TASK_CHEETAH: {    // suite/cheetah.py:52-57,79-93
      if (o) {
        for (int i = 1; i < m.nq; i++) o[k++] = qpos[i];
        for (int i = 0; i < m.nv; i++) o[k++] = qvel[i];
      }
      r = tolerance(sens[T.ip[0]], T.fp[0], INF, T.fp[0], SIG_LINEAR, 0);
      break;
    }
    case TASK_WALKER: {     // suite/walker.py:78-98,131-154
      int torso = T.ip[0];
      double height = xpos[3 * torso + 2];
      if (o) {
        for (int b = 1; b < m.nbody; b++) { o[k++] = xmat[9 * b]; o[k++] = xmat[9 * b + 2]; }
        o[k++] = height;
        for (int i = 0; i < m.nv; i++) o[k++] = qvel[i];
      }
      double standing = tolerance(height, 1.2, INF, 0.6, SIG_GAUSSIAN, .1);
      double upright = (1 + xmat[9 * torso + 8]) / 2;
      double stand = (3 * standing + upright) / 4;
      double ms = T.fp[0];
      if (ms == 0) r = stand;
      else {
        double mv = tolerance(sens[T.ip[1]], ms, INF, ms / 2, SIG_LINEAR, .5);
        r = stand * (5 * mv + 1) / 6;
      }
      break;
    }
    case TASK_HUMANOID: {   // suite/humanoid.py:93-129,168-207; suite/humanoid_CMU.py:65-180
      int torso = T.ip[0], head = T.ip[1];
      const ImgView R = xmat.at(9 * torso);
      const ImgView vel = sens.at(T.ip[6]);
      if (o) {
        for (int i = 7; i < m.nq; i++) o[k++] = qpos[i];
        o[k++] = xpos[3 * head + 2];
        for (int e = 0; e < 4; e++) {
          int b = T.ip[2 + e];
          double v[3] = {xpos[3 * b] - xpos[3 * torso], xpos[3 * b + 1] - xpos[3 * torso + 1],
                         xpos[3 * b + 2] - xpos[3 * torso + 2]};
          for (int cidx = 0; cidx < 3; cidx++)
            o[k++] = v[0] * R[cidx] + v[1] * R[3 + cidx] + v[2] * R[6 + cidx];
        }
        o[k++] = R[6]; o[k++] = R[7]; o[k++] = R[8];
        for (int i = 0; i < 3; i++) o[k++] = vel[i];
        for (int i = 0; i < m.nv; i++) o[k++] = qvel[i];
      }
      double standing = tolerance(xpos[3 * head + 2], 1.4, INF, 1.4 / 4, SIG_GAUSSIAN, .1);
      // ip[7]: element of the torso frame read as "upright": zz for humanoid.py:96-98, zy
      // for humanoid_CMU.py:68-70 (its root body is pitched by 90 degrees)
      double upright = tolerance(R[T.ip[7]], 0.9, INF, 1.9, SIG_LINEAR, 0);
      double sc = 0;
      for (int i = 0; i < m.nu; i++) sc += tolerance(ctrl[i], 0, 0, 1, SIG_QUADRATIC, 0);
      sc = (4 + sc / m.nu) / 5;
      double ms = T.fp[0];
      if (ms == 0) {
        double dm = (tolerance(vel[0], 0, 0, 2, SIG_GAUSSIAN, .1) +
                     tolerance(vel[1], 0, 0, 2, SIG_GAUSSIAN, .1)) / 2;
        r = sc * standing * upright * dm;
      } else {
        double cv = sqrt(vel[0] * vel[0] + vel[1] * vel[1]);
        double mv = (5 * tolerance(cv, ms, INF, ms, SIG_LINEAR, 0) + 1) / 6;
        r = sc * standing * upright * mv;
      }
      break;
    }
    case TASK_HOPPER: {     // suite/hopper.py:80-99,122-148
      // ip: torso body, foot body, speed sensor adr, touch_toe adr, touch_heel adr; fp: hopping
      const ImgView xipos = d.at(L.xipos);
      double height = xipos[3 * T.ip[0] + 2] - xipos[3 * T.ip[1] + 2];
      if (o) {
        for (int i = 1; i < m.nq; i++) o[k++] = qpos[i];
        for (int i = 0; i < m.nv; i++) o[k++] = qvel[i];
        o[k++] = log1p(sens[T.ip[3]]); o[k++] = log1p(sens[T.ip[4]]);
      }
      double standing = tolerance(height, 0.6, 2, 0, SIG_GAUSSIAN, .1);
      if (T.fp[0] != 0) {
        r = standing * tolerance(sens[T.ip[2]], 2, INF, 1, SIG_LINEAR, .5);
      } else {
        double sc = 0;
        for (int i = 0; i < m.nu; i++) sc += tolerance(ctrl[i], 0, 0, 1, SIG_QUADRATIC, 0);
        r = standing * (sc / m.nu + 4) / 5;
      }
      break;
    }
    case TASK_FINGER_TURN: {  // suite/finger.py:87-126,196-215 (turn_easy / turn_hard)
      // ip as TASK_FINGER plus ip[9] = sensordata address of the target site's framepos;
      // fp[1] = target radius
      double tx = sens[T.ip[9]] - sens[T.ip[6]], tz = sens[T.ip[9] + 2] - sens[T.ip[6] + 2];
      double px = sens[T.ip[5]] - sens[T.ip[6]], pz = sens[T.ip[5] + 2] - sens[T.ip[6] + 2];
      double dist = sqrt((tx - px) * (tx - px) + (tz - pz) * (tz - pz)) - T.fp[1];
      if (o) {
        o[k++] = sens[T.ip[0]]; o[k++] = sens[T.ip[1]]; o[k++] = px; o[k++] = pz;
        o[k++] = sens[T.ip[2]]; o[k++] = sens[T.ip[3]]; o[k++] = sens[T.ip[4]];
        o[k++] = log1p(sens[T.ip[7]]); o[k++] = log1p(sens[T.ip[8]]);
        o[k++] = tx; o[k++] = tz; o[k++] = dist;
      }
      r = dist <= 0 ? 1.0 : 0.0;
      break;
    }
    case TASK_FINGER: {     // suite/finger.py:87-112,139-156 (spin)
      // ip: sensordata addresses of proximal, distal, proximal_velocity, distal_velocity,
      // hinge_velocity, tip, spinner, touchtop, touchbottom; fp: spin velocity
      if (o) {
        o[k++] = sens[T.ip[0]]; o[k++] = sens[T.ip[1]];
        o[k++] = sens[T.ip[5]] - sens[T.ip[6]];
        o[k++] = sens[T.ip[5] + 2] - sens[T.ip[6] + 2];
        o[k++] = sens[T.ip[2]]; o[k++] = sens[T.ip[3]]; o[k++] = sens[T.ip[4]];
        o[k++] = log1p(sens[T.ip[7]]); o[k++] = log1p(sens[T.ip[8]]);
      }
      r = sens[T.ip[4]] <= -T.fp[0] ? 1.0 : 0.0;
      break;
    }
    case TASK_REACHER: {    // suite/reacher.py:63-70,103-114; ip: target geom, finger geom;
      // fp: radii = target size + finger size
      const ImgView gx = d.at(L.geom_xpos);
      double dx = gx[3 * T.ip[0]] - gx[3 * T.ip[1]], dy = gx[3 * T.ip[0] + 1] - gx[3 * T.ip[1] + 1];
      if (o) {
        for (int i = 0; i < m.nq; i++) o[k++] = qpos[i];
        o[k++] = dx; o[k++] = dy;
        for (int i = 0; i < m.nv; i++) o[k++] = qvel[i];
      }
      r = tolerance(sqrt(dx * dx + dy * dy), 0, T.fp[0], 0, SIG_GAUSSIAN, .1);
      break;
    }
    case TASK_POINT_MASS: { // suite/point_mass.py:59-67,110-127; ip: target geom, pointmass
      // geom; fp: target size
      const ImgView gx = d.at(L.geom_xpos);
      double dd = 0;
      for (int i = 0; i < 3; i++) {
        double e = gx[3 * T.ip[0] + i] - gx[3 * T.ip[1] + i];
        dd += e * e;
      }
      if (o) {
        for (int i = 0; i < m.nq; i++) o[k++] = qpos[i];
        for (int i = 0; i < m.nv; i++) o[k++] = qvel[i];
      }
      double near_target = tolerance(sqrt(dd), 0, T.fp[0], T.fp[0], SIG_GAUSSIAN, .1);
      double cr = 0;
      for (int i = 0; i < m.nu; i++) cr += tolerance(ctrl[i], 0, 0, 1, SIG_QUADRATIC, 0);
      r = near_target * (cr / m.nu + 4) / 5;
      break;
    }
    case TASK_SWIMMER: {    // suite/swimmer.py:139-160,199-213
      // ip: target geom, nose geom, head body, sensordata address of head_vel; fp: target size
      const ImgView gx = d.at(L.geom_xpos);
      const ImgView R = xmat.at(9 * T.ip[2]);
      double v[3], tt[2];
      for (int i = 0; i < 3; i++) v[i] = gx[3 * T.ip[0] + i] - gx[3 * T.ip[1] + i];
      for (int cidx = 0; cidx < 2; cidx++)
        tt[cidx] = v[0] * R[cidx] + v[1] * R[3 + cidx] + v[2] * R[6 + cidx];
      if (o) {
        for (int i = 3; i < m.nq; i++) o[k++] = qpos[i];
        o[k++] = tt[0]; o[k++] = tt[1];
        for (int b = 0; b < m.nbody - 1; b++) {
          int a = T.ip[3] + 6 * b;
          o[k++] = sens[a]; o[k++] = sens[a + 1]; o[k++] = sens[a + 5];
        }
      }
      r = tolerance(sqrt(tt[0] * tt[0] + tt[1] * tt[1]), 0, T.fp[0], 5 * T.fp[0], SIG_LONG_TAIL, .1);
      break;
    }
    case TASK_FISH_UPRIGHT: case TASK_FISH_SWIM: {
      // suite/fish.py:72-94 (observables), 114-125 (Upright), 155-171 (Swim)
      // ip: torso body, first tail / fin joint qpos address, number of those joints,
      // target geom, mouth geom; fp: radii = mouth size + target size
      const bool swim = T.kind == TASK_FISH_SWIM;
      const double up = xmat[9 * T.ip[0] + 8];
      double tt[3] = {0, 0, 0};
      if (swim) {
        const ImgView gx = d.at(L.geom_xpos);
        const ImgView R = d.at(L.geom_xmat).at(9 * T.ip[4]);
        double v[3];
        for (int i = 0; i < 3; i++) v[i] = gx[3 * T.ip[3] + i] - gx[3 * T.ip[4] + i];
        for (int cidx = 0; cidx < 3; cidx++)
          tt[cidx] = v[0] * R[cidx] + v[1] * R[3 + cidx] + v[2] * R[6 + cidx];
      }
      if (o) {
        for (int i = 0; i < T.ip[2]; i++) o[k++] = qpos[T.ip[1] + i];
        o[k++] = up;
        if (swim) for (int i = 0; i < 3; i++) o[k++] = tt[i];
        for (int i = 0; i < m.nv; i++) o[k++] = qvel[i];
      }
      if (swim) {
        double in_target = tolerance(sqrt(tt[0] * tt[0] + tt[1] * tt[1] + tt[2] * tt[2]), 0,
                                     T.fp[0], 2 * T.fp[0], SIG_GAUSSIAN, .1);
        r = (7 * in_target + 0.5 * (up + 1)) / 8;
      } else {
        r = tolerance(up, 1, 1, 1, SIG_GAUSSIAN, .1);
      }
      break;
    }
    case TASK_BALL_IN_CUP: {  // suite/ball_in_cup.py:49-63,99-110
      // ip: target site, ball body; fp: target half sizes (x, z) minus the ball radius
      const ImgView sx = d.at(L.site_xpos), xp = d.at(L.xpos);
      if (o) {
        for (int i = 0; i < m.nq; i++) o[k++] = qpos[i];
        for (int i = 0; i < m.nv; i++) o[k++] = qvel[i];
      }
      double ex = fabs(sx[3 * T.ip[0]] - xp[3 * T.ip[1]]);
      double ez = fabs(sx[3 * T.ip[0] + 2] - xp[3 * T.ip[1] + 2]);
      r = (ex < T.fp[0] && ez < T.fp[1]) ? 1.0 : 0.0;
      break;
    }
    case TASK_MANIPULATOR: {  // suite/manipulator.py:133-163 (observables), 258-297
      // ip[0..7]: qpos (= dof) address of the arm joints, [8] first touch sensor address,
      // [9] hand body, [10] object body, [11] target body, [12] first object joint address,
      // [13] object site, [14] target site; fp: _CLOSE
      const ImgView sx = d.at(L.site_xpos), xp = d.at(L.xpos), xq = d.at(L.xquat);
      if (o) {
        for (int i = 0; i < 8; i++) { o[k++] = sin(qpos[T.ip[i]]); o[k++] = cos(qpos[T.ip[i]]); }
        for (int i = 0; i < 8; i++) o[k++] = qvel[T.ip[i]];
        for (int i = 0; i < 5; i++) o[k++] = log1p(sens[T.ip[8] + i]);
        for (int q = 0; q < 3; q++) {
          const int b = T.ip[9 + q];
          o[k++] = xp[3 * b]; o[k++] = xp[3 * b + 2]; o[k++] = xq[4 * b]; o[k++] = xq[4 * b + 2];
          if (q == 1) for (int i = 0; i < 3; i++) o[k++] = qvel[T.ip[12] + i];
        }
      }
      auto close = [&](int s1, int s2) {        // _is_close(site_distance(s1, s2))
        double dd = 0;
        for (int i = 0; i < 3; i++) {
          double e = sx[3 * s1 + i] - sx[3 * s2 + i];
          dd += e * e;
        }
        return tolerance(sqrt(dd), 0, T.fp[0], 2 * T.fp[0], SIG_GAUSSIAN, .1);
      };
      if (T.ip[15]) {
        // _peg_reward (manipulator.py:276-286).  ip[16..21]: sites peg_grasp, grasp, peg_pinch,
        // pinch, target_peg_tip, peg_tip
        const double grasping = (close(T.ip[16], T.ip[17]) + close(T.ip[18], T.ip[19])) / 2;
        const double bringing = (close(T.ip[13], T.ip[14]) + close(T.ip[20], T.ip[21])) / 2;
        r = fmax(bringing, grasping / 3);
      } else {
        r = close(T.ip[13], T.ip[14]);        // _ball_reward
      }
      break;
    }
    case TASK_STACKER: {      // suite/stacker.py:187-214
      // ip[0..7]: arm joint addresses, [8] first touch sensor, [9] hand body, [10] number of
      // boxes, [11] first box body, [12] first box joint (dof) address, [13] target body,
      // [14] grasp site, [15] target site, [16] first box site; fp: box size, _CLOSE
      const ImgView sx = d.at(L.site_xpos), xp = d.at(L.xpos), xq = d.at(L.xquat);
      const int nb = T.ip[10];
      if (o) {
        for (int i = 0; i < 8; i++) { o[k++] = sin(qpos[T.ip[i]]); o[k++] = cos(qpos[T.ip[i]]); }
        for (int i = 0; i < 8; i++) o[k++] = qvel[T.ip[i]];
        for (int i = 0; i < 5; i++) o[k++] = log1p(sens[T.ip[8] + i]);
        { const int b = T.ip[9];
          o[k++] = xp[3 * b]; o[k++] = xp[3 * b + 2]; o[k++] = xq[4 * b]; o[k++] = xq[4 * b + 2]; }
        for (int q = 0; q < nb; q++) {
          const int b = T.ip[11] + q;
          o[k++] = xp[3 * b]; o[k++] = xp[3 * b + 2]; o[k++] = xq[4 * b]; o[k++] = xq[4 * b + 2];
        }
        // stacker.py:133-136 names the box joints x, y, z; the model declares them x, z, y
        for (int q = 0; q < nb; q++) {
          const int a = T.ip[12] + 3 * q;
          o[k++] = qvel[a]; o[k++] = qvel[a + 2]; o[k++] = qvel[a + 1];
        }
        o[k++] = xp[3 * T.ip[13]]; o[k++] = xp[3 * T.ip[13] + 2];
      }
      auto dist = [&](int s1, int s2) {
        double dd = 0;
        for (int i = 0; i < 3; i++) {
          double e = sx[3 * s1 + i] - sx[3 * s2 + i];
          dd += e * e;
        }
        return sqrt(dd);
      };
      double dmin = dist(T.ip[16], T.ip[15]);
      for (int q = 1; q < nb; q++) dmin = fmin(dmin, dist(T.ip[16] + q, T.ip[15]));
      const double box_is_close = tolerance(dmin, 0, 0, 2 * T.fp[0], SIG_GAUSSIAN, .1);
      const double hand_is_far = tolerance(dist(T.ip[14], T.ip[15]), .1, INF, T.fp[1],
                                           SIG_GAUSSIAN, .1);
      r = box_is_close * hand_is_far;
      break;
    }
    case TASK_LQR: {        // suite/lqr.py:236-250; fp: control cost coefficient
      if (o) {
        for (int i = 0; i < m.nq; i++) o[k++] = qpos[i];
        for (int i = 0; i < m.nv; i++) o[k++] = qvel[i];
      }
      double sc = 0, cc = 0;
      for (int i = 0; i < m.nq; i++) sc += qpos[i] * qpos[i];
      for (int i = 0; i < m.nu; i++) cc += ctrl[i] * ctrl[i];
      r = 1 - (0.5 * sc + 0.5 * cc * T.fp[0]);
      break;
    }
    case TASK_QUADRUPED: case TASK_QUADRUPED_ESCAPE: case TASK_QUADRUPED_FETCH: {
      // suite/quadruped.py:163-216,280-288 (observations), 291-306,343-357 (Move reward),
      // 399-418 (Escape).  ip: torso body, velocimeter adr, accelerometer adr, gyro adr,
      // first force/torque adr, workspace site, first rangefinder adr;
      // fp: desired speed, cos(deviation angle), terrain half size
      // Fetch (quadruped.py:218-245,445-479) adds ip[7] ball body, [8] target site, [9] ball
      // dof address, [10] root dof address; fp[3] workspace + ball radius, [4] arena
      // radius, [5] target radius
      const int torso = T.ip[0];
      const ImgView R = xmat.at(9 * torso), vel = sens.at(T.ip[1]);
      const bool escape = T.kind == TASK_QUADRUPED_ESCAPE;
      const bool fetch = T.kind == TASK_QUADRUPED_FETCH;
      if (o) {
        const ImgView act = d.at(L.act);
        for (int j = 0; j < m.njnt; j++) if (m.jnt_type[j] == JNT_HINGE) o[k++] = qpos[m.jnt_qposadr[j]];
        for (int j = 0; j < m.njnt; j++) if (m.jnt_type[j] == JNT_HINGE) o[k++] = qvel[m.jnt_dofadr[j]];
        for (int i = 0; i < m.na; i++) o[k++] = act[i];
        for (int i = 0; i < 3; i++) o[k++] = vel[i];
        o[k++] = R[8];
        for (int i = 0; i < 3; i++) o[k++] = sens[T.ip[2] + i];
        for (int i = 0; i < 3; i++) o[k++] = sens[T.ip[3] + i];
        for (int i = 0; i < 3 * m.nsens_ft; i++) o[k++] = asinh(sens[T.ip[4] + i]);
        if (escape) {
          for (int cidx = 0; cidx < 3; cidx++)
            o[k++] = -(xpos[3 * torso] * R[cidx] + xpos[3 * torso + 1] * R[3 + cidx] +
                       xpos[3 * torso + 2] * R[6 + cidx]);
          for (int i = 0; i < m.nsens_rf; i++) {
            double rf = sens[T.ip[6] + i];
            o[k++] = rf == -1.0 ? 1.0 : tanh(rf);
          }
        }
        if (fetch) {
          const ImgView tg = d.at(L.site_xpos + 3 * T.ip[8]);
          double rows[4][3];
          for (int i = 0; i < 3; i++) {
            rows[0][i] = xpos[3 * T.ip[7] + i] - xpos[3 * torso + i];
            rows[1][i] = qvel[T.ip[9] + i] - qvel[T.ip[10] + i];
            rows[2][i] = qvel[T.ip[9] + 3 + i];
            rows[3][i] = tg[i] - xpos[3 * torso + i];
          }
          for (int q = 0; q < 4; q++)
            for (int cidx = 0; cidx < 3; cidx++)
              o[k++] = rows[q][0] * R[cidx] + rows[q][1] * R[3 + cidx] + rows[q][2] * R[6 + cidx];
        }
      }
      const double dev = T.fp[1];
      double upright = tolerance(R[8], dev, INF, 1 + dev, SIG_LINEAR, 0);
      if (escape) {
        const ImgView ws = d.at(L.site_xpos + 3 * T.ip[5]);
        double dist = sqrt(ws[0] * ws[0] + ws[1] * ws[1] + ws[2] * ws[2]);
        r = upright * tolerance(dist, T.fp[2], INF, T.fp[2], SIG_LINEAR, 0);
      } else if (fetch) {
        const ImgView ws = d.at(L.site_xpos + 3 * T.ip[5]), tg = d.at(L.site_xpos + 3 * T.ip[8]);
        const double bx = xpos[3 * T.ip[7]], by = xpos[3 * T.ip[7] + 1];
        double reach = tolerance(sqrt((ws[0] - bx) * (ws[0] - bx) + (ws[1] - by) * (ws[1] - by)),
                                 0, T.fp[3], T.fp[4], SIG_LINEAR, 0);
        double fetched = tolerance(sqrt((tg[0] - bx) * (tg[0] - bx) + (tg[1] - by) * (tg[1] - by)),
                                   0, T.fp[5], T.fp[4], SIG_LINEAR, 0);
        r = upright * (reach * (0.5 + 0.5 * fetched));
      } else {
        r = upright * tolerance(vel[0], T.fp[0], INF, T.fp[0], SIG_LINEAR, .5);
      }
      break;
    }
    default: break;
  }
  if (rew_out) rew_out[env] = r;
}

#endif  // BP_TASK_CUH_
