/* CPU ORACLE — TEST INFRASTRUCTURE ONLY (see oracle.h).  fp64, one env per OData.
 *
 * Stage list follows the reference's call sequence Physics.step() ->
 * mj_step2 + mj_step1 (/root/reference/dm_control/mujoco/engine.py:149-166) and
 * SURVEY.md §3.1 / Appendix A ([MJ-doc] = restated from MuJoCo's documentation,
 * not from the reference tree, which holds no engine source).
 */
#include "oracle.h"

#include <math.h>
#include <pthread.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#define MINVAL 1e-15
#define MAXVAL 1e10
#define PI 3.14159265358979323846

enum { GEOM_PLANE, GEOM_HFIELD, GEOM_SPHERE, GEOM_CAPSULE, GEOM_ELLIPSOID,
       GEOM_CYLINDER, GEOM_BOX, GEOM_MESH };
enum { JNT_FREE, JNT_BALL, JNT_SLIDE, JNT_HINGE };
enum { DSBL_CONSTRAINT = 1, DSBL_EQUALITY = 2, DSBL_FRICTIONLOSS = 4,
       DSBL_LIMIT = 8, DSBL_CONTACT = 16, DSBL_PASSIVE = 32, DSBL_GRAVITY = 64,
       DSBL_CLAMPCTRL = 128, DSBL_WARMSTART = 256, DSBL_FILTERPARENT = 512,
       DSBL_ACTUATION = 1024, DSBL_REFSAFE = 2048 };
enum { ENBL_OVERRIDE = 1, ENBL_ENERGY = 2 };
enum { CN_EQUALITY, CN_FRICTION_DOF, CN_FRICTION_TENDON, CN_LIMIT_JOINT,
       CN_LIMIT_TENDON, CN_CONTACT_FRICTIONLESS, CN_CONTACT_PYRAMIDAL,
       CN_CONTACT_ELLIPTIC };
enum { ST_SATISFIED, ST_QUADRATIC, ST_LINEARNEG, ST_LINEARPOS, ST_CONE };
enum { W_INERTIA, W_CONTACTFULL, W_CNSTRFULL, W_VGEOMFULL, W_BADQPOS,
       W_BADQVEL, W_BADQACC, W_BADCTRL };
enum { SENS_TOUCH = 0, SENS_ACCELEROMETER = 1, SENS_VELOCIMETER = 2,
       SENS_GYRO = 3, SENS_FORCE = 4, SENS_TORQUE = 5, SENS_RANGEFINDER = 7,
       SENS_JOINTPOS = 8,
       SENS_JOINTVEL = 9, SENS_ACTUATORFRC = 14, SENS_FRAMEPOS = 23,
       SENS_FRAMEXAXIS = 25, SENS_FRAMEYAXIS = 26, SENS_FRAMEZAXIS = 27,
       SENS_SUBTREECOM = 32, SENS_SUBTREELINVEL = 33 };
enum { OBJ_BODY = 1, OBJ_XBODY = 2, OBJ_GEOM = 5, OBJ_SITE = 6 };
enum { STAGE_POS = 1, STAGE_VEL = 2, STAGE_ACC = 3 };

/* ------------------------------------------------------------------ model */

typedef struct {
  char name[40];
  int32_t dtype, ndim, dims[4];
  int64_t offset, count;
} BlobEntry;

#define MODEL_INT_SCALARS(X) \
  X(nq) X(nv) X(nu) X(na) X(nbody) X(njnt) X(ngeom) X(nsite) X(nM) X(npair) \
  X(nconmax) X(njmax) X(nsensor) X(nsensordata) X(nkey) X(ntendon) X(neq) \
  X(nhfield) X(nhfielddata) X(nhfpair) X(nspatial) \
  X(opt_integrator) X(opt_cone) X(opt_solver) X(opt_iterations) \
  X(opt_ls_iterations) X(opt_disableflags) X(opt_enableflags)
#define MODEL_F64_SCALARS(X) \
  X(opt_timestep) X(opt_impratio) X(opt_tolerance) X(opt_ls_tolerance) \
  X(stat_meaninertia) X(opt_density) X(opt_viscosity)
#define MODEL_INT_ARRAYS(X) \
  X(body_parentid) X(body_rootid) X(body_weldid) X(body_jntnum) X(body_jntadr) \
  X(body_dofnum) X(body_dofadr) X(jnt_type) X(jnt_qposadr) X(jnt_dofadr) \
  X(jnt_bodyid) X(jnt_limited) X(dof_bodyid) X(dof_jntid) X(dof_parentid) \
  X(dof_Madr) X(geom_type) X(geom_condim) X(geom_priority) X(geom_bodyid) X(site_bodyid) \
  X(site_type) X(pair_geom1) X(pair_geom2) X(pair_parentchild) \
  X(actuator_trntype) X(actuator_dyntype) X(actuator_gaintype) \
  X(actuator_biastype) X(actuator_trnid) X(actuator_ctrllimited) \
  X(actuator_forcelimited) X(sensor_type) X(sensor_objtype) X(sensor_objid) \
  X(sensor_dim) X(sensor_adr) X(sensor_needstage) X(tendon_adr) X(tendon_num) \
  X(wrap_objid) X(eq_type) X(eq_obj1id) X(eq_obj2id) X(eq_active) \
  X(geom_dataid) X(hfield_nrow) X(hfield_ncol) X(hfield_adr) X(wrap_type) \
  X(tendon_limited) X(tendon_Jadr)
#define MODEL_F64_ARRAYS(X) \
  X(opt_gravity) X(body_pos) X(body_quat) X(body_ipos) X(body_iquat) \
  X(body_mass) X(body_subtreemass) X(body_inertia) X(body_invweight0) \
  X(jnt_pos) X(jnt_axis) X(jnt_stiffness) X(jnt_range) X(jnt_margin) \
  X(jnt_solref) X(jnt_solimp) X(dof_armature) X(dof_damping) \
  X(dof_frictionloss) X(dof_solref) X(dof_solimp) X(dof_invweight0) \
  X(geom_size) X(geom_rbound) X(geom_pos) X(geom_quat) X(geom_friction) \
  X(geom_margin) X(geom_gap) X(geom_solmix) X(geom_solref) X(geom_solimp) \
  X(site_pos) X(site_quat) X(site_size) X(qpos0) X(qpos_spring) \
  X(actuator_dynprm) X(actuator_gainprm) X(actuator_biasprm) \
  X(actuator_ctrlrange) X(actuator_forcerange) X(actuator_gear) \
  X(sensor_cutoff) X(key_qpos) X(key_qvel) X(key_act) X(key_time) X(wrap_prm) \
  X(tendon_length0) X(tendon_invweight0) X(tendon_stiffness) X(tendon_damping) X(tendon_range) X(tendon_margin) \
  X(tendon_solref_lim) X(tendon_solimp_lim) X(eq_solref) X(eq_solimp) X(eq_data) \
  X(hfield_size) X(hfield_data)

struct OModel {
  char* blob;
  int64_t nbytes;
  int nfields;
  BlobEntry* entries;
#define X(n) int n;
  MODEL_INT_SCALARS(X)
#undef X
#define X(n) double n;
  MODEL_F64_SCALARS(X)
#undef X
#define X(n) int32_t* n;
  MODEL_INT_ARRAYS(X)
#undef X
#define X(n) double* n;
  MODEL_F64_ARRAYS(X)
#undef X
};

static BlobEntry* find_entry(const OModel* m, const char* name) {
  for (int i = 0; i < m->nfields; i++)
    if (strncmp(m->entries[i].name, name, 40) == 0) return &m->entries[i];
  return NULL;
}

void* om_field(OModel* m, const char* name, int64_t* count, int* dtype) {
  BlobEntry* e = find_entry(m, name);
  if (!e) return NULL;
  if (count) *count = e->count;
  if (dtype) *dtype = e->dtype;
  return m->blob + e->offset;
}

OModel* om_create(const void* blob, int64_t nbytes, char* err, int errsz) {
  if (nbytes < 16 || memcmp(blob, "BPMODEL1", 8) != 0) {
    if (err) snprintf(err, errsz, "bad model blob magic");
    return NULL;
  }
  OModel* m = (OModel*)calloc(1, sizeof(OModel));
  m->blob = (char*)malloc(nbytes);
  memcpy(m->blob, blob, nbytes);
  m->nbytes = nbytes;
  m->nfields = *(int32_t*)(m->blob + 8);
  m->entries = (BlobEntry*)(m->blob + 16);
  const char* missing = NULL;
#define X(n) { BlobEntry* e = find_entry(m, #n); if (!e) missing = #n; \
               else m->n = *(int32_t*)(m->blob + e->offset); }
  MODEL_INT_SCALARS(X)
#undef X
#define X(n) { BlobEntry* e = find_entry(m, #n); if (!e) missing = #n; \
               else m->n = *(double*)(m->blob + e->offset); }
  MODEL_F64_SCALARS(X)
#undef X
#define X(n) { BlobEntry* e = find_entry(m, #n); if (!e) missing = #n; \
               else m->n = (int32_t*)(m->blob + e->offset); }
  MODEL_INT_ARRAYS(X)
#undef X
#define X(n) { BlobEntry* e = find_entry(m, #n); if (!e) missing = #n; \
               else m->n = (double*)(m->blob + e->offset); }
  MODEL_F64_ARRAYS(X)
#undef X
  if (missing) {
    if (err) snprintf(err, errsz, "model blob lacks field %s", missing);
    om_free(m);
    return NULL;
  }
  return m;
}

void om_free(OModel* m) {
  if (!m) return;
  free(m->blob);
  free(m);
}

/* ------------------------------------------------------------------- data */

typedef struct {
  double dist, pos[3], frame[9], includemargin, friction[5], solref[2],
      solimp[5], mu;
  int dim, geom1, geom2, exclude, efc_address;
} OContact;

/* name, type, element count (expression over m) */
#define DATA_FIELDS(X) \
  X(qpos, double, m->nq) X(qvel, double, m->nv) X(act, double, m->na) \
  X(ctrl, double, m->nu) X(qacc, double, m->nv) X(qacc_warmstart, double, m->nv) \
  X(act_dot, double, m->na) X(qfrc_applied, double, m->nv) \
  X(xfrc_applied, double, 6 * m->nbody) X(time, double, 1) \
  X(energy, double, 2) X(xpos, double, 3 * m->nbody) \
  X(xquat, double, 4 * m->nbody) X(xmat, double, 9 * m->nbody) \
  X(xipos, double, 3 * m->nbody) X(ximat, double, 9 * m->nbody) \
  X(xanchor, double, 3 * m->njnt) X(xaxis, double, 3 * m->njnt) \
  X(geom_xpos, double, 3 * m->ngeom) X(geom_xmat, double, 9 * m->ngeom) \
  X(site_xpos, double, 3 * m->nsite) X(site_xmat, double, 9 * m->nsite) \
  X(subtree_com, double, 3 * m->nbody) X(cinert, double, 10 * m->nbody) \
  X(cdof, double, 6 * m->nv) X(cdof_dot, double, 6 * m->nv) \
  X(cvel, double, 6 * m->nbody) X(cacc, double, 6 * m->nbody) \
  X(cfrc_int, double, 6 * m->nbody) X(cfrc_ext, double, 6 * m->nbody) \
  X(crb, double, 10 * m->nbody) X(subtree_linvel, double, 3 * m->nbody) \
  X(subtree_angmom, double, 3 * m->nbody) \
  X(qM, double, m->nM) X(qLD, double, m->nM) X(qLDiagInv, double, m->nv) \
  X(qfrc_bias, double, m->nv) X(qfrc_passive, double, m->nv) \
  X(qfrc_actuator, double, m->nv) X(qfrc_smooth, double, m->nv) \
  X(qacc_smooth, double, m->nv) X(qfrc_constraint, double, m->nv) \
  X(actuator_length, double, m->nu) X(actuator_velocity, double, m->nu) \
  X(actuator_moment, double, m->nu * m->nv) X(actuator_force, double, m->nu) \
  X(ten_length, double, m->ntendon) X(ten_velocity, double, m->ntendon) \
  X(ten_J, double, m->nspatial * m->nv) \
  X(sensordata, double, m->nsensordata) \
  X(efc_J, double, m->njmax * m->nv) X(efc_pos, double, m->njmax) \
  X(efc_margin, double, m->njmax) X(efc_frictionloss, double, m->njmax) \
  X(efc_diagApprox, double, m->njmax) X(efc_KBIP, double, 4 * m->njmax) \
  X(efc_D, double, m->njmax) X(efc_R, double, m->njmax) \
  X(efc_vel, double, m->njmax) X(efc_aref, double, m->njmax) \
  X(efc_force, double, m->njmax) X(efc_type, int32_t, m->njmax) \
  X(efc_id, int32_t, m->njmax) X(efc_state, int32_t, m->njmax) \
  X(contact_geom1, int32_t, m->nconmax) X(contact_geom2, int32_t, m->nconmax) \
  X(contact_dim, int32_t, m->nconmax) X(contact_dist, double, m->nconmax) \
  X(contact_pos, double, 3 * m->nconmax) X(contact_frame, double, 9 * m->nconmax) \
  X(contact_efc_address, int32_t, m->nconmax) \
  X(ncon, int32_t, 1) X(nefc, int32_t, 1) X(ne, int32_t, 1) X(nf, int32_t, 1) \
  X(solver_iter, int32_t, 1) X(warning, int32_t, 8) \
  X(solver_improvement, double, 1) X(solver_gradient, double, 1)

struct OData {
  const OModel* m;
#define X(n, t, c) t* n; int64_t n##_count;
  DATA_FIELDS(X)
#undef X
  OContact* contact;
  double* hfield;   /* this env's elevation data (Escape writes it per episode) */
  double *geom_pos, *site_pos;   /* this env's copies (per-episode target moves) */
  double *body_pos, *body_quat;
  double* wrap_prm;   /* this env's tendon coefficients (point_mass.py:85-101 randomize_gains) */
  /* solver scratch */
  double *s_Ma, *s_jar, *s_grad, *s_Mgrad, *s_search, *s_Mv, *s_jv, *s_H,
      *s_Mdense, *s_fluid, *s_AR, *s_Mgold;
};

static int64_t wrap_count(const OModel* m) {
  BlobEntry* e = find_entry(m, "wrap_prm");
  return e ? e->count : 0;
}

OData* od_create(const OModel* m) {
  OData* d = (OData*)calloc(1, sizeof(OData));
  d->m = m;
#define X(n, t, c) d->n##_count = (c); \
  d->n = (t*)calloc((size_t)((c) > 0 ? (c) : 1), sizeof(t));
  DATA_FIELDS(X)
#undef X
  d->contact = (OContact*)calloc(m->nconmax > 0 ? m->nconmax : 1, sizeof(OContact));
  d->hfield = (double*)calloc(m->nhfielddata > 0 ? m->nhfielddata : 1, sizeof(double));
  for (int i = 0; i < m->nhfielddata; i++) d->hfield[i] = m->hfield_data[i];
  d->geom_pos = (double*)calloc(m->ngeom > 0 ? 3 * m->ngeom : 1, sizeof(double));
  d->site_pos = (double*)calloc(m->nsite > 0 ? 3 * m->nsite : 1, sizeof(double));
  for (int i = 0; i < 3 * m->ngeom; i++) d->geom_pos[i] = m->geom_pos[i];
  for (int i = 0; i < 3 * m->nsite; i++) d->site_pos[i] = m->site_pos[i];
  d->body_pos = (double*)calloc(3 * m->nbody + 1, sizeof(double));
  d->body_quat = (double*)calloc(4 * m->nbody + 1, sizeof(double));
  for (int i = 0; i < 3 * m->nbody; i++) d->body_pos[i] = m->body_pos[i];
  for (int i = 0; i < 4 * m->nbody; i++) d->body_quat[i] = m->body_quat[i];
  d->wrap_prm = (double*)calloc(wrap_count(m) + 1, sizeof(double));
  for (int64_t i = 0; i < wrap_count(m); i++) d->wrap_prm[i] = m->wrap_prm[i];
  int nv = m->nv > 0 ? m->nv : 1, nj = m->njmax > 0 ? m->njmax : 1;
  d->s_Ma = calloc(nv, 8); d->s_jar = calloc(nj, 8); d->s_grad = calloc(nv, 8);
  d->s_Mgrad = calloc(nv, 8); d->s_search = calloc(nv, 8); d->s_Mv = calloc(nv, 8);
  d->s_jv = calloc(nj, 8); d->s_H = calloc((size_t)nv * nv, 8);
  d->s_Mdense = calloc((size_t)nv * nv, 8);
  d->s_fluid = calloc((size_t)6 * (m->nbody > 0 ? m->nbody : 1), 8);
  d->s_AR = calloc(m->opt_solver == 0 ? (size_t)nj * nj : 1, 8);     /* PGS only */
  d->s_Mgold = calloc(nv, 8);
  o_reset(m, d);
  return d;
}

void od_free(OData* d) {
  if (!d) return;
#define X(n, t, c) free(d->n);
  DATA_FIELDS(X)
#undef X
  free(d->contact);
  free(d->hfield); free(d->geom_pos); free(d->site_pos);
  free(d->body_pos); free(d->body_quat); free(d->wrap_prm);
  free(d->s_Ma); free(d->s_jar); free(d->s_grad); free(d->s_Mgrad);
  free(d->s_search); free(d->s_Mv); free(d->s_jv); free(d->s_H);
  free(d->s_Mdense); free(d->s_fluid); free(d->s_AR); free(d->s_Mgold);
  free(d);
}

void* od_field(OData* d, const char* name, int64_t* count, int* dtype) {
  if (strcmp(name, "geom_pos") == 0 || strcmp(name, "site_pos") == 0) {
    int g = name[0] == 'g';
    if (count) *count = 3 * (g ? d->m->ngeom : d->m->nsite);
    if (dtype) *dtype = 0;
    return g ? d->geom_pos : d->site_pos;
  }
  if (strcmp(name, "body_pos") == 0 || strcmp(name, "body_quat") == 0) {
    int q = name[5] == 'q';
    if (count) *count = (q ? 4 : 3) * d->m->nbody;
    if (dtype) *dtype = 0;
    return q ? d->body_quat : d->body_pos;
  }
  if (strcmp(name, "wrap_prm") == 0) {
    if (count) *count = wrap_count(d->m);
    if (dtype) *dtype = 0;
    return d->wrap_prm;
  }
  if (strcmp(name, "hfield_data") == 0) {
    if (count) *count = d->m->nhfielddata;
    if (dtype) *dtype = 0;
    return d->hfield;
  }
#define X(n, t, c) if (strcmp(name, #n) == 0) { \
    if (count) *count = d->n##_count; \
    if (dtype) { *dtype = (sizeof(t) == 8) ? 0 : 1; } \
    return d->n; }
  DATA_FIELDS(X)
#undef X
  return NULL;
}

void o_reset(const OModel* m, OData* d) {
#define X(n, t, c) memset(d->n, 0, sizeof(t) * (size_t)((c) > 0 ? (c) : 1));
  DATA_FIELDS(X)
#undef X
  memcpy(d->qpos, m->qpos0, sizeof(double) * m->nq);
}

/* ------------------------------------------------------------- small math */

static inline double dot3(const double* a, const double* b) {
  return a[0] * b[0] + a[1] * b[1] + a[2] * b[2];
}
static inline void cross3(double* r, const double* a, const double* b) {
  double x = a[1] * b[2] - a[2] * b[1], y = a[2] * b[0] - a[0] * b[2],
         z = a[0] * b[1] - a[1] * b[0];
  r[0] = x; r[1] = y; r[2] = z;
}
static inline double norm3(const double* a) { return sqrt(dot3(a, a)); }
static inline double normalize3(double* a) {
  double n = norm3(a);
  if (n < MINVAL) { a[0] = 1; a[1] = 0; a[2] = 0; }
  else { a[0] /= n; a[1] /= n; a[2] /= n; }
  return n;
}
static inline void normalize4(double* q) {
  double n = sqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
  if (n < MINVAL) { q[0] = 1; q[1] = q[2] = q[3] = 0; }
  else { q[0] /= n; q[1] /= n; q[2] /= n; q[3] /= n; }
}
static inline void quat_mul(double* r, const double* a, const double* b) {
  double w = a[0] * b[0] - a[1] * b[1] - a[2] * b[2] - a[3] * b[3];
  double x = a[0] * b[1] + a[1] * b[0] + a[2] * b[3] - a[3] * b[2];
  double y = a[0] * b[2] - a[1] * b[3] + a[2] * b[0] + a[3] * b[1];
  double z = a[0] * b[3] + a[1] * b[2] - a[2] * b[1] + a[3] * b[0];
  r[0] = w; r[1] = x; r[2] = y; r[3] = z;
}
static inline void quat2mat(double* R, const double* q) {
  double w = q[0], x = q[1], y = q[2], z = q[3];
  R[0] = w * w + x * x - y * y - z * z; R[1] = 2 * (x * y - w * z);
  R[2] = 2 * (x * z + w * y); R[3] = 2 * (x * y + w * z);
  R[4] = w * w - x * x + y * y - z * z; R[5] = 2 * (y * z - w * x);
  R[6] = 2 * (x * z - w * y); R[7] = 2 * (y * z + w * x);
  R[8] = w * w - x * x - y * y + z * z;
}
static inline void mat_vec3(double* r, const double* R, const double* v) {
  double x = R[0] * v[0] + R[1] * v[1] + R[2] * v[2];
  double y = R[3] * v[0] + R[4] * v[1] + R[5] * v[2];
  double z = R[6] * v[0] + R[7] * v[1] + R[8] * v[2];
  r[0] = x; r[1] = y; r[2] = z;
}
static inline void matT_vec3(double* r, const double* R, const double* v) {
  double x = R[0] * v[0] + R[3] * v[1] + R[6] * v[2];
  double y = R[1] * v[0] + R[4] * v[1] + R[7] * v[2];
  double z = R[2] * v[0] + R[5] * v[1] + R[8] * v[2];
  r[0] = x; r[1] = y; r[2] = z;
}
static inline void axis_angle_quat(double* q, const double* axis, double ang) {
  double s = sin(0.5 * ang);
  q[0] = cos(0.5 * ang); q[1] = axis[0] * s; q[2] = axis[1] * s; q[3] = axis[2] * s;
}
/* integrate a quaternion by a body-frame angular velocity over dt */
static void quat_integrate(double* q, const double* w, double dt) {
  double ax[3] = {w[0], w[1], w[2]};
  double n = norm3(ax);
  if (n < MINVAL) return;
  ax[0] /= n; ax[1] /= n; ax[2] /= n;
  double dq[4], r[4];
  axis_angle_quat(dq, ax, n * dt);
  quat_mul(r, q, dq);
  normalize4(r);
  memcpy(q, r, 32);
}

/* spatial algebra: motion vectors [w; v], force vectors [torque; force], both
 * referred to the kinematic tree's centre of mass */
static void cross_motion(double* r, const double* v, const double* s) {
  double a[3], b[3], c[3];
  cross3(a, v, s);          /* w_v x w_s */
  cross3(b, v, s + 3);      /* w_v x v_s */
  cross3(c, v + 3, s);      /* v_v x w_s */
  r[0] = a[0]; r[1] = a[1]; r[2] = a[2];
  r[3] = b[0] + c[0]; r[4] = b[1] + c[1]; r[5] = b[2] + c[2];
}
static void cross_force(double* r, const double* v, const double* f) {
  double a[3], b[3], c[3];
  cross3(a, v, f);          /* w x torque */
  cross3(b, v + 3, f + 3);  /* v x force  */
  cross3(c, v, f + 3);      /* w x force  */
  r[0] = a[0] + b[0]; r[1] = a[1] + b[1]; r[2] = a[2] + b[2];
  r[3] = c[0]; r[4] = c[1]; r[5] = c[2];
}
/* 10-number spatial inertia about the reference point:
 * [Ixx Iyy Izz Ixy Ixz Iyz  m*dx m*dy m*dz  m] */
static void inert_mul(double* r, const double* I, const double* v) {
  const double* h = I + 6;
  double t[3], f[3];
  t[0] = I[0] * v[0] + I[3] * v[1] + I[4] * v[2];
  t[1] = I[3] * v[0] + I[1] * v[1] + I[5] * v[2];
  t[2] = I[4] * v[0] + I[5] * v[1] + I[2] * v[2];
  double hv[3], wh[3];
  cross3(hv, h, v + 3);     /* (m d) x v  */
  cross3(wh, v, h);         /* w x (m d)  */
  f[0] = I[9] * v[3] + wh[0]; f[1] = I[9] * v[4] + wh[1]; f[2] = I[9] * v[5] + wh[2];
  r[0] = t[0] + hv[0]; r[1] = t[1] + hv[1]; r[2] = t[2] + hv[2];
  r[3] = f[0]; r[4] = f[1]; r[5] = f[2];
}
static inline double dot6(const double* a, const double* b) {
  return a[0] * b[0] + a[1] * b[1] + a[2] * b[2] + a[3] * b[3] + a[4] * b[4] +
         a[5] * b[5];
}

static void warn(OData* d, int w) { d->warning[w]++; }

/* ------------------------------------------------ position stage [A.1-A.4] */

static void o_kinematics(const OModel* m, OData* d) {
  /* normalise quaternions stored in qpos */
  for (int j = 0; j < m->njnt; j++) {
    if (m->jnt_type[j] == JNT_FREE) normalize4(d->qpos + m->jnt_qposadr[j] + 3);
    else if (m->jnt_type[j] == JNT_BALL) normalize4(d->qpos + m->jnt_qposadr[j]);
  }
  d->xquat[0] = 1;
  d->xmat[0] = d->xmat[4] = d->xmat[8] = 1;
  for (int b = 1; b < m->nbody; b++) {
    int p = m->body_parentid[b];
    double pos[3], quat[4], R[9], t[3];
    mat_vec3(t, d->xmat + 9 * p, d->body_pos + 3 * b);
    for (int k = 0; k < 3; k++) pos[k] = d->xpos[3 * p + k] + t[k];
    quat_mul(quat, d->xquat + 4 * p, d->body_quat + 4 * b);
    for (int k = 0; k < m->body_jntnum[b]; k++) {
      int j = m->body_jntadr[b] + k, qa = m->jnt_qposadr[j];
      int jt = m->jnt_type[j];
      if (jt == JNT_FREE) {
        memcpy(pos, d->qpos + qa, 24);
        memcpy(quat, d->qpos + qa + 3, 32);
        memcpy(d->xanchor + 3 * j, pos, 24);
        d->xaxis[3 * j] = 0; d->xaxis[3 * j + 1] = 0; d->xaxis[3 * j + 2] = 1;
        continue;
      }
      quat2mat(R, quat);
      mat_vec3(t, R, m->jnt_pos + 3 * j);
      double* anchor = d->xanchor + 3 * j;
      for (int c = 0; c < 3; c++) anchor[c] = pos[c] + t[c];
      mat_vec3(d->xaxis + 3 * j, R, m->jnt_axis + 3 * j);
      if (jt == JNT_SLIDE) {
        double dq = d->qpos[qa] - m->qpos0[qa];
        for (int c = 0; c < 3; c++) pos[c] += d->xaxis[3 * j + c] * dq;
      } else {
        double qr[4], qn[4];
        if (jt == JNT_HINGE)
          axis_angle_quat(qr, m->jnt_axis + 3 * j, d->qpos[qa] - m->qpos0[qa]);
        else
          memcpy(qr, d->qpos + qa, 32);
        quat_mul(qn, quat, qr);
        memcpy(quat, qn, 32);
        quat2mat(R, quat);
        mat_vec3(t, R, m->jnt_pos + 3 * j);
        for (int c = 0; c < 3; c++) pos[c] = anchor[c] - t[c];
      }
    }
    normalize4(quat);
    memcpy(d->xpos + 3 * b, pos, 24);
    memcpy(d->xquat + 4 * b, quat, 32);
    quat2mat(d->xmat + 9 * b, quat);
  }
  for (int b = 0; b < m->nbody; b++) {
    double t[3], q[4];
    mat_vec3(t, d->xmat + 9 * b, m->body_ipos + 3 * b);
    for (int c = 0; c < 3; c++) d->xipos[3 * b + c] = d->xpos[3 * b + c] + t[c];
    quat_mul(q, d->xquat + 4 * b, m->body_iquat + 4 * b);
    quat2mat(d->ximat + 9 * b, q);
  }
  for (int g = 0; g < m->ngeom; g++) {
    int b = m->geom_bodyid[g];
    double t[3], q[4];
    mat_vec3(t, d->xmat + 9 * b, d->geom_pos + 3 * g);
    for (int c = 0; c < 3; c++) d->geom_xpos[3 * g + c] = d->xpos[3 * b + c] + t[c];
    quat_mul(q, d->xquat + 4 * b, m->geom_quat + 4 * g);
    quat2mat(d->geom_xmat + 9 * g, q);
  }
  for (int s = 0; s < m->nsite; s++) {
    int b = m->site_bodyid[s];
    double t[3], q[4];
    mat_vec3(t, d->xmat + 9 * b, d->site_pos + 3 * s);
    for (int c = 0; c < 3; c++) d->site_xpos[3 * s + c] = d->xpos[3 * b + c] + t[c];
    quat_mul(q, d->xquat + 4 * b, m->site_quat + 4 * s);
    quat2mat(d->site_xmat + 9 * s, q);
  }
}

static void o_com_pos(const OModel* m, OData* d) {
  int nb = m->nbody;
  for (int b = 0; b < nb; b++)
    for (int c = 0; c < 3; c++)
      d->subtree_com[3 * b + c] = m->body_mass[b] * d->xipos[3 * b + c];
  for (int b = nb - 1; b > 0; b--) {
    int p = m->body_parentid[b];
    for (int c = 0; c < 3; c++) d->subtree_com[3 * p + c] += d->subtree_com[3 * b + c];
  }
  for (int b = 0; b < nb; b++) {
    if (m->body_subtreemass[b] < MINVAL)
      memcpy(d->subtree_com + 3 * b, d->xipos + 3 * b, 24);
    else
      for (int c = 0; c < 3; c++) d->subtree_com[3 * b + c] /= m->body_subtreemass[b];
  }
  /* body inertias about the tree COM, world axes */
  for (int b = 1; b < nb; b++) {
    const double* R = d->ximat + 9 * b;
    const double* I = m->body_inertia + 3 * b;
    const double* com = d->subtree_com + 3 * m->body_rootid[b];
    double mass = m->body_mass[b], dif[3];
    for (int c = 0; c < 3; c++) dif[c] = d->xipos[3 * b + c] - com[c];
    double T[9];
    for (int r = 0; r < 3; r++)
      for (int c = 0; c < 3; c++)
        T[3 * r + c] = R[3 * r] * I[0] * R[3 * c] + R[3 * r + 1] * I[1] * R[3 * c + 1] +
                       R[3 * r + 2] * I[2] * R[3 * c + 2];
    double* ci = d->cinert + 10 * b;
    ci[0] = T[0] + mass * (dif[1] * dif[1] + dif[2] * dif[2]);
    ci[1] = T[4] + mass * (dif[0] * dif[0] + dif[2] * dif[2]);
    ci[2] = T[8] + mass * (dif[0] * dif[0] + dif[1] * dif[1]);
    ci[3] = T[1] - mass * dif[0] * dif[1];
    ci[4] = T[2] - mass * dif[0] * dif[2];
    ci[5] = T[5] - mass * dif[1] * dif[2];
    ci[6] = mass * dif[0]; ci[7] = mass * dif[1]; ci[8] = mass * dif[2];
    ci[9] = mass;
  }
  /* motion axes about the tree COM */
  for (int j = 0; j < m->njnt; j++) {
    int b = m->jnt_bodyid[j], da = m->jnt_dofadr[j];
    const double* com = d->subtree_com + 3 * m->body_rootid[b];
    double off[3];
    for (int c = 0; c < 3; c++) off[c] = com[c] - d->xanchor[3 * j + c];
    int jt = m->jnt_type[j];
    if (jt == JNT_SLIDE) {
      double* cd = d->cdof + 6 * da;
      cd[0] = cd[1] = cd[2] = 0;
      memcpy(cd + 3, d->xaxis + 3 * j, 24);
    } else if (jt == JNT_HINGE) {
      double* cd = d->cdof + 6 * da;
      memcpy(cd, d->xaxis + 3 * j, 24);
      cross3(cd + 3, cd, off);
    } else {
      if (jt == JNT_FREE) {
        for (int k = 0; k < 3; k++) {
          double* cd = d->cdof + 6 * (da + k);
          memset(cd, 0, 48);
          cd[3 + k] = 1;
        }
        da += 3;
      }
      const double* R = d->xmat + 9 * b;
      for (int k = 0; k < 3; k++) {
        double* cd = d->cdof + 6 * (da + k);
        cd[0] = R[k]; cd[1] = R[3 + k]; cd[2] = R[6 + k];
        cross3(cd + 3, cd, off);
      }
    }
  }
}

static void o_crb(const OModel* m, OData* d) {
  int nb = m->nbody, nv = m->nv;
  memcpy(d->crb, d->cinert, sizeof(double) * 10 * nb);
  for (int b = nb - 1; b > 0; b--) {
    int p = m->body_parentid[b];
    if (p > 0)
      for (int c = 0; c < 10; c++) d->crb[10 * p + c] += d->crb[10 * b + c];
  }
  memset(d->qM, 0, sizeof(double) * m->nM);
  for (int i = 0; i < nv; i++) {
    double buf[6];
    inert_mul(buf, d->crb + 10 * m->dof_bodyid[i], d->cdof + 6 * i);
    int adr = m->dof_Madr[i];
    d->qM[adr] = m->dof_armature[i];
    for (int j = i; j >= 0; j = m->dof_parentid[j])
      d->qM[adr++] += dot6(d->cdof + 6 * j, buf);
  }
}

/* in-place sparse L^T D L of a qM-layout matrix [A.4] */
static void factor_sparse(const OModel* m, double* LD, double* diaginv) {
  int nv = m->nv;
  for (int k = nv - 1; k >= 0; k--) {
    int akk = m->dof_Madr[k];
    int aki = akk + 1;
    double dkk = LD[akk];
    for (int i = m->dof_parentid[k]; i >= 0; i = m->dof_parentid[i], aki++) {
      double t = LD[aki] / dkk;
      int aij = m->dof_Madr[i], akj = aki;
      for (int j = i; j >= 0; j = m->dof_parentid[j]) LD[aij++] -= t * LD[akj++];
      LD[aki] = t;
    }
    diaginv[k] = 1.0 / dkk;
  }
}

static void solve_sparse(const OModel* m, const double* LD, const double* diaginv,
                         double* x) {
  int nv = m->nv;
  for (int i = nv - 1; i >= 0; i--) {          /* x <- L^-T x */
    int a = m->dof_Madr[i] + 1;
    for (int j = m->dof_parentid[i]; j >= 0; j = m->dof_parentid[j])
      x[j] -= LD[a++] * x[i];
  }
  for (int i = 0; i < nv; i++) x[i] *= diaginv[i];
  for (int i = 0; i < nv; i++) {               /* x <- L^-1 x */
    int a = m->dof_Madr[i] + 1;
    for (int j = m->dof_parentid[i]; j >= 0; j = m->dof_parentid[j])
      x[i] -= LD[a++] * x[j];
  }
}

static void o_factor_m(const OModel* m, OData* d) {
  memcpy(d->qLD, d->qM, sizeof(double) * m->nM);
  factor_sparse(m, d->qLD, d->qLDiagInv);
}

static void mul_m(const OModel* m, const double* qM, double* res, const double* v) {
  for (int i = 0; i < m->nv; i++) res[i] = 0;
  for (int i = 0; i < m->nv; i++) {
    int a = m->dof_Madr[i];
    res[i] += qM[a] * v[i];
    a++;
    for (int j = m->dof_parentid[i]; j >= 0; j = m->dof_parentid[j], a++) {
      res[i] += qM[a] * v[j];
      res[j] += qM[a] * v[i];
    }
  }
}

/* --------------------------------------------------- velocity stage [A.5] */

static void o_com_vel(const OModel* m, OData* d) {
  memset(d->cvel, 0, 48);
  for (int b = 1; b < m->nbody; b++) {
    double cv[6];
    memcpy(cv, d->cvel + 6 * m->body_parentid[b], 48);
    for (int k = 0; k < m->body_jntnum[b]; k++) {
      int j = m->body_jntadr[b] + k, da = m->jnt_dofadr[j], jt = m->jnt_type[j];
      if (jt == JNT_FREE) {
        for (int i = 0; i < 3; i++) {
          memset(d->cdof_dot + 6 * (da + i), 0, 48);
          for (int c = 0; c < 6; c++) cv[c] += d->cdof[6 * (da + i) + c] * d->qvel[da + i];
        }
        da += 3;
      }
      int n = (jt == JNT_FREE || jt == JNT_BALL) ? 3 : 1;
      for (int i = 0; i < n; i++)
        cross_motion(d->cdof_dot + 6 * (da + i), cv, d->cdof + 6 * (da + i));
      for (int i = 0; i < n; i++)
        for (int c = 0; c < 6; c++) cv[c] += d->cdof[6 * (da + i) + c] * d->qvel[da + i];
    }
    memcpy(d->cvel + 6 * b, cv, 48);
  }
}

static void o_passive(const OModel* m, OData* d) {
  memset(d->qfrc_passive, 0, sizeof(double) * m->nv);
  if (m->opt_disableflags & DSBL_PASSIVE) return;
  for (int j = 0; j < m->njnt; j++) {
    double k = m->jnt_stiffness[j];
    if (k == 0) continue;
    int jt = m->jnt_type[j], qa = m->jnt_qposadr[j], da = m->jnt_dofadr[j];
    if (jt == JNT_SLIDE || jt == JNT_HINGE)
      d->qfrc_passive[da] -= k * (d->qpos[qa] - m->qpos_spring[qa]);
  }
  for (int i = 0; i < m->nv; i++) d->qfrc_passive[i] -= m->dof_damping[i] * d->qvel[i];
  /* tendon spring-dampers [MJ-doc: mj_passive]: the spring's rest length is the tendon's
   * length at qpos0 */
  for (int t = 0; t < m->ntendon; t++) {
    double k = m->tendon_stiffness[t], b = m->tendon_damping[t];
    if (k == 0 && b == 0) continue;
    double frc = -k * (d->ten_length[t] - m->tendon_length0[t]) - b * d->ten_velocity[t];
    for (int w = m->tendon_adr[t]; w < m->tendon_adr[t] + m->tendon_num[t]; w++)
      d->qfrc_passive[m->jnt_dofadr[m->wrap_objid[w]]] += d->wrap_prm[w] * frc;
  }
  /* fluid forces [MJ-doc: inertia-box model of mj_passive]: every body is replaced by the
   * box with its mass and inertia; viscous force and torque of the equivalent sphere, plus
   * quadratic drag per box face (density).  Wind is not served.  The wrench acts at the
   * body's centre of mass; here it is moved to the subtree centre the cdofs refer to. */
  if (m->opt_density > 0 || m->opt_viscosity > 0) {
    double* wr = d->s_fluid;
    for (int b = 1; b < m->nbody; b++) {
      double* w = wr + 6 * b;
      memset(w, 0, 48);
      double mass = m->body_mass[b];
      if (mass < MINVAL) continue;
      const double* I = m->body_inertia + 3 * b;
      double box[3] = {sqrt(fmax(MINVAL, I[1] + I[2] - I[0]) / mass * 6.0),
                       sqrt(fmax(MINVAL, I[0] + I[2] - I[1]) / mass * 6.0),
                       sqrt(fmax(MINVAL, I[0] + I[1] - I[2]) / mass * 6.0)};
      const double* com = d->subtree_com + 3 * m->body_rootid[b];
      const double *cv = d->cvel + 6 * b, *R = d->ximat + 9 * b;
      double dif[3] = {d->xipos[3 * b] - com[0], d->xipos[3 * b + 1] - com[1],
                       d->xipos[3 * b + 2] - com[2]};
      double t3[3], vlin[3], lw[3], lv[3], fa[3] = {0, 0, 0}, fl[3] = {0, 0, 0};
      cross3(t3, cv, dif);
      for (int k = 0; k < 3; k++) vlin[k] = cv[3 + k] + t3[k];
      matT_vec3(lw, R, cv);
      matT_vec3(lv, R, vlin);
      if (m->opt_viscosity > 0) {
        double diam = (box[0] + box[1] + box[2]) / 3.0;
        for (int k = 0; k < 3; k++) {
          fa[k] = -PI * diam * diam * diam * m->opt_viscosity * lw[k];
          fl[k] = -3.0 * PI * diam * m->opt_viscosity * lv[k];
        }
      }
      if (m->opt_density > 0) {
        double rho = m->opt_density;
        fl[0] -= 0.5 * rho * box[1] * box[2] * fabs(lv[0]) * lv[0];
        fl[1] -= 0.5 * rho * box[0] * box[2] * fabs(lv[1]) * lv[1];
        fl[2] -= 0.5 * rho * box[0] * box[1] * fabs(lv[2]) * lv[2];
        double b0 = box[0] * box[0], b1 = box[1] * box[1], b2 = box[2] * box[2];
        fa[0] -= rho * box[0] * (b1 * b1 + b2 * b2) * fabs(lw[0]) * lw[0] / 64.0;
        fa[1] -= rho * box[1] * (b0 * b0 + b2 * b2) * fabs(lw[1]) * lw[1] / 64.0;
        fa[2] -= rho * box[2] * (b0 * b0 + b1 * b1) * fabs(lw[2]) * lw[2] / 64.0;
      }
      double ga[3], gl[3];
      mat_vec3(ga, R, fa);
      mat_vec3(gl, R, fl);
      cross3(t3, dif, gl);
      for (int k = 0; k < 3; k++) { w[k] = ga[k] + t3[k]; w[3 + k] = gl[k]; }
    }
    for (int i = 0; i < m->nv; i++) {
      double acc = 0;
      for (int b = 1; b < m->nbody; b++) {
        int anc = 0;
        for (int p = b; p > 0; p = m->body_parentid[p])
          if (p == m->dof_bodyid[i]) { anc = 1; break; }
        if (anc) acc += dot6(d->cdof + 6 * i, wr + 6 * b);
      }
      d->qfrc_passive[i] += acc;
    }
  }
}

/* recursive Newton-Euler; flg_acc adds cdof*qacc.  result = inverse-dynamics
 * joint forces */
static void o_rne(const OModel* m, OData* d, int flg_acc, double* result) {
  int nb = m->nbody;
  double* cacc = (double*)calloc(6 * nb, 8);
  double* cfrc = (double*)calloc(6 * nb, 8);
  if (!(m->opt_disableflags & DSBL_GRAVITY))
    for (int c = 0; c < 3; c++) cacc[3 + c] = -m->opt_gravity[c];
  for (int b = 1; b < nb; b++) {
    double* a = cacc + 6 * b;
    memcpy(a, cacc + 6 * m->body_parentid[b], 48);
    int da = m->body_dofadr[b];
    for (int k = 0; k < m->body_dofnum[b]; k++) {
      for (int c = 0; c < 6; c++) {
        a[c] += d->cdof_dot[6 * (da + k) + c] * d->qvel[da + k];
        if (flg_acc) a[c] += d->cdof[6 * (da + k) + c] * d->qacc[da + k];
      }
    }
    double Ia[6], Iv[6], vxIv[6];
    inert_mul(Ia, d->cinert + 10 * b, a);
    inert_mul(Iv, d->cinert + 10 * b, d->cvel + 6 * b);
    cross_force(vxIv, d->cvel + 6 * b, Iv);
    for (int c = 0; c < 6; c++) cfrc[6 * b + c] = Ia[c] + vxIv[c];
  }
  for (int b = nb - 1; b > 0; b--) {
    int p = m->body_parentid[b];
    for (int c = 0; c < 6; c++) cfrc[6 * p + c] += cfrc[6 * b + c];
  }
  for (int i = 0; i < m->nv; i++)
    result[i] = dot6(d->cdof + 6 * i, cfrc + 6 * m->dof_bodyid[i]);
  free(cacc); free(cfrc);
}

/* ------------------------------------------------------ actuation [A.6] */

static void o_jac(const OModel* m, const OData* d, double* jp, double* jr,
                  const double* point, int body);

/* mj_tendon.  Fixed tendons: length = sum coef * qpos[joint], the Jacobian is the constant
 * coef row.  Spatial tendons (straight segments between sites, no wrapping geoms or
 * pulleys): length = sum |p1 - p0|, Jacobian = sum (jacp(p1) - jacp(p0))^T dir, kept in
 * ten_J (one dense row per spatial tendon). */
static void o_tendon(const OModel* m, OData* d) {
  int nv = m->nv;
  for (int t = 0; t < m->ntendon; t++) {
    double len = 0;
    int w0 = m->tendon_adr[t], w1 = w0 + m->tendon_num[t], ja = m->tendon_Jadr[t];
    if (ja < 0) {
      for (int w = w0; w < w1; w++)
        len += d->wrap_prm[w] * d->qpos[m->jnt_qposadr[m->wrap_objid[w]]];
    } else {
      double* J = d->ten_J + (size_t)ja * nv;
      double* jp0 = (double*)malloc(sizeof(double) * 6 * (nv + 1));
      double* jp1 = jp0 + 3 * nv;
      memset(J, 0, sizeof(double) * nv);
      for (int w = w0; w + 1 < w1; w++) {
        int s0 = m->wrap_objid[w], s1 = m->wrap_objid[w + 1];
        const double *p0 = d->site_xpos + 3 * s0, *p1 = d->site_xpos + 3 * s1;
        double dif[3] = {p1[0] - p0[0], p1[1] - p0[1], p1[2] - p0[2]};
        double dist = sqrt(dif[0] * dif[0] + dif[1] * dif[1] + dif[2] * dif[2]);
        if (dist < MINVAL) { dif[0] = 1; dif[1] = 0; dif[2] = 0; }
        else for (int k = 0; k < 3; k++) dif[k] /= dist;
        len += dist;
        o_jac(m, d, jp0, NULL, p0, m->site_bodyid[s0]);
        o_jac(m, d, jp1, NULL, p1, m->site_bodyid[s1]);
        for (int i = 0; i < nv; i++)
          for (int k = 0; k < 3; k++) J[i] += (jp1[k * nv + i] - jp0[k * nv + i]) * dif[k];
      }
      free(jp0);
    }
    d->ten_length[t] = len;
  }
}

static void o_transmission(const OModel* m, OData* d) {
  int nv = m->nv;
  memset(d->actuator_moment, 0, sizeof(double) * m->nu * nv);
  for (int i = 0; i < m->nu; i++) {
    int j = m->actuator_trnid[2 * i];
    double g = m->actuator_gear[6 * i];
    if (m->actuator_trntype[i] == 3) {          /* tendon transmission */
      d->actuator_length[i] = g * d->ten_length[j];
      for (int w = m->tendon_adr[j]; w < m->tendon_adr[j] + m->tendon_num[j]; w++)
        d->actuator_moment[i * nv + m->jnt_dofadr[m->wrap_objid[w]]] += g * d->wrap_prm[w];
      continue;
    }
    /* joint transmission, scalar joints only in this subset */
    d->actuator_length[i] = g * d->qpos[m->jnt_qposadr[j]];
    d->actuator_moment[i * nv + m->jnt_dofadr[j]] = g;
  }
}

static void o_fwd_actuation(const OModel* m, OData* d) {
  int nv = m->nv, nu = m->nu;
  memset(d->qfrc_actuator, 0, sizeof(double) * nv);
  memset(d->actuator_force, 0, sizeof(double) * nu);
  for (int i = 0; i < m->na; i++) d->act_dot[i] = 0;
  if (nu == 0 || (m->opt_disableflags & DSBL_ACTUATION)) return;
  for (int i = 0; i < nu; i++)
    if (isnan(d->ctrl[i]) || fabs(d->ctrl[i]) > MAXVAL) {
      warn(d, W_BADCTRL);
      memset(d->ctrl, 0, sizeof(double) * nu);
      break;
    }
  int ia = 0;
  for (int i = 0; i < nu; i++) {
    double c = d->ctrl[i];
    if (m->actuator_ctrllimited[i] && !(m->opt_disableflags & DSBL_CLAMPCTRL)) {
      double lo = m->actuator_ctrlrange[2 * i], hi = m->actuator_ctrlrange[2 * i + 1];
      c = c < lo ? lo : (c > hi ? hi : c);
    }
    double in = c;
    int dyn = m->actuator_dyntype[i];
    if (dyn != 0) {
      double a = d->act[ia];
      if (dyn == 1) d->act_dot[ia] = c;                             /* integrator */
      else d->act_dot[ia] = (c - a) / fmax(MINVAL, m->actuator_dynprm[3 * i]);
      in = a;
      ia++;
    }
    double f = m->actuator_gainprm[3 * i] * in;
    if (m->actuator_biastype[i] == 1)
      f += m->actuator_biasprm[3 * i] +
           m->actuator_biasprm[3 * i + 1] * d->actuator_length[i] +
           m->actuator_biasprm[3 * i + 2] * d->actuator_velocity[i];
    if (m->actuator_forcelimited[i]) {
      double lo = m->actuator_forcerange[2 * i], hi = m->actuator_forcerange[2 * i + 1];
      f = f < lo ? lo : (f > hi ? hi : f);
    }
    d->actuator_force[i] = f;
    for (int k = 0; k < nv; k++) d->qfrc_actuator[k] += d->actuator_moment[i * nv + k] * f;
  }
}

static void o_fwd_acceleration(const OModel* m, OData* d) {
  for (int i = 0; i < m->nv; i++) {
    d->qfrc_smooth[i] = d->qfrc_passive[i] - d->qfrc_bias[i] + d->qfrc_actuator[i] +
                        d->qfrc_applied[i];
    d->qacc_smooth[i] = d->qfrc_smooth[i];
  }
  solve_sparse(m, d->qLD, d->qLDiagInv, d->qacc_smooth);
}

/* ------------------------------------------------------- collision [A.7] */

typedef struct { double dist, pos[3], frame[9]; } RawCon;

static int col_plane_sphere_raw(const double* ppos, const double* pn,
                                const double* c, double r, double margin,
                                RawCon* out) {
  double dif[3] = {c[0] - ppos[0], c[1] - ppos[1], c[2] - ppos[2]};
  double dist = dot3(dif, pn) - r;
  if (dist > margin) return 0;
  out->dist = dist;
  for (int k = 0; k < 3; k++) {
    out->pos[k] = c[k] - pn[k] * (r + 0.5 * dist);
    out->frame[k] = pn[k];
  }
  memset(out->frame + 3, 0, 48);
  return 1;
}

static int col_sphere_sphere_raw(const double* c1, double r1, const double* c2,
                                 double r2, double margin, RawCon* out) {
  double n[3] = {c2[0] - c1[0], c2[1] - c1[1], c2[2] - c1[2]};
  double len = norm3(n);
  double dist = len - r1 - r2;
  if (dist > margin) return 0;
  if (len < MINVAL) { n[0] = 1; n[1] = 0; n[2] = 0; }
  else { n[0] /= len; n[1] /= len; n[2] /= len; }
  out->dist = dist;
  for (int k = 0; k < 3; k++) {
    out->pos[k] = c1[k] + n[k] * (r1 + 0.5 * dist);
    out->frame[k] = n[k];
  }
  memset(out->frame + 3, 0, 48);
  return 1;
}

static void geom_axis(const double* xmat, int col, double* ax) {
  ax[0] = xmat[col]; ax[1] = xmat[3 + col]; ax[2] = xmat[6 + col];
}

/* plane - cylinder [MJ-doc: analytic, up to four points]: the deepest rim point of the
 * disk facing the plane, the matching point of the other disk, and two more points of
 * the first disk at +-120 degrees when that disk lies nearly flat */
static int col_plane_cylinder_raw(const double* ppos, const double* n, const double* cpos,
                                  const double* cmat, const double* size, double margin,
                                  RawCon* out) {
  double axis[3], vec[3];
  geom_axis(cmat, 2, axis);
  double prjaxis = dot3(n, axis);
  if (prjaxis > 0) { for (int k = 0; k < 3; k++) axis[k] = -axis[k]; prjaxis = -prjaxis; }
  double dif[3] = {cpos[0] - ppos[0], cpos[1] - ppos[1], cpos[2] - ppos[2]};
  double dist0 = dot3(dif, n);
  for (int k = 0; k < 3; k++) vec[k] = axis[k] * prjaxis - n[k];
  double len = norm3(vec);
  if (len >= MINVAL) for (int k = 0; k < 3; k++) vec[k] *= size[0] / len;
  else { geom_axis(cmat, 0, vec); for (int k = 0; k < 3; k++) vec[k] *= size[0]; }
  double prjvec = dot3(vec, n);
  for (int k = 0; k < 3; k++) axis[k] *= size[1];
  prjaxis *= size[1];
  if (dist0 + prjaxis + prjvec > margin) return 0;
  int cnt = 0;
  double dd = dist0 + prjaxis + prjvec;
  out[cnt].dist = dd;
  for (int k = 0; k < 3; k++) {
    out[cnt].pos[k] = cpos[k] + vec[k] + axis[k] - n[k] * dd * 0.5;
    out[cnt].frame[k] = n[k];
  }
  memset(out[cnt].frame + 3, 0, 48);
  cnt++;
  dd = dist0 - prjaxis + prjvec;
  if (dd <= margin) {
    out[cnt].dist = dd;
    for (int k = 0; k < 3; k++) {
      out[cnt].pos[k] = cpos[k] + vec[k] - axis[k] - n[k] * dd * 0.5;
      out[cnt].frame[k] = n[k];
    }
    memset(out[cnt].frame + 3, 0, 48);
    cnt++;
  }
  double prjvec1 = -prjvec * 0.5;
  dd = dist0 + prjaxis + prjvec1;
  if (dd <= margin) {
    double vec1[3];
    cross3(vec1, vec, axis);
    normalize3(vec1);
    for (int k = 0; k < 3; k++) vec1[k] *= size[0] * sqrt(3.0) * 0.5;
    for (int sgn = 1; sgn >= -1; sgn -= 2) {
      out[cnt].dist = dd;
      for (int k = 0; k < 3; k++) {
        out[cnt].pos[k] = cpos[k] + sgn * vec1[k] + axis[k] - vec[k] * 0.5 - n[k] * dd * 0.5;
        out[cnt].frame[k] = n[k];
      }
      memset(out[cnt].frame + 3, 0, 48);
      cnt++;
    }
  }
  return cnt;
}

/* closest point of a solid cylinder (centre p, unit axis a, radius r, half length h) to q;
 * returns the distance (0 inside) */
static double cyl_closest(const double* p, const double* a, double r, double h,
                          const double* q, double* cl) {
  double dq[3] = {q[0] - p[0], q[1] - p[1], q[2] - p[2]};
  double x = dot3(dq, a);
  double w[3] = {dq[0] - x * a[0], dq[1] - x * a[1], dq[2] - x * a[2]};
  double rho = norm3(w);
  double xc = x < -h ? -h : (x > h ? h : x);
  double sc = rho > r ? r / rho : 1.0;
  for (int k = 0; k < 3; k++) cl[k] = p[k] + xc * a[k] + sc * w[k];
  double e[3] = {q[0] - cl[0], q[1] - cl[1], q[2] - cl[2]};
  return norm3(e);
}

/* sphere (centre q, radius rs; geom1) against a cylinder (geom2).  Outside: along the
 * closest-point direction.  Centre inside the cylinder: pushed out through the nearer of
 * the lateral surface and the caps (MuJoCo resolves deep convex overlaps with MPR; the
 * suite models never reach this branch) */
static int col_sphere_cylinder_raw(const double* q, double rs, const double* p,
                                   const double* a, double r, double h, double margin,
                                   RawCon* out) {
  double cl[3], nrm[3], dist;
  double dd = cyl_closest(p, a, r, h, q, cl);
  if (dd > MINVAL) {
    for (int k = 0; k < 3; k++) nrm[k] = (cl[k] - q[k]) / dd;
    dist = dd - rs;
  } else {
    double dq[3] = {q[0] - p[0], q[1] - p[1], q[2] - p[2]};
    double x = dot3(dq, a);
    double w[3] = {dq[0] - x * a[0], dq[1] - x * a[1], dq[2] - x * a[2]};
    double rho = norm3(w);
    if (r - rho < h - fabs(x) && rho > MINVAL) {
      for (int k = 0; k < 3; k++) nrm[k] = -w[k] / rho;
      dist = -(r - rho) - rs;
    } else {
      double sg = x >= 0 ? 1.0 : -1.0;
      for (int k = 0; k < 3; k++) nrm[k] = -sg * a[k];
      dist = -(h - fabs(x)) - rs;
    }
  }
  if (dist > margin) return 0;
  out->dist = dist;
  for (int k = 0; k < 3; k++) {
    out->pos[k] = q[k] + nrm[k] * (rs + 0.5 * dist);
    out->frame[k] = nrm[k];
  }
  memset(out->frame + 3, 0, 48);
  return 1;
}

/* capsule (segment pc + t ac, |t| <= l, radius rc; geom1) against a cylinder: the
 * distance from the segment point to the cylinder is convex in t, so the closest point
 * is found by bisection on the sign of its derivative, then handled as a sphere */
static int col_capsule_cylinder_raw(const double* pc, const double* ac, double rc, double l,
                                    const double* p, const double* a, double r, double h,
                                    double margin, RawCon* out) {
  double lo = -l, hi = l, q[3], cl[3];
  for (int it = 0; it < 64; it++) {
    double t = 0.5 * (lo + hi);
    for (int k = 0; k < 3; k++) q[k] = pc[k] + t * ac[k];
    cyl_closest(p, a, r, h, q, cl);
    double g = (q[0] - cl[0]) * ac[0] + (q[1] - cl[1]) * ac[1] + (q[2] - cl[2]) * ac[2];
    if (g >= 0) hi = t; else lo = t;
  }
  double t = 0.5 * (lo + hi);
  for (int k = 0; k < 3; k++) q[k] = pc[k] + t * ac[k];
  int n = col_sphere_cylinder_raw(q, rc, p, a, r, h, margin, out);
  if (n) memcpy(out->frame + 3, ac, 24);
  return n;
}

/* ---- ellipsoids.  MuJoCo sends these pairs through its general convex routine (MPR,
 * tolerance 1e-6); here the closest point is computed directly.  Point q in the
 * ellipsoid's frame, semi-axes a: the closest surface point is x_i = a_i^2 q_i / (t + a_i^2)
 * with t >= 0 the root of F(t) = sum (a_i q_i / (t + a_i^2))^2 - 1 (convex, decreasing:
 * Newton from the left never overshoots).  *t carries the multiplier in and out so that
 * neighbouring queries start close.  Returns the distance, 0 if q is inside. */
static double ell_closest(const double* a, const double* q, double* x, double* t) {
  double f0 = 0;
  for (int k = 0; k < 3; k++) f0 += (q[k] / a[k]) * (q[k] / a[k]);
  if (f0 <= 1) { for (int k = 0; k < 3; k++) x[k] = q[k]; *t = 0; return 0; }
  double tt = *t > 0 ? *t : 0;
  for (int it = 0; it < 100; it++) {
    double F = -1, dF = 0;
    for (int k = 0; k < 3; k++) {
      double den = tt + a[k] * a[k], r = a[k] * q[k] / den;
      F += r * r; dF -= 2 * r * r / den;
    }
    if (F < 0 && it == 0 && tt > 0) { tt = 0; continue; }   /* warm start overshot: restart */
    double step = -F / dF;
    tt += step;
    if (fabs(step) <= 1e-16 * (tt + a[0] * a[0])) break;
  }
  double dd = 0;
  for (int k = 0; k < 3; k++) {
    x[k] = a[k] * a[k] * q[k] / (tt + a[k] * a[k]);
    dd += (q[k] - x[k]) * (q[k] - x[k]);
  }
  *t = tt;
  return sqrt(dd);
}

/* sphere (world centre qw, radius rs; geom1) against an ellipsoid (geom2: centre p, frame
 * R row-major, semi-axes a) */
static int col_sphere_ellipsoid_raw(const double* qw, double rs, const double* p,
                                    const double* R, const double* a, double margin,
                                    double* tmul, RawCon* out) {
  double dq[3] = {qw[0] - p[0], qw[1] - p[1], qw[2] - p[2]}, q[3], x[3], nl[3], dist;
  matT_vec3(q, R, dq);
  double dd = ell_closest(a, q, x, tmul);
  if (dd > MINVAL) {
    for (int k = 0; k < 3; k++) nl[k] = (x[k] - q[k]) / dd;       /* sphere -> ellipsoid */
    dist = dd - rs;
  } else {
    /* centre inside: pushed out along the surface normal of the radially scaled point */
    double s = 0;
    for (int k = 0; k < 3; k++) s += (q[k] / a[k]) * (q[k] / a[k]);
    s = sqrt(s);
    double g[3], gl = 0, pen = 0;
    for (int k = 0; k < 3; k++) {
      double xs = s > MINVAL ? q[k] / s : (k == 2 ? a[2] : 0);
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
  out->dist = dist;
  for (int k = 0; k < 3; k++) {
    out->pos[k] = qw[k] + nw[k] * (rs + 0.5 * dist);
    out->frame[k] = nw[k];
  }
  memset(out->frame + 3, 0, 48);
  return 1;
}

/* capsule (geom1) against an ellipsoid: the distance from the segment point pc + t ac to
 * the ellipsoid is convex in t; its derivative g(t) = n(t) . ac is driven to zero by the
 * Illinois variant of regula falsi on [-l, l], then the point is handled as a sphere */
static double cap_ell_slope(const double* pc, const double* ac, double t, const double* p,
                            const double* R, const double* a, double* tmul) {
  double dq[3] = {pc[0] + t * ac[0] - p[0], pc[1] + t * ac[1] - p[1], pc[2] + t * ac[2] - p[2]};
  double q[3], x[3], al[3];
  matT_vec3(q, R, dq);
  matT_vec3(al, R, ac);
  double dd = ell_closest(a, q, x, tmul);
  if (dd <= MINVAL) return 0;
  return ((q[0] - x[0]) * al[0] + (q[1] - x[1]) * al[1] + (q[2] - x[2]) * al[2]) / dd;
}

static int col_capsule_ellipsoid_raw(const double* pc, const double* ac, double rc, double l,
                                     const double* p, const double* R, const double* a,
                                     double margin, RawCon* out) {
  double tmul = 0, lo = -l, hi = l;
  double glo = cap_ell_slope(pc, ac, lo, p, R, a, &tmul);
  double ghi = cap_ell_slope(pc, ac, hi, p, R, a, &tmul);
  double t;
  if (glo >= 0) t = lo;
  else if (ghi <= 0) t = hi;
  else {
    int side = 0;
    t = 0.5 * (lo + hi);
    for (int it = 0; it < 100; it++) {
      t = (lo * ghi - hi * glo) / (ghi - glo);
      double g = cap_ell_slope(pc, ac, t, p, R, a, &tmul);
      if (fabs(g) < 1e-15 || hi - lo < 1e-15 * l) break;
      if (g > 0) { hi = t; ghi = g; if (side == 1) glo *= 0.5; side = 1; }
      else { lo = t; glo = g; if (side == -1) ghi *= 0.5; side = -1; }
    }
  }
  double q[3] = {pc[0] + t * ac[0], pc[1] + t * ac[1], pc[2] + t * ac[2]};
  int n = col_sphere_ellipsoid_raw(q, rc, p, R, a, margin, &tmul, out);
  if (n) memcpy(out->frame + 3, ac, 24);
  return n;
}

/* plane against an ellipsoid: the support point of the ellipsoid along -n */
static int col_plane_ellipsoid_raw(const double* ppos, const double* n, const double* p,
                                   const double* R, const double* a, double margin,
                                   RawCon* out) {
  double dl[3], x[3], xw[3], s = 0;
  matT_vec3(dl, R, n);
  for (int k = 0; k < 3; k++) s += a[k] * a[k] * dl[k] * dl[k];
  s = sqrt(s);
  for (int k = 0; k < 3; k++) x[k] = -a[k] * a[k] * dl[k] / s;
  mat_vec3(xw, R, x);
  double dif[3] = {p[0] + xw[0] - ppos[0], p[1] + xw[1] - ppos[1], p[2] + xw[2] - ppos[2]};
  double dist = dot3(dif, n);
  if (dist > margin) return 0;
  out->dist = dist;
  for (int k = 0; k < 3; k++) {
    out->pos[k] = p[k] + xw[k] - n[k] * dist * 0.5;
    out->frame[k] = n[k];
  }
  memset(out->frame + 3, 0, 48);
  return 1;
}

/* ------------------------------------------------ height fields [A.7h]
 * MuJoCo walks the grid cells under the geom's bounding sphere, builds a triangular prism
 * per half cell and sends prism-geom to its general convex routine (libccd MPR, closed).
 * Here the same cell walk feeds closed-form tests against the prism's top triangle; the
 * ground below the triangle counts as solid.  All of it in the height field's frame.
 *   cell (r, c), half 0: a = v(r,c), b = v(r,c+1),   c = v(r+1,c+1)
 *                half 1: a = v(r,c), b = v(r+1,c+1), c = v(r+1,c)        (both CCW from above)
 * Features shared by neighbouring triangles are tested by ONE owner so that a contact is
 * not reported twice: half 0 owns vertex a and edges ab, ca; half 1 owns edge ca.
 *   sphere    closest point of the triangle (face always, edge / vertex when owned)
 *   capsule   both end spheres + the segment's interior against the owned edges
 *   ellipsoid, cylinder   support point(s) along -n against the face + the owned vertex
 *             against the solid (edges of the terrain are not tested: documented gap)    */
typedef struct {
  const double* data;
  int nrow, ncol;
  double rx, ry, zs, base, dx, dy;
} HField;

static void hf_init(const OModel* m, const OData* d, int hid, HField* h) {
  h->data = d->hfield + m->hfield_adr[hid];
  h->nrow = m->hfield_nrow[hid]; h->ncol = m->hfield_ncol[hid];
  h->rx = m->hfield_size[4 * hid]; h->ry = m->hfield_size[4 * hid + 1];
  h->zs = m->hfield_size[4 * hid + 2]; h->base = m->hfield_size[4 * hid + 3];
  h->dx = 2 * h->rx / (h->ncol - 1); h->dy = 2 * h->ry / (h->nrow - 1);
}

static void hf_vertex(const HField* h, int r, int c, double* v) {
  v[0] = c * h->dx - h->rx; v[1] = r * h->dy - h->ry;
  v[2] = h->data[r * h->ncol + c] * h->zs;
}

static void hf_triangle(const HField* h, int r, int c, int half, double* a, double* b,
                        double* cc) {
  hf_vertex(h, r, c, a);
  if (half == 0) { hf_vertex(h, r, c + 1, b); hf_vertex(h, r + 1, c + 1, cc); }
  else { hf_vertex(h, r + 1, c + 1, b); hf_vertex(h, r + 1, c, cc); }
}

static int hf_cell(double u, double rad, double dd, int nn) {
  double f = floor((u + rad) / dd);
  if (f < 0) f = 0;
  if (f > nn - 2) f = nn - 2;
  return (int)f;
}

/* elevation of the surface over (x, y); *tid = id of the triangle that holds the point,
 * (cell index << 1) | half */
static double hf_surface(const HField* h, double x, double y, int* tid) {
  int c = hf_cell(x, h->rx, h->dx, h->ncol), r = hf_cell(y, h->ry, h->dy, h->nrow);
  double fx = (x + h->rx) / h->dx - c, fy = (y + h->ry) / h->dy - r;
  fx = fx < 0 ? 0 : (fx > 1 ? 1 : fx); fy = fy < 0 ? 0 : (fy > 1 ? 1 : fy);
  const double* dt = h->data;
  double z00 = dt[r * h->ncol + c], z10 = dt[r * h->ncol + c + 1],
         z01 = dt[(r + 1) * h->ncol + c], z11 = dt[(r + 1) * h->ncol + c + 1];
  int half = fx >= fy ? 0 : 1;
  double zsurf = half == 0 ? z00 + fx * (z10 - z00) + fy * (z11 - z10)
                           : z00 + fx * (z11 - z01) + fy * (z01 - z00);
  if (tid) *tid = ((r * h->ncol + c) << 1) | half;
  return zsurf * h->zs;
}

/* closest point of triangle abc to p (Voronoi regions of the triangle's features);
 * returns the region: 0 face, 1 / 2 / 3 vertex a / b / c, 4 edge ab, 5 edge bc, 6 edge ca */
static int tri_closest(const double* p, const double* a, const double* b, const double* c,
                       double* q) {
  double ab[3], ac[3], ap[3], bp[3], cp[3];
  for (int k = 0; k < 3; k++) {
    ab[k] = b[k] - a[k]; ac[k] = c[k] - a[k];
    ap[k] = p[k] - a[k]; bp[k] = p[k] - b[k]; cp[k] = p[k] - c[k];
  }
  double d1 = dot3(ab, ap), d2 = dot3(ac, ap);
  if (d1 <= 0 && d2 <= 0) { memcpy(q, a, 24); return 1; }
  double d3 = dot3(ab, bp), d4 = dot3(ac, bp);
  if (d3 >= 0 && d4 <= d3) { memcpy(q, b, 24); return 2; }
  double vc = d1 * d4 - d3 * d2;
  if (vc <= 0 && d1 >= 0 && d3 <= 0) {
    double v = d1 / (d1 - d3);
    for (int k = 0; k < 3; k++) q[k] = a[k] + v * ab[k];
    return 4;
  }
  double d5 = dot3(ab, cp), d6 = dot3(ac, cp);
  if (d6 >= 0 && d5 <= d6) { memcpy(q, c, 24); return 3; }
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

static void raw_set(RawCon* o, double dist, const double* pos, const double* n,
                    const double* hint) {
  o->dist = dist;
  memcpy(o->pos, pos, 24);
  memcpy(o->frame, n, 24);
  memset(o->frame + 3, 0, 48);
  if (hint) memcpy(o->frame + 3, hint, 24);
}

/* sphere (centre q, radius r) against triangle `tid` with unit normal nt.  A centre above
 * the surface touches the closest point of the triangle (face always, edges / vertices
 * when owned); a centre that has sunk below the surface is pushed out through the face of
 * the one triangle that holds it, so no sideways normals from neighbouring edges */
static int hf_sphere_tri(const HField* h, int tid, const double* a, const double* b,
                         const double* c, const double* nt, const double* q, double r,
                         double margin, const double* hint, RawCon* out) {
  double aq[3] = {q[0] - a[0], q[1] - a[1], q[2] - a[2]}, cl[3], nrm[3], dist;
  double s = dot3(nt, aq);
  int holder;
  if (q[2] < hf_surface(h, q[0], q[1], &holder)) {
    if (holder != tid) return 0;
    memcpy(nrm, nt, 24);
    dist = s - r;
  } else {
    if (s < 0) return 0;
    int reg = tri_closest(q, a, b, c, cl);
    if (!(((tid & 1) ? HF_OWN1 : HF_OWN0) >> reg & 1)) return 0;
    double dv[3] = {q[0] - cl[0], q[1] - cl[1], q[2] - cl[2]};
    double dd = norm3(dv);
    if (dd < MINVAL) memcpy(nrm, nt, 24);
    else for (int k = 0; k < 3; k++) nrm[k] = dv[k] / dd;
    dist = dd - r;
  }
  if (dist > margin) return 0;
  double pos[3];
  for (int k = 0; k < 3; k++) pos[k] = q[k] - nrm[k] * (r + 0.5 * dist);
  raw_set(out, dist, pos, nrm, hint);
  return 1;
}

/* interior of the capsule's segment (p +- l ax) against the edge e0 -> e1; a hit at the
 * start vertex counts when vertex0 is set (the edge that owns its first vertex) */
static int hf_capsule_edge(const HField* h, const double* p, const double* ax, double l,
                           double r, const double* e0, const double* e1, int vertex0,
                           double margin, RawCon* out) {
  double d1[3], d2[3], rr[3];
  for (int k = 0; k < 3; k++) {
    d1[k] = 2 * l * ax[k]; d2[k] = e1[k] - e0[k];
    rr[k] = p[k] - l * ax[k] - e0[k];
  }
  double A = dot3(d1, d1), E = dot3(d2, d2), F = dot3(d2, rr), C = dot3(d1, rr),
         B = dot3(d1, d2);
  double den = A * E - B * B, s, t;
  if (den > 1e-12 * A * E) {
    s = (B * F - C * E) / den;
    s = s < 0 ? 0 : (s > 1 ? 1 : s);
  } else s = 0.5;
  t = (B * s + F) / E;
  if (t < 0) { t = 0; s = -C / A; s = s < 0 ? 0 : (s > 1 ? 1 : s); }
  else if (t > 1) { t = 1; s = (B - C) / A; s = s < 0 ? 0 : (s > 1 ? 1 : s); }
  if (!(s > 0 && s < 1)) return 0;                 /* the ends are tested as spheres */
  if (!(t < 1 && (t > 0 || vertex0))) return 0;
  double c1[3], c2[3], dv[3];
  for (int k = 0; k < 3; k++) {
    c1[k] = p[k] - l * ax[k] + s * d1[k]; c2[k] = e0[k] + t * d2[k];
    dv[k] = c1[k] - c2[k];
  }
  double dd = norm3(dv);
  if (dd < MINVAL || dv[2] <= 0) return 0;         /* the axis has sunk below the edge */
  if (c1[2] < hf_surface(h, c1[0], c1[1], NULL)) return 0;       /* ... or the surface */
  double dist = dd - r;
  if (dist > margin) return 0;
  double nrm[3], pos[3];
  for (int k = 0; k < 3; k++) { nrm[k] = dv[k] / dd; pos[k] = c1[k] - nrm[k] * (r + 0.5 * dist); }
  raw_set(out, dist, pos, nrm, ax);
  return 1;
}

/* a surface point x of a convex geom against the face: inside the prism's footprint and
 * below (or within margin of) the face */
static int hf_point_face(const double* a, const double* b, const double* c, const double* nt,
                         const double* x, double margin, RawCon* out) {
  double cl[3];
  if (tri_closest(x, a, b, c, cl) != 0) return 0;
  double ax[3] = {x[0] - a[0], x[1] - a[1], x[2] - a[2]};
  double dist = dot3(nt, ax);
  if (dist > margin) return 0;
  double pos[3];
  for (int k = 0; k < 3; k++) pos[k] = x[k] - nt[k] * 0.5 * dist;
  raw_set(out, dist, pos, nt, NULL);
  return 1;
}

/* geom (type gt, centre gp, frame gR row-major, sizes gs; all in the field's frame) against
 * one triangle; returns the number of contacts written (at most 4), normals pointing
 * from the terrain to the geom */
static int hf_tri_geom(const HField* h, int tid, const double* a, const double* b,
                       const double* c, int gt, const double* gp, const double* gR,
                       const double* gs, double margin, RawCon* out) {
  const int half = tid & 1;
  double ab[3], ac[3], nt[3];
  for (int k = 0; k < 3; k++) { ab[k] = b[k] - a[k]; ac[k] = c[k] - a[k]; }
  cross3(nt, ab, ac);
  normalize3(nt);
  int n = 0;
  if (gt == GEOM_SPHERE)
    return hf_sphere_tri(h, tid, a, b, c, nt, gp, gs[0], margin, NULL, out);
  if (gt == GEOM_CAPSULE) {
    double ax[3], e[3];
    geom_axis(gR, 2, ax);
    for (int sgn = 1; sgn >= -1; sgn -= 2) {
      for (int k = 0; k < 3; k++) e[k] = gp[k] + sgn * ax[k] * gs[1];
      n += hf_sphere_tri(h, tid, a, b, c, nt, e, gs[0], margin, ax, out + n);
    }
    if (half == 0) n += hf_capsule_edge(h, gp, ax, gs[1], gs[0], a, b, 1, margin, out + n);
    n += hf_capsule_edge(h, gp, ax, gs[1], gs[0], c, a, 0, margin, out + n);
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
    n += hf_point_face(a, b, c, nt, xw, margin, out + n);
    if (half == 0) {
      double tmul = 0;
      RawCon tmp;
      if (col_sphere_ellipsoid_raw(a, 0.0, gp, gR, gs, margin, &tmul, &tmp)) out[n++] = tmp;
    }
    return n;
  }
  if (gt == GEOM_CYLINDER) {
    double axis[3], vec[3], x[3];
    geom_axis(gR, 2, axis);
    double prj = dot3(nt, axis);
    for (int k = 0; k < 3; k++) vec[k] = axis[k] * prj - nt[k];
    double len = norm3(vec);
    if (len >= MINVAL) for (int k = 0; k < 3; k++) vec[k] *= gs[0] / len;
    else { geom_axis(gR, 0, vec); for (int k = 0; k < 3; k++) vec[k] *= gs[0]; }
    for (int sgn = 1; sgn >= -1; sgn -= 2) {
      for (int k = 0; k < 3; k++) x[k] = gp[k] + vec[k] + sgn * axis[k] * gs[1];
      n += hf_point_face(a, b, c, nt, x, margin, out + n);
    }
    if (half == 0) {
      RawCon tmp;
      if (col_sphere_cylinder_raw(a, 0.0, gp, axis, gs[0], gs[1], margin, &tmp)) out[n++] = tmp;
    }
    return n;
  }
  return 0;
}

/* cells of the grid under [lo, hi] along one axis (half extent rad, spacing dd, nn
 * vertices): first cell and one past the last */
static void hf_range(double lo, double hi, double rad, double dd, int nn, int* i0, int* i1) {
  double f0 = floor((lo + rad) / dd), f1 = ceil((hi + rad) / dd);
  if (f0 < 0) f0 = 0;
  if (f1 > nn - 1) f1 = nn - 1;
  *i0 = (int)f0; *i1 = (int)f1;
}

/* ---- box narrowphase: sphere / capsule / box against a box ---------------------------
 * [MJ-doc] MuJoCo has dedicated routines for these pairs (mjc_SphereBox, mjc_CapsuleBox,
 * mjc_BoxBox: "up to 8 contacts for box-box"); their source is not in the reference, so the
 * geometry is restated here from first principles and the CUDA path is held to THIS
 * restatement (DESIGN.md section 2 lists it among the departures):
 *   sphere - box   closest point of the solid box; a centre inside the box leaves through
 *                  the nearest face
 *   capsule - box  the squared distance from a point of the capsule's segment to the box is
 *                  convex along the segment: the two ends of its set of minimisers (found by
 *                  bisection on the sign of the derivative) and the two segment ends are the
 *                  candidate sphere centres; the deepest hit and the hit farthest from it
 *                  along the segment are kept (<= 2 contacts, ascending along the axis)
 *   box - box      separating-axis test over the 15 axes, faces preferred over edge pairs;
 *                  face case: the incident face clipped against the reference face
 *                  (<= 8 contacts, in clipping order); edge case: closest points of the two
 *                  edges (1 contact)
 * Normals point from geom 1 to geom 2. */
static int col_sphere_box_raw(const double* c, double r, const double* pb, const double* Rb,
                              const double* h, double margin, RawCon* out) {
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
    nl[0] = nl[1] = nl[2] = 0;
    nl[best] = d[best] >= 0 ? -1.0 : 1.0;
    dist = -depth - r;
  }
  if (dist > margin) return 0;
  for (int k = 0; k < 3; k++) pl[k] = d[k] + nl[k] * (r + 0.5 * dist);
  out->dist = dist;
  mat_vec3(out->pos, Rb, pl);
  for (int k = 0; k < 3; k++) out->pos[k] += pb[k];
  mat_vec3(out->frame, Rb, nl);
  memset(out->frame + 3, 0, 48);
  return 1;
}

/* derivative (up to a factor 2) and value of the squared distance from c0 + t u to the box */
static double capbox_dg(const double* c0, const double* u, const double* h, double t) {
  double s = 0;
  for (int k = 0; k < 3; k++) {
    double v = c0[k] + t * u[k];
    double e = v > h[k] ? v - h[k] : (v < -h[k] ? v + h[k] : 0.0);
    s += e * u[k];
  }
  return s;
}
static double capbox_g(const double* c0, const double* u, const double* h, double t) {
  double s = 0;
  for (int k = 0; k < 3; k++) {
    double v = c0[k] + t * u[k];
    double e = v > h[k] ? v - h[k] : (v < -h[k] ? v + h[k] : 0.0);
    s += e * e;
  }
  return s;
}

static int col_capsule_box_raw(const double* p1, const double* a1, double r, double l,
                               const double* pb, const double* Rb, const double* h,
                               double margin, RawCon* out) {
  double dif[3] = {p1[0] - pb[0], p1[1] - pb[1], p1[2] - pb[2]}, c0[3], u[3];
  matT_vec3(c0, Rb, dif);
  matT_vec3(u, Rb, a1);
  double tlo, thi;
  if (capbox_dg(c0, u, h, -l) >= 0) tlo = -l;
  else if (capbox_dg(c0, u, h, l) < 0) tlo = l;
  else {
    double lo = -l, hi = l;
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
  for (int i = 0; i < nc; i++) {
    double q[3] = {p1[0] + cand[i] * a1[0], p1[1] + cand[i] * a1[1], p1[2] + cand[i] * a1[2]};
    if (col_sphere_box_raw(q, r, pb, Rb, h, margin, &hit[nh])) {
      memcpy(hit[nh].frame + 3, a1, 24);       /* tangent hint: capsule axis */
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

/* clips the polygon (x, y, z)[n] against s * coord <= lim (coord = x or y); returns n' */
static int boxbox_clip(const double* xi, const double* yi, const double* zi, int n, int usey,
                       double sgn, double lim, double* xo, double* yo, double* zo) {
  int m = 0;
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

static int col_box_box_raw(const double* pa, const double* Ra, const double* ha,
                           const double* pb, const double* Rb, const double* hb,
                           double margin, RawCon* out) {
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
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 3; j++) {
      int i1 = (i + 1) % 3, i2 = (i + 2) % 3, j1 = (j + 1) % 3, j2 = (j + 2) % 3;
      double len2 = 1.0 - R[3 * i + j] * R[3 * i + j];
      if (len2 < 1e-12) continue;               /* parallel edges: covered by the faces */
      double len = sqrt(len2);
      double tl = fabs(t[i2] * R[3 * i1 + j] - t[i1] * R[3 * i2 + j]);
      double ra = ha[i1] * Q[3 * i2 + j] + ha[i2] * Q[3 * i1 + j];
      double rb = hb[j1] * Q[3 * i + j2] + hb[j2] * Q[3 * i + j1];
      double s = (tl - ra - rb) / len;
      if (s > margin) return 0;
      if (s > best + 0.05 * fabs(best) + 1e-10) { best = s; code = 6 + 3 * i + j; }
    }
  if (code < 6) {
    const int refA = code < 3, ai = refA ? code : code - 3;
    const double *pr = refA ? pa : pb, *Rr = refA ? Ra : Rb, *hr = refA ? ha : hb;
    const double *pi = refA ? pb : pa, *Ri = refA ? Rb : Ra, *hi = refA ? hb : ha;
    double ax[3], nref[3], dc[3] = {pi[0] - pr[0], pi[1] - pr[1], pi[2] - pr[2]};
    geom_axis(Rr, ai, ax);
    double sg = dot3(dc, ax) > 0 ? 1.0 : -1.0;
    for (int k = 0; k < 3; k++) nref[k] = sg * ax[k];
    int kf = 0;
    double bestdot = -1;
    for (int k = 0; k < 3; k++) {
      double ak[3];
      geom_axis(Ri, k, ak);
      double dk = fabs(dot3(nref, ak));
      if (dk > bestdot + 1e-10) { bestdot = dk; kf = k; }
    }
    double af[3], a1[3], a2[3], fc[3];
    const int k1 = (kf + 1) % 3, k2 = (kf + 2) % 3, r1 = (ai + 1) % 3, r2 = (ai + 2) % 3;
    geom_axis(Ri, kf, af); geom_axis(Ri, k1, a1); geom_axis(Ri, k2, a2);
    double sf = dot3(nref, af) > 0 ? -1.0 : 1.0;
    for (int k = 0; k < 3; k++) fc[k] = pi[k] + sf * hi[kf] * af[k] - pr[k];
    double t1[3], t2[3];
    geom_axis(Rr, r1, t1); geom_axis(Rr, r2, t2);
    static const double S1[4] = {1, -1, -1, 1}, S2[4] = {1, 1, -1, -1};
    double x[2][9], y[2][9], z[2][9];
    for (int v = 0; v < 4; v++) {
      double w[3];
      for (int k = 0; k < 3; k++) w[k] = fc[k] + S1[v] * hi[k1] * a1[k] + S2[v] * hi[k2] * a2[k];
      x[0][v] = dot3(w, t1); y[0][v] = dot3(w, t2); z[0][v] = dot3(w, nref) - hr[ai];
    }
    int n = 4;
    n = boxbox_clip(x[0], y[0], z[0], n, 0, 1.0, hr[r1], x[1], y[1], z[1]);
    n = boxbox_clip(x[1], y[1], z[1], n, 0, -1.0, hr[r1], x[0], y[0], z[0]);
    n = boxbox_clip(x[0], y[0], z[0], n, 1, 1.0, hr[r2], x[1], y[1], z[1]);
    n = boxbox_clip(x[1], y[1], z[1], n, 1, -1.0, hr[r2], x[0], y[0], z[0]);
    int cnt = 0;
    for (int v = 0; v < n && cnt < 8; v++) {
      if (z[0][v] > margin) continue;
      out[cnt].dist = z[0][v];
      for (int k = 0; k < 3; k++) {
        out[cnt].pos[k] = pr[k] + x[0][v] * t1[k] + y[0][v] * t2[k] +
                          (hr[ai] + 0.5 * z[0][v]) * nref[k];
        out[cnt].frame[k] = refA ? nref[k] : -nref[k];
      }
      memset(out[cnt].frame + 3, 0, 48);
      cnt++;
    }
    return cnt;
  }
  /* edge (axis i of A) against edge (axis j of B) */
  const int i = (code - 6) / 3, j = (code - 6) % 3;
  double ua[3], ub[3], L[3];
  geom_axis(Ra, i, ua); geom_axis(Rb, j, ub);
  cross3(L, ua, ub);
  normalize3(L);
  if (dot3(L, dw) < 0) { L[0] = -L[0]; L[1] = -L[1]; L[2] = -L[2]; }
  double ea[3] = {pa[0], pa[1], pa[2]}, eb[3] = {pb[0], pb[1], pb[2]};
  for (int k = 0; k < 3; k++) {
    double ak[3], bk[3];
    if (k != i) {
      geom_axis(Ra, k, ak);
      double sgk = dot3(L, ak) > 0 ? 1.0 : -1.0;
      for (int q = 0; q < 3; q++) ea[q] += sgk * ha[k] * ak[q];
    }
    if (k != j) {
      geom_axis(Rb, k, bk);
      double sgk = dot3(L, bk) > 0 ? -1.0 : 1.0;
      for (int q = 0; q < 3; q++) eb[q] += sgk * hb[k] * bk[q];
    }
  }
  double d[3] = {eb[0] - ea[0], eb[1] - ea[1], eb[2] - ea[2]};
  double uaub = dot3(ua, ub), q1 = dot3(ua, d), q2 = -dot3(ub, d), den = 1.0 - uaub * uaub;
  double al = (q1 + uaub * q2) / den, be = (uaub * q1 + q2) / den;
  al = al < -ha[i] ? -ha[i] : (al > ha[i] ? ha[i] : al);
  be = be < -hb[j] ? -hb[j] : (be > hb[j] ? hb[j] : be);
  double ca[3], cb[3], sepv[3];
  for (int k = 0; k < 3; k++) {
    ca[k] = ea[k] + al * ua[k]; cb[k] = eb[k] + be * ub[k]; sepv[k] = cb[k] - ca[k];
  }
  double dist = dot3(sepv, L);
  if (dist > margin) return 0;
  out->dist = dist;
  for (int k = 0; k < 3; k++) { out->pos[k] = 0.5 * (ca[k] + cb[k]); out->frame[k] = L[k]; }
  memset(out->frame + 3, 0, 48);
  return 1;
}

static int collide_pair(const OModel* m, const OData* d, int g1, int g2,
                        double margin, RawCon* out) {
  int t1 = m->geom_type[g1], t2 = m->geom_type[g2];
  const double *p1 = d->geom_xpos + 3 * g1, *p2 = d->geom_xpos + 3 * g2;
  const double *R1 = d->geom_xmat + 9 * g1, *R2 = d->geom_xmat + 9 * g2;
  const double *s1 = m->geom_size + 3 * g1, *s2 = m->geom_size + 3 * g2;
  if (t1 == GEOM_PLANE) {
    double n[3];
    geom_axis(R1, 2, n);
    if (t2 == GEOM_SPHERE) return col_plane_sphere_raw(p1, n, p2, s2[0], margin, out);
    if (t2 == GEOM_CAPSULE) {
      double ax[3], e[3];
      geom_axis(R2, 2, ax);
      int cnt = 0;
      for (int sgn = 1; sgn >= -1; sgn -= 2) {
        for (int k = 0; k < 3; k++) e[k] = p2[k] + sgn * ax[k] * s2[1];
        if (col_plane_sphere_raw(p1, n, e, s2[0], margin, out + cnt)) {
          memcpy(out[cnt].frame + 3, ax, 24);   /* tangent hint: capsule axis */
          cnt++;
        }
      }
      return cnt;
    }
    if (t2 == GEOM_ELLIPSOID) return col_plane_ellipsoid_raw(p1, n, p2, R2, s2, margin, out);
    if (t2 == GEOM_CYLINDER) return col_plane_cylinder_raw(p1, n, p2, R2, s2, margin, out);
    if (t2 == GEOM_BOX) {
      double dif[3] = {p2[0] - p1[0], p2[1] - p1[1], p2[2] - p1[2]};
      double dist = dot3(dif, n);
      int cnt = 0;
      for (int i = 0; i < 8 && cnt < 4; i++) {
        double v[3] = {(i & 1) ? s2[0] : -s2[0], (i & 2) ? s2[1] : -s2[1],
                       (i & 4) ? s2[2] : -s2[2]};
        double c[3];
        mat_vec3(c, R2, v);
        double ld = dot3(n, c);
        if (dist + ld > margin) continue;
        out[cnt].dist = dist + ld;
        for (int k = 0; k < 3; k++) {
          out[cnt].pos[k] = p2[k] + c[k] - n[k] * 0.5 * (dist + ld);
          out[cnt].frame[k] = n[k];
        }
        memset(out[cnt].frame + 3, 0, 48);
        cnt++;
      }
      return cnt;
    }
    return 0;
  }
  if (t1 == GEOM_SPHERE && t2 == GEOM_SPHERE)
    return col_sphere_sphere_raw(p1, s1[0], p2, s2[0], margin, out);
  if (t1 == GEOM_SPHERE && t2 == GEOM_CAPSULE) {
    double ax[3], dif[3] = {p1[0] - p2[0], p1[1] - p2[1], p1[2] - p2[2]};
    geom_axis(R2, 2, ax);
    double x = dot3(ax, dif);
    x = x < -s2[1] ? -s2[1] : (x > s2[1] ? s2[1] : x);
    double c[3] = {p2[0] + ax[0] * x, p2[1] + ax[1] * x, p2[2] + ax[2] * x};
    return col_sphere_sphere_raw(p1, s1[0], c, s2[0], margin, out);
  }
  if (t2 == GEOM_ELLIPSOID && t1 == GEOM_SPHERE) {
    double tmul = 0;
    return col_sphere_ellipsoid_raw(p1, s1[0], p2, R2, s2, margin, &tmul, out);
  }
  if (t2 == GEOM_ELLIPSOID && t1 == GEOM_CAPSULE) {
    double a1[3];
    geom_axis(R1, 2, a1);
    return col_capsule_ellipsoid_raw(p1, a1, s1[0], s1[1], p2, R2, s2, margin, out);
  }
  if (t2 == GEOM_CYLINDER && (t1 == GEOM_SPHERE || t1 == GEOM_CAPSULE || t1 == GEOM_CYLINDER)) {
    double a2[3], a1[3];
    geom_axis(R2, 2, a2);
    if (t1 == GEOM_SPHERE)
      return col_sphere_cylinder_raw(p1, s1[0], p2, a2, s2[0], s2[1], margin, out);
    /* cylinder - cylinder: geom1 is replaced by its enclosing capsule, which is exact
     * whenever the two are apart (all the suite models need) and approximate in contact */
    geom_axis(R1, 2, a1);
    return col_capsule_cylinder_raw(p1, a1, s1[0], s1[1], p2, a2, s2[0], s2[1], margin, out);
  }
  if (t2 == GEOM_BOX) {
    if (t1 == GEOM_SPHERE) return col_sphere_box_raw(p1, s1[0], p2, R2, s2, margin, out);
    if (t1 == GEOM_CAPSULE) {
      double a1[3];
      geom_axis(R1, 2, a1);
      return col_capsule_box_raw(p1, a1, s1[0], s1[1], p2, R2, s2, margin, out);
    }
    if (t1 == GEOM_BOX) return col_box_box_raw(p1, R1, s1, p2, R2, s2, margin, out);
    return 0;
  }
  if (t1 == GEOM_CAPSULE && t2 == GEOM_CAPSULE) {
    /* closest points of two segments p1 + x1*a1, p2 + x2*a2, |xi| <= li */
    double a1[3], a2[3], dif[3] = {p1[0] - p2[0], p1[1] - p2[1], p1[2] - p2[2]};
    geom_axis(R1, 2, a1); geom_axis(R2, 2, a2);
    double l1 = s1[1], l2 = s2[1];
    double b = dot3(a1, a2), u = -dot3(a1, dif), v = dot3(a2, dif);
    double det = 1.0 - b * b;
    double x1, x2;
    if (fabs(det) >= 1e-12) {
      x1 = (u + b * v) / det;
      x2 = (v + b * u) / det;
      if (x1 > l1) { x1 = l1; x2 = v + b * x1; }
      else if (x1 < -l1) { x1 = -l1; x2 = v + b * x1; }
      if (x2 > l2) {
        x2 = l2; x1 = u + b * x2;
        x1 = x1 < -l1 ? -l1 : (x1 > l1 ? l1 : x1);
      } else if (x2 < -l2) {
        x2 = -l2; x1 = u + b * x2;
        x1 = x1 < -l1 ? -l1 : (x1 > l1 ? l1 : x1);
      }
    } else {
      /* parallel axes: centre of the overlap interval */
      x1 = u; x1 = x1 < -l1 ? -l1 : (x1 > l1 ? l1 : x1);
      x2 = v + b * x1; x2 = x2 < -l2 ? -l2 : (x2 > l2 ? l2 : x2);
      x1 = u + b * x2; x1 = x1 < -l1 ? -l1 : (x1 > l1 ? l1 : x1);
    }
    double c1[3], c2[3];
    for (int k = 0; k < 3; k++) { c1[k] = p1[k] + a1[k] * x1; c2[k] = p2[k] + a2[k] * x2; }
    return col_sphere_sphere_raw(c1, s1[0], c2, s2[0], margin, out);
  }
  return 0;
}

static void make_frame(double* f) {
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

/* appends one contact; returns 0 when the buffer is full */
static int add_contact(const OModel* m, OData* d, int g1, int g2, double margin, double gap,
                       const RawCon* raw) {
  if (d->ncon[0] >= m->nconmax) { warn(d, W_CONTACTFULL); return 0; }
  OContact* c = d->contact + d->ncon[0];
  int id = d->ncon[0]++;
  c->dist = raw->dist;
  memcpy(c->pos, raw->pos, 24);
  memcpy(c->frame, raw->frame, 72);
  make_frame(c->frame);
  c->includemargin = margin - gap;
  c->geom1 = g1; c->geom2 = g2;
  /* [MJ-doc: contact parameters] equal priority: max condim, max friction, solmix-weighted
   * solref / solimp; otherwise everything from the geom with the higher priority */
  int dim1 = m->geom_condim[g1], dim2 = m->geom_condim[g2];
  int pr1 = m->geom_priority[g1], pr2 = m->geom_priority[g2];
  c->dim = pr1 > pr2 ? dim1 : pr2 > pr1 ? dim2 : (dim1 > dim2 ? dim1 : dim2);
  const double *f1 = m->geom_friction + 3 * g1, *f2 = m->geom_friction + 3 * g2;
  double fr[3];
  for (int k = 0; k < 3; k++) fr[k] = pr1 > pr2 ? f1[k] : pr2 > pr1 ? f2[k] : fmax(f1[k], f2[k]);
  c->friction[0] = c->friction[1] = fr[0];
  c->friction[2] = fr[1];
  c->friction[3] = c->friction[4] = fr[2];
  double mix1 = m->geom_solmix[g1], mix2 = m->geom_solmix[g2], mix;
  if (mix1 >= MINVAL && mix2 >= MINVAL) mix = mix1 / (mix1 + mix2);
  else if (mix1 < MINVAL && mix2 < MINVAL) mix = 0.5;
  else mix = mix1 < MINVAL ? 0.0 : 1.0;
  if (pr1 != pr2) mix = pr1 > pr2 ? 1.0 : 0.0;
  for (int k = 0; k < 2; k++)
    c->solref[k] = mix * m->geom_solref[2 * g1 + k] + (1 - mix) * m->geom_solref[2 * g2 + k];
  for (int k = 0; k < 5; k++)
    c->solimp[k] = mix * m->geom_solimp[5 * g1 + k] + (1 - mix) * m->geom_solimp[5 * g2 + k];
  c->exclude = c->dist >= c->includemargin;
  c->efc_address = -1;
  d->contact_geom1[id] = g1; d->contact_geom2[id] = g2;
  d->contact_dim[id] = c->dim; d->contact_dist[id] = c->dist;
  d->contact_efc_address[id] = -1;
  memcpy(d->contact_pos + 3 * id, c->pos, 24);
  memcpy(d->contact_frame + 9 * id, c->frame, 72);
  return 1;
}

/* height field g1 against geom g2: the cells under the geom's bounding sphere, row by row,
 * two triangles each; contacts in that order, rotated back to the world frame */
static int collide_hfield(const OModel* m, OData* d, int g1, int g2, double margin,
                          double gap) {
  HField h;
  hf_init(m, d, m->geom_dataid[g1], &h);
  const double *p1 = d->geom_xpos + 3 * g1, *R1 = d->geom_xmat + 9 * g1;
  const double *p2 = d->geom_xpos + 3 * g2, *R2 = d->geom_xmat + 9 * g2;
  double dif[3] = {p2[0] - p1[0], p2[1] - p1[1], p2[2] - p1[2]}, gp[3], gR[9];
  matT_vec3(gp, R1, dif);
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 3; j++)
      gR[3 * i + j] = R1[i] * R2[j] + R1[3 + i] * R2[3 + j] + R1[6 + i] * R2[6 + j];
  double rb = m->geom_rbound[g2] + margin;
  if (gp[0] - rb > h.rx || gp[0] + rb < -h.rx || gp[1] - rb > h.ry || gp[1] + rb < -h.ry ||
      gp[2] - rb > h.zs || gp[2] + rb < -h.base) return 1;
  int c0, c1, r0, r1;
  hf_range(gp[0] - rb, gp[0] + rb, h.rx, h.dx, h.ncol, &c0, &c1);
  hf_range(gp[1] - rb, gp[1] + rb, h.ry, h.dy, h.nrow, &r0, &r1);
  for (int r = r0; r < r1; r++)
    for (int c = c0; c < c1; c++)
      for (int half = 0; half < 2; half++) {
        double a[3], b[3], cc[3];
        hf_triangle(&h, r, c, half, a, b, cc);
        if (fmax(a[2], fmax(b[2], cc[2])) < gp[2] - rb) continue;
        RawCon raw[4];
        int n = hf_tri_geom(&h, ((r * h.ncol + c) << 1) | half, a, b, cc, m->geom_type[g2],
                            gp, gR, m->geom_size + 3 * g2, margin, raw);
        for (int i = 0; i < n; i++) {
          RawCon w = raw[i];
          mat_vec3(w.pos, R1, raw[i].pos);
          for (int k = 0; k < 3; k++) w.pos[k] += p1[k];
          mat_vec3(w.frame, R1, raw[i].frame);
          mat_vec3(w.frame + 3, R1, raw[i].frame + 3);
          if (!add_contact(m, d, g1, g2, margin, gap, &w)) return 0;
        }
      }
  return 1;
}

static void o_collision(const OModel* m, OData* d) {
  d->ncon[0] = 0;
  if (m->opt_disableflags & (DSBL_CONTACT | DSBL_CONSTRAINT)) return;
  int filterparent = !(m->opt_disableflags & DSBL_FILTERPARENT);
  for (int p = 0; p < m->npair; p++) {
    int g1 = m->pair_geom1[p], g2 = m->pair_geom2[p];
    if (filterparent && m->pair_parentchild[p]) continue;
    double margin = fmax(m->geom_margin[g1], m->geom_margin[g2]);
    double gap = fmax(m->geom_gap[g1], m->geom_gap[g2]);
    if (m->geom_type[g1] == GEOM_HFIELD) {      /* the last nhfpair pairs of the list */
      if (!collide_hfield(m, d, g1, g2, margin, gap)) return;
      continue;
    }
    /* bounding-sphere rejection */
    const double *p1 = d->geom_xpos + 3 * g1, *p2 = d->geom_xpos + 3 * g2;
    if (m->geom_type[g1] == GEOM_PLANE) {
      double n[3], dif[3] = {p2[0] - p1[0], p2[1] - p1[1], p2[2] - p1[2]};
      geom_axis(d->geom_xmat + 9 * g1, 2, n);
      if (dot3(dif, n) > m->geom_rbound[g2] + margin) continue;
    } else {
      double dif[3] = {p2[0] - p1[0], p2[1] - p1[1], p2[2] - p1[2]};
      double bound = m->geom_rbound[g1] + m->geom_rbound[g2] + margin;
      if (dot3(dif, dif) > bound * bound) continue;
    }
    RawCon raw[8];
    int n = collide_pair(m, d, g1, g2, margin, raw);
    for (int i = 0; i < n; i++)
      if (!add_contact(m, d, g1, g2, margin, gap, raw + i)) return;
  }
}

/* --------------------------------------------- constraint rows [A.8] */

/* translational / rotational Jacobian (3 x nv each) of a world point on a body */
static void o_jac(const OModel* m, const OData* d, double* jp, double* jr,
                  const double* point, int body) {
  int nv = m->nv;
  if (jp) memset(jp, 0, sizeof(double) * 3 * nv);
  if (jr) memset(jr, 0, sizeof(double) * 3 * nv);
  while (body > 0 && m->body_dofnum[body] == 0) body = m->body_parentid[body];
  if (body == 0) return;
  const double* com = d->subtree_com + 3 * m->body_rootid[body];
  double off[3] = {point[0] - com[0], point[1] - com[1], point[2] - com[2]};
  for (int i = m->body_dofadr[body] + m->body_dofnum[body] - 1; i >= 0;
       i = m->dof_parentid[i]) {
    const double* cd = d->cdof + 6 * i;
    if (jr) { jr[i] = cd[0]; jr[nv + i] = cd[1]; jr[2 * nv + i] = cd[2]; }
    if (jp) {
      double t[3];
      cross3(t, cd, off);
      jp[i] = cd[3] + t[0]; jp[nv + i] = cd[4] + t[1]; jp[2 * nv + i] = cd[5] + t[2];
    }
  }
}

static int add_row(const OModel* m, OData* d, int type, int id, double pos,
                   double margin, double floss) {
  if (d->nefc[0] >= m->njmax) { warn(d, W_CNSTRFULL); return -1; }
  int r = d->nefc[0]++;
  memset(d->efc_J + (size_t)r * m->nv, 0, sizeof(double) * m->nv);
  d->efc_type[r] = type; d->efc_id[r] = id; d->efc_pos[r] = pos;
  d->efc_margin[r] = margin; d->efc_frictionloss[r] = floss;
  return r;
}

static void impedance(const double* si, double pos, double margin, double* imp,
                      double* impP) {
  if (si[0] == si[1] || si[2] <= MINVAL) { *imp = 0.5 * (si[0] + si[1]); *impP = 0; return; }
  double x = (pos - margin) / si[2], sgn = 1;
  if (x < 0) { x = -x; sgn = -1; }
  if (x >= 1 || x <= 0) { *imp = x >= 1 ? si[1] : si[0]; *impP = 0; return; }
  double y, yP, p = si[4], mid = si[3];
  if (p == 1) { y = x; yP = 1; }
  else if (x <= mid) {
    double a = 1 / pow(mid, p - 1);
    y = a * pow(x, p); yP = p * a * pow(x, p - 1);
  } else {
    double b = 1 / pow(1 - mid, p - 1);
    y = 1 - b * pow(1 - x, p); yP = p * b * pow(1 - x, p - 1);
  }
  *imp = si[0] + y * (si[1] - si[0]);
  *impP = yP * sgn * (si[1] - si[0]) / si[2];
}

static void o_make_constraint(const OModel* m, OData* d) {
  int nv = m->nv;
  d->nefc[0] = d->ne[0] = d->nf[0] = 0;
  if (m->opt_disableflags & DSBL_CONSTRAINT) return;
  /* equalities between joints or (fixed) tendons:
   * (x1 - x1_0) - poly(x2 - x2_0) = 0, poly(0) = data[0]; without a second object the
   * residual is x1 - x1_0 - data[0] */
  if (!(m->opt_disableflags & DSBL_EQUALITY))
    for (int e = 0; e < m->neq; e++) {
      if (!m->eq_active[e]) continue;
      int o1 = m->eq_obj1id[e], o2 = m->eq_obj2id[e], tend = m->eq_type[e] == 3;
      const double* c = m->eq_data + 5 * e;
      double x1 = tend ? d->ten_length[o1] - m->tendon_length0[o1]
                       : d->qpos[m->jnt_qposadr[o1]] - m->qpos0[m->jnt_qposadr[o1]];
      double pos, deriv = 0;
      if (o2 >= 0) {
        double x2 = tend ? d->ten_length[o2] - m->tendon_length0[o2]
                         : d->qpos[m->jnt_qposadr[o2]] - m->qpos0[m->jnt_qposadr[o2]];
        pos = x1 - (c[0] + x2 * (c[1] + x2 * (c[2] + x2 * (c[3] + x2 * c[4]))));
        deriv = c[1] + x2 * (2 * c[2] + x2 * (3 * c[3] + x2 * 4 * c[4]));
      } else pos = x1 - c[0];
      int r = add_row(m, d, CN_EQUALITY, e, pos, 0, 0);
      if (r < 0) return;
      double* J = d->efc_J + (size_t)r * nv;
      if (tend) {
        for (int w = m->tendon_adr[o1]; w < m->tendon_adr[o1] + m->tendon_num[o1]; w++)
          J[m->jnt_dofadr[m->wrap_objid[w]]] += d->wrap_prm[w];
        if (o2 >= 0)
          for (int w = m->tendon_adr[o2]; w < m->tendon_adr[o2] + m->tendon_num[o2]; w++)
            J[m->jnt_dofadr[m->wrap_objid[w]]] -= deriv * d->wrap_prm[w];
      } else {
        J[m->jnt_dofadr[o1]] += 1;
        if (o2 >= 0) J[m->jnt_dofadr[o2]] -= deriv;
      }
      d->ne[0]++;
    }
  /* friction loss */
  if (!(m->opt_disableflags & DSBL_FRICTIONLOSS))
    for (int i = 0; i < nv; i++)
      if (m->dof_frictionloss[i] > 0) {
        int r = add_row(m, d, CN_FRICTION_DOF, i, 0, 0, m->dof_frictionloss[i]);
        if (r < 0) return;
        d->efc_J[(size_t)r * nv + i] = 1;
        d->nf[0]++;
      }
  /* joint limits (scalar joints) */
  if (!(m->opt_disableflags & DSBL_LIMIT))
    for (int j = 0; j < m->njnt; j++) {
      if (!m->jnt_limited[j]) continue;
      int jt = m->jnt_type[j];
      if (jt != JNT_SLIDE && jt != JNT_HINGE) continue;
      double q = d->qpos[m->jnt_qposadr[j]], margin = m->jnt_margin[j];
      for (int side = -1; side <= 1; side += 2) {
        double dist = side * (m->jnt_range[2 * j + (side + 1) / 2] - q);
        if (dist < margin) {
          int r = add_row(m, d, CN_LIMIT_JOINT, j, dist, margin, 0);
          if (r < 0) return;
          d->efc_J[(size_t)r * nv + m->jnt_dofadr[j]] = -side;
        }
      }
    }
  /* tendon limits: the row is -side * (tendon Jacobian) */
  if (!(m->opt_disableflags & DSBL_LIMIT))
    for (int t = 0; t < m->ntendon; t++) {
      if (!m->tendon_limited[t]) continue;
      double len = d->ten_length[t], margin = m->tendon_margin[t];
      for (int side = -1; side <= 1; side += 2) {
        double dist = side * (m->tendon_range[2 * t + (side + 1) / 2] - len);
        if (dist < margin) {
          int r = add_row(m, d, CN_LIMIT_TENDON, t, dist, margin, 0);
          if (r < 0) return;
          double* row = d->efc_J + (size_t)r * nv;
          if (m->tendon_Jadr[t] >= 0)
            for (int i = 0; i < nv; i++) row[i] = -side * d->ten_J[(size_t)m->tendon_Jadr[t] * nv + i];
          else
            for (int w = m->tendon_adr[t]; w < m->tendon_adr[t] + m->tendon_num[t]; w++)
              row[m->jnt_dofadr[m->wrap_objid[w]]] += -side * d->wrap_prm[w];
        }
      }
    }
  /* contacts */
  double* jp1 = malloc(sizeof(double) * 12 * (nv + 1));
  double *jr1 = jp1 + 3 * nv, *jp2 = jr1 + 3 * nv, *jr2 = jp2 + 3 * nv;
  for (int ci = 0; ci < d->ncon[0]; ci++) {
    OContact* c = d->contact + ci;
    c->efc_address = -1;
    d->contact_efc_address[ci] = -1;
    if (c->exclude) continue;
    int b1 = m->geom_bodyid[c->geom1], b2 = m->geom_bodyid[c->geom2];
    o_jac(m, d, jp1, jr1, c->pos, b1);
    o_jac(m, d, jp2, jr2, c->pos, b2);
    /* contact-frame Jacobian rows: 0..2 translational, 3..5 rotational */
    double Jc[6][64 * 2];
    for (int k = 0; k < 3; k++)
      for (int i = 0; i < nv; i++) {
        double tp = 0, tr = 0;
        for (int a = 0; a < 3; a++) {
          tp += c->frame[3 * k + a] * (jp2[a * nv + i] - jp1[a * nv + i]);
          tr += c->frame[3 * k + a] * (jr2[a * nv + i] - jr1[a * nv + i]);
        }
        Jc[k][i] = tp; Jc[3 + k][i] = tr;
      }
    int dim = c->dim;
    /* a contact gets all of its rows or none [MJ-doc: mj_addConstraint reserves the block] */
    int need = dim == 1 ? 1 : (m->opt_cone == 0 ? 2 * (dim - 1) : dim);
    if (d->nefc[0] + need > m->njmax) { warn(d, W_CNSTRFULL); break; }
    if (dim == 1) {
      int r = add_row(m, d, CN_CONTACT_FRICTIONLESS, ci, c->dist, c->includemargin, 0);
      if (r < 0) break;
      memcpy(d->efc_J + (size_t)r * nv, Jc[0], sizeof(double) * nv);
      c->efc_address = r;
    } else if (m->opt_cone == 0) {
      for (int k = 1; k < dim; k++)
        for (int sgn = 1; sgn >= -1; sgn -= 2) {
          int r = add_row(m, d, CN_CONTACT_PYRAMIDAL, ci, c->dist, c->includemargin, 0);
          if (r < 0) goto done;
          if (c->efc_address < 0) c->efc_address = r;
          double mu = c->friction[k - 1];
          for (int i = 0; i < nv; i++)
            d->efc_J[(size_t)r * nv + i] = Jc[0][i] + sgn * mu * Jc[k][i];
        }
    } else {
      for (int k = 0; k < dim; k++) {
        int r = add_row(m, d, CN_CONTACT_ELLIPTIC, ci, c->dist, c->includemargin, 0);
        if (r < 0) goto done;
        if (c->efc_address < 0) c->efc_address = r;
        memcpy(d->efc_J + (size_t)r * nv, Jc[k], sizeof(double) * nv);
      }
    }
    d->contact_efc_address[ci] = c->efc_address;
  }
done:
  free(jp1);
  /* per-row solver parameters: diagApprox, impedance, R, D, KBIP, aref */
  int nefc = d->nefc[0];
  double dt = m->opt_timestep;
  for (int r = 0; r < nefc; r++) {
    int type = d->efc_type[r], id = d->efc_id[r];
    const double *solref, *solimp;
    double dA;
    if (type == CN_EQUALITY) {
      solref = m->eq_solref + 2 * id; solimp = m->eq_solimp + 5 * id;
      int o1 = m->eq_obj1id[id], o2 = m->eq_obj2id[id];
      if (m->eq_type[id] == 3)
        dA = m->tendon_invweight0[o1] + (o2 >= 0 ? m->tendon_invweight0[o2] : 0);
      else
        dA = m->dof_invweight0[m->jnt_dofadr[o1]] +
             (o2 >= 0 ? m->dof_invweight0[m->jnt_dofadr[o2]] : 0);
    } else if (type == CN_FRICTION_DOF) {
      solref = m->dof_solref + 2 * id; solimp = m->dof_solimp + 5 * id;
      dA = m->dof_invweight0[id];
    } else if (type == CN_LIMIT_JOINT) {
      solref = m->jnt_solref + 2 * id; solimp = m->jnt_solimp + 5 * id;
      dA = m->dof_invweight0[m->jnt_dofadr[id]];
    } else if (type == CN_LIMIT_TENDON) {
      solref = m->tendon_solref_lim + 2 * id; solimp = m->tendon_solimp_lim + 5 * id;
      dA = m->tendon_invweight0[id];
    } else {
      OContact* c = d->contact + id;
      solref = c->solref; solimp = c->solimp;
      int b1 = m->geom_bodyid[c->geom1], b2 = m->geom_bodyid[c->geom2];
      double tran = m->body_invweight0[2 * b1] + m->body_invweight0[2 * b2];
      double rot = m->body_invweight0[2 * b1 + 1] + m->body_invweight0[2 * b2 + 1];
      int k = r - c->efc_address;
      if (type == CN_CONTACT_FRICTIONLESS) dA = tran;
      else if (type == CN_CONTACT_PYRAMIDAL) {
        double fri = c->friction[k / 2];
        dA = tran + fri * fri * (k < 4 ? tran : rot);
      } else dA = k < 3 ? tran : rot;
    }
    d->efc_diagApprox[r] = dA;
    double sr0 = solref[0], sr1 = solref[1];
    if (!(m->opt_disableflags & DSBL_REFSAFE) && sr0 > 0 && sr0 < 2 * dt) sr0 = 2 * dt;
    double imp, impP;
    impedance(solimp, d->efc_pos[r], d->efc_margin[r], &imp, &impP);
    if (imp < 0.0001) imp = 0.0001;
    if (imp > 0.9999) imp = 0.9999;
    double R = fmax(MINVAL, (1 - imp) * dA / imp);
    double dmax = solimp[1] < 0.0001 ? 0.0001 : (solimp[1] > 0.9999 ? 0.9999 : solimp[1]);
    double K, B;
    if (type == CN_FRICTION_DOF) { K = 0; B = 0; }
    else if (sr0 > 0) {
      K = 1 / fmax(MINVAL, dmax * dmax * sr0 * sr0 * sr1 * sr1);
      B = 2 / fmax(MINVAL, dmax * sr0);
    } else { K = -sr0 / fmax(MINVAL, dmax * dmax); B = -sr1 / fmax(MINVAL, dmax); }
    d->efc_R[r] = R;
    d->efc_KBIP[4 * r] = K; d->efc_KBIP[4 * r + 1] = B;
    d->efc_KBIP[4 * r + 2] = imp; d->efc_KBIP[4 * r + 3] = impP;
  }
  /* friction dimensions of frictional contacts */
  for (int ci = 0; ci < d->ncon[0]; ci++) {
    OContact* c = d->contact + ci;
    int a = c->efc_address;
    if (a < 0 || c->dim == 1) continue;
    if (m->opt_cone == 0) {
      c->mu = c->friction[0] / sqrt(m->opt_impratio);
      double Rpy = 2 * c->mu * c->mu * d->efc_R[a];
      for (int k = 0; k < 2 * (c->dim - 1); k++) d->efc_R[a + k] = Rpy;
    } else {
      d->efc_R[a + 1] = d->efc_R[a] / fmax(MINVAL, m->opt_impratio);
      c->mu = c->friction[0] * sqrt(d->efc_R[a + 1] / d->efc_R[a]);
      for (int k = 2; k < c->dim; k++)
        d->efc_R[a + k] = d->efc_R[a + 1] * c->friction[0] * c->friction[0] /
                          (c->friction[k - 1] * c->friction[k - 1]);
    }
  }
  for (int r = 0; r < nefc; r++) {
    d->efc_D[r] = 1 / d->efc_R[r];
    double v = 0;
    for (int i = 0; i < nv; i++) v += d->efc_J[(size_t)r * nv + i] * d->qvel[i];
    d->efc_vel[r] = v;
  }
}

/* reference acceleration: needs qvel, hence evaluated with the velocity stage */
static void o_reference_constraint(const OModel* m, OData* d) {
  int nv = m->nv;
  for (int r = 0; r < d->nefc[0]; r++) {
    double v = 0;
    for (int i = 0; i < nv; i++) v += d->efc_J[(size_t)r * nv + i] * d->qvel[i];
    d->efc_vel[r] = v;
    d->efc_aref[r] = -d->efc_KBIP[4 * r + 1] * v -
                     d->efc_KBIP[4 * r] * d->efc_KBIP[4 * r + 2] *
                         (d->efc_pos[r] - d->efc_margin[r]);
  }
}

/* ------------------------------------------------ Newton solver [A.9] */

typedef struct { double cost, d0, d1; } LSPoint;

/* zone of an elliptic contact at residual x[0..dim): 0 top, 1 bottom, 2 middle */
typedef struct { int zone; double N, T, U[6]; } ConeZone;
static ConeZone cone_zone(const OContact* c, const double* x) {
  ConeZone z;
  z.U[0] = x[0] * c->mu;
  double tt = 0;
  for (int j = 1; j < c->dim; j++) { z.U[j] = x[j] * c->friction[j - 1]; tt += z.U[j] * z.U[j]; }
  z.N = z.U[0]; z.T = sqrt(tt);
  if (z.N >= c->mu * z.T || (z.T <= 0 && z.N >= 0)) z.zone = 0;
  else if (c->mu * z.N + z.T <= 0 || (z.T <= 0 && z.N < 0)) z.zone = 1;
  else z.zone = 2;
  return z;
}

/* constraint cost / force / state at acceleration-space residual jar */
static double update_constraint(const OModel* m, OData* d, const double* jar) {
  int nv = m->nv, nefc = d->nefc[0];
  double cost = 0;
  for (int r = 0; r < nefc; r++) {
    double D = d->efc_D[r], x = jar[r];
    int type = d->efc_type[r];
    if (type == CN_EQUALITY) {
      d->efc_force[r] = -D * x; d->efc_state[r] = ST_QUADRATIC; cost += 0.5 * D * x * x;
    } else if (type == CN_FRICTION_DOF || type == CN_FRICTION_TENDON) {
      double f = d->efc_frictionloss[r], R = d->efc_R[r];
      if (x <= -R * f) { d->efc_force[r] = f; d->efc_state[r] = ST_LINEARNEG; cost += -f * (0.5 * R * f + x); }
      else if (x >= R * f) { d->efc_force[r] = -f; d->efc_state[r] = ST_LINEARPOS; cost += -f * (0.5 * R * f - x); }
      else { d->efc_force[r] = -D * x; d->efc_state[r] = ST_QUADRATIC; cost += 0.5 * D * x * x; }
    } else if (type == CN_CONTACT_ELLIPTIC) {
      continue;                                  /* whole contacts, below */
    } else {
      if (x >= 0) { d->efc_force[r] = 0; d->efc_state[r] = ST_SATISFIED; }
      else { d->efc_force[r] = -D * x; d->efc_state[r] = ST_QUADRATIC; cost += 0.5 * D * x * x; }
    }
  }
  /* elliptic cones [A.9]: in the scaled coordinates U = (mu x0, fri_j x_j), N = U0 and
   * T = |U_1..| the dual cone is N >= mu T (top zone: no force), its polar is
   * mu N + T <= 0 (bottom zone: plain quadratic in all rows) and in between the cost is
   * the squared distance to the cone, 0.5 Dm (N - mu T)^2 */
  if (m->opt_cone == 1)
    for (int ci = 0; ci < d->ncon[0]; ci++) {
      const OContact* c = d->contact + ci;
      int a = c->efc_address, dim = c->dim;
      if (a < 0 || dim == 1) continue;
      ConeZone z = cone_zone(c, jar + a);
      if (z.zone == 0) {
        for (int j = 0; j < dim; j++) { d->efc_force[a + j] = 0; d->efc_state[a + j] = ST_SATISFIED; }
      } else if (z.zone == 1) {
        for (int j = 0; j < dim; j++) {
          double D = d->efc_D[a + j], x = jar[a + j];
          d->efc_force[a + j] = -D * x; d->efc_state[a + j] = ST_QUADRATIC;
          cost += 0.5 * D * x * x;
        }
      } else {
        double Dm = d->efc_D[a] / (c->mu * c->mu * (1 + c->mu * c->mu));
        double NmT = z.N - c->mu * z.T;
        cost += 0.5 * Dm * NmT * NmT;
        double f0 = -Dm * NmT * c->mu;
        d->efc_force[a] = f0; d->efc_state[a] = ST_CONE;
        for (int j = 1; j < dim; j++) {
          d->efc_force[a + j] = -f0 / z.T * z.U[j] * c->friction[j - 1];
          d->efc_state[a + j] = ST_CONE;
        }
      }
    }
  for (int i = 0; i < nv; i++) {
    double s = 0;
    for (int r = 0; r < nefc; r++) s += d->efc_J[(size_t)r * nv + i] * d->efc_force[r];
    d->qfrc_constraint[i] = s;
  }
  return cost;
}

static LSPoint ls_eval(const OModel* m, const OData* d, double alpha,
                       const double* quad_gauss) {
  int nefc = d->nefc[0];
  double c0 = quad_gauss[0], c1 = quad_gauss[1], c2 = quad_gauss[2];
  for (int r = 0; r < nefc; r++) {
    double D = d->efc_D[r], j0 = d->s_jar[r], jv = d->s_jv[r];
    double x = j0 + alpha * jv;
    int type = d->efc_type[r];
    double q0 = 0.5 * D * j0 * j0, q1 = D * j0 * jv, q2 = 0.5 * D * jv * jv;
    if (type == CN_EQUALITY) { c0 += q0; c1 += q1; c2 += q2; }
    else if (type == CN_FRICTION_DOF || type == CN_FRICTION_TENDON) {
      double f = d->efc_frictionloss[r], R = d->efc_R[r];
      if (x <= -R * f) { c0 += -f * (0.5 * R * f + j0); c1 += -f * jv; }
      else if (x >= R * f) { c0 += -f * (0.5 * R * f - j0); c1 += f * jv; }
      else { c0 += q0; c1 += q1; c2 += q2; }
    } else if (type == CN_CONTACT_ELLIPTIC) continue;
    else if (x < 0) { c0 += q0; c1 += q1; c2 += q2; }
  }
  double e0 = 0, e1 = 0, e2 = 0;       /* middle-zone cones: not quadratic in alpha */
  if (m->opt_cone == 1)
    for (int ci = 0; ci < d->ncon[0]; ci++) {
      const OContact* c = d->contact + ci;
      int a = c->efc_address, dim = c->dim;
      if (a < 0 || dim == 1) continue;
      double mu = c->mu;
      double U0 = d->s_jar[a] * mu, V0 = d->s_jv[a] * mu, UU = 0, UV = 0, VV = 0;
      for (int j = 1; j < dim; j++) {
        double u = d->s_jar[a + j] * c->friction[j - 1], v = d->s_jv[a + j] * c->friction[j - 1];
        UU += u * u; UV += u * v; VV += v * v;
      }
      double N = U0 + alpha * V0, Tsqr = UU + alpha * (2 * UV + alpha * VV);
      double T = Tsqr > 0 ? sqrt(Tsqr) : 0;
      if (N >= mu * T || (T <= 0 && N >= 0)) continue;
      if (mu * N + T <= 0 || (T <= 0 && N < 0)) {
        for (int j = 0; j < dim; j++) {
          double D = d->efc_D[a + j], j0 = d->s_jar[a + j], jv = d->s_jv[a + j];
          c0 += 0.5 * D * j0 * j0; c1 += D * j0 * jv; c2 += 0.5 * D * jv * jv;
        }
      } else {
        double Dm = d->efc_D[a] / (mu * mu * (1 + mu * mu));
        double N1 = V0, T1 = (UV + alpha * VV) / T;
        double T2 = VV / T - (UV + alpha * VV) * T1 / (T * T);
        double NmT = N - mu * T, NmT1 = N1 - mu * T1;
        e0 += 0.5 * Dm * NmT * NmT;
        e1 += Dm * NmT * NmT1;
        e2 += Dm * (NmT1 * NmT1 - NmT * mu * T2);
      }
    }
  LSPoint p;
  p.cost = alpha * alpha * c2 + alpha * c1 + c0 + e0;
  p.d0 = 2 * alpha * c2 + c1 + e1;
  p.d1 = 2 * c2 + e2;
  return p;
}

/* exact 1-D minimisation along `search`: safeguarded Newton on the (monotone,
 * piecewise-linear) derivative; returns the step length */
static double linesearch(const OModel* m, OData* d, double scale) {
  int nv = m->nv;
  double snorm = 0;
  for (int i = 0; i < nv; i++) snorm += d->s_search[i] * d->s_search[i];
  snorm = sqrt(snorm);
  if (snorm < MINVAL) return 0;
  double gtol = m->opt_tolerance * m->opt_ls_tolerance * snorm / scale;
  double qg[3] = {0, 0, 0};
  for (int i = 0; i < nv; i++) {
    qg[0] += 0.5 * (d->s_Ma[i] - d->qfrc_smooth[i]) * (d->qacc[i] - d->qacc_smooth[i]);
    qg[1] += d->s_search[i] * (d->s_Ma[i] - d->qfrc_smooth[i]);
    qg[2] += 0.5 * d->s_search[i] * d->s_Mv[i];
  }
  LSPoint p = ls_eval(m, d, 0, qg);
  if (p.d1 < MINVAL) return 0;
  double lo = 0, hi = 0;
  int have_hi = 0;
  double a = -p.d0 / p.d1;
  for (int it = 0; it < m->opt_ls_iterations; it++) {
    p = ls_eval(m, d, a, qg);
    if (fabs(p.d0) < gtol) break;
    if (p.d0 < 0) lo = a; else { hi = a; have_hi = 1; }
    double an = a - p.d0 / p.d1;
    if (have_hi && !(an > lo && an < hi)) an = 0.5 * (lo + hi);
    a = an;
  }
  return a;
}

/* dense Cholesky H = L L^T in place (lower), returns 0 on failure */
static int chol_factor(double* H, int n) {
  for (int j = 0; j < n; j++) {
    double s = H[j * n + j];
    for (int k = 0; k < j; k++) s -= H[j * n + k] * H[j * n + k];
    if (s < MINVAL) return 0;
    s = sqrt(s);
    H[j * n + j] = s;
    for (int i = j + 1; i < n; i++) {
      double t = H[i * n + j];
      for (int k = 0; k < j; k++) t -= H[i * n + k] * H[j * n + k];
      H[i * n + j] = t / s;
    }
  }
  return 1;
}
static void chol_solve(const double* L, int n, double* x) {
  for (int i = 0; i < n; i++) {
    double t = x[i];
    for (int k = 0; k < i; k++) t -= L[i * n + k] * x[k];
    x[i] = t / L[i * n + i];
  }
  for (int i = n - 1; i >= 0; i--) {
    double t = x[i];
    for (int k = i + 1; k < n; k++) t -= L[k * n + i] * x[k];
    x[i] = t / L[i * n + i];
  }
}

static void dense_m(const OModel* m, const double* qM, double* Md) {
  int nv = m->nv;
  memset(Md, 0, sizeof(double) * nv * nv);
  for (int i = 0; i < nv; i++) {
    int a = m->dof_Madr[i];
    for (int j = i; j >= 0; j = m->dof_parentid[j], a++) {
      Md[i * nv + j] = qM[a]; Md[j * nv + i] = qM[a];
    }
  }
}

static void newton_gradient(const OModel* m, OData* d) {
  int nv = m->nv, nefc = d->nefc[0];
  for (int i = 0; i < nv; i++)
    d->s_grad[i] = d->s_Ma[i] - d->qfrc_smooth[i] - d->qfrc_constraint[i];
  memcpy(d->s_H, d->s_Mdense, sizeof(double) * nv * nv);
  for (int r = 0; r < nefc; r++) {
    if (d->efc_state[r] != ST_QUADRATIC) continue;
    const double* J = d->efc_J + (size_t)r * nv;
    double D = d->efc_D[r];
    for (int i = 0; i < nv; i++) {
      if (J[i] == 0) continue;
      for (int j = 0; j <= i; j++) d->s_H[i * nv + j] += D * J[i] * J[j];
    }
  }
  /* middle-zone cones: H += Jc^T Hc Jc with Hc = Dm S (g g^T + kappa (I - u u^T)_t) S,
   * S = diag(mu, fri), g = (1, -mu u), u = U_t / T, kappa = mu^2 - mu N / T */
  if (m->opt_cone == 1)
    for (int ci = 0; ci < d->ncon[0]; ci++) {
      const OContact* c = d->contact + ci;
      int a = c->efc_address, dim = c->dim;
      if (a < 0 || dim == 1 || d->efc_state[a] != ST_CONE) continue;
      ConeZone z = cone_zone(c, d->s_jar + a);
      double mu = c->mu, Dm = d->efc_D[a] / (mu * mu * (1 + mu * mu));
      double kappa = mu * mu - mu * z.N / z.T;
      double S[6], g[6], Hc[36];
      S[0] = mu; g[0] = 1;
      for (int j = 1; j < dim; j++) { S[j] = c->friction[j - 1]; g[j] = -mu * z.U[j] / z.T; }
      for (int p = 0; p < dim; p++)
        for (int q = 0; q < dim; q++) {
          double h = g[p] * g[q];
          if (p > 0 && q > 0)
            h += kappa * ((p == q ? 1.0 : 0.0) - z.U[p] * z.U[q] / (z.T * z.T));
          Hc[p * dim + q] = Dm * S[p] * S[q] * h;
        }
      const double* J = d->efc_J + (size_t)a * nv;
      for (int i = 0; i < nv; i++) {
        double y[6];
        int nz = 0;
        for (int p = 0; p < dim; p++) {
          y[p] = 0;
          for (int q = 0; q < dim; q++) y[p] += Hc[p * dim + q] * J[q * nv + i];
          nz |= J[p * nv + i] != 0;
        }
        if (!nz) continue;
        for (int j = 0; j <= i; j++) {
          double sum = 0;
          for (int p = 0; p < dim; p++) sum += y[p] * J[p * nv + j];
          d->s_H[i * nv + j] += sum;
        }
      }
    }
  if (!chol_factor(d->s_H, nv)) {
    /* not expected (M is PD); fall back to the unconstrained metric */
    memcpy(d->s_H, d->s_Mdense, sizeof(double) * nv * nv);
    chol_factor(d->s_H, nv);
  }
  memcpy(d->s_Mgrad, d->s_grad, sizeof(double) * nv);
  chol_solve(d->s_H, nv, d->s_Mgrad);
}

static double total_cost(const OModel* m, OData* d, double ccost) {
  double g = 0;
  for (int i = 0; i < m->nv; i++)
    g += 0.5 * (d->s_Ma[i] - d->qfrc_smooth[i]) * (d->qacc[i] - d->qacc_smooth[i]);
  return ccost + g;
}

static void jar_from_qacc(const OModel* m, OData* d) {
  int nv = m->nv;
  for (int r = 0; r < d->nefc[0]; r++) {
    double s = 0;
    for (int i = 0; i < nv; i++) s += d->efc_J[(size_t)r * nv + i] * d->qacc[i];
    d->s_jar[r] = s - d->efc_aref[r];
  }
}

/* CG [MJ-doc A.9: "same cost, Polak-Ribiere with M^-1 preconditioner"]: the gradient of the
 * primal cost preconditioned by the sparse factor of M instead of the Newton Hessian */
static void cg_gradient(const OModel* m, OData* d) {
  int nv = m->nv;
  for (int i = 0; i < nv; i++)
    d->s_grad[i] = d->s_Ma[i] - d->qfrc_smooth[i] - d->qfrc_constraint[i];
  memcpy(d->s_Mgrad, d->s_grad, sizeof(double) * nv);
  solve_sparse(m, d->qLD, d->qLDiagInv, d->s_Mgrad);
}

/* PGS [MJ-doc A.9: "dual, needs efc_AR = J M^-1 J^T + R, Gauss-Seidel with projection"]:
 * minimise 0.5 f^T AR f + f^T b over the constraint set, b = J qacc_smooth - aref, one
 * row at a time in row order: equality rows unconstrained, friction-loss rows clamped to
 * +-frictionloss, limit / contact rows (frictionless, pyramid edges) to f >= 0.  The warm
 * start maps qacc_warmstart to forces through the primal constraint update and is kept
 * when its dual cost is below that of f = 0 (zero).  Stops when the scaled cost decrease
 * of a sweep falls below opt.tolerance.  Pyramidal cones only (validate()). */
static void pgs_solve(const OModel* m, OData* d) {
  int nv = m->nv, nefc = d->nefc[0];
  double *AR = d->s_AR, *b = d->s_jv, *x = d->s_grad, *f = d->efc_force;
  for (int i = 0; i < nefc; i++) {
    memcpy(x, d->efc_J + (size_t)i * nv, sizeof(double) * nv);
    solve_sparse(m, d->qLD, d->qLDiagInv, x);
    for (int j = 0; j < nefc; j++) {
      double s = 0;
      for (int k = 0; k < nv; k++) s += d->efc_J[(size_t)j * nv + k] * x[k];
      AR[(size_t)i * nefc + j] = s;
    }
    AR[(size_t)i * nefc + i] += d->efc_R[i];
  }
  for (int r = 0; r < nefc; r++) {
    double s = 0;
    for (int i = 0; i < nv; i++) s += d->efc_J[(size_t)r * nv + i] * d->qacc_smooth[i];
    b[r] = s - d->efc_aref[r];
  }
  int warm = 0;
  if (!(m->opt_disableflags & DSBL_WARMSTART)) {
    memcpy(d->qacc, d->qacc_warmstart, sizeof(double) * nv);
    jar_from_qacc(m, d);
    update_constraint(m, d, d->s_jar);
    double cost = 0;
    for (int i = 0; i < nefc; i++) {
      double s = 0;
      for (int j = 0; j < nefc; j++) s += AR[(size_t)i * nefc + j] * f[j];
      cost += f[i] * (0.5 * s + b[i]);
    }
    warm = cost < 0;
  }
  if (!warm) memset(f, 0, sizeof(double) * nefc);
  double scale = 1.0 / (m->stat_meaninertia * (nv > 1 ? nv : 1));
  int iter = 0;
  while (iter < m->opt_iterations) {
    double improvement = 0;
    for (int i = 0; i < nefc; i++) {
      double res = b[i];
      for (int j = 0; j < nefc; j++) res += AR[(size_t)i * nefc + j] * f[j];
      double aii = AR[(size_t)i * nefc + i], old = f[i], fn = old - res / aii;
      int t = d->efc_type[i];
      if (t == CN_FRICTION_DOF || t == CN_FRICTION_TENDON) {
        double fl = d->efc_frictionloss[i];
        fn = fn < -fl ? -fl : (fn > fl ? fl : fn);
      } else if (t != CN_EQUALITY) {
        if (fn < 0) fn = 0;
      }
      double delta = fn - old;
      f[i] = fn;
      improvement -= 0.5 * delta * delta * aii + delta * res;
    }
    improvement *= scale;
    d->solver_improvement[0] = improvement; d->solver_gradient[0] = 0;
    iter++;
    if (improvement < m->opt_tolerance) break;
  }
  for (int i = 0; i < nefc; i++) {
    int t = d->efc_type[i];
    if (t == CN_EQUALITY) d->efc_state[i] = ST_QUADRATIC;
    else if (t == CN_FRICTION_DOF || t == CN_FRICTION_TENDON) {
      double fl = d->efc_frictionloss[i];
      d->efc_state[i] = f[i] >= fl ? ST_LINEARNEG : (f[i] <= -fl ? ST_LINEARPOS : ST_QUADRATIC);
    } else d->efc_state[i] = f[i] > 0 ? ST_QUADRATIC : ST_SATISFIED;
  }
  for (int i = 0; i < nv; i++) {
    double s = 0;
    for (int r = 0; r < nefc; r++) s += d->efc_J[(size_t)r * nv + i] * f[r];
    d->qfrc_constraint[i] = s;
    x[i] = s;
  }
  solve_sparse(m, d->qLD, d->qLDiagInv, x);
  for (int i = 0; i < nv; i++) d->qacc[i] = d->qacc_smooth[i] + x[i];
  d->solver_iter[0] = iter;
  memcpy(d->qacc_warmstart, d->qacc, sizeof(double) * nv);
}

static void o_fwd_constraint(const OModel* m, OData* d) {
  int nv = m->nv, nefc = d->nefc[0];
  const int cg = m->opt_solver == 1;
  d->solver_iter[0] = 0;
  if (nefc == 0) {
    memcpy(d->qacc, d->qacc_smooth, sizeof(double) * nv);
    memcpy(d->qacc_warmstart, d->qacc_smooth, sizeof(double) * nv);
    memset(d->qfrc_constraint, 0, sizeof(double) * nv);
    return;
  }
  if (m->opt_solver == 0) { pgs_solve(m, d); return; }
  dense_m(m, d->qM, d->s_Mdense);
  /* warm start: the cheaper of qacc_warmstart and qacc_smooth */
  double cost;
  if (!(m->opt_disableflags & DSBL_WARMSTART)) {
    memcpy(d->qacc, d->qacc_warmstart, sizeof(double) * nv);
    mul_m(m, d->qM, d->s_Ma, d->qacc);
    jar_from_qacc(m, d);
    double cw = total_cost(m, d, update_constraint(m, d, d->s_jar));
    memcpy(d->qacc, d->qacc_smooth, sizeof(double) * nv);
    memcpy(d->s_Ma, d->qfrc_smooth, sizeof(double) * nv);
    jar_from_qacc(m, d);
    double cs = update_constraint(m, d, d->s_jar);
    if (cw < cs) memcpy(d->qacc, d->qacc_warmstart, sizeof(double) * nv);
  } else {
    memcpy(d->qacc, d->qacc_smooth, sizeof(double) * nv);
  }
  mul_m(m, d->qM, d->s_Ma, d->qacc);
  jar_from_qacc(m, d);
  cost = total_cost(m, d, update_constraint(m, d, d->s_jar));
  if (cg) cg_gradient(m, d); else newton_gradient(m, d);
  for (int i = 0; i < nv; i++) d->s_search[i] = -d->s_Mgrad[i];
  double scale = 1.0 / (m->stat_meaninertia * (nv > 1 ? nv : 1));
  int iter = 0;
  while (iter < m->opt_iterations) {
    mul_m(m, d->qM, d->s_Mv, d->s_search);
    for (int r = 0; r < nefc; r++) {
      double s = 0;
      for (int i = 0; i < nv; i++) s += d->efc_J[(size_t)r * nv + i] * d->s_search[i];
      d->s_jv[r] = s;
    }
    double alpha = linesearch(m, d, scale);
    if (alpha == 0) break;
    for (int i = 0; i < nv; i++) {
      d->qacc[i] += alpha * d->s_search[i];
      d->s_Ma[i] += alpha * d->s_Mv[i];
    }
    for (int r = 0; r < nefc; r++) d->s_jar[r] += alpha * d->s_jv[r];
    double old = cost, cg_den = 0;
    if (cg) {                     /* Polak-Ribiere needs the previous gradient pair */
      for (int i = 0; i < nv; i++) cg_den += d->s_grad[i] * d->s_Mgrad[i];
      memcpy(d->s_Mgold, d->s_Mgrad, sizeof(double) * nv);
    }
    cost = total_cost(m, d, update_constraint(m, d, d->s_jar));
    if (cg) cg_gradient(m, d); else newton_gradient(m, d);
    double gn = 0;
    for (int i = 0; i < nv; i++) gn += d->s_grad[i] * d->s_grad[i];
    double improvement = scale * (old - cost), gradient = scale * sqrt(gn);
    d->solver_improvement[0] = improvement; d->solver_gradient[0] = gradient;
    iter++;
    if (improvement < m->opt_tolerance || gradient < m->opt_tolerance) break;
    if (cg) {
      double num = 0;
      for (int i = 0; i < nv; i++) num += d->s_grad[i] * (d->s_Mgrad[i] - d->s_Mgold[i]);
      double beta = num / (cg_den > MINVAL ? cg_den : MINVAL);
      if (beta < 0) beta = 0;
      for (int i = 0; i < nv; i++) d->s_search[i] = -d->s_Mgrad[i] + beta * d->s_search[i];
    } else {
      for (int i = 0; i < nv; i++) d->s_search[i] = -d->s_Mgrad[i];
    }
  }
  d->solver_iter[0] = iter;
  memcpy(d->qacc_warmstart, d->qacc, sizeof(double) * nv);
}

/* ------------------------------------------------------- sensors [A.11] */

static void o_subtree_vel(const OModel* m, OData* d) {
  int nb = m->nbody;
  for (int b = 0; b < nb; b++) {
    const double* cv = d->cvel + 6 * b;
    const double* com = d->subtree_com + 3 * m->body_rootid[b];
    double dif[3], t[3];
    for (int c = 0; c < 3; c++) dif[c] = d->xipos[3 * b + c] - com[c];
    cross3(t, cv, dif);
    for (int c = 0; c < 3; c++)
      d->subtree_linvel[3 * b + c] = m->body_mass[b] * (cv[3 + c] + t[c]);
  }
  for (int b = nb - 1; b > 0; b--) {
    int p = m->body_parentid[b];
    for (int c = 0; c < 3; c++) d->subtree_linvel[3 * p + c] += d->subtree_linvel[3 * b + c];
  }
  for (int b = 0; b < nb; b++) {
    double sm = m->body_subtreemass[b] > MINVAL ? m->body_subtreemass[b] : MINVAL;
    for (int c = 0; c < 3; c++) d->subtree_linvel[3 * b + c] /= sm;
  }
}

static void contact_force(const OModel* m, const OData* d, int ci, double* f6) {
  const OContact* c = d->contact + ci;
  memset(f6, 0, 48);
  int a = c->efc_address;
  if (a < 0) return;
  if (c->dim == 1) { f6[0] = d->efc_force[a]; return; }
  if (m->opt_cone == 0) {
    for (int k = 0; k < 2 * (c->dim - 1); k++) f6[0] += d->efc_force[a + k];
    for (int k = 1; k < c->dim; k++)
      f6[k] = c->friction[k - 1] * (d->efc_force[a + 2 * k - 2] - d->efc_force[a + 2 * k - 1]);
  } else {
    for (int k = 0; k < c->dim; k++) f6[k] = d->efc_force[a + k];
  }
}

static void o_rne_post_constraint(const OModel* m, OData* d) {
  int nb = m->nbody;
  memset(d->cfrc_ext, 0, sizeof(double) * 6 * nb);
  for (int ci = 0; ci < d->ncon[0]; ci++) {
    const OContact* c = d->contact + ci;
    if (c->efc_address < 0) continue;
    double f6[6], fw[3], tw[3];
    contact_force(m, d, ci, f6);
    matT_vec3(fw, c->frame, f6);
    matT_vec3(tw, c->frame, f6 + 3);
    for (int side = 0; side < 2; side++) {
      int b = m->geom_bodyid[side ? c->geom2 : c->geom1];
      if (b == 0) continue;
      double sgn = side ? 1 : -1;
      const double* com = d->subtree_com + 3 * m->body_rootid[b];
      double dif[3] = {c->pos[0] - com[0], c->pos[1] - com[1], c->pos[2] - com[2]}, t[3];
      cross3(t, dif, fw);
      for (int k = 0; k < 3; k++) {
        d->cfrc_ext[6 * b + k] += sgn * (tw[k] + t[k]);
        d->cfrc_ext[6 * b + 3 + k] += sgn * fw[k];
      }
    }
  }
  memset(d->cacc, 0, 48);
  if (!(m->opt_disableflags & DSBL_GRAVITY))
    for (int c = 0; c < 3; c++) d->cacc[3 + c] = -m->opt_gravity[c];
  memset(d->cfrc_int, 0, 48);
  for (int b = 1; b < nb; b++) {
    double* a = d->cacc + 6 * b;
    memcpy(a, d->cacc + 6 * m->body_parentid[b], 48);
    int da = m->body_dofadr[b];
    for (int k = 0; k < m->body_dofnum[b]; k++)
      for (int c = 0; c < 6; c++)
        a[c] += d->cdof_dot[6 * (da + k) + c] * d->qvel[da + k] +
                d->cdof[6 * (da + k) + c] * d->qacc[da + k];
    double Ia[6], Iv[6], vxIv[6];
    inert_mul(Ia, d->cinert + 10 * b, a);
    inert_mul(Iv, d->cinert + 10 * b, d->cvel + 6 * b);
    cross_force(vxIv, d->cvel + 6 * b, Iv);
    for (int c = 0; c < 6; c++) d->cfrc_int[6 * b + c] = Ia[c] + vxIv[c] - d->cfrc_ext[6 * b + c];
  }
  for (int b = nb - 1; b > 0; b--) {
    int p = m->body_parentid[b];
    if (p > 0)
      for (int c = 0; c < 6; c++) d->cfrc_int[6 * p + c] += d->cfrc_int[6 * b + c];
  }
}

/* ------------------------------------------------------- rays [mj_ray]
 * Distance along a unit ray to the nearest geom surface, -1 without a hit; the
 * rangefinder sensor casts it from its site along the site's z axis and skips the geoms
 * of the site's own body [MJ-doc: mj_ray, mju_rayGeom].  Solutions of a quadratic: the
 * smallest non-negative root, so a ray that starts inside a geom reports its exit. */
static double ray_quad(double a, double b, double c) {   /* a x^2 + 2 b x + c = 0 */
  double det = b * b - a * c;
  if (det < 0 || a < MINVAL) return -1;
  det = sqrt(det);
  double x0 = (-b - det) / a, x1 = (-b + det) / a;
  if (x0 >= 0) return x0;
  if (x1 >= 0) return x1;
  return -1;
}

static double ray_min(double x, double y) {   /* smaller non-negative of the two */
  if (x < 0) return y;
  if (y < 0) return x;
  return x < y ? x : y;
}

/* ray against the top surface of one triangle (Moeller-Trumbore, both faces) */
static double ray_triangle(const double* o, const double* v, const double* a,
                           const double* b, const double* c) {
  double e1[3], e2[3], pv[3], tv[3], qv[3];
  for (int k = 0; k < 3; k++) { e1[k] = b[k] - a[k]; e2[k] = c[k] - a[k]; tv[k] = o[k] - a[k]; }
  cross3(pv, v, e2);
  double det = dot3(e1, pv);
  if (fabs(det) < MINVAL) return -1;
  double u = dot3(tv, pv) / det;
  if (u < 0 || u > 1) return -1;
  cross3(qv, tv, e1);
  double w = dot3(v, qv) / det;
  if (w < 0 || u + w > 1) return -1;
  double t = dot3(e2, qv) / det;
  return t >= 0 ? t : -1;
}

/* ray (origin o, unit direction v, field frame) against the height field: the box is
 * entered through a wall or the bottom (solid), or the surface is hit further on.  The
 * grid is swept slab by slab along the axis the ray advances fastest on. */
static double ray_hfield(const HField* h, const double* o, const double* v) {
  double lo[3] = {-h->rx, -h->ry, -h->base}, hi[3] = {h->rx, h->ry, h->zs};
  double t0 = 0, t1 = 1e300;
  int axis0 = -1;
  for (int k = 0; k < 3; k++) {
    if (fabs(v[k]) < MINVAL) {
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
    /* entering through a wall (below the local surface) or through the bottom */
    if (axis0 == 2) return t0;
    double x = o[0] + t0 * v[0], y = o[1] + t0 * v[1], z = o[2] + t0 * v[2];
    if (z <= hf_surface(h, x, y, NULL)) return t0;
  }
  int am = fabs(v[0]) >= fabs(v[1]) ? 0 : 1, an = 1 - am;
  double radm = am == 0 ? h->rx : h->ry, radn = am == 0 ? h->ry : h->rx;
  double ddm = am == 0 ? h->dx : h->dy, ddn = am == 0 ? h->dy : h->dx;
  int nm = am == 0 ? h->ncol : h->nrow, nn = am == 0 ? h->nrow : h->ncol;
  int i0 = hf_cell(o[am] + t0 * v[am], radm, ddm, nm);
  int i1 = hf_cell(o[am] + t1 * v[am], radm, ddm, nm);
  int step = i1 >= i0 ? 1 : -1;
  for (int i = i0; i != i1 + step; i += step) {
    double ta = t0, tb = t1;
    if (fabs(v[am]) >= MINVAL) {
      double sa = (i * ddm - radm - o[am]) / v[am], sb = ((i + 1) * ddm - radm - o[am]) / v[am];
      if (sa > sb) { double s = sa; sa = sb; sb = s; }
      if (sa > ta) ta = sa;
      if (sb < tb) tb = sb;
    }
    if (ta > tb) continue;
    double na = o[an] + ta * v[an], nb = o[an] + tb * v[an];
    int j0 = hf_cell(na < nb ? na : nb, radn, ddn, nn), j1 = hf_cell(na < nb ? nb : na, radn, ddn, nn);
    for (int j = j0; j <= j1; j++) {
      int r = am == 0 ? j : i, c = am == 0 ? i : j;
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

static double ray_geom(const OModel* m, const OData* d, int g, const double* pnt,
                       const double* vec) {
  const double *pos = d->geom_xpos + 3 * g, *R = d->geom_xmat + 9 * g;
  const double* sz = m->geom_size + 3 * g;
  double dif[3] = {pnt[0] - pos[0], pnt[1] - pos[1], pnt[2] - pos[2]}, o[3], v[3];
  matT_vec3(o, R, dif);
  matT_vec3(v, R, vec);
  switch (m->geom_type[g]) {
    case GEOM_PLANE: {
      if (v[2] > -MINVAL) return -1;            /* only the front face, from above */
      double x = -o[2] / v[2];
      if (x < 0) return -1;
      double p0 = o[0] + x * v[0], p1 = o[1] + x * v[1];
      if ((sz[0] <= 0 || fabs(p0) <= sz[0]) && (sz[1] <= 0 || fabs(p1) <= sz[1])) return x;
      return -1;
    }
    case GEOM_HFIELD: {
      HField h;
      hf_init(m, d, m->geom_dataid[g], &h);
      return ray_hfield(&h, o, v);
    }
    case GEOM_SPHERE:
      return ray_quad(dot3(v, v), dot3(v, o), dot3(o, o) - sz[0] * sz[0]);
    case GEOM_CAPSULE: {
      /* side wall where |z| <= half length, else the end spheres */
      double a = v[0] * v[0] + v[1] * v[1], b = v[0] * o[0] + v[1] * o[1],
             c = o[0] * o[0] + o[1] * o[1] - sz[0] * sz[0];
      double x = ray_quad(a, b, c);
      if (x >= 0 && fabs(o[2] + x * v[2]) > sz[1]) x = -1;
      for (int sgn = -1; sgn <= 1; sgn += 2) {
        double oz = o[2] - sgn * sz[1];
        double dd = dot3(v, v), db = v[0] * o[0] + v[1] * o[1] + v[2] * oz,
               dc = o[0] * o[0] + o[1] * o[1] + oz * oz - sz[0] * sz[0];
        double det = db * db - dd * dc;
        if (det < 0) continue;
        det = sqrt(det);
        /* the near root if it lies on the cap's outer half, else the far one */
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
      if (fabs(v[2]) >= MINVAL)
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
      for (int k = 0; k < 3; k++) {
        if (fabs(v[k]) < MINVAL) continue;
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

static double o_ray(const OModel* m, const OData* d, const double* pnt, const double* vec,
                    int bodyexclude) {
  double best = -1;
  for (int g = 0; g < m->ngeom; g++) {
    if (m->geom_bodyid[g] == bodyexclude) continue;
    best = ray_min(best, ray_geom(m, d, g, pnt, vec));
  }
  return best;
}

static int inside_site(const OModel* m, const OData* d, int s, const double* p) {
  double dif[3] = {p[0] - d->site_xpos[3 * s], p[1] - d->site_xpos[3 * s + 1],
                   p[2] - d->site_xpos[3 * s + 2]}, l[3];
  matT_vec3(l, d->site_xmat + 9 * s, dif);
  const double* sz = m->site_size + 3 * s;
  switch (m->site_type[s]) {
    case GEOM_SPHERE: return dot3(l, l) <= sz[0] * sz[0];
    case GEOM_CAPSULE: {
      double z = l[2] < -sz[1] ? -sz[1] : (l[2] > sz[1] ? sz[1] : l[2]);
      double dz = l[2] - z;
      return l[0] * l[0] + l[1] * l[1] + dz * dz <= sz[0] * sz[0];
    }
    case GEOM_ELLIPSOID:
      return l[0] * l[0] / (sz[0] * sz[0]) + l[1] * l[1] / (sz[1] * sz[1]) +
             l[2] * l[2] / (sz[2] * sz[2]) <= 1;
    case GEOM_CYLINDER:
      return fabs(l[2]) <= sz[1] && l[0] * l[0] + l[1] * l[1] <= sz[0] * sz[0];
    case GEOM_BOX:
      return fabs(l[0]) <= sz[0] && fabs(l[1]) <= sz[1] && fabs(l[2]) <= sz[2];
  }
  return 0;
}

static void obj_frame(const OModel* m, const OData* d, int objtype, int id,
                      const double** pos, const double** mat, int* body) {
  if (objtype == OBJ_SITE) { *pos = d->site_xpos + 3 * id; *mat = d->site_xmat + 9 * id; *body = m->site_bodyid[id]; }
  else if (objtype == OBJ_GEOM) { *pos = d->geom_xpos + 3 * id; *mat = d->geom_xmat + 9 * id; *body = m->geom_bodyid[id]; }
  else if (objtype == OBJ_XBODY) { *pos = d->xpos + 3 * id; *mat = d->xmat + 9 * id; *body = id; }
  else { *pos = d->xipos + 3 * id; *mat = d->ximat + 9 * id; *body = id; }
}

static void o_sensors(const OModel* m, OData* d, int stage) {
  int need_subtree = 0, need_rne = 0;
  for (int s = 0; s < m->nsensor; s++) {
    if (m->sensor_needstage[s] != stage) continue;
    int t = m->sensor_type[s];
    if (t == SENS_SUBTREELINVEL) need_subtree = 1;
    if (t == SENS_ACCELEROMETER || t == SENS_FORCE || t == SENS_TORQUE) need_rne = 1;
  }
  if (need_subtree) o_subtree_vel(m, d);
  if (need_rne) o_rne_post_constraint(m, d);
  for (int s = 0; s < m->nsensor; s++) {
    if (m->sensor_needstage[s] != stage) continue;
    int t = m->sensor_type[s], id = m->sensor_objid[s];
    double* out = d->sensordata + m->sensor_adr[s];
    const double *pos, *mat;
    int body;
    switch (t) {
      case SENS_JOINTPOS: out[0] = d->qpos[m->jnt_qposadr[id]]; break;
      case SENS_JOINTVEL: out[0] = d->qvel[m->jnt_dofadr[id]]; break;
      case SENS_ACTUATORFRC: out[0] = d->actuator_force[id]; break;
      case SENS_SUBTREECOM: memcpy(out, d->subtree_com + 3 * id, 24); break;
      case SENS_RANGEFINDER: {
        double zax[3];
        geom_axis(d->site_xmat + 9 * id, 2, zax);
        out[0] = o_ray(m, d, d->site_xpos + 3 * id, zax, m->site_bodyid[id]);
        break;
      }
      case SENS_SUBTREELINVEL: memcpy(out, d->subtree_linvel + 3 * id, 24); break;
      case SENS_FRAMEPOS:
        obj_frame(m, d, m->sensor_objtype[s], id, &pos, &mat, &body);
        memcpy(out, pos, 24);
        break;
      case SENS_FRAMEXAXIS: case SENS_FRAMEYAXIS: case SENS_FRAMEZAXIS:
        obj_frame(m, d, m->sensor_objtype[s], id, &pos, &mat, &body);
        geom_axis(mat, t - SENS_FRAMEXAXIS, out);
        break;
      case SENS_VELOCIMETER: case SENS_GYRO: case SENS_ACCELEROMETER: {
        obj_frame(m, d, OBJ_SITE, id, &pos, &mat, &body);
        const double* com = d->subtree_com + 3 * m->body_rootid[body];
        const double* cv = d->cvel + 6 * body;
        double dif[3] = {pos[0] - com[0], pos[1] - com[1], pos[2] - com[2]};
        double t3[3], vlin[3], wl[3], vl[3];
        cross3(t3, cv, dif);
        for (int c = 0; c < 3; c++) vlin[c] = cv[3 + c] + t3[c];
        matT_vec3(wl, mat, cv);
        matT_vec3(vl, mat, vlin);
        if (t == SENS_GYRO) memcpy(out, wl, 24);
        else if (t == SENS_VELOCIMETER) memcpy(out, vl, 24);
        else {
          const double* ca = d->cacc + 6 * body;
          double alin[3], al[3], cor[3];
          cross3(t3, ca, dif);
          for (int c = 0; c < 3; c++) alin[c] = ca[3 + c] + t3[c];
          matT_vec3(al, mat, alin);
          cross3(cor, wl, vl);
          for (int c = 0; c < 3; c++) out[c] = al[c] + cor[c];
        }
        break;
      }
      case SENS_FORCE: case SENS_TORQUE: {
        obj_frame(m, d, OBJ_SITE, id, &pos, &mat, &body);
        const double* com = d->subtree_com + 3 * m->body_rootid[body];
        const double* cf = d->cfrc_int + 6 * body;
        if (t == SENS_FORCE) matT_vec3(out, mat, cf + 3);
        else {
          double dif[3] = {pos[0] - com[0], pos[1] - com[1], pos[2] - com[2]}, t3[3], tq[3];
          cross3(t3, dif, cf + 3);
          for (int c = 0; c < 3; c++) tq[c] = cf[c] - t3[c];
          matT_vec3(out, mat, tq);
        }
        break;
      }
      case SENS_TOUCH: {
        int sb = m->site_bodyid[id];
        out[0] = 0;
        for (int ci = 0; ci < d->ncon[0]; ci++) {
          const OContact* c = d->contact + ci;
          if (c->efc_address < 0) continue;
          if (m->geom_bodyid[c->geom1] != sb && m->geom_bodyid[c->geom2] != sb) continue;
          double f6[6];
          contact_force(m, d, ci, f6);
          if (f6[0] <= 0) continue;
          if (inside_site(m, d, id, c->pos)) out[0] += f6[0];
        }
        break;
      }
    }
    double cut = m->sensor_cutoff[s];
    if (cut > 0)
      for (int c = 0; c < m->sensor_dim[s]; c++) {
        if (t == SENS_TOUCH) { if (out[c] > cut) out[c] = cut; }
        else { if (out[c] > cut) out[c] = cut; else if (out[c] < -cut) out[c] = -cut; }
      }
  }
}

static void o_energy_pos(const OModel* m, OData* d) {
  if (!(m->opt_enableflags & ENBL_ENERGY)) { d->energy[0] = d->energy[1] = 0; return; }
  double e = 0;
  if (!(m->opt_disableflags & DSBL_GRAVITY))
    for (int b = 1; b < m->nbody; b++)
      e -= m->body_mass[b] * dot3(m->opt_gravity, d->xipos + 3 * b);
  if (!(m->opt_disableflags & DSBL_PASSIVE))
    for (int j = 0; j < m->njnt; j++) {
      int jt = m->jnt_type[j], qa = m->jnt_qposadr[j];
      if (jt == JNT_SLIDE || jt == JNT_HINGE) {
        double dq = d->qpos[qa] - m->qpos_spring[qa];
        e += 0.5 * m->jnt_stiffness[j] * dq * dq;
      }
    }
  d->energy[0] = e;
}

static void o_energy_vel(const OModel* m, OData* d) {
  if (!(m->opt_enableflags & ENBL_ENERGY)) return;
  double* t = (double*)calloc(m->nv > 0 ? m->nv : 1, 8);
  mul_m(m, d->qM, t, d->qvel);
  double e = 0;
  for (int i = 0; i < m->nv; i++) e += 0.5 * t[i] * d->qvel[i];
  d->energy[1] = e;
  free(t);
}

/* ---------------------------------------------------- pipeline (§3.1) */

static int is_bad(double x) { return isnan(x) || x > MAXVAL || x < -MAXVAL; }

static int check_array(const OModel* m, OData* d, const double* x, int n, int w) {
  for (int i = 0; i < n; i++)
    if (is_bad(x[i])) {
      o_reset(m, d);
      d->warning[w] = 1;
      return 1;
    }
  return 0;
}

static void fwd_position(const OModel* m, OData* d) {
  o_kinematics(m, d);
  o_com_pos(m, d);
  o_tendon(m, d);
  o_transmission(m, d);
  o_crb(m, d);
  o_factor_m(m, d);
  o_collision(m, d);
  o_make_constraint(m, d);
}

static void fwd_velocity(const OModel* m, OData* d) {
  int nv = m->nv;
  for (int i = 0; i < m->nu; i++) {
    double v = 0;
    for (int k = 0; k < nv; k++) v += d->actuator_moment[i * nv + k] * d->qvel[k];
    d->actuator_velocity[i] = v;
  }
  for (int t = 0; t < m->ntendon; t++) {
    double v = 0;
    if (m->tendon_Jadr[t] >= 0)
      for (int i = 0; i < nv; i++) v += d->ten_J[(size_t)m->tendon_Jadr[t] * nv + i] * d->qvel[i];
    else
      for (int w = m->tendon_adr[t]; w < m->tendon_adr[t] + m->tendon_num[t]; w++)
        v += d->wrap_prm[w] * d->qvel[m->jnt_dofadr[m->wrap_objid[w]]];
    d->ten_velocity[t] = v;
  }
  o_com_vel(m, d);
  o_passive(m, d);
  o_reference_constraint(m, d);
  o_rne(m, d, 0, d->qfrc_bias);
}

static void forward_skip(const OModel* m, OData* d, int skipsensor) {
  fwd_position(m, d);
  if (!skipsensor) o_sensors(m, d, STAGE_POS);
  o_energy_pos(m, d);
  fwd_velocity(m, d);
  if (!skipsensor) o_sensors(m, d, STAGE_VEL);
  o_energy_vel(m, d);
  o_fwd_actuation(m, d);
  o_fwd_acceleration(m, d);
  o_fwd_constraint(m, d);
  if (!skipsensor) o_sensors(m, d, STAGE_ACC);
}

void o_forward(const OModel* m, OData* d) { forward_skip(m, d, 0); }

static void integrate_pos(const OModel* m, double* qpos, const double* qvel, double dt) {
  for (int j = 0; j < m->njnt; j++) {
    int qa = m->jnt_qposadr[j], da = m->jnt_dofadr[j];
    switch (m->jnt_type[j]) {
      case JNT_FREE:
        for (int c = 0; c < 3; c++) qpos[qa + c] += dt * qvel[da + c];
        quat_integrate(qpos + qa + 3, qvel + da + 3, dt);
        break;
      case JNT_BALL: quat_integrate(qpos + qa, qvel + da, dt); break;
      default: qpos[qa] += dt * qvel[da];
    }
  }
}

/* qvel += dt*qacc ; qpos integrates with qvel_for_pos (NULL = updated qvel) */
static void advance(const OModel* m, OData* d, const double* act_dot,
                    const double* qacc, const double* qvel_for_pos) {
  double dt = m->opt_timestep;
  for (int i = 0; i < m->na; i++) d->act[i] += dt * act_dot[i];
  for (int i = 0; i < m->nv; i++) d->qvel[i] += dt * qacc[i];
  integrate_pos(m, d->qpos, qvel_for_pos ? qvel_for_pos : d->qvel, dt);
  d->time[0] += dt;
}

static void o_euler(const OModel* m, OData* d) {
  int nv = m->nv, damped = 0;
  for (int i = 0; i < nv; i++) if (m->dof_damping[i] > 0) { damped = 1; break; }
  double* qacc = (double*)malloc(sizeof(double) * (nv + 1));
  if (!damped) memcpy(qacc, d->qacc, sizeof(double) * nv);
  else {
    /* implicit joint damping: (M + dt*diag(damping)) a = f_smooth + f_constraint */
    double* MM = (double*)malloc(sizeof(double) * (m->nM + 1));
    double* di = (double*)malloc(sizeof(double) * (nv + 1));
    memcpy(MM, d->qM, sizeof(double) * m->nM);
    for (int i = 0; i < nv; i++) MM[m->dof_Madr[i]] += m->opt_timestep * m->dof_damping[i];
    factor_sparse(m, MM, di);
    for (int i = 0; i < nv; i++) qacc[i] = d->qfrc_smooth[i] + d->qfrc_constraint[i];
    solve_sparse(m, MM, di, qacc);
    free(MM); free(di);
  }
  advance(m, d, d->act_dot, qacc, NULL);
  free(qacc);
}

static void o_rk4(const OModel* m, OData* d) {
  static const double A[3][3] = {{0.5, 0, 0}, {0, 0.5, 0}, {0, 0, 1}};
  static const double B[4] = {1.0 / 6, 1.0 / 3, 1.0 / 3, 1.0 / 6};
  static const double C[3] = {0.5, 0.5, 1.0};
  int nq = m->nq, nv = m->nv, na = m->na;
  double h = m->opt_timestep, t0 = d->time[0];
  double* buf = (double*)calloc((size_t)(nq + nv + na + 4 * (2 * nv + na) + 2 * nv + na + 8), 8);
  double *q0 = buf, *v0 = q0 + nq, *a0 = v0 + nv;
  double* F[4];
  F[0] = a0 + na;
  for (int i = 1; i < 4; i++) F[i] = F[i - 1] + 2 * nv + na;
  double* dX = F[3] + 2 * nv + na;
  memcpy(q0, d->qpos, 8 * nq); memcpy(v0, d->qvel, 8 * nv); memcpy(a0, d->act, 8 * na);
  memcpy(F[0], d->qvel, 8 * nv); memcpy(F[0] + nv, d->qacc, 8 * nv);
  memcpy(F[0] + 2 * nv, d->act_dot, 8 * na);
  for (int i = 1; i < 4; i++) {
    for (int k = 0; k < 2 * nv + na; k++) {
      dX[k] = 0;
      for (int j = 0; j < i; j++) dX[k] += A[i - 1][j] * F[j][k];
    }
    memcpy(d->qpos, q0, 8 * nq);
    integrate_pos(m, d->qpos, dX, h);
    for (int k = 0; k < nv; k++) d->qvel[k] = v0[k] + h * dX[nv + k];
    for (int k = 0; k < na; k++) d->act[k] = a0[k] + h * dX[2 * nv + k];
    d->time[0] = t0 + h * C[i - 1];
    forward_skip(m, d, 1);
    memcpy(F[i], d->qvel, 8 * nv); memcpy(F[i] + nv, d->qacc, 8 * nv);
    memcpy(F[i] + 2 * nv, d->act_dot, 8 * na);
  }
  for (int k = 0; k < 2 * nv + na; k++) {
    dX[k] = 0;
    for (int j = 0; j < 4; j++) dX[k] += B[j] * F[j][k];
  }
  memcpy(d->qpos, q0, 8 * nq); memcpy(d->qvel, v0, 8 * nv); memcpy(d->act, a0, 8 * na);
  d->time[0] = t0;
  advance(m, d, dX + 2 * nv, dX + nv, dX);
  free(buf);
}

static void check_acc(const OModel* m, OData* d) {
  if (check_array(m, d, d->qacc, m->nv, W_BADQACC)) o_forward(m, d);
}

void o_step(const OModel* m, OData* d) {
  check_array(m, d, d->qpos, m->nq, W_BADQPOS);
  check_array(m, d, d->qvel, m->nv, W_BADQVEL);
  o_forward(m, d);
  check_acc(m, d);
  if (m->opt_integrator == 1) o_rk4(m, d); else o_euler(m, d);
}

void o_step1(const OModel* m, OData* d) {
  check_array(m, d, d->qpos, m->nq, W_BADQPOS);
  check_array(m, d, d->qvel, m->nv, W_BADQVEL);
  fwd_position(m, d);
  o_sensors(m, d, STAGE_POS);
  o_energy_pos(m, d);
  fwd_velocity(m, d);
  o_sensors(m, d, STAGE_VEL);
  o_energy_vel(m, d);
}

void o_step2(const OModel* m, OData* d) {
  o_fwd_actuation(m, d);
  o_fwd_acceleration(m, d);
  o_fwd_constraint(m, d);
  o_sensors(m, d, STAGE_ACC);
  check_acc(m, d);
  o_euler(m, d);
}

/* engine.py:149-166 */
void o_physics_step(const OModel* m, OData* d) {
  if (m->opt_integrator == 0) o_step2(m, d); else o_step(m, d);
  o_step1(m, d);
}

/* mj_contactForce (wrapper/core.py:747-769): wrench of contact `ci` in the contact frame,
 * (normal, tangent, tangent | torsional, roll, roll) */
void o_contact_force(const OModel* m, const OData* d, int ci, double* f6) {
  if (ci < 0 || ci >= d->ncon[0]) { memset(f6, 0, 48); return; }
  contact_force(m, d, ci, f6);
}

/* mj_objectVelocity (wrapper/core.py:713-745): 6D velocity (angular; linear) of an object in
 * an object-centred frame, world-aligned or (flg_local) object-aligned.  objtype as mjtObj:
 * 1 body (inertial frame), 2 xbody, 5 geom, 6 site. */
void o_object_velocity(const OModel* m, const OData* d, int objtype, int id, double* res,
                       int flg_local) {
  const double *pos, *mat;
  int body;
  obj_frame(m, d, objtype, id, &pos, &mat, &body);
  const double* com = d->subtree_com + 3 * m->body_rootid[body];
  const double* cv = d->cvel + 6 * body;
  double dif[3] = {pos[0] - com[0], pos[1] - com[1], pos[2] - com[2]}, t3[3], vlin[3];
  cross3(t3, cv, dif);
  for (int c = 0; c < 3; c++) vlin[c] = cv[3 + c] + t3[c];
  if (flg_local) { matT_vec3(res, mat, cv); matT_vec3(res + 3, mat, vlin); }
  else { memcpy(res, cv, 24); memcpy(res + 3, vlin, 24); }
}

/* -------------------------------------------------- CPU-baseline driver */

typedef struct {
  const OModel* m; int lo, hi, nsteps, nsub, ctrl_per_step, nenv;
  double *qpos, *qvel, *act, *warm, *time; const double* ctrl;
  const double* hfield;   /* [nenv][nhfielddata] per-env elevation data, or NULL */
} BatchJob;

static void* batch_worker(void* arg) {
  BatchJob* j = (BatchJob*)arg;
  const OModel* m = j->m;
  OData* d = od_create(m);
  for (int e = j->lo; e < j->hi; e++) {
    o_reset(m, d);
    if (j->hfield)
      memcpy(d->hfield, j->hfield + (size_t)e * m->nhfielddata, 8 * (size_t)m->nhfielddata);
    memcpy(d->qpos, j->qpos + (size_t)e * m->nq, 8 * m->nq);
    memcpy(d->qvel, j->qvel + (size_t)e * m->nv, 8 * m->nv);
    if (m->na) memcpy(d->act, j->act + (size_t)e * m->na, 8 * m->na);
    memcpy(d->qacc_warmstart, j->warm + (size_t)e * m->nv, 8 * m->nv);
    d->time[0] = j->time[e];
    o_step1(m, d);   /* derived quantities in sync with the loaded state */
    for (int s = 0; s < j->nsteps; s++) {
      const double* c = j->ctrl + ((size_t)(j->ctrl_per_step ? s : 0) * j->nenv + e) * m->nu;
      memcpy(d->ctrl, c, 8 * m->nu);
      for (int k = 0; k < j->nsub; k++) o_physics_step(m, d);
    }
    memcpy(j->qpos + (size_t)e * m->nq, d->qpos, 8 * m->nq);
    memcpy(j->qvel + (size_t)e * m->nv, d->qvel, 8 * m->nv);
    if (m->na) memcpy(j->act + (size_t)e * m->na, d->act, 8 * m->na);
    memcpy(j->warm + (size_t)e * m->nv, d->qacc_warmstart, 8 * m->nv);
    j->time[e] = d->time[0];
  }
  od_free(d);
  return NULL;
}

void o_run_batch_hf(const OModel* m, int nenv, int nsteps, int nsub, double* qpos,
                    double* qvel, double* act, double* warm, double* time,
                    const double* ctrl, int ctrl_per_step, int nthreads,
                    const double* hfield) {
  if (nthreads < 1) nthreads = 1;
  if (nthreads > nenv) nthreads = nenv;
  pthread_t* th = (pthread_t*)malloc(sizeof(pthread_t) * nthreads);
  BatchJob* jobs = (BatchJob*)malloc(sizeof(BatchJob) * nthreads);
  for (int t = 0; t < nthreads; t++) {
    BatchJob j = {m, (int)((long)nenv * t / nthreads), (int)((long)nenv * (t + 1) / nthreads),
                  nsteps, nsub, ctrl_per_step, nenv, qpos, qvel, act, warm, time, ctrl,
                  hfield};
    jobs[t] = j;
    pthread_create(&th[t], NULL, batch_worker, &jobs[t]);
  }
  for (int t = 0; t < nthreads; t++) pthread_join(th[t], NULL);
  free(th); free(jobs);
}

void o_run_batch(const OModel* m, int nenv, int nsteps, int nsub, double* qpos,
                 double* qvel, double* act, double* warm, double* time,
                 const double* ctrl, int ctrl_per_step, int nthreads) {
  o_run_batch_hf(m, nenv, nsteps, nsub, qpos, qvel, act, warm, time, ctrl, ctrl_per_step,
                 nthreads, NULL);
}
