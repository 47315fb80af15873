// Shared host/device types of the batched physics backend.
//
// The model mirrors the subset of `mjModel` the suite tasks touch (field names
// follow mjModel so the reference's `named.model.*` indexing,
// /root/reference/dm_control/mujoco/index.py:113-173, keeps working); the data
// "image" mirrors `mjData` for ONE environment.  One warp owns one environment:
// it pulls the env's image from HBM into shared memory, advances it, and writes it
// back (see DESIGN.md "Data layout").
#ifndef BP_TYPES_H_
#define BP_TYPES_H_
#include <stdint.h>

#define BP_MINVAL 1e-15
#define BP_MAXVAL 1e10

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
       W_BADQVEL, W_BADQACC, W_BADCTRL, W_COUNT };
enum { SENS_TOUCH = 0, SENS_ACCELEROMETER = 1, SENS_VELOCIMETER = 2,
       SENS_GYRO = 3, SENS_FORCE = 4, SENS_TORQUE = 5, SENS_RANGEFINDER = 7,
       SENS_JOINTPOS = 8,
       SENS_JOINTVEL = 9, SENS_ACTUATORFRC = 14, SENS_FRAMEPOS = 23,
       SENS_FRAMEXAXIS = 25, SENS_FRAMEYAXIS = 26, SENS_FRAMEZAXIS = 27,
       SENS_SUBTREECOM = 32, SENS_SUBTREELINVEL = 33 };
enum { OBJ_BODY = 1, OBJ_XBODY = 2, OBJ_GEOM = 5, OBJ_SITE = 6 };
enum { STAGE_POS = 1, STAGE_VEL = 2, STAGE_ACC = 3 };
enum { BP_MODE_STEP = 0, BP_MODE_FORWARD = 1, BP_MODE_STEP1 = 2,
       BP_MODE_STEP2 = 3, BP_MODE_MJSTEP = 4 };

// ---- model fields that arrive in the blob (dm_control_b200/model.py) ----------
#define BP_MODEL_INT_SCALARS(X) \
  X(nq) X(nv) X(nu) X(na) X(nbody) X(njnt) X(ngeom) X(nsite) X(nM) X(npair) \
  X(nconmax) X(njmax) X(nsensor) X(nsensordata) X(nkey) X(ntendon) X(neq) \
  X(nhfield) X(nhfielddata) X(nhfpair) X(nspatial) \
  X(opt_integrator) X(opt_cone) X(opt_solver) X(opt_iterations) \
  X(opt_ls_iterations) X(opt_disableflags) X(opt_enableflags)
#define BP_MODEL_F64_SCALARS(X) \
  X(opt_timestep) X(opt_impratio) X(opt_tolerance) X(opt_ls_tolerance) \
  X(stat_meaninertia) X(opt_density) X(opt_viscosity)
#define BP_MODEL_INT_ARRAYS(X) \
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
#define BP_MODEL_F64_ARRAYS(X) \
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
  X(sensor_cutoff) X(wrap_prm) X(tendon_length0) X(tendon_invweight0) X(tendon_stiffness) X(tendon_damping) X(tendon_range) X(tendon_margin) \
  X(tendon_solref_lim) X(tendon_solimp_lim) X(eq_solref) \
  X(eq_solimp) X(eq_data) X(hfield_size) X(hfield_data)
// ---- arrays derived on the host when the model is created (bp_api.cu) ---------
//   body_level_*: bodies grouped by tree depth (parents complete before children)
//   body_child_*: children of each body, DESCENDING id (the accumulation order of a
//                 backward pass over body ids)
//   dof_depth:    number of ancestor dofs;  M_col: column dof of each qM entry
//   dof_level_*:  dofs grouped by depth
//   fric_dof / lim_jnt: rows that exist statically (friction loss) / candidates
//   tri_i, tri_j: (row, col) of the e-th lower-triangle entry of an nv x nv matrix
#define BP_MODEL_DERIVED_INT_ARRAYS(X) \
  X(body_level_adr) X(body_level_list) X(body_child_adr) X(body_child_list) \
  X(dof_depth) X(M_col) X(dof_level_adr) X(dof_level_list) X(fric_dof) \
  X(lim_jnt) X(tri_i) X(tri_j) X(dof_actadr) X(dof_actlist)
//   dof_actcoef: moment of actuator dof_actlist[k] on its dof (gear, times the tendon
//                coefficient for tendon transmissions)
#define BP_MODEL_DERIVED_F64_ARRAYS(X) X(dof_actcoef)

// Read-only model array: element reads go through the non-coherent path (ld.global.nc), which
// also tells the compiler that stores to an env's image cannot alias them.
template <class T>
struct BpRO {
  const T* p;
  BpRO() = default;
  __host__ __device__ BpRO(const T* q) : p(q) {}
#ifdef __CUDA_ARCH__
  __device__ __forceinline__ T operator[](int i) const { return __ldg(p + i); }
  __device__ __forceinline__ T operator[](long long i) const { return __ldg(p + i); }
#else
  T operator[](int i) const { return p[i]; }
  T operator[](long long i) const { return p[i]; }
#endif
  __host__ __device__ operator const T*() const { return p; }
  __host__ __device__ BpRO operator+(long long i) const { return BpRO(p + i); }
  __host__ __device__ BpRO operator+(int i) const { return BpRO(p + i); }
};

struct BpModel {
#define X(n) int n;
  BP_MODEL_INT_SCALARS(X)
#undef X
#define X(n) double n;
  BP_MODEL_F64_SCALARS(X)
#undef X
#define X(n) BpRO<int> n;
  BP_MODEL_INT_ARRAYS(X)
  BP_MODEL_DERIVED_INT_ARRAYS(X)
#undef X
#define X(n) BpRO<double> n;
  BP_MODEL_F64_ARRAYS(X)
  BP_MODEL_DERIVED_F64_ARRAYS(X)
#undef X
  const unsigned long long* body_dofmask;  // bit i set <=> dof i moves the body
  int nlevel, ndoflevel, nfric, nlim, ntri, nvp, neq_active;  // nvp: odd row stride >= nv
  int any_damping, need_subtreevel, need_rnepost_stage;  // sensor prerequisites
  int nhinge, nsens_ft, nsens_rf;   // hinge joints; force + torque sensors; rangefinders
  int lean_static;   // structure fits the LEAN instantiation (Ctx<G, S, 1>, bp_math.cuh)
  int mid_static;    // structure fits the MID instantiation (Ctx<G, S, 2>)
};

// ---- per-environment data image ---------------------------------------------
// (name, element count).  F64 fields first, then I32 fields, then SCRATCH (shared
// memory only, never written to HBM).  njmax/nconmax are the batch's, not the
// model's (bp_batch_create).
#define BP_DATA_F64(X) \
  X(qpos, m.nq) X(qvel, m.nv) X(act, m.na) X(ctrl, m.nu) X(qacc, m.nv) \
  X(qacc_warmstart, m.nv) X(act_dot, m.na) X(qfrc_applied, m.nv) X(time, 1) \
  X(energy, 2) X(xpos, 3 * m.nbody) X(xquat, 4 * m.nbody) X(xmat, 9 * m.nbody) \
  X(xipos, 3 * m.nbody) X(ximat, 9 * m.nbody) X(xanchor, 3 * m.njnt) \
  X(xaxis, 3 * m.njnt) X(geom_xpos, 3 * m.ngeom) X(geom_xmat, 9 * m.ngeom) \
  X(site_xpos, 3 * m.nsite) X(site_xmat, 9 * m.nsite) \
  X(subtree_com, 3 * m.nbody) X(cinert, 10 * m.nbody) X(cdof, 6 * m.nv) \
  X(cdof_dot, 6 * m.nv) X(cvel, 6 * m.nbody) X(cacc, 6 * m.nbody) \
  X(cfrc_int, 6 * m.nbody) X(cfrc_ext, 6 * m.nbody) X(crb, 10 * m.nbody) \
  X(subtree_linvel, 3 * m.nbody) X(qM, m.nM) X(qLD, m.nM) X(qLDiagInv, m.nv) \
  X(qfrc_bias, m.nv) X(qfrc_passive, m.nv) X(qfrc_actuator, m.nv) \
  X(qfrc_smooth, m.nv) X(qacc_smooth, m.nv) X(qfrc_constraint, m.nv) \
  X(actuator_length, m.nu) X(actuator_velocity, m.nu) X(actuator_force, m.nu) \
  X(ten_length, m.ntendon) X(ten_velocity, m.ntendon) X(ten_J, m.nspatial * m.nv) \
  X(sensordata, m.nsensordata) X(efc_J, njmax * m.nvp) X(efc_pos, njmax) \
  X(efc_margin, njmax) X(efc_frictionloss, njmax) X(efc_diagApprox, njmax) \
  X(efc_KBIP, 4 * njmax) X(efc_D, njmax) X(efc_R, njmax) X(efc_vel, njmax) \
  X(efc_aref, njmax) X(efc_force, njmax) X(contact_dist, nconmax) \
  X(contact_pos, 3 * nconmax) X(contact_frame, 9 * nconmax) \
  X(contact_includemargin, nconmax) X(contact_friction, 5 * nconmax) \
  X(contact_solref, 2 * nconmax) X(contact_solimp, 5 * nconmax) \
  X(contact_mu, nconmax) X(solver_stat, 2)
#define BP_DATA_I32(X) \
  X(efc_type, njmax) X(efc_id, njmax) X(efc_state, njmax) \
  X(contact_geom1, nconmax) X(contact_geom2, nconmax) X(contact_dim, nconmax) \
  X(contact_efc_address, nconmax) X(contact_exclude, nconmax) X(ncon, 1) \
  X(nefc, 1) X(ne, 1) X(nf, 1) X(solver_iter, 1) X(warning, 8)
#define BP_DATA_SCRATCH(X) \
  X(s_Ma, m.nv) X(s_jar, njmax) X(s_grad, m.nv) X(s_Mgrad, m.nv) \
  X(s_search, m.nv) X(s_Mv, m.nv) X(s_jv, njmax) X(s_H, m.nv * m.nvp + m.nv) \
  X(s_Md, m.nv * m.nvp) X(s_Hdinv, m.nv) X(s_MM, m.nM) X(s_MMdinv, m.nv) \
  X(s_rdof, njmax) X(s_rw, njmax) X(s_tmp, 2 * m.nv + m.na) X(s_cacc, 6 * m.nbody) X(s_cfrc, 6 * m.nbody) X(s_rk, m.nq + m.nv + m.na + 5 * (2 * m.nv + m.na) + 1) \
  X(s_AR, m.opt_solver == 0 ? njmax * njmax : 0)   /* PGS only: efc_AR = J M^-1 J^T + R */

struct BpLayout {   // offsets in doubles from the start of the env's image
#define X(n, c) int n;
  BP_DATA_F64(X)
  BP_DATA_I32(X)     // offsets in int32 units from `ints`
  BP_DATA_SCRATCH(X)
#undef X
  int ints;          // start of the int32 region (in doubles)
  int persist;       // doubles written to / read from HBM per env (incl. ints)
  int total;         // doubles of shared memory per env (persist + scratch)
  int njmax, nconmax;
};

// suite task descriptor: kind + ids / constants resolved by name on the host
struct BpTask {
  int kind;
  int ip[24];
  double fp[8];
};

struct BpParams {
  BpModel m;
  BpLayout L;
  double* image;     // [nenv][stride] (or [nenv/32][stride][32] when interleave == 32)
  const unsigned char* mask;   // optional [nenv]; 0 = leave the env untouched
  long long nenv;
  long long stride;  // doubles per env in `image` (persist for the smem variant, total otherwise)
  int interleave;    // 1: env images contiguous; 32: images of a warp's 32 envs interleaved
  int lanes;         // lanes per environment (group size G)
  int mode, nsub, disable_actuation;
  int hot_doubles;   // per-env shared-memory arena of the global-memory shapes (0 = none)
  // height-field elevation, one copy per env (float32 as in mjModel.hfield_data): the
  // Escape task writes a different terrain into every episode (quadruped.py:371-382)
  const float* hfield;     // [nenv][nhfielddata], null when the model has no height field
  // per-env copies of model frames that tasks move per episode (reacher.py:96-101 target
  // geom, finger.py turn target site); null: every env reads the model's
  const double* env_geom_pos;   // [nenv][3 * ngeom]
  const double* env_site_pos;   // [nenv][3 * nsite]
  const double* env_body_pos;   // [nenv][3 * nbody]   (manipulator.py:218-224 target body)
  const double* env_body_quat;  // [nenv][4 * nbody]
  const double* env_wrap_prm;   // [nenv][env_wrap_n]   (point_mass.py:85-101 randomize_gains)
  int env_wrap_n;
  // fused Environment.step (bp_env_step): dense [nenv][nu] controls scattered into the
  // images before the first sub-step, task observation / reward written after the last
  const double* ctrl_in;   // null: images keep their ctrl
  double* obs_out;         // null: no observation
  double* rew_out;         // null: no epilogue at all
  int task_nobs;
  BpTask task;
  // End-to-end step (bp_env_step_host) with pinned, mapped host result buffers: every block
  // copies the rows of ITS environments from obs_out / rew_out to these host-mapped
  // addresses with warp-contiguous stores as soon as it is done, so the PCIe transfer of the
  // results overlaps the blocks still computing and no D2H copy follows the kernel.
  double* obs_host;        // null: results stay on the device
  double* rew_host;
};
// arena layout: packed Hessian / Cholesky factor + rhs row (ntri + nv), inverse pivots (nv), Mgrad (nv)



#endif  // BP_TYPES_H_
