// C ABI of libbphys.so (include/bphys.h): model upload, batch allocation, launches.
// Host side only; the kernels live in bp_step.cu / bp_task.cu.
#include <cuda_runtime.h>
#include <stdio.h>
#include <string.h>
#include <math.h>

#include <map>
#include <string>
#include <vector>

#include "../../include/bphys.h"
#include "bp_types.h"

extern "C" int bp_launch_step_g32(const BpParams* P, cudaStream_t stream);
extern "C" int bp_launch_step_g16(const BpParams* P, cudaStream_t stream);
extern "C" int bp_launch_step_g8(const BpParams* P, cudaStream_t stream);
extern "C" int bp_launch_step_g4(const BpParams* P, cudaStream_t stream);
extern "C" int bp_launch_step_g2(const BpParams* P, cudaStream_t stream);
extern "C" int bp_launch_step_g1(const BpParams* P, cudaStream_t stream);
extern "C" int bp_launch_step_g16_lean(const BpParams* P, cudaStream_t stream);
extern "C" int bp_launch_step_g8_lean(const BpParams* P, cudaStream_t stream);
extern "C" int bp_launch_step_g4_lean(const BpParams* P, cudaStream_t stream);
extern "C" int bp_launch_step_g16_mid(const BpParams* P, cudaStream_t stream);
extern "C" int bp_launch_step_g8_mid(const BpParams* P, cudaStream_t stream);
extern "C" int bp_launch_step_static(const BpParams* P, unsigned long long hash, int plain,
                                     cudaStream_t stream);
extern "C" int bp_launch_field_copy(const BpParams* P, long long off, int count, int is_int,
                                    void* dense, long long env0, long long n, int to_image,
                                    cudaStream_t stream);
extern "C" int bp_launch_reset(const BpParams* P, const double* key_qpos,
                               const double* key_qvel, const double* key_act,
                               double key_time, cudaStream_t stream);
extern "C" int bp_launch_task(const BpParams* P, const BpTask* T, double* obs,
                              double* reward, cudaStream_t stream);
extern "C" int bp_task_nobs_impl(const BpModel* m, int task);

static thread_local std::string g_err;
static int fail(int code, const std::string& msg) { g_err = msg; return code; }
// for the library's other translation units (bp_noalloc.cu); not part of include/bphys.h
extern "C" void bp_set_last_error_(const char* msg) { g_err = msg ? msg : ""; }
#define CUDA_OK(expr)                                                        \
  do {                                                                       \
    cudaError_t e_ = (expr);                                                 \
    if (e_ != cudaSuccess)                                                   \
      return fail(BP_ECUDA, std::string(#expr) + ": " + cudaGetErrorString(e_)); \
  } while (0)

struct BlobEntry {
  char name[40];
  int32_t dtype, ndim, dims[4];
  int64_t offset, count;
};

struct HostField {
  int dtype;                     // 0 f64, 1 i32
  std::vector<double> f;
  std::vector<int32_t> i;
  size_t dev_off = 0;            // byte offset inside the device buffer
  size_t nbytes() const { return dtype == 0 ? f.size() * 8 : i.size() * 4; }
  const void* data() const { return dtype == 0 ? (const void*)f.data() : (const void*)i.data(); }
};

struct FieldInfo { std::string name; int64_t offset_bytes, count; int dtype; };

struct bp_model {
  int device = 0;
  std::map<std::string, HostField> fields;   // blob fields + derived arrays
  std::vector<unsigned long long> dofmask;
  char* dev = nullptr;                        // one device buffer for every array
  size_t dev_bytes = 0;
  BpModel M;                                  // scalars + device pointers
  unsigned long long struct_hash = 0;         // of everything a static instance bakes in
};

static const std::vector<int32_t>& I(bp_model* m, const char* n) { return m->fields[n].i; }
static const std::vector<double>& F(bp_model* m, const char* n) { return m->fields[n].f; }
static void put_i(bp_model* m, const char* n, const std::vector<int32_t>& v) {
  HostField h; h.dtype = 1; h.i = v; if (h.i.empty()) h.i.push_back(0); m->fields[n] = h;
}

static void put_f(bp_model* m, const char* n, const std::vector<double>& v) {
  HostField h; h.dtype = 0; h.f = v; if (h.f.empty()) h.f.push_back(0); m->fields[n] = h;
}

// scalars are refreshed from the host fields (after bp_model_set_field too)
static int refresh_scalars(bp_model* m) {
  BpModel& M = m->M;
#define X(n) { auto it = m->fields.find(#n); \
    if (it == m->fields.end()) return fail(BP_EINVAL, "model blob lacks field " #n); \
    M.n = it->second.dtype == 1 ? it->second.i[0] : (int)it->second.f[0]; }
  BP_MODEL_INT_SCALARS(X)
#undef X
#define X(n) { auto it = m->fields.find(#n); \
    if (it == m->fields.end()) return fail(BP_EINVAL, "model blob lacks field " #n); \
    M.n = it->second.dtype == 0 ? it->second.f[0] : (double)it->second.i[0]; }
  BP_MODEL_F64_SCALARS(X)
#undef X
  M.any_damping = 0;
  for (double d : F(m, "dof_damping")) if (d > 0) M.any_damping = 1;
  if (M.nv == 0) M.any_damping = 0;
  return BP_OK;
}

static void compute_struct_hash(bp_model* m);

static int derive(bp_model* m) {
  BpModel& M = m->M;
  const int nb = M.nbody, nv = M.nv;
  if (nv > 64) return fail(BP_EINVAL, "models with nv > 64 are not supported");
  auto& par = I(m, "body_parentid");
  std::vector<int32_t> depth(nb, 0);
  int maxd = 0;
  for (int b = 1; b < nb; b++) { depth[b] = depth[par[b]] + 1; if (depth[b] > maxd) maxd = depth[b]; }
  M.nlevel = maxd + 1;
  std::vector<int32_t> ladr(M.nlevel + 1, 0), llist;
  for (int l = 0; l < M.nlevel; l++) {
    ladr[l] = (int)llist.size();
    for (int b = 0; b < nb; b++) if (depth[b] == l) llist.push_back(b);
  }
  ladr[M.nlevel] = (int)llist.size();
  put_i(m, "body_level_adr", ladr); put_i(m, "body_level_list", llist);
  std::vector<int32_t> cadr(nb + 1, 0), clist;
  for (int b = 0; b < nb; b++) {
    cadr[b] = (int)clist.size();
    for (int ch = nb - 1; ch > 0; ch--) if (par[ch] == b && ch != b) clist.push_back(ch);
  }
  cadr[nb] = (int)clist.size();
  put_i(m, "body_child_adr", cadr); put_i(m, "body_child_list", clist);
  auto& dpar = I(m, "dof_parentid");
  auto& madr = I(m, "dof_Madr");
  std::vector<int32_t> ddepth(nv > 0 ? nv : 1, 0), mcol(M.nM > 0 ? M.nM : 1, 0);
  int maxdd = 0;
  for (int i = 0; i < nv; i++) {
    int a = madr[i], d = 0;
    for (int j = i; j >= 0; j = dpar[j]) { mcol[a++] = j; d++; }
    ddepth[i] = d - 1;
    if (d - 1 > maxdd) maxdd = d - 1;
  }
  M.ndoflevel = nv > 0 ? maxdd + 1 : 0;
  put_i(m, "dof_depth", ddepth); put_i(m, "M_col", mcol);
  std::vector<int32_t> dladr(M.ndoflevel + 1, 0), dllist;
  for (int l = 0; l < M.ndoflevel; l++) {
    dladr[l] = (int)dllist.size();
    for (int i = 0; i < nv; i++) if (ddepth[i] == l) dllist.push_back(i);
  }
  dladr[M.ndoflevel] = (int)dllist.size();
  put_i(m, "dof_level_adr", dladr); put_i(m, "dof_level_list", dllist);
  // dof ancestry masks per body
  auto& bdofadr = I(m, "body_dofadr");
  auto& bdofnum = I(m, "body_dofnum");
  m->dofmask.assign(nb, 0ull);
  for (int b = 1; b < nb; b++) {
    unsigned long long mk = m->dofmask[par[b]];
    for (int k = 0; k < bdofnum[b]; k++) mk |= 1ull << (bdofadr[b] + k);
    m->dofmask[b] = mk;
  }
  std::vector<int32_t> fric, lim;
  auto& fl = F(m, "dof_frictionloss");
  for (int i = 0; i < nv; i++) if (fl[i] > 0) fric.push_back(i);
  auto& jl = I(m, "jnt_limited");
  auto& jt = I(m, "jnt_type");
  for (int j = 0; j < M.njnt; j++)
    if (jl[j] && (jt[j] == JNT_SLIDE || jt[j] == JNT_HINGE)) lim.push_back(j);
  M.nfric = (int)fric.size(); M.nlim = (int)lim.size();
  put_i(m, "fric_dof", fric); put_i(m, "lim_jnt", lim);
  std::vector<int32_t> ti, tj;
  for (int i = 0; i < nv; i++) for (int j = 0; j <= i; j++) { ti.push_back(i); tj.push_back(j); }
  M.ntri = (int)ti.size();
  put_i(m, "tri_i", ti); put_i(m, "tri_j", tj);
  M.nvp = nv | 1;
  // actuators attached to each dof (joint transmission on scalar joints)
  auto& trnid = I(m, "actuator_trnid");
  auto& trntype = I(m, "actuator_trntype");
  auto& jdof = I(m, "jnt_dofadr");
  std::vector<int32_t> aadr(nv + 1, 0), alist;
  std::vector<double> acoef;
  auto& gear = F(m, "actuator_gear");
  auto& tadr = I(m, "tendon_adr");
  auto& tnum = I(m, "tendon_num");
  auto& wobj = I(m, "wrap_objid");
  auto& wprm = F(m, "wrap_prm");
  for (int u = 0; u < M.nu; u++)
    if (trntype[u] != 0 && trntype[u] != 3)
      return fail(BP_EINVAL, "only joint and tendon transmissions are supported");
  for (int d = 0; d < nv; d++) {
    aadr[d] = (int)alist.size();
    for (int u = 0; u < M.nu; u++) {
      if (trntype[u] == 0) {
        if (jdof[trnid[2 * u]] == d) { alist.push_back(u); acoef.push_back(gear[6 * u]); }
      } else {
        int t = trnid[2 * u];
        double cf = 0; bool hit = false;
        for (int w = tadr[t]; w < tadr[t] + tnum[t]; w++)
          if (jdof[wobj[w]] == d) { cf += gear[6 * u] * wprm[w]; hit = true; }
        if (hit) { alist.push_back(u); acoef.push_back(cf); }
      }
    }
  }
  aadr[nv] = (int)alist.size();
  put_f(m, "dof_actcoef", acoef);
  M.neq_active = 0;
  if (M.neq > 0) { auto& ea = I(m, "eq_active"); for (int e = 0; e < M.neq; e++) M.neq_active += ea[e] != 0; }
  put_i(m, "dof_actadr", aadr); put_i(m, "dof_actlist", alist);
  M.need_subtreevel = 0; M.need_rnepost_stage = 0;
  M.nhinge = 0; M.nsens_ft = 0; M.nsens_rf = 0;
  for (int j = 0; j < M.njnt; j++) M.nhinge += jt[j] == JNT_HINGE;
  auto& st = I(m, "sensor_type");
  auto& ss = I(m, "sensor_needstage");
  for (int s = 0; s < M.nsensor; s++) {
    if (st[s] == SENS_SUBTREELINVEL) M.need_subtreevel = 1;
    if (st[s] == SENS_FORCE || st[s] == SENS_TORQUE) M.nsens_ft++;
    if (st[s] == SENS_RANGEFINDER) M.nsens_rf++;
    if (st[s] == SENS_ACCELEROMETER || st[s] == SENS_FORCE || st[s] == SENS_TORQUE)
      M.need_rnepost_stage = ss[s];
  }
  // LEAN shape: plane / sphere / capsule geoms of equal priority, no tendons, equalities,
  // height fields, rangefinders or actuator dynamics (cone and fluid are checked per launch)
  M.lean_static = M.ntendon == 0 && M.neq == 0 && M.nhfield == 0 && M.nsens_rf == 0 && M.na == 0;
  auto& gt = I(m, "geom_type");
  auto& gpr = I(m, "geom_priority");
  for (int g = 0; g < M.ngeom; g++)
    if ((gt[g] != GEOM_PLANE && gt[g] != GEOM_SPHERE && gt[g] != GEOM_CAPSULE) || gpr[g] != 0)
      M.lean_static = 0;
  // MID shape: everything but spatial tendons, tendon limits / springs, box geoms, priorities
  M.mid_static = M.nspatial == 0;
  for (int g = 0; g < M.ngeom; g++)
    if (gt[g] == GEOM_BOX || gpr[g] != 0) M.mid_static = 0;
  if (M.ntendon > 0) {
    auto& tl = I(m, "tendon_limited");
    auto& tk = F(m, "tendon_stiffness");
    auto& tb = F(m, "tendon_damping");
    for (int t = 0; t < M.ntendon; t++)
      if (tl[t] != 0 || tk[t] != 0 || tb[t] != 0) M.mid_static = 0;
  }
  compute_struct_hash(m);
  return BP_OK;
}

// What the kernels implement.  Everything else must fail loudly instead of running other
// physics (the reference surfaces engine errors, wrapper/core_test.py:97-103).
//   pairs: the narrowphase table of csrc/bp_collide.cuh / bp_hfield.cuh
#ifndef BP_HAVE_BOX_PAIRS
#define BP_HAVE_BOX_PAIRS 1      // sphere / capsule / box against box
#endif
#ifndef BP_HAVE_PGS_CG
#define BP_HAVE_PGS_CG 1         // opt_solver 0 (PGS) and 1 (CG)
#endif
static bool pair_supported(int t1, int t2) {
  if (t1 > t2) { int t = t1; t1 = t2; t2 = t; }
  if (!BP_HAVE_BOX_PAIRS && t2 == GEOM_BOX && t1 != GEOM_PLANE) return false;
  if (t1 == GEOM_PLANE)
    return t2 == GEOM_SPHERE || t2 == GEOM_CAPSULE || t2 == GEOM_ELLIPSOID ||
           t2 == GEOM_CYLINDER || t2 == GEOM_BOX;
  if (t1 == GEOM_HFIELD)
    return t2 == GEOM_SPHERE || t2 == GEOM_CAPSULE || t2 == GEOM_ELLIPSOID || t2 == GEOM_CYLINDER;
  if (t1 == GEOM_SPHERE)
    return t2 == GEOM_SPHERE || t2 == GEOM_CAPSULE || t2 == GEOM_ELLIPSOID ||
           t2 == GEOM_CYLINDER || t2 == GEOM_BOX;
  if (t1 == GEOM_CAPSULE)
    return t2 == GEOM_CAPSULE || t2 == GEOM_ELLIPSOID || t2 == GEOM_CYLINDER || t2 == GEOM_BOX;
  if (t1 == GEOM_CYLINDER) return t2 == GEOM_CYLINDER;
  if (t1 == GEOM_BOX) return t2 == GEOM_BOX;
  return false;
}

static int validate(bp_model* m) {
  const BpModel& M = m->M;
  static const char* tn[] = {"plane", "hfield", "sphere", "capsule", "ellipsoid", "cylinder",
                             "box", "mesh"};
  auto& gt = I(m, "geom_type");
  for (int g = 0; g < M.ngeom; g++)
    if (gt[g] < 0 || gt[g] >= GEOM_MESH)
      return fail(BP_EINVAL, "geom " + std::to_string(g) + ": mesh geoms are not supported");
  if (M.opt_solver < 0 || M.opt_solver > 2)
    return fail(BP_EINVAL, "opt_solver must be 0 (PGS), 1 (CG) or 2 (Newton)");
  if (!BP_HAVE_PGS_CG && M.opt_solver != 2)
    return fail(BP_EINVAL, "only the Newton solver is implemented (opt_solver = PGS / CG)");
  if (M.opt_solver == 0 && M.opt_cone != 0)
    return fail(BP_EINVAL, "the PGS solver is implemented for pyramidal cones only");
  auto ns = m->fields.find("opt_noslip_iterations");
  if (ns != m->fields.end() && !ns->second.i.empty() && ns->second.i[0] > 0)
    return fail(BP_EINVAL, "noslip_iterations > 0 is not supported (the noslip post-solver is "
                           "not implemented)");
  // colliding pairs without a narrowphase: only acceptable while contacts cannot occur
  const bool contacts_on = !(M.opt_disableflags & (DSBL_CONTACT | DSBL_CONSTRAINT));
  if (contacts_on) {
    auto& p1 = I(m, "pair_geom1");
    auto& p2 = I(m, "pair_geom2");
    auto& pc = I(m, "pair_parentchild");
    const bool filt = !(M.opt_disableflags & DSBL_FILTERPARENT);
    for (int p = 0; p < M.npair; p++) {
      if (filt && pc[p]) continue;      // parent-child pairs are filtered out at run time
      const int t1 = gt[p1[p]], t2 = gt[p2[p]];
      if (!pair_supported(t1, t2))
        return fail(BP_EINVAL, std::string("geom pair (") + std::to_string(p1[p]) + ", " +
                    std::to_string(p2[p]) + ") of types " + tn[t1] + " - " + tn[t2] +
                    " has no narrowphase: exclude the pair, or disable contacts");
    }
  }
  return BP_OK;
}

// ---- static (ahead-of-time specialised) instances -------------------------------------
// A static instance (csrc/static/<name>.cuh, tools/gen_static_model.py) bakes in the model's
// structure: the structural int scalars, every int array (blob and derived), body_dofmask and
// the derived counts.  struct_hash identifies that structure; a launch uses a static instance
// only while the live model still hashes to the instance's value (run-time edits of e.g.
// dof_frictionloss or jnt_limited change derived arrays and hence the hash).
static const char* kStaticScalars[] = {
    "nq", "nv", "nu", "na", "nbody", "njnt", "ngeom", "nsite", "nM", "npair", "nsensor",
    "nsensordata", "nkey", "ntendon", "neq", "nhfield", "nhfielddata", "nhfpair", "nspatial"};

static void hash_mix(unsigned long long& h, const void* data, size_t n) {
  const unsigned char* p = (const unsigned char*)data;
  for (size_t i = 0; i < n; i++) { h ^= p[i]; h *= 1099511628211ull; }
}

static std::vector<std::string> static_int_arrays() {
  std::vector<std::string> v;
#define X(n) v.push_back(#n);
  BP_MODEL_INT_ARRAYS(X)
  BP_MODEL_DERIVED_INT_ARRAYS(X)
#undef X
  return v;
}

static void compute_struct_hash(bp_model* m) {
  unsigned long long h = 1469598103934665603ull;
  const BpModel& M = m->M;
  for (const char* n : kStaticScalars) {
    int v = m->fields[n].dtype == 1 ? m->fields[n].i[0] : (int)m->fields[n].f[0];
    hash_mix(h, &v, 4);
  }
  for (auto& n : static_int_arrays()) {
    auto& a = m->fields[n].i;
    int cnt = (int)a.size();
    hash_mix(h, &cnt, 4);
    hash_mix(h, a.data(), a.size() * 4);
  }
  hash_mix(h, m->dofmask.data(), m->dofmask.size() * 8);
  int derived[] = {M.nlevel, M.ndoflevel, M.nfric, M.nlim, M.ntri, M.nvp, M.neq_active,
                   M.any_damping, M.need_subtreevel, M.need_rnepost_stage, M.nhinge, M.nsens_ft,
                   M.nsens_rf, M.lean_static, M.mid_static};
  hash_mix(h, derived, sizeof derived);
  m->struct_hash = h;
}

static int upload(bp_model* m) {
  size_t off = 0;
  for (auto& kv : m->fields) { kv.second.dev_off = off; off += (kv.second.nbytes() + 15) / 16 * 16; }
  size_t mask_off = off;
  off += (m->dofmask.size() * 8 + 15) / 16 * 16 + 16;
  if (!m->dev || off > m->dev_bytes) {
    if (m->dev) cudaFree(m->dev);
    m->dev = nullptr;
    CUDA_OK(cudaMalloc((void**)&m->dev, off));
    m->dev_bytes = off;
  }
  std::vector<char> stage(off, 0);
  for (auto& kv : m->fields) memcpy(stage.data() + kv.second.dev_off, kv.second.data(), kv.second.nbytes());
  if (!m->dofmask.empty()) memcpy(stage.data() + mask_off, m->dofmask.data(), m->dofmask.size() * 8);
  CUDA_OK(cudaMemcpy(m->dev, stage.data(), off, cudaMemcpyHostToDevice));
  BpModel& M = m->M;
#define X(n) M.n = (const int*)(m->dev + m->fields[#n].dev_off);
  BP_MODEL_INT_ARRAYS(X)
  BP_MODEL_DERIVED_INT_ARRAYS(X)
#undef X
#define X(n) M.n = (const double*)(m->dev + m->fields[#n].dev_off);
  BP_MODEL_F64_ARRAYS(X)
  BP_MODEL_DERIVED_F64_ARRAYS(X)
#undef X
  M.body_dofmask = (const unsigned long long*)(m->dev + mask_off);
  return BP_OK;
}

extern "C" int bp_version(void) { return 100; }
extern "C" const char* bp_last_error(void) { return g_err.c_str(); }
extern "C" int bp_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
  return n;
}

static void make_layout(const BpModel& m, int njmax, int nconmax, BpLayout* L,
                        std::vector<FieldInfo>* fields);

// Host-only: parses a model blob, derives what bp_model_create derives and writes the C++
// header of a static instance (no CUDA call: runs in the build container).
extern "C" int bp_static_header(const void* blob, int64_t nbytes, int njmax, int nconmax,
                                const char* name, const char* path) {
  if (!blob || nbytes < 16 || memcmp(blob, "BPMODEL1", 8) != 0 || !name || !path)
    return fail(BP_EINVAL, "bad arguments to bp_static_header");
  const char* base = (const char*)blob;
  int nfields = *(const int32_t*)(base + 8);
  const BlobEntry* ent = (const BlobEntry*)(base + 16);
  bp_model mm;
  bp_model* m = &mm;
  for (int k = 0; k < nfields; k++) {
    HostField h;
    h.dtype = ent[k].dtype;
    if (h.dtype == 0) h.f.assign((const double*)(base + ent[k].offset), (const double*)(base + ent[k].offset) + ent[k].count);
    else h.i.assign((const int32_t*)(base + ent[k].offset), (const int32_t*)(base + ent[k].offset) + ent[k].count);
    if (h.dtype == 0 && h.f.empty()) h.f.push_back(0);
    if (h.dtype == 1 && h.i.empty()) h.i.push_back(0);
    char nm[41]; memcpy(nm, ent[k].name, 40); nm[40] = 0;
    m->fields[nm] = h;
  }
  if (refresh_scalars(m) != BP_OK || derive(m) != BP_OK) return BP_EINVAL;
  const BpModel& M = m->M;
  if (njmax <= 0) njmax = M.njmax;
  if (nconmax <= 0) nconmax = M.nconmax;
  BpLayout L;
  make_layout(M, njmax, nconmax, &L, nullptr);
  FILE* f = fopen(path, "w");
  if (!f) return fail(BP_EINVAL, std::string("cannot write ") + path);
  fprintf(f, "// GENERATED by tools/gen_static_model.py (bp_static_header) -- do not edit.\n");
  fprintf(f, "// Static instance \"%s\": structure of the compiled model and the data layout for\n", name);
  fprintf(f, "// njmax = %d, nconmax = %d as compile-time constants (csrc/bp_math.cuh Ctx<.., MT>).\n", njmax, nconmax);
  fprintf(f, "#ifndef BP_STATIC_%s_CUH_\n#define BP_STATIC_%s_CUH_\n#include \"../bp_types.h\"\n", name, name);
  fprintf(f, "namespace bp_static_%s {\n", name);
  auto arrays = static_int_arrays();
  for (auto& n : arrays) {
    auto& a = m->fields[n].i;
    fprintf(f, "__device__ constexpr int %s_[%d] = {", n.c_str(), (int)a.size());
    for (size_t i = 0; i < a.size(); i++) fprintf(f, "%s%d", i ? ", " : "", a[i]);
    fprintf(f, "};\n");
  }
  fprintf(f, "__device__ constexpr unsigned long long body_dofmask_[%d] = {", (int)(m->dofmask.empty() ? 1 : m->dofmask.size()));
  for (size_t i = 0; i < m->dofmask.size(); i++) fprintf(f, "%s%lluull", i ? ", " : "", m->dofmask[i]);
  if (m->dofmask.empty()) fprintf(f, "0ull");
  fprintf(f, "};\n");
  fprintf(f, "struct Layout {\n");
#define X(n, c) fprintf(f, "  static constexpr int %s = %d;\n", #n, L.n);
  BP_DATA_F64(X)
  BP_DATA_I32(X)
  BP_DATA_SCRATCH(X)
#undef X
  fprintf(f, "  static constexpr int ints = %d, persist = %d, total = %d, njmax = %d, nconmax = %d;\n};\n",
          L.ints, L.persist, L.total, L.njmax, L.nconmax);
  fprintf(f, "struct Model {\n");
  fprintf(f, "#define X(n) BpRO<double> n;\n  BP_MODEL_F64_ARRAYS(X) BP_MODEL_DERIVED_F64_ARRAYS(X)\n#undef X\n");
  fprintf(f, "#define X(n) double n;\n  BP_MODEL_F64_SCALARS(X)\n#undef X\n");
  fprintf(f, "  int opt_integrator, opt_cone, opt_solver, opt_iterations, opt_ls_iterations, opt_disableflags, opt_enableflags;\n");
  fprintf(f, "  int nconmax, njmax;\n");
  for (const char* n : kStaticScalars) {
    int v = m->fields[n].dtype == 1 ? m->fields[n].i[0] : (int)m->fields[n].f[0];
    fprintf(f, "  static constexpr int %s = %d;\n", n, v);
  }
  fprintf(f, "  static constexpr int nlevel = %d, ndoflevel = %d, nfric = %d, nlim = %d, ntri = %d, nvp = %d, neq_active = %d;\n",
          M.nlevel, M.ndoflevel, M.nfric, M.nlim, M.ntri, M.nvp, M.neq_active);
  fprintf(f, "  static constexpr int any_damping = %d, need_subtreevel = %d, need_rnepost_stage = %d, nhinge = %d, nsens_ft = %d, nsens_rf = %d, lean_static = %d, mid_static = %d;\n",
          M.any_damping, M.need_subtreevel, M.need_rnepost_stage, M.nhinge, M.nsens_ft, M.nsens_rf, M.lean_static, M.mid_static);
  for (auto& n : arrays) {
    fprintf(f, "  struct A_%s { __device__ __forceinline__ int operator[](int i) const { return %s_[i]; } } %s;\n",
            n.c_str(), n.c_str(), n.c_str());
  }
  fprintf(f, "  struct A_body_dofmask { __device__ __forceinline__ unsigned long long operator[](int i) const { return body_dofmask_[i]; } } body_dofmask;\n");
  fprintf(f, "  __device__ __forceinline__ explicit Model(const BpModel& d)\n      : ");
  bool first = true;
#define X(n) fprintf(f, "%s%s(d.%s)", first ? "" : ", ", #n, #n); first = false;
  BP_MODEL_F64_ARRAYS(X)
  BP_MODEL_DERIVED_F64_ARRAYS(X)
  BP_MODEL_F64_SCALARS(X)
#undef X
  fprintf(f, ",\n        opt_integrator(d.opt_integrator), opt_cone(d.opt_cone), opt_solver(d.opt_solver), opt_iterations(d.opt_iterations), opt_ls_iterations(d.opt_ls_iterations), opt_disableflags(d.opt_disableflags), opt_enableflags(d.opt_enableflags), nconmax(d.nconmax), njmax(d.njmax) {}\n};\n");
  fprintf(f, "struct Traits {\n  using Model = bp_static_%s::Model;\n  using Layout = bp_static_%s::Layout;\n", name, name);
  fprintf(f, "  static constexpr unsigned long long HASH = %lluull;\n  static constexpr int njmax = %d, nconmax = %d;\n};\n", m->struct_hash, njmax, nconmax);
  fprintf(f, "}  // namespace bp_static_%s\n#endif\n", name);
  fclose(f);
  return BP_OK;
}

extern "C" unsigned long long bp_model_struct_hash(bp_model* m) { return m ? m->struct_hash : 0; }

extern "C" bp_model* bp_model_create(const void* blob, int64_t nbytes, int device,
                                     char* err, int errsz) {
  auto bail = [&](const std::string& msg) -> bp_model* {
    g_err = msg;
    if (err && errsz > 0) snprintf(err, errsz, "%s", msg.c_str());
    return nullptr;
  };
  if (!blob || nbytes < 16 || memcmp(blob, "BPMODEL1", 8) != 0) return bail("bad model blob magic");
  if (bp_device_count() <= 0) return bail("no CUDA device: libbphys has no CPU fallback");
  if (cudaSetDevice(device) != cudaSuccess) return bail("cudaSetDevice failed");
  const char* base = (const char*)blob;
  int nfields = *(const int32_t*)(base + 8);
  const BlobEntry* ent = (const BlobEntry*)(base + 16);
  bp_model* m = new bp_model();
  m->device = device;
  for (int k = 0; k < nfields; k++) {
    HostField h;
    h.dtype = ent[k].dtype;
    if (ent[k].offset < 0 || ent[k].offset + ent[k].count * (h.dtype == 0 ? 8 : 4) > nbytes) {
      delete m; return bail("model blob truncated");
    }
    if (h.dtype == 0) h.f.assign((const double*)(base + ent[k].offset), (const double*)(base + ent[k].offset) + ent[k].count);
    else h.i.assign((const int32_t*)(base + ent[k].offset), (const int32_t*)(base + ent[k].offset) + ent[k].count);
    if (h.dtype == 0 && h.f.empty()) h.f.push_back(0);
    if (h.dtype == 1 && h.i.empty()) h.i.push_back(0);
    char nm[41]; memcpy(nm, ent[k].name, 40); nm[40] = 0;
    m->fields[nm] = h;
  }
  const char* missing = nullptr;
#define X(n) if (!m->fields.count(#n)) missing = #n;
  BP_MODEL_INT_ARRAYS(X)
  BP_MODEL_F64_ARRAYS(X)
#undef X
  if (missing) { std::string s = std::string("model blob lacks field ") + missing; delete m; return bail(s); }
  if (refresh_scalars(m) != BP_OK || derive(m) != BP_OK || validate(m) != BP_OK ||
      upload(m) != BP_OK) {
    std::string s = g_err; if (m->dev) cudaFree(m->dev); delete m; return bail(s);
  }
  return m;
}

extern "C" void bp_model_free(bp_model* m) {
  if (!m) return;
  if (m->dev) cudaFree(m->dev);
  delete m;
}

// geom_rbound as the compiler derives it (mjcf_compiler.py _rbound) -- recomputed when a
// geom_size / geom_type edit arrives at run time (the broadphase reads it)
static void recompute_rbound(bp_model* m) {
  auto& gt = I(m, "geom_type");
  auto& gs = F(m, "geom_size");
  auto& rb = m->fields["geom_rbound"].f;
  for (int g = 0; g < m->M.ngeom; g++) {
    const double* z = &gs[3 * g];
    double r = 0;
    switch (gt[g]) {
      case GEOM_SPHERE: r = z[0]; break;
      case GEOM_CAPSULE: r = z[0] + z[1]; break;
      case GEOM_CYLINDER: r = sqrt(z[0] * z[0] + z[1] * z[1]); break;
      case GEOM_BOX: r = sqrt(z[0] * z[0] + z[1] * z[1] + z[2] * z[2]); break;
      case GEOM_ELLIPSOID: r = z[0] > z[1] ? (z[0] > z[2] ? z[0] : z[2]) : (z[1] > z[2] ? z[1] : z[2]); break;
      default: r = 0;
    }
    rb[g] = r;
  }
}

// Run-time model edit.  The arrays derived on the host (dof_actcoef from actuator_gear /
// wrap_prm, fric_dof, lim_jnt, neq_active, the LEAN / MID shape flags, geom_rbound) are
// rebuilt and the whole model is uploaded again, so no stale copy survives an edit; an edit
// that produces a model the kernels cannot run (validate) is rolled back and rejected.
// Structural fields (sizes, tree topology, addresses) cannot be edited.
extern "C" int bp_model_set_field(bp_model* m, const char* name, const void* data,
                                  int64_t nbytes) {
  if (!m || !name || !data) return fail(BP_EINVAL, "null argument");
  auto it = m->fields.find(name);
  if (it == m->fields.end()) return fail(BP_ENOFIELD, std::string("no model field ") + name);
  static const char* structural[] = {
      "nq", "nv", "nu", "na", "nbody", "njnt", "ngeom", "nsite", "nM", "npair", "nconmax",
      "njmax", "nsensor", "nsensordata", "nkey", "ntendon", "neq", "nhfield", "nhfielddata",
      "nhfpair", "nspatial", "opt_solver", "body_parentid", "body_rootid", "body_weldid", "body_jntnum",
      "body_jntadr", "body_dofnum", "body_dofadr", "jnt_type", "jnt_qposadr", "jnt_dofadr",
      "jnt_bodyid", "dof_bodyid", "dof_jntid", "dof_parentid", "dof_Madr", "geom_bodyid",
      "site_bodyid", "pair_geom1", "pair_geom2", "pair_parentchild", "actuator_trntype",
      "actuator_trnid", "sensor_type", "sensor_objtype", "sensor_objid", "sensor_dim",
      "sensor_adr", "sensor_needstage", "tendon_adr", "tendon_num", "wrap_objid", "wrap_type",
      "tendon_Jadr", "eq_type", "eq_obj1id", "eq_obj2id", "geom_dataid", "hfield_nrow",
      "hfield_ncol", "hfield_adr"};
  for (const char* sname : structural)
    if (strcmp(sname, name) == 0)
      return fail(BP_EINVAL, std::string("model field ") + name +
                  " is structural and cannot be edited at run time");
  HostField& h = it->second;
  if ((int64_t)h.nbytes() != nbytes) return fail(BP_EINVAL, std::string("size mismatch for ") + name);
  HostField saved = h;
  memcpy(h.dtype == 0 ? (void*)h.f.data() : (void*)h.i.data(), data, nbytes);
  std::vector<double> saved_rbound;
  const bool resize = strcmp(name, "geom_size") == 0 || strcmp(name, "geom_type") == 0;
  if (resize) { saved_rbound = m->fields["geom_rbound"].f; recompute_rbound(m); }
  cudaSetDevice(m->device);
  int e = refresh_scalars(m);
  if (e == BP_OK) e = derive(m);
  if (e == BP_OK) e = validate(m);
  if (e != BP_OK) {                      // roll back: the model stays what it was
    std::string msg = g_err;
    m->fields[name] = saved;
    if (resize) m->fields["geom_rbound"].f = saved_rbound;
    refresh_scalars(m); derive(m);
    return fail(e, msg);
  }
  // launches already queued on batch streams may still read the old buffer
  CUDA_OK(cudaDeviceSynchronize());
  return upload(m);
}

extern "C" int bp_model_get_field(bp_model* m, const char* name, void* data, int64_t nbytes) {
  if (!m || !name || !data) return fail(BP_EINVAL, "null argument");
  auto it = m->fields.find(name);
  if (it == m->fields.end()) return fail(BP_ENOFIELD, std::string("no model field ") + name);
  if ((int64_t)it->second.nbytes() != nbytes) return fail(BP_EINVAL, "size mismatch");
  memcpy(data, it->second.data(), nbytes);
  return BP_OK;
}

// ------------------------------------------------------------------ batch

struct bp_batch {
  bp_model* model;
  int64_t nenv;
  BpLayout L;
  double* image = nullptr;
  unsigned char* mask = nullptr;
  bool use_mask = false;
  cudaStream_t stream = nullptr;
  std::vector<FieldInfo> fields;
  double *ctrl_stage = nullptr, *obs_dev = nullptr, *rew_dev = nullptr;  // e2e staging
  int obs_cap = 0;
  int lanes = 32, interleave = 1;
  int64_t stride = 0, nenv_alloc = 0;
  void* stage = nullptr;           // dense staging buffer for host <-> image copies
  size_t stage_bytes = 0;
  float* hfield = nullptr;         // [nenv_alloc][nhfielddata] per-env elevation data
  double *env_geom_pos = nullptr, *env_site_pos = nullptr;   // per-env model frames (lazy)
  double *env_body_pos = nullptr, *env_body_quat = nullptr;
  double* env_wrap_prm = nullptr;  // per-env tendon coefficients (lazy)
  BpTask task;
  int last_static = 0;             // the last launch was served by a static instance
};

static void make_layout(const BpModel& m, int njmax, int nconmax, BpLayout* L,
                        std::vector<FieldInfo>* fields) {
  // alignment inside an env's image: an array of at least one 128-byte line (16 doubles / 32
  // ints) starts on a line, smaller arrays stay packed, and the image is a whole number of
  // lines -- so the big arrays (J, H, the contact / body tables) touch as few lines as their
  // size allows.  Measured with the static instances, same box (profiles/r2_tuning.md):
  // cheetah +2.9 %, humanoid +2.7 %, walker +2.6 %, quadruped-escape +1.7 %, finger -0.9 %;
  // aligning the small arrays as well costs the small models more lines than it saves.
  // BPHYS_ALIGN=1 packs everything, 16 puts every array on its own line (static instances
  // are generated for the default).
  int al = 0;
  if (const char* e = getenv("BPHYS_ALIGN")) { int v = atoi(e); if (v == 1 || v == 2 || v == 4 || v == 8 || v == 16) al = v; }
  auto up = [&](int x, int cnt, int cap) {
    int a = 1;
    if (al > 0) a = cnt > 1 ? al * (cap / 16) : 1;
    else if (cnt >= cap) a = cap;
    return (x + a - 1) / a * a;
  };
  int off = 0;
#define X(n, cnt) { int c_ = (cnt); off = up(off, c_, 16); L->n = off; \
    if (fields) fields->push_back({#n, (int64_t)off * 8, c_, 0}); off += c_; }
  BP_DATA_F64(X)
#undef X
  off = (off + 15) / 16 * 16;
  L->ints = off;
  int ioff = 0;
#define X(n, cnt) { int c_ = (cnt); ioff = up(ioff, c_, 32); L->n = ioff; \
    if (fields) fields->push_back({#n, (int64_t)L->ints * 8 + (int64_t)ioff * 4, c_, 1}); ioff += c_; }
  BP_DATA_I32(X)
#undef X
  off += (ioff + 1) / 2;
  off = (off + 15) / 16 * 16;
  L->persist = off;
#define X(n, cnt) { int c_ = (cnt); off = up(off, c_, 16); L->n = off; off += c_; }
  BP_DATA_SCRATCH(X)
#undef X
  off = (off + 15) / 16 * 16;
  L->total = off;
  L->njmax = njmax; L->nconmax = nconmax;
}

extern "C" bp_batch* bp_batch_create(bp_model* m, int64_t nenv, int njmax, int nconmax,
                                     int lanes, char* err, int errsz) {
  auto bail = [&](const std::string& msg) -> bp_batch* {
    g_err = msg;
    if (err && errsz > 0) snprintf(err, errsz, "%s", msg.c_str());
    return nullptr;
  };
  if (!m || nenv <= 0) return bail("bad arguments to bp_batch_create");
  if (njmax <= 0) njmax = m->M.njmax;
  if (nconmax <= 0) nconmax = m->M.nconmax;
  cudaSetDevice(m->device);
  bp_batch* b = new bp_batch();
  b->model = m; b->nenv = nenv;
  memset(&b->task, 0, sizeof b->task);
  make_layout(m->M, njmax, nconmax, &b->L, &b->fields);
  if (lanes <= 0) lanes = 32;
  if (lanes != 1 && lanes != 2 && lanes != 4 && lanes != 8 && lanes != 16 && lanes != 32) {
    delete b; return bail("lanes per environment must be 1, 2, 4, 8, 16 or 32");
  }
  b->lanes = lanes;
  b->interleave = lanes == 1 ? 32 : 1;
  b->stride = lanes == 32 ? b->L.persist : b->L.total;
  b->nenv_alloc = (nenv + 31) / 32 * 32;
  int maxsm = 0;
  cudaDeviceGetAttribute(&maxsm, cudaDevAttrMaxSharedMemoryPerBlockOptin, m->device);
  if (lanes == 32 && (size_t)b->L.total * 8 > (size_t)maxsm) {
    char msg[256];
    snprintf(msg, sizeof msg, "one environment needs %zu B of shared memory (njmax=%d, "
             "nconmax=%d) but an SM offers %d B: lower njmax/nconmax",
             (size_t)b->L.total * 8, njmax, nconmax, maxsm);
    delete b; return bail(msg);
  }
  size_t bytes = (size_t)b->nenv_alloc * b->stride * 8;
  if (cudaMalloc((void**)&b->image, bytes) != cudaSuccess ||
      cudaMalloc((void**)&b->mask, nenv) != cudaSuccess ||
      cudaStreamCreateWithFlags(&b->stream, cudaStreamNonBlocking) != cudaSuccess) {
    std::string s = std::string("device allocation failed: ") + cudaGetErrorString(cudaGetLastError());
    if (b->image) cudaFree(b->image);
    if (b->mask) cudaFree(b->mask);
    delete b; return bail(s);
  }
  cudaMemsetAsync(b->image, 0, bytes, b->stream);
  cudaMemsetAsync(b->mask, 1, nenv, b->stream);
  if (m->M.nhfielddata > 0) {
    // every env starts from the model's elevation data (mjModel.hfield_data is float32)
    size_t nh = (size_t)m->M.nhfielddata;
    std::vector<float> h0(nh);
    auto& src = m->fields["hfield_data"].f;
    for (size_t i = 0; i < nh; i++) h0[i] = (float)src[i];
    if (cudaMalloc((void**)&b->hfield, (size_t)b->nenv_alloc * nh * 4) != cudaSuccess) {
      cudaFree(b->image); cudaFree(b->mask);
      delete b; return bail("device allocation of the height-field data failed");
    }
    bool zero = true;
    for (float v : h0) if (v != 0) zero = false;
    if (zero) cudaMemsetAsync(b->hfield, 0, (size_t)b->nenv_alloc * nh * 4, b->stream);
    else
      for (int64_t e = 0; e < b->nenv_alloc; e++)
        cudaMemcpyAsync(b->hfield + e * nh, h0.data(), nh * 4, cudaMemcpyHostToDevice, b->stream);
    cudaStreamSynchronize(b->stream);
  }
  bp_reset(b, -1);
  cudaStreamSynchronize(b->stream);
  return b;
}

extern "C" void bp_batch_free(bp_batch* b) {
  if (!b) return;
  cudaSetDevice(b->model->device);
  if (b->stream) { cudaStreamSynchronize(b->stream); cudaStreamDestroy(b->stream); }
  cudaFree(b->image); cudaFree(b->mask);
  if (b->hfield) cudaFree(b->hfield);
  if (b->env_geom_pos) cudaFree(b->env_geom_pos);
  if (b->env_site_pos) cudaFree(b->env_site_pos);
  if (b->env_body_pos) cudaFree(b->env_body_pos);
  if (b->env_body_quat) cudaFree(b->env_body_quat);
  if (b->env_wrap_prm) cudaFree(b->env_wrap_prm);
  if (b->ctrl_stage) cudaFree(b->ctrl_stage);
  if (b->stage) cudaFree(b->stage);
  if (b->obs_dev) cudaFree(b->obs_dev);
  if (b->rew_dev) cudaFree(b->rew_dev);
  delete b;
}

extern "C" void* bp_batch_image(bp_batch* b) { return b ? b->image : nullptr; }
extern "C" int64_t bp_batch_stride(bp_batch* b) { return b ? b->stride : 0; }
extern "C" int bp_batch_lanes(bp_batch* b) { return b ? b->lanes : 0; }
extern "C" int bp_batch_interleave(bp_batch* b) { return b ? b->interleave : 0; }
extern "C" int64_t bp_batch_nenv(bp_batch* b) { return b ? b->nenv : 0; }
extern "C" int64_t bp_batch_smem_bytes(bp_batch* b) { return b ? (int64_t)b->L.total * 8 : 0; }
extern "C" void* bp_batch_stream(bp_batch* b) { return b ? (void*)b->stream : nullptr; }
extern "C" int bp_field_count(bp_batch* b) { return b ? (int)b->fields.size() : 0; }
extern "C" const char* bp_field_name(bp_batch* b, int i) {
  return (b && i >= 0 && i < (int)b->fields.size()) ? b->fields[i].name.c_str() : nullptr;
}
static const FieldInfo* find_field(bp_batch* b, const char* name) {
  for (auto& f : b->fields) if (f.name == name) return &f;
  return nullptr;
}
extern "C" int bp_field_info(bp_batch* b, const char* name, int64_t* off, int64_t* count,
                             int* dtype) {
  if (!b || !name) return fail(BP_EINVAL, "null argument");
  const FieldInfo* f = find_field(b, name);
  if (!f) return fail(BP_ENOFIELD, std::string("no data field ") + name);
  if (off) *off = f->offset_bytes;
  if (count) *count = f->count;
  if (dtype) *dtype = f->dtype;
  return BP_OK;
}

static BpParams make_params(bp_batch* b, int mode, int nsub, int disable_actuation);

// dense [n][count] device buffer <-> the field inside the (possibly interleaved) image
static int field_copy_dev(bp_batch* b, const FieldInfo* f, void* dense, int64_t env0, int64_t n,
                          bool to_image) {
  BpParams P = make_params(b, 0, 0, 0);
  long long off = f->dtype == 0 ? f->offset_bytes / 8 : (f->offset_bytes - (int64_t)b->L.ints * 8) / 4;
  int e = bp_launch_field_copy(&P, off, (int)f->count, f->dtype, dense, env0, n, to_image ? 1 : 0,
                               b->stream);
  if (e != 0) return fail(BP_ECUDA, std::string("field copy: ") + cudaGetErrorString((cudaError_t)e));
  return BP_OK;
}

static int field_copy(bp_batch* b, const char* name, void* host, int64_t env0, int64_t n,
                      bool to_device) {
  if (!b || !name || !host) return fail(BP_EINVAL, "null argument");
  const FieldInfo* f = find_field(b, name);
  if (!f) return fail(BP_ENOFIELD, std::string("no data field ") + name);
  if (env0 < 0 || n < 0 || env0 + n > b->nenv) return fail(BP_EINVAL, "env range out of bounds");
  if (n == 0 || f->count == 0) return BP_OK;
  cudaSetDevice(b->model->device);
  size_t bytes = (size_t)n * f->count * (f->dtype == 0 ? 8 : 4);
  if (bytes > b->stage_bytes) {
    if (b->stage) { cudaStreamSynchronize(b->stream); cudaFree(b->stage); b->stage = nullptr; }
    CUDA_OK(cudaMalloc(&b->stage, bytes));
    b->stage_bytes = bytes;
  }
  if (to_device) {
    CUDA_OK(cudaMemcpyAsync(b->stage, host, bytes, cudaMemcpyHostToDevice, b->stream));
    int e = field_copy_dev(b, f, b->stage, env0, n, true);
    if (e != BP_OK) return e;
  } else {
    int e = field_copy_dev(b, f, b->stage, env0, n, false);
    if (e != BP_OK) return e;
    CUDA_OK(cudaMemcpyAsync(host, b->stage, bytes, cudaMemcpyDeviceToHost, b->stream));
  }
  CUDA_OK(cudaStreamSynchronize(b->stream));
  return BP_OK;
}
extern "C" int bp_field_write(bp_batch* b, const char* name, const void* host, int64_t env0,
                              int64_t n) {
  return field_copy(b, name, (void*)host, env0, n, true);
}
extern "C" int bp_field_read(bp_batch* b, const char* name, void* host, int64_t env0,
                             int64_t n) {
  return field_copy(b, name, host, env0, n, false);
}

extern "C" int bp_field_copy_device(bp_batch* b, const char* name, void* dev, int to_image) {
  if (!b || !name || !dev) return fail(BP_EINVAL, "null argument");
  const FieldInfo* f = find_field(b, name);
  if (!f) return fail(BP_ENOFIELD, std::string("no data field ") + name);
  if (f->count == 0) return BP_OK;
  cudaSetDevice(b->model->device);
  return field_copy_dev(b, f, dev, 0, b->nenv, to_image != 0);
}

extern "C" int bp_dfma_bench(int blocks, int threads, int iters, double* scratch, float* ms,
                             cudaStream_t stream);
extern "C" int bp_fp64_peak(int device, double* tflops) {
  if (!tflops) return fail(BP_EINVAL, "null argument");
  CUDA_OK(cudaSetDevice(device));
  int sms = 0;
  CUDA_OK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));
  double* scratch = nullptr;
  CUDA_OK(cudaMalloc((void**)&scratch, 64));
  const int blocks = sms * 8, threads = 256, iters = 1 << 16;
  float ms = 0;
  int e = bp_dfma_bench(blocks, threads, iters, scratch, &ms, 0);
  cudaFree(scratch);
  if (e != 0) return fail(BP_ECUDA, std::string("dfma bench: ") + cudaGetErrorString((cudaError_t)e));
  *tflops = (double)blocks * threads * iters * 8.0 * 2.0 / (ms * 1e-3) / 1e12;
  return BP_OK;
}

extern "C" int bp_set_mask(bp_batch* b, const uint8_t* mask_host) {
  if (!b) return fail(BP_EINVAL, "null batch");
  cudaSetDevice(b->model->device);
  b->use_mask = mask_host != nullptr;
  if (mask_host) {
    CUDA_OK(cudaMemcpyAsync(b->mask, mask_host, b->nenv, cudaMemcpyHostToDevice, b->stream));
    CUDA_OK(cudaStreamSynchronize(b->stream));
  }
  return BP_OK;
}

static BpParams make_params(bp_batch* b, int mode, int nsub, int disable_actuation) {
  BpParams P;
  memset(&P, 0, sizeof P);
  P.m = b->model->M; P.L = b->L; P.image = b->image;
  P.mask = b->use_mask ? b->mask : nullptr;
  P.nenv = b->nenv; P.mode = mode; P.nsub = nsub; P.disable_actuation = disable_actuation;
  P.stride = b->stride; P.interleave = b->interleave; P.lanes = b->lanes;
  P.hfield = b->hfield;
  P.env_geom_pos = b->env_geom_pos; P.env_site_pos = b->env_site_pos;
  P.env_body_pos = b->env_body_pos; P.env_body_quat = b->env_body_quat;
  P.env_wrap_prm = b->env_wrap_prm;
  P.env_wrap_n = (int)b->model->fields["wrap_prm"].f.size();
  return P;
}

// Per-env copy of a model array that a task edits per episode.  Allocated on the first
// write, every env starting from the model's values.
static int env_model_copy(bp_batch* b, const char* name, double* host, int64_t env0, int64_t n,
                          bool to_device) {
  if (!b || !name || !host) return fail(BP_EINVAL, "null argument");
  double** slot; int cnt;
  if (strcmp(name, "geom_pos") == 0) { slot = &b->env_geom_pos; cnt = 3 * b->model->M.ngeom; }
  else if (strcmp(name, "site_pos") == 0) { slot = &b->env_site_pos; cnt = 3 * b->model->M.nsite; }
  else if (strcmp(name, "body_pos") == 0) { slot = &b->env_body_pos; cnt = 3 * b->model->M.nbody; }
  else if (strcmp(name, "body_quat") == 0) { slot = &b->env_body_quat; cnt = 4 * b->model->M.nbody; }
  else if (strcmp(name, "wrap_prm") == 0) {
    slot = &b->env_wrap_prm; cnt = (int)b->model->fields["wrap_prm"].f.size();
  }
  else return fail(BP_ENOFIELD, std::string("no per-env model field ") + name +
                   " (geom_pos, site_pos, body_pos, body_quat, wrap_prm; hfield_data has "
                   "bp_hfield_write)");
  if (env0 < 0 || n < 0 || env0 + n > b->nenv) return fail(BP_EINVAL, "env range out of bounds");
  if (cnt == 0 || n == 0) return BP_OK;
  cudaSetDevice(b->model->device);
  if (!*slot) {
    auto& src = b->model->fields[name].f;
    std::vector<double> init((size_t)b->nenv_alloc * cnt);
    for (int64_t e = 0; e < b->nenv_alloc; e++) memcpy(&init[(size_t)e * cnt], src.data(), cnt * 8);
    CUDA_OK(cudaStreamSynchronize(b->stream));
    CUDA_OK(cudaMalloc((void**)slot, init.size() * 8));
    CUDA_OK(cudaMemcpy(*slot, init.data(), init.size() * 8, cudaMemcpyHostToDevice));
  }
  if (to_device) CUDA_OK(cudaMemcpyAsync(*slot + env0 * cnt, host, (size_t)n * cnt * 8,
                                         cudaMemcpyHostToDevice, b->stream));
  else CUDA_OK(cudaMemcpyAsync(host, *slot + env0 * cnt, (size_t)n * cnt * 8,
                               cudaMemcpyDeviceToHost, b->stream));
  CUDA_OK(cudaStreamSynchronize(b->stream));
  return BP_OK;
}
extern "C" int bp_env_model_write(bp_batch* b, const char* name, const double* host,
                                  int64_t env0, int64_t n) {
  return env_model_copy(b, name, (double*)host, env0, n, true);
}
extern "C" int bp_env_model_read(bp_batch* b, const char* name, double* host, int64_t env0,
                                 int64_t n) {
  return env_model_copy(b, name, host, env0, n, false);
}

// per-env model data: elevation of every height field of envs [env0, env0 + n)
static int hfield_copy(bp_batch* b, float* host, int64_t env0, int64_t n, bool to_device) {
  if (!b || !host) return fail(BP_EINVAL, "null argument");
  if (!b->hfield) return fail(BP_ENOFIELD, "the model has no height field");
  if (env0 < 0 || n < 0 || env0 + n > b->nenv) return fail(BP_EINVAL, "env range out of bounds");
  cudaSetDevice(b->model->device);
  size_t nh = (size_t)b->model->M.nhfielddata;
  if (to_device) CUDA_OK(cudaMemcpyAsync(b->hfield + env0 * nh, host, (size_t)n * nh * 4,
                                         cudaMemcpyHostToDevice, b->stream));
  else CUDA_OK(cudaMemcpyAsync(host, b->hfield + env0 * nh, (size_t)n * nh * 4,
                               cudaMemcpyDeviceToHost, b->stream));
  CUDA_OK(cudaStreamSynchronize(b->stream));
  return BP_OK;
}
extern "C" int bp_hfield_write(bp_batch* b, const float* host, int64_t env0, int64_t n) {
  return hfield_copy(b, (float*)host, env0, n, true);
}
extern "C" int bp_hfield_read(bp_batch* b, float* host, int64_t env0, int64_t n) {
  return hfield_copy(b, host, env0, n, false);
}

static int launch(bp_batch* b, int mode, int nsub, int disable_actuation,
                  const double* ctrl_in = nullptr, double* obs_out = nullptr,
                  double* rew_out = nullptr, double* obs_host = nullptr,
                  double* rew_host = nullptr) {
  if (!b) return fail(BP_EINVAL, "null batch");
  cudaSetDevice(b->model->device);
  BpParams P = make_params(b, mode, nsub, disable_actuation);
  P.ctrl_in = ctrl_in; P.obs_out = obs_out; P.rew_out = rew_out;
  P.obs_host = obs_host; P.rew_host = rew_host;
  if (rew_out) {
    P.task = b->task;
    P.task_nobs = bp_task_nobs_impl(&b->model->M, b->task.kind);
    if (P.task_nobs < 0) return fail(BP_EINVAL, "unknown task kind");
  }
  int e;
  const BpModel& M = b->model->M;
  // BPHYS_NO_LEAN=1 / BPHYS_SPLIT=1 keep the full instantiation
  const bool plain = M.opt_solver == 2 && M.opt_cone == 0 && M.opt_density == 0 && M.opt_viscosity == 0 &&
                     !P.env_geom_pos && !P.env_site_pos && !P.env_body_pos && !P.env_body_quat &&
                     !P.env_wrap_prm &&
                     !getenv("BPHYS_NO_LEAN") && !getenv("BPHYS_SPLIT");
  // a static (ahead-of-time specialised) instance when one serves this model's structure,
  // this batch's layout and lane width, and the shape the instance was built for still applies
  // (csrc/bp_step_static.cu); BPHYS_NO_STATIC=1 / BPHYS_SPLIT=1 keep the generic kernels
  e = -1;
  if (!getenv("BPHYS_NO_STATIC") && !getenv("BPHYS_SPLIT") && b->lanes != 32)
    e = bp_launch_step_static(&P, b->model->struct_hash, plain ? 1 : 0, b->stream);
  b->last_static = e != -1;
  if (e != -1) {
  } else if (plain && M.lean_static && (b->lanes == 4 || b->lanes == 8 || b->lanes == 16)) {
    e = b->lanes == 4 ? bp_launch_step_g4_lean(&P, b->stream)
        : b->lanes == 8 ? bp_launch_step_g8_lean(&P, b->stream)
                        : bp_launch_step_g16_lean(&P, b->stream);
  } else if (plain && M.mid_static && (b->lanes == 8 || b->lanes == 16)) {
    e = b->lanes == 8 ? bp_launch_step_g8_mid(&P, b->stream)
                      : bp_launch_step_g16_mid(&P, b->stream);
  } else
  switch (b->lanes) {
    case 1: e = bp_launch_step_g1(&P, b->stream); break;
    case 2: e = bp_launch_step_g2(&P, b->stream); break;
    case 4: e = bp_launch_step_g4(&P, b->stream); break;
    case 8: e = bp_launch_step_g8(&P, b->stream); break;
    case 16: e = bp_launch_step_g16(&P, b->stream); break;
    default: e = bp_launch_step_g32(&P, b->stream); break;
  }
  if (e != 0) return fail(BP_ECUDA, std::string("step launch: ") + cudaGetErrorString((cudaError_t)e));
  return BP_OK;
}
extern "C" int bp_batch_last_static(bp_batch* b) { return b ? b->last_static : 0; }
extern "C" int bp_step(bp_batch* b, int nsub) { return launch(b, BP_MODE_STEP, nsub, 0); }
extern "C" int bp_mj_step(bp_batch* b, int nsub) { return launch(b, BP_MODE_MJSTEP, nsub, 0); }
extern "C" int bp_forward(bp_batch* b, int da) { return launch(b, BP_MODE_FORWARD, 1, da); }
extern "C" int bp_step1(bp_batch* b) { return launch(b, BP_MODE_STEP1, 1, 0); }
extern "C" int bp_step2(bp_batch* b) { return launch(b, BP_MODE_STEP2, 1, 0); }

extern "C" int bp_env_step(bp_batch* b, int nsub, const double* ctrl_dev, double* obs_dev,
                           double* reward_dev) {
  if (!b || !reward_dev) return fail(BP_EINVAL, "null argument");
  return launch(b, BP_MODE_STEP, nsub, 0, b->model->M.nu > 0 ? ctrl_dev : nullptr, obs_dev,
                reward_dev);
}

extern "C" int bp_reset(bp_batch* b, int key) {
  if (!b) return fail(BP_EINVAL, "null batch");
  bp_model* m = b->model;
  cudaSetDevice(m->device);
  BpParams P = make_params(b, 0, 0, 0);
  const double *kq = nullptr, *kv = nullptr, *ka = nullptr;
  double kt = 0;
  if (key >= 0) {
    if (key >= m->M.nkey) return fail(BP_EINVAL, "keyframe index out of range");
    auto dev = [&](const char* n, int stride) {
      return (const double*)(m->dev + m->fields[n].dev_off) + (size_t)key * stride;
    };
    if (!m->fields.count("key_qpos")) return fail(BP_ENOFIELD, "model has no keyframes");
    kq = dev("key_qpos", m->M.nq); kv = dev("key_qvel", m->M.nv); ka = dev("key_act", m->M.na);
    kt = m->fields["key_time"].f[key];
  }
  int e = bp_launch_reset(&P, kq, kv, ka, kt, b->stream);
  if (e != 0) return fail(BP_ECUDA, std::string("reset launch: ") + cudaGetErrorString((cudaError_t)e));
  return BP_OK;
}

extern "C" int bp_set_task(bp_batch* b, int kind, const int32_t* ip, int nip,
                           const double* fp, int nfp) {
  if (!b || nip < 0 || nip > 24 || nfp < 0 || nfp > 8) return fail(BP_EINVAL, "bad task arguments");
  BpTask t;
  memset(&t, 0, sizeof t);
  t.kind = kind;
  for (int i = 0; i < nip; i++) t.ip[i] = ip[i];
  for (int i = 0; i < nfp; i++) t.fp[i] = fp[i];
  if (bp_task_nobs_impl(&b->model->M, kind) < 0) return fail(BP_EINVAL, "unknown task kind");
  b->task = t;
  return BP_OK;
}

extern "C" int bp_task_nobs(bp_batch* b) {
  return b ? bp_task_nobs_impl(&b->model->M, b->task.kind) : -1;
}

extern "C" int bp_obs_reward(bp_batch* b, double* obs, double* reward) {
  if (!b) return fail(BP_EINVAL, "null batch");
  cudaSetDevice(b->model->device);
  BpParams P = make_params(b, 0, 0, 0);
  int e = bp_launch_task(&P, &b->task, obs, reward, b->stream);
  if (e < 0) return fail(BP_EINVAL, "unknown task kind");
  if (e != 0) return fail(BP_ECUDA, std::string("task launch: ") + cudaGetErrorString((cudaError_t)e));
  return BP_OK;
}

static int ensure_task_buffers(bp_batch* b, int nobs) {
  int nu = b->model->M.nu;
  if (!b->ctrl_stage) {
    CUDA_OK(cudaMalloc((void**)&b->ctrl_stage, (size_t)b->nenv * (nu > 0 ? nu : 1) * 8));
    CUDA_OK(cudaMalloc((void**)&b->rew_dev, (size_t)b->nenv * 8));
  }
  if (nobs > b->obs_cap) {
    if (b->obs_dev) cudaFree(b->obs_dev);
    b->obs_dev = nullptr;
    CUDA_OK(cudaMalloc((void**)&b->obs_dev, (size_t)b->nenv * nobs * 8));
    b->obs_cap = nobs;
  }
  return BP_OK;
}

extern "C" int bp_obs_reward_host(bp_batch* b, double* obs_host, double* reward_host) {
  if (!b) return fail(BP_EINVAL, "null batch");
  cudaSetDevice(b->model->device);
  int nobs = bp_task_nobs_impl(&b->model->M, b->task.kind);
  if (nobs < 0) return fail(BP_EINVAL, "unknown task kind");
  int e = ensure_task_buffers(b, nobs);
  if (e != BP_OK) return e;
  e = bp_obs_reward(b, nobs > 0 ? b->obs_dev : nullptr, b->rew_dev);
  if (e != BP_OK) return e;
  if (obs_host && nobs > 0)
    CUDA_OK(cudaMemcpyAsync(obs_host, b->obs_dev, (size_t)b->nenv * nobs * 8,
                            cudaMemcpyDeviceToHost, b->stream));
  if (reward_host)
    CUDA_OK(cudaMemcpyAsync(reward_host, b->rew_dev, (size_t)b->nenv * 8,
                            cudaMemcpyDeviceToHost, b->stream));
  CUDA_OK(cudaStreamSynchronize(b->stream));
  return BP_OK;
}

extern "C" int bp_env_step_host(bp_batch* b, int nsub, const double* ctrl_host,
                                double* obs_host, double* reward_host) {
  if (!b || !ctrl_host) return fail(BP_EINVAL, "null argument");
  bp_model* m = b->model;
  cudaSetDevice(m->device);
  int nobs = bp_task_nobs_impl(&m->M, b->task.kind);
  if (nobs < 0) return fail(BP_EINVAL, "unknown task kind");
  int nu = m->M.nu;
  int eb = ensure_task_buffers(b, nobs);
  if (eb != BP_OK) return eb;
  // A pinned (mapped) host control buffer is read by the kernel itself over PCIe instead
  // of being staged by a copy (BPHYS_NO_ZEROCOPY=1 turns that off).
  static const bool zerocopy = getenv("BPHYS_NO_ZEROCOPY") == nullptr;
  const double* ctrl_src = b->ctrl_stage;
  if (zerocopy && nu > 0) {
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, ctrl_host) == cudaSuccess && a.type == cudaMemoryTypeHost &&
        a.devicePointer != nullptr)
      ctrl_src = (const double*)a.devicePointer;
    else
      cudaGetLastError();
  }
  if (nu > 0 && ctrl_src == b->ctrl_stage)
    CUDA_OK(cudaMemcpyAsync(b->ctrl_stage, ctrl_host, (size_t)b->nenv * nu * 8,
                            cudaMemcpyHostToDevice, b->stream));
  // Pinned (mapped) result buffers are filled by the kernel itself, block by block, while
  // other blocks still compute (BpParams.obs_host); pageable ones by D2H copies afterwards.
  // The global-memory shapes support it (every thread of a lock-step block reaches the end).
  double *obs_map = nullptr, *rew_map = nullptr;
  static const bool direct = getenv("BPHYS_NO_DIRECT_OUT") == nullptr;
  if (direct && b->lanes != 32 && !b->use_mask && !getenv("BPHYS_SPLIT")) {
    auto mapped = [](const void* p) -> double* {
      cudaPointerAttributes a;
      if (p && cudaPointerGetAttributes(&a, p) == cudaSuccess && a.type == cudaMemoryTypeHost &&
          a.devicePointer != nullptr)
        return (double*)a.devicePointer;
      cudaGetLastError();
      return nullptr;
    };
    if (nobs > 0) obs_map = mapped(obs_host);
    rew_map = mapped(reward_host);
  }
  int e = launch(b, BP_MODE_STEP, nsub, 0, nu > 0 ? ctrl_src : nullptr,
                 nobs > 0 ? b->obs_dev : nullptr, b->rew_dev, obs_map, rew_map);
  if (e != BP_OK) return e;
  if (obs_host && nobs > 0 && !obs_map)
    CUDA_OK(cudaMemcpyAsync(obs_host, b->obs_dev, (size_t)b->nenv * nobs * 8,
                            cudaMemcpyDeviceToHost, b->stream));
  if (reward_host && !rew_map)
    CUDA_OK(cudaMemcpyAsync(reward_host, b->rew_dev, (size_t)b->nenv * 8,
                            cudaMemcpyDeviceToHost, b->stream));
  CUDA_OK(cudaStreamSynchronize(b->stream));
  return BP_OK;
}

// mj_copyData (wrapper/core.py:653-685 MjData.copy / engine.py:332-339 Physics.copy): every
// env image, the per-env model arrays (height fields, frames) and the bound task of `src`
// into `dst`, which must have been created with the same model sizes, nenv, njmax / nconmax
// and lanes.  Device-to-device; synchronises.
extern "C" int bp_batch_copy(bp_batch* dst, bp_batch* src) {
  if (!dst || !src) return fail(BP_EINVAL, "null batch");
  if (dst->nenv != src->nenv || dst->stride != src->stride || dst->lanes != src->lanes ||
      dst->L.total != src->L.total || dst->L.persist != src->L.persist ||
      dst->model->M.nhfielddata != src->model->M.nhfielddata)
    return fail(BP_EINVAL, "bp_batch_copy: the two batches differ in shape");
  cudaSetDevice(src->model->device);
  CUDA_OK(cudaStreamSynchronize(src->stream));
  CUDA_OK(cudaStreamSynchronize(dst->stream));
  CUDA_OK(cudaMemcpy(dst->image, src->image, (size_t)src->nenv_alloc * src->stride * 8,
                     cudaMemcpyDeviceToDevice));
  if (src->hfield)
    CUDA_OK(cudaMemcpy(dst->hfield, src->hfield,
                       (size_t)src->nenv_alloc * src->model->M.nhfielddata * 4,
                       cudaMemcpyDeviceToDevice));
  struct { double** d; double** s; int cnt; } slots[] = {
      {&dst->env_geom_pos, &src->env_geom_pos, 3 * src->model->M.ngeom},
      {&dst->env_site_pos, &src->env_site_pos, 3 * src->model->M.nsite},
      {&dst->env_body_pos, &src->env_body_pos, 3 * src->model->M.nbody},
      {&dst->env_body_quat, &src->env_body_quat, 4 * src->model->M.nbody},
      {&dst->env_wrap_prm, &src->env_wrap_prm, (int)src->model->fields["wrap_prm"].f.size()}};
  for (auto& sl : slots) {
    if (!*sl.s) {
      if (*sl.d) { cudaFree(*sl.d); *sl.d = nullptr; }
      continue;
    }
    size_t bytes = (size_t)src->nenv_alloc * sl.cnt * 8;
    if (!*sl.d) CUDA_OK(cudaMalloc((void**)sl.d, bytes));
    CUDA_OK(cudaMemcpy(*sl.d, *sl.s, bytes, cudaMemcpyDeviceToDevice));
  }
  dst->task = src->task;
  return BP_OK;
}

extern "C" int bp_launch_contact_force(const BpParams* P, int ci, double* out, cudaStream_t stream);
extern "C" int bp_launch_object_velocity(const BpParams* P, int objtype, int id, int flg_local,
                                         double* out, cudaStream_t stream);

// [nenv][6] results of a per-env query kernel -> host, through the batch's staging buffer
static int query6(bp_batch* b, double* out_host, int (*go)(const BpParams*, double*, void*),
                  void* arg) {
  cudaSetDevice(b->model->device);
  size_t bytes = (size_t)b->nenv * 6 * 8;
  if (bytes > b->stage_bytes) {
    if (b->stage) { cudaStreamSynchronize(b->stream); cudaFree(b->stage); b->stage = nullptr; }
    CUDA_OK(cudaMalloc(&b->stage, bytes));
    b->stage_bytes = bytes;
  }
  BpParams P = make_params(b, 0, 0, 0);
  int e = go(&P, (double*)b->stage, arg);
  if (e != 0) return fail(BP_ECUDA, std::string("query launch: ") + cudaGetErrorString((cudaError_t)e));
  CUDA_OK(cudaMemcpyAsync(out_host, b->stage, bytes, cudaMemcpyDeviceToHost, b->stream));
  CUDA_OK(cudaStreamSynchronize(b->stream));
  return BP_OK;
}

extern "C" int bp_contact_force(bp_batch* b, int contact_id, double* out_host) {
  if (!b || !out_host) return fail(BP_EINVAL, "null argument");
  if (contact_id < 0 || contact_id >= b->L.nconmax)
    return fail(BP_EINVAL, "contact_id must be an integer in range [0, " +
                std::to_string(b->L.nconmax - 1) + "] (inclusive), got: " +
                std::to_string(contact_id) + ".");
  struct A { int ci; cudaStream_t s; } a{contact_id, b->stream};
  return query6(b, out_host, [](const BpParams* P, double* out, void* arg) {
    A* q = (A*)arg; return bp_launch_contact_force(P, q->ci, out, q->s); }, &a);
}

extern "C" int bp_object_velocity(bp_batch* b, int objtype, int objid, int flg_local,
                                  double* out_host) {
  if (!b || !out_host) return fail(BP_EINVAL, "null argument");
  const BpModel& M = b->model->M;
  int n = objtype == OBJ_SITE ? M.nsite : objtype == OBJ_GEOM ? M.ngeom :
          (objtype == OBJ_BODY || objtype == OBJ_XBODY) ? M.nbody : -1;
  if (n < 0) return fail(BP_EINVAL, "object type must be body (1), xbody (2), geom (5) or site (6)");
  if (objid < 0 || objid >= n) return fail(BP_EINVAL, "object id out of range");
  struct A { int t, id, loc; cudaStream_t s; } a{objtype, objid, flg_local, b->stream};
  return query6(b, out_host, [](const BpParams* P, double* out, void* arg) {
    A* q = (A*)arg; return bp_launch_object_velocity(P, q->t, q->id, q->loc, out, q->s); }, &a);
}

extern "C" int bp_warnings(bp_batch* b, int32_t* out_host) {
  return bp_field_read(b, "warning", out_host, 0, b ? b->nenv : 0);
}

extern "C" int bp_sync(bp_batch* b) {
  if (!b) return fail(BP_EINVAL, "null batch");
  cudaSetDevice(b->model->device);
  CUDA_OK(cudaStreamSynchronize(b->stream));
  CUDA_OK(cudaGetLastError());
  return BP_OK;
}

extern "C" int bp_launch_tolerance(const double* x, double* out, long long n, double lower,
                                   double upper, double margin, int sig, double v1,
                                   cudaStream_t stream);
extern "C" int bp_rewards_tolerance(int device, const double* x_host, double* out_host,
                                    int64_t n, double lower, double upper, double margin,
                                    int sigmoid, double value_at_margin) {
  if (!x_host || !out_host || n < 0) return fail(BP_EINVAL, "bad arguments");
  if (lower > upper) return fail(BP_EINVAL, "Lower bound must be <= upper bound.");
  if (margin < 0) return fail(BP_EINVAL, "`margin` must be non-negative.");
  if (sigmoid < 0 || sigmoid > 3) return fail(BP_EINVAL, "Unknown sigmoid type");
  if (n == 0) return BP_OK;
  CUDA_OK(cudaSetDevice(device));
  double *dx = nullptr, *dout = nullptr;
  CUDA_OK(cudaMalloc((void**)&dx, n * 8));
  CUDA_OK(cudaMalloc((void**)&dout, n * 8));
  CUDA_OK(cudaMemcpy(dx, x_host, n * 8, cudaMemcpyHostToDevice));
  int e = bp_launch_tolerance(dx, dout, n, lower, upper, margin, sigmoid, value_at_margin, 0);
  cudaError_t ce = cudaMemcpy(out_host, dout, n * 8, cudaMemcpyDeviceToHost);
  cudaFree(dx); cudaFree(dout);
  if (e != 0 || ce != cudaSuccess) return fail(BP_ECUDA, "tolerance kernel failed");
  return BP_OK;
}
