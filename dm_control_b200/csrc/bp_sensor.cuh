// mj_sensorPos/Vel/Acc, mj_subtreeVel, mj_rnePostConstraint, energy
// (read by the suite's observation/reward functions,
// e.g. /root/reference/dm_control/suite/cheetah.py:55-57).
#ifndef BP_SENSOR_CUH_
#define BP_SENSOR_CUH_
#include "bp_dyn.cuh"
#include "bp_hfield.cuh"

BP_STAGE void bp_subtree_vel(const C& c) {
  BP_MODEL(m);
  int nb = m.nbody;
  auto lv = DF(subtree_linvel), cvel = DF(cvel), com = DF(subtree_com), xipos = DF(xipos);
#pragma unroll 1
  for (int b = c.lane; b < nb; b += C::G) {
    auto cm = com + 3 * m.body_rootid[b];
    double dif[3], t[3], cv[6];
    for (int k = 0; k < 6; k++) cv[k] = cvel[6 * b + k];
    for (int k = 0; k < 3; k++) dif[k] = xipos[3 * b + k] - cm[k];
    cross3(t, cv, dif);
    for (int k = 0; k < 3; k++) lv[3 * b + k] = m.body_mass[b] * (cv[3 + k] + t[k]);
  }
  SYNC();
  bp_tree_gather<3>(c, lv, 0);
  for (int w = c.lane; w < 3 * nb; w += C::G) {
    double sm = m.body_subtreemass[w / 3];
    lv[w] /= (sm > BP_MINVAL ? sm : BP_MINVAL);
  }
  SYNC();
}

template <class C>
__device__ __forceinline__ void bp_contact_force(const C& c, int ci, double* f6) {
  for (int k = 0; k < 6; k++) f6[k] = 0;
  int a = IF(contact_efc_address)[ci];
  if (a < 0) return;
  int dim = IF(contact_dim)[ci];
  auto ef = DF(efc_force);
  if (dim == 1) { f6[0] = ef[a]; return; }
  if (bp_mv(c).opt_cone == 0) {
    for (int k = 0; k < 2 * (dim - 1); k++) f6[0] += ef[a + k];
    for (int k = 1; k < dim; k++)
      f6[k] = DF(contact_friction)[5 * ci + k - 1] * (ef[a + 2 * k - 2] - ef[a + 2 * k - 1]);
  } else {
    for (int k = 0; k < dim; k++) f6[k] = ef[a + k];
  }
}

BP_STAGE void bp_rne_post_constraint(const C& c) {
  BP_MODEL(m);
  int nb = m.nbody, ncon = IF(ncon)[0];
  auto ext = DF(cfrc_ext), com = DF(subtree_com);
  // each body gathers its own contacts in contact order (no atomics, fixed order)
#pragma unroll 1
  for (int b = c.lane; b < nb; b += C::G) {
    double acc[6] = {0, 0, 0, 0, 0, 0};
    if (b > 0) {
#pragma unroll 1
      for (int ci = 0; ci < ncon; ci++) {
        if (IF(contact_efc_address)[ci] < 0) continue;
        int b1 = m.geom_bodyid[IF(contact_geom1)[ci]], b2 = m.geom_bodyid[IF(contact_geom2)[ci]];
        if (b1 != b && b2 != b) continue;
        double f6[6], fw[3], tw[3], fr[9];
        bp_contact_force(c, ci, f6);
        for (int k = 0; k < 9; k++) fr[k] = DF(contact_frame)[9 * ci + k];
        matT_vec3(fw, fr, f6);
        matT_vec3(tw, fr, f6 + 3);
        auto pos = DF(contact_pos) + 3 * ci;
        auto cm = com + 3 * m.body_rootid[b];
        double dif[3] = {pos[0] - cm[0], pos[1] - cm[1], pos[2] - cm[2]}, t[3];
        cross3(t, dif, fw);
        for (int side = 0; side < 2; side++) {
          if ((side ? b2 : b1) != b) continue;
          double sgn = side ? 1 : -1;
          for (int k = 0; k < 3; k++) { acc[k] += sgn * (tw[k] + t[k]); acc[3 + k] += sgn * fw[k]; }
        }
      }
    }
    for (int k = 0; k < 6; k++) ext[6 * b + k] = acc[k];
  }
  SYNC();
  bp_rne(c, true, DF(cacc), DF(cfrc_int), ext, c.null(), 1);
}

template <class C>
__device__ __forceinline__ bool bp_inside_site(const C& c, int s, const double* p) {
  BP_MODEL(m);
  auto sp = DF(site_xpos) + 3 * s;
  double dif[3] = {p[0] - sp[0], p[1] - sp[1], p[2] - sp[2]}, l[3];
  matT_vec3(l, DF(site_xmat) + 9 * s, dif);
  const double* sz = m.site_size + 3 * s;
  switch (m.site_type[s]) {
    case GEOM_SPHERE: return dot3(l, l) <= sz[0] * sz[0];
    case GEOM_CAPSULE: {
      double z = clampd(l[2], -sz[1], sz[1]), dz = l[2] - z;
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
  return false;
}

// frame of a sensor's object copied to registers: pos[3], mat[9], owning body
template <class C>
__device__ __forceinline__ int bp_obj_frame(const C& c, int objtype, int id, double* pos,
                                            double* mat) {
  BP_MODEL(m);
  typename C::DP p, R;
  int body;
  if (objtype == OBJ_SITE) { p = DF(site_xpos) + 3 * id; R = DF(site_xmat) + 9 * id; body = m.site_bodyid[id]; }
  else if (objtype == OBJ_GEOM) { p = DF(geom_xpos) + 3 * id; R = DF(geom_xmat) + 9 * id; body = m.geom_bodyid[id]; }
  else if (objtype == OBJ_XBODY) { p = DF(xpos) + 3 * id; R = DF(xmat) + 9 * id; body = id; }
  else { p = DF(xipos) + 3 * id; R = DF(ximat) + 9 * id; body = id; }
  for (int k = 0; k < 3; k++) pos[k] = p[k];
  for (int k = 0; k < 9; k++) mat[k] = R[k];
  return body;
}

// mj_ray: nearest geom surface along the unit ray (pnt, vec), geoms of `bodyexclude`
// skipped; -1 without a hit (the rangefinder's reading, quadruped.py:212-217)
BP_STAGE double bp_ray(const C& c, const double* pnt, const double* vec, int bodyexclude) {
  BP_MODEL(m);
  auto gpos = DF(geom_xpos), gmat = DF(geom_xmat);
  double best = -1;
#pragma unroll 1
  for (int g = 0; g < m.ngeom; g++) {
    if (m.geom_bodyid[g] == bodyexclude) continue;
    double R[9], dif[3], o[3], v[3], sz[3];
    for (int k = 0; k < 9; k++) R[k] = gmat[9 * g + k];
    for (int k = 0; k < 3; k++) { dif[k] = pnt[k] - gpos[3 * g + k]; sz[k] = m.geom_size[3 * g + k]; }
    matT_vec3(o, R, dif);
    matT_vec3(v, R, vec);
    const int gt = m.geom_type[g];
    double x;
    if (gt == GEOM_HFIELD) {
      if (c.hf == nullptr) continue;
      HField h;
      hf_init(m, c.hf, m.geom_dataid[g], h);
      x = ray_hfield(h, o, v);
    } else {
      x = ray_prim(gt, sz, o, v);
    }
    best = ray_min(best, x);
  }
  return best;
}

BP_STAGE void bp_sensors(const C& c, int stage) {
  BP_MODEL(m);
  if (m.nsensor == 0) return;
  if (stage == STAGE_VEL && m.need_subtreevel) bp_subtree_vel(c);
  if (stage == m.need_rnepost_stage) bp_rne_post_constraint(c);
  auto sd = DF(sensordata);
#pragma unroll 1
  for (int s = c.lane; s < m.nsensor; s += C::G) {
    if (m.sensor_needstage[s] != stage) continue;
    int t = m.sensor_type[s], id = m.sensor_objid[s], dim = m.sensor_dim[s];
    double out[3] = {0, 0, 0}, pos[3], mat[9];
    switch (t) {
      case SENS_JOINTPOS: out[0] = DF(qpos)[m.jnt_qposadr[id]]; break;
      case SENS_JOINTVEL: out[0] = DF(qvel)[m.jnt_dofadr[id]]; break;
      case SENS_ACTUATORFRC: out[0] = DF(actuator_force)[id]; break;
      case SENS_SUBTREECOM: for (int k = 0; k < 3; k++) out[k] = DF(subtree_com)[3 * id + k]; break;
      case SENS_RANGEFINDER: if constexpr (!C::LEAN) {
        int body = bp_obj_frame(c, OBJ_SITE, id, pos, mat);
        double zax[3] = {mat[2], mat[5], mat[8]};
        out[0] = bp_ray(c, pos, zax, body);
      } break;
      case SENS_SUBTREELINVEL: for (int k = 0; k < 3; k++) out[k] = DF(subtree_linvel)[3 * id + k]; break;
      case SENS_FRAMEPOS:
        bp_obj_frame(c, m.sensor_objtype[s], id, pos, mat);
        for (int k = 0; k < 3; k++) out[k] = pos[k];
        break;
      case SENS_FRAMEXAXIS: case SENS_FRAMEYAXIS: case SENS_FRAMEZAXIS:
        bp_obj_frame(c, m.sensor_objtype[s], id, pos, mat);
        for (int k = 0; k < 3; k++) out[k] = mat[3 * k + (t - SENS_FRAMEXAXIS)];
        break;
      case SENS_VELOCIMETER: case SENS_GYRO: case SENS_ACCELEROMETER: {
        int body = bp_obj_frame(c, OBJ_SITE, id, pos, mat);
        auto cm = DF(subtree_com) + 3 * m.body_rootid[body];
        double cv[6];
        for (int k = 0; k < 6; k++) cv[k] = DF(cvel)[6 * body + k];
        double dif[3] = {pos[0] - cm[0], pos[1] - cm[1], pos[2] - cm[2]};
        double t3[3], vlin[3], wl[3], vl[3];
        cross3(t3, cv, dif);
        for (int k = 0; k < 3; k++) vlin[k] = cv[3 + k] + t3[k];
        matT_vec3(wl, mat, cv);
        matT_vec3(vl, mat, vlin);
        if (t == SENS_GYRO) { for (int k = 0; k < 3; k++) out[k] = wl[k]; }
        else if (t == SENS_VELOCIMETER) { for (int k = 0; k < 3; k++) out[k] = vl[k]; }
        else {
          double ca[6];
          for (int k = 0; k < 6; k++) ca[k] = DF(cacc)[6 * body + k];
          double alin[3], al[3], cor[3];
          cross3(t3, ca, dif);
          for (int k = 0; k < 3; k++) alin[k] = ca[3 + k] + t3[k];
          matT_vec3(al, mat, alin);
          cross3(cor, wl, vl);
          for (int k = 0; k < 3; k++) out[k] = al[k] + cor[k];
        }
        break;
      }
      case SENS_FORCE: case SENS_TORQUE: {
        int body = bp_obj_frame(c, OBJ_SITE, id, pos, mat);
        auto cm = DF(subtree_com) + 3 * m.body_rootid[body];
        double cf[6];
        for (int k = 0; k < 6; k++) cf[k] = DF(cfrc_int)[6 * body + k];
        if (t == SENS_FORCE) matT_vec3(out, mat, cf + 3);
        else {
          double dif[3] = {pos[0] - cm[0], pos[1] - cm[1], pos[2] - cm[2]}, t3[3], tq[3];
          cross3(t3, dif, cf + 3);
          for (int k = 0; k < 3; k++) tq[k] = cf[k] - t3[k];
          matT_vec3(out, mat, tq);
        }
        break;
      }
      case SENS_TOUCH: {
        int sb = m.site_bodyid[id], ncon = IF(ncon)[0];
        double tot = 0;
#pragma unroll 1
        for (int ci = 0; ci < ncon; ci++) {
          if (IF(contact_efc_address)[ci] < 0) continue;
          if (m.geom_bodyid[IF(contact_geom1)[ci]] != sb &&
              m.geom_bodyid[IF(contact_geom2)[ci]] != sb) continue;
          double f6[6], cp[3];
          bp_contact_force(c, ci, f6);
          if (f6[0] <= 0) continue;
          for (int k = 0; k < 3; k++) cp[k] = DF(contact_pos)[3 * ci + k];
          if (bp_inside_site(c, id, cp)) tot += f6[0];
        }
        out[0] = tot;
        break;
      }
    }
    double cut = m.sensor_cutoff[s];
    if (cut > 0)
      for (int k = 0; k < 3; k++) {
        if (t == SENS_TOUCH) { if (out[k] > cut) out[k] = cut; }
        else out[k] = clampd(out[k], -cut, cut);
      }
    int adr = m.sensor_adr[s];
    for (int k = 0; k < dim && k < 3; k++) sd[adr + k] = out[k];
  }
  SYNC();
}

BP_STAGE void bp_energy_pos(const C& c) {
  BP_MODEL(m);
  auto en = DF(energy);
  if (!(m.opt_enableflags & ENBL_ENERGY)) {
    for (int i = c.lane; i < 2; i += C::G) en[i] = 0;
    SYNC();
    return;
  }
  double e = 0;
  if (!(m.opt_disableflags & DSBL_GRAVITY))
    for (int b = 1 + c.lane; b < m.nbody; b += C::G)
      e -= m.body_mass[b] * dot3(m.opt_gravity, DF(xipos) + 3 * b);
  if (!(m.opt_disableflags & DSBL_PASSIVE))
    for (int j = c.lane; j < m.njnt; j += C::G) {
      int jt = m.jnt_type[j], qa = m.jnt_qposadr[j];
      if (jt == JNT_SLIDE || jt == JNT_HINGE) {
        double dq = DF(qpos)[qa] - m.qpos_spring[qa];
        e += 0.5 * m.jnt_stiffness[j] * dq * dq;
      }
    }
  e = group_sum(c, e);
  if (c.lane == 0) en[0] = e;
  SYNC();
}

// kinetic energy 0.5 v^T M v from the sparse qM
BP_STAGE void bp_energy_vel(const C& c) {
  BP_MODEL(m);
  if (!(m.opt_enableflags & ENBL_ENERGY)) return;
  auto qM = DF(qM), v = DF(qvel);
  double e = 0;
#pragma unroll 1
  for (int i = c.lane; i < m.nv; i += C::G) {
    int a = m.dof_Madr[i];
    double s = 0.5 * qM[a++] * v[i];
    for (int j = m.dof_parentid[i]; j >= 0; j = m.dof_parentid[j]) s += qM[a++] * v[j];
    e += s * v[i];
  }
  e = group_sum(c, e);
  if (c.lane == 0) DF(energy)[1] = e;
  SYNC();
}

#endif  // BP_SENSOR_CUH_
