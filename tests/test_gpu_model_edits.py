"""bp_model_create / bp_model_set_field: what the kernels cannot run is refused, and run-time
model edits rebuild the host-derived arrays (`dof_actcoef`, `fric_dof`, `lim_jnt`,
`geom_rbound`, the LEAN / MID shape flags went stale after an edit)."""

import numpy as np
import pytest

import parity_util as PU
from dm_control_b200 import binding
from dm_control_b200 import mjcf_compiler as C
from dm_control_b200 import model as M
from test_rejections_cpu import _xml

pytestmark = pytest.mark.gpu


def test_model_create_refuses_what_the_kernels_do_not_implement():
  cases = [(_xml("ellipsoid", "ellipsoid"), "no narrowphase"),
           (_xml(opt='noslip_iterations="2"'), "noslip_iterations")]
  if not C.HAVE_PGS_CG:
    cases.append((_xml(opt='solver="PGS"'), "only the Newton solver"))
  for xml, msg in cases:
    m = C.compile_xml(xml, validate=False)
    with pytest.raises(binding.Error, match=msg):
      binding.Model(m)


def test_reenabling_contacts_on_a_model_with_unserved_pairs_is_rolled_back():
  m = C.compile_xml(_xml("ellipsoid", "ellipsoid", flags='<flag contact="disable"/>'))
  model = binding.Model(m)
  flags = int(m.opt_disableflags)
  with pytest.raises(binding.Error, match="no narrowphase"):
    model.set_field("opt_disableflags", flags & ~M.DSBL["contact"])
  out = np.zeros(1, dtype=np.int32)
  binding.check(binding.load().bp_model_get_field(model.ptr, b"opt_disableflags",
                                                  out.ctypes.data, 4))
  assert int(out[0]) == flags
  with pytest.raises(binding.Error, match="structural"):
    model.set_field("body_parentid", m.body_parentid)
  model.free()


def _compare(m_edit, model, nenv=24, nsteps=3, seed=4, lanes=8):
  """forward + steps on the GPU model (already edited through set_field) against oracles
  built from the edited FlatModel."""
  from oracle.oracle_py import Oracle, OracleModel
  qpos, qvel, ctrl = PU.random_states(m_edit, nenv, seed)
  batch = binding.Batch(model, nenv, 64, 16, lanes)
  batch.write("qpos", qpos); batch.write("qvel", qvel); batch.write("ctrl", ctrl)
  mo = m_edit.copy()
  mo.arrays["njmax"][:] = 64; mo.arrays["nconmax"][:] = 16
  om = OracleModel(mo)
  oracles = []
  for e in range(nenv):
    o = Oracle(om); o.reset()
    o.qpos[:] = qpos[e]; o.qvel[:] = qvel[e]; o.ctrl[:] = ctrl[e]
    oracles.append(o)
  names = (PU.F64_FIELDS + PU.EFC_F64 + PU.EFC_I32 + ["efc_J"] + [n for n, _ in PU.CON_F64] +
           PU.CON_I32 + PU.INT_SCALARS)
  cmp = PU.Comparison()
  for step in range(nsteps + 1):
    if step == 0:
      batch.forward(); [o.forward() for o in oracles]
    else:
      batch.step(1); [o.step() for o in oracles]
    batch.sync()
    gpu = {n: batch.read(n) for n in names}
    for e, o in enumerate(oracles):
      PU.compare_env(cmp, m_edit, o, gpu, e, m_edit.nv | 1)
  stats = dict(nefc=float(gpu["nefc"].mean()), nf=float(gpu["nf"].mean()),
               ncon=float(gpu["ncon"].mean()))
  batch.free()
  return cmp, stats


def _assert(cmp):
  bad = {k: v for k, v in cmp.int_mismatch.items() if v}
  assert not bad, cmp.report()
  assert cmp.worst() <= 1e-8 and cmp.worst_elem() <= 1.0, cmp.report()


def test_run_time_edits_reach_the_derived_arrays():
  m = PU.load_model("cheetah")
  model = binding.Model(m)
  # actuator_gear feeds dof_actcoef (qfrc_actuator) as well as actuator_length
  gear = m.actuator_gear.copy(); gear[:, 0] *= np.linspace(0.5, 2.0, m.nu)
  model.set_field("actuator_gear", gear)
  # friction loss appears on two dofs: new friction rows (fric_dof), LEAN shape still fine
  fl = m.dof_frictionloss.copy(); fl[3] = 0.4; fl[6] = 0.2
  model.set_field("dof_frictionloss", fl)
  # one joint loses its limit (lim_jnt)
  lim = m.jnt_limited.copy(); lim[4] = 0
  model.set_field("jnt_limited", lim)
  # fatter feet: geom_rbound must follow geom_size or the broadphase drops contacts
  size = m.geom_size.copy()
  for g in range(m.ngeom):
    if m.geom_type[g] == M.GEOM_CAPSULE:
      size[g, 0] *= 2.5
  model.set_field("geom_size", size)
  me = model.flat.copy()      # binding keeps the host copy in step
  np.testing.assert_array_equal(me.geom_size, size)
  np.testing.assert_allclose(me.geom_rbound, [C._rbound(int(t), s) for t, s in
                                              zip(me.geom_type, me.geom_size)], rtol=1e-15)
  cmp, stats = _compare(me, model)
  _assert(cmp)
  assert stats["nf"] == 2 and stats["ncon"] > 0.5
  model.free()
