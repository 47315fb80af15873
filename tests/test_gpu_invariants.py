"""Engine-independent checks of the CUDA path ("extra checks"): properties a
correct mj_forward / mj_step must have whatever the reference engine's digits are —
solver optimality (KKT), contact complementarity, consistency of the constraint force
with J, symmetric positive-definite mass matrix, momentum of a free-floating body.
They are evaluated at BASELINE.json's full batch sizes, where the oracle is too slow."""

import numpy as np
import pytest

import parity_util as PU

pytestmark = pytest.mark.gpu

CN_LIMIT, CN_FRICTIONLESS, CN_PYRAMIDAL = 3, 4, 5          # csrc/bp_types.h efc types
FULL = {"cheetah": dict(nenv=16384, njmax=64, nconmax=16, lanes=8),
        "humanoid": dict(nenv=8192, njmax=120, nconmax=40, lanes=16)}


def dense_mass_matrix(m, qM):
  """[nenv, nv, nv] from the sparse qM layout (row i: itself then its ancestors)."""
  n = qM.shape[0]
  M = np.zeros((n, m.nv, m.nv))
  for i in range(m.nv):
    a, j = m.dof_Madr[i], i
    while j >= 0:
      M[:, i, j] = qM[:, a]; M[:, j, i] = qM[:, a]
      a += 1; j = m.dof_parentid[j]
  return M


def settled_batch(domain, steps=60):
  from dm_control_b200 import binding
  cfg = FULL[domain]
  m = PU.load_model(domain)
  model = binding.Model(m)
  b = binding.Batch(model, cfg["nenv"], cfg["njmax"], cfg["nconmax"], cfg["lanes"])
  rng = np.random.RandomState(3)
  qpos, qvel, ctrl = PU.random_states(m, cfg["nenv"], 21)
  b.write("qpos", qpos); b.write("qvel", 0.2 * qvel)
  b.forward()
  for k in range(steps // 10):
    b.write("ctrl", rng.uniform(-1, 1, (cfg["nenv"], m.nu)))
    b.step(10)
  b.forward()                 # qacc / efc_force of the current state
  b.sync()
  return m, model, b


@pytest.mark.parametrize("domain", ["cheetah", "humanoid"])
def test_solver_kkt_and_complementarity_at_full_size(domain):
  m, model, b = settled_batch(domain)
  n = b.nenv
  nefc = b.read("nefc")[:, 0]
  assert nefc.max() > 0
  assert (b.warnings()[:, 4:7] == 0).all()          # no BADQPOS / BADQVEL / BADQACC
  nvp = m.nv | 1
  J = b.read("efc_J").reshape(n, -1, nvp)[:, :, :m.nv]
  force, aref = b.read("efc_force"), b.read("efc_aref")
  etype = b.read("efc_type")
  qacc, qfs, qfc = b.read("qacc"), b.read("qfrc_smooth"), b.read("qfrc_constraint")
  M = dense_mass_matrix(m, b.read("qM"))
  rows = np.arange(J.shape[1])[None, :] < nefc[:, None]
  # mass matrix symmetric positive definite
  assert np.linalg.eigvalsh(M).min() > 0
  # qfrc_constraint = J^T efc_force over the active rows
  jtf = np.einsum("nri,nr->ni", J, np.where(rows, force, 0.0))
  np.testing.assert_allclose(jtf, qfc, rtol=1e-10, atol=1e-9)
  # stationarity of the primal problem: M qacc = qfrc_smooth + J^T f, to solver tolerance
  scale = 1.0 / (m.stat_meaninertia * max(m.nv, 1))
  grad = np.einsum("nij,nj->ni", M, qacc) - qfs - qfc
  resid = scale * np.linalg.norm(grad, axis=1)
  print(domain, "KKT residual max %.3g  p99 %.3g  median %.3g" % (
      resid.max(), np.quantile(resid, 0.99), np.median(resid)))
  assert resid.max() < 1e-8          # opt.tolerance; measured 4e-13 .. 2e-12
  # unilateral rows (limits, contacts): force >= 0 and zero wherever the row is satisfied
  uni = rows & np.isin(etype, (CN_LIMIT, CN_FRICTIONLESS, CN_PYRAMIDAL))
  jar = np.einsum("nri,ni->nr", J, qacc) - aref
  assert (force[uni] >= 0).all()
  assert (np.abs(force[uni & (jar > 1e-9)]) == 0).all()
  # contact rows come in whole pyramids: nefc counts agree with the contact list
  ncon = b.read("ncon")[:, 0]
  assert (ncon <= FULL[domain]["nconmax"]).all() and (nefc <= FULL[domain]["njmax"]).all()
  b.free(); model.free()


def com_velocity_drift(integrator, dt_div, gravity=False, n=256, steps=40):
  from dm_control_b200 import binding
  from dm_control_b200 import model as MD
  m = PU.load_model("humanoid")
  flags = int(m.opt_disableflags) | MD.DSBL["contact"] | (0 if gravity else MD.DSBL["gravity"])
  m.arrays["opt_disableflags"][:] = flags
  m.arrays["opt_integrator"][:] = integrator
  m.arrays["opt_timestep"][:] = m.opt_timestep / dt_div
  model = binding.Model(m)
  b = binding.Batch(model, n, 120, 40, 16)
  qpos, qvel, ctrl = PU.random_states(m, n, 8)
  b.write("qpos", qpos); b.write("qvel", qvel); b.write("ctrl", ctrl)
  b.forward()
  adr = m.sensor_adr[m.name2id("torso_subtreelinvel", "sensor")]
  v0 = b.read("sensordata")[:, adr:adr + 3]
  (b.mj_step if integrator == 1 else b.step)(steps * dt_div)
  b.forward()
  v1 = b.read("sensordata")[:, adr:adr + 3]
  assert (b.read("ncon") == 0).all() and (b.warnings()[:, 4:7] == 0).all()
  b.free(); model.free()
  return v1 - v0, steps * dt_div * m.opt_timestep


def test_free_body_momentum_without_external_forces():
  """Humanoid with gravity and contacts disabled: joint springs, dampers, limits and
  actuators are internal forces, so the centre-of-mass velocity (subtree_linvel of the
  root) is conserved by the continuous dynamics.  The discrete drift must therefore
  vanish at the integrator's order: first order in dt for semi-implicit Euler, far
  smaller for RK4 (the CPU oracle shows the same figures: 2-8e-2, /4, 1e-4-1e-3)."""
  euler, _ = com_velocity_drift(0, 1)
  euler4, _ = com_velocity_drift(0, 4)
  rk4, _ = com_velocity_drift(1, 1)
  e1, e4, r = (np.abs(x).max(axis=1) for x in (euler, euler4, rk4))
  print("COM velocity drift: euler %.3g  euler dt/4 %.3g  rk4 %.3g (medians)" % (
      np.median(e1), np.median(e4), np.median(r)))
  assert 2.5 < np.median(e1 / e4) < 6.0
  assert np.median(r) < 0.05 * np.median(e1) and r.max() < 0.1 * e1.max()
  # with gravity back on the COM falls at g: the sensor is live, the check not vacuous
  grav, T = com_velocity_drift(1, 1, gravity=True)
  np.testing.assert_allclose(grav[:, 2], -9.81 * T, rtol=0, atol=3 * r.max() + 1e-9)
