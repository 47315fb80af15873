"""Pins the CPU oracle against every known answer the reference holds for the
physics path.  Citations are under /root/reference/dm_control/."""

import os
import numpy as np
import pytest

from dm_control_b200 import mjcf_compiler as C
from oracle.oracle_py import Oracle

# mujoco/README.md:9-49 — the only digit-level value the reference publishes.
README_XML = """
<mujoco>
  <worldbody>
    <light name="top" pos="0 0 1.5"/>
    <geom name="floor" type="plane" size="1 1 .1"/>
    <body name="box" pos="0 0 .3">
      <joint name="up_down" type="slide" axis="0 0 1"/>
      <geom name="box" type="box" size=".2 .2 .2" rgba="1 0 0 1"/>
      <geom name="sphere" pos=".2 .2 .2" size=".1" rgba="0 1 0 1"/>
    </body>
  </worldbody>
</mujoco>"""


def test_readme_golden_geom_positions():
  m = C.compile_xml(README_XML)
  o = Oracle(m)
  o.reset()
  o.qpos[m.name2id("up_down", "joint")] = 0.5
  o.forward()                                    # reset_context exit
  gx = o.geom_xpos.reshape(-1, 3)
  np.testing.assert_allclose(gx, [[0, 0, 0], [0, 0, .8], [.2, .2, 1.]], atol=1e-15)
  while o.time[0] < 1.0:
    o.step()
  z = o.geom_xpos.reshape(-1, 3)[[m.name2id("box", "geom"),
                                  m.name2id("sphere", "geom")], 2]
  # README prints 8 decimals: [ 0.19996362  0.39996362]
  np.testing.assert_allclose(z, [0.19996362, 0.39996362], atol=5e-9)


# wrapper/core_test.py:328-367 (testDisableFlags) — cube on plane with a touch
# sensor: 100 steps -> qvel ~ 0, touch ~ 9.81 N.
CUBE_XML = """
<mujoco>
  <option{opt}/>
  <worldbody>
    <geom name="floor" type="plane" pos="0 0 0" size="10 10 0.1"/>
    <body name="cube" pos="0 0 0.1">
      <geom type="box" size="0.1 0.1 0.1" mass="1"/>
      <site name="cube_site" type="box" size="0.1 0.1 0.1"/>
      <joint type="slide"/>
    </body>
  </worldbody>
  <sensor>
    <touch name="touch_sensor" site="cube_site"/>
  </sensor>
</mujoco>"""


def test_touch_sensor_reads_weight():
  m = C.compile_xml(CUBE_XML.format(opt=""))
  o = Oracle(m)
  o.forward()
  for _ in range(100):
    o.mj_step()
  o.forward()
  np.testing.assert_almost_equal(o.qvel[0], 0, decimal=4)
  np.testing.assert_almost_equal(o.sensordata[0], 9.81, decimal=2)


def test_disable_contact_and_gravity():
  xml = CUBE_XML.format(opt="").replace(
      "<option/>", '<option><flag contact="disable" gravity="disable"/></option>')
  o = Oracle(C.compile_xml(xml))
  for _ in range(10):
    o.mj_step()
  o.forward()
  assert o.sensordata[0] == 0 and o.qvel[0] == 0
  xml = CUBE_XML.format(opt="").replace(
      "<option/>", '<option><flag contact="disable"/></option>')
  o = Oracle(C.compile_xml(xml))
  for _ in range(10):
    o.mj_step()
  assert o.qvel[0] < -0.1                       # core_test.py:360-367: falls


# core_test.py:461-484 (testContactForce): box on floor, 500 steps, the summed
# normal force equals the weight to 7 places.
def test_contact_force_equals_weight():
  xml = """
  <mujoco>
    <worldbody>
      <geom name="floor" type="plane" size="1 1 1"/>
      <body name="block" pos="0 0 0.1">
        <freejoint/>
        <geom type="box" size="0.1 0.1 0.1" mass="2.5"/>
      </body>
    </worldbody>
  </mujoco>"""
  m = C.compile_xml(xml)
  o = Oracle(m)
  for _ in range(500):
    o.mj_step()
  o.forward()
  ncon, nefc = int(o.ncon[0]), int(o.nefc[0])
  assert ncon == 4 and nefc == 16
  normal = o.efc_force[:nefc].sum()      # pyramid edges: normal = sum of rows
  np.testing.assert_almost_equal(normal, 2.5 * 9.81, decimal=7)
  # time advances by exactly `timestep` each step (core_test.py:159-172)
  t = o.time[0]
  o.mj_step()
  assert o.time[0] == t + m.opt_timestep


# core_test.py:408-459 (testObjectVelocity): cart at 1 m/s, pole hinge at 1 rad/s.
def test_object_velocity_cartpole():
  xml = """
  <mujoco>
    <worldbody>
      <body name="cart">
        <joint name="slider" type="slide" axis="1 0 0"/>
        <geom type="box" size=".2 .1 .1" mass="1"/>
        <body name="pole">
          <joint name="hinge" type="hinge" axis="0 1 0"/>
          <geom name="pole" type="capsule" fromto="0 0 0 0 0 1" size=".05" mass=".1"/>
          <site name="tip" pos="0 0 1"/>
          <site name="mid" pos="0 0 .5"/>
        </body>
      </body>
    </worldbody>
    <sensor>
      <velocimeter name="tip_vel" site="tip"/>
      <gyro name="tip_gyro" site="tip"/>
      <velocimeter name="mid_vel" site="mid"/>
    </sensor>
  </mujoco>"""
  m = C.compile_xml(xml)
  o = Oracle(m)
  o.qvel[:] = [1.0, 1.0]
  o.step1()
  # world-frame == site-frame at qpos=0: tip moves at 1 + 1*1 = 2 m/s ... the
  # reference's numbers use slider=.5: check the same linear structure here.
  np.testing.assert_allclose(o.sensordata[0:3], [2.0, 0, 0], atol=1e-12)
  np.testing.assert_allclose(o.sensordata[3:6], [0, 1.0, 0], atol=1e-12)
  np.testing.assert_allclose(o.sensordata[6:9], [1.5, 0, 0], atol=1e-12)
  o.qvel[:] = [0.5, 1.0]
  o.step1()
  np.testing.assert_allclose(o.sensordata[0:3], [1.5, 0, 0], atol=1e-12)
  # rotated pole: local-frame velocity of the tip about the hinge only
  o.qpos[:] = [0, np.pi]
  o.qvel[:] = [0.0, 0.5]
  o.step1()
  np.testing.assert_allclose(o.sensordata[0:3], [0.5, 0, 0], atol=1e-12)


# engine_test.py:491-504: accelerometer reads +9.81 along z after reset;
# actuator_force is zero with actuation disabled and gain*ctrl otherwise.
def test_accelerometer_and_actuation():
  xml = """
  <mujoco>
    <worldbody>
      <geom type="plane" size="1 1 .1"/>
      <body name="b" pos="0 0 .1">
        <joint name="j" type="slide" axis="0 0 1"/>
        <geom type="sphere" size=".1"/>
        <site name="s"/>
      </body>
    </worldbody>
    <actuator><motor name="m" joint="j" gear="1"/></actuator>
    <sensor><accelerometer name="acc" site="s"/></sensor>
  </mujoco>"""
  m = C.compile_xml(xml)
  o = Oracle(m)
  for _ in range(300):
    o.mj_step()
  o.forward()
  np.testing.assert_allclose(o.sensordata[:3], [0, 0, 9.81], atol=1e-5)
  o.ctrl[0] = 1.0
  o.forward()
  assert o.actuator_force[0] == 1.0


def test_bad_ctrl_and_bad_qpos_warnings():
  """engine_test.py:400-436: NaN ctrl -> BADCTRL; inf/nan/huge qpos -> BADQPOS."""
  m = C.compile_xml(CUBE_XML.format(opt="").replace(
      "</worldbody>", "</worldbody><actuator><motor joint='j0'/></actuator>"
  ).replace('<joint type="slide"/>', '<joint name="j0" type="slide"/>'))
  o = Oracle(m)
  o.ctrl[0] = np.nan
  o.mj_step()
  assert o.warning[7] == 1
  for bad in (np.inf, np.nan, 1e15):
    o = Oracle(m)
    o.qpos[0] = bad
    o.mj_step()
    assert o.warning[4] == 1
    assert np.isfinite(o.qpos).all()


def test_determinism_and_copy():
  """engine_test.py:462-484: identical stepping gives exactly equal state."""
  m = C.compile_xml(README_XML)
  a, b = Oracle(m), Oracle(m)
  for o in (a, b):
    o.qpos[0] = 0.25
    o.forward()
  for _ in range(50):
    a.step(); b.step()
  assert a.time[0] == b.time[0]
  np.testing.assert_array_equal(a.qpos, b.qpos)
  np.testing.assert_array_equal(a.xpos, b.xpos)


PENDULUM2 = """
<mujoco>
  <option timestep="0.002" integrator="RK4"><flag energy="enable" contact="disable"/></option>
  <worldbody>
    <body pos="0 0 2">
      <joint name="a" type="hinge" axis="0 1 0"/>
      <geom type="capsule" fromto="0 0 0 0 0 -1" size=".05" mass="1"/>
      <body pos="0 0 -1">
        <joint name="b" type="hinge" axis="0 1 0"/>
        <geom type="capsule" fromto="0 0 0 0 0 -1" size=".04" mass="1"/>
      </body>
    </body>
  </worldbody>
</mujoco>"""


def test_energy_conservation_rk4():
  """Engine-independent: undamped double pendulum under RK4 conserves energy."""
  m = C.compile_xml(PENDULUM2)
  o = Oracle(m)
  o.qpos[:] = [1.0, -0.5]
  o.forward()
  e0 = o.energy.sum()
  for _ in range(2000):
    o.mj_step()
  o.forward()
  assert abs(o.energy.sum() - e0) < 1e-7 * max(1, abs(e0))


def test_mass_matrix_and_bias_consistency():
  """CRB inertia equals the dense J^T I J sum of the host compiler, is PD, and
  M*qacc_smooth == qfrc_smooth (SURVEY §8c extra checks)."""
  m = C.compile_xml(PENDULUM2.replace('integrator="RK4"', ""))
  o = Oracle(m)
  rng = np.random.RandomState(0)
  o.qpos[:] = rng.uniform(-2, 2, m.nq)
  o.qvel[:] = rng.uniform(-2, 2, m.nv)
  o.forward()
  kin = C.host_kinematics(m, np.array(o.qpos))
  Md = C.host_mass_matrix(m, kin)
  Mo = np.zeros((m.nv, m.nv))
  for i in range(m.nv):
    a, j = m.dof_Madr[i], i
    while j >= 0:
      Mo[i, j] = Mo[j, i] = o.qM[a]
      a += 1
      j = m.dof_parentid[j]
  np.testing.assert_allclose(Mo, Md, rtol=1e-12, atol=1e-14)
  assert np.all(np.linalg.eigvalsh(Mo) > 0)
  np.testing.assert_allclose(Mo @ o.qacc_smooth, o.qfrc_smooth, rtol=1e-11, atol=1e-12)
  # bias force against a finite-difference Lagrangian: d/dt(dL/dv) - dL/dq
  def lagr_terms(q, v):
    o.qpos[:] = q; o.qvel[:] = v; o.forward()
    return o.energy[1] - o.energy[0]
  q, v, eps = np.array(o.qpos), np.array(o.qvel), 1e-6
  bias = np.array(o.qfrc_bias)
  Mq = Mo.copy()
  dLdq = np.array([(lagr_terms(q + eps * np.eye(2)[k], v) -
                    lagr_terms(q - eps * np.eye(2)[k], v)) / (2 * eps)
                   for k in range(2)])
  # Mdot*v via finite difference of M along v
  def Mat(qq):
    o.qpos[:] = qq; o.qvel[:] = v; o.forward()
    Mx = np.zeros((2, 2))
    for i in range(2):
      a, j = m.dof_Madr[i], i
      while j >= 0:
        Mx[i, j] = Mx[j, i] = o.qM[a]; a += 1; j = m.dof_parentid[j]
    return Mx
  Mdot = (Mat(q + eps * v) - Mat(q - eps * v)) / (2 * eps)
  np.testing.assert_allclose(bias, Mdot @ v - dLdq, rtol=1e-6, atol=1e-6)


_REF_HUMANOID = "/root/reference/dm_control/mujoco/testing/assets/humanoid.xml"


@pytest.mark.skipif(not os.path.exists(_REF_HUMANOID), reason="reference tree not present")
def test_compiler_object_ids_match_reference_expectations():
  """core_test.py:226-235 testNamesIds: ids MuJoCo's compiler assigns in the reference's
  own test humanoid (body / joint ordering of the MJCF compiler)."""
  from dm_control_b200 import mjcf_compiler as C
  # structure only (the asset asks for solver="PGS"): no simulation here
  m = C.compile_xml(open(_REF_HUMANOID).read(), validate=False)
  assert m.name2id("right_foot", "body") == 6
  assert m.name2id("left_knee", "joint") == 11
  assert m.names["body"][6] == "right_foot" and m.names["joint"][11] == "left_knee"
  with pytest.raises(ValueError, match="exists"):
    m.name2id("nonexistent_body_name", "body")


@pytest.mark.parametrize("n_bodies,n_actuators", [(2, 1), (6, 2)])
def test_lqr_riccati_cost(n_bodies, n_actuators):
  """suite/lqr_test.py:30-59 (lqr_2_1, lqr_6_2): under the optimal linear policy of the
  discrete Riccati equation (suite/lqr_solver.py, built from the engine's mass matrix,
  joint stiffness and actuator moment) the measured total cost equals .5 x0' P x0 to
  rtol 1e-3 -- a known answer that the mass matrix, spring forces, actuation and the
  Euler integrator of the oracle have to reproduce together."""
  import math
  from dm_control_b200 import mjcf_compiler as C
  from dm_control_b200.suite import lqr
  rng = np.random.RandomState(3)
  m = C.compile_xml(lqr.make_model(n_bodies, n_actuators, rng))
  o = Oracle(m)
  o.forward()
  np.testing.assert_allclose(o.actuator_moment.reshape(m.nu, m.nv),
                             np.eye(n_bodies)[:n_actuators], atol=0)
  p, k, beta = lqr.solve(m, o.qM.copy())
  tolerance = 1e-3
  n_steps = int(math.ceil(math.log10(tolerance) / math.log10(beta)))
  unit = rng.randn(m.nq)
  o.reset()
  o.qpos[:] = np.sqrt(2) * unit / np.linalg.norm(unit)        # lqr.py:229-233
  o.forward()
  x0 = np.hstack((o.qpos, o.qvel)).copy()
  total_loss, reward = 0.0, None
  for _ in range(n_steps):
    x = np.hstack((o.qpos, o.qvel))
    total_loss += 1 - (reward or 0.0)
    o.ctrl[:] = k.dot(x)
    o.step()
    reward = 1 - (0.5 * o.qpos.dot(o.qpos) + 0.5 * o.ctrl.dot(o.ctrl) * lqr.CONTROL_COST_COEF)
  np.testing.assert_allclose(.5 * x0.dot(p).dot(x0), total_loss, rtol=tolerance)



def test_box_stack_at_rest_carries_the_weight_above_it():
  """Known answer for the box narrowphase (sphere / capsule / box against box), in the
  style of core_test.py:461-484: at rest the summed normal force of every support equals
  the weight it carries -- floor: b1 + b2 + sphere, b1 <- b2 + sphere, b2 <- sphere, and a
  capsule lying on a box rests on two contacts."""
  from box_models import BOX_STACK
  m = C.compile_xml(BOX_STACK)
  o = Oracle(m)
  for _ in range(1500):
    o.mj_step()
  o.forward()
  ncon = int(o.ncon[0])
  g1, g2 = o.contact_geom1[:ncon], o.contact_geom2[:ncon]
  adr, dim = o.contact_efc_address[:ncon], o.contact_dim[:ncon]

  def support(a, b):
    idx = [i for i in range(ncon) if (g1[i], g2[i]) == (a, b)]
    return len(idx), sum(o.efc_force[adr[i]:adr[i] + 2 * (dim[i] - 1)].sum() for i in idx)

  g = 9.81
  assert support(0, 1) == (4, pytest.approx((2.5 + 1.5 + 0.5) * g, abs=1e-5))    # floor - b1
  assert support(1, 2) == (4, pytest.approx((1.5 + 0.5) * g, abs=1e-5))          # b1 - b2
  assert support(3, 2) == (1, pytest.approx(0.5 * g, abs=1e-5))                  # sphere - b2
  assert support(0, 4) == (4, pytest.approx((1.0 + 0.3) * g, abs=1e-5))          # floor - c
  assert support(5, 4) == (2, pytest.approx(0.3 * g, abs=1e-5))                  # capsule - c
  assert ncon == 15


@pytest.mark.parametrize("condim,expected_torques", [(3, [False, False, False]),
                                                      (4, [True, False, False]),
                                                      (6, [True, True, True])])
def test_contact_torque_sparsity_by_condim(condim, expected_torques):
  """core_test.py:494-530 testContactTorque: a spinning ball on the floor; sliding friction
  only gives no contact torque, condim 4 adds the torsional one, condim 6 all three."""
  xml = """
  <mujoco>
    <worldbody>
      <geom name='floor' type='plane' size='1 1 1'/>
      <body name='ball' pos='0 0 .1'>
        <freejoint/>
        <geom name='ball' size='.1' friction='1 .1 .1'/>
      </body>
    </worldbody>
  </mujoco>""".replace("<geom name=", "<geom condim='%d' name=" % condim)   # buffers sized for it
  m = C.compile_xml(xml)
  assert (np.asarray(m.geom_condim) == condim).all()
  o = Oracle(m)
  o.qvel[3:] = (1., 1., 1.)
  for _ in range(10):
    o.mj_step()
  assert int(o.ncon[0]) == 1
  force, torque = o.contact_force(0)
  np.testing.assert_array_equal(torque != 0, expected_torques)
  assert force[0] > 0


def test_free_humanoid_momentum_drift_vanishes_at_the_integrators_order():
  """Engine-independent, on the ORACLE (the GPU side holds the same invariant in
  test_gpu_invariants.py): without gravity and contacts every force of the suite humanoid is
  internal, so the continuous dynamics conserve the centre-of-mass velocity; the discrete
  drift is first order in dt for semi-implicit Euler and far smaller for RK4."""
  import parity_util as PU
  from dm_control_b200 import model as MD

  def drift(integrator, dt_div, n=8, steps=40):
    m = PU.load_model("humanoid")
    m.arrays["opt_disableflags"][...] = int(m.opt_disableflags) | MD.DSBL["contact"] | MD.DSBL["gravity"]
    m.arrays["opt_integrator"][...] = integrator
    m.arrays["opt_timestep"][...] = m.opt_timestep / dt_div
    qpos, qvel, ctrl = PU.random_states(m, n, 8)
    adr = m.sensor_adr[m.name2id("torso_subtreelinvel", "sensor")]
    out = []
    for e in range(n):
      o = Oracle(m)
      o.reset(); o.qpos[:] = qpos[e]; o.qvel[:] = qvel[e]; o.ctrl[:] = ctrl[e]
      o.forward()
      v0 = np.array(o.sensordata[adr:adr + 3])
      for _ in range(steps * dt_div):
        o.mj_step()
      o.forward()
      assert int(o.ncon[0]) == 0
      out.append(np.abs(np.array(o.sensordata[adr:adr + 3]) - v0).max())
    return np.array(out)

  e1, e4, r = drift(0, 1), drift(0, 4), drift(1, 1)
  assert 2.5 < np.median(e1 / e4) < 6.0
  assert np.median(r) < 0.05 * np.median(e1)


def test_efc_J_is_the_derivative_of_efc_pos():
  """Finite differences: for limit and contact rows, efc_pos(q + eps dq) - efc_pos(q) =
  eps J dq (dq in dof space, integrated with the engine's own position integration)."""
  import parity_util as PU
  m = PU.load_model("cheetah")
  qpos, qvel, ctrl = PU.random_states(m, 6, 5)
  om = C and __import__("oracle.oracle_py", fromlist=["OracleModel"]).OracleModel(m)
  checked = 0
  for e in range(6):
    o = Oracle(om)
    o.reset(); o.qpos[:] = qpos[e]; o.forward()
    nefc, nv = int(o.nefc[0]), m.nv
    if nefc == 0:
      continue
    J = np.array(o.efc_J[:nefc * nv]).reshape(nefc, nv)
    pos0, ids0 = np.array(o.efc_pos[:nefc]), np.array(o.efc_id[:nefc])
    types0 = np.array(o.efc_type[:nefc])
    rng = np.random.RandomState(e)
    dq = rng.randn(nv)
    eps = 1e-7
    o2 = Oracle(om)
    o2.reset(); o2.qpos[:] = qpos[e] + eps * dq       # cheetah: slide / hinge joints only
    o2.forward()
    if int(o2.nefc[0]) != nefc or not np.array_equal(np.array(o2.efc_id[:nefc]), ids0):
      continue
    num = (np.array(o2.efc_pos[:nefc]) - pos0) / eps
    # pyramidal contact rows share the contact's distance: d pos = Jn dq, and the rows hold
    # Jn +- mu Jt, so consecutive row PAIRS average to Jn
    ana = J @ dq
    for r in range(nefc):
      if types0[r] == 3:                                # joint limit
        np.testing.assert_allclose(num[r], ana[r], rtol=1e-5, atol=1e-6); checked += 1
    con = np.flatnonzero(types0 == 6)
    for r in con[::2]:
      np.testing.assert_allclose(num[r], 0.5 * (ana[r] + ana[r + 1]), rtol=1e-4, atol=1e-5)
      checked += 1
  assert checked > 5
