"""mjcf.Physics.bind on the batched backend (/root/reference/dm_control/mjcf/physics.py:469-690;
the cases follow mjcf/physics_test.py: attribute names by element kind, reads equal to named
indexing, writes that reach the simulation, dirty -> forward on the next derived read, empty
and mixed bindings)."""

import numpy as np
import pytest

import parity_util as PU
from dm_control_b200 import mjcf

pytestmark = pytest.mark.gpu


@pytest.fixture
def physics():
  p = mjcf.Physics(PU.load_model("cheetah"), 5)
  yield p
  p.free()


def test_bind_reads_equal_named_indexing(physics):
  m = physics.model
  joints = mjcf.elements(m, "joint")
  geoms = mjcf.elements(m, "geom")
  b = physics.bind(joints[3])
  np.testing.assert_array_equal(b.qpos, physics.named.data.qpos[joints[3].full_identifier])
  assert b.qpos.shape == (5, 1) and b.element_id == 3
  np.testing.assert_array_equal(b.range, m.jnt_range[3])                 # model: no env axis
  np.testing.assert_array_equal(b.damping, m.dof_damping[m.jnt_dofadr[3]:m.jnt_dofadr[3] + 1])
  bg = physics.bind(geoms[:3])
  names = [g.full_identifier for g in geoms[:3]]
  np.testing.assert_array_equal(bg.xpos, physics.named.data.geom_xpos[names])
  assert bg.xpos.shape == (5, 3, 3)
  np.testing.assert_array_equal(bg.size, m.geom_size[[m.name2id(n, "geom") for n in names]])
  np.testing.assert_array_equal(bg["xpos", "z"], physics.named.data.geom_xpos[names, "z"])
  np.testing.assert_array_equal(bg.element_id, [m.name2id(n, "geom") for n in names])
  ba = physics.bind(mjcf.elements(m, "actuator"))
  assert ba.ctrl.shape == (5, m.nu) and ba.force.shape == (5, m.nu)
  np.testing.assert_array_equal(ba.gear, m.actuator_gear)
  bs = physics.bind(mjcf.Element("subtreelinvel", "torso_subtreelinvel"))
  assert bs.sensordata.shape == (5, 3)
  torso = physics.bind(mjcf.Element("body", "torso"))
  np.testing.assert_array_equal(torso.xpos, physics.named.data.xpos["torso"])
  assert "xquat" in dir(torso) and "mass" in dir(torso)
  with pytest.raises(AttributeError, match="does not have attribute"):
    torso.qpos


def test_bind_writes_reach_the_simulation_and_derived_reads_follow(physics):
  m = physics.model
  rootz = physics.bind(mjcf.Element("joint", "rootz"))
  torso = physics.bind(mjcf.Element("body", "torso"))
  z0 = torso.xpos[:, 2].copy()
  assert not physics.is_dirty
  rootz.qpos = np.linspace(0.1, 0.5, 5)              # one value per env
  assert physics.is_dirty
  z1 = torso.xpos[:, 2]                               # derived read: forward() runs first
  assert not physics.is_dirty
  np.testing.assert_allclose(z1 - z0, np.linspace(0.1, 0.5, 5), atol=1e-12)
  q = rootz.qpos
  q[2] = 0.9                                          # item assignment writes through
  np.testing.assert_array_equal(physics.named.data.qpos["rootz"][:, 0], [0.1, 0.2, 0.9, 0.4, 0.5])
  act = physics.bind(mjcf.elements(m, "actuator")[:2])
  act.ctrl = [[0.3, -0.2]] * 5
  np.testing.assert_array_equal(physics.data.ctrl[:, :2], [[0.3, -0.2]] * 5)
  physics.step()
  assert not physics.is_dirty
  # a model write is shared by every env and reaches the device model
  geom = physics.bind(mjcf.Element("geom", "torso"))
  geom.friction = [0.7, 0.1, 0.1]
  np.testing.assert_array_equal(physics.named.model.geom_friction["torso"], [0.7, 0.1, 0.1])
  assert physics.is_dirty


def test_empty_none_and_mixed_bindings(physics):
  m = physics.model
  assert physics.bind(None) is None
  assert physics.bind([]).xpos.shape == (0,)
  with pytest.raises(ValueError, match="empty binding"):
    physics.bind([]).xpos = 1
  with pytest.raises(ValueError, match="multiple element types"):
    physics.bind([mjcf.Element("joint", "rootz"), mjcf.Element("geom", "torso")])
  assert physics.bind(mjcf.Element("joint", "rootz")) is physics.bind(mjcf.Element("joint", "rootz"))

  class RealLooking:                   # what a PyMJCF element exposes
    tag = "geom"
    full_identifier = "torso"
    class spec:
      namespace = "geom"
  np.testing.assert_array_equal(physics.bind(RealLooking()).xpos,
                                physics.named.data.geom_xpos["torso"])


def test_from_mjcf_model_protocol():
  class Root:
    def to_xml_string(self):
      return """<mujoco><worldbody><body name="b" pos="0 0 1"><joint name="j" type="slide"
        axis="0 0 1"/><geom name="g" size=".1"/></body></worldbody></mujoco>"""
    def get_assets(self):
      return {}
  p = mjcf.Physics.from_mjcf_model(Root(), nenv=3)
  b = p.bind(mjcf.Element("joint", "j"))
  b.qpos = [0.1, 0.2, 0.3]
  np.testing.assert_allclose(p.bind(mjcf.Element("geom", "g")).xpos[:, 2], [1.1, 1.2, 1.3])   # single element: [nenv, 3]
  p.free()
