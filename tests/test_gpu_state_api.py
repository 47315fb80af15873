"""Physics.copy / pickling / resumable state, contact_force and object_velocity on the batched
backend (engine.py:247-266,332-339; wrapper/core.py:653-685,713-769)."""

import pickle

import numpy as np
import pytest

import parity_util as PU
from dm_control_b200 import suite
from dm_control_b200.physics import BatchedPhysics

pytestmark = pytest.mark.gpu


def _physics(domain, n=16, seed=2, steps=5):
  m = PU.load_model(domain)
  p = BatchedPhysics(m, n)
  qpos, qvel, ctrl = PU.random_states(m, n, seed)
  with p.reset_context():
    p.data.qpos = qpos; p.data.qvel = qvel
  p.set_control(ctrl)
  p.step(steps)
  return p, ctrl


@pytest.mark.parametrize("domain", ["cheetah", "humanoid", "cartpole"])
def test_copy_then_steps_are_bit_identical(domain):
  """engine_test.py:462-484 testCopyOrPickle: a copy steps exactly like the original; with
  share_model the model is the same object, otherwise a separate equal one."""
  p, ctrl = _physics(domain)
  for share in (False, True):
    q = p.copy(share_model=share)
    assert (q._dev_model is p._dev_model) == share
    np.testing.assert_array_equal(q.data.qpos, p.data.qpos)
    np.testing.assert_array_equal(q.data.xpos, p.data.xpos)
    p2 = p.copy()
    q.step(10); p2.step(10)
    for f in ("qpos", "qvel", "qacc", "qacc_warmstart", "sensordata", "efc_force", "time"):
      np.testing.assert_array_equal(getattr(q.data, f), getattr(p2.data, f), err_msg=f)
    q.free(); p2.free()
  p.free()


@pytest.mark.parametrize("domain", ["cheetah", "quadruped_escape", "reacher"])
def test_pickle_round_trip_resumes_bit_exactly(domain):
  """qacc_warmstart and time travel with the state, per-env terrains
  and target frames too: the unpickled physics continues as the original does."""
  if domain == "cheetah":
    p, ctrl = _physics(domain)
  else:
    env = suite.load("quadruped" if domain == "quadruped_escape" else domain,
                     "escape" if domain == "quadruped_escape" else "easy",
                     task_kwargs={"random": 3}, nenv=6)
    env.reset()
    rng = np.random.RandomState(0)
    for _ in range(3):
      env.step(rng.uniform(-1, 1, (6, env.action_spec().shape[0])))
    p = env.physics
  blob = pickle.dumps(p)
  q = pickle.loads(blob)
  np.testing.assert_array_equal(q.data.qacc_warmstart, p.data.qacc_warmstart)
  np.testing.assert_array_equal(q.data.time, p.data.time)
  p.step(7); q.step(7)
  for f in ("qpos", "qvel", "qacc", "sensordata", "time", "ncon"):
    np.testing.assert_array_equal(getattr(q.data, f), getattr(p.data, f), err_msg=f)
  q.free(); p.free()


def test_contact_force_and_object_velocity_match_oracle():
  from oracle.oracle_py import Oracle, OracleModel
  m = PU.load_model("humanoid")
  n = 24
  p = BatchedPhysics(m, n, njmax=120, nconmax=40)
  qpos, qvel, ctrl = PU.random_states(m, n, 7)
  qpos[:, 2] = 0.6                          # low enough to touch the floor
  with p.reset_context():
    p.data.qpos = qpos; p.data.qvel = qvel
  p.set_control(ctrl)
  p.step(3)
  p.forward()
  mo = m.copy(); mo.arrays["njmax"][:] = 120; mo.arrays["nconmax"][:] = 40
  om = OracleModel(mo)
  oracles = []
  for e in range(n):
    o = Oracle(om); o.reset()
    o.qpos[:] = qpos[e]; o.qvel[:] = qvel[e]; o.ctrl[:] = ctrl[e]
    o.forward()
    for _ in range(3):
      o.step()
    o.forward()
    oracles.append(o)
  ncon = p.data.ncon
  assert ncon.max() >= 2
  for ci in range(int(ncon.max())):
    got = p._batch.contact_force(ci)
    for e, o in enumerate(oracles):
      if ci < ncon[e]:
        np.testing.assert_allclose(got[e], o.contact_force(ci), rtol=1e-8, atol=1e-9)
      else:
        assert np.isnan(got[e]).all()
  with pytest.raises(ValueError, match="contact_id"):
    p.data.contact_force(int(ncon.min()))           # not a contact in every env
  hand = m.name2id("left_hand", "body")
  for otype, oid, kind in ((1, hand, "body"), (2, hand, "xbody"), (5, 3, "geom")):
    for local in (False, True):
      got = p.data.object_velocity(oid, kind, local_frame=local)
      for e, o in enumerate(oracles):
        np.testing.assert_allclose(got[e], o.object_velocity(oid, otype, local),
                                   rtol=1e-8, atol=1e-9)
  np.testing.assert_array_equal(p.data.object_velocity("left_hand", "body"),
                                p.data.object_velocity(hand, 1))
  p.free()


def test_from_byte_string_and_reload(tmp_path):
  """engine.py:429-433,447-488: a Physics built from a model's byte string / binary file steps
  bit for bit like the one it came from; reload_from_xml_string swaps the model behind the
  same object and leaves it in the state of a fresh one."""
  p, ctrl = _physics("walker")
  blob = p.model.to_bytes()
  path = str(tmp_path / "walker.bpb")
  p.model.save_binary(path)
  for q in (BatchedPhysics.from_byte_string(blob, nenv=p.nenv),
            BatchedPhysics.from_binary_path(path, nenv=p.nenv)):
    q.set_full_state(p.get_full_state())
    p2 = p.copy()
    q.step(7); p2.step(7)
    for f in ("qpos", "qvel", "qacc", "sensordata", "time"):
      np.testing.assert_array_equal(getattr(q.data, f), getattr(p2.data, f), err_msg=f)
    assert q.named.data.xpos["torso"].shape == (p.nenv, 3)
    q.free(); p2.free()
  xml = """<mujoco><worldbody><body name="b"><joint name="j" type="hinge" axis="0 1 0"/>
  <geom name="g" type="capsule" fromto="0 0 0 0 0 -1" size=".05"/></body></worldbody></mujoco>"""
  p.reload_from_xml_string(xml)
  assert p.model.nq == 1 and p.data.qpos.shape == (p.nenv, 1) and np.all(p.data.time == 0)
  fresh = BatchedPhysics.from_xml_string(xml, nenv=p.nenv)
  p.step(5); fresh.step(5)
  np.testing.assert_array_equal(p.data.qpos, fresh.data.qpos)
  assert p.named.data.xpos["b"].shape == (p.nenv, 3)
  p.free(); fresh.free()
