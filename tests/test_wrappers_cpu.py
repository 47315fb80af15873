"""Host-only tests of dm_control_b200/wrappers.py with a fake batched environment
(modelled on the reference's suite/wrappers/action_scale_test.py and on the worker
contract of scripts/vec_env.py:334-352)."""

import collections

import numpy as np
import pytest

from dm_control_b200 import specs, wrappers


class FakeEnv:
  """Batched env with per-env episodes of `length[e]` steps; records the actions it gets."""

  def __init__(self, nenv=3, nu=2, length=3, lo=-2.0, hi=4.0):
    self.nenv, self.nu = nenv, nu
    self.length = np.broadcast_to(np.asarray(length), (nenv,)).copy()
    self._spec = specs.BoundedArray((nu,), np.float64, lo, hi, name="action")
    self.actions, self.resets, self._t = [], 0, None
    self.env_resets = np.zeros(nenv, dtype=int)
    self.task = type("T", (), {"_seed": None, "_rngs": None})()
    self.closed = False

  def action_spec(self): return self._spec

  def observation_spec(self):
    return collections.OrderedDict(position=specs.Array((2,), np.float64),
                                   height=specs.Array((), np.float64))

  def _obs(self):
    base = np.arange(self.nenv, dtype=np.float64)[:, None]
    return collections.OrderedDict(position=np.concatenate([base, base + self._t[:, None]], axis=1),
                                   height=base[:, 0] * 10 + self._t)

  def reset(self):
    self.resets += 1; self._t = np.zeros(self.nenv)
    self.env_resets += 1
    return specs.TimeStep(np.full(self.nenv, specs.StepType.FIRST), None, None, self._obs())

  def reset_envs(self, mask):
    mask = np.asarray(mask, dtype=bool)
    self._t[mask] = 0
    self.env_resets[mask] += 1
    o = self._obs()
    return np.concatenate([np.asarray(v).reshape(self.nenv, -1) for v in o.values()], axis=1)

  def step(self, action):
    if self._t is None:
      return self.reset()
    self.actions.append(np.array(action)); self._t = self._t + 1
    st = np.where(self._t >= self.length, specs.StepType.LAST, specs.StepType.MID)
    return specs.TimeStep(st, 0.5 * self._t, np.ones(self.nenv), self._obs())

  def close(self): self.closed = True


@pytest.mark.parametrize("minimum,maximum,scaled,orig", [
    (-1.0, 1.0, [[-1, -1], [0, 1]], [[-2, -2], [1, 4]]),
    ([0.0, -1.0], [1.0, 1.0], [[0.5, 0.0]], [[1.0, 1.0]]),
])
def test_action_scale_maps_bounds(minimum, maximum, scaled, orig):
  env = FakeEnv(nenv=len(scaled))
  w = wrappers.ActionScale(env, minimum, maximum)
  w.reset()
  w.step(np.array(scaled, dtype=np.float64))
  np.testing.assert_allclose(env.actions[-1], np.array(orig, dtype=np.float64))
  spec = w.action_spec()
  np.testing.assert_array_equal(spec.minimum, np.broadcast_to(minimum, (2,)))
  np.testing.assert_array_equal(spec.maximum, np.broadcast_to(maximum, (2,)))
  assert w.nenv == env.nenv                       # attribute forwarding


def test_action_scale_rejects_bad_bounds():
  with pytest.raises(ValueError, match="finite"):
    wrappers.ActionScale(FakeEnv(), -np.inf, 1)
  with pytest.raises(ValueError, match="broadcastable"):
    wrappers.ActionScale(FakeEnv(), [0, 0, 0], 1)
  with pytest.raises(ValueError, match="finite"):
    wrappers.ActionScale(FakeEnv(hi=np.inf), -1, 1)
  env = FakeEnv()
  env._spec = specs.Array((2,), np.float64)
  with pytest.raises(ValueError, match="BoundedArray"):
    wrappers.ActionScale(env, -1, 1)


def test_action_noise_matches_reference_formula():
  """suite/wrappers/action_noise_test.py: noise std = scale * range, per env from the
  task's generators, clipped to the bounds; zero scale leaves the actions alone."""
  env = FakeEnv(nenv=4, nu=2, lo=-2.0, hi=4.0)
  env.task.random = [np.random.RandomState(10 + e) for e in range(4)]
  env.reset()
  w = wrappers.ActionNoise(env, scale=0.1)
  act = np.array([[0.0, 1.0], [3.99, -1.99], [4.0, -2.0], [1.0, 1.0]])
  w.step(act)
  expect = np.stack([np.random.RandomState(10 + e).normal(scale=0.1 * 6.0 * np.ones(2))
                     for e in range(4)]) + act
  np.testing.assert_allclose(env.actions[-1], np.clip(expect, -2.0, 4.0))
  assert env.actions[-1].min() >= -2.0 and env.actions[-1].max() <= 4.0
  wrappers.ActionNoise(env, scale=0.0).step(act)
  np.testing.assert_array_equal(env.actions[-1], act)
  with pytest.raises(ValueError):
    wrappers.ActionNoise(FakeEnv(hi=np.inf))
  assert w.action_spec() is env.action_spec()


def test_vec_env_auto_reset_and_terminal_observation():
  env = FakeEnv(nenv=3, length=2)
  v = wrappers.BatchedVecEnv(env)
  assert v.num_envs == 3
  obs = v.reset()
  assert obs.shape == (3, 3)                      # position(2) + height(1), flattened per env
  a = np.zeros((3, 2))
  obs, rew, done, infos = v.step(a)
  assert not done.any() and rew.tolist() == [0.5] * 3 and infos == [{}, {}, {}]
  v.step_async(a)
  with pytest.raises(RuntimeError):
    v.step_async(a)
  obs, rew, done, infos = v.step_wait()
  assert done.all() and rew.tolist() == [1.0] * 3
  # final observation is handed back in info, the returned one is the reset observation
  np.testing.assert_array_equal(infos[1]["terminal_observation"], [1.0, 3.0, 12.0])
  np.testing.assert_array_equal(obs[1], [1.0, 1.0, 10.0])
  assert env.resets == 1 and env.env_resets.tolist() == [2, 2, 2]
  with pytest.raises(RuntimeError):
    v.step_wait()
  with pytest.raises(ValueError):
    v.step_async(np.zeros((2, 2)))
  assert v.seed(7) == [7, 8, 9] and env.task._seed == [7, 8, 9]
  assert v.get_attr("nu") == [2, 2, 2] and v.get_attr("nu", indices=1) == [2]
  v.close()
  assert env.closed and v.closed


def test_vec_env_resets_only_the_envs_that_are_done():
  """scripts/vec_env.py:346-352 is per worker: an env that finishes is reset, the others go on
  (step_wait used to reset the whole batch when any env was done)."""
  env = FakeEnv(nenv=3, length=[2, 3, 2])
  v = wrappers.BatchedVecEnv(env)
  v.reset()
  a = np.zeros((3, 2))
  v.step(a)
  obs, rew, done, infos = v.step(a)
  assert done.tolist() == [True, False, True]
  assert "terminal_observation" in infos[0] and infos[1] == {}
  np.testing.assert_array_equal(obs[:, 1], [0.0, 3.0, 2.0])      # env 1 is still at t = 2
  assert env.env_resets.tolist() == [2, 1, 2]
  obs, rew, done, infos = v.step(a)
  assert done.tolist() == [False, True, False] and rew.tolist() == [0.5, 1.5, 0.5]
