"""Per-env auto-reset and the VecEnv facades over the CUDA env (VERDICT r1 items 7, 8;
/root/reference/dm_control/rl/control.py:97-98, scripts/vec_env.py:346-352)."""

import numpy as np
import pytest

from dm_control_b200 import specs, suite, wrappers

pytestmark = pytest.mark.gpu


def _load(domain, task, n, seed, limit):
  return suite.load(domain, task, task_kwargs={"random": seed, "time_limit": limit}, nenv=n)


@pytest.mark.parametrize("domain,task,limit", [("cartpole", "swingup", 0.2),
                                               ("cheetah", "run", 0.12),
                                               ("reacher", "easy", 0.3)])
def test_environment_resets_each_env_on_the_step_after_its_own_last(domain, task, limit):
  """Episodes staggered by resetting half the batch mid-episode: afterwards every env ends,
  returns LAST and is reset (FIRST, action ignored) on its own clock, and the envs that are
  NOT being reset advance exactly as in an undisturbed twin batch."""
  n = 12
  env = _load(domain, task, n, 3, limit)
  twin = _load(domain, task, n, 3, limit)
  nu = env.action_spec().shape[0]
  rng = np.random.RandomState(0)
  steps = int(round(limit / env.control_timestep()))
  ts, tt = env.reset(), twin.reset()
  half = np.arange(n) < n // 2
  k0 = steps // 3
  for k in range(k0):
    a = rng.uniform(-1, 1, (n, nu))
    ts, tt = env.step(a), twin.step(a)
  for name in ts.observation:
    np.testing.assert_array_equal(ts.observation[name], tt.observation[name])
  env.reset_envs(half)                       # the first half starts a new episode now
  seen_first = np.zeros(n, dtype=int)
  seen_last = np.zeros(n, dtype=int)
  for k in range(k0, 3 * steps + 4):
    a = rng.uniform(-1, 1, (n, nu))
    ts = env.step(a)
    st = np.asarray(ts.step_type)
    seen_first += st == specs.StepType.FIRST
    seen_last += st == specs.StepType.LAST
    # the undisturbed half ends exactly when the twin does
    if k < steps:
      tt = twin.step(a)
      for name in ts.observation:
        np.testing.assert_array_equal(np.asarray(ts.observation[name])[~half],
                                      np.asarray(tt.observation[name])[~half], err_msg=name)
      np.testing.assert_array_equal(st[~half], np.asarray(tt.step_type)[~half])
      np.testing.assert_array_equal(ts.reward[~half], tt.reward[~half])
    # the two halves never end on the same step
    assert not ((st[half] == specs.StepType.LAST).any() and (st[~half] == specs.StepType.LAST).any())
    first = st == specs.StepType.FIRST
    assert np.isnan(ts.reward[first]).all() and np.isfinite(ts.reward[~first]).all()
  assert (seen_last >= 2).all() and (seen_first >= 2).all()
  env.close(); twin.close()


def test_lqr_envs_terminate_and_restart_individually():
  """lqr.py:256-259: an env whose state norm falls below 1e-6 ends with discount 0; only
  that env is reset (the batch used to be reset as a whole)."""
  n = 6
  env = suite.load("lqr", "lqr_2_1", task_kwargs={"random": 5}, nenv=n)
  ts = env.reset()
  q = env.physics.data.qpos
  q[:3] *= 1e-9                                   # three envs start on the goal
  with env.physics.reset_context():               # derived quantities follow the new state
    env.physics.data.qpos = q
  ts = env.step(np.zeros((n, 1)))
  st = np.asarray(ts.step_type)
  assert (st[:3] == specs.StepType.LAST).all() and (st[3:] == specs.StepType.MID).all()
  assert (ts.discount[:3] == 0).all() and (ts.discount[3:] == 1).all()
  before = env.physics.data.qpos[3:].copy()
  ts = env.step(np.zeros((n, 1)))
  st = np.asarray(ts.step_type)
  assert (st[:3] == specs.StepType.FIRST).all() and (st[3:] == specs.StepType.MID).all()
  assert np.linalg.norm(env.physics.data.qpos[:3], axis=1).min() > 1.0      # fresh draws
  assert not np.array_equal(env.physics.data.qpos[3:], before)               # the others moved on
  env.close()


@pytest.mark.parametrize("domain,task,limit", [("cartpole", "balance", 0.15),
                                               ("walker", "walk", 0.2)])
def test_torch_vec_env_equals_numpy_vec_env(domain, task, limit):
  """TorchVecEnv (CUDA tensors in and out, one bp_env_step launch per step) against
  BatchedVecEnv (host arrays through bp_env_step_host): same seeds, same actions -> the same
  observations, rewards, dones and terminal observations through two staggered episode ends."""
  import torch
  n = 10
  e1, e2 = _load(domain, task, n, 9, limit), _load(domain, task, n, 9, limit)
  v1, v2 = wrappers.BatchedVecEnv(e1), wrappers.TorchVecEnv(e2, check_every=7)
  o1, o2 = v1.reset(), v2.reset()
  assert o2.is_cuda and o2.dtype == torch.float64
  np.testing.assert_array_equal(o1, o2.cpu().numpy())
  stagger = np.arange(n) % 2 == 0
  rng = np.random.RandomState(1)
  nu = e1.action_spec().shape[0]
  steps = int(round(limit / e1.control_timestep()))
  ndone = 0
  for k in range(3 * steps):
    if k == steps // 2:                       # stagger the clocks of half of the envs
      e1.reset_envs(stagger); e2.reset_envs(stagger)
    a = rng.uniform(-1, 1, (n, nu))
    o1, r1, d1, info = v1.step(a)
    o2, r2, d2, term = v2.step(torch.from_numpy(a).cuda())
    np.testing.assert_array_equal(o1, o2.cpu().numpy())
    np.testing.assert_array_equal(r1, r2.cpu().numpy())
    np.testing.assert_array_equal(d1, d2.cpu().numpy())
    for i in np.flatnonzero(d1):
      np.testing.assert_array_equal(info[i]["terminal_observation"], term[i].cpu().numpy())
    ndone += int(d1.sum())
    assert not (d1[stagger].any() and d1[~stagger].any())
  assert ndone >= 2 * n
  v1.close(); v2.close()
