"""Batched `control.Environment` (reference: /root/reference/dm_control/rl/control.py:33-193).

Same episode logic — reset on the step after LAST (control.py:97-98), sub-step loop
(:100-102), reward / observation / discount / time limit (:104-123) — for `nenv`
environments at once.  The sub-step loop, the task's observation and its reward run
on the device inside ONE C-ABI call (`bp_env_step_host`): actions go host->device,
observations and rewards come back.
"""

import collections

import numpy as np

from dm_control_b200 import specs
from dm_control_b200.physics import PhysicsError   # noqa: F401  (re-export)


def compute_n_steps(control_timestep, physics_timestep, tolerance=1e-8):
  """control.py:164-190."""
  if control_timestep < physics_timestep:
    raise ValueError("Control timestep ({}) cannot be smaller than physics timestep ({}).".
                     format(control_timestep, physics_timestep))
  ratio = control_timestep / physics_timestep
  if abs(ratio - round(ratio)) > tolerance:
    raise ValueError("Control timestep ({}) must be an integer multiple of physics timestep "
                     "({})".format(control_timestep, physics_timestep))
  return int(round(ratio))


def flatten_observation(observation, output_key="observations"):
  """control.py:368-393, batched: values are [nenv, ...] and are flattened per env."""
  keys = observation.keys() if isinstance(observation, collections.OrderedDict) \
      else sorted(observation.keys())
  parts = [np.asarray(observation[k]).reshape(len(observation[k]), -1) for k in keys]
  return type(observation)([(output_key, np.concatenate(parts, axis=1))])


class Environment:
  """`nenv` copies of one suite task stepping in lockstep."""

  def __init__(self, physics, task, time_limit=float("inf"), control_timestep=None,
               n_sub_steps=None, flat_observation=False):
    self._physics, self._task = physics, task
    self._flat_observation = flat_observation
    if n_sub_steps is not None and control_timestep is not None:
      raise ValueError("Both n_sub_steps and control_timestep were supplied.")
    elif n_sub_steps is not None:
      self._n_sub_steps = n_sub_steps
    elif control_timestep is not None:
      self._n_sub_steps = compute_n_steps(control_timestep, physics.timestep())
    else:
      self._n_sub_steps = 1
    self._step_limit = float("inf") if time_limit == float("inf") else \
        time_limit / (physics.timestep() * self._n_sub_steps)
    task.bind(physics)
    n = physics.nenv
    # per-env episode clocks: an env is reset on the step after ITS LAST (control.py:97-98)
    self._step_count = np.zeros(n, dtype=np.int64)
    self._reset_next = np.ones(n, dtype=bool)
    self._obs = np.empty((n, task.nobs), dtype=np.float64)
    self._rew = np.empty((n,), dtype=np.float64)
    # Physics.check_invalid_state every k-th step (and at every episode end): a sync and a
    # [nenv, 8] device read each time (SURVEY.md 8b "lazy / every-k checking")
    self.check_every = 1

  @property
  def nenv(self): return self._physics.nenv
  @property
  def physics(self): return self._physics
  @property
  def task(self): return self._task

  def _observation(self, flat):
    obs = collections.OrderedDict()
    k = 0
    for name, width in self._task.obs_layout:
      v = flat[:, k:k + width].copy()
      obs[name] = v[:, 0] if width == 1 and name in self._task.scalar_obs else v
      k += width
    return flatten_observation(obs) if self._flat_observation else obs

  def reset(self):
    """control.py:77-92 for the whole batch."""
    self._reset_next[:] = False
    self._step_count[:] = 0
    with self._physics.reset_context():
      self._task.initialize_episode(self._physics)
    self._physics._batch.obs_reward_host(self._obs, self._rew)
    n = self.nenv
    return specs.TimeStep(np.full(n, specs.StepType.FIRST, dtype=np.int32), None, None,
                          self._observation(self._obs))

  def reset_envs(self, mask):
    """Resets the envs of `mask` only (masked mj_resetData + the task's initialize_episode
    for those envs, each from its own generator); the others keep their state.  Returns the
    flat observation array [nenv, nobs] with the reset envs' rows renewed."""
    mask = np.asarray(mask, dtype=bool)
    if mask.any():
      with self._physics.reset_context(mask=mask):
        self._task.initialize_episode(self._physics, mask)
      self._physics._batch.obs_reward_host(self._obs, self._rew)
      self._step_count[mask] = 0
      self._reset_next[mask] = False
    return self._obs

  def step(self, action):
    """control.py:94-123 for every env; `action` is [nenv, nu].  An env whose previous
    TimeStep was LAST is reset instead of stepped (its action is ignored) and returns FIRST
    with reward and discount NaN (the reference's None); the others step normally."""
    if self._reset_next.all():
      return self.reset()
    action = np.ascontiguousarray(
        np.broadcast_to(np.asarray(action, dtype=np.float64),
                        (self.nenv, self._physics.model.nu)))
    first = self._reset_next.copy()
    batch = self._physics._batch
    if first.any():
      self.reset_envs(first)
      batch.set_mask(~first)               # only the running envs take the step
      try:
        batch.env_step_host(self._n_sub_steps, action, self._obs, self._rew)
      finally:
        batch.set_mask(None)
      batch.obs_reward_host(self._obs, self._rew)     # rows of every env, reset ones included
    else:
      batch.env_step_host(self._n_sub_steps, action, self._obs, self._rew)
    running = ~first
    self._step_count[running] += 1
    n = self.nenv
    reward = self._rew.copy()
    obs = self._observation(self._obs)
    step_type = np.full(n, specs.StepType.MID, dtype=np.int32)
    discount = np.ones(n)
    timeout = running & (self._step_count >= self._step_limit)
    step_type[timeout] = specs.StepType.LAST
    # control.py:110-117 get_termination: a discount ends the episode of THAT env
    terminal = self._task.get_termination(self._physics, self._obs)
    if terminal is not None:
      done = running & ~np.isnan(terminal) & ~timeout
      step_type[done] = specs.StepType.LAST
      discount[done] = terminal[done]
    last = step_type == specs.StepType.LAST
    self._reset_next |= last
    step_type[first] = specs.StepType.FIRST
    reward[first] = np.nan
    discount[first] = np.nan
    if last.any() or self.check_every <= 1 or int(self._step_count.max()) % self.check_every == 0:
      self._physics.check_invalid_state()
    return specs.TimeStep(step_type, reward, discount, obs)

  def action_spec(self):
    return self._physics.action_spec()

  def observation_spec(self):
    """control.py:133-149: inferred from an observation (per-env shapes)."""
    spec = collections.OrderedDict()
    for name, width in self._task.obs_layout:
      shape = () if width == 1 and name in self._task.scalar_obs else (width,)
      spec[name] = specs.Array(shape, np.float64, name=name)
    return spec

  def control_timestep(self):
    return self._physics.timestep() * self._n_sub_steps

  def close(self):
    self._physics.free()
