"""Callers on the other side of the hot path.

* `ActionScale` — /root/reference/dm_control/suite/wrappers/action_scale.py:29-110 for the
  batched Environment: actions in [minimum, maximum] are mapped affinely onto the wrapped
  env's action bounds (the same transform for every env of the batch).
* `ActionNoise` — /root/reference/dm_control/suite/wrappers/action_noise.py:25-72: Gaussian
  noise with std = scale * (max - min) per action dimension, drawn per env from that
  env's task generator (`env.task.random`, one RandomState per env), then clipped.
* `BatchedVecEnv` — the `VecEnv` interface of the fork's `SubprocVecEnv`
  (/root/reference/dm_control/scripts/vec_env.py:22-194,396-472): `reset`, `step_async`,
  `step_wait`, `step`, `seed`, `close`, with the worker's auto-reset contract
  (`vec_env.py:347-352`: when an episode ends the final observation is returned in
  `info["terminal_observation"]` and the env is reset).  One device batch replaces the
  pool of worker processes: `num_envs` = the batch's env count, observations are the flat
  `[num_envs, nobs]` array the device kernel wrote.
"""

import numpy as np

from dm_control_b200 import specs

_ACTION_SPEC_MUST_BE_BOUNDED_ARRAY = (
    "`env.action_spec()` must return a single `BoundedArray`, got: {}.")
_MUST_BE_FINITE = "All values in `{name}` must be finite, got: {bounds}."
_MUST_BROADCAST = "`{name}` must be broadcastable to shape {shape}, got: {bounds}."


class ActionScale:
  """Wraps a batched control environment to rescale actions to a specific range."""

  def __init__(self, env, minimum, maximum):
    action_spec = env.action_spec()
    if not isinstance(action_spec, specs.BoundedArray):
      raise ValueError(_ACTION_SPEC_MUST_BE_BOUNDED_ARRAY.format(action_spec))
    minimum = np.array(minimum)
    maximum = np.array(maximum)
    shape = action_spec.shape
    orig_minimum, orig_maximum = action_spec.minimum, action_spec.maximum

    def validate(bounds, name):
      if not np.all(np.isfinite(bounds)):
        raise ValueError(_MUST_BE_FINITE.format(name=name, bounds=bounds))
      try:
        np.broadcast_to(bounds, shape)
      except ValueError:
        raise ValueError(_MUST_BROADCAST.format(name=name, bounds=bounds, shape=shape))

    validate(minimum, "minimum")
    validate(maximum, "maximum")
    validate(orig_minimum, "env.action_spec().minimum")
    validate(orig_maximum, "env.action_spec().maximum")
    scale = (orig_maximum - orig_minimum) / (maximum - minimum)
    self._transform = lambda a: (orig_minimum + scale * (np.asarray(a) - minimum)).astype(
        action_spec.dtype, copy=False)
    self._action_spec = specs.BoundedArray(
        shape=shape, dtype=np.result_type(minimum, maximum, action_spec.dtype),
        minimum=np.broadcast_to(minimum, shape), maximum=np.broadcast_to(maximum, shape),
        name=action_spec.name)
    self._env = env

  def step(self, action):
    return self._env.step(self._transform(action))

  def reset(self):
    return self._env.reset()

  def observation_spec(self):
    return self._env.observation_spec()

  def action_spec(self):
    return self._action_spec

  def __getattr__(self, name):
    return getattr(self._env, name)


_BOUNDS_MUST_BE_FINITE = (
    "All bounds in `env.action_spec()` must be finite, got: {action_spec}")


class ActionNoise:
  """Wraps a batched control environment and adds Gaussian noise to actions."""

  def __init__(self, env, scale=0.01):
    action_spec = env.action_spec()
    if not (np.all(np.isfinite(action_spec.minimum)) and
            np.all(np.isfinite(action_spec.maximum))):
      raise ValueError(_BOUNDS_MUST_BE_FINITE.format(action_spec=action_spec))
    self._minimum = action_spec.minimum
    self._maximum = action_spec.maximum
    self._noise_std = scale * (action_spec.maximum - action_spec.minimum)
    self._env = env

  def step(self, action):
    action = np.asarray(action, dtype=np.float64)
    noise = np.stack([np.broadcast_to(rng.normal(scale=self._noise_std), action.shape[1:])
                      for rng in self._env.task.random])
    noisy_action = action + noise
    np.clip(noisy_action, self._minimum, self._maximum, out=noisy_action)
    return self._env.step(noisy_action)

  def reset(self):
    return self._env.reset()

  def observation_spec(self):
    return self._env.observation_spec()

  def action_spec(self):
    return self._env.action_spec()

  def __getattr__(self, name):
    return getattr(self._env, name)


def _flat(observation):
  parts = [np.asarray(v).reshape(len(v), -1) for v in observation.values()]
  return np.concatenate(parts, axis=1)


class BatchedVecEnv:
  """`VecEnv`-style facade over one batched Environment (scripts/vec_env.py:22-194).  Every
  env has its own episode clock: an env that finishes is reset on the spot -- and only that
  env -- its last observation goes to infos[i]["terminal_observation"] and the returned
  observation row is the first of its next episode (vec_env.py:346-352)."""

  def __init__(self, env):
    self._env = env
    self.num_envs = env.nenv
    spec = env.action_spec()
    self.action_space = spec
    self.observation_space = env.observation_spec()
    self._pending = None
    self.closed = False

  def reset(self):
    return _flat(self._env.reset().observation)

  def step_async(self, actions):
    if self._pending is not None:
      raise RuntimeError("step_async called twice without step_wait")
    actions = np.asarray(actions, dtype=np.float64)
    if actions.shape != (self.num_envs,) + tuple(self.action_space.shape):
      raise ValueError("actions have shape {}, expected {}".format(
          actions.shape, (self.num_envs,) + tuple(self.action_space.shape)))
    self._pending = actions

  def step_wait(self):
    if self._pending is None:
      raise RuntimeError("step_wait called without step_async")
    actions, self._pending = self._pending, None
    ts = self._env.step(actions)
    obs = _flat(ts.observation)
    n = self.num_envs
    reward = np.zeros(n) if ts.reward is None else np.asarray(ts.reward, dtype=np.float64)
    done = np.asarray(ts.step_type) == specs.StepType.LAST
    infos = [{} for _ in range(n)]
    if done.any():                       # vec_env.py:347-352: only the envs that are done
      for i in np.flatnonzero(done):
        infos[i]["terminal_observation"] = obs[i].copy()
      obs = obs.copy()
      obs[done] = self._env.reset_envs(done)[done]
    return obs, reward, done, infos

  def step(self, actions):
    self.step_async(actions)
    return self.step_wait()

  def seed(self, seed=None):
    """vec_env.py:462-465: env i is seeded seed + i; takes effect at the next reset."""
    if seed is None:
      return [None] * self.num_envs
    seeds = [int(seed) + i for i in range(self.num_envs)]
    self._env.task._seed = seeds
    self._env.task._rngs = [np.random.RandomState(s) for s in seeds]
    return seeds

  def get_attr(self, attr_name, indices=None):
    value = getattr(self._env, attr_name)
    return [value for _ in self._indices(indices)]

  def env_method(self, method_name, *args, indices=None, **kwargs):
    out = getattr(self._env, method_name)(*args, **kwargs)
    return [out for _ in self._indices(indices)]

  def _indices(self, indices):
    if indices is None:
      return range(self.num_envs)
    if isinstance(indices, int):
      return [indices]
    return indices

  def close(self):
    if not self.closed:
      self._env.close()
      self.closed = True

  @property
  def unwrapped(self):
    return self


class TorchVecEnv:
  """Device-resident VecEnv for learners that keep their data on the GPU (the fork's PPO,
  scripts/rl/ppo/ppo.py, consumes SubprocVecEnv output as torch tensors): actions come in as
  a CUDA float64 tensor [num_envs, nu] and observations / rewards / dones go out as CUDA
  tensors -- no host copy of either on the step path.  One `bp_env_step` launch per step on
  the batch's stream, ordered against torch's current stream with stream waits.

  Episode ends: time limits are counted on the host (no device read); a task with its own
  termination (lqr) is checked on the device and read back as one flag.  Envs that are done
  are reset on the spot -- only those -- by the task's own initialize_episode on the host
  (per-env generators); `terminal_observation` is returned as a tensor [num_envs, nobs]
  whose rows are valid where `done` is set.  Physics.check_invalid_state runs every
  `check_every` steps and at every episode end."""

  def __init__(self, env, check_every=100):
    import torch
    self._torch = torch
    self._env = env
    self.num_envs = env.nenv
    self.action_space = env.action_spec()
    self.observation_space = env.observation_spec()
    self.check_every = int(check_every)
    physics = env.physics
    self._batch = physics._batch
    self._dev = torch.device("cuda", physics._dev_model.device)
    n = self.num_envs
    self._obs = torch.empty((n, env.task.nobs), dtype=torch.float64, device=self._dev)
    self._rew = torch.empty((n,), dtype=torch.float64, device=self._dev)
    self._stream = torch.cuda.ExternalStream(self._batch.stream, device=self._dev)
    self._steps = 0
    self.closed = False

  def _sync_out(self):
    self._torch.cuda.current_stream(self._dev).wait_stream(self._stream)

  def _observe(self):
    self._stream.wait_stream(self._torch.cuda.current_stream(self._dev))
    self._batch.obs_reward(self._obs.data_ptr() if self._obs.numel() else None,
                           self._rew.data_ptr())
    self._sync_out()

  def reset(self):
    self._env.reset()
    self._observe()
    return self._obs.clone()

  def step(self, actions):
    torch = self._torch
    env = self._env
    if actions.shape != (self.num_envs,) + tuple(self.action_space.shape):
      raise ValueError("actions have shape {}, expected {}".format(
          tuple(actions.shape), (self.num_envs,) + tuple(self.action_space.shape)))
    if actions.device != self._dev or actions.dtype != torch.float64 or \
        not actions.is_contiguous():
      actions = actions.to(device=self._dev, dtype=torch.float64).contiguous()
    self._stream.wait_stream(torch.cuda.current_stream(self._dev))
    self._batch.env_step(env._n_sub_steps, actions.data_ptr(),
                         self._obs.data_ptr() if self._obs.numel() else None,
                         self._rew.data_ptr())
    self._sync_out()      # torch's stream is now ordered after the launch: `actions` may be reused
    self._steps += 1
    env._step_count += 1
    done_host = env._step_count >= env._step_limit
    obs, reward = self._obs.clone(), self._rew.clone()
    done = torch.as_tensor(done_host, device=self._dev)
    if env.task.tdef.domain == "lqr":                 # task termination: checked on the device
      from dm_control_b200.suite import lqr
      term = torch.linalg.vector_norm(obs, dim=1) < lqr.TERMINAL_TOL
      if bool(term.any()):                             # one flag crosses to the host
        done = done | term
        done_host = done.cpu().numpy()
    terminal_observation = obs
    if done_host.any() or self._steps % self.check_every == 0:
      env.physics.check_invalid_state()
    if done_host.any():
      env.reset_envs(done_host)                        # masked reset, host-side episode init
      self._observe()
      rows = torch.as_tensor(done_host, device=self._dev)
      obs = torch.where(rows[:, None], self._obs, obs)
    return obs, reward, done, terminal_observation

  def close(self):
    if not self.closed:
      self._torch.cuda.synchronize(self._dev)
      self._stream = None
      self._env.close()
      self.closed = True
