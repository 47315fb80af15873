"""GPU tests of the host interface above the C ABI: BatchedPhysics (engine.py's
Physics), named indexing, the batched control.Environment, suite.load and the task
observation / reward kernels (checked against oracle/task_oracle.py).

Modelled on the reference's own tests for the path: mujoco/engine_test.py (step,
reset, reset_context, named indexing, invalid state), rl/control_test.py
(TimeStep sequencing, sub-steps), suite/tests/domains_test.py (determinism,
observation/reward sanity for every task).
"""

import numpy as np
import pytest

import parity_util as PU
from oracle import task_oracle

pytestmark = pytest.mark.gpu

SERVED = [("cartpole", "swingup"), ("cartpole", "swingup_sparse"), ("cartpole", "balance"),
          ("cartpole", "balance_sparse"), ("acrobot", "swingup"), ("acrobot", "swingup_sparse"),
          ("pendulum", "swingup"), ("cheetah", "run"), ("walker", "stand"), ("walker", "walk"),
          ("walker", "run"), ("hopper", "stand"), ("hopper", "hop"), ("finger", "spin"),
          ("humanoid", "stand"), ("humanoid", "walk"), ("humanoid", "run"),
          ("quadruped", "walk"), ("quadruped", "run"), ("quadruped", "escape"), ("quadruped", "fetch"),
          ("lqr", "lqr_2_1"), ("lqr", "lqr_6_2"), ("reacher", "easy"), ("reacher", "hard"),
          ("point_mass", "easy"), ("point_mass", "hard"), ("finger", "turn_easy"), ("finger", "turn_hard"),
          ("humanoid_CMU", "stand"), ("humanoid_CMU", "run"),
          ("swimmer", "swimmer6"), ("swimmer", "swimmer15"),
          ("fish", "upright"), ("fish", "swim"),
          ("ball_in_cup", "catch"), ("manipulator", "bring_ball"), ("cartpole", "two_poles"), ("cartpole", "three_poles"),
          ("manipulator", "bring_peg"), ("manipulator", "insert_ball"), ("manipulator", "insert_peg"),
          ("stacker", "stack_2"), ("stacker", "stack_4")]
FIELDS = ("qpos", "qvel", "act", "ctrl", "xmat", "xpos", "xquat", "xipos", "site_xpos", "geom_xpos",
          "geom_xmat", "sensordata")


def _fields(physics):
  return {k: np.asarray(getattr(physics.data, k)) for k in FIELDS}


@pytest.mark.parametrize("domain,task", SERVED)
def test_obs_reward_kernel_matches_task_oracle(domain, task):
  """suite/<domain>.py get_observation / get_reward restated in numpy vs the kernel."""
  from dm_control_b200 import suite
  from dm_control_b200.suite import tasks as T
  n = 24
  env = suite.load(domain, task, task_kwargs={"random": 11}, nenv=n)
  tdef = T.REGISTRY[(domain, task)]
  ip, fp = tdef.params(env.physics.model)
  speed = fp[0] if domain in ("walker", "humanoid", "quadruped", "humanoid_CMU") else None
  rng = np.random.RandomState(3)
  ts = env.reset()
  spec = env.action_spec()
  for k in range(6):
    obs_flat = np.concatenate([np.asarray(v).reshape(n, -1) for v in ts.observation.values()],
                              axis=1)
    want_obs, want_r = task_oracle.obs_reward(domain, task, env.physics.model,
                                              _fields(env.physics), speed)
    np.testing.assert_allclose(obs_flat, want_obs, rtol=1e-12, atol=1e-13)
    if ts.reward is not None:
      np.testing.assert_allclose(ts.reward, want_r, rtol=1e-12, atol=1e-13)
      if domain != "lqr":                                             # quadratic cost
        assert ((ts.reward >= 0) & (ts.reward <= 1)).all()            # domains_test.py
    assert np.isfinite(obs_flat).all()
    a = rng.uniform(np.maximum(spec.minimum, -1), np.minimum(spec.maximum, 1),
                    (n, spec.shape[0]))
    ts = env.step(a)
  env.close()


@pytest.mark.parametrize("domain,task", [("cartpole", "swingup"), ("cheetah", "run"),
                                         ("walker", "walk"), ("humanoid", "stand")])
def test_same_seed_same_trajectory(domain, task):
  """suite/tests/domains_test.py test_task_runs / determinism: two envs built with the
  same seed produce bit-identical TimeSteps; a different seed does not."""
  from dm_control_b200 import suite

  def rollout(seed):
    env = suite.load(domain, task, task_kwargs={"random": seed}, nenv=8)
    rng = np.random.RandomState(0)
    out = []
    ts = env.reset()
    for _ in range(5):
      out.append(np.concatenate([np.asarray(v).reshape(8, -1) for v in ts.observation.values()],
                                axis=1))
      ts = env.step(rng.uniform(-1, 1, (8, env.action_spec().shape[0])))
    out.append(ts.reward)
    env.close()
    return out
  a, b, c = rollout(5), rollout(5), rollout(6)
  for x, y in zip(a, b):
    assert np.array_equal(x, y)
  assert not np.array_equal(a[0], c[0])
  # per-env streams: env i of seed 5 == env i-1 of seed 6 at reset (RandomState(seed+i))
  if domain != "humanoid":       # humanoid's rejection loop also matches but is slower
    np.testing.assert_array_equal(a[0][1:], c[0][:-1])


def test_timestep_sequence_and_time_limit():
  """rl/control_test.py:49-86: FIRST, MID..., LAST at the time limit, then FIRST."""
  from dm_control_b200 import suite
  from dm_control_b200.specs import StepType
  env = suite.load("pendulum", "swingup", task_kwargs={"random": 1, "time_limit": 0.1},
                   nenv=5)
  nstep = int(round(0.1 / env.control_timestep()))
  ts = env.reset()
  assert (ts.step_type == StepType.FIRST).all() and ts.reward is None and ts.discount is None
  a = np.zeros((5, 1))
  for k in range(nstep - 1):
    ts = env.step(a)
    assert (ts.step_type == StepType.MID).all()
    assert ts.reward.shape == (5,) and (ts.discount == 1).all()
  ts = env.step(a)
  assert (ts.step_type == StepType.LAST).all()
  np.testing.assert_allclose(env.physics.time(), 0.1, rtol=0, atol=1e-12)
  ts = env.step(a)
  assert (ts.step_type == StepType.FIRST).all()
  assert (env.physics.time() == 0).all()
  env.close()


def test_sub_steps_advance_time_exactly():
  """control.py:100-102: one env.step = n_sub_steps physics steps."""
  from dm_control_b200 import suite
  env = suite.load("walker", "walk", task_kwargs={"random": 0}, nenv=4)
  dt = env.physics.timestep()
  assert env.control_timestep() == pytest.approx(0.025)
  n = int(round(0.025 / dt))
  env.reset()
  env.step(np.zeros((4, env.action_spec().shape[0])))
  np.testing.assert_allclose(env.physics.time(), n * dt, rtol=0, atol=1e-15)
  # and Physics.step(k) is k single steps
  p = env.physics
  q0 = p.state()
  p.step(3)
  q3 = p.state()
  p.set_state(q0); p.data.time = 0.0
  p.forward()
  for _ in range(3):
    p.step()
  np.testing.assert_array_equal(p.state(), q3)
  env.close()


def test_named_indexing_and_reset_context():
  """engine_test.py testNamedViews / testResetContext on the batched Physics."""
  from dm_control_b200.physics import BatchedPhysics
  m = PU.load_model("cartpole")
  p = BatchedPhysics(m, nenv=6)
  with p.reset_context():
    p.named.data.qpos["slider"] = np.arange(6) * 0.1
    p.named.data.qpos["hinge_1"] = 0.3
  np.testing.assert_allclose(p.named.data.qpos["slider"][:, 0], np.arange(6) * 0.1)
  # derived quantities are fresh on exit of the context (engine.py:291-295)
  cart_x = p.named.data.xpos["cart", "x"]
  np.testing.assert_allclose(cart_x, np.arange(6) * 0.1, atol=1e-15)
  zz = p.named.data.xmat["pole_1", "zz"]
  np.testing.assert_allclose(zz, np.cos(0.3), atol=1e-15)
  # per-env masked reset leaves the other envs alone
  p.step(5)
  before = p.state()
  mask = np.array([1, 0, 0, 1, 0, 0], dtype=bool)
  with p.reset_context(mask=mask):
    pass
  after = p.state()
  np.testing.assert_array_equal(after[~mask], before[~mask])
  assert (after[mask][:, m.nq:] == 0).all() and (p.time()[mask] == 0).all()
  assert (p.time()[~mask] > 0).all()
  with pytest.raises(IndexError):
    p.named.data.qpos["no_such_joint"]
  p.free()


def test_invalid_state_raises_physics_error():
  """engine_test.py testNanControl / check_invalid_state (engine.py:307-330)."""
  from dm_control_b200.physics import BatchedPhysics, PhysicsError
  m = PU.load_model("cartpole")
  p = BatchedPhysics(m, nenv=4)
  p.reset()
  ctrl = np.zeros((4, m.nu)); ctrl[2] = np.nan
  p.set_control(ctrl)
  with pytest.raises(PhysicsError, match="BADCTRL"):
    p.step()
  # the other envs kept integrating; a second step with clean controls is quiet
  p.set_control(np.zeros((4, m.nu)))
  p.step()
  # diverged positions reset the env and raise BADQPOS
  with p.reset_context():
    pass
  q = p.data.qpos
  q[1, 0] = 1e15
  p.data.qpos = q
  with pytest.raises(PhysicsError, match="BADQPOS"):
    p.step()
  assert (np.abs(p.data.qpos) < 1e3).all()          # mj_resetData happened in-kernel
  p.free()


def test_set_state_shape_error_and_action_spec():
  from dm_control_b200.physics import BatchedPhysics
  m = PU.load_model("cheetah")
  p = BatchedPhysics(m, nenv=2)
  with pytest.raises(ValueError):
    p.set_state(np.zeros(3))
  with pytest.raises(ValueError):
    p.set_control(np.zeros((2, m.nu + 1)))
  spec = p.action_spec()
  assert spec.shape == (m.nu,) and (spec.minimum == -1).all() and (spec.maximum == 1).all()
  p.free()


def test_device_resident_stepping_matches_host_stepping():
  """bp_field_copy_device + bp_obs_reward (device buffers) == bp_env_step_host."""
  import torch
  from dm_control_b200 import suite
  n = 64
  e1 = suite.load("cheetah", "run", task_kwargs={"random": 2}, nenv=n)
  e2 = suite.load("cheetah", "run", task_kwargs={"random": 2}, nenv=n)
  e1.reset(); e2.reset()
  b = e2.physics._batch
  nu = e2.physics.model.nu
  stream = torch.cuda.ExternalStream(b.stream)
  obs = torch.empty((n, b.nobs), device="cuda", dtype=torch.float64)
  rew = torch.empty((n,), device="cuda", dtype=torch.float64)
  rng = np.random.RandomState(4)
  for _ in range(4):
    a = rng.uniform(-1, 1, (n, nu))
    ts = e1.step(a)
    d_a = torch.from_numpy(a).cuda()
    torch.cuda.synchronize()
    b.copy_from_device("ctrl", d_a.data_ptr())
    b.step(1)
    b.obs_reward(obs.data_ptr(), rew.data_ptr())
    b.sync()
    np.testing.assert_array_equal(rew.cpu().numpy(), ts.reward)
    np.testing.assert_array_equal(obs.cpu().numpy()[:, :nu + 2],
                                  ts.observation["position"])
  del stream
  e1.close(); e2.close()


def test_fused_env_step_equals_separate_launches():
  """bp_env_step (ctrl scatter + sub-steps + obs/reward in one launch) is bit-identical
  to bp_field_copy_device + bp_step + bp_obs_reward, for every execution shape."""
  import torch
  from dm_control_b200 import binding
  from dm_control_b200.suite import tasks as T
  for domain, task, lanes in (("cheetah", "run", 4), ("cheetah", "run", 1), ("walker", "walk", 32),
                              ("humanoid", "run", 8)):
    tdef = T.REGISTRY[(domain, task)]
    m = PU.load_model(tdef.model)
    kw = dict(njmax=120, nconmax=40) if domain == "humanoid" else dict(njmax=0, nconmax=0)
    n = 37
    qpos, qvel, ctrl = PU.random_states(m, n, 6)
    outs = []
    for fused in (False, True):
      model = binding.Model(m)
      b = binding.Batch(model, n, kw["njmax"], kw["nconmax"], lanes)
      ip, fp = tdef.params(m)
      b.set_task(tdef.kind, ip, fp)
      b.write("qpos", qpos); b.write("qvel", qvel)
      b.forward()
      obs = torch.zeros((n, b.nobs), device="cuda", dtype=torch.float64)
      rew = torch.zeros((n,), device="cuda", dtype=torch.float64)
      nsub = T.n_sub_steps(tdef, m)
      for k in range(3):
        a = torch.from_numpy(np.roll(ctrl, k, axis=0).copy()).cuda()
        torch.cuda.synchronize()
        if fused:
          b.env_step(nsub, a.data_ptr(), obs.data_ptr(), rew.data_ptr())
        else:
          b.copy_from_device("ctrl", a.data_ptr()); b.step(nsub)
          b.obs_reward(obs.data_ptr(), rew.data_ptr())
        b.sync()
      outs.append((obs.cpu().numpy(), rew.cpu().numpy(), b.read("qpos"), b.read("sensordata")))
      b.free(); model.free()
    for x, y in zip(*outs):
      np.testing.assert_array_equal(x, y)


def test_config_c1_cartpole_trace_through_suite_load():
  """BASELINE.json configs[0] / SURVEY.md 8d C1: cartpole-swingup, task seed 0, 1000 steps of
  RandomState(0) actions through suite.load -> Environment.step, against the committed
  CPU-oracle trace (tests/golden/c1_cartpole_swingup.npz, tools/make_golden.py)."""
  import os
  from dm_control_b200 import suite
  g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden",
                           "c1_cartpole_swingup.npz"))
  env = suite.load("cartpole", "swingup", task_kwargs={"random": 0}, nenv=1)
  ts = env.reset()
  np.testing.assert_array_equal(env.physics.data.qpos[0], g["qpos0"])   # same episode start
  worst = 0.0
  for k in range(len(g["actions"])):
    ts = env.step(g["actions"][k][None, :])
    obs = np.concatenate([np.asarray(v).reshape(1, -1) for v in ts.observation.values()], axis=1)[0]
    err = np.max(np.abs(obs - g["obs"][k]) / (1e-6 + np.abs(g["obs"][k])))
    worst = max(worst, err, abs(ts.reward[0] - g["reward"][k]))
    if k in (0, 9, 99):
      assert worst <= 1e-8, (k, worst)            # single / short-horizon bar
  print("C1 1000-step trace: worst relative obs / abs reward error %.3e" % worst)
  assert worst < 1e-5                              # long-horizon divergence, reported
  assert abs(np.sum(g["reward"]) - 27.6088597862409) < 1e-9
  assert (ts.step_type == 2).all()                 # 1000 steps = the 10 s time limit
  env.close()


@pytest.mark.parametrize("level", ["lqr_2_1", "lqr_6_2"])
def test_lqr_optimal_policy_on_device(level):
  """suite/lqr_test.py:30-59 through suite.load on the GPU: the Riccati-optimal policy's
  measured cost equals .5 x0' P x0 (rtol 1e-3) for every env of the batch."""
  import math
  from dm_control_b200 import suite
  from dm_control_b200.suite import lqr
  n = 6
  env = suite.load("lqr", level, task_kwargs={"random": 4}, nenv=n)
  ts = env.reset()
  m = env.physics.model
  p, k, beta = lqr.solve(m, env.physics.data.qM[0])
  n_steps = int(math.ceil(math.log10(1e-3) / math.log10(beta)))
  x0 = np.hstack((ts.observation["position"], ts.observation["velocity"]))
  total = np.zeros(n)
  for _ in range(n_steps):
    x = np.hstack((ts.observation["position"], ts.observation["velocity"]))
    total += 1 - (ts.reward if ts.reward is not None else 0.0)
    ts = env.step(x.dot(k.T))
  expected = .5 * np.einsum("ni,ij,nj->n", x0, p, x0)
  np.testing.assert_allclose(expected, total, rtol=1e-3)
  assert len(np.unique(np.round(expected, 6))) == n          # six different start states
  env.close()



@pytest.mark.gpu
@pytest.mark.parametrize("name,nenv", [("cheetah-run", 1000), ("humanoid-run", 300),
                                       ("pendulum-swingup", 77)])
def test_env_step_host_direct_results_equal_copied_results(name, nenv):
  """bp_env_step_host with pinned (mapped) result buffers -- filled by the kernel's blocks
  themselves -- against pageable buffers filled by D2H copies: identical bits, also for the
  ragged last block."""
  import torch
  import bench
  from dm_control_b200 import binding
  outs = []
  for pinned in (True, False):
    tdef, m, model, batch, nsub = bench.make_batch(name, nenv, 0, seed=5)
    rng = np.random.RandomState(3)
    obs = torch.zeros((nenv, batch.nobs), dtype=torch.float64)
    rew = torch.zeros((nenv,), dtype=torch.float64)
    if pinned:
      obs, rew = obs.pin_memory(), rew.pin_memory()
    seq = []
    for k in range(3):
      ctrl = rng.uniform(-1, 1, (nenv, m.nu))
      batch.env_step_host(nsub, ctrl, obs.numpy(), rew.numpy())
      seq.append((obs.numpy().copy(), rew.numpy().copy()))
    outs.append(seq)
    batch.free(); model.free()
  for (o1, r1), (o2, r2) in zip(*outs):
    assert np.all(np.isfinite(o1)) and np.abs(o1).max() > 0
    np.testing.assert_array_equal(o1, o2)
    np.testing.assert_array_equal(r1, r2)


@pytest.mark.gpu
def test_point_mass_hard_randomizes_gains_per_env():
  """suite/point_mass.py:85-101: the hard task draws, per episode, two unit actuation
  directions that are not 'too parallel' and writes them into model.wrap_prm -- here per env
  (bp_env_model_write), in the reference's draw order after the joint randomizer."""
  from dm_control_b200 import suite
  n = 16
  env = suite.load("point_mass", "hard", task_kwargs={"random": 3}, nenv=n)
  env.reset()
  w = env.physics.env_model_field("wrap_prm").reshape(n, 2, 2)
  np.testing.assert_allclose(np.linalg.norm(w, axis=2), 1.0, atol=1e-12)
  assert np.all(np.abs(np.sum(w[:, 0] * w[:, 1], axis=1)) <= 0.9)
  assert len(np.unique(np.round(w[:, 0, 0], 9))) == n
  rng = np.random.RandomState(3)          # env 0: two limited slide joints, then the gains
  rng.uniform(0, 1); rng.uniform(0, 1)
  d1 = rng.randn(2); d1 /= np.linalg.norm(d1)
  np.testing.assert_allclose(w[0, 0], d1, atol=1e-12)
  # control 0 pushes along this env's first direction
  env.physics.data.qvel = np.zeros((n, 2))
  env.physics.set_control(np.tile([1.0, 0.0], (n, 1)))
  env.physics.forward()
  f = env.physics.data.qfrc_actuator
  np.testing.assert_allclose(f / np.linalg.norm(f, axis=1, keepdims=True), w[:, 0], atol=1e-10)
  # a second episode draws new gains; the easy task leaves the model's
  env.reset()
  assert not np.allclose(env.physics.env_model_field("wrap_prm").reshape(n, 2, 2), w)
  env.close()
  easy = suite.load("point_mass", "easy", task_kwargs={"random": 3}, nenv=4)
  easy.reset()
  assert not easy.physics._batch.has_env_model("wrap_prm")
  easy.close()
