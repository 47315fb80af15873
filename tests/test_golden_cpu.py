"""CPU checks of the test oracles against the committed golden fixtures."""

import os

import numpy as np
import pytest

import parity_util as PU
from oracle import task_oracle
from oracle.oracle_py import Oracle, OracleModel

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
SIG = {0: "gaussian", 1: "linear", 2: "quadratic", 3: "long_tail"}


def test_tolerance_restatement_matches_reference_rewards_py():
  # fixture = outputs of /root/reference/dm_control/utils/rewards.py itself
  g = np.load(os.path.join(GOLDEN, "rewards_tolerance.npz"))
  for (lo, hi, mg, sg, v1), want in zip(g["cases"], g["out"]):
    got = task_oracle.tolerance(g["x"], (lo, hi), mg, SIG[int(sg)], v1)
    np.testing.assert_allclose(got, want, rtol=0, atol=1e-15)


@pytest.mark.parametrize("domain", ["pendulum", "cartpole", "cheetah", "finger", "humanoid",
                                    "quadruped_escape"])
def test_oracle_reproduces_committed_trajectories(domain):
  g = np.load(os.path.join(GOLDEN, "traj_%s.npz" % domain))
  m = PU.load_model(domain)
  if "sizes" in g.files:
    m.arrays["njmax"][:] = g["sizes"][0]; m.arrays["nconmax"][:] = g["sizes"][1]
  om = OracleModel(m)
  steps = list(g["steps"])
  for e in range(2):                         # two envs keep the CPU suite fast
    o = Oracle(om)
    if "hfield_seed" in g.files:
      o.hfield_data[:] = PU.escape_terrain(np.random.RandomState(int(g["hfield_seed"][e])))
    o.reset()
    o.qpos[:] = g["qpos0"][e]; o.qvel[:] = g["qvel0"][e]; o.ctrl[:] = g["ctrl"][e]
    o.forward()
    for s in range(max(steps) + 1):
      if s > 0:
        o.step()
      if s in steps:
        k = steps.index(s)
        np.testing.assert_allclose(o.qpos, g["qpos"][e, k], rtol=1e-10, atol=1e-12)
        np.testing.assert_array_equal(o.ncon, g["ncon"][e, k])
        np.testing.assert_array_equal(o.nefc, g["nefc"][e, k])


def test_c1_trace_fixture_is_what_the_generator_makes():
  """The committed C1 trace equals a fresh run of its generator for the first 50 steps
  (oracle + task oracle + the product's own episode initialisation)."""
  import sys
  sys.path.insert(0, os.path.join(os.path.dirname(GOLDEN), "..", "tools"))
  g = np.load(os.path.join(GOLDEN, "c1_cartpole_swingup.npz"))
  from oracle.oracle_py import Oracle, OracleModel
  m = PU.load_model("cartpole")
  o = Oracle(OracleModel(m))
  o.reset()
  o.qpos[:] = g["qpos0"]; o.qvel[:] = g["qvel0"]
  o.forward()
  fields = ("qpos", "qvel", "ctrl", "xmat", "xpos", "site_xpos", "sensordata")
  for k in range(50):
    o.ctrl[:] = g["actions"][k]
    o.step()
    d = {f: np.array(getattr(o, f), dtype=np.float64).reshape(1, -1) for f in fields}
    obs, r = task_oracle.obs_reward("cartpole", "swingup", m, d)
    np.testing.assert_allclose(obs[0], g["obs"][k], rtol=1e-12, atol=1e-14)
    np.testing.assert_allclose(r[0], g["reward"][k], rtol=1e-12, atol=1e-14)
