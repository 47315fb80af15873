"""CPU checks of the test oracles against the committed golden fixtures."""

import os

import numpy as np
import pytest

import parity_util as PU
from oracle import task_oracle
from oracle.oracle_py import Oracle, OracleModel

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
SIG = {0: "gaussian", 1: "linear", 2: "quadratic"}


def test_tolerance_restatement_matches_reference_rewards_py():
  # fixture = outputs of /root/reference/dm_control/utils/rewards.py itself
  g = np.load(os.path.join(GOLDEN, "rewards_tolerance.npz"))
  for (lo, hi, mg, sg, v1), want in zip(g["cases"], g["out"]):
    got = task_oracle.tolerance(g["x"], (lo, hi), mg, SIG[int(sg)], v1)
    np.testing.assert_allclose(got, want, rtol=0, atol=1e-15)


@pytest.mark.parametrize("domain", ["pendulum", "cartpole", "cheetah", "humanoid"])
def test_oracle_reproduces_committed_trajectories(domain):
  g = np.load(os.path.join(GOLDEN, "traj_%s.npz" % domain))
  m = PU.load_model(domain)
  om = OracleModel(m)
  steps = list(g["steps"])
  for e in range(2):                         # two envs keep the CPU suite fast
    o = Oracle(om)
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
