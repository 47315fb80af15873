"""bp_rewards_tolerance_noalloc (include/bphys.h): the entry point that allocates no device
memory, against the existing bp_rewards_tolerance, through the C ABI.  Named test_zz_* so
that pytest collects it after every other file.  It must pass
alone in a fresh process and after the other GPU tests: it depends on no state of theirs
and the scratch array is per process."""

import numpy as np
import pytest

from dm_control_b200 import binding

CASES = [  # (bounds, margin, sigmoid, value_at_margin)
    ((0.0, 0.0), 0.0, "gaussian", 0.1),
    ((-0.25, 0.25), 2.0, "gaussian", 0.1),
    ((1.2, float("inf")), 0.6, "linear", 0.5),
    ((0.0, 0.0), 1.0, "quadratic", 0.0),
    ((0.0, 0.05), 0.25, "long_tail", 0.1),
]


@pytest.mark.gpu
def test_noalloc_matches_existing_entry_point_on_two_calls_in_a_row():
  nmax = binding.load().bp_rewards_tolerance_noalloc_max()
  assert nmax == 65536
  rng = np.random.RandomState(11)
  # call 1: a large input that fills the whole scratch; call 2: a shorter one with other
  # parameters, so stale contents of call 1 would show up as wrong values
  xs = [rng.uniform(-3, 3, size=nmax), rng.uniform(-3, 3, size=1000)]
  for bounds, margin, sig, v1 in CASES:
    for x in xs:
      want = binding.rewards_tolerance(x, bounds, margin, sig, v1)
      got = binding.rewards_tolerance_noalloc(x, bounds, margin, sig, v1)
      # the other tolerance tests compare to 1e-12; same kernel arithmetic => bit-exact
      np.testing.assert_array_equal(got, want)
  # immediately repeated call with identical arguments gives identical output
  a = binding.rewards_tolerance_noalloc(xs[1], (0.0, 0.0), 2.0, "gaussian", 0.1)
  b = binding.rewards_tolerance_noalloc(xs[1], (0.0, 0.0), 2.0, "gaussian", 0.1)
  np.testing.assert_array_equal(a, b)


@pytest.mark.gpu
def test_noalloc_rejects_sizes_beyond_the_scratch_and_bad_arguments():
  nmax = binding.load().bp_rewards_tolerance_noalloc_max()
  with pytest.raises(binding.Error, match="exceeds the fixed scratch"):
    binding.rewards_tolerance_noalloc(np.zeros(nmax + 1))
  with pytest.raises(binding.Error, match="Lower bound"):
    binding.rewards_tolerance_noalloc(np.zeros(4), bounds=(1.0, 0.0))
  # an empty input is fine and the entry point still works after the errors
  assert binding.rewards_tolerance_noalloc(np.zeros(0)).shape == (0,)
  x = np.linspace(-1, 1, 17)
  np.testing.assert_array_equal(binding.rewards_tolerance_noalloc(x, (0, 0), 1.0),
                                binding.rewards_tolerance(x, (0, 0), 1.0))
