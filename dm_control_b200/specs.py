"""Minimal stand-ins for the `dm_env` types the reference returns (dm_env is not
installed here): TimeStep / StepType with the same field names
(/root/reference/dm_control/rl/control.py:88-92,120-123) and Array / BoundedArray
specs (rl/control.py:368-393, mujoco/engine.py:1018-1028)."""

import collections
import enum

import numpy as np


class StepType(enum.IntEnum):
  FIRST = 0
  MID = 1
  LAST = 2


class TimeStep(collections.namedtuple(
    "TimeStep", ["step_type", "reward", "discount", "observation"])):
  """Batched: step_type is an int array [nenv] (StepType values); reward and
  discount are float arrays [nenv] (None on FIRST, as in the reference)."""
  __slots__ = ()

  def first(self):
    return np.asarray(self.step_type) == StepType.FIRST

  def mid(self):
    return np.asarray(self.step_type) == StepType.MID

  def last(self):
    return np.asarray(self.step_type) == StepType.LAST


class Array:
  def __init__(self, shape, dtype, name=None):
    self.shape, self.dtype, self.name = tuple(shape), np.dtype(dtype), name

  def __repr__(self):
    return "Array(shape={}, dtype={}, name={!r})".format(self.shape, self.dtype, self.name)


class BoundedArray(Array):
  def __init__(self, shape, dtype, minimum, maximum, name=None):
    super().__init__(shape, dtype, name)
    self.minimum = np.broadcast_to(np.asarray(minimum, dtype=dtype), shape)
    self.maximum = np.broadcast_to(np.asarray(maximum, dtype=dtype), shape)
