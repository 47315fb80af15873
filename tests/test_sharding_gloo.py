"""N>1 host logic on CPU: world_size-2 `gloo` processes exercising the env sharding,
per-env seeding and the episode-statistics all-gather of dm_control_b200/sharding.py
(the only collective of the path; NCCL on the GPU box)."""

import os
import socket
import types

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import parity_util as PU
from dm_control_b200 import sharding


def test_shard_partitions_the_env_axis():
  for total in (0, 1, 7, 16384, 65536 + 3):
    for w in (1, 2, 3, 8):
      spans = [sharding.shard(total, r, w) for r in range(w)]
      assert spans[0][0] == 0 and sum(c for _, c in spans) == total
      for (s0, c0), (s1, _) in zip(spans, spans[1:]):
        assert s0 + c0 == s1
      assert max(c for _, c in spans) - min(c for _, c in spans) <= 1
  with pytest.raises(ValueError):
    sharding.shard(8, 2, 2)
  assert sharding.env_seeds(100, 10, 1, 2) == [105, 106, 107, 108, 109]


class _FakeBatch:
  nobs = 3

  def set_task(self, *a):
    pass


def _fake_physics(m, n):
  data = types.SimpleNamespace(qpos=np.tile(m.qpos0, (n, 1)), qvel=np.zeros((n, m.nv)))
  return types.SimpleNamespace(model=m, nenv=n, data=data, _batch=_FakeBatch())


def _initial_qpos(seeds):
  from dm_control_b200 import suite
  from dm_control_b200.suite import tasks as T
  m = PU.load_model("pendulum")
  task = suite.Task(T.REGISTRY[("pendulum", "swingup")], random=seeds)
  phys = _fake_physics(m, len(seeds))
  task.bind(phys)
  task.initialize_episode(phys)
  return phys.data.qpos.copy()


def _worker(rank, world, port, total, q):
  os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank),
                    WORLD_SIZE=str(world), LOCAL_RANK=str(rank))
  dist.init_process_group("gloo", rank=rank, world_size=world)
  try:
    assert sharding.world() == (rank, world, rank)
    seeds = sharding.env_seeds(40, total, rank, world)
    qpos = _initial_qpos(seeds)                           # this rank's episode starts
    start, count = sharding.shard(total, rank, world)
    stats = np.concatenate([qpos[:, :1], np.arange(start, start + count)[:, None] * 1.0],
                           axis=1)
    allstats = sharding.gather_episode_stats(torch.from_numpy(stats), total)
    tmax = sharding.max_over_ranks(1.0 + rank)
    if rank == 0:
      q.put((allstats, tmax))
  finally:
    dist.destroy_process_group()


def test_two_ranks_reproduce_the_single_process_job():
  total = 11                                             # ragged: 6 + 5
  s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
  ctx = mp.get_context("spawn")
  q = ctx.Queue()
  procs = [ctx.Process(target=_worker, args=(r, 2, port, total, q)) for r in range(2)]
  for p in procs:
    p.start()
  allstats, tmax = q.get(timeout=120)
  for p in procs:
    p.join(120)
    assert p.exitcode == 0
  assert tmax == 2.0                                      # max over ranks
  np.testing.assert_array_equal(allstats[:, 1], np.arange(total))   # global env order
  single = _initial_qpos(sharding.env_seeds(40, total, 0, 1))
  np.testing.assert_array_equal(allstats[:, 0], single[:, 0])       # same envs, any N


def test_gather_without_process_group_is_identity():
  x = np.arange(6.0).reshape(3, 2)
  np.testing.assert_array_equal(sharding.gather_episode_stats(x), x)
  assert sharding.max_over_ranks(3.5) == 3.5
