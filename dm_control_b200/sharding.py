"""Env-batch sharding over the GPUs of one box: one process per GPU, no per-step
collective (BASELINE.json north_star).  Environments are independent, so rank r owns
a contiguous slice of the global env axis and steps it alone; the only exchange is the
optional all-gather of episode statistics (NCCL on the GPUs, gloo in the CPU tests).

A job of E global envs seeded with `seed` draws env i's episode randomness from
`RandomState(seed + i)` whatever the world size, so an N-GPU run reproduces the 1-GPU
run env for env (tests/test_sharding_gloo.py).
"""

import os

import numpy as np


def world():
  """(rank, world_size, local_rank) from the torchrun environment."""
  return (int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")),
          int(os.environ.get("LOCAL_RANK", "0")))


def shard(total_envs, rank, world_size):
  """Contiguous slice [start, start + count) of the global env axis owned by `rank`;
  the first `total_envs % world_size` ranks hold one env more."""
  if not 0 <= rank < world_size:
    raise ValueError("rank {} outside world of {}".format(rank, world_size))
  if total_envs < 0:
    raise ValueError("total_envs must be >= 0")
  base, rem = divmod(int(total_envs), int(world_size))
  start = rank * base + min(rank, rem)
  return start, base + (1 if rank < rem else 0)


def env_seeds(seed, total_envs, rank, world_size):
  """Per-env seeds of this rank's slice (suite.Task seeds env i with seed + i)."""
  start, count = shard(total_envs, rank, world_size)
  return [int(seed) + start + i for i in range(count)]


def load_shard(domain_name, task_name, total_envs, seed, rank=None, world_size=None,
               device=None, **kw):
  """`suite.load` for this rank's slice of a `total_envs` job."""
  from dm_control_b200 import suite
  r, w, local = world()
  rank = r if rank is None else rank
  world_size = w if world_size is None else world_size
  seeds = env_seeds(seed, total_envs, rank, world_size)
  return suite.load(domain_name, task_name,
                    task_kwargs={"random": seeds, "model_seed": seed}, nenv=len(seeds),
                    device=local if device is None else device, **kw)


def gather_episode_stats(local_stats, total_envs=None):
  """All-gather of per-env episode statistics ([count, k] float64 tensor or array on this
  rank) into the global [total_envs, k] array, in global env order.  The one collective
  of the data path; a no-op copy when torch.distributed is not initialised."""
  import torch
  import torch.distributed as dist
  t = torch.as_tensor(local_stats)
  if t.dim() == 1:
    t = t[:, None]
  if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
    return t.detach().cpu().numpy()
  w = dist.get_world_size()
  counts = torch.zeros(w, dtype=torch.int64, device=t.device)
  counts[dist.get_rank()] = t.shape[0]
  dist.all_reduce(counts)
  cmax = int(counts.max())
  pad = torch.zeros((cmax, t.shape[1]), dtype=t.dtype, device=t.device)
  pad[:t.shape[0]] = t
  parts = [torch.empty_like(pad) for _ in range(w)]
  dist.all_gather(parts, pad)
  out = torch.cat([p[:int(c)] for p, c in zip(parts, counts)], dim=0)
  if total_envs is not None and out.shape[0] != total_envs:
    raise RuntimeError("gathered {} envs, expected {}".format(out.shape[0], total_envs))
  return out.cpu().numpy()


def max_over_ranks(x, device=None):
  """Device-side elapsed time as the job sees it: max over ranks."""
  import torch
  import torch.distributed as dist
  if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
    return float(x)
  t = torch.tensor([float(x)], dtype=torch.float64, device=device or "cpu")
  dist.all_reduce(t, op=dist.ReduceOp.MAX)
  return float(t.item())
