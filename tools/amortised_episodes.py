"""Amortised throughput over whole episodes, resets included: the
device-resident TorchVecEnv (dm_control_b200/wrappers.py) steps `nenv` envs of a suite task
through `episodes` full episodes with random actions generated on the device; every episode
end runs the task's initialize_episode for all envs (host RNG streams, cheetah: + 200 settle
steps on the device).  Prints one JSON line.

  python tools/amortised_episodes.py cheetah run 16384 2
"""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

from dm_control_b200 import suite, wrappers  # noqa: E402


def main(domain="cheetah", task="run", nenv=16384, episodes=2):
  env = suite.load(domain, task, task_kwargs={"random": 0}, nenv=nenv)
  v = wrappers.TorchVecEnv(env, check_every=100)
  nu = env.action_spec().shape[0]
  steps = int(round(env._step_limit))
  t0 = time.perf_counter()
  v.reset()
  torch.cuda.synchronize()
  t_reset = time.perf_counter() - t0
  gen = torch.Generator(device="cuda"); gen.manual_seed(0)
  ret = torch.zeros(nenv, dtype=torch.float64, device="cuda")
  ndone = 0
  t0 = time.perf_counter()
  for k in range(episodes * steps):
    a = torch.rand((nenv, nu), generator=gen, device="cuda", dtype=torch.float64) * 2 - 1
    obs, rew, done, _ = v.step(a)
    ret += rew
    if (k + 1) % steps == 0:
      ndone += int(done.sum())
  torch.cuda.synchronize()
  dt = time.perf_counter() - t0
  out = {"task": "%s-%s" % (domain, task), "envs": nenv, "episodes": episodes,
         "steps_per_episode": steps, "n_sub_steps": env._n_sub_steps,
         "amortised_env_steps_per_s": nenv * episodes * steps / dt,
         "wall_s": dt, "first_reset_s": t_reset, "envs_done": ndone,
         "mean_return_per_episode": float(ret.mean()) / episodes,
         "note": "TorchVecEnv: actions and results stay on the device; each episode end "
                 "includes the host-side initialize_episode of every env"}
  print(json.dumps(out))
  v.close()


if __name__ == "__main__":
  a = sys.argv[1:]
  main(a[0] if a else "cheetah", a[1] if len(a) > 1 else "run",
       int(a[2]) if len(a) > 2 else 16384, int(a[3]) if len(a) > 3 else 2)
