"""quadruped-escape rollout on the GPU with random actions: contact / solver statistics,
warnings and reward sanity (run on a GPU box; prints one JSON line)."""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from dm_control_b200 import suite   # noqa: E402


def main():
  nenv = int(sys.argv[1]) if len(sys.argv) > 1 else 256
  nstep = int(sys.argv[2]) if len(sys.argv) > 2 else 200
  task = sys.argv[3] if len(sys.argv) > 3 else "escape"
  kw = {}
  if len(sys.argv) > 4:
    kw = dict(njmax=int(sys.argv[4]), nconmax=int(sys.argv[5]))
  t0 = time.time()
  env = suite.load("quadruped", task, task_kwargs={"random": 0}, nenv=nenv, **kw)
  ts = env.reset()
  t_reset = time.time() - t0
  spec = env.action_spec()
  rng = np.random.RandomState(0)
  stats = {"ncon": [], "nefc": [], "iter": []}
  rew = []
  errors = 0
  t0 = time.time()
  for k in range(nstep):
    a = rng.uniform(spec.minimum, spec.maximum, (nenv, spec.shape[0]))
    try:
      ts = env.step(a)
    except Exception as e:   # PhysicsError: report and go on
      errors += 1
      if errors < 3:
        print("step", k, repr(e)[:300])
      continue
    rew.append(float(np.mean(ts.reward)))
    if k % 10 == 0:
      d = env.physics.data
      stats["ncon"].append(d.ncon); stats["nefc"].append(d.nefc); stats["iter"].append(d.solver_iter)
  dt = time.time() - t0
  d = env.physics.data
  out = {"task": task, "nenv": nenv, "steps": nstep, "reset_s": round(t_reset, 2),
         "env_steps_per_s_host_loop": nenv * nstep / dt,
         "errors": errors, "reward_mean": float(np.mean(rew)), "reward_max": float(np.max(rew)),
         "z_mean": float(np.mean(d.qpos[:, 2])), "finite": bool(np.isfinite(d.qpos).all()),
         "warnings": env.physics._batch.warnings().sum(axis=0).tolist()}
  for k, v in stats.items():
    v = np.concatenate(v)
    out[k] = {"mean": float(v.mean()), "p99": float(np.percentile(v, 99)), "max": int(v.max())}
  rf = np.asarray(ts.observation.get("rangefinder", np.zeros((1, 1))))
  out["rangefinder_hit_frac"] = float(np.mean(rf < 1.0))
  print(json.dumps(out))
  env.close()


if __name__ == "__main__":
  main()
