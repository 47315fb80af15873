"""fp64 operations the CPU oracle actually executes per physics sub-step (SURVEY.md 8d,
"binding rule"), measured with the op-count build of the oracle (oracle/opcount.h) on the
bench workloads: random initial joint angles, `settle` sub-steps, then `steps` sub-steps
with actions ~ U(-1, 1) redrawn every env-step, averaged over `nenv` environments.

  python tools/count_ops.py            -> profiles/r1_oracle_opcount.json

flops = add + mul + div + sqrt + transcendental (an FMA counts as 2; comparisons are
listed but not counted).  TEST INFRASTRUCTURE: runs on the CPU, never on the product path.
"""

import ctypes
import json
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from oracle import oracle_py  # noqa: E402

CLASSES = ("add", "mul", "div", "sqrt", "transcendental", "compare")
WORK = {"cheetah-run": 200, "walker-walk": 500, "humanoid-run": 500, "hopper-hop": 200,
        "finger-spin": 100, "quadruped-walk": 400, "quadruped-escape": 400,
        "cartpole-swingup": 0, "acrobot-swingup": 0, "pendulum-swingup": 0}


# r2: the contact-rich variants of the bench (actions 0.1 * U(-1, 1): the robots rest on the
# ground) -> profiles/r2_oracle_opcount.json
WORK_REST = {"cheetah-run@rest": 400, "humanoid-run@rest": 500}


def main(nenv=32, steps=200, work=WORK, scale=1.0, out_name="r1_oracle_opcount.json"):
  so = os.path.join(ROOT, "oracle", "_build", "liboracle_ops.so")
  subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "oracle"),
                         "_build/liboracle_ops.so"])
  oracle_py.build = lambda force=False: so
  oracle_py._lib = None
  L = oracle_py.lib()
  L.o_opcount_read.argtypes = [ctypes.POINTER(ctypes.c_ulonglong)]
  import bench
  from dm_control_b200.suite import tasks as T
  out = {}
  for key, settle in work.items():
    name = key.split("@")[0]
    tdef, m = bench.load_task(name)
    nsub = T.n_sub_steps(tdef, m)
    qpos = bench.initial_state(name, m, nenv, 0)
    hf = bench.initial_hfield(name, m, nenv, 0)
    om = oracle_py.OracleModel(m)
    rng = np.random.RandomState(1)
    tot = np.zeros(len(CLASSES))
    stats = np.zeros(3)
    for e in range(nenv):
      o = oracle_py.Oracle(om)
      if hf is not None:
        o.hfield_data[:] = hf[e]
      o.reset(); o.qpos[:] = qpos[e]; o.forward()
      for k in range(settle + steps):
        if k % nsub == 0:
          o.ctrl[:] = scale * rng.uniform(-1, 1, m.nu)
        if k == settle:
          L.o_opcount_reset()
        o.step()
        if k >= settle:
          stats += (o.ncon[0], o.nefc[0], o.solver_iter[0])
      buf = (ctypes.c_ulonglong * len(CLASSES))()
      L.o_opcount_read(buf)
      tot += np.array(list(buf), dtype=np.float64)
    per = tot / (nenv * steps)
    flops = float(per[:5].sum())
    st = stats / (nenv * steps)
    out[key] = {"flops_per_substep": flops, "nsub": nsub,
                 "flops_per_env_step": flops * nsub,
                 "classes": dict(zip(CLASSES, [float(x) for x in per])),
                 "mean_ncon": float(st[0]), "mean_nefc": float(st[1]),
                 "mean_solver_iter": float(st[2]),
                 "sample": "%d envs x %d sub-steps after %d settle sub-steps" % (nenv, steps, settle)}
    print("%-18s %10.0f flops/sub-step  (add %.0f mul %.0f div %.0f sqrt %.0f trans %.0f) "
          "ncon %.2f nefc %.2f iter %.2f" % ((key, flops) + tuple(per[:5]) + tuple(st)))
  path = os.path.join(ROOT, "profiles", out_name)
  with open(path, "w") as f:
    json.dump(out, f, indent=1, sort_keys=True)
  print("wrote", path)


if __name__ == "__main__":
  if "--rest" in sys.argv:
    main(work=WORK_REST, scale=0.1, out_name="r2_oracle_opcount.json")
  else:
    main()
