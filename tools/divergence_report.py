"""1000-step trajectory and return divergence, GPU vs CPU oracle (BASELINE.json north_star:
"Because the dynamics are chaotic, trajectory and return divergence over 1000 steps is
reported rather than held to a tolerance").

For every BASELINE config: `nenv` envs built as bench.py builds them, the task's settle steps
on the GPU, then the oracle takes over a copy of every env's state and both sides take the same
1000 random actions.  Per env-step: the qpos error of every env; per env: the return (sum of the
task's reward, GPU: the fused reward kernel, oracle: oracle/task_oracle.py).  Writes
profiles/r2_divergence.json and .md.  TEST INFRASTRUCTURE (runs the oracle).

  python tools/divergence_report.py [nenv] [steps]
"""
import concurrent.futures
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import torch  # noqa: E402

import bench  # noqa: E402
from oracle import task_oracle  # noqa: E402
from oracle.oracle_py import Oracle, OracleModel  # noqa: E402

TASKS = ["cartpole-swingup", "acrobot-swingup", "pendulum-swingup", "cheetah-run", "walker-walk",
         "humanoid-run", "quadruped-escape", "finger-spin"]
FIELDS = ("qpos", "qvel", "ctrl", "xmat", "xpos", "xquat", "site_xpos", "site_xmat",
          "sensordata", "subtree_com", "geom_xpos", "act", "cvel", "xipos")
MARKS = (1, 10, 100, 300, 1000)


def run(name, nenv, steps):
  w = bench.WORKLOADS[name]
  tdef, m, model, batch, nsub = bench.make_batch(name, nenv, 0, seed=1000)
  dev = "cuda:0"
  gen = torch.Generator(device=dev); gen.manual_seed(0)
  for k in range(w["settle"]):
    c = torch.rand((nenv, m.nu), generator=gen, device=dev, dtype=torch.float64) * 2 - 1
    torch.cuda.synchronize()
    batch.copy_from_device("ctrl", c.data_ptr())
    batch.step(1 if name.startswith("cheetah") else nsub)
  batch.sync()
  state = {f: batch.read(f) for f in ("qpos", "qvel", "act", "qacc_warmstart", "time")}
  hfield = batch.read_hfield() if int(m.nhfielddata) > 0 else None
  mo = m.copy()
  mo.arrays["njmax"][:] = w["njmax"]; mo.arrays["nconmax"][:] = w["nconmax"]
  om = OracleModel(mo)
  oracles = []
  for e in range(nenv):
    o = Oracle(om)
    if hfield is not None:
      o.hfield_data[:] = hfield[e]
    o.reset()
    o.qpos[:] = state["qpos"][e]; o.qvel[:] = state["qvel"][e]
    if m.na:
      o.act[:] = state["act"][e]
    o.qacc_warmstart[:] = state["qacc_warmstart"][e]
    o.time[:] = state["time"][e]
    o.step1()
    oracles.append(o)
  _, fp = tdef.params(m)
  speed = fp[0] if tdef.domain in ("walker", "humanoid", "quadruped", "humanoid_CMU") else None
  obs = torch.empty((nenv, batch.nobs), device=dev, dtype=torch.float64)
  rew = torch.empty((nenv,), device=dev, dtype=torch.float64)
  pool = concurrent.futures.ThreadPoolExecutor(os.cpu_count() or 8)

  def advance(args):
    o, c = args
    o.ctrl[:] = c
    for _ in range(nsub):
      o.step()

  ret_g, ret_o = np.zeros(nenv), np.zeros(nenv)
  curve = {}
  first_int = None        # first step at which some env's contact count differs
  for step in range(1, steps + 1):
    c = torch.rand((nenv, m.nu), generator=gen, device=dev, dtype=torch.float64) * 2 - 1
    torch.cuda.synchronize()
    batch.env_step(nsub, c.data_ptr(), obs.data_ptr(), rew.data_ptr())
    ch = c.cpu().numpy()
    list(pool.map(advance, [(o, ch[e]) for e, o in enumerate(oracles)]))
    batch.sync()
    ret_g += rew.cpu().numpy()
    d = {f: np.stack([np.array(getattr(o, f), dtype=np.float64) for o in oracles]) for f in FIELDS}
    _, r = task_oracle.obs_reward(tdef.domain, tdef.task, m, d, speed)
    ret_o += r
    if step in MARKS or step == steps:
      q = batch.read("qpos")
      err = np.max(np.abs(q - d["qpos"]), axis=1)
      curve[step] = {"median": float(np.median(err)), "p90": float(np.percentile(err, 90)),
                     "max": float(err.max()),
                     "envs_within_1e-6": float(np.mean(err < 1e-6))}
    if first_int is None and step % 10 == 0:
      nc = batch.read("ncon")[:, 0]
      if not np.array_equal(nc, np.array([int(o.ncon[0]) for o in oracles])):
        first_int = step
  pool.shutdown()
  dret = np.abs(ret_g - ret_o)
  out = {"workload": bench.workload_name(name, nenv, nsub), "envs": nenv, "env_steps": steps,
         "n_sub_steps": nsub, "qpos_error": curve,
         "first_step_with_a_different_contact_count_(checked_every_10)": first_int,
         "return": {"mean_gpu": float(ret_g.mean()), "mean_oracle": float(ret_o.mean()),
                    "abs_diff_median": float(np.median(dret)), "abs_diff_p90":
                    float(np.percentile(dret, 90)), "abs_diff_max": float(dret.max()),
                    "abs_diff_of_means": float(abs(ret_g.mean() - ret_o.mean()))}}
  batch.free(); model.free()
  return out


def main():
  nenv = int(sys.argv[1]) if len(sys.argv) > 1 else 256
  steps = int(sys.argv[2]) if len(sys.argv) > 2 else 1000
  res = {}
  for name in TASKS:
    res[name] = run(name, nenv, steps)
    print(name, json.dumps(res[name]["qpos_error"]), json.dumps(res[name]["return"]), flush=True)
  os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
  with open(os.path.join(ROOT, "gpurun_out", "r2_divergence.json"), "w") as f:
    json.dump(res, f, indent=1)
  lines = ["# 1000-step trajectory and return divergence, GPU vs oracle (%d envs per config)" % nenv, "",
           "qpos error = max over the env's coordinates of |GPU - oracle|; median / max over envs.",
           "", "| config | step 1 | step 10 | step 100 | step 300 | step %d | envs still within 1e-6 at the end | "
           "mean return GPU / oracle | median / max abs return diff per env |" % steps, "|---|---|---|---|---|---|---|---|---|"]
  for name, r in res.items():
    c = r["qpos_error"]
    cell = lambda s: "%.1e / %.1e" % (c[s]["median"], c[s]["max"]) if s in c else "-"
    lines.append("| %s | %s | %s | %s | %s | %s | %.0f %% | %.3f / %.3f | %.2e / %.2e |" % (
        name, cell(1), cell(10), cell(100), cell(300), cell(steps),
        100 * c[steps]["envs_within_1e-6"], r["return"]["mean_gpu"], r["return"]["mean_oracle"],
        r["return"]["abs_diff_median"], r["return"]["abs_diff_max"]))
  with open(os.path.join(ROOT, "gpurun_out", "r2_divergence.md"), "w") as f:
    f.write("\n".join(lines) + "\n")
  print("\n".join(lines))


if __name__ == "__main__":
  main()
