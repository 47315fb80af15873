"""Histogram of solver iterations / contacts / constraint rows over the envs of a workload."""
import sys, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
name = sys.argv[1] if len(sys.argv) > 1 else "humanoid-run"
nenv = bench.WORKLOADS[name]["nenv"]
tdef, m, model, batch, nsub = bench.make_batch(name, nenv, 0, 1000)
rng = np.random.RandomState(0)
for k in range(bench.WORKLOADS[name]["settle"] + 20):
  batch.write("ctrl", rng.uniform(-1, 1, (nenv, m.nu)))
  batch.step(nsub)
batch.sync()
for f in ("solver_iter", "ncon", "nefc"):
  v = batch.read(f)[:, 0]
  print(f, "mean %.2f" % v.mean(), "hist", np.bincount(v)[:40].tolist())
it = batch.read("solver_iter")[:, 0]
for blk in (4, 8, 56, 112):
  n = len(it) // blk * blk
  print("mean of max over groups of %d envs: %.2f" % (blk, it[:n].reshape(-1, blk).max(axis=1).mean()))
