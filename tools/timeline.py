"""Stage wall times of block 0 (debug build with -DBP_TIMELINE): cycles between barriers."""
import ctypes, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from dm_control_b200 import binding
name = sys.argv[1]; lanes = int(sys.argv[2])
bench.LANES["value"] = lanes
nenv = bench.WORKLOADS[name]["nenv"]
tdef, m, model, batch, nsub = bench.make_batch(name, nenv, 0, 1000)
L = binding.load()
out = (ctypes.c_ulonglong * 32)()
rng = np.random.RandomState(0)
for k in range(bench.WORKLOADS[name]["settle"] + 10):
  batch.write("ctrl", rng.uniform(-1, 1, (nenv, m.nu))); batch.step(nsub)
dbg = getattr(L, "bp_debug_timeline_g%d" % lanes)
dbg(out, 1)
N = 20
for k in range(N):
  batch.write("ctrl", rng.uniform(-1, 1, (nenv, m.nu))); batch.step(nsub)
dbg(out, 0)
names = {0: "kin..factorM", 1: "collision", 2: "makeConstraint", 3: "velocity(fwd)", 4: "act+acc",
         5: "solver", 6: "acc sensors+euler", 7: "velocity(step1)", 8: "newton iters"}
tot = sum(out[i] for i in range(16))
for i in range(16):
  if out[16 + i]:
    print("%-18s %6.2f%%  %10.0f cycles/sub-step  (%d barriers)" % (names.get(i, str(i)), 100.0 * out[i] / tot, out[i] / (N * nsub), out[16 + i]))
print("total cycles/sub-step %.0f; newton iterations per sub-step (block 0): %.2f" % (tot / (N * nsub), out[16 + 8] / (N * nsub)))
