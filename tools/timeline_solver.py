"""Cycles of block 0's first group inside one Newton iteration, by sub-stage (debug build
-DBP_TIMELINE, generic LEAN kernels: BPHYS_NO_STATIC=1 BPHYS_LIB_PATH=variants/libbphys_tl.so)."""
import ctypes, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ["BPHYS_NO_STATIC"] = "1"
import bench
from dm_control_b200 import binding
name = sys.argv[1]; lanes = int(sys.argv[2]); nenv = int(sys.argv[3]) if len(sys.argv) > 3 else bench.WORKLOADS[name]["nenv"]
bench.LANES["value"] = lanes
tdef, m, model, batch, nsub = bench.make_batch(name, nenv, 0, 1000)
L = binding.load()
out = (ctypes.c_ulonglong * 64)()
rng = np.random.RandomState(0)
for k in range(bench.WORKLOADS[name]["settle"] + 10):
  batch.write("ctrl", rng.uniform(-1, 1, (nenv, m.nu))); batch.step(nsub)
dbg = getattr(L, "bp_debug_tl2_g%d_lean" % lanes)
dbg(out, 1)
N = 200
for k in range(N):
  batch.write("ctrl", rng.uniform(-1, 1, (nenv, m.nu))); batch.step(nsub)
dbg(out, 0)
names = ["mul_md + mul_j (Mv, jv)", "linesearch", "update + constraint update + cost", "gradient (J^T f, M^-1 grad)",
         "H: copy M + single-entry rows", "H: contact rows", "cholesky factor", "cholesky solve", "solve only (factor reused)"]
tot = sum(out[i] for i in range(32))
for i, nm in enumerate(names):
  if out[32 + i]:
    print("%-40s %6.2f%%  %8.0f cycles per call  (%d calls)" % (nm, 100.0 * out[i] / tot, out[i] / out[32 + i], out[32 + i]))
print("envs %d: total %.0f cycles per physics sub-step in these sub-stages (first group of block 0)" % (nenv, tot / (N * nsub)))
