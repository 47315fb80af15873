"""Prints per-field GPU-vs-oracle errors for the suite models (run on a GPU box)."""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import parity_util as PU   # noqa: E402


def main():
  lanes_list = [int(x) for x in os.environ.get("BP_LANES", "32,1,2,4,8,16").split(",")]
  domains = sys.argv[1:] or ["pendulum", "cartpole", "acrobot", "cheetah", "walker",
                             "hopper", "humanoid"]
  os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
  with open(os.path.join(ROOT, "gpurun_out", "parity_report.txt"), "w") as f:
    for d, lanes in [(d, l) for d in domains for l in lanes_list]:
      t = time.time()
      kw = dict(njmax=120, nconmax=40) if d in ("humanoid", "quadruped") else {}
      if d == "quadruped_escape":
        kw = dict(njmax=300, nconmax=64)
      if d == "humanoid_CMU":
        kw = dict(njmax=200, nconmax=48)
      try:
        res, stats = PU.run_parity(d, nenv=int(os.environ.get('BP_NENV', 64)), nsteps=20, seed=int(os.environ.get('BP_SEED', 1)),
                                   check_steps=[0, 1, 5, 20], lanes=lanes, **kw)
        d = '%s/G%d' % (d, lanes)
      except Exception as e:   # keep going: this is a report, not a test
        msg = "%s/G%d: FAILED %r" % (d, lanes, e)
        print(msg); f.write(msg + "\n")
        continue
      for step, cmp in sorted(res.items()):
        bad_int = {k: v for k, v in cmp.int_mismatch.items() if v}
        msg = "%s step %d worst %.3e int-mismatch %s" % (d, step, cmp.worst(), bad_int)
        print(msg); f.write(msg + "\n")
        f.write(cmp.report() + "\n")
      msg = "%s stats %s (%.1fs)" % (d, stats, time.time() - t)
      print(msg); f.write(msg + "\n")
      f.flush()


if __name__ == "__main__":
  main()
