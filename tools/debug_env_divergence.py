"""Finds the env and step where a GPU batch and the oracle part ways (run on a GPU box).
usage: debug_env_divergence.py DOMAIN [seed nenv nsteps lanes njmax nconmax]"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import parity_util as PU   # noqa: E402


def main():
  from dm_control_b200 import binding
  from oracle.oracle_py import Oracle, OracleModel
  a = sys.argv[1:]
  domain = a[0]
  seed, nenv, nsteps, lanes, njmax, nconmax = [int(x) for x in a[1:7]]
  m = PU.load_model(domain)
  qpos, qvel, ctrl = PU.random_states(m, nenv, seed)
  model = binding.Model(m)
  batch = binding.Batch(model, nenv, njmax, nconmax, lanes)
  batch.write("qpos", qpos); batch.write("qvel", qvel); batch.write("ctrl", ctrl)
  m = m.copy(); m.arrays["njmax"][:] = njmax; m.arrays["nconmax"][:] = nconmax
  om = OracleModel(m)
  oracles = []
  for e in range(nenv):
    o = Oracle(om); o.reset()
    o.qpos[:] = qpos[e]; o.qvel[:] = qvel[e]; o.ctrl[:] = ctrl[e]
    oracles.append(o)
  names = m.names["geom"]
  for step in range(nsteps + 1):
    if step == 0:
      batch.forward(); [o.forward() for o in oracles]
    else:
      batch.step(1); [o.step() for o in oracles]
    if os.environ.get("BP_ONLY_LAST") and step < nsteps:
      continue
    batch.sync()
    if os.environ.get("BP_READ_ALL"):
      allnames = (PU.F64_FIELDS + PU.EFC_F64 + PU.EFC_I32 + ["efc_J"] + [n for n, _ in PU.CON_F64] +
                  PU.CON_I32 + PU.INT_SCALARS + PU.TENDON_F64)
      for nm in allnames[:int(os.environ["BP_READ_ALL"])]:
        batch.read(nm)
    if os.environ.get("BP_READ_ORACLE"):
      for o in oracles:
        for nm in PU.F64_FIELDS + PU.EFC_F64 + PU.TENDON_F64:
          getattr(o, nm)
    g = {n: batch.read(n) for n in ("qpos", "qacc", "ncon", "nefc", "contact_dist",
                                    "contact_geom1", "contact_geom2", "contact_frame",
                                    "contact_pos", "efc_force", "solver_iter")}
    err = np.array([max(PU.rel_err(g["qacc"][e], o.qacc), PU.rel_err(g["qpos"][e], o.qpos),
                        PU.rel_err(g["contact_frame"][e, :9 * int(o.ncon[0])],
                                   o.contact_frame[:9 * int(o.ncon[0])]))
                    for e, o in enumerate(oracles)])
    e = int(np.argmax(err)); o = oracles[e]
    print("step %d worst qacc err %.2e env %d ncon gpu %d oracle %d iter %d/%d" %
          (step, err[e], e, g["ncon"][e, 0], o.ncon[0], g["solver_iter"][e, 0], o.solver_iter[0]))
    if err[e] > 1e-9:
      nc = int(o.ncon[0])
      for c in range(nc):
        print("   con %d %s - %s dist gpu %.12g oracle %.12g  normal err %.1e pos err %.1e" %
              (c, names[o.contact_geom1[c]], names[o.contact_geom2[c]], g["contact_dist"][e, c],
               o.contact_dist[c],
               np.abs(g["contact_frame"][e, 9 * c:9 * c + 9] - o.contact_frame[9 * c:9 * c + 9]).max(),
               np.abs(g["contact_pos"][e, 3 * c:3 * c + 3] - o.contact_pos[3 * c:3 * c + 3]).max()))
      for c in range(nc):
        print("   frame gpu", g["contact_frame"][e, 9 * c:9 * c + 9])
        print("   frame ora", np.array(o.contact_frame[9 * c:9 * c + 9]))
      if err[e] > 1e-3:
        break


if __name__ == "__main__":
  main()
