import sys, os
sys.path.insert(0, "tests"); sys.path.insert(0, ".")
import numpy as np
import parity_util as PU
from dm_control_b200 import binding
from oracle.oracle_py import Oracle, OracleModel
dom = sys.argv[1] if len(sys.argv) > 1 else "stacker_4"
m = PU.load_model(dom)
nenv = 48
qpos, qvel, ctrl = PU.random_states(m, nenv, 5)
model = binding.Model(m); b = binding.Batch(model, nenv, 160, 40, 8)
b.write("qpos", qpos); b.write("qvel", qvel); b.write("ctrl", ctrl)
mo = m.copy(); mo.arrays["njmax"][:] = 160; mo.arrays["nconmax"][:] = 40
om = OracleModel(mo)
os_ = []
for e in range(nenv):
  o = Oracle(om); o.reset(); o.qpos[:] = qpos[e]; o.qvel[:] = qvel[e]; o.ctrl[:] = ctrl[e]; os_.append(o)
names = ["plane","hfield","sphere","capsule","ellipsoid","cylinder","box"]
for step in range(4):
  if step == 0:
    b.forward(); [o.forward() for o in os_]
  else:
    b.step(1); [o.step() for o in os_]
  b.sync()
  cp = b.read("contact_pos"); cd = b.read("contact_dist"); g1 = b.read("contact_geom1"); g2 = b.read("contact_geom2"); nc = b.read("ncon")[:, 0]
  cf = b.read("contact_frame")
  shown = 0
  for e, o in enumerate(os_):
    n = int(o.ncon[0])
    if n != nc[e]:
      print("step", step, "env", e, "ncon differs", n, nc[e]); continue
    for i in range(n):
      dp = np.abs(cp[e, 3*i:3*i+3] - np.array(o.contact_pos[3*i:3*i+3])).max()
      if dp > 1e-9 and shown < 6:
        shown += 1
        a, bb = int(o.contact_geom1[i]), int(o.contact_geom2[i])
        print("step", step, "env", e, "contact", i, "geoms", a, bb, names[m.geom_type[a]], names[m.geom_type[bb]],
              "pos gpu", cp[e, 3*i:3*i+3], "oracle", np.array(o.contact_pos[3*i:3*i+3]), "dist", cd[e, i], o.contact_dist[i],
              "n gpu", cf[e, 9*i:9*i+3], "n oracle", np.array(o.contact_frame[9*i:9*i+3]))
