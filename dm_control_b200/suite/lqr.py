"""The LQR domain (reference: /root/reference/dm_control/suite/lqr.py, lqr.xml): a chain of
`n_bodies` masses on slide joints with random spring stiffness, the first `n_actuators`
driven by motors; cost = sum(position^2) + c * sum(control^2).

The model is generated, not loaded from a cache: stiffness and damping are drawn from the
task's RandomState when the environment is built (lqr.py:142-200).  The spatial tendons the
reference strings between consecutive bodies are "for visualisation purposes only"
(lqr.py:189) and are left out; lighting, cameras and materials are too.
"""

import numpy as np

CONTROL_COST_COEF = 0.1          # lqr.py:33
TERMINAL_TOL = 1e-6              # lqr.py:220

# suite/lqr.xml without the rendering includes
_BASE = """<mujoco model="LQR">
  <option timestep=".03"/>
  <default>
    <joint type="slide" axis="0 1 0"/>
    <geom type="sphere" size=".1"/>
    <site size=".01"/>
  </default>
  <option>
    <flag constraint="disable"/>
  </option>
  <worldbody>
    <geom name="floor" size="4 1 .2" type="plane"/>
    <geom name="origin" pos="2 0 .05" size="2 .003 .05" type="box"/>
    %s
  </worldbody>
  <actuator>
    %s
  </actuator>
</mujoco>
"""


def make_model(n_bodies, n_actuators, random, stiffness_range=(15, 25), damping_range=(0, 0)):
  """lqr.py:142-200 `_make_model`: same draws in the same order (stiffness, then damping,
  body by body)."""
  if n_bodies < 1 or n_actuators < 1:
    raise ValueError("At least 1 body and 1 actuator required.")
  if n_actuators > n_bodies:
    raise ValueError("At most 1 actuator per body.")
  bodies, close, motors = "", "", ""
  for body in range(n_bodies):
    stiffness = random.uniform(stiffness_range[0], stiffness_range[1])
    damping = random.uniform(damping_range[0], damping_range[1])
    pos = ".25 0 .1" if body == 0 else ".25 0 0"
    bodies += ('<body name="body_%d" pos="%s"><joint name="joint_%d" stiffness="%r" '
               'damping="%r"/><geom name="geom_%d"/><site name="site_%d"/>'
               % (body, pos, body, stiffness, damping, body, body))
    close += "</body>"
    if body < n_actuators:
      motors += '<motor name="motor_%d" joint="joint_%d"/>' % (body, body)
  return _BASE % (bodies + close, motors)


def full_mass_matrix(m, qM):
  """mj_fullM (lqr_solver.py:49): dense nv x nv matrix from the sparse qM of one env."""
  nv = int(m.nv)
  out = np.zeros((nv, nv))
  for i in range(nv):
    a = int(m.dof_Madr[i])
    j = i
    while j >= 0:
      out[i, j] = out[j, i] = qM[a]
      a += 1
      j = int(m.dof_parentid[j])
  return out


def solve(m, qM, control_cost_coef=CONTROL_COST_COEF):
  """suite/lqr_solver.py:30-98 for a compiled model: (p, k, beta) of the discrete Riccati
  equation; V(x) = .5 x' p x, u = k x, state -> 0 like beta^n."""
  from scipy import linalg as scipy_linalg
  n, nu = int(m.nq), int(m.nu)
  mass = full_mass_matrix(m, qM)
  stiffness = np.diag(np.asarray(m.jnt_stiffness).ravel())
  damping = np.diag(np.asarray(m.dof_damping).ravel())
  dt = float(m.opt_timestep)
  j = np.linalg.solve(-mass, np.hstack((stiffness, damping)))
  a = np.eye(2 * n) + dt * np.vstack((dt * j + np.hstack((np.zeros((n, n)), np.eye(n))), j))
  moment = np.zeros((nu, n))                      # data.actuator_moment of joint motors
  for u in range(nu):
    moment[u, int(m.jnt_dofadr[int(m.actuator_trnid[u, 0])])] = m.actuator_gear[u, 0]
  b = moment.T
  bc = np.linalg.solve(mass, b)
  b = dt * np.vstack((dt * bc, bc))
  q = np.diag(np.hstack([np.ones(n), np.zeros(n)]))
  r = control_cost_coef * np.eye(nu)
  p = scipy_linalg.solve_discrete_are(a, b, q, r)
  k = -np.linalg.solve(b.T.dot(p.dot(b)) + r, b.T.dot(p.dot(a)))
  beta = np.abs(np.linalg.eigvals(a + b.dot(k))).max()
  if beta >= 1.0:
    raise RuntimeError("Controlled system is unstable.")
  return p, k, beta
