"""Shared helpers of the GPU-vs-oracle parity tests (test infrastructure).

States are seeded with numpy so that both sides see identical inputs; the oracle
(oracle/mjstep_oracle.c) is the checker, the CUDA path behind include/bphys.h is
the thing checked.
"""

import os

import numpy as np

from dm_control_b200.model import FlatModel, JNT_FREE, JNT_BALL

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
MODELS = os.path.join(ROOT, "dm_control_b200", "suite", "models")

# fields that exist on both sides, by dtype
F64_FIELDS = ["qpos", "qvel", "act", "qacc", "qacc_warmstart", "act_dot", "time",
              "energy", "xpos", "xquat", "xmat", "xipos", "ximat", "xanchor",
              "xaxis", "geom_xpos", "geom_xmat", "site_xpos", "site_xmat",
              "subtree_com", "cinert", "cdof", "cdof_dot", "cvel", "crb", "qM", "qLD",
              "qLDiagInv", "qfrc_bias", "qfrc_passive", "qfrc_actuator",
              "qfrc_smooth", "qacc_smooth", "qfrc_constraint", "actuator_length",
              "actuator_velocity", "actuator_force", "sensordata", "cacc",
              "cfrc_int", "cfrc_ext", "subtree_linvel"]
EFC_F64 = ["efc_pos", "efc_margin", "efc_frictionloss", "efc_diagApprox", "efc_D",
           "efc_R", "efc_vel", "efc_aref", "efc_force"]
EFC_I32 = ["efc_type", "efc_id", "efc_state"]
CON_F64 = [("contact_dist", 1), ("contact_pos", 3), ("contact_frame", 9)]
CON_I32 = ["contact_geom1", "contact_geom2", "contact_dim", "contact_efc_address"]
INT_SCALARS = ["ncon", "nefc", "ne", "nf", "solver_iter", "warning"]
TENDON_F64 = ["ten_length", "ten_velocity", "ten_J"]


def load_model(domain):
  return FlatModel.load(os.path.join(MODELS, domain + ".npz"))


def random_states(m, nenv, seed, vel_scale=1.0, spread=1.0):
  """Seeded (qpos, qvel, ctrl) covering joint limits, contacts and free bodies."""
  rng = np.random.RandomState(seed)
  qpos = np.tile(m.qpos0, (nenv, 1))
  for j in range(m.njnt):
    jt, qa = m.jnt_type[j], m.jnt_qposadr[j]
    if jt == JNT_FREE:
      qpos[:, qa:qa + 2] += rng.uniform(-.5, .5, (nenv, 2))
      qpos[:, qa + 2] = rng.uniform(0.1, 1.6, nenv)
      q = rng.randn(nenv, 4)
      qpos[:, qa + 3:qa + 7] = q / np.linalg.norm(q, axis=1, keepdims=True)
    elif jt == JNT_BALL:
      q = rng.randn(nenv, 4)
      qpos[:, qa:qa + 4] = q / np.linalg.norm(q, axis=1, keepdims=True)
    elif m.jnt_limited[j]:
      lo, hi = m.jnt_range[j]
      w = (hi - lo) * 0.05            # a little beyond the limits: active limit rows
      qpos[:, qa] = rng.uniform(lo - w, hi + w, nenv)
    else:
      qpos[:, qa] += spread * rng.uniform(-1.0, 1.0, nenv)
  qvel = vel_scale * rng.randn(nenv, m.nv)
  ctrl = rng.uniform(-1, 1, (nenv, m.nu))
  return qpos, qvel, ctrl


def escape_terrain(rng, res=201, half_size=30.0):
  """suite/quadruped.py:366-380 (Escape.initialize_episode): sinusoidal bowl times smooth
  random bumps, as float32 like mjModel.hfield_data."""
  from scipy import ndimage
  row_grid, col_grid = np.ogrid[-1:1:res * 1j, -1:1:res * 1j]
  radius = np.clip(np.sqrt(col_grid ** 2 + row_grid ** 2), .04, 1)
  bowl_shape = .5 - np.cos(2 * np.pi * radius) / 2
  bump_res = int(2 * half_size / 2)
  bumps = rng.uniform(.15, 1, (bump_res, bump_res))
  smooth_bumps = ndimage.zoom(bumps, res / float(bump_res))
  return (bowl_shape * smooth_bumps).ravel().astype(np.float32)


def terrain_states(m, qpos, hfield, seed):
  """Moves the free root of every env to a random spot of its terrain, a little above
  or slightly inside the surface, so that height-field contacts of every kind occur."""
  rng = np.random.RandomState(seed + 77)
  nrow, ncol = int(m.hfield_nrow[0]), int(m.hfield_ncol[0])
  rx, ry, zs, _ = m.hfield_size[0]
  for e in range(qpos.shape[0]):
    x, y = rng.uniform(-0.8 * rx, 0.8 * rx), rng.uniform(-0.8 * ry, 0.8 * ry)
    c = int(round((x + rx) / (2 * rx) * (ncol - 1))); r = int(round((y + ry) / (2 * ry) * (nrow - 1)))
    z = float(hfield[e, r * ncol + c]) * zs
    qpos[e, 0:3] = x, y, z + rng.uniform(0.05, 0.45)


def rel_err(a, b):
  """max |a-b| scaled by the magnitude of the reference array b (>= 1e-12)."""
  a = np.asarray(a, dtype=np.float64); b = np.asarray(b, dtype=np.float64)
  if a.size == 0:
    return 0.0
  if not (np.all(np.isfinite(a)) and np.all(np.isfinite(b))):
    return np.inf
  return float(np.max(np.abs(a - b)) / max(np.max(np.abs(b)), 1e-6))


# Element-wise bar (BASELINE.json north_star: "within 1e-8 relative"): every element must
# satisfy |a - b| <= ELEM_RTOL * |b| + ELEM_ATOL * max(1, max|b|), where b is the oracle's
# array of that field for that env.  The absolute floor is what fp64 cancellation leaves on
# an entry that is small next to its neighbours (a 1e-3 entry of an array whose largest
# entry is 1e3 is held to 1e-11 + 1e-7 absolute, not to 1e-5 as the array-normalised
# rel_err alone would allow).  elem_err returns max over elements of |a - b| / bar: <= 1 passes.
ELEM_RTOL = 1e-8
ELEM_ATOL = 1e-10


def elem_err(a, b):
  a = np.asarray(a, dtype=np.float64); b = np.asarray(b, dtype=np.float64)
  if a.size == 0:
    return 0.0
  if not (np.all(np.isfinite(a)) and np.all(np.isfinite(b))):
    return np.inf
  bar = ELEM_RTOL * np.abs(b) + ELEM_ATOL * max(1.0, float(np.max(np.abs(b))))
  return float(np.max(np.abs(a - b) / bar))


class Comparison:
  """Accumulates per-field worst errors over envs and steps."""

  def __init__(self):
    self.f64 = {}
    self.elem = {}
    self.int_mismatch = {}

  def add_f64(self, name, a, b):
    self.f64[name] = max(self.f64.get(name, 0.0), rel_err(a, b))
    self.elem[name] = max(self.elem.get(name, 0.0), elem_err(a, b))

  def worst_elem(self):
    return max(self.elem.values()) if self.elem else 0.0

  def add_int(self, name, a, b):
    bad = int(np.sum(np.asarray(a) != np.asarray(b)))
    self.int_mismatch[name] = self.int_mismatch.get(name, 0) + bad

  def worst(self):
    return max(self.f64.values()) if self.f64 else 0.0

  def report(self):
    lines = ["  %-22s %.3e   elementwise %.3e of its bar" % (k, v, self.elem.get(k, 0.0))
             for k, v in sorted(self.f64.items(), key=lambda kv: -kv[1])]
    lines += ["  %-22s %d mismatches" % (k, v) for k, v in
              sorted(self.int_mismatch.items())]
    return "\n".join(lines)


def compare_env(cmp, m, oracle, gpu, e, nvp, skip=()):
  """Compares every shared field of env `e` (gpu: dict name -> [nenv, n])."""
  nv = m.nv
  for name in F64_FIELDS:
    if name in skip:
      continue
    cmp.add_f64(name, gpu[name][e], getattr(oracle, name))
  for name in INT_SCALARS:
    if name in skip:
      continue
    cmp.add_int(name, gpu[name][e], getattr(oracle, name))
  if m.ntendon:
    for name in TENDON_F64:
      cmp.add_f64(name, gpu[name][e], getattr(oracle, name))
  nefc, ncon = int(oracle.nefc[0]), int(oracle.ncon[0])
  if int(gpu["nefc"][e, 0]) == nefc and int(gpu["ncon"][e, 0]) == ncon:
    for name in EFC_F64:
      if name not in skip:
        cmp.add_f64(name, gpu[name][e, :nefc], getattr(oracle, name)[:nefc])
    for name in EFC_I32:
      if name not in skip:
        cmp.add_int(name, gpu[name][e, :nefc], getattr(oracle, name)[:nefc])
    gj = gpu["efc_J"][e].reshape(-1, nvp)[:nefc, :nv]
    cmp.add_f64("efc_J", gj, oracle.efc_J[:nefc * nv].reshape(nefc, nv))
    for name, w in CON_F64:
      cmp.add_f64(name, gpu[name][e, :ncon * w], getattr(oracle, name)[:ncon * w])
    for name in CON_I32:
      cmp.add_int(name, gpu[name][e, :ncon], getattr(oracle, name)[:ncon])


def run_parity(domain, nenv=32, nsteps=3, seed=0, njmax=0, nconmax=0, check_steps=None,
               vel_scale=1.0, spread=1.0, mutate=None, lanes=0, env_model=None,
               iter_slack=False, state_hook=None, flat_model=None):
  """forward() then `nsteps` Physics.step() on both sides; returns (Comparison per
  phase, stats).  The GPU side goes through the C ABI (dm_control_b200.binding).

  iter_slack: the Newton solver stops when improvement or gradient falls below
  opt.tolerance; on a stiff model an env can sit on that threshold and stop one iteration
  apart on the two sides, which leaves a tolerance-sized (not rounding-sized) difference.
  With iter_slack those envs are taken out of the strict comparison from the step their
  solver_iter first differs and collected in `cmp.loose`; stats["loose"] counts them."""
  from dm_control_b200 import binding
  from oracle.oracle_py import Oracle, OracleModel
  m = flat_model.copy() if flat_model is not None else load_model(domain)
  if mutate:
    mutate(m)
  qpos, qvel, ctrl = random_states(m, nenv, seed, vel_scale, spread)
  if state_hook:
    state_hook(m, qpos, qvel, ctrl)
  hfield = None
  if int(m.nhfielddata) > 0:       # one terrain per env (quadruped.py:371-382)
    hfield = np.stack([escape_terrain(np.random.RandomState(seed + 1000 + e))
                       for e in range(nenv)])
    terrain_states(m, qpos, hfield, seed)
  model = binding.Model(m)
  batch = binding.Batch(model, nenv, njmax, nconmax, lanes)
  nvp = m.nv | 1
  batch.write("qpos", qpos); batch.write("qvel", qvel); batch.write("ctrl", ctrl)
  if hfield is not None:
    batch.write_hfield(hfield)
  for k, v in (env_model or {}).items():     # per-env model frames (reacher's target)
    batch.write_env_model(k, v)
  if njmax or nconmax:             # the oracle sizes its buffers from the model
    m = m.copy()
    if njmax: m.arrays["njmax"][:] = njmax
    if nconmax: m.arrays["nconmax"][:] = nconmax
  om = OracleModel(m)
  oracles = []
  for e in range(nenv):
    o = Oracle(om)
    if hfield is not None:
      o.hfield_data[:] = hfield[e]
    for k, v in (env_model or {}).items():
      getattr(o, k)[:] = np.asarray(v[e]).ravel()
    o.reset()
    o.qpos[:] = qpos[e]; o.qvel[:] = qvel[e]; o.ctrl[:] = ctrl[e]
    oracles.append(o)
  names = (F64_FIELDS + EFC_F64 + EFC_I32 + ["efc_J"] + [n for n, _ in CON_F64] +
           CON_I32 + INT_SCALARS + (TENDON_F64 if m.ntendon else []))
  out = {}
  check_steps = set(range(nsteps + 1)) if check_steps is None else set(check_steps)
  stats = {"ncon": [], "nefc": [], "iter": [], "loose": 0}
  loose = np.zeros(nenv, dtype=bool)
  for step in range(nsteps + 1):
    if step == 0:
      batch.forward()
      for o in oracles:
        o.forward()
    else:
      batch.step(1)
      for o in oracles:
        o.step()
    if iter_slack:
      batch.sync()
      it = batch.read("solver_iter")[:, 0]
      loose |= it != np.array([int(o.solver_iter[0]) for o in oracles])
      stats["loose"] = int(loose.sum())
    if step in check_steps:
      batch.sync()
      gpu = {n: batch.read(n) for n in names}
      cmp = Comparison()
      cmp.loose = Comparison()
      for e, o in enumerate(oracles):
        compare_env(cmp.loose if loose[e] else cmp, m, o, gpu, e, nvp)
      out[step] = cmp
      stats["ncon"].append(float(np.mean(gpu["ncon"])))
      stats["nefc"].append(float(np.mean(gpu["nefc"])))
      stats["iter"].append(float(np.mean(gpu["solver_iter"])))
  batch.free(); model.free()
  return out, stats
