"""`suite.load(domain, task)` for the device backend.

Mirrors /root/reference/dm_control/suite/__init__.py:93-150 (`load`, `build_environment`)
for the domains the batched backend serves (tasks.py).  Environments are batched:
`load(..., nenv=N)` returns a `control.Environment` whose TimeSteps carry a leading
env axis.  Models come from the compiled caches in `suite/models/` (built from the
reference's XML by tools/compile_suite_models.py); episode initialisation restates
each task's `initialize_episode` with one `numpy.random.RandomState` per env.
"""

import os

import numpy as np

from dm_control_b200 import control
from dm_control_b200 import model as M
from dm_control_b200.physics import BatchedPhysics, PhysicsError
from dm_control_b200.suite import tasks as T

_MODELS = os.path.join(os.path.dirname(os.path.abspath(__file__)), "models")
ALL_TASKS = T.ALL_TASKS
BENCHMARKING = T.ALL_TASKS


_RANDOMIZER_PLANS = {}


def _randomizer_plan(m):
  """The draws of randomizers.py:35-86 for one model, in joint order, with runs of consecutive
  scalar uniform draws (limited hinge / slide joints, unlimited hinges) merged into one
  vectorised `uniform(lo, hi)` call: RandomState.uniform with array bounds computes
  lo + (hi - lo) * random_sample() element by element from the same stream, so the values
  equal those of the reference's one-call-per-joint loop bit for bit (pinned by
  tests/test_golden_cpu.py against the reference's own randomizers.py)."""
  # cached per model; `set_model_field` replaces the edited array object, which invalidates it
  a = m.arrays
  key = (a["jnt_range"], a["jnt_limited"], a["jnt_type"])
  plan = _RANDOMIZER_PLANS.get(id(m))
  if plan is not None and plan[0] is m and all(x is y for x, y in zip(plan[1], key)):
    return plan[2]
  ops, run = [], []

  def flush():
    if run:
      ops.append(("u", np.array([r[0] for r in run]), np.array([r[1] for r in run]),
                  np.array([r[2] for r in run])))
      del run[:]
  for j in range(m.njnt):
    jt, qa = int(m.jnt_type[j]), int(m.jnt_qposadr[j])
    lo, hi = (float(x) for x in m.jnt_range[j])
    if m.jnt_limited[j]:
      if jt in (M.JNT_HINGE, M.JNT_SLIDE):
        run.append((qa, lo, hi))
      elif jt == M.JNT_BALL:
        flush(); ops.append(("ball_limited", qa, hi))
    elif jt == M.JNT_HINGE:
      run.append((qa, -np.pi, np.pi))
    elif jt == M.JNT_BALL:
      flush(); ops.append(("ball", qa))
    elif jt == M.JNT_FREE:
      flush(); ops.append(("free", qa))
  flush()
  _RANDOMIZER_PLANS[id(m)] = (m, key, ops)
  return ops


def _randomize_limited_and_rotational_joints(m, qpos, rng):
  """suite/utils/randomizers.py:35-86 for one env (qpos: [nq], edited in place)."""
  for op in _randomizer_plan(m):
    kind, qa = op[0], op[1]
    if kind == "u":
      qpos[qa] = rng.uniform(op[2], op[3])
    elif kind == "ball_limited":     # randomizers.py:23-33 random_limited_quaternion
      axis = rng.randn(3)
      axis /= np.linalg.norm(axis)
      angle = rng.rand() * op[2]
      qpos[qa] = np.cos(angle / 2)             # mju_axisAngle2Quat
      qpos[qa + 1:qa + 4] = axis * np.sin(angle / 2)
    elif kind == "ball":
      q = rng.randn(4)
      qpos[qa:qa + 4] = q / np.linalg.norm(q)
    else:                            # free joint: a random orientation, position untouched
      q = rng.rand(4)
      qpos[qa + 3:qa + 7] = q / np.linalg.norm(q)


class Task:
  """Host side of a suite task: RNG streams, episode initialisation, obs layout.
  Observation and reward themselves run on the device (csrc/bp_task.cu)."""

  def __init__(self, tdef, random=None):
    self.tdef = tdef
    self._seed = random
    self._rngs = None
    self._mask = None
    self.scalar_obs = {"height", "head_height", "torso_upright", "dist_to_target"}

  def bind(self, physics):
    m = physics.model
    n = physics.nenv
    if isinstance(self._seed, (list, tuple, np.ndarray)):
      self._rngs = [r if isinstance(r, np.random.RandomState) else np.random.RandomState(r)
                    for r in self._seed]
    elif isinstance(self._seed, np.random.RandomState):
      # base.Task accepts a RandomState (suite/base.py:40-50); the batch needs one stream per
      # env, drawn from it
      self._rngs = [np.random.RandomState(int(s))
                    for s in self._seed.randint(0, 2 ** 31 - 1, size=n)]
    elif self._seed is None:
      self._rngs = [np.random.RandomState(None) for _ in range(n)]
    else:
      self._rngs = [np.random.RandomState(int(self._seed) + i) for i in range(n)]
    ip, fp = self.tdef.params(m)
    physics._batch.set_task(self.tdef.kind, ip, fp)
    self.nobs = physics._batch.nobs
    self.obs_layout = self.tdef.obs(m)

  @property
  def random(self):
    return self._rngs

  def get_termination(self, physics, obs):
    """base.Task.get_termination: None (no task-defined termination) or a [nenv] array,
    NaN where the episode goes on and the terminal discount where it ends.  Only the LQR
    levels terminate (suite/lqr.py:256-259: state norm below 1e-6 -> discount 0)."""
    if self.tdef.domain != "lqr":
      return None
    from dm_control_b200.suite import lqr
    return np.where(np.linalg.norm(obs, axis=1) < lqr.TERMINAL_TOL, 0.0, np.nan)

  def _each(self):
    """(env, generator) of the envs this initialisation covers (all, or the masked ones)."""
    if self._mask is None:
      return list(enumerate(self._rngs))
    return [(int(e), self._rngs[e]) for e in np.flatnonzero(self._mask)]

  def _todo(self, n):
    return np.ones(n, dtype=bool) if self._mask is None else self._mask.copy()

  def initialize_episode(self, physics, mask=None):
    """`mask` ([nenv] bool): per-env auto-reset -- only those envs draw from their
    generators and are written (the caller holds physics.reset_context(mask=mask), inside
    which data writes, model-frame writes and steps reach the masked envs only)."""
    self._mask = None if mask is None else np.asarray(mask, dtype=bool)
    try:
      self._initialize_episode(physics)
    finally:
      self._mask = None

  def _initialize_episode(self, physics):
    m = physics.model
    n = physics.nenv
    qpos = physics.data.qpos
    qvel = np.zeros((n, m.nv))
    d, t = self.tdef.domain, self.tdef.task
    if d == "cartpole":           # suite/cartpole.py:177-195
      sl = m.jnt_qposadr[m.name2id("slider", "joint")]
      h1 = m.jnt_qposadr[m.name2id("hinge_1", "joint")]
      for e, rng in self._each():
        if t.startswith("swingup") or t.endswith("_poles"):
          qpos[e, sl] = .01 * rng.randn()
          qpos[e, h1] = np.pi + .01 * rng.randn()
          qpos[e, 2:] = .1 * rng.randn(m.nv - 2)
        else:
          qpos[e, sl] = rng.uniform(-.1, .1)
          qpos[e, 1:] = rng.uniform(-.034, .034, m.nv - 1)
        qvel[e] = 0.01 * rng.randn(m.nv)
    elif d == "acrobot":          # suite/acrobot.py:96-106
      ids = [m.jnt_qposadr[m.name2id(k, "joint")] for k in ("shoulder", "elbow")]
      for e, rng in self._each():
        qpos[e, ids] = rng.uniform(-np.pi, np.pi, 2)
    elif d == "pendulum":         # suite/pendulum.py:80-90
      h = m.jnt_qposadr[m.name2id("hinge", "joint")]
      for e, rng in self._each():
        qpos[e, h] = rng.uniform(-np.pi, np.pi)
    elif d == "cheetah":          # suite/cheetah.py:63-77
      lim = m.jnt_limited == 1
      lo, hi = m.jnt_range[lim].T
      for e, rng in self._each():
        qpos[e, lim] = rng.uniform(lo, hi)
      physics.data.qpos = qpos
      # cheetah.py:73-75 steps straight away: the first mj_step2 of the settle run still sees
      # the quantities reset() derived at qpos0 (actuation disabled), as in the reference
      for _ in range(4):
        physics.step(50)                    # 200 physics steps, as the reference
      physics.data.time = 0.0
      return
    elif d == "hopper":           # suite/hopper.py:116-120
      for e, rng in self._each():
        _randomize_limited_and_rotational_joints(m, qpos[e], rng)
    elif d == "walker":           # suite/walker.py:118-129
      for e, rng in self._each():
        _randomize_limited_and_rotational_joints(m, qpos[e], rng)
    elif d == "quadruped":
      self._initialize_quadruped(physics, t)
      return
    elif d == "point_mass":       # suite/point_mass.py:85-107
      wprm = np.tile(np.asarray(m.wrap_prm, dtype=np.float64), (n, 1))
      for e, rng in self._each():
        _randomize_limited_and_rotational_joints(m, qpos[e], rng)
        if t == "hard":           # each control actuates a random combination of the joints
          dir1 = rng.randn(2)
          dir1 /= np.linalg.norm(dir1)
          parallel = True         # a second direction that is not 'too parallel' to the first
          while parallel:
            dir2 = rng.randn(2)
            dir2 /= np.linalg.norm(dir2)
            parallel = abs(np.dot(dir1, dir2)) > 0.9
          wprm[e, [0, 1]] = dir1
          wprm[e, [2, 3]] = dir2
      if t == "hard":             # per-env model.wrap_prm (bp_env_model_write)
        physics.set_env_model_field("wrap_prm", wprm)
    elif d == "reacher":          # suite/reacher.py:91-103
      tgt = m.name2id("target", "geom")
      size = np.array(m.geom_size, dtype=np.float64)
      size[tgt, 0] = .05 if t == "easy" else .015
      physics.set_model_field("geom_size", size)
      gpos = np.tile(np.asarray(m.geom_pos, dtype=np.float64), (n, 1, 1))
      for e, rng in self._each():
        _randomize_limited_and_rotational_joints(m, qpos[e], rng)
        angle = rng.uniform(0, 2 * np.pi)
        radius = rng.uniform(.05, .20)
        gpos[e, tgt, 0] = radius * np.sin(angle)
        gpos[e, tgt, 1] = radius * np.cos(angle)
      physics.set_env_model_field("geom_pos", gpos)
    elif d == "swimmer":          # suite/swimmer.py:176-197
      tgt = m.name2id("target", "geom")
      gpos = np.tile(np.asarray(m.geom_pos, dtype=np.float64), (n, 1, 1))
      for e, rng in self._each():
        _randomize_limited_and_rotational_joints(m, qpos[e], rng)
        close_target = rng.rand() < .2
        target_box = .3 if close_target else 2
        gpos[e, tgt, :2] = rng.uniform(-target_box, target_box, size=2)
      physics.set_env_model_field("geom_pos", gpos)
    elif d == "manipulator":      # suite/manipulator.py:186-247
      _initialize_manipulator(self, physics, qpos, qvel, use_peg="peg" in t,
                              insert=t.startswith("insert"))
      return
    elif d == "stacker":          # suite/stacker.py:139-176
      _initialize_stacker(self, physics, qpos, int(t[-1]))
      return
    elif d == "ball_in_cup":      # suite/ball_in_cup.py:69-84: a ball position without contacts
      bx = m.jnt_qposadr[m.name2id("ball_x", "joint")]
      bz = m.jnt_qposadr[m.name2id("ball_z", "joint")]
      todo = self._todo(n)
      while todo.any():
        for e in np.flatnonzero(todo):
          qpos[e, bx] = self._rngs[e].uniform(-.2, .2)
          qpos[e, bz] = self._rngs[e].uniform(.2, .5)
        physics.data.qpos = qpos
        physics.after_reset()
        todo &= physics.data.ncon > 0
      return
    elif d == "fish":             # suite/fish.py:107-116,139-150
      tgt = m.name2id("target", "geom")
      root = m.jnt_qposadr[m.name2id("root", "joint")]
      gpos = np.tile(np.asarray(m.geom_pos, dtype=np.float64), (n, 1, 1))
      for e, rng in self._each():
        quat = rng.randn(4)
        qpos[e, root + 3:root + 7] = quat / np.linalg.norm(quat)
        for j in range(7):
          qpos[e, 7 + j] = rng.uniform(-.2, .2)
        if t == "swim":
          gpos[e, tgt, 0] = rng.uniform(-.4, .4)
          gpos[e, tgt, 1] = rng.uniform(-.4, .4)
          gpos[e, tgt, 2] = rng.uniform(.1, .3)
      if t == "swim":
        physics.set_env_model_field("geom_pos", gpos)
    elif d == "lqr":              # suite/lqr.py:229-233: a random point of a sphere
      for e, rng in self._each():
        unit = rng.randn(m.nq)
        qpos[e] = np.sqrt(2) * unit / np.linalg.norm(unit)
    elif d in ("humanoid", "finger", "humanoid_CMU"):
      # suite/humanoid.py:152-166, suite/finger.py:139-144,201-212: rejection on contacts
      if d == "finger" and t == "spin":   # finger.py:142: the same edit for every env
        damping = np.array(m.dof_damping, dtype=np.float64)
        damping[m.jnt_dofadr[m.name2id("hinge", "joint")]] = .03
        physics.set_model_field("dof_damping", damping)
      elif d == "finger":         # finger.py:172-181 (turn): a target per env
        hinge = m.name2id("hinge", "joint")
        anchor = physics.data.xanchor.reshape(n, -1, 3)[:, hinge]     # valid after reset()
        radius = float(np.sum(m.geom_size[m.name2id("cap1", "geom")]))
        tgt = m.name2id("target", "site")
        spos = np.tile(np.asarray(m.site_pos, dtype=np.float64), (n, 1, 1))
        for e, rng in self._each():
          angle = rng.uniform(-np.pi, np.pi)
          spos[e, tgt, 0] = anchor[e, 0] + radius * np.sin(angle)
          spos[e, tgt, 2] = anchor[e, 2] + radius * np.cos(angle)
        physics.set_env_model_field("site_pos", spos)
        size = np.array(m.site_size, dtype=np.float64)
        size[tgt, 0] = self.tdef.params(m)[1][1]
        physics.set_model_field("site_size", size)
      todo = self._todo(n)
      while todo.any():
        for e in np.flatnonzero(todo):
          _randomize_limited_and_rotational_joints(m, qpos[e], self._rngs[e])
        physics.data.qpos = qpos
        physics.after_reset()
        todo &= physics.data.ncon > 0
      return
    physics.data.qpos = qpos
    physics.data.qvel = qvel


def _initialize_stacker(self, physics, qpos, n_boxes):
  """suite/stacker.py:139-176 for every env at once: arm angles, the target's height (a
  random level of the stack) and x, the boxes' poses, redrawn from the env's own generator
  until nothing touches."""
  from dm_control_b200.suite.tasks import MANIPULATOR_ARM_JOINTS
  m, n = physics.model, physics.nenv
  arm = [m.name2id(j, "joint") for j in MANIPULATOR_ARM_JOINTS]
  arm_adr = np.array([m.jnt_qposadr[j] for j in arm])
  limited = np.asarray(m.jnt_limited)[arm].astype(bool)
  rng_ = np.asarray(m.jnt_range)[arm]
  lower = np.where(limited, rng_[:, 0], -np.pi)
  upper = np.where(limited, rng_[:, 1], np.pi)
  finger = m.jnt_qposadr[m.name2id("finger", "joint")]
  thumb = m.jnt_qposadr[m.name2id("thumb", "joint")]
  target = m.name2id("target", "body")
  box_size = float(m.geom_size[m.name2id("target", "geom"), 0])
  box_adr = [[m.jnt_qposadr[m.name2id("box%d_%s" % (b, k), "joint")] for k in "xzy"]
             for b in range(n_boxes)]
  bpos = physics.env_model_field("body_pos") if physics._batch.has_env_model("body_pos") else \
      np.tile(np.asarray(m.body_pos, dtype=np.float64), (n, 1, 1))
  todo = self._todo(n)
  while todo.any():
    for e in np.flatnonzero(todo):
      rng = self._rngs[e]
      qpos[e, arm_adr] = rng.uniform(lower, upper)
      qpos[e, finger] = qpos[e, thumb]
      target_height = 2 * rng.randint(n_boxes) + 1
      bpos[e, target, 2] = box_size * target_height
      bpos[e, target, 0] = rng.uniform(-.37, .37)
      for adr in box_adr:
        qpos[e, adr[0]] = rng.uniform(.1, .3)
        qpos[e, adr[1]] = rng.uniform(0, .7)
        qpos[e, adr[2]] = rng.uniform(0, 2 * np.pi)
    physics.set_env_model_field("body_pos", bpos)
    physics.data.qpos = qpos
    physics.after_reset()
    todo &= physics.data.ncon > 0


def _initialize_manipulator(self, physics, qpos, qvel, use_peg=False, insert=False):
  """suite/manipulator.py:186-247 for every env at once: each env draws from its own
  generator in the reference's order until its configuration is collision-free; the
  'in_hand' envs read the grasp site after a forward pass with their new arm angles."""
  from dm_control_b200.suite.tasks import MANIPULATOR_ARM_JOINTS
  m, n = physics.model, physics.nenv
  arm = [m.name2id(j, "joint") for j in MANIPULATOR_ARM_JOINTS]
  arm_adr = np.array([m.jnt_qposadr[j] for j in arm])
  limited = np.asarray(m.jnt_limited)[arm].astype(bool)
  rng_ = np.asarray(m.jnt_range)[arm]
  lower = np.where(limited, rng_[:, 0], -np.pi)
  upper = np.where(limited, rng_[:, 1], np.pi)
  finger = m.jnt_qposadr[m.name2id("finger", "joint")]
  thumb = m.jnt_qposadr[m.name2id("thumb", "joint")]
  oname = "peg" if use_peg else "ball"
  target = m.name2id("target_" + oname, "body")
  receptacle = m.name2id("slot" if use_peg else "cup", "body") if insert else -1
  obj = [m.jnt_qposadr[m.name2id(oname + "_" + k, "joint")] for k in "xzy"]
  obj_xvel = m.jnt_dofadr[m.name2id(oname + "_x", "joint")]
  grasp = m.name2id("grasp", "site")
  if physics._batch.has_env_model("body_pos"):     # partial reset: the others keep theirs
    bpos, bquat = physics.env_model_field("body_pos"), physics.env_model_field("body_quat")
  else:
    bpos = np.tile(np.asarray(m.body_pos, dtype=np.float64), (n, 1, 1))
    bquat = np.tile(np.asarray(m.body_quat, dtype=np.float64), (n, 1, 1))
  todo = self._todo(n)
  while todo.any():
    idx = np.flatnonzero(todo)
    kind, pose = {}, {}
    for e in idx:
      rng = self._rngs[e]
      qpos[e, arm_adr] = rng.uniform(lower, upper)
      qpos[e, finger] = qpos[e, thumb]
      tx, tz = rng.uniform(-.4, .4), rng.uniform(.1, .4)
      if insert:                 # manipulator.py:213-218: the receptacle moves with the target
        ta = rng.uniform(-np.pi / 3, np.pi / 3)
        bpos[e, receptacle, 0], bpos[e, receptacle, 2] = tx, tz
        bquat[e, receptacle, 0], bquat[e, receptacle, 2] = np.cos(ta / 2), np.sin(ta / 2)
      else:
        ta = rng.uniform(-np.pi, np.pi)
      bpos[e, target, 0], bpos[e, target, 2] = tx, tz
      bquat[e, target, 0], bquat[e, target, 2] = np.cos(ta / 2), np.sin(ta / 2)
      kind[e] = rng.choice(["in_hand", "in_target", "uniform"], p=[.1, .1, .8])
      pose[e] = (tx, tz, ta)
    physics.set_env_model_field("body_pos", bpos)
    physics.set_env_model_field("body_quat", bquat)
    if any(k == "in_hand" for k in kind.values()):
      physics.data.qpos = qpos
      physics.after_reset()
      sx = np.asarray(physics.data.site_xpos).reshape(n, -1, 3)
      sm = np.asarray(physics.data.site_xmat).reshape(n, -1, 9)
    for e in idx:
      rng = self._rngs[e]
      if kind[e] == "in_target":
        ox, oz, oa = pose[e]
      elif kind[e] == "in_hand":
        ox, oz = sx[e, grasp, 0], sx[e, grasp, 2]
        oa = np.pi - np.arctan2(sm[e, grasp, 6], sm[e, grasp, 0])   # ('zx', 'xx')
      else:
        ox, oz = rng.uniform(-.5, .5), rng.uniform(0, .7)
        oa = rng.uniform(0, 2 * np.pi)
        qvel[e, obj_xvel] = rng.uniform(-5, 5)
      qpos[e, obj] = ox, oz, oa
    physics.data.qpos = qpos
    physics.data.qvel = qvel
    physics.after_reset()
    todo &= physics.data.ncon > 0


_ZOOM_OPERATORS = {}


def _zoom_operator(src, dst):
  """scipy.ndimage.zoom (cubic spline) of a src x src image to dst x dst is linear and
  separable: zoom(b) = W b W^T with W = zoom(identity, (dst / src, 1)).  The matrix form agrees
  with ndimage.zoom to 2e-15 (rounding order) and turns 8192 spline evaluations per reset
  (5 ms each: VERDICT r1 "Host side") into one batched matrix product."""
  key = (src, dst)
  if key not in _ZOOM_OPERATORS:
    from scipy import ndimage
    _ZOOM_OPERATORS[key] = ndimage.zoom(np.eye(src), (dst / float(src), 1))
  return _ZOOM_OPERATORS[key]


def _escape_terrains(m, rngs):
  """suite/quadruped.py:366-380 for a list of RNG streams: sinusoidal bowl times smooth random
  bumps, one terrain per stream (each stream draws what the reference draws: bump_res^2
  uniforms), [len(rngs), res * res] float32 like mjModel.hfield_data."""
  res = int(m.hfield_nrow[0])
  assert res == int(m.hfield_ncol[0])
  row_grid, col_grid = np.ogrid[-1:1:res * 1j, -1:1:res * 1j]
  radius = np.clip(np.sqrt(col_grid ** 2 + row_grid ** 2), .04, 1)
  bowl_shape = .5 - np.cos(2 * np.pi * radius) / 2
  terrain_size = 2 * m.hfield_size[0, 0]
  bump_res = int(terrain_size / 2)                    # _TERRAIN_BUMP_SCALE = 2
  bumps = np.stack([rng.uniform(.15, 1, (bump_res, bump_res))   # _TERRAIN_SMOOTHNESS = .15
                    for rng in rngs])
  w = _zoom_operator(bump_res, res)
  smooth_bumps = np.matmul(w, np.matmul(bumps, w.T))
  return (bowl_shape * smooth_bumps).reshape(len(rngs), -1).astype(np.float32)


def _escape_terrain(m, rng):
  """One terrain (float64 as the reference computes it before the float32 store)."""
  return _escape_terrains(m, [rng])[0].astype(np.float64)


def _initialize_quadruped(self, physics, task_name):
  """suite/quadruped.py:324-334 (Move) and 360-397 (Escape) for every env at once; the
  height search of `_find_non_contacting_height` (quadruped.py:245-277) raises all envs
  that still touch something by 1 cm per pass."""
  m, n = physics.model, physics.nenv
  if task_name == "escape":
    # a new terrain for every env being reset, in chunks (512 terrains = 83 MB of float32)
    envs = [e for e, _ in self._each()]
    for k in range(0, len(envs), 512):
      chunk = envs[k:k + 512]
      terrains = _escape_terrains(m, [self._rngs[e] for e in chunk])
      if chunk == list(range(chunk[0], chunk[0] + len(chunk))):
        physics.set_hfield(terrains, env0=chunk[0])
      else:
        for row, e in zip(terrains, chunk):
          physics.set_hfield(row[None], env0=e)
  xy = np.zeros((n, 2))
  if task_name == "fetch":              # quadruped.py:431-436: random azimuth and position
    spawn_radius = 0.9 * m.geom_size[m.name2id("floor", "geom"), 0]
    orientation = np.zeros((n, 4))
    for e, rng in self._each():
      azimuth = rng.uniform(0, 2 * np.pi)
      orientation[e] = (np.cos(azimuth / 2), 0, 0, np.sin(azimuth / 2))
      xy[e] = rng.uniform(-spawn_radius, spawn_radius, size=(2,))
  else:
    orientation = np.tile([1.0, 0, 0, 0], (n, 1))
    for e, rng in self._each():
      orientation[e] = rng.randn(4)
    orientation /= np.linalg.norm(orientation, axis=1, keepdims=True)
  root = m.jnt_qposadr[m.name2id("root", "joint")]
  qpos = np.tile(np.asarray(m.qpos0, dtype=np.float64), (n, 1))
  if self._mask is not None:            # the other envs keep their state
    cur = physics.data.qpos
    qpos[~self._mask] = cur[~self._mask]
  z = np.zeros(n)                       # start embedded in the floor
  todo = self._todo(n)
  for _ in range(10001):
    qpos[:, root:root + 3] = np.stack([xy[:, 0], xy[:, 1], z], axis=1)
    qpos[:, root + 3:root + 7] = orientation
    physics.data.qpos = qpos
    physics.data.qvel = np.zeros((n, m.nv))
    try:
      physics.after_reset()
    except PhysicsError:
      pass                              # a full contact buffer: keep raising (quadruped.py:266-270)
    todo &= physics.data.ncon > 0
    if not todo.any():
      if task_name == "fetch":          # quadruped.py:438-443: initial ball state
        ball = m.jnt_qposadr[m.name2id("ball_root", "joint")]
        ballv = m.jnt_dofadr[m.name2id("ball_root", "joint")]
        qvel = np.zeros((n, m.nv))
        for e, rng in self._each():
          qpos[e, ball:ball + 2] = rng.uniform(-spawn_radius, spawn_radius, size=(2,))
          qpos[e, ball + 2] = 2
          qvel[e, ballv:ballv + 2] = 5 * rng.randn(2)
        physics.data.qpos = qpos
        physics.data.qvel = qvel
      return
    z[todo] += 0.01
  raise RuntimeError("Failed to find a non-contacting configuration.")


Task._initialize_quadruped = _initialize_quadruped


def _load_model(name):
  return M.FlatModel.load(os.path.join(_MODELS, name + ".npz"))


def build_environment(domain_name, task_name, task_kwargs=None, environment_kwargs=None,
                      nenv=1, device=0, lanes=None, njmax=None, nconmax=None):
  """suite/__init__.py:117-150."""
  key = (domain_name, task_name)
  if domain_name not in {d for d, _ in T.REGISTRY}:
    raise ValueError("Domain {!r} does not exist.".format(domain_name))
  if key not in T.REGISTRY:
    raise ValueError("Level {!r} does not exist in domain {!r}.".format(task_name, domain_name))
  tdef = T.REGISTRY[key]
  task_kwargs = dict(task_kwargs or {})
  environment_kwargs = dict(environment_kwargs or {})
  seed = task_kwargs.get("random", None)
  if domain_name == "lqr":
    # lqr.py:93-101: the model itself is random (spring stiffness), drawn from the task's
    # RandomState before the task uses it; one model serves every env of the batch
    from dm_control_b200 import mjcf_compiler
    from dm_control_b200.suite import lqr
    # a sharded job passes its job-level seed as `model_seed` so that every rank and every
    # world size generates the SAME model (sharding.load_shard)
    model_seed = task_kwargs.pop("model_seed", None)
    if model_seed is not None:
      model_rng = np.random.RandomState(model_seed)
    elif isinstance(seed, np.random.RandomState):
      model_rng = seed
    else:
      model_rng = np.random.RandomState(seed)
    m = mjcf_compiler.compile_xml(lqr.make_model(tdef.model[0], tdef.model[1], model_rng))
  else:
    m = _load_model(tdef.model)
  if njmax is None and nconmax is None and tdef.model in T.STATIC_SIZES:
    njmax, nconmax = T.STATIC_SIZES[tdef.model]
    if lanes is None:
      lanes = T.STATIC_LANES.get(tdef.model)
  physics = BatchedPhysics(m, nenv, device=device, lanes=lanes, njmax=njmax, nconmax=nconmax)
  task = Task(tdef, random=task_kwargs.pop("random", None))
  task_kwargs.pop("model_seed", None)
  time_limit = task_kwargs.pop("time_limit", tdef.time_limit)
  if task_kwargs:
    raise TypeError("unexpected task_kwargs: {}".format(sorted(task_kwargs)))
  return control.Environment(physics, task, time_limit=time_limit,
                             control_timestep=tdef.control_timestep, **environment_kwargs)


def load(domain_name, task_name, task_kwargs=None, environment_kwargs=None,
         visualize_reward=False, **batch_kwargs):
  """suite/__init__.py:93-114.  `visualize_reward` needs a renderer: unavailable."""
  del visualize_reward
  return build_environment(domain_name, task_name, task_kwargs, environment_kwargs,
                           **batch_kwargs)
