"""Generates the committed fixtures under tests/golden/ (run in the build container).

1. rewards_tolerance.npz — outputs of the REFERENCE's own pure-numpy
   `dm_control/utils/rewards.py` (imported by file path from /root/reference; it is
   importable without MuJoCo) for the sigmoids / bounds / margins the served tasks use.
2. traj_<domain>.npz — seeded trajectories of the CPU oracle (the reference's physics
   engine is a closed binary that is absent): states at steps
   0, 1, 10, 100 for 8 envs per suite model, with contact / constraint counts.
"""
import importlib.util
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
OUT = os.path.join(ROOT, "tests", "golden")

CASES = [  # (lower, upper, margin, sigmoid, value_at_margin)
    (0.0, 0.0, 2.0, "gaussian", 0.1), (0.0, 0.0, 1.0, "quadratic", 0.0),
    (0.0, 0.0, 5.0, "gaussian", 0.1), (10.0, np.inf, 10.0, "linear", 0.0),
    (1.2, np.inf, 0.6, "gaussian", 0.1), (1.0, np.inf, 0.5, "linear", 0.5),
    (1.4, np.inf, 0.35, "gaussian", 0.1), (0.9, np.inf, 1.9, "linear", 0.0),
    (-0.25, 0.25, 0.0, "gaussian", 0.1), (0.995, 1.0, 0.0, "gaussian", 0.1),
    (0.0, 0.2, 1.0, "gaussian", 0.1), (0.0, 0.0, 2.0, "gaussian", 0.1),
    (0.0, 0.1, 0.5, "long_tail", 0.1), (-1.0, 1.0, 3.0, "long_tail", 0.3)]


def rewards_fixture():
  spec = importlib.util.spec_from_file_location(
      "ref_rewards", "/root/reference/dm_control/utils/rewards.py")
  mod = importlib.util.module_from_spec(spec)
  spec.loader.exec_module(mod)
  x = np.concatenate([np.linspace(-12, 12, 97), [0.0, 0.2, 0.995, 1.0, 1.2, 1.4, 10.0]])
  out = np.stack([mod.tolerance(x, bounds=(lo, hi), margin=mg, sigmoid=sg, value_at_margin=v1)
                  for lo, hi, mg, sg, v1 in CASES])
  np.savez(os.path.join(OUT, "rewards_tolerance.npz"), x=x, out=out,
           cases=np.array([(lo, hi, mg, {"gaussian": 0, "linear": 1, "quadratic": 2, "long_tail": 3}[sg], v1)
                           for lo, hi, mg, sg, v1 in CASES]))
  print("rewards_tolerance.npz", out.shape)


def traj_fixture(domain, nenv=8, steps=(0, 1, 10, 100), sizes=None):
  import parity_util as PU
  from oracle.oracle_py import Oracle, OracleModel
  m = PU.load_model(domain)
  qpos, qvel, ctrl = PU.random_states(m, nenv, seed=11)
  hfield, extra = None, {}
  if int(m.nhfielddata) > 0:
    # terrains are regenerated from their seeds (parity_util.escape_terrain), not stored
    seeds = 11 + 1000 + np.arange(nenv)
    hfield = np.stack([PU.escape_terrain(np.random.RandomState(int(sd))) for sd in seeds])
    PU.terrain_states(m, qpos, hfield, 11)
    extra["hfield_seed"] = seeds
  if sizes:
    m = m.copy()
    m.arrays["njmax"][:] = sizes[0]; m.arrays["nconmax"][:] = sizes[1]
    extra["sizes"] = np.array(sizes)
  om = OracleModel(m)
  rec = {k: [] for k in ("qpos", "qvel", "qacc", "ncon", "nefc", "solver_iter", "sensordata")}
  for e in range(nenv):
    o = Oracle(om)
    if hfield is not None:
      o.hfield_data[:] = hfield[e]
    o.reset()
    o.qpos[:] = qpos[e]; o.qvel[:] = qvel[e]; o.ctrl[:] = ctrl[e]
    o.forward()
    per = {k: [] for k in rec}
    for s in range(max(steps) + 1):
      if s > 0:
        o.step()
      if s in steps:
        for k in rec:
          per[k].append(np.array(getattr(o, k)))
    for k in rec:
      rec[k].append(np.stack(per[k]))
  np.savez_compressed(os.path.join(OUT, "traj_%s.npz" % domain), qpos0=qpos, qvel0=qvel,
                      ctrl=ctrl, steps=np.array(steps), **extra,
                      **{k: np.stack(v) for k, v in rec.items()})     # [env][step][n]
  print("traj_%s.npz" % domain, rec["qpos"][0].shape)


def c1_trace(steps=1000):
  """cartpole-swingup, 1 env, task seed 0, actions
  RandomState(0).uniform(-1, 1, (1000, 1)), through the CPU oracle + task oracle: per-step
  qpos, qvel, observation and reward.  The episode start uses the same host code as the
  product (suite.Task.initialize_episode) on a stand-in physics object."""
  import types
  import parity_util as PU
  from oracle import task_oracle
  from oracle.oracle_py import Oracle, OracleModel
  from dm_control_b200 import suite
  from dm_control_b200.suite import tasks as T
  m = PU.load_model("cartpole")
  task = suite.Task(T.REGISTRY[("cartpole", "swingup")], random=0)
  phys = types.SimpleNamespace(
      model=m, nenv=1, _batch=types.SimpleNamespace(nobs=5, set_task=lambda *a: None),
      data=types.SimpleNamespace(qpos=np.tile(m.qpos0, (1, 1)), qvel=np.zeros((1, m.nv))))
  task.bind(phys)
  task.initialize_episode(phys)
  o = Oracle(OracleModel(m))
  o.reset()
  o.qpos[:] = phys.data.qpos[0]; o.qvel[:] = phys.data.qvel[0]
  o.forward()
  actions = np.random.RandomState(0).uniform(-1, 1, size=(steps, 1))
  fields = ("qpos", "qvel", "ctrl", "xmat", "xpos", "site_xpos", "sensordata")
  rec = {k: [] for k in ("qpos", "qvel", "obs", "reward")}
  for k in range(steps):
    o.ctrl[:] = actions[k]
    o.step()
    d = {f: np.array(getattr(o, f), dtype=np.float64).reshape(1, -1) for f in fields}
    obs, r = task_oracle.obs_reward("cartpole", "swingup", m, d)
    rec["qpos"].append(np.array(o.qpos)); rec["qvel"].append(np.array(o.qvel))
    rec["obs"].append(obs[0]); rec["reward"].append(r[0])
  np.savez_compressed(os.path.join(OUT, "c1_cartpole_swingup.npz"), actions=actions,
                      qpos0=phys.data.qpos[0], qvel0=phys.data.qvel[0],
                      **{k: np.stack(v) for k, v in rec.items()})
  print("c1_cartpole_swingup.npz", np.stack(rec["obs"]).shape, "reward sum", np.sum(rec["reward"]))


def load_reference_randomizers():
  """Imports the REFERENCE's suite/utils/randomizers.py by file path.  It is pure numpy but
  for the `mjbindings` import (enum values of mjtJoint and mju_axisAngle2Quat), which is
  stubbed here with MuJoCo's documented enum order and the axis-angle formula."""
  import types
  mjb = types.ModuleType("dm_control.mujoco.wrapper.mjbindings")
  mjb.enums = types.SimpleNamespace(mjtJoint=types.SimpleNamespace(
      mjJNT_FREE=0, mjJNT_BALL=1, mjJNT_SLIDE=2, mjJNT_HINGE=3))

  def mju_axisAngle2Quat(res, axis, angle):
    res[0] = np.cos(angle / 2)
    res[1:4] = np.asarray(axis) * np.sin(angle / 2)
  mjb.mjlib = types.SimpleNamespace(mju_axisAngle2Quat=mju_axisAngle2Quat)
  saved = {k: sys.modules.get(k) for k in ("dm_control", "dm_control.mujoco",
                                           "dm_control.mujoco.wrapper",
                                           "dm_control.mujoco.wrapper.mjbindings")}
  pk = types.ModuleType("dm_control"); mj = types.ModuleType("dm_control.mujoco")
  wr = types.ModuleType("dm_control.mujoco.wrapper")
  wr.mjbindings = mjb; mj.wrapper = wr; pk.mujoco = mj
  sys.modules.update({"dm_control": pk, "dm_control.mujoco": mj,
                      "dm_control.mujoco.wrapper": wr,
                      "dm_control.mujoco.wrapper.mjbindings": mjb})
  try:
    spec = importlib.util.spec_from_file_location(
        "ref_randomizers", "/root/reference/dm_control/suite/utils/randomizers.py")
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
  finally:
    for k, v in saved.items():
      if v is None:
        sys.modules.pop(k, None)
      else:
        sys.modules[k] = v
  return mod


class FakePhysics:
  """The slice of `Physics` that randomizers.py touches: model.njnt / id2name / jnt_type /
  jnt_limited / jnt_range and named.data.qpos[joint_name] (read: a view, write: assign)."""

  def __init__(self, m):
    self.qpos = np.array(m.qpos0, dtype=np.float64)
    width = {0: 7, 1: 4, 2: 1, 3: 1}
    names = [n if n else "joint%d" % j for j, n in enumerate(m.names["joint"])]
    outer = self

    class _Qpos:
      def _slice(self, name):
        j = names.index(name)
        a = int(m.jnt_qposadr[j])
        return slice(a, a + width[int(m.jnt_type[j])])

      def __getitem__(self, name):
        return outer.qpos[self._slice(name)]

      def __setitem__(self, name, value):
        outer.qpos[self._slice(name)] = value

    self.named = type("N", (), {"data": type("D", (), {"qpos": _Qpos()})()})()
    self.model = type("Mdl", (), {
        "njnt": int(m.njnt), "jnt_type": np.asarray(m.jnt_type),
        "jnt_limited": np.asarray(m.jnt_limited), "jnt_range": np.asarray(m.jnt_range),
        "id2name": staticmethod(lambda i, kind: names[i])})()


RANDOMIZER_MODELS = ["cartpole", "acrobot", "pendulum", "cheetah", "walker", "hopper",
                     "humanoid", "humanoid_CMU", "finger", "reacher", "point_mass", "swimmer6",
                     "fish", "ball_in_cup", "quadruped", "manipulator_ball"]


def randomizers_fixture():
  """qpos after the REFERENCE's randomize_limited_and_rotational_joints for every served
  model and three seeds -> tests/golden/randomizers.npz (plus a model with a limited ball
  joint, which no suite model has)."""
  import parity_util as PU
  mod = load_reference_randomizers()
  out = {}
  for dom in RANDOMIZER_MODELS:
    m = PU.load_model(dom)
    for seed in (0, 1, 2):
      ph = FakePhysics(m)
      mod.randomize_limited_and_rotational_joints(ph, np.random.RandomState(seed))
      out["%s/%d" % (dom, seed)] = ph.qpos
  np.savez(os.path.join(OUT, "randomizers.npz"), **out)
  print("randomizers.npz", len(out), "cases")


def ref_test_humanoid_model():
  """The reference's own test model that asks for RK4 + PGS
  (/root/reference/dm_control/mujoco/testing/assets/humanoid.xml:9), compiled here because
  the GPU box has no reference tree."""
  from dm_control_b200 import mjcf_compiler as C
  m = C.compile_file("/root/reference/dm_control/mujoco/testing/assets/humanoid.xml")
  m.save(os.path.join(OUT, "model_ref_test_humanoid.npz"))
  print("model_ref_test_humanoid.npz solver", int(m.opt_solver), "integrator", int(m.opt_integrator))


def suite_tasks_fixture():
  """suite_tasks.json -- what the reference's own suite modules declare for every task, read
  from their source with `ast` (the modules cannot be imported: they load the MuJoCo
  binary): time limit and control timestep of each `@SUITE.add` function, and the
  observation keys every Task class assigns in `get_observation`, in source order."""
  import ast
  import json
  suite_dir = "/root/reference/dm_control/suite"

  def resolve(node, consts):
    if node is None:
      return None
    if isinstance(node, ast.Name):
      return consts.get(node.id, node.id)
    try:
      return ast.literal_eval(node)
    except Exception:
      return ast.unparse(node)

  def obs_keys(fn):
    keys = []
    for sub in ast.walk(fn):
      if (isinstance(sub, ast.Subscript) and isinstance(sub.value, ast.Name) and
          sub.value.id == "obs" and isinstance(sub.ctx, ast.Store)):
        try:
          keys.append((sub.lineno, ast.literal_eval(sub.slice)))
        except Exception:
          pass
    return [k for _, k in sorted(keys)]

  res = {}
  for f in sorted(os.listdir(suite_dir)):
    if not f.endswith(".py") or f.startswith("_") or f.endswith("_test.py"):
      continue
    tree = ast.parse(open(os.path.join(suite_dir, f)).read())
    consts = {}
    for n in tree.body:
      if isinstance(n, ast.Assign) and len(n.targets) == 1 and isinstance(n.targets[0], ast.Name):
        try:
          consts[n.targets[0].id] = ast.literal_eval(n.value)
        except Exception:
          if ast.unparse(n.value) == "float('inf')":
            consts[n.targets[0].id] = "inf"
    tasks, obs = {}, {}
    helpers = {n.name: obs_keys(n) for n in tree.body
               if isinstance(n, ast.FunctionDef) and obs_keys(n)}
    for n in tree.body:
      if isinstance(n, ast.FunctionDef) and any(
          isinstance(d, ast.Call) and getattr(d.func, "attr", "") == "add"
          for d in n.decorator_list):
        tl = ct = None
        for a, d in zip(n.args.args[::-1], n.args.defaults[::-1]):
          if a.arg == "time_limit":
            tl = resolve(d, consts)
        envs = [c for c in ast.walk(n)
                if isinstance(c, ast.Call) and getattr(c.func, "attr", "") == "Environment"]
        if not envs:      # built by a module-level helper (swimmer._make_swimmer)
          envs = [c for c in ast.walk(tree)
                  if isinstance(c, ast.Call) and getattr(c.func, "attr", "") == "Environment"]
        for c in envs:
          for k in c.keywords:
            if k.arg == "control_timestep":
              ct = resolve(k.value, consts)
        tasks[n.name] = {"time_limit": tl, "control_timestep": ct}
      if isinstance(n, ast.ClassDef):
        for m in n.body:
          if isinstance(m, ast.FunctionDef) and m.name == "get_observation":
            keys = obs_keys(m)
            # observations started by a module-level helper (quadruped._common_observations)
            for sub in ast.walk(m):
              if isinstance(sub, ast.Call) and isinstance(sub.func, ast.Name) and \
                  sub.func.id in helpers:
                keys = helpers[sub.func.id] + keys
            obs[n.name] = keys
    if tasks:
      res[f[:-3]] = {"tasks": tasks, "observations": obs}
  with open(os.path.join(OUT, "suite_tasks.json"), "w") as fh:
    json.dump(res, fh, indent=1, sort_keys=True)
  print("suite_tasks.json", sum(len(v["tasks"]) for v in res.values()), "tasks of", len(res), "modules")


if __name__ == "__main__":
  suite_tasks_fixture()
  randomizers_fixture()
  ref_test_humanoid_model()
  os.makedirs(OUT, exist_ok=True)
  rewards_fixture()
  for d in ("pendulum", "cartpole", "acrobot", "cheetah", "walker", "hopper", "humanoid"):
    traj_fixture(d)
  traj_fixture("finger", nenv=32)       # a tenth of the random states start in contact
  traj_fixture("quadruped", sizes=(120, 40))
  traj_fixture("quadruped_escape", sizes=(300, 64))
  c1_trace()
