"""Registry of the suite tasks served by the device backend.

Each entry restates the constants of the reference's task module
(/root/reference/dm_control/suite/<domain>.py; lines cited per entry) and resolves
the object ids its observation/reward kernel (csrc/bp_task.cu) needs by name, the
way the reference resolves `named.data.*` keys (mujoco/index.py:436-510).
"""

import collections

import numpy as np

# task kinds of csrc/bp_task.cu
NONE, CARTPOLE, ACROBOT, PENDULUM, CHEETAH, WALKER, HUMANOID, HOPPER = range(8)
FINGER = 9
QUADRUPED, QUADRUPED_ESCAPE = 10, 11
LQR = 12
REACHER, POINT_MASS = 8, 13
FINGER_TURN = 14
SWIMMER = 15
FISH_UPRIGHT, FISH_SWIM = 16, 17
BALL_IN_CUP = 18
MANIPULATOR = 19
QUADRUPED_FETCH = 20
STACKER = 21

TaskDef = collections.namedtuple(
    "TaskDef", "domain task model kind time_limit control_timestep params obs")


def _sens_adr(m, name):
  return int(m.sensor_adr[m.name2id(name, "sensor")])


def _cartpole(sparse):
  def f(m):
    return [int(m.jnt_qposadr[m.name2id("slider", "joint")]), int(sparse)], []
  return f


def _acrobot(sparse):
  def f(m):
    tgt = m.name2id("target", "site")
    return ([m.name2id("upper_arm", "body"), m.name2id("lower_arm", "body"), tgt,
             m.name2id("tip", "site")],
            [float(m.site_size[tgt, 0]), 0.0 if sparse else 1.0])
  return f


def _pendulum(m):
  # suite/pendulum.py:31-33 _ANGLE_BOUND = 8 degrees
  return ([m.name2id("pole", "body"), int(m.jnt_dofadr[m.name2id("hinge", "joint")])],
          [float(np.cos(np.deg2rad(8)))])


def _cheetah(m):
  return [_sens_adr(m, "torso_subtreelinvel")], [10.0]      # cheetah.py:31 _RUN_SPEED


def _walker(speed):
  def f(m):
    return [m.name2id("torso", "body"), _sens_adr(m, "torso_subtreelinvel")], [speed]
  return f


def _humanoid(speed, cmu=False):
  def f(m):
    names = ("thorax", "head", "lhand", "lfoot", "rhand", "rfoot") if cmu else \
        ("torso", "head", "left_hand", "left_foot", "right_hand", "right_foot")
    ids = [m.name2id(n, "body") for n in names]
    # "upright": xmat[torso, 'zz'] (humanoid.py:96-98) / xmat[thorax, 'zy'] (humanoid_CMU.py:68-70)
    return ids + [_sens_adr(m, "thorax_subtreelinvel" if cmu else "torso_subtreelinvel"),
                  7 if cmu else 8], [speed]
  return f


def _hopper(hopping):
  def f(m):
    return ([m.name2id("torso", "body"), m.name2id("foot", "body"),
             _sens_adr(m, "torso_subtreelinvel"), _sens_adr(m, "touch_toe"),
             _sens_adr(m, "touch_heel")], [1.0 if hopping else 0.0])
  return f


def _finger(m):
  names = ("proximal", "distal", "proximal_velocity", "distal_velocity", "hinge_velocity",
           "tip", "spinner", "touchtop", "touchbottom")
  return [_sens_adr(m, n) for n in names], [15.0]           # finger.py:38 _SPIN_VELOCITY


def _finger_turn(target_radius):
  def f(m):
    ip, _ = _finger(m)
    return ip + [_sens_adr(m, "target")], [15.0, target_radius]
  return f


def _quadruped(speed=0.0, deviation_angle=0.0, escape=False, fetch=False):
  """quadruped.py:291-306 _upright_reward, 343-357 Move.get_reward, 406-418 Escape."""
  def f(m):
    from dm_control_b200 import model as M
    types = np.asarray(m.sensor_type)
    ft = np.flatnonzero((types == M.SENS["force"]) | (types == M.SENS["torque"]))
    adr = np.asarray(m.sensor_adr)
    # force_torque() reads these sensors in model order; the kernel reads one window
    assert np.array_equal(adr[ft], adr[ft[0]] + 3 * np.arange(len(ft)))
    ip = [m.name2id("torso", "body"), _sens_adr(m, "velocimeter"), _sens_adr(m, "imu_accel"),
          _sens_adr(m, "imu_gyro"), int(adr[ft[0]]), m.name2id("workspace", "site"), 0]
    fp = [speed, float(np.cos(np.deg2rad(deviation_angle))), 0.0]
    if escape:
      rf = np.flatnonzero(types == M.SENS["rangefinder"])
      assert np.array_equal(adr[rf], adr[rf[0]] + np.arange(len(rf)))
      ip[6] = int(adr[rf[0]])
      fp[2] = float(m.hfield_size[0, 0])          # quadruped.py:408 terrain_size
    if fetch:                                     # quadruped.py:455-479
      ball = m.name2id("ball", "geom")
      ip += [m.name2id("ball", "body"), m.name2id("target", "site"),
             int(m.jnt_dofadr[m.name2id("ball_root", "joint")]),
             int(m.jnt_dofadr[m.name2id("root", "joint")])]
      fp += [float(m.site_size[m.name2id("workspace", "site"), 0] + m.geom_size[ball, 0]),
             float(m.geom_size[m.name2id("floor", "geom"), 0] * np.sqrt(2)),
             float(m.site_size[m.name2id("target", "site"), 0])]
    return ip, fp
  return f


def _obs_quadruped(escape, fetch=False):
  def f(m):
    from dm_control_b200 import model as M
    nh = int(np.sum(np.asarray(m.jnt_type) == M.JNT_HINGE))
    types = np.asarray(m.sensor_type)
    nft = int(np.sum((types == M.SENS["force"]) | (types == M.SENS["torque"])))
    obs = [("egocentric_state", 2 * nh + m.na), ("torso_velocity", 3), ("torso_upright", 1),
           ("imu", 6), ("force_torque", 3 * nft)]
    if escape:
      obs += [("origin", 3), ("rangefinder", int(np.sum(types == M.SENS["rangefinder"])))]
    if fetch:
      obs += [("ball_state", 9), ("target_position", 3)]
    return obs
  return f


def _reacher(target_size):
  def f(m):
    # reacher.py:112-114: radii = the task's target size + the finger's
    finger = m.name2id("finger", "geom")
    return ([m.name2id("target", "geom"), finger],
            [float(target_size + m.geom_size[finger, 0])])
  return f


def _point_mass(m):
  tgt = m.name2id("target", "geom")
  return [tgt, m.name2id("pointmass", "geom")], [float(m.geom_size[tgt, 0])]


def _swimmer(m):
  tgt = m.name2id("target", "geom")
  # body_velocities() reads sensordata[12:] as (velocimeter, gyro) pairs (swimmer.py:151-155)
  assert _sens_adr(m, "head_vel") == 12 and m.nsensordata == 12 + 6 * (m.nbody - 1)
  return ([tgt, m.name2id("nose", "geom"), m.name2id("head", "body"), 12],
          [float(m.geom_size[tgt, 0])])


def _obs_swimmer(m):
  return [("joints", m.nq - 3), ("to_target", 2), ("body_velocities", 3 * (m.nbody - 1))]


def _ball_in_cup(m):
  # ball_in_cup.py:57-62: |target - ball| < target size - ball size, on x and z
  tgt = m.name2id("target", "site")
  ball = float(m.geom_size[m.name2id("ball", "geom"), 0])
  return ([tgt, m.name2id("ball", "body")],
          [float(m.site_size[tgt, 0]) - ball, float(m.site_size[tgt, 2]) - ball])


def _obs_ball_in_cup(m):
  return [("position", m.nq), ("velocity", m.nv)]


MANIPULATOR_ARM_JOINTS = ["arm_root", "arm_shoulder", "arm_elbow", "arm_wrist",
                          "finger", "fingertip", "thumb", "thumbtip"]   # manipulator.py:36-37


def _manipulator(use_peg=False):
  """manipulator.py:70-130: bring_ball / bring_peg / insert_ball / insert_peg share the
  observation; object = ball | peg, target = target_ball | target_peg."""
  def f(m):
    obj_name = "peg" if use_peg else "ball"
    arm = [int(m.jnt_qposadr[m.name2id(j, "joint")]) for j in MANIPULATOR_ARM_JOINTS]
    assert arm == [int(m.jnt_dofadr[m.name2id(j, "joint")]) for j in MANIPULATOR_ARM_JOINTS]
    touch = [_sens_adr(m, s) for s in ("palm_touch", "finger_touch", "thumb_touch",
                                       "fingertip_touch", "thumbtip_touch")]
    assert touch == list(range(touch[0], touch[0] + 5))
    obj = [int(m.jnt_dofadr[m.name2id(obj_name + "_" + k, "joint")]) for k in "xzy"]
    assert obj == list(range(obj[0], obj[0] + 3))
    ip = arm + [touch[0], m.name2id("hand", "body"), m.name2id(obj_name, "body"),
                m.name2id("target_" + obj_name, "body"), obj[0], m.name2id(obj_name, "site"),
                m.name2id("target_" + obj_name, "site"), int(use_peg)]
    if use_peg:          # _peg_reward (manipulator.py:276-286)
      ip += [m.name2id(k, "site") for k in ("peg_grasp", "grasp", "peg_pinch", "pinch",
                                            "target_peg_tip", "peg_tip")]
    return ip, [.01]
  return f


def _stacker(n_boxes):
  """stacker.py:187-214."""
  def f(m):
    arm = [int(m.jnt_qposadr[m.name2id(j, "joint")]) for j in MANIPULATOR_ARM_JOINTS]
    touch = [_sens_adr(m, s) for s in ("palm_touch", "finger_touch", "thumb_touch",
                                       "fingertip_touch", "thumbtip_touch")]
    assert touch == list(range(touch[0], touch[0] + 5))
    boxes = [m.name2id("box%d" % b, "body") for b in range(n_boxes)]
    assert boxes == list(range(boxes[0], boxes[0] + n_boxes))
    joints = [int(m.jnt_dofadr[m.name2id("box%d_%s" % (b, k), "joint")])
              for b in range(n_boxes) for k in "xzy"]         # stacker.py:133-136 order
    assert joints == list(range(joints[0], joints[0] + 3 * n_boxes))
    sites = [m.name2id("box%d" % b, "site") for b in range(n_boxes)]
    assert sites == list(range(sites[0], sites[0] + n_boxes))
    return (arm + [touch[0], m.name2id("hand", "body"), n_boxes, boxes[0], joints[0],
                   m.name2id("target", "body"), m.name2id("grasp", "site"),
                   m.name2id("target", "site"), sites[0]],
            [float(m.geom_size[m.name2id("target", "geom"), 0]), .01])
  return f


def _obs_stacker(n_boxes):
  def f(m):
    return [("arm_pos", 16), ("arm_vel", 8), ("touch", 5), ("hand_pos", 4),
            ("box_pos", 4 * n_boxes), ("box_vel", 3 * n_boxes), ("target_pos", 2)]
  return f


def _obs_manipulator(m):
  return [("arm_pos", 16), ("arm_vel", 8), ("touch", 5), ("hand_pos", 4), ("object_pos", 4),
          ("object_vel", 3), ("target_pos", 4)]


def _fish(m):
  # fish.py:30-36 _JOINTS: the seven tail / fin hinges follow the free root joint
  joints = ["tail1", "tail_twist", "tail2", "finright_roll", "finright_pitch",
            "finleft_roll", "finleft_pitch"]
  adr = [int(m.jnt_qposadr[m.name2id(j, "joint")]) for j in joints]
  assert adr == list(range(7, 14)) and m.nq == 14
  tgt, mouth = m.name2id("target", "geom"), m.name2id("mouth", "geom")
  return ([m.name2id("torso", "body"), adr[0], len(adr), tgt, mouth],
          [float(m.geom_size[mouth, 0] + m.geom_size[tgt, 0])])


def _obs_fish(swim):
  def f(m):
    obs = [("joint_angles", 7), ("upright", 1)]
    if swim:
      obs.append(("target", 3))
    return obs + [("velocity", m.nv)]
  return f


def _obs_hopper(m):
  return [("position", m.nq - 1), ("velocity", m.nv), ("touch", 2)]


def _obs_cartpole(m):
  return [("position", 1 + 2 * (m.nbody - 2)), ("velocity", m.nv)]


_TASKS = [
    # cartpole.py:30 time limit 10 s, no control_timestep (62-99)
    TaskDef("cartpole", "swingup", "cartpole", CARTPOLE, 10, None, _cartpole(False), _obs_cartpole),
    TaskDef("cartpole", "swingup_sparse", "cartpole", CARTPOLE, 10, None, _cartpole(True), _obs_cartpole),
    TaskDef("cartpole", "balance", "cartpole", CARTPOLE, 10, None, _cartpole(False), _obs_cartpole),
    TaskDef("cartpole", "balance_sparse", "cartpole", CARTPOLE, 10, None, _cartpole(True), _obs_cartpole),
    # cartpole.py:84-103: swing-up with a chain of poles (_make_model, 105-127)
    TaskDef("cartpole", "two_poles", "cartpole2", CARTPOLE, 10, None, _cartpole(False), _obs_cartpole),
    TaskDef("cartpole", "three_poles", "cartpole3", CARTPOLE, 10, None, _cartpole(False), _obs_cartpole),
    # acrobot.py:28 time limit 10 s
    TaskDef("acrobot", "swingup", "acrobot", ACROBOT, 10, None, _acrobot(False),
            lambda m: [("orientations", 4), ("velocity", m.nv)]),
    TaskDef("acrobot", "swingup_sparse", "acrobot", ACROBOT, 10, None, _acrobot(True),
            lambda m: [("orientations", 4), ("velocity", m.nv)]),
    # pendulum.py:29 time limit 20 s
    TaskDef("pendulum", "swingup", "pendulum", PENDULUM, 20, None, _pendulum,
            lambda m: [("orientation", 2), ("velocity", 1)]),
    # cheetah.py:28 time limit 10 s
    TaskDef("cheetah", "run", "cheetah", CHEETAH, 10, None, _cheetah,
            lambda m: [("position", m.nq - 1), ("velocity", m.nv)]),
    # walker.py:29-36: 25 s, control timestep .025, speeds 0 / 1 / 8
    TaskDef("walker", "stand", "walker", WALKER, 25, .025, _walker(0.0),
            lambda m: [("orientations", 2 * (m.nbody - 1)), ("height", 1), ("velocity", m.nv)]),
    TaskDef("walker", "walk", "walker", WALKER, 25, .025, _walker(1.0),
            lambda m: [("orientations", 2 * (m.nbody - 1)), ("height", 1), ("velocity", m.nv)]),
    TaskDef("walker", "run", "walker", WALKER, 25, .025, _walker(8.0),
            lambda m: [("orientations", 2 * (m.nbody - 1)), ("height", 1), ("velocity", m.nv)]),
    # hopper.py:32-41: 20 s, control timestep .02, stand height .6, hop speed 2
    TaskDef("hopper", "stand", "hopper", HOPPER, 20, .02, _hopper(False), _obs_hopper),
    TaskDef("hopper", "hop", "hopper", HOPPER, 20, .02, _hopper(True), _obs_hopper),
    # finger.py:28-33: 20 s, control timestep .02; turn targets of radius .07 / .03 sit on
    # a per-env site (bp_env_model_write "site_pos")
    TaskDef("finger", "spin", "finger", FINGER, 20, .02, _finger,
            lambda m: [("position", 4), ("velocity", 3), ("touch", 2)]),
    TaskDef("finger", "turn_easy", "finger", FINGER_TURN, 20, .02, _finger_turn(.07),
            lambda m: [("position", 4), ("velocity", 3), ("touch", 2), ("target_position", 2),
                       ("dist_to_target", 1)]),
    TaskDef("finger", "turn_hard", "finger", FINGER_TURN, 20, .02, _finger_turn(.03),
            lambda m: [("position", 4), ("velocity", 3), ("touch", 2), ("target_position", 2),
                       ("dist_to_target", 1)]),
    # quadruped.py:36-44: 20 s, control timestep .02, walk speed 0.5, run speed 5; escape
    # uses the terrain model and a 20 degree upright deviation (fetch: walls, ball and
    # condim-6 contacts are not served)
    TaskDef("quadruped", "walk", "quadruped", QUADRUPED, 20, .02, _quadruped(0.5),
            _obs_quadruped(False)),
    TaskDef("quadruped", "run", "quadruped", QUADRUPED, 20, .02, _quadruped(5.0),
            _obs_quadruped(False)),
    TaskDef("quadruped", "escape", "quadruped_escape", QUADRUPED_ESCAPE, 20, .02,
            _quadruped(0.0, 20.0, True), _obs_quadruped(True)),
    # quadruped.py:133-142 fetch: walls, a ball with a free joint (condim 6, priority 1,
    # direct solref) and a target site
    TaskDef("quadruped", "fetch", "quadruped_fetch", QUADRUPED_FETCH, 20, .02,
            _quadruped(fetch=True), _obs_quadruped(False, True)),
    # reacher.py:29-57: 20 s, target size .05 (easy) / .015 (hard); the target geom moves
    # per episode and per env (bp_env_model_write)
    TaskDef("reacher", "easy", "reacher", REACHER, 20, None, _reacher(.05),
            lambda m: [("position", m.nq), ("to_target", 2), ("velocity", m.nv)]),
    TaskDef("reacher", "hard", "reacher", REACHER, 20, None, _reacher(.015),
            lambda m: [("position", m.nq), ("to_target", 2), ("velocity", m.nv)]),
    # point_mass.py:29-43: 20 s; `hard` rewires the tendon coefficients per episode and per
    # env (model.wrap_prm through bp_env_model_write)
    TaskDef("point_mass", "easy", "point_mass", POINT_MASS, 20, None, _point_mass,
            lambda m: [("position", m.nq), ("velocity", m.nv)]),
    TaskDef("point_mass", "hard", "point_mass", POINT_MASS, 20, None, _point_mass,
            lambda m: [("position", m.nq), ("velocity", m.nv)]),
    # swimmer.py:28-30: 30 s, control timestep .03; fluid drag (option density), the target
    # geom moves per episode and per env
    # ball_in_cup.py:29-30: 20 s, control timestep .02; the string is a spatial tendon with
    # a length limit
    TaskDef("ball_in_cup", "catch", "ball_in_cup", BALL_IN_CUP, 20, .02, _ball_in_cup,
            _obs_ball_in_cup),
    # manipulator.py:32-33,70-82: 10 s, control timestep .01; fully observable bring_ball
    # (the model keeps the ball and its target: make_model, 47-67)
    TaskDef("manipulator", "bring_ball", "manipulator_ball", MANIPULATOR, 10, .01,
            _manipulator(False), _obs_manipulator),
    # manipulator.py:85-130: the peg prop and the receptacles (cup / slot: box geoms against
    # the ball / peg capsules and the hand)
    TaskDef("manipulator", "bring_peg", "manipulator_peg", MANIPULATOR, 10, .01,
            _manipulator(True), _obs_manipulator),
    TaskDef("manipulator", "insert_ball", "manipulator_insert_ball", MANIPULATOR, 10, .01,
            _manipulator(False), _obs_manipulator),
    TaskDef("manipulator", "insert_peg", "manipulator_insert_peg", MANIPULATOR, 10, .01,
            _manipulator(True), _obs_manipulator),
    # stacker.py:31-32,52-80: 10 s, control timestep .01; boxes stacked on a target
    TaskDef("stacker", "stack_2", "stacker_2", STACKER, 10, .01, _stacker(2), _obs_stacker(2)),
    TaskDef("stacker", "stack_4", "stacker_4", STACKER, 10, .01, _stacker(4), _obs_stacker(4)),
    # fish.py:28-29: 40 s, control timestep .04; gravity and constraints disabled, fluid
    # density 5000, a tendon spring between the fins, position servos
    TaskDef("fish", "upright", "fish", FISH_UPRIGHT, 40, .04, _fish, _obs_fish(False)),
    TaskDef("fish", "swim", "fish", FISH_SWIM, 40, .04, _fish, _obs_fish(True)),
    TaskDef("swimmer", "swimmer6", "swimmer6", SWIMMER, 30, .03, _swimmer, _obs_swimmer),
    TaskDef("swimmer", "swimmer15", "swimmer15", SWIMMER, 30, .03, _swimmer, _obs_swimmer),
    # lqr.py:31-60: no time limit, control cost 0.1; "model" = (n_bodies, n_actuators): the
    # model is generated from the task's seed when the environment is built (suite/lqr.py)
    TaskDef("lqr", "lqr_2_1", (2, 1), LQR, float("inf"), None, lambda m: ([], [0.1]),
            lambda m: [("position", m.nq), ("velocity", m.nv)]),
    TaskDef("lqr", "lqr_6_2", (6, 2), LQR, float("inf"), None, lambda m: ([], [0.1]),
            lambda m: [("position", m.nq), ("velocity", m.nv)]),
    # humanoid.py:28-35: 25 s, control timestep .025, speeds 0 / 1 / 10
    TaskDef("humanoid", "stand", "humanoid", HUMANOID, 25, .025, _humanoid(0.0), None),
    TaskDef("humanoid", "walk", "humanoid", HUMANOID, 25, .025, _humanoid(1.0), None),
    TaskDef("humanoid", "run", "humanoid", HUMANOID, 25, .025, _humanoid(10.0), None),
    # humanoid_CMU.py:28-36: 20 s, control timestep .02, speeds 0 / 10 (nv = 62, 56 motors)
    TaskDef("humanoid_CMU", "stand", "humanoid_CMU", HUMANOID, 20, .02, _humanoid(0.0, True), None),
    TaskDef("humanoid_CMU", "run", "humanoid_CMU", HUMANOID, 20, .02, _humanoid(10.0, True), None),
]


def _obs_humanoid(m):
  return [("joint_angles", m.nq - 7), ("head_height", 1), ("extremities", 12),
          ("torso_vertical", 3), ("com_velocity", 3), ("velocity", m.nv)]


REGISTRY = {}
for _t in _TASKS:
  if _t.obs is None:
    _t = _t._replace(obs=_obs_humanoid)
  REGISTRY[(_t.domain, _t.task)] = _t

# (domain, task) pairs in the order of the reference's suite.BENCHMARKING subset we serve
ALL_TASKS = tuple(sorted(REGISTRY))


# Batch sizes (njmax, nconmax) of the models that have a static (ahead-of-time specialised)
# kernel instance (csrc/static/*.cuh, tools/gen_static_model.py TRAITS): suite.load uses them
# by default so that its batches are served by those instances.
STATIC_SIZES = {"cheetah": (64, 16), "walker": (48, 12), "humanoid": (120, 40),
                "quadruped_escape": (184, 42), "finger": (24, 6),
                "cartpole": (4, 1), "acrobot": (1, 1), "pendulum": (1, 1)}
# lanes per env of the instances that differ from physics.default_lanes: the contact-free
# models run one thread per env
STATIC_LANES = {"cartpole": 1, "acrobot": 1, "pendulum": 1}


def n_sub_steps(tdef, m):
  """rl/control.py:164-190 compute_n_steps."""
  if tdef.control_timestep is None:
    return 1
  ratio = tdef.control_timestep / m.opt_timestep
  if tdef.control_timestep < m.opt_timestep or abs(ratio - round(ratio)) > 1e-8:
    raise ValueError("Control timestep ({}) must be an integer multiple of physics "
                     "timestep ({})".format(tdef.control_timestep, m.opt_timestep))
  return int(round(ratio))
