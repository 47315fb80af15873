"""`BatchedPhysics`: dm_control's `mujoco.Physics` with a leading environment axis.

Mirrors the reference facade (/root/reference/dm_control/mujoco/engine.py:108-573)
method by method — step / forward / reset / after_reset / reset_context /
set_control / control / position / velocity / activation / state / get_state /
set_state / time / timestep / check_invalid_state / named — over a device batch
behind the C ABI (binding.py).  State arrays carry a leading env dimension.
There is no CPU path: constructing it without libbphys.so + a CUDA device raises.
"""

import contextlib

import numpy as np

from dm_control_b200 import binding
from dm_control_b200 import index
from dm_control_b200 import mjcf_compiler
from dm_control_b200 import model as M

_INVALID_PHYSICS_STATE = ("Physics state is invalid. Warning(s) raised: {warning_names} "
                          "in environment(s) {envs}")   # engine.py:58-60


class PhysicsError(RuntimeError):
  """Raised if the state of the physics simulation becomes divergent
  (rl/control.py:196-199)."""


# default execution shape per model size: lanes per environment (DESIGN.md §Shapes;
# chosen from the sweep in profiles/)
def default_lanes(m):
  """Execution shape from the lane sweeps (profiles/r1_tuning.md): 8 lanes per env for the
  planar models, 16 for humanoid-sized ones (both run the 72-register build); the finger
  (nv = 3) runs 4 lanes, which is also what its static instance is built for."""
  return 4 if m.nv <= 3 else 8 if m.nv <= 12 else 16


# model arrays that tasks edit per episode and that a batch therefore holds per env
ENV_MODEL_FIELDS = ("geom_pos", "site_pos", "body_pos", "body_quat", "wrap_prm")


def default_sizes(m):
  """(njmax, nconmax) that fit typical suite episodes; overflow raises the same
  CNSTRFULL / CONTACTFULL warnings the reference surfaces (core_test.py:97-103)."""
  nconmax = int(min(m.nconmax, max(8, 2 * m.ngeom)))
  njmax = int(min(m.njmax, 4 * nconmax + int(np.sum(m.jnt_limited != 0)) +
                  int(np.sum(m.dof_frictionloss > 0))))
  return max(njmax, 1), max(nconmax, 1)


class _Data:
  """`physics.data.<field>`: host copies [nenv, n] (reads) / whole-field writes."""

  def __init__(self, physics):
    object.__setattr__(self, "_p", physics)

  def __getattr__(self, name):
    b = self._p._batch
    if name in b.fields:
      a = b.read(name)
      return a[:, 0] if name in ("time", "ncon", "nefc", "ne", "nf", "solver_iter") else a
    raise AttributeError(name)

  _OBJ = {"body": 1, "xbody": 2, "geom": 5, "site": 6}

  def object_velocity(self, object_id, object_type, local_frame=False):
    """MjData.object_velocity (wrapper/core.py:713-745) for every env: [nenv, 2, 3] stacked
    (linear_velocity, angular_velocity); object_type 'body' | 'xbody' | 'geom' | 'site' (or
    the mjtObj value), object_id an id or a name."""
    p = self._p
    if not isinstance(object_type, (int, np.integer)):
      if object_type not in self._OBJ:
        raise binding.Error("{!r} is not a valid object type name.".format(object_type))
      kind = "body" if object_type == "xbody" else object_type
      object_type = self._OBJ[object_type]
    else:
      kind = {1: "body", 2: "body", 5: "geom", 6: "site"}.get(int(object_type), "body")
    if not isinstance(object_id, (int, np.integer)):
      object_id = p.model.name2id(object_id, kind)
    return p._batch.object_velocity(object_id, object_type, local_frame)

  def contact_force(self, contact_id):
    """MjData.contact_force (wrapper/core.py:747-769) for every env: [nenv, 2, 3] stacked
    (force, torque) in the contact's frame.  contact_id must be a valid index in EVERY env
    (0 <= contact_id < min ncon), else ValueError as in the reference."""
    p = self._p
    ncon = int(p._batch.read("ncon").min())
    if not 0 <= contact_id < ncon:
      raise ValueError("`contact_id` must be an integer in range [0, {max_valid}] "
                       "(inclusive), got: {actual}.".format(max_valid=ncon - 1,
                                                             actual=contact_id))
    return p._batch.contact_force(contact_id)

  def __setattr__(self, name, value):
    b = self._p._batch
    if name not in b.fields:
      raise AttributeError(name)
    cnt = b.fields[name][1]
    a = np.asarray(value, dtype=b.fields[name][2])
    if a.ndim == 0 or a.shape == (cnt,):
      a = np.broadcast_to(a.reshape(-1) if a.ndim else a, (b.nenv, cnt))
    a = np.ascontiguousarray(a).reshape(b.nenv, cnt)
    mask = self._p._reset_mask
    if mask is not None and cnt:
      # inside reset_context(mask=...): only the envs being reset are written
      cur = b.read(name)
      rows = np.asarray(mask, dtype=bool)
      cur[rows] = a[rows]
      a = cur
    b.write(name, a)


class _Named:
  def __init__(self, physics):
    self.data = index.data_indexer(physics)
    self.model = index.model_indexer(physics)


class BatchedPhysics:
  """`nenv` independent simulations of one compiled model on one GPU."""

  def __init__(self, flat_model, nenv, device=0, lanes=None, njmax=None, nconmax=None):
    self.model = flat_model
    self.nenv = int(nenv)
    dj, dc = default_sizes(flat_model)
    self._ctor = dict(device=device, njmax=njmax or dj, nconmax=nconmax or dc,
                      lanes=lanes if lanes is not None else default_lanes(flat_model))
    self._dev_model = binding.Model(flat_model, device)
    self._batch = binding.Batch(self._dev_model, self.nenv, self._ctor["njmax"],
                                self._ctor["nconmax"], self._ctor["lanes"])
    self._axes = index.make_axes(flat_model)
    self.data = _Data(self)
    self.named = _Named(self)
    self._warnings_seen = np.zeros((self.nenv, 8), dtype=np.int64)
    self._reset_mask = None
    self.reset()

  # ---- construction (engine.py:332-470 from_xml_string / from_xml_path / ...) ----
  @classmethod
  def from_xml_string(cls, xml_string, assets=None, **kw):
    return cls(mjcf_compiler.compile_xml(xml_string, assets), **kw)

  @classmethod
  def from_xml_path(cls, path, **kw):
    return cls(mjcf_compiler.compile_file(path), **kw)

  @classmethod
  def from_flat_model(cls, flat_model, **kw):
    return cls(flat_model, **kw)

  from_model = from_flat_model          # engine.py:405-409

  @classmethod
  def from_byte_string(cls, byte_string, **kw):
    """engine.py:429-433: from a compiled model's `to_bytes()` (model.FlatModel.to_bytes: this
    library's own binary layout, not MuJoCo's closed MJB)."""
    return cls(M.FlatModel.from_byte_string(byte_string), **kw)

  @classmethod
  def from_binary_path(cls, file_path, **kw):
    """engine.py:447-458, for files written by `physics.model.save_binary(path)`."""
    return cls(M.FlatModel.from_binary_path(file_path), **kw)

  def _reload_from_model(self, flat_model):
    """engine.py:341-359: a new model behind the same Physics object, state as after
    construction."""
    device, nenv = self._ctor["device"], self.nenv
    self.free()
    # buffer sizes and lane width follow the new model
    BatchedPhysics.__init__(self, flat_model, nenv, device=device)

  def reload_from_xml_string(self, xml_string, assets=None):
    """engine.py:460-476."""
    self._reload_from_model(mjcf_compiler.compile_xml(xml_string, assets))

  def reload_from_xml_path(self, file_path):
    """engine.py:478-488."""
    self._reload_from_model(mjcf_compiler.compile_file(file_path))

  # ---- simulation (engine.py:141-166, 297-305) ---------------------------------------
  def set_control(self, control):
    """Sets the control signal of every env: [nenv, nu] (or [nu], broadcast)."""
    c = np.asarray(control, dtype=np.float64)
    if c.shape[-1:] != (self.model.nu,):
      raise ValueError("control has shape {}, expected (..., {})".format(c.shape, self.model.nu))
    self.data.ctrl = c

  def step(self, n_sub_steps=1):
    """Physics.step() of engine.py:149-166, `n_sub_steps` times in one launch.  Inside
    reset_context(mask=...) only the envs being reset advance (cheetah.py:73-75 settles the
    model inside initialize_episode)."""
    self._masked(lambda: self._batch.step(int(n_sub_steps)))
    self.check_invalid_state()

  def forward(self):
    self._masked(lambda: self._batch.forward(False))
    self.check_invalid_state()

  def _masked(self, call):
    if self._reset_mask is None:
      return call()
    self._batch.set_mask(self._reset_mask)
    try:
      call()
      self._batch.sync()
    finally:
      self._batch.set_mask(None)

  def reset(self, keyframe_id=None, mask=None):
    """engine.py:268-289: mj_resetData(+Keyframe), then forward with actuation
    disabled.  `mask` ([nenv] bool) restricts the reset to some envs."""
    if keyframe_id is not None and not 0 <= keyframe_id < self.model.nkey:
      raise ValueError("keyframe_id must be between 0 and {}".format(self.model.nkey - 1))
    self._batch.set_mask(mask)
    try:
      self._batch.reset(-1 if keyframe_id is None else keyframe_id)
      self._batch.forward(True)
      self._batch.sync()
    finally:
      self._batch.set_mask(None)
    if mask is None:
      self._warnings_seen[:] = 0
    else:
      self._warnings_seen[np.asarray(mask, dtype=bool)] = 0

  def after_reset(self):
    """engine.py:291-295: forward() with actuation disabled."""
    self._batch.set_mask(self._reset_mask)
    try:
      self._batch.forward(True)
      self._batch.sync()
    finally:
      self._batch.set_mask(None)
    self.check_invalid_state()

  @contextlib.contextmanager
  def reset_context(self, mask=None):
    """engine.py:247-266: reset, let the caller edit the state, then after_reset."""
    self.reset(mask=mask)
    self._reset_mask = mask
    try:
      yield self
    finally:
      self.after_reset()
      self._reset_mask = None

  def check_invalid_state(self):
    """engine.py:307-330: new warning counts become a PhysicsError."""
    w = self._batch.warnings().astype(np.int64)
    new = w > self._warnings_seen
    self._warnings_seen = w
    if new.any():
      names = [M.WARNINGS[i] for i in np.flatnonzero(new.any(axis=0))]
      envs = np.flatnonzero(new.any(axis=1))
      raise PhysicsError(_INVALID_PHYSICS_STATE.format(
          warning_names=", ".join(names), envs=envs[:16].tolist()))

  # ---- state accessors (engine.py:217-245, 540-573) ---------------------------------------
  def control(self): return self.data.ctrl
  def position(self): return self.data.qpos
  def velocity(self): return self.data.qvel
  def activation(self): return self.data.act
  def time(self): return self.data.time
  def timestep(self): return self.model.opt_timestep

  def state(self):
    return np.concatenate([self.data.qpos, self.data.qvel, self.data.act], axis=1)

  get_state = state

  # ---- copies and resumable state (engine.py:247-266, 332-339; wrapper/core.py:653-685) ----
  FULL_STATE_FIELDS = ("qpos", "qvel", "act", "ctrl", "qacc_warmstart", "qfrc_applied", "time")

  def copy(self, share_model=False):
    """Physics.copy (engine.py:332-339): a new BatchedPhysics with a bit copy of every env's
    data (mj_copyData: bp_batch_copy).  share_model=True makes the copy use this object's
    device model instead of uploading its own."""
    cls = type(self)
    new = cls.__new__(cls)
    new.model = self.model if share_model else self.model.copy()
    new.nenv = self.nenv
    new._ctor = dict(self._ctor)
    new._dev_model = self._dev_model if share_model else \
        binding.Model(new.model, self._ctor["device"])
    new._owns_model = not share_model
    new._batch = binding.Batch(new._dev_model, self.nenv, self._ctor["njmax"],
                               self._ctor["nconmax"], self._ctor["lanes"])
    new._batch.copy_from(self._batch)
    new._axes = self._axes
    new.data = _Data(new)
    new.named = _Named(new)
    new._warnings_seen = self._warnings_seen.copy()
    new._reset_mask = None
    return new

  def get_full_state(self):
    """Everything a bit-exact resume needs: qpos, qvel, act, ctrl, the
    solver's warm start, applied forces and time, plus the per-env model data tasks edit."""
    st = {f: getattr(self.data, f) for f in self.FULL_STATE_FIELDS}
    for name in ENV_MODEL_FIELDS:
      if self._batch.has_env_model(name):
        st["env_" + name] = self._batch.read_env_model(name)
    if int(self.model.nhfielddata) > 0:
      st["hfield"] = self._batch.read_hfield()
    return st

  def set_full_state(self, st):
    """Restores get_full_state() and recomputes what Physics.step() leaves behind (mj_step1 of
    the restored state), so that stepping on continues bit for bit."""
    for name in ENV_MODEL_FIELDS:
      if "env_" + name in st:
        self._batch.write_env_model(name, st["env_" + name])
    if "hfield" in st:
      self._batch.write_hfield(st["hfield"])
    for f in self.FULL_STATE_FIELDS:
      setattr(self.data, f, st[f])
    self._batch.step1()
    self._batch.sync()

  def __getstate__(self):
    return {"model": self.model, "nenv": self.nenv, "ctor": self._ctor,
            "state": self.get_full_state(), "warnings_seen": self._warnings_seen}

  def __setstate__(self, d):
    self.__init__(d["model"], d["nenv"], **d["ctor"])
    self.set_full_state(d["state"])
    self._warnings_seen = d["warnings_seen"]

  def set_state(self, physics_state):
    s = np.asarray(physics_state, dtype=np.float64)
    nq, nv, na = self.model.nq, self.model.nv, self.model.na
    if s.shape[-1] != nq + nv + na:
      raise ValueError("Input physics state has shape {}. Expected (..., {}).".format(
          s.shape, nq + nv + na))
    s = np.broadcast_to(s, (self.nenv, nq + nv + na))
    self.data.qpos = s[:, :nq]
    self.data.qvel = s[:, nq:nq + nv]
    if na:
      self.data.act = s[:, nq + nv:]

  def set_model_field(self, name, value):
    """Run-time model edit shared by all envs (core.py:581-620, finger.py:140-142)."""
    self._dev_model.set_field(name, value)

  def set_env_model_field(self, name, value, env0=0):
    """Per-env run-time model edit (reacher.py:96-101 `geom_pos['target']`): value is
    [n, ...] with the model field's trailing shape; supported: geom_pos, site_pos,
    body_pos, body_quat (manipulator.py:218-224), wrap_prm (point_mass.py:85-101).  Inside reset_context(mask=...) a
    full-batch value only reaches the envs being reset."""
    if self._reset_mask is not None and env0 == 0 and len(value) == self.nenv:
      cur = self._batch.read_env_model(name)
      rows = np.asarray(self._reset_mask, dtype=bool)
      cur[rows] = np.asarray(value, dtype=np.float64).reshape(cur.shape)[rows]
      value = cur
    self._batch.write_env_model(name, value, env0)

  def env_model_field(self, name):
    return self._batch.read_env_model(name)

  def set_hfield(self, data, env0=0):
    """Per-env `model.hfield_data` (quadruped.py:381): [n, nhfielddata], stored as float32
    like mjModel's."""
    self._batch.write_hfield(np.asarray(data, dtype=np.float32), env0)

  def hfield(self):
    return self._batch.read_hfield()

  @contextlib.contextmanager
  def disable(self, *flags):
    """model.disable('actuation', ...) context of core.py:581-620."""
    old = self.model.opt_disableflags
    new = old
    for f in flags:
      if f not in M.DSBL:
        raise ValueError("'{}' is not a valid flag name; valid names: {}".format(
            f, ", ".join(M.DISABLE_FLAGS)))
      new |= M.DSBL[f]
    self.set_model_field("opt_disableflags", new)
    try:
      yield
    finally:
      self.set_model_field("opt_disableflags", old)

  def device_view(self, name):
    """Zero-copy torch view [nenv, n] of a data field (contiguous images only)."""
    return self._batch.view(name)

  def action_spec(self):
    """engine.py:1018-1028: bounds from ctrlrange where ctrllimited, else +-inf."""
    from dm_control_b200 import specs
    m = self.model
    lim = np.asarray(m.actuator_ctrllimited).astype(bool)
    rng = np.asarray(m.actuator_ctrlrange).reshape(-1, 2)
    lo = np.where(lim, rng[:, 0], -np.inf)
    hi = np.where(lim, rng[:, 1], np.inf)
    return specs.BoundedArray(shape=(m.nu,), dtype=np.float64, minimum=lo, maximum=hi)

  def free(self):
    self._batch.free()
    if getattr(self, "_owns_model", True):
      self._dev_model.free()
