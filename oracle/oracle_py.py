"""ctypes binding of the CPU oracle.  TEST INFRASTRUCTURE ONLY: imported by
tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs; never by the dm_control_b200 product path."""

import ctypes
import os
import subprocess

import numpy as np

_DIR = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_DIR, "_build", "liboracle.so")


def build(force=False):
  src = os.path.join(_DIR, "mjstep_oracle.c")
  if (force or not os.path.exists(_SO) or
      os.path.getmtime(_SO) < os.path.getmtime(src)):
    subprocess.check_call(["make", "-s", "-C", _DIR, "-B" if force else "-s"])
  return _SO


_lib = None


def lib():
  global _lib
  if _lib is None:
    L = ctypes.CDLL(build())
    vp, i64p, ip = ctypes.c_void_p, ctypes.POINTER(ctypes.c_int64), \
        ctypes.POINTER(ctypes.c_int)
    L.om_create.restype = vp
    L.om_create.argtypes = [ctypes.c_char_p, ctypes.c_int64, ctypes.c_char_p,
                            ctypes.c_int]
    L.om_free.argtypes = [vp]
    L.od_create.restype = vp
    L.od_create.argtypes = [vp]
    L.od_free.argtypes = [vp]
    L.od_field.restype = vp
    L.od_field.argtypes = [vp, ctypes.c_char_p, i64p, ip]
    L.om_field.restype = vp
    L.om_field.argtypes = [vp, ctypes.c_char_p, i64p, ip]
    for f in ("o_reset", "o_forward", "o_step", "o_step1", "o_step2",
              "o_physics_step"):
      getattr(L, f).argtypes = [vp, vp]
      getattr(L, f).restype = None
    dp = ctypes.POINTER(ctypes.c_double)
    L.o_run_batch.argtypes = [vp, ctypes.c_int, ctypes.c_int, ctypes.c_int, dp,
                              dp, dp, dp, dp, dp, ctypes.c_int, ctypes.c_int]
    L.o_run_batch.restype = None
    L.o_run_batch_hf.argtypes = L.o_run_batch.argtypes + [dp]
    L.o_run_batch_hf.restype = None
    L.o_contact_force.argtypes = [vp, vp, ctypes.c_int, dp]
    L.o_contact_force.restype = None
    L.o_object_velocity.argtypes = [vp, vp, ctypes.c_int, ctypes.c_int, dp, ctypes.c_int]
    L.o_object_velocity.restype = None
    _lib = L
  return _lib


class OracleModel:
  def __init__(self, flat_model):
    self.flat = flat_model
    blob = flat_model.to_blob()
    err = ctypes.create_string_buffer(256)
    self.ptr = lib().om_create(blob, len(blob), err, 256)
    if not self.ptr:
      raise RuntimeError(err.value.decode())

  def field(self, name):
    n, dt = ctypes.c_int64(), ctypes.c_int()
    p = lib().om_field(self.ptr, name.encode(), ctypes.byref(n), ctypes.byref(dt))
    if not p:
      raise KeyError(name)
    ct = ctypes.c_double if dt.value == 0 else ctypes.c_int32
    return np.ctypeslib.as_array(ctypes.cast(p, ctypes.POINTER(ct)),
                                 shape=(n.value,))

  def __del__(self):
    if getattr(self, "ptr", None) and _lib is not None:
      _lib.om_free(self.ptr)
      self.ptr = None


class Oracle:
  """One environment: model + data, numpy views on every data field."""

  def __init__(self, flat_model):
    self.model = flat_model if isinstance(flat_model, OracleModel) else \
        OracleModel(flat_model)
    self.m = self.model.flat
    self.ptr = lib().od_create(self.model.ptr)
    self._views = {}

  def __getattr__(self, name):
    if name.startswith("_") or name in ("model", "m", "ptr"):
      raise AttributeError(name)
    v = self._views.get(name)
    if v is None:
      n, dt = ctypes.c_int64(), ctypes.c_int()
      p = lib().od_field(self.ptr, name.encode(), ctypes.byref(n),
                         ctypes.byref(dt))
      if not p:
        raise AttributeError(name)
      ct = ctypes.c_double if dt.value == 0 else ctypes.c_int32
      v = np.ctypeslib.as_array(ctypes.cast(p, ctypes.POINTER(ct)),
                                shape=(max(n.value, 0),))
      self._views[name] = v
    return v

  def reset(self): lib().o_reset(self.model.ptr, self.ptr)
  def forward(self): lib().o_forward(self.model.ptr, self.ptr)
  def mj_step(self): lib().o_step(self.model.ptr, self.ptr)
  def step1(self): lib().o_step1(self.model.ptr, self.ptr)
  def step2(self): lib().o_step2(self.model.ptr, self.ptr)
  def step(self): lib().o_physics_step(self.model.ptr, self.ptr)

  def contact_force(self, contact_id):
    """wrapper/core.py:747-769: 2 x 3 (force, torque) in the contact frame."""
    out = np.zeros(6)
    lib().o_contact_force(self.model.ptr, self.ptr, int(contact_id),
                          out.ctypes.data_as(ctypes.POINTER(ctypes.c_double)))
    return out.reshape(2, 3)

  def object_velocity(self, object_id, object_type, local_frame=False):
    """wrapper/core.py:713-745: 2 x 3 (linear, angular); object_type is an mjtObj value
    (1 body, 2 xbody, 5 geom, 6 site)."""
    out = np.zeros(6)
    lib().o_object_velocity(self.model.ptr, self.ptr, int(object_type), int(object_id),
                            out.ctypes.data_as(ctypes.POINTER(ctypes.c_double)),
                            int(bool(local_frame)))
    return out.reshape(2, 3)[::-1]

  def __del__(self):
    if getattr(self, "ptr", None) and _lib is not None:
      _lib.od_free(self.ptr)
      self.ptr = None


def run_batch(omodel, qpos, qvel, act, warm, time, ctrl, nsteps, nsub,
              nthreads, hfield=None):
  """In-place batched stepping on host cores (cpu_baseline driver)."""
  dp = ctypes.POINTER(ctypes.c_double)
  nenv = qpos.shape[0]
  per_step = 1 if ctrl.ndim == 3 else 0
  arrs = [np.ascontiguousarray(a, dtype=np.float64) for a in
          (qpos, qvel, act, warm, time, ctrl)]
  hf = None
  if hfield is not None:
    hf = np.ascontiguousarray(hfield, dtype=np.float64)
    assert hf.shape == (nenv, int(omodel.flat.nhfielddata))
  lib().o_run_batch_hf(omodel.ptr, nenv, nsteps, nsub,
                       *[a.ctypes.data_as(dp) for a in arrs], per_step, nthreads,
                       hf.ctypes.data_as(dp) if hf is not None else None)
  return arrs[:5]
