"""ctypes binding of libbphys.so (include/bphys.h).

Plays the role of the reference's `mjbindings` + `wrapper/util.py` loader
(/root/reference/dm_control/mujoco/wrapper/util.py:107-120: library path from an
environment variable, `ctypes.cdll.LoadLibrary`, argtypes declared per function as
autowrap/c_declarations.py:407-427 does).  There is NO CPU fallback: if the library
or a CUDA device is missing, `load()` / `Model()` raise.
"""

import ctypes
import os

import numpy as np

ENV_VAR = "BPHYS_LIB_PATH"          # analogue of MJLIB_PATH (util.py:38)
_HERE = os.path.dirname(os.path.abspath(__file__))
DEFAULT_PATH = os.path.join(_HERE, "libbphys.so")

_lib = None


class Error(RuntimeError):
  """Raised for library-level failures (wrapper/core.py:56 `Error`)."""


def lib_path():
  return os.environ.get(ENV_VAR, DEFAULT_PATH)


_vp, _i64, _i32 = ctypes.c_void_p, ctypes.c_int64, ctypes.c_int
_SIGNATURES = {
    # name: (restype, argtypes)
    "bp_version": (_i32, []),
    "bp_last_error": (ctypes.c_char_p, []),
    "bp_device_count": (_i32, []),
    "bp_model_create": (_vp, [ctypes.c_char_p, _i64, _i32, ctypes.c_char_p, _i32]),
    "bp_model_free": (None, [_vp]),
    "bp_model_set_field": (_i32, [_vp, ctypes.c_char_p, _vp, _i64]),
    "bp_model_get_field": (_i32, [_vp, ctypes.c_char_p, _vp, _i64]),
    "bp_batch_create": (_vp, [_vp, _i64, _i32, _i32, _i32, ctypes.c_char_p, _i32]),
    "bp_batch_lanes": (_i32, [_vp]),
    "bp_batch_interleave": (_i32, [_vp]),
    "bp_batch_free": (None, [_vp]),
    "bp_batch_image": (_vp, [_vp]),
    "bp_batch_stride": (_i64, [_vp]),
    "bp_batch_nenv": (_i64, [_vp]),
    "bp_batch_smem_bytes": (_i64, [_vp]),
    "bp_field_count": (_i32, [_vp]),
    "bp_field_name": (ctypes.c_char_p, [_vp, _i32]),
    "bp_field_info": (_i32, [_vp, ctypes.c_char_p, ctypes.POINTER(_i64),
                             ctypes.POINTER(_i64), ctypes.POINTER(_i32)]),
    "bp_field_write": (_i32, [_vp, ctypes.c_char_p, _vp, _i64, _i64]),
    "bp_field_read": (_i32, [_vp, ctypes.c_char_p, _vp, _i64, _i64]),
    "bp_field_copy_device": (_i32, [_vp, ctypes.c_char_p, _vp, _i32]),
    "bp_fp64_peak": (_i32, [_i32, ctypes.POINTER(ctypes.c_double)]),
    "bp_set_mask": (_i32, [_vp, _vp]),
    "bp_hfield_write": (_i32, [_vp, _vp, _i64, _i64]),
    "bp_hfield_read": (_i32, [_vp, _vp, _i64, _i64]),
    "bp_env_model_write": (_i32, [_vp, ctypes.c_char_p, _vp, _i64, _i64]),
    "bp_env_model_read": (_i32, [_vp, ctypes.c_char_p, _vp, _i64, _i64]),
    "bp_step": (_i32, [_vp, _i32]),
    "bp_mj_step": (_i32, [_vp, _i32]),
    "bp_forward": (_i32, [_vp, _i32]),
    "bp_step1": (_i32, [_vp]),
    "bp_step2": (_i32, [_vp]),
    "bp_reset": (_i32, [_vp, _i32]),
    "bp_set_task": (_i32, [_vp, _i32, _vp, _i32, _vp, _i32]),
    "bp_task_nobs": (_i32, [_vp]),
    "bp_obs_reward": (_i32, [_vp, _vp, _vp]),
    "bp_obs_reward_host": (_i32, [_vp, _vp, _vp]),
    "bp_env_step": (_i32, [_vp, _i32, _vp, _vp, _vp]),
    "bp_env_step_host": (_i32, [_vp, _i32, _vp, _vp, _vp]),
    "bp_rewards_tolerance": (_i32, [_i32, _vp, _vp, _i64, ctypes.c_double, ctypes.c_double,
                                    ctypes.c_double, _i32, ctypes.c_double]),
    "bp_rewards_tolerance_noalloc": (_i32, [_i32, _vp, _vp, _i64, ctypes.c_double,
                                            ctypes.c_double, ctypes.c_double, _i32,
                                            ctypes.c_double]),
    "bp_rewards_tolerance_noalloc_max": (_i32, []),
    "bp_batch_copy": (_i32, [_vp, _vp]),
    "bp_batch_last_static": (_i32, [_vp]),
    "bp_contact_force": (_i32, [_vp, _i32, _vp]),
    "bp_object_velocity": (_i32, [_vp, _i32, _i32, _i32, _vp]),
    "bp_warnings": (_i32, [_vp, _vp]),
    "bp_sync": (_i32, [_vp]),
    "bp_batch_stream": (_vp, [_vp]),
}
EXPORTS = tuple(sorted(_SIGNATURES))


def load():
  """Loads libbphys.so and declares every signature of include/bphys.h."""
  global _lib
  if _lib is None:
    path = lib_path()
    if not os.path.exists(path):
      raise Error("Cannot find the batched-physics library at {!r}; build it with "
                  "`python __graft_entry__.py` or set {}.".format(path, ENV_VAR))
    L = ctypes.cdll.LoadLibrary(path)
    for name, (res, args) in _SIGNATURES.items():
      fn = getattr(L, name)
      fn.restype, fn.argtypes = res, args
    _lib = L
  return _lib


def fp64_peak_tflops(device=0):
  out = ctypes.c_double()
  check(load().bp_fp64_peak(device, ctypes.byref(out)))
  return out.value


SIGMOIDS = {"gaussian": 0, "linear": 1, "quadratic": 2, "long_tail": 3}


def rewards_tolerance(x, bounds=(0.0, 0.0), margin=0.0, sigmoid="gaussian",
                      value_at_margin=0.1, device=0):
  """rewards.tolerance (utils/rewards.py:93-135) on the device."""
  if sigmoid not in SIGMOIDS:
    raise ValueError("Unknown sigmoid type {!r}.".format(sigmoid))
  a = np.ascontiguousarray(x, dtype=np.float64)
  out = np.empty_like(a)
  check(load().bp_rewards_tolerance(device, _ptr(a), _ptr(out), a.size, float(bounds[0]),
                                    float(bounds[1]), float(margin), SIGMOIDS[sigmoid],
                                    float(value_at_margin)))
  return out


def rewards_tolerance_noalloc(x, bounds=(0.0, 0.0), margin=0.0, sigmoid="gaussian",
                              value_at_margin=0.1, device=0):
  """rewards_tolerance() through the entry point that allocates no device memory (fixed
  scratch of bp_rewards_tolerance_noalloc_max() elements; larger inputs raise)."""
  if sigmoid not in SIGMOIDS:
    raise ValueError("Unknown sigmoid type {!r}.".format(sigmoid))
  a = np.ascontiguousarray(x, dtype=np.float64)
  out = np.empty_like(a)
  check(load().bp_rewards_tolerance_noalloc(device, _ptr(a), _ptr(out), a.size,
                                            float(bounds[0]), float(bounds[1]), float(margin),
                                            SIGMOIDS[sigmoid], float(value_at_margin)))
  return out


def last_error():
  return load().bp_last_error().decode()


def check(status):
  if status != 0:
    raise Error("libbphys error {}: {}".format(status, last_error()))


def _ptr(a):
  return a.ctypes.data_as(ctypes.c_void_p)


class Model:
  """Device-resident compiled model (cf. `MjModel`, wrapper/core.py:444-627)."""

  def __init__(self, flat_model, device=0):
    L = load()
    self.flat = flat_model
    self.device = device
    blob = flat_model.to_blob()
    err = ctypes.create_string_buffer(1000)       # core.py:41 _ERROR_BUFSIZE
    self.ptr = L.bp_model_create(blob, len(blob), device, err, len(err))
    if not self.ptr:
      raise Error(err.value.decode())

  def set_field(self, name, value):
    """Writes a whole mjModel field on the device and in the host copy."""
    old = self.flat.arrays[name]
    a = np.ascontiguousarray(np.broadcast_to(np.asarray(value, dtype=old.dtype),
                                             old.shape))
    check(load().bp_model_set_field(self.ptr, name.encode(), _ptr(a), a.nbytes))
    self.flat.arrays[name] = a.copy()
    if name in ("geom_size", "geom_type"):      # the library re-derives the bounding radii
      rb = np.empty_like(self.flat.arrays["geom_rbound"])
      check(load().bp_model_get_field(self.ptr, b"geom_rbound", _ptr(rb), rb.nbytes))
      self.flat.arrays["geom_rbound"] = rb

  def free(self):
    if getattr(self, "ptr", None):
      load().bp_model_free(self.ptr)
      self.ptr = None

  def __del__(self):
    try:
      self.free()
    except Exception:   # interpreter shutdown
      pass


class Batch:
  """`nenv` mjData images on the device (cf. `MjData`, wrapper/core.py:630-780)."""

  def __init__(self, model, nenv, njmax=0, nconmax=0, lanes=0):
    L = load()
    self.model = model
    err = ctypes.create_string_buffer(1000)
    self.ptr = L.bp_batch_create(model.ptr, nenv, njmax, nconmax, lanes, err, len(err))
    if not self.ptr:
      raise Error(err.value.decode())
    self.nenv = nenv
    self.stride = L.bp_batch_stride(self.ptr)
    self.image_ptr = L.bp_batch_image(self.ptr)
    self.smem_bytes = L.bp_batch_smem_bytes(self.ptr)
    self.lanes = L.bp_batch_lanes(self.ptr)
    self.interleave = L.bp_batch_interleave(self.ptr)
    self.fields = {}
    for i in range(L.bp_field_count(self.ptr)):
      name = L.bp_field_name(self.ptr, i)
      off, cnt, dt = _i64(), _i64(), _i32()
      check(L.bp_field_info(self.ptr, name, ctypes.byref(off), ctypes.byref(cnt),
                            ctypes.byref(dt)))
      self.fields[name.decode()] = (off.value, cnt.value,
                                    np.float64 if dt.value == 0 else np.int32)

  # ---- host copies ------------------------------------------------------------
  def read(self, name, env0=0, n=None):
    n = self.nenv - env0 if n is None else n
    _, cnt, dt = self.fields[name]
    out = np.empty((n, cnt), dtype=dt)
    if cnt:
      check(load().bp_field_read(self.ptr, name.encode(), _ptr(out), env0, n))
    return out

  def write(self, name, value, env0=0):
    _, cnt, dt = self.fields[name]
    if cnt == 0:
      return
    a = np.ascontiguousarray(value, dtype=dt)
    if a.ndim == 1:
      a = np.ascontiguousarray(np.broadcast_to(a, (self.nenv - env0, cnt)))
    a = a.reshape(-1, cnt)
    if cnt:
      check(load().bp_field_write(self.ptr, name.encode(), _ptr(a), env0, a.shape[0]))

  def write_env_model(self, name, value, env0=0):
    """Per-env copy of a model array ("geom_pos", "site_pos", "body_pos", "body_quat",
    "wrap_prm"): value [n, count]."""
    cnt = self.model.flat.arrays[name].size
    a = np.ascontiguousarray(value, dtype=np.float64).reshape(-1, cnt)
    check(load().bp_env_model_write(self.ptr, name.encode(), _ptr(a), env0, a.shape[0]))
    self._env_model_written = set(getattr(self, "_env_model_written", ())) | {name}

  def read_env_model(self, name, env0=0, n=None):
    n = self.nenv - env0 if n is None else n
    shape = self.model.flat.arrays[name].shape
    out = np.empty((n,) + tuple(shape), dtype=np.float64)
    check(load().bp_env_model_read(self.ptr, name.encode(), _ptr(out), env0, n))
    return out

  def has_env_model(self, name):
    """True once a per-env copy of the model array exists (made on the first write)."""
    return name in getattr(self, "_env_model_written", ())

  def write_hfield(self, data, env0=0):
    """Per-env `model.hfield_data` (quadruped.py:381): data [n, nhfielddata] float32."""
    a = np.ascontiguousarray(data, dtype=np.float32)
    a = a.reshape(-1, int(self.model.flat.nhfielddata))
    check(load().bp_hfield_write(self.ptr, _ptr(a), env0, a.shape[0]))

  def read_hfield(self, env0=0, n=None):
    n = self.nenv - env0 if n is None else n
    out = np.empty((n, int(self.model.flat.nhfielddata)), dtype=np.float32)
    check(load().bp_hfield_read(self.ptr, _ptr(out), env0, n))
    return out

  # ---- zero-copy device views ---------------------------------------------------
  def view(self, name):
    """torch view [nenv, count] of a field inside the image (no copy).  Only for
    contiguous images (interleave == 1); use read_device() otherwise."""
    import torch
    if self.interleave != 1:
      raise Error("zero-copy views need a contiguous image (lanes > 1)")
    off, cnt, dt = self.fields[name]
    img = self.image_tensor()
    if dt == np.float64:
      return img.as_strided((self.nenv, cnt), (self.stride, 1), off // 8)
    return img.view(torch.int32).as_strided((self.nenv, cnt), (self.stride * 2, 1),
                                            off // 4)

  def image_tensor(self):
    import torch
    if getattr(self, "_image", None) is None:
      holder = _CudaArray(self.image_ptr, (self.nenv, self.stride), self)
      self._image = torch.as_tensor(holder, device="cuda:%d" % self.model.device)
    return self._image

  def read_device(self, name):
    """Dense torch copy [nenv, count] of a field (works for every layout)."""
    import torch
    _, cnt, dt = self.fields[name]
    out = torch.empty((self.nenv, cnt), device="cuda:%d" % self.model.device,
                      dtype=torch.float64 if dt == np.float64 else torch.int32)
    if cnt:
      self.copy_to_device(name, out.data_ptr())
    return out

  def copy_from_device(self, name, dev_ptr):
    check(load().bp_field_copy_device(self.ptr, name.encode(), dev_ptr, 1))

  def copy_to_device(self, name, dev_ptr):
    check(load().bp_field_copy_device(self.ptr, name.encode(), dev_ptr, 0))

  # ---- pipeline -------------------------------------------------------------------
  def set_mask(self, mask):
    if mask is None:
      check(load().bp_set_mask(self.ptr, None))
    else:
      a = np.ascontiguousarray(mask, dtype=np.uint8)
      assert a.shape == (self.nenv,)
      check(load().bp_set_mask(self.ptr, _ptr(a)))

  def step(self, nsub=1): check(load().bp_step(self.ptr, nsub))
  def mj_step(self, nsub=1): check(load().bp_mj_step(self.ptr, nsub))
  def forward(self, disable_actuation=False):
    check(load().bp_forward(self.ptr, int(disable_actuation)))
  def step1(self): check(load().bp_step1(self.ptr))
  def step2(self): check(load().bp_step2(self.ptr))
  def reset(self, key=-1): check(load().bp_reset(self.ptr, key))
  def sync(self): check(load().bp_sync(self.ptr))

  def set_task(self, kind, ip=(), fp=()):
    ia = np.ascontiguousarray(ip, dtype=np.int32)
    fa = np.ascontiguousarray(fp, dtype=np.float64)
    check(load().bp_set_task(self.ptr, kind, _ptr(ia), len(ia), _ptr(fa), len(fa)))
    self.nobs = load().bp_task_nobs(self.ptr)

  def obs_reward(self, obs_ptr, reward_ptr):
    check(load().bp_obs_reward(self.ptr, obs_ptr, reward_ptr))

  def obs_reward_host(self, obs, reward):
    check(load().bp_obs_reward_host(self.ptr, _ptr(obs), _ptr(reward)))

  def env_step(self, nsub, ctrl_ptr, obs_ptr, reward_ptr):
    """Fused Environment.step on device buffers (one launch); pointers are ints."""
    check(load().bp_env_step(self.ptr, nsub, ctrl_ptr, obs_ptr, reward_ptr))

  def env_step_host(self, nsub, ctrl, obs, reward):
    check(load().bp_env_step_host(self.ptr, nsub, _ptr(ctrl), _ptr(obs), _ptr(reward)))

  @property
  def last_static(self):
    """Whether the last launch ran a static instance of the kernel (csrc/bp_step_static.cu)."""
    return bool(load().bp_batch_last_static(self.ptr))

  def copy_from(self, other):
    """mj_copyData: this batch becomes a bit copy of `other` (same shapes)."""
    check(load().bp_batch_copy(self.ptr, other.ptr))
    self.nobs = getattr(other, "nobs", None)
    self._env_model_written = set(getattr(other, "_env_model_written", ()))

  def contact_force(self, contact_id):
    """[nenv, 2, 3] (force, torque) of contact `contact_id` in its frame; NaN rows where an
    env has fewer contacts (wrapper/core.py:747-769)."""
    out = np.empty((self.nenv, 6), dtype=np.float64)
    check(load().bp_contact_force(self.ptr, int(contact_id), _ptr(out)))
    return out.reshape(self.nenv, 2, 3)

  def object_velocity(self, object_id, object_type, local_frame=False):
    """[nenv, 2, 3] (linear, angular) velocity of an object (wrapper/core.py:713-745)."""
    out = np.empty((self.nenv, 6), dtype=np.float64)
    check(load().bp_object_velocity(self.ptr, int(object_type), int(object_id),
                                    int(bool(local_frame)), _ptr(out)))
    return out.reshape(self.nenv, 2, 3)[:, ::-1]

  def warnings(self):
    out = np.empty((self.nenv, 8), dtype=np.int32)
    check(load().bp_warnings(self.ptr, _ptr(out)))
    return out

  @property
  def stream(self):
    return load().bp_batch_stream(self.ptr)

  def free(self):
    if getattr(self, "ptr", None):
      self._image = None
      load().bp_batch_free(self.ptr)
      self.ptr = None

  def __del__(self):
    try:
      self.free()
    except Exception:
      pass


class _CudaArray:
  """Exposes the library-owned image through __cuda_array_interface__."""

  def __init__(self, ptr, shape, owner):
    self._owner = owner
    self.__cuda_array_interface__ = {
        "shape": tuple(shape), "typestr": "<f8", "data": (int(ptr), False),
        "version": 2, "strides": None}
