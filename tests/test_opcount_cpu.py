"""The op-count build of the oracle (oracle/opcount.h, SURVEY.md 8d binding rule) is the
same arithmetic as the plain C build: identical states after a few steps, and a count
that moves with the work done."""

import ctypes
import os
import subprocess

import numpy as np

import parity_util as PU
from oracle import oracle_py

ORACLE = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle")


def test_counting_build_matches_plain_build_and_counts():
  subprocess.check_call(["make", "-s", "-C", ORACLE, "_build/liboracle_ops.so"])
  L = ctypes.CDLL(os.path.join(ORACLE, "_build", "liboracle_ops.so"))
  vp = ctypes.c_void_p
  L.om_create.restype = vp; L.od_create.restype = vp; L.od_field.restype = vp
  L.om_create.argtypes = [ctypes.c_char_p, ctypes.c_int64, ctypes.c_char_p, ctypes.c_int]
  L.od_create.argtypes = [vp]
  L.od_field.argtypes = [vp, ctypes.c_char_p, ctypes.POINTER(ctypes.c_int64),
                         ctypes.POINTER(ctypes.c_int)]
  for f in ("o_reset", "o_forward", "o_physics_step"):
    getattr(L, f).argtypes = [vp, vp]
  m = PU.load_model("cheetah")
  qpos, qvel, ctrl = PU.random_states(m, 1, 4)
  blob = m.to_blob()
  om = L.om_create(blob, len(blob), None, 0)
  od = L.od_create(om)

  def field(name, n):
    cnt, dt = ctypes.c_int64(), ctypes.c_int()
    p = L.od_field(od, name.encode(), ctypes.byref(cnt), ctypes.byref(dt))
    return np.ctypeslib.as_array(ctypes.cast(p, ctypes.POINTER(ctypes.c_double)), (n,))

  L.o_reset(om, od)
  field("qpos", m.nq)[:] = qpos[0]; field("qvel", m.nv)[:] = qvel[0]
  field("ctrl", m.nu)[:] = ctrl[0]
  ref = oracle_py.Oracle(oracle_py.OracleModel(m))
  ref.reset(); ref.qpos[:] = qpos[0]; ref.qvel[:] = qvel[0]; ref.ctrl[:] = ctrl[0]
  L.o_forward(om, od); ref.forward()
  counts = []
  buf = (ctypes.c_ulonglong * 6)()
  for k in range(5):
    L.o_opcount_reset()
    L.o_physics_step(om, od); ref.step()
    L.o_opcount_read(buf)
    counts.append(sum(buf[:5]))
  # same source, same compiler family: the two builds agree to rounding of contracted ops
  np.testing.assert_allclose(field("qpos", m.nq), ref.qpos, rtol=1e-12, atol=1e-13)
  np.testing.assert_allclose(field("qacc", m.nv), ref.qacc, rtol=1e-9, atol=1e-9)
  assert all(3e3 < c < 1e5 for c in counts), counts       # cheetah: ~1.2e4 per sub-step
