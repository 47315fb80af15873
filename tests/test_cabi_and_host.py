"""CPU-side checks: the C-ABI library loads and exports every symbol include/bphys.h
declares, the product refuses to run without a GPU (no CPU fallback), and the host
logic (model blob, task registry, named indexing tables, episode logic helpers)
behaves like the reference's (citations under /root/reference/dm_control/)."""

import os
import re

import numpy as np
import pytest

from dm_control_b200 import binding
from dm_control_b200 import control
from dm_control_b200 import index
from dm_control_b200 import model as M
from dm_control_b200.suite import tasks as T

import parity_util as PU
from conftest import needs_reference, REFERENCE_SUITE

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
  text = open(os.path.join(ROOT, "include", "bphys.h")).read()
  text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
  return sorted(set(re.findall(r"\b(bp_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
  lib = binding.load()
  names = _declared_symbols()
  assert len(names) >= 30
  for n in names:
    assert hasattr(lib, n), "libbphys.so lacks %s" % n
  # and the ctypes table covers the header (binding.py <-> bphys.h stay in sync)
  assert set(names) == set(binding.EXPORTS)
  assert lib.bp_version() == 100


def test_library_holds_every_execution_shape():
  """One launcher per lane width, plus the LEAN (G = 4, 8, 16) and MID (G = 8, 16)
  instantiations of the step kernel and the per-stage launchers of the split pipeline
  (csrc/bp_pipeline.cuh, bp_api.cu launch())."""
  lib = binding.load()
  shapes = ["bp_launch_step_g%d" % g for g in (1, 2, 4, 8, 16, 32)]
  shapes += ["bp_launch_step_g%d_lean" % g for g in (4, 8, 16)]
  shapes += ["bp_launch_step_g%d_mid" % g for g in (8, 16)]
  shapes += ["bp_launch_seg_g%d" % g for g in (4, 8, 16)]
  for n in shapes:
    assert hasattr(lib, n), "libbphys.so lacks %s" % n


def test_no_cpu_fallback_without_device():
  lib = binding.load()
  if lib.bp_device_count() > 0:
    pytest.skip("a CUDA device is present")
  with pytest.raises(binding.Error, match="no CUDA device"):
    binding.Model(PU.load_model("pendulum"))


def test_bad_blob_is_rejected_with_message():
  lib = binding.load()
  import ctypes
  err = ctypes.create_string_buffer(1000)
  assert not lib.bp_model_create(b"NOTAMODEL" * 4, 36, 0, err, 1000)
  assert b"magic" in err.value            # core.py:314-323: NULL + message


def test_blob_roundtrip_and_model_dims():
  # SURVEY §8.0 dimension table (nq/nv/nu, nbody, ngeom)
  dims = {"cartpole": (2, 2, 1, 3, 5), "acrobot": (2, 2, 1, 3, 4), "pendulum": (1, 1, 1, 2, 4),
          "cheetah": (9, 9, 6, 8, 9), "walker": (9, 9, 6, 8, 8), "humanoid": (28, 27, 21, 17, 20)}
  for d, (nq, nv, nu, nb, ng) in dims.items():
    m = PU.load_model(d)
    assert (m.nq, m.nv, m.nu, m.nbody, m.ngeom) == (nq, nv, nu, nb, ng), d
    blob = m.to_blob()
    assert blob[:8] == M.MAGIC and len(blob) % 8 == 0


@needs_reference
def test_compiled_caches_match_fresh_compile():
  from dm_control_b200 import mjcf_compiler as C
  for d in ("cartpole", "cheetah", "humanoid"):
    fresh = C.compile_file(os.path.join(REFERENCE_SUITE, d + ".xml"))
    cached = PU.load_model(d)
    assert sorted(fresh.arrays) == sorted(cached.arrays)
    for k in fresh.arrays:
      np.testing.assert_array_equal(fresh.arrays[k], cached.arrays[k], err_msg=d + "." + k)
    assert fresh.names == cached.names


def test_task_registry_and_substeps():
  # walker 0.025 / 0.0025 = 10 (walker.py:30,66), humanoid 0.025 / 0.005 = 5
  for name, nsub, nobs in (("walker", 10, 24), ("humanoid", 5, 67), ("cheetah", 1, 17),
                           ("cartpole", 1, 5), ("acrobot", 1, 6), ("pendulum", 1, 3)):
    tdef = next(t for (d, _), t in T.REGISTRY.items() if d == name)
    m = PU.load_model(tdef.model)
    assert T.n_sub_steps(tdef, m) == nsub
    assert sum(w for _, w in tdef.obs(m)) == nobs          # SURVEY §8.0 nobs column
    ip, fp = tdef.params(m)
    assert len(ip) <= 16 and len(fp) <= 8


def test_compute_n_steps_matches_reference_semantics():
  # rl/control.py:164-190 and control_test.py
  assert control.compute_n_steps(0.2, 0.1) == 2
  assert control.compute_n_steps(.003, .001) == 3
  with pytest.raises(ValueError):
    control.compute_n_steps(0.1, 0.2)
  with pytest.raises(ValueError):
    control.compute_n_steps(0.25, 0.1)


def test_named_axes_follow_reference_tables():
  m = PU.load_model("humanoid")
  axes = index.make_axes(m)
  # free root joint owns qpos[0:7] and qvel[0:6] (index.py:94-99 ragged axes)
  assert axes["nq"].convert("root") == slice(0, 7)
  assert axes["nv"].convert("root") == slice(0, 6)
  assert axes["mat"].convert("zz") == 8 and axes["xyz"].convert("z") == 2
  assert axes["nbody"].convert("torso") == m.name2id("torso", "body")
  np.testing.assert_array_equal(axes["nbody"].convert(["head", "torso"]),
                                [m.name2id("head", "body"), m.name2id("torso", "body")])
  with pytest.raises(IndexError):
    axes["nbody"].convert("no_such_body")


def test_flatten_observation_orders_keys():
  import collections
  obs = collections.OrderedDict([("b", np.ones((3, 2))), ("a", np.zeros((3, 1)))])
  flat = control.flatten_observation(obs)
  assert list(flat) == ["observations"] and flat["observations"].shape == (3, 3)
  np.testing.assert_array_equal(flat["observations"][0], [1, 1, 0])   # insertion order kept


def test_model_binary_roundtrip(tmp_path):
  """MjModel.to_bytes / from_byte_string / save_binary / from_binary_path
  (wrapper/core.py:460-468,506-519; core_test.py:82-95 testSerialize... pattern): a model
  that went through its byte string has the same arrays, names and scalars."""
  for domain in ("cheetah", "quadruped_escape", "point_mass"):
    m = PU.load_model(domain)
    b = m.to_bytes()
    assert b[:8] == b"BPMODEL1" and b.startswith(m.to_blob())
    for m2 in (M.FlatModel.from_byte_string(b),):
      assert sorted(m2.arrays) == sorted(m.arrays)
      for k, a in m.arrays.items():
        assert m2.arrays[k].dtype == a.dtype and m2.arrays[k].shape == a.shape, k
        np.testing.assert_array_equal(m2.arrays[k], a, err_msg=k)
      assert m2.names == m.names and m2.model_name == m.model_name
      assert m2.nq == m.nq and isinstance(m2.nq, int) and m2.to_bytes() == b
    path = str(tmp_path / (domain + ".bpb"))
    m.save_binary(path)
    assert M.FlatModel.from_binary_path(path).to_bytes() == b
  with pytest.raises(ValueError):
    M.FlatModel.from_byte_string(b"MJB\0" + b"\0" * 64)
  with pytest.raises(ValueError):
    M.FlatModel.from_byte_string(m.to_blob())          # blob without its name table


def test_committed_static_headers_match_the_library(tmp_path):
  """csrc/static/<model>.cuh are GENERATED by the library's own bp_static_header (host-only)
  through tools/gen_static_model.py; a layout or derivation change without regenerating them
  would leave instances that no longer match any live model (their hash / layout check would
  send every launch to the generic kernels).  Regenerate into a scratch directory and compare,
  and check that suite.load's default sizes are the ones the instances were generated for."""
  import ctypes
  import importlib.util
  root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
  spec = importlib.util.spec_from_file_location("gen_static", os.path.join(root, "tools",
                                                                           "gen_static_model.py"))
  gen = importlib.util.module_from_spec(spec)
  spec.loader.exec_module(gen)
  lib = ctypes.CDLL(binding.lib_path())
  lib.bp_static_header.argtypes = [ctypes.c_char_p, ctypes.c_int64, ctypes.c_int, ctypes.c_int,
                                   ctypes.c_char_p, ctypes.c_char_p]
  for name, (model, njmax, nconmax) in gen.TRAITS.items():
    blob = PU.load_model(model).to_blob()
    out = str(tmp_path / (name + ".cuh"))
    assert lib.bp_static_header(blob, len(blob), njmax, nconmax, name.encode(), out.encode()) == 0
    committed = os.path.join(root, "dm_control_b200", "csrc", "static", name + ".cuh")
    assert open(out).read() == open(committed).read(), \
        "%s is stale: run tools/gen_static_model.py and rebuild" % committed
    assert T.STATIC_SIZES[model] == (njmax, nconmax)
  for name, lanes, _ in gen.INSTANCES:
    assert os.path.exists(os.path.join(root, "dm_control_b200", "csrc",
                                       "bp_static_%s_g%d.cu" % (name, lanes)))
    if lanes == 1:
      assert T.STATIC_LANES[gen.TRAITS[name][0]] == 1
