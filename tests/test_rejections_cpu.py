"""Models the step kernels cannot simulate are refused with an error instead of being run
with other physics (the reference surfaces engine errors rather than
swallowing them, /root/reference/dm_control/mujoco/wrapper/core_test.py:97-103).  CPU side:
the MJCF compiler's check_supported(); the same rules in bp_model_create / bp_model_set_field
are exercised on the GPU in tests/test_gpu_env.py."""

import numpy as np
import pytest

from dm_control_b200 import mjcf_compiler as C
from dm_control_b200 import model as M

TWO_BODIES = """
<mujoco>
  <option {opt}>{flags}</option>
  <worldbody>
    <geom name="floor" type="plane" size="1 1 .1"/>
    <body name="a" pos="0 0 .3"><freejoint/><geom name="ga" type="{t1}" size=".1 .08 .06"/></body>
    <body name="b" pos=".1 0 .6"><freejoint/><geom name="gb" type="{t2}" size=".1 .08 .06"/></body>
  </worldbody>
</mujoco>"""


def _xml(t1="sphere", t2="sphere", opt="", flags=""):
  return TWO_BODIES.format(t1=t1, t2=t2, opt=opt, flags=flags)


def test_supported_model_compiles():
  m = C.compile_xml(_xml("sphere", "capsule"))
  assert int(m.npair) == 3


@pytest.mark.skipif(C.HAVE_PGS_CG, reason="PGS / CG are implemented")
@pytest.mark.parametrize("solver", ["PGS", "CG"])
def test_solver_other_than_newton_is_rejected(solver):
  with pytest.raises(ValueError, match="solver='%s' is not implemented" % solver):
    C.compile_xml(_xml(opt='solver="%s"' % solver))


def test_noslip_iterations_are_rejected():
  with pytest.raises(ValueError, match="noslip_iterations"):
    C.compile_xml(_xml(opt='noslip_iterations="3"'))


def test_mesh_geoms_are_rejected():
  xml = """<mujoco><asset><mesh name="m" vertex="0 0 0 1 0 0 0 1 0 0 0 1"/></asset>
    <worldbody><body><freejoint/><geom type="mesh" mesh="m"/></body></worldbody></mujoco>"""
  with pytest.raises(ValueError, match="mesh geoms are not supported"):
    C.compile_xml(xml)


def test_pair_without_narrowphase_is_rejected_unless_contacts_are_off_or_dropped():
  with pytest.raises(ValueError, match="no narrowphase.*'ga' \\(ellipsoid\\) - 'gb' \\(ellipsoid\\)"):
    C.compile_xml(_xml("ellipsoid", "ellipsoid"))
  # the model's own flags rule contacts out (cartpole.xml, swimmer.xml style): accepted
  m = C.compile_xml(_xml("ellipsoid", "ellipsoid", flags='<flag contact="disable"/>'))
  assert int(m.npair) == 3
  # explicit decision of the caller: the pair leaves the list, the others stay
  m = C.compile_xml(_xml("ellipsoid", "ellipsoid"), unsupported_pairs="drop")
  assert int(m.npair) == 2
  types = {(int(m.geom_type[a]), int(m.geom_type[b]))
           for a, b in zip(m.pair_geom1, m.pair_geom2)}
  assert types == {(M.GEOM_PLANE, M.GEOM_ELLIPSOID)}


def test_filtered_parent_child_pairs_do_not_count():
  xml = """<mujoco><worldbody><body name="p"><freejoint/><geom type="ellipsoid" size=".1 .2 .3"/>
    <body name="c" pos="0 0 .5"><joint type="hinge"/><geom type="ellipsoid" size=".1 .2 .3"/></body>
    </body></worldbody></mujoco>"""
  m = C.compile_xml(xml)
  assert int(m.npair) == 1 and int(m.pair_parentchild[0]) == 1
  with pytest.raises(ValueError, match="no narrowphase"):
    C.compile_xml(xml.replace("<worldbody>", '<option><flag filterparent="disable"/></option><worldbody>'))


def test_pair_table_matches_the_library():
  """pair_supported() is the compiler's copy of csrc/bp_api.cu pair_supported(): the types
  the narrowphase serves are symmetric and every plane / primitive pair of the suite is in."""
  for t1 in range(M.GEOM_MESH):
    for t2 in range(M.GEOM_MESH):
      assert C.pair_supported(t1, t2) == C.pair_supported(t2, t1)
  for t in (M.GEOM_SPHERE, M.GEOM_CAPSULE, M.GEOM_ELLIPSOID, M.GEOM_CYLINDER, M.GEOM_BOX):
    assert C.pair_supported(M.GEOM_PLANE, t)
  assert not C.pair_supported(M.GEOM_ELLIPSOID, M.GEOM_ELLIPSOID)
  assert not C.pair_supported(M.GEOM_PLANE, M.GEOM_PLANE)
  assert C.pair_supported(M.GEOM_SPHERE, M.GEOM_BOX) == C.HAVE_BOX_PAIRS


def test_pairs_are_listed_body_pair_major():
  """[MJ-doc] contact order: body pairs ascending, then the geoms of body1 x body2."""
  xml = """<mujoco><worldbody>
    <body name="a"><freejoint/><geom name="a1" size=".1"/><geom name="a2" size=".1" pos=".3 0 0"/></body>
    <body name="b" pos="0 1 0"><freejoint/><geom name="b1" size=".1"/><geom name="b2" size=".1" pos=".3 0 0"/></body>
    <body name="c" pos="0 2 0"><freejoint/><geom name="c1" size=".1"/></body>
    </worldbody></mujoco>"""
  m = C.compile_xml(xml)
  got = [(m.names["geom"][a], m.names["geom"][b]) for a, b in zip(m.pair_geom1, m.pair_geom2)]
  assert got == [("a1", "b1"), ("a1", "b2"), ("a2", "b1"), ("a2", "b2"), ("a1", "c1"),
                 ("a2", "c1"), ("b1", "c1"), ("b2", "c1")]
