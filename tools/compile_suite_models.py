"""Compiles the suite's MJCF files into FlatModel caches (dm_control_b200/suite/models).

Run in the build container, where /root/reference is mounted; the GPU box has no
reference tree, so the product and the tests load these compiled models instead
(the analogue of shipping MJB files, wrapper/core.py:510-519).
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from dm_control_b200 import mjcf_compiler as C   # noqa: E402

SUITE = "/root/reference/dm_control/suite"
DOMAINS = ["cartpole", "acrobot", "pendulum", "cheetah", "walker", "humanoid",
           "hopper", "reacher", "finger", "swimmer", "point_mass", "humanoid_CMU", "fish", "ball_in_cup"]
OUT = os.path.join(ROOT, "dm_control_b200", "suite", "models")


def quadruped_xml(floor_size=None, terrain=False, rangefinders=False, walls_and_ball=False):
  """suite/quadruped.py:52-93 make_model: the same element removals on the XML tree."""
  import xml.etree.ElementTree as ET
  root = ET.fromstring(open(os.path.join(SUITE, "quadruped.xml")).read())
  parent = {c: p for p in root.iter() for c in p}

  def drop(el):
    parent[el].remove(el)

  def find(tag, name):
    return next(e for e in root.iter(tag) if e.get("name") == name)

  if floor_size is not None:
    find("geom", "floor").set("size", "%g %g .5" % (floor_size, floor_size))
  if not walls_and_ball:
    for w in ("wall_px", "wall_py", "wall_nx", "wall_ny"):
      drop(find("geom", w))
    drop(find("body", "ball"))
    drop(find("site", "target"))
  if not terrain:
    drop(find("geom", "terrain"))
  if not rangefinders:
    for rf in list(root.iter("rangefinder")):
      drop(rf)
  return ET.tostring(root, encoding="unicode")


def swimmer_xml(n_bodies):
  """suite/swimmer.py:87-133 _make_model / _make_body: the same elements appended to
  swimmer.xml (camera moves are rendering only and left out)."""
  import xml.etree.ElementTree as ET
  root = ET.fromstring(open(os.path.join(SUITE, "swimmer.xml")).read())
  parent = root.find("./worldbody/body")
  actuator = ET.SubElement(root, "actuator")
  sensor = ET.SubElement(root, "sensor")
  for i in range(n_bodies - 1):
    child = ET.Element("body", name="segment_%d" % i, pos="0 .1 0")
    ET.SubElement(child, "geom", {"class": "visual", "name": "visual_%d" % i})
    ET.SubElement(child, "geom", {"class": "inertial", "name": "inertial_%d" % i})
    child.append(ET.Element("site", name="site_%d" % i))
    limit = 360.0 / n_bodies
    child.append(ET.Element("joint", {"name": "joint_%d" % i, "range": "%r %r" % (-limit, limit)}))
    actuator.append(ET.Element("motor", name="motor_%d" % i, joint="joint_%d" % i))
    sensor.append(ET.Element("velocimeter", name="velocimeter_%d" % i, site="site_%d" % i))
    sensor.append(ET.Element("gyro", name="gyro_%d" % i, site="site_%d" % i))
    parent.append(child)
    parent = child
  return ET.tostring(root, encoding="unicode")


class _Dir(dict):
  def __contains__(self, k):
    return os.path.exists(os.path.join(SUITE, k))

  def __getitem__(self, k):
    with open(os.path.join(SUITE, k)) as f:
      return f.read()


def cartpole_xml(n_poles):
  """suite/cartpole.py:105-127 _make_model: a chain of pole bodies below the first one and
  the floor moved down (camera moves are rendering only and left out)."""
  import xml.etree.ElementTree as ET
  root = ET.fromstring(open(os.path.join(SUITE, "cartpole.xml")).read())
  parent = root.find("./worldbody/body/body")
  for i in range(2, n_poles + 1):
    child = ET.SubElement(parent, "body", name="pole_%d" % i, pos="0 0 1", childclass="pole")
    ET.SubElement(child, "joint", name="hinge_%d" % i)
    ET.SubElement(child, "geom", name="pole_%d" % i)
    parent = child
  root.find("./worldbody/geom").set("pos", "0 0 %s" % (1 - n_poles - .05))
  return ET.tostring(root, encoding="unicode")


def manipulator_xml(use_peg=False, insert=False):
  """suite/manipulator.py:47-67 make_model: the props the task does not use are removed."""
  import xml.etree.ElementTree as ET
  root = ET.fromstring(open(os.path.join(SUITE, "manipulator.xml")).read())
  required = ["peg", "target_peg"] + (["slot"] if insert else []) if use_peg else \
             ["ball", "target_ball"] + (["cup"] if insert else [])
  wb = root.find("worldbody")
  for b in list(wb.findall("body")):
    if b.get("name") in ("ball", "target_ball", "cup", "peg", "target_peg", "slot") and \
       b.get("name") not in required:
      wb.remove(b)
  return ET.tostring(root, encoding="unicode")


def stacker_xml(n_boxes):
  """suite/stacker.py:38-49 make_model: the unused boxes are removed."""
  import xml.etree.ElementTree as ET
  root = ET.fromstring(open(os.path.join(SUITE, "stacker.xml")).read())
  wb = root.find("worldbody")
  for b in list(wb.findall("body")):
    if b.get("name") in ["box%d" % k for k in range(n_boxes, 4)]:
      wb.remove(b)
  return ET.tostring(root, encoding="unicode")


def manipulators():
  for name, kw in (("manipulator_ball", {}), ("manipulator_peg", dict(use_peg=True)),
                   ("manipulator_insert_ball", dict(insert=True)),
                   ("manipulator_insert_peg", dict(use_peg=True, insert=True))):
    m = C.compile_xml(manipulator_xml(**kw), _Dir())
    m.save(os.path.join(OUT, name + ".npz"))
    print(name, "nq", m.nq, "nbody", m.nbody, "npair", m.npair, "neq", m.neq)
  for n in (2, 4):
    m = C.compile_xml(stacker_xml(n), _Dir())
    m.save(os.path.join(OUT, "stacker_%d.npz" % n))
    print("stacker_%d nq" % n, m.nq, "nbody", m.nbody, "npair", m.npair, "njmax", m.njmax)


def cartpoles():
  for n in (2, 3):
    m = C.compile_xml(cartpole_xml(n), _Dir())
    m.save(os.path.join(OUT, "cartpole%d.npz" % n))
    print("cartpole%d nq" % n, m.nq, "nbody", m.nbody, "npair", m.npair)


def swimmers():
  for n in (6, 15):
    m = C.compile_xml(swimmer_xml(n), _Dir())
    m.save(os.path.join(OUT, "swimmer%d.npz" % n))
    print("swimmer%d nq" % n, m.nq, "nv", m.nv, "nu", m.nu, "nsensordata", m.nsensordata,
          "density", m.opt_density)


def main():
  os.makedirs(OUT, exist_ok=True)
  for d in DOMAINS:
    # humanoid_CMU: the two hand ellipsoids are the one live pair without a narrowphase
    # (ellipsoid - ellipsoid); it is dropped explicitly (DESIGN.md §2 departures)
    m = C.compile_file(os.path.join(SUITE, d + ".xml"),
                       unsupported_pairs="drop" if d == "humanoid_CMU" else "error")
    path = os.path.join(OUT, d + ".npz")
    m.save(path)
    print(d, "nq", m.nq, "nv", m.nv, "nbody", m.nbody, "->", os.path.getsize(path), "B")


def quadruped():
  # quadruped.py:96-105 walk (floor 20 * 0.5); run differs only in the floor size, which
  # no computation depends on (infinite plane)
  assets = C.compile_file.__globals__  # noqa: F841 (includes resolved relative to SUITE)
  xml = quadruped_xml(floor_size=10)

  class _Dir(dict):
    def __contains__(self, k):
      return os.path.exists(os.path.join(SUITE, k))

    def __getitem__(self, k):
      with open(os.path.join(SUITE, k)) as f:
        return f.read()

  m = C.compile_xml(xml, _Dir())
  path = os.path.join(OUT, "quadruped.npz")
  m.save(path)
  print("quadruped nq", m.nq, "nv", m.nv, "nu", m.nu, "na", m.na, "nbody", m.nbody, "ngeom",
        m.ngeom, "ntendon", m.ntendon, "neq", m.neq, "nsensordata", m.nsensordata)
  # quadruped.py:121-130 escape: floor 40, terrain height field, rangefinders
  m = C.compile_xml(quadruped_xml(floor_size=40, terrain=True, rangefinders=True), _Dir())
  m.save(os.path.join(OUT, "quadruped_escape.npz"))
  # quadruped.py:133-142 fetch: walls, ball and target kept, the XML's own floor size
  mf = C.compile_xml(quadruped_xml(walls_and_ball=True), _Dir())
  mf.save(os.path.join(OUT, "quadruped_fetch.npz"))
  print("quadruped_fetch ngeom", mf.ngeom, "npair", mf.npair, "nq", mf.nq, "njmax", mf.njmax,
        "nconmax", mf.nconmax)
  print("quadruped_escape ngeom", m.ngeom, "npair", m.npair, "nhfpair", m.nhfpair,
        "nhfielddata", m.nhfielddata, "nsensordata", m.nsensordata)


if __name__ == "__main__":
  main()
  quadruped()
  swimmers()
  cartpoles()
  manipulators()
