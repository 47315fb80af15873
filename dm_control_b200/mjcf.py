"""`mjcf.Physics.bind` / `from_mjcf_model` on the batched backend.

Mirrors the element-binding surface of /root/reference/dm_control/mjcf/physics.py:469-690
(`Physics.from_mjcf_model`, `bind`, `Binding`, the attribute tables of :65-152 and the dirty /
forward protocol of :520-534) over `BatchedPhysics`: `physics.bind(elements).attr` reads and
writes the rows of `named.data.<field>` / `named.model.<field>` that belong to those elements.
Data attributes carry a leading env axis ([nenv, n_elements(, cols)]); model attributes are
shared by all envs and have none.

The PyMJCF object model itself (mjcf/element.py, parser, attribute types) is out of scope: an
"element" here is anything with a `.tag` (or `.namespace`) and a `.full_identifier` (or
`.name`) -- which real `mjcf.Element`s have -- and `Element(tag, name)` / `elements(model)`
provide such handles for a compiled model.
"""

import collections
import re

import numpy as np

from dm_control_b200 import index
from dm_control_b200.physics import BatchedPhysics

# mjcf/constants.py:42-45
MJDATA_TRIGGERS_DIRTY = ("qpos", "qvel", "act", "ctrl", "qfrc_applied", "xfrc_applied")
MJMODEL_DOESNT_TRIGGER_DIRTY = ("rgba", "matid", "emission", "specular", "shininess",
                                "reflectance")

# namespace -> (array sizes whose leading dimension belongs to it, prefixes to strip):
# the table of mjcf/physics.py:130-152 for the element kinds the compiled models hold
_NAMESPACES = {
    "actuator": (("na", "nu"), ("actuator",)),
    "body": (("nbody",), ("body",)),
    "equality": (("neq",), ("eq",)),
    "geom": (("ngeom",), ("geom",)),
    "joint": (("nq", "nv", "njnt"), ("jnt", "dof")),
    "sensor": (("nsensor", "nsensordata"), ("sensor",)),
    "site": (("nsite",), ("site",)),
    "tendon": (("ntendon",), ("tendon", "ten")),
}
_MODEL_SIZES = dict(index._MODEL_PREFIX, **{"tendon_": "ntendon", "eq_": "neq"})
# MJCF tag -> namespace (element.spec.namespace of the reference's schema)
_TAG_NAMESPACE = {
    "worldbody": "body", "body": "body", "joint": "joint", "freejoint": "joint", "geom": "geom",
    "site": "site", "motor": "actuator", "position": "actuator", "velocity": "actuator",
    "general": "actuator", "cylinder": "actuator", "muscle": "actuator", "fixed": "tendon",
    "spatial": "tendon", "weld": "equality", "connect": "equality",
    "touch": "sensor", "accelerometer": "sensor", "velocimeter": "sensor", "gyro": "sensor",
    "force": "sensor", "torque": "sensor", "rangefinder": "sensor", "jointpos": "sensor",
    "jointvel": "sensor", "actuatorfrc": "sensor", "framepos": "sensor", "framexaxis": "sensor",
    "frameyaxis": "sensor", "framezaxis": "sensor", "subtreecom": "sensor",
    "subtreelinvel": "sensor",
}
_ACTUATOR_STATE_FIELDS = frozenset(("act", "act_dot"))     # leading dimension na


class Element(collections.namedtuple("Element", "tag full_identifier")):
  """Handle of one named element of a compiled model: what `bind` needs of an mjcf.Element."""
  __slots__ = ()

  @property
  def namespace(self):
    return _TAG_NAMESPACE.get(self.tag, self.tag)


def elements(flat_model, namespace):
  """Every named element of `namespace` ('body', 'joint', 'geom', 'site', 'actuator', 'sensor',
  'tendon') of a compiled model, in id order."""
  return [Element(namespace, n) for n in flat_model.names[namespace] if n]


def _get_namespace(element):
  ns = getattr(element, "namespace", None)
  if ns is None and hasattr(element, "spec"):          # a real mjcf.Element
    ns = element.spec.namespace.split("/")[0]
  if ns is None or getattr(element, "tag", None) == "worldbody":
    ns = _TAG_NAMESPACE.get(getattr(element, "tag", None), ns)
  if ns is None:
    raise ValueError("cannot tell the element type of {!r}".format(element))
  return ns


def _identifier(element):
  name = getattr(element, "full_identifier", None)
  if name is None:
    name = getattr(element, "name", None)
  if getattr(element, "tag", None) == "worldbody":
    name = "world"
  if not name:
    raise ValueError("only named elements can be bound: {!r}".format(element))
  return name


def names_from_elements(mjcf_elements):
  """mjcf/physics.py:155-190: (namespace, name or list of names)."""
  if isinstance(mjcf_elements, collections.abc.Iterable) and not isinstance(mjcf_elements, Element):
    tup = tuple(mjcf_elements)
    if not tup:
      return None, None
    ns = _get_namespace(tup[0])
    for e in tup:
      if _get_namespace(e) != ns:
        raise ValueError("Cannot bind to a collection containing multiple element types "
                         "({!r} != {!r}).".format(_get_namespace(e), ns))
    return ns, [_identifier(e) for e in tup]
  return _get_namespace(mjcf_elements), _identifier(mjcf_elements)


def _attributes(physics, namespace):
  """attribute name -> ('data' | 'model', field): fields whose leading dimension belongs to the
  namespace, prefix stripped (mjcf/physics.py:65-98 _get_attributes)."""
  sizes, prefixes = _NAMESPACES[namespace]
  strip = re.compile(r"\A({})_".format("|".join(prefixes)))
  out = {}
  for field, (row, _) in index.DATA_AXES.items():
    if row in sizes and field in physics._batch.fields:
      out[strip.sub("", field)] = ("data", field)
  for field in physics.model.arrays:
    size = next((v for p, v in _MODEL_SIZES.items() if field.startswith(p)), None)
    if size in sizes:
      out.setdefault(strip.sub("", field), ("model", field))
  return out


class BoundArray(np.ndarray):
  """Array returned by a Binding attribute: a host copy that writes through to the physics
  when assigned into (the role of SynchronizingArrayWrapper, mjcf/physics.py:183-268)."""

  def __new__(cls, values, binding, name):
    obj = np.array(values).view(cls)
    obj._binding, obj._name, obj._root = binding, name, True
    return obj

  def __array_finalize__(self, obj):
    self._binding = getattr(obj, "_binding", None)
    self._name = getattr(obj, "_name", None)
    self._root = False          # views and results of arithmetic do not write through

  def __setitem__(self, key, value):
    super().__setitem__(key, value)
    if self._root and self._binding is not None:
      setattr(self._binding, self._name, np.asarray(self))

  def copy(self, order="C"):
    return np.array(self, order=order)


class Binding:
  """mjcf/physics.py:271-430: attribute access to the rows of the bound element(s)."""

  __slots__ = ("_physics", "_namespace", "_named_index", "_attributes")

  def __init__(self, physics, namespace, named_index):
    if namespace not in _NAMESPACES:
      raise ValueError("elements of type {!r} cannot be bound to physics".format(namespace))
    object.__setattr__(self, "_physics", physics)
    object.__setattr__(self, "_namespace", namespace)
    object.__setattr__(self, "_named_index", named_index)
    object.__setattr__(self, "_attributes", _attributes(physics, namespace))

  def __dir__(self):
    return sorted(set(dir(type(self))).union(self._attributes))

  @property
  def element_id(self):
    m = self._physics.model
    if isinstance(self._named_index, list):
      return np.array([m.name2id(n, self._namespace) for n in self._named_index])
    return m.name2id(self._named_index, self._namespace)

  def _lookup(self, name):
    try:
      kind, field = self._attributes[name]
    except KeyError:
      raise AttributeError("bound element <{}> does not have attribute {!r}".format(
          self._namespace, name))
    indexer = getattr(self._physics.named.data if kind == "data" else
                      self._physics.named.model, field)
    key = self._named_index
    if field in _ACTUATOR_STATE_FIELDS:        # stateless actuators have no activation row
      m = self._physics.model
      keep = lambda n: m.name2id(n, "actuator") >= m.nu - m.na
      key = [n for n in key if keep(n)] if isinstance(key, list) else (key if keep(key) else [])
    return kind, field, indexer, key

  def _triggers_dirty(self, kind, name):
    return name in MJDATA_TRIGGERS_DIRTY if kind == "data" else \
        name not in MJMODEL_DOESNT_TRIGGER_DIRTY

  def __getattr__(self, name):
    kind, field, indexer, key = self._lookup(name)
    p = self._physics
    if p.is_dirty and not self._triggers_dirty(kind, name):
      p.forward()
    return BoundArray(indexer[key], self, name)

  def __setattr__(self, name, value):
    kind, field, indexer, key = self._lookup(name)
    p = self._physics
    if p.is_dirty and not self._triggers_dirty(kind, name):
      p.forward()
    indexer[key] = value
    if self._triggers_dirty(kind, name):
      p.mark_as_dirty()

  def __getitem__(self, idx):
    """binding['xpos', 'z'] / binding['qpos']: mjcf/physics.py:395-420."""
    name, cols = idx if isinstance(idx, tuple) else (idx, None)
    kind, field, indexer, key = self._lookup(name)
    p = self._physics
    if p.is_dirty and not self._triggers_dirty(kind, name):
      p.forward()
    return indexer[key] if cols is None else indexer[key, cols]

  def __setitem__(self, idx, value):
    name, cols = idx if isinstance(idx, tuple) else (idx, None)
    kind, field, indexer, key = self._lookup(name)
    if cols is None:
      indexer[key] = value
    else:
      indexer[key, cols] = value
    if self._triggers_dirty(kind, name):
      self._physics.mark_as_dirty()


class _EmptyBinding:
  """mjcf/physics.py:433-449."""
  __slots__ = ("_arr",)

  def __init__(self):
    object.__setattr__(self, "_arr", np.empty((0)))

  def __getattr__(self, name):
    return self._arr

  def __setattr__(self, name, value):
    raise ValueError("Cannot assign a value to an empty binding.")


class Physics(BatchedPhysics):
  """`BatchedPhysics` that supports binding to MJCF elements (mjcf/physics.py:469-690)."""

  def __init__(self, *args, **kw):
    self._bindings = {(): _EmptyBinding()}
    self._dirty = False
    super().__init__(*args, **kw)

  @classmethod
  def from_mjcf_model(cls, mjcf_model, **kw):
    """mjcf/physics.py:472-497: from anything that renders itself as MJCF
    (`to_xml_string()` and, optionally, `get_assets()` -- the mjcf.RootElement protocol),
    or from an XML string."""
    if isinstance(mjcf_model, (str, bytes)):
      return cls.from_xml_string(mjcf_model, **kw)
    xml = mjcf_model.to_xml_string()
    assets = mjcf_model.get_assets() if hasattr(mjcf_model, "get_assets") else None
    return cls.from_xml_string(xml, assets=assets, **kw)

  def _reload_from_model(self, flat_model):
    """mjcf/physics.py:499-518: a reload drops every cached binding."""
    self._bindings = {(): _EmptyBinding()}
    self._dirty = False
    super()._reload_from_model(flat_model)

  @property
  def is_dirty(self):
    """Whether derived quantities need to be recalculated (mjcf/physics.py:520-523)."""
    return self._dirty

  def mark_as_dirty(self):
    self._dirty = True

  def forward(self):
    super().forward()
    self._dirty = False

  def step(self, n_sub_steps=1):
    super().step(n_sub_steps)
    self._dirty = False

  def bind(self, mjcf_elements):
    """mjcf/physics.py:536-690: a Binding for an element or a collection of elements of one
    kind; `None` and empty collections give the empty binding."""
    if mjcf_elements is None:
      return None
    namespace, named_index = names_from_elements(mjcf_elements)
    if namespace is None:
      return self._bindings[()]
    key = (namespace, tuple(named_index) if isinstance(named_index, list) else named_index)
    b = self._bindings.get(key)
    if b is None:
      b = self._bindings[key] = Binding(self, namespace, named_index)
    return b
