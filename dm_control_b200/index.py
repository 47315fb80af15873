"""Named indexing with a leading environment axis.

Restates the key conversion of the reference's `FieldIndexer`
(/root/reference/dm_control/mujoco/index.py:94-173 naming tables, :436-510 key
conversion, :584-646 struct indexer): rows are addressed by element name (ragged
for qpos/qvel/sensordata, index.py:94-99), columns by 'x','y','z' / 'qw'.. / 'xx'..
(index.py:101-112).  A *data* field of the batch has shape [nenv, rows(, cols)];
the key addresses axes 1.. and the env axis is kept whole, so
`named.data.qpos['hinge']` is an [nenv, 1] array.  Model fields have no env axis.
"""

import collections

import numpy as np

_COLS = {"xyz": ["x", "y", "z"], "quat": ["qw", "qx", "qy", "qz"],
         "mat": ["xx", "xy", "xz", "yx", "yy", "yz", "zx", "zy", "zz"]}

# field -> (row axis, column kind or None or int width)
DATA_AXES = {
    "qpos": ("nq", None), "qvel": ("nv", None), "qacc": ("nv", None),
    "qacc_warmstart": ("nv", None), "qfrc_applied": ("nv", None),
    "qfrc_bias": ("nv", None), "qfrc_passive": ("nv", None),
    "qfrc_actuator": ("nv", None), "qfrc_smooth": ("nv", None),
    "qacc_smooth": ("nv", None), "qfrc_constraint": ("nv", None),
    "qLDiagInv": ("nv", None), "act": ("na", None), "act_dot": ("na", None),
    "ctrl": ("nu", None), "actuator_length": ("nu", None),
    "actuator_velocity": ("nu", None), "actuator_force": ("nu", None),
    "ten_length": ("ntendon", None), "ten_velocity": ("ntendon", None),
    "xpos": ("nbody", "xyz"), "xquat": ("nbody", "quat"), "xmat": ("nbody", "mat"),
    "xipos": ("nbody", "xyz"), "ximat": ("nbody", "mat"),
    "subtree_com": ("nbody", "xyz"), "subtree_linvel": ("nbody", "xyz"),
    "cinert": ("nbody", 10), "crb": ("nbody", 10), "cvel": ("nbody", 6),
    "cacc": ("nbody", 6), "cfrc_int": ("nbody", 6), "cfrc_ext": ("nbody", 6),
    "cdof": ("nv", 6), "cdof_dot": ("nv", 6),
    "xanchor": ("njnt", "xyz"), "xaxis": ("njnt", "xyz"),
    "geom_xpos": ("ngeom", "xyz"), "geom_xmat": ("ngeom", "mat"),
    "site_xpos": ("nsite", "xyz"), "site_xmat": ("nsite", "mat"),
    "sensordata": ("nsensordata", None),
}
_MODEL_PREFIX = {"body_": "nbody", "jnt_": "njnt", "dof_": "nv", "geom_": "ngeom",
                 "site_": "nsite", "actuator_": "nu", "sensor_": "nsensor"}
_MODEL_COLS = {"body_pos": "xyz", "body_ipos": "xyz", "body_inertia": "xyz",
               "jnt_pos": "xyz", "jnt_axis": "xyz", "geom_size": "xyz", "geom_pos": "xyz",
               "site_size": "xyz", "site_pos": "xyz", "body_quat": "quat",
               "body_iquat": "quat", "geom_quat": "quat", "site_quat": "quat"}


class Axis:
  """Row or column axis: names -> index arrays (index.py:176-262 Axis classes)."""

  def __init__(self, names=None, sizes=None):
    self.names = list(names) if names is not None else None
    if names is None:
      self.lookup = None
    elif sizes is None:
      self.lookup = {n: i for i, n in enumerate(names) if n}
      self.ragged = False
    else:
      adr = np.concatenate([[0], np.cumsum(sizes)]).astype(int)
      self.lookup = {n: slice(int(adr[i]), int(adr[i + 1])) for i, n in enumerate(names) if n}
      self.ragged = True

  def convert(self, key):
    if self.lookup is None or key is None:
      return key
    if isinstance(key, str):
      try:
        return self.lookup[key] if not self.ragged else self.lookup[key]
      except KeyError:
        raise IndexError("invalid name {!r}; valid names: {}".format(key, sorted(self.lookup)))
    if isinstance(key, (list, tuple, np.ndarray)):
      arr = np.asarray(key)
      if arr.dtype.kind in "US":
        out = []
        for k in arr.ravel():
          v = self.convert(str(k))
          out.extend(range(v.start, v.stop) if isinstance(v, slice) else [v])
        return np.array(out, dtype=int)
    return key


def make_axes(model):
  """Size name -> Axis (index.py:264-299 make_axis_indexers)."""
  names = model.names
  jn = names["joint"]
  axes = collections.defaultdict(Axis)
  axes["nbody"] = Axis(names["body"])
  axes["njnt"] = Axis(jn)
  axes["ngeom"] = Axis(names["geom"])
  axes["nsite"] = Axis(names["site"])
  axes["nu"] = Axis(names["actuator"])
  axes["nsensor"] = Axis(names["sensor"])
  axes["ntendon"] = Axis(names["tendon"])
  if model.njnt:
    axes["nq"] = Axis(jn, np.diff(np.r_[model.jnt_qposadr, model.nq]))
    axes["nv"] = Axis(jn, np.diff(np.r_[model.jnt_dofadr, model.nv]))
  if model.nsensor:
    axes["nsensordata"] = Axis(names["sensor"],
                               np.diff(np.r_[model.sensor_adr, model.nsensordata]))
  if model.na:
    axes["na"] = Axis(names["actuator"][-model.na:])
  for k, v in _COLS.items():
    axes[k] = Axis(v)
  return axes


class FieldIndexer:
  """`named.data.<field>[key]` / `named.model.<field>[key]` (index.py:436-582)."""

  def __init__(self, getter, setter, row_axis, col_axis, ncols, batched):
    self._get, self._set = getter, setter
    self._row, self._col, self._ncols, self._batched = row_axis, col_axis, ncols, batched

  def _convert(self, key):
    if not isinstance(key, tuple):
      key = (key,)
    if len(key) > 2:
      raise IndexError("Index tuple has {} elements, but array has 2 named axes"
                       .format(len(key)))
    row = self._row.convert(key[0])
    out = [row]
    if len(key) == 2:
      if self._ncols is None:
        raise IndexError("field has a single named axis")
      out.append(self._col.convert(key[1]))
    return tuple(out)

  def _array(self):
    a = self._get()
    if self._ncols is not None:
      a = a.reshape(a.shape[:-1] + (-1, self._ncols)) if self._batched else \
          a.reshape(-1, self._ncols)
    return a

  def __getitem__(self, key):
    a = self._array()
    k = self._convert(key)
    return a[(slice(None),) + k] if self._batched else a[k]

  def __setitem__(self, key, value):
    a = self._array()
    k = self._convert(key)
    if self._batched:
      sel = (slice(None),) + k
      value = np.asarray(value)
      tgt = a[sel].shape
      # one value per env for a single-element key: [nenv] -> [nenv, 1, ...]
      if value.ndim == 1 and len(tgt) > 1 and value.shape[0] == tgt[0] and \
          all(s == 1 for s in tgt[1:]):
        value = value.reshape(tgt)
      a[sel] = value
    else:
      a[k] = value
    self._set(a.reshape(a.shape[0], -1) if self._batched else a)

  def __array__(self, dtype=None):
    a = self._array()
    return a.astype(dtype) if dtype is not None else a

  def __repr__(self):
    return "FieldIndexer(shape={})".format(self._array().shape)


class StructIndexer:
  """Attribute access to FieldIndexers (index.py:584-646 struct_indexer)."""

  def __init__(self, make):
    object.__setattr__(self, "_make", make)

  def __getattr__(self, name):
    fi = self._make(name)
    if fi is None:
      raise AttributeError(name)
    return fi

  def __setattr__(self, name, value):
    raise AttributeError("fields are assigned through indexing, e.g. named.data.%s[:] = v" % name)


def data_indexer(physics):
  axes = physics._axes

  def make(name):
    if name not in DATA_AXES or name not in physics._batch.fields:
      return None
    row, col = DATA_AXES[name]
    ncols = len(_COLS[col]) if isinstance(col, str) else col
    return FieldIndexer(lambda: physics._batch.read(name),
                        lambda a: physics._batch.write(name, a),
                        axes[row], axes[col] if isinstance(col, str) else Axis(), ncols, True)
  return StructIndexer(make)


def model_indexer(physics):
  axes = physics._axes
  flat = physics.model

  def make(name):
    if name not in flat.arrays:
      return None
    row = next((v for p, v in _MODEL_PREFIX.items() if name.startswith(p)), None)
    a = flat.arrays[name]
    col = _MODEL_COLS.get(name)
    ncols = len(_COLS[col]) if col else (a.shape[1] if a.ndim == 2 else None)
    return FieldIndexer(lambda: flat.arrays[name].copy(),
                        lambda v: physics.set_model_field(name, v),
                        axes[row] if row else Axis(), axes[col] if col else Axis(),
                        ncols, False)
  return StructIndexer(make)
