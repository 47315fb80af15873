"""Flat structure-of-arrays model (`FlatModel`) and its C-ABI blob encoding.

`FlatModel` plays the role of the reference's `MjModel`
(/root/reference/dm_control/mujoco/wrapper/core.py:444-627): field names mirror
`mjModel` so `named.model.*` indexing keeps working.  It is produced by
`mjcf_compiler.compile_xml` (the stand-in for `mj_loadXML`, core.py:312-328) and
shipped to the native library as a self-describing blob (the analogue of the MJB
byte string of core.py:460-468).

Blob layout (little endian):
    char   magic[8]  = b"BPMODEL1"
    int32  nfields, int32 reserved
    nfields x { char name[40]; int32 dtype (0=f64, 1=i32); int32 ndim;
                int32 dims[4]; int64 offset; int64 count }
    payload (each field 8-byte aligned, offsets relative to blob start)
"""

import io
import struct

import numpy as np

MAGIC = b"BPMODEL1"
NAMES_MAGIC = b"BPNAMES1"
_ENTRY = struct.Struct("<40sii4iqq")

# --- enums (values follow MuJoCo 2.0's mjmodel.h; the generated
# mjbindings/enums.py the reference would import is absent) -----

GEOM_PLANE, GEOM_HFIELD, GEOM_SPHERE, GEOM_CAPSULE, GEOM_ELLIPSOID, \
    GEOM_CYLINDER, GEOM_BOX, GEOM_MESH = range(8)
GEOM_TYPES = {"plane": 0, "hfield": 1, "sphere": 2, "capsule": 3,
              "ellipsoid": 4, "cylinder": 5, "box": 6, "mesh": 7}

JNT_FREE, JNT_BALL, JNT_SLIDE, JNT_HINGE = range(4)
JNT_TYPES = {"free": 0, "ball": 1, "slide": 2, "hinge": 3}

INT_EULER, INT_RK4 = 0, 1
CONE_PYRAMIDAL, CONE_ELLIPTIC = 0, 1
SOL_PGS, SOL_CG, SOL_NEWTON = 0, 1, 2

DISABLE_FLAGS = ["constraint", "equality", "frictionloss", "limit", "contact",
                 "passive", "gravity", "clampctrl", "warmstart", "filterparent",
                 "actuation", "refsafe"]
ENABLE_FLAGS = ["override", "energy", "fwdinv", "sensornoise"]
DSBL = {n: 1 << i for i, n in enumerate(DISABLE_FLAGS)}
ENBL = {n: 1 << i for i, n in enumerate(ENABLE_FLAGS)}

OBJ_TYPES = ["unknown", "body", "xbody", "joint", "dof", "geom", "site",
             "camera", "light", "mesh", "hfield", "texture", "material", "pair",
             "exclude", "equality", "tendon", "actuator", "sensor", "numeric",
             "text", "tuple", "key"]
OBJ = {n: i for i, n in enumerate(OBJ_TYPES)}

SENSOR_TYPES = ["touch", "accelerometer", "velocimeter", "gyro", "force",
                "torque", "magnetometer", "rangefinder", "jointpos", "jointvel",
                "tendonpos", "tendonvel", "actuatorpos", "actuatorvel",
                "actuatorfrc", "ballquat", "ballangvel", "jointlimitpos",
                "jointlimitvel", "jointlimitfrc", "tendonlimitpos",
                "tendonlimitvel", "tendonlimitfrc", "framepos", "framequat",
                "framexaxis", "frameyaxis", "framezaxis", "framelinvel",
                "frameangvel", "framelinacc", "frameangacc", "subtreecom",
                "subtreelinvel", "subtreeangmom", "user"]
SENS = {n: i for i, n in enumerate(SENSOR_TYPES)}
STAGE_NONE, STAGE_POS, STAGE_VEL, STAGE_ACC = range(4)

WRAP_NONE, WRAP_JOINT, WRAP_PULLEY, WRAP_SITE, WRAP_SPHERE, WRAP_CYLINDER = range(6)
TRN_JOINT, TRN_JOINTINPARENT, TRN_SLIDERCRANK, TRN_TENDON, TRN_SITE = range(5)
EQ_CONNECT, EQ_WELD, EQ_JOINT, EQ_TENDON, EQ_DISTANCE = range(5)
DYN_NONE, DYN_INTEGRATOR, DYN_FILTER, DYN_MUSCLE, DYN_USER = range(5)
GAIN_FIXED, GAIN_MUSCLE, GAIN_USER = range(3)
BIAS_NONE, BIAS_AFFINE, BIAS_MUSCLE, BIAS_USER = range(4)

CNSTR_EQUALITY, CNSTR_FRICTION_DOF, CNSTR_FRICTION_TENDON, CNSTR_LIMIT_JOINT, \
    CNSTR_LIMIT_TENDON, CNSTR_CONTACT_FRICTIONLESS, CNSTR_CONTACT_PYRAMIDAL, \
    CNSTR_CONTACT_ELLIPTIC = range(8)

WARNINGS = ["INERTIA", "CONTACTFULL", "CNSTRFULL", "VGEOMFULL", "BADQPOS",
            "BADQVEL", "BADQACC", "BADCTRL"]

MINVAL = 1e-15
MAXVAL = 1e10

# object-name tables kept on the host for named indexing (index.py:264-289)
NAME_CATEGORIES = ["body", "joint", "geom", "site", "camera", "light", "mesh",
                   "hfield", "tendon", "actuator", "sensor", "equality", "key",
                   "numeric"]


class FlatModel:
  """Host-side compiled model: a bag of numpy arrays + name tables."""

  def __init__(self):
    self.arrays = {}      # name -> np.ndarray (float64 or int32)
    self.names = {c: [] for c in NAME_CATEGORIES}
    self.model_name = ""

  # attribute access mirrors `model.nq`, `model.jnt_range`, ... of the reference
  def __getattr__(self, key):
    arrays = self.__dict__.get("arrays", {})
    if key in arrays:
      a = arrays[key]
      if a.ndim == 0 or (a.shape == (1,) and key in _SCALARS):
        return a.reshape(()).item()
      return a
    raise AttributeError(key)

  def set(self, key, value, dtype=None):
    a = np.asarray(value, dtype=dtype)
    scalar = a.ndim == 0
    a = np.ascontiguousarray(a)
    if a.dtype.kind in "iub":
      a = a.astype(np.int32)
    else:
      a = a.astype(np.float64)
    if scalar:
      a = a.reshape(1)
      _SCALARS.add(key)
    self.arrays[key] = a

  def name2id(self, name, object_type):
    """core.py:535-557."""
    try:
      return self.names[object_type].index(name)
    except (KeyError, ValueError):
      raise ValueError("No {} with name {!r} exists.".format(object_type, name))

  def id2name(self, object_id, object_type):
    """core.py:559-579."""
    return self.names[object_type][object_id]

  # ---- blob encoding -------------------------------------------------------
  def to_blob(self):
    fields = sorted(self.arrays.items())
    head = 16 + _ENTRY.size * len(fields)
    off = (head + 7) // 8 * 8
    entries, payload = [], []
    for name, a in fields:
      dims = list(a.shape) + [0] * (4 - a.ndim)
      dt = 0 if a.dtype == np.float64 else 1
      raw = a.tobytes()
      entries.append(_ENTRY.pack(name.encode()[:39], dt, a.ndim, *dims, off,
                                 a.size))
      pad = (-len(raw)) % 8
      payload.append(raw + b"\0" * pad)
      off += len(raw) + pad
    out = io.BytesIO()
    out.write(MAGIC)
    out.write(struct.pack("<ii", len(fields), 0))
    for e in entries:
      out.write(e)
    out.write(b"\0" * ((head + 7) // 8 * 8 - head))
    for p in payload:
      out.write(p)
    return out.getvalue()

  # ---- model binaries: MjModel.to_bytes / from_byte_string / save_binary /
  # from_binary_path (wrapper/core.py:460-468, 506-519).  The byte string is the C-ABI blob
  # above followed by a name trailer (b"BPNAMES1", int64 length, JSON); MuJoCo's own MJB
  # layout is closed and undocumented in the reference tree, so real .mjb files cannot be
  # read -- a model saved here loads here, bit for bit.
  def to_bytes(self):
    import json
    meta = json.dumps({"names": self.names, "model_name": self.model_name,
                       "scalars": sorted(k for k in self.arrays if k in _SCALARS)}).encode()
    return self.to_blob() + NAMES_MAGIC + struct.pack("<q", len(meta)) + meta

  @classmethod
  def from_byte_string(cls, byte_string):
    import json
    b = bytes(byte_string)
    if len(b) < 16 or b[:8] != MAGIC:
      raise ValueError("not a compiled-model byte string (bad magic)")
    nfields = struct.unpack_from("<i", b, 8)[0]
    m = cls()
    end = 16 + _ENTRY.size * nfields
    for k in range(nfields):
      name, dt, ndim, d0, d1, d2, d3, off, count = _ENTRY.unpack_from(b, 16 + k * _ENTRY.size)
      dtype = np.float64 if dt == 0 else np.int32
      a = np.frombuffer(b, dtype=dtype, count=count, offset=off).copy()
      m.arrays[name.rstrip(b"\0").decode()] = a.reshape((d0, d1, d2, d3)[:ndim])
      end = max(end, off + (count * a.itemsize + 7) // 8 * 8)
    if b[end:end + 8] != NAMES_MAGIC:
      raise ValueError("compiled-model byte string without its name table")
    n = struct.unpack_from("<q", b, end + 8)[0]
    meta = json.loads(b[end + 16:end + 16 + n].decode())
    m.names = {c: list(v) for c, v in meta["names"].items()}
    m.model_name = meta["model_name"]
    _SCALARS.update(meta["scalars"])
    return m

  def save_binary(self, binary_path):
    with open(binary_path, "wb") as f:
      f.write(self.to_bytes())

  @classmethod
  def from_binary_path(cls, binary_path):
    with open(binary_path, "rb") as f:
      return cls.from_byte_string(f.read())

  # ---- persistence (the compiled-model cache shipped with the suite) -------
  def save(self, path):
    meta = {"__names__" + c: np.array(v, dtype=object) for c, v in
            self.names.items()}
    scal = np.array(sorted(k for k in self.arrays if k in _SCALARS),
                    dtype=object)
    np.savez_compressed(path, __scalars__=scal,
                        __model_name__=np.array(self.model_name, dtype=object),
                        **meta, **self.arrays)

  @classmethod
  def load(cls, path):
    m = cls()
    with np.load(path, allow_pickle=True) as z:
      for k in z.files:
        if k.startswith("__names__"):
          m.names[k[len("__names__"):]] = [str(s) for s in z[k]]
        elif k == "__scalars__":
          _SCALARS.update(str(s) for s in z[k])
        elif k == "__model_name__":
          m.model_name = str(z[k])
        else:
          m.arrays[k] = z[k]
    return m

  def copy(self):
    m = FlatModel()
    m.arrays = {k: v.copy() for k, v in self.arrays.items()}
    m.names = {k: list(v) for k, v in self.names.items()}
    m.model_name = self.model_name
    return m


_SCALARS = set()
