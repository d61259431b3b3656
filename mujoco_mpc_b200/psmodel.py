"""PsModel: the flat, read-only model description that crosses the C ABI.

Field names and semantics follow MuJoCo 3.1.4's public ``mjModel`` layout (the struct the
reference planner receives in ``SamplingPlanner::Initialize(mjModel*, const Task&)``,
/root/reference/mjpc/planners/sampling/planner.cc:40-75) restricted to what the Predictive
Sampling path touches, so that a reference build fills a ``PsModel`` by copying pointers
out of its ``mjModel``.  This file is the single source of truth: ``tools/gen_header.py``
renders ``include/ps_model.h`` from ``FIELDS`` and the ctypes mirror below is built from the
same table, so the two cannot drift.
"""
import ctypes
import numpy as np

# (name, kind, shape) ; kind: 'i' int32 scalar, 'd' double scalar, 'I' int32 array,
# 'D' double array, 'S' char array.  shape is a tuple of size-field names / ints.
FIELDS = [
    # ---- sizes
    ("nq", "i", ()), ("nv", "i", ()), ("nu", "i", ()), ("na", "i", ()),
    ("nbody", "i", ()), ("njnt", "i", ()), ("ngeom", "i", ()), ("nsite", "i", ()),
    ("nmocap", "i", ()), ("nsensor", "i", ()), ("nsensordata", "i", ()),
    ("nuser_sensor", "i", ()), ("nnumeric", "i", ()), ("nnumericdata", "i", ()),
    ("ntext", "i", ()), ("ntextdata", "i", ()), ("nkey", "i", ()),
    ("nuserdata", "i", ()), ("nnames", "i", ()),
    # ---- options (mjOption)
    ("opt_timestep", "d", ()), ("opt_impratio", "d", ()), ("opt_tolerance", "d", ()),
    ("opt_ls_tolerance", "d", ()), ("opt_density", "d", ()), ("opt_viscosity", "d", ()),
    ("opt_integrator", "i", ()), ("opt_cone", "i", ()), ("opt_solver", "i", ()),
    ("opt_iterations", "i", ()), ("opt_ls_iterations", "i", ()),
    ("opt_disableflags", "i", ()), ("opt_enableflags", "i", ()),
    ("stat_meaninertia", "d", ()),
    ("opt_gravity", "D", (3,)), ("opt_wind", "D", (3,)),
    # ---- default generalized coordinates
    ("qpos0", "D", ("nq",)), ("qpos_spring", "D", ("nq",)),
    # ---- bodies
    ("body_parentid", "I", ("nbody",)), ("body_rootid", "I", ("nbody",)),
    ("body_jntnum", "I", ("nbody",)), ("body_jntadr", "I", ("nbody",)),
    ("body_dofnum", "I", ("nbody",)), ("body_dofadr", "I", ("nbody",)),
    ("body_geomnum", "I", ("nbody",)), ("body_geomadr", "I", ("nbody",)),
    ("body_mocapid", "I", ("nbody",)),
    ("body_pos", "D", ("nbody", 3)), ("body_quat", "D", ("nbody", 4)),
    ("body_ipos", "D", ("nbody", 3)), ("body_iquat", "D", ("nbody", 4)),
    ("body_mass", "D", ("nbody",)), ("body_subtreemass", "D", ("nbody",)),
    ("body_inertia", "D", ("nbody", 3)), ("body_invweight0", "D", ("nbody", 2)),
    # ---- joints
    ("jnt_type", "I", ("njnt",)), ("jnt_qposadr", "I", ("njnt",)),
    ("jnt_dofadr", "I", ("njnt",)), ("jnt_bodyid", "I", ("njnt",)),
    ("jnt_limited", "I", ("njnt",)),
    ("jnt_pos", "D", ("njnt", 3)), ("jnt_axis", "D", ("njnt", 3)),
    ("jnt_stiffness", "D", ("njnt",)), ("jnt_range", "D", ("njnt", 2)),
    ("jnt_margin", "D", ("njnt",)), ("jnt_solref", "D", ("njnt", 2)),
    ("jnt_solimp", "D", ("njnt", 5)),
    # ---- dofs
    ("dof_bodyid", "I", ("nv",)), ("dof_jntid", "I", ("nv",)),
    ("dof_parentid", "I", ("nv",)),
    ("dof_armature", "D", ("nv",)), ("dof_damping", "D", ("nv",)),
    ("dof_frictionloss", "D", ("nv",)), ("dof_solref", "D", ("nv", 2)),
    ("dof_solimp", "D", ("nv", 5)), ("dof_invweight0", "D", ("nv",)),
    # ---- geoms
    ("geom_type", "I", ("ngeom",)), ("geom_contype", "I", ("ngeom",)),
    ("geom_conaffinity", "I", ("ngeom",)), ("geom_condim", "I", ("ngeom",)),
    ("geom_bodyid", "I", ("ngeom",)), ("geom_priority", "I", ("ngeom",)),
    ("geom_group", "I", ("ngeom",)),
    ("geom_size", "D", ("ngeom", 3)), ("geom_pos", "D", ("ngeom", 3)),
    ("geom_quat", "D", ("ngeom", 4)), ("geom_friction", "D", ("ngeom", 3)),
    ("geom_solmix", "D", ("ngeom",)), ("geom_solref", "D", ("ngeom", 2)),
    ("geom_solimp", "D", ("ngeom", 5)), ("geom_margin", "D", ("ngeom",)),
    ("geom_gap", "D", ("ngeom",)), ("geom_rbound", "D", ("ngeom",)),
    # ---- sites
    ("site_bodyid", "I", ("nsite",)), ("site_pos", "D", ("nsite", 3)),
    ("site_quat", "D", ("nsite", 4)),
    # ---- actuators
    ("actuator_trntype", "I", ("nu",)), ("actuator_dyntype", "I", ("nu",)),
    ("actuator_gaintype", "I", ("nu",)), ("actuator_biastype", "I", ("nu",)),
    ("actuator_trnid", "I", ("nu", 2)), ("actuator_actadr", "I", ("nu",)),
    ("actuator_actnum", "I", ("nu",)), ("actuator_ctrllimited", "I", ("nu",)),
    ("actuator_forcelimited", "I", ("nu",)), ("actuator_actlimited", "I", ("nu",)),
    ("actuator_dynprm", "D", ("nu", 3)), ("actuator_gainprm", "D", ("nu", 3)),
    ("actuator_biasprm", "D", ("nu", 3)), ("actuator_ctrlrange", "D", ("nu", 2)),
    ("actuator_forcerange", "D", ("nu", 2)), ("actuator_actrange", "D", ("nu", 2)),
    ("actuator_gear", "D", ("nu",)),
    # ---- sensors
    ("sensor_type", "I", ("nsensor",)), ("sensor_objtype", "I", ("nsensor",)),
    ("sensor_objid", "I", ("nsensor",)), ("sensor_dim", "I", ("nsensor",)),
    ("sensor_adr", "I", ("nsensor",)),
    ("sensor_user", "D", ("nsensor", "nuser_sensor")),
    # ---- custom numeric / text, keyframes
    ("numeric_adr", "I", ("nnumeric",)), ("numeric_size", "I", ("nnumeric",)),
    ("numeric_data", "D", ("nnumericdata",)),
    ("text_adr", "I", ("ntext",)), ("text_size", "I", ("ntext",)),
    ("text_data", "S", ("ntextdata",)),
    ("key_qpos", "D", ("nkey", "nq")), ("key_qvel", "D", ("nkey", "nv")),
    ("key_act", "D", ("nkey", "na")), ("key_ctrl", "D", ("nkey", "nu")),
    # ---- names
    ("name_bodyadr", "I", ("nbody",)), ("name_jntadr", "I", ("njnt",)),
    ("name_geomadr", "I", ("ngeom",)), ("name_siteadr", "I", ("nsite",)),
    ("name_actuatoradr", "I", ("nu",)), ("name_sensoradr", "I", ("nsensor",)),
    ("name_numericadr", "I", ("nnumeric",)), ("name_textadr", "I", ("ntext",)),
    ("name_keyadr", "I", ("nkey",)),
    ("names", "S", ("nnames",)),
]

# enums (values follow MuJoCo 3.1.4 mjmodel.h)
JNT_FREE, JNT_BALL, JNT_SLIDE, JNT_HINGE = 0, 1, 2, 3
GEOM_PLANE, GEOM_HFIELD, GEOM_SPHERE, GEOM_CAPSULE, GEOM_ELLIPSOID, GEOM_CYLINDER, GEOM_BOX, GEOM_MESH = range(8)
INT_EULER, INT_RK4, INT_IMPLICIT, INT_IMPLICITFAST = 0, 1, 2, 3
CONE_PYRAMIDAL, CONE_ELLIPTIC = 0, 1
SOL_PGS, SOL_CG, SOL_NEWTON = 0, 1, 2
DSBL_CONSTRAINT, DSBL_EQUALITY, DSBL_FRICTIONLOSS, DSBL_LIMIT, DSBL_CONTACT, DSBL_PASSIVE, \
    DSBL_GRAVITY, DSBL_CLAMPCTRL, DSBL_WARMSTART, DSBL_FILTERPARENT, DSBL_ACTUATION, \
    DSBL_REFSAFE, DSBL_SENSOR, DSBL_MIDPHASE, DSBL_EULERDAMP = [1 << k for k in range(15)]
DYN_NONE, DYN_INTEGRATOR, DYN_FILTER, DYN_FILTEREXACT = 0, 1, 2, 3
GAIN_FIXED, GAIN_AFFINE = 0, 1
BIAS_NONE, BIAS_AFFINE = 0, 1
TRN_JOINT = 0
# mjtSensor subset
SENS_TOUCH, SENS_ACCELEROMETER = 0, 1
SENS_JOINTPOS, SENS_JOINTVEL = 8, 9
SENS_FRAMEPOS, SENS_FRAMEQUAT, SENS_FRAMEXAXIS, SENS_FRAMEYAXIS, SENS_FRAMEZAXIS, \
    SENS_FRAMELINVEL, SENS_FRAMEANGVEL, SENS_FRAMELINACC, SENS_FRAMEANGACC = range(25, 34)
SENS_SUBTREECOM, SENS_SUBTREELINVEL, SENS_SUBTREEANGMOM = 34, 35, 36
SENS_USER = 39
# mjtObj subset
OBJ_UNKNOWN, OBJ_BODY, OBJ_XBODY, OBJ_JOINT, OBJ_DOF, OBJ_GEOM, OBJ_SITE = range(7)
OBJ_ACTUATOR, OBJ_SENSOR, OBJ_NUMERIC, OBJ_TEXT, OBJ_KEY = 19, 20, 21, 22, 25

_CT = {"i": ctypes.c_int, "d": ctypes.c_double,
       "I": ctypes.POINTER(ctypes.c_int), "D": ctypes.POINTER(ctypes.c_double),
       "S": ctypes.c_char_p}


class PsModelC(ctypes.Structure):
    _fields_ = [(n, _CT[k]) for n, k, _ in FIELDS]


class PsModel:
    """Host-side model: numpy arrays keyed by field name + a ctypes view for the C ABI."""

    def __init__(self, values: dict):
        self.v = {}
        for name, kind, shape in FIELDS:
            if name not in values:
                raise KeyError(f"PsModel field missing: {name}")
            x = values[name]
            if kind == "i":
                self.v[name] = int(x)
            elif kind == "d":
                self.v[name] = float(x)
        for name, kind, shape in FIELDS:
            if kind in "ids":
                continue
            x = values[name]
            if kind == "S":
                b = bytes(x)
                assert len(b) == self.v[shape[0]], (name, len(b), self.v[shape[0]])
                self.v[name] = b
                continue
            dims = tuple(self.v[s] if isinstance(s, str) else s for s in shape)
            dt = np.int32 if kind == "I" else np.float64
            a = np.ascontiguousarray(np.asarray(x, dtype=dt).reshape(dims))
            self.v[name] = a
        self._c = None

    def __getattr__(self, k):
        v = self.__dict__.get("v")
        if v is not None and k in v:
            return v[k]
        raise AttributeError(k)

    def c_struct(self) -> PsModelC:
        if self._c is None:
            c = PsModelC()
            self._keep = []
            for name, kind, _ in FIELDS:
                x = self.v[name]
                if kind in "id":
                    setattr(c, name, x)
                elif kind == "S":
                    buf = ctypes.create_string_buffer(x, len(x) + 1)
                    self._keep.append(buf)
                    setattr(c, name, ctypes.cast(buf, ctypes.c_char_p))
                else:
                    # zero-size arrays still get a valid (dummy) pointer
                    arr = x if x.size else np.zeros(1, dtype=x.dtype)
                    self._keep.append(arr)
                    setattr(c, name, arr.ctypes.data_as(_CT[kind]))
            self._c = c
        return self._c

    # ---- name lookup (mj_name2id equivalent) -------------------------------------------
    _ADR = {"body": "name_bodyadr", "joint": "name_jntadr", "geom": "name_geomadr",
            "site": "name_siteadr", "actuator": "name_actuatoradr", "sensor": "name_sensoradr",
            "numeric": "name_numericadr", "text": "name_textadr", "key": "name_keyadr"}

    def name_of(self, kind: str, i: int) -> str:
        adr = int(self.v[self._ADR[kind]][i])
        end = self.v["names"].index(b"\0", adr)
        return self.v["names"][adr:end].decode()

    def name2id(self, kind: str, name: str) -> int:
        adrs = self.v[self._ADR[kind]]
        for i in range(len(adrs)):
            if self.name_of(kind, i) == name:
                return i
        return -1

    def numeric(self, name: str, default=None):
        i = self.name2id("numeric", name)
        if i < 0:
            return default
        a = int(self.numeric_adr[i])
        return self.numeric_data[a:a + int(self.numeric_size[i])]

    def text(self, name: str, default=None):
        i = self.name2id("text", name)
        if i < 0:
            return default
        a = int(self.text_adr[i])
        return self.text_data[a:a + int(self.text_size[i]) - 1].decode()
