"""A small MJCF -> PsModel compiler for the model class the Predictive Sampling path supports.

The reference obtains its ``mjModel`` from MuJoCo 3.1.4's XML compiler (``mj_loadXML`` at
/root/reference/mjpc/agent.cc:242).  MuJoCo is not available in this environment, so this
module restates the subset of the compiler the six benchmarked tasks need: default classes,
kinematic tree, primitive-geom inertias, explicit <inertial>, hinge/slide/free joints,
motor/general/position actuators, the sensors the residuals read, custom numeric/text,
keyframes and the ``mj_setConst`` quantities (invweight0, meaninertia) the constraint
solver uses.  It is input preparation, not part of the planning hot path.
"""
import math
import xml.etree.ElementTree as ET
import numpy as np

from . import psmodel as P
from .psmodel import PsModel

MINVAL = 1e-15


# ---------------------------------------------------------------------------- math helpers
def _floats(s, n=None):
    v = [float(x) for x in s.replace(",", " ").split()]
    if n is not None and len(v) != n:
        raise ValueError(f"expected {n} numbers, got '{s}'")
    return v


def quat_mul(a, b):
    return np.array([
        a[0] * b[0] - a[1] * b[1] - a[2] * b[2] - a[3] * b[3],
        a[0] * b[1] + a[1] * b[0] + a[2] * b[3] - a[3] * b[2],
        a[0] * b[2] - a[1] * b[3] + a[2] * b[0] + a[3] * b[1],
        a[0] * b[3] + a[1] * b[2] - a[2] * b[1] + a[3] * b[0]])


def quat2mat(q):
    w, x, y, z = q
    return np.array([
        [w * w + x * x - y * y - z * z, 2 * (x * y - w * z), 2 * (x * z + w * y)],
        [2 * (x * y + w * z), w * w - x * x + y * y - z * z, 2 * (y * z - w * x)],
        [2 * (x * z - w * y), 2 * (y * z + w * x), w * w - x * x - y * y + z * z]])


def mat2quat(R):
    # robust conversion (Shepperd)
    t = np.trace(R)
    if t > 0:
        s = math.sqrt(t + 1.0) * 2
        q = [0.25 * s, (R[2, 1] - R[1, 2]) / s, (R[0, 2] - R[2, 0]) / s, (R[1, 0] - R[0, 1]) / s]
    elif R[0, 0] > R[1, 1] and R[0, 0] > R[2, 2]:
        s = math.sqrt(1.0 + R[0, 0] - R[1, 1] - R[2, 2]) * 2
        q = [(R[2, 1] - R[1, 2]) / s, 0.25 * s, (R[0, 1] + R[1, 0]) / s, (R[0, 2] + R[2, 0]) / s]
    elif R[1, 1] > R[2, 2]:
        s = math.sqrt(1.0 + R[1, 1] - R[0, 0] - R[2, 2]) * 2
        q = [(R[0, 2] - R[2, 0]) / s, (R[0, 1] + R[1, 0]) / s, 0.25 * s, (R[1, 2] + R[2, 1]) / s]
    else:
        s = math.sqrt(1.0 + R[2, 2] - R[0, 0] - R[1, 1]) * 2
        q = [(R[1, 0] - R[0, 1]) / s, (R[0, 2] + R[2, 0]) / s, (R[1, 2] + R[2, 1]) / s, 0.25 * s]
    q = np.array(q)
    return q / np.linalg.norm(q)


def axisangle2quat(axis, ang):
    axis = np.asarray(axis, float)
    n = np.linalg.norm(axis)
    if n < MINVAL:
        return np.array([1.0, 0, 0, 0])
    axis = axis / n
    return np.concatenate([[math.cos(ang / 2)], axis * math.sin(ang / 2)])


def z2quat(vec):
    """Quaternion rotating the z axis onto ``vec`` (MuJoCo compiler convention)."""
    vec = np.asarray(vec, float)
    vec = vec / np.linalg.norm(vec)
    axis = np.cross([0, 0, 1.0], vec)
    s = np.linalg.norm(axis)
    if s < 1e-10:
        axis = np.array([1.0, 0, 0])
    else:
        axis = axis / s
    ang = math.atan2(s, vec[2])
    return np.concatenate([[math.cos(ang / 2)], axis * math.sin(ang / 2)])


def euler2quat(e, seq="xyz"):
    q = np.array([1.0, 0, 0, 0])
    for ang, ax in zip(e, seq):
        a = {"x": [1, 0, 0], "y": [0, 1, 0], "z": [0, 0, 1]}[ax.lower()]
        r = axisangle2quat(a, ang)
        # lower case = intrinsic (rotating frame): post-multiply ; upper = extrinsic
        q = quat_mul(q, r) if ax.islower() else quat_mul(r, q)
    return q


# ---------------------------------------------------------------------------- defaults
class Defaults:
    """One default class: per-tag attribute dictionaries, chained to a parent class."""

    def __init__(self, parent=None):
        self.parent = parent
        self.tags = {}

    def get(self, tag):
        base = dict(self.parent.get(tag)) if self.parent else {}
        base.update(self.tags.get(tag, {}))
        return base


_ACT_TAGS = ("motor", "general", "position", "velocity")


def _parse_defaults(elem, parent, table):
    name = elem.get("class", "main")
    d = Defaults(parent)
    table[name] = d
    for ch in elem:
        if ch.tag == "default":
            _parse_defaults(ch, d, table)
        else:
            tag = ch.tag
            # actuator shortcuts share the 'general' default (MuJoCo stores one mjCDef actuator)
            if tag in _ACT_TAGS:
                tag = "general"
                attrs = _actuator_shortcut(ch.tag, dict(ch.attrib))
            else:
                attrs = dict(ch.attrib)
            d.tags.setdefault(tag, {}).update(attrs)
    return d


def _actuator_shortcut(tag, a):
    """Translate motor/position shortcuts into general-actuator attributes."""
    a = dict(a)
    if tag == "motor":
        a.setdefault("gainprm", "1")
        a.setdefault("biastype", "none")
        a.setdefault("dyntype", "none")
        a.setdefault("gaintype", "fixed")
    elif tag == "position":
        if "kp" in a:
            kp = float(a.pop("kp"))
            a["gainprm"] = f"{kp!r}"
            a["biasprm"] = f"0 {-kp!r} {-float(a.pop('kv', 0.0))!r}"
        a.setdefault("biastype", "affine")
        a.setdefault("gaintype", "fixed")
        a.setdefault("dyntype", "none")
    return a


# ---------------------------------------------------------------------------- geom inertia
def _geom_volume_inertia(gtype, size):
    """Volume and unit-density-normalised principal inertia (per unit mass) of a primitive."""
    if gtype == P.GEOM_SPHERE:
        r = size[0]
        vol = 4.0 / 3.0 * math.pi * r ** 3
        I = np.array([2 * r * r / 5] * 3)
    elif gtype == P.GEOM_CAPSULE:
        r, h = size[0], 2 * size[1]
        vol = math.pi * (r * r * h + 4.0 / 3.0 * r ** 3)
        sm = 4 * r / (4 * r + 3 * h)          # sphere mass fraction
        cm = 1 - sm
        ixy = cm * (3 * r * r + h * h) / 12 + 2 * sm * r * r / 5 + sm * h * (3 * r + 2 * h) / 8
        izz = cm * r * r / 2 + 2 * sm * r * r / 5
        I = np.array([ixy, ixy, izz])
    elif gtype == P.GEOM_CYLINDER:
        r, h = size[0], 2 * size[1]
        vol = math.pi * r * r * h
        I = np.array([(3 * r * r + h * h) / 12, (3 * r * r + h * h) / 12, r * r / 2])
    elif gtype == P.GEOM_BOX:
        a, b, c = size
        vol = 8 * a * b * c
        I = np.array([(b * b + c * c) / 3, (a * a + c * c) / 3, (a * a + b * b) / 3])
    elif gtype == P.GEOM_ELLIPSOID:
        a, b, c = size
        vol = 4.0 / 3.0 * math.pi * a * b * c
        I = np.array([(b * b + c * c) / 5, (a * a + c * c) / 5, (a * a + b * b) / 5])
    else:  # plane etc: massless
        vol, I = 0.0, np.zeros(3)
    return vol, I


def _rbound(gtype, size):
    if gtype == P.GEOM_SPHERE:
        return size[0]
    if gtype == P.GEOM_CAPSULE:
        return size[0] + size[1]
    if gtype == P.GEOM_CYLINDER:
        return math.hypot(size[0], size[1])
    if gtype == P.GEOM_BOX:
        return float(np.linalg.norm(size))
    if gtype == P.GEOM_ELLIPSOID:
        return max(size)
    return 0.0


_GEOM_TYPES = {"plane": P.GEOM_PLANE, "sphere": P.GEOM_SPHERE, "capsule": P.GEOM_CAPSULE,
               "ellipsoid": P.GEOM_ELLIPSOID, "cylinder": P.GEOM_CYLINDER, "box": P.GEOM_BOX,
               "mesh": P.GEOM_MESH}
_JNT_TYPES = {"free": P.JNT_FREE, "ball": P.JNT_BALL, "slide": P.JNT_SLIDE, "hinge": P.JNT_HINGE}
_INTEGRATORS = {"Euler": P.INT_EULER, "RK4": P.INT_RK4, "implicit": P.INT_IMPLICIT,
                "implicitfast": P.INT_IMPLICITFAST}
_DSBL = {"constraint": P.DSBL_CONSTRAINT, "equality": P.DSBL_EQUALITY,
         "frictionloss": P.DSBL_FRICTIONLOSS, "limit": P.DSBL_LIMIT, "contact": P.DSBL_CONTACT,
         "passive": P.DSBL_PASSIVE, "gravity": P.DSBL_GRAVITY, "clampctrl": P.DSBL_CLAMPCTRL,
         "warmstart": P.DSBL_WARMSTART, "filterparent": P.DSBL_FILTERPARENT,
         "actuation": P.DSBL_ACTUATION, "refsafe": P.DSBL_REFSAFE, "sensor": P.DSBL_SENSOR,
         "midphase": P.DSBL_MIDPHASE, "eulerdamp": P.DSBL_EULERDAMP}
_SENSORS = {  # tag -> (type, dim, how the object is named)
    "user": (P.SENS_USER, None, None),
    "touch": (P.SENS_TOUCH, 1, "site"), "accelerometer": (P.SENS_ACCELEROMETER, 3, "site"),
    "jointpos": (P.SENS_JOINTPOS, 1, "joint"), "jointvel": (P.SENS_JOINTVEL, 1, "joint"),
    "framepos": (P.SENS_FRAMEPOS, 3, "obj"), "framequat": (P.SENS_FRAMEQUAT, 4, "obj"),
    "framexaxis": (P.SENS_FRAMEXAXIS, 3, "obj"), "frameyaxis": (P.SENS_FRAMEYAXIS, 3, "obj"),
    "framezaxis": (P.SENS_FRAMEZAXIS, 3, "obj"), "framelinvel": (P.SENS_FRAMELINVEL, 3, "obj"),
    "frameangvel": (P.SENS_FRAMEANGVEL, 3, "obj"),
    "subtreecom": (P.SENS_SUBTREECOM, 3, "body"), "subtreelinvel": (P.SENS_SUBTREELINVEL, 3, "body"),
    "subtreeangmom": (P.SENS_SUBTREEANGMOM, 3, "body"),
}
_OBJTYPES = {"body": P.OBJ_BODY, "xbody": P.OBJ_XBODY, "geom": P.OBJ_GEOM, "site": P.OBJ_SITE,
             "joint": P.OBJ_JOINT}


# ---------------------------------------------------------------------------- compiler
class _Names:
    def __init__(self):
        self.buf = bytearray()

    def add(self, s):
        adr = len(self.buf)
        self.buf += (s or "").encode() + b"\0"
        return adr


class Compiler:
    def __init__(self, root):
        self.root = root
        comp = root.find("compiler")
        self.radian = comp is not None and comp.get("angle", "degree") == "radian"
        self.eulerseq = comp.get("eulerseq", "xyz") if comp is not None else "xyz"
        self.autolimits = (comp.get("autolimits", "true") == "true") if comp is not None else True
        self.defaults = {}
        main = Defaults(None)
        self.defaults["main"] = main
        for d in root.findall("default"):
            # every top-level <default> contributes to the main class
            for ch in d:
                if ch.tag == "default":
                    _parse_defaults(ch, main, self.defaults)
                else:
                    tag = "general" if ch.tag in _ACT_TAGS else ch.tag
                    attrs = _actuator_shortcut(ch.tag, dict(ch.attrib)) if ch.tag in _ACT_TAGS \
                        else dict(ch.attrib)
                    main.tags.setdefault(tag, {}).update(attrs)
        self.bodies, self.joints, self.geoms, self.sites = [], [], [], []

    # -- attribute helpers
    def ang(self, x):
        return x if self.radian else x * math.pi / 180.0

    def attrs(self, elem, tag, childclass):
        cls = elem.get("class") or childclass or "main"
        a = self.defaults[cls].get(tag)
        a.update(elem.attrib)
        return a

    def orientation(self, a):
        if "quat" in a:
            q = np.array(_floats(a["quat"], 4))
            return q / np.linalg.norm(q)
        if "euler" in a:
            return euler2quat([self.ang(x) for x in _floats(a["euler"], 3)], self.eulerseq)
        if "axisangle" in a:
            v = _floats(a["axisangle"], 4)
            return axisangle2quat(v[:3], self.ang(v[3]))
        if "zaxis" in a:
            return z2quat(_floats(a["zaxis"], 3))
        if "xyaxes" in a:
            v = _floats(a["xyaxes"], 6)
            x = np.array(v[:3]); x /= np.linalg.norm(x)
            y = np.array(v[3:]); y -= x * np.dot(x, y); y /= np.linalg.norm(y)
            return mat2quat(np.stack([x, y, np.cross(x, y)], axis=1))
        return np.array([1.0, 0, 0, 0])

    # -- tree walk
    def walk_body(self, elem, parent, childclass):
        bid = len(self.bodies)
        if elem.tag == "worldbody":
            b = dict(name="world", parent=0, pos=np.zeros(3), quat=np.array([1.0, 0, 0, 0]),
                     mocap=False, inertial=None, joints=[], geoms=[])
        else:
            childclass = elem.get("childclass", childclass)
            b = dict(name=elem.get("name", ""), parent=parent,
                     pos=np.array(_floats(elem.get("pos", "0 0 0"), 3)),
                     quat=self.orientation(elem.attrib), mocap=elem.get("mocap") == "true",
                     inertial=None, joints=[], geoms=[])
        self.bodies.append(b)
        # MuJoCo orders each body's own elements by type; children bodies are depth-first
        for ch in elem:
            if ch.tag == "inertial":
                a = dict(ch.attrib)
                ine = dict(pos=np.array(_floats(a.get("pos", "0 0 0"), 3)), mass=float(a["mass"]),
                           quat=self.orientation(a))
                if "diaginertia" in a:
                    ine["diag"] = np.array(_floats(a["diaginertia"], 3))
                else:
                    f = _floats(a["fullinertia"], 6)
                    ine["full"] = np.array([[f[0], f[3], f[4]], [f[3], f[1], f[5]], [f[4], f[5], f[2]]])
                b["inertial"] = ine
            elif ch.tag in ("joint", "freejoint"):
                self.add_joint(ch, bid, childclass)
            elif ch.tag == "geom":
                self.add_geom(ch, bid, childclass)
            elif ch.tag == "site":
                self.add_site(ch, bid, childclass)
        for ch in elem:
            if ch.tag == "body":
                self.walk_body(ch, bid, childclass)

    def add_joint(self, ch, bid, childclass):
        if ch.tag == "freejoint":
            a = {"type": "free", "name": ch.get("name", "")}
            # freejoint ignores defaults (MuJoCo: no class, no damping/armature)
        else:
            a = self.attrs(ch, "joint", childclass)
        jt = _JNT_TYPES[a.get("type", "hinge")]
        is_ang = jt == P.JNT_HINGE
        rng = _floats(a.get("range", "0 0"), 2)
        if is_ang:
            rng = [self.ang(x) for x in rng]
        lim = a.get("limited", "auto")
        limited = (lim == "true") or (lim == "auto" and self.autolimits and rng[0] < rng[1])
        axis = np.array(_floats(a.get("axis", "0 0 1"), 3))
        if jt in (P.JNT_HINGE, P.JNT_SLIDE):
            axis = axis / np.linalg.norm(axis)
        ref = float(a.get("ref", 0.0))
        sref = float(a.get("springref", 0.0))
        j = dict(name=a.get("name", ""), type=jt, body=bid,
                 pos=np.array(_floats(a.get("pos", "0 0 0"), 3)), axis=axis,
                 range=rng, limited=limited, ref=self.ang(ref) if is_ang else ref,
                 springref=self.ang(sref) if is_ang else sref,
                 stiffness=float(a.get("stiffness", 0.0)), damping=float(a.get("damping", 0.0)),
                 armature=float(a.get("armature", 0.0)), frictionloss=float(a.get("frictionloss", 0.0)),
                 margin=float(a.get("margin", 0.0)),
                 solreflimit=_floats(a.get("solreflimit", "0.02 1"), 2),
                 solimplimit=_floats(a.get("solimplimit", "0.9 0.95 0.001 0.5 2")),
                 solreffriction=_floats(a.get("solreffriction", "0.02 1"), 2),
                 solimpfriction=_floats(a.get("solimpfriction", "0.9 0.95 0.001 0.5 2")))
        for k in ("solimplimit", "solimpfriction"):
            d = [0.9, 0.95, 0.001, 0.5, 2.0]
            d[:len(j[k])] = j[k]
            j[k] = d
        self.bodies[bid]["joints"].append(len(self.joints))
        self.joints.append(j)

    def add_geom(self, ch, bid, childclass):
        a = self.attrs(ch, "geom", childclass)
        gt = _GEOM_TYPES[a.get("type", "sphere")]
        if gt == P.GEOM_MESH:
            # visual-only meshes (contype=conaffinity=0, explicit <inertial> on the body) carry no
            # physics; they are dropped (documented in DESIGN.md)
            return
        size = _floats(a.get("size", "0"))
        size = (size + [0.0, 0.0, 0.0])[:3]
        pos = np.array(_floats(a.get("pos", "0 0 0"), 3))
        quat = self.orientation(a)
        if "fromto" in a:
            ft = np.array(_floats(a["fromto"], 6))
            vec = ft[:3] - ft[3:]
            half = np.linalg.norm(vec) / 2
            pos = (ft[:3] + ft[3:]) / 2
            quat = z2quat(vec)
            if gt in (P.GEOM_CAPSULE, P.GEOM_CYLINDER):
                size[1] = half
            else:
                size[2] = half
        vol, In = _geom_volume_inertia(gt, size)
        if "mass" in a:
            mass = float(a["mass"])
        else:
            mass = float(a.get("density", 1000.0)) * vol
        fr = [1.0, 0.005, 0.0001]
        f_in = _floats(a["friction"]) if "friction" in a else []
        fr[:len(f_in)] = f_in
        simp = [0.9, 0.95, 0.001, 0.5, 2.0]
        s_in = _floats(a["solimp"]) if "solimp" in a else []
        simp[:len(s_in)] = s_in
        g = dict(name=a.get("name", ""), type=gt, body=bid, size=np.array(size), pos=pos, quat=quat,
                 mass=mass, inertia=In * mass, friction=fr,
                 contype=int(a.get("contype", 1)), conaffinity=int(a.get("conaffinity", 1)),
                 condim=int(a.get("condim", 3)), priority=int(a.get("priority", 0)),
                 group=int(a.get("group", 0)), solmix=float(a.get("solmix", 1.0)),
                 solref=_floats(a.get("solref", "0.02 1"), 2), solimp=simp,
                 margin=float(a.get("margin", 0.0)), gap=float(a.get("gap", 0.0)))
        self.bodies[bid]["geoms"].append(len(self.geoms))
        self.geoms.append(g)

    def add_site(self, ch, bid, childclass):
        a = self.attrs(ch, "site", childclass)
        self.sites.append(dict(name=a.get("name", ""), body=bid,
                               pos=np.array(_floats(a.get("pos", "0 0 0"), 3)),
                               quat=self.orientation(a)))

    # -- body inertial frame from <inertial> or geoms
    def body_inertial(self, b):
        if b["inertial"] is not None:
            ine = b["inertial"]
            if "diag" in ine:
                return ine["mass"], ine["pos"], ine["quat"], ine["diag"]
            w, V = np.linalg.eigh(ine["full"])
            w, V = w[::-1], V[:, ::-1]
            if np.linalg.det(V) < 0:
                V[:, 2] = -V[:, 2]
            return ine["mass"], ine["pos"], mat2quat(V), w
        gs = [self.geoms[g] for g in b["geoms"]]
        massive = [g for g in gs if g["mass"] > 0]
        if not massive:
            return 0.0, np.zeros(3), np.array([1.0, 0, 0, 0]), np.zeros(3)
        if len(gs) == 1:
            g = gs[0]
            return g["mass"], g["pos"].copy(), g["quat"].copy(), g["inertia"].copy()
        mass = sum(g["mass"] for g in massive)
        com = sum(g["mass"] * g["pos"] for g in massive) / mass
        I = np.zeros((3, 3))
        for g in massive:
            R = quat2mat(g["quat"])
            d = g["pos"] - com
            I += R @ np.diag(g["inertia"]) @ R.T + g["mass"] * (np.dot(d, d) * np.eye(3) - np.outer(d, d))
        w, V = np.linalg.eigh(I)
        w, V = w[::-1], V[:, ::-1]
        if np.linalg.det(V) < 0:
            V[:, 2] = -V[:, 2]
        return mass, com, mat2quat(V), w

    # -- actuators / sensors / custom / keys are read straight from the XML sections
    def compile(self):
        root = self.root
        for wb in root.findall("worldbody"):
            if not self.bodies:
                self.walk_body(wb, 0, None)
            else:  # additional <worldbody> sections append to the world body
                for ch in wb:
                    if ch.tag == "geom":
                        self.add_geom(ch, 0, None)
                    elif ch.tag == "site":
                        self.add_site(ch, 0, None)
                for ch in wb:
                    if ch.tag == "body":
                        self.walk_body(ch, 0, None)
        # geoms/sites must be grouped by body id (MuJoCo numbering): stable sort
        order = sorted(range(len(self.geoms)), key=lambda i: self.geoms[i]["body"])
        remap = {old: new for new, old in enumerate(order)}
        self.geoms = [self.geoms[i] for i in order]
        for b in self.bodies:
            b["geoms"] = sorted(remap[g] for g in b["geoms"])
        self.sites.sort(key=lambda s: s["body"])
        return self.assemble()

    def assemble(self):
        root, B, J, G, S = self.root, self.bodies, self.joints, self.geoms, self.sites
        names = _Names()
        v = {}
        nbody = len(B)
        # ---- options
        opt = root.find("option")
        oa = dict(opt.attrib) if opt is not None else {}
        dis = 0
        if opt is not None:
            for fl in opt.findall("flag"):
                for k, val in fl.attrib.items():
                    if k in _DSBL and val == "disable":
                        dis |= _DSBL[k]
        v.update(opt_timestep=float(oa.get("timestep", 0.002)), opt_impratio=float(oa.get("impratio", 1.0)),
                 opt_tolerance=float(oa.get("tolerance", 1e-8)), opt_ls_tolerance=float(oa.get("ls_tolerance", 0.01)),
                 opt_density=float(oa.get("density", 0.0)), opt_viscosity=float(oa.get("viscosity", 0.0)),
                 opt_integrator=_INTEGRATORS[oa.get("integrator", "Euler")],
                 opt_cone=P.CONE_ELLIPTIC if oa.get("cone", "pyramidal") == "elliptic" else P.CONE_PYRAMIDAL,
                 opt_solver={"PGS": P.SOL_PGS, "CG": P.SOL_CG, "Newton": P.SOL_NEWTON}[oa.get("solver", "Newton")],
                 opt_iterations=int(oa.get("iterations", 100)), opt_ls_iterations=int(oa.get("ls_iterations", 50)),
                 opt_disableflags=dis, opt_enableflags=0,
                 opt_gravity=_floats(oa.get("gravity", "0 0 -9.81"), 3), opt_wind=_floats(oa.get("wind", "0 0 0"), 3))
        # ---- bodies, joints, dofs
        qpos0, qspring = [], []
        jnt_qposadr, jnt_dofadr = [], []
        dof_body, dof_jnt, dof_parent, dof_arm, dof_damp, dof_fl, dof_sref, dof_simp = [], [], [], [], [], [], [], []
        body_jntadr, body_dofadr, body_dofnum, body_lastdof = [], [], [], []
        for bi, b in enumerate(B):
            par_last = body_lastdof[b["parent"]] if bi > 0 else -1
            body_jntadr.append(b["joints"][0] if b["joints"] else -1)
            body_dofadr.append(len(dof_body) if b["joints"] else -1)
            last = par_last
            nd0 = len(dof_body)
            for ji in b["joints"]:
                j = J[ji]
                jnt_qposadr.append(len(qpos0)); jnt_dofadr.append(len(dof_body))
                if j["type"] == P.JNT_FREE:
                    q = list(b["pos"]) + list(b["quat"]); nd = 6
                elif j["type"] == P.JNT_BALL:
                    q = [1.0, 0, 0, 0]; nd = 3
                else:
                    q = [j["ref"]]; nd = 1
                qpos0 += q
                qspring += q if j["type"] in (P.JNT_FREE, P.JNT_BALL) else [j["springref"]]
                for _ in range(nd):
                    dof_body.append(bi); dof_jnt.append(ji); dof_parent.append(last)
                    last = len(dof_body) - 1
                    free = j["type"] == P.JNT_FREE
                    dof_arm.append(0.0 if free else j["armature"]); dof_damp.append(0.0 if free else j["damping"])
                    dof_fl.append(0.0 if free else j["frictionloss"])
                    dof_sref.append(j["solreffriction"]); dof_simp.append(j["solimpfriction"])
            body_dofnum.append(len(dof_body) - nd0)
            body_lastdof.append(last)
        nq, nv, njnt = len(qpos0), len(dof_body), len(J)
        inert = [self.body_inertial(b) for b in B]
        inert[0] = (0.0, np.zeros(3), np.array([1.0, 0, 0, 0]), np.zeros(3))  # world is massless
        rootid = []
        for bi, b in enumerate(B):
            r = bi
            while r != 0 and B[r]["parent"] != 0:
                r = B[r]["parent"]
            rootid.append(r)
        subm = [m for m, _, _, _ in inert]
        for bi in range(nbody - 1, 0, -1):
            subm[B[bi]["parent"]] += subm[bi]
        mocapid, nmocap = [], 0
        for b in B:
            if b["mocap"]:
                mocapid.append(nmocap); nmocap += 1
            else:
                mocapid.append(-1)
        geomadr = [(b["geoms"][0] if b["geoms"] else -1) for b in B]
        v.update(nq=nq, nv=nv, nbody=nbody, njnt=njnt, ngeom=len(G), nsite=len(S), nmocap=nmocap,
                 qpos0=qpos0, qpos_spring=qspring,
                 body_parentid=[b["parent"] for b in B], body_rootid=rootid,
                 body_jntnum=[len(b["joints"]) for b in B], body_jntadr=body_jntadr,
                 body_dofnum=body_dofnum, body_dofadr=body_dofadr,
                 body_geomnum=[len(b["geoms"]) for b in B], body_geomadr=geomadr, body_mocapid=mocapid,
                 body_pos=[b["pos"] for b in B], body_quat=[b["quat"] for b in B],
                 body_ipos=[i[1] for i in inert], body_iquat=[i[2] for i in inert],
                 body_mass=[i[0] for i in inert], body_subtreemass=subm,
                 body_inertia=[i[3] for i in inert],
                 jnt_type=[j["type"] for j in J], jnt_qposadr=jnt_qposadr, jnt_dofadr=jnt_dofadr,
                 jnt_bodyid=[j["body"] for j in J], jnt_limited=[int(j["limited"]) for j in J],
                 jnt_pos=[j["pos"] for j in J], jnt_axis=[j["axis"] for j in J],
                 jnt_stiffness=[j["stiffness"] for j in J], jnt_range=[j["range"] for j in J],
                 jnt_margin=[j["margin"] for j in J], jnt_solref=[j["solreflimit"] for j in J],
                 jnt_solimp=[j["solimplimit"] for j in J],
                 dof_bodyid=dof_body, dof_jntid=dof_jnt, dof_parentid=dof_parent, dof_armature=dof_arm,
                 dof_damping=dof_damp, dof_frictionloss=dof_fl, dof_solref=dof_sref, dof_solimp=dof_simp,
                 geom_type=[g["type"] for g in G], geom_contype=[g["contype"] for g in G],
                 geom_conaffinity=[g["conaffinity"] for g in G], geom_condim=[g["condim"] for g in G],
                 geom_bodyid=[g["body"] for g in G], geom_priority=[g["priority"] for g in G],
                 geom_group=[g["group"] for g in G], geom_size=[g["size"] for g in G],
                 geom_pos=[g["pos"] for g in G], geom_quat=[g["quat"] for g in G],
                 geom_friction=[g["friction"] for g in G], geom_solmix=[g["solmix"] for g in G],
                 geom_solref=[g["solref"] for g in G], geom_solimp=[g["solimp"] for g in G],
                 geom_margin=[g["margin"] for g in G], geom_gap=[g["gap"] for g in G],
                 geom_rbound=[_rbound(g["type"], g["size"]) for g in G],
                 site_bodyid=[s["body"] for s in S], site_pos=[s["pos"] for s in S],
                 site_quat=[s["quat"] for s in S])
        v["name_bodyadr"] = [names.add(b["name"]) for b in B]
        v["name_jntadr"] = [names.add(j["name"]) for j in J]
        v["name_geomadr"] = [names.add(g["name"]) for g in G]
        v["name_siteadr"] = [names.add(s["name"]) for s in S]
        self._assemble_rest(v, names)
        return v

    def _find(self, lst, name, what):
        for i, x in enumerate(lst):
            if x["name"] == name:
                return i
        raise KeyError(f"{what} '{name}' not found")

    def _assemble_rest(self, v, names):
        root, B, J, G, S = self.root, self.bodies, self.joints, self.geoms, self.sites
        # ---- actuators
        A = []
        for sec in root.findall("actuator"):
            for ch in sec:
                cls = ch.get("class") or "main"
                a = self.defaults[cls].get("general")
                a.update(_actuator_shortcut(ch.tag, dict(ch.attrib)))
                A.append(a)
        nu, na = len(A), 0
        rows = {k: [] for k in ("dyntype", "gaintype", "biastype", "trnid", "actadr", "actnum", "ctrllimited",
                                "forcelimited", "actlimited", "dynprm", "gainprm", "biasprm", "ctrlrange",
                                "forcerange", "actrange", "gear")}

        def pad3(s, d):
            x = _floats(s) if s is not None else []
            out = list(d); out[:min(3, len(x))] = x[:3]
            return out

        def lim(a, key, rkey):
            r = _floats(a.get(rkey, "0 0"), 2)
            f = a.get(key, "auto")
            return int(f == "true" or (f == "auto" and self.autolimits and r[0] < r[1])), r

        for a in A:
            dyn = {"none": P.DYN_NONE, "integrator": P.DYN_INTEGRATOR, "filter": P.DYN_FILTER,
                   "filterexact": P.DYN_FILTEREXACT}[a.get("dyntype", "none")]
            rows["dyntype"].append(dyn)
            rows["gaintype"].append({"fixed": P.GAIN_FIXED, "affine": P.GAIN_AFFINE}[a.get("gaintype", "fixed")])
            rows["biastype"].append({"none": P.BIAS_NONE, "affine": P.BIAS_AFFINE}[a.get("biastype", "none")])
            rows["trnid"].append([self._find(J, a["joint"], "joint"), -1])
            n = 0 if dyn == P.DYN_NONE else 1
            rows["actadr"].append(na if n else -1); rows["actnum"].append(n); na += n
            cl, cr = lim(a, "ctrllimited", "ctrlrange"); rows["ctrllimited"].append(cl); rows["ctrlrange"].append(cr)
            fl, fr = lim(a, "forcelimited", "forcerange"); rows["forcelimited"].append(fl); rows["forcerange"].append(fr)
            al, ar = lim(a, "actlimited", "actrange"); rows["actlimited"].append(al); rows["actrange"].append(ar)
            rows["dynprm"].append(pad3(a.get("dynprm"), [1, 0, 0]))
            rows["gainprm"].append(pad3(a.get("gainprm"), [1, 0, 0]))
            rows["biasprm"].append(pad3(a.get("biasprm"), [0, 0, 0]))
            rows["gear"].append(_floats(a.get("gear", "1"))[0])
        v.update(nu=nu, na=na, actuator_trntype=[P.TRN_JOINT] * nu, **{"actuator_" + k: r for k, r in rows.items()})
        v["name_actuatoradr"] = [names.add(a.get("name", "")) for a in A]
        # ---- sensors
        SN = []
        for sec in root.findall("sensor"):
            for ch in sec:
                if ch.tag not in _SENSORS:
                    raise ValueError(f"unsupported sensor <{ch.tag}>")
                st, dim, how = _SENSORS[ch.tag]
                s = dict(name=ch.get("name", ""), type=st, dim=dim, objtype=P.OBJ_UNKNOWN, objid=-1, user=[])
                if ch.tag == "user":
                    s["dim"] = int(ch.get("dim"))
                    s["user"] = _floats(ch.get("user", ""))
                elif how == "site":
                    s["objtype"], s["objid"] = P.OBJ_SITE, self._find(S, ch.get("site"), "site")
                elif how == "joint":
                    s["objtype"], s["objid"] = P.OBJ_JOINT, self._find(J, ch.get("joint"), "joint")
                elif how == "body":
                    s["objtype"], s["objid"] = P.OBJ_BODY, self._find(B, ch.get("body"), "body")
                else:
                    ot = ch.get("objtype")
                    lst = {"body": B, "xbody": B, "geom": G, "site": S}[ot]
                    s["objtype"], s["objid"] = _OBJTYPES[ot], self._find(lst, ch.get("objname"), ot)
                SN.append(s)
        nus = max([len(s["user"]) for s in SN] + [0])
        adr, sadr = 0, []
        for s in SN:
            sadr.append(adr); adr += s["dim"]
        v.update(nsensor=len(SN), nsensordata=adr, nuser_sensor=nus,
                 sensor_type=[s["type"] for s in SN], sensor_objtype=[s["objtype"] for s in SN],
                 sensor_objid=[s["objid"] for s in SN], sensor_dim=[s["dim"] for s in SN], sensor_adr=sadr,
                 sensor_user=[(s["user"] + [0.0] * nus)[:nus] for s in SN])
        v["name_sensoradr"] = [names.add(s["name"]) for s in SN]
        # ---- custom numeric / text
        nums, texts = [], []
        for sec in root.findall("custom"):
            for ch in sec:
                if ch.tag == "numeric":
                    d = _floats(ch.get("data"))
                    size = int(ch.get("size", len(d)))
                    nums.append((ch.get("name"), (d + [0.0] * size)[:size]))
                elif ch.tag == "text":
                    texts.append((ch.get("name"), ch.get("data")))
        v.update(nnumeric=len(nums), nnumericdata=sum(len(d) for _, d in nums),
                 numeric_adr=list(np.cumsum([0] + [len(d) for _, d in nums])[:-1]),
                 numeric_size=[len(d) for _, d in nums], numeric_data=[x for _, d in nums for x in d])
        tb = bytearray(); tadr = []
        for _, t in texts:
            tadr.append(len(tb)); tb += t.encode() + b"\0"
        v.update(ntext=len(texts), ntextdata=len(tb), text_adr=tadr, text_size=[len(t) + 1 for _, t in texts],
                 text_data=bytes(tb))
        v["name_numericadr"] = [names.add(n) for n, _ in nums]
        v["name_textadr"] = [names.add(n) for n, _ in texts]
        # ---- keyframes
        K = [k for sec in root.findall("keyframe") for k in sec.findall("key")]
        nq, nv = v["nq"], v["nv"]
        v.update(nkey=len(K),
                 key_qpos=[_floats(k.get("qpos")) if k.get("qpos") else list(v["qpos0"]) for k in K],
                 key_qvel=[_floats(k.get("qvel")) if k.get("qvel") else [0.0] * nv for k in K],
                 key_act=[_floats(k.get("act")) if k.get("act") else [0.0] * na for k in K],
                 key_ctrl=[_floats(k.get("ctrl")) if k.get("ctrl") else [0.0] * nu for k in K])
        v["name_keyadr"] = [names.add(k.get("name", "")) for k in K]
        size = root.find("size")
        v["nuserdata"] = int(size.get("nuserdata", 0)) if size is not None else 0
        v["names"] = bytes(names.buf)
        v["nnames"] = len(names.buf)


# ---------------------------------------------------------------------------- mj_setConst subset
def kinematics0(v):
    """World frames of all bodies at qpos0 (joints at reference => pure static chain)."""
    nb = v["nbody"]
    xpos, xmat = np.zeros((nb, 3)), np.zeros((nb, 3, 3))
    xquat = np.zeros((nb, 4))
    xquat[0] = [1, 0, 0, 0]; xmat[0] = np.eye(3)
    bp, bq, par = np.asarray(v["body_pos"], float), np.asarray(v["body_quat"], float), v["body_parentid"]
    for i in range(1, nb):
        p = par[i]
        xpos[i] = xpos[p] + xmat[p] @ bp[i]
        xquat[i] = quat_mul(xquat[p], bq[i]); xquat[i] /= np.linalg.norm(xquat[i])
        xmat[i] = quat2mat(xquat[i])
    return xpos, xquat, xmat


def mass_matrix0(v):
    """Joint-space inertia at qpos0 from body Jacobians: M = sum_b m Jp'Jp + Jr' I Jr + armature.

    Deliberately NOT the composite-rigid-body recursion the engines use, so it doubles as an
    independent check of them.  Returns (M, Jp_com[b], Jr[b])."""
    nb, nv = v["nbody"], v["nv"]
    xpos, xquat, xmat = kinematics0(v)
    ipos = np.asarray(v["body_ipos"], float); iquat = np.asarray(v["body_iquat"], float)
    xipos = np.array([xpos[i] + xmat[i] @ ipos[i] for i in range(nb)])
    ximat = np.array([quat2mat(quat_mul(xquat[i], iquat[i])) for i in range(nb)])
    # per-dof axis/anchor in world
    jt, jb = v["jnt_type"], v["jnt_bodyid"]
    jpos, jax = np.asarray(v["jnt_pos"], float).reshape(-1, 3), np.asarray(v["jnt_axis"], float).reshape(-1, 3)
    dof = []  # (kind, axis, anchor) kind: 'lin' or 'rot'
    for d in range(nv):
        j = v["dof_jntid"][d]; b = jb[j]; k = d - v["jnt_dofadr"][j]
        if jt[j] == P.JNT_FREE:
            dof.append(("lin", np.eye(3)[k], None) if k < 3 else ("rot", xmat[b][:, k - 3], xpos[b]))
        elif jt[j] == P.JNT_BALL:
            dof.append(("rot", xmat[b][:, k], xpos[b] + xmat[b] @ jpos[j]))
        elif jt[j] == P.JNT_SLIDE:
            dof.append(("lin", xmat[b] @ jax[j], None))
        else:
            dof.append(("rot", xmat[b] @ jax[j], xpos[b] + xmat[b] @ jpos[j]))
    Jp = np.zeros((nb, 3, nv)); Jr = np.zeros((nb, 3, nv))
    for b in range(1, nb):
        # dofs affecting body b: walk up the dof chain from the body's last dof
        bb = b
        while bb != 0 and v["body_dofnum"][bb] == 0:
            bb = v["body_parentid"][bb]
        if bb == 0:
            continue
        d = v["body_dofadr"][bb] + v["body_dofnum"][bb] - 1
        while d >= 0:
            kind, ax, anc = dof[d]
            if kind == "lin":
                Jp[b][:, d] = ax
            else:
                Jp[b][:, d] = np.cross(ax, xipos[b] - anc); Jr[b][:, d] = ax
            d = v["dof_parentid"][d]
    M = np.diag(np.asarray(v["dof_armature"], float)) if nv else np.zeros((0, 0))
    mass, inertia = np.asarray(v["body_mass"], float), np.asarray(v["body_inertia"], float)
    for b in range(1, nb):
        Iw = ximat[b] @ np.diag(inertia[b]) @ ximat[b].T
        M = M + mass[b] * Jp[b].T @ Jp[b] + Jr[b].T @ Iw @ Jr[b]
    return M, Jp, Jr


def set_const(v):
    nb, nv = v["nbody"], v["nv"]
    M, Jp, Jr = mass_matrix0(v)
    v["stat_meaninertia"] = float(np.trace(M) / nv) if nv else 1.0
    Minv = np.linalg.inv(M) if nv else M
    dinv = np.zeros(nv)
    for j in range(v["njnt"]):
        a = v["jnt_dofadr"][j]
        t = v["jnt_type"][j]
        if t == P.JNT_FREE:
            dinv[a:a + 3] = np.mean(np.diag(Minv)[a:a + 3]); dinv[a + 3:a + 6] = np.mean(np.diag(Minv)[a + 3:a + 6])
        elif t == P.JNT_BALL:
            dinv[a:a + 3] = np.mean(np.diag(Minv)[a:a + 3])
        else:
            dinv[a] = Minv[a, a]
    v["dof_invweight0"] = dinv
    biw = np.zeros((nb, 2))
    for b in range(1, nb):
        if np.any(Jp[b]) or np.any(Jr[b]):
            biw[b, 0] = np.trace(Jp[b] @ Minv @ Jp[b].T) / 3
            biw[b, 1] = np.trace(Jr[b] @ Minv @ Jr[b].T) / 3
    v["body_invweight0"] = biw


def compile_xml(text: str) -> PsModel:
    root = ET.fromstring(text)
    v = Compiler(root).compile()
    set_const(v)
    return PsModel(v)


def load(path: str) -> PsModel:
    with open(path) as f:
        return compile_xml(f.read())
