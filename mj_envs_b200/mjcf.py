"""MJCF model compiler for the Adroit hand-manipulation scenes.

Flattens the MJCF feature subset used by the reference's six XML files
(hand_manipulation_suite/assets/DAPG_{assets,Adroit,hammer,door,pen,relocate}.xml,
enumerated in SURVEY.md App. A.6) into structure-of-arrays numpy buffers
(`Model`), computes the compile-time constants the constraint solver needs
(dof/body/tendon `invweight0`, mean inertia, bounding radii), builds the static
candidate collision-pair list with MuJoCo's contype/conaffinity/parent/weld
filter semantics and serialises everything to a
self-describing binary blob that the C oracle and the CUDA library both read.

This is host-side, run-once code: nothing here is on the step hot path.
"""
from __future__ import annotations

import os
import struct
import xml.etree.ElementTree as ET
from dataclasses import dataclass, field

import numpy as np

# geom type codes follow MuJoCo's ordering (normal points from lower to higher type)
GEOM_PLANE, GEOM_SPHERE, GEOM_CAPSULE, GEOM_CYLINDER, GEOM_BOX = 0, 2, 3, 5, 6
_GEOM_TYPES = {"plane": GEOM_PLANE, "sphere": GEOM_SPHERE, "capsule": GEOM_CAPSULE,
               "cylinder": GEOM_CYLINDER, "box": GEOM_BOX}
JNT_SLIDE, JNT_HINGE = 2, 3

TASK_IDS = {"hammer": 0, "door": 1, "pen": 2, "relocate": 3}


# --------------------------------------------------------------------------- math
def _vec(s, n=None, default=None):
    if s is None:
        return None if default is None else np.array(default, dtype=np.float64)
    v = np.array([float(x) for x in s.split()], dtype=np.float64)
    if n is not None and v.size < n:
        v = np.concatenate([v, np.zeros(n - v.size)])
    return v


def quat_mul(a, b):
    return np.array([
        a[0] * b[0] - a[1] * b[1] - a[2] * b[2] - a[3] * b[3],
        a[0] * b[1] + a[1] * b[0] + a[2] * b[3] - a[3] * b[2],
        a[0] * b[2] - a[1] * b[3] + a[2] * b[0] + a[3] * b[1],
        a[0] * b[3] + a[1] * b[2] - a[2] * b[1] + a[3] * b[0]])


def quat_to_mat(q):
    w, x, y, z = q
    return np.array([
        [w * w + x * x - y * y - z * z, 2 * (x * y - w * z), 2 * (x * z + w * y)],
        [2 * (x * y + w * z), w * w - x * x + y * y - z * z, 2 * (y * z - w * x)],
        [2 * (x * z - w * y), 2 * (y * z + w * x), w * w - x * x - y * y + z * z]])


def mat_to_quat(R):
    t = np.trace(R)
    if t > 0:
        s = np.sqrt(t + 1.0) * 2
        q = np.array([0.25 * s, (R[2, 1] - R[1, 2]) / s, (R[0, 2] - R[2, 0]) / s, (R[1, 0] - R[0, 1]) / s])
    elif R[0, 0] > R[1, 1] and R[0, 0] > R[2, 2]:
        s = np.sqrt(1.0 + R[0, 0] - R[1, 1] - R[2, 2]) * 2
        q = np.array([(R[2, 1] - R[1, 2]) / s, 0.25 * s, (R[0, 1] + R[1, 0]) / s, (R[0, 2] + R[2, 0]) / s])
    elif R[1, 1] > R[2, 2]:
        s = np.sqrt(1.0 + R[1, 1] - R[0, 0] - R[2, 2]) * 2
        q = np.array([(R[0, 2] - R[2, 0]) / s, (R[0, 1] + R[1, 0]) / s, 0.25 * s, (R[1, 2] + R[2, 1]) / s])
    else:
        s = np.sqrt(1.0 + R[2, 2] - R[0, 0] - R[1, 1]) * 2
        q = np.array([(R[1, 0] - R[0, 1]) / s, (R[0, 2] + R[2, 0]) / s, (R[1, 2] + R[2, 1]) / s, 0.25 * s])
    return q / np.linalg.norm(q)


def axis_angle_quat(axis, ang):
    return np.concatenate([[np.cos(ang / 2)], np.sin(ang / 2) * np.asarray(axis, dtype=np.float64)])


def euler_to_quat(e):
    """compiler eulerseq="xyz" (MuJoCo default): intrinsic rotations x, then y, then z."""
    q = np.array([1.0, 0, 0, 0])
    for i in range(3):
        ax = np.zeros(3)
        ax[i] = 1.0
        q = quat_mul(q, axis_angle_quat(ax, e[i]))
    return q


def _orient(attr):
    if "quat" in attr:
        q = _vec(attr["quat"])
        return q / np.linalg.norm(q)
    if "euler" in attr:
        return euler_to_quat(_vec(attr["euler"]))
    return np.array([1.0, 0, 0, 0])


# --------------------------------------------------------------------------- model
@dataclass
class Model:
    """Flattened scene. All arrays are numpy; float64 / int32."""
    task: str = ""
    names: dict = field(default_factory=dict)     # kind -> list of names
    A: dict = field(default_factory=dict)         # array name -> ndarray (the blob payload)

    def __getattr__(self, k):
        A = object.__getattribute__(self, "A")
        if k in A:
            return A[k]
        raise AttributeError(k)

    def name2id(self, kind, name):
        return self.names[kind].index(name)

    def with_variation(self, variation_type):
        """Copy of a hammer scene whose reset draws the reference's `variation_type` (hammer_v0.py:110-129):
        'mass' (body_mass[Object] ~ U(0.05, 2.5)), 'pos' (geom_pos x of head / neck), 'size' (geom_size[head][0:2]).
        The drawn values become per-env parameters par[3], par[4]; nothing is recompiled (inertia tensor,
        invweight0 and geom_rbound stay stale, as in the reference -- SURVEY.md App. A.8 Q9)."""
        codes = {"mass": 1, "pos": 2, "size": 3}
        if variation_type not in codes or self.task != "hammer":
            raise Exception(f"Unsupported variation type {variation_type}")   # hammer_v0.py:129
        A = {k: v.copy() for k, v in self.A.items()}
        head, neck, obj = self.name2id("geom", "head"), self.name2id("geom", "neck"), self.name2id("body", "Object")
        A["task_int"][8:12] = [codes[variation_type], obj, head, neck]
        A["task_int"][12:14] = [A["render_geom_index"][head], A["render_geom_index"][neck]]   # their rows of the render table
        par0 = np.zeros(5)
        par0[:3] = A["par0"][:3]
        if variation_type == "mass":
            par0[3] = A["body_mass"][obj]
        elif variation_type == "pos":
            par0[3], par0[4] = A["geom_pos"].reshape(-1, 3)[head, 0], A["geom_pos"].reshape(-1, 3)[neck, 0]
        else:
            par0[3], par0[4] = A["geom_size"].reshape(-1, 3)[head, 0], A["geom_size"].reshape(-1, 3)[head, 1]
        A["par0"] = par0
        A["sizes"][7] = 5
        return Model(task=self.task, names=self.names, A=A)

    # sizes
    @property
    def nv(self): return int(self.A["sizes"][0])
    @property
    def nbody(self): return int(self.A["sizes"][1])
    @property
    def ngeom(self): return int(self.A["sizes"][2])
    @property
    def nsite(self): return int(self.A["sizes"][3])
    @property
    def nu(self): return int(self.A["sizes"][4])
    @property
    def ntendon(self): return int(self.A["sizes"][5])
    @property
    def npair(self): return int(self.A["sizes"][6])
    @property
    def npar(self): return int(self.A["sizes"][7])
    @property
    def obs_dim(self): return int(self.A["sizes"][8])
    @property
    def frame_skip(self): return int(self.A["sizes"][9])

    # ------------------------------------------------------------------ blob io
    MAGIC = b"ADRB200\x01"

    def to_blob(self) -> bytes:
        """Self-describing blob: MAGIC, nfields, then per field
        (name[32], dtype u32 (0=int32,1=float64), count u32, byte offset u64), then payload (8-byte aligned)."""
        fields = []
        payload = bytearray()
        for name in sorted(self.A):
            arr = np.ascontiguousarray(self.A[name])
            if arr.dtype.kind in "iub":
                arr = arr.astype(np.int32)
                dt = 0
            else:
                arr = arr.astype(np.float64)
                dt = 1
            while len(payload) % 8:
                payload.append(0)
            fields.append((name, dt, arr.size, len(payload)))
            payload += arr.tobytes()
        head = bytearray(self.MAGIC) + struct.pack("<Q", len(fields))
        table_len = len(fields) * 48
        base = len(head) + table_len
        base_pad = (8 - base % 8) % 8
        for name, dt, cnt, off in fields:
            nb = name.encode()
            assert len(nb) < 32
            head += nb + b"\0" * (32 - len(nb)) + struct.pack("<IIQ", dt, cnt, base + base_pad + off)
        meta = _encode_names(self.task, self.names)
        return bytes(head) + b"\0" * base_pad + bytes(payload) + meta

    @classmethod
    def from_blob(cls, blob: bytes) -> "Model":
        assert blob[:8] == cls.MAGIC, "bad magic"
        (nf,) = struct.unpack_from("<Q", blob, 8)
        A = {}
        end = 0
        for i in range(nf):
            o = 16 + 48 * i
            name = blob[o:o + 32].split(b"\0")[0].decode()
            dt, cnt, off = struct.unpack_from("<IIQ", blob, o + 32)
            dtype = np.int32 if dt == 0 else np.float64
            A[name] = np.frombuffer(blob, dtype=dtype, count=cnt, offset=off).copy()
            end = max(end, off + cnt * (4 if dt == 0 else 8))
        task, names = _decode_names(blob[end:])
        m = cls(task=task, names=names, A=A)
        m._reshape()
        return m

    def _reshape(self):
        A = self.A
        for k, w in _SHAPES.items():
            if k in A and w > 1:
                A[k] = A[k].reshape(-1, w)


_SHAPES = {"body_pos": 3, "body_quat": 4, "body_ipos": 3, "body_iquat": 4, "body_inertia": 3,
           "jnt_pos": 3, "jnt_axis": 3, "jnt_range": 2, "geom_pos": 3, "geom_quat": 4, "geom_size": 3,
           "geom_friction": 3, "site_pos": 3, "site_quat": 4, "site_size": 3, "act_ctrlrange": 2,
           "tendon_range": 2, "tendon_dof": 2, "tendon_coef": 2, "pair_geom": 2, "pair_friction": 5,
           "body_invweight0": 2, "body_par": 2, "act_bias": 3, "geom_rgba": 4, "cam_pos": 3, "cam_quat": 4, "light_pos": 3,
           "render_int": 4, "render_real": 19, "light_diffuse": 3}


def _encode_names(task, names):
    s = task + "\n" + "\n".join(k + ":" + ",".join(v) for k, v in sorted(names.items()))
    return b"NAMES\0" + s.encode()


def _decode_names(b):
    i = b.find(b"NAMES\0")
    if i < 0:
        return "", {}
    lines = b[i + 6:].decode().split("\n")
    names = {}
    for ln in lines[1:]:
        k, _, v = ln.partition(":")
        names[k] = v.split(",") if v else []
    return lines[0], names


# --------------------------------------------------------------------------- parsing
def _expand_includes(elem, basedir):
    out = []
    for ch in list(elem):
        if ch.tag == "include":
            inc = ET.parse(os.path.join(basedir, ch.attrib["file"])).getroot()
            _expand_includes(inc, basedir)
            out.extend(list(inc))
        else:
            _expand_includes(ch, basedir)
            out.append(ch)
    for ch in list(elem):
        elem.remove(ch)
    for ch in out:
        elem.append(ch)


class _Defaults:
    """Default classes with MuJoCo's copy-at-creation inheritance."""

    def __init__(self):
        self.cls = {"main": {}}

    def process(self, elem, parent="main"):
        # <default> with no class at top level is "main"
        name = elem.attrib.get("class", "main" if parent == "main" and "class" not in elem.attrib else None)
        if name not in self.cls:
            self.cls[name] = {k: dict(v) for k, v in self.cls[parent].items()}
        for ch in elem:
            if ch.tag == "default":
                self.process(ch, name)
            else:
                self.cls[name].setdefault(ch.tag, {}).update(ch.attrib)

    def resolve(self, tag, attrib, childclass=None):
        c = attrib.get("class", childclass or "main")
        out = dict(self.cls.get(c, {}).get(tag, {}))
        out.update(attrib)
        return out


def _geom_volume_inertia(gtype, size):
    """Volume and unit-density inertia diagonal (about the geom centre, geom frame)."""
    if gtype == GEOM_SPHERE:
        r = size[0]
        v = 4.0 / 3.0 * np.pi * r ** 3
        i = 0.4 * r * r
        return v, v * np.array([i, i, i])
    if gtype == GEOM_BOX:
        a, b, c = size
        v = 8 * a * b * c
        return v, v / 3.0 * np.array([b * b + c * c, a * a + c * c, a * a + b * b])
    if gtype == GEOM_CYLINDER:
        r, h = size[0], size[1]
        v = np.pi * r * r * 2 * h
        return v, v * np.array([(3 * r * r + 4 * h * h) / 12.0, (3 * r * r + 4 * h * h) / 12.0, r * r / 2.0])
    if gtype == GEOM_CAPSULE:
        r, h = size[0], size[1]
        vc = np.pi * r * r * 2 * h
        vs = 4.0 / 3.0 * np.pi * r ** 3
        # cylinder part + two hemispheres (parallel-axis for the caps)
        ixx = vc * (3 * r * r + 4 * h * h) / 12.0 + vs * (0.4 * r * r + h * h + 0.75 * r * h)
        izz = vc * r * r / 2.0 + vs * 0.4 * r * r
        return vc + vs, np.array([ixx, ixx, izz])
    return 0.0, np.zeros(3)


def load_mjcf(path: str, task: str | None = None) -> Model:
    """Parse + compile one of the Adroit task scenes."""
    basedir = os.path.dirname(os.path.abspath(path))
    root = ET.parse(path).getroot()
    _expand_includes(root, basedir)
    if task is None:
        for t in TASK_IDS:
            if t in os.path.basename(path).lower():
                task = t
    assert task in TASK_IDS, f"cannot infer Adroit task from {path}"

    dfl = _Defaults()
    opt = {"timestep": 0.002, "iterations": 100, "noslip_iterations": 0, "tolerance": 1e-8,
           "noslip_tolerance": 1e-6, "impratio": 1.0, "mpr_iterations": 50, "mpr_tolerance": 1e-6}
    bodies, joints, geoms, sites, cams, lights = [], [], [], [], [], []
    pairs_x, excludes, tendons, actuators, sensors = [], [], [], [], []
    materials = {}
    textures = {}

    bodies.append(dict(name="world", parent=-1, pos=np.zeros(3), quat=np.array([1.0, 0, 0, 0]),
                       inertial=None, mocap=False))

    def add_body(elem, bid):
        for ch in elem:
            if ch.tag == "body":
                b = dict(name=ch.attrib.get("name", f"body{len(bodies)}"), parent=bid,
                         pos=_vec(ch.attrib.get("pos"), 3, [0, 0, 0]), quat=_orient(ch.attrib),
                         inertial=None, mocap=ch.attrib.get("mocap", "false") == "true")
                bodies.append(b)
                add_body(ch, len(bodies) - 1)
            elif ch.tag == "inertial":
                a = ch.attrib
                bodies[bid]["inertial"] = dict(pos=_vec(a.get("pos"), 3, [0, 0, 0]), quat=_orient(a),
                                               mass=float(a["mass"]), diag=_vec(a["diaginertia"]))
            elif ch.tag == "joint":
                a = dfl.resolve("joint", ch.attrib)
                jt = a.get("type", "hinge")
                assert jt in ("hinge", "slide"), f"joint type {jt} is outside the Adroit subset"
                ax = _vec(a.get("axis"), 3, [0, 0, 1])
                joints.append(dict(name=a.get("name", f"jnt{len(joints)}"), body=bid,
                                   type=JNT_HINGE if jt == "hinge" else JNT_SLIDE,
                                   pos=_vec(a.get("pos"), 3, [0, 0, 0]), axis=ax / np.linalg.norm(ax),
                                   range=_vec(a.get("range"), 2, [0, 0]), limited=a.get("limited", "false") == "true",
                                   margin=float(a.get("margin", 0)), damping=float(a.get("damping", 0)),
                                   armature=float(a.get("armature", 0)), frictionloss=float(a.get("frictionloss", 0)),
                                   ref=float(a.get("ref", 0))))
            elif ch.tag == "geom":
                a = dfl.resolve("geom", ch.attrib)
                gt = a.get("type", "sphere")
                rgba = _vec(a.get("rgba")) if "rgba" in a else materials.get(a.get("material"), np.array([0.5, 0.5, 0.5, 1.0]))
                if gt == "mesh":
                    assert int(a.get("contype", 1)) == 0 and int(a.get("conaffinity", 1)) == 0, \
                        "collidable meshes are outside the Adroit subset"
                    continue  # visual only (assets absent)
                fr = _vec(a.get("friction"), None, [1, 0.005, 0.0001])
                fr = np.concatenate([fr, [1, 0.005, 0.0001][fr.size:]])
                geoms.append(dict(name=a.get("name", f"geom{len(geoms)}"), body=bid, type=_GEOM_TYPES[gt],
                                  pos=_vec(a.get("pos"), 3, [0, 0, 0]), quat=_orient(a),
                                  size=_vec(a.get("size"), 3, [0, 0, 0]), contype=int(a.get("contype", 1)),
                                  conaffinity=int(a.get("conaffinity", 1)), condim=int(a.get("condim", 3)),
                                  friction=fr, margin=float(a.get("margin", 0)), gap=float(a.get("gap", 0)),
                                  density=float(a.get("density", 1000)), group=int(a.get("group", 0)), rgba=rgba))
            elif ch.tag == "site":
                a = dfl.resolve("site", ch.attrib)
                sites.append(dict(name=a.get("name", f"site{len(sites)}"), body=bid,
                                  type=a.get("type", "sphere"), pos=_vec(a.get("pos"), 3, [0, 0, 0]),
                                  quat=_orient(a), size=_vec(a.get("size"), 3, [0.005, 0, 0]),
                                  rgba=_vec(a.get("rgba"), 4, [0.5, 0.5, 0.5, 1.0]), group=int(a.get("group", 0))))
            elif ch.tag == "camera":
                cams.append(dict(name=ch.attrib.get("name", ""), body=bid,
                                 pos=_vec(ch.attrib.get("pos"), 3, [0, 0, 0]), quat=_orient(ch.attrib)))
            elif ch.tag == "light":
                lights.append(dict(pos=_vec(ch.attrib.get("pos"), 3, [0, 0, 0]), body=bid,
                                   diffuse=_vec(ch.attrib.get("diffuse"), 3, [0.7, 0.7, 0.7])))

    # document order matters for defaults (copy-at-creation), so walk top-level sequentially
    for sec in root:
        if sec.tag == "default":
            dfl.process(sec)
        elif sec.tag == "option":
            for k in ("timestep", "tolerance", "noslip_tolerance", "impratio"):
                if k in sec.attrib:
                    opt[k] = float(sec.attrib[k])
            for k in ("iterations", "noslip_iterations"):
                if k in sec.attrib:
                    opt[k] = int(sec.attrib[k])
        elif sec.tag == "asset":
            for ch in sec:
                if ch.tag == "texture" and ch.attrib.get("builtin", "none") != "none":
                    r1, r2 = _vec(ch.attrib.get("rgb1"), 3, [0.8, 0.8, 0.8]), _vec(ch.attrib.get("rgb2"), 3, [0.5, 0.5, 0.5])
                    textures[ch.attrib.get("name", "")] = r1 if ch.attrib["builtin"] == "flat" else 0.5 * (r1 + r2)
                if ch.tag == "material":
                    rgba = _vec(ch.attrib.get("rgba"), 4, [1, 1, 1, 1]).copy()
                    if ch.attrib.get("texture") in textures:   # builtin texture: its mean colour stands in (no GL here)
                        rgba[:3] *= textures[ch.attrib["texture"]]
                    materials[ch.attrib["name"]] = rgba
    # MuJoCo resolves element defaults when the element is parsed; worldbody appears before the
    # included defaults file in the task XMLs, but MuJoCo's parser handles <default> sections first
    # within one file pass of each top-level section type.  All body elements in these scenes use
    # classes defined in DAPG_assets.xml, so resolve bodies after every <default> is known.
    for sec in root:
        if sec.tag == "worldbody":
            add_body(sec, 0)
        elif sec.tag == "contact":
            for ch in sec:
                if ch.tag == "pair":
                    pairs_x.append(dict(g1=ch.attrib["geom1"], g2=ch.attrib["geom2"],
                                        condim=int(ch.attrib.get("condim", -1))))
                elif ch.tag == "exclude":
                    excludes.append((ch.attrib["body1"], ch.attrib["body2"]))
        elif sec.tag == "tendon":
            for ch in sec:
                assert ch.tag == "fixed"
                a = dfl.resolve("tendon", ch.attrib)
                tendons.append(dict(name=a["name"], range=_vec(a.get("range"), 2, [0, 0]),
                                    limited=a.get("limited", "false") == "true",
                                    margin=float(a.get("margin", 0)),
                                    joints=[(j.attrib["joint"], float(j.attrib["coef"])) for j in ch]))
        elif sec.tag == "actuator":
            for ch in sec:
                assert ch.tag == "general"
                a = dfl.resolve("general", ch.attrib)
                assert a.get("biastype", "none") == "affine" and "joint" in a
                actuators.append(dict(name=a["name"], joint=a["joint"],
                                      ctrlrange=_vec(a.get("ctrlrange"), 2, [0, 0]),
                                      ctrllimited=a.get("ctrllimited", "false") == "true",
                                      gain=_vec(a.get("gainprm"), 3, [1, 0, 0]), bias=_vec(a.get("biasprm"), 3, [0, 0, 0]),
                                      gear=_vec(a.get("gear"), 1, [1])[0]))
        elif sec.tag == "sensor":
            for ch in sec:
                sensors.append(dict(type=ch.tag, name=ch.attrib.get("name", ""), site=ch.attrib.get("site")))

    # MuJoCo numbers the elements body by body (mjCModel::MakeLists: a body's own joints / geoms / sites / cameras /
    # lights, then its children), not in document order: an element declared AFTER a child body (the world's
    # `uprightwall` geom, relocate's `target` and pen's `eps_ball` sites) still precedes every element of that child.
    # Ids are visible to the reference through placeholder ids (relocate's first reset_model writes
    # site_pos[0], which is `target` only under this numbering), so reproduce it.
    for lst in (joints, geoms, sites, cams, lights):
        lst.sort(key=lambda e: e["body"])          # stable: document order within a body
    return _compile(task, opt, bodies, joints, geoms, sites, cams, lights, pairs_x, excludes, tendons,
                    actuators, sensors)


# --------------------------------------------------------------------------- compile
def _compile(task, opt, bodies, joints, geoms_all, sites, cams, lights, pairs_x, excludes, tendons,
             actuators, sensors) -> Model:
    nb, nv = len(bodies), len(joints)
    bname = [b["name"] for b in bodies]
    jname = [j["name"] for j in joints]

    # ---- body inertias (explicit, or from geoms: SURVEY.md App. A.5)
    body_mass = np.zeros(nb)
    body_ipos = np.zeros((nb, 3))
    body_iquat = np.tile([1.0, 0, 0, 0], (nb, 1))
    body_inertia = np.zeros((nb, 3))
    for i, b in enumerate(bodies):
        if i == 0:
            continue
        if b["inertial"] is not None:
            it = b["inertial"]
            body_mass[i], body_ipos[i], body_iquat[i], body_inertia[i] = it["mass"], it["pos"], it["quat"], it["diag"]
            continue
        gs = [g for g in geoms_all if g["body"] == i]
        ms, cs, Is = [], [], []
        for g in gs:
            v, iu = _geom_volume_inertia(g["type"], g["size"])
            if v <= 0:
                continue
            R = quat_to_mat(g["quat"])
            ms.append(g["density"] * v)
            cs.append(g["pos"])
            Is.append(R @ np.diag(g["density"] * iu) @ R.T)
        if not ms:
            continue
        m = sum(ms)
        com = sum(mi * ci for mi, ci in zip(ms, cs)) / m
        I = np.zeros((3, 3))
        for mi, ci, Ii in zip(ms, cs, Is):
            d = ci - com
            I += Ii + mi * (np.dot(d, d) * np.eye(3) - np.outer(d, d))
        w, V = np.linalg.eigh(I)
        order = np.argsort(-w)
        w, V = w[order], V[:, order]
        if np.linalg.det(V) < 0:
            V[:, 2] = -V[:, 2]
        body_mass[i], body_ipos[i], body_iquat[i], body_inertia[i] = m, com, mat_to_quat(V), w

    # ---- tree bookkeeping
    body_parent = np.array([b["parent"] for b in bodies], dtype=np.int32)
    body_parent[0] = 0
    body_jntadr = np.full(nb, -1, dtype=np.int32)
    body_jntnum = np.zeros(nb, dtype=np.int32)
    for j, jn in enumerate(joints):
        if body_jntadr[jn["body"]] < 0:
            body_jntadr[jn["body"]] = j
        body_jntnum[jn["body"]] += 1
    # joints were appended in depth-first body order, so they are contiguous per body
    for i in range(nb):
        if body_jntnum[i]:
            assert all(joints[body_jntadr[i] + k]["body"] == i for k in range(body_jntnum[i]))
    body_weld = np.zeros(nb, dtype=np.int32)
    body_root = np.zeros(nb, dtype=np.int32)
    for i in range(1, nb):
        body_weld[i] = i if body_jntnum[i] else body_weld[body_parent[i]]
        body_root[i] = i if body_parent[i] == 0 else body_root[body_parent[i]]
    dof_body = np.array([j["body"] for j in joints], dtype=np.int32)
    dof_parent = np.full(nv, -1, dtype=np.int32)
    last_dof_of_body = np.full(nb, -1, dtype=np.int32)   # last dof on the chain up to & incl. body
    for i in range(1, nb):
        last = last_dof_of_body[body_parent[i]]
        for k in range(body_jntnum[i]):
            d = body_jntadr[i] + k
            dof_parent[d] = last
            last = d
        last_dof_of_body[i] = last
    body_depth = np.zeros(nb, dtype=np.int32)
    for i in range(1, nb):
        body_depth[i] = body_depth[body_parent[i]] + 1

    # ---- collidable geoms only (visual-only primitives are dropped from the physics model)
    keep = [g for g in geoms_all if g["contype"] or g["conaffinity"]]
    gname = [g["name"] for g in keep]
    ng = len(keep)

    def rbound(g):
        s, t = g["size"], g["type"]
        if t == GEOM_SPHERE: return s[0]
        if t == GEOM_CAPSULE: return s[0] + s[1]
        if t == GEOM_CYLINDER: return float(np.hypot(s[0], s[1]))
        if t == GEOM_BOX: return float(np.linalg.norm(s))
        return 0.0

    # ---- candidate pairs
    excl = {(bname.index(a), bname.index(b)) for a, b in excludes}
    excl |= {(b, a) for a, b in excl}
    pair_rows = []

    def mix(g1, g2, condim=None):
        fr = np.maximum(g1["friction"], g2["friction"])
        return dict(condim=condim if condim is not None and condim > 0 else max(g1["condim"], g2["condim"]),
                    friction=np.array([fr[0], fr[0], fr[1], fr[2], fr[2]]),
                    margin=max(g1["margin"], g2["margin"]), gap=max(g1["gap"], g2["gap"]))

    for i in range(ng):
        for j in range(i + 1, ng):
            g1, g2 = keep[i], keep[j]
            b1, b2 = g1["body"], g2["body"]
            if b1 == b2:
                continue
            w1, w2 = body_weld[b1], body_weld[b2]
            if w1 == w2:
                continue                                   # same weld group (incl. both static)
            if w1 != 0 and w2 != 0 and (body_weld[body_parent[w1]] == w2 or body_weld[body_parent[w2]] == w1):
                continue                                   # parent-child filter
            if not ((g1["contype"] & g2["conaffinity"]) or (g2["contype"] & g1["conaffinity"])):
                continue
            if (b1, b2) in excl:
                continue
            if g1["type"] == GEOM_PLANE and g2["type"] == GEOM_PLANE:
                continue
            pair_rows.append((min(b1, b2), max(b1, b2), 0, i, j, mix(g1, g2)))
    n_dynamic = len(pair_rows)
    for k, p in enumerate(pairs_x):
        i, j = gname.index(p["g1"]), gname.index(p["g2"])
        b1, b2 = keep[i]["body"], keep[j]["body"]
        pair_rows.append((min(b1, b2), max(b1, b2), 1 + k, i, j, mix(keep[i], keep[j], p["condim"])))
    # contact order: by body pair, dynamic before explicit, then geom ids (noslip sweeps follow it)
    pair_rows.sort(key=lambda r: (r[0], r[1], r[2], r[3], r[4]))
    pair_geom = np.zeros((len(pair_rows), 2), dtype=np.int32)
    pair_condim = np.zeros(len(pair_rows), dtype=np.int32)
    pair_friction = np.zeros((len(pair_rows), 5))
    pair_margin = np.zeros(len(pair_rows))
    for k, (_, _, _, i, j, mx) in enumerate(pair_rows):
        # lower geom type first: the contact normal points from geom1 to geom2
        if keep[i]["type"] > keep[j]["type"]:
            i, j = j, i
        pair_geom[k] = (i, j)
        pair_condim[k] = mx["condim"]
        pair_friction[k] = mx["friction"]
        pair_margin[k] = mx["margin"] - mx["gap"]

    # ---- tendons (fixed, ≤2 joints each in these scenes)
    nt = len(tendons)
    tendon_dof = np.full((nt, 2), -1, dtype=np.int32)
    tendon_coef = np.zeros((nt, 2))
    for t, td in enumerate(tendons):
        assert len(td["joints"]) <= 2
        for k, (jn, c) in enumerate(td["joints"]):
            tendon_dof[t, k] = jname.index(jn)
            tendon_coef[t, k] = c

    # ---- actuators
    nu = len(actuators)
    act_dof = np.array([jname.index(a["joint"]) for a in actuators], dtype=np.int32)
    act_ctrlrange = np.array([a["ctrlrange"] for a in actuators]).reshape(nu, 2)
    act_gain = np.array([a["gain"][0] * a["gear"] for a in actuators])
    act_bias = np.array([[a["bias"][0], a["bias"][1], a["bias"][2]] for a in actuators]).reshape(nu, 3)
    for a in actuators:
        assert a["ctrllimited"], "unclamped ctrl is outside the Adroit subset"
        assert a["gear"] == 1.0

    A = {}
    A["body_parent"] = body_parent
    A["body_pos"] = np.array([b["pos"] for b in bodies])
    A["body_quat"] = np.array([b["quat"] for b in bodies])
    A["body_ipos"], A["body_iquat"] = body_ipos, np.array([q / np.linalg.norm(q) for q in body_iquat])
    A["body_mass"], A["body_inertia"] = body_mass, body_inertia
    A["body_jntadr"], A["body_jntnum"] = body_jntadr, body_jntnum
    A["body_weld"], A["body_root"], A["body_depth"] = body_weld, body_root, body_depth
    A["jnt_type"] = np.array([j["type"] for j in joints], dtype=np.int32)
    A["jnt_pos"] = np.array([j["pos"] for j in joints])
    A["jnt_axis"] = np.array([j["axis"] for j in joints])
    A["jnt_range"] = np.array([j["range"] for j in joints])
    A["jnt_limited"] = np.array([j["limited"] for j in joints], dtype=np.int32)
    A["jnt_margin"] = np.array([j["margin"] for j in joints])
    A["dof_body"], A["dof_parent"] = dof_body, dof_parent
    A["dof_damping"] = np.array([j["damping"] for j in joints])
    A["dof_armature"] = np.array([j["armature"] for j in joints])
    A["dof_frictionloss"] = np.array([j["frictionloss"] for j in joints])
    for j in joints:
        assert j["ref"] == 0.0
    A["geom_type"] = np.array([g["type"] for g in keep], dtype=np.int32)
    A["geom_body"] = np.array([g["body"] for g in keep], dtype=np.int32)
    A["geom_pos"] = np.array([g["pos"] for g in keep])
    A["geom_quat"] = np.array([g["quat"] for g in keep])
    A["geom_size"] = np.array([g["size"] for g in keep])
    A["geom_rbound"] = np.array([rbound(g) for g in keep])
    A["geom_rgba"] = np.array([g["rgba"] for g in keep])
    A["site_body"] = np.array([s["body"] for s in sites], dtype=np.int32)
    A["site_pos"] = np.array([s["pos"] for s in sites])
    A["site_quat"] = np.array([s["quat"] for s in sites])
    A["site_size"] = np.array([s["size"] for s in sites])
    A["site_type"] = np.array([{"sphere": GEOM_SPHERE, "box": GEOM_BOX, "cylinder": GEOM_CYLINDER}[s["type"]]
                               for s in sites], dtype=np.int32)
    A["pair_geom"], A["pair_condim"] = pair_geom, pair_condim
    A["pair_friction"], A["pair_margin"] = pair_friction, pair_margin
    A["tendon_dof"], A["tendon_coef"] = tendon_dof, tendon_coef
    A["tendon_range"] = np.array([t["range"] for t in tendons]).reshape(nt, 2)
    A["tendon_limited"] = np.array([t["limited"] for t in tendons], dtype=np.int32)
    A["tendon_margin"] = np.array([t["margin"] for t in tendons])
    A["act_dof"], A["act_ctrlrange"], A["act_gain"] = act_dof, act_ctrlrange, act_gain
    A["act_bias"] = act_bias
    A["cam_pos"] = np.array([c["pos"] for c in cams]).reshape(-1, 3)
    A["cam_quat"] = np.array([c["quat"] for c in cams]).reshape(-1, 4)
    A["light_pos"] = np.array([l["pos"] for l in lights]).reshape(-1, 3)

    names = dict(body=bname, joint=jname, geom=gname, site=[s["name"] for s in sites],
                 actuator=[a["name"] for a in actuators], tendon=[t["name"] for t in tendons],
                 sensor=[s["name"] for s in sensors], camera=[c["name"] for c in cams])

    # ---- options: [timestep, gx, gy, gz, tolerance, noslip_tol, impratio, solref0, solref1,
    #                solimp0..4, mpr_tol]
    A["opt_real"] = np.array([opt["timestep"], 0, 0, -9.81, opt["tolerance"], opt["noslip_tolerance"],
                              opt["impratio"], 0.02, 1.0, 0.9, 0.95, 0.001, 0.5, 2.0, opt["mpr_tolerance"]])
    A["opt_int"] = np.array([opt["iterations"], opt["noslip_iterations"], opt["mpr_iterations"], 50], dtype=np.int32)

    # ---- task layer tables (reference *_v0.py)
    sid = lambda n: names["site"].index(n)
    bid = lambda n: bname.index(n)
    task_int = np.full(16, -1, dtype=np.int32)
    body_par = np.full((nb, 2), -1, dtype=np.int32)        # per-env param offset of body_pos / body_quat
    site_par = np.full(len(sites), -1, dtype=np.int32)     # per-env param offset of site_pos
    if task == "hammer":     # hammer_v0.py:44-48,96-104,109
        frame_skip, obs_dim, horizon = 5, 46, 200
        task_int[:6] = [sid("S_target"), sid("S_grasp"), bid("Object"), sid("tool"), sid("nail_goal"), bid("nail")]
        body_par[bid("nail_board"), 0] = 0
        npar = 3
    elif task == "door":     # door_v0.py:49-52,91-101,107-109
        frame_skip, obs_dim, horizon = 1, 39, 200
        task_int[:4] = [jname.index("door_hinge"), sid("S_grasp"), sid("S_handle"), bid("frame")]
        body_par[bid("frame"), 0] = 0
        npar = 3
    elif task == "pen":      # pen_v0.py:48-58,105-113,122
        frame_skip, obs_dim, horizon = 5, 45, 100
        task_int[:8] = [bid("target"), sid("S_grasp"), bid("Object"), sid("eps_ball"), sid("object_top"),
                        sid("object_bottom"), sid("target_top"), sid("target_bottom")]
        body_par[bid("target"), 1] = 0
        npar = 4
    else:                    # relocate_v0.py:36-38,78-83,89-93
        frame_skip, obs_dim, horizon = 5, 39, 200
        task_int[:3] = [sid("target"), sid("S_grasp"), bid("Object")]
        body_par[bid("Object"), 0] = 0
        site_par[sid("target")] = 3
        npar = 6
    A["task_int"] = task_int
    A["body_par"] = body_par
    A["site_par"] = site_par
    par0 = np.zeros(npar)
    for b in range(nb):
        if body_par[b, 0] >= 0:
            par0[body_par[b, 0]:body_par[b, 0] + 3] = A["body_pos"][b]
        if body_par[b, 1] >= 0:
            par0[body_par[b, 1]:body_par[b, 1] + 4] = A["body_quat"][b]
    for s in range(len(sites)):
        if site_par[s] >= 0:
            par0[site_par[s]:site_par[s] + 3] = A["site_pos"][s]
    A["par0"] = par0
    # touch sensor for S_nail (hammer_v0.py:102; DAPG_hammer.xml:34-36)
    touch_site = -1
    for s in sensors:
        if s["type"] == "touch" and s["name"] == "S_nail":
            touch_site = sid(s["site"])
    A["touch_site"] = np.array([touch_site], dtype=np.int32)

    # ---- render table for the pixel path (reference headless_observer.py).  Meshes and texture
    # files are absent and no GL exists here, so the CUDA ray-caster draws every primitive geom (collision shapes
    # included: they stand in for the missing visual meshes) and the visible sites (group <= 2; translucent ones are alpha-blended) with
    # their material colours.  render_int rows: type, body, per-env parameter offset of the local position (-1 = none),
    # is_site; render_real rows: local pos[3], local mat[9], size[3], rgba[4].
    rg_int, rg_real = [], []
    rg_of_geom = {}
    for g in geoms_all:
        if g["rgba"][3] <= 0.0:
            continue
        rg_of_geom[id(g)] = len(rg_int)
        rg_int.append([g["type"], g["body"], -1, 0])
        rg_real.append(np.concatenate([g["pos"], quat_to_mat(g["quat"]).ravel(), g["size"], g["rgba"][:3], [1.0]]))   # geoms are drawn opaque
    for si, st in enumerate(sites):
        if st["group"] > 2 or st["rgba"][3] < 0.05:
            continue
        ty = {"sphere": GEOM_SPHERE, "box": GEOM_BOX, "cylinder": GEOM_CYLINDER}[st["type"]]
        rg_int.append([ty, st["body"], int(site_par[si]), 1])
        rg_real.append(np.concatenate([st["pos"], quat_to_mat(st["quat"]).ravel(), st["size"], st["rgba"]]))           # sites keep their alpha
    A["render_int"] = np.array(rg_int, dtype=np.int32).reshape(-1, 4)
    A["render_real"] = np.array(rg_real).reshape(-1, 19)
    A["render_geom_index"] = np.array([rg_of_geom.get(id(g), -1) for g in keep], dtype=np.int32)   # collidable geom -> render row
    A["light_diffuse"] = np.array([l["diffuse"] for l in lights]).reshape(-1, 3)
    A["light_body"] = np.array([l["body"] for l in lights], dtype=np.int32)
    A["cam_body"] = np.array([c["body"] for c in cams], dtype=np.int32)
    # which body the reference hands to HeadlessObserver: hammer -1 (= last body), door / relocate the
    # placeholder 0 (door_v0.py:41, relocate_v0.py:32), pen its own setup with the target body (pen_v0.py:163-177)
    A["render_obj_body"] = np.array([{"hammer": nb - 1, "door": 0, "relocate": 0, "pen": bid("target") if task == "pen" else 0}[task]],
                                    dtype=np.int32)
    A["sizes"] = np.array([nv, nb, ng, len(sites), nu, nt, len(pair_rows), npar, obs_dim, frame_skip,
                           TASK_IDS[task], horizon, n_dynamic, len(cams), 0, 0], dtype=np.int32)

    m = Model(task=task, names=names, A=A)
    _set_const(m)
    _set_render_camera(m)
    return m


# --------------------------------------------------------------------------- mj_setConst equivalent
def fk_numpy(m: Model, qpos, par=None):
    """Reference forward kinematics in plain numpy (used for compile-time constants and as an
    independent cross-check of the oracle / CUDA kinematics)."""
    A = m.A
    nb = m.nbody
    par = A["par0"] if par is None else par
    xpos = np.zeros((nb, 3))
    xquat = np.tile([1.0, 0, 0, 0], (nb, 1))
    xanchor = np.zeros((m.nv, 3))
    xaxis = np.zeros((m.nv, 3))
    for b in range(1, nb):
        p = A["body_parent"][b]
        bp = A["body_pos"][b] if A["body_par"][b, 0] < 0 else par[A["body_par"][b, 0]:A["body_par"][b, 0] + 3]
        bq = A["body_quat"][b] if A["body_par"][b, 1] < 0 else par[A["body_par"][b, 1]:A["body_par"][b, 1] + 4]
        bq = np.asarray(bq) / np.linalg.norm(bq)
        pos = xpos[p] + quat_to_mat(xquat[p]) @ bp
        quat = quat_mul(xquat[p], bq)
        for k in range(A["body_jntnum"][b]):
            j = A["body_jntadr"][b] + k
            R = quat_to_mat(quat)
            xanchor[j] = pos + R @ A["jnt_pos"][j]
            xaxis[j] = R @ A["jnt_axis"][j]
            if A["jnt_type"][j] == JNT_SLIDE:
                pos = pos + xaxis[j] * qpos[j]
            else:
                quat = quat_mul(quat, axis_angle_quat(A["jnt_axis"][j], qpos[j]))
                pos = xanchor[j] - quat_to_mat(quat) @ A["jnt_pos"][j]
        xpos[b], xquat[b] = pos, quat / np.linalg.norm(quat)
    return xpos, xquat, xanchor, xaxis


def jac_numpy(m: Model, xanchor, xaxis, body, point):
    """6 x nv Jacobian [translational; rotational] of `point` attached to `body`."""
    A = m.A
    J = np.zeros((6, m.nv))
    b = body
    while b > 0:
        for k in range(A["body_jntnum"][b]):
            j = A["body_jntadr"][b] + k
            if A["jnt_type"][j] == JNT_SLIDE:
                J[:3, j] = xaxis[j]
            else:
                J[:3, j] = np.cross(xaxis[j], point - xanchor[j])
                J[3:, j] = xaxis[j]
        b = A["body_parent"][b]
    return J


def mass_matrix_numpy(m: Model, qpos, par=None):
    A = m.A
    xpos, xquat, xanchor, xaxis = fk_numpy(m, qpos, par)
    M = np.diag(A["dof_armature"].astype(np.float64).copy())
    for b in range(1, m.nbody):
        if A["body_mass"][b] <= 0:
            continue
        R = quat_to_mat(xquat[b])
        c = xpos[b] + R @ A["body_ipos"][b]
        Ri = R @ quat_to_mat(A["body_iquat"][b])
        Iw = Ri @ np.diag(A["body_inertia"][b]) @ Ri.T
        J = jac_numpy(m, xanchor, xaxis, b, c)
        M += A["body_mass"][b] * J[:3].T @ J[:3] + J[3:].T @ Iw @ J[3:]
    return M


def _set_render_camera(m: Model):
    """Free-camera constants of the reference's headless pixel path (headless_observer.py:20-31,59-67; pen_v0.py:163-177),
    evaluated once at qpos0 like the reference does at construction:
      lookat    = per-axis median of the geom world positions  [EXT mujoco-py MjRenderContext._init_camera]; the mesh geoms
                  are absent here, so the median runs over the primitive geoms of the render table
      azimuth   = 90, distance = 4.5                             (headless_observer.py:29-30)
      elevation = -45 + deg(acos(v_x / v_z)) / 2, v = xpos[obj body] - cam_xpos[last camera]   (:61-65)
      fovy      = 45 (MuJoCo default), frame 640x480, centre crop 128x128 (:38,44)
    np.arccos outside [-1, 1] gives NaN in the reference (MuJoCo then renders with a garbage camera); this build clamps
    the argument and records that it did so in render_cam[7].
    render_cam = [lookat x, y, z, azimuth, elevation, distance, fovy, clamped]."""
    A = m.A
    xpos, xquat, _, _ = fk_numpy(m, np.zeros(m.nv))
    ri, rr = A["render_int"], A["render_real"]
    gp = np.array([xpos[b] + quat_to_mat(xquat[b]) @ r[:3] for (_, b, _, site), r in zip(ri, rr) if not site])
    lookat = np.median(gp, axis=0)
    cb = int(A["cam_body"][-1])
    cam = xpos[cb] + quat_to_mat(xquat[cb]) @ A["cam_pos"][-1]
    v = xpos[int(A["render_obj_body"][0])] - cam
    ratio = v[0] / v[2] if v[2] != 0 else 0.0
    clamped = float(abs(ratio) > 1)
    elev = -45.0 + np.rad2deg(np.arccos(np.clip(ratio, -1.0, 1.0))) / 2
    A["render_cam"] = np.array([lookat[0], lookat[1], lookat[2], 90.0, elev, 4.5, 45.0, clamped])


def _set_const(m: Model):
    """dof/body/tendon invweight0 and mean inertia at qpos0 (MuJoCo mj_setConst semantics)."""
    A = m.A
    nv = m.nv
    q0 = np.zeros(nv)
    M = mass_matrix_numpy(m, q0)
    Minv = np.linalg.inv(M)
    A["dof_invweight0"] = np.diag(Minv).copy()
    xpos, xquat, xanchor, xaxis = fk_numpy(m, q0)
    biw = np.zeros((m.nbody, 2))
    for b in range(1, m.nbody):
        if A["body_weld"][b] == 0:
            continue
        c = xpos[b] + quat_to_mat(xquat[b]) @ A["body_ipos"][b]
        J = jac_numpy(m, xanchor, xaxis, b, c)
        Ab = J @ Minv @ J.T
        biw[b] = [np.trace(Ab[:3, :3]) / 3.0, np.trace(Ab[3:, 3:]) / 3.0]
    A["body_invweight0"] = biw
    tiw = np.zeros(m.ntendon)
    for t in range(m.ntendon):
        Jt = np.zeros(nv)
        for k in range(2):
            if A["tendon_dof"][t, k] >= 0:
                Jt[A["tendon_dof"][t, k]] += A["tendon_coef"][t, k]
        tiw[t] = Jt @ Minv @ Jt
    A["tendon_invweight0"] = tiw
    A["meaninertia"] = np.array([np.trace(M) / nv])


# --------------------------------------------------------------------------- packaged models
_COMPILED_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "compiled")


def load_compiled(task: str) -> Model:
    """Load the pre-compiled scene shipped with the package (built by tools/compile_models.py)."""
    task = task.replace("-v0", "")
    with open(os.path.join(_COMPILED_DIR, f"{task}.adroit"), "rb") as f:
        return Model.from_blob(f.read())
