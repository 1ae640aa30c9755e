"""Offline MJCF model compiler: MuJoCo XML -> constant tables for the oracle and the CUDA engine.

Replaces what the reference gets from ``mj_loadXML`` (libs/gymcpp/mujoco/mujoco_env.h:34) for exactly the
feature subset the four task assets use (libs/gymcpp/mujoco/assets/{half_cheetah,hopper,ant,humanoid}.xml):
one level of <default>, bodies with slide/hinge/free joints, plane/sphere/capsule geoms (pos/quat/axisangle/
fromto), inertiafromgeom, settotalmass, motors on joints, fixed tendons without spring/limit (dynamically
inert, asserted and dropped).  MuJoCo 3.2.0 is not available in this environment, so the derived constants
(inertial frames, invweight0, meaninertia, collision pair list) restate MuJoCo's published compile rules;
tests/test_model_compile.py pins the known answers the reference carries (body masses in
src/ac_ppo_continuous_action.cpp:496-497, shape asserts in the task headers).

Run ``python -m ppo_cpp_b200.mjcf_compile <assets dir> <out header>`` to regenerate csrc/model_tables.h.
"""
from __future__ import annotations

import math
import sys
import xml.etree.ElementTree as ET
from dataclasses import dataclass, field

import numpy as np

MJ_MINVAL = 1e-15
GEOM_PLANE, GEOM_SPHERE, GEOM_CAPSULE = 0, 2, 3  # MuJoCo mjtGeom values (ordering matters for pair swap)
JNT_FREE, JNT_SLIDE, JNT_HINGE = 0, 2, 3  # mjtJoint values


# ----------------------------------------------------------------------------- small math helpers
def _floats(s, n=None):
    v = [float(x) for x in s.split()]
    if n is not None and len(v) < n:
        v = v + [0.0] * (n - len(v))
    return v


def quat_mul(a, b):
    w1, x1, y1, z1 = a
    w2, x2, y2, z2 = b
    return np.array([
        w1 * w2 - x1 * x2 - y1 * y2 - z1 * z2,
        w1 * x2 + x1 * w2 + y1 * z2 - z1 * y2,
        w1 * y2 - x1 * z2 + y1 * w2 + z1 * x2,
        w1 * z2 + x1 * y2 - y1 * x2 + z1 * w2,
    ])


def quat_to_mat(q):
    w, x, y, z = q
    return np.array([
        [w * w + x * x - y * y - z * z, 2 * (x * y - w * z), 2 * (x * z + w * y)],
        [2 * (x * y + w * z), w * w - x * x + y * y - z * z, 2 * (y * z - w * x)],
        [2 * (x * z - w * y), 2 * (y * z + w * x), w * w - x * x - y * y + z * z],
    ])


def quat_normalize(q):
    q = np.asarray(q, dtype=np.float64)
    n = np.linalg.norm(q)
    if n < MJ_MINVAL:
        return np.array([1.0, 0, 0, 0])
    return q / n


def axisangle_to_quat(axis, angle):
    axis = np.asarray(axis, dtype=np.float64)
    n = np.linalg.norm(axis)
    if n < MJ_MINVAL:
        return np.array([1.0, 0, 0, 0])
    axis = axis / n
    return np.concatenate([[math.cos(angle / 2)], axis * math.sin(angle / 2)])


def z_to_quat(vec):
    """Quaternion rotating the z axis onto ``vec`` (MuJoCo's rule for fromto geoms)."""
    vec = np.asarray(vec, dtype=np.float64)
    vec = vec / np.linalg.norm(vec)
    z = np.array([0.0, 0.0, 1.0])
    axis = np.cross(z, vec)
    s = np.linalg.norm(axis)
    if s < 1e-10:
        axis = np.array([1.0, 0.0, 0.0])
    else:
        axis = axis / s
    ang = math.atan2(s, vec[2])
    return np.concatenate([[math.cos(ang / 2)], axis * math.sin(ang / 2)])


# ----------------------------------------------------------------------------- parsed model
@dataclass
class Model:
    name: str = ""
    timestep: float = 0.002
    gravity: tuple = (0.0, 0.0, -9.81)
    integrator: int = 0  # 0 Euler, 1 RK4
    solver: int = 2  # 0 PGS, 2 Newton (MuJoCo default)
    iterations: int = 100
    tolerance: float = 1e-8
    # bodies
    body_parent: list = field(default_factory=list)
    body_pos: list = field(default_factory=list)
    body_quat: list = field(default_factory=list)
    body_name: list = field(default_factory=list)
    body_jntadr: list = field(default_factory=list)
    body_jntnum: list = field(default_factory=list)
    # joints
    jnt_type: list = field(default_factory=list)
    jnt_body: list = field(default_factory=list)
    jnt_pos: list = field(default_factory=list)
    jnt_axis: list = field(default_factory=list)
    jnt_limited: list = field(default_factory=list)
    jnt_range: list = field(default_factory=list)
    jnt_stiffness: list = field(default_factory=list)
    jnt_damping: list = field(default_factory=list)
    jnt_armature: list = field(default_factory=list)
    jnt_margin: list = field(default_factory=list)
    jnt_ref: list = field(default_factory=list)
    jnt_springref: list = field(default_factory=list)
    jnt_solref: list = field(default_factory=list)
    jnt_solimp: list = field(default_factory=list)
    jnt_name: list = field(default_factory=list)
    # geoms
    geom_type: list = field(default_factory=list)
    geom_body: list = field(default_factory=list)
    geom_size: list = field(default_factory=list)
    geom_pos: list = field(default_factory=list)
    geom_quat: list = field(default_factory=list)
    geom_density: list = field(default_factory=list)
    geom_friction: list = field(default_factory=list)
    geom_margin: list = field(default_factory=list)
    geom_gap: list = field(default_factory=list)
    geom_condim: list = field(default_factory=list)
    geom_contype: list = field(default_factory=list)
    geom_conaffinity: list = field(default_factory=list)
    geom_solref: list = field(default_factory=list)
    geom_solimp: list = field(default_factory=list)
    geom_solmix: list = field(default_factory=list)
    geom_name: list = field(default_factory=list)
    # actuators
    act_jnt: list = field(default_factory=list)
    act_gear: list = field(default_factory=list)
    act_ctrllimited: list = field(default_factory=list)
    act_ctrlrange: list = field(default_factory=list)
    settotalmass: float = -1.0


def _merged(defaults, elem):
    d = dict(defaults)
    d.update(elem.attrib)
    return d


def parse_mjcf(path: str) -> Model:
    root = ET.parse(path).getroot()
    m = Model(name=root.attrib.get("model", ""))
    comp = root.find("compiler")
    comp = comp.attrib if comp is not None else {}
    degree = comp.get("angle", "degree") == "degree"
    assert comp.get("inertiafromgeom", "auto") in ("true", "auto")
    assert comp.get("coordinate", "local") == "local"
    m.settotalmass = float(comp.get("settotalmass", -1))
    ang = (math.pi / 180.0) if degree else 1.0

    opt = root.find("option")
    if opt is not None:
        m.timestep = float(opt.attrib.get("timestep", m.timestep))
        if "gravity" in opt.attrib:
            m.gravity = tuple(_floats(opt.attrib["gravity"]))
        m.integrator = {"Euler": 0, "RK4": 1}[opt.attrib.get("integrator", "Euler")]
        m.solver = {"PGS": 0, "CG": 1, "Newton": 2}[opt.attrib.get("solver", "Newton")]
        m.iterations = int(opt.attrib.get("iterations", m.iterations))
        m.tolerance = float(opt.attrib.get("tolerance", m.tolerance))
    assert m.solver in (0, 2)

    dflt = root.find("default")
    d_joint, d_geom, d_motor = {}, {}, {}
    if dflt is not None:
        assert dflt.find("default") is None, "nested default classes are not supported"
        for tag, dst in (("joint", d_joint), ("geom", d_geom), ("motor", d_motor)):
            e = dflt.find(tag)
            if e is not None:
                dst.update(e.attrib)

    joint_ids = {}

    def add_geom(e, body_id):
        a = _merged(d_geom, e)
        typ = {"plane": GEOM_PLANE, "sphere": GEOM_SPHERE, "capsule": GEOM_CAPSULE}[a.get("type", "sphere")]
        size = _floats(a.get("size", "0"), 3)
        pos = np.array(_floats(a.get("pos", "0 0 0")))
        quat = np.array([1.0, 0, 0, 0])
        if "quat" in a:
            quat = quat_normalize(_floats(a["quat"]))
        elif "axisangle" in a:
            v = _floats(a["axisangle"])
            quat = axisangle_to_quat(v[:3], v[3] * ang)
        if "fromto" in a:
            ft = np.array(_floats(a["fromto"]))
            assert typ == GEOM_CAPSULE
            # MuJoCo: pos = midpoint, z axis along (from - to), half-length from the segment
            vec = ft[0:3] - ft[3:6]
            pos = 0.5 * (ft[0:3] + ft[3:6])
            size = [size[0], 0.5 * float(np.linalg.norm(vec)), 0.0]
            quat = z_to_quat(vec)
        m.geom_type.append(typ)
        m.geom_body.append(body_id)
        m.geom_size.append(size)
        m.geom_pos.append(pos)
        m.geom_quat.append(quat)
        m.geom_density.append(float(a.get("density", 1000.0)))
        fr = _floats(a.get("friction", "1 0.005 0.0001"))
        fr = fr + [0.005, 0.0001][len(fr) - 1:] if len(fr) < 3 else fr
        m.geom_friction.append(fr[:3])
        m.geom_margin.append(float(a.get("margin", 0.0)))
        m.geom_gap.append(float(a.get("gap", 0.0)))
        m.geom_condim.append(int(a.get("condim", 3)))
        m.geom_contype.append(int(a.get("contype", 1)))
        m.geom_conaffinity.append(int(a.get("conaffinity", 1)))
        m.geom_solref.append(_floats(a.get("solref", "0.02 1")))
        si = _floats(a.get("solimp", "0.9 0.95 0.001 0.5 2"))
        si = si + [0.9, 0.95, 0.001, 0.5, 2.0][len(si):]
        m.geom_solimp.append(si)
        m.geom_solmix.append(float(a.get("solmix", 1.0)))
        m.geom_name.append(a.get("name", f"geom{len(m.geom_name)}"))

    def add_joint(e, body_id):
        a = _merged(d_joint, e)
        typ = {"free": JNT_FREE, "slide": JNT_SLIDE, "hinge": JNT_HINGE}[a.get("type", "hinge")]
        is_ang = typ == JNT_HINGE
        rng = _floats(a.get("range", "0 0"))
        if is_ang:
            rng = [r * ang for r in rng]
        limited = a.get("limited", "auto")
        if limited == "auto":
            limited = "true" if "range" in a else "false"
        if typ == JNT_FREE:
            limited = "false"
        axis = np.array(_floats(a.get("axis", "0 0 1")))
        if typ != JNT_FREE:
            axis = axis / np.linalg.norm(axis)
        m.jnt_type.append(typ)
        m.jnt_body.append(body_id)
        m.jnt_pos.append(np.array(_floats(a.get("pos", "0 0 0"))))
        m.jnt_axis.append(axis)
        m.jnt_limited.append(limited == "true")
        m.jnt_range.append(rng)
        # free joints ignore stiffness/damping/armature defaults in these assets (explicitly zero there)
        m.jnt_stiffness.append(float(a.get("stiffness", 0.0)))
        m.jnt_damping.append(float(a.get("damping", 0.0)))
        m.jnt_armature.append(float(a.get("armature", 0.0)))
        m.jnt_margin.append(float(a.get("margin", 0.0)))
        m.jnt_ref.append(float(a.get("ref", 0.0)) * (ang if is_ang else 1.0))
        m.jnt_springref.append(float(a.get("springref", 0.0)) * (ang if is_ang else 1.0))
        m.jnt_solref.append(_floats(a.get("solreflimit", "0.02 1")))
        si = _floats(a.get("solimplimit", "0.9 0.95 0.001 0.5 2"))
        si = si + [0.9, 0.95, 0.001, 0.5, 2.0][len(si):]
        m.jnt_solimp.append(si)
        name = a.get("name", f"jnt{len(m.jnt_name)}")
        joint_ids[name] = len(m.jnt_name)
        m.jnt_name.append(name)

    def add_body(e, parent_id):
        body_id = len(m.body_parent)
        m.body_parent.append(parent_id)
        m.body_name.append(e.attrib.get("name", "world" if parent_id < 0 else f"body{body_id}"))
        m.body_pos.append(np.array(_floats(e.attrib.get("pos", "0 0 0"))))
        m.body_quat.append(quat_normalize(_floats(e.attrib.get("quat", "1 0 0 0"))))
        m.body_jntadr.append(len(m.jnt_type))
        for j in e.findall("joint"):
            add_joint(j, body_id)
        if e.find("freejoint") is not None:
            raise NotImplementedError("freejoint tag")
        m.body_jntnum.append(len(m.jnt_type) - m.body_jntadr[-1])
        for g in e.findall("geom"):
            add_geom(g, body_id)
        for b in e.findall("body"):
            add_body(b, body_id)

    add_body(root.find("worldbody"), -1)
    m.body_parent[0] = 0

    for t in root.findall("tendon/*"):
        assert t.tag == "fixed"
        for k in ("stiffness", "damping", "limited", "range", "frictionloss", "springlength"):
            assert k not in t.attrib, "only dynamically inert fixed tendons are supported"

    act = root.find("actuator")
    for e in (act if act is not None else []):
        assert e.tag == "motor"
        a = _merged(d_motor, e)
        m.act_jnt.append(joint_ids[a["joint"]])
        m.act_gear.append(_floats(a.get("gear", "1"))[0])
        lim = a.get("ctrllimited", "auto")
        if lim == "auto":
            lim = "true" if "ctrlrange" in a else "false"
        m.act_ctrllimited.append(lim == "true")
        m.act_ctrlrange.append(_floats(a.get("ctrlrange", "0 0")))
    return m


# ----------------------------------------------------------------------------- derived constants
def _geom_mass_inertia(typ, size, density):
    """Mass and diagonal inertia (geom frame, about the geom centre) for sphere/capsule at uniform density."""
    if typ == GEOM_SPHERE:
        r = size[0]
        mass = density * 4.0 / 3.0 * math.pi * r ** 3
        i = 2.0 * mass * r * r / 5.0
        return mass, np.array([i, i, i])
    if typ == GEOM_CAPSULE:
        r, half = size[0], size[1]
        height = 2.0 * half
        vol = math.pi * (r * r * height + 4.0 * r ** 3 / 3.0)
        mass = density * vol
        sphere_mass = mass * 4.0 * r / (4.0 * r + 3.0 * height)
        cyl_mass = mass - sphere_mass
        ixx = cyl_mass * (3.0 * r * r + height * height) / 12.0
        izz = cyl_mass * r * r / 2.0
        sph_i = 2.0 * sphere_mass * r * r / 5.0
        ixx += sph_i + sphere_mass * height * (3.0 * r + 2.0 * height) / 8.0
        izz += sph_i
        return mass, np.array([ixx, ixx, izz])
    return 0.0, np.zeros(3)  # planes are massless (world body)


def compile_model(m: Model) -> dict:
    nbody = len(m.body_parent)
    njnt = len(m.jnt_type)
    ngeom = len(m.geom_type)
    nu = len(m.act_jnt)

    # ---- inertial properties from geoms (inertiafromgeom) ----
    body_mass = np.zeros(nbody)
    body_ipos = np.zeros((nbody, 3))
    body_inertia = np.zeros((nbody, 3, 3))  # about the body com, in body-frame axes
    for b in range(1, nbody):
        gs = [g for g in range(ngeom) if m.geom_body[g] == b]
        mi = [_geom_mass_inertia(m.geom_type[g], m.geom_size[g], m.geom_density[g]) for g in gs]
        mass = sum(x[0] for x in mi)
        assert mass > 0, f"body {m.body_name[b]} has no mass"
        com = sum(x[0] * np.asarray(m.geom_pos[g]) for x, g in zip(mi, gs)) / mass
        inertia = np.zeros((3, 3))
        for (gm, gi), g in zip(mi, gs):
            rot = quat_to_mat(m.geom_quat[g])
            d = np.asarray(m.geom_pos[g]) - com
            inertia += rot @ np.diag(gi) @ rot.T + gm * (d @ d * np.eye(3) - np.outer(d, d))
        body_mass[b], body_ipos[b], body_inertia[b] = mass, com, inertia
    if m.settotalmass > 0:
        scale = m.settotalmass / body_mass.sum()
        body_mass *= scale
        body_inertia *= scale

    # ---- dof / qpos layout ----
    jnt_qposadr, jnt_dofadr = [], []
    dof_jnt, dof_body = [], []
    nq = nv = 0
    for j in range(njnt):
        jnt_qposadr.append(nq)
        jnt_dofadr.append(nv)
        n_q, n_v = (7, 6) if m.jnt_type[j] == JNT_FREE else (1, 1)
        nq += n_q
        for _ in range(n_v):
            dof_jnt.append(j)
            dof_body.append(m.jnt_body[j])
        nv += n_v
    body_dofadr, body_dofnum = [], []
    for b in range(nbody):
        ds = [d for d in range(nv) if dof_body[d] == b]
        body_dofadr.append(ds[0] if ds else -1)
        body_dofnum.append(len(ds))

    def last_dof_at_or_above(b):
        while b > 0:
            if body_dofnum[b] > 0:
                return body_dofadr[b] + body_dofnum[b] - 1
            b = m.body_parent[b]
        return -1

    dof_parent = []
    for d in range(nv):
        b = dof_body[d]
        if d > body_dofadr[b]:
            dof_parent.append(d - 1)
        else:
            dof_parent.append(last_dof_at_or_above(m.body_parent[b]))

    qpos0 = np.zeros(nq)
    for j in range(njnt):
        a = jnt_qposadr[j]
        if m.jnt_type[j] == JNT_FREE:
            b = m.jnt_body[j]
            assert m.body_parent[b] == 0
            qpos0[a:a + 3] = m.body_pos[b]
            qpos0[a + 3:a + 7] = m.body_quat[b]
        else:
            qpos0[a] = m.jnt_ref[j]

    # weld ids (bodies without joints are welded to the nearest ancestor that has some) and root ids
    body_weld = [0] * nbody
    body_root = [0] * nbody
    for b in range(1, nbody):
        body_weld[b] = b if body_dofnum[b] > 0 else body_weld[m.body_parent[b]]
        body_root[b] = b if m.body_parent[b] == 0 else body_root[m.body_parent[b]]
    subtree_size = [1] * nbody
    for b in range(nbody - 1, 0, -1):
        subtree_size[m.body_parent[b]] += subtree_size[b]
    body_depth = [0] * nbody
    for b in range(1, nbody):
        body_depth[b] = body_depth[m.body_parent[b]] + 1

    # ---- collision pair list: (contype & conaffinity), same-weld and parent-child filters ----
    pairs = []
    for g1 in range(ngeom):
        for g2 in range(g1 + 1, ngeom):
            b1, b2 = m.geom_body[g1], m.geom_body[g2]
            if not ((m.geom_contype[g1] & m.geom_conaffinity[g2]) or (m.geom_contype[g2] & m.geom_conaffinity[g1])):
                continue
            w1, w2 = body_weld[b1], body_weld[b2]
            if w1 == w2:
                continue
            pw1, pw2 = body_weld[m.body_parent[w1]], body_weld[m.body_parent[w2]]
            if w1 != 0 and w2 != 0 and (w1 == pw2 or w2 == pw1):
                continue
            a, b = (g1, g2) if m.geom_type[g1] <= m.geom_type[g2] else (g2, g1)
            assert not (m.geom_type[a] == GEOM_PLANE and m.geom_type[b] == GEOM_PLANE)
            pairs.append((min(b1, b2), max(b1, b2), g1, g2, a, b))
    pairs.sort()

    def mix_pair(g1, g2):
        # equal priority: condim=max, friction=max, solref/solimp solmix-weighted, margin/gap=max
        s1, s2 = m.geom_solmix[g1], m.geom_solmix[g2]
        mix = s1 / (s1 + s2) if (s1 >= MJ_MINVAL and s2 >= MJ_MINVAL) else 0.5
        fr = np.maximum(m.geom_friction[g1], m.geom_friction[g2])
        solref = mix * np.asarray(m.geom_solref[g1]) + (1 - mix) * np.asarray(m.geom_solref[g2])
        solimp = mix * np.asarray(m.geom_solimp[g1]) + (1 - mix) * np.asarray(m.geom_solimp[g2])
        return dict(condim=max(m.geom_condim[g1], m.geom_condim[g2]), friction=fr, solref=solref, solimp=solimp,
                    margin=max(m.geom_margin[g1], m.geom_margin[g2]), gap=max(m.geom_gap[g1], m.geom_gap[g2]))

    pair_tab = []
    for (_, _, _, _, a, b) in pairs:
        p = mix_pair(a, b)
        assert p["condim"] in (1, 3)
        p.update(g1=a, g2=b)
        pair_tab.append(p)

    out = dict(
        name=m.name, nq=nq, nv=nv, nu=nu, nbody=nbody, njnt=njnt, ngeom=ngeom, npair=len(pair_tab),
        timestep=m.timestep, gravity=list(m.gravity), integrator=m.integrator, solver=m.solver,
        iterations=m.iterations, tolerance=m.tolerance,
        body_parent=list(m.body_parent), body_pos=np.array(m.body_pos), body_quat=np.array(m.body_quat),
        body_mass=body_mass, body_ipos=body_ipos, body_inertia=body_inertia,
        body_jntadr=list(m.body_jntadr), body_jntnum=list(m.body_jntnum),
        body_dofadr=body_dofadr, body_dofnum=body_dofnum, body_weld=body_weld, body_root=body_root,
        body_subtree=subtree_size, body_depth=body_depth, body_name=list(m.body_name),
        jnt_type=list(m.jnt_type), jnt_body=list(m.jnt_body), jnt_pos=np.array(m.jnt_pos),
        jnt_axis=np.array(m.jnt_axis), jnt_limited=[int(x) for x in m.jnt_limited],
        jnt_range=np.array(m.jnt_range), jnt_stiffness=np.array(m.jnt_stiffness),
        jnt_margin=np.array(m.jnt_margin), jnt_springref=np.array(m.jnt_springref),
        jnt_solref=np.array(m.jnt_solref), jnt_solimp=np.array(m.jnt_solimp),
        jnt_qposadr=jnt_qposadr, jnt_dofadr=jnt_dofadr, jnt_name=list(m.jnt_name),
        dof_jnt=dof_jnt, dof_body=dof_body, dof_parent=dof_parent,
        dof_armature=np.array([m.jnt_armature[j] for j in dof_jnt]),
        dof_damping=np.array([m.jnt_damping[j] for j in dof_jnt]),
        qpos0=qpos0,
        geom_type=list(m.geom_type), geom_body=list(m.geom_body), geom_size=np.array(m.geom_size),
        geom_pos=np.array(m.geom_pos), geom_quat=np.array(m.geom_quat), geom_name=list(m.geom_name),
        pair_g1=[p["g1"] for p in pair_tab], pair_g2=[p["g2"] for p in pair_tab],
        pair_condim=[p["condim"] for p in pair_tab],
        pair_friction=np.array([p["friction"] for p in pair_tab]).reshape(-1, 3),
        pair_solref=np.array([p["solref"] for p in pair_tab]).reshape(-1, 2),
        pair_solimp=np.array([p["solimp"] for p in pair_tab]).reshape(-1, 5),
        pair_margin=np.array([p["margin"] for p in pair_tab]), pair_gap=np.array([p["gap"] for p in pair_tab]),
        act_dof=[jnt_dofadr[j] for j in m.act_jnt], act_gear=np.array(m.act_gear),
        act_ctrllimited=[int(x) for x in m.act_ctrllimited], act_ctrlrange=np.array(m.act_ctrlrange),
    )
    for j in m.act_jnt:
        assert m.jnt_type[j] != JNT_FREE
    _set_const(out)
    return out


def kinematics_np(md: dict, qpos: np.ndarray):
    """World poses of bodies plus joint anchors/axes at ``qpos`` (numpy, used only at compile time)."""
    nbody = md["nbody"]
    xpos = np.zeros((nbody, 3))
    xquat = np.zeros((nbody, 4))
    xquat[0, 0] = 1.0
    xanchor = np.zeros((md["njnt"], 3))
    xaxis = np.zeros((md["njnt"], 3))
    for b in range(1, nbody):
        p = md["body_parent"][b]
        ja, jn = md["body_jntadr"][b], md["body_jntnum"][b]
        if jn == 1 and md["jnt_type"][ja] == JNT_FREE:
            a = md["jnt_qposadr"][ja]
            pos = qpos[a:a + 3].copy()
            quat = quat_normalize(qpos[a + 3:a + 7])
            xanchor[ja] = pos
            xaxis[ja] = (0, 0, 1)
        else:
            pmat = quat_to_mat(xquat[p])
            pos = xpos[p] + pmat @ md["body_pos"][b]
            quat = quat_mul(xquat[p], md["body_quat"][b])
            for j in range(ja, ja + jn):
                mat = quat_to_mat(quat)
                xanchor[j] = mat @ md["jnt_pos"][j] + pos
                xaxis[j] = mat @ md["jnt_axis"][j]
                q = qpos[md["jnt_qposadr"][j]] - md["qpos0"][md["jnt_qposadr"][j]]
                if md["jnt_type"][j] == JNT_SLIDE:
                    pos = pos + xaxis[j] * q
                else:
                    quat = quat_mul(quat, axisangle_to_quat(md["jnt_axis"][j], q))
                    # keep the anchor fixed in the world: correct for off-centre rotation
                    pos = xanchor[j] - quat_to_mat(quat) @ md["jnt_pos"][j]
        xpos[b], xquat[b] = pos, quat_normalize(quat)
    return xpos, xquat, xanchor, xaxis


def _jac_point(md, xanchor, xaxis, xquat, body, point):
    """3 x nv translational and rotational Jacobians of ``point`` fixed to ``body`` (world frame)."""
    nv = md["nv"]
    jp, jr = np.zeros((3, nv)), np.zeros((3, nv))
    b = body
    while b > 0:
        for j in range(md["body_jntadr"][b], md["body_jntadr"][b] + md["body_jntnum"][b]):
            d = md["jnt_dofadr"][j]
            t = md["jnt_type"][j]
            if t == JNT_FREE:
                jp[:, d:d + 3] = np.eye(3)
                mat = quat_to_mat(xquat[b])
                for k in range(3):
                    ax = mat[:, k]
                    jr[:, d + 3 + k] = ax
                    jp[:, d + 3 + k] = np.cross(ax, point - xanchor[j])
            elif t == JNT_SLIDE:
                jp[:, d] = xaxis[j]
            else:
                jr[:, d] = xaxis[j]
                jp[:, d] = np.cross(xaxis[j], point - xanchor[j])
        b = md["body_parent"][b]
    return jp, jr


def mass_matrix_np(md: dict, qpos: np.ndarray):
    xpos, xquat, xanchor, xaxis = kinematics_np(md, qpos)
    nv = md["nv"]
    M = np.diag(np.asarray(md["dof_armature"], dtype=np.float64)).copy()
    jacs = {}
    for b in range(1, md["nbody"]):
        mat = quat_to_mat(xquat[b])
        com = xpos[b] + mat @ md["body_ipos"][b]
        jp, jr = _jac_point(md, xanchor, xaxis, xquat, b, com)
        iw = mat @ md["body_inertia"][b] @ mat.T
        M += md["body_mass"][b] * jp.T @ jp + jr.T @ iw @ jr
        jacs[b] = (jp, jr)
    return M, jacs


def _set_const(md: dict):
    """dof_invweight0 / body_invweight0 / meaninertia at qpos0 (MuJoCo's mj_setConst rules)."""
    nv, nbody = md["nv"], md["nbody"]
    M, jacs = mass_matrix_np(md, md["qpos0"])
    Minv = np.linalg.inv(M)
    md["meaninertia"] = float(np.trace(M) / nv)
    dof_inv = np.diag(Minv).copy()
    for j in range(md["njnt"]):
        if md["jnt_type"][j] == JNT_FREE:
            d = md["jnt_dofadr"][j]
            dof_inv[d:d + 3] = dof_inv[d:d + 3].mean()
            dof_inv[d + 3:d + 6] = dof_inv[d + 3:d + 6].mean()
    md["dof_invweight0"] = dof_inv
    binv = np.zeros((nbody, 2))
    for b in range(1, nbody):
        jp, jr = jacs[b]
        J = np.vstack([jp, jr])
        A = J @ Minv @ J.T
        binv[b, 0] = (A[0, 0] + A[1, 1] + A[2, 2]) / 3.0
        binv[b, 1] = (A[3, 3] + A[4, 4] + A[5, 5]) / 3.0
    md["body_invweight0"] = binv


# ----------------------------------------------------------------------------- C header emission
_INT_FIELDS = ["body_parent", "body_jntadr", "body_jntnum", "body_dofadr", "body_dofnum", "body_weld", "body_root",
               "body_subtree", "body_depth", "jnt_type", "jnt_body", "jnt_limited", "jnt_qposadr", "jnt_dofadr",
               "dof_jnt", "dof_body", "dof_parent", "geom_type", "geom_body", "pair_g1", "pair_g2", "pair_condim",
               "act_dof", "act_ctrllimited"]
_DBL_FIELDS = ["body_pos", "body_quat", "body_mass", "body_ipos", "body_inertia", "body_invweight0", "jnt_pos",
               "jnt_axis", "jnt_range", "jnt_stiffness", "jnt_margin", "jnt_springref", "jnt_solref", "jnt_solimp",
               "dof_armature", "dof_damping", "dof_invweight0", "qpos0", "geom_size", "geom_pos", "geom_quat",
               "pair_friction", "pair_solref", "pair_solimp", "pair_margin", "pair_gap", "act_gear",
               "act_ctrlrange"]


def _c_array(a):
    a = np.asarray(a)
    if a.ndim == 1:
        if a.dtype.kind in "iub":
            return "{" + ", ".join(str(int(x)) for x in a) + "}"
        return "{" + ", ".join(repr(float(x)) for x in a) + "}"
    return "{" + ", ".join(_c_array(x) for x in a) + "}"


def emit_c(md: dict, symbol: str) -> str:
    lines = [f"static const pb_model {symbol} = {{"]
    lines.append(f'  .name = "{md["name"]}",')
    for k in ("nq", "nv", "nu", "nbody", "njnt", "ngeom", "npair", "integrator", "solver", "iterations"):
        lines.append(f"  .{k} = {int(md[k])},")
    for k in ("timestep", "tolerance", "meaninertia"):
        lines.append(f"  .{k} = {float(md[k])!r},")
    lines.append(f"  .gravity = {_c_array(md['gravity'])},")
    order = ["body_parent", "body_jntadr", "body_jntnum", "body_dofadr", "body_dofnum", "body_weld", "body_root",
             "body_subtree", "body_depth", "body_pos", "body_quat", "body_mass", "body_ipos", "body_inertia",
             "body_invweight0", "jnt_type", "jnt_body", "jnt_limited", "jnt_qposadr", "jnt_dofadr", "jnt_pos",
             "jnt_axis", "jnt_range", "jnt_stiffness", "jnt_margin", "jnt_springref", "jnt_solref", "jnt_solimp",
             "dof_jnt", "dof_body", "dof_parent", "dof_armature", "dof_damping", "dof_invweight0", "qpos0",
             "geom_type", "geom_body", "geom_size", "geom_pos", "geom_quat", "pair_g1", "pair_g2", "pair_condim",
             "pair_friction", "pair_solref", "pair_solimp", "pair_margin", "pair_gap", "act_dof", "act_ctrllimited",
             "act_gear", "act_ctrlrange"]  # pb_model declaration order (C++ designated initialisers require it)
    assert set(order) == set(_INT_FIELDS) | set(_DBL_FIELDS)
    for k in order:
        if k in _INT_FIELDS:
            if len(md[k]):
                lines.append(f"  .{k} = {_c_array(np.asarray(md[k], dtype=np.int64))},")
            continue
        a = np.asarray(md[k], dtype=np.float64)
        if a.size:
            if k == "body_inertia":
                a = a.reshape(len(a), 9)
            lines.append(f"  .{k} = {_c_array(a)},")
    lines.append("};")
    return "\n".join(lines)


ASSETS = [("half_cheetah.xml", "PB_MODEL_HALFCHEETAH"), ("hopper.xml", "PB_MODEL_HOPPER"),
          ("ant.xml", "PB_MODEL_ANT"), ("humanoid.xml", "PB_MODEL_HUMANOID")]


def main(argv):
    assets_dir, out = argv[1], argv[2]
    parts = ["/* GENERATED by ppo_cpp_b200/mjcf_compile.py from the MJCF task assets - do not edit.",
             " * Constants only (no reference code): kinematic tree, inertial frames, limits, collision pairs. */",
             "#ifndef PB_MODEL_TABLES_H", "#define PB_MODEL_TABLES_H", '#include "pb_model.h"', ""]
    for fname, sym in ASSETS:
        md = compile_model(parse_mjcf(f"{assets_dir}/{fname}"))
        parts.append(emit_c(md, sym))
        parts.append("")
    parts.append("#endif")
    with open(out, "w") as f:
        f.write("\n".join(parts) + "\n")


if __name__ == "__main__":
    main(sys.argv)
